// Neighbour searches for sm_100a: ball query, 3-NN and k-NN with tiled distance evaluation.
//
// Replaces ball_query_kernel_fast (reference ball_query_gpu.cu:15-51) and three_nn_kernel_fast
// (interpolate_gpu.cu:16-59), and torch.cdist+topk of KNN (models/layers/group.py:12-28).
// In the reference one thread walks all N support points with stride-12-byte global loads.  Here
// the support cloud is staged tile by tile in shared memory as float4 (coalesced global reads,
// one broadcast LDS.128 per support point per warp), each thread owns one query and evaluates the
// reference's exact fp32 expression (apb_sqdist), so indices are bit-identical.  Ball-query rows
// are assembled in shared memory and written back coalesced.
#include <cfloat>
#include "apb_common.cuh"

namespace {

constexpr int QT = 128;      // queries (threads) per CTA
constexpr int TILE = 2048;   // support points per shared-memory tile (32 KB as float4)

__device__ __forceinline__ void stage_tile(float4 *tile, const float *xyz, int base, int count) {
    // xyz is packed (x,y,z); 3*count consecutive floats are read coalesced and scattered into float4 slots
    const float *src = xyz + (size_t)base * 3;
    for (int i = threadIdx.x; i < count * 3; i += blockDim.x) {
        const float v = __ldg(src + i);
        const int pnt = i / 3, comp = i - pnt * 3;
        reinterpret_cast<float *>(tile + pnt)[comp] = v;
    }
}

// ---- ball query -------------------------------------------------------------------------------
// rows: [QT][K] ints in shared memory; thread t owns row t.
__global__ void __launch_bounds__(QT) ball_query_kernel(const float *__restrict__ new_xyz,
                                                         const float *__restrict__ xyz,
                                                         int32_t *__restrict__ idx, int N, int M, float radius,
                                                         int K) {
    extern __shared__ float4 smem_f4[];
    float4 *tile = smem_f4;
    int *rows = reinterpret_cast<int *>(tile + TILE);
    const int b = blockIdx.y;
    const int q = blockIdx.x * QT + threadIdx.x;
    const bool active = q < M;
    xyz += (size_t)b * N * 3;
    const float radius2 = __fmul_rn(radius, radius);  // ball_query_gpu.cu:29
    float qx = 0.f, qy = 0.f, qz = 0.f;
    if (active) {
        const float *qp = new_xyz + ((size_t)b * M + q) * 3;
        qx = qp[0]; qy = qp[1]; qz = qp[2];
    }
    int *row = rows + threadIdx.x * (K + 1);  // +1: odd stride spreads banks when K is even
    int cnt = active ? 0 : K;

    for (int base = 0; base < N; base += TILE) {
        const int count = min(TILE, N - base);
        __syncthreads();
        stage_tile(tile, xyz, base, count);
        __syncthreads();
        if (cnt < K) {
            for (int k0 = 0; k0 < count; k0 += 32) {
                const int kend = min(k0 + 32, count);
                for (int k = k0; k < kend; ++k) {
                    const float4 p = tile[k];
                    const float d2 = apb_sqdist(qx - p.x, qy - p.y, qz - p.z);
                    if (d2 < radius2 && cnt < K) {
                        row[cnt] = base + k;
                        ++cnt;
                    }
                }
                if (__all_sync(__activemask(), cnt >= K)) break;
            }
        }
        if (__syncthreads_and(cnt >= K)) break;
    }
    __syncthreads();
    // pad with the first hit (or 0 when the ball is empty: the reference's pre-zeroed buffer)
    if (active) {
        const int first = cnt > 0 ? row[0] : 0;
        for (int l = cnt; l < K; ++l) row[l] = first;
    }
    __syncthreads();
    // coalesced write-back of the CTA's QT x K block
    const int q0 = blockIdx.x * QT;
    const int nq = min(QT, M - q0);
    int32_t *out = idx + ((size_t)b * M + q0) * K;
    for (int i = threadIdx.x; i < nq * K; i += blockDim.x) {
        const int r = i / K, c = i - r * K;
        out[i] = rows[r * (K + 1) + c];
    }
}

// ---- three nearest neighbours -----------------------------------------------------------------
__global__ void __launch_bounds__(QT) three_nn_kernel(const float *__restrict__ unknown,
                                                       const float *__restrict__ known,
                                                       float *__restrict__ dist2, int32_t *__restrict__ idx,
                                                       int n, int m) {
    __shared__ float4 tile[TILE];
    const int b = blockIdx.y;
    const int q = blockIdx.x * QT + threadIdx.x;
    const bool active = q < n;
    known += (size_t)b * m * 3;
    float ux = 0.f, uy = 0.f, uz = 0.f;
    if (active) {
        const float *up = unknown + ((size_t)b * n + q) * 3;
        ux = up[0]; uy = up[1]; uz = up[2];
    }
    // The reference keeps the bests as double 1e40 and compares the float distance against them
    // with strict '<' (interpolate_gpu.cu:37-56).  float +inf is equivalent: a float d satisfies
    // d < 1e40 iff d < +inf, and (float)1e40 == +inf on the final store.
    float b1 = INFINITY, b2 = INFINITY, b3 = INFINITY;
    int i1 = 0, i2 = 0, i3 = 0;
    for (int base = 0; base < m; base += TILE) {
        const int count = min(TILE, m - base);
        __syncthreads();
        stage_tile(tile, known, base, count);
        __syncthreads();
        if (active) {
#pragma unroll 4
            for (int k = 0; k < count; ++k) {
                const float4 p = tile[k];
                const float d = apb_sqdist(ux - p.x, uy - p.y, uz - p.z);
                if (d < b3) {
                    const int kk = base + k;
                    if (d < b1) {
                        b3 = b2; i3 = i2; b2 = b1; i2 = i1; b1 = d; i1 = kk;
                    } else if (d < b2) {
                        b3 = b2; i3 = i2; b2 = d; i2 = kk;
                    } else {
                        b3 = d; i3 = kk;
                    }
                }
            }
        }
    }
    if (active) {
        float *od = dist2 + ((size_t)b * n + q) * 3;
        int32_t *oi = idx + ((size_t)b * n + q) * 3;
        od[0] = b1; od[1] = b2; od[2] = b3;
        oi[0] = i1; oi[1] = i2; oi[2] = i3;
    }
}

// ---- k nearest neighbours ---------------------------------------------------------------------
// One warp per query.  The current K best (distance, index) pairs are kept sorted across the warp,
// R = ceil(K/32) entries per lane (entry e lives in lane e%32, register e/32).  Each lane evaluates
// one candidate per step; candidates that beat the current worst are inserted one at a time with a
// ballot (insertion position) + shuffle-up (shift), so steady-state cost is the distance loop.
template <int R>
__global__ void __launch_bounds__(256) knn_kernel(const float *__restrict__ new_xyz,
                                                   const float *__restrict__ xyz, int32_t *__restrict__ idx,
                                                   float *__restrict__ dist2, int N, int M, int K) {
    __shared__ float4 tile[TILE];
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int q = blockIdx.x * 8 + warp;
    const bool active = q < M;
    xyz += (size_t)b * N * 3;
    float qx = 0.f, qy = 0.f, qz = 0.f;
    if (active) {
        const float *qp = new_xyz + ((size_t)b * M + q) * 3;
        qx = qp[0]; qy = qp[1]; qz = qp[2];
    }
    float bd[R];
    int bi[R];
#pragma unroll
    for (int r = 0; r < R; ++r) { bd[r] = INFINITY; bi[r] = 0x7fffffff; }
    // worst = entry K-1
    const int wl = (K - 1) & 31, wr = (K - 1) >> 5;
    float worst_d = INFINITY;
    int worst_i = 0x7fffffff;

    for (int base = 0; base < N; base += TILE) {
        const int count = min(TILE, N - base);
        __syncthreads();
        stage_tile(tile, xyz, base, count);
        __syncthreads();
        if (!active) continue;
        for (int k0 = 0; k0 < count; k0 += 32) {
            const int k = k0 + lane;
            float d = INFINITY;
            int ki = 0x7fffffff;
            if (k < count) {
                const float4 p = tile[k];
                d = apb_sqdist(qx - p.x, qy - p.y, qz - p.z);
                ki = base + k;
            }
            unsigned hits = __ballot_sync(0xffffffffu, d < worst_d || (d == worst_d && ki < worst_i));
            while (hits) {
                const int src = __ffs(hits) - 1;
                hits &= hits - 1;
                const float nd = __shfl_sync(0xffffffffu, d, src);
                const int ni = __shfl_sync(0xffffffffu, ki, src);
                if (!(nd < worst_d || (nd == worst_d && ni < worst_i))) continue;  // worst moved meanwhile
                float carry_d = nd;  // element entering register level r at its insertion point
                int carry_i = ni;
                bool inserted = false;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    // number of entries of this level that sort before the new one
                    const bool less = bd[r] < nd || (bd[r] == nd && bi[r] < ni);
                    const unsigned lm = __ballot_sync(0xffffffffu, less);
                    const int pos = __popc(lm);  // entries are sorted, so 'less' is a prefix
                    // value falling off the end of this level (lane 31) feeds the next level
                    const float out_d = __shfl_sync(0xffffffffu, bd[r], 31);
                    const int out_i = __shfl_sync(0xffffffffu, bi[r], 31);
                    const float up_d = __shfl_up_sync(0xffffffffu, bd[r], 1);
                    const int up_i = __shfl_up_sync(0xffffffffu, bi[r], 1);
                    if (inserted) {  // whole level shifts by one, carry enters at lane 0
                        bd[r] = lane == 0 ? carry_d : up_d;
                        bi[r] = lane == 0 ? carry_i : up_i;
                        carry_d = out_d; carry_i = out_i;
                    } else if (pos < 32) {
                        if (lane == pos) { bd[r] = nd; bi[r] = ni; }
                        else if (lane > pos) { bd[r] = up_d; bi[r] = up_i; }
                        carry_d = out_d; carry_i = out_i;
                        inserted = true;
                    }
                }
                worst_d = __shfl_sync(0xffffffffu, bd[wr], wl);
                worst_i = __shfl_sync(0xffffffffu, bi[wr], wl);
            }
        }
    }
    if (active) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int e = r * 32 + lane;
            if (e < K) {
                const bool filled = bi[r] != 0x7fffffff;
                idx[((size_t)b * M + q) * K + e] = filled ? bi[r] : 0;
                if (dist2) dist2[((size_t)b * M + q) * K + e] = bd[r];
            }
        }
    }
}

}  // namespace

extern "C" int apb_ball_query_f32(const float *new_xyz, const float *xyz, int32_t *idx, int B, int N, int M,
                                  float radius, int nsample, apb_stream_t stream) {
    if (B < 0 || N < 0 || M < 0 || nsample < 0) return APB_E_INVALID;
    if (B == 0 || M == 0 || nsample == 0) return APB_OK;
    if (!new_xyz || !idx || (N > 0 && !xyz)) return APB_E_INVALID;
    const size_t smem = sizeof(float4) * TILE + sizeof(int) * QT * (nsample + 1);
    if (smem > 227 * 1024) return APB_E_UNSUPPORTED;
    APB_CUDA(cudaFuncSetAttribute((const void *)ball_query_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
    dim3 grid(apb_div_up(M, QT), B);
    ball_query_kernel<<<grid, QT, smem, (cudaStream_t)stream>>>(new_xyz, xyz, idx, N, M, radius, nsample);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_three_nn_f32(const float *unknown, const float *known, float *dist2, int32_t *idx, int B,
                                int n, int m, apb_stream_t stream) {
    if (B < 0 || n < 0 || m < 0) return APB_E_INVALID;
    if (B == 0 || n == 0) return APB_OK;
    if (!unknown || !dist2 || !idx || (m > 0 && !known)) return APB_E_INVALID;
    dim3 grid(apb_div_up(n, QT), B);
    three_nn_kernel<<<grid, QT, 0, (cudaStream_t)stream>>>(unknown, known, dist2, idx, n, m);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_knn_f32(const float *new_xyz, const float *xyz, int32_t *idx, float *dist2, int B, int N,
                           int M, int K, apb_stream_t stream) {
    if (B < 0 || N < 0 || M < 0 || K < 0) return APB_E_INVALID;
    if (B == 0 || M == 0 || K == 0) return APB_OK;
    if (!new_xyz || !idx || (N > 0 && !xyz)) return APB_E_INVALID;
    if (K > 128) return APB_E_UNSUPPORTED;
    dim3 grid(apb_div_up(M, 8), B);
    cudaStream_t st = (cudaStream_t)stream;
    const int R = (K + 31) / 32;
    switch (R) {
        case 1: knn_kernel<1><<<grid, 256, 0, st>>>(new_xyz, xyz, idx, dist2, N, M, K); break;
        case 2: knn_kernel<2><<<grid, 256, 0, st>>>(new_xyz, xyz, idx, dist2, N, M, K); break;
        case 3: knn_kernel<3><<<grid, 256, 0, st>>>(new_xyz, xyz, idx, dist2, N, M, K); break;
        default: knn_kernel<4><<<grid, 256, 0, st>>>(new_xyz, xyz, idx, dist2, N, M, K); break;
    }
    APB_CHECK_LAUNCH();
    return APB_OK;
}
