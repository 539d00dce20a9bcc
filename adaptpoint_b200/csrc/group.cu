// Index-driven copies for sm_100a: grouping / gather / three_interpolate, forward and backward.
//
// Replaces group_points_kernel_fast / group_points_grad_kernel_fast (reference
// group_points_gpu.cu:53-72 / 14-31), gather_points_* (sampling_gpu.cu:15-31 / 53-70) and
// three_interpolate_* (interpolate_gpu.cu:84-104 / 127-149).
//
// Forward grouping is write-dominated (K-fold expansion).  The reference launches one thread per
// output element, re-reads idx once per channel and issues 4-byte scattered loads that each pull a
// 32-byte sector.  Here a CTA stages CH feature rows (channel-major, N floats each) in shared
// memory with coalesced 16-byte loads, reads each int4 of idx ONCE for CH channels, gathers from
// shared memory (bank-conflict bound, no sector waste) and emits float4 streaming stores, so HBM
// sees the algorithmic bytes: C*N + M*K reads (idx re-reads hit L2) + C*M*K writes.
#include "apb_common.cuh"

namespace {

constexpr int GT = 256;  // threads per CTA

// out[b,c,p] = points[b,c,idx[b,p]]   p in [0,P) (P = M*K for grouping, m for gather)
// grid: (position chunks, channel groups, B); dynamic smem: CH*N floats.
template <int CH>
__global__ void __launch_bounds__(GT) group_rows_kernel(const float *__restrict__ points,
                                                         const int32_t *__restrict__ idx,
                                                         float *__restrict__ out, int C, int N, long long P,
                                                         int chunk) {
    extern __shared__ float rows[];  // [CH][N]
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CH;
    const int nch = min(CH, C - c0);
    const float *src = points + ((size_t)b * C + c0) * N;  // nch consecutive rows are contiguous
    const size_t total = (size_t)nch * N;
    if (((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
        const size_t nv = total >> 2;
        for (size_t i = threadIdx.x; i < nv; i += GT)
            reinterpret_cast<float4 *>(rows)[i] = __ldg(reinterpret_cast<const float4 *>(src) + i);
        for (size_t i = (nv << 2) + threadIdx.x; i < total; i += GT) rows[i] = __ldg(src + i);
    } else {
        for (size_t i = threadIdx.x; i < total; i += GT) rows[i] = __ldg(src + i);
    }
    __syncthreads();

    const int32_t *ix = idx + (size_t)b * P;
    float *o = out + ((size_t)b * C + c0) * P;
    const long long p_begin = (long long)blockIdx.x * chunk;
    const long long p_end = min(P, p_begin + chunk);
    const bool vec_ok = ((P & 3) == 0) && ((reinterpret_cast<uintptr_t>(ix) & 15) == 0) &&
                        ((reinterpret_cast<uintptr_t>(o) & 15) == 0);
    if (vec_ok) {  // chunk is a multiple of 4
        for (long long p = p_begin + (long long)threadIdx.x * 4; p < p_end; p += GT * 4) {
            const int4 i4 = apb_ld_nc_i4(ix + p);
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                if (ch < nch) {
                    const float *r = rows + (size_t)ch * N;
                    float4 v;
                    v.x = r[i4.x]; v.y = r[i4.y]; v.z = r[i4.z]; v.w = r[i4.w];
                    apb_st_stream_f4(o + (size_t)ch * P + p, v);
                }
            }
        }
    } else {
        for (long long p = p_begin + threadIdx.x; p < p_end; p += GT) {
            const int i = __ldg(ix + p);
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (ch < nch) o[(size_t)ch * P + p] = rows[(size_t)ch * N + i];
        }
    }
}

// Fallback when one row does not fit in shared memory (N > ~56k): direct gather through L1/L2.
__global__ void __launch_bounds__(GT) group_direct_kernel(const float *__restrict__ points,
                                                           const int32_t *__restrict__ idx,
                                                           float *__restrict__ out, int C, int N, long long P) {
    const int b = blockIdx.z, c = blockIdx.y;
    const float *r = points + ((size_t)b * C + c) * N;
    const int32_t *ix = idx + (size_t)b * P;
    float *o = out + ((size_t)b * C + c) * P;
    for (long long p = (long long)blockIdx.x * GT + threadIdx.x; p < P; p += (long long)gridDim.x * GT)
        o[p] = __ldg(r + __ldg(ix + p));
}

// grad_points[b,c,idx[b,p]] += grad_out[b,c,p]     (grad_points pre-zeroed by the caller)
// fp32 reductions at L2 (RED.ADD.F32), one idx read per CHB channels.
constexpr int CHB = 4;
__global__ void __launch_bounds__(GT) group_grad_kernel(const float *__restrict__ grad_out,
                                                         const int32_t *__restrict__ idx,
                                                         float *__restrict__ grad_points, int C, int N,
                                                         long long P) {
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CHB;
    const int nch = min(CHB, C - c0);
    const int32_t *ix = idx + (size_t)b * P;
    const float *g = grad_out + ((size_t)b * C + c0) * P;
    float *gp = grad_points + ((size_t)b * C + c0) * N;
    const bool vec_ok = ((P & 3) == 0) && ((reinterpret_cast<uintptr_t>(ix) & 15) == 0) &&
                        ((reinterpret_cast<uintptr_t>(g) & 15) == 0);
    if (vec_ok) {
        for (long long p = ((long long)blockIdx.x * GT + threadIdx.x) * 4; p < P; p += (long long)gridDim.x * GT * 4) {
            const int4 i4 = apb_ld_nc_i4(ix + p);
#pragma unroll
            for (int ch = 0; ch < CHB; ++ch) {
                if (ch < nch) {
                    const float4 v = apb_ld_nc_f4(g + (size_t)ch * P + p);
                    float *r = gp + (size_t)ch * N;
                    atomicAdd(r + i4.x, v.x);
                    atomicAdd(r + i4.y, v.y);
                    atomicAdd(r + i4.z, v.z);
                    atomicAdd(r + i4.w, v.w);
                }
            }
        }
    } else {
        for (long long p = (long long)blockIdx.x * GT + threadIdx.x; p < P; p += (long long)gridDim.x * GT) {
            const int i = __ldg(ix + p);
            for (int ch = 0; ch < nch; ++ch) atomicAdd(gp + (size_t)ch * N + i, __ldg(g + (size_t)ch * P + p));
        }
    }
}

// out[b,c,i] = w1*p[i1] (+) fma(w0,p[i0]) (+) fma(w2,p[i2])   -- the reference's contracted order
// (interpolate_gpu.cu:103 as compiled: FMUL on term 1, FFMA term 0, FFMA term 2).
// grid: (point chunks, channel groups, B); smem: CH rows of m floats.
template <int CH>
__global__ void __launch_bounds__(GT) three_interp_kernel(const float *__restrict__ points,
                                                           const int32_t *__restrict__ idx,
                                                           const float *__restrict__ weight,
                                                           float *__restrict__ out, int C, int m, int n) {
    extern __shared__ float rows[];
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CH;
    const int nch = min(CH, C - c0);
    const float *src = points + ((size_t)b * C + c0) * m;
    for (size_t i = threadIdx.x; i < (size_t)nch * m; i += GT) rows[i] = __ldg(src + i);
    __syncthreads();
    const int32_t *ix = idx + (size_t)b * n * 3;
    const float *w = weight + (size_t)b * n * 3;
    float *o = out + ((size_t)b * C + c0) * n;
    for (int i = blockIdx.x * GT + threadIdx.x; i < n; i += gridDim.x * GT) {
        const int i0 = __ldg(ix + i * 3 + 0), i1 = __ldg(ix + i * 3 + 1), i2 = __ldg(ix + i * 3 + 2);
        const float w0 = __ldg(w + i * 3 + 0), w1 = __ldg(w + i * 3 + 1), w2 = __ldg(w + i * 3 + 2);
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) {
            if (ch < nch) {
                const float *r = rows + (size_t)ch * m;
                float t = __fmul_rn(w1, r[i1]);
                t = __fmaf_rn(w0, r[i0], t);
                t = __fmaf_rn(w2, r[i2], t);
                o[(size_t)ch * n + i] = t;
            }
        }
    }
}

__global__ void __launch_bounds__(GT) three_interp_direct_kernel(const float *__restrict__ points,
                                                                  const int32_t *__restrict__ idx,
                                                                  const float *__restrict__ weight,
                                                                  float *__restrict__ out, int C, int m, int n) {
    const int b = blockIdx.z, c = blockIdx.y;
    const float *r = points + ((size_t)b * C + c) * m;
    const int32_t *ix = idx + (size_t)b * n * 3;
    const float *w = weight + (size_t)b * n * 3;
    float *o = out + ((size_t)b * C + c) * n;
    for (int i = blockIdx.x * GT + threadIdx.x; i < n; i += gridDim.x * GT) {
        float t = __fmul_rn(__ldg(w + i * 3 + 1), __ldg(r + __ldg(ix + i * 3 + 1)));
        t = __fmaf_rn(__ldg(w + i * 3 + 0), __ldg(r + __ldg(ix + i * 3 + 0)), t);
        t = __fmaf_rn(__ldg(w + i * 3 + 2), __ldg(r + __ldg(ix + i * 3 + 2)), t);
        o[i] = t;
    }
}

// grad_points[b,c,idx[b,i,j]] += grad_out[b,c,i] * weight[b,i,j]   (pre-zeroed accumulator)
__global__ void __launch_bounds__(GT) three_interp_grad_kernel(const float *__restrict__ grad_out,
                                                                const int32_t *__restrict__ idx,
                                                                const float *__restrict__ weight,
                                                                float *__restrict__ grad_points, int C, int n,
                                                                int m) {
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CHB;
    const int nch = min(CHB, C - c0);
    const int32_t *ix = idx + (size_t)b * n * 3;
    const float *w = weight + (size_t)b * n * 3;
    const float *g = grad_out + ((size_t)b * C + c0) * n;
    float *gp = grad_points + ((size_t)b * C + c0) * m;
    for (int i = blockIdx.x * GT + threadIdx.x; i < n; i += gridDim.x * GT) {
        const int i0 = __ldg(ix + i * 3 + 0), i1 = __ldg(ix + i * 3 + 1), i2 = __ldg(ix + i * 3 + 2);
        const float w0 = __ldg(w + i * 3 + 0), w1 = __ldg(w + i * 3 + 1), w2 = __ldg(w + i * 3 + 2);
        for (int ch = 0; ch < nch; ++ch) {
            const float gv = __ldg(g + (size_t)ch * n + i);
            float *r = gp + (size_t)ch * m;
            atomicAdd(r + i0, __fmul_rn(gv, w0));
            atomicAdd(r + i1, __fmul_rn(gv, w1));
            atomicAdd(r + i2, __fmul_rn(gv, w2));
        }
    }
}

constexpr size_t kSmemBudget = 200 * 1024;

template <int CH>
int launch_group_rows(const float *points, const int32_t *idx, float *out, int B, int C, int N, long long P,
                      cudaStream_t st) {
    const size_t smem = (size_t)CH * N * sizeof(float);
    APB_CUDA(cudaFuncSetAttribute((const void *)group_rows_kernel<CH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
    // positions per CTA: enough output per staged row (>= 8x the row bytes) but keep the grid wide
    long long chunk = (long long)8 * N;
    if (chunk < 4096) chunk = 4096;
    const int cgroups = apb_div_up(C, CH);
    while (chunk > 4096 && (long long)apb_div_up(P, chunk) * cgroups * B < 4 * APB_NUM_SMS) chunk >>= 1;
    chunk = (chunk + 1023) / 1024 * 1024;  // multiple of GT*4
    dim3 grid(apb_div_up(P, chunk), cgroups, B);
    group_rows_kernel<CH><<<grid, GT, smem, st>>>(points, idx, out, C, N, P, (int)chunk);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

int group_forward(const float *points, const int32_t *idx, float *out, int B, int C, int N, long long P,
                  cudaStream_t st) {
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!points || !idx || !out || N <= 0) return APB_E_INVALID;
    if (C > 65535 || B > 65535) return APB_E_UNSUPPORTED;
    const size_t row = (size_t)N * sizeof(float);
    if (row > kSmemBudget) {
        dim3 grid((unsigned)min((long long)apb_div_up(P, GT), (long long)8 * APB_NUM_SMS), C, B);
        group_direct_kernel<<<grid, GT, 0, st>>>(points, idx, out, C, N, P);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    const int fit = (int)(kSmemBudget / row);
    if (fit >= 4 && C >= 4) return launch_group_rows<4>(points, idx, out, B, C, N, P, st);
    if (fit >= 2 && C >= 2) return launch_group_rows<2>(points, idx, out, B, C, N, P, st);
    return launch_group_rows<1>(points, idx, out, B, C, N, P, st);
}

int group_backward(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C, int N,
                   long long P, cudaStream_t st) {
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!grad_out || !idx || !grad_points || N <= 0) return APB_E_INVALID;
    if (C > 65535 * CHB || B > 65535) return APB_E_UNSUPPORTED;
    dim3 grid((unsigned)min((long long)apb_div_up(P, GT * 4), (long long)4 * APB_NUM_SMS), apb_div_up(C, CHB), B);
    group_grad_kernel<<<grid, GT, 0, st>>>(grad_out, idx, grad_points, C, N, P);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

}  // namespace

extern "C" int apb_group_f32(const float *points, const int32_t *idx, float *out, int B, int C, int N, int M,
                             int K, apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || M < 0 || K < 0) return APB_E_INVALID;
    return group_forward(points, idx, out, B, C, N, (long long)M * K, (cudaStream_t)stream);
}
extern "C" int apb_group_grad_f32(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C,
                                  int N, int M, int K, apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || M < 0 || K < 0) return APB_E_INVALID;
    return group_backward(grad_out, idx, grad_points, B, C, N, (long long)M * K, (cudaStream_t)stream);
}
extern "C" int apb_gather_f32(const float *points, const int32_t *idx, float *out, int B, int C, int N, int m,
                              apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || m < 0) return APB_E_INVALID;
    return group_forward(points, idx, out, B, C, N, (long long)m, (cudaStream_t)stream);
}
extern "C" int apb_gather_grad_f32(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C,
                                   int N, int m, apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || m < 0) return APB_E_INVALID;
    return group_backward(grad_out, idx, grad_points, B, C, N, (long long)m, (cudaStream_t)stream);
}

extern "C" int apb_three_interp_f32(const float *points, const int32_t *idx, const float *weight, float *out,
                                    int B, int C, int m, int n, apb_stream_t stream) {
    if (B < 0 || C < 0 || m < 0 || n < 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || n == 0) return APB_OK;
    if (!points || !idx || !weight || !out || m <= 0) return APB_E_INVALID;
    if (B > 65535) return APB_E_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)m * sizeof(float);
    const int xblocks = min(apb_div_up(n, GT), 4 * APB_NUM_SMS);
    if (row * 4 <= kSmemBudget && C >= 4) {
        const size_t smem = row * 4;
        APB_CUDA(cudaFuncSetAttribute((const void *)three_interp_kernel<4>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dim3 grid(min(xblocks, max(1, apb_div_up(n, 8 * GT))), apb_div_up(C, 4), B);
        three_interp_kernel<4><<<grid, GT, smem, st>>>(points, idx, weight, out, C, m, n);
    } else if (row <= kSmemBudget) {
        APB_CUDA(cudaFuncSetAttribute((const void *)three_interp_kernel<1>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)row));
        dim3 grid(min(xblocks, max(1, apb_div_up(n, 8 * GT))), C, B);
        three_interp_kernel<1><<<grid, GT, row, st>>>(points, idx, weight, out, C, m, n);
    } else {
        dim3 grid(xblocks, C, B);
        three_interp_direct_kernel<<<grid, GT, 0, st>>>(points, idx, weight, out, C, m, n);
    }
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_three_interp_grad_f32(const float *grad_out, const int32_t *idx, const float *weight,
                                         float *grad_points, int B, int C, int n, int m, apb_stream_t stream) {
    if (B < 0 || C < 0 || m < 0 || n < 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || n == 0) return APB_OK;
    if (!grad_out || !idx || !weight || !grad_points || m <= 0) return APB_E_INVALID;
    if (B > 65535) return APB_E_UNSUPPORTED;
    dim3 grid(min(apb_div_up(n, GT), 4 * APB_NUM_SMS), apb_div_up(C, CHB), B);
    three_interp_grad_kernel<<<grid, GT, 0, (cudaStream_t)stream>>>(grad_out, idx, weight, grad_points, C, n, m);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
