// Training-mode set abstraction for ONE-conv blocks (LocalAggregation, and SetAbstraction with sa_layers = 1: every
// block of PointNeXt-B/L/XL) without the (B, C, M, K) tensors -- forward AND backward.
//
// Reference op chain (openpoints/models/backbone/pointnext.py:66-77, :164-166; layers/group.py:235-255, :323-325;
// layers/conv.py:24-60):  ball query -> group xyz, subtract query (/ radius) -> group features -> cat -> Conv2d 1x1
// -> BatchNorm2d (batch statistics) -> ReLU -> max over K, five (B,C,M,K) activations forward and five more backward.
//
// Algebra.  The conv is linear in its grouped input, so with  U = Wf . f  per SUPPORT point (B*N rows, a dense GEMM:
// tc_cm_gemm.cu, point-major store) the pre-BatchNorm activation of row r = (b, m, k) is
//        y[r, c] = U[b, n_r, c] + Wp[c, :] . dp_r ,      n_r = idx[b, m, k],   dp_r = p[n_r] - q[m]   (/ radius folded in Wp)
// and BatchNorm + ReLU are monotone per channel, so  max_k act(BN(y)) = act(BN(max_k y or min_k y)).
//   forward  (group_ystats_kernel): one pass over idx gathers rows of U, forms y on the fly and keeps, per (b, c, m),
//            max / min over k with their k, and per channel sum y, sum y^2 (the batch statistics).  Nothing of size
//            B*C*M*K is written.
//   backward: with batch statistics the BatchNorm gradient is dense in r:
//                dy[r, c] = gs[b, c, m] * [k == arg] + a[c] + bc[c] * y[r, c]
//            (gs = gamma*rstd*g after the ReLU mask; a, bc from the two channel sums of the BatchNorm backward).  The
//            first term touches one row per (b, c, m) (sa1_bwd_sparse_kernel).  The dense part needs no pass over r
//            either: summed over the rows that gathered support point n it is
//                dU[b, n, c] += cnt[n] * a[c] + bc[c] * (cnt[n] * U[b, n, c] + Wp[c, :] . sdp[n])
//            with cnt[n] = how often n occurs in idx[b] and sdp[n] = sum of dp over those rows (sa1_bwd_count_kernel, one
//            pass over idx with 4 atomics per entry; sa1_bwd_dense_kernel, one pass over the B*N*C rows of U), and
//                dWp[c, j] = sparse + a[c] * S1[j] + bc[c] * (sum_n U[n, c] * sdp[n, j] + Wp[c, :] . S2[:, j])
//            with S1 = sum_r dp, S2 = sum_r dp dp^T.  dgrad / wgrad of U = Wf . f then run on B*N rows instead of B*M*K
//            (32x fewer for LocalAggregation, 8x for a stride-4 SetAbstraction).
// In eval mode (running statistics: the feedback-loss pass of AdaptPoint) a = bc = 0 and only the sparse part remains;
// it also carries the gradient to the coordinates (p[n] += gs * Wp[c, :], q[m] -= ...).
#include <cuda_runtime.h>
#include <cfloat>
#include "apb_common.cuh"

namespace {

constexpr int GY_QB = 32;      // queries per CTA
constexpr int GY_CC = 128;     // channels per pass (4 per lane)

// vmax/vmin (B, C, M), arg (B, C, M) = kmax | kmin << 16 (all three nullptr: statistics only), stats (C, 2) f64 += sum y,
// sum y^2 (nullptr: skip)
__global__ void __launch_bounds__(256) group_ystats_kernel(const float *__restrict__ U, const float *__restrict__ xyz,
                                                           const float *__restrict__ new_xyz,
                                                           const int32_t *__restrict__ idx, const float *__restrict__ Wp,
                                                           float *__restrict__ vmax, float *__restrict__ vmin,
                                                           int32_t *__restrict__ arg, double *__restrict__ stats, int N,
                                                           int M, int K, int C) {
    extern __shared__ __align__(16) unsigned char gy_smem[];   // 3 staging planes [GY_CC][GY_QB + 1] + f64 sums [GY_CC][2]
    float (*st_max)[GY_QB + 1] = reinterpret_cast<float (*)[GY_QB + 1]>(gy_smem);
    float (*st_min)[GY_QB + 1] = st_max + GY_CC;
    int (*st_arg)[GY_QB + 1] = reinterpret_cast<int (*)[GY_QB + 1]>(st_min + GY_CC);
    double (*st_sum)[2] = reinterpret_cast<double (*)[2]>(st_arg + GY_CC);
    const int b = blockIdx.y;
    const int m0 = blockIdx.x * GY_QB;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float *Ub = U + (size_t)b * N * C;
    const int nq = min(GY_QB, M - m0);
    for (int cbase = 0; cbase < C; cbase += GY_CC) {
        const int c0 = cbase + lane * 4;
        const bool cok = c0 < C;   // C % 4 == 0
        if (threadIdx.x < GY_CC) {
            st_sum[threadIdx.x][0] = 0.0;
            st_sum[threadIdx.x][1] = 0.0;
        }
        __syncthreads();
        float4 wx = make_float4(0, 0, 0, 0), wy = wx, wz = wx;
        if (cok) {   // Wp (C, 3) row-major
            const float *w = Wp + (size_t)c0 * 3;
            wx = make_float4(w[0], w[3], w[6], w[9]);
            wy = make_float4(w[1], w[4], w[7], w[10]);
            wz = make_float4(w[2], w[5], w[8], w[11]);
        }
        float s1[4] = {0, 0, 0, 0}, s2[4] = {0, 0, 0, 0};
        for (int qi = warp; qi < nq; qi += 8) {
            const int m = m0 + qi;
            const float *qq = new_xyz + ((size_t)b * M + m) * 3;
            const float qx = __ldg(qq), qy = __ldg(qq + 1), qz = __ldg(qq + 2);
            const int32_t *ix = idx + ((size_t)b * M + m) * K;
            float mx[4] = {-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}, mn[4] = {FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX};
            int kx[4] = {0, 0, 0, 0}, kn[4] = {0, 0, 0, 0};
            for (int k = 0; k < K; ++k) {
                const int n = __ldg(ix + k);
                const float *pp = xyz + ((size_t)b * N + n) * 3;
                const float dx = __ldg(pp) - qx, dy = __ldg(pp + 1) - qy, dz = __ldg(pp + 2) - qz;
                if (cok) {
                    const float4 u = __ldg(reinterpret_cast<const float4 *>(Ub + (size_t)n * C + c0));
                    float y[4];
                    y[0] = fmaf(wz.x, dz, fmaf(wy.x, dy, fmaf(wx.x, dx, u.x)));
                    y[1] = fmaf(wz.y, dz, fmaf(wy.y, dy, fmaf(wx.y, dx, u.y)));
                    y[2] = fmaf(wz.z, dz, fmaf(wy.z, dy, fmaf(wx.z, dx, u.z)));
                    y[3] = fmaf(wz.w, dz, fmaf(wy.w, dy, fmaf(wx.w, dx, u.w)));
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        s1[e] += y[e];
                        s2[e] = fmaf(y[e], y[e], s2[e]);
                        if (y[e] > mx[e]) { mx[e] = y[e]; kx[e] = k; }   // strict: first k wins ties (torch.max order)
                        if (y[e] < mn[e]) { mn[e] = y[e]; kn[e] = k; }
                    }
                }
            }
            if (cok && vmax != nullptr) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    st_max[lane * 4 + e][qi] = mx[e];
                    st_min[lane * 4 + e][qi] = mn[e];
                    st_arg[lane * 4 + e][qi] = kx[e] | (kn[e] << 16);
                }
            }
        }
        if (stats != nullptr && cok) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                atomicAdd(&st_sum[lane * 4 + e][0], (double)s1[e]);
                atomicAdd(&st_sum[lane * 4 + e][1], (double)s2[e]);
            }
        }
        __syncthreads();
        const int cc = min(GY_CC, C - cbase);
        for (int i = threadIdx.x; vmax != nullptr && i < cc * GY_QB; i += 256) {
            const int c = i / GY_QB, qi = i - c * GY_QB;
            if (qi < nq) {
                const size_t o = ((size_t)b * C + cbase + c) * M + m0 + qi;
                vmax[o] = st_max[c][qi];
                vmin[o] = st_min[c][qi];
                arg[o] = st_arg[c][qi];
            }
        }
        if (stats != nullptr && threadIdx.x < cc) {
            atomicAdd(stats + 2 * (size_t)(cbase + threadIdx.x), st_sum[threadIdx.x][0]);
            atomicAdd(stats + 2 * (size_t)(cbase + threadIdx.x) + 1, st_sum[threadIdx.x][1]);
        }
        __syncthreads();
    }
}

// cnt (B, N) f32, sdp (B, N, 3) f32 pre-zeroed; mom (12) f64 pre-zeroed: S1[3], S2 row-major [9]
__global__ void __launch_bounds__(256) sa1_bwd_count_kernel(const float *__restrict__ xyz, const float *__restrict__ new_xyz,
                                                            const int32_t *__restrict__ idx, float *__restrict__ cnt,
                                                            float *__restrict__ sdp, double *__restrict__ mom, int N,
                                                            int M, int K) {
    __shared__ float red[12][8];
    const int b = blockIdx.y;
    const long long total = (long long)M * K;
    float acc[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[i] = 0.f;
    for (long long r = (long long)blockIdx.x * 256 + threadIdx.x; r < total; r += (long long)gridDim.x * 256) {
        const int m = (int)(r / K);
        const int n = __ldg(idx + (size_t)b * total + r);
        const float *pp = xyz + ((size_t)b * N + n) * 3;
        const float *qq = new_xyz + ((size_t)b * M + m) * 3;
        const float d0 = __ldg(pp) - __ldg(qq), d1 = __ldg(pp + 1) - __ldg(qq + 1), d2 = __ldg(pp + 2) - __ldg(qq + 2);
        atomicAdd(cnt + (size_t)b * N + n, 1.f);
        float *s = sdp + ((size_t)b * N + n) * 3;
        atomicAdd(s, d0);
        atomicAdd(s + 1, d1);
        atomicAdd(s + 2, d2);
        acc[0] += d0; acc[1] += d1; acc[2] += d2;
        acc[3] = fmaf(d0, d0, acc[3]); acc[4] = fmaf(d0, d1, acc[4]); acc[5] = fmaf(d0, d2, acc[5]);
        acc[6] = fmaf(d1, d0, acc[6]); acc[7] = fmaf(d1, d1, acc[7]); acc[8] = fmaf(d1, d2, acc[8]);
        acc[9] = fmaf(d2, d0, acc[9]); acc[10] = fmaf(d2, d1, acc[10]); acc[11] = fmaf(d2, d2, acc[11]);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 12; ++i) {
        float v = acc[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) red[i][warp] = v;
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        double v = 0.0;
        for (int w = 0; w < 8; ++w) v += (double)red[threadIdx.x][w];
        atomicAdd(mom + threadIdx.x, v);
    }
}

// One row per (b, c, m): gs (B, C, M) = gradient reaching the selected pre-BatchNorm value; ksel (B, C, M) its k.
// dU (B, C, N) channel-major pre-zeroed; dwp (C, 3) f64 pre-zeroed (or nullptr); dxyz (B, N, 3), dq (B, M, 3) pre-zeroed
// or nullptr (no coordinate gradient wanted).
__global__ void __launch_bounds__(256) sa1_bwd_sparse_kernel(const float *__restrict__ gs, const int32_t *__restrict__ ksel,
                                                             const int32_t *__restrict__ idx, const float *__restrict__ xyz,
                                                             const float *__restrict__ new_xyz,
                                                             const float *__restrict__ Wp, float *__restrict__ dU,
                                                             double *__restrict__ dwp, float *__restrict__ dxyz,
                                                             float *__restrict__ dq, int N, int M, int K, int C) {
    __shared__ float red[3][8];
    const int c = blockIdx.y, b = blockIdx.z;
    const float w0 = Wp[c * 3 + 0], w1 = Wp[c * 3 + 1], w2 = Wp[c * 3 + 2];
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
    for (int m = blockIdx.x * 256 + threadIdx.x; m < M; m += gridDim.x * 256) {
        const size_t o = ((size_t)b * C + c) * M + m;
        const float v = gs[o];
        if (v == 0.f) continue;
        const int n = __ldg(idx + ((size_t)b * M + m) * K + ksel[o]);
        atomicAdd(dU + ((size_t)b * C + c) * N + n, v);
        const float *pp = xyz + ((size_t)b * N + n) * 3;
        const float *qq = new_xyz + ((size_t)b * M + m) * 3;
        a0 = fmaf(v, __ldg(pp) - __ldg(qq), a0);
        a1 = fmaf(v, __ldg(pp + 1) - __ldg(qq + 1), a1);
        a2 = fmaf(v, __ldg(pp + 2) - __ldg(qq + 2), a2);
        if (dxyz != nullptr) {
            float *dp = dxyz + ((size_t)b * N + n) * 3;
            atomicAdd(dp, v * w0); atomicAdd(dp + 1, v * w1); atomicAdd(dp + 2, v * w2);
        }
        if (dq != nullptr) {
            float *dd = dq + ((size_t)b * M + m) * 3;
            atomicAdd(dd, -v * w0); atomicAdd(dd + 1, -v * w1); atomicAdd(dd + 2, -v * w2);
        }
    }
    if (dwp == nullptr) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a0 += __shfl_xor_sync(0xffffffffu, a0, o);
        a1 += __shfl_xor_sync(0xffffffffu, a1, o);
        a2 += __shfl_xor_sync(0xffffffffu, a2, o);
    }
    if (lane == 0) { red[0][warp] = a0; red[1][warp] = a1; red[2][warp] = a2; }
    __syncthreads();
    if (threadIdx.x < 3) {
        double v = 0.0;
        for (int w = 0; w < 8; ++w) v += (double)red[threadIdx.x][w];
        if (v != 0.0) atomicAdd(dwp + c * 3 + threadIdx.x, v);
    }
}

// dU[b, c, n] += cnt[n]*a[c] + bc[c]*(cnt[n]*U[b, n, c] + Wp[c, :].sdp[b, n]);  usdp (C, 3) f64 += sum_n U[b, n, c]*sdp[b, n, :]
// tile = 32 points x 32 channels: U is read point-major (lanes = channels), dU updated channel-major (lanes = points).
__global__ void __launch_bounds__(256) sa1_bwd_dense_kernel(const float *__restrict__ U, const float *__restrict__ cnt,
                                                            const float *__restrict__ sdp, const float *__restrict__ a,
                                                            const float *__restrict__ bc, const float *__restrict__ Wp,
                                                            float *__restrict__ dU, double *__restrict__ usdp, int N,
                                                            int C) {
    __shared__ float tile[32][33];
    __shared__ float s_cnt[32], s_sdp[32][3];
    const int b = blockIdx.z, n0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 8 rows of 32
    if (threadIdx.x < 32) {
        const int n = n0 + threadIdx.x;
        s_cnt[threadIdx.x] = n < N ? cnt[(size_t)b * N + n] : 0.f;
#pragma unroll
        for (int j = 0; j < 3; ++j) s_sdp[threadIdx.x][j] = n < N ? sdp[((size_t)b * N + n) * 3 + j] : 0.f;
    }
    __syncthreads();
    const int c = c0 + tx;
    const bool cok = c < C;
    const float ac = cok ? a[c] : 0.f, bcc = cok ? bc[c] : 0.f;
    const float w0 = cok ? Wp[c * 3] : 0.f, w1 = cok ? Wp[c * 3 + 1] : 0.f, w2 = cok ? Wp[c * 3 + 2] : 0.f;
    float u0 = 0.f, u1 = 0.f, u2 = 0.f;
    for (int r = ty; r < 32; r += 8) {
        const int n = n0 + r;
        float d = 0.f;
        if (n < N && cok) {
            const float u = __ldg(U + ((size_t)b * N + n) * C + c);
            const float cn = s_cnt[r];
            d = fmaf(cn, ac, bcc * fmaf(cn, u, fmaf(w0, s_sdp[r][0], fmaf(w1, s_sdp[r][1], w2 * s_sdp[r][2]))));
            u0 = fmaf(u, s_sdp[r][0], u0);
            u1 = fmaf(u, s_sdp[r][1], u1);
            u2 = fmaf(u, s_sdp[r][2], u2);
        }
        tile[r][tx] = d;
    }
    __syncthreads();
    for (int cc = ty; cc < 32; cc += 8) {
        const int n = n0 + tx;
        if (n < N && c0 + cc < C) dU[((size_t)b * C + c0 + cc) * N + n] += tile[tx][cc];
    }
    // per-channel partial of U^T sdp over this tile's points: reduce the 8 row groups through shared memory
    __syncthreads();
    float *red = &tile[0][0];   // [3][8][32] floats = 768 <= 32*33
    red[(0 * 8 + ty) * 32 + tx] = u0;
    red[(1 * 8 + ty) * 32 + tx] = u1;
    red[(2 * 8 + ty) * 32 + tx] = u2;
    __syncthreads();
    if (threadIdx.x < 96) {
        const int j = threadIdx.x >> 5, t = threadIdx.x & 31;
        double v = 0.0;
        for (int g = 0; g < 8; ++g) v += (double)red[(j * 8 + g) * 32 + t];
        if (c0 + t < C && v != 0.0) atomicAdd(usdp + (size_t)(c0 + t) * 3 + j, v);
    }
}

}  // namespace

extern "C" int apb_group_ystats_f32(const float *u_pm, const float *xyz, const float *new_xyz, const int32_t *idx,
                                    const float *wp, float *vmax, float *vmin, int32_t *arg, double *stats, int B,
                                    int N, int M, int K, int C, apb_stream_t stream) {
    if (B < 0 || N <= 0 || M < 0 || K <= 0 || C <= 0) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!u_pm || !xyz || !new_xyz || !idx || !wp) return APB_E_INVALID;
    if ((!vmax || !vmin || !arg) && (vmax || vmin || arg || !stats)) return APB_E_INVALID;   // statistics-only pass: all three NULL
    if ((C & 3) != 0 || K > 65535 || B > 65535 || (reinterpret_cast<uintptr_t>(u_pm) & 15) != 0) return APB_E_UNSUPPORTED;
    const size_t smem = (size_t)3 * GY_CC * (GY_QB + 1) * 4 + (size_t)GY_CC * 2 * 8;
    APB_CUDA(cudaFuncSetAttribute((const void *)group_ystats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    group_ystats_kernel<<<dim3(apb_div_up(M, GY_QB), B), 256, smem, (cudaStream_t)stream>>>(u_pm, xyz, new_xyz, idx, wp,
                                                                                            vmax, vmin, arg, stats, N, M,
                                                                                            K, C);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_sa1_bwd_count_f32(const float *xyz, const float *new_xyz, const int32_t *idx, float *cnt, float *sdp,
                                     double *mom, int B, int N, int M, int K, apb_stream_t stream) {
    if (B < 0 || N <= 0 || M < 0 || K <= 0) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!xyz || !new_xyz || !idx || !cnt || !sdp || !mom) return APB_E_INVALID;
    if (B > 65535) return APB_E_UNSUPPORTED;
    int gx = apb_div_up((long long)M * K, 256 * 4);
    if (gx > 4 * APB_NUM_SMS) gx = 4 * APB_NUM_SMS;
    sa1_bwd_count_kernel<<<dim3(gx, B), 256, 0, (cudaStream_t)stream>>>(xyz, new_xyz, idx, cnt, sdp, mom, N, M, K);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_sa1_bwd_sparse_f32(const float *gs, const int32_t *ksel, const int32_t *idx, const float *xyz,
                                      const float *new_xyz, const float *wp, float *du_cm, double *dwp, float *dxyz,
                                      float *dq, int B, int N, int M, int K, int C, apb_stream_t stream) {
    if (B < 0 || N <= 0 || M < 0 || K <= 0 || C <= 0) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!gs || !ksel || !idx || !xyz || !new_xyz || !wp || !du_cm) return APB_E_INVALID;
    if (B > 65535 || C > 65535) return APB_E_UNSUPPORTED;
    int gx = apb_div_up(M, 256);
    if (gx > 64) gx = 64;
    sa1_bwd_sparse_kernel<<<dim3(gx, C, B), 256, 0, (cudaStream_t)stream>>>(gs, ksel, idx, xyz, new_xyz, wp, du_cm, dwp,
                                                                             dxyz, dq, N, M, K, C);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_sa1_bwd_dense_f32(const float *u_pm, const float *cnt, const float *sdp, const float *a, const float *bc,
                                     const float *wp, float *du_cm, double *usdp, int B, int N, int C,
                                     apb_stream_t stream) {
    if (B < 0 || N <= 0 || C <= 0) return APB_E_INVALID;
    if (B == 0) return APB_OK;
    if (!u_pm || !cnt || !sdp || !a || !bc || !wp || !du_cm || !usdp) return APB_E_INVALID;
    if (B > 65535 || apb_div_up(C, 32) > 65535) return APB_E_UNSUPPORTED;
    sa1_bwd_dense_kernel<<<dim3(apb_div_up(N, 32), apb_div_up(C, 32), B), 256, 0, (cudaStream_t)stream>>>(
        u_pm, cnt, sdp, a, bc, wp, du_cm, usdp, N, C);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

// ---- batch statistics of y[r, c] = U[n_r, c] + Wp[c, :] . dp_r WITHOUT a pass over the rows r ------------------------------
// With cnt[n] / sdp[n] / S1 / S2 of apb_sa1_bwd_count_f32 (they depend on idx only):
//   sum_r y    = sum_n cnt[n] U[n, c]                                  + Wp[c, :] . S1
//   sum_r y^2  = sum_n (cnt[n] U[n, c]^2 + 2 U[n, c] Wp[c, :] . sdp[n]) + Wp[c, :] S2 Wp[c, :]^T
// one pass over the B*N*C values of U instead of B*M*K*C gathered ones (16x fewer for a stride-2 block with K = 32).
namespace {

__global__ void __launch_bounds__(256) sa_ustats_kernel(const float *__restrict__ U, const float *__restrict__ cnt,
                                                        const float *__restrict__ sdp, const float *__restrict__ Wp,
                                                        const double *__restrict__ mom, double *__restrict__ stats, int N,
                                                        int C) {
    __shared__ float red[2][8][32];
    __shared__ float s_cnt[32], s_sdp[32][3];
    const int b = blockIdx.z, n0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    if (threadIdx.x < 32) {
        const int n = n0 + threadIdx.x;
        s_cnt[threadIdx.x] = n < N ? cnt[(size_t)b * N + n] : 0.f;
#pragma unroll
        for (int j = 0; j < 3; ++j) s_sdp[threadIdx.x][j] = n < N ? sdp[((size_t)b * N + n) * 3 + j] : 0.f;
    }
    __syncthreads();
    const int c = c0 + tx;
    const bool cok = c < C;
    const float w0 = cok ? Wp[c * 3] : 0.f, w1 = cok ? Wp[c * 3 + 1] : 0.f, w2 = cok ? Wp[c * 3 + 2] : 0.f;
    float a = 0.f, q = 0.f;
    for (int r = ty; r < 32; r += 8) {
        const int n = n0 + r;
        if (n < N && cok) {
            const float u = __ldg(U + ((size_t)b * N + n) * C + c);
            const float cn = s_cnt[r];
            const float ws = fmaf(w0, s_sdp[r][0], fmaf(w1, s_sdp[r][1], w2 * s_sdp[r][2]));
            a = fmaf(cn, u, a);
            q = fmaf(u, fmaf(cn, u, 2.f * ws), q);
        }
    }
    red[0][ty][tx] = a;
    red[1][ty][tx] = q;
    __syncthreads();
    if (threadIdx.x < 64) {
        const int k = threadIdx.x >> 5, t = threadIdx.x & 31;
        double v = 0.0;
        for (int g = 0; g < 8; ++g) v += (double)red[k][g][t];
        const int cc = c0 + t;
        if (cc < C) {
            if (blockIdx.x == 0 && b == 0) {   // the terms that do not depend on U, once per channel
                const double x0 = Wp[cc * 3], x1 = Wp[cc * 3 + 1], x2 = Wp[cc * 3 + 2];
                if (k == 0) v += x0 * mom[0] + x1 * mom[1] + x2 * mom[2];
                else
                    v += x0 * (x0 * mom[3] + x1 * mom[4] + x2 * mom[5]) + x1 * (x0 * mom[6] + x1 * mom[7] + x2 * mom[8]) +
                         x2 * (x0 * mom[9] + x1 * mom[10] + x2 * mom[11]);
            }
            atomicAdd(stats + 2 * (size_t)cc + k, v);
        }
    }
}

}  // namespace

extern "C" int apb_sa_ustats_f32(const float *u_pm, const float *cnt, const float *sdp, const float *wp, const double *mom,
                                 double *stats, int B, int N, int C, apb_stream_t stream) {
    if (B < 0 || N <= 0 || C <= 0) return APB_E_INVALID;
    if (B == 0) return APB_OK;
    if (!u_pm || !cnt || !sdp || !wp || !mom || !stats) return APB_E_INVALID;
    if (B > 65535 || apb_div_up(C, 32) > 65535) return APB_E_UNSUPPORTED;
    sa_ustats_kernel<<<dim3(apb_div_up(N, 32), apb_div_up(C, 32), B), 256, 0, (cudaStream_t)stream>>>(u_pm, cnt, sdp, wp, mom,
                                                                                                       stats, N, C);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
