// Spectral normalisation of the discriminator's weights in 3 launches forward / 2 backward.
//
// Reference: PointDiscriminator1 wraps its three 1x1 convs and four linears in torch spectral_norm
// (openpoints/models_adaptpoint/point_discriminator.py:23-34, 149-191).  In training mode every forward runs one power
// iteration per layer:  u <- normalize(W v);  v <- normalize(W^T u);  sigma = u^T W v;  W_sn = W / sigma, which torch
// issues as ~14 small kernels (mv, norm, clamp, div, clone, dot, div) and ~8 more in the backward pass.  With 7 layers and
// 3 discriminator passes per AdaptPoint iteration that is ~400 launches per step for ~10 MFLOP of work.
//   sn_mv_kernel     t = W v, accumulates |t|^2                                   (warp per row)
//   sn_mtv_kernel    u = t / max(|t|, eps) (written back), s = W^T u, accumulates |s|^2   (block per 32 columns)
//   sn_finish_kernel v = s / max(|s|, eps) (written back), sigma = |s|, W_sn = W / sigma
// (u^T W v with v = W^T u / |W^T u| is |W^T u| = |s|, so the third matrix pass of the reference is not needed.)
// Backward, with u, v constants as in torch:  dW = (G - <G, W_sn> u v^T) / sigma.
#include "apb_common.cuh"

namespace {

__device__ __forceinline__ float sn_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// scratch: [0] = |t|^2, [1] = |s|^2, [2] = <G, W_sn> (backward), [4 .. 4+w) = s
__global__ void __launch_bounds__(256) sn_mv_kernel(const float *__restrict__ W, const float *__restrict__ v,
                                                    float *__restrict__ t, float *__restrict__ scratch, int h, int w) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * 8 + warp;
    float acc = 0.f;
    if (i < h) {
        const float *row = W + (size_t)i * w;
        for (int j = lane; j < w; j += 32) acc = fmaf(__ldg(row + j), __ldg(v + j), acc);
    }
    acc = sn_warp_sum(acc);
    if (i < h && lane == 0) {
        t[i] = acc;
        atomicAdd(scratch, acc * acc);
    }
}

__global__ void __launch_bounds__(256) sn_mtv_kernel(const float *__restrict__ W, const float *__restrict__ t,
                                                     float *__restrict__ u, float *__restrict__ u_saved,
                                                     float *__restrict__ scratch, float eps, int h, int w) {
    __shared__ float red[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + tx;
    const float rn = 1.f / fmaxf(sqrtf(scratch[0]), eps);
    float acc = 0.f;
    for (int i = ty; i < h; i += 8) {
        const float ui = __ldg(t + i) * rn;
        if (j < w) acc = fmaf(__ldg(W + (size_t)i * w + j), ui, acc);
        if (blockIdx.x == 0 && tx == 0) {
            u[i] = ui;
            u_saved[i] = ui;
        }
    }
    red[ty][tx] = acc;
    __syncthreads();
    if (ty == 0) {
        float s = 0.f;
#pragma unroll
        for (int g = 0; g < 8; ++g) s += red[g][tx];
        if (j < w) scratch[4 + j] = s;
        const float sq = sn_warp_sum(j < w ? s * s : 0.f);
        if (tx == 0) atomicAdd(scratch + 1, sq);
    }
}

__global__ void __launch_bounds__(256) sn_finish_kernel(const float *__restrict__ W, const float *__restrict__ scratch,
                                                        float *__restrict__ v, float *__restrict__ v_saved,
                                                        float *__restrict__ sigma_out, float *__restrict__ out, float eps,
                                                        long long n, int w) {
    const float norm_s = sqrtf(scratch[1]);
    const float rn = 1.f / fmaxf(norm_s, eps);
    // sigma = u^T W v = |s|^2 / max(|s|, eps)
    const float sigma = scratch[1] * rn;
    const float inv = 1.f / sigma;
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid < w) {
        const float vj = scratch[4 + gid] * rn;
        v[gid] = vj;
        v_saved[gid] = vj;
    }
    if (gid == 0) sigma_out[0] = sigma;
    for (long long i = gid; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = __ldg(W + i) * inv;
}

__global__ void __launch_bounds__(256) sn_bwd_dot_kernel(const float *__restrict__ G, const float *__restrict__ Wsn,
                                                         float *__restrict__ scratch, long long n) {
    __shared__ float red[8];
    float acc = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        acc = fmaf(__ldg(G + i), __ldg(Wsn + i), acc);
    acc = sn_warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
#pragma unroll
        for (int g = 0; g < 8; ++g) s += red[g];
        atomicAdd(scratch + 2, s);
    }
}

__global__ void __launch_bounds__(256) sn_bwd_kernel(const float *__restrict__ G, const float *__restrict__ u,
                                                     const float *__restrict__ v, const float *__restrict__ sigma,
                                                     const float *__restrict__ scratch, float *__restrict__ dW,
                                                     long long n, int w) {
    const float inv = 1.f / sigma[0], dot = scratch[2];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / w), c = (int)(i - (long long)r * w);
        dW[i] = (__ldg(G + i) - dot * __ldg(u + r) * __ldg(v + c)) * inv;
    }
}

}  // namespace

extern "C" int apb_spectral_norm_fwd_f32(const float *w_mat, float *u, float *v, float *t_ws, float *scratch,
                                         float *u_saved, float *v_saved, float *sigma, float *w_sn, float eps, int h,
                                         int w, apb_stream_t stream) {
    if (h <= 0 || w <= 0) return APB_E_INVALID;
    if (!w_mat || !u || !v || !t_ws || !scratch || !u_saved || !v_saved || !sigma || !w_sn) return APB_E_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const long long n = (long long)h * w;
    sn_mv_kernel<<<apb_div_up(h, 8), 256, 0, st>>>(w_mat, v, t_ws, scratch, h, w);
    APB_CHECK_LAUNCH();
    sn_mtv_kernel<<<apb_div_up(w, 32), 256, 0, st>>>(w_mat, t_ws, u, u_saved, scratch, eps, h, w);
    APB_CHECK_LAUNCH();
    int grid = apb_div_up(n, 256 * 4);
    if (grid < apb_div_up(w, 256)) grid = apb_div_up(w, 256);
    if (grid > 2 * APB_NUM_SMS) grid = 2 * APB_NUM_SMS > apb_div_up(w, 256) ? 2 * APB_NUM_SMS : apb_div_up(w, 256);
    sn_finish_kernel<<<grid, 256, 0, st>>>(w_mat, scratch, v, v_saved, sigma, w_sn, eps, n, w);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_spectral_norm_bwd_f32(const float *grad, const float *w_sn, const float *u_saved, const float *v_saved,
                                         const float *sigma, float *scratch, float *grad_w, int h, int w,
                                         apb_stream_t stream) {
    if (h <= 0 || w <= 0) return APB_E_INVALID;
    if (!grad || !w_sn || !u_saved || !v_saved || !sigma || !scratch || !grad_w) return APB_E_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    const long long n = (long long)h * w;
    int grid = apb_div_up(n, 256 * 4);
    if (grid > 2 * APB_NUM_SMS) grid = 2 * APB_NUM_SMS;
    sn_bwd_dot_kernel<<<grid, 256, 0, st>>>(grad, w_sn, scratch, n);
    APB_CHECK_LAUNCH();
    sn_bwd_kernel<<<grid, 256, 0, st>>>(grad, u_saved, v_saved, sigma, scratch, grad_w, n, w);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
