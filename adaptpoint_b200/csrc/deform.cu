// Fused AdaptPoint / PointWOLF deformation for sm_100a: per-anchor local affine transform, kernel
// regression over the anchors, per-cloud re-centre + max-norm rescale, and the imitator's point mask.
//
// Replaces the ~25 small torch ops of AdaptPoint_Augmentor.forward lines 148-173 + local_transformaton
// (apply part, line 294) + kernel_regression (204-234) + normalize (313-327) of the reference's
// openpoints/models_adaptpoint/generator_component4_15.py, and the same code in
// openpoints/online_aug/pointwolf.py:27-178, all of which materialise (B,M,N,3) tensors several times.
//
// Per point p_n and anchor a_m (M anchors, M <= 8):
//   t_mn = (p_n - a_m) . A_m + trl_m + a_m          A_m = R_m S_m (3x3, row-vector convention)
//   w_mn = exp(-0.5 * |(a_m - p_n) * axis|^2 / sigma^2)
//   q_n  = sum_m w_mn t_mn / sum_m w_mn
//   out_n = mask_n * s * (q_n - mean_n q),   s = 0.999999 / max_n |q_n - mean q|
// One CTA per cloud; q is cached in shared memory when it fits (N <= 12288), otherwise recomputed in
// each of the three passes (sum -> max -> write).  HBM traffic is the algorithmic 12N read + 4N mask read +
// 12N write per cloud.  The backward kernel returns d/dA_m, d/dtrl_m (12 floats per anchor) and d/dmask;
// the tiny (B,M,9) parameter -> (R,S,trl) math stays in torch autograd.  xyz / anchors carry no gradient
// (they are data in the reference's training loop, train_autoaug.py:136-148).
#include <cfloat>
#include "apb_common.cuh"

namespace {

constexpr int DT = 512;        // threads per CTA
constexpr int MAXM = 8;        // anchors
constexpr int QCACHE_MAX = 12288;  // points whose q fits the shared-memory cache (144 KB)

struct DeformCtx {
    float a[MAXM][3];
    float A[MAXM][9];
    float trl[MAXM][3];
    float ax[3];
    float neg_half_inv_s2;
    int M;
};

__device__ __forceinline__ void load_ctx(DeformCtx *c, const float *anchor, const float *affine,
                                         const float *proj_axis, float sigma, int b, int M) {
    for (int i = threadIdx.x; i < M * 3; i += blockDim.x) c->a[i / 3][i % 3] = anchor[(size_t)b * M * 3 + i];
    for (int i = threadIdx.x; i < M * 12; i += blockDim.x) {
        const int m = i / 12, j = i % 12;
        const float v = affine[(size_t)b * M * 12 + i];
        if (j < 9) c->A[m][j] = v; else c->trl[m][j - 9] = v;
    }
    if (threadIdx.x < 3) c->ax[threadIdx.x] = proj_axis[(size_t)b * 3 + threadIdx.x];
    if (threadIdx.x == 0) {
        c->neg_half_inv_s2 = -0.5f / (sigma * sigma);
        c->M = M;
    }
}

// q_n and W_n = sum_m w_mn
__device__ __forceinline__ void eval_q(const DeformCtx *c, float px, float py, float pz, float &qx, float &qy,
                                       float &qz, float &W) {
    float sx = 0.f, sy = 0.f, sz = 0.f, sw = 0.f;
    for (int m = 0; m < c->M; ++m) {
        const float vx = px - c->a[m][0], vy = py - c->a[m][1], vz = pz - c->a[m][2];
        const float dx = vx * c->ax[0], dy = vy * c->ax[1], dz = vz * c->ax[2];
        const float w = expf((dx * dx + dy * dy + dz * dz) * c->neg_half_inv_s2);
        const float *A = c->A[m];
        const float tx = vx * A[0] + vy * A[3] + vz * A[6] + c->trl[m][0] + c->a[m][0];
        const float ty = vx * A[1] + vy * A[4] + vz * A[7] + c->trl[m][1] + c->a[m][1];
        const float tz = vx * A[2] + vy * A[5] + vz * A[8] + c->trl[m][2] + c->a[m][2];
        sx += w * tx; sy += w * ty; sz += w * tz; sw += w;
    }
    const float inv = 1.0f / sw;
    qx = sx * inv; qy = sy * inv; qz = sz * inv; W = sw;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide sum of NV values per thread; result valid in every thread (red: >= NV*32 floats of smem)
template <int NV>
__device__ __forceinline__ void block_sum(float (&v)[NV], float *red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
    __syncthreads();
    if (lane == 0)
#pragma unroll
        for (int i = 0; i < NV; ++i) red[i * 32 + warp] = v[i];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        float t = lane < nw ? red[i * 32 + lane] : 0.f;
        v[i] = warp_sum(t);
    }
}

// stats layout per cloud (8 floats): mean xyz, scale s, r* (max norm), n* (int bits), pad, pad
__global__ void __launch_bounds__(DT) deform_fwd_kernel(const float *__restrict__ xyz,
                                                         const float *__restrict__ anchor,
                                                         const float *__restrict__ affine,
                                                         const float *__restrict__ proj_axis,
                                                         const float *__restrict__ mask, float sigma,
                                                         float *__restrict__ out, float *__restrict__ stats, int N,
                                                         int M, int cache_q) {
    extern __shared__ float qc[];  // [3][N] when cache_q
    __shared__ DeformCtx ctx;
    __shared__ float red[3 * 32];
    __shared__ unsigned long long best;
    const int b = blockIdx.x;
    const float *p = xyz + (size_t)b * N * 3;
    load_ctx(&ctx, anchor, affine, proj_axis, sigma, b, M);
    if (threadIdx.x == 0) best = 0ull;
    __syncthreads();

    float s3[3] = {0.f, 0.f, 0.f};
    for (int n = threadIdx.x; n < N; n += DT) {
        float qx, qy, qz, W;
        eval_q(&ctx, p[n * 3], p[n * 3 + 1], p[n * 3 + 2], qx, qy, qz, W);
        if (cache_q) { qc[n] = qx; qc[N + n] = qy; qc[2 * N + n] = qz; }
        s3[0] += qx; s3[1] += qy; s3[2] += qz;
    }
    block_sum<3>(s3, red);
    const float invN = 1.0f / (float)N;
    const float mx = s3[0] * invN, my = s3[1] * invN, mz = s3[2] * invN;

    // max |q - mean| with the smallest index among equal maxima (torch.max picks one; ties are measure-zero)
    unsigned long long loc = 0ull;
    for (int n = threadIdx.x; n < N; n += DT) {
        float qx, qy, qz, W;
        if (cache_q) { qx = qc[n]; qy = qc[N + n]; qz = qc[2 * N + n]; }
        else eval_q(&ctx, p[n * 3], p[n * 3 + 1], p[n * 3 + 2], qx, qy, qz, W);
        const float cx = qx - mx, cy = qy - my, cz = qz - mz;
        const float r = sqrtf(cx * cx + cy * cy + cz * cz);
        const unsigned long long key = ((unsigned long long)__float_as_uint(r) << 32) | (unsigned)(0x7fffffff - n);
        loc = key > loc ? key : loc;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, loc, o);
        loc = other > loc ? other : loc;
    }
    if ((threadIdx.x & 31) == 0) atomicMax(&best, loc);
    __syncthreads();
    const float rmax = __uint_as_float((unsigned)(best >> 32));
    const int nstar = 0x7fffffff - (int)(unsigned)(best & 0xffffffffu);
    const float s = (1.0f / rmax) * 0.999999f;
    if (threadIdx.x == 0 && stats) {
        float *st = stats + (size_t)b * 8;
        st[0] = mx; st[1] = my; st[2] = mz; st[3] = s; st[4] = rmax; st[5] = __int_as_float(nstar);
        st[6] = 0.f; st[7] = 0.f;
    }
    float *o = out + (size_t)b * N * 3;
    const float *mk = mask ? mask + (size_t)b * N : nullptr;
    for (int n = threadIdx.x; n < N; n += DT) {
        float qx, qy, qz, W;
        if (cache_q) { qx = qc[n]; qy = qc[N + n]; qz = qc[2 * N + n]; }
        else eval_q(&ctx, p[n * 3], p[n * 3 + 1], p[n * 3 + 2], qx, qy, qz, W);
        const float f = mk ? mk[n] * s : s;
        o[n * 3 + 0] = f * (qx - mx);
        o[n * 3 + 1] = f * (qy - my);
        o[n * 3 + 2] = f * (qz - mz);
    }
}

__global__ void __launch_bounds__(DT) deform_bwd_kernel(const float *__restrict__ xyz,
                                                         const float *__restrict__ anchor,
                                                         const float *__restrict__ affine,
                                                         const float *__restrict__ proj_axis,
                                                         const float *__restrict__ mask, float sigma,
                                                         const float *__restrict__ stats,
                                                         const float *__restrict__ grad_out,
                                                         float *__restrict__ grad_affine,
                                                         float *__restrict__ grad_mask, int N, int M) {
    __shared__ DeformCtx ctx;
    __shared__ float red[12 * 32];
    const int b = blockIdx.x;
    const float *p = xyz + (size_t)b * N * 3;
    const float *g = grad_out + (size_t)b * N * 3;
    const float *mk = mask ? mask + (size_t)b * N : nullptr;
    load_ctx(&ctx, anchor, affine, proj_axis, sigma, b, M);
    __syncthreads();
    const float *st = stats + (size_t)b * 8;
    const float mx = st[0], my = st[1], mz = st[2], s = st[3], rmax = st[4];
    const int nstar = __float_as_int(st[5]);

    // pass 1: S1 = sum h.c, H = sum h  (h = mask * g), and d/dmask
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int n = threadIdx.x; n < N; n += DT) {
        float qx, qy, qz, W;
        eval_q(&ctx, p[n * 3], p[n * 3 + 1], p[n * 3 + 2], qx, qy, qz, W);
        const float cx = qx - mx, cy = qy - my, cz = qz - mz;
        const float gx = g[n * 3], gy = g[n * 3 + 1], gz = g[n * 3 + 2];
        const float gc = gx * cx + gy * cy + gz * cz;
        const float mv = mk ? mk[n] : 1.0f;
        if (grad_mask) grad_mask[(size_t)b * N + n] = s * gc;
        acc[0] += mv * gc; acc[1] += mv * gx; acc[2] += mv * gy; acc[3] += mv * gz;
    }
    block_sum<4>(acc, red);
    const float S1 = acc[0];
    float csx, csy, csz;
    {
        float W;
        eval_q(&ctx, p[nstar * 3], p[nstar * 3 + 1], p[nstar * 3 + 2], csx, csy, csz, W);
        csx -= mx; csy -= my; csz -= mz;
    }
    // d/dc at n* gets S1 * ds/dc = S1 * (-0.999999 / r*^3) c*
    const float coef = S1 * (-0.999999f / (rmax * rmax * rmax));
    const float ex = coef * csx, ey = coef * csy, ez = coef * csz;
    const float invN = 1.0f / (float)N;
    // mean over n of dc_n
    const float dmx = (s * acc[1] + ex) * invN, dmy = (s * acc[2] + ey) * invN, dmz = (s * acc[3] + ez) * invN;

    // pass 2: per anchor, 12 sums over the points
    for (int m = 0; m < M; ++m) {
        float a12[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) a12[i] = 0.f;
        for (int n = threadIdx.x; n < N; n += DT) {
            const float px = p[n * 3], py = p[n * 3 + 1], pz = p[n * 3 + 2];
            float wsum = 0.f, wm = 0.f;
            for (int k = 0; k < M; ++k) {
                const float dx = (px - ctx.a[k][0]) * ctx.ax[0], dy = (py - ctx.a[k][1]) * ctx.ax[1],
                            dz = (pz - ctx.a[k][2]) * ctx.ax[2];
                const float w = expf((dx * dx + dy * dy + dz * dz) * ctx.neg_half_inv_s2);
                wsum += w;
                wm = k == m ? w : wm;
            }
            const float mv = mk ? mk[n] : 1.0f;
            float dqx = s * mv * g[n * 3] - dmx, dqy = s * mv * g[n * 3 + 1] - dmy, dqz = s * mv * g[n * 3 + 2] - dmz;
            if (n == nstar) { dqx += ex; dqy += ey; dqz += ez; }
            const float wn = wm / wsum;
            dqx *= wn; dqy *= wn; dqz *= wn;
            const float vx = px - ctx.a[m][0], vy = py - ctx.a[m][1], vz = pz - ctx.a[m][2];
            a12[0] += vx * dqx; a12[1] += vx * dqy; a12[2] += vx * dqz;
            a12[3] += vy * dqx; a12[4] += vy * dqy; a12[5] += vy * dqz;
            a12[6] += vz * dqx; a12[7] += vz * dqy; a12[8] += vz * dqz;
            a12[9] += dqx; a12[10] += dqy; a12[11] += dqz;
        }
        block_sum<12>(a12, red);
        if (threadIdx.x < 12) {
            float v = 0.f;
#pragma unroll
            for (int i = 0; i < 12; ++i) v = threadIdx.x == i ? a12[i] : v;
            grad_affine[((size_t)b * M + m) * 12 + threadIdx.x] = v;
        }
    }
}

}  // namespace

extern "C" int apb_deform_mask_fwd_f32(const float *xyz, const float *anchor, const float *affine,
                                       const float *proj_axis, const float *mask, float sigma, float *out,
                                       float *stats, int B, int N, int M, apb_stream_t stream) {
    if (B < 0 || N < 0 || M < 0) return APB_E_INVALID;
    if (B == 0 || N == 0) return APB_OK;
    if (!xyz || !anchor || !affine || !proj_axis || !out || M == 0 || !(sigma > 0.f)) return APB_E_INVALID;
    if (M > MAXM) return APB_E_UNSUPPORTED;
    const int cache_q = N <= QCACHE_MAX;
    const size_t smem = cache_q ? (size_t)3 * N * sizeof(float) : 0;
    APB_CUDA(cudaFuncSetAttribute((const void *)deform_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)((size_t)3 * QCACHE_MAX * sizeof(float))));
    deform_fwd_kernel<<<B, DT, smem, (cudaStream_t)stream>>>(xyz, anchor, affine, proj_axis, mask, sigma, out, stats,
                                                            N, M, cache_q);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_deform_mask_bwd_f32(const float *xyz, const float *anchor, const float *affine,
                                       const float *proj_axis, const float *mask, float sigma, const float *stats,
                                       const float *grad_out, float *grad_affine, float *grad_mask, int B, int N,
                                       int M, apb_stream_t stream) {
    if (B < 0 || N < 0 || M < 0) return APB_E_INVALID;
    if (B == 0 || N == 0) return APB_OK;
    if (!xyz || !anchor || !affine || !proj_axis || !stats || !grad_out || !grad_affine || M == 0 || !(sigma > 0.f))
        return APB_E_INVALID;
    if (M > MAXM) return APB_E_UNSUPPORTED;
    deform_bwd_kernel<<<B, DT, 0, (cudaStream_t)stream>>>(xyz, anchor, affine, proj_axis, mask, sigma, stats,
                                                         grad_out, grad_affine, grad_mask, N, M);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
