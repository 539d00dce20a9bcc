// Multi-head self-attention for the imitator's anchor / mask branches (reference
// generator_component4_15.py:434-480, Anchor_selfattention: softmax(q k^T / sqrt(d)) v with d = 16, 4 heads,
// M = 4 anchors or M = N = 2048 points).  The reference materialises the (B,heads,M,M) score matrix (2.1 GB at
// B=32, M=2048) four times; here scores live in registers only (flash-style online softmax), in fp32 on the CUDA
// cores: with d = 16 the work per score is 32 FMAs + one exp2, i.e. the kernel is bound by fp32 issue and MUFU, not
// by a contraction the tensor cores could take over without dropping below fp32 accuracy.
//
// Layout: q, k, v, o, dO, dq, dk, dv are (B, M, H*16) token-major (what nn.Linear produces / consumes), so no
// permute/contiguous copies surround the call.  lse (B,H,M) is the log2-domain log-sum-exp of the scaled scores.
//
// forward : thread = query, keys/values streamed through shared memory tiles and read as broadcast LDS.128.
// backward: pass A (thread = query) recomputes p and accumulates dq, also emits delta = dO . O;
//           pass B (thread = key) recomputes p and accumulates dk, dv from tiles of (q, dO, lse, delta).
#include "apb_common.cuh"

namespace {

constexpr int D = 16;
constexpr int QT = 128;  // threads per CTA = rows owned by the CTA
constexpr int KT = 128;  // rows of the streamed operand per shared-memory tile
constexpr int KB = 8;    // scores formed per online-softmax block

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ void load_row(const float *p, float (&r)[D]) {
#pragma unroll
    for (int c = 0; c < D; c += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4 *>(p + c));
        r[c] = t.x; r[c + 1] = t.y; r[c + 2] = t.z; r[c + 3] = t.w;
    }
}
__device__ __forceinline__ void store_row(float *p, const float (&r)[D]) {
#pragma unroll
    for (int c = 0; c < D; c += 4) *reinterpret_cast<float4 *>(p + c) = make_float4(r[c], r[c + 1], r[c + 2], r[c + 3]);
}
__device__ __forceinline__ float dot_smem(const float (&a)[D], const float *s) {
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < D; c += 4) {
        const float4 t = *reinterpret_cast<const float4 *>(s + c);
        acc = fmaf(a[c], t.x, acc); acc = fmaf(a[c + 1], t.y, acc);
        acc = fmaf(a[c + 2], t.z, acc); acc = fmaf(a[c + 3], t.w, acc);
    }
    return acc;
}
__device__ __forceinline__ void axpy_smem(float (&acc)[D], float a, const float *s) {
#pragma unroll
    for (int c = 0; c < D; c += 4) {
        const float4 t = *reinterpret_cast<const float4 *>(s + c);
        acc[c] = fmaf(a, t.x, acc[c]); acc[c + 1] = fmaf(a, t.y, acc[c + 1]);
        acc[c + 2] = fmaf(a, t.z, acc[c + 2]); acc[c + 3] = fmaf(a, t.w, acc[c + 3]);
    }
}

// Stage rows [r0, r0+KT) of head h of a (B,M,H*D) tensor into dst[KT][D] (zero beyond M), scaled by `mul`.
__device__ __forceinline__ void stage_tile(float *dst, const float *src, int b, int h, int H, int M, int r0, float mul) {
    for (int t = threadIdx.x; t < KT * (D / 4); t += QT) {
        const int r = t >> 2, c = (t & 3) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r0 + r < M) {
            v = __ldg(reinterpret_cast<const float4 *>(src + (((size_t)b * M + r0 + r) * H + h) * D + c));
            v.x *= mul; v.y *= mul; v.z *= mul; v.w *= mul;
        }
        *reinterpret_cast<float4 *>(dst + r * D + c) = v;
    }
}

__global__ void __launch_bounds__(QT) attn_fwd_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                       const float *__restrict__ v, float *__restrict__ o,
                                                       float *__restrict__ lse, int M, int H, float scale2) {
    __shared__ __align__(16) float ks[KT * D];
    __shared__ __align__(16) float vs[KT * D];
    const int b = blockIdx.z, h = blockIdx.y;
    const int i = blockIdx.x * QT + threadIdx.x;
    const int ic = min(i, M - 1);
    float qr[D], acc[D];
    load_row(q + (((size_t)b * M + ic) * H + h) * D, qr);
#pragma unroll
    for (int c = 0; c < D; ++c) {
        qr[c] *= scale2;  // scores in the log2 domain: (q.k) / sqrt(d) * log2(e)
        acc[c] = 0.f;
    }
    float m = -INFINITY, l = 0.f;
    for (int j0 = 0; j0 < M; j0 += KT) {
        __syncthreads();
        stage_tile(ks, k, b, h, H, M, j0, 1.f);
        stage_tile(vs, v, b, h, H, M, j0, 1.f);
        __syncthreads();
        const int nk = min(KT, M - j0);
        for (int jb = 0; jb < nk; jb += KB) {
            float s[KB];
            float mb = -INFINITY;
#pragma unroll
            for (int u = 0; u < KB; ++u) {
                s[u] = jb + u < nk ? dot_smem(qr, ks + (jb + u) * D) : -INFINITY;
                mb = fmaxf(mb, s[u]);
            }
            const float mn = fmaxf(m, mb);
            const float corr = ex2(m - mn);  // first block: m = -inf, mn finite -> 0
            m = mn;
            l *= corr;
#pragma unroll
            for (int c = 0; c < D; ++c) acc[c] *= corr;
#pragma unroll
            for (int u = 0; u < KB; ++u) {
                const float p = ex2(s[u] - mn);
                l += p;
                axpy_smem(acc, p, vs + (jb + u) * D);
            }
        }
    }
    if (i < M) {
        const float inv = 1.f / l;
#pragma unroll
        for (int c = 0; c < D; ++c) acc[c] *= inv;
        store_row(o + (((size_t)b * M + i) * H + h) * D, acc);
        lse[((size_t)b * H + h) * M + i] = m + log2f(l);
    }
}

// pass A: dq_i = scale * sum_j p_ij (dO_i.v_j - delta_i) k_j ; delta_i = dO_i . O_i
__global__ void __launch_bounds__(QT) attn_bwd_dq_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                          const float *__restrict__ v, const float *__restrict__ o,
                                                          const float *__restrict__ dout, const float *__restrict__ lse,
                                                          float *__restrict__ dq, float *__restrict__ delta, int M,
                                                          int H, float scale, float scale2) {
    __shared__ __align__(16) float ks[KT * D];
    __shared__ __align__(16) float vs[KT * D];
    const int b = blockIdx.z, h = blockIdx.y;
    const int i = blockIdx.x * QT + threadIdx.x;
    const int ic = min(i, M - 1);
    const size_t row = (((size_t)b * M + ic) * H + h) * D;
    float qr[D], dor[D], acc[D];
    load_row(q + row, qr);
    load_row(dout + row, dor);
    float dl = 0.f;
    {
        float orow[D];
        load_row(o + row, orow);
#pragma unroll
        for (int c = 0; c < D; ++c) dl = fmaf(dor[c], orow[c], dl);
    }
#pragma unroll
    for (int c = 0; c < D; ++c) {
        qr[c] *= scale2;
        acc[c] = 0.f;
    }
    const float L = lse[((size_t)b * H + h) * M + ic];
    for (int j0 = 0; j0 < M; j0 += KT) {
        __syncthreads();
        stage_tile(ks, k, b, h, H, M, j0, 1.f);
        stage_tile(vs, v, b, h, H, M, j0, 1.f);
        __syncthreads();
        const int nk = min(KT, M - j0);
#pragma unroll 4
        for (int j = 0; j < nk; ++j) {
            const float p = ex2(dot_smem(qr, ks + j * D) - L);
            const float dp = dot_smem(dor, vs + j * D);
            axpy_smem(acc, p * (dp - dl), ks + j * D);
        }
    }
    if (i < M) {
#pragma unroll
        for (int c = 0; c < D; ++c) acc[c] *= scale;
        store_row(dq + row, acc);
        delta[((size_t)b * H + h) * M + i] = dl;
    }
}

// pass B: dv_j = sum_i p_ij dO_i ; dk_j = scale * sum_i p_ij (dO_i.v_j - delta_i) q_i
__global__ void __launch_bounds__(QT) attn_bwd_dkv_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                           const float *__restrict__ v, const float *__restrict__ dout,
                                                           const float *__restrict__ lse, const float *__restrict__ delta,
                                                           float *__restrict__ dk, float *__restrict__ dv, int M, int H,
                                                           float scale2) {
    __shared__ __align__(16) float qs[KT * D];    // q * scale2
    __shared__ __align__(16) float dos[KT * D];
    __shared__ __align__(8) float2 ld[KT];        // (lse, delta)
    const int b = blockIdx.z, h = blockIdx.y;
    const int j = blockIdx.x * QT + threadIdx.x;
    const int jc = min(j, M - 1);
    const size_t row = (((size_t)b * M + jc) * H + h) * D;
    float kr[D], vr[D], dkr[D], dvr[D];
    load_row(k + row, kr);
    load_row(v + row, vr);
#pragma unroll
    for (int c = 0; c < D; ++c) dkr[c] = dvr[c] = 0.f;
    for (int i0 = 0; i0 < M; i0 += KT) {
        __syncthreads();
        stage_tile(qs, q, b, h, H, M, i0, scale2);
        stage_tile(dos, dout, b, h, H, M, i0, 1.f);
        for (int t = threadIdx.x; t < KT; t += QT) {
            const bool ok = i0 + t < M;
            const size_t li = ((size_t)b * H + h) * M + i0 + t;
            ld[t] = ok ? make_float2(__ldg(lse + li), __ldg(delta + li)) : make_float2(INFINITY, 0.f);  // p = 0
        }
        __syncthreads();
        const int nq = min(KT, M - i0);
#pragma unroll 4
        for (int i = 0; i < nq; ++i) {
            const float2 Ld = ld[i];
            const float p = ex2(dot_smem(kr, qs + i * D) - Ld.x);
            const float dp = dot_smem(vr, dos + i * D);
            axpy_smem(dvr, p, dos + i * D);
            axpy_smem(dkr, p * (dp - Ld.y), qs + i * D);
        }
    }
    if (j < M) {
        const float ln2 = 0.6931471805599453f;  // qs carries scale*log2(e): dk = scale * sum ds q = ln2 * sum ds qs
#pragma unroll
        for (int c = 0; c < D; ++c) dkr[c] *= ln2;
        store_row(dk + row, dkr);
        store_row(dv + row, dvr);
    }
}

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

extern "C" int apb_attn_fwd_f32(const float *q, const float *k, const float *v, float *o, float *lse, int B, int M,
                                int H, int head_dim, apb_stream_t stream) {
    if (B < 0 || M < 0 || H < 0) return APB_E_INVALID;
    if (B == 0 || M == 0 || H == 0) return APB_OK;
    if (!q || !k || !v || !o || !lse) return APB_E_INVALID;
    if (head_dim != D || B > 65535 || H > 65535) return APB_E_UNSUPPORTED;
    if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(o)) return APB_E_INVALID;
    const float scale2 = 1.4426950408889634f / sqrtf((float)D);
    dim3 grid(apb_div_up(M, QT), H, B);
    attn_fwd_kernel<<<grid, QT, 0, (cudaStream_t)stream>>>(q, k, v, o, lse, M, H, scale2);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_attn_bwd_f32(const float *q, const float *k, const float *v, const float *o, const float *dout,
                                const float *lse, float *dq, float *dk, float *dv, float *delta, int B, int M, int H,
                                int head_dim, apb_stream_t stream) {
    if (B < 0 || M < 0 || H < 0) return APB_E_INVALID;
    if (B == 0 || M == 0 || H == 0) return APB_OK;
    if (!q || !k || !v || !o || !dout || !lse || !dq || !dk || !dv || !delta) return APB_E_INVALID;
    if (head_dim != D || B > 65535 || H > 65535) return APB_E_UNSUPPORTED;
    if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(o) || !aligned16(dout) || !aligned16(dq) ||
        !aligned16(dk) || !aligned16(dv))
        return APB_E_INVALID;
    const float scale = 1.f / sqrtf((float)D);
    const float scale2 = 1.4426950408889634f * scale;
    dim3 grid(apb_div_up(M, QT), H, B);
    cudaStream_t st = (cudaStream_t)stream;
    attn_bwd_dq_kernel<<<grid, QT, 0, st>>>(q, k, v, o, dout, lse, dq, delta, M, H, scale, scale2);
    APB_CHECK_LAUNCH();
    attn_bwd_dkv_kernel<<<grid, QT, 0, st>>>(q, k, v, dout, lse, delta, dk, dv, M, H, scale2);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
