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

// Packed fp32 pairs: on sm_100 a three-register FFMA issues every other cycle per scheduler (register-port bound), an
// FFMA2 (fma.rn.f32x2, two independent fp32 FMAs on a 64-bit register pair) issues at the same rate, so the fp32 FMA
// peak is only reachable through FFMA2.  A 16-float row is held as 8 pairs; results are ordinary IEEE fp32 FMAs.
typedef unsigned long long f2;
constexpr int D2 = D / 2;
__device__ __forceinline__ f2 pk(float lo, float hi) {
    f2 r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float hsum(f2 a) {
    float lo, hi;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a));
    return lo + hi;
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
    f2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
    f2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
    f2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

__device__ __forceinline__ void load_row(const float *p, f2 (&r)[D2]) {
#pragma unroll
    for (int c = 0; c < D2; c += 2) {
        const float4 t = __ldg(reinterpret_cast<const float4 *>(p + 2 * c));
        r[c] = pk(t.x, t.y);
        r[c + 1] = pk(t.z, t.w);
    }
}
__device__ __forceinline__ void store_row(float *p, const f2 (&r)[D2]) {
#pragma unroll
    for (int c = 0; c < D2; c += 2) {
        ulonglong2 t;
        t.x = r[c];
        t.y = r[c + 1];
        *reinterpret_cast<ulonglong2 *>(p + 2 * c) = t;
    }
}
__device__ __forceinline__ void scale_row(f2 (&r)[D2], float a) {
    const f2 a2 = pk(a, a);
#pragma unroll
    for (int c = 0; c < D2; ++c) r[c] = mul2(r[c], a2);
}
__device__ __forceinline__ void zero_row(f2 (&r)[D2]) {
#pragma unroll
    for (int c = 0; c < D2; ++c) r[c] = 0ull;
}
// a . s  (s: 16 floats in shared memory, 16-byte aligned): two chains of 4 FFMA2
__device__ __forceinline__ float dot_smem(const f2 (&a)[D2], const float *s) {
    const ulonglong2 *s2 = reinterpret_cast<const ulonglong2 *>(s);
    ulonglong2 t = s2[0];
    f2 e = mul2(a[0], t.x), o = mul2(a[1], t.y);
#pragma unroll
    for (int c = 1; c < D2 / 2; ++c) {
        t = s2[c];
        e = fma2(a[2 * c], t.x, e);
        o = fma2(a[2 * c + 1], t.y, o);
    }
    return hsum(add2(e, o));
}
__device__ __forceinline__ float dot_reg(const f2 (&a)[D2], const f2 (&b)[D2]) {
    f2 e = mul2(a[0], b[0]), o = mul2(a[1], b[1]);
#pragma unroll
    for (int c = 2; c < D2; c += 2) {
        e = fma2(a[c], b[c], e);
        o = fma2(a[c + 1], b[c + 1], o);
    }
    return hsum(add2(e, o));
}
// acc += a * s
__device__ __forceinline__ void axpy_smem(f2 (&acc)[D2], float a, const float *s) {
    const ulonglong2 *s2 = reinterpret_cast<const ulonglong2 *>(s);
    const f2 a2 = pk(a, a);
#pragma unroll
    for (int c = 0; c < D2 / 2; ++c) {
        const ulonglong2 t = s2[c];
        acc[2 * c] = fma2(a2, t.x, acc[2 * c]);
        acc[2 * c + 1] = fma2(a2, t.y, acc[2 * c + 1]);
    }
}

__device__ __forceinline__ void lds_row(f2 (&r)[D2], const float *s) {
    const ulonglong2 *s2 = reinterpret_cast<const ulonglong2 *>(s);
#pragma unroll
    for (int c = 0; c < D2 / 2; ++c) {
        const ulonglong2 t = s2[c];
        r[2 * c] = t.x;
        r[2 * c + 1] = t.y;
    }
}
__device__ __forceinline__ void axpy_reg(f2 (&acc)[D2], float a, const f2 (&x)[D2]) {
    const f2 a2 = pk(a, a);
#pragma unroll
    for (int c = 0; c < D2; ++c) acc[c] = fma2(a2, x[c], acc[c]);
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

// R rows (queries, or keys in pass B) per thread: every streamed shared-memory row is loaded ONCE per thread and used
// for R rows, which halves the LDS.128 traffic per score at R = 2 (the kernels are bound by shared-memory return
// bandwidth, not by the FMA pipe, once the math is FFMA2).  Row r of a thread is  blockIdx.x*QT*R + r*QT + tid.
template <int R>
__global__ void __launch_bounds__(QT) attn_fwd_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                       const float *__restrict__ v, float *__restrict__ o,
                                                       float *__restrict__ lse, int M, int H, float scale2) {
    __shared__ __align__(16) float ks[KT * D];
    __shared__ __align__(16) float vs[KT * D];
    const int b = blockIdx.z, h = blockIdx.y;
    int i[R];
    f2 qr[R][D2], acc[R][D2];
    float m[R], l[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        i[r] = (blockIdx.x * R + r) * QT + threadIdx.x;
        load_row(q + (((size_t)b * M + min(i[r], M - 1)) * H + h) * D, qr[r]);
        scale_row(qr[r], scale2);  // scores in the log2 domain: (q.k) / sqrt(d) * log2(e)
        zero_row(acc[r]);
        m[r] = -INFINITY;
        l[r] = 0.f;
    }
    for (int j0 = 0; j0 < M; j0 += KT) {
        __syncthreads();
        stage_tile(ks, k, b, h, H, M, j0, 1.f);
        stage_tile(vs, v, b, h, H, M, j0, 1.f);
        __syncthreads();
        const int nk = min(KT, M - j0);
        for (int jb = 0; jb < nk; jb += KB) {
            float s[R][KB];
            float mn[R];
#pragma unroll
            for (int r = 0; r < R; ++r) mn[r] = m[r];
#pragma unroll
            for (int u = 0; u < KB; ++u) {
                f2 row[D2];
                lds_row(row, ks + (jb + u) * D);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    s[r][u] = jb + u < nk ? dot_reg(qr[r], row) : -INFINITY;
                    mn[r] = fmaxf(mn[r], s[r][u]);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float corr = ex2(m[r] - mn[r]);  // first block: m = -inf, mn finite -> 0
                m[r] = mn[r];
                l[r] *= corr;
                scale_row(acc[r], corr);
            }
#pragma unroll
            for (int u = 0; u < KB; ++u) {
                f2 row[D2];
                lds_row(row, vs + (jb + u) * D);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float p = ex2(s[r][u] - mn[r]);
                    l[r] += p;
                    axpy_reg(acc[r], p, row);
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (i[r] < M) {
            scale_row(acc[r], 1.f / l[r]);
            store_row(o + (((size_t)b * M + i[r]) * H + h) * D, acc[r]);
            lse[((size_t)b * H + h) * M + i[r]] = m[r] + log2f(l[r]);
        }
    }
}

// pass A: dq_i = scale * sum_j p_ij (dO_i.v_j - delta_i) k_j ; delta_i = dO_i . O_i
template <int R>
__global__ void __launch_bounds__(QT) attn_bwd_dq_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                          const float *__restrict__ v, const float *__restrict__ o,
                                                          const float *__restrict__ dout, const float *__restrict__ lse,
                                                          float *__restrict__ dq, float *__restrict__ delta, int M,
                                                          int H, float scale, float scale2) {
    __shared__ __align__(16) float ks[KT * D];
    __shared__ __align__(16) float vs[KT * D];
    const int b = blockIdx.z, h = blockIdx.y;
    int i[R];
    size_t row[R];
    f2 qr[R][D2], dor[R][D2], acc[R][D2];
    float dl[R], L[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        i[r] = (blockIdx.x * R + r) * QT + threadIdx.x;
        const int ic = min(i[r], M - 1);
        row[r] = (((size_t)b * M + ic) * H + h) * D;
        load_row(q + row[r], qr[r]);
        load_row(dout + row[r], dor[r]);
        {
            f2 orow[D2];
            load_row(o + row[r], orow);
            dl[r] = dot_reg(dor[r], orow);
        }
        scale_row(qr[r], scale2);
        zero_row(acc[r]);
        L[r] = lse[((size_t)b * H + h) * M + ic];
    }
    for (int j0 = 0; j0 < M; j0 += KT) {
        __syncthreads();
        stage_tile(ks, k, b, h, H, M, j0, 1.f);
        stage_tile(vs, v, b, h, H, M, j0, 1.f);
        __syncthreads();
        const int nk = min(KT, M - j0);
#pragma unroll 4
        for (int j = 0; j < nk; ++j) {
            f2 kr[D2], vr[D2];
            lds_row(kr, ks + j * D);
            lds_row(vr, vs + j * D);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float p = ex2(dot_reg(qr[r], kr) - L[r]);
                const float dp = dot_reg(dor[r], vr);
                axpy_reg(acc[r], p * (dp - dl[r]), kr);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (i[r] < M) {
            scale_row(acc[r], scale);
            store_row(dq + row[r], acc[r]);
            delta[((size_t)b * H + h) * M + i[r]] = dl[r];
        }
    }
}

// pass B: dv_j = sum_i p_ij dO_i ; dk_j = scale * sum_i p_ij (dO_i.v_j - delta_i) q_i
template <int R>
__global__ void __launch_bounds__(QT) attn_bwd_dkv_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                           const float *__restrict__ v, const float *__restrict__ dout,
                                                           const float *__restrict__ lse, const float *__restrict__ delta,
                                                           float *__restrict__ dk, float *__restrict__ dv, int M, int H,
                                                           float scale2) {
    __shared__ __align__(16) float qs[KT * D];    // q * scale2
    __shared__ __align__(16) float dos[KT * D];
    __shared__ __align__(8) float2 ld[KT];        // (lse, delta)
    const int b = blockIdx.z, h = blockIdx.y;
    int j[R];
    size_t row[R];
    f2 kr[R][D2], vr[R][D2], dkr[R][D2], dvr[R][D2];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        j[r] = (blockIdx.x * R + r) * QT + threadIdx.x;
        row[r] = (((size_t)b * M + min(j[r], M - 1)) * H + h) * D;
        load_row(k + row[r], kr[r]);
        load_row(v + row[r], vr[r]);
        zero_row(dkr[r]);
        zero_row(dvr[r]);
    }
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
#pragma unroll 2
        for (int i = 0; i < nq; ++i) {
            const float2 Ld = ld[i];
            f2 qrow[D2], dorow[D2];
            lds_row(qrow, qs + i * D);
            lds_row(dorow, dos + i * D);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float p = ex2(dot_reg(kr[r], qrow) - Ld.x);
                const float dp = dot_reg(vr[r], dorow);
                axpy_reg(dvr[r], p, dorow);
                axpy_reg(dkr[r], p * (dp - Ld.y), qrow);
            }
        }
    }
    const float ln2 = 0.6931471805599453f;  // qs carries scale*log2(e): dk = scale * sum ds q = ln2 * sum ds qs
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (j[r] < M) {
            scale_row(dkr[r], ln2);
            store_row(dk + row[r], dkr[r]);
            store_row(dv + row[r], dvr[r]);
        }
    }
}


// ---- few tokens, any head width: the anchor head of the imitator (generator_component4_15.py:534-585 Producefactor:
// 4 anchors, dim 1024, 4 heads => head_dim 256).  One CTA per (head, batch); the M x M score matrix (M <= 32) lives in
// shared memory; P is kept for the backward pass ((B,H,M,M) floats).
constexpr int SM_MAXM = 32;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(256) attn_small_fwd_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                              const float *__restrict__ v, float *__restrict__ o,
                                                              float *__restrict__ pout, int M, int H, int Dh,
                                                              float scale) {
    __shared__ float S[SM_MAXM][SM_MAXM + 1];
    const int h = blockIdx.x, b = blockIdx.y;
    const int C = H * Dh;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
    const float *qb = q + (size_t)b * M * C + h * Dh, *kb = k + (size_t)b * M * C + h * Dh;
    const float *vb = v + (size_t)b * M * C + h * Dh;
    for (int pr = warp; pr < M * M; pr += nwarp) {
        const int i = pr / M, j = pr - i * M;
        float acc = 0.f;
        for (int d = lane; d < Dh; d += 32) acc = fmaf(qb[(size_t)i * C + d], kb[(size_t)j * C + d], acc);
        acc = warp_sum(acc);
        if (lane == 0) S[i][j] = acc * scale;
    }
    __syncthreads();
    for (int i = warp; i < M; i += nwarp) {
        const float x = lane < M ? S[i][lane] : -INFINITY;
        float mx = x;
#pragma unroll
        for (int of = 16; of > 0; of >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, of));
        const float e = lane < M ? expf(x - mx) : 0.f;
        const float den = warp_sum(e);
        if (lane < M) {
            const float pv = e / den;
            S[i][lane] = pv;
            pout[(((size_t)b * H + h) * M + i) * M + lane] = pv;
        }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < M * Dh; t += blockDim.x) {
        const int i = t / Dh, d = t - i * Dh;
        float acc = 0.f;
        for (int j = 0; j < M; ++j) acc = fmaf(S[i][j], vb[(size_t)j * C + d], acc);
        o[(size_t)b * M * C + (size_t)i * C + h * Dh + d] = acc;
    }
}

__global__ void __launch_bounds__(256) attn_small_bwd_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                              const float *__restrict__ v, const float *__restrict__ pin,
                                                              const float *__restrict__ dout, float *__restrict__ dq,
                                                              float *__restrict__ dk, float *__restrict__ dv, int M,
                                                              int H, int Dh, float scale) {
    __shared__ float P[SM_MAXM][SM_MAXM + 1];
    __shared__ float dS[SM_MAXM][SM_MAXM + 1];
    const int h = blockIdx.x, b = blockIdx.y;
    const int C = H * Dh;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
    const size_t base = (size_t)b * M * C + h * Dh;
    const float *qb = q + base, *kb = k + base, *vb = v + base, *gb = dout + base;
    for (int t = threadIdx.x; t < M * M; t += blockDim.x)
        P[t / M][t % M] = pin[(((size_t)b * H + h) * M) * M + t];
    // dP[i][j] = dO_i . V_j
    for (int pr = warp; pr < M * M; pr += nwarp) {
        const int i = pr / M, j = pr - i * M;
        float acc = 0.f;
        for (int d = lane; d < Dh; d += 32) acc = fmaf(gb[(size_t)i * C + d], vb[(size_t)j * C + d], acc);
        acc = warp_sum(acc);
        if (lane == 0) dS[i][j] = acc;
    }
    __syncthreads();
    // dS = P o (dP - rowsum(P o dP)) * scale
    for (int i = warp; i < M; i += nwarp) {
        const float pv = lane < M ? P[i][lane] : 0.f, dp = lane < M ? dS[i][lane] : 0.f;
        const float delta = warp_sum(pv * dp);
        if (lane < M) dS[i][lane] = pv * (dp - delta) * scale;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < M * Dh; t += blockDim.x) {
        const int i = t / Dh, d = t - i * Dh;
        float aq = 0.f, ak = 0.f, av = 0.f;
        for (int j = 0; j < M; ++j) {
            aq = fmaf(dS[i][j], kb[(size_t)j * C + d], aq);   // dQ_i = sum_j dS_ij K_j
            ak = fmaf(dS[j][i], qb[(size_t)j * C + d], ak);   // dK_i = sum_j dS_ji Q_j
            av = fmaf(P[j][i], gb[(size_t)j * C + d], av);    // dV_i = sum_j P_ji dO_j
        }
        const size_t off = base + (size_t)i * C + d;
        dq[off] = aq;
        dk[off] = ak;
        dv[off] = av;
    }
}

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

constexpr int MMA_MIN_TOKENS = 33;   // attention_mma.cu (3xTF32 warp-level MMA) takes every call with more tokens

}  // namespace

// attention_mma.cu
int apb_attn_mma_fwd(const float *q, const float *k, const float *v, float *o, float *lse, int B, int M, int H,
                     cudaStream_t st);
int apb_attn_mma_bwd(const float *q, const float *k, const float *v, const float *o, const float *dout, const float *lse,
                     float *dq, float *dk, float *dv, float *delta, int B, int M, int H, cudaStream_t st);

extern "C" int apb_attn_fwd_f32(const float *q, const float *k, const float *v, float *o, float *lse, int B, int M,
                                int H, int head_dim, apb_stream_t stream) {
    if (B < 0 || M < 0 || H < 0) return APB_E_INVALID;
    if (B == 0 || M == 0 || H == 0) return APB_OK;
    if (!q || !k || !v || !o || !lse) return APB_E_INVALID;
    if (head_dim != D || B > 65535 || H > 65535) return APB_E_UNSUPPORTED;
    if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(o)) return APB_E_INVALID;
    if (M >= MMA_MIN_TOKENS) return apb_attn_mma_fwd(q, k, v, o, lse, B, M, H, (cudaStream_t)stream);
    const float scale2 = 1.4426950408889634f / sqrtf((float)D);
    if (M > QT) {  // two queries per thread
        dim3 grid(apb_div_up(M, 2 * QT), H, B);
        attn_fwd_kernel<2><<<grid, QT, 0, (cudaStream_t)stream>>>(q, k, v, o, lse, M, H, scale2);
    } else {
        dim3 grid(apb_div_up(M, QT), H, B);
        attn_fwd_kernel<1><<<grid, QT, 0, (cudaStream_t)stream>>>(q, k, v, o, lse, M, H, scale2);
    }
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
    if (M >= MMA_MIN_TOKENS)
        return apb_attn_mma_bwd(q, k, v, o, dout, lse, dq, dk, dv, delta, B, M, H, (cudaStream_t)stream);
    const float scale = 1.f / sqrtf((float)D);
    const float scale2 = 1.4426950408889634f * scale;
    cudaStream_t st = (cudaStream_t)stream;
    if (M > QT) {  // two rows per thread
        dim3 grid(apb_div_up(M, 2 * QT), H, B);
        attn_bwd_dq_kernel<2><<<grid, QT, 0, st>>>(q, k, v, o, dout, lse, dq, delta, M, H, scale, scale2);
        APB_CHECK_LAUNCH();
        attn_bwd_dkv_kernel<2><<<grid, QT, 0, st>>>(q, k, v, dout, lse, delta, dk, dv, M, H, scale2);
        APB_CHECK_LAUNCH();
    } else {
        dim3 grid(apb_div_up(M, QT), H, B);
        attn_bwd_dq_kernel<1><<<grid, QT, 0, st>>>(q, k, v, o, dout, lse, dq, delta, M, H, scale, scale2);
        APB_CHECK_LAUNCH();
        attn_bwd_dkv_kernel<1><<<grid, QT, 0, st>>>(q, k, v, dout, lse, delta, dk, dv, M, H, scale2);
        APB_CHECK_LAUNCH();
    }
    return APB_OK;
}

extern "C" int apb_attn_small_fwd_f32(const float *q, const float *k, const float *v, float *o, float *p, int B, int M,
                                      int H, int head_dim, apb_stream_t stream) {
    if (B < 0 || M < 0 || H < 0 || head_dim <= 0) return APB_E_INVALID;
    if (B == 0 || M == 0 || H == 0) return APB_OK;
    if (!q || !k || !v || !o || !p) return APB_E_INVALID;
    if (M > SM_MAXM || B > 65535) return APB_E_UNSUPPORTED;
    attn_small_fwd_kernel<<<dim3(H, B), 256, 0, (cudaStream_t)stream>>>(q, k, v, o, p, M, H, head_dim,
                                                                        1.f / sqrtf((float)head_dim));
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_attn_small_bwd_f32(const float *q, const float *k, const float *v, const float *p, const float *dout,
                                      float *dq, float *dk, float *dv, int B, int M, int H, int head_dim,
                                      apb_stream_t stream) {
    if (B < 0 || M < 0 || H < 0 || head_dim <= 0) return APB_E_INVALID;
    if (B == 0 || M == 0 || H == 0) return APB_OK;
    if (!q || !k || !v || !p || !dout || !dq || !dk || !dv) return APB_E_INVALID;
    if (M > SM_MAXM || B > 65535) return APB_E_UNSUPPORTED;
    attn_small_bwd_kernel<<<dim3(H, B), 256, 0, (cudaStream_t)stream>>>(q, k, v, p, dout, dq, dk, dv, M, H, head_dim,
                                                                        1.f / sqrtf((float)head_dim));
    APB_CHECK_LAUNCH();
    return APB_OK;
}
