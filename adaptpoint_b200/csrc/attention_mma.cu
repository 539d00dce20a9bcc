// Multi-head self-attention, head_dim 16, on the warp-level tensor-core path with fp32-level accuracy (3xTF32).
//
// Same contract as attention.cu (reference generator_component4_15.py:434-480, Anchor_selfattention:
// softmax(q k^T / sqrt(d)) v; q, k, v, o, dO, dq, dk, dv token-major (B, M, H*16); lse (B,H,M) log2-domain), same
// flash-style structure (scores never leave registers, backward = a dq pass + a dk/dv pass that recompute p).
// What changes is where the multiply-adds run: round 1 evaluated every score with 32 fp32 FMAs on the CUDA cores
// (FMA pipe 66 % busy, 2.9 ms backward per step).  Here each 16x8 block of scores and each 16x8x16 update is a group of
// mma.sync.m16n8k8 TF32 instructions.  Plain TF32 would lose the fp32 accuracy the reference has (its matmuls are fp32;
// the parity tests hold 1e-5 / 5e-5), so every operand is split x = hi + lo with hi = tf32(x), lo = tf32(x - hi) and a
// product is evaluated as hi*hi + hi*lo + lo*hi (the lo*lo term is below fp32 rounding): three MMAs, ~2^-21 relative.
//
// Fragment bookkeeping (m16n8k8, g = lane / 4, t = lane % 4): A rows {g, g+8}, k {t, t+4}; B k {t, t+4}, n g;
// C rows {g, g+8}, columns {2t, 2t+1}.  A score block leaves the MMA in the C layout and re-enters the next MMA as an A
// operand; instead of shuffling, the summation index is RELABELLED: C column 2t is used as k = t and column 2t+1 as
// k = t+4, and the B operand of that MMA is loaded with the same relabelling (rows key0+2t, key0+2t+1).
// Shared-memory rows are padded to 20 floats, which makes both B-load patterns bank-conflict free.
//
// tcgen05 would need the hi/lo split of all five operands staged through shared memory by producer warps and the
// softmax between two dependent UMMAs (an FA4-class pipeline); with K = 16 per contraction the warp-level path already
// moves the bound from the fp32 pipe to exp2 + fragment bookkeeping.
#include "apb_common.cuh"

namespace {

constexpr int AD = 16;        // head_dim
constexpr int LD = 20;        // padded shared-memory row (floats)
constexpr int TK = 64;        // rows of the streamed operand per shared-memory tile
constexpr int NW = 8;         // warps per CTA; each owns 16 rows of the resident operand
constexpr int ROWS = NW * 16;

__device__ __forceinline__ float ex2f(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint32_t tf32_hi(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void split(float x, uint32_t &hi, uint32_t &lo) {
    hi = tf32_hi(x);
    lo = tf32_hi(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma8(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// c += A.B with A = ah + al, B = bh + bl (3xTF32)
__device__ __forceinline__ void mma3(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], uint32_t bh0,
                                     uint32_t bh1, uint32_t bl0, uint32_t bl1) {
    mma8(c, al, bh0, bh1);
    mma8(c, ah, bl0, bl1);
    mma8(c, ah, bh0, bh1);
}

// Rows r0+g, r0+g+8 of a token-major (.., H*16) matrix as A fragments for the two k-steps (d 0-7, 8-15), split hi/lo.
__device__ __forceinline__ void load_a_frags(const float *base, int rs, int r0, int M, int g, int t, float mul,
                                             uint32_t (&hi)[2][4], uint32_t (&lo)[2][4]) {
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = r0 + g + ((i & 1) ? 8 : 0);
            const int col = 8 * ks + t + ((i & 2) ? 4 : 0);
            const float v = row < M ? __ldg(base + (size_t)row * rs + col) * mul : 0.f;
            split(v, hi[ks][i], lo[ks][i]);
        }
    }
}

// Stage rows [row0, row0+TK) of a token-major matrix into shared memory as hi / lo planes [TK][LD].
__device__ __forceinline__ void stage_tile(const float *base, int rs, int row0, int M, float *sh, float *sl) {
    for (int i = threadIdx.x; i < TK * 4; i += NW * 32) {
        const int r = i >> 2, c4 = (i & 3) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < M) v = __ldg(reinterpret_cast<const float4 *>(base + (size_t)(row0 + r) * rs + c4));
        const float x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            uint32_t h, l;
            split(x[e], h, l);
            sh[r * LD + c4 + e] = __uint_as_float(h);
            sl[r * LD + c4 + e] = __uint_as_float(l);
        }
    }
}

// S block (16 x 8) = A(16 x 16) . T[row0 .. row0+8)^T with T staged in shared memory (B operand: k = d, n = row)
__device__ __forceinline__ void score_block(float (&c)[4], const uint32_t (&ah)[2][4], const uint32_t (&al)[2][4],
                                            const float *th, const float *tl, int row0, int g, int t) {
    c[0] = c[1] = c[2] = c[3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
        const int o = (row0 + g) * LD + 8 * ks + t;
        mma3(c, ah[ks], al[ks], __float_as_uint(th[o]), __float_as_uint(th[o + 4]), __float_as_uint(tl[o]),
             __float_as_uint(tl[o + 4]));
    }
}

// acc(16 x 16) += P(16 x 8) . T[row0 .. row0+8)  (P in the C layout, relabelled as an A operand; B: k = row, n = d)
__device__ __forceinline__ void update_block(float (&acc)[2][4], const float (&p)[4], const float *th, const float *tl,
                                             int row0, int g, int t) {
    uint32_t ph[4], pl[4];
    split(p[0], ph[0], pl[0]);   // (g,   2t)   -> a0 (g,   k = t)
    split(p[2], ph[1], pl[1]);   // (g+8, 2t)   -> a1 (g+8, k = t)
    split(p[1], ph[2], pl[2]);   // (g,   2t+1) -> a2 (g,   k = t+4)
    split(p[3], ph[3], pl[3]);   // (g+8, 2t+1) -> a3 (g+8, k = t+4)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int o0 = (row0 + 2 * t) * LD + 8 * j + g, o1 = o0 + LD;
        mma3(acc[j], ph, pl, __float_as_uint(th[o0]), __float_as_uint(th[o1]), __float_as_uint(tl[o0]),
             __float_as_uint(tl[o1]));
    }
}

__device__ __forceinline__ float quad_max(float v) {
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
    return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NW * 32) attn_mma_fwd_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                               const float *__restrict__ v, float *__restrict__ o,
                                                               float *__restrict__ lse, int M, int H, float scale2) {
    __shared__ __align__(16) float kh[TK * LD], kl[TK * LD], vh[TK * LD], vl[TK * LD];
    const int h = blockIdx.y, b = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int rs = H * AD;
    const size_t base = (size_t)b * M * rs + h * AD;
    const int r0 = blockIdx.x * ROWS + warp * 16;
    uint32_t qh[2][4], ql[2][4];
    load_a_frags(q + base, rs, r0, M, g, t, scale2, qh, ql);
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
    for (int k0 = 0; k0 < M; k0 += TK) {
        __syncthreads();
        stage_tile(k + base, rs, k0, M, kh, kl);
        stage_tile(v + base, rs, k0, M, vh, vl);
        __syncthreads();
        const int nk = min(TK, M - k0);
        for (int kb = 0; kb < nk; kb += 32) {   // 32 keys per online-softmax step
            float s[4][4];
            float bm[2] = {-INFINITY, -INFINITY};
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
                if (kb + 8 * nt < nk) {
                    score_block(s[nt], qh, ql, kh, kl, kb + 8 * nt, g, t);
                    const int key = k0 + kb + 8 * nt + 2 * t;
                    if (key >= M) s[nt][0] = s[nt][2] = -INFINITY;
                    if (key + 1 >= M) s[nt][1] = s[nt][3] = -INFINITY;
                } else {
                    s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = -INFINITY;
                }
                bm[0] = fmaxf(bm[0], fmaxf(s[nt][0], s[nt][1]));
                bm[1] = fmaxf(bm[1], fmaxf(s[nt][2], s[nt][3]));
            }
            float alpha[2];
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                const float mn = fmaxf(mrow[r], quad_max(bm[r]));
                alpha[r] = ex2f(mrow[r] - mn);
                mrow[r] = mn;
                lrow[r] *= alpha[r];
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                acc[j][0] *= alpha[0]; acc[j][1] *= alpha[0];
                acc[j][2] *= alpha[1]; acc[j][3] *= alpha[1];
            }
            // The tensor core accumulates with truncation; a 2048-key chain inside one accumulator would bias the result
            // by ~1e-5.  Each 32-key step therefore accumulates into a fresh block accumulator that is added to the
            // running one with an ordinary (round-to-nearest) fp32 add.
            float blk[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
                if (kb + 8 * nt < nk) {
                    float p[4];
                    p[0] = ex2f(s[nt][0] - mrow[0]); p[1] = ex2f(s[nt][1] - mrow[0]);
                    p[2] = ex2f(s[nt][2] - mrow[1]); p[3] = ex2f(s[nt][3] - mrow[1]);
                    lrow[0] += p[0] + p[1];
                    lrow[1] += p[2] + p[3];
                    update_block(blk, p, vh, vl, kb + 8 * nt, g, t);
                }
            }
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[j][e] += blk[j][e];
        }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int row = r0 + g + 8 * r;
        const float l = quad_sum(lrow[r]);
        if (row < M) {
            const float inv = 1.f / l;
#pragma unroll
            for (int j = 0; j < 2; ++j)
                *reinterpret_cast<float2 *>(o + base + (size_t)row * rs + 8 * j + 2 * t) =
                    make_float2(acc[j][2 * r] * inv, acc[j][2 * r + 1] * inv);
            if (t == 0) lse[((size_t)b * H + h) * M + row] = mrow[r] + log2f(l);
        }
    }
}

// dq pass: thread block = 128 queries; also writes delta = rowsum(dO o O)
__global__ void __launch_bounds__(NW * 32) attn_mma_bwd_dq_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                                  const float *__restrict__ v, const float *__restrict__ o,
                                                                  const float *__restrict__ dout,
                                                                  const float *__restrict__ lse, float *__restrict__ dq,
                                                                  float *__restrict__ delta, int M, int H, float scale,
                                                                  float scale2) {
    __shared__ __align__(16) float kh[TK * LD], kl[TK * LD], vh[TK * LD], vl[TK * LD];
    const int h = blockIdx.y, b = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int rs = H * AD;
    const size_t base = (size_t)b * M * rs + h * AD;
    const int r0 = blockIdx.x * ROWS + warp * 16;
    uint32_t qh[2][4], ql[2][4], gh[2][4], gl[2][4];
    load_a_frags(q + base, rs, r0, M, g, t, scale2, qh, ql);
    load_a_frags(dout + base, rs, r0, M, g, t, 1.f, gh, gl);
    float lrow[2], drow[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int row = r0 + g + 8 * r;
        float d = 0.f;
        if (row < M) {   // this lane's share of the row: columns 4t .. 4t+3
            const float4 a = __ldg(reinterpret_cast<const float4 *>(o + base + (size_t)row * rs + 4 * t));
            const float4 e = __ldg(reinterpret_cast<const float4 *>(dout + base + (size_t)row * rs + 4 * t));
            d = a.x * e.x + a.y * e.y + a.z * e.z + a.w * e.w;
        }
        d = quad_sum(d);
        drow[r] = d;
        lrow[r] = row < M ? __ldg(lse + ((size_t)b * H + h) * M + row) : 0.f;
        if (row < M && t == 0) delta[((size_t)b * H + h) * M + row] = d;
    }
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    for (int k0 = 0; k0 < M; k0 += TK) {
        __syncthreads();
        stage_tile(k + base, rs, k0, M, kh, kl);
        stage_tile(v + base, rs, k0, M, vh, vl);
        __syncthreads();
        const int nk = min(TK, M - k0);
        float blk[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};   // per-tile chain, see the forward kernel
        for (int kb = 0; kb < nk; kb += 8) {
            float s[4], dp[4], ds[4];
            score_block(s, qh, ql, kh, kl, kb, g, t);
            score_block(dp, gh, gl, vh, vl, kb, g, t);
            const int key = k0 + kb + 2 * t;
            const bool ok0 = key < M, ok1 = key + 1 < M;
            ds[0] = ok0 ? ex2f(s[0] - lrow[0]) * (dp[0] - drow[0]) : 0.f;
            ds[1] = ok1 ? ex2f(s[1] - lrow[0]) * (dp[1] - drow[0]) : 0.f;
            ds[2] = ok0 ? ex2f(s[2] - lrow[1]) * (dp[2] - drow[1]) : 0.f;
            ds[3] = ok1 ? ex2f(s[3] - lrow[1]) * (dp[3] - drow[1]) : 0.f;
            update_block(blk, ds, kh, kl, kb, g, t);
        }
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[j][e] += blk[j][e];
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int row = r0 + g + 8 * r;
        if (row < M) {
#pragma unroll
            for (int j = 0; j < 2; ++j)
                *reinterpret_cast<float2 *>(dq + base + (size_t)row * rs + 8 * j + 2 * t) =
                    make_float2(acc[j][2 * r] * scale, acc[j][2 * r + 1] * scale);
        }
    }
}

// dk / dv pass: thread block = 128 keys; queries, dO, lse, delta stream through shared memory
__global__ void __launch_bounds__(NW * 32) attn_mma_bwd_dkv_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                                   const float *__restrict__ v,
                                                                   const float *__restrict__ dout,
                                                                   const float *__restrict__ lse,
                                                                   const float *__restrict__ delta, float *__restrict__ dk,
                                                                   float *__restrict__ dv, int M, int H, float scale,
                                                                   float scale2) {
    __shared__ __align__(16) float qh[TK * LD], ql[TK * LD], gh[TK * LD], gl[TK * LD];
    __shared__ float s_lse[TK], s_del[TK];
    const int h = blockIdx.y, b = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int rs = H * AD;
    const size_t base = (size_t)b * M * rs + h * AD;
    const int r0 = blockIdx.x * ROWS + warp * 16;
    uint32_t kfh[2][4], kfl[2][4], vfh[2][4], vfl[2][4];
    load_a_frags(k + base, rs, r0, M, g, t, scale2, kfh, kfl);
    load_a_frags(v + base, rs, r0, M, g, t, 1.f, vfh, vfl);
    float adk[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}}, adv[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    for (int q0 = 0; q0 < M; q0 += TK) {
        __syncthreads();
        stage_tile(q + base, rs, q0, M, qh, ql);
        stage_tile(dout + base, rs, q0, M, gh, gl);
        if (threadIdx.x < TK) {
            const int row = q0 + threadIdx.x;
            s_lse[threadIdx.x] = row < M ? __ldg(lse + ((size_t)b * H + h) * M + row) : 0.f;
            s_del[threadIdx.x] = row < M ? __ldg(delta + ((size_t)b * H + h) * M + row) : 0.f;
        }
        __syncthreads();
        const int nq = min(TK, M - q0);
        float bk[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}}, bv[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        for (int qb = 0; qb < nq; qb += 8) {
            float s[4], dp[4], p[4], ds[4];
            score_block(s, kfh, kfl, qh, ql, qb, g, t);     // S^T: rows = keys, columns = queries qb+2t, qb+2t+1
            score_block(dp, vfh, vfl, gh, gl, qb, g, t);    // dP^T = V . dO^T
            const int c0 = qb + 2 * t, c1 = c0 + 1;
            const bool ok0 = q0 + c0 < M, ok1 = q0 + c1 < M;
            const float l0 = s_lse[c0], l1 = s_lse[c1], d0 = s_del[c0], d1 = s_del[c1];
            p[0] = ok0 ? ex2f(s[0] - l0) : 0.f; p[1] = ok1 ? ex2f(s[1] - l1) : 0.f;
            p[2] = ok0 ? ex2f(s[2] - l0) : 0.f; p[3] = ok1 ? ex2f(s[3] - l1) : 0.f;
            ds[0] = p[0] * (dp[0] - d0); ds[1] = p[1] * (dp[1] - d1);
            ds[2] = p[2] * (dp[2] - d0); ds[3] = p[3] * (dp[3] - d1);
            update_block(bv, p, gh, gl, qb, g, t);          // dV += P^T . dO
            update_block(bk, ds, qh, ql, qb, g, t);         // dK += dS^T . Q
        }
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                adk[j][e] += bk[j][e];
                adv[j][e] += bv[j][e];
            }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int row = r0 + g + 8 * r;
        if (row < M) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                *reinterpret_cast<float2 *>(dk + base + (size_t)row * rs + 8 * j + 2 * t) =
                    make_float2(adk[j][2 * r] * scale, adk[j][2 * r + 1] * scale);
                *reinterpret_cast<float2 *>(dv + base + (size_t)row * rs + 8 * j + 2 * t) =
                    make_float2(adv[j][2 * r], adv[j][2 * r + 1]);
            }
        }
    }
}

}  // namespace

// Called by apb_attn_{fwd,bwd}_f32 (attention.cu) for M > 32 tokens.
int apb_attn_mma_fwd(const float *q, const float *k, const float *v, float *o, float *lse, int B, int M, int H,
                     cudaStream_t st) {
    const float scale2 = 1.4426950408889634f / sqrtf((float)AD);
    attn_mma_fwd_kernel<<<dim3(apb_div_up(M, ROWS), H, B), NW * 32, 0, st>>>(q, k, v, o, lse, M, H, scale2);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

int apb_attn_mma_bwd(const float *q, const float *k, const float *v, const float *o, const float *dout, const float *lse,
                     float *dq, float *dk, float *dv, float *delta, int B, int M, int H, cudaStream_t st) {
    const float scale = 1.f / sqrtf((float)AD);
    const float scale2 = 1.4426950408889634f * scale;
    const dim3 grid(apb_div_up(M, ROWS), H, B);
    attn_mma_bwd_dq_kernel<<<grid, NW * 32, 0, st>>>(q, k, v, o, dout, lse, dq, delta, M, H, scale, scale2);
    APB_CHECK_LAUNCH();
    attn_mma_bwd_dkv_kernel<<<grid, NW * 32, 0, st>>>(q, k, v, dout, lse, delta, dk, dv, M, H, scale, scale2);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
