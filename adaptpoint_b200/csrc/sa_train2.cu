// Training-mode set abstraction for TWO-conv blocks (PointNeXt-S SetAbstraction, sa_layers = 2) without the (B, C, M, K)
// tensors -- forward AND backward -- on tcgen05 / TMEM / TMA.
//
// Reference op chain (openpoints/models/backbone/pointnext.py:164-166; layers/group.py:235-255, :323-325;
// layers/conv.py:24-60):
//   ball query -> group xyz, subtract query (/ radius) -> group features -> cat -> Conv2d 1x1 -> BatchNorm2d -> ReLU
//   -> Conv2d 1x1 -> BatchNorm2d [-> ReLU] -> max over K
// which writes eight (B, C, M, K) activations forward and as many gradients backward.
//
// Algebra (rows r = (b, m, k), n_r = idx[b, m, k], dp_r = p[n_r] - q[m], / radius folded into Wp):
//   y1[r, :] = U[n_r, :] + Wp . dp_r          U = W1f . f per SUPPORT point (dense GEMM on B*N rows, tc_cm_gemm.cu)
//   h1 = relu(s1 * y1 + t1)                   (s1, t1) = BatchNorm 1 (batch or running statistics)
//   y2 = W2 . h1                              B*M*K rows on the tensor cores
//   out = act(BN2(max_k y2 or min_k y2))      BatchNorm + ReLU are monotone per channel (norm.cu, bnmax_finalize)
// forward, pass 1 (sa_train.cu group_ystats_kernel, statistics only): sum y1, sum y1^2 -> (s1, t1).
// forward, pass 2 (sa2_rows_kernel<false>): producer warps gather rows of U, form h1 in registers and store it as the TF32
//   A tile (128 rows x 32 channels per stage, 128B swizzle); W2 chunks arrive by TMA; the accumulator lives in TMEM;
//   the epilogue keeps, per (b, c2, m), max / min over the K = 32 rows of a query with their k (CREDUX + ballot) and per
//   channel sum y2, sum y2^2 (register butterfly over the 32 lanes).  Nothing of size B*C*M*K is written.
// backward.  With batch statistics the BatchNorm-2 gradient is dense in r:  dy2 = gs*[k == arg] + a2 + b2*y2, so
//     dh1[r, :] = W2^T (gs*[k == arg])[r, :] + W2^T a2 + (W2^T diag(b2) W2) h1[r, :]
//   = one GEMM with the A tile [ one-hot gs rows | h1 rows ] against [ W2^T | G ] (sa2_rows_kernel<true>; eval mode: the
//   first block only).  Its epilogue recomputes the ReLU mask from U, forms dz1 = (dh1 + v) * [z1 > 0], scatters
//   s1 * dz1 into dU (128-bit RED), accumulates the BatchNorm-1 sums (sum dz1, sum dz1 * xhat1), sum dz1 * dp (-> dWp),
//   sum h1, and the coordinate gradients.  The dense BatchNorm-1 terms (a1 + b1 * y1) are linear in U and dp, so they take
//   the closed per-support-point form of the one-conv path (sa_train.cu: sa1_bwd_count / sa1_bwd_dense).
//   dW2 = sum_r dy2[r]^T h1[r] = S + a2 (sum h1)^T + diag(b2) W2 H  with  S = sum_r (gs one-hot)[r]^T h1[r],
//   H = sum_r h1[r] h1[r]^T:  sa2_wgrad_kernel accumulates [S ; H] = [one-hot gs | h1]^T h1 over all rows in TMEM, the
//   operands written by the producer warps as K-major (row-contiguous) TF32 tiles.
// Numerics: TF32 operands, fp32 accumulation (the precision class of tcconv.py's training path).
#include <cuda.h>
#include <cfloat>
#include <climits>
#include "apb_common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int BM = 128;            // rows per tile (UMMA M)
constexpr int A_BYTES = BM * 128;  // A chunk: 128 rows x 32 tf32
constexpr int NGROUPS = 2;         // A-producer groups of 128 threads (alternate tiles)
constexpr int PROD0 = 64;
constexpr int EPI0 = PROD0 + 128 * NGROUPS;
constexpr int NTHREADS = EPI0 + 128;
constexpr int MAX_STAGES = 6;
constexpr int NPRM = 8;            // s1, t1, wx, wy, wz, v, mean1, rstd1
constexpr int NSUM = 6;            // backward channel sums: dz, dz*xhat, dz*dpx, dz*dpy, dz*dpz, h1

struct S2Params {
    long long rows;        // B*M*K
    int kchunks;           // K chunks (of 32) per tile: forward C1p/32; backward C2p/32 (+ C1p/32 in training mode)
    int kc_sparse;         // backward: leading chunks built from (gs, ksel); forward: 0
    int NCP, wrows, stages;
    int Cn;                // valid accumulator columns: forward C2, backward C1
    int C1, C2, c1pad, N, M, K, relu1;
    const float *U;        // (B, N, C1) point-major
    const float *xyz, *new_xyz;
    const int32_t *idx;    // (B, M, K)
    const float *Wp;       // (C1, 3) row-major, already divided by the radius
    const float *s1, *t1;  // (C1)
    // forward outputs
    float *vmax, *vmin;    // (B, C2, M)
    int32_t *arg;          // (B, C2, M) kmax | kmin << 16
    double *stats;         // (C2, 2) += sum y2, sum y2^2 (nullptr: eval mode)
    // backward inputs
    const float *gs_pm;    // (B*M, C2) gradient reaching the selected y2 value
    const uint8_t *ksel_pm;  // (B*M, C2) its k
    const float *vvec;     // (C1) W2^T a2 (nullptr: eval mode)
    const float *mean1, *rstd1;
    // backward outputs
    float *dU_pm;          // (B, N, C1) pre-zeroed, += s1 * dz1
    double *sums1;         // (C1, 2) += sum dz1, sum dz1*xhat1   (nullptr: not wanted)
    double *sdzdp;         // (C1, 3) += sum dz1 * dp
    double *sh1;           // (C1)    += sum h1
    float *dxyz, *dq;      // (B, N, 3), (B, M, 3) pre-zeroed or nullptr
};

__device__ __forceinline__ uint8_t *align1024(uint8_t *p) {
    return reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 1023) & ~(uintptr_t)1023);
}

// x[j] of lane l = value of (row l, column j).  Returns, in lane l, the sum over the 32 rows of column l
// (recursive halving: 31 shuffles instead of 32 five-step reductions).  Destroys x.
template <int S>
__device__ __forceinline__ void colsum_step(float (&x)[32], int lane) {
    const bool up = (lane & S) != 0;
#pragma unroll
    for (int i = 0; i < S; ++i) {
        const float send = up ? x[i] : x[i + S];
        const float keep = up ? x[i + S] : x[i];
        x[i] = keep + __shfl_xor_sync(0xffffffffu, send, S);
    }
}
__device__ __forceinline__ float colsum32(float (&x)[32], int lane) {
    colsum_step<16>(x, lane);
    colsum_step<8>(x, lane);
    colsum_step<4>(x, lane);
    colsum_step<2>(x, lane);
    colsum_step<1>(x, lane);
    return x[0];
}

__device__ __forceinline__ void red_add_v4(float *dst, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// h1 of 4 consecutive channels of one row: relu(s1 * (u + Wp . dp) + t1); prm = [NPRM][c1pad] in shared memory
__device__ __forceinline__ float4 h1_quad(const float *prm, int c1pad, int c, float4 u, float dx, float dy, float dz,
                                          int relu1) {
    const float4 s = *reinterpret_cast<const float4 *>(prm + 0 * c1pad + c);
    const float4 t = *reinterpret_cast<const float4 *>(prm + 1 * c1pad + c);
    const float4 wx = *reinterpret_cast<const float4 *>(prm + 2 * c1pad + c);
    const float4 wy = *reinterpret_cast<const float4 *>(prm + 3 * c1pad + c);
    const float4 wz = *reinterpret_cast<const float4 *>(prm + 4 * c1pad + c);
    float4 h;
    h.x = fmaf(s.x, fmaf(wz.x, dz, fmaf(wy.x, dy, fmaf(wx.x, dx, u.x))), t.x);
    h.y = fmaf(s.y, fmaf(wz.y, dz, fmaf(wy.y, dy, fmaf(wx.y, dx, u.y))), t.y);
    h.z = fmaf(s.z, fmaf(wz.z, dz, fmaf(wy.z, dy, fmaf(wx.z, dx, u.z))), t.z);
    h.w = fmaf(s.w, fmaf(wz.w, dz, fmaf(wy.w, dy, fmaf(wx.w, dx, u.w))), t.w);
    if (relu1) {
        h.x = fmaxf(h.x, 0.f); h.y = fmaxf(h.y, 0.f); h.z = fmaxf(h.z, 0.f); h.w = fmaxf(h.w, 0.f);
    }
    return h;
}

// one-hot gradient of 4 consecutive channels of row (q, k): gs[q, d] where ksel[q, d] == k, else 0
__device__ __forceinline__ float4 onehot_quad(const float *gs_pm, const uint8_t *ksel_pm, long long q, int C2, int d,
                                              int k) {
    const float4 g = __ldg(reinterpret_cast<const float4 *>(gs_pm + q * C2 + d));
    const uint32_t ks = __ldg(reinterpret_cast<const uint32_t *>(ksel_pm + q * C2 + d));
    float4 r;
    r.x = (int)(ks & 0xff) == k ? g.x : 0.f;
    r.y = (int)((ks >> 8) & 0xff) == k ? g.y : 0.f;
    r.z = (int)((ks >> 16) & 0xff) == k ? g.z : 0.f;
    r.w = (int)(ks >> 24) == k ? g.w : 0.f;
    return r;
}

// ---- sa2_rows_kernel: persistent, one CTA per SM, 448 threads, warp-specialised ------------------------------------
//    warp 0      TMA producer for the weight chunk of every pipeline stage
//    warp 1      TMEM allocator + single-thread tcgen05.mma issuer (UMMA 128 x N x 8, tf32 -> fp32)
//    warps 2-9   A-tile producers, thread <-> row of the tile (2 groups of 128 threads, alternating tiles)
//    warps 10-13 epilogue, thread <-> row (TMEM lane); a warp's 32 rows are the K = 32 neighbours of one query
template <bool BWD>
__global__ void __launch_bounds__(NTHREADS, 1)
sa2_rows_kernel(const __grid_constant__ CUtensorMap wmap, const S2Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = align1024(smem_raw);
    const int stage_bytes = A_BYTES + p.wrows * 128;
    uint8_t *tail = smem + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);  // [MAX_STAGES]
    uint64_t *empty_bar = full_bar + MAX_STAGES;              // [MAX_STAGES]
    uint64_t *tfull_bar = empty_bar + MAX_STAGES;             // [2]
    uint64_t *tempty_bar = tfull_bar + 2;                     // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);
    float *prm = reinterpret_cast<float *>(tail + 256);       // [NPRM][c1pad]
    const int c1pad = p.c1pad;
    float *wacc = prm + NPRM * c1pad;                         // forward [4][2][NCP]; backward [4][NSUM][c1pad]
    const int wacc_per_warp = BWD ? NSUM * c1pad : 2 * p.NCP;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int acc_stages = p.NCP <= 256 ? 2 : 1;
    const uint32_t spg = p.stages / NGROUPS;  // smem stages per producer group
    const long long ntiles = (p.rows + BM - 1) / BM;

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&wmap);
        for (int s = 0; s < p.stages; ++s) {
            tc::mbar_init(&full_bar[s], 128 + 1);  // 128 A-producer arrivals + the TMA thread's expect_tx arrive
            tc::mbar_init(&empty_bar[s], 1);       // one tcgen05.commit
        }
        for (int a = 0; a < 2; ++a) {
            tc::mbar_init(&tfull_bar[a], 1);
            tc::mbar_init(&tempty_bar[a], 128);
        }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
    for (int c = tid; c < c1pad; c += NTHREADS) {
        const bool ok = c < p.C1;
        prm[0 * c1pad + c] = ok ? p.s1[c] : 0.f;
        prm[1 * c1pad + c] = ok ? p.t1[c] : 0.f;
        prm[2 * c1pad + c] = ok ? p.Wp[c * 3 + 0] : 0.f;
        prm[3 * c1pad + c] = ok ? p.Wp[c * 3 + 1] : 0.f;
        prm[4 * c1pad + c] = ok ? p.Wp[c * 3 + 2] : 0.f;
        prm[5 * c1pad + c] = (BWD && ok && p.vvec != nullptr) ? p.vvec[c] : 0.f;
        prm[6 * c1pad + c] = (BWD && ok && p.mean1 != nullptr) ? p.mean1[c] : 0.f;
        prm[7 * c1pad + c] = (BWD && ok && p.rstd1 != nullptr) ? p.rstd1[c] : 0.f;
    }
    for (int i = tid; i < 4 * wacc_per_warp; i += NTHREADS) wacc[i] = 0.f;
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer: weight chunk of every stage =====================
        if (lane == 0) {
            uint32_t lt = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
                const int grp = lt % NGROUPS;
                uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;  // chunk counter of this group's own stage ring
                for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                    const int s = grp * spg + it % spg;
                    const uint32_t ph = (it / spg) & 1;
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t *wdst = smem + (size_t)s * stage_bytes + A_BYTES;
                    tc::mbar_arrive_expect_tx(&full_bar[s], (uint32_t)p.wrows * 128);
                    tc::tma_load_2d(wdst, &wmap, kc * 32, 0, &full_bar[s]);
                    if (p.wrows > 256) tc::tma_load_2d(wdst + 256 * 128, &wmap, kc * 32, 256, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const int n_lo = p.NCP <= 256 ? p.NCP : 256, n_hi = p.NCP > 256 ? p.NCP - 256 : 16;
            const uint32_t idesc_lo = tc::idesc_tf32_f32(BM, n_lo);
            const uint32_t idesc_hi = tc::idesc_tf32_f32(BM, n_hi);
            uint32_t lt = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
                const int grp = lt % NGROUPS;
                uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;
                const int a = lt % acc_stages;
                const uint32_t aph = (lt / acc_stages) & 1;
                tc::mbar_wait(&tempty_bar[a], aph ^ 1);
                tc::tc_fence_after();
                const uint32_t tacc = tmem_base + (uint32_t)a * 256;
                for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                    const int s = grp * spg + it % spg;
                    const uint32_t ph = (it / spg) & 1;
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_addr = tc::smem_u32(smem + (size_t)s * stage_bytes);
                    const uint32_t w_addr = a_addr + A_BYTES;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {  // 4 x 32 bytes per chunk: UMMA_K = 8 tf32
                        const uint64_t ad = tc::smem_desc_sw128(a_addr + k * 32);
                        const uint64_t bd = tc::smem_desc_sw128(w_addr + k * 32);
                        tc::umma_tf32(tacc, ad, bd, idesc_lo, (kc | k) != 0);
                        if (p.NCP > 256) {
                            const uint64_t bd2 = tc::smem_desc_sw128(w_addr + 256 * 128 + k * 32);
                            tc::umma_tf32(tacc + 256, ad, bd2, idesc_hi, (kc | k) != 0);
                        }
                    }
                    tc::umma_commit(&empty_bar[s]);  // smem stage reusable once these MMAs have read it
                }
                tc::umma_commit(&tfull_bar[a]);      // accumulator complete -> epilogue
            }
        }
    } else if (warp < 2 + 4 * NGROUPS) {
        // ===================== A-tile producers: group g fills tiles lt % NGROUPS == g, thread <-> row ==================
        const int grp = (warp - 2) >> 2;
        const int pt = (tid - PROD0) & 127;
        uint32_t lt = 0;
        // row metadata (support-point row of U, dp = p[idx] - q), fetched one tile ahead so that the dependent
        // idx -> xyz loads overlap the previous tile's chunk work
        long long nx_urow = -1, nx_q = 0;
        int nx_k = 0;
        float nx_dx = 0.f, nx_dy = 0.f, nx_dz = 0.f;
        auto fetch_meta = [&](long long tile) {
            nx_urow = -1;
            nx_dx = nx_dy = nx_dz = 0.f;
            const long long grow = tile * BM + pt;
            if (tile < ntiles && grow < p.rows) {
                const long long q = grow / p.K;  // b*M + m
                const long long b = q / p.M;
                const int n = __ldg(p.idx + grow);
                const float *pp = p.xyz + (b * p.N + n) * 3;
                const float *qq = p.new_xyz + q * 3;
                nx_dx = __ldg(pp + 0) - __ldg(qq + 0);
                nx_dy = __ldg(pp + 1) - __ldg(qq + 1);
                nx_dz = __ldg(pp + 2) - __ldg(qq + 2);
                nx_urow = b * p.N + n;
                nx_q = q;
                nx_k = (int)(grow - q * p.K);
            }
        };
        fetch_meta((long long)blockIdx.x + (long long)grp * gridDim.x);
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            if ((int)(lt % NGROUPS) != grp) continue;
            uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;  // this group's own stage ring (in-order => parity-safe)
            const long long urow = nx_urow, q = nx_q;
            const int krow = nx_k;
            const float dx = nx_dx, dy = nx_dy, dz = nx_dz;
            fetch_meta(tile + (long long)NGROUPS * gridDim.x);
            const float *urowp = p.U + (urow < 0 ? 0 : urow) * p.C1;
            for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                const int s = grp * spg + it % spg;
                const uint32_t ph = (it / spg) & 1;
                const uint32_t arow = tc::smem_u32(smem + (size_t)s * stage_bytes) + pt * 128;
                float4 v[8];
                if (BWD && kc < p.kc_sparse) {
                    const int d0 = kc * 32;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int d = d0 + 4 * j;
                        v[j] = (urow >= 0 && d < p.C2) ? onehot_quad(p.gs_pm, p.ksel_pm, q, p.C2, d, krow)
                                                       : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                } else {
                    const int c0 = (kc - (BWD ? p.kc_sparse : 0)) * 32;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int c = c0 + 4 * j;
                        v[j] = (urow >= 0 && c < p.C1) ? __ldg(reinterpret_cast<const float4 *>(urowp + c))
                                                       : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int c = c0 + 4 * j;
                        if (urow >= 0 && c < p.C1) v[j] = h1_quad(prm, c1pad, c, v[j], dx, dy, dz, p.relu1);
                    }
                }
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    tc::st_shared_v4(arow + ((j ^ (pt & 7)) << 4), tc::to_tf32(v[j].x), tc::to_tf32(v[j].y),
                                     tc::to_tf32(v[j].z), tc::to_tf32(v[j].w));
                tc::fence_proxy_async_smem();
                tc::mbar_arrive(&full_bar[s]);
            }
        }
    } else {
        // ===================== epilogue (128 threads; warp w reads TMEM lanes 32*(w%4)..) =====================
        const int q4 = warp & 3;
        float *myacc = wacc + q4 * wacc_per_warp;
        uint32_t lt = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            const int a = lt % acc_stages;
            const uint32_t aph = (lt / acc_stages) & 1;
            const long long row0 = tile * BM + q4 * 32;      // the 32 rows of this warp: the K neighbours of ONE query
            const bool qvalid = row0 < p.rows;
            const long long q = row0 / p.K;
            const long long b = q / p.M, m = q - b * p.M;
            // backward: row metadata, loaded before the accumulator wait
            long long urow = 0;
            float dx = 0.f, dy = 0.f, dz = 0.f;
            if (BWD && qvalid) {
                const int n = __ldg(p.idx + row0 + lane);
                urow = b * p.N + n;
                const float *pp = p.xyz + urow * 3;
                const float *qq = p.new_xyz + q * 3;
                dx = __ldg(pp + 0) - __ldg(qq + 0);
                dy = __ldg(pp + 1) - __ldg(qq + 1);
                dz = __ldg(pp + 2) - __ldg(qq + 2);
            }
            tc::mbar_wait(&tfull_bar[a], aph);
            tc::tc_fence_after();
            const uint32_t tacc = tmem_base + (uint32_t)a * 256 + ((uint32_t)(q4 * 32) << 16);
            if (!BWD) {
                for (int cb = 0; cb < p.NCP; cb += 32) {
                    uint32_t v[32];
                    tc::tmem_ld_32x32(tacc + cb, v);
                    tc::tmem_ld_wait();
                    int mx = INT_MIN, mn = INT_MAX, kx = 0, kn = 0;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int o = tc::float_to_ordered(__uint_as_float(v[j]));
                        const int rmax = __reduce_max_sync(0xffffffffu, o);
                        const int rmin = __reduce_min_sync(0xffffffffu, o);
                        const uint32_t bx = __ballot_sync(0xffffffffu, o == rmax);
                        const uint32_t bn = __ballot_sync(0xffffffffu, o == rmin);
                        if (lane == j) {
                            mx = rmax; mn = rmin;
                            kx = __ffs(bx) - 1;      // first k wins ties (torch.max order)
                            kn = __ffs(bn) - 1;
                        }
                    }
                    float x[32], x2[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        x[j] = __uint_as_float(v[j]);
                        x2[j] = x[j] * x[j];
                    }
                    const float s1 = colsum32(x, lane);
                    const float s2 = colsum32(x2, lane);
                    const int c = cb + lane;
                    if (qvalid && c < p.Cn) {
                        const size_t o = ((size_t)b * p.C2 + c) * p.M + m;
                        p.vmax[o] = tc::ordered_to_float(mx);
                        p.vmin[o] = tc::ordered_to_float(mn);
                        p.arg[o] = kx | (kn << 16);
                        myacc[c] += s1;
                        myacc[p.NCP + c] += s2;
                    }
                }
            } else {
                float gx = 0.f, gy = 0.f, gz = 0.f;
                const bool want_sums = p.sums1 != nullptr;
                for (int cb = 0; cb < p.NCP; cb += 32) {
                    uint32_t v[32];
                    tc::tmem_ld_32x32(tacc + cb, v);
                    tc::tmem_ld_wait();
                    float y1[32], dzv[32];
                    const float *up = p.U + urow * p.C1 + cb;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 u = (qvalid && cb + 4 * j < p.C1) ? __ldg(reinterpret_cast<const float4 *>(up) + j)
                                                                       : make_float4(0.f, 0.f, 0.f, 0.f);
                        y1[4 * j + 0] = u.x; y1[4 * j + 1] = u.y; y1[4 * j + 2] = u.z; y1[4 * j + 3] = u.w;
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int c = cb + j;   // < c1pad: padding channels carry zero parameters
                        const float y = fmaf(prm[4 * c1pad + c], dz, fmaf(prm[3 * c1pad + c], dy, fmaf(prm[2 * c1pad + c], dx, y1[j])));
                        const float z = fmaf(prm[0 * c1pad + c], y, prm[1 * c1pad + c]);
                        const bool live = qvalid && c < p.C1 && (!p.relu1 || z > 0.f);
                        const float d = live ? __uint_as_float(v[j]) + prm[5 * c1pad + c] : 0.f;
                        const float ds = prm[0 * c1pad + c] * d;
                        gx = fmaf(ds, prm[2 * c1pad + c], gx);
                        gy = fmaf(ds, prm[3 * c1pad + c], gy);
                        gz = fmaf(ds, prm[4 * c1pad + c], gz);
                        y1[j] = y;
                        dzv[j] = d;
                    }
                    if (qvalid) {
                        float *dst = p.dU_pm + urow * p.C1 + cb;
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            if (cb + j < p.C1)
                                red_add_v4(dst + j, prm[0 * c1pad + cb + j] * dzv[j], prm[0 * c1pad + cb + j + 1] * dzv[j + 1],
                                           prm[0 * c1pad + cb + j + 2] * dzv[j + 2], prm[0 * c1pad + cb + j + 3] * dzv[j + 3]);
                    }
                    if (want_sums) {
                        const int c = cb + lane;
                        float t[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) t[j] = dzv[j];
                        const float a0 = colsum32(t, lane);
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            t[j] = dzv[j] * (y1[j] - prm[6 * c1pad + cb + j]) * prm[7 * c1pad + cb + j];
                        const float a1 = colsum32(t, lane);
#pragma unroll
                        for (int j = 0; j < 32; ++j) t[j] = dzv[j] * dx;
                        const float a2 = colsum32(t, lane);
#pragma unroll
                        for (int j = 0; j < 32; ++j) t[j] = dzv[j] * dy;
                        const float a3 = colsum32(t, lane);
#pragma unroll
                        for (int j = 0; j < 32; ++j) t[j] = dzv[j] * dz;
                        const float a4 = colsum32(t, lane);
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float z = fmaf(prm[0 * c1pad + cb + j], y1[j], prm[1 * c1pad + cb + j]);
                            t[j] = (qvalid && cb + j < p.C1) ? (p.relu1 ? fmaxf(z, 0.f) : z) : 0.f;
                        }
                        const float a5 = colsum32(t, lane);
                        if (c < c1pad) {
                            myacc[0 * c1pad + c] += a0;
                            myacc[1 * c1pad + c] += a1;
                            myacc[2 * c1pad + c] += a2;
                            myacc[3 * c1pad + c] += a3;
                            myacc[4 * c1pad + c] += a4;
                            myacc[5 * c1pad + c] += a5;
                        }
                    }
                }
                if (p.dxyz != nullptr && qvalid) {
                    float *d = p.dxyz + urow * 3;
                    atomicAdd(d, gx); atomicAdd(d + 1, gy); atomicAdd(d + 2, gz);
                }
                if (p.dq != nullptr) {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        gx += __shfl_xor_sync(0xffffffffu, gx, o);
                        gy += __shfl_xor_sync(0xffffffffu, gy, o);
                        gz += __shfl_xor_sync(0xffffffffu, gz, o);
                    }
                    if (lane == 0 && qvalid) {
                        float *d = p.dq + q * 3;
                        atomicAdd(d, -gx); atomicAdd(d + 1, -gy); atomicAdd(d + 2, -gz);
                    }
                }
            }
            tc::tc_fence_before();
            tc::mbar_arrive(&tempty_bar[a]);
        }
        // flush this warp's channel sums
        if (!BWD) {
            if (p.stats != nullptr)
                for (int c = lane; c < p.Cn; c += 32) {
                    atomicAdd(p.stats + 2 * c, (double)myacc[c]);
                    atomicAdd(p.stats + 2 * c + 1, (double)myacc[p.NCP + c]);
                }
        } else if (p.sums1 != nullptr) {
            for (int c = lane; c < p.C1; c += 32) {
                atomicAdd(p.sums1 + 2 * c, (double)myacc[0 * c1pad + c]);
                atomicAdd(p.sums1 + 2 * c + 1, (double)myacc[1 * c1pad + c]);
                if (p.sdzdp != nullptr) {
                    atomicAdd(p.sdzdp + 3 * c + 0, (double)myacc[2 * c1pad + c]);
                    atomicAdd(p.sdzdp + 3 * c + 1, (double)myacc[3 * c1pad + c]);
                    atomicAdd(p.sdzdp + 3 * c + 2, (double)myacc[4 * c1pad + c]);
                }
                if (p.sh1 != nullptr) atomicAdd(p.sh1 + c, (double)myacc[5 * c1pad + c]);
            }
        }
    }

    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, 512);
    }
}

// ---- sa2_wgrad_kernel: D[(C2 + C1) x C1] += [one-hot gs | h1]^T . h1 over all B*M*K rows ----------------------------
// Both operands are K-major with K = rows: a chunk is [channels][32 rows] (128 bytes per channel, 128B swizzle), written by
// the 8 producer warps with lane <-> row (conflict-free scalar stores).  grid = (row splits, slabs of 128 A-channels);
// every CTA keeps ONE accumulator (128 lanes x C1p columns) in TMEM for all its tiles and adds it to D with 128-bit REDs.
constexpr int WG_THREADS = 288;   // warps 0-7 producers (0-3 also the epilogue), warp 8 MMA issuer / TMEM allocator
constexpr int WG_MAX_STAGES = 8;

struct W2Params {
    long long rows;
    int C1, C2, c1pad, N, M, K, relu1, stages, tmem_cols, Mtot;   // Mtot = C2 + C1 (eval mode: C2)
    const float *U, *xyz, *new_xyz;
    const int32_t *idx;
    const float *Wp, *s1, *t1;
    const float *gs_pm;
    const uint8_t *ksel_pm;
    float *D;   // (Mtot, c1pad) pre-zeroed
};

__global__ void __launch_bounds__(WG_THREADS, 1) sa2_wgrad_kernel(const W2Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = align1024(smem_raw);
    const int c1pad = p.c1pad;
    const int stage_bytes = A_BYTES + c1pad * 128;
    uint8_t *tail = smem + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);   // [WG_MAX_STAGES]
    uint64_t *empty_bar = full_bar + WG_MAX_STAGES;            // [WG_MAX_STAGES]
    uint64_t *done_bar = empty_bar + WG_MAX_STAGES;            // [1]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(done_bar + 1);
    float *prm = reinterpret_cast<float *>(tail + 256);        // [5][c1pad]: s1, t1, wx, wy, wz

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int slab = blockIdx.y;
    const long long ntiles = (p.rows + BM - 1) / BM;

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            tc::mbar_init(&full_bar[s], 256);
            tc::mbar_init(&empty_bar[s], 1);
        }
        tc::mbar_init(done_bar, 1);
        tc::fence_barrier_init();
    }
    if (warp == 8) tc::tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
    for (int c = tid; c < c1pad; c += WG_THREADS) {
        const bool ok = c < p.C1;
        prm[0 * c1pad + c] = ok ? p.s1[c] : 0.f;
        prm[1 * c1pad + c] = ok ? p.t1[c] : 0.f;
        prm[2 * c1pad + c] = ok ? p.Wp[c * 3 + 0] : 0.f;
        prm[3 * c1pad + c] = ok ? p.Wp[c * 3 + 1] : 0.f;
        prm[4 * c1pad + c] = ok ? p.Wp[c * 3 + 2] : 0.f;
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        if (lane == 0) {
            const uint32_t idesc = tc::idesc_tf32_f32(BM, c1pad);
            uint32_t it = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                for (int kq = 0; kq < 4; ++kq, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_addr = tc::smem_u32(smem + (size_t)s * stage_bytes);
                    const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        tc::umma_tf32(tmem_base, tc::smem_desc_sw128(a_addr + k * 32), tc::smem_desc_sw128(b_addr + k * 32),
                                      idesc, (it | (uint32_t)k) != 0);
                    tc::umma_commit(&empty_bar[s]);
                }
            }
            tc::umma_commit(done_bar);
        }
    } else {
        // ---- producers: warp w writes B' channels [w*CB, (w+1)*CB) and A' slab channels [w*16, w*16+16) of every chunk
        const int CB = c1pad / 8;
        uint32_t it = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            for (int kq = 0; kq < 4; ++kq, ++it) {
                const int s = it % p.stages;
                const uint32_t ph = (it / p.stages) & 1;
                const long long grow = tile * BM + kq * 32 + lane;
                const bool valid = grow < p.rows;
                long long q = 0, urow = 0;
                int krow = 0;
                float dx = 0.f, dy = 0.f, dz = 0.f;
                if (valid) {
                    q = grow / p.K;
                    const long long b = q / p.M;
                    const int n = __ldg(p.idx + grow);
                    urow = b * p.N + n;
                    const float *pp = p.xyz + urow * 3;
                    const float *qq = p.new_xyz + q * 3;
                    dx = __ldg(pp + 0) - __ldg(qq + 0);
                    dy = __ldg(pp + 1) - __ldg(qq + 1);
                    dz = __ldg(pp + 2) - __ldg(qq + 2);
                    krow = (int)(grow - q * p.K);
                }
                const float *urowp = p.U + urow * p.C1;
                // values first (global loads in flight), then wait for the stage, then the stores
                float4 bv[8];   // CB <= 32 channels -> up to 8 quads
                float4 av[4];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = warp * CB + 4 * j;
                    bv[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (4 * j < CB && valid && c < p.C1)
                        bv[j] = h1_quad(prm, c1pad, c, __ldg(reinterpret_cast<const float4 *>(urowp + c)), dx, dy, dz, p.relu1);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int ch = slab * BM + warp * 16 + 4 * j;
                    av[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (valid) {
                        if (ch < p.C2) av[j] = onehot_quad(p.gs_pm, p.ksel_pm, q, p.C2, ch, krow);
                        else if (ch < p.Mtot)
                            av[j] = h1_quad(prm, c1pad, ch - p.C2, __ldg(reinterpret_cast<const float4 *>(urowp + ch - p.C2)), dx,
                                            dy, dz, p.relu1);
                    }
                }
                tc::mbar_wait(&empty_bar[s], ph ^ 1);
                const uint32_t a_base = tc::smem_u32(smem + (size_t)s * stage_bytes);
                const uint32_t b_base = a_base + A_BYTES;
                const uint32_t unit = (uint32_t)(lane >> 2), sub = (uint32_t)(lane & 3) * 4;
                auto put = [&](uint32_t base, int ch, float v) {
                    const uint32_t addr = base + (uint32_t)ch * 128 + ((unit ^ (uint32_t)(ch & 7)) << 4) + sub;
                    asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(tc::to_tf32(v)) : "memory");
                };
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (4 * j < CB) {
                        const int c = warp * CB + 4 * j;
                        put(b_base, c, bv[j].x); put(b_base, c + 1, bv[j].y); put(b_base, c + 2, bv[j].z); put(b_base, c + 3, bv[j].w);
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int ch = warp * 16 + 4 * j;
                    put(a_base, ch, av[j].x); put(a_base, ch + 1, av[j].y); put(a_base, ch + 2, av[j].z); put(a_base, ch + 3, av[j].w);
                }
                tc::fence_proxy_async_smem();
                tc::mbar_arrive(&full_bar[s]);
            }
        }
        if (warp < 4) {
            // ---- epilogue: TMEM lane = slab channel, columns = C1 channels
            tc::mbar_wait(done_bar, 0);
            tc::tc_fence_after();
            const int row = slab * BM + warp * 32 + lane;
            const uint32_t tacc = tmem_base + ((uint32_t)(warp * 32) << 16);
            for (int cb = 0; cb < c1pad; cb += 32) {
                uint32_t v[32];
                tc::tmem_ld_32x32(tacc + cb, v);
                tc::tmem_ld_wait();
                if (row < p.Mtot) {
                    float *dst = p.D + (size_t)row * c1pad + cb;
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        red_add_v4(dst + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                   __uint_as_float(v[j + 3]));
                }
            }
            tc::tc_fence_before();
        }
    }
    __syncthreads();
    if (warp == 8) {
        __syncwarp();
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
    }
}

// ---- host side -----------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn s2_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

// K-major fp32 weights W (rows, Kp) read as TF32; box = 32 columns x box_rows, 128B swizzle; rows past `rows` read as zero
int s2_weight_map(CUtensorMap *map, const float *W, int rows, int Kp, int box_rows) {
    EncodeTiledFn fn = s2_encode_fn();
    if (!fn) return APB_E_UNSUPPORTED;
    const cuuint64_t gdim[2] = {(cuuint64_t)Kp, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)Kp * 4};
    const cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 2, const_cast<float *>(W), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? APB_OK : APB_E_INVALID;
}

template <bool BWD>
int launch_rows(S2Params p, const float *W, int wrows_valid, int Kp, cudaStream_t st) {
    if (p.rows <= 0) return APB_OK;
    if ((Kp & 31) != 0 || Kp <= 0 || (reinterpret_cast<uintptr_t>(W) & 15) != 0) return APB_E_INVALID;
    p.kchunks = Kp / 32;
    p.NCP = (p.Cn + 15) / 16 * 16;
    if (p.NCP > 512) return APB_E_UNSUPPORTED;
    p.wrows = p.NCP <= 256 ? p.NCP : 512;
    const int stage_bytes = A_BYTES + p.wrows * 128;
    const size_t wacc = (size_t)4 * (BWD ? NSUM * p.c1pad : 2 * p.NCP) * 4;
    const size_t extra = 1024 /*align slack*/ + 256 /*barriers*/ + (size_t)NPRM * p.c1pad * 4 + wacc;
    int stages = (int)((227 * 1024 - extra) / stage_bytes);
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    stages = stages / NGROUPS * NGROUPS;  // each producer group owns its own in-order ring
    if (stages < NGROUPS) return APB_E_UNSUPPORTED;
    p.stages = stages;
    size_t smem = (size_t)stages * stage_bytes + extra;
    if (smem < 120 * 1024) smem = 120 * 1024;   // one CTA per SM: every CTA allocates all 512 TMEM columns
    CUtensorMap map;
    const int rc = s2_weight_map(&map, W, wrows_valid, Kp, p.wrows > 256 ? 256 : p.wrows);
    if (rc != APB_OK) return rc;
    auto kern = sa2_rows_kernel<BWD>;
    APB_CUDA(cudaFuncSetAttribute((const void *)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long ntiles = (p.rows + BM - 1) / BM;
    kern<<<(unsigned)(ntiles < APB_NUM_SMS ? ntiles : APB_NUM_SMS), NTHREADS, smem, st>>>(map, p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

bool s2_shapes_ok(int B, int N, int M, int K, int C1, int C2) {
    return B > 0 && N > 0 && M > 0 && K == 32 && C1 > 0 && C2 > 0 && (C1 & 3) == 0 && (C2 & 3) == 0 && C1 <= 512 && C2 <= 512;
}

}  // namespace

extern "C" int apb_sa2_fwd_tf32(const float *u_pm, const float *xyz, const float *new_xyz, const int32_t *idx,
                                const float *wp, const float *s1, const float *t1, int relu1, const float *w2, float *vmax,
                                float *vmin, int32_t *arg, double *stats, int B, int N, int M, int K, int C1, int C2,
                                int K2p, apb_stream_t stream) {
    if (B < 0 || N <= 0 || M < 0 || K <= 0 || C1 <= 0 || C2 <= 0 || K2p < C1) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!u_pm || !xyz || !new_xyz || !idx || !wp || !s1 || !t1 || !w2 || !vmax || !vmin || !arg) return APB_E_INVALID;
    if (!s2_shapes_ok(B, N, M, K, C1, C2) || (reinterpret_cast<uintptr_t>(u_pm) & 15) != 0) return APB_E_UNSUPPORTED;
    S2Params p{};
    p.rows = (long long)B * M * K;
    p.Cn = C2;
    p.C1 = C1; p.C2 = C2; p.c1pad = K2p; p.N = N; p.M = M; p.K = K; p.relu1 = relu1;
    p.U = u_pm; p.xyz = xyz; p.new_xyz = new_xyz; p.idx = idx; p.Wp = wp; p.s1 = s1; p.t1 = t1;
    p.vmax = vmax; p.vmin = vmin; p.arg = arg; p.stats = stats;
    return launch_rows<false>(p, w2, C2, K2p, (cudaStream_t)stream);
}

// wb (C1, Kp) = [ W2^T zero-padded to kc_sparse*32 columns | G zero-padded ]  (eval mode: the first block only, vvec = nullptr)
extern "C" int apb_sa2_bwd_rows_tf32(const float *u_pm, const float *xyz, const float *new_xyz, const int32_t *idx,
                                     const float *wp, const float *s1, const float *t1, int relu1, const float *gs_pm,
                                     const uint8_t *ksel_pm, const float *wb, int Kp, int kc_sparse, const float *vvec,
                                     const float *mean1, const float *rstd1, float *du_pm, double *sums1, double *sdzdp,
                                     double *sh1, float *dxyz, float *dq, int B, int N, int M, int K, int C1, int C2,
                                     apb_stream_t stream) {
    if (B < 0 || N <= 0 || M < 0 || K <= 0 || C1 <= 0 || C2 <= 0 || kc_sparse * 32 < C2 || Kp < kc_sparse * 32) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!u_pm || !xyz || !new_xyz || !idx || !wp || !s1 || !t1 || !gs_pm || !ksel_pm || !wb || !du_pm) return APB_E_INVALID;
    if (sums1 != nullptr && (!mean1 || !rstd1)) return APB_E_INVALID;
    if (!s2_shapes_ok(B, N, M, K, C1, C2) || (reinterpret_cast<uintptr_t>(u_pm) & 15) != 0 ||
        (reinterpret_cast<uintptr_t>(gs_pm) & 15) != 0 || (reinterpret_cast<uintptr_t>(ksel_pm) & 3) != 0 ||
        (reinterpret_cast<uintptr_t>(du_pm) & 15) != 0)
        return APB_E_UNSUPPORTED;
    const int c1pad = (C1 + 31) / 32 * 32;
    if (Kp != kc_sparse * 32 && Kp != kc_sparse * 32 + c1pad) return APB_E_INVALID;
    S2Params p{};
    p.rows = (long long)B * M * K;
    p.Cn = C1;
    p.kc_sparse = kc_sparse;
    p.C1 = C1; p.C2 = C2; p.c1pad = c1pad; p.N = N; p.M = M; p.K = K; p.relu1 = relu1;
    p.U = u_pm; p.xyz = xyz; p.new_xyz = new_xyz; p.idx = idx; p.Wp = wp; p.s1 = s1; p.t1 = t1;
    p.gs_pm = gs_pm; p.ksel_pm = ksel_pm; p.vvec = vvec; p.mean1 = mean1; p.rstd1 = rstd1;
    p.dU_pm = du_pm; p.sums1 = sums1; p.sdzdp = sdzdp; p.sh1 = sh1; p.dxyz = dxyz; p.dq = dq;
    return launch_rows<true>(p, wb, C1, Kp, (cudaStream_t)stream);
}

// d_out (C2 [+ C1], c1pad) pre-zeroed: rows [0, C2) = S = sum_r onehot_gs[r]^T h1[r]; with_gram: rows [C2, C2 + C1) = H
extern "C" int apb_sa2_wgrad_tf32(const float *u_pm, const float *xyz, const float *new_xyz, const int32_t *idx,
                                  const float *wp, const float *s1, const float *t1, int relu1, const float *gs_pm,
                                  const uint8_t *ksel_pm, float *d_out, int with_gram, int B, int N, int M, int K, int C1,
                                  int C2, apb_stream_t stream) {
    if (B < 0 || N <= 0 || M < 0 || K <= 0 || C1 <= 0 || C2 <= 0) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!u_pm || !xyz || !new_xyz || !idx || !wp || !s1 || !t1 || !gs_pm || !ksel_pm || !d_out) return APB_E_INVALID;
    if (!s2_shapes_ok(B, N, M, K, C1, C2) || C1 > 256 || (reinterpret_cast<uintptr_t>(u_pm) & 15) != 0 ||
        (reinterpret_cast<uintptr_t>(gs_pm) & 15) != 0 || (reinterpret_cast<uintptr_t>(ksel_pm) & 3) != 0 ||
        (reinterpret_cast<uintptr_t>(d_out) & 15) != 0)
        return APB_E_UNSUPPORTED;
    W2Params p{};
    p.rows = (long long)B * M * K;
    p.C1 = C1; p.C2 = C2; p.c1pad = (C1 + 31) / 32 * 32; p.N = N; p.M = M; p.K = K; p.relu1 = relu1;
    p.Mtot = C2 + (with_gram ? C1 : 0);
    p.U = u_pm; p.xyz = xyz; p.new_xyz = new_xyz; p.idx = idx; p.Wp = wp; p.s1 = s1; p.t1 = t1;
    p.gs_pm = gs_pm; p.ksel_pm = ksel_pm; p.D = d_out;
    int cols = 32;
    while (cols < p.c1pad) cols *= 2;
    p.tmem_cols = cols;
    const int stage_bytes = A_BYTES + p.c1pad * 128;
    const size_t extra = 1024 + 256 + (size_t)5 * p.c1pad * 4;
    int stages = (int)((227 * 1024 - extra) / stage_bytes);
    if (stages > WG_MAX_STAGES) stages = WG_MAX_STAGES;
    if (stages < 2) return APB_E_UNSUPPORTED;
    p.stages = stages;
    const size_t smem = (size_t)stages * stage_bytes + extra;
    APB_CUDA(cudaFuncSetAttribute((const void *)sa2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long ntiles = (p.rows + BM - 1) / BM;
    const int nslab = (p.Mtot + BM - 1) / BM;
    long long gx = APB_NUM_SMS / nslab;
    if (gx < 1) gx = 1;
    if (gx > ntiles) gx = ntiles;
    sa2_wgrad_kernel<<<dim3((unsigned)gx, (unsigned)nslab), WG_THREADS, smem, (cudaStream_t)stream>>>(p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
