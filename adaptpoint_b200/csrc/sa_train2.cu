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
// forward, pass 1 (sa_train.cu sa_ustats_kernel): sum y1, sum y1^2 in closed form from U and the idx-only moments -> (s1, t1).
// forward, pass 2 (sa2_fwd_kernel): producer warps gather rows of U (8 lanes per row, coalesced), form h1 in registers and
//   store it as a TF32 tile (128 rows x 32 channels per stage, 128B swizzle); W2 chunks arrive by TMA.  The GEMM is issued
//   TRANSPOSED -- D^T[channel, row] = W2 . h1^T, weights as the M = 128 operand, the row tile as N = 128 -- so that in
//   TMEM a lane is a channel and the 32 columns of a column block are the K = 32 neighbours of one query: the epilogue
//   thread of a channel scans them sequentially in registers for max / min with their k and sum y2, sum y2^2, with no
//   cross-lane traffic at all.  (Fewer than 128 channels: the weight rows are replicated across the lane quarters so that
//   every epilogue warp has work -- each takes a share of the queries.)  Nothing of size B*C*M*K is written.
// backward.  With batch statistics the BatchNorm-2 gradient is dense in r:  dy2 = gs*[k == arg] + a2 + b2*y2, so
//     dh1[r, :] = W2^T (gs*[k == arg])[r, :] + W2^T a2 + (W2^T diag(b2) W2) h1[r, :]
//   = one GEMM with the A tile [ one-hot gs rows | h1 rows ] against [ W2^T | G ] (sa2_bwd_kernel, lane = row; eval mode:
//   the first block only).  Its epilogue (two groups of 4 warps, one per TMEM accumulator stage) recomputes the ReLU mask from U, forms dz1 = (dh1 + v) * [z1 > 0], scatters
//   s1 * dz1 into dU (128-bit RED), accumulates the BatchNorm-1 sums (sum dz1, sum dz1 * xhat1), sum dz1 * dp (-> dWp),
//   sum h1, and the coordinate gradients.  The dense BatchNorm-1 terms (a1 + b1 * y1) are linear in U and dp, so they take
//   the closed per-support-point form of the one-conv path (sa_train.cu: sa1_bwd_count / sa1_bwd_dense).
//   dW2 = sum_r dy2[r]^T h1[r] = S + a2 (sum h1)^T + diag(b2) W2 H  with  S = sum_r (gs one-hot)[r]^T h1[r],
//   H = sum_r h1[r] h1[r]^T:  sa2_wgrad_kernel accumulates [S ; H] = [one-hot gs | h1]^T h1 over all rows in TMEM; both
//   operands are MN-major (K = rows): the producer warps write [32 rows][32 channels] boxes in the SWIZZLE_128B_BASE32B
//   layout (the one TMA's 128B_ATOM_32B mode produces for tc_cm_gemm.cu).
// Numerics: TF32 operands, fp32 accumulation (the precision class of tcconv.py's training path).
#include <cuda.h>
#include <cfloat>
#include <climits>
#include "apb_common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int BM = 128;            // rows per tile (UMMA M)
constexpr int A_BYTES = BM * 128;  // A chunk: 128 rows x 32 tf32
constexpr int NGROUPS = 2;         // forward: A-producer groups of 128 threads (alternate tiles)
constexpr int PROD0 = 64;
constexpr int EPI0 = PROD0 + 128 * NGROUPS;
constexpr int NTHREADS = EPI0 + 128;
constexpr int MAX_STAGES = 6;
// backward rows kernel: two producer groups, two epilogue groups (one per TMEM accumulator stage), 16-channel epilogue blocks
constexpr int BW_THREADS = PROD0 + 128 * NGROUPS + 256;
constexpr int EB = 16;
constexpr int NPRM = 8;            // s1, t1, wx, wy, wz, v, mean1, rstd1
constexpr int NSUM = 6;            // backward channel sums: dz, dz*xhat, dz*dpx, dz*dpy, dz*dpz, h1

struct S2Params {
    long long rows;        // B*M*K
    int kchunks;           // K chunks (of 32) per tile: forward C1p/32; backward C2p/32 (+ C1p/32 in training mode)
    int kc_sparse;         // backward: leading chunks built from (gs, ksel); forward: 0
    int NCP, stages;       // backward: accumulator columns = C1 rounded up to 16 (<= 256)
    int nmb;               // forward: 128-lane channel blocks of the accumulator: ceil(C2 / 128), <= 4
    int Cq;                // forward, nmb == 1: channels per replica (32, 64 or 128; 128 / Cq replicas share the queries)
    int Cn;                // valid accumulator channels: forward C2, backward C1
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
    int dbg;               // benchmarks only (apb_tune_sa2): 1 skip the dU REDs, 2 the epilogue's U loads, 4 the sums,
                           // 8 the producers' U loads, 16 their h1 arithmetic, 32 the epilogue's row metadata
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

// 16-column variant: x[j] of lane l = (row l, column j), j < 16.  Returns in every lane l the sum over the 32 rows of
// column l & 15.
template <int S>
__device__ __forceinline__ void colsum16_step(float (&x)[16], int lane) {
    const bool up = (lane & S) != 0;
#pragma unroll
    for (int i = 0; i < S; ++i) {
        const float send = up ? x[i] : x[i + S];
        const float keep = up ? x[i + S] : x[i];
        x[i] = keep + __shfl_xor_sync(0xffffffffu, send, S);
    }
}
__device__ __forceinline__ float colsum16(float (&x)[16], int lane) {
    colsum16_step<8>(x, lane);
    colsum16_step<4>(x, lane);
    colsum16_step<2>(x, lane);
    colsum16_step<1>(x, lane);
    return x[0] + __shfl_xor_sync(0xffffffffu, x[0], 16);
}

__device__ __forceinline__ void red_add_v4(float *dst, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// h1 of 4 consecutive channels of one row: relu(s1 * (u + Wp . dp) + t1); prm = [NPRM][c1pad] in shared memory
__device__ __forceinline__ float4 h1_quad(const float *prm, int c1pad, int c, float4 u, float dx, float dy, float dz,
                                          int relu1) {
    const float4 s = tc::lds_f4_ro(prm + 0 * c1pad + c);
    const float4 t = tc::lds_f4_ro(prm + 1 * c1pad + c);
    const float4 wx = tc::lds_f4_ro(prm + 2 * c1pad + c);
    const float4 wy = tc::lds_f4_ro(prm + 3 * c1pad + c);
    const float4 wz = tc::lds_f4_ro(prm + 4 * c1pad + c);
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


__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- sa2_fwd_kernel: persistent, one CTA per SM, 448 threads, warp-specialised ---------------------------------------
//    warp 0      TMA producer for the weight chunk of every pipeline stage (the M = 128 operand)
//    warp 1      TMEM allocator + single-thread tcgen05.mma issuer (UMMA 128 x 128 x 8, tf32 -> fp32)
//    warps 2-9   row-tile producers (2 groups of 128 threads, alternating tiles; 8 lanes per row)
//    warps 10-13 epilogue, thread <-> channel (TMEM lane); a column block of 32 = the K = 32 neighbours of one query
__global__ void __launch_bounds__(NTHREADS, 1)
sa2_fwd_kernel(const __grid_constant__ CUtensorMap wmap, const S2Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = align1024(smem_raw);
    const int w_bytes = p.nmb * BM * 128;
    const int stage_bytes = A_BYTES + w_bytes;
    uint8_t *tail = smem + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);  // [MAX_STAGES]
    uint64_t *empty_bar = full_bar + MAX_STAGES;              // [MAX_STAGES]
    uint64_t *tfull_bar = empty_bar + MAX_STAGES;             // [2]
    uint64_t *tempty_bar = tfull_bar + 2;                     // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);
    float *prm = reinterpret_cast<float *>(tail + 256);       // [5][c1pad]
    const int c1pad = p.c1pad;
    int *meta_row = reinterpret_cast<int *>(prm + 5 * c1pad);                  // [NGROUPS][2][BM]
    float *meta_dp = reinterpret_cast<float *>(meta_row + NGROUPS * 2 * BM);   // [NGROUPS][2][3][BM]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int acc_stages = p.nmb <= 2 ? 2 : 1;
    const uint32_t spg = p.stages / NGROUPS;  // smem stages per producer group
    const long long ntiles = (p.rows + BM - 1) / BM;

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&wmap);
        for (int s = 0; s < p.stages; ++s) {
            tc::mbar_init(&full_bar[s], 128 + 1);  // 128 producer arrivals + the TMA thread's expect_tx arrive
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
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer: weight chunk of every stage =====================
        if (lane == 0) {
            const int nrep = p.nmb == 1 ? BM / p.Cq : 1;
            uint32_t lt = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
                const int grp = lt % NGROUPS;
                uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;  // chunk counter of this group's own stage ring
                for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                    const int s = grp * spg + it % spg;
                    const uint32_t ph = (it / spg) & 1;
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t *wdst = smem + (size_t)s * stage_bytes + A_BYTES;
                    tc::mbar_arrive_expect_tx(&full_bar[s], (uint32_t)w_bytes);
                    if (p.nmb == 1) {
                        for (int r = 0; r < nrep; ++r)   // the same Cq weight rows under every lane group of Cq lanes
                            tc::tma_load_2d(wdst + (size_t)r * p.Cq * 128, &wmap, kc * 32, 0, &full_bar[s]);
                    } else {
                        for (int mb = 0; mb < p.nmb; ++mb)
                            tc::tma_load_2d(wdst + (size_t)mb * BM * 128, &wmap, kc * 32, mb * BM, &full_bar[s]);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: D^T (128 channels x 128 rows) += W chunk . tile chunk^T =====================
        if (lane == 0) {
            const uint32_t idesc = tc::idesc_tf32_f32(BM, BM);
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
                    const uint32_t t_addr = tc::smem_u32(smem + (size_t)s * stage_bytes);
                    const uint32_t w_addr = t_addr + A_BYTES;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {  // 4 x 32 bytes per chunk: UMMA_K = 8 tf32
                        const uint64_t bd = tc::smem_desc_sw128(t_addr + k * 32);
                        for (int mb = 0; mb < p.nmb; ++mb)
                            tc::umma_tf32(tacc + mb * BM, tc::smem_desc_sw128(w_addr + mb * BM * 128 + k * 32), bd, idesc,
                                          (kc | k) != 0);
                    }
                    tc::umma_commit(&empty_bar[s]);  // smem stage reusable once these MMAs have read it
                }
                tc::umma_commit(&tfull_bar[a]);      // accumulator complete -> epilogue
            }
        }
    } else if (warp < 2 + 4 * NGROUPS) {
        // ===================== row-tile producers: group g fills tiles lt % NGROUPS == g =====================
        const int grp = (warp - 2) >> 2;
        const int pt = (tid - PROD0) & 127;  // thread of the group; also the row whose metadata it fetches
        int *g_meta_row = meta_row + grp * 2 * BM;
        float *g_meta_dp = meta_dp + grp * 6 * BM;
        uint32_t lt = 0;
        // row metadata (support-point row of U, dp = p[idx] - q), fetched one tile ahead so that the dependent
        // idx -> xyz loads overlap the previous tile's chunk work
        int nx_urow = -1;
        float nx_dx = 0.f, nx_dy = 0.f, nx_dz = 0.f;
        auto fetch_meta = [&](long long tile) {
            nx_urow = -1;
            nx_dx = nx_dy = nx_dz = 0.f;
            const long long grow = tile * BM + pt;
            if (tile < ntiles && grow < p.rows) {
                const long long q = grow >> 5;  // b*M + m (K == 32)
                const long long b = (int)q / p.M;
                const int n = __ldg(p.idx + grow);
                const float *pp = p.xyz + (b * p.N + n) * 3;
                const float *qq = p.new_xyz + q * 3;
                nx_dx = __ldg(pp + 0) - __ldg(qq + 0);
                nx_dy = __ldg(pp + 1) - __ldg(qq + 1);
                nx_dz = __ldg(pp + 2) - __ldg(qq + 2);
                nx_urow = (int)(b * p.N + n);
            }
        };
        fetch_meta((long long)blockIdx.x + (long long)grp * gridDim.x);
        const int pw = pt >> 5, l8 = lane & 7, lr = lane >> 3;   // 8 lanes per row, 4 channels (16 B) per lane
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            if ((int)(lt % NGROUPS) != grp) continue;
            uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;  // this group's own stage ring (in-order => parity-safe)
            const int mb = (lt / NGROUPS) & 1;
            tc::sts_i(g_meta_row + mb * BM + pt, nx_urow);
            tc::sts_f(g_meta_dp + (mb * 3 + 0) * BM + pt, nx_dx);
            tc::sts_f(g_meta_dp + (mb * 3 + 1) * BM + pt, nx_dy);
            tc::sts_f(g_meta_dp + (mb * 3 + 2) * BM + pt, nx_dz);
            named_bar_sync(1 + grp, 128);
            fetch_meta(tile + (long long)NGROUPS * gridDim.x);  // next tile of this group, consumed next round
            int ur[8];
#pragma unroll
            for (int step = 0; step < 8; ++step) ur[step] = tc::lds_i(g_meta_row + mb * BM + pw * 32 + step * 4 + lr);
            for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                const int s = grp * spg + it % spg;
                const uint32_t ph = (it / spg) & 1;
                const uint32_t abase = tc::smem_u32(smem + (size_t)s * stage_bytes);
                const int c0 = kc * 32 + l8 * 4;
                const bool cvalid = c0 < p.C1;   // C1 % 4 == 0: a lane's 4 channels are all valid or all padding
                float4 v[8];
#pragma unroll
                for (int step = 0; step < 8; ++step)   // rows of U for the 8 steps, issued up front
                    v[step] = (ur[step] >= 0 && cvalid)
                                  ? __ldg(reinterpret_cast<const float4 *>(p.U + (size_t)ur[step] * p.C1 + c0))
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
                tc::mbar_wait(&empty_bar[s], ph ^ 1);
#pragma unroll
                for (int step = 0; step < 8; ++step) {
                    const int r = pw * 32 + step * 4 + lr;
                    if (ur[step] >= 0 && cvalid)
                        v[step] = h1_quad(prm, c1pad, c0, v[step], tc::lds_f(g_meta_dp + (mb * 3 + 0) * BM + r),
                                          tc::lds_f(g_meta_dp + (mb * 3 + 1) * BM + r), tc::lds_f(g_meta_dp + (mb * 3 + 2) * BM + r), p.relu1);
                    tc::st_shared_v4(abase + r * 128 + ((l8 ^ (r & 7)) << 4), tc::to_tf32(v[step].x), tc::to_tf32(v[step].y),
                                     tc::to_tf32(v[step].z), tc::to_tf32(v[step].w));
                }
                tc::fence_proxy_async_smem();
                tc::mbar_arrive(&full_bar[s]);
            }
        }
    } else {
        // ===================== epilogue (128 threads; warp w reads TMEM lanes 32*(w%4)..) =====================
        const int q4 = warp & 3;
        const int lane_g = q4 * 32 + lane;                 // TMEM lane == channel of the 128-lane block
        const int nrep = p.nmb == 1 ? BM / p.Cq : 1;
        const int cq = p.nmb == 1 ? lane_g % p.Cq : lane_g;
        const int share = p.nmb == 1 ? lane_g / p.Cq : 0;  // which of the nrep query shares this lane group takes
        float acc[4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = 0.f;
        uint32_t lt = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            const int a = lt % acc_stages;
            const uint32_t aph = (lt / acc_stages) & 1;
            tc::mbar_wait(&tfull_bar[a], aph);
            tc::tc_fence_after();
            const uint32_t tacc = tmem_base + (uint32_t)a * 256 + ((uint32_t)(q4 * 32) << 16);
#pragma unroll
            for (int mb = 0; mb < 4; ++mb) {
                if (mb >= p.nmb) break;
                const int c = mb * BM + cq;
                const bool cvalid = c < p.Cn;
                for (int qb = share; qb < 4; qb += nrep) {
                    const long long row0 = tile * BM + qb * 32;     // the 32 rows of this block: the K neighbours of ONE query
                    const bool qvalid = row0 < p.rows;
                    uint32_t v[32];
                    tc::tmem_ld_32x32(tacc + mb * BM + qb * 32, v);
                    tc::tmem_ld_wait();
                    float mx = -FLT_MAX, mn = FLT_MAX, s = 0.f, s2 = 0.f;
                    int kx = 0, kn = 0;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float y = __uint_as_float(v[j]);
                        s += y;
                        s2 = fmaf(y, y, s2);
                        if (y > mx) { mx = y; kx = j; }   // strict: first k wins ties (torch.max order)
                        if (y < mn) { mn = y; kn = j; }
                    }
                    if (qvalid && cvalid) {
                        const long long q = row0 >> 5;
                        const long long b = (int)q / p.M, m = q - b * p.M;
                        const size_t o = ((size_t)b * p.C2 + c) * p.M + m;
                        p.vmax[o] = mx;
                        p.vmin[o] = mn;
                        p.arg[o] = kx | (kn << 16);
                        acc[mb][0] += s;
                        acc[mb][1] += s2;
                    }
                }
            }
            tc::tc_fence_before();
            tc::mbar_arrive(&tempty_bar[a]);
        }
        if (p.stats != nullptr) {
#pragma unroll
            for (int mb = 0; mb < 4; ++mb) {
                if (mb >= p.nmb) break;
                const int c = mb * BM + cq;
                if (c < p.Cn) {
                    atomicAdd(p.stats + 2 * c, (double)acc[mb][0]);
                    atomicAdd(p.stats + 2 * c + 1, (double)acc[mb][1]);
                }
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

// ---- sa2_bwd_kernel: persistent, one CTA per SM, 576 threads, warp-specialised ---------------------------------------
//    warp 0       TMA producer for the [W2^T | G] chunk of every pipeline stage (the N operand, C1 rows)
//    warp 1       TMEM allocator + single-thread tcgen05.mma issuer (UMMA 128 x NCP x 8, tf32 -> fp32)
//    warps 2-9    row-tile producers (2 groups of 128 threads, alternating tiles, 8 lanes per row): one-hot gs chunks,
//                 then h1 chunks
//    warps 10-13  epilogue of the even tiles (TMEM accumulator stage 0), thread <-> row (TMEM lane)
//    warps 14-17  epilogue of the odd tiles  (TMEM accumulator stage 1)
// The epilogue works on column blocks of 16 channels so that the whole kernel fits 112 registers per thread.
// COORDS: also the gradient to the coordinates (dxyz, dq)
template <bool COORDS>
__global__ void __launch_bounds__(BW_THREADS, 1)
sa2_bwd_kernel(const __grid_constant__ CUtensorMap wmap, const S2Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = align1024(smem_raw);
    const int w_bytes = p.NCP * 128;
    const int stage_bytes = A_BYTES + w_bytes;
    uint8_t *tail = smem + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);  // [MAX_STAGES]
    uint64_t *empty_bar = full_bar + MAX_STAGES;              // [MAX_STAGES]
    uint64_t *tfull_bar = empty_bar + MAX_STAGES;             // [2]
    uint64_t *tempty_bar = tfull_bar + 2;                     // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);
    float *prm = reinterpret_cast<float *>(tail + 256);       // [NPRM][c1pad]
    const int c1pad = p.c1pad;
    int *meta_row = reinterpret_cast<int *>(prm + NPRM * c1pad);               // producers: [NGROUPS][2][BM]
    float *meta_dp = reinterpret_cast<float *>(meta_row + NGROUPS * 2 * BM);   // [NGROUPS][2][3][BM]
    float *sacc = meta_dp + NGROUPS * 6 * BM;                                  // [NSUM][c1pad] channel sums of the CTA

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t spg = p.stages / NGROUPS;   // smem stages per producer group
    const long long ntiles = (p.rows + BM - 1) / BM;
    const bool want_sums = p.sums1 != nullptr;

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&wmap);
        for (int s = 0; s < p.stages; ++s) {
            tc::mbar_init(&full_bar[s], 128 + 1);  // 128 producer arrivals + the TMA thread's expect_tx arrive
            tc::mbar_init(&empty_bar[s], 1);       // one tcgen05.commit
        }
        for (int a = 0; a < 2; ++a) {
            tc::mbar_init(&tfull_bar[a], 1);
            tc::mbar_init(&tempty_bar[a], 128);    // the 128 threads of the stage's epilogue group
        }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
    for (int c = tid; c < c1pad; c += BW_THREADS) {
        const bool ok = c < p.C1;
        prm[0 * c1pad + c] = ok ? p.s1[c] : 0.f;
        prm[1 * c1pad + c] = ok ? p.t1[c] : 0.f;
        prm[2 * c1pad + c] = ok ? p.Wp[c * 3 + 0] : 0.f;
        prm[3 * c1pad + c] = ok ? p.Wp[c * 3 + 1] : 0.f;
        prm[4 * c1pad + c] = ok ? p.Wp[c * 3 + 2] : 0.f;
        prm[5 * c1pad + c] = (ok && p.vvec != nullptr) ? p.vvec[c] : 0.f;
        prm[6 * c1pad + c] = (ok && p.mean1 != nullptr) ? p.mean1[c] : 0.f;
        prm[7 * c1pad + c] = (ok && p.rstd1 != nullptr) ? p.rstd1[c] : 0.f;
    }
    for (int i = tid; i < NSUM * c1pad; i += BW_THREADS) sacc[i] = 0.f;
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t lt = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
                const int grp = lt % NGROUPS;
                uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;   // chunk counter of this group's own stage ring
                for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                    const int s = grp * spg + it % spg;
                    const uint32_t ph = (it / spg) & 1;
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t *wdst = smem + (size_t)s * stage_bytes + A_BYTES;
                    tc::mbar_arrive_expect_tx(&full_bar[s], (uint32_t)w_bytes);
                    tc::tma_load_2d(wdst, &wmap, kc * 32, 0, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = tc::idesc_tf32_f32(BM, p.NCP);
            uint32_t lt = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
                const int grp = lt % NGROUPS;
                uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;
                const int a = lt & 1;
                const uint32_t aph = (lt >> 1) & 1;
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
                    for (int k = 0; k < 4; ++k)   // 4 x 32 bytes per chunk: UMMA_K = 8 tf32
                        tc::umma_tf32(tacc, tc::smem_desc_sw128(a_addr + k * 32), tc::smem_desc_sw128(w_addr + k * 32), idesc,
                                      (kc | k) != 0);
                    tc::umma_commit(&empty_bar[s]);  // smem stage reusable once these MMAs have read it
                }
                tc::umma_commit(&tfull_bar[a]);      // accumulator complete -> epilogue
            }
        }
    } else if (warp < 2 + 4 * NGROUPS) {
        // ===================== row-tile producers: group g fills tiles lt % NGROUPS == g =====================
        const int grp = (warp - 2) >> 2;
        const int pt = (tid - PROD0) & 127;   // also the row whose metadata this thread fetches
        int *g_meta_row = meta_row + grp * 2 * BM;
        float *g_meta_dp = meta_dp + grp * 6 * BM;
        int nx_urow = -1;
        float nx_dx = 0.f, nx_dy = 0.f, nx_dz = 0.f;
        auto fetch_meta = [&](long long tile) {
            nx_urow = -1;
            nx_dx = nx_dy = nx_dz = 0.f;
            const long long grow = tile * BM + pt;
            if (tile < ntiles && grow < p.rows) {
                const long long q = grow >> 5;  // b*M + m (K == 32)
                const long long b = (int)q / p.M;
                const int n = __ldg(p.idx + grow);
                const float *pp = p.xyz + (b * p.N + n) * 3;
                const float *qq = p.new_xyz + q * 3;
                nx_dx = __ldg(pp + 0) - __ldg(qq + 0);
                nx_dy = __ldg(pp + 1) - __ldg(qq + 1);
                nx_dz = __ldg(pp + 2) - __ldg(qq + 2);
                nx_urow = (int)(b * p.N + n);
            }
        };
        fetch_meta((long long)blockIdx.x + (long long)grp * gridDim.x);
        const int pw = pt >> 5, l8 = lane & 7, lr = lane >> 3;
        uint32_t lt = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            if ((int)(lt % NGROUPS) != grp) continue;
            uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;   // this group's own stage ring (in-order => parity-safe)
            const int mb = (lt / NGROUPS) & 1;
            tc::sts_i(g_meta_row + mb * BM + pt, nx_urow);
            tc::sts_f(g_meta_dp + (mb * 3 + 0) * BM + pt, nx_dx);
            tc::sts_f(g_meta_dp + (mb * 3 + 1) * BM + pt, nx_dy);
            tc::sts_f(g_meta_dp + (mb * 3 + 2) * BM + pt, nx_dz);
            named_bar_sync(1 + grp, 128);
            fetch_meta(tile + (long long)NGROUPS * gridDim.x);   // next tile of this group, consumed next round
            int ur[8];
#pragma unroll
            for (int step = 0; step < 8; ++step) ur[step] = tc::lds_i(g_meta_row + mb * BM + pw * 32 + step * 4 + lr);
            for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                const int s = grp * spg + it % spg;
                const uint32_t ph = (it / spg) & 1;
                const uint32_t abase = tc::smem_u32(smem + (size_t)s * stage_bytes);
                float4 v[8];
                if (kc < p.kc_sparse) {
                    const int d0 = kc * 32 + l8 * 4;
                    // K == 32: the 32 rows pw*32.. of the tile are the neighbours k = 0..31 of query tile*4 + pw: one load of
                    // that query's gs / ksel quad, 8 one-hot variants
                    const bool ok = ur[0] >= 0 && d0 < p.C2;
                    const long long qoff = (tile * 4 + pw) * p.C2 + d0;
                    const float4 g = ok ? __ldg(reinterpret_cast<const float4 *>(p.gs_pm + qoff)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    const uint32_t kq = ok ? __ldg(reinterpret_cast<const uint32_t *>(p.ksel_pm + qoff)) : 0xffffffffu;
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
#pragma unroll
                    for (int step = 0; step < 8; ++step) {
                        const int k = step * 4 + lr;
                        v[step].x = (int)(kq & 0xff) == k ? g.x : 0.f;
                        v[step].y = (int)((kq >> 8) & 0xff) == k ? g.y : 0.f;
                        v[step].z = (int)((kq >> 16) & 0xff) == k ? g.z : 0.f;
                        v[step].w = (int)(kq >> 24) == k ? g.w : 0.f;
                    }
                } else {
                    const int c0 = (kc - p.kc_sparse) * 32 + l8 * 4;
                    const bool cvalid = c0 < p.C1;   // C1 % 4 == 0: a lane's 4 channels are all valid or all padding
#pragma unroll
                    for (int step = 0; step < 8; ++step)   // rows of U for the 8 steps, issued up front
                        v[step] = (ur[step] >= 0 && cvalid && !(p.dbg & 8))
                                      ? __ldg(reinterpret_cast<const float4 *>(p.U + (size_t)ur[step] * p.C1 + c0))
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
#pragma unroll
                    for (int step = 0; step < 8; ++step) {
                        const int r = pw * 32 + step * 4 + lr;
                        if (ur[step] >= 0 && cvalid && !(p.dbg & 16))
                            v[step] = h1_quad(prm, c1pad, c0, v[step], tc::lds_f(g_meta_dp + (mb * 3 + 0) * BM + r),
                                              tc::lds_f(g_meta_dp + (mb * 3 + 1) * BM + r), tc::lds_f(g_meta_dp + (mb * 3 + 2) * BM + r),
                                              p.relu1);
                    }
                }
#pragma unroll
                for (int step = 0; step < 8; ++step) {
                    const int r = pw * 32 + step * 4 + lr;
                    tc::st_shared_v4(abase + r * 128 + ((l8 ^ (r & 7)) << 4), tc::to_tf32(v[step].x), tc::to_tf32(v[step].y),
                                     tc::to_tf32(v[step].z), tc::to_tf32(v[step].w));
                }
                tc::fence_proxy_async_smem();
                tc::mbar_arrive(&full_bar[s]);
            }
        }
    } else {
        // ===================== epilogue: group g = accumulator stage g = tiles with lt % 2 == g =====================
        const int grp = (warp - (2 + 4 * NGROUPS)) >> 2;
        const int q4 = warp & 3;     // TMEM lane quarter of this warp (warps 10..13 -> 2,3,0,1; 14..17 -> 2,3,0,1)
        uint32_t lt = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            if ((int)(lt & 1) != grp) continue;
            const uint32_t aph = (lt >> 1) & 1;
            const long long row0 = tile * BM + q4 * 32;      // the 32 rows of this warp: the K neighbours of ONE query
            const bool qvalid = row0 < p.rows;
            const long long q = row0 >> 5;
            const long long b = (int)q / p.M;
            long long urow = 0;
            float dx = 0.f, dy = 0.f, dz = 0.f;
            if (qvalid && !(p.dbg & 32)) {   // row metadata, loaded before the accumulator wait
                const int n = __ldg(p.idx + row0 + lane);
                urow = b * p.N + n;
                const float *pp = p.xyz + urow * 3;
                const float *qq = p.new_xyz + q * 3;
                dx = __ldg(pp + 0) - __ldg(qq + 0);
                dy = __ldg(pp + 1) - __ldg(qq + 1);
                dz = __ldg(pp + 2) - __ldg(qq + 2);
            }
            tc::mbar_wait(&tfull_bar[grp], aph);
            tc::tc_fence_after();
            const uint32_t tacc = tmem_base + (uint32_t)grp * 256 + ((uint32_t)(q4 * 32) << 16);
            float gx = 0.f, gy = 0.f, gz = 0.f;
            for (int cb = 0; cb < p.NCP; cb += EB) {
                float y1[EB], dzv[EB];
                const float *up = p.U + urow * p.C1 + cb;
#pragma unroll
                for (int j = 0; j < EB / 4; ++j) {   // this row of U, issued before the TMEM load completes
                    const float4 u = (qvalid && cb + 4 * j < p.C1 && !(p.dbg & 2)) ? __ldg(reinterpret_cast<const float4 *>(up) + j)
                                                                                   : make_float4(0.f, 0.f, 0.f, 0.f);
                    y1[4 * j + 0] = u.x; y1[4 * j + 1] = u.y; y1[4 * j + 2] = u.z; y1[4 * j + 3] = u.w;
                }
                uint32_t v[EB];
                tc::tmem_ld_32x16(tacc + cb, v);
                tc::tmem_ld_wait();
#pragma unroll
                for (int j4 = 0; j4 < EB; j4 += 4) {
                    const int c = cb + j4;   // < c1pad: padding channels carry zero parameters
                    const float4 ps = tc::lds_f4_ro(prm + 0 * c1pad + c);
                    const float4 pt1 = tc::lds_f4_ro(prm + 1 * c1pad + c);
                    const float4 pwx = tc::lds_f4_ro(prm + 2 * c1pad + c);
                    const float4 pwy = tc::lds_f4_ro(prm + 3 * c1pad + c);
                    const float4 pwz = tc::lds_f4_ro(prm + 4 * c1pad + c);
                    const float4 pvv = tc::lds_f4_ro(prm + 5 * c1pad + c);
                    const float s_[4] = {ps.x, ps.y, ps.z, ps.w}, t_[4] = {pt1.x, pt1.y, pt1.z, pt1.w};
                    const float wx_[4] = {pwx.x, pwx.y, pwx.z, pwx.w}, wy_[4] = {pwy.x, pwy.y, pwy.z, pwy.w};
                    const float wz_[4] = {pwz.x, pwz.y, pwz.z, pwz.w}, vv_[4] = {pvv.x, pvv.y, pvv.z, pvv.w};
                    float ds[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int j = j4 + e;
                        const float y = fmaf(wz_[e], dz, fmaf(wy_[e], dy, fmaf(wx_[e], dx, y1[j])));
                        const float z = fmaf(s_[e], y, t_[e]);
                        const bool live = qvalid && (c + e) < p.C1 && (!p.relu1 || z > 0.f);
                        const float d = live ? __uint_as_float(v[j]) + vv_[e] : 0.f;
                        ds[e] = s_[e] * d;
                        if (COORDS) {
                            gx = fmaf(ds[e], wx_[e], gx);
                            gy = fmaf(ds[e], wy_[e], gy);
                            gz = fmaf(ds[e], wz_[e], gz);
                        }
                        y1[j] = y;
                        dzv[j] = d;
                    }
                    if (qvalid && c < p.C1 && !(p.dbg & 1)) red_add_v4(p.dU_pm + urow * p.C1 + c, ds[0], ds[1], ds[2], ds[3]);
                }
                if (want_sums && !(p.dbg & 4) && cb < c1pad) {
                    float t[EB];
#pragma unroll
                    for (int j = 0; j < EB; ++j) t[j] = dzv[j];
                    const float a0 = colsum16(t, lane);
#pragma unroll
                    for (int j = 0; j < EB; ++j)
                        t[j] = dzv[j] * (y1[j] - tc::lds_f_ro(prm + 6 * c1pad + cb + j)) * tc::lds_f_ro(prm + 7 * c1pad + cb + j);
                    const float a1 = colsum16(t, lane);
#pragma unroll
                    for (int j = 0; j < EB; ++j) t[j] = dzv[j] * dx;
                    const float a2 = colsum16(t, lane);
#pragma unroll
                    for (int j = 0; j < EB; ++j) t[j] = dzv[j] * dy;
                    const float a3 = colsum16(t, lane);
#pragma unroll
                    for (int j = 0; j < EB; ++j) t[j] = dzv[j] * dz;
                    const float a4 = colsum16(t, lane);
#pragma unroll
                    for (int j = 0; j < EB; ++j) {
                        const float z = fmaf(tc::lds_f_ro(prm + 0 * c1pad + cb + j), y1[j], tc::lds_f_ro(prm + 1 * c1pad + cb + j));
                        t[j] = (qvalid && cb + j < p.C1) ? (p.relu1 ? fmaxf(z, 0.f) : z) : 0.f;
                    }
                    const float a5 = colsum16(t, lane);
                    if (lane < EB) {   // lane l holds the sums of channel cb + l
                        const int c = cb + lane;
                        tc::red_shared_f32(sacc + 0 * c1pad + c, a0);
                        tc::red_shared_f32(sacc + 1 * c1pad + c, a1);
                        tc::red_shared_f32(sacc + 2 * c1pad + c, a2);
                        tc::red_shared_f32(sacc + 3 * c1pad + c, a3);
                        tc::red_shared_f32(sacc + 4 * c1pad + c, a4);
                        tc::red_shared_f32(sacc + 5 * c1pad + c, a5);
                    }
                }
            }
            tc::tc_fence_before();
            tc::mbar_arrive(&tempty_bar[grp]);
            if (COORDS) {
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
        }
        if (want_sums) {   // all epilogue warps done -> flush the CTA's channel sums
            named_bar_sync(3, 256);
            for (int c = tid - (PROD0 + 128 * NGROUPS); c < p.C1; c += 256) {
                atomicAdd(p.sums1 + 2 * c, (double)sacc[0 * c1pad + c]);
                atomicAdd(p.sums1 + 2 * c + 1, (double)sacc[1 * c1pad + c]);
                if (p.sdzdp != nullptr) {
                    atomicAdd(p.sdzdp + 3 * c + 0, (double)sacc[2 * c1pad + c]);
                    atomicAdd(p.sdzdp + 3 * c + 1, (double)sacc[3 * c1pad + c]);
                    atomicAdd(p.sdzdp + 3 * c + 2, (double)sacc[4 * c1pad + c]);
                }
                if (p.sh1 != nullptr) atomicAdd(p.sh1 + c, (double)sacc[5 * c1pad + c]);
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
// Both operands are MN-major with K = rows.  A chunk (pipeline stage) is 32 rows: boxes of [32 rows][32 channels] (4 KB,
// SWIZZLE_128B_BASE32B: the 32-byte unit index of a 128-byte row is XORed with row & 3) -- 4 per 128-channel slab of the
// A operand, c1pad / 32 for the B operand (h1).  Every pipeline stage has its own fixed set of producer warps (8 / stages
// of them, splitting the boxes; parity waits on an mbarrier are only safe for a producer that is never more than one
// phase ahead), 8 lanes per row, coalesced U rows, 16-byte shared stores.  grid = (row splits, slab groups); a CTA keeps
// the accumulators of its `nsl` slabs (128 lanes x c1pad columns each) in TMEM for all its tiles and adds them to D with
// 128-bit REDs at the end.
constexpr int WG_THREADS = 288;   // warps 0-7 producers (0-3 also the epilogue), warp 8 MMA issuer / TMEM allocator
constexpr int WG_MAX_STAGES = 8;
constexpr int WG_BOX = 4096;

struct W2Params {
    long long rows;
    int C1, C2, c1pad, N, M, K, relu1, stages, tmem_cols, Mtot;   // Mtot = C2 + C1 (eval mode: C2)
    int nsl;          // 128-channel slabs of the A operand per CTA
    const float *U, *xyz, *new_xyz;
    const int32_t *idx;
    const float *Wp, *s1, *t1;
    const float *gs_pm;
    const uint8_t *ksel_pm;
    float *D;   // (Mtot, c1pad) pre-zeroed
};

// MN-major 32-bit operand, SWIZZLE_128B_BASE32B (same descriptor as tc_cm_gemm.cu's activation operand)
__device__ __forceinline__ uint64_t w2_desc_mn(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((WG_BOX >> 4) & 0x3FFF) << 16;      // leading byte offset: next 32-channel box along M / N
    d |= (uint64_t)(512 >> 4) << 32;                    // stride byte offset: next group of 4 k-rows
    d |= (uint64_t)1 << 46;                             // descriptor version (sm_100)
    d |= (uint64_t)1 << 61;                             // layout type 1 = SWIZZLE_128B_BASE32B
    return d;
}

__global__ void __launch_bounds__(WG_THREADS, 1) sa2_wgrad_kernel(const W2Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = align1024(smem_raw);
    const int c1pad = p.c1pad;
    const int nboxA = 4 * p.nsl, nboxB = c1pad / 32;
    const int stage_bytes = (nboxA + nboxB) * WG_BOX;
    uint8_t *tail = smem + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);   // [WG_MAX_STAGES]
    uint64_t *empty_bar = full_bar + WG_MAX_STAGES;            // [WG_MAX_STAGES]
    uint64_t *done_bar = empty_bar + WG_MAX_STAGES;            // [1]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(done_bar + 1);
    float *prm = reinterpret_cast<float *>(tail + 256);        // [5][c1pad]: s1, t1, wx, wy, wz

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int slab0 = blockIdx.y * p.nsl;
    const long long ntiles = (p.rows + BM - 1) / BM;
    const long long my_tiles = (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x;   // >= 1 (grid.x <= ntiles)
    const uint32_t nchunks = (uint32_t)my_tiles * 4;
    const int wps = p.stages >= 8 ? 1 : 8 / p.stages;   // producer warps per stage

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            tc::mbar_init(&full_bar[s], 32 * wps);
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
            const uint32_t idesc = tc::idesc_tf32_f32(BM, c1pad) | (1u << 15) | (1u << 16);   // A and B MN-major
            for (uint32_t it = 0; it < nchunks; ++it) {
                const int s = it % p.stages;
                const uint32_t ph = (it / p.stages) & 1;
                tc::mbar_wait(&full_bar[s], ph);
                tc::tc_fence_after();
                const uint32_t a_addr = tc::smem_u32(smem + (size_t)s * stage_bytes);
                const uint32_t b_addr = a_addr + nboxA * WG_BOX;
#pragma unroll
                for (int k = 0; k < 4; ++k) {   // UMMA_K = 8 rows = 1024 bytes of every box
                    const uint64_t bd = w2_desc_mn(b_addr + k * 1024);
                    for (int sl = 0; sl < p.nsl; ++sl)
                        tc::umma_tf32(tmem_base + sl * c1pad, w2_desc_mn(a_addr + sl * 4 * WG_BOX + k * 1024), bd, idesc,
                                      (it | (uint32_t)k) != 0);
                }
                tc::umma_commit(&empty_bar[s]);
            }
            tc::umma_commit(done_bar);
        }
    } else {
        // ---- producers: chunk `it` (32 rows) goes to stage it % stages and is built by that stage's `wps` warps
        const int l8 = lane & 7, lr = lane >> 3;
        const int my_stage = warp / wps, part = warp % wps;
        const uint32_t first = my_stage < p.stages ? (uint32_t)my_stage : nchunks;   // spare warps build nothing
        auto chunk_row0 = [&](uint32_t it) { return ((long long)blockIdx.x + (long long)(it >> 2) * gridDim.x) * BM + (it & 3) * 32; };
        int nidx[8];   // idx of this warp's next chunk, fetched one chunk ahead
        auto fetch_idx = [&](uint32_t it) {
#pragma unroll
            for (int step = 0; step < 8; ++step) {
                nidx[step] = -1;
                if (it < nchunks) {
                    const long long grow = chunk_row0(it) + step * 4 + lr;
                    if (grow < p.rows) nidx[step] = __ldg(p.idx + grow);
                }
            }
        };
        fetch_idx(first);
        for (uint32_t it = first; it < nchunks; it += p.stages) {
            const int s = my_stage;
            const uint32_t ph = (it / p.stages) & 1;
            const long long row0 = chunk_row0(it);
            // K == 32: the 32 rows of a chunk are the neighbours k = 0..31 of ONE query
            const long long q = row0 >> 5;
            const long long bq = (int)q / p.M;
            const float *qq = p.new_xyz + (row0 < p.rows ? q : 0) * 3;   // (chunks past the last row of the last tile)
            const float qx = __ldg(qq + 0), qy = __ldg(qq + 1), qz = __ldg(qq + 2);
            int urow[8];
            float dx[8], dy[8], dz[8];
#pragma unroll
            for (int step = 0; step < 8; ++step) {
                urow[step] = -1;
                dx[step] = dy[step] = dz[step] = 0.f;
                if (nidx[step] >= 0) {
                    urow[step] = (int)(bq * p.N + nidx[step]);
                    const float *pp = p.xyz + (size_t)urow[step] * 3;
                    dx[step] = __ldg(pp + 0) - qx;
                    dy[step] = __ldg(pp + 1) - qy;
                    dz[step] = __ldg(pp + 2) - qz;
                }
            }
            fetch_idx(it + p.stages);
            const uint32_t a_base = tc::smem_u32(smem + (size_t)s * stage_bytes);
            bool waited = false;
            for (int bx = part; bx < nboxA + nboxB; bx += wps) {
                // channels of this lane's 16 bytes: A boxes walk the concatenated [one-hot gs (C2) | h1 (C1)] channel list
                // of the CTA's slabs, B boxes the h1 channels
                const int ch = bx < nboxA ? (slab0 * 4 + bx) * 32 + l8 * 4 : p.C2 + (bx - nboxA) * 32 + l8 * 4;
                float4 v[8];
                if (ch < p.C2) {
                    // the chunk's 32 rows share one query: one load of its gs / ksel quad, 8 one-hot variants
                    const bool ok = row0 < p.rows;
                    const float4 g = ok ? __ldg(reinterpret_cast<const float4 *>(p.gs_pm + q * p.C2 + ch)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    const uint32_t kq = ok ? __ldg(reinterpret_cast<const uint32_t *>(p.ksel_pm + q * p.C2 + ch)) : 0xffffffffu;
#pragma unroll
                    for (int step = 0; step < 8; ++step) {
                        const int k = step * 4 + lr;
                        v[step].x = (int)(kq & 0xff) == k ? g.x : 0.f;
                        v[step].y = (int)((kq >> 8) & 0xff) == k ? g.y : 0.f;
                        v[step].z = (int)((kq >> 16) & 0xff) == k ? g.z : 0.f;
                        v[step].w = (int)(kq >> 24) == k ? g.w : 0.f;
                    }
                } else {
                    const int c = ch - p.C2;
                    const bool cvalid = c < p.C1 && (bx >= nboxA || ch < p.Mtot);
#pragma unroll
                    for (int step = 0; step < 8; ++step)
                        v[step] = (urow[step] >= 0 && cvalid)
                                      ? __ldg(reinterpret_cast<const float4 *>(p.U + (size_t)urow[step] * p.C1 + c))
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int step = 0; step < 8; ++step)
                        if (urow[step] >= 0 && cvalid)
                            v[step] = h1_quad(prm, c1pad, c, v[step], dx[step], dy[step], dz[step], p.relu1);
                }
                if (!waited) {
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    waited = true;
                }
                const uint32_t box = a_base + bx * WG_BOX;
#pragma unroll
                for (int step = 0; step < 8; ++step) {
                    const int r = step * 4 + lr;
                    tc::st_shared_v4(box + r * 128 + (((l8 >> 1) ^ (r & 3)) << 5) + ((l8 & 1) << 4), tc::to_tf32(v[step].x),
                                     tc::to_tf32(v[step].y), tc::to_tf32(v[step].z), tc::to_tf32(v[step].w));
                }
            }
            if (!waited) tc::mbar_wait(&empty_bar[s], ph ^ 1);   // (a warp with no box of this stage still keeps the phase)
            tc::fence_proxy_async_smem();
            tc::mbar_arrive(&full_bar[s]);
        }
        if (warp < 4) {
            // ---- epilogue: TMEM lane = channel of the slab, columns = C1 channels
            tc::mbar_wait(done_bar, 0);
            tc::tc_fence_after();
            for (int sl = 0; sl < p.nsl; ++sl) {
                const int row = (slab0 + sl) * BM + warp * 32 + lane;
                const uint32_t tacc = tmem_base + sl * c1pad + ((uint32_t)(warp * 32) << 16);
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

int launch_fwd(S2Params p, const float *W, int Kp, cudaStream_t st) {
    if (p.rows <= 0) return APB_OK;
    if ((Kp & 31) != 0 || Kp <= 0 || (reinterpret_cast<uintptr_t>(W) & 15) != 0) return APB_E_INVALID;
    p.kchunks = Kp / 32;
    p.nmb = (p.Cn + BM - 1) / BM;
    if (p.nmb > 4) return APB_E_UNSUPPORTED;
    p.Cq = p.nmb > 1 ? BM : (p.Cn <= 32 ? 32 : (p.Cn <= 64 ? 64 : BM));
    const int stage_bytes = A_BYTES + p.nmb * BM * 128;
    const size_t extra = 1024 /*align slack*/ + 256 /*barriers*/ + (size_t)5 * p.c1pad * 4 + (size_t)NGROUPS * 8 * BM * 4;
    int stages = (int)((227 * 1024 - extra) / stage_bytes);
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    stages = stages / NGROUPS * NGROUPS;  // each producer group owns its own in-order ring
    if (stages < NGROUPS) return APB_E_UNSUPPORTED;
    p.stages = stages;
    size_t smem = (size_t)stages * stage_bytes + extra;
    if (smem < 120 * 1024) smem = 120 * 1024;   // one CTA per SM: every CTA allocates all 512 TMEM columns
    CUtensorMap map;
    const int rc = s2_weight_map(&map, W, p.Cn, Kp, p.Cq);
    if (rc != APB_OK) return rc;
    APB_CUDA(cudaFuncSetAttribute((const void *)sa2_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long ntiles = (p.rows + BM - 1) / BM;
    sa2_fwd_kernel<<<(unsigned)(ntiles < APB_NUM_SMS ? ntiles : APB_NUM_SMS), NTHREADS, smem, st>>>(map, p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

template <bool COORDS>
int launch_bwd(S2Params p, const float *W, int Kp, cudaStream_t st) {
    if (p.rows <= 0) return APB_OK;
    if ((Kp & 31) != 0 || Kp <= 0 || (reinterpret_cast<uintptr_t>(W) & 15) != 0) return APB_E_INVALID;
    p.kchunks = Kp / 32;
    p.NCP = (p.C1 + 15) / 16 * 16;
    if (p.NCP > 256) return APB_E_UNSUPPORTED;   // two accumulator stages of <= 256 columns
    const int stage_bytes = A_BYTES + p.NCP * 128;
    const size_t extra = 1024 /*align slack*/ + 256 /*barriers*/ + (size_t)NPRM * p.c1pad * 4 +
                         (size_t)NGROUPS * 8 * BM * 4 /*metadata*/ + (size_t)NSUM * p.c1pad * 4 /*channel sums*/;
    int stages = (int)((227 * 1024 - extra) / stage_bytes);
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    stages = stages / NGROUPS * NGROUPS;   // each producer group owns its own in-order ring
    if (stages < NGROUPS) return APB_E_UNSUPPORTED;
    p.stages = stages;
    size_t smem = (size_t)stages * stage_bytes + extra;
    if (smem < 120 * 1024) smem = 120 * 1024;   // one CTA per SM: every CTA allocates all 512 TMEM columns
    CUtensorMap map;
    const int rc = s2_weight_map(&map, W, p.C1, Kp, p.NCP);
    if (rc != APB_OK) return rc;
    auto kern = sa2_bwd_kernel<COORDS>;
    APB_CUDA(cudaFuncSetAttribute((const void *)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long ntiles = (p.rows + BM - 1) / BM;
    kern<<<(unsigned)(ntiles < APB_NUM_SMS ? ntiles : APB_NUM_SMS), BW_THREADS, smem, st>>>(map, p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

int g_sa2_dbg = 0;

bool s2_shapes_ok(int B, int N, int M, int K, int C1, int C2) {
    return B > 0 && N > 0 && M > 0 && K == 32 && C1 > 0 && C2 > 0 && (C1 & 3) == 0 && (C2 & 3) == 0 && C1 <= 512 &&
           C2 <= 512 && (long long)B * N < (1ll << 31) && (long long)B * M < (1ll << 31);
}

}  // namespace

// Benchmarks only: switch parts of the backward rows kernel off to time the rest (results are then wrong).
extern "C" int apb_tune_sa2(int flags) {
    g_sa2_dbg = flags;
    return APB_OK;
}

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
    return launch_fwd(p, w2, K2p, (cudaStream_t)stream);
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
    if (!s2_shapes_ok(B, N, M, K, C1, C2) || C1 > 256 || (reinterpret_cast<uintptr_t>(u_pm) & 15) != 0 ||
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
    p.dbg = g_sa2_dbg;
    if (dxyz != nullptr || dq != nullptr) return launch_bwd<true>(p, wb, Kp, (cudaStream_t)stream);
    return launch_bwd<false>(p, wb, Kp, (cudaStream_t)stream);
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
    const int nslab = (p.Mtot + BM - 1) / BM;
    int nsl = 512 / p.c1pad;                  // accumulators of one CTA: nsl x c1pad TMEM columns
    if (nsl > 4) nsl = 4;
    if (nsl > nslab) nsl = nslab;
    p.nsl = nsl;
    int cols = 32;
    while (cols < nsl * p.c1pad) cols *= 2;
    p.tmem_cols = cols;
    const int stage_bytes = (4 * nsl + p.c1pad / 32) * WG_BOX;
    const size_t extra = 1024 + 256 + (size_t)5 * p.c1pad * 4;
    int stages = (int)((227 * 1024 - extra) / stage_bytes);
    if (stages > WG_MAX_STAGES) stages = WG_MAX_STAGES;
    if (stages < 2) return APB_E_UNSUPPORTED;
    p.stages = stages;
    size_t smem = (size_t)stages * stage_bytes + extra;
    if (cols > 256 && smem < 120 * 1024) smem = 120 * 1024;   // 512 TMEM columns: one CTA per SM
    APB_CUDA(cudaFuncSetAttribute((const void *)sa2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long ntiles = (p.rows + BM - 1) / BM;
    const int ngrp = (nslab + nsl - 1) / nsl;
    long long gx = APB_NUM_SMS / ngrp;
    if (gx < 1) gx = 1;
    if (gx > ntiles) gx = ntiles;
    sa2_wgrad_kernel<<<dim3((unsigned)gx, (unsigned)ngrp), WG_THREADS, smem, (cudaStream_t)stream>>>(p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
