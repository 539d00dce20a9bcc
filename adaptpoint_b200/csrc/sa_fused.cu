// The shared MLP of the set-abstraction path on the sm_100a tensor cores (tcgen05 + TMEM + TMA).
//
// Replaces, for BatchNorm-folded (eval-mode) blocks, the reference's op chain
//   QueryAndGroup -> cat([dp, fj]) -> Conv2d 1x1 (+BN+ReLU) [-> Conv2d 1x1 (+BN)(+ReLU)] -> max over K
// (openpoints/models/backbone/pointnext.py:164-166 SetAbstraction, :66-77 LocalAggregation;
//  layers/group.py:235-255, 323-325; layers/conv.py:24-60), which materialises (B, C, M, K) tensors five times,
// and the plain 1x1 Conv1d/Conv2d (+BN)(+ReLU) layers (conv.py:8-21).
//
// Design (DESIGN.md "Kernel 4"):
//  * The first conv of a block is linear in its grouped input, so it is evaluated ONCE PER SUPPORT POINT instead
//    of once per (query, neighbour):   W1 . [dp/r ; f[idx]]  =  (W1f . f)[idx]  +  (W1p / r) . (p[idx] - q).
//    U = W1f . f is a dense GEMM over the B*N points (tc_rows_gemm_kernel, channel-major loader, point-major
//    store); the xyz part is 3 FMAs per output in fp32, with no cancellation.
//  * One-layer blocks (LocalAggregation, sa_layers=1) then need no per-neighbour GEMM at all: group_max_kernel
//    gathers rows of U, adds the xyz term, applies the folded BN (+ReLU) and max-pools over K.
//  * Two-layer blocks (PointNeXt-S SetAbstraction) run the second conv on all B*M*K rows on the tensor cores:
//    tc_rows_gemm_kernel with the GATHER loader builds the bf16 A tile  relu(s1*(U[idx] + Wp.dp) + t1)  directly
//    in the 128B-swizzled K-major layout tcgen05.mma reads, W2 tiles arrive by TMA, the fp32 accumulator lives in
//    TMEM, and the epilogue applies the folded BN2 and max-pools the 32 neighbour rows of a query with one
//    CREDUX per channel -- the (B,C,M,K) activations never exist.
//
// tc_rows_gemm_kernel: persistent, one CTA per SM, 448 threads, warp-specialised:
//    warp 0    TMA producer for the weight chunk of every pipeline stage
//    warp 1    TMEM allocator + single-thread tcgen05.mma issuer (UMMA 128 x N x 16, bf16 -> fp32)
//    warps 2-9 A-tile producers (2 groups of 128 threads, alternating tiles): global -> registers (fp32 math) -> bf16 -> swizzled smem
//    warps 10-13 epilogue: tcgen05.ld accumulator -> affine/ReLU -> store or K-max -> global
//  smem ring of STAGES x (A chunk 128x64 bf16 = 16 KB | W chunk NCPx64 bf16), full/empty mbarriers;
//  TMEM: 512 columns, double-buffered accumulator when Cout <= 256, so the epilogue of tile t overlaps the
//  loads + MMAs of tile t+1.
#include <cuda.h>
#include <cstdio>
#include "apb_common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int BM = 128;       // rows per tile (UMMA M)
constexpr int A_BYTES = BM * 128;  // A chunk: 128 rows x one 128-byte swizzle span (64 bf16 or 32 tf32)
constexpr int NGROUPS = 2;    // A-producer groups of 128 threads (alternate tiles)
constexpr int PROD0 = 64;     // first producer thread
constexpr int EPI0 = PROD0 + 128 * NGROUPS;  // first epilogue thread (warp index % 4 == 2: TMEM lane quarters rotate)
constexpr int NTHREADS = EPI0 + 128;
constexpr int MAX_STAGES = 6;

enum Loader { LOAD_CHANNEL_MAJOR = 0, LOAD_GATHER = 1 };
enum Epilogue { EPI_STORE_CM = 0, EPI_STORE_PM = 1, EPI_KMAX = 2 };

struct TcParams {
    long long rows;   // valid rows: B*P (channel-major loader) or B*M*K (gather loader)
    int kchunks;      // padded reduction length / 64
    int Cout;         // valid output channels (all slices together)
    int nslice;       // output channels per CTA column (blockIdx.y); == NCP.  One slice unless the row tiles alone
                      // cannot fill the SMs (launch_tc), then Cout is cut into slices of <= 256 channels
    int NCP;          // Cout rounded up to 16 (UMMA N granularity), <= 512
    int wrows;        // weight rows resident per stage: NCP if <= 256 else 512
    int stages;       // smem ring depth
    // channel-major loader: X (B, Cin, P) f32
    const float *X;
    int Cin, P;
    // gather loader: rows (b,m,k) -> support point idx[b,m,k]
    const float *U;        // (B, N, C1) f32 point-major: feature part of the first conv, per support point
    const float *xyz;      // (B, N, 3)
    const float *new_xyz;  // (B, M, 3)
    const int32_t *idx;    // (B, M, K)
    const float *Wp;       // (3, C1): xyz part of the first conv, already divided by the radius; row j = coord j
    const float *s1, *t1;  // (C1) folded BatchNorm of the first conv
    int C1, N, M, K, relu1;
    // epilogue
    const float *scale, *shift;  // (Cout) folded BatchNorm / bias; nullptr = identity
    int relu;
    float *Y;  // EPI_STORE_CM: (B, Cout, P);  EPI_STORE_PM: (rows, Cout);  EPI_KMAX: (B, Cout, M)
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// TF32 = false: bf16 operands (64 per 128-byte chunk), kind::f16.  TF32 = true: fp32 operands rounded to tf32 (32 per
// chunk), kind::tf32 -- the training path, where bf16 operand rounding measurably perturbs the gradients of this
// max-pool / ReLU-routed network (tests/test_tcconv_gpu.py).
template <int LOADER, int EPI, bool TF32>
__global__ void __launch_bounds__(NTHREADS, 1)
tc_rows_gemm_kernel(const __grid_constant__ CUtensorMap wmap, const TcParams p) {
    constexpr int BKE = TF32 ? 32 : 64;  // K elements per 128-byte chunk
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // 1024-byte alignment is required by the 128B swizzle; dynamic smem base is only guaranteed 16 B
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int stage_bytes = A_BYTES + p.wrows * 128;
    uint8_t *tail = smem + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);  // [MAX_STAGES]
    uint64_t *empty_bar = full_bar + MAX_STAGES;              // [MAX_STAGES]
    uint64_t *tfull_bar = empty_bar + MAX_STAGES;             // [2]
    uint64_t *tempty_bar = tfull_bar + 2;                     // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);
    float *prm = reinterpret_cast<float *>(tail + 256);       // gather loader: s1,t1,wx,wy,wz [5][C1pad]
    const int c1pad = p.kchunks * BKE;
    int *meta_row = reinterpret_cast<int *>(prm + 5 * (LOADER == LOAD_GATHER ? c1pad : 0));  // [NGROUPS][2][BM]
    float *meta_dp = reinterpret_cast<float *>(meta_row + NGROUPS * 2 * BM);                 // [NGROUPS][2][3][BM]
    float *ep_ss = meta_dp + NGROUPS * 6 * BM;  // epilogue scale [512] | shift [512]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int acc_stages = p.NCP <= 256 ? 2 : 1;
    const uint32_t spg = p.stages / NGROUPS;  // smem stages per producer group
    const long long ntiles = (p.rows + BM - 1) / BM;
    const int n0 = blockIdx.y * p.nslice;            // first output channel of this CTA's slice
    const int Cout = min(p.nslice, p.Cout - n0);     // valid channels of the slice

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
    if (LOADER == LOAD_GATHER && tid >= PROD0 && tid < EPI0) {
        for (int c = tid - PROD0; c < c1pad; c += 128 * NGROUPS) {
            const bool ok = c < p.C1;
            prm[0 * c1pad + c] = ok ? p.s1[c] : 0.f;
            prm[1 * c1pad + c] = ok ? p.t1[c] : 0.f;
            prm[2 * c1pad + c] = ok ? p.Wp[0 * p.C1 + c] : 0.f;
            prm[3 * c1pad + c] = ok ? p.Wp[1 * p.C1 + c] : 0.f;
            prm[4 * c1pad + c] = ok ? p.Wp[2 * p.C1 + c] : 0.f;
        }
    }
    if (tid >= EPI0) {
        for (int c = tid - EPI0; c < 512; c += 128) {
            const bool ok = p.scale != nullptr && c < Cout;
            ep_ss[c] = ok ? p.scale[n0 + c] : 1.f;
            ep_ss[512 + c] = ok ? p.shift[n0 + c] : 0.f;
        }
    }
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
                    tc::tma_load_2d(wdst, &wmap, kc * BKE, n0, &full_bar[s]);
                    if (p.wrows > 256) tc::tma_load_2d(wdst + 256 * 128, &wmap, kc * BKE, n0 + 256, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const int n_lo = p.NCP <= 256 ? p.NCP : 256, n_hi = p.NCP > 256 ? p.NCP - 256 : 16;
            const uint32_t idesc_lo = TF32 ? tc::idesc_tf32_f32(BM, n_lo) : tc::idesc_bf16_f32(BM, n_lo);
            const uint32_t idesc_hi = TF32 ? tc::idesc_tf32_f32(BM, n_hi) : tc::idesc_bf16_f32(BM, n_hi);
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
                    for (int k = 0; k < 4; ++k) {  // 4 x 32 bytes per chunk: UMMA_K = 16 bf16 or 8 tf32
                        const uint64_t ad = tc::smem_desc_sw128(a_addr + k * 32);
                        const uint64_t bd = tc::smem_desc_sw128(w_addr + k * 32);
                        if (TF32) tc::umma_tf32(tacc, ad, bd, idesc_lo, (kc | k) != 0);
                        else tc::umma_bf16(tacc, ad, bd, idesc_lo, (kc | k) != 0);
                        if (p.NCP > 256) {
                            const uint64_t bd2 = tc::smem_desc_sw128(w_addr + 256 * 128 + k * 32);
                            if (TF32) tc::umma_tf32(tacc + 256, ad, bd2, idesc_hi, (kc | k) != 0);
                            else tc::umma_bf16(tacc + 256, ad, bd2, idesc_hi, (kc | k) != 0);
                        }
                    }
                    tc::umma_commit(&empty_bar[s]);  // smem stage reusable once these MMAs have read it
                }
                tc::umma_commit(&tfull_bar[a]);      // accumulator complete -> epilogue
            }
        }
    } else if (warp < 2 + 4 * NGROUPS) {
        // ===================== A-tile producers: NGROUPS groups of 128 threads, group g fills tiles lt % NGROUPS == g
        const int grp = (warp - 2) >> 2;
        const int pt = (tid - PROD0) & 127;  // row of the tile (channel-major loader) / thread of the group
        int *g_meta_row = meta_row + grp * 2 * BM;
        float *g_meta_dp = meta_dp + grp * 6 * BM;
        uint32_t lt = 0;

        // gather loader: metadata of a tile's row `pt` (support-point row of U, dp = p[idx] - q), prefetched into
        // registers one tile ahead so the idx -> xyz dependent loads overlap the previous tile's chunk work
        int nx_urow = -1;
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
                nx_urow = (int)(b * p.N + n);
            }
        };
        if (LOADER == LOAD_GATHER) fetch_meta((long long)blockIdx.x + (long long)grp * gridDim.x);

        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            if ((int)(lt % NGROUPS) != grp) continue;
            uint32_t it = (lt / NGROUPS) * (uint32_t)p.kchunks;  // this group's own stage ring (in-order => parity-safe)
            if (LOADER == LOAD_CHANNEL_MAJOR) {
                // thread <-> row: lanes read 32 consecutive points of one channel (coalesced 128 B)
                const long long grow = tile * BM + pt;
                const bool valid = grow < p.rows;
                const long long b = valid ? grow / p.P : 0;
                const long long pp = valid ? grow - b * p.P : 0;
                const float *src = p.X + (b * p.Cin) * p.P + pp;
                for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                    const int s = grp * spg + it % spg;
                    const uint32_t ph = (it / spg) & 1;
                    const uint32_t arow = tc::smem_u32(smem + (size_t)s * stage_bytes) + pt * 128;
                    // all loads of the chunk are issued before the first use (16-32 KB in flight per group)
                    float v[BKE];
                    const int cbase = kc * BKE;
                    if (valid && cbase + BKE <= p.Cin) {
#pragma unroll
                        for (int e = 0; e < BKE; ++e) v[e] = __ldg(src + (size_t)(cbase + e) * p.P);
                    } else {
#pragma unroll
                        for (int e = 0; e < BKE; ++e)
                            v[e] = (valid && cbase + e < p.Cin) ? __ldg(src + (size_t)(cbase + e) * p.P) : 0.f;
                    }
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    if (TF32) {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            tc::st_shared_v4(arow + ((j ^ (pt & 7)) << 4), tc::to_tf32(v[4 * j + 0]),
                                             tc::to_tf32(v[4 * j + 1]), tc::to_tf32(v[4 * j + 2]),
                                             tc::to_tf32(v[4 * j + 3]));
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            tc::st_shared_v4(arow + ((j ^ (pt & 7)) << 4),
                                             tc::pack_bf16x2(v[(8 * j + 0) % BKE], v[(8 * j + 1) % BKE]),
                                             tc::pack_bf16x2(v[(8 * j + 2) % BKE], v[(8 * j + 3) % BKE]),
                                             tc::pack_bf16x2(v[(8 * j + 4) % BKE], v[(8 * j + 5) % BKE]),
                                             tc::pack_bf16x2(v[(8 * j + 6) % BKE], v[(8 * j + 7) % BKE]));
                    }
                    tc::fence_proxy_async_smem();
                    tc::mbar_arrive(&full_bar[s]);
                }
            } else {
                const int mb = (lt / NGROUPS) & 1;
                tc::sts_i(g_meta_row + mb * BM + pt, nx_urow);
                tc::sts_f(g_meta_dp + (mb * 3 + 0) * BM + pt, nx_dx);
                tc::sts_f(g_meta_dp + (mb * 3 + 1) * BM + pt, nx_dy);
                tc::sts_f(g_meta_dp + (mb * 3 + 2) * BM + pt, nx_dz);
                named_bar_sync(1 + grp, 128);
                fetch_meta(tile + (long long)NGROUPS * gridDim.x);  // next tile of this group, consumed next round
                // 8 lanes per row, 8 channels (32 B of U) per lane; a warp covers 4 rows per step
                const int pw = pt >> 5, l8 = lane & 7, lr = lane >> 3;
                for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                    const int s = grp * spg + it % spg;
                    const uint32_t ph = (it / spg) & 1;
                    const int c0 = kc * BKE + l8 * 8;
                    const bool cvalid = c0 < p.C1;  // C1 % 8 == 0: a lane's 8 channels are all valid or all padding
                    // rows of U for the 8 steps, issued up front
                    float4 u0[8], u1[8];
                    int ur[8];
#pragma unroll
                    for (int step = 0; step < 8; ++step) {
                        const int r = pw * 32 + step * 4 + lr;
                        ur[step] = tc::lds_i(g_meta_row + mb * BM + r);
                        const float4 *up = reinterpret_cast<const float4 *>(
                            p.U + (size_t)(ur[step] < 0 ? 0 : ur[step]) * p.C1 + (cvalid ? c0 : 0));
                        u0[step] = __ldg(up);
                        u1[step] = __ldg(up + 1);
                    }
                    float ps[8], pt1[8], wx[8], wy[8], wz[8];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float4 a = tc::lds_f4_ro(prm + 0 * c1pad + c0 + 4 * h);
                        const float4 bq = tc::lds_f4_ro(prm + 1 * c1pad + c0 + 4 * h);
                        const float4 cx = tc::lds_f4_ro(prm + 2 * c1pad + c0 + 4 * h);
                        const float4 cy = tc::lds_f4_ro(prm + 3 * c1pad + c0 + 4 * h);
                        const float4 cz = tc::lds_f4_ro(prm + 4 * c1pad + c0 + 4 * h);
                        ps[4 * h + 0] = a.x; ps[4 * h + 1] = a.y; ps[4 * h + 2] = a.z; ps[4 * h + 3] = a.w;
                        pt1[4 * h + 0] = bq.x; pt1[4 * h + 1] = bq.y; pt1[4 * h + 2] = bq.z; pt1[4 * h + 3] = bq.w;
                        wx[4 * h + 0] = cx.x; wx[4 * h + 1] = cx.y; wx[4 * h + 2] = cx.z; wx[4 * h + 3] = cx.w;
                        wy[4 * h + 0] = cy.x; wy[4 * h + 1] = cy.y; wy[4 * h + 2] = cy.z; wy[4 * h + 3] = cy.w;
                        wz[4 * h + 0] = cz.x; wz[4 * h + 1] = cz.y; wz[4 * h + 2] = cz.z; wz[4 * h + 3] = cz.w;
                    }
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    const uint32_t abase = tc::smem_u32(smem + (size_t)s * stage_bytes);
#pragma unroll
                    for (int step = 0; step < 8; ++step) {
                        const int r = pw * 32 + step * 4 + lr;
                        const float dx = tc::lds_f(g_meta_dp + (mb * 3 + 0) * BM + r);
                        const float dy = tc::lds_f(g_meta_dp + (mb * 3 + 1) * BM + r);
                        const float dz = tc::lds_f(g_meta_dp + (mb * 3 + 2) * BM + r);
                        float v[8] = {u0[step].x, u0[step].y, u0[step].z, u0[step].w,
                                      u1[step].x, u1[step].y, u1[step].z, u1[step].w};
                        const bool live = ur[step] >= 0 && cvalid;
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            float y = fmaf(wx[e], dx, v[e]);
                            y = fmaf(wy[e], dy, y);
                            y = fmaf(wz[e], dz, y);
                            y = fmaf(ps[e], y, pt1[e]);
                            y = p.relu1 ? fmaxf(y, 0.f) : y;
                            v[e] = live ? y : 0.f;
                        }
                        tc::st_shared_v4(abase + r * 128 + ((l8 ^ (r & 7)) << 4), tc::pack_bf16x2(v[0], v[1]),
                                         tc::pack_bf16x2(v[2], v[3]), tc::pack_bf16x2(v[4], v[5]),
                                         tc::pack_bf16x2(v[6], v[7]));
                    }
                    tc::fence_proxy_async_smem();
                    tc::mbar_arrive(&full_bar[s]);
                }
            }
        }
    } else {
        // ===================== epilogue (128 threads; warp w reads TMEM lanes 32*(w%4)..) =====================
        const int q4 = warp & 3;  // TMEM lane quarter of this warp == 32-row group of the tile
        uint32_t lt = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
            const int a = lt % acc_stages;
            const uint32_t aph = (lt / acc_stages) & 1;
            tc::mbar_wait(&tfull_bar[a], aph);
            tc::tc_fence_after();
            const uint32_t tacc = tmem_base + (uint32_t)a * 256 + ((uint32_t)(q4 * 32) << 16);
            const long long grow = tile * BM + q4 * 32 + lane;
            const bool valid = grow < p.rows;
            if (EPI == EPI_KMAX) {
                // the 32 rows of this warp are the K = 32 neighbours of ONE query
                const long long q = (tile * BM + q4 * 32) / p.K;  // b*M + m
                const bool qvalid = (tile * BM + q4 * 32) < p.rows;
                const long long b = q / p.M, m = q - b * p.M;
                float *out = p.Y + (b * p.Cout + n0) * p.M + m;
                for (int cb = 0; cb < p.NCP; cb += 32) {
                    uint32_t v[32];
                    tc::tmem_ld_32x32(tacc + cb, v);
                    tc::tmem_ld_wait();
                    int mine = INT_MIN;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float x = fmaf(tc::lds_f_ro(ep_ss + cb + j), __uint_as_float(v[j]), tc::lds_f_ro(ep_ss + 512 + cb + j));
                        const int r = __reduce_max_sync(0xffffffffu, tc::float_to_ordered(x));
                        if (lane == j) mine = r;
                    }
                    const int c = cb + lane;
                    if (qvalid && c < Cout) {
                        float x = tc::ordered_to_float(mine);
                        if (p.relu) x = fmaxf(x, 0.f);
                        out[(size_t)c * p.M] = x;
                    }
                }
            } else {
                const long long b = valid ? grow / p.P : 0;
                const long long pp = valid ? grow - b * p.P : 0;
                float *ycm = p.Y + (b * p.Cout + n0) * p.P + pp;  // EPI_STORE_CM: lanes = consecutive points (coalesced)
                for (int cb = 0; cb < p.NCP; cb += 32) {
                    uint32_t v[32];
                    tc::tmem_ld_32x32(tacc + cb, v);
                    tc::tmem_ld_wait();
                    float x[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float t = fmaf(tc::lds_f_ro(ep_ss + cb + j), __uint_as_float(v[j]), tc::lds_f_ro(ep_ss + 512 + cb + j));
                        x[j] = p.relu ? fmaxf(t, 0.f) : t;
                    }
                    if (EPI == EPI_STORE_CM) {
                        if (valid) {
                            if (cb + 32 <= Cout) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) {
                                    *ycm = x[j];
                                    ycm += p.P;
                                }
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j) {
                                    if (cb + j < Cout) *ycm = x[j];
                                    ycm += p.P;
                                }
                            }
                        }
                    } else if (valid) {
                        float *y = p.Y + grow * p.Cout + n0 + cb;  // one 128-byte line per thread
                        if (cb + 32 <= Cout && (p.Cout & 3) == 0) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4)
                                *reinterpret_cast<float4 *>(y + j) = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (cb + j < Cout) y[j] = x[j];
                        }
                    }
                }
            }
            tc::tc_fence_before();
            tc::mbar_arrive(&tempty_bar[a]);
        }
    }

    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, 512);
    }
}

// ---- one-layer blocks: gather rows of U, add the xyz term, folded BN (+ReLU), max over K -------------------
// out[b,c,m] = act( max_k ( s[c] * (U[b, idx[b,m,k], c] + Wp[:,c] . (p[idx] - q)) + t[c] ) )
// CTA = 8 warps, QB = 32 consecutive queries; a warp walks the K neighbours of a query with its lanes across
// channels (coalesced 16 B/lane reads of the U row), results are transposed through shared memory so that the
// channel-major output is written as 128-byte rows.
constexpr int GM_QB = 32;
__global__ void __launch_bounds__(256) group_max_kernel(const float *__restrict__ U, const float *__restrict__ xyz,
                                                        const float *__restrict__ new_xyz,
                                                        const int32_t *__restrict__ idx, const float *__restrict__ Wp,
                                                        const float *__restrict__ s, const float *__restrict__ t,
                                                        int relu, float *__restrict__ out, int N, int M, int K,
                                                        int C) {
    extern __shared__ float stage[];  // [C][GM_QB + 1]
    const int b = blockIdx.y;
    const int m0 = blockIdx.x * GM_QB;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float *Ub = U + (size_t)b * N * C;
    for (int qi = warp; qi < GM_QB; qi += 8) {
        const int m = m0 + qi;
        if (m >= M) break;
        const float *qq = new_xyz + ((size_t)b * M + m) * 3;
        const float qx = __ldg(qq), qy = __ldg(qq + 1), qz = __ldg(qq + 2);
        const int32_t *ix = idx + ((size_t)b * M + m) * K;
        for (int c0 = lane * 4; c0 < C; c0 += 128) {
            const float4 sc = *reinterpret_cast<const float4 *>(s + c0);
            const float4 sh = *reinterpret_cast<const float4 *>(t + c0);
            const float4 wx = *reinterpret_cast<const float4 *>(Wp + 0 * C + c0);
            const float4 wy = *reinterpret_cast<const float4 *>(Wp + 1 * C + c0);
            const float4 wz = *reinterpret_cast<const float4 *>(Wp + 2 * C + c0);
            float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
            for (int k = 0; k < K; ++k) {
                const int n = __ldg(ix + k);
                const float *pp = xyz + ((size_t)b * N + n) * 3;
                const float dx = __ldg(pp) - qx, dy = __ldg(pp + 1) - qy, dz = __ldg(pp + 2) - qz;
                const float4 u = __ldg(reinterpret_cast<const float4 *>(Ub + (size_t)n * C + c0));
                float y;
                y = fmaf(wz.x, dz, fmaf(wy.x, dy, fmaf(wx.x, dx, u.x))); best.x = fmaxf(best.x, fmaf(sc.x, y, sh.x));
                y = fmaf(wz.y, dz, fmaf(wy.y, dy, fmaf(wx.y, dx, u.y))); best.y = fmaxf(best.y, fmaf(sc.y, y, sh.y));
                y = fmaf(wz.z, dz, fmaf(wy.z, dy, fmaf(wx.z, dx, u.z))); best.z = fmaxf(best.z, fmaf(sc.z, y, sh.z));
                y = fmaf(wz.w, dz, fmaf(wy.w, dy, fmaf(wx.w, dx, u.w))); best.w = fmaxf(best.w, fmaf(sc.w, y, sh.w));
            }
            if (relu) {
                best.x = fmaxf(best.x, 0.f); best.y = fmaxf(best.y, 0.f);
                best.z = fmaxf(best.z, 0.f); best.w = fmaxf(best.w, 0.f);
            }
            stage[(c0 + 0) * (GM_QB + 1) + qi] = best.x;
            stage[(c0 + 1) * (GM_QB + 1) + qi] = best.y;
            stage[(c0 + 2) * (GM_QB + 1) + qi] = best.z;
            stage[(c0 + 3) * (GM_QB + 1) + qi] = best.w;
        }
    }
    __syncthreads();
    const int nq = min(GM_QB, M - m0);
    for (int i = threadIdx.x; i < C * GM_QB; i += 256) {
        const int c = i / GM_QB, qi = i - c * GM_QB;
        if (qi < nq) out[((size_t)b * C + c) * M + m0 + qi] = stage[c * (GM_QB + 1) + qi];
    }
}

// ---- host side -------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
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

// Tensor map over K-major weights W (Cout rows, Kp columns; bf16, or fp32 read as tf32), box = one 128-byte span of
// columns x box_rows, 128B swizzle.
int encode_weight_map(CUtensorMap *map, const void *W, int Cout, int Kp, int box_rows, bool tf32) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return APB_E_UNSUPPORTED;
    const cuuint64_t gdim[2] = {(cuuint64_t)Kp, (cuuint64_t)Cout};
    const cuuint64_t gstride[1] = {(cuuint64_t)Kp * (tf32 ? 4 : 2)};
    const cuuint32_t box[2] = {(cuuint32_t)(tf32 ? 32 : 64), (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, tf32 ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                          const_cast<void *>(W), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? APB_OK : APB_E_INVALID;
}

template <int LOADER, int EPI, bool TF32 = false>
int launch_tc(TcParams p, const void *Wb, int Kp, cudaStream_t st) {
    constexpr int BKE = TF32 ? 32 : 64;
    if (p.rows <= 0) return APB_OK;
    if (p.Cout <= 0 || p.Cout > 512 || Kp <= 0 || (Kp % BKE) != 0) return APB_E_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(Wb) & 15) != 0) return APB_E_INVALID;
    p.kchunks = Kp / BKE;
    p.NCP = (p.Cout + 15) / 16 * 16;
    // Few row tiles (deep stages: B*P <= a few thousand rows) leave most SMs idle and make one CTA stream the whole
    // weight matrix: cut Cout into slices of >= 64 channels until the grid covers the SMs.
    const long long ntiles = (p.rows + BM - 1) / BM;
    int nsplit = 1;
    if (ntiles * 2 <= APB_NUM_SMS && p.Cout >= 128) {
        nsplit = (int)(APB_NUM_SMS / ntiles);
        if (nsplit > p.Cout / 64) nsplit = p.Cout / 64;
        if (nsplit < 1) nsplit = 1;
    }
    if (nsplit > 1) {
        p.NCP = ((p.Cout + nsplit - 1) / nsplit + 15) / 16 * 16;
        nsplit = (p.Cout + p.NCP - 1) / p.NCP;
    }
    p.nslice = p.NCP;
    p.wrows = p.NCP <= 256 ? p.NCP : 512;
    const int stage_bytes = A_BYTES + p.wrows * 128;
    const size_t extra = 1024 /*align slack*/ + 256 /*barriers*/ +
                         (LOADER == LOAD_GATHER ? (size_t)5 * Kp * 4 : 0) + (size_t)NGROUPS * 8 * BM * 4 /*meta*/ + 4096 /*epilogue scale|shift*/;
    int stages = (int)((227 * 1024 - extra) / stage_bytes);
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    stages = stages / NGROUPS * NGROUPS;  // each producer group owns its own in-order ring
    if (stages < NGROUPS) return APB_E_UNSUPPORTED;
    p.stages = stages;
    const size_t smem = (size_t)stages * stage_bytes + extra;
    CUtensorMap map;
    const int rc = encode_weight_map(&map, Wb, p.Cout, Kp, p.wrows > 256 ? 256 : p.wrows, TF32);
    if (rc != APB_OK) return rc;
    auto kern = tc_rows_gemm_kernel<LOADER, EPI, TF32>;
    APB_CUDA(cudaFuncSetAttribute((const void *)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const dim3 grid((unsigned)(ntiles < APB_NUM_SMS ? ntiles : APB_NUM_SMS), (unsigned)nsplit);
    kern<<<grid, NTHREADS, smem, st>>>(map, p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

}  // namespace

extern "C" int apb_pwconv_fwd_bf16(const float *x, const uint16_t *w_bf16, const float *scale, const float *shift,
                                   int relu, float *y, int B, int Cin, int Cout, int P, int Kp,
                                   apb_stream_t stream) {
    if (B < 0 || Cin <= 0 || Cout <= 0 || P < 0 || Kp < Cin) return APB_E_INVALID;
    if (B == 0 || P == 0) return APB_OK;
    if (!x || !w_bf16 || !y || ((scale == nullptr) != (shift == nullptr))) return APB_E_INVALID;
    TcParams p{};
    p.rows = (long long)B * P;
    p.Cout = Cout;
    p.X = x; p.Cin = Cin; p.P = P;
    p.scale = scale; p.shift = shift; p.relu = relu;
    p.Y = y;
    return launch_tc<LOAD_CHANNEL_MAJOR, EPI_STORE_CM>(p, w_bf16, Kp, (cudaStream_t)stream);
}

// Register-staged TF32 row GEMM (round 1).  apb_pwconv_fwd_tf32 (tc_cm_gemm.cu, both operands by TMA) falls back to it
// when the activation rows are not 16-byte aligned (P % 4 != 0), which TMA cannot address.
extern "C" int apb_pwconv_fwd_tf32_rows(const float *x, const float *w, const float *scale, const float *shift,
                                        int relu, float *y, int B, int Cin, int Cout, int P, int Kp,
                                        apb_stream_t stream) {
    if (B < 0 || Cin <= 0 || Cout <= 0 || P < 0 || Kp < Cin) return APB_E_INVALID;
    if (B == 0 || P == 0) return APB_OK;
    if (!x || !w || !y || ((scale == nullptr) != (shift == nullptr))) return APB_E_INVALID;
    TcParams p{};
    p.rows = (long long)B * P;
    p.Cout = Cout;
    p.X = x; p.Cin = Cin; p.P = P;
    p.scale = scale; p.shift = shift; p.relu = relu;
    p.Y = y;
    return launch_tc<LOAD_CHANNEL_MAJOR, EPI_STORE_CM, true>(p, w, Kp, (cudaStream_t)stream);
}

extern "C" int apb_sa_fused_fwd_bf16(const float *xyz, const float *new_xyz, const float *feat, const int32_t *idx,
                                     const uint16_t *w1f_bf16, const float *w1p, const float *s1, const float *t1,
                                     int relu1, const uint16_t *w2_bf16, const float *s2, const float *t2, int relu2,
                                     float *u_ws, float *out, int B, int N, int M, int K, int Cin, int C1, int K1p,
                                     int C2, int K2p, apb_stream_t stream) {
    if (B < 0 || N <= 0 || M < 0 || K <= 0 || Cin <= 0 || C1 <= 0 || K1p < Cin) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!xyz || !new_xyz || !feat || !idx || !w1f_bf16 || !w1p || !s1 || !t1 || !u_ws || !out) return APB_E_INVALID;
    if ((C1 & 7) != 0 || C1 > 512) return APB_E_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    // (1) U = W1f . f for every support point: (B*N) x Cin -> C1, point-major fp32
    TcParams g{};
    g.rows = (long long)B * N;
    g.Cout = C1;
    g.X = feat; g.Cin = Cin; g.P = N;
    g.Y = u_ws;
    int rc = launch_tc<LOAD_CHANNEL_MAJOR, EPI_STORE_PM>(g, w1f_bf16, K1p, st);
    if (rc != APB_OK) return rc;
    if (!w2_bf16) {
        // (2a) one-layer block: gather + xyz term + folded BN (+ReLU) + max over K
        if (B > 65535) return APB_E_UNSUPPORTED;
        const size_t smem = (size_t)C1 * (GM_QB + 1) * sizeof(float);
        APB_CUDA(cudaFuncSetAttribute((const void *)group_max_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem));
        dim3 grid(apb_div_up(M, GM_QB), B);
        group_max_kernel<<<grid, 256, smem, st>>>(u_ws, xyz, new_xyz, idx, w1p, s1, t1, relu1, out, N, M, K, C1);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    // (2b) two-layer block: second conv over all B*M*K rows on the tensor cores, max over the K rows of a query
    if (K != 32) return APB_E_UNSUPPORTED;
    if (C2 <= 0 || K2p < C1 || !s2 || !t2) return APB_E_INVALID;
    TcParams p{};
    p.rows = (long long)B * M * K;
    p.Cout = C2;
    p.U = u_ws; p.xyz = xyz; p.new_xyz = new_xyz; p.idx = idx;
    p.Wp = w1p; p.s1 = s1; p.t1 = t1; p.C1 = C1; p.N = N; p.M = M; p.K = K; p.relu1 = relu1;
    p.scale = s2; p.shift = t2; p.relu = relu2;
    p.Y = out;
    return launch_tc<LOAD_GATHER, EPI_KMAX>(p, w2_bf16, K2p, st);
}
