// Farthest-point sampling for sm_100a: register-resident points, one CTA or one thread-block
// cluster per cloud, one barrier per round.
//
// Replaces furthest_point_sampling_kernel (reference sampling_gpu.cu:100-216).  The reference keeps
// the running min-distance array and the coordinates in global memory (20 B/point/round of L2/HBM
// traffic), runs ONE CTA per cloud and does 10 __syncthreads per round.  Here every point
// (x,y,z,min-dist) lives in registers for the whole call, a cloud is spread over C CTAs of a
// cluster when it is too large for one SM (or when SMs would otherwise idle), and a round costs
//   per-thread strict-'>' scan -> 2 CREDUX (warp argmax) -> 5-word slot store (DSMEM for C>1)
//   -> ONE barrier (bar.sync, or barrier.cluster for C>1) -> every warp redundantly reduces the
//   <=128 slots (2 CREDUX) and shuffles the winner's coordinates.
// The compulsory DRAM traffic is 12N+4m bytes per cloud.
//
// Bit-exact tie rule.  In the reference, thread tid scans k = tid, tid+bs, ... with strict '>' and
// the tree reduction lets the LOWER slot win ties, so among equal maxima the winner has the smallest
// bit-reversed (k mod bs), then the smallest k.  We reduce the total-order key
//   (value as s32 bits  [max],  rank' = brev_L(k mod bs) << QB | (k div bs)  [min])
// and lay the points of a thread out in ascending rank' so that the in-thread scan is a plain
// strict '>' exactly like the reference's.  bs = apb_fps_block_size(N) = reference opt_n_threads.
#include <cooperative_groups.h>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include "apb_common.cuh"
#include "tc_ptx.cuh"

namespace cg = cooperative_groups;

namespace {

struct FpsParams {
    const float *xyz;
    float *temp;
    int32_t *idx;
    int N, m;
    int L;     // log2(bs)
    int logT;  // log2(C*NT)
    int QB;    // bits reserved for (k div bs) in rank'
    int C;     // CTAs per cloud (cluster size)
};

__host__ __device__ constexpr int brev_small(int v, int D) {
    // bit reversal of v over log2(D) bits, D in {1,2,4,8}
    return D == 1 ? 0 : D == 2 ? v : D == 4 ? (((v & 1) << 1) | (v >> 1)) : (((v & 1) << 2) | (v & 2) | (v >> 2));
}

// st.async: store to a (possibly remote) CTA's shared memory and count the bytes on that CTA's mbarrier.
__device__ __forceinline__ void st_async_v4(uint32_t raddr, int a, int b, int c, int d, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%2,%3,%4}, [%5];" ::"r"(
                     raddr),
                 "r"(a), "r"(b), "r"(c), "r"(d), "r"(rbar)
                 : "memory");
}
__device__ __forceinline__ void st_async_b32(uint32_t raddr, int a, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr),
                 "r"(a), "r"(rbar)
                 : "memory");
}
// Wait for a phase of a local mbarrier whose bytes are delivered by st.async from any CTA of the cluster
// (acquire at cluster scope).  A peer that died would otherwise hang the GPU: trap after ~2^26 polls.
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (spin > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t laddr, int rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(laddr), "r"(rank));
    return r;
}

// MULTI = false: one CTA per cloud, slots exchanged with STS + bar.sync.
// MULTI = true : p.C CTAs of one cluster per cloud; every warp's winner is pushed into every CTA's slot
//                table with st.async, which also counts bytes on that CTA's mbarrier, so a round needs no
//                cluster-wide barrier: each CTA just waits until its own table is complete.
template <int NT, int D, int Q, bool MULTI>
__global__ void __launch_bounds__(NT, 1) fps_kernel(const FpsParams p) {
    constexpr int PPT = D * Q;
    constexpr int NW = NT / 32;
    constexpr int MAXIT = MULTI ? (8 * NW + 31) / 32 : 1;  // slot-table passes of one warp (C <= 8)
    extern __shared__ __align__(16) float smem[];
    __shared__ uint64_t bars[2];
    float *sx = smem, *sy = sx + PPT * NT, *sz = sy + PPT * NT;

    const int C = MULTI ? p.C : 1;
    const int nslots = C * NW;
    int4 *slotA = reinterpret_cast<int4 *>(sz + PPT * NT);      // [2 parity][nslots] (value, rank', x, y)
    int *slotB = reinterpret_cast<int *>(slotA + 2 * nslots);   // [2 parity][nslots] z
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned crank = 0;
    if (MULTI) crank = cg::this_cluster().block_rank();
    const int cloud = blockIdx.y;
    const int N = p.N, L = p.L, bs = 1 << L;
    const float *xyz = p.xyz + (size_t)cloud * N * 3;
    float *temp = p.temp ? p.temp + (size_t)cloud * N : nullptr;
    int32_t *out = p.idx + (size_t)cloud * p.m;

    // ---- point -> (thread, slot) layout ------------------------------------------------------
    const int g = crank * NT + tid;  // thread id within the cloud's team of T = C*NT threads
    int t0, qbase;
    unsigned rb;
    if (p.logT >= L) {  // T >= bs : thread owns residue t0, a window of Q consecutive (k div bs)
        t0 = g & (bs - 1);
        qbase = (g >> L) * Q;
        rb = L ? (__brev((unsigned)t0) >> (32 - L)) : 0u;
    } else {  // T < bs : thread owns D residues g + lo*T, all (k div bs)
        t0 = g;
        qbase = 0;
        rb = (p.logT ? (__brev((unsigned)g) >> (32 - p.logT)) : 0u) * D;
    }
    const int Tthreads = 1 << p.logT;

    float x[PPT], y[PPT], z[PPT], md[PPT];
#pragma unroll
    for (int lp = 0; lp < D; ++lp) {
#pragma unroll
        for (int sq = 0; sq < Q; ++sq) {
            const int s = lp * Q + sq;
            const long long k = (long long)(qbase + sq) * bs + t0 + (long long)brev_small(lp, D) * Tthreads;
            if (k < N) {
                x[s] = xyz[k * 3 + 0];
                y[s] = xyz[k * 3 + 1];
                z[s] = xyz[k * 3 + 2];
                md[s] = temp ? temp[k] : 1e10f;
            } else {
                x[s] = y[s] = z[s] = 0.f;
                md[s] = -1.f;  // never beats the reference's initial best = -1 under strict '>'
            }
            sx[s * NT + tid] = x[s];
            sy[s * NT + tid] = y[s];
            sz[s * NT + tid] = z[s];
        }
    }
    float cx = xyz[0], cy = xyz[1], cz = xyz[2];
    if (g == 0) out[0] = 0;
    if (MULTI) {
        if (tid == 0) {
            tc::mbar_init(&bars[0], 1);
            tc::mbar_init(&bars[1], 1);
            tc::fence_barrier_init();
        }
        cg::this_cluster().sync();
    } else {
        __syncthreads();
    }
    const uint32_t slotA_s = tc::smem_u32(slotA), slotB_s = tc::smem_u32(slotB);
    const uint32_t bars_s = tc::smem_u32(bars);
    const int si = crank * NW + warp;  // this warp's slot

    const unsigned qmask = (1u << p.QB) - 1u;
    for (int j = 1; j < p.m; ++j) {
        const int par = j & 1;
        if (MULTI && tid == 0) tc::mbar_arrive_expect_tx(&bars[par], (uint32_t)nslots * 20u);
        // per-thread arg-max, first (lowest rank') point wins ties: pairwise tree with "left wins"
        float v[PPT];
        int id[PPT];
#pragma unroll
        for (int s = 0; s < PPT; ++s) {
            const float d = apb_sqdist(x[s] - cx, y[s] - cy, z[s] - cz);
            md[s] = fminf(d, md[s]);
            v[s] = md[s];
            id[s] = s;
        }
#pragma unroll
        for (int w = 1; w < PPT; w *= 2) {
#pragma unroll
            for (int s = 0; s + w < PPT; s += 2 * w) {
                const bool gt = v[s + w] > v[s];
                id[s] = gt ? id[s + w] : id[s];
                v[s] = gt ? v[s + w] : v[s];
            }
        }
        const bool any = v[0] > -1.f;  // the reference starts from best = -1, besti = 0
        const float best = any ? v[0] : -1.f;
        const int bests = any ? id[0] : 0;
        const unsigned rank = ((rb + (unsigned)(bests / Q)) << p.QB) | (unsigned)(qbase + bests % Q);
        const int vbits = __float_as_int(best);
        const int wv = __reduce_max_sync(0xffffffffu, vbits);
        const unsigned cand = vbits == wv ? rank : 0xffffffffu;
        const unsigned wr = __reduce_min_sync(0xffffffffu, cand);
        if (cand == wr) {  // exactly one lane: rank' is unique per (thread, slot)
            const int wx = __float_as_int(sx[bests * NT + tid]);
            const int wy = __float_as_int(sy[bests * NT + tid]);
            const int wz = __float_as_int(sz[bests * NT + tid]);
            if (MULTI) {
                const uint32_t la = slotA_s + (uint32_t)(par * nslots + si) * 16u;
                const uint32_t lb = slotB_s + (uint32_t)(par * nslots + si) * 4u;
                const uint32_t lbar = bars_s + (uint32_t)par * 8u;
                for (int c = 0; c < C; ++c) {
                    const uint32_t rbar = map_to_cta(lbar, c);
                    st_async_v4(map_to_cta(la, c), wv, (int)wr, wx, wy, rbar);
                    st_async_b32(map_to_cta(lb, c), wz, rbar);
                }
            } else {
                slotA[par * nslots + si] = make_int4(wv, (int)wr, wx, wy);
                slotB[par * nslots + si] = wz;
            }
        }
        // bars[par] is used in rounds par, par+2, ...: its ((j-1)/2)-th phase
        if (MULTI) mbar_wait_cluster(bars_s + (uint32_t)par * 8u, (uint32_t)((j - 1) >> 1) & 1u); else __syncthreads();

        // every warp reduces the slot table redundantly
        int bv = INT_MIN, bx = 0, by = 0, bz = 0;
        unsigned br = 0xffffffffu;
#pragma unroll
        for (int it = 0; it < MAXIT; ++it) {
            const int i = lane + 32 * it;
            if (i < nslots) {
                const int4 a = slotA[par * nslots + i];
                const int zz = slotB[par * nslots + i];
                const bool better = a.x > bv || (a.x == bv && (unsigned)a.y < br);
                bv = better ? a.x : bv;
                br = better ? (unsigned)a.y : br;
                bx = better ? a.z : bx;
                by = better ? a.w : by;
                bz = better ? zz : bz;
            }
        }
        const int gv = __reduce_max_sync(0xffffffffu, bv);
        const unsigned gcand = bv == gv ? br : 0xffffffffu;
        const unsigned gr = __reduce_min_sync(0xffffffffu, gcand);
        const bool win = gcand == gr && gr != 0xffffffffu;  // one lane (rank' unique over the cloud)
        cx = __int_as_float((int)__reduce_or_sync(0xffffffffu, win ? (unsigned)bx : 0u));
        cy = __int_as_float((int)__reduce_or_sync(0xffffffffu, win ? (unsigned)by : 0u));
        cz = __int_as_float((int)__reduce_or_sync(0xffffffffu, win ? (unsigned)bz : 0u));
        if (g == 0) {
            const unsigned hi = gr >> p.QB;
            const int t = L ? (int)(__brev(hi) >> (32 - L)) : 0;
            out[j] = (int)((gr & qmask) << L) + t;
        }
    }

    if (temp) {
#pragma unroll
        for (int lp = 0; lp < D; ++lp)
#pragma unroll
            for (int sq = 0; sq < Q; ++sq) {
                const long long k = (long long)(qbase + sq) * bs + t0 + (long long)brev_small(lp, D) * Tthreads;
                if (k < N) temp[k] = md[lp * Q + sq];
            }
    }
    if (MULTI) cg::this_cluster().sync();  // no CTA leaves while a peer may still be storing into it
}

typedef void (*fps_fn_t)(const FpsParams);
struct FpsVariant {
    int NT, D, Q;
    fps_fn_t fn;        // one CTA per cloud
    fps_fn_t fn_multi;  // one cluster per cloud (NT >= 256 only)
};
#define V(NT, D, Q) {NT, D, Q, fps_kernel<NT, D, Q, false>, (NT) >= 256 ? fps_kernel<(NT) >= 256 ? NT : 256, D, Q, true> : nullptr}
const FpsVariant kVariants[] = {
    V(32, 1, 2),   V(64, 1, 2),   V(128, 1, 2),
    V(256, 1, 2),  V(256, 2, 2),  V(256, 4, 1),  V(256, 4, 2),  V(256, 4, 3),  V(256, 4, 4),
    V(512, 1, 1),  V(512, 1, 2),  V(512, 1, 3),  V(512, 1, 4),  V(512, 1, 6),  V(512, 1, 8),
    V(512, 1, 12), V(512, 1, 16), V(512, 1, 20), V(512, 1, 24),
    V(512, 2, 1),  V(512, 2, 2),  V(512, 2, 4),
    V(1024, 1, 1), V(1024, 1, 2), V(1024, 1, 4),
    V(256, 1, 32), V(256, 1, 40), V(256, 1, 48), V(256, 1, 56),
};
#undef V

int ilog2(int v) {
    int l = 0;
    while ((1 << (l + 1)) <= v) ++l;
    return l;
}

const FpsVariant *find_variant(int NT, int D, int Qneed) {
    const FpsVariant *bestv = nullptr;
    for (const FpsVariant &v : kVariants)
        if (v.NT == NT && v.D == D && v.Q >= Qneed && (!bestv || v.Q < bestv->Q)) bestv = &v;
    return bestv;
}

// Work split for one cloud: threads per CTA, CTAs per cloud.  Candidates are scored with a
// simple per-round latency model (cycles): point work + barrier + slot reduce; the model constants
// were fitted on B200 (profiles/fps_config_sweep_r01.md).
struct FpsPlan {
    const FpsVariant *v;
    int C, logT, QB;
};

bool plan_for(int N, int bs, int NT, int C, FpsPlan *plan) {
    const int L = ilog2(bs);
    const int T = NT * C, logT = ilog2(T);
    const int cnt = (N + bs - 1) / bs;
    int D, Qneed, qspan;
    if (T >= bs) {
        D = 1;
        const int e = T / bs;
        Qneed = (cnt + e - 1) / e;
        qspan = e;
    } else {
        D = bs / T;
        Qneed = cnt;
        qspan = 1;
    }
    if (D > 4) return false;
    const FpsVariant *v = find_variant(NT, D, Qneed);
    if (!v) return false;
    int QB = 0;
    while ((1 << QB) < qspan * v->Q) ++QB;
    if (L + QB > 30) return false;
    plan->v = v;
    plan->C = C;
    plan->logT = logT;
    plan->QB = QB;
    return true;
}

double plan_cost(const FpsPlan &pl, int B) {
    const int NT = pl.v->NT, ppt = pl.v->D * pl.v->Q, NW = NT / 32;
    // issue-bound point loop: ~9 instr/point, NW warps over 4 schedulers
    double work = 9.0 * ppt * (NW > 4 ? NW / 4.0 : 1.0) + 40.0;
    double sync = pl.C > 1 ? 520.0 + 12.0 * pl.C : 60.0 + 3.0 * NW;
    double red = 110.0 + 20.0 * ((pl.C * NW + 31) / 32);
    // clusters that do not fit on the chip at once run in waves
    const int per_wave = pl.C > 1 ? (132 / pl.C) : 148;
    const int waves = (B + per_wave - 1) / per_wave;
    return (work + sync + red) * waves;
}

}  // namespace

extern "C" int apb_fps_block_size(int n) {
    // reference cuda_utils.h:10-14, same double log arithmetic
    if (n <= 0) return 1;
    const int pow_2 = (int)(std::log(static_cast<double>(n)) / std::log(2.0));
    int v = 1 << pow_2;
    if (v > 1024) v = 1024;
    if (v < 1) v = 1;
    return v;
}

extern "C" int apb_fps_f32(const float *xyz, float *temp, int32_t *idx, int B, int N, int m,
                           apb_stream_t stream) {
    if (B < 0 || N < 0 || m < 0) return APB_E_INVALID;
    if (B == 0 || m == 0) return APB_OK;
    if (!xyz || !idx || N == 0) return APB_E_INVALID;
    const int bs = apb_fps_block_size(N);

    FpsPlan best{};
    double best_cost = 1e300;
    int force_nt = 0, force_c = 0;
    if (const char *f = getenv("APB_FPS_FORCE")) sscanf(f, "%d,%d", &force_nt, &force_c);
    const int nts[] = {32, 64, 128, 256, 512, 1024};
    const int cs[] = {1, 2, 4, 8};
    for (int NT : nts)
        for (int C : cs) {
            if (force_nt && (NT != force_nt || C != force_c)) continue;
            if (NT < 256 && (C > 1 || NT != (bs < 32 ? 32 : bs))) continue;  // tiny clouds: T == bs
            FpsPlan pl;
            if (!plan_for(N, bs, NT, C, &pl)) continue;
            const double c = plan_cost(pl, B);
            if (c < best_cost) {
                best_cost = c;
                best = pl;
            }
        }
    if (!best.v) return APB_E_UNSUPPORTED;

    const int NT = best.v->NT, PPT = best.v->D * best.v->Q, C = best.C;
    const size_t smem = (size_t)3 * PPT * NT * sizeof(float) + (size_t)2 * 5 * C * (NT / 32) * sizeof(int);
    const fps_fn_t fn = C > 1 ? best.v->fn_multi : best.v->fn;
    if (!fn) return APB_E_UNSUPPORTED;
    APB_CUDA(cudaFuncSetAttribute((const void *)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

    FpsParams p{xyz, temp, idx, N, m, ilog2(bs), best.logT, best.QB, C};
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(C, B, 1);
    cfg.blockDim = dim3(NT, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    APB_CUDA(cudaLaunchKernelEx(&cfg, fn, p));
    return APB_OK;
}
