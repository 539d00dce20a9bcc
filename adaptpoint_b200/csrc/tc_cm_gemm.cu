// 1x1 convolution over channel-major activations on the sm_100a tensor cores, BOTH operands fed by TMA (TF32):
//     Y[b, :, p] = W . X[b, :, p]  (+ per-channel affine, ReLU)          X (B, Cin, P) f32, W (Cout, Kp) f32, Y (B, Cout, P)
// This is every shared-MLP / Conv1d / Conv2d 1x1 layer of the training path (reference: nn.Conv1d/Conv2d through cuDNN,
// openpoints/models/layers/conv.py:8-21, blocks conv.py:24-103), its data gradient (the same kernel on W^T), and -- through
// the epilogue variants -- the BatchNorm statistics pass that follows the conv (norm.py:57-71) and the conv -> ReLU ->
// max-over-points head of the discriminator (openpoints/models_adaptpoint/point_discriminator.py:166-191).
//
// Round-1 kernel (sa_fused.cu tc_rows_gemm_kernel) moved the activation tile global -> registers -> swizzled smem with four
// producer warps, which bound it at 14-22 % tensor-pipe activity.  Here no thread touches an operand:
//   * A tile = 128 consecutive points of one cloud x KC channels.  The activation is channel-major, i.e. MN-major for this
//     contraction (the reduction runs over channels), so it is loaded as FOUR TMA boxes {32 points, KC channels} and
//     consumed by tcgen05.mma through an MN-major shared-memory descriptor.  32-bit MN-major operands have exactly one
//     legal shared-memory layout, "128B swizzle with 32-byte atoms" (UMMA layout type 1 = CUTLASS
//     Layout_MN_SW128_32B_Atom, Swizzle<2,5,2>: the 32-byte chunk index of a 128-byte row is XORed with row & 3), which
//     is what TMA writes under CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B: atom = 32 points x 4 channels = 512 B, so
//     SBO (next 4 channels) = 512 B and LBO (next 32 points) = one box = KC*128 B.
//   * B tile = the weight slice, K-major, TMA boxes {32 k, NCP rows}; when the whole slice fits in shared memory it is
//     loaded ONCE per CTA and stays resident, so the per-tile L2 -> SM traffic is the activation only.
//   * accumulators: 2 x (128 lanes x <=256 columns) fp32 in TMEM, so the epilogue of tile t overlaps the loads and MMAs of
//     tile t+1; Cout > 256 is cut into column slices (blockIdx.y), the activation tile is then re-read from L2.
//   * epilogue (8 warps: TMEM lane quarter = warp % 4, two warps per quarter take alternate 32-column blocks):
//     tcgen05.ld -> affine (+ReLU) -> coalesced channel-major store (lanes = consecutive points) or point-major rows;
//     optionally sum / sum-of-squares per channel (BatchNorm batch statistics: 32x32 transpose through padded shared
//     memory, shared-memory partials, one fp64 atomic per channel per CTA) or max + arg-max over the points
//     (CREDUX + ballot, running maxima in shared memory, 64-bit atomicMax per channel when the CTA leaves a cloud).
// Persistent: one CTA per SM, contiguous tile ranges, 320 threads: warp 0 = TMA, warp 1 = TMEM alloc + MMA issue,
// warps 2-9 = epilogue.  HBM-bound by design: algorithmic bytes 4*B*P*(Cin + Cout) (PMAX: 4*B*P*Cin).
#include <cuda.h>
#include <climits>
#include "apb_common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int CM_BM = 128;     // points per tile (UMMA M)
constexpr int CM_KC = 32;      // channels per pipeline stage (4 UMMA K-steps of 8 tf32)
constexpr int CM_A_BYTES = CM_BM * CM_KC * 4;   // 16 KB: 4 boxes of {32 points, 32 channels}
constexpr int CM_BOX_BYTES = 32 * CM_KC * 4;    // one box = one MN atom column: 4 KB (= LBO)
constexpr int CM_EPI_WARPS = 8;   // two per TMEM lane quarter, interleaved over the 32-column blocks
constexpr int CM_EPI_THREADS = CM_EPI_WARPS * 32;
constexpr int CM_THREADS = 64 + CM_EPI_THREADS;
constexpr int CM_MAX_STAGES = 8;

enum CmEpilogue { CM_STORE_CM = 0, CM_STORE_PM = 1, CM_PMAX = 2 };

struct CmParams {
    int B, Cin, Cout, P;
    int kchunks;        // Kp / 32
    int NCP;            // output channels per CTA column (multiple of 16, <= 256)
    int tiles_per_b;    // ceil(P / 128)
    long long ntiles;   // B * tiles_per_b
    int stages;         // activation ring depth
    int w_resident;     // 1: the weight slice (kchunks x NCP x 128 B) stays in shared memory for the whole kernel
    const float *scale, *shift;   // (Cout) or nullptr
    int relu;
    float *Y;           // CM_STORE_CM: (B, Cout, P); CM_STORE_PM: (B*P, Cout)
    double *stats;      // nullptr or (Cout, 2): += sum, sum of squares of the (affine, pre-ReLU) outputs
    unsigned long long *pmax;   // CM_PMAX: (B, Cout) packed (ordered value << 32 | ~point), pre-zeroed
};

// MN-major 32-bit operand, SWIZZLE_128B_BASE32B: atoms of (32 elements along M) x (4 k-rows of 128 B)
__device__ __forceinline__ uint64_t smem_desc_mn_sw128_32b(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;   // leading byte offset: next 32-point atom along M
    d |= (uint64_t)(512 >> 4) << 32;                    // stride byte offset: next group of 4 k-rows
    d |= (uint64_t)1 << 46;                             // descriptor version (sm_100)
    d |= (uint64_t)1 << 61;                             // layout type 1 = SWIZZLE_128B_BASE32B
    return d;
}

__device__ __forceinline__ void cm_named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

template <int EPI>
__global__ void __launch_bounds__(CM_THREADS, 1)
tc_cm_gemm_kernel(const __grid_constant__ CUtensorMap xmap, const __grid_constant__ CUtensorMap wmap, const CmParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int w_chunk_bytes = p.NCP * 128;
    const int stage_bytes = CM_A_BYTES + (p.w_resident ? 0 : w_chunk_bytes);
    uint8_t *wres = smem + (size_t)p.stages * stage_bytes;                       // resident weights (if any)
    uint8_t *tail = wres + (p.w_resident ? (size_t)p.kchunks * w_chunk_bytes : 0);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);       // [CM_MAX_STAGES]
    uint64_t *empty_bar = full_bar + CM_MAX_STAGES;                // [CM_MAX_STAGES]
    uint64_t *tfull_bar = empty_bar + CM_MAX_STAGES;               // [2]
    uint64_t *tempty_bar = tfull_bar + 2;                          // [2]
    uint64_t *wfull_bar = tempty_bar + 2;                          // [1]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(wfull_bar + 1);
    float *ep_ss = reinterpret_cast<float *>(tail + 256);          // scale [256] | shift [256]
    double *st_acc = reinterpret_cast<double *>(ep_ss + 512);      // sum [256] | sumsq [256]: fp64, so the order in which
                                                                   // the warps' fp32 partials meet does not show
    unsigned long long *pm_acc = reinterpret_cast<unsigned long long *>(st_acc + 512);   // [256]
    float *tr_buf = reinterpret_cast<float *>(pm_acc + 256);       // [CM_EPI_WARPS][32][33] statistics transpose scratch

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n0 = blockIdx.y * p.NCP;
    const int Cout = min(p.NCP, p.Cout - n0);
    // contiguous tile range of this CTA
    const long long t_begin = p.ntiles * blockIdx.x / gridDim.x;
    const long long t_end = p.ntiles * (blockIdx.x + 1) / gridDim.x;

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&xmap);
        tc::tma_prefetch_desc(&wmap);
        for (int s = 0; s < p.stages; ++s) {
            tc::mbar_init(&full_bar[s], 1);
            tc::mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            tc::mbar_init(&tfull_bar[a], 1);
            tc::mbar_init(&tempty_bar[a], CM_EPI_THREADS);
        }
        tc::mbar_init(wfull_bar, 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
    if (tid >= 64) {
        for (int c = tid - 64; c < 256; c += CM_EPI_THREADS) {
            const bool ok = c < Cout;
            ep_ss[c] = (ok && p.scale) ? p.scale[n0 + c] : 1.f;
            ep_ss[256 + c] = (ok && p.shift) ? p.shift[n0 + c] : 0.f;
            st_acc[c] = 0.0;
            st_acc[256 + c] = 0.0;
            pm_acc[c] = 0ull;
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0 && t_begin < t_end) {
            if (p.w_resident) {
                tc::mbar_arrive_expect_tx(wfull_bar, (uint32_t)(p.kchunks * w_chunk_bytes));
                for (int kc = 0; kc < p.kchunks; ++kc)
                    tc::tma_load_2d(wres + (size_t)kc * w_chunk_bytes, &wmap, kc * CM_KC, n0, wfull_bar);
            }
            uint32_t it = 0;
            for (long long tile = t_begin; tile < t_end; ++tile) {
                const int b = (int)(tile / p.tiles_per_b);
                const int p0 = (int)(tile - (long long)b * p.tiles_per_b) * CM_BM;
                for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t *dst = smem + (size_t)s * stage_bytes;
                    tc::mbar_arrive_expect_tx(&full_bar[s], (uint32_t)stage_bytes);
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        tc::tma_load_3d(dst + i * CM_BOX_BYTES, &xmap, p0 + 32 * i, kc * CM_KC, b, &full_bar[s]);
                    if (!p.w_resident) tc::tma_load_2d(dst + CM_A_BYTES, &wmap, kc * CM_KC, n0, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0 && t_begin < t_end) {
            // A: MN-major (bit 15), B: K-major
            const uint32_t idesc = tc::idesc_tf32_f32(CM_BM, p.NCP) | (1u << 15);
            if (p.w_resident) {
                tc::mbar_wait(wfull_bar, 0);
                tc::tc_fence_after();
            }
            uint32_t it = 0, lt = 0;
            for (long long tile = t_begin; tile < t_end; ++tile, ++lt) {
                const int a = lt & 1;
                const uint32_t aph = (lt >> 1) & 1;
                tc::mbar_wait(&tempty_bar[a], aph ^ 1);
                tc::tc_fence_after();
                const uint32_t tacc = tmem_base + (uint32_t)a * 256;
                for (int kc = 0; kc < p.kchunks; ++kc, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_addr = tc::smem_u32(smem + (size_t)s * stage_bytes);
                    const uint32_t w_addr = p.w_resident ? tc::smem_u32(wres + (size_t)kc * w_chunk_bytes)
                                                         : a_addr + CM_A_BYTES;
#pragma unroll
                    for (int k = 0; k < CM_KC / 8; ++k) {   // UMMA_K = 8 tf32: one group of 8 channel rows (1024 B)
                        const uint64_t ad = smem_desc_mn_sw128_32b(a_addr + k * 1024, CM_BOX_BYTES);
                        const uint64_t bd = tc::smem_desc_sw128(w_addr + k * 32);
                        tc::umma_tf32(tacc, ad, bd, idesc, (kc | k) != 0);
                    }
                    tc::umma_commit(&empty_bar[s]);
                }
                tc::umma_commit(&tfull_bar[a]);
            }
        }
    } else {
        // ===================== epilogue: warp w owns TMEM lanes 32*(w%4).., column blocks cb = 32*half, +64, ... ======
        const int q4 = warp & 3;
        const int half = (warp - 2) >> 2;
        const int et = tid - 64;
        float *trw = tr_buf + (warp - 2) * (32 * 33);
        uint32_t lt = 0;
        int cur_b = -1;
        for (long long tile = t_begin; tile < t_end; ++tile, ++lt) {
            const int a = lt & 1;
            const uint32_t aph = (lt >> 1) & 1;
            const int b = (int)(tile / p.tiles_per_b);
            const int p0 = (int)(tile - (long long)b * p.tiles_per_b) * CM_BM;
            const int pp = p0 + q4 * 32 + lane;
            const bool valid = pp < p.P;
            if (EPI == CM_PMAX && b != cur_b) {
                if (cur_b >= 0) {   // the CTA leaves cloud cur_b: publish its running maxima
                    cm_named_bar(1, CM_EPI_THREADS);
                    for (int c = et; c < Cout; c += CM_EPI_THREADS) {
                        atomicMax(p.pmax + (size_t)cur_b * p.Cout + n0 + c, pm_acc[c]);
                        pm_acc[c] = 0ull;
                    }
                    cm_named_bar(1, CM_EPI_THREADS);
                }
                cur_b = b;
            }
            tc::mbar_wait(&tfull_bar[a], aph);
            tc::tc_fence_after();
            const uint32_t tacc = tmem_base + (uint32_t)a * 256 + ((uint32_t)(q4 * 32) << 16);
            float *ycm = p.Y + ((size_t)b * p.Cout + n0) * p.P + pp;
            for (int cb = half * 32; cb < p.NCP; cb += 64) {
                uint32_t v[32];
                tc::tmem_ld_32x32(tacc + cb, v);
                tc::tmem_ld_wait();
                float x[32];
#pragma unroll
                for (int j = 0; j < 32; j += 4) {   // (explicit ld.shared: see tc_ptx.cuh)
                    const float4 sc = tc::lds_f4_ro(ep_ss + cb + j), sh = tc::lds_f4_ro(ep_ss + 256 + cb + j);
                    x[j] = fmaf(sc.x, __uint_as_float(v[j]), sh.x);
                    x[j + 1] = fmaf(sc.y, __uint_as_float(v[j + 1]), sh.y);
                    x[j + 2] = fmaf(sc.z, __uint_as_float(v[j + 2]), sh.z);
                    x[j + 3] = fmaf(sc.w, __uint_as_float(v[j + 3]), sh.w);
                }
                if (p.stats != nullptr) {
                    // per-channel sum and sum of squares over this warp's 32 rows: 32x32 transpose through padded shared
                    // memory (conflict-free both ways), lane j then owns column cb + j
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 32; ++j) tc::sts_f(trw + lane * 33 + j, valid ? x[j] : 0.f);
                    __syncwarp();
                    float s1 = 0.f, s2 = 0.f;
#pragma unroll
                    for (int r = 0; r < 32; ++r) {
                        const float t = tc::lds_f(trw + r * 33 + lane);
                        s1 += t;
                        s2 = fmaf(t, t, s2);
                    }
                    if (cb + lane < Cout) {
                        tc::red_shared_f64(&st_acc[cb + lane], (double)s1);
                        tc::red_shared_f64(&st_acc[256 + cb + lane], (double)s2);
                    }
                }
                if (EPI == CM_PMAX) {
                    int best = INT_MIN, barg = 0;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int key = valid ? tc::float_to_ordered(x[j]) : INT_MIN;
                        const int m = __reduce_max_sync(0xffffffffu, key);
                        const unsigned hit = __ballot_sync(0xffffffffu, key == m);
                        if (lane == j) {
                            best = m;
                            barg = p0 + q4 * 32 + (__ffs(hit) - 1);
                        }
                    }
                    if (cb + lane < Cout) {
                        const unsigned long long packed =
                            ((unsigned long long)((unsigned)best ^ 0x80000000u) << 32) | (unsigned)(0xFFFFFFFFu - (unsigned)barg);
                        atomicMax(&pm_acc[cb + lane], packed);
                    }
                } else {
                    if (p.relu) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
                    }
                    if (EPI == CM_STORE_CM) {
                        if (valid) {
                            float *yp = ycm + (size_t)cb * p.P;
                            if (cb + 32 <= Cout) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) yp[(size_t)j * p.P] = x[j];
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (cb + j < Cout) yp[(size_t)j * p.P] = x[j];
                            }
                        }
                    } else if (valid) {
                        float *y = p.Y + ((size_t)b * p.P + pp) * p.Cout + n0 + cb;
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
        cm_named_bar(1, CM_EPI_THREADS);
        if (EPI == CM_PMAX && cur_b >= 0) {
            for (int c = et; c < Cout; c += CM_EPI_THREADS) atomicMax(p.pmax + (size_t)cur_b * p.Cout + n0 + c, pm_acc[c]);
        }
        if (p.stats != nullptr && t_begin < t_end) {
            for (int c = et; c < Cout; c += CM_EPI_THREADS) {
                atomicAdd(p.stats + 2 * (size_t)(n0 + c), st_acc[c]);
                atomicAdd(p.stats + 2 * (size_t)(n0 + c) + 1, st_acc[256 + c]);
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

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn cm_encode_fn() {
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

template <int EPI>
int launch_cm(CmParams p, const float *x, const float *w, int Kp, cudaStream_t st) {
    EncodeTiledFn fn = cm_encode_fn();
    if (!fn) return APB_E_UNSUPPORTED;
    p.kchunks = Kp / CM_KC;
    p.tiles_per_b = (p.P + CM_BM - 1) / CM_BM;
    p.ntiles = (long long)p.B * p.tiles_per_b;
    // column slices of <= 256 channels; when the row tiles alone cannot fill the SMs, thinner slices (>= 64 channels)
    int gy = (p.Cout + 255) / 256;
    if (p.ntiles * gy * 2 <= APB_NUM_SMS && p.Cout >= 128) {
        gy = (int)(APB_NUM_SMS / p.ntiles);
        if (gy > p.Cout / 64) gy = p.Cout / 64;
        if (gy < 1) gy = 1;
    }
    p.NCP = ((p.Cout + gy - 1) / gy + 15) / 16 * 16;
    gy = (p.Cout + p.NCP - 1) / p.NCP;
    const int w_chunk_bytes = p.NCP * 128;
    const size_t extra = 1024 /*align*/ + 256 /*barriers*/ + 2048 /*scale|shift*/ + 4096 /*stats f64*/ + 2048 /*pmax*/ +
                         (p.stats ? (size_t)CM_EPI_WARPS * 32 * 33 * 4 : 0) /*statistics transpose scratch*/;
    const size_t budget = 227 * 1024 - extra;
    const size_t w_all = (size_t)p.kchunks * w_chunk_bytes;
    p.w_resident = (w_all + 3 * (size_t)CM_A_BYTES <= budget) ? 1 : 0;
    const int stage_bytes = CM_A_BYTES + (p.w_resident ? 0 : w_chunk_bytes);
    int stages = (int)((budget - (p.w_resident ? w_all : 0)) / stage_bytes);
    if (stages > CM_MAX_STAGES) stages = CM_MAX_STAGES;
    if (stages < 2) return APB_E_UNSUPPORTED;
    p.stages = stages;
    const size_t smem = (size_t)stages * stage_bytes + (p.w_resident ? w_all : 0) + extra;

    CUtensorMap xmap, wmap;
    {   // X (B, Cin, P) fp32 read as TF32: box {32 points, 32 channels, 1 cloud}; out-of-range points / channels read 0
        const cuuint64_t gdim[3] = {(cuuint64_t)p.P, (cuuint64_t)p.Cin, (cuuint64_t)p.B};
        const cuuint64_t gstride[2] = {(cuuint64_t)p.P * 4, (cuuint64_t)p.Cin * p.P * 4};
        const cuuint32_t box[3] = {32, (cuuint32_t)CM_KC, 1};
        const cuuint32_t estr[3] = {1, 1, 1};
        if (fn(&xmap, CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 3, const_cast<float *>(x), gdim, gstride, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return APB_E_INVALID;
    }
    {   // W (Cout, Kp) fp32 read as TF32, K-major: box {32 k, NCP rows}; rows >= Cout read 0
        const cuuint64_t gdim[2] = {(cuuint64_t)Kp, (cuuint64_t)p.Cout};
        const cuuint64_t gstride[1] = {(cuuint64_t)Kp * 4};
        const cuuint32_t box[2] = {32, (cuuint32_t)p.NCP};
        const cuuint32_t estr[2] = {1, 1};
        if (fn(&wmap, CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 2, const_cast<float *>(w), gdim, gstride, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return APB_E_INVALID;
    }
    auto kern = tc_cm_gemm_kernel<EPI>;
    APB_CUDA(cudaFuncSetAttribute((const void *)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long gx = APB_NUM_SMS / gy;
    if (gx < 1) gx = 1;
    if (gx > p.ntiles) gx = p.ntiles;
    kern<<<dim3((unsigned)gx, (unsigned)gy), CM_THREADS, smem, st>>>(xmap, wmap, p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}


// ---- tiny layers in exact fp32 on the CUDA cores ------------------------------------------------------------------------
// Layers with <= 16 output channels or <= 8 input channels and at most 2048 weights, over >= 256 points (the imitator's
// mask head 64 -> 3, 6 -> 2 and the xyz embeddings 3 -> 64 on 2048-point clouds) hold no FLOPs worth a tensor core -- they are pure streaming -- but their
// outputs feed a tau = 0.1 Gumbel soft-max and tanh / sigmoid heads, where TF32 operand rounding is amplified 10x.  cuDNN
// makes the same call for such shapes (no TF32 kernel is selected), so computing them in fp32 keeps the product path as
// close to the reference's default precision as the large layers are.  One thread per point, lanes = consecutive points.
constexpr int TINY_COUT = 16, TINY_CIN = 8;

__global__ void __launch_bounds__(256) pwconv_small_cout_kernel(const float *__restrict__ x, const float *__restrict__ w,
                                                                const float *__restrict__ scale,
                                                                const float *__restrict__ shift, int relu,
                                                                float *__restrict__ y, int Cin, int Cout, int P, int Kp) {
    extern __shared__ float sw[];   // [Cout][Cin]
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < Cout * Cin; i += 256) sw[i] = w[(size_t)(i / Cin) * Kp + (i % Cin)];
    __syncthreads();
    const int pp = blockIdx.x * 256 + threadIdx.x;
    if (pp >= P) return;
    const float *xb = x + (size_t)b * Cin * P + pp;
    float acc[TINY_COUT];
#pragma unroll
    for (int c = 0; c < TINY_COUT; ++c) acc[c] = 0.f;
    for (int k = 0; k < Cin; ++k) {
        const float v = __ldg(xb + (size_t)k * P);
#pragma unroll
        for (int c = 0; c < TINY_COUT; ++c)
            if (c < Cout) acc[c] = fmaf(sw[c * Cin + k], v, acc[c]);
    }
    float *yb = y + (size_t)b * Cout * P + pp;
#pragma unroll
    for (int c = 0; c < TINY_COUT; ++c) {
        if (c < Cout) {
            float v = acc[c];
            if (scale) v *= scale[c];
            if (shift) v += shift[c];
            yb[(size_t)c * P] = relu ? fmaxf(v, 0.f) : v;
        }
    }
}

__global__ void __launch_bounds__(256) pwconv_small_cin_kernel(const float *__restrict__ x, const float *__restrict__ w,
                                                               const float *__restrict__ scale,
                                                               const float *__restrict__ shift, int relu,
                                                               float *__restrict__ y, int Cin, int Cout, int P, int Kp) {
    extern __shared__ float sw[];   // [Cout][Cin]
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < Cout * Cin; i += 256) sw[i] = w[(size_t)(i / Cin) * Kp + (i % Cin)];
    __syncthreads();
    const int pp = blockIdx.x * 256 + threadIdx.x;
    if (pp >= P) return;
    const float *xb = x + (size_t)b * Cin * P + pp;
    float v[TINY_CIN];
#pragma unroll
    for (int k = 0; k < TINY_CIN; ++k) v[k] = k < Cin ? __ldg(xb + (size_t)k * P) : 0.f;
    float *yb = y + (size_t)b * Cout * P + pp;
    for (int c = 0; c < Cout; ++c) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < TINY_CIN; ++k)
            if (k < Cin) acc = fmaf(sw[c * Cin + k], v[k], acc);
        if (scale) acc *= scale[c];
        if (shift) acc += shift[c];
        yb[(size_t)c * P] = relu ? fmaxf(acc, 0.f) : acc;
    }
}

int launch_tiny(const float *x, const float *w, const float *scale, const float *shift, int relu, float *y, int B, int Cin,
                int Cout, int P, int Kp, cudaStream_t st) {
    if (B > 65535 || (size_t)Cin * Cout * 4 > 96 * 1024) return APB_E_UNSUPPORTED;
    const size_t smem = (size_t)Cin * Cout * 4;
    const dim3 grid(apb_div_up(P, 256), B);
    if (Cout <= TINY_COUT) {
        APB_CUDA(cudaFuncSetAttribute((const void *)pwconv_small_cout_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        pwconv_small_cout_kernel<<<grid, 256, smem, st>>>(x, w, scale, shift, relu, y, Cin, Cout, P, Kp);
    } else {
        APB_CUDA(cudaFuncSetAttribute((const void *)pwconv_small_cin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        pwconv_small_cin_kernel<<<grid, 256, smem, st>>>(x, w, scale, shift, relu, y, Cin, Cout, P, Kp);
    }
    APB_CHECK_LAUNCH();
    return APB_OK;
}

bool cm_supported(const void *x, const void *w, int P, int Kp) {
    return (P % 4) == 0 && (Kp % CM_KC) == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w)) & 15) == 0;
}

}  // namespace

// sa_fused.cu: the round-1 register-staged kernel, kept for shapes TMA cannot address (P % 4 != 0)
extern "C" int apb_pwconv_fwd_tf32_rows(const float *x, const float *w, const float *scale, const float *shift, int relu,
                                        float *y, int B, int Cin, int Cout, int P, int Kp, apb_stream_t stream);

extern "C" int apb_pwconv_fwd_tf32(const float *x, const float *w, const float *scale, const float *shift, int relu,
                                   float *y, int B, int Cin, int Cout, int P, int Kp, apb_stream_t stream) {
    if (B < 0 || Cin <= 0 || Cout <= 0 || P < 0 || Kp < Cin) return APB_E_INVALID;
    if (B == 0 || P == 0) return APB_OK;
    if (!x || !w || !y) return APB_E_INVALID;
    // (one thread per point: only worth it when there are points to spread over the SMs and the weight block is small)
    if ((Cout <= TINY_COUT || Cin <= TINY_CIN) && Cin * Cout <= 2048 && P >= 256 && B <= 65535)
        return launch_tiny(x, w, scale, shift, relu, y, B, Cin, Cout, P, Kp, (cudaStream_t)stream);
    if (!cm_supported(x, w, P, Kp)) {
        if (Cout > 512 || (scale == nullptr) != (shift == nullptr)) return APB_E_UNSUPPORTED;
        return apb_pwconv_fwd_tf32_rows(x, w, scale, shift, relu, y, B, Cin, Cout, P, Kp, stream);
    }
    CmParams p{};
    p.B = B; p.Cin = Cin; p.Cout = Cout; p.P = P;
    p.scale = scale; p.shift = shift; p.relu = relu; p.Y = y;
    return launch_cm<CM_STORE_CM>(p, x, w, Kp, (cudaStream_t)stream);
}

extern "C" int apb_pwconv_ex_tf32(const float *x, const float *w, const float *scale, const float *shift, int relu,
                                  float *y, int point_major, double *stats, unsigned long long *pmax, int B, int Cin,
                                  int Cout, int P, int Kp, apb_stream_t stream) {
    if (B < 0 || Cin <= 0 || Cout <= 0 || P < 0 || Kp < Cin) return APB_E_INVALID;
    if (B == 0 || P == 0) return APB_OK;
    if (!x || !w || (!y && !pmax)) return APB_E_INVALID;
    if (!cm_supported(x, w, P, Kp)) return APB_E_UNSUPPORTED;
    CmParams p{};
    p.B = B; p.Cin = Cin; p.Cout = Cout; p.P = P;
    p.scale = scale; p.shift = shift; p.relu = relu; p.Y = y; p.stats = stats; p.pmax = pmax;
    cudaStream_t st = (cudaStream_t)stream;
    if (pmax) return launch_cm<CM_PMAX>(p, x, w, Kp, st);
    if (point_major) return launch_cm<CM_STORE_PM>(p, x, w, Kp, st);
    return launch_cm<CM_STORE_CM>(p, x, w, Kp, st);
}

// packed (ordered value << 32 | ~point) -> value, point
__global__ void pmax_decode_kernel(const unsigned long long *__restrict__ packed, float *__restrict__ val,
                                   int32_t *__restrict__ arg, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long v = packed[i];
    val[i] = tc::ordered_to_float((int)((unsigned)(v >> 32) ^ 0x80000000u));
    arg[i] = (int32_t)(0xFFFFFFFFu - (unsigned)(v & 0xFFFFFFFFull));
}

extern "C" int apb_pmax_decode(const unsigned long long *packed, float *val, int32_t *arg, long long n,
                               apb_stream_t stream) {
    if (n < 0) return APB_E_INVALID;
    if (n == 0) return APB_OK;
    if (!packed || !val || !arg) return APB_E_INVALID;
    pmax_decode_kernel<<<apb_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(packed, val, arg, n);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
