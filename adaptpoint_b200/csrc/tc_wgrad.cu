// Weight gradient of a 1x1 convolution on the sm_100a tensor cores, TMA-fed, TF32:
//   gW[co, ci] += sum_{b,p} gY[b, co, p] * X[b, ci, p]
// (what torch computes through cuDNN's wgrad for nn.Conv1d/Conv2d kernel 1 in the reference,
// openpoints/models/layers/conv.py:8-21; no native counterpart in the reference).
//
// Both operands are channel-major (B, C, P) fp32, i.e. K-major for this contraction (K = the points), so they go
// global -> shared memory by TMA (3-D tensor maps {P, C, B}, box {32 points, rows, 1}, 128-byte swizzle, TF32
// rounding done by the TMA unit) and straight into tcgen05.mma.kind::tf32 -- no thread touches the operands.
// Split-K over (b, 32-point chunks): grid = (ksplit, groups of MT Cout tiles of 128, Cin tiles of nb <= 256); a CTA keeps
// MT accumulators of 128 x nb in TMEM (MT*nb <= 512 columns), so one x tile feeds up to four gY tiles and the operands
// are re-read (Cout/128/MT) and (Cin/nb) times instead of (Cout/128) and (Cin/nb): (MT, nb) is chosen to minimise
// that traffic.  Each CTA adds its slice to gW with fp32 RED at L2 (gW is pre-zeroed by the caller, as the reference
// does for its gradient buffers).  HBM-bound: algorithmic bytes 4*B*P*(Cout + Cin).
#include <cuda.h>
#include "apb_common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int WG_BM = 128;  // Cout rows per tile (UMMA M)
constexpr int WG_BK = 32;   // points per chunk = 128 bytes of fp32 = one swizzle span
constexpr int WG_THREADS = 192;
constexpr int WG_MAX_STAGES = 8;
int g_wg_force_mt = 0, g_wg_force_kb = 0;
long long g_wg_min_chunks = 16;

struct WgParams {
    int Cout, Cin, P, B;
    int mt;          // Cout tiles of 128 rows per CTA (1, 2 or 4)
    int kb;          // 32-point boxes per operand per pipeline stage (1, 2 or 4): fewer, larger stages for thin layers
    int arows;       // rows of one gY box: 128, or Cout rounded up to 8 when Cout < 128 (rows beyond are never read back)
    int nb;          // Cin rows per tile (multiple of 16, mt*nb <= 512)
    int pchunks;     // ceil(P / 32)
    long long total; // B * pchunks
    int stages;
    float *gW;       // (Cout, ldw) row pitch ldw >= Cin
    int ldw;
};

__global__ void __launch_bounds__(WG_THREADS, 1)
tc_wgrad_kernel(const __grid_constant__ CUtensorMap gy_map, const __grid_constant__ CUtensorMap x_map,
                const WgParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int a_box = p.arows * 128, b_box = p.nb * 128;
    const int a_bytes = p.kb * p.mt * a_box;
    const int stage_bytes = a_bytes + p.kb * b_box;
    uint8_t *tail = smem + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);
    uint64_t *empty_bar = full_bar + WG_MAX_STAGES;
    uint64_t *done_bar = empty_bar + WG_MAX_STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(done_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int mg = blockIdx.y, nt = blockIdx.z;   // group of p.mt Cout tiles, Cin tile
    // K slice of this CTA
    const long long per = (p.total + gridDim.x - 1) / gridDim.x;
    const long long k_begin = (long long)blockIdx.x * per;
    const long long k_end = k_begin + per < p.total ? k_begin + per : p.total;
    const bool has_work = k_begin < k_end;

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&gy_map);
        tc::tma_prefetch_desc(&x_map);
        for (int s = 0; s < p.stages; ++s) {
            tc::mbar_init(&full_bar[s], 1);
            tc::mbar_init(&empty_bar[s], 1);
        }
        tc::mbar_init(done_bar, 1);
        tc::fence_barrier_init();
    }
    const uint32_t tmem_cols = p.mt * p.nb <= 256 ? 256 : 512;   // two co-resident CTAs (thin layers) must both fit
    if (warp == 1) tc::tmem_alloc(tmem_slot, tmem_cols);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (has_work) {
        if (warp == 0) {
            if (lane == 0) {
                uint32_t it = 0;
                for (long long kk = k_begin; kk < k_end; ++kk, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
                    const int b = (int)(kk / p.pchunks);
                    const int pc = (int)(kk - (long long)b * p.pchunks);
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t *dst = smem + (size_t)s * stage_bytes;
                    tc::mbar_arrive_expect_tx(&full_bar[s], (uint32_t)stage_bytes);
                    for (int j = 0; j < p.kb; ++j) {
                        const int pt = (pc * p.kb + j) * WG_BK;   // points beyond P read zeros
                        for (int i = 0; i < p.mt; ++i)            // rows beyond Cout read zeros
                            tc::tma_load_3d(dst + (j * p.mt + i) * a_box, &gy_map, pt, (mg * p.mt + i) * WG_BM, b, &full_bar[s]);
                        tc::tma_load_3d(dst + a_bytes + j * b_box, &x_map, pt, nt * p.nb, b, &full_bar[s]);
                    }
                }
            }
        } else if (warp == 1) {
            if (lane == 0) {
                const uint32_t idesc = tc::idesc_tf32_f32(WG_BM, p.nb);
                uint32_t it = 0;
                for (long long kk = k_begin; kk < k_end; ++kk, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_addr = tc::smem_u32(smem + (size_t)s * stage_bytes);
                    const uint32_t b_addr = a_addr + a_bytes;
                    for (int j = 0; j < p.kb; ++j) {
                        for (int i = 0; i < p.mt; ++i) {
#pragma unroll
                            for (int k = 0; k < WG_BK / 8; ++k)  // UMMA_K = 8 tf32 = 32 bytes
                                tc::umma_tf32(tmem_base + (uint32_t)(i * p.nb),
                                              tc::smem_desc_sw128(a_addr + (j * p.mt + i) * a_box + k * 32),
                                              tc::smem_desc_sw128(b_addr + j * b_box + k * 32), idesc, (it | j | k) != 0);
                        }
                    }
                    tc::umma_commit(&empty_bar[s]);
                }
                tc::umma_commit(done_bar);
            }
        } else {
            // epilogue: warp w reads TMEM lanes 32*(w%4).. = Cout rows of the tile
            const int q4 = warp & 3;
            tc::mbar_wait(done_bar, 0);
            tc::tc_fence_after();
            for (int i = 0; i < p.mt; ++i) {
                const int co = (mg * p.mt + i) * WG_BM + q4 * 32 + lane;
                if ((mg * p.mt + i) * WG_BM >= p.Cout) break;   // whole tile beyond Cout (uniform over the CTA)
                const uint32_t tacc = tmem_base + (uint32_t)(i * p.nb) + ((uint32_t)(q4 * 32) << 16);
                for (int cb = 0; cb < p.nb; cb += 32) {
                    uint32_t v[32];
                    tc::tmem_ld_32x32(tacc + cb, v);
                    tc::tmem_ld_wait();
                    if (co < p.Cout) {
                        float *dst = p.gW + (size_t)co * p.ldw + nt * p.nb + cb;
                        // valid columns of this block (a padded pitch may take the zero columns Cin .. ldw-1 as well)
                        const int ncol = min(min(32, p.nb - cb), p.ldw - nt * p.nb - cb);
                        if ((p.ldw & 3) == 0 && (reinterpret_cast<uintptr_t>(p.gW) & 15) == 0) {   // rows are 16-byte aligned: 128-bit reductions, 4 columns per instruction
#pragma unroll
                            for (int j = 0; j < 32; j += 4) {
                                if (j + 4 <= ncol)
                                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j),
                                                 "f"(__uint_as_float(v[j])), "f"(__uint_as_float(v[j + 1])),
                                                 "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3]))
                                                 : "memory");
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (j < ncol) atomicAdd(dst + j, __uint_as_float(v[j]));
                        }
                    }
                }
            }
            tc::tc_fence_before();
        }
    }
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, tmem_cols);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn wg_encode_fn() {
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

// (B, C, P) fp32 channel-major tensor, box = 32 points x rows x 1 batch, read as TF32
int encode_cm_map(CUtensorMap *map, const float *t, int B, int C, int P, int rows) {
    EncodeTiledFn fn = wg_encode_fn();
    if (!fn) return APB_E_UNSUPPORTED;
    const cuuint64_t gdim[3] = {(cuuint64_t)P, (cuuint64_t)C, (cuuint64_t)B};
    const cuuint64_t gstride[2] = {(cuuint64_t)P * 4, (cuuint64_t)C * P * 4};
    const cuuint32_t box[3] = {(cuuint32_t)WG_BK, (cuuint32_t)rows, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 3, const_cast<float *>(t), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? APB_OK : APB_E_INVALID;
}

}  // namespace

// Benchmarks only: force MT (0 = choose by traffic) and the minimum K chunks per CTA of the split-K heuristic.
extern "C" int apb_tune_wgrad(int force_mt, int min_chunks) {
    if (force_mt >= 10) {   // 10 + kb: force the boxes per stage instead (10 = automatic)
        const int kb = force_mt - 10;
        if (!(kb == 0 || kb == 1 || kb == 2 || kb == 4)) return APB_E_INVALID;
        g_wg_force_kb = kb;
        return APB_OK;
    }
    if (!(force_mt == 0 || force_mt == 1 || force_mt == 2 || force_mt == 4) || min_chunks < 0) return APB_E_INVALID;
    g_wg_force_mt = force_mt;
    g_wg_min_chunks = min_chunks > 0 ? min_chunks : 16;
    return APB_OK;
}

extern "C" int apb_pwconv_wgrad_ld_tf32(const float *grad_y, const float *x, float *grad_w, int ldw, int B, int Cin,
                                        int Cout, int P, apb_stream_t stream);

extern "C" int apb_pwconv_wgrad_tf32(const float *grad_y, const float *x, float *grad_w, int B, int Cin, int Cout,
                                     int P, apb_stream_t stream) {
    return apb_pwconv_wgrad_ld_tf32(grad_y, x, grad_w, Cin, B, Cin, Cout, P, stream);
}

extern "C" int apb_pwconv_wgrad_ld_tf32(const float *grad_y, const float *x, float *grad_w, int ldw, int B, int Cin,
                                        int Cout, int P, apb_stream_t stream) {
    if (B < 0 || Cin <= 0 || Cout <= 0 || P < 0 || ldw < Cin) return APB_E_INVALID;
    if (B == 0 || P == 0) return APB_OK;
    if (!grad_y || !x || !grad_w) return APB_E_INVALID;
    if ((P & 3) != 0 || ((reinterpret_cast<uintptr_t>(grad_y) | reinterpret_cast<uintptr_t>(x)) & 15) != 0)
        return APB_E_UNSUPPORTED;  // TMA needs 16-byte aligned rows
    WgParams p{};
    p.Cout = Cout; p.Cin = Cin; p.P = P; p.B = B;
    const int cin16 = (Cin + 15) / 16 * 16;
    const int mtiles = (Cout + WG_BM - 1) / WG_BM;
    // (MT, nb): MT accumulators of 128 x nb in the 512 TMEM columns.  Operand traffic per K element is proportional to
    // Cin * ceil(mtiles / MT) + Cout * ntiles; take the cheapest of MT in {1, 2, 4}.
    int best_mt = 1, best_nb = 0, ntiles = 0;
    long long best_cost = -1;
    for (int mt = 1; mt <= 4; mt *= 2) {
        if (mt > 1 && mt / 2 >= mtiles) break;
        const int cap = (512 / mt) < 256 ? (512 / mt) : 256;
        const int nt_ = (cin16 + cap - 1) / cap;
        const int nb_ = ((cin16 + nt_ - 1) / nt_ + 15) / 16 * 16;
        const long long cost = (long long)Cin * ((mtiles + mt - 1) / mt) + (long long)Cout * nt_;
        // measured (profiles/wgrad_sweep_r02_i.jsonl): MT > 1 loses 30-50 % on every shape of the step (fewer pipeline
        // stages, more split-K atomics), so MT = 1 unless a benchmark forces it
        if (g_wg_force_mt ? mt == g_wg_force_mt : mt == 1) {
            best_cost = cost; best_mt = mt; best_nb = nb_; ntiles = nt_;
        }
    }
    if (best_nb == 0) return APB_E_UNSUPPORTED;
    p.mt = best_mt;
    p.nb = best_nb;
    const int mgroups = (mtiles + p.mt - 1) / p.mt;
    p.arows = (mtiles > 1 || Cout >= WG_BM) ? WG_BM : (Cout + 7) / 8 * 8;
    // boxes per stage: thin layers move < 10 KB per 32-point chunk and were bound by the per-stage round trip
    // (0.6 us per chunk whatever its size); give them up to 4 boxes per stage while >= 3 stages still fit
    p.kb = 1;
    for (int kb = g_wg_force_kb ? g_wg_force_kb : 4; kb > 1; kb >>= 1) {
        if ((200 * 1024) / (kb * (p.mt * p.arows * 128 + p.nb * 128)) >= 3 || g_wg_force_kb) {
            p.kb = kb;
            break;
        }
    }
    p.pchunks = (P + WG_BK * p.kb - 1) / (WG_BK * p.kb);
    p.total = (long long)B * p.pchunks;
    const int stage_bytes = p.kb * (p.mt * p.arows * 128 + p.nb * 128);
    int stages = (200 * 1024) / stage_bytes;
    if (stages > WG_MAX_STAGES) stages = WG_MAX_STAGES;
    if (stages < 2) return APB_E_UNSUPPORTED;
    p.stages = stages;
    p.gW = grad_w;
    p.ldw = ldw;
    // a thin gY box (arows < 128) is still read as 128 rows by the UMMA: keep those reads inside the allocation
    size_t smem = (size_t)stages * stage_bytes + 1024 + 256 + (p.arows < WG_BM ? WG_BM * 128 : 0);
    if (p.mt * p.nb > 256 && smem < 120 * 1024) smem = 120 * 1024;   // 512 TMEM columns: one CTA per SM
    CUtensorMap gy_map, x_map;
    int rc = encode_cm_map(&gy_map, grad_y, B, Cout, P, p.arows);
    if (rc != APB_OK) return rc;
    rc = encode_cm_map(&x_map, x, B, Cin, P, p.nb);
    if (rc != APB_OK) return rc;
    APB_CUDA(cudaFuncSetAttribute((const void *)tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long ksplit = (long long)APB_NUM_SMS / (mgroups * ntiles);  // one wave: 1 CTA per SM (smem-bound)
    // every split adds a full 128 x nb tile to gW with fp32 atomics: a CTA must have enough K to pay for that (a
    // (256 -> 512, P = 64) layer split 37 ways spent 186 us in its 4.8 M atomics, round 2 trace)
    const long long min_chunks = g_wg_min_chunks;
    if (ksplit > p.total / min_chunks) ksplit = p.total / min_chunks;
    if (ksplit < 1) ksplit = 1;
    if (ksplit > p.total) ksplit = p.total;
    dim3 grid((unsigned)ksplit, mgroups, ntiles);
    tc_wgrad_kernel<<<grid, WG_THREADS, smem, (cudaStream_t)stream>>>(gy_map, x_map, p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
