// Index-driven copies for sm_100a: grouping / gather / three_interpolate, forward and backward.
//
// Replaces group_points_kernel_fast / group_points_grad_kernel_fast (reference
// group_points_gpu.cu:53-72 / 14-31), gather_points_* (sampling_gpu.cu:15-31 / 53-70) and
// three_interpolate_* (interpolate_gpu.cu:84-104 / 127-149).
//
// Forward grouping is write-dominated (K-fold expansion).  The reference launches one thread per
// output element, re-reads idx once per channel and issues 4-byte scattered loads that each pull a
// 32-byte sector.  Here a CTA stages CH feature rows (channel-major, N floats each) in shared
// memory with coalesced 16-byte loads, reads each int4 of idx ONCE for CH channels, gathers from
// shared memory (bank-conflict bound, no sector waste) and emits float4 streaming stores, so HBM
// sees the algorithmic bytes: C*N + M*K reads (idx re-reads hit L2) + C*M*K writes.
#include "apb_common.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int GT = 256;  // threads per CTA (backward / interpolation kernels)

// One-dimensional TMA bulk copy global -> shared, completion signalled on an mbarrier.
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     tc::smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(tc::smem_u32(bar))
                 : "memory");
}

// out[b,c,p] = points[b,c,idx[b,p]]   p in [0,P) (P = M*K for grouping, m for gather)
// grid: (position chunks, channel groups, B); dynamic smem: CH*N floats.
// The CH rows are contiguous in HBM: when 16-byte aligned they arrive as TMA bulk copies (<= 32 KB
// pieces issued by the lanes of warp 0, one mbarrier), so the whole row set is in flight at once and
// no register or LSU slot is spent on staging; the gather loop keeps two idx quads in flight.
template <int CH, int NT>
__global__ void __launch_bounds__(NT) group_rows_kernel(const float *__restrict__ points,
                                                         const int32_t *__restrict__ idx,
                                                         float *__restrict__ out, int C, int N, long long P,
                                                         int chunk) {
    extern __shared__ __align__(128) float rows[];  // [CH][N]
    __shared__ uint64_t bar;
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CH;
    const int nch = min(CH, C - c0);
    const float *src = points + ((size_t)b * C + c0) * N;  // nch consecutive rows are contiguous
    const size_t total = (size_t)nch * N;
    const size_t bytes = total * sizeof(float);
    if (((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((bytes & 15) == 0)) {
        constexpr uint32_t PIECE = 32768;
        if (threadIdx.x == 0) {
            tc::mbar_init(&bar, 1);
            tc::fence_barrier_init();
            tc::mbar_arrive_expect_tx(&bar, (uint32_t)bytes);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            for (size_t off = (size_t)threadIdx.x * PIECE; off < bytes; off += (size_t)32 * PIECE)
                bulk_g2s(reinterpret_cast<char *>(rows) + off, reinterpret_cast<const char *>(src) + off,
                         (uint32_t)min((size_t)PIECE, bytes - off), &bar);
        }
        tc::mbar_wait(&bar, 0);
    } else {
        for (size_t i = threadIdx.x; i < total; i += NT) rows[i] = __ldg(src + i);
        __syncthreads();
    }

    const int32_t *ix = idx + (size_t)b * P;
    float *o = out + ((size_t)b * C + c0) * P;
    const long long p_begin = (long long)blockIdx.x * chunk;
    const long long p_end = min(P, p_begin + chunk);
    const bool vec_ok = ((P & 3) == 0) && ((reinterpret_cast<uintptr_t>(ix) & 15) == 0) &&
                        ((reinterpret_cast<uintptr_t>(o) & 15) == 0);
    if (vec_ok) {  // chunk is a multiple of 4
        long long p = p_begin + (long long)threadIdx.x * 4;
        int4 nxt = make_int4(0, 0, 0, 0);
        if (p < p_end) nxt = apb_ld_nc_i4(ix + p);
        for (; p < p_end; p += NT * 4) {
            const int4 i4 = nxt;
            if (p + NT * 4 < p_end) nxt = apb_ld_nc_i4(ix + p + NT * 4);
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                if (ch < nch) {
                    const float *r = rows + (size_t)ch * N;
                    float4 v;
                    v.x = r[i4.x]; v.y = r[i4.y]; v.z = r[i4.z]; v.w = r[i4.w];
                    apb_st_stream_f4(o + (size_t)ch * P + p, v);
                }
            }
        }
    } else {
        for (long long p = p_begin + threadIdx.x; p < p_end; p += NT) {
            const int i = __ldg(ix + p);
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (ch < nch) o[(size_t)ch * P + p] = rows[(size_t)ch * N + i];
        }
    }
}

// Rows longer than one SM's shared memory (N > 51200): nseg CTAs share one chunk of positions, CTA s stages slice
// s of the row (points [s*seg, (s+1)*seg)), reads the chunk's idx and stores only the elements whose source lies in
// its slice (4-byte stores, merged with the sibling CTAs' stores in L2: siblings are adjacent in launch order, so
// they run at the same time).  The gather itself stays in shared memory, so L2 never sees 4-byte reads that cost a
// 32-byte sector each.  (A thread-block-cluster variant that gathered from the peers' slices through DSMEM measured
// 1.7x SLOWER than the direct L2 gather: profiles/ops_group_cluster_dsmem_r01_q.jsonl.)
// grid: (position chunks * nseg, C, B); dynamic smem: seg floats.
template <int NT>
__global__ void __launch_bounds__(NT, 2) group_rows_seg_kernel(const float *__restrict__ points,
                                                                const int32_t *__restrict__ idx,
                                                                float *__restrict__ out, int C, int N, long long P,
                                                                int seg, int nseg, int chunk) {
    extern __shared__ __align__(128) float rows[];  // [seg]
    __shared__ uint64_t bar;
    const int b = blockIdx.z, c = blockIdx.y;
    const int s = blockIdx.x % nseg;
    const int s0 = s * seg;
    const int cnt = max(0, min(seg, N - s0));
    const float *src = points + ((size_t)b * C + c) * N + s0;
    const size_t bytes = (size_t)cnt * sizeof(float);
    if (cnt > 0 && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((bytes & 15) == 0)) {
        constexpr uint32_t PIECE = 32768;
        if (threadIdx.x == 0) {
            tc::mbar_init(&bar, 1);
            tc::fence_barrier_init();
            tc::mbar_arrive_expect_tx(&bar, (uint32_t)bytes);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            for (size_t off = (size_t)threadIdx.x * PIECE; off < bytes; off += (size_t)32 * PIECE)
                bulk_g2s(reinterpret_cast<char *>(rows) + off, reinterpret_cast<const char *>(src) + off,
                         (uint32_t)min((size_t)PIECE, bytes - off), &bar);
        }
        tc::mbar_wait(&bar, 0);
    } else {
        for (int i = threadIdx.x; i < cnt; i += NT) rows[i] = __ldg(src + i);
        __syncthreads();
    }
    const int32_t *ix = idx + (size_t)b * P;
    float *o = out + ((size_t)b * C + c) * P;
    const long long p_begin = (long long)(blockIdx.x / nseg) * chunk;
    const long long p_end = min(P, p_begin + chunk);
    const unsigned ucnt = (unsigned)cnt;
    if (((P & 3) == 0) && ((reinterpret_cast<uintptr_t>(ix) & 15) == 0)) {  // chunk is a multiple of 4
        long long p = p_begin + (long long)threadIdx.x * 4;
        int4 nxt = make_int4(0, 0, 0, 0);
        if (p < p_end) nxt = apb_ld_nc_i4(ix + p);
        for (; p < p_end; p += NT * 4) {
            const int4 i4 = nxt;
            if (p + NT * 4 < p_end) nxt = apb_ld_nc_i4(ix + p + NT * 4);
            const unsigned a0 = (unsigned)(i4.x - s0), a1 = (unsigned)(i4.y - s0), a2 = (unsigned)(i4.z - s0),
                           a3 = (unsigned)(i4.w - s0);
            if (a0 < ucnt) o[p + 0] = rows[a0];
            if (a1 < ucnt) o[p + 1] = rows[a1];
            if (a2 < ucnt) o[p + 2] = rows[a2];
            if (a3 < ucnt) o[p + 3] = rows[a3];
        }
    } else {
        for (long long p = p_begin + threadIdx.x; p < p_end; p += NT) {
            const unsigned a = (unsigned)(__ldg(ix + p) - s0);
            if (a < ucnt) o[p] = rows[a];
        }
    }
}

// Fallback when one row does not fit in shared memory (N > ~56k): direct gather through L1/L2.
__global__ void __launch_bounds__(GT) group_direct_kernel(const float *__restrict__ points,
                                                           const int32_t *__restrict__ idx,
                                                           float *__restrict__ out, int C, int N, long long P) {
    const int b = blockIdx.z, c = blockIdx.y;
    const float *r = points + ((size_t)b * C + c) * N;
    const int32_t *ix = idx + (size_t)b * P;
    float *o = out + ((size_t)b * C + c) * P;
    for (long long p = (long long)blockIdx.x * GT + threadIdx.x; p < P; p += (long long)gridDim.x * GT)
        o[p] = __ldg(r + __ldg(ix + p));
}

// grad_points[b,c,idx[b,p]] += grad_out[b,c,p]     (grad_points pre-zeroed by the caller)
// fp32 reductions at L2 (RED.ADD.F32), one idx read per CHB channels.
constexpr int CHB = 4;
__global__ void __launch_bounds__(GT) group_grad_kernel(const float *__restrict__ grad_out,
                                                         const int32_t *__restrict__ idx,
                                                         float *__restrict__ grad_points, int C, int N,
                                                         long long P) {
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CHB;
    const int nch = min(CHB, C - c0);
    const int32_t *ix = idx + (size_t)b * P;
    const float *g = grad_out + ((size_t)b * C + c0) * P;
    float *gp = grad_points + ((size_t)b * C + c0) * N;
    const bool vec_ok = ((P & 3) == 0) && ((reinterpret_cast<uintptr_t>(ix) & 15) == 0) &&
                        ((reinterpret_cast<uintptr_t>(g) & 15) == 0);
    if (vec_ok) {
        for (long long p = ((long long)blockIdx.x * GT + threadIdx.x) * 4; p < P; p += (long long)gridDim.x * GT * 4) {
            const int4 i4 = apb_ld_nc_i4(ix + p);
#pragma unroll
            for (int ch = 0; ch < CHB; ++ch) {
                if (ch < nch) {
                    const float4 v = apb_ld_nc_f4(g + (size_t)ch * P + p);
                    float *r = gp + (size_t)ch * N;
                    atomicAdd(r + i4.x, v.x);
                    atomicAdd(r + i4.y, v.y);
                    atomicAdd(r + i4.z, v.z);
                    atomicAdd(r + i4.w, v.w);
                }
            }
        }
    } else {
        for (long long p = (long long)blockIdx.x * GT + threadIdx.x; p < P; p += (long long)gridDim.x * GT) {
            const int i = __ldg(ix + p);
            for (int ch = 0; ch < nch; ++ch) atomicAdd(gp + (size_t)ch * N + i, __ldg(g + (size_t)ch * P + p));
        }
    }
}

// Coalesced fire-and-forget accumulation of a shared-memory accumulator into global memory (RED at L2, 16 B
// per request when aligned); zero entries are skipped.  dst/src cover `total` consecutive floats.
__device__ __forceinline__ void red_add_f4(float *p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
template <int NT>
__device__ __forceinline__ void flush_rows(float *__restrict__ dst, const float *rows, size_t total) {
    if (((reinterpret_cast<uintptr_t>(dst) & 15) == 0) && ((total & 3) == 0) && ((tc::smem_u32(rows) & 15) == 0)) {
        const float4 *r4 = reinterpret_cast<const float4 *>(rows);
        for (size_t i = threadIdx.x; i < total / 4; i += NT) {
            const float4 v = r4[i];
            if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) red_add_f4(dst + i * 4, v);
        }
    } else {
        for (size_t i = threadIdx.x; i < total; i += NT) {
            const float v = rows[i];
            if (v != 0.f) atomicAdd(dst + i, v);
        }
    }
}

// Backward of grouping with the accumulator rows in shared memory: a CTA owns CH channel rows of one cloud
// (CH*N floats, zeroed), streams its chunk of grad_out (float4) and idx (int4, read once for the CH channels),
// accumulates with shared-memory atomics (no L2 round trip per element, no sector amplification) and flushes the
// rows with coalesced 16-byte reductions.  HBM sees C*P + P reads and ~C*N reduction traffic = the algorithmic bytes.
// grid: (position chunks, channel groups, B); dynamic smem: CH*N floats.
template <int CH, int NT>
__global__ void __launch_bounds__(NT) group_grad_rows_kernel(const float *__restrict__ grad_out,
                                                              const int32_t *__restrict__ idx,
                                                              float *__restrict__ grad_points, int C, int N,
                                                              long long P, long long chunk, int seg, int nseg) {
    extern __shared__ __align__(128) float rows[];  // [CH][cnt]
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CH;
    const int nch = min(CH, C - c0);
    // rows longer than shared memory: nseg sibling CTAs (adjacent in launch order) read the same chunk and each
    // accumulates only the sources that fall in its slice [s0, s0+cnt) of the row; nseg == 1 otherwise
    const int s0 = (int)(blockIdx.x % nseg) * seg;
    const int cnt = max(0, min(seg, N - s0));
    const unsigned ucnt = (unsigned)cnt;
    const size_t total = (size_t)nch * cnt;
    for (size_t i = threadIdx.x; i < total; i += NT) rows[i] = 0.f;
    __syncthreads();
    const int32_t *ix = idx + (size_t)b * P;
    const float *g = grad_out + ((size_t)b * C + c0) * P;
    const long long p_begin = (long long)(blockIdx.x / nseg) * chunk;
    const long long p_end = min(P, p_begin + chunk);
    const bool vec_ok = ((P & 3) == 0) && ((reinterpret_cast<uintptr_t>(ix) & 15) == 0) &&
                        ((reinterpret_cast<uintptr_t>(g) & 15) == 0);
    if (vec_ok) {  // chunk is a multiple of 4
        for (long long p = p_begin + (long long)threadIdx.x * 4; p < p_end; p += NT * 4) {
            const int4 i4 = apb_ld_nc_i4(ix + p);
            const unsigned a0 = (unsigned)(i4.x - s0), a1 = (unsigned)(i4.y - s0), a2 = (unsigned)(i4.z - s0),
                           a3 = (unsigned)(i4.w - s0);
            float4 v[CH];
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (ch < nch) v[ch] = apb_ld_nc_f4(g + (size_t)ch * P + p);
            // runs of one index inside the thread's four slots are summed in registers first: ball-query rows are
            // padded with their first index (ball_query_gpu.cu:41-47), and same-address shared atomics serialise
            const bool e1 = i4.y == i4.x, e2 = i4.z == i4.y, e3 = i4.w == i4.z;
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                if (ch < nch) {
                    float *r = rows + (size_t)ch * cnt;
                    float x0 = v[ch].x, x1 = v[ch].y, x2 = v[ch].z, x3 = v[ch].w;
                    if (e3) x2 += x3;
                    if (e2) x1 += x2;
                    if (e1) x0 += x1;
                    if (a0 < ucnt) atomicAdd(r + a0, x0);
                    if (!e1 && a1 < ucnt) atomicAdd(r + a1, x1);
                    if (!e2 && a2 < ucnt) atomicAdd(r + a2, x2);
                    if (!e3 && a3 < ucnt) atomicAdd(r + a3, x3);
                }
            }
        }
    } else {
        for (long long p = p_begin + threadIdx.x; p < p_end; p += NT) {
            const unsigned a = (unsigned)(__ldg(ix + p) - s0);
            if (a < ucnt)
                for (int ch = 0; ch < nch; ++ch) atomicAdd(rows + (size_t)ch * cnt + a, __ldg(g + (size_t)ch * P + p));
        }
    }
    __syncthreads();
    for (int ch = 0; ch < nch; ++ch)
        flush_rows<NT>(grad_points + ((size_t)b * C + c0 + ch) * N + s0, rows + (size_t)ch * cnt, (size_t)cnt);
}

// out[b,c,i] = w1*p[i1] (+) fma(w0,p[i0]) (+) fma(w2,p[i2])   -- the reference's contracted order
// (interpolate_gpu.cu:103 as compiled: FMUL on term 1, FFMA term 0, FFMA term 2).
// grid: (point chunks, channel groups, B); smem: CH rows of m floats.
template <int CH>
__global__ void __launch_bounds__(GT) three_interp_kernel(const float *__restrict__ points,
                                                           const int32_t *__restrict__ idx,
                                                           const float *__restrict__ weight,
                                                           float *__restrict__ out, int C, int m, int n) {
    extern __shared__ float rows[];
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CH;
    const int nch = min(CH, C - c0);
    const float *src = points + ((size_t)b * C + c0) * m;
    for (size_t i = threadIdx.x; i < (size_t)nch * m; i += GT) rows[i] = __ldg(src + i);
    __syncthreads();
    const int32_t *ix = idx + (size_t)b * n * 3;
    const float *w = weight + (size_t)b * n * 3;
    float *o = out + ((size_t)b * C + c0) * n;
    for (int i = blockIdx.x * GT + threadIdx.x; i < n; i += gridDim.x * GT) {
        const int i0 = __ldg(ix + i * 3 + 0), i1 = __ldg(ix + i * 3 + 1), i2 = __ldg(ix + i * 3 + 2);
        const float w0 = __ldg(w + i * 3 + 0), w1 = __ldg(w + i * 3 + 1), w2 = __ldg(w + i * 3 + 2);
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) {
            if (ch < nch) {
                const float *r = rows + (size_t)ch * m;
                float t = __fmul_rn(w1, r[i1]);
                t = __fmaf_rn(w0, r[i0], t);
                t = __fmaf_rn(w2, r[i2], t);
                o[(size_t)ch * n + i] = t;
            }
        }
    }
}

__global__ void __launch_bounds__(GT) three_interp_direct_kernel(const float *__restrict__ points,
                                                                  const int32_t *__restrict__ idx,
                                                                  const float *__restrict__ weight,
                                                                  float *__restrict__ out, int C, int m, int n) {
    const int b = blockIdx.z, c = blockIdx.y;
    const float *r = points + ((size_t)b * C + c) * m;
    const int32_t *ix = idx + (size_t)b * n * 3;
    const float *w = weight + (size_t)b * n * 3;
    float *o = out + ((size_t)b * C + c) * n;
    for (int i = blockIdx.x * GT + threadIdx.x; i < n; i += gridDim.x * GT) {
        float t = __fmul_rn(__ldg(w + i * 3 + 1), __ldg(r + __ldg(ix + i * 3 + 1)));
        t = __fmaf_rn(__ldg(w + i * 3 + 0), __ldg(r + __ldg(ix + i * 3 + 0)), t);
        t = __fmaf_rn(__ldg(w + i * 3 + 2), __ldg(r + __ldg(ix + i * 3 + 2)), t);
        o[i] = t;
    }
}

// grad_points[b,c,idx[b,i,j]] += grad_out[b,c,i] * weight[b,i,j]   (pre-zeroed accumulator)
__global__ void __launch_bounds__(GT) three_interp_grad_kernel(const float *__restrict__ grad_out,
                                                                const int32_t *__restrict__ idx,
                                                                const float *__restrict__ weight,
                                                                float *__restrict__ grad_points, int C, int n,
                                                                int m) {
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CHB;
    const int nch = min(CHB, C - c0);
    const int32_t *ix = idx + (size_t)b * n * 3;
    const float *w = weight + (size_t)b * n * 3;
    const float *g = grad_out + ((size_t)b * C + c0) * n;
    float *gp = grad_points + ((size_t)b * C + c0) * m;
    for (int i = blockIdx.x * GT + threadIdx.x; i < n; i += gridDim.x * GT) {
        const int i0 = __ldg(ix + i * 3 + 0), i1 = __ldg(ix + i * 3 + 1), i2 = __ldg(ix + i * 3 + 2);
        const float w0 = __ldg(w + i * 3 + 0), w1 = __ldg(w + i * 3 + 1), w2 = __ldg(w + i * 3 + 2);
        for (int ch = 0; ch < nch; ++ch) {
            const float gv = __ldg(g + (size_t)ch * n + i);
            float *r = gp + (size_t)ch * m;
            atomicAdd(r + i0, __fmul_rn(gv, w0));
            atomicAdd(r + i1, __fmul_rn(gv, w1));
            atomicAdd(r + i2, __fmul_rn(gv, w2));
        }
    }
}

// Same scheme for the backward of three_interpolate: CH accumulator rows of m floats in shared memory.
template <int CH, int NT>
__global__ void __launch_bounds__(NT) three_interp_grad_rows_kernel(const float *__restrict__ grad_out,
                                                                     const int32_t *__restrict__ idx,
                                                                     const float *__restrict__ weight,
                                                                     float *__restrict__ grad_points, int C, int n,
                                                                     int m, int chunk) {
    extern __shared__ __align__(128) float rows[];  // [CH][m]
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CH;
    const int nch = min(CH, C - c0);
    const size_t total = (size_t)nch * m;
    for (size_t i = threadIdx.x; i < total; i += NT) rows[i] = 0.f;
    __syncthreads();
    const int32_t *ix = idx + (size_t)b * n * 3;
    const float *w = weight + (size_t)b * n * 3;
    const float *g = grad_out + ((size_t)b * C + c0) * n;
    const int i_end = min(n, (int)(blockIdx.x + 1) * chunk);
    for (int i = blockIdx.x * chunk + threadIdx.x; i < i_end; i += NT) {
        const int i0 = __ldg(ix + i * 3 + 0), i1 = __ldg(ix + i * 3 + 1), i2 = __ldg(ix + i * 3 + 2);
        const float w0 = __ldg(w + i * 3 + 0), w1 = __ldg(w + i * 3 + 1), w2 = __ldg(w + i * 3 + 2);
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) {
            if (ch < nch) {
                const float gv = __ldg(g + (size_t)ch * n + i);
                float *r = rows + (size_t)ch * m;
                atomicAdd(r + i0, __fmul_rn(gv, w0));
                atomicAdd(r + i1, __fmul_rn(gv, w1));
                atomicAdd(r + i2, __fmul_rn(gv, w2));
            }
        }
    }
    __syncthreads();
    flush_rows<NT>(grad_points + ((size_t)b * C + c0) * m, rows, total);
}

constexpr size_t kSmemBudget = 200 * 1024;

// Slice size of the sliced (row > shared memory) kernels; benchmark override apb_tune_group_seg (KB, 0 = automatic).
size_t g_seg_budget = 200 * 1024;

// Benchmark-only override of the forward work split (apb_tune_group): 0 = automatic.
int g_tune_ch = 0, g_tune_nt = 0;

template <int CH, int NT>
int launch_group_rows(const float *points, const int32_t *idx, float *out, int B, int C, int N, long long P,
                      cudaStream_t st) {
    const size_t smem = (size_t)CH * N * sizeof(float);
    APB_CUDA(cudaFuncSetAttribute((const void *)group_rows_kernel<CH, NT>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // positions per CTA: enough output per staged row (>= 8x the row bytes) but keep the grid wide
    long long chunk = (long long)8 * N;
    if (chunk < 4096) chunk = 4096;
    const int cgroups = apb_div_up(C, CH);
    while (chunk > 4096 && (long long)apb_div_up(P, chunk) * cgroups * B < 4 * APB_NUM_SMS) chunk >>= 1;
    chunk = (chunk + 4095) / 4096 * 4096;  // multiple of NT*4 for every NT
    dim3 grid(apb_div_up(P, chunk), cgroups, B);
    group_rows_kernel<CH, NT><<<grid, NT, smem, st>>>(points, idx, out, C, N, P, (int)chunk);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

template <int CH>
int launch_group_rows_nt(int nt, const float *points, const int32_t *idx, float *out, int B, int C, int N,
                         long long P, cudaStream_t st) {
    if (nt >= 1024) return launch_group_rows<CH, 1024>(points, idx, out, B, C, N, P, st);
    if (nt >= 512) return launch_group_rows<CH, 512>(points, idx, out, B, C, N, P, st);
    return launch_group_rows<CH, 256>(points, idx, out, B, C, N, P, st);
}

int launch_group_rows_seg(const float *points, const int32_t *idx, float *out, int B, int C, int N, long long P,
                          cudaStream_t st) {
    constexpr int NT = 1024;
    const int nseg = apb_div_up((long long)N * 4, g_seg_budget);
    const int seg = (apb_div_up(N, nseg) + 3) / 4 * 4;
    const size_t smem = (size_t)seg * sizeof(float);
    APB_CUDA(cudaFuncSetAttribute((const void *)group_rows_seg_kernel<NT>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // every CTA stages seg floats for `chunk` positions: chunk >= 4*seg keeps the staging under 1/4 of the stores
    long long chunk = (4LL * seg + 4095) / 4096 * 4096;
    while (chunk > 4096 && (long long)apb_div_up(P, chunk) * nseg * C * B < 2 * APB_NUM_SMS)
        chunk = (chunk / 2 + 4095) / 4096 * 4096;
    dim3 grid((unsigned)(apb_div_up(P, chunk) * nseg), C, B);
    group_rows_seg_kernel<NT><<<grid, NT, smem, st>>>(points, idx, out, C, N, P, seg, nseg, (int)chunk);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

int group_forward(const float *points, const int32_t *idx, float *out, int B, int C, int N, long long P,
                  cudaStream_t st) {
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!points || !idx || !out || N <= 0) return APB_E_INVALID;
    if (C > 65535 || B > 65535) return APB_E_UNSUPPORTED;
    const size_t row = (size_t)N * sizeof(float);
    if (row > kSmemBudget) {
        // expanding calls only: a gather (P <= 2N) reads each slice for too few outputs to pay for staging it
        if (row <= 16 * kSmemBudget && g_tune_ch >= 0 && P > 2LL * N)
            return launch_group_rows_seg(points, idx, out, B, C, N, P, st);
        dim3 grid((unsigned)min((long long)apb_div_up(P, GT), (long long)8 * APB_NUM_SMS), C, B);
        group_direct_kernel<<<grid, GT, 0, st>>>(points, idx, out, C, N, P);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    // Work split (fitted on B200, profiles/ops_group_tune_r01.jsonl).  Expanding calls (grouping, P >> N) are
    // store-bound: stage as many channels per idx read as fit and give the CTA enough warps to cover the
    // idx latency (1024 threads once the row set owns more than a quarter of the SM's shared memory).
    // Load-bound calls (gather, P <= 2N) instead want several small CTAs per SM so one CTA's row copy
    // overlaps another's gathers.
    const int fit = (int)(kSmemBudget / row);
    int ch = 1, nt = 256;
    if (P > 2LL * N) {
        if (C >= 4 && fit >= 4) ch = 4;
        else if (C >= 4 && fit >= 2) ch = 2;
        const size_t smem = ch * row;
        nt = smem > 50 * 1024 ? 1024 : (smem > 25 * 1024 ? 512 : 256);
        if (ch == 1 && smem > 16 * 1024) nt = smem > 100 * 1024 ? 1024 : 512;
    } else {
        constexpr size_t kThird = 64 * 1024;
        if (C >= 4 && 4 * row <= kThird) ch = 4;
        else if (C >= 2 && 2 * row <= kThird) ch = 2;
        nt = ch * row > 16 * 1024 ? 512 : 256;
        if (row > 100 * 1024) nt = 1024;
    }
    if (g_tune_ch > 0 && g_tune_ch <= fit && g_tune_ch <= C) ch = g_tune_ch;
    if (g_tune_nt > 0) nt = g_tune_nt;
    if (ch >= 4) return launch_group_rows_nt<4>(nt, points, idx, out, B, C, N, P, st);
    if (ch >= 2) return launch_group_rows_nt<2>(nt, points, idx, out, B, C, N, P, st);
    return launch_group_rows_nt<1>(nt, points, idx, out, B, C, N, P, st);
}

// Grouping fused with the max / min over the K neighbours of each query:
//   vmax[b,c,j] = max_k points[b,c,idx[b,j,k]]  (amax[b,c,j] = the POINT index attaining it, first k on ties), same for min.
// This is what "group -> affine normalise -> max over neighbours" (generator_component4_15.py:368-431, PointsetGrouper) and
// "group -> max" (Producefactor, :560-563) need: alpha*(g - centre) + beta is monotone in g, so the max over k of the
// normalised tensor is the normalised max (alpha >= 0) or min (alpha < 0), bit for bit, and the (B,C,M,K) tensor, its
// four elementwise passes and their backward passes disappear; the backward is a scatter of (B,C,M) values.
template <int CH, int NT>
__global__ void __launch_bounds__(NT) group_minmax_kernel(const float *__restrict__ points,
                                                           const int32_t *__restrict__ idx, float *__restrict__ vmax,
                                                           float *__restrict__ vmin, int32_t *__restrict__ amax,
                                                           int32_t *__restrict__ amin, int C, int N, int M, int K,
                                                           int mchunk) {
    extern __shared__ __align__(128) float rows[];  // [CH][N]
    __shared__ uint64_t bar;
    const int b = blockIdx.z;
    const int c0 = blockIdx.y * CH;
    const int nch = min(CH, C - c0);
    const float *src = points + ((size_t)b * C + c0) * N;
    const size_t total = (size_t)nch * N;
    const size_t bytes = total * sizeof(float);
    if (((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((bytes & 15) == 0)) {
        constexpr uint32_t PIECE = 32768;
        if (threadIdx.x == 0) {
            tc::mbar_init(&bar, 1);
            tc::fence_barrier_init();
            tc::mbar_arrive_expect_tx(&bar, (uint32_t)bytes);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            for (size_t off = (size_t)threadIdx.x * PIECE; off < bytes; off += (size_t)32 * PIECE)
                bulk_g2s(reinterpret_cast<char *>(rows) + off, reinterpret_cast<const char *>(src) + off,
                         (uint32_t)min((size_t)PIECE, bytes - off), &bar);
        }
        tc::mbar_wait(&bar, 0);
    } else {
        for (size_t i = threadIdx.x; i < total; i += NT) rows[i] = __ldg(src + i);
        __syncthreads();
    }
    const int j_end = min(M, (int)(blockIdx.x + 1) * mchunk);
    const bool vec = (K & 3) == 0 && ((reinterpret_cast<uintptr_t>(idx) & 15) == 0);
    for (int j = blockIdx.x * mchunk + threadIdx.x; j < j_end; j += NT) {
        const int32_t *ix = idx + ((size_t)b * M + j) * K;
        float mx[CH], mn[CH];
        int ax[CH], an[CH];
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) {
            mx[ch] = -INFINITY; mn[ch] = INFINITY; ax[ch] = 0; an[ch] = 0;
        }
        auto upd = [&](int i) {
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                if (ch < nch) {
                    const float v = rows[(size_t)ch * N + i];
                    if (v > mx[ch]) { mx[ch] = v; ax[ch] = i; }
                    if (v < mn[ch]) { mn[ch] = v; an[ch] = i; }
                }
            }
        };
        if (vec) {
            for (int k = 0; k < K; k += 4) {
                const int4 i4 = *reinterpret_cast<const int4 *>(ix + k);
                upd(i4.x); upd(i4.y); upd(i4.z); upd(i4.w);
            }
        } else {
            for (int k = 0; k < K; ++k) upd(__ldg(ix + k));
        }
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) {
            if (ch < nch) {
                const size_t o = ((size_t)b * C + c0 + ch) * M + j;
                vmax[o] = mx[ch]; vmin[o] = mn[ch]; amax[o] = ax[ch]; amin[o] = an[ch];
            }
        }
    }
}

template <int CH>
int launch_group_minmax(const float *points, const int32_t *idx, float *vmax, float *vmin, int32_t *amax,
                        int32_t *amin, int B, int C, int N, int M, int K, cudaStream_t st) {
    constexpr int NT = 256;
    const size_t smem = (size_t)CH * N * sizeof(float);
    APB_CUDA(cudaFuncSetAttribute((const void *)group_minmax_kernel<CH, NT>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int cgroups = apb_div_up(C, CH);
    int mchunk = M;
    while (mchunk > 2 * NT && (long long)apb_div_up(M, mchunk) * cgroups * B < 2 * APB_NUM_SMS) mchunk = (mchunk + 1) / 2;
    dim3 grid(apb_div_up(M, mchunk), cgroups, B);
    group_minmax_kernel<CH, NT><<<grid, NT, smem, st>>>(points, idx, vmax, vmin, amax, amin, C, N, M, K, mchunk);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

template <int CH, int NT>
int launch_group_grad_rows(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C, int N,
                           long long P, cudaStream_t st) {
    const int nseg = (size_t)CH * N * 4 <= kSmemBudget ? 1 : apb_div_up((long long)CH * N * 4, g_seg_budget);  // > 1 only for CH == 1, N > 51200
    const int seg = nseg == 1 ? N : (apb_div_up(N, nseg) + 3) / 4 * 4;
    const size_t smem = (size_t)CH * seg * sizeof(float);
    APB_CUDA(cudaFuncSetAttribute((const void *)group_grad_rows_kernel<CH, NT>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // positions per CTA: every chunk pays a zero + flush of its CH rows, so split P only while the grid is too
    // narrow to fill the SMs and a chunk still carries >= 2x the row length
    const int cgroups = apb_div_up(C, CH);
    long long chunk = (P + 3) / 4 * 4;
    while (chunk >= 4LL * seg && chunk > NT * 8 &&
           (long long)apb_div_up(P, chunk) * nseg * cgroups * B < 2 * APB_NUM_SMS)
        chunk = ((chunk + 1) / 2 + 3) / 4 * 4;
    dim3 grid((unsigned)(apb_div_up(P, chunk) * nseg), cgroups, B);
    group_grad_rows_kernel<CH, NT><<<grid, NT, smem, st>>>(grad_out, idx, grad_points, C, N, P, chunk, seg, nseg);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

template <int CH>
int launch_group_grad_rows_nt(int nt, const float *grad_out, const int32_t *idx, float *grad_points, int B, int C,
                              int N, long long P, cudaStream_t st) {
    if (nt >= 1024) return launch_group_grad_rows<CH, 1024>(grad_out, idx, grad_points, B, C, N, P, st);
    if (nt >= 512) return launch_group_grad_rows<CH, 512>(grad_out, idx, grad_points, B, C, N, P, st);
    return launch_group_grad_rows<CH, 256>(grad_out, idx, grad_points, B, C, N, P, st);
}

// Benchmark-only override of the backward path (apb_tune_group_grad): 0 = automatic, 1 = L2 atomics, 2 = shared rows.
int g_tune_grad_mode = 0;

int group_backward(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C, int N,
                   long long P, cudaStream_t st) {
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!grad_out || !idx || !grad_points || N <= 0) return APB_E_INVALID;
    if (C > 65535 || B > 65535) return APB_E_UNSUPPORTED;
    const size_t row = (size_t)N * sizeof(float);
    // Expanding calls (grouping backward, P >> N): accumulate in shared memory.  Sparse calls (gather backward,
    // P <= 2N) touch each accumulator at most a few times: zeroing + flushing whole rows would cost more than the
    // scattered L2 reductions, so they keep the direct kernel, as do rows that do not fit in shared memory.
    const bool rows_path = (g_tune_grad_mode == 2) || (g_tune_grad_mode == 0 && P > 2LL * N);
    if (rows_path && row > kSmemBudget && row <= 16 * kSmemBudget)
        return launch_group_grad_rows<1, 1024>(grad_out, idx, grad_points, B, C, N, P, st);
    if (rows_path && row <= kSmemBudget) {
        const int fit = (int)(kSmemBudget / row);
        int ch = 1;
        if (C >= 4 && fit >= 4) ch = 4;
        else if (C >= 2 && fit >= 2) ch = 2;
        if (g_tune_ch > 0 && g_tune_ch <= fit && g_tune_ch <= C) ch = g_tune_ch;
        const size_t smem = ch * row;
        int nt = smem > 50 * 1024 ? 1024 : (smem > 25 * 1024 ? 512 : 256);
        if (g_tune_nt > 0) nt = g_tune_nt;
        if (ch >= 4) return launch_group_grad_rows_nt<4>(nt, grad_out, idx, grad_points, B, C, N, P, st);
        if (ch >= 2) return launch_group_grad_rows_nt<2>(nt, grad_out, idx, grad_points, B, C, N, P, st);
        return launch_group_grad_rows_nt<1>(nt, grad_out, idx, grad_points, B, C, N, P, st);
    }
    dim3 grid((unsigned)min((long long)apb_div_up(P, GT * 4), (long long)4 * APB_NUM_SMS), apb_div_up(C, CHB), B);
    group_grad_kernel<<<grid, GT, 0, st>>>(grad_out, idx, grad_points, C, N, P);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

}  // namespace

extern "C" int apb_tune_group(int channels_per_cta, int threads) {
    if (channels_per_cta != -1 && channels_per_cta != 0 && channels_per_cta != 1 && channels_per_cta != 2 &&
        channels_per_cta != 4)
        return APB_E_INVALID;
    if (threads != 0 && threads != 256 && threads != 512 && threads != 1024) return APB_E_INVALID;
    g_tune_ch = channels_per_cta;
    g_tune_nt = threads;
    return APB_OK;
}
extern "C" int apb_tune_group_seg(int kb) {
    if (kb != 0 && (kb < 16 || kb > 200)) return APB_E_INVALID;
    g_seg_budget = (size_t)(kb ? kb : 200) * 1024;
    return APB_OK;
}
extern "C" int apb_tune_group_grad(int mode) {
    if (mode < 0 || mode > 2) return APB_E_INVALID;
    g_tune_grad_mode = mode;
    return APB_OK;
}
extern "C" int apb_group_f32(const float *points, const int32_t *idx, float *out, int B, int C, int N, int M,
                             int K, apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || M < 0 || K < 0) return APB_E_INVALID;
    return group_forward(points, idx, out, B, C, N, (long long)M * K, (cudaStream_t)stream);
}
extern "C" int apb_group_minmax_f32(const float *points, const int32_t *idx, float *vmax, float *vmin, int32_t *amax,
                                    int32_t *amin, int B, int C, int N, int M, int K, apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || M < 0 || K < 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || M == 0) return APB_OK;
    if (!points || !idx || !vmax || !vmin || !amax || !amin || N <= 0 || K <= 0) return APB_E_INVALID;
    if (C > 65535 || B > 65535) return APB_E_UNSUPPORTED;
    const size_t row = (size_t)N * sizeof(float);
    if (row > kSmemBudget) return APB_E_UNSUPPORTED;  // callers fall back to apb_group_f32 + a reduction
    cudaStream_t st = (cudaStream_t)stream;
    constexpr size_t kThird = 64 * 1024;
    if (C >= 4 && 4 * row <= kThird) return launch_group_minmax<4>(points, idx, vmax, vmin, amax, amin, B, C, N, M, K, st);
    if (C >= 2 && 2 * row <= kThird) return launch_group_minmax<2>(points, idx, vmax, vmin, amax, amin, B, C, N, M, K, st);
    return launch_group_minmax<1>(points, idx, vmax, vmin, amax, amin, B, C, N, M, K, st);
}
extern "C" int apb_group_grad_f32(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C,
                                  int N, int M, int K, apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || M < 0 || K < 0) return APB_E_INVALID;
    return group_backward(grad_out, idx, grad_points, B, C, N, (long long)M * K, (cudaStream_t)stream);
}
extern "C" int apb_gather_f32(const float *points, const int32_t *idx, float *out, int B, int C, int N, int m,
                              apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || m < 0) return APB_E_INVALID;
    return group_forward(points, idx, out, B, C, N, (long long)m, (cudaStream_t)stream);
}
extern "C" int apb_gather_grad_f32(const float *grad_out, const int32_t *idx, float *grad_points, int B, int C,
                                   int N, int m, apb_stream_t stream) {
    if (B < 0 || C < 0 || N < 0 || m < 0) return APB_E_INVALID;
    return group_backward(grad_out, idx, grad_points, B, C, N, (long long)m, (cudaStream_t)stream);
}

extern "C" int apb_three_interp_f32(const float *points, const int32_t *idx, const float *weight, float *out,
                                    int B, int C, int m, int n, apb_stream_t stream) {
    if (B < 0 || C < 0 || m < 0 || n < 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || n == 0) return APB_OK;
    if (!points || !idx || !weight || !out || m <= 0) return APB_E_INVALID;
    if (B > 65535) return APB_E_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)m * sizeof(float);
    const int xblocks = min(apb_div_up(n, GT), 4 * APB_NUM_SMS);
    // A CTA stages CH rows of m floats and must produce enough output to pay for them: at least 4 output points per
    // staged point (round 2's sweep: 8*GT points per CTA re-staged 128 KB for 32 KB of output at m = 8192 and ran
    // BELOW the reference kernel), fewer only when the grid would otherwise not cover the SMs.
    auto xgrid = [&](int groups) {
        long long per = (long long)4 * m;
        if (per < 8 * GT) per = 8 * GT;
        while (per > 8 * GT && (long long)apb_div_up(n, per) * groups * B < 2 * APB_NUM_SMS) per >>= 1;
        return (int)min((long long)xblocks, max(1LL, (long long)apb_div_up(n, per)));
    };
    if (row * 4 <= kSmemBudget && C >= 4) {
        const size_t smem = row * 4;
        APB_CUDA(cudaFuncSetAttribute((const void *)three_interp_kernel<4>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dim3 grid(xgrid(apb_div_up(C, 4)), apb_div_up(C, 4), B);
        three_interp_kernel<4><<<grid, GT, smem, st>>>(points, idx, weight, out, C, m, n);
    } else if (row <= kSmemBudget) {
        APB_CUDA(cudaFuncSetAttribute((const void *)three_interp_kernel<1>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)row));
        dim3 grid(xgrid(C), C, B);
        three_interp_kernel<1><<<grid, GT, row, st>>>(points, idx, weight, out, C, m, n);
    } else {
        dim3 grid(xblocks, C, B);
        three_interp_direct_kernel<<<grid, GT, 0, st>>>(points, idx, weight, out, C, m, n);
    }
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_three_interp_grad_f32(const float *grad_out, const int32_t *idx, const float *weight,
                                         float *grad_points, int B, int C, int n, int m, apb_stream_t stream) {
    if (B < 0 || C < 0 || m < 0 || n < 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || n == 0) return APB_OK;
    if (!grad_out || !idx || !weight || !grad_points || m <= 0) return APB_E_INVALID;
    if (B > 65535) return APB_E_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)m * sizeof(float);
    if (g_tune_grad_mode != 1 && n >= m && row <= kSmemBudget && C <= 65535) {
        // accumulator rows in shared memory (see group_grad_rows_kernel)
        constexpr int NT = 256;
        const int ch = (C >= 4 && 4 * row <= 64 * 1024) ? 4 : 1;
        const int cgroups = apb_div_up(C, ch);
        int chunk = n;
        while (chunk >= 8 * m && chunk > NT * 4 && (long long)apb_div_up(n, chunk) * cgroups * B < 2 * APB_NUM_SMS)
            chunk = (chunk + 1) / 2;
        dim3 grid(apb_div_up(n, chunk), cgroups, B);
        const size_t smem = ch * row;
        if (ch == 4) {
            APB_CUDA(cudaFuncSetAttribute((const void *)three_interp_grad_rows_kernel<4, NT>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            three_interp_grad_rows_kernel<4, NT><<<grid, NT, smem, st>>>(grad_out, idx, weight, grad_points, C, n, m, chunk);
        } else {
            APB_CUDA(cudaFuncSetAttribute((const void *)three_interp_grad_rows_kernel<1, NT>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            three_interp_grad_rows_kernel<1, NT><<<grid, NT, smem, st>>>(grad_out, idx, weight, grad_points, C, n, m, chunk);
        }
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    dim3 grid(min(apb_div_up(n, GT), 4 * APB_NUM_SMS), apb_div_up(C, CHB), B);
    three_interp_grad_kernel<<<grid, GT, 0, st>>>(grad_out, idx, weight, grad_points, C, n, m);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
