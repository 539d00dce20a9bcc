// Ball query through a uniform cell grid: O(M * neighbourhood) distance evaluations instead of O(M * N).
//
// Reference semantics (openpoints/cpp/pointnet2_batch/src/ball_query_gpu.cu:15-51): for every query, the FIRST nsample
// support indices IN INDEX ORDER with d2 < r^2 (strict, r^2 = radius*radius in fp32), padded with the first one; an
// empty ball leaves the pre-zeroed row.  The reference (and query.cu's tiled kernel) walk all N support points per
// query; at N >= 16 k that is > 98 % wasted pairs (a ball of radius 0.1 holds ~40 of 16 384 shell points).
// Here the support cloud is binned once per call into cells of side >= r (per cloud, <= 32 cells per axis), so every
// neighbour of a query lies in the 27 cells around it:
//   bq_bbox / bq_count / bq_scan / bq_fill   build cell_start (B, ncell+1) and the cell-ordered copy of the points
//                                            (x, y, z, original index) -- counting sort, order inside a cell arbitrary;
//   bq_grid_query                            one warp per query scans the 27 cells (coalesced float4 reads, the
//                                            reference's exact distance expression), appends hits to a shared-memory
//                                            buffer, and selects the nsample SMALLEST indices by rank; once nsample hits
//                                            are known their largest index T prunes every later candidate with one
//                                            integer compare (idx >= T can never be among the first nsample).
// The result is bit-identical to the index-order walk because "first nsample in index order" == "nsample smallest
// indices among the hits".  Work: 27 * (points per cell) pair evaluations per query; bytes negligible; fp32-ALU bound.
#include <cfloat>
#include <climits>
#include "apb_common.cuh"

namespace {

constexpr int GMAX = 32;                 // cells per axis (<= 32768 cells per cloud)
constexpr int NCELL_MAX = GMAX * GMAX * GMAX;
constexpr int CAP = 128;                 // hit buffer per warp
constexpr int QW = 8;                    // warps (queries) per CTA

struct Grid {        // per cloud, derived from its bounding box
    float mn[3], inv[3];
    int g[3];
};

__device__ __forceinline__ Grid make_grid(const float *bbox, float radius) {
    Grid gr;
    const float cs0 = radius * 1.0001f;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float lo = bbox[d], hi = bbox[3 + d];
        const float ext = fmaxf(hi - lo, 0.f);
        float cs = fmaxf(cs0, ext / GMAX * 1.0001f);
        if (!(cs > 0.f)) cs = 1.f;       // radius 0 and a degenerate axis: one cell
        int g = (int)floorf(ext / cs) + 1;
        gr.g[d] = min(max(g, 1), GMAX);
        gr.mn[d] = lo;
        gr.inv[d] = 1.f / cs;
    }
    return gr;
}
__device__ __forceinline__ int cell_coord(const Grid &gr, int d, float x) {
    const float t = floorf((x - gr.mn[d]) * gr.inv[d]);
    return (int)fminf(fmaxf(t, 0.f), (float)(gr.g[d] - 1));   // monotone clamp: neighbours stay neighbours
}

// bbox (B, 6): min xyz, max xyz; also zeroes the cloud's cell counters (cell_start doubles as the counter array)
__global__ void __launch_bounds__(256) bq_bbox_kernel(const float *__restrict__ xyz, float *__restrict__ bbox,
                                                      int *__restrict__ cell_start, int N) {
    __shared__ float red[6][8];
    const int b = blockIdx.x;
    const float *p = xyz + (size_t)b * N * 3;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = threadIdx.x; i < N; i += 256) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const float v = __ldg(p + (size_t)i * 3 + d);
            lo[d] = fminf(lo[d], v);
            hi[d] = fmaxf(hi[d], v);
        }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[d] = fminf(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
            hi[d] = fmaxf(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], o));
        }
        if (lane == 0) { red[d][warp] = lo[d]; red[3 + d][warp] = hi[d]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = red[threadIdx.x][0];
        for (int w = 1; w < 8; ++w) v = threadIdx.x < 3 ? fminf(v, red[threadIdx.x][w]) : fmaxf(v, red[threadIdx.x][w]);
        bbox[b * 6 + threadIdx.x] = v;
    }
    int *cs = cell_start + (size_t)b * (NCELL_MAX + 1);
    for (int i = threadIdx.x; i <= NCELL_MAX; i += 256) cs[i] = 0;
}

__global__ void __launch_bounds__(256) bq_count_kernel(const float *__restrict__ xyz, const float *__restrict__ bbox,
                                                       int *__restrict__ cell_start, int *__restrict__ cell_of, int N,
                                                       float radius) {
    const int b = blockIdx.y;
    const Grid gr = make_grid(bbox + b * 6, radius);
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= N) return;
    const float *p = xyz + ((size_t)b * N + i) * 3;
    const int c = (cell_coord(gr, 2, __ldg(p + 2)) * gr.g[1] + cell_coord(gr, 1, __ldg(p + 1))) * gr.g[0] +
                  cell_coord(gr, 0, __ldg(p));
    cell_of[(size_t)b * N + i] = c;
    atomicAdd(cell_start + (size_t)b * (NCELL_MAX + 1) + c + 1, 1);   // counts shifted by one: scan gives starts
}

// in-place inclusive scan of counts[1..ncell] -> cell_start; cursor (B, NCELL_MAX) = a copy of the starts
__global__ void __launch_bounds__(1024) bq_scan_kernel(int *__restrict__ cell_start, int *__restrict__ cursor) {
    __shared__ int part[1024];
    const int b = blockIdx.x;
    int *cs = cell_start + (size_t)b * (NCELL_MAX + 1);
    constexpr int PER = NCELL_MAX / 1024;   // 32 consecutive cells per thread
    int local[PER];
    int sum = 0;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        sum += cs[1 + threadIdx.x * PER + j];
        local[j] = sum;
    }
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    const int base = part[threadIdx.x] - sum;
    int *cur = cursor + (size_t)b * NCELL_MAX;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const int cell = threadIdx.x * PER + j;
        cs[1 + cell] = base + local[j];                       // end of `cell` == start of `cell + 1`
        cur[cell] = base + (j ? local[j - 1] : 0);            // start of `cell`
    }
}

__global__ void __launch_bounds__(256) bq_fill_kernel(const float *__restrict__ xyz, const int *__restrict__ cell_of,
                                                      int *__restrict__ cursor, float4 *__restrict__ sorted, int N) {
    const int b = blockIdx.y;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= N) return;
    const int c = cell_of[(size_t)b * N + i];
    const int slot = atomicAdd(cursor + (size_t)b * NCELL_MAX + c, 1);
    const float *p = xyz + ((size_t)b * N + i) * 3;
    sorted[(size_t)b * N + slot] = make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), __int_as_float(i));
}

// One warp per query.  hits: per-warp buffers [2][CAP] in shared memory.
__global__ void __launch_bounds__(QW * 32) bq_grid_query_kernel(const float *__restrict__ new_xyz,
                                                                const float *__restrict__ bbox,
                                                                const int *__restrict__ cell_start,
                                                                const float4 *__restrict__ sorted,
                                                                int32_t *__restrict__ idx, int N, int M, float radius,
                                                                int K) {
    __shared__ int bufs[QW][2][CAP];
    const int b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * QW + warp;
    if (q >= M) return;
    const Grid gr = make_grid(bbox + b * 6, radius);
    const float radius2 = __fmul_rn(radius, radius);   // ball_query_gpu.cu:29
    const float *qp = new_xyz + ((size_t)b * M + q) * 3;
    const float qx = __ldg(qp), qy = __ldg(qp + 1), qz = __ldg(qp + 2);
    const int cx = cell_coord(gr, 0, qx), cy = cell_coord(gr, 1, qy), cz = cell_coord(gr, 2, qz);
    const int *cs = cell_start + (size_t)b * (NCELL_MAX + 1);
    const float4 *pts = sorted + (size_t)b * N;
    int *cur = bufs[warp][0], *alt = bufs[warp][1];
    int cnt = 0, T = INT_MAX;

    // keep the K smallest of cur[0..cnt) in ascending order (rank = number of smaller entries; indices are distinct)
    auto select = [&]() {
        __syncwarp();
        for (int i = lane; i < cnt; i += 32) {
            const int v = cur[i];
            int rank = 0;
            for (int j = 0; j < cnt; ++j) rank += cur[j] < v;
            if (rank < K) alt[rank] = v;
        }
        __syncwarp();
        int *t = cur; cur = alt; alt = t;
        cnt = min(cnt, K);
        if (cnt == K) T = cur[K - 1];
        __syncwarp();
    };

    for (int dz = -1; dz <= 1; ++dz) {
        const int z = cz + dz;
        if (z < 0 || z >= gr.g[2]) continue;
        for (int dy = -1; dy <= 1; ++dy) {
            const int y = cy + dy;
            if (y < 0 || y >= gr.g[1]) continue;
            // the three x-neighbours are consecutive cells: one contiguous range of the cell-ordered array
            const int x0 = max(cx - 1, 0), x1 = min(cx + 1, gr.g[0] - 1);
            const int row = (z * gr.g[1] + y) * gr.g[0];
            const int s = __ldg(cs + row + x0), e = __ldg(cs + row + x1 + 1);
            for (int base = s; base < e; base += 32) {
                const int j = base + lane;
                bool hit = false;
                int id = 0;
                if (j < e) {
                    const float4 c = __ldg(pts + j);
                    id = __float_as_int(c.w);
                    hit = id < T && apb_sqdist(qx - c.x, qy - c.y, qz - c.z) < radius2;
                }
                const unsigned mask = __ballot_sync(0xffffffffu, hit);
                if (hit) cur[cnt + __popc(mask & ((1u << lane) - 1))] = id;
                cnt += __popc(mask);
                if (cnt > CAP - 32) select();
            }
        }
    }
    select();
    int32_t *out = idx + ((size_t)b * M + q) * K;
    const int first = cnt > 0 ? cur[0] : 0;     // empty ball: the reference's pre-zeroed row
    for (int k = lane; k < K; k += 32) out[k] = k < cnt ? cur[k] : first;
}

}  // namespace

extern "C" long long apb_ball_query_grid_ws_bytes(int B, int N) {
    if (B < 0 || N < 0) return -1;
    // bbox | cell_start | cursor | cell_of | sorted (16-byte aligned blocks)
    long long bytes = 0;
    bytes += ((long long)B * 6 * 4 + 255) / 256 * 256;
    bytes += ((long long)B * (NCELL_MAX + 1) * 4 + 255) / 256 * 256;
    bytes += ((long long)B * NCELL_MAX * 4 + 255) / 256 * 256;
    bytes += ((long long)B * N * 4 + 255) / 256 * 256;
    bytes += (long long)B * N * 16;
    return bytes;
}

extern "C" int apb_ball_query_grid_f32(const float *new_xyz, const float *xyz, int32_t *idx, void *workspace,
                                       long long ws_bytes, int B, int N, int M, float radius, int nsample,
                                       apb_stream_t stream) {
    if (B < 0 || N < 0 || M < 0 || nsample <= 0) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!new_xyz || !idx || !xyz || !workspace || N == 0) return APB_E_INVALID;
    if (nsample > CAP - 32 || B > 65535) return APB_E_UNSUPPORTED;
    if (ws_bytes < apb_ball_query_grid_ws_bytes(B, N) || (reinterpret_cast<uintptr_t>(workspace) & 255) != 0)
        return APB_E_INVALID;
    cudaStream_t st = (cudaStream_t)stream;
    char *w = static_cast<char *>(workspace);
    auto take = [&](long long bytes) { char *p = w; w += (bytes + 255) / 256 * 256; return p; };
    float *bbox = reinterpret_cast<float *>(take((long long)B * 6 * 4));
    int *cell_start = reinterpret_cast<int *>(take((long long)B * (NCELL_MAX + 1) * 4));
    int *cursor = reinterpret_cast<int *>(take((long long)B * NCELL_MAX * 4));
    int *cell_of = reinterpret_cast<int *>(take((long long)B * N * 4));
    float4 *sorted = reinterpret_cast<float4 *>(take((long long)B * N * 16));
    bq_bbox_kernel<<<B, 256, 0, st>>>(xyz, bbox, cell_start, N);
    APB_CHECK_LAUNCH();
    bq_count_kernel<<<dim3(apb_div_up(N, 256), B), 256, 0, st>>>(xyz, bbox, cell_start, cell_of, N, radius);
    APB_CHECK_LAUNCH();
    bq_scan_kernel<<<B, 1024, 0, st>>>(cell_start, cursor);
    APB_CHECK_LAUNCH();
    bq_fill_kernel<<<dim3(apb_div_up(N, 256), B), 256, 0, st>>>(xyz, cell_of, cursor, sorted, N);
    APB_CHECK_LAUNCH();
    bq_grid_query_kernel<<<dim3(apb_div_up(M, QW), B), QW * 32, 0, st>>>(new_xyz, bbox, cell_start, sorted, idx, N, M,
                                                                        radius, nsample);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

// ---- the same query for callers that may not allocate (no counterpart in
// ---- the reference) ------------------------------------------------------------------------------------------------
// The cell-grid workspace lives in a file-scope device array instead of caller memory: nothing here (or below it) calls
// cudaMalloc / cudaFree or any allocating library routine, so the entry point can be used under CUDA graph capture, in
// serving loops and in harnesses that forbid allocation.  Every GPU that loads libapb200.so holds the array until exit.
// All calls on a device share it: it keeps the last call's contents (the build kernels re-initialise every field they
// read, as for caller workspaces with arbitrary contents), and two calls that overlap on the GPU (different streams
// without an ordering between them) RACE on it.
#define APB_BQ_STATIC_WS_BYTES (16ll << 20)
static __device__ __align__(256) unsigned char g_bq_static_ws[APB_BQ_STATIC_WS_BYTES];

namespace {
constexpr int BQ_MAX_DEVICES = 64;
void *g_bq_static_ptr[BQ_MAX_DEVICES];   // device address of g_bq_static_ws per device, resolved on first use
}  // namespace

extern "C" long long apb_ball_query_grid_static_ws_bytes(void) { return APB_BQ_STATIC_WS_BYTES; }

extern "C" int apb_ball_query_grid_static_f32(const float *new_xyz, const float *xyz, int32_t *idx, int B, int N, int M,
                                              float radius, int nsample, apb_stream_t stream) {
    if (B < 0 || N < 0 || M < 0 || nsample <= 0) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    const long long need = apb_ball_query_grid_ws_bytes(B, N);
    if (need < 0) return APB_E_INVALID;
    if (need > APB_BQ_STATIC_WS_BYTES) return APB_E_SCRATCH_LIMIT;
    int dev = -1;
    APB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= BQ_MAX_DEVICES) return APB_E_UNSUPPORTED;
    // In host code `g_bq_static_ws` is only a stand-in for the symbol: its device address on the CURRENT device comes
    // from the runtime (once per device; a benign race between host threads stores the same value twice).
    void *ws = __atomic_load_n(&g_bq_static_ptr[dev], __ATOMIC_ACQUIRE);
    if (ws == nullptr) {
        APB_CUDA(cudaGetSymbolAddress(&ws, g_bq_static_ws));
        if (ws == nullptr) return APB_E_INVALID;
        __atomic_store_n(&g_bq_static_ptr[dev], ws, __ATOMIC_RELEASE);
    }
    return apb_ball_query_grid_f32(new_xyz, xyz, idx, ws, APB_BQ_STATIC_WS_BYTES, B, N, M, radius, nsample, stream);
}
