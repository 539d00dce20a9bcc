// One-shot all-reduce of SMALL buffers over NVLink/NVSwitch peer memory (SyncBN statistics: 2*C values).
//
// The reference synchronises BatchNorm statistics with torch's SyncBatchNorm (main.py:27 turns it on whenever
// world_size > 1), i.e. one NCCL all-gather/all-reduce of a few hundred values per BN layer per direction: ~75 tiny
// collectives per AdaptPoint iteration whose cost is pure launch + protocol latency.  Here every rank owns a
// workspace that all peers have mapped (CUDA IPC); an all-reduce is ONE kernel of one CTA:
//   1. store my values into slot [parity][my rank] of EVERY rank's workspace (P2P stores through NVLink),
//   2. __threadfence_system, then publish flag [parity][my rank] = seq+1 on every rank (st.release.sys),
//   3. wait until my own workspace holds flag seq+1 from every rank (the spin reads LOCAL memory only),
//   4. sum the world slots in rank order (identical bits on every rank) back into the buffer.
// `seq` lives in the workspace and is advanced by the kernel itself, so the launch is replayable inside a CUDA graph.
// Two parities suffice: a rank can run at most one all-reduce ahead of its slowest peer (it needs that peer's flag
// for call k+1, which the peer publishes only after it has finished reading call k).
// A peer that died would leave the others spinning: the wait traps after ~60 s so the box is never hung.
#include "apb_common.cuh"

namespace {

constexpr int kMaxWorld = 8;
constexpr int kMaxElems = 8192;      // 8-byte words per slot
constexpr int kFlagStride = 32;      // uint32 words between flags (128 B apart)

struct PeerWs {                      // layout of one rank's workspace
    unsigned long long seq;
    unsigned int pad[30];
    unsigned int flags[2][kMaxWorld][kFlagStride];
    double data[2][kMaxWorld][kMaxElems];
};

struct PeerPtrs {
    PeerWs *ws[kMaxWorld];
};

__device__ __forceinline__ void st_release_sys(unsigned int *p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <typename T>
__global__ void __launch_bounds__(256) peer_allreduce_kernel(T *__restrict__ buf, int n, PeerPtrs peers, int world,
                                                             int rank) {
    PeerWs *mine = peers.ws[rank];
    const unsigned long long seq = *reinterpret_cast<volatile unsigned long long *>(&mine->seq);
    const int par = (int)(seq & 1);
    const unsigned int want = (unsigned int)(seq + 1);
    // 1. my contribution into every rank's slot [par][rank]
    for (int i = threadIdx.x; i < n * world; i += blockDim.x) {
        const int r = i / n, k = i - r * n;
        reinterpret_cast<T *>(peers.ws[r]->data[par][rank])[k] = buf[k];
    }
    __threadfence_system();
    __syncthreads();
    // 2. publish, 3. wait for everyone (local spin)
    if (threadIdx.x < world) {
        st_release_sys(&peers.ws[threadIdx.x]->flags[par][rank][0], want);
        const long long t0 = clock64();
        while ((int)(ld_acquire_sys(&mine->flags[par][threadIdx.x][0]) - want) < 0) {
            if (clock64() - t0 > 120000000000LL) __trap();  // ~60 s
        }
    }
    __syncthreads();
    // 4. rank-ordered sum (reads bypass L1: the slots were written by other GPUs)
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
        T s = __ldcv(reinterpret_cast<const T *>(mine->data[par][0]) + k);
        for (int r = 1; r < world; ++r) s += __ldcv(reinterpret_cast<const T *>(mine->data[par][r]) + k);
        buf[k] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) *reinterpret_cast<volatile unsigned long long *>(&mine->seq) = seq + 1;
}

}  // namespace

extern "C" int apb_peer_ws_bytes(void) { return (int)sizeof(PeerWs); }
extern "C" int apb_peer_max_elems(void) { return kMaxElems; }

// Allocates this rank's workspace (zeroed) and returns its device pointer and a 64-byte CUDA IPC handle.
extern "C" int apb_peer_ws_create(void **ws_out, unsigned char *handle_out64) {
    if (!ws_out || !handle_out64) return APB_E_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void *p = nullptr;
    APB_CUDA(cudaMalloc(&p, sizeof(PeerWs)));
    APB_CUDA(cudaMemset(p, 0, sizeof(PeerWs)));
    APB_CUDA(cudaDeviceSynchronize());
    cudaIpcMemHandle_t h;
    APB_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(handle_out64, &h, 64);
    *ws_out = p;
    return APB_OK;
}

// Maps a peer's workspace into this process (enables peer access lazily).
extern "C" int apb_peer_ws_open(const unsigned char *handle64, void **ws_out) {
    if (!handle64 || !ws_out) return APB_E_INVALID;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    APB_CUDA(cudaIpcOpenMemHandle(ws_out, h, cudaIpcMemLazyEnablePeerAccess));
    return APB_OK;
}

extern "C" int apb_peer_ws_close(void *ws, int own) {
    if (!ws) return APB_OK;
    if (own) APB_CUDA(cudaFree(ws));
    else APB_CUDA(cudaIpcCloseMemHandle(ws));
    return APB_OK;
}

// In-place sum over ranks of buf[0..n) (dtype 0 = f32, 1 = f64); ws[r] = rank r's workspace as mapped HERE.
extern "C" int apb_peer_allreduce(void *buf, int n, int dtype, void *const *ws, int world, int rank,
                                  apb_stream_t stream) {
    if (n == 0) return APB_OK;
    if (!buf || !ws || n < 0 || world < 1 || world > kMaxWorld || rank < 0 || rank >= world) return APB_E_INVALID;
    if (n > kMaxElems) return APB_E_UNSUPPORTED;
    PeerPtrs p{};
    for (int r = 0; r < world; ++r) {
        if (!ws[r]) return APB_E_INVALID;
        p.ws[r] = static_cast<PeerWs *>(ws[r]);
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == 1) peer_allreduce_kernel<double><<<1, 256, 0, st>>>(static_cast<double *>(buf), n, p, world, rank);
    else if (dtype == 0) peer_allreduce_kernel<float><<<1, 256, 0, st>>>(static_cast<float *>(buf), n, p, world, rank);
    else return APB_E_INVALID;
    APB_CHECK_LAUNCH();
    return APB_OK;
}
