// grouping_operation on bf16 features (SURVEY.md 8b `apb_group_bf16`; the reference kernel it mirrors is
// openpoints/cpp/pointnet2_batch/src/group_points_gpu.cu:53-72, fp32 only):
//     out[b, c, j, s] = feat[b, c, idx[b, j, s]]        feat (B,C,N) bf16, idx (B,M,K) int32, out (B,C,M,K) bf16
// Pure byte movement, HBM-bound: algorithmic bytes 2*B*C*N + 4*B*M*K + 2*B*C*M*K -- half the fp32 kernel's on the feature
// side.  A CTA stages CH feature rows in shared memory (2 bytes per point: rows of up to 100 k points fit, twice the fp32
// limit), reads eight idx entries as two int4 once for all CH channels, gathers from shared memory and emits ONE 16-byte
// store of eight bf16 per channel (the "bf16x8" access of BASELINE.json's north_star).
#include "apb_common.cuh"

namespace {

constexpr int GB_NT = 512;

template <int CH>
__global__ void __launch_bounds__(GB_NT) group_bf16_rows_kernel(const uint16_t *__restrict__ feat,
                                                                const int32_t *__restrict__ idx,
                                                                uint16_t *__restrict__ out, int C, int N, long long P,
                                                                int chunk) {
    extern __shared__ __align__(16) uint16_t rows[];   // [CH][N]
    const int b = blockIdx.z, c0 = blockIdx.y * CH;
    const int nch = min(CH, C - c0);
    const uint16_t *src = feat + ((size_t)b * C + c0) * N;
    const size_t total = (size_t)nch * N;
    if (((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((total & 7) == 0)) {
        const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
        uint4 *d4 = reinterpret_cast<uint4 *>(rows);
        for (size_t i = threadIdx.x; i < total / 8; i += GB_NT) d4[i] = __ldg(s4 + i);
    } else {
        for (size_t i = threadIdx.x; i < total; i += GB_NT) rows[i] = __ldg(src + i);
    }
    __syncthreads();
    const int32_t *ix = idx + (size_t)b * P;
    uint16_t *o = out + ((size_t)b * C + c0) * P;
    const long long p_begin = (long long)blockIdx.x * chunk, p_end = min(P, p_begin + chunk);
    const bool vec_ok = ((P & 7) == 0) && ((reinterpret_cast<uintptr_t>(ix) & 15) == 0) &&
                        ((reinterpret_cast<uintptr_t>(o) & 15) == 0);
    if (vec_ok) {   // chunk is a multiple of 8
        for (long long p = p_begin + (long long)threadIdx.x * 8; p < p_end; p += (long long)GB_NT * 8) {
            const int4 a = apb_ld_nc_i4(ix + p), c = apb_ld_nc_i4(ix + p + 4);
#pragma unroll
            for (int ch = 0; ch < CH; ++ch) {
                if (ch < nch) {
                    const uint16_t *r = rows + (size_t)ch * N;
                    uint4 v;
                    v.x = (uint32_t)r[a.x] | ((uint32_t)r[a.y] << 16);
                    v.y = (uint32_t)r[a.z] | ((uint32_t)r[a.w] << 16);
                    v.z = (uint32_t)r[c.x] | ((uint32_t)r[c.y] << 16);
                    v.w = (uint32_t)r[c.z] | ((uint32_t)r[c.w] << 16);
                    asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(o + (size_t)ch * P + p), "r"(v.x),
                                 "r"(v.y), "r"(v.z), "r"(v.w)
                                 : "memory");
                }
            }
        }
    } else {
        for (long long p = p_begin + threadIdx.x; p < p_end; p += GB_NT) {
            const int i = __ldg(ix + p);
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (ch < nch) o[(size_t)ch * P + p] = rows[(size_t)ch * N + i];
        }
    }
}

// rows too long for shared memory: gather straight from L2
__global__ void __launch_bounds__(256) group_bf16_direct_kernel(const uint16_t *__restrict__ feat,
                                                                const int32_t *__restrict__ idx,
                                                                uint16_t *__restrict__ out, int C, int N, long long P) {
    const int c = blockIdx.y, b = blockIdx.z;
    const uint16_t *r = feat + ((size_t)b * C + c) * N;
    const int32_t *ix = idx + (size_t)b * P;
    uint16_t *o = out + ((size_t)b * C + c) * P;
    for (long long p = (long long)blockIdx.x * 256 + threadIdx.x; p < P; p += (long long)gridDim.x * 256)
        o[p] = __ldg(r + __ldg(ix + p));
}

template <int CH>
int launch_rows(const uint16_t *feat, const int32_t *idx, uint16_t *out, int B, int C, int N, long long P,
                cudaStream_t st) {
    const size_t smem = (size_t)CH * N * sizeof(uint16_t);
    APB_CUDA(cudaFuncSetAttribute((const void *)group_bf16_rows_kernel<CH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
    long long chunk = (long long)8 * N;
    if (chunk < 4096) chunk = 4096;
    const int cgroups = apb_div_up(C, CH);
    while (chunk > 4096 && (long long)apb_div_up(P, chunk) * cgroups * B < 4 * APB_NUM_SMS) chunk >>= 1;
    chunk = (chunk + 4095) / 4096 * 4096;
    group_bf16_rows_kernel<CH><<<dim3(apb_div_up(P, chunk), cgroups, B), GB_NT, smem, st>>>(feat, idx, out, C, N, P,
                                                                                           (int)chunk);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

}  // namespace

extern "C" int apb_group_bf16(const uint16_t *features, const int32_t *idx, uint16_t *out, int B, int C, int N, int M,
                              int K, apb_stream_t stream) {
    if (B < 0 || C < 0 || N <= 0 || M < 0 || K < 0) return APB_E_INVALID;
    const long long P = (long long)M * K;
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!features || !idx || !out) return APB_E_INVALID;
    if (C > 65535 || B > 65535) return APB_E_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)N * sizeof(uint16_t), budget = 200 * 1024;
    if (row > budget) {
        dim3 grid((unsigned)min((long long)apb_div_up(P, 256), (long long)8 * APB_NUM_SMS), C, B);
        group_bf16_direct_kernel<<<grid, 256, 0, st>>>(features, idx, out, C, N, P);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    if (C >= 4 && 4 * row <= budget) return launch_rows<4>(features, idx, out, B, C, N, P, st);
    if (C >= 2 && 2 * row <= budget) return launch_rows<2>(features, idx, out, B, C, N, P, st);
    return launch_rows<1>(features, idx, out, B, C, N, P, st);
}
