// Shared device/host helpers for libapb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/apb200.h"

#define APB_NUM_SMS 148  // B200: 2 dies x 74 SMs

#define APB_CHECK_LAUNCH()                         \
    do {                                           \
        cudaError_t e__ = cudaGetLastError();      \
        if (e__ != cudaSuccess) return (int)e__;   \
    } while (0)

#define APB_CUDA(call)                             \
    do {                                           \
        cudaError_t e__ = (call);                  \
        if (e__ != cudaSuccess) return (int)e__;   \
    } while (0)

// Squared distance in the reference's (nvcc-contracted) operation order:
//   t = dy*dy ; t = fma(dx,dx,t) ; t = fma(dz,dz,t)
// (sampling_gpu.cu:140, ball_query_gpu.cu:40, interpolate_gpu.cu:42 as compiled with -O2 / --fmad=true;
// verified in the reference's sm_100a SASS).  Explicit intrinsics so no compiler flag can change it.
__device__ __forceinline__ float apb_sqdist(float dx, float dy, float dz) {
    float t = __fmul_rn(dy, dy);
    t = __fmaf_rn(dx, dx, t);
    t = __fmaf_rn(dz, dz, t);
    return t;
}

static inline int apb_div_up(long long a, long long b) { return (int)((a + b - 1) / b); }

// 128-bit streaming store (output written once, never re-read by this kernel).
__device__ __forceinline__ void apb_st_stream_f4(float *p, float4 v) {
    asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ int4 apb_ld_nc_i4(const int *p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ float4 apb_ld_nc_f4(const float *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
