// Shared device/host helpers for libpyrecover_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/pyrecover_b200.h"

// ------------------------------------------------------------------ error plumbing ----
void pr_set_error(const char *fmt, ...);   // api.cu

#define PR_CUDA(call)                                                                     \
    do {                                                                                  \
        cudaError_t _e = (call);                                                          \
        if (_e != cudaSuccess) {                                                          \
            pr_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,             \
                         cudaGetErrorString(_e));                                         \
            return PR_ERR_CUDA;                                                           \
        }                                                                                 \
    } while (0)

#define PR_REQUIRE(cond, ...)                                                             \
    do {                                                                                  \
        if (!(cond)) {                                                                    \
            pr_set_error(__VA_ARGS__);                                                    \
            return PR_ERR_INVALID;                                                        \
        }                                                                                 \
    } while (0)

int pr_sm_count();   // api.cu: SM count of the current device (cached per device)
void pr_count_launch(int n = 1);   // api.cu: bumps the counter pr_launch_count() reports

static inline bool pr_aligned16(const void *p) { return (((uintptr_t)p) & 15u) == 0; }

#ifdef __CUDACC__
// ------------------------------------------------------------------ bf16 <-> fp32 -----
// Bit-level conversions, identical to oracle/pr_oracle.c (RNE, quiet NaN).
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

__device__ __forceinline__ uint32_t f2bf_bits(float f) {
    uint32_t u = __float_as_uint(f);
    if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0u;
    u += 0x7fffu + ((u >> 16) & 1u);
    return u >> 16;
}
__device__ __forceinline__ uint32_t pack_bf2(float lo, float hi) {
    // cvt.rn.bf16x2.f32 d, a, b : a -> upper half, b -> lower half.  Round-to-nearest-even,
    // bit-identical to f2bf_bits for every non-NaN input; a NaN becomes 0x7fff here and 0x7fc0
    // in the oracle (both quiet NaNs -- the tests compare NaN-ness, not payloads).
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ float round_bf(float f) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(f), "f"(f));
    return __uint_as_float(r & 0xffff0000u);
}

// ------------------------------------------------------------------ 16-byte ld/st -----
// Streaming accesses: every byte on this path is touched once per launch, so keep it out
// of L1 and mark it evict-first in L2.
__device__ __forceinline__ uint4 ld_stream(const void *p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ uint4 ld_stream_nc(const void *p) {   // read-only for the whole launch
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(void *p, const uint4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ------------------------------------------------------------------ system-scope flags
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
#endif  // __CUDACC__
