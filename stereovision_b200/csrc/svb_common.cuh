// Shared host/device helpers for libsvb.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/svb.h"

void svb_set_error(const char* fmt, ...);

#define SVB_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            svb_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return SVB_ERR_CUDA;                                                         \
        }                                                                                \
    } while (0)

#define SVB_LAUNCH_CHECK()                                                               \
    do {                                                                                 \
        cudaError_t e__ = cudaGetLastError();                                            \
        if (e__ != cudaSuccess) {                                                        \
            svb_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
            return SVB_ERR_CUDA;                                                         \
        }                                                                                \
    } while (0)

static inline int svb_ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// Grow-only device buffer owned by a handle.
struct SvbBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) return SVB_OK;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) {
            svb_set_error("cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
            cudaGetLastError();
            return e == cudaErrorMemoryAllocation ? SVB_ERR_NOMEM : SVB_ERR_CUDA;
        }
        cap = bytes;
        return SVB_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

#ifdef __CUDACC__
// ---- packed u16x2 helpers (map 1:1 to VIADD.16x2 / VIMNMX.U16x2 / VIADDMNMX.U16x2 on sm_100a)
__device__ __forceinline__ uint32_t pk_min(uint32_t a, uint32_t b) { return __vminu2(a, b); }
__device__ __forceinline__ uint32_t pk_addmin(uint32_t a, uint32_t b, uint32_t c) { return __viaddmin_u16x2(a, b, c); }
__device__ __forceinline__ uint32_t pk_add(uint32_t a, uint32_t b) { return __vadd2(a, b); }
__device__ __forceinline__ uint32_t pk_dup(uint32_t v) { return (v & 0xffffu) * 0x10001u; }
__device__ __forceinline__ uint32_t pk_hmin(uint32_t a) { return min(a & 0xffffu, a >> 16); }
// (hi16 of a -> lo, lo16 of b -> hi)
__device__ __forceinline__ uint32_t pk_shift_in(uint32_t a, uint32_t b) { return __byte_perm(a, b, 0x5432); }
#endif
