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

// Grow-only pinned host buffer owned by a handle (staging for pageable caller buffers).
struct SvbPinBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) return SVB_OK;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocDefault);
        if (e != cudaSuccess) {
            svb_set_error("cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
            cudaGetLastError();
            return SVB_ERR_NOMEM;
        }
        cap = bytes;
        return SVB_OK;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

// A few host threads that move data between the caller's PAGEABLE arrays and pinned staging (core.cu).  The reference's
// call site hands NumPy arrays to compute(); a cudaMemcpy from pageable memory is a single-threaded staged copy inside
// the driver (~9 GB/s here) that blocks the calling thread, which is what `e2e_pageable` used to pay.
#include <atomic>
#include <functional>
struct SvbTaskGroup { std::atomic<int> pending{0}; };
void svb_pool_submit(SvbTaskGroup* g, std::function<void()> fn);
void svb_pool_wait(SvbTaskGroup* g);                     // returns when every task submitted to g has run
int svb_pool_threads();
bool svb_host_pageable(const void* p);                   // not pinned / registered / device memory

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

#ifdef __CUDACC__
// ---- TMA (cp.async.bulk.tensor) + mbarrier helpers, sm_100a
#include <cuda.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_inval(uint64_t* bar)
{
    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
// 2-D tile global -> shared, completion counted in bytes on `bar` (x = inner coordinate)
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tmap, int x, int y, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
}
#endif

// Host: encode a 2-D u32 tensor map (row pitch in bytes must be a multiple of 16); the driver
// entry point is resolved through the runtime, libsvb does not link libcuda.
int svb_make_tmap_2d_u32(void* tmap_out /* CUtensorMap */, const void* base, uint64_t inner, uint64_t rows,
                         uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_rows);
