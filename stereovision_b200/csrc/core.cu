// core.cu -- error state, device selection.
#include "svb_common.cuh"
#include <string.h>

static thread_local char g_err[512] = "";

void svb_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* svb_last_error(void) { return g_err; }
extern "C" int svb_version(void) { return 100; }

extern "C" int svb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int svb_init(int device)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        svb_set_error("no CUDA device available (%s); libsvb has no CPU fallback",
                      e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
        cudaGetLastError();
        return SVB_ERR_CUDA;
    }
    if (device < 0 || device >= n) { svb_set_error("device %d out of range [0,%d)", device, n); return SVB_ERR_PARAM; }
    SVB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    SVB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        svb_set_error("device %d is sm_%d%d; libsvb is built for sm_100a only", device, prop.major, prop.minor);
        return SVB_ERR_CUDA;
    }
    SVB_CUDA(cudaFree(0));
    return SVB_OK;
}

// ---- host thread pool for pageable <-> pinned copies
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>
#include <sched.h>
#include <unistd.h>
namespace {
struct SvbPool {
    std::mutex m;
    std::condition_variable cv;
    std::deque<std::pair<SvbTaskGroup*, std::function<void()>>> q;
    std::vector<std::thread> th;
    SvbPool()
    {
        cpu_set_t set;
        int n = sched_getaffinity(0, sizeof(set), &set) == 0 ? CPU_COUNT(&set) : (int)std::thread::hardware_concurrency();
        n = n / 4 < 2 ? 2 : (n / 4 > 6 ? 6 : n / 4);         // a quarter of the cores this process may use, 2..6
        for (int i = 0; i < n; ++i) th.emplace_back([this] { run(); });
    }
    void run()
    {
        for (;;) {
            std::unique_lock<std::mutex> l(m);
            cv.wait(l, [&] { return !q.empty(); });
            auto t = std::move(q.front());
            q.pop_front();
            l.unlock();
            t.second();
            t.first->pending.fetch_sub(1, std::memory_order_release);
        }
    }
};
// created at the first pageable call and never destroyed: its threads sleep on the condition variable until the
// process ends (joining them from a static destructor would race with interpreter shutdown)
// (a forked child inherits the object but not its threads: it gets a pool of its own)
SvbPool& svb_pool()
{
    static std::mutex mk;
    static SvbPool* p = nullptr;
    static pid_t owner = 0;
    std::lock_guard<std::mutex> l(mk);
    if (!p || owner != getpid()) { p = new SvbPool(); owner = getpid(); }
    return *p;
}
}  // namespace

void svb_pool_submit(SvbTaskGroup* g, std::function<void()> fn)
{
    SvbPool& p = svb_pool();
    g->pending.fetch_add(1, std::memory_order_relaxed);
    {
        std::lock_guard<std::mutex> l(p.m);
        p.q.emplace_back(g, std::move(fn));
    }
    p.cv.notify_one();
}
void svb_pool_wait(SvbTaskGroup* g)
{
    while (g->pending.load(std::memory_order_acquire) > 0) std::this_thread::yield();
}
int svb_pool_threads() { return (int)svb_pool().th.size(); }
bool svb_host_pageable(const void* ptr)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, ptr) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

// ---- pageable host arrays <-> device through two pinned chunks
namespace {
struct SvbCopyStage {
    std::mutex m;
    static constexpr size_t CH = (size_t)8 << 20;
    char* pin[2] = {nullptr, nullptr};
    cudaEvent_t dev_ev[16][2] = {};                          // events belong to a device: one pair per device
    cudaEvent_t* ev = nullptr;                               // the current device's pair
    cudaEvent_t busy[2] = {nullptr, nullptr};                // last event recorded after a DMA that READS pin[i]
    int ensure()
    {
        int dev = 0;
        SVB_CUDA(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 16) { svb_set_error("device %d out of range", dev); return SVB_ERR_PARAM; }
        for (int i = 0; i < 2; ++i) {
            if (!pin[i]) SVB_CUDA(cudaHostAlloc((void**)&pin[i], CH, cudaHostAllocPortable));
            if (!dev_ev[dev][i]) SVB_CUDA(cudaEventCreateWithFlags(&dev_ev[dev][i], cudaEventDisableTiming));
        }
        ev = dev_ev[dev];
        return SVB_OK;
    }
};
SvbCopyStage& svb_copy_stage() { static SvbCopyStage* s = new SvbCopyStage(); return *s; }
void svb_parallel_memcpy(char* dst, const char* src, size_t bytes)
{
    const int parts = svb_pool_threads();
    const size_t per = ((bytes + parts - 1) / parts + 4095) & ~(size_t)4095;
    SvbTaskGroup g;
    for (size_t o = per; o < bytes; o += per)
        svb_pool_submit(&g, [=] { memcpy(dst + o, src + o, bytes - o < per ? bytes - o : per); });
    memcpy(dst, src, bytes < per ? bytes : per);            // the calling thread takes the first part
    svb_pool_wait(&g);
}
}  // namespace

extern "C" int svb_copy_h2d(void* dst_dev, const void* src_host, size_t bytes, void* stream)
{
    if (!bytes) return SVB_OK;
    if (!dst_dev || !src_host) { svb_set_error("copy_h2d: null argument"); return SVB_ERR_PARAM; }
    cudaStream_t st = (cudaStream_t)stream;
    if (!svb_host_pageable(src_host)) {
        SVB_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, st));
        return SVB_OK;
    }
    SvbCopyStage& s = svb_copy_stage();
    std::lock_guard<std::mutex> l(s.m);
    int rc = s.ensure();
    if (rc != SVB_OK) return rc;
    int k = 0;
    for (size_t o = 0; o < bytes; o += SvbCopyStage::CH, k ^= 1) {
        const size_t len = bytes - o < SvbCopyStage::CH ? bytes - o : SvbCopyStage::CH;
        if (s.busy[k]) SVB_CUDA(cudaEventSynchronize(s.busy[k]));    // the chunk that last left this buffer
        svb_parallel_memcpy(s.pin[k], (const char*)src_host + o, len);
        SVB_CUDA(cudaMemcpyAsync((char*)dst_dev + o, s.pin[k], len, cudaMemcpyHostToDevice, st));
        SVB_CUDA(cudaEventRecord(s.ev[k], st));
        s.busy[k] = s.ev[k];
    }
    return SVB_OK;
}

extern "C" int svb_copy_d2h(void* dst_host, const void* src_dev, size_t bytes, void* stream)
{
    if (!bytes) return SVB_OK;
    if (!dst_host || !src_dev) { svb_set_error("copy_d2h: null argument"); return SVB_ERR_PARAM; }
    cudaStream_t st = (cudaStream_t)stream;
    if (!svb_host_pageable(dst_host)) {
        SVB_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, st));
        SVB_CUDA(cudaStreamSynchronize(st));
        return SVB_OK;
    }
    SvbCopyStage& s = svb_copy_stage();
    std::lock_guard<std::mutex> l(s.m);
    int rc = s.ensure();
    if (rc != SVB_OK) return rc;
    for (int i = 0; i < 2; ++i)
        if (s.busy[i]) { SVB_CUDA(cudaEventSynchronize(s.busy[i])); s.busy[i] = nullptr; }   // uploads still reading the buffers
    int k = 0;
    size_t prev_o = 0, prev_len = 0;
    for (size_t o = 0; o < bytes; o += SvbCopyStage::CH, k ^= 1) {
        const size_t len = bytes - o < SvbCopyStage::CH ? bytes - o : SvbCopyStage::CH;
        SVB_CUDA(cudaMemcpyAsync(s.pin[k], (const char*)src_dev + o, len, cudaMemcpyDeviceToHost, st));
        SVB_CUDA(cudaEventRecord(s.ev[k], st));
        if (prev_len) {                                       // the previous chunk leaves its buffer while this one arrives
            SVB_CUDA(cudaEventSynchronize(s.ev[k ^ 1]));
            svb_parallel_memcpy((char*)dst_host + prev_o, s.pin[k ^ 1], prev_len);
        }
        prev_o = o; prev_len = len;
    }
    SVB_CUDA(cudaEventSynchronize(s.ev[k ^ 1]));
    svb_parallel_memcpy((char*)dst_host + prev_o, s.pin[k ^ 1], prev_len);
    return SVB_OK;
}

#include <cuda.h>
typedef CUresult (*svb_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int svb_make_tmap_2d_u32(void* tmap_out, const void* base, uint64_t inner, uint64_t rows, uint64_t pitch_bytes,
                         uint32_t box_inner, uint32_t box_rows)
{
    static svb_encode_tiled_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        SVB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (!p || q != cudaDriverEntryPointSuccess) { svb_set_error("cuTensorMapEncodeTiled is not available"); return SVB_ERR_CUDA; }
        fn = (svb_encode_tiled_fn)p;
    }
    const cuuint64_t dims[2] = {inner, rows};
    const cuuint64_t strides[1] = {pitch_bytes};
    const cuuint32_t box[2] = {box_inner, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = fn((CUtensorMap*)tmap_out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        svb_set_error("cuTensorMapEncodeTiled failed (%d): inner %llu rows %llu pitch %llu box %ux%u", (int)r,
                      (unsigned long long)inner, (unsigned long long)rows, (unsigned long long)pitch_bytes, box_inner, box_rows);
        return SVB_ERR_CUDA;
    }
    return SVB_OK;
}
