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
