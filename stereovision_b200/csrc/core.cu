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
