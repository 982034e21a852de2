// display-side reductions -- the normalisations and colour-map lookups that turn a depth /
// disparity map into the BGR image the reference shows or saves, written to a headless device buffer.
//   depth_map.py:936-937     ((depth_clipped - min) / (max - min) * 255).astype(u8) -> COLORMAP_TURBO
//   depth_map.py:1187-1195   same for the displayed depth (zeros where depth <= 0)
//   depth_map.py:1207-1209   cv2.normalize(disparity, 0, 255, NORM_MINMAX).astype(u8) -> COLORMAP_JET
//   fused_depth_map.py:1016, 1700, 3026-3028, 3052-3054   u8 / clip(x, 0, 255).astype(u8) -> COLORMAP_JET
//   fused_depth_map.py:1997-2004   legend range: min / max over the non-zero pixels of the fused map
#include "svb_common.cuh"
#include <float.h>
#include <math.h>
#include "colormap_tables.inc"

namespace {

__constant__ unsigned char c_lut[2][768];
bool g_lut_loaded[64] = {};

int load_luts()
{
    int dev = 0;
    SVB_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && g_lut_loaded[dev]) return SVB_OK;
    SVB_CUDA(cudaMemcpyToSymbol(c_lut, kColormapJET, 768, 0));
    SVB_CUDA(cudaMemcpyToSymbol(c_lut, kColormapTURBO, 768, 768));
    if (dev >= 0 && dev < 64) g_lut_loaded[dev] = true;
    return SVB_OK;
}

enum { SRC_U8 = 0, SRC_CLIP_F32, SRC_DISP_F32, SRC_DISP_F64, SRC_DEPTH_F32, SRC_DEPTH_F64, SRC_MINMAX_F32 };

struct DisplayArgs {
    const void* src;
    size_t n;                   // pixels per image
    double fxB, min_depth, max_depth;
    const unsigned* range;      // SRC_MINMAX_F32: ordered keys {min, max} per image (stride in words below)
    int range_stride;
    double lo, hi;              // SRC_MINMAX_F32 target range
    unsigned char* out_u8;      // nullable
    unsigned char* out_bgr;     // nullable
};

__device__ __forceinline__ float key_to_float(unsigned k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// NumPy's float -> uint8 astype on x86-64: truncate to int32, keep the low byte
__device__ __forceinline__ unsigned char cast_u8(float v) { return (unsigned char)(__float2int_rz(v) & 255); }
__device__ __forceinline__ unsigned char cast_u8(double v) { return (unsigned char)(__double2int_rz(v) & 255); }

template <int SRC>
__device__ __forceinline__ unsigned char display_index(const DisplayArgs& a, size_t i, float sc, float sh)
{
    if (SRC == SRC_U8) return ((const unsigned char*)a.src)[i];
    if (SRC == SRC_CLIP_F32) return cast_u8(fminf(fmaxf(((const float*)a.src)[i], 0.f), 255.f));
    if (SRC == SRC_MINMAX_F32) return cast_u8(__fmaf_rn(((const float*)a.src)[i], sc, sh));
    if (SRC == SRC_DISP_F32 || SRC == SRC_DEPTH_F32) {
        // float32 throughout (NEP 50: the Python scalars stay weak)
        const float mn = (float)a.min_depth, mx = (float)a.max_depth, len = (float)(a.max_depth - a.min_depth);
        float z = ((const float*)a.src)[i];
        if (SRC == SRC_DISP_F32) z = fminf(fmaxf(__fdiv_rn((float)a.fxB, __fadd_rn(z, 1e-6f)), mn), mx);
        else z = z > 0.f ? fminf(fmaxf(z, mn), mx) : 0.f;
        return cast_u8(__fmul_rn(__fdiv_rn(__fsub_rn(z, mn), len), 255.f));
    }
    // float64: fx taken from a float64 camera matrix promotes the whole expression (depth_map.py:917, 926)
    const double mn = a.min_depth, mx = a.max_depth, len = a.max_depth - a.min_depth;
    double z;
    if (SRC == SRC_DISP_F64) z = fmin(fmax(__ddiv_rn(a.fxB, (double)__fadd_rn(((const float*)a.src)[i], 1e-6f)), mn), mx);
    else { z = ((const double*)a.src)[i]; z = z > 0.0 ? fmin(fmax(z, mn), mx) : 0.0; }
    return cast_u8(__dmul_rn(__ddiv_rn(__dsub_rn(z, mn), len), 255.0));
}

// 4 consecutive pixels per thread: one 32-bit index store and three 32-bit BGR stores
template <int SRC>
__global__ void __launch_bounds__(256) k_display(DisplayArgs a, int cmap)
{
    __shared__ unsigned char lut[768];
    for (int i = threadIdx.x; i < 768; i += blockDim.x) lut[i] = c_lut[cmap][i];
    __syncthreads();
    const int b = blockIdx.y;
    float sc = 0.f, sh = 0.f;
    if (SRC == SRC_MINMAX_F32) {
        // cv2.normalize(..., NORM_MINMAX) with a CV_32F result: scale and shift rounded to float32, dst = fma(src, scale, shift)
        const double smin = (double)key_to_float(a.range[b * a.range_stride]), smax = (double)key_to_float(a.range[b * a.range_stride + 1]);
        const double scale = (a.hi - a.lo) * (smax - smin > DBL_EPSILON ? 1.0 / (smax - smin) : 0.0);
        sc = (float)scale;
        sh = (float)a.lo - (float)(smin * (double)sc);
    }
    const size_t base = (size_t)b * a.n;
    const size_t quads = (a.n + 3) / 4;
    const bool aligned = (a.n & 3) == 0;
    for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < quads; q += (size_t)gridDim.x * blockDim.x) {
        const size_t i0 = q * 4;
        unsigned char v[4];
        const int cnt = (int)min((size_t)4, a.n - i0);
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = k < cnt ? display_index<SRC>(a, base + i0 + k, sc, sh) : 0;
        if (a.out_u8) {
            if (aligned) *(uchar4*)(a.out_u8 + base + i0) = make_uchar4(v[0], v[1], v[2], v[3]);
            else for (int k = 0; k < cnt; ++k) a.out_u8[base + i0 + k] = v[k];
        }
        if (a.out_bgr) {
            unsigned char c[12];
#pragma unroll
            for (int k = 0; k < 4; ++k) { c[3 * k] = lut[3 * v[k]]; c[3 * k + 1] = lut[3 * v[k] + 1]; c[3 * k + 2] = lut[3 * v[k] + 2]; }
            unsigned char* o = a.out_bgr + (base + i0) * 3;
            if (aligned) {
                uint32_t* o32 = (uint32_t*)o;
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    o32[k] = c[4 * k] | (c[4 * k + 1] << 8) | (c[4 * k + 2] << 16) | ((uint32_t)c[4 * k + 3] << 24);
            } else {
                for (int k = 0; k < cnt * 3; ++k) o[k] = c[k];
            }
        }
    }
}

__global__ void k_range_init(unsigned* r, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) r[i] = (i & 1) ? 0u : 0xffffffffu;
}

__device__ __forceinline__ unsigned float_to_key(float f)
{
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void k_range_f32(const float* __restrict__ src, size_t n, unsigned* __restrict__ range)
{
    const int b = blockIdx.y;
    const float* p = src + (size_t)b * n;
    unsigned kmin = 0xffffffffu, kmax = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const unsigned k = float_to_key(p[i]);
        kmin = min(kmin, k);
        kmax = max(kmax, k);
    }
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    if ((threadIdx.x & 31) == 0 && kmax) {
        atomicMin(&range[2 * b], kmin);
        atomicMax(&range[2 * b + 1], kmax);
    }
}

// min / max over the non-zero pixels; {0, 0} when there are none
__global__ void k_range_nonzero_u8(const unsigned char* __restrict__ src, size_t n, unsigned* __restrict__ range)
{
    const int b = blockIdx.y;
    const unsigned char* p = src + (size_t)b * n;
    unsigned mn = 0xffffffffu, mx = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const unsigned v = p[i];
        if (v) { mn = min(mn, v); mx = max(mx, v); }
    }
    mn = __reduce_min_sync(0xffffffffu, mn);
    mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0 && mx) {
        atomicMin(&range[2 * b], mn);
        atomicMax(&range[2 * b + 1], mx);
    }
}

__global__ void k_range_export_u8(const unsigned* __restrict__ range, int B, int* __restrict__ out)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) {
        const bool any = range[2 * b + 1] != 0;
        out[2 * b] = any ? (int)range[2 * b] : 0;
        out[2 * b + 1] = any ? (int)range[2 * b + 1] : 0;
    }
}

int reduce_grid(int B, size_t n) { return (int)max((size_t)1, min((size_t)svb_ceil_div(148 * 8, B), (n + 1023) / 1024)); }

template <int SRC>
int launch_display(const DisplayArgs& a, int B, int cmap, cudaStream_t s)
{
    const size_t quads = (a.n + 3) / 4;
    const int nb = (int)max((size_t)1, min((size_t)svb_ceil_div(148 * 8, B), (quads + 255) / 256));
    k_display<SRC><<<dim3(nb, B), 256, 0, s>>>(a, cmap);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

int cmap_slot(int colormap)
{
    if (colormap == 2) return 0;       // cv2.COLORMAP_JET
    if (colormap == 20) return 1;      // cv2.COLORMAP_TURBO
    return -1;
}

}  // namespace

extern "C" size_t svb_display_scratch_bytes(int B) { return (size_t)B * 2 * sizeof(unsigned) + 256; }

extern "C" int svb_colormap_dev(const void* src_dev, int src_kind, int B, size_t npix, int colormap, const double* params5,
                                uint8_t* index_dev, uint8_t* bgr_dev, void* scratch_dev, size_t scratch_bytes, void* stream)
{
    if (!src_dev || B < 1 || npix < 1 || (!index_dev && !bgr_dev)) { svb_set_error("svb_colormap_dev: bad argument"); return SVB_ERR_PARAM; }
    const int slot = cmap_slot(colormap);
    if (slot < 0) { svb_set_error("svb_colormap_dev: colormap %d (only JET = 2 and TURBO = 20)", colormap); return SVB_ERR_UNSUPPORTED; }
    if (src_kind < SRC_U8 || src_kind > SRC_MINMAX_F32) { svb_set_error("svb_colormap_dev: source kind %d", src_kind); return SVB_ERR_PARAM; }
    if (src_kind >= SRC_DISP_F32 && !params5) { svb_set_error("svb_colormap_dev: parameters missing"); return SVB_ERR_PARAM; }
    int rc = load_luts();
    if (rc != SVB_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    DisplayArgs a = {};
    a.src = src_dev;
    a.n = npix;
    a.out_u8 = index_dev;
    a.out_bgr = bgr_dev;
    if (params5) { a.fxB = params5[0]; a.min_depth = params5[1]; a.max_depth = params5[2]; a.lo = params5[3]; a.hi = params5[4]; }
    switch (src_kind) {
    case SRC_U8: return launch_display<SRC_U8>(a, B, slot, s);
    case SRC_CLIP_F32: return launch_display<SRC_CLIP_F32>(a, B, slot, s);
    case SRC_DISP_F32: return launch_display<SRC_DISP_F32>(a, B, slot, s);
    case SRC_DISP_F64: return launch_display<SRC_DISP_F64>(a, B, slot, s);
    case SRC_DEPTH_F32: return launch_display<SRC_DEPTH_F32>(a, B, slot, s);
    case SRC_DEPTH_F64: return launch_display<SRC_DEPTH_F64>(a, B, slot, s);
    default: break;
    }
    if (!scratch_dev || scratch_bytes < svb_display_scratch_bytes(B)) { svb_set_error("svb_colormap_dev: scratch too small"); return SVB_ERR_PARAM; }
    unsigned* range = (unsigned*)(((uintptr_t)scratch_dev + 127) & ~(uintptr_t)127);
    k_range_init<<<svb_ceil_div(2 * B, 128), 128, 0, s>>>(range, 2 * B);
    k_range_f32<<<dim3(reduce_grid(B, npix), B), 256, 0, s>>>((const float*)src_dev, npix, range);
    a.range = range;
    a.range_stride = 2;
    return launch_display<SRC_MINMAX_F32>(a, B, slot, s);
}

extern "C" int svb_range_nonzero_u8_dev(const uint8_t* src_dev, int B, size_t npix, int* minmax_dev, void* scratch_dev,
                                        size_t scratch_bytes, void* stream)
{
    if (!src_dev || !minmax_dev || !scratch_dev || B < 1 || npix < 1) { svb_set_error("svb_range_nonzero_u8_dev: bad argument"); return SVB_ERR_PARAM; }
    if (scratch_bytes < svb_display_scratch_bytes(B)) { svb_set_error("svb_range_nonzero_u8_dev: scratch too small"); return SVB_ERR_PARAM; }
    cudaStream_t s = (cudaStream_t)stream;
    unsigned* range = (unsigned*)(((uintptr_t)scratch_dev + 127) & ~(uintptr_t)127);
    k_range_init<<<svb_ceil_div(2 * B, 128), 128, 0, s>>>(range, 2 * B);
    k_range_nonzero_u8<<<dim3(reduce_grid(B, npix), B), 256, 0, s>>>(src_dev, npix, range);
    k_range_export_u8<<<svb_ceil_div(B, 128), 128, 0, s>>>(range, B, minmax_dev);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
