// imgproc.cu -- the OpenCV imgproc/calib3d helpers on the reference's per-frame path,
// as sm_100a kernels (bit-exact for integer outputs):
//   remap (+ fused BGR2GRAY)   depth_map.py:815-826, 871-879; fused_depth_map.py:480-491, 979-984
//   medianBlur 3 (s16) / 5 (f32)            inside StereoSGBM; depth_map.py:912
//   filterSpeckles (connected components)   inside StereoSGBM
//   disparity -> depth                      depth_map.py:909-933
//   reprojectImageTo3D                      BASELINE.json (Q from depth_map.py:625)
#include "svb_common.cuh"
#include <mutex>
#include <float.h>
#include <algorithm>

// ---------------------------------------------------------------------------------------
// Exact medians by forgetful selection: keep ceil(N/2)+1 candidates, repeatedly drop the
// min and the max and take in the next sample.  Everything is unrolled so the candidates
// stay in registers.
// ---------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ void cex(T& a, T& b)
{
    T lo = a < b ? a : b, hi = a < b ? b : a;
    a = lo; b = hi;
}
template <> __device__ __forceinline__ void cex<float>(float& a, float& b)
{
    float lo = fminf(a, b), hi = fmaxf(a, b);
    a = lo; b = hi;
}
// min of v[0..N) -> v[0], max -> v[N-1]
template <int N, typename T> __device__ __forceinline__ void minmax(T (&v)[14])
{
#pragma unroll
    for (int i = 0; i < N / 2; ++i) cex(v[i], v[N - 1 - i]);
#pragma unroll
    for (int i = 1; i < (N + 1) / 2; ++i) cex(v[0], v[i]);
#pragma unroll
    for (int i = N / 2; i < N - 1; ++i) cex(v[i], v[N - 1]);
}
// 5x5 median on f32.  FROM_S16: the source is the s16 x16 disparity and is scaled by 1/16
// on load (exact), fusing depth_map.py:909 with :912.  DEPTH: also emit
// depth_map.py:926-933 (depth = fxB / (d + 1e-6), clip, validity mask) in the epilogue.
// DEPTH 2: the same in float64, which is what NumPy computes when fx comes from a float64 camera matrix (:917).
// Two vertically adjacent outputs per thread (rows 2t, 2t + 1): their windows share four rows = 20 samples.  A
// shared sample with 13 larger shared samples cannot be the median of either window (rank 13 of 25), nor one with 13
// smaller: the 7 smallest and 7 largest of the 20 are dropped ONCE (forgetful selection, 7 rounds), and each output
// is the 6th of the 6 survivors + its own 5 samples (5 rounds): ~96 compare-exchanges per output instead of ~150,
// 15 loads instead of 25.  (The kernel is bound by the selection network, not by memory.)  Exact: order statistics.
template <bool FROM_S16, int DEPTH>
__global__ void __launch_bounds__(256) k_median5_f32(const void* __restrict__ src_, int H, int W,
                                                     float* __restrict__ dst, void* __restrict__ depth_,
                                                     double fxB64, float min_disp, double min_depth64, double max_depth64)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y0 = 2 * blockIdx.y;
    if (x >= W) return;
    const size_t img = (size_t)blockIdx.z * H * W;
    int xx[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) xx[i] = min(max(x + i - 2, 0), W - 1);
    auto ld = [&](int row, int i) -> float {
        const size_t o = img + (size_t)min(max(row, 0), H - 1) * W + xx[i];
        return FROM_S16 ? (float)((const int16_t*)src_)[o] * 0.0625f : ((const float*)src_)[o];
    };
    float v[14];
    {
        float c[20];                                         // rows y0 - 1 .. y0 + 2 (clamped): in both windows
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int i = 0; i < 5; ++i) c[j * 5 + i] = ld(y0 - 1 + j, i);
#pragma unroll
        for (int i = 0; i < 14; ++i) v[i] = c[i];
        minmax<14>(v); v[0] = c[14];
        minmax<13>(v); v[0] = c[15];
        minmax<12>(v); v[0] = c[16];
        minmax<11>(v); v[0] = c[17];
        minmax<10>(v); v[0] = c[18];
        minmax<9>(v); v[0] = c[19];
        minmax<8>(v);                                        // survivors: v[1..6]
    }
    const float fxB = (float)fxB64, min_depth = (float)min_depth64, max_depth = (float)max_depth64;
#pragma unroll
    for (int o2 = 0; o2 < 2; ++o2) {
        const int y = y0 + o2;
        if (y >= H) break;
        float e[5], w[14];
#pragma unroll
        for (int i = 0; i < 5; ++i) e[i] = ld(o2 ? y0 + 3 : y0 - 2, i);
#pragma unroll
        for (int i = 0; i < 6; ++i) w[i] = v[1 + i];
        w[6] = e[0];
        minmax<7>(w); w[0] = e[1];
        minmax<6>(w); w[0] = e[2];
        minmax<5>(w); w[0] = e[3];
        minmax<4>(w); w[0] = e[4];
        minmax<3>(w);
        const float d = w[1];
        const size_t o = img + (size_t)y * W + x;
        dst[o] = d;
        if (DEPTH == 1) {
            // NumPy float32 semantics of depth_map.py:926-933
            const float z = __fdiv_rn(fxB, __fadd_rn(d, 1e-6f));
            const float zc = fminf(fmaxf(z, min_depth), max_depth);
            const bool valid = (d > min_disp) && (zc >= min_depth) && (zc <= max_depth);
            ((float*)depth_)[o] = valid ? zc : 0.f;
        }
        if (DEPTH == 2) {
            const double z = __ddiv_rn(fxB64, (double)__fadd_rn(d, 1e-6f));
            const double zc = fmin(fmax(z, min_depth64), max_depth64);
            ((double*)depth_)[o] = (d > min_disp) ? zc : 0.0;
        }
    }
}

// The same selection for the s16 disparity, TWO horizontally adjacent pixels per compare-exchange: the median commutes
// with the monotone map s16 -> float(s16) / 16 (depth_map.py:909-912 convert first, the result is the same float), so
// the order statistics are taken on packed s16x2 words (one VIMNMX.S16x2 per min / max of two pixels) and only the
// four medians of a thread's 2 x 2 output block are converted.  Halves the selection network's instructions per pixel.
struct s16x2 { uint32_t v; };
template <> __device__ __forceinline__ void cex<s16x2>(s16x2& a, s16x2& b)
{
    const uint32_t lo = __vmins2(a.v, b.v), hi = __vmaxs2(a.v, b.v);
    a.v = lo; b.v = hi;
}
template <int DEPTH>
__global__ void __launch_bounds__(256) k_median5_s16x2(const int16_t* __restrict__ src, int H, int W,
                                                       float* __restrict__ dst, void* __restrict__ depth_,
                                                       double fxB64, float min_disp, double min_depth64, double max_depth64)
{
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x), y0 = 2 * blockIdx.y;
    if (x >= W) return;
    const size_t img = (size_t)blockIdx.z * H * W;
    int xx[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) xx[i] = min(max(x + i - 2, 0), W - 1);
    // window row `row` of the pixel pair (x, x + 1): word i = (column x + i - 2, column x + i - 1), replicate border
    // (interior pairs of a 4-byte aligned row: three 32-bit loads and two permutes instead of six 16-bit loads)
    const bool interior = x >= 2 && x + 3 < W;
    auto ldrow = [&](int row, s16x2* o) {
        const size_t ro = img + (size_t)min(max(row, 0), H - 1) * W;
        const unsigned short* r = reinterpret_cast<const unsigned short*>(src) + ro;
        if (interior && !(reinterpret_cast<uintptr_t>(r + x) & 3)) {
            const uint32_t* q = reinterpret_cast<const uint32_t*>(r + x - 2);
            const uint32_t a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            o[0].v = a; o[1].v = __byte_perm(a, b, 0x5432); o[2].v = b; o[3].v = __byte_perm(b, c, 0x5432); o[4].v = c;
        } else {
            uint32_t p[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) p[i] = __ldg(r + xx[i]);
#pragma unroll
            for (int i = 0; i < 5; ++i) o[i].v = p[i] | (p[i + 1] << 16);
        }
    };
    s16x2 v[14];
    {
        s16x2 c[20];                                         // rows y0 - 1 .. y0 + 2: in both windows
#pragma unroll
        for (int j = 0; j < 4; ++j) ldrow(y0 - 1 + j, c + 5 * j);
#pragma unroll
        for (int i = 0; i < 14; ++i) v[i] = c[i];
        minmax<14>(v); v[0] = c[14];
        minmax<13>(v); v[0] = c[15];
        minmax<12>(v); v[0] = c[16];
        minmax<11>(v); v[0] = c[17];
        minmax<10>(v); v[0] = c[18];
        minmax<9>(v); v[0] = c[19];
        minmax<8>(v);                                        // survivors: v[1..6]
    }
    const float fxB = (float)fxB64, min_depth = (float)min_depth64, max_depth = (float)max_depth64;
#pragma unroll
    for (int o2 = 0; o2 < 2; ++o2) {
        const int y = y0 + o2;
        if (y >= H) break;
        s16x2 e[5], w[14];
        ldrow(o2 ? y0 + 3 : y0 - 2, e);
#pragma unroll
        for (int i = 0; i < 6; ++i) w[i] = v[1 + i];
        w[6] = e[0];
        minmax<7>(w); w[0] = e[1];
        minmax<6>(w); w[0] = e[2];
        minmax<5>(w); w[0] = e[3];
        minmax<4>(w); w[0] = e[4];
        minmax<3>(w);
        float dd[2], zf[2];
        double zd[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int d16 = h ? (int)w[1].v >> 16 : (int)(int16_t)(w[1].v & 0xffffu);
            const float d = (float)d16 * 0.0625f;            // exact
            dd[h] = d;
            if (DEPTH == 1) {
                // NumPy float32 semantics of depth_map.py:926-933
                const float z = __fdiv_rn(fxB, __fadd_rn(d, 1e-6f));
                const float zc = fminf(fmaxf(z, min_depth), max_depth);
                const bool valid = (d > min_disp) && (zc >= min_depth) && (zc <= max_depth);
                zf[h] = valid ? zc : 0.f;
            }
            if (DEPTH == 2) {
                const double z = __ddiv_rn(fxB64, (double)__fadd_rn(d, 1e-6f));
                const double zc = fmin(fmax(z, min_depth64), max_depth64);
                zd[h] = (d > min_disp) ? zc : 0.0;
            }
        }
        const size_t o = img + (size_t)y * W + x;
        const bool al = !(reinterpret_cast<uintptr_t>(dst + o) & 7) &&
                        (DEPTH == 0 || !(reinterpret_cast<uintptr_t>(depth_) + o * (DEPTH == 2 ? 8 : 4) & (DEPTH == 2 ? 15 : 7)));
        if (x + 1 < W && al) {                               // both pixels, 8- / 16-byte aligned: one store each
            *reinterpret_cast<float2*>(dst + o) = make_float2(dd[0], dd[1]);
            if (DEPTH == 1) *reinterpret_cast<float2*>((float*)depth_ + o) = make_float2(zf[0], zf[1]);
            if (DEPTH == 2) *reinterpret_cast<double2*>((double*)depth_ + o) = make_double2(zd[0], zd[1]);
        } else {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (x + h >= W) break;
                dst[o + h] = dd[h];
                if (DEPTH == 1) ((float*)depth_)[o + h] = zf[h];
                if (DEPTH == 2) ((double*)depth_)[o + h] = zd[h];
            }
        }
    }
}

// 3x3 median of the s16 disparity (inside StereoSGBM, A.5) for TWO horizontally adjacent pixels per thread on packed
// s16x2 words: forgetful selection (keep 6 of the 9 samples, drop min and max, take in the next) with one VIMNMX.S16x2 per
// min / max of two pixels; replicate border by clamped loads.
__global__ void __launch_bounds__(256) k_median3_s16x2(const int16_t* __restrict__ src, int H, int W, int16_t* __restrict__ dst)
{
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x), y = blockIdx.y;
    if (x >= W) return;
    const size_t img = (size_t)blockIdx.z * H * W;
    int xx[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) xx[i] = min(max(x + i - 1, 0), W - 1);
    s16x2 s[9];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const unsigned short* r = reinterpret_cast<const unsigned short*>(src) + img + (size_t)min(max(y + j - 1, 0), H - 1) * W;
        uint32_t p[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) p[i] = __ldg(r + xx[i]);
#pragma unroll
        for (int i = 0; i < 3; ++i) s[j * 3 + i].v = p[i] | (p[i + 1] << 16);   // (column x + i - 1, column x + i)
    }
    s16x2 v[14];
#pragma unroll
    for (int i = 0; i < 6; ++i) v[i] = s[i];
    minmax<6>(v); v[0] = s[6];
    minmax<5>(v); v[0] = s[7];
    minmax<4>(v); v[0] = s[8];
    minmax<3>(v);
    int16_t* o = dst + img + (size_t)y * W + x;
    if (x + 1 < W && !(reinterpret_cast<uintptr_t>(o) & 3)) *reinterpret_cast<uint32_t*>(o) = v[1].v;
    else {
        o[0] = (int16_t)(v[1].v & 0xffffu);
        if (x + 1 < W) o[1] = (int16_t)(v[1].v >> 16);
    }
}

extern "C" int svb_median3_s16_dev(const int16_t* src, int B, int H, int W, int16_t* dst, void* stream)
{
    if (!src || !dst || B <= 0 || H <= 0 || W <= 0) { svb_set_error("median3: bad argument"); return SVB_ERR_PARAM; }
    k_median3_s16x2<<<dim3(svb_ceil_div(W, 512), H, B), 256, 0, (cudaStream_t)stream>>>(src, H, W, dst);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_median5_f32_dev(const float* src, int B, int H, int W, float* dst, void* stream)
{
    if (!src || !dst || B <= 0 || H <= 0 || W <= 0) { svb_set_error("median5: bad argument"); return SVB_ERR_PARAM; }
    k_median5_f32<false, 0><<<dim3(svb_ceil_div(W, 256), (H + 1) / 2, B), 256, 0, (cudaStream_t)stream>>>(
        src, H, W, dst, nullptr, 0.0, 0.f, 0.0, 0.0);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_disp_to_depth_dev(const int16_t* disp16, int B, int H, int W, double fx, double baseline,
                                     double min_disp, double min_depth, double max_depth,
                                     float* disparity, float* depth, void* stream)
{
    if (!disp16 || !disparity || !depth || B <= 0 || H <= 0 || W <= 0) {
        svb_set_error("disp_to_depth: bad argument"); return SVB_ERR_PARAM;
    }
    // Python computes fx*BASELINE in double, NumPy then casts the scalar to float32
    k_median5_s16x2<1><<<dim3(svb_ceil_div(W, 512), (H + 1) / 2, B), 256, 0, (cudaStream_t)stream>>>(
        disp16, H, W, disparity, (void*)depth, fx * baseline, (float)min_disp, min_depth, max_depth);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_disp_to_depth_f64_dev(const int16_t* disp16, int B, int H, int W, double fx, double baseline,
                                         double min_disp, double min_depth, double max_depth,
                                         float* disparity, double* depth, void* stream)
{
    if (!disp16 || !disparity || !depth || B <= 0 || H <= 0 || W <= 0) {
        svb_set_error("disp_to_depth_f64: bad argument"); return SVB_ERR_PARAM;
    }
    k_median5_s16x2<2><<<dim3(svb_ceil_div(W, 512), (H + 1) / 2, B), 256, 0, (cudaStream_t)stream>>>(
        disp16, H, W, disparity, (void*)depth, fx * baseline, (float)min_disp, min_depth, max_depth);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ---------------------------------------------------------------------------------------
// filterSpeckles = connected-component labelling (4-neighbourhood, |a-b| <= maxDiff, both
// != newVal) + size threshold.  Horizontal runs are labelled without atomics; union-find with
// atomicMin hooking (roots are the smallest linear index of their component) joins the runs
// vertically, so the outcome is deterministic.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ int ccl_find(const int32_t* label, int a)
{
    int p = label[a];
    while (p != a) { a = p; p = label[a]; }
    return a;
}
__device__ __forceinline__ void ccl_union(int32_t* label, int a, int b)
{
    while (true) {
        a = ccl_find(label, a);
        b = ccl_find(label, b);
        if (a == b) return;
        if (a < b) { int t = a; a = b; b = t; }      // a > b: hook a under b
        const int old = atomicMin(&label[a], b);
        if (old == a) return;
        a = old;
    }
}
// Rows first: a warp walks one image row in 32-pixel chunks and labels every pixel with the
// linear index of the first pixel of its horizontal run (ballot + clz, no atomics).  Only the
// run heads are union-find nodes afterwards.
__global__ void __launch_bounds__(256) k_ccl_runs(const int16_t* __restrict__ img, int rows, int W, int newVal,
                                                  int maxDiff, int32_t* __restrict__ label, int32_t* __restrict__ count)
{
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const size_t base = (size_t)row * W;
    int carry = -1, lastv = newVal;
    for (int x0 = 0; x0 < W; x0 += 32) {
        const int x = x0 + lane;
        const int v = x < W ? (int)img[base + x] : newVal;
        int left = __shfl_up_sync(0xffffffffu, v, 1);
        if (lane == 0) left = lastv;
        const bool valid = v != newVal;
        const bool start = valid && !(left != newVal && abs(v - left) <= maxDiff);
        const unsigned smask = __ballot_sync(0xffffffffu, start);
        const unsigned m = smask & ((2u << lane) - 1u);
        const int rs = m ? x0 + 31 - __clz(m) : carry;
        if (x < W) {
            label[base + x] = valid ? (int32_t)(base + rs) : -1;
            count[base + x] = 0;
        }
        carry = __shfl_sync(0xffffffffu, valid ? rs : -1, 31);
        lastv = __shfl_sync(0xffffffffu, v, 31);
    }
}
// Vertical links: one union per (run, run above) contact stretch -- a pixel whose left neighbour is in
// the same run on both rows and is itself linked upwards leaves the union to that neighbour.
__global__ void __launch_bounds__(256) k_ccl_vmerge(const int16_t* __restrict__ img, int H, int W, int newVal,
                                                    int maxDiff, int32_t* __restrict__ label)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y + 1;
    if (x >= W) return;
    const int p = (int)(((size_t)blockIdx.z * H + y) * W + x);
    const int v = img[p];
    if (v == newVal) return;
    const int u = img[p - W];
    if (u == newVal || abs(v - u) > maxDiff) return;
    if (x > 0) {
        const int vl = img[p - 1], ul = img[p - W - 1];
        if (vl != newVal && ul != newVal && abs(v - vl) <= maxDiff && abs(u - ul) <= maxDiff && abs(vl - ul) <= maxDiff) return;
    }
    ccl_union(label, p, p - W);
}
// Component sizes: the last pixel of every run adds the run length to its root, and points the run
// head straight at the root.
__global__ void __launch_bounds__(256) k_ccl_count(const int16_t* __restrict__ img, int W, int newVal, int maxDiff,
                                                   int32_t* __restrict__ label, int32_t* __restrict__ count)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= W) return;
    const int p = (int)(((size_t)blockIdx.z * gridDim.y + blockIdx.y) * W + x);
    const int v = img[p];
    if (v == newVal) return;
    if (x + 1 < W) {
        const int vn = img[p + 1];
        if (vn != newVal && abs(vn - v) <= maxDiff) return;           // not the end of its run
    }
    bool head = x == 0;
    if (!head) { const int vl = img[p - 1]; head = vl == newVal || abs(v - vl) > maxDiff; }
    const int rs = head ? p : label[p];
    const int root = ccl_find(label, rs);
    atomicAdd(&count[root], p - rs + 1);
    if (root != rs) label[rs] = root;
}
__global__ void __launch_bounds__(256) k_ccl_apply(int16_t* __restrict__ img, size_t n, int newVal, int maxSize,
                                                   const int32_t* __restrict__ label, const int32_t* __restrict__ count)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (label[i] >= 0) {
        const int root = ccl_find(label, (int)i);
        if (count[root] <= maxSize) img[i] = (int16_t)newVal;
    }
}

extern "C" int svb_filter_speckles_s16_dev(int16_t* img, int B, int H, int W, int newVal, int maxSpeckleSize,
                                           int maxDiff, int32_t* scratch, void* stream)
{
    if (!img || !scratch || B <= 0 || H <= 0 || W <= 0) { svb_set_error("filterSpeckles: bad argument"); return SVB_ERR_PARAM; }
    const size_t n = (size_t)B * H * W;
    if (n >= ((size_t)1 << 31)) { svb_set_error("filterSpeckles: batch too large for 32-bit labels"); return SVB_ERR_UNSUPPORTED; }
    cudaStream_t st = (cudaStream_t)stream;
    int32_t* label = scratch;
    int32_t* count = scratch + n;
    const int rows = B * H;
    k_ccl_runs<<<svb_ceil_div(rows, 8), 256, 0, st>>>(img, rows, W, newVal, maxDiff, label, count);
    SVB_LAUNCH_CHECK();
    if (H > 1) {
        k_ccl_vmerge<<<dim3(svb_ceil_div(W, 256), H - 1, B), 256, 0, st>>>(img, H, W, newVal, maxDiff, label);
        SVB_LAUNCH_CHECK();
    }
    k_ccl_count<<<dim3(svb_ceil_div(W, 256), H, B), 256, 0, st>>>(img, W, newVal, maxDiff, label, count);
    SVB_LAUNCH_CHECK();
    k_ccl_apply<<<svb_ceil_div(n, 256), 256, 0, st>>>(img, n, newVal, maxSpeckleSize, label, count);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ---------------------------------------------------------------------------------------
// svb_filter_speckles_s16_static: the same filter for callers that cannot allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The labels + counts live in a file-scope device array of SVB_SPECKLE_STATIC_MAX_PIXELS pixels
// (2 x int32 each = 16 MiB, held by every GPU that loads the library until the process exits).  In host code the
// array's NAME is only a host-side stand-in: its device address is fetched once per device, with that device
// current, through cudaGetSymbolAddress and handed to the kernels.  Nothing here calls cudaMalloc / cudaFree or
// an allocating library routine.  The array keeps the previous call's contents; nothing has to be cleared by
// the host because k_ccl_runs writes every label and zeroes every count of the n pixels before they are read.
// Every call (any stream, any thread) shares the array: calls that overlap on the GPU race.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) int32_t g_speckle_scratch[2 * (size_t)SVB_SPECKLE_STATIC_MAX_PIXELS];
#define SVB_STATIC_MAX_DEVICES 64
static int32_t* g_speckle_scratch_dev[SVB_STATIC_MAX_DEVICES] = {};    // device address per device ordinal
static std::mutex g_speckle_scratch_mu;

extern "C" int svb_filter_speckles_s16_static(int16_t* img, int B, int H, int W, int newVal, int maxSpeckleSize,
                                              int maxDiff, void* stream)
{
    if (!img || B <= 0 || H <= 0 || W <= 0) { svb_set_error("filterSpeckles (static): bad argument"); return SVB_ERR_PARAM; }
    const size_t n = (size_t)B * H * W;
    if (n > (size_t)SVB_SPECKLE_STATIC_MAX_PIXELS) {
        svb_set_error("filterSpeckles (static): B*H*W = %zu pixels exceeds the built-in scratch of %d pixels "
                      "(SVB_SPECKLE_STATIC_MAX_PIXELS); use svb_filter_speckles_s16_dev with caller scratch",
                      n, SVB_SPECKLE_STATIC_MAX_PIXELS);
        return SVB_ERR_UNSUPPORTED;
    }
    int dev = -1;
    SVB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= SVB_STATIC_MAX_DEVICES) { svb_set_error("filterSpeckles (static): device ordinal %d unsupported", dev); return SVB_ERR_UNSUPPORTED; }
    int32_t* scratch;
    {
        std::lock_guard<std::mutex> lk(g_speckle_scratch_mu);
        if (!g_speckle_scratch_dev[dev]) {                   // set-up, once per device, this device current
            void* ptr = nullptr;
            SVB_CUDA(cudaGetSymbolAddress(&ptr, g_speckle_scratch));
            if (!ptr) { svb_set_error("filterSpeckles (static): cudaGetSymbolAddress returned a null address"); return SVB_ERR_CUDA; }
            g_speckle_scratch_dev[dev] = (int32_t*)ptr;
        }
        scratch = g_speckle_scratch_dev[dev];
    }
    return svb_filter_speckles_s16_dev(img, B, H, W, newVal, maxSpeckleSize, maxDiff, scratch, stream);
}

// ---------------------------------------------------------------------------------------
// remap, INTER_LINEAR with fixed-point maps (A.7): src tap (sx,sy) = map1, 5+5-bit fraction
// in map2; weights (32-fy)(32-fx)*32 ... sum to 32768; (acc + 16384) >> 15; taps outside the
// source contribute 0.  One thread per output pixel; C = 1 or 3 interleaved channels; the
// maps are shared by the whole batch (blockIdx.z) and stay L2-resident.
// ---------------------------------------------------------------------------------------
// one destination pixel of cv2.remap (A.7): fixed-point bilinear tap of the 2x2 source block, BORDER_CONSTANT 0
template <int C>
__device__ __forceinline__ void remap_pixel(const uint8_t* __restrict__ img, const uint8_t* lim, int Hs, int Ws, int sx,
                                            int sy, int f, int (&out)[C])
{
    const int fx = f & 31, fy = f >> 5;
    if constexpr (C == 3) {
        // Interior pixel: the two taps of a source row are 6 contiguous bytes; fetch them as aligned 32-bit words
        // (three loads and two funnel shifts per row instead of six byte loads -- the kernel is bound by the number
        // of load instructions, not by bytes).  `lim` = last address from which a 12-byte read stays inside src.
        // (32-bit byte offsets from the image's 4-byte aligned base: one 64-bit address per source row)
        const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(img) & 3);
        const uint32_t* base = reinterpret_cast<const uint32_t*>(img - mis);
        const uint32_t o0 = ((uint32_t)sy * (uint32_t)Ws + (uint32_t)sx) * 3u + mis, o1 = o0 + (uint32_t)Ws * 3u;
        const size_t room = (size_t)(lim - reinterpret_cast<const uint8_t*>(base));   // (the rest of the batch: may exceed 32 bits)
        const uint32_t limoff = room > 0xffffffffull ? 0xffffffffu : (uint32_t)room;
        if ((unsigned)sx < (unsigned)(Ws - 1) && (unsigned)sy < (unsigned)(Hs - 1) && o1 <= limoff) {   // all four taps inside
            uint32_t lo[2], hi[2];
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                const uint32_t o = r ? o1 : o0;
                const uint32_t* w = base + (o >> 2);
                const uint32_t sh = (o & 3u) * 8u;
                const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
                lo[r] = __funnelshift_r(w0, w1, sh);
                hi[r] = __funnelshift_r(w1, w2, sh);
            }
            // tap x0 = bytes 0..2, tap x1 = bytes 3..5 of (lo, hi).  Two exact stages instead of four products per
            // channel: h_r = (32 - fx) a_r0 + fx a_r1 as byte dot products (the weight words select the channel's two
            // bytes, so no byte is extracted), v = (32 - fy) h_0 + fy h_1, and (32 v + 16384) >> 15 = (v + 512) >> 10.
            // Non-negative weights that sum to 32768 keep the result in [0, 255]: no clamp.
            const uint32_t gx = 32u - (uint32_t)fx, ux = (uint32_t)fx;
            const uint32_t wB = gx | (ux << 24), wGl = gx << 8, wRl = gx << 16, wRh = ux << 8;
            const int gy = 32 - fy;
            int h[2][3];
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                h[r][0] = __dp4a(lo[r], wB, 0u);
                h[r][1] = __dp4a(hi[r], ux, __dp4a(lo[r], wGl, 0u));
                h[r][2] = __dp4a(hi[r], wRh, __dp4a(lo[r], wRl, 0u));
            }
#pragma unroll
            for (int c = 0; c < 3; ++c) out[c] = (gy * h[0][c] + fy * h[1][c] + 512) >> 10;
            return;
        }
    }
    const int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32;
    const int w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
    const bool x0 = sx >= 0 && sx < Ws, x1 = sx + 1 >= 0 && sx + 1 < Ws;
    const bool y0 = sy >= 0 && sy < Hs, y1 = sy + 1 >= 0 && sy + 1 < Hs;
#pragma unroll
    for (int c = 0; c < C; ++c) {
        int acc = 16384;
        if (y0) {
            const uint8_t* r = img + ((size_t)sy * Ws + sx) * C + c;
            if (x0) acc += w00 * r[0];
            if (x1) acc += w01 * r[C];
        }
        if (y1) {
            const uint8_t* r = img + ((size_t)(sy + 1) * Ws + sx) * C + c;
            if (x0) acc += w10 * r[0];
            if (x1) acc += w11 * r[C];
        }
        out[c] = min(max(acc >> 15, 0), 255);
    }
}
__device__ __forceinline__ int gray_of(int b, int g, int r) { return (3735 * b + 19235 * g + 9798 * r + 16384) >> 15; }

template <int C, bool GRAY>
__global__ void __launch_bounds__(256) k_remap_u8(const uint8_t* __restrict__ src, int Hs, int Ws,
                                                  const int16_t* __restrict__ map1, const uint16_t* __restrict__ map2,
                                                  int Ho, int Wo, uint8_t* __restrict__ dst)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= Wo) return;
    const size_t mo = (size_t)y * Wo + x;
    const short2 m1 = reinterpret_cast<const short2*>(map1)[mo];
    int out[C];
    remap_pixel<C>(src + (size_t)blockIdx.z * Hs * Ws * C, src + (size_t)gridDim.z * Hs * Ws * C - 16, Hs, Ws, m1.x, m1.y,
                   map2[mo] & 1023, out);
    const size_t o = ((size_t)blockIdx.z * Ho + y) * Wo + x;
    if (GRAY) {
        dst[o] = (uint8_t)gray_of(out[0], out[1], out[C - 1]);
    } else {
#pragma unroll
        for (int c = 0; c < C; ++c) dst[o * C + c] = (uint8_t)out[c];
    }
}

// Four destination pixels per thread (Wo % 4 == 0, 16-byte aligned maps and destination): the maps come in as one
// 128-bit and one 64-bit load, the result leaves as 32-bit words (3 for BGR, 1 for gray / single channel); the
// source taps stay byte gathers through the read-only path (neighbouring pixels share their cache lines).
template <int C, bool GRAY>
__global__ void __launch_bounds__(256) k_remap_u8x4(const uint8_t* __restrict__ src, int Hs, int Ws,
                                                    const int16_t* __restrict__ map1, const uint16_t* __restrict__ map2,
                                                    int Ho, int Wo, uint8_t* __restrict__ dst)
{
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x >= Wo) return;
    const size_t mo = (size_t)y * Wo + x;
    const int4 m1 = *reinterpret_cast<const int4*>(map1 + 2 * mo);          // 4 x (sx | sy << 16)
    const uint2 m2 = *reinterpret_cast<const uint2*>(map2 + mo);            // 4 x u16
    const uint8_t* img = src + (size_t)blockIdx.z * Hs * Ws * C;
    const uint8_t* lim = src + (size_t)gridDim.z * Hs * Ws * C - 16;       // word loads stay inside the batch
    const int mm[4] = {m1.x, m1.y, m1.z, m1.w};
    const unsigned ff[4] = {m2.x & 0xffffu, m2.x >> 16, m2.y & 0xffffu, m2.y >> 16};
    constexpr int CO = GRAY ? 1 : C;
    uint8_t px[4 * CO];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        int out[C];
        remap_pixel<C>(img, lim, Hs, Ws, (int)(short)(mm[q] & 0xffff), mm[q] >> 16, (int)(ff[q] & 1023u), out);
        if (GRAY) px[q] = (uint8_t)gray_of(out[0], out[1], out[C - 1]);
        else {
#pragma unroll
            for (int c = 0; c < C; ++c) px[q * C + c] = (uint8_t)out[c];
        }
    }
    uint32_t* o = reinterpret_cast<uint32_t*>(dst + (((size_t)blockIdx.z * Ho + y) * Wo + x) * CO);
#pragma unroll
    for (int w = 0; w < CO; ++w)
        o[w] = (uint32_t)px[4 * w] | ((uint32_t)px[4 * w + 1] << 8) | ((uint32_t)px[4 * w + 2] << 16) | ((uint32_t)px[4 * w + 3] << 24);
}

extern "C" int svb_remap_u8_dev(const uint8_t* src, int B, int Hs, int Ws, int C, const int16_t* map1,
                                const uint16_t* map2, int Ho, int Wo, uint8_t* dst, int to_gray, void* stream)
{
    if (!src || !map1 || !map2 || !dst || B <= 0 || Hs <= 0 || Ws <= 0 || Ho <= 0 || Wo <= 0) {
        svb_set_error("remap: bad argument"); return SVB_ERR_PARAM;
    }
    if (C != 1 && C != 3) { svb_set_error("remap: channels must be 1 or 3 (got %d)", C); return SVB_ERR_UNSUPPORTED; }
    if (to_gray && C != 3) { svb_set_error("remap: to_gray needs 3 channels"); return SVB_ERR_PARAM; }
    dim3 grid(svb_ceil_div(Wo, 256), Ho, B);
    cudaStream_t st = (cudaStream_t)stream;
    const bool x4 = Wo % 4 == 0 && ((uintptr_t)map1 & 15) == 0 && ((uintptr_t)map2 & 7) == 0 && ((uintptr_t)dst & 3) == 0;
    if (x4) {
        dim3 g4(svb_ceil_div(Wo / 4, 128), Ho, B);
        if (C == 1) k_remap_u8x4<1, false><<<g4, 128, 0, st>>>(src, Hs, Ws, map1, map2, Ho, Wo, dst);
        else if (to_gray) k_remap_u8x4<3, true><<<g4, 128, 0, st>>>(src, Hs, Ws, map1, map2, Ho, Wo, dst);
        else k_remap_u8x4<3, false><<<g4, 128, 0, st>>>(src, Hs, Ws, map1, map2, Ho, Wo, dst);
    } else if (C == 1) k_remap_u8<1, false><<<grid, 256, 0, st>>>(src, Hs, Ws, map1, map2, Ho, Wo, dst);
    else if (to_gray) k_remap_u8<3, true><<<grid, 256, 0, st>>>(src, Hs, Ws, map1, map2, Ho, Wo, dst);
    else k_remap_u8<3, false><<<grid, 256, 0, st>>>(src, Hs, Ws, map1, map2, Ho, Wo, dst);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// BGR -> gray (A.8): four pixels per thread -- three 32-bit loads, one 32-bit store -- when both pointers allow it
__global__ void __launch_bounds__(256) k_bgr2gray(const uint8_t* __restrict__ bgr, size_t n, uint8_t* __restrict__ gray)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t* p = bgr + 3 * i;
    gray[i] = (uint8_t)((3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15);
}
__global__ void __launch_bounds__(256) k_bgr2gray_x4(const uint32_t* __restrict__ bgr, size_t n4, uint32_t* __restrict__ gray)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const uint32_t a = __ldg(bgr + 3 * i), b = __ldg(bgr + 3 * i + 1), c = __ldg(bgr + 3 * i + 2);   // B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
    auto gr = [](uint32_t B, uint32_t G, uint32_t R) { return (3735u * B + 19235u * G + 9798u * R + 16384u) >> 15; };
    const uint32_t g0 = gr(a & 0xff, (a >> 8) & 0xff, (a >> 16) & 0xff);
    const uint32_t g1 = gr(a >> 24, b & 0xff, (b >> 8) & 0xff);
    const uint32_t g2 = gr((b >> 16) & 0xff, b >> 24, c & 0xff);
    const uint32_t g3 = gr((c >> 8) & 0xff, (c >> 16) & 0xff, c >> 24);
    gray[i] = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
}
extern "C" int svb_bgr2gray_dev(const uint8_t* bgr, size_t npix, uint8_t* gray, void* stream)
{
    if (!bgr || !gray || npix == 0) { svb_set_error("bgr2gray: bad argument"); return SVB_ERR_PARAM; }
    const size_t n4 = (((uintptr_t)bgr | (uintptr_t)gray) & 3) == 0 ? npix / 4 : 0;
    if (n4) k_bgr2gray_x4<<<svb_ceil_div(n4, 256), 256, 0, (cudaStream_t)stream>>>((const uint32_t*)bgr, n4, (uint32_t*)gray);
    if (npix > 4 * n4)
        k_bgr2gray<<<svb_ceil_div(npix - 4 * n4, 256), 256, 0, (cudaStream_t)stream>>>(bgr + 12 * n4, npix - 4 * n4, gray + 4 * n4);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ---------------------------------------------------------------------------------------
// reprojectImageTo3D: [X Y Z W]^T = Q [x y d 1]^T, out = (X,Y,Z)/W, double accumulate.
// handleMissingValues: Z = 10000 where d equals the image's minimum disparity.
// ---------------------------------------------------------------------------------------
struct QMat { double q[16]; };
template <bool S16> __device__ __forceinline__ float load_disp(const void* p, size_t i)
{
    return S16 ? (float)((const int16_t*)p)[i] : ((const float*)p)[i];
}
template <bool S16>
__global__ void __launch_bounds__(256) k_min_per_image(const void* __restrict__ disp, size_t npix, float* __restrict__ minv)
{
    // minv[b] pre-set to +inf bit pattern; non-negative-float trick does not hold for negative
    // disparities, so reduce in a block and finish with an integer CAS-free ordered-int atomic.
    const size_t base = (size_t)blockIdx.y * npix;
    float m = FLT_MAX;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x)
        m = fminf(m, load_disp<S16>(disp, base + i));
    for (int o = 16; o; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) {
        // order-preserving map float -> int for atomicMin
        int bits = __float_as_int(m);
        bits = bits >= 0 ? bits : bits ^ 0x7fffffff;
        atomicMin(reinterpret_cast<int*>(minv) + blockIdx.y, bits);
    }
}
template <bool S16>
__global__ void __launch_bounds__(256) k_reproject(const void* __restrict__ disp, int H, int W, QMat Q, int handle,
                                                   const float* __restrict__ minv, float* __restrict__ out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const size_t i = ((size_t)blockIdx.z * H + y) * W + x;
    const double d = load_disp<S16>(disp, i);
    const double X = Q.q[0] * x + Q.q[1] * y + Q.q[2] * d + Q.q[3];
    const double Y = Q.q[4] * x + Q.q[5] * y + Q.q[6] * d + Q.q[7];
    double Z = Q.q[8] * x + Q.q[9] * y + Q.q[10] * d + Q.q[11];
    const double Wd = Q.q[12] * x + Q.q[13] * y + Q.q[14] * d + Q.q[15];
    const double iw = 1.0 / Wd;
    float zf = (float)(Z * iw);
    if (handle) {
        int bits = reinterpret_cast<const int*>(minv)[blockIdx.z];
        bits = bits >= 0 ? bits : bits ^ 0x7fffffff;
        if (fabs(d - (double)__int_as_float(bits)) <= FLT_EPSILON) zf = 10000.f;
    }
    out[3 * i] = (float)(X * iw);
    out[3 * i + 1] = (float)(Y * iw);
    out[3 * i + 2] = zf;
}
__global__ void k_fill_i32(int* p, int n, int v) { int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = v; }

extern "C" int svb_reproject3d_dev(const void* disp, int is_s16, int B, int H, int W, const double* Q16,
                                   int handleMissingValues, float* minscratch_dev, float* out, void* stream)
{
    if (!disp || !Q16 || !out || B <= 0 || H <= 0 || W <= 0) { svb_set_error("reproject3d: bad argument"); return SVB_ERR_PARAM; }
    if (handleMissingValues && !minscratch_dev) { svb_set_error("reproject3d: handleMissingValues needs B floats of scratch"); return SVB_ERR_PARAM; }
    QMat Q;
    for (int i = 0; i < 16; ++i) Q.q[i] = Q16[i];
    cudaStream_t st = (cudaStream_t)stream;
    if (handleMissingValues) {
        k_fill_i32<<<svb_ceil_div(B, 256), 256, 0, st>>>((int*)minscratch_dev, B, 0x7f7fffff);
        dim3 g(std::min(svb_ceil_div((size_t)H * W, 1024), 512), B);
        if (is_s16) k_min_per_image<true><<<g, 256, 0, st>>>(disp, (size_t)H * W, minscratch_dev);
        else k_min_per_image<false><<<g, 256, 0, st>>>(disp, (size_t)H * W, minscratch_dev);
        SVB_LAUNCH_CHECK();
    }
    dim3 grid(svb_ceil_div(W, 256), H, B);
    if (is_s16) k_reproject<true><<<grid, 256, 0, st>>>(disp, H, W, Q, handleMissingValues, minscratch_dev, out);
    else k_reproject<false><<<grid, 256, 0, st>>>(disp, H, W, Q, handleMissingValues, minscratch_dev, out);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------------------------------
// cv2.initUndistortRectifyMap(K, dist, R, P, size, CV_16SC2)  (depth_map.py:633-638,
// fused_depth_map.py:404-409).  One thread per destination pixel, float64 like
// OpenCV: back-project through ir = inv(P[:, :3] * R) (inverted on the host), rational +
// tangential + thin-prism distortion, project with K, then the fixed-point form cv2.remap
// consumes: map1 = (iu >> 5, iv >> 5) as s16, map2 = (iv & 31) * 32 + (iu & 31).
// ------------------------------------------------------------------------------------------
struct MapArgs {
    double ir[9];
    double k[12];             // k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4
    double fx, fy, u0, v0;
};

__device__ __forceinline__ int cv_round_sat(double v)
{
    // cvRound = round half to even (lrint); saturate_cast<int>
    if (!(v > -2147483648.0)) return INT_MIN;
    if (v >= 2147483647.0) return INT_MAX;
    return __double2int_rn(v);
}

__global__ void __launch_bounds__(256) k_init_undistort_rectify_map(MapArgs a, int H, int W, short2* __restrict__ map1,
                                                                     unsigned short* __restrict__ map2)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= W) return;
    // explicit _rn operations: no FMA contraction, the same roundings as the scalar C++ expression
    const double _x = __dadd_rn(__dmul_rn((double)j, a.ir[0]), __dadd_rn(__dmul_rn((double)i, a.ir[1]), a.ir[2]));
    const double _y = __dadd_rn(__dmul_rn((double)j, a.ir[3]), __dadd_rn(__dmul_rn((double)i, a.ir[4]), a.ir[5]));
    const double _w = __dadd_rn(__dmul_rn((double)j, a.ir[6]), __dadd_rn(__dmul_rn((double)i, a.ir[7]), a.ir[8]));
    const double w = __ddiv_rn(1.0, _w);
    const double x = __dmul_rn(_x, w), y = __dmul_rn(_y, w);
    const double x2 = __dmul_rn(x, x), y2 = __dmul_rn(y, y);
    const double r2 = __dadd_rn(x2, y2), _2xy = __dmul_rn(__dmul_rn(2.0, x), y);
    const double k1 = a.k[0], k2 = a.k[1], p1 = a.k[2], p2 = a.k[3], k3 = a.k[4], k4 = a.k[5], k5 = a.k[6], k6 = a.k[7];
    const double s1 = a.k[8], s2 = a.k[9], s3 = a.k[10], s4 = a.k[11];
    const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k3, r2), k2), r2), k1), r2));
    const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k6, r2), k5), r2), k4), r2));
    const double kr = __ddiv_rn(num, den);
    const double r4 = __dmul_rn(r2, r2);
    double xd = __dadd_rn(__dmul_rn(x, kr), __dmul_rn(p1, _2xy));
    xd = __dadd_rn(xd, __dmul_rn(p2, __dadd_rn(r2, __dmul_rn(2.0, x2))));
    xd = __dadd_rn(__dadd_rn(xd, __dmul_rn(s1, r2)), __dmul_rn(s2, r4));
    double yd = __dadd_rn(__dmul_rn(y, kr), __dmul_rn(p1, __dadd_rn(r2, __dmul_rn(2.0, y2))));
    yd = __dadd_rn(yd, __dmul_rn(p2, _2xy));
    yd = __dadd_rn(__dadd_rn(yd, __dmul_rn(s3, r2)), __dmul_rn(s4, r4));
    const double u = __dadd_rn(__dmul_rn(a.fx, xd), a.u0), v = __dadd_rn(__dmul_rn(a.fy, yd), a.v0);
    const int iu = cv_round_sat(__dmul_rn(u, 32.0)), iv = cv_round_sat(__dmul_rn(v, 32.0));
    const size_t o = (size_t)i * W + j;
    map1[o] = make_short2((short)(iu >> 5), (short)(iv >> 5));
    map2[o] = (unsigned short)((iv & 31) * 32 + (iu & 31));
}

extern "C" int svb_init_undistort_rectify_map_dev(const double* K9, const double* dist, int ndist, const double* R9,
                                                  const double* P, int pcols, int W, int H, int16_t* map1_dev,
                                                  uint16_t* map2_dev, void* stream)
{
    if (!K9 || !P || !map1_dev || !map2_dev || W < 1 || H < 1 || (pcols != 3 && pcols != 4) || ndist < 0 ||
        (ndist && !dist)) {
        svb_set_error("svb_init_undistort_rectify_map_dev: bad argument");
        return SVB_ERR_PARAM;
    }
    if (ndist != 0 && ndist != 4 && ndist != 5 && ndist != 8 && ndist != 12 && ndist != 14) {
        svb_set_error("distCoeffs must have 4, 5, 8, 12 or 14 elements (got %d)", ndist);   // cv2's own check
        return SVB_ERR_PARAM;
    }
    if (ndist == 14 && (dist[12] != 0.0 || dist[13] != 0.0)) {
        svb_set_error("tilted sensor model (tauX, tauY) is not implemented");
        return SVB_ERR_UNSUPPORTED;
    }
    MapArgs a;
    for (int q = 0; q < 12; ++q) a.k[q] = q < ndist ? dist[q] : 0.0;
    a.fx = K9[0]; a.fy = K9[4]; a.u0 = K9[2]; a.v0 = K9[5];
    // M = P[:, :3] * R, ir = M^-1 (adjugate / determinant in double)
    double M[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            double s = 0;
            for (int q = 0; q < 3; ++q) s += P[r * pcols + q] * (R9 ? R9[q * 3 + c] : (q == c ? 1.0 : 0.0));
            M[r * 3 + c] = s;
        }
    const double det = M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) +
                       M[2] * (M[3] * M[7] - M[4] * M[6]);
    if (det == 0.0) { svb_set_error("newCameraMatrix * R is singular"); return SVB_ERR_PARAM; }
    const double id = 1.0 / det;
    a.ir[0] = (M[4] * M[8] - M[5] * M[7]) * id; a.ir[1] = (M[2] * M[7] - M[1] * M[8]) * id; a.ir[2] = (M[1] * M[5] - M[2] * M[4]) * id;
    a.ir[3] = (M[5] * M[6] - M[3] * M[8]) * id; a.ir[4] = (M[0] * M[8] - M[2] * M[6]) * id; a.ir[5] = (M[2] * M[3] - M[0] * M[5]) * id;
    a.ir[6] = (M[3] * M[7] - M[4] * M[6]) * id; a.ir[7] = (M[1] * M[6] - M[0] * M[7]) * id; a.ir[8] = (M[0] * M[4] - M[1] * M[3]) * id;
    k_init_undistort_rectify_map<<<dim3(svb_ceil_div(W, 256), H), 256, 0, (cudaStream_t)stream>>>(
        a, H, W, (short2*)map1_dev, map2_dev);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ---------------------------------------------------------------------------------------
// cv2.resize(src, dsize, interpolation=INTER_LINEAR): u8 (1 or 3 interleaved channels, bit-exact) and f32.
// Reference call sites: fused_depth_map.py:473-476 (both camera frames, EVERY processed frame, PROCESSING_SCALE
// 0.33), depth_map.py:776, 809-811 (capture size != calibration size), fused_depth_map.py:1214-1217 (mono map to
// the stereo shape, float32).  Arithmetic as restated and pinned in oracle/imgproc_oracle.c (orc_resize_*):
// source coordinate in double without FMA contraction, cast to float, horizontal fraction reset where the tap
// clamps, vertical taps clipped WITHOUT resetting the fraction, 11-bit weights (round half even), 32-bit
// horizontal sums, (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2.
// One thread per destination pixel; the four source taps of neighbouring threads share sectors (downscaling
// reads at most scale^2 source bytes per output byte, once: the kernel is a streaming read).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void resize_coord_u8(int d, double scale, int ssize, bool clamp, int& s, int& w0, int& w1)
{
    float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
    s = (int)floorf(f);
    f = __fsub_rn(f, (float)s);
    if (clamp) {
        if (s < 0) { f = 0.f; s = 0; }
        if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
    }
    w0 = __float2int_rn(__fmul_rn(__fsub_rn(1.f, f), 2048.f));
    w1 = __float2int_rn(__fmul_rn(f, 2048.f));
}

template <int C>
__global__ void __launch_bounds__(256) k_resize_u8(const uint8_t* __restrict__ src, int Hs, int Ws,
                                                   uint8_t* __restrict__ dst, int Hd, int Wd, double sx, double sy)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= Wd) return;
    int xs, a0, a1, ys, b0, b1;
    resize_coord_u8(x, sx, Ws, true, xs, a0, a1);
    resize_coord_u8(y, sy, Hs, false, ys, b0, b1);
    const int x1 = min(xs + 1, Ws - 1);
    const int y0 = min(max(ys, 0), Hs - 1), y1 = min(max(ys + 1, 0), Hs - 1);
    const uint8_t* img = src + (size_t)blockIdx.z * Hs * Ws * C;
    const uint8_t *r0 = img + (size_t)y0 * Ws * C, *r1 = img + (size_t)y1 * Ws * C;
    uint8_t* o = dst + (((size_t)blockIdx.z * Hd + y) * Wd + x) * C;
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const int S0 = r0[xs * C + c] * a0 + r0[x1 * C + c] * a1;
        const int S1 = r1[xs * C + c] * a0 + r1[x1 * C + c] * a1;
        o[c] = (uint8_t)((((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2);
    }
}

// float32, one channel: coordinate and fraction in double, weight rounded to float once (what this cv2 build does,
// oracle/imgproc_oracle.c); products and sums as separate float operations.
__global__ void __launch_bounds__(256) k_resize_f32(const float* __restrict__ src, int Hs, int Ws,
                                                    float* __restrict__ dst, int Hd, int Wd, double sx, double sy)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= Wd) return;
    const double fy = __dadd_rn(__dmul_rn((double)y + 0.5, sy), -0.5);
    const int ys = (int)floor(fy);
    const float wy1 = (float)(fy - (double)ys), wy0 = __fsub_rn(1.f, wy1);
    const int y0 = min(max(ys, 0), Hs - 1), y1 = min(max(ys + 1, 0), Hs - 1);
    double fx = __dadd_rn(__dmul_rn((double)x + 0.5, sx), -0.5);
    int xs = (int)floor(fx);
    double fr = fx - (double)xs;
    if (xs < 0) { fr = 0; xs = 0; }
    if (xs >= Ws - 1) { fr = 0; xs = Ws - 1; }
    const int x1 = min(xs + 1, Ws - 1);
    const float wx1 = (float)fr, wx0 = __fsub_rn(1.f, wx1);
    const float* img = src + (size_t)blockIdx.z * Hs * Ws;
    const float *p0 = img + (size_t)y0 * Ws, *p1 = img + (size_t)y1 * Ws;
    const float r0 = __fadd_rn(__fmul_rn(p0[xs], wx0), __fmul_rn(p0[x1], wx1));
    const float r1 = __fadd_rn(__fmul_rn(p1[xs], wx0), __fmul_rn(p1[x1], wx1));
    dst[((size_t)blockIdx.z * Hd + y) * Wd + x] = __fadd_rn(__fmul_rn(r0, wy0), __fmul_rn(r1, wy1));
}

static int resize_check(const void* src, const void* dst, int B, int Hs, int Ws, int Hd, int Wd)
{
    if (!src || !dst) { svb_set_error("resize: null argument"); return SVB_ERR_PARAM; }
    if (B <= 0 || Hs <= 0 || Ws <= 0) { svb_set_error("resize: empty source (cv2: !ssize.empty())"); return SVB_ERR_PARAM; }
    if (Hd <= 0 || Wd <= 0) { svb_set_error("resize: empty destination size (cv2: !dsize.empty())"); return SVB_ERR_PARAM; }
    if (Hd > 65535 || B > 65535) { svb_set_error("resize: more than 65535 rows or images unsupported"); return SVB_ERR_UNSUPPORTED; }
    return SVB_OK;
}

extern "C" int svb_resize_u8_dev(const uint8_t* src, int B, int Hs, int Ws, int C, uint8_t* dst, int Hd, int Wd, void* stream)
{
    int rc = resize_check(src, dst, B, Hs, Ws, Hd, Wd);
    if (rc != SVB_OK) return rc;
    if (C != 1 && C != 3) { svb_set_error("resize: 1 or 3 channels (got %d)", C); return SVB_ERR_UNSUPPORTED; }
    const double sx = 1.0 / ((double)Wd / Ws), sy = 1.0 / ((double)Hd / Hs);      // cv2: scale = 1 / inv_scale
    const dim3 grid(svb_ceil_div(Wd, 256), Hd, B);
    if (C == 1) k_resize_u8<1><<<grid, 256, 0, (cudaStream_t)stream>>>(src, Hs, Ws, dst, Hd, Wd, sx, sy);
    else k_resize_u8<3><<<grid, 256, 0, (cudaStream_t)stream>>>(src, Hs, Ws, dst, Hd, Wd, sx, sy);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_resize_f32_dev(const float* src, int B, int Hs, int Ws, float* dst, int Hd, int Wd, void* stream)
{
    int rc = resize_check(src, dst, B, Hs, Ws, Hd, Wd);
    if (rc != SVB_OK) return rc;
    const double sx = 1.0 / ((double)Wd / Ws), sy = 1.0 / ((double)Hd / Hs);
    k_resize_f32<<<dim3(svb_ceil_div(Wd, 256), Hd, B), 256, 0, (cudaStream_t)stream>>>(src, Hs, Ws, dst, Hd, Wd, sx, sy);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
