// Fusion chain of fused_depth_map.py on the device:
//   normalise + confidence            fused_depth_map.py:1010-1029
//   calibrate_midas_to_stereo         fused_depth_map.py:1169-1257   (np.percentile -> exact radix select)
//   normalize_to_stereo_range         fused_depth_map.py:1503-1554
//   fuse_depth_maps                   fused_depth_map.py:1560-1698   (GaussianBlur 15, bilateralFilter 9/75/75)
//   detect_camera_occlusion stats     fused_depth_map.py:170-240
// All of it is streaming float work bound by HBM; nothing here synchronises the stream: the
// data-dependent scalars (percentiles, min/max, branch choices) stay in device memory and are
// read by the kernel that needs them.
#include "svb_common.cuh"
#include <math.h>
#include <float.h>

// ------------------------------------------------------------------------------------------
// a13: disparity -> (normalised 0..254 as float, confidence 0/1)
// ------------------------------------------------------------------------------------------
__global__ void k_norm_conf(const float* __restrict__ disp, size_t n, float lo, float hi, float numD,
                            float c_lo, float c_hi, float* __restrict__ norm, float* __restrict__ conf)
{
    auto nrm = [&](float d) {
        const float c = fminf(fmaxf(d, lo), hi);
        const float v = __fmul_rn(__fdiv_rn(__fsub_rn(c, lo), numD), 255.0f);
        return (float)(unsigned char)(int)v;
    };
    auto cnf = [&](float d) { return (d > c_lo && d < c_hi) ? 1.0f : 0.0f; };
    const bool al16 = ((reinterpret_cast<uintptr_t>(disp) | reinterpret_cast<uintptr_t>(norm) | reinterpret_cast<uintptr_t>(conf)) & 15) == 0;
    const size_t n4 = al16 ? n / 4 : 0, stride = (size_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = t0; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(disp)[i];
        reinterpret_cast<float4*>(norm)[i] = make_float4(nrm(a.x), nrm(a.y), nrm(a.z), nrm(a.w));
        reinterpret_cast<float4*>(conf)[i] = make_float4(cnf(a.x), cnf(a.y), cnf(a.z), cnf(a.w));
    }
    for (size_t i = 4 * n4 + t0; i < n; i += stride) {
        const float d = disp[i];
        norm[i] = nrm(d);
        conf[i] = cnf(d);
    }
}

extern "C" int svb_normalize_disparity_dev(const float* disp_dev, int B, int H, int W, int min_disp, int num_disp,
                                           float* norm_dev, float* conf_dev, void* stream)
{
    if (!disp_dev || !norm_dev || !conf_dev || B < 1 || H < 1 || W < 1 || num_disp < 1) {
        svb_set_error("svb_normalize_disparity_dev: bad argument");
        return SVB_ERR_PARAM;
    }
    const size_t n = (size_t)B * H * W;
    const int grid = (int)min((size_t)148 * 16, (n + 255) / 256);
    k_norm_conf<<<grid, 256, 0, (cudaStream_t)stream>>>(disp_dev, n, (float)min_disp, (float)(min_disp + num_disp - 1),
                                                        (float)num_disp, (float)(min_disp + 1),
                                                        (float)(min_disp + num_disp - 1), norm_dev, conf_dev);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------------------------------
// Exact order statistics: 3-pass radix select (11 + 11 + 10 bits) on order-preserving keys.
// One "selection" = the values of one image under a mask; each asks for two percentiles =
// four ranks (floor and floor+1 of each virtual index).  Two selections run side by side
// (blockIdx.z), one image per blockIdx.y.
// ------------------------------------------------------------------------------------------
#define SEL_NR 4
#define SEL_BINS 2048

enum { MASK_NONE = 0, MASK_GT = 1, MASK_SELF_POS = 2, MASK_GT_IF_ENOUGH = 3 };

struct SelDesc {
    const float* vals;      // [B][n]
    const float* mask;      // [B][n] or null
    int mode;               // MASK_*
    float thr;
};

struct SelState {           // per (selection, image)
    unsigned hist[SEL_NR][SEL_BINS];
    unsigned pre[SEL_NR];   // key prefix found so far
    unsigned rem[SEL_NR];   // rank inside the prefix
    unsigned count;         // selected elements
    int f64;                // percentiles interpolate in double (list q) or float (scalar q)
    int few;                // MASK_GT_IF_ENOUGH fell back to "no mask"
    double gam[2];
    float val[SEL_NR];
    double pct[2];          // the two percentiles
};

__device__ __forceinline__ unsigned f2key(float f)
{
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(unsigned k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void k_count_gt(const float* __restrict__ m, size_t n, float thr, unsigned* __restrict__ cnt)
{
    const int b = blockIdx.y;
    const float* p = m + (size_t)b * n;
    unsigned c = 0;
    const size_t n4 = (reinterpret_cast<uintptr_t>(p) & 15) == 0 ? n / 4 : 0, stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = t0; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(p)[i];
        c += (a.x > thr) + (a.y > thr) + (a.z > thr) + (a.w > thr);
    }
    for (size_t i = 4 * n4 + t0; i < n; i += stride) c += p[i] > thr;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&cnt[b], c);
}

template <int PASS>
__global__ void __launch_bounds__(512) k_sel_hist(SelDesc d0, SelDesc d1, size_t n, SelState* __restrict__ st,
                                                  const unsigned* __restrict__ relcnt, int B)
{
    constexpr int NH = PASS == 0 ? 1 : SEL_NR;
    __shared__ unsigned sh[NH][SEL_BINS];
    const int b = blockIdx.y, z = blockIdx.z;
    const SelDesc d = z ? d1 : d0;
    SelState* s = st + (size_t)z * B + b;
    for (int i = threadIdx.x; i < NH * SEL_BINS; i += blockDim.x) (&sh[0][0])[i] = 0;
    int mode = d.mode;
    if (mode == MASK_GT_IF_ENOUGH) mode = relcnt[b] < 100u ? MASK_NONE : MASK_GT;
    unsigned pre[SEL_NR];
    bool dup[SEL_NR];
    if (PASS > 0) {
#pragma unroll
        for (int r = 0; r < SEL_NR; ++r) {
            pre[r] = s->pre[r];
            dup[r] = false;
#pragma unroll
            for (int q = 0; q < r; ++q) dup[r] = dup[r] || pre[q] == pre[r];   // same prefix: count once, copy later
        }
    }
    __syncthreads();
    const float* v = d.vals + (size_t)b * n;
    const float* m = d.mask ? d.mask + (size_t)b * n : nullptr;
    auto count = [&](float x, float mv) {
        bool sel = true;
        if (mode == MASK_GT) sel = mv > d.thr;
        else if (mode == MASK_SELF_POS) sel = x > 0.0f;
        if (!sel) return;
        const unsigned k = f2key(x);
        if (PASS == 0) {
            atomicAdd(&sh[0][k >> 21], 1u);
        } else {
            const unsigned hi = PASS == 1 ? k >> 21 : k >> 10;
            const unsigned dig = PASS == 1 ? (k >> 10) & 0x7ffu : k & 0x3ffu;
#pragma unroll
            for (int r = 0; r < SEL_NR; ++r)
                if (!dup[r] && hi == pre[r]) atomicAdd(&sh[r][dig], 1u);
        }
    };
    // 128-bit loads where the image starts on a 16-byte boundary (16 bytes in flight per thread instead of 4: the pass is a
    // stream over the image and was latency-bound at a third of the HBM rate); the order of the counts does not matter
    const bool use_m = mode == MASK_GT;
    const bool al16 = ((reinterpret_cast<uintptr_t>(v) | (use_m ? reinterpret_cast<uintptr_t>(m) : 0)) & 15) == 0;
    const size_t n4 = al16 ? n / 4 : 0, stride = (size_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const float4* v4 = reinterpret_cast<const float4*>(v);
    const float4* m4 = reinterpret_cast<const float4*>(m);
    for (size_t i = t0; i < n4; i += stride) {
        const float4 a = v4[i];
        const float4 q = use_m ? m4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        count(a.x, q.x); count(a.y, q.y); count(a.z, q.z); count(a.w, q.w);
    }
    for (size_t i = 4 * n4 + t0; i < n; i += stride) count(v[i], use_m ? m[i] : 0.f);
    __syncthreads();
    for (int i = threadIdx.x; i < NH * SEL_BINS; i += blockDim.x) {
        const unsigned c = (&sh[0][0])[i];
        if (c) atomicAdd(&(&s->hist[0][0])[i], c);
    }
}

// One block (256 threads) per (selection, image): find each rank's bin in this pass's histogram.
template <int PASS>
__global__ void __launch_bounds__(256) k_sel_pick(SelState* __restrict__ st, const unsigned* __restrict__ relcnt,
                                                  int mode0, int mode1, int B)
{
    __shared__ unsigned sh[SEL_NR][SEL_BINS];
    __shared__ unsigned part[SEL_NR][8];
    __shared__ unsigned tot;
    const int b = blockIdx.x, z = blockIdx.y;
    SelState* s = st + (size_t)z * B + b;
    const int mode = z ? mode1 : mode0;
    constexpr int NH = PASS == 0 ? 1 : SEL_NR;
    for (int i = threadIdx.x; i < NH * SEL_BINS; i += blockDim.x) {
        (&sh[0][0])[i] = (&s->hist[0][0])[i];
        (&s->hist[0][0])[i] = 0;                         // ready for the next pass / next call
    }
    __syncthreads();
    if (PASS == 0) {
        if (threadIdx.x == 0) tot = 0;
        __syncthreads();
        unsigned c = 0;
        for (int i = threadIdx.x; i < SEL_BINS; i += blockDim.x) c += sh[0][i];
        c = __reduce_add_sync(0xffffffffu, c);
        if ((threadIdx.x & 31) == 0) atomicAdd(&tot, c);
        __syncthreads();
        if (threadIdx.x == 0) {
            // np.percentile(a, q), method 'linear': virtual index (n-1)*q/100 in q's dtype
            const unsigned n = tot;
            const bool few = mode == MASK_GT_IF_ENOUGH && relcnt[b] < 100u;
            const bool f64 = mode == MASK_GT_IF_ENOUGH && !few;     // np.percentile(v, [10, 90])
            const float ql = f64 ? 10.f : 5.f, qh = f64 ? 90.f : 95.f;
            s->count = n;
            s->f64 = f64;
            s->few = few;
            for (int q = 0; q < 2; ++q) {
                unsigned lo = 0, hi = 0;
                double g = 0.0;
                if (n > 0) {
                    if (f64) {
                        const double qq = (double)(q ? qh : ql) / 100.0;
                        const double vi = __dmul_rn((double)(n - 1), qq);
                        if (vi >= (double)(n - 1)) lo = hi = n - 1;
                        else { lo = (unsigned)floor(vi); hi = lo + 1; g = vi - floor(vi); }
                    } else {
                        const float qq = __fdiv_rn(q ? qh : ql, 100.0f);
                        const float vi = __fmul_rn((float)(n - 1), qq);
                        if (vi >= (float)(n - 1)) lo = hi = n - 1;
                        else { lo = (unsigned)floorf(vi); hi = lo + 1; g = (double)__fsub_rn(vi, floorf(vi)); }
                    }
                }
                s->rem[2 * q] = lo;
                s->rem[2 * q + 1] = hi;
                s->gam[q] = g;
            }
        }
        __syncthreads();
    }
    // rank r is handled by warp r: lanes own 64 consecutive bins
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp < SEL_NR) {
        const int r = warp;
        int src = r;                                       // histogram of the first rank with the same prefix
        if (PASS > 0)
            for (int q = r - 1; q >= 0; --q)
                if (s->pre[q] == s->pre[r]) src = q;
        const unsigned* h = sh[PASS == 0 ? 0 : src];
        constexpr int PER = SEL_BINS / 32;
        unsigned sum = 0;
        for (int i = 0; i < PER; ++i) sum += h[lane * PER + i];
        unsigned inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        const unsigned exc = inc - sum;
        const unsigned k = s->rem[r];
        const bool mine = k >= exc && k < inc;
        if (mine) {
            unsigned acc = exc;
            int bin = lane * PER;
            for (int i = 0; i < PER; ++i) {
                const unsigned c = h[lane * PER + i];
                if (k < acc + c) { bin = lane * PER + i; break; }
                acc += c;
            }
            part[r][0] = (unsigned)bin;
            part[r][1] = k - acc;
        }
        __syncwarp();
        const unsigned any = __ballot_sync(0xffffffffu, mine);
        if (lane == 0) {
            if (any) {
                const unsigned bin = part[r][0];
                const unsigned pre = PASS == 0 ? bin : (PASS == 1 ? (s->pre[r] << 11) | bin : (s->pre[r] << 10) | bin);
                part[r][2] = pre;
            } else {
                part[r][1] = 0;
                part[r][2] = PASS == 0 ? 0u : (PASS == 1 ? s->pre[r] << 11 : s->pre[r] << 10);   // empty selection
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < SEL_NR) {
        s->pre[threadIdx.x] = part[threadIdx.x][2];
        s->rem[threadIdx.x] = part[threadIdx.x][1];
        if (PASS == 2) s->val[threadIdx.x] = key2f(part[threadIdx.x][2]);
    }
    if (PASS == 2) {
        __syncthreads();
        if (threadIdx.x < 2) {
            // numpy _lerp: a + (b-a)*t, replaced by b - (b-a)*(1-t) where t >= 0.5
            const int q = threadIdx.x;
            const float x = key2f(part[2 * q][2]), y = key2f(part[2 * q + 1][2]);
            const float df = __fsub_rn(y, x);
            double res;
            if (s->f64) {
                const double g = s->gam[q];
                res = g < 0.5 ? __dadd_rn((double)x, __dmul_rn((double)df, g))
                              : __dsub_rn((double)y, __dmul_rn((double)df, __dsub_rn(1.0, g)));
            } else {
                const float g = (float)s->gam[q];
                res = g < 0.5f ? (double)__fadd_rn(x, __fmul_rn(df, g))
                               : (double)__fsub_rn(y, __fmul_rn(df, __fsub_rn(1.0f, g)));
            }
            s->pct[q] = s->count ? res : (double)NAN;
        }
    }
}

static int run_select(const SelDesc& d0, const SelDesc& d1, int B, size_t n, SelState* st, const unsigned* relcnt,
                      cudaStream_t s)
{
    const int nb = (int)min((size_t)svb_ceil_div(148 * 4, 2 * B) + 1, (n + 2047) / 2048);
    dim3 grid(max(nb, 1), B, 2);
    k_sel_hist<0><<<grid, 512, 0, s>>>(d0, d1, n, st, relcnt, B);
    k_sel_pick<0><<<dim3(B, 2), 256, 0, s>>>(st, relcnt, d0.mode, d1.mode, B);
    k_sel_hist<1><<<grid, 512, 0, s>>>(d0, d1, n, st, relcnt, B);
    k_sel_pick<1><<<dim3(B, 2), 256, 0, s>>>(st, relcnt, d0.mode, d1.mode, B);
    k_sel_hist<2><<<grid, 512, 0, s>>>(d0, d1, n, st, relcnt, B);
    k_sel_pick<2><<<dim3(B, 2), 256, 0, s>>>(st, relcnt, d0.mode, d1.mode, B);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ---- range mapping: out = f(src; percentiles).  kind 0: calibrate_midas_to_stereo, 1: normalize_to_stereo_range.
// selection 0 = stereo disparity, selection 1 = the map being calibrated.
__global__ void k_range_apply(const float* __restrict__ src, size_t n, const SelState* __restrict__ st, int B, int kind,
                              float* __restrict__ out)
{
    const int b = blockIdx.y;
    const SelState* ss = st + b;
    const SelState* sm = st + B + b;
    const float* p = src + (size_t)b * n;
    float* o = out + (size_t)b * n;
    int mode;                 // 0 constant, 1 float32 min/max normalisation, 2 float64 affine
    float cst = 0.f, mmin = 0.f, den = 1.f, smin = 0.f, srng = 0.f;
    double scale = 1.0, offset = 0.0;
    if (kind == 0 && !ss->few) {
        const double s0 = ss->pct[0], s1 = ss->pct[1], m0 = sm->pct[0], m1 = sm->pct[1];
        scale = (m1 - m0) < 1e-6 ? 1.0 : __ddiv_rn(s1 - s0, __dadd_rn(m1 - m0, 1e-8));
        offset = __dsub_rn(s0, __dmul_rn(m0, scale));
        mode = 2;
    } else {
        float s0 = (float)ss->pct[0], s1 = (float)ss->pct[1];
        if (kind == 1 && ss->count == 0) { s0 = 0.f; s1 = 255.f; }
        const float m0 = (float)sm->pct[0], m1 = (float)sm->pct[1];
        const float dm = __fsub_rn(m1, m0);
        if (dm < 1e-6f) {
            cst = __fdiv_rn(__fadd_rn(s0, s1), 2.0f);
            mode = 0;
        } else {
            mmin = m0;
            den = __fadd_rn(dm, 1e-8f);
            smin = s0;
            srng = __fsub_rn(s1, s0);
            mode = 1;
        }
    }
    auto map = [&](float x) {
        if (mode == 2) return (float)__dadd_rn(__dmul_rn((double)x, scale), offset);
        if (mode == 1) return __fadd_rn(smin, __fmul_rn(__fdiv_rn(__fsub_rn(x, mmin), den), srng));
        return cst;
    };
    const bool al16 = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(o)) & 15) == 0;
    const size_t n4 = al16 ? n / 4 : 0, stride = (size_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = t0; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(p)[i];
        reinterpret_cast<float4*>(o)[i] = make_float4(map(a.x), map(a.y), map(a.z), map(a.w));
    }
    for (size_t i = 4 * n4 + t0; i < n; i += stride) o[i] = map(p[i]);
}

extern "C" size_t svb_range_scratch_bytes(int B) { return (size_t)2 * B * sizeof(SelState) + (size_t)B * sizeof(unsigned) + 64; }

static int range_map(int kind, const float* src, const float* stereo, const float* conf, int B, int H, int W, float* out,
                     void* scratch, size_t scratch_bytes, void* stream)
{
    if (!src || !stereo || !out || !scratch || B < 1 || H < 1 || W < 1 || (kind == 0 && !conf)) {
        svb_set_error("svb range mapping: bad argument");
        return SVB_ERR_PARAM;
    }
    if (scratch_bytes < svb_range_scratch_bytes(B)) {
        svb_set_error("svb range mapping: scratch too small (%zu < %zu)", scratch_bytes, svb_range_scratch_bytes(B));
        return SVB_ERR_PARAM;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n = (size_t)H * W;
    SelState* st = (SelState*)scratch;
    unsigned* relcnt = (unsigned*)((char*)scratch + (size_t)2 * B * sizeof(SelState));
    SVB_CUDA(cudaMemsetAsync(scratch, 0, svb_range_scratch_bytes(B), s));
    const int nb = (int)min((size_t)svb_ceil_div(148 * 8, B), (n + 1023) / 1024);
    SelDesc d0, d1;
    if (kind == 0) {
        k_count_gt<<<dim3(nb, B), 256, 0, s>>>(conf, n, 0.7f, relcnt);
        d0 = SelDesc{stereo, conf, MASK_GT_IF_ENOUGH, 0.7f};
        d1 = SelDesc{src, conf, MASK_GT_IF_ENOUGH, 0.7f};
    } else {
        d0 = SelDesc{stereo, nullptr, MASK_SELF_POS, 0.f};
        d1 = SelDesc{src, nullptr, MASK_NONE, 0.f};
    }
    int rc = run_select(d0, d1, B, n, st, relcnt, s);
    if (rc != SVB_OK) return rc;
    k_range_apply<<<dim3(nb, B), 256, 0, s>>>(src, n, st, B, kind, out);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_calibrate_midas_dev(const float* midas_dev, const float* stereo_disp_dev, const float* stereo_conf_dev,
                                       int B, int H, int W, float* out_dev, void* scratch_dev, size_t scratch_bytes,
                                       void* stream)
{
    return range_map(0, midas_dev, stereo_disp_dev, stereo_conf_dev, B, H, W, out_dev, scratch_dev, scratch_bytes, stream);
}

extern "C" int svb_normalize_to_stereo_range_dev(const float* depth_dev, const float* stereo_disp_dev, int B, int H, int W,
                                                 float* out_dev, void* scratch_dev, size_t scratch_bytes, void* stream)
{
    return range_map(1, depth_dev, stereo_disp_dev, nullptr, B, H, W, out_dev, scratch_dev, scratch_bytes, stream);
}

// np.percentile(vals[mask], q_lo/q_hi) on its own (tests, and callers that want the numbers).
extern "C" int svb_percentiles_f32_dev(const float* vals_dev, const float* mask_dev, float mask_thr, int mask_mode, int B,
                                       size_t n, double* out_dev /*[B][2]: 5/95 */, unsigned* count_dev,
                                       void* scratch_dev, size_t scratch_bytes, void* stream);

__global__ void k_sel_export(const SelState* __restrict__ st, int B, double* __restrict__ out, unsigned* __restrict__ cnt)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    out[2 * b] = st[b].pct[0];
    out[2 * b + 1] = st[b].pct[1];
    if (cnt) cnt[b] = st[b].count;
}

extern "C" int svb_percentiles_f32_dev(const float* vals_dev, const float* mask_dev, float mask_thr, int mask_mode, int B,
                                       size_t n, double* out_dev, unsigned* count_dev, void* scratch_dev,
                                       size_t scratch_bytes, void* stream)
{
    if (!vals_dev || !out_dev || !scratch_dev || B < 1 || n < 1 || mask_mode < 0 || mask_mode > 3 ||
        ((mask_mode == MASK_GT || mask_mode == MASK_GT_IF_ENOUGH) && !mask_dev)) {
        svb_set_error("svb_percentiles_f32_dev: bad argument");
        return SVB_ERR_PARAM;
    }
    if (scratch_bytes < svb_range_scratch_bytes(B)) {
        svb_set_error("svb_percentiles_f32_dev: scratch too small");
        return SVB_ERR_PARAM;
    }
    cudaStream_t s = (cudaStream_t)stream;
    SelState* st = (SelState*)scratch_dev;
    unsigned* relcnt = (unsigned*)((char*)scratch_dev + (size_t)2 * B * sizeof(SelState));
    SVB_CUDA(cudaMemsetAsync(scratch_dev, 0, svb_range_scratch_bytes(B), s));
    if (mask_mode == MASK_GT_IF_ENOUGH) {
        const int nb = (int)min((size_t)svb_ceil_div(148 * 8, B), (n + 1023) / 1024);
        k_count_gt<<<dim3(nb, B), 256, 0, s>>>(mask_dev, n, mask_thr, relcnt);
    }
    SelDesc d{vals_dev, mask_dev, mask_mode, mask_thr};
    int rc = run_select(d, d, B, n, st, relcnt, s);
    if (rc != SVB_OK) return rc;
    k_sel_export<<<svb_ceil_div(B, 128), 128, 0, s>>>(st, B, out_dev, count_dev);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------------------------------
// a16: fuse_depth_maps.  Kernels:
//   k_fill_rows   (1 - conf) * w_midas, horizontal Gaussian taps (BORDER_REFLECT_101) -> tmp
//   k_blend       vertical taps + clip -> fill weight; stereo*w, MiDaS fill, flow fill -> fused f32,
//                 per-image min/max (ordered-int atomics)
//   k_bilat_lut   cv2's 4096-bin exp table for this image's value range
//   k_bilateral   circular 9x9 window, table lookups with linear interpolation, clip -> u8
// ------------------------------------------------------------------------------------------
#define FUSE_MAXK 31
#define FUSE_MAXD 15
#define LUT_BINS 4096

struct GaussTaps { int k; float w[FUSE_MAXK]; };
struct BilatTaps { int d; float gc; float sw[FUSE_MAXD * FUSE_MAXD]; };   // sw < 0: outside the circle

struct FuseState {          // per image
    unsigned kmin, kmax;    // ordered keys of min / max of the fused map
    float scale;            // LUT_BINS / (max - min)
    int filter;             // run the bilateral filter (max > 10 and a non-degenerate range)
    float lut[LUT_BINS + 2];
};

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}

// Horizontal pass of the confidence blur.  A block owns FR_T consecutive columns of one row: it stages v = (1 - conf) * wm for
// them and their k - 1 neighbours (reflect-101 applied here, once per staged value) in shared memory; every thread then runs
// its k taps from there in cv2's order.  K: the kernel size as a compile-time constant (0 = g.k): taps unrolled, weights as
// constant-bank operands (instances for 15 -- the reference's midas_blend_radius is used as the kernel size -- and 31).
#define FR_T 128
template <int K>
__global__ void __launch_bounds__(FR_T) k_fill_rows(const float* __restrict__ conf, int H, int W, float wm, GaussTaps g,
                                                    float* __restrict__ tmp)
{
    __shared__ float sv[FR_T + FUSE_MAXK - 1];
    const int k = K ? K : g.k, r = k / 2;
    const int b = blockIdx.z, y = blockIdx.y, x0 = blockIdx.x * FR_T;
    const float* row = conf + ((size_t)b * H + y) * W;
    for (int i = threadIdx.x; i < FR_T + 2 * r; i += FR_T) {
        const int xx = min(x0 - r + i, W - 1 + r);           // (columns past the row's last tap are never used)
        sv[i] = __fmul_rn(__fsub_rn(1.0f, row[reflect101(xx, W)]), wm);
    }
    __syncthreads();
    const int x = x0 + threadIdx.x;
    if (x >= W) return;
    float s = __fmul_rn(g.w[0], sv[threadIdx.x]);
#pragma unroll
    for (int i = 1; i < (K ? K : k); ++i) s = __fadd_rn(s, __fmul_rn(g.w[i], sv[threadIdx.x + i]));
    tmp[((size_t)b * H + y) * W + x] = s;
}

// Per-image running min / max (ordered keys) from a whole block.  The target only moves one way, so a value that cannot
// improve the current one (a plain L2 read, possibly stale -- safe for a monotone target) skips its atomic.  ONE thread per
// block does that read: with one per warp, the half a million warps of a call all read the same two words of one L2 slice
// and every block's last step queued behind them.  Call from all threads of the block.
__device__ __forceinline__ void block_minmax_update(unsigned kmin, unsigned kmax, unsigned* gmin, unsigned* gmax)
{
    __shared__ unsigned s_mn[32], s_mx[32];
    const int tid = (threadIdx.z * blockDim.y + threadIdx.y) * blockDim.x + threadIdx.x;
    const int nw = (blockDim.x * blockDim.y * blockDim.z + 31) >> 5;
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    if ((tid & 31) == 0) { s_mn[tid >> 5] = kmin; s_mx[tid >> 5] = kmax; }
    __syncthreads();
    if (tid < 32) {
        kmin = __reduce_min_sync(0xffffffffu, tid < nw ? s_mn[tid] : 0xffffffffu);
        kmax = __reduce_max_sync(0xffffffffu, tid < nw ? s_mx[tid] : 0u);
        if (tid == 0) {
            if (kmin < __ldcg(gmin)) atomicMin(gmin, kmin);
            if (kmax > __ldcg(gmax)) atomicMax(gmax, kmax);
        }
    }
}

struct BlendArgs {
    const float* stereo; const float* conf; const float* midas; const float* flow; const float* tmp;
    float ws, thr_conf, hole, wf, wf1;   // wf1 = (float)(1 - w_flow)
    int base;                            // 0 stereo, 1 midas, 2 flow
    int midas_fill, flow_fill;
};

// K: as in k_fill_rows; rows whose window lies inside the image walk down the column without the reflection
template <int K>
__global__ void k_blend(BlendArgs a, int H, int W, GaussTaps g, float* __restrict__ fused, FuseState* __restrict__ fs)
{
    const int b = blockIdx.z, y = blockIdx.y, x = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned kmin = 0xffffffffu, kmax = 0;
    if (x < W) {
        const size_t i = ((size_t)b * H + y) * W + x;
        float f;
        if (a.base == 0) {
            f = __fmul_rn(a.stereo[i], a.ws);
            if (a.midas_fill) {
                const float c = a.conf ? a.conf[i] : 1.0f;
                if (c < a.thr_conf) {
                    const int k = K ? K : g.k, r = k / 2;
                    float s;
                    if (y - r >= 0 && y + r < H) {
                        const float* col = a.tmp + ((size_t)b * H + (y - r)) * W + x;
                        s = __fmul_rn(g.w[0], col[0]);
#pragma unroll
                        for (int t = 1; t < (K ? K : k); ++t) s = __fadd_rn(s, __fmul_rn(g.w[t], col[(size_t)t * W]));
                    } else {
                        s = 0.f;
                        for (int t = 0; t < k; ++t) {
                            const float v = a.tmp[((size_t)b * H + reflect101(y - r + t, H)) * W + x];
                            s = t == 0 ? __fmul_rn(g.w[0], v) : __fadd_rn(s, __fmul_rn(g.w[t], v));
                        }
                    }
                    const float fw = fminf(fmaxf(s, 0.0f), 1.0f);
                    if (fw > 0.1f)
                        f = __fadd_rn(__fmul_rn(f, __fsub_rn(1.0f, fw)), __fmul_rn(a.midas[i], fw));
                }
            }
        } else if (a.base == 1) {
            f = a.midas[i];
        } else {
            f = a.flow[i];
        }
        if (a.base != 2 && a.flow_fill && (f < a.hole || f == 0.0f))
            f = __fadd_rn(__fmul_rn(f, a.wf1), __fmul_rn(a.flow[i], a.wf));
        fused[i] = f;
        kmin = kmax = f2key(f);
    }
    // (half a million same-address atomics per call had made this kernel latency-bound: ncu 8.7 % issue-active at 410 GB/s)
    block_minmax_update(kmin, kmax, &fs[b].kmin, &fs[b].kmax);
}

__global__ void k_fuse_init(FuseState* fs, int B)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) { fs[b].kmin = 0xffffffffu; fs[b].kmax = 0; }
}

__global__ void k_bilat_lut(FuseState* __restrict__ fs, float gc, float min_max /* filter only when max > this */)
{
    FuseState* s = fs + blockIdx.x;
    const float mn = key2f(s->kmin), mx = key2f(s->kmax);
    const float len = __fsub_rn(mx, mn);
    const bool filter = mx > min_max && len > FLT_EPSILON;
    const float scale = filter ? __fdiv_rn((float)LUT_BINS, len) : 0.f;
    __syncthreads();
    if (threadIdx.x == 0) { s->filter = filter; s->scale = scale; }
    if (!filter) return;
    // cv2 stops evaluating exp() once it has underflowed to 0; exp is monotone here, so the
    // entries after the first zero are zero anyway.
    for (int i = threadIdx.x; i < LUT_BINS + 2; i += blockDim.x) {
        const double v = (double)i / (double)scale;
        s->lut[i] = (float)exp(v * v * (double)gc);
    }
}

#define BT_X 32
#define BT_Y 8
// DD: the window diameter as a compile-time constant (0 = bt.d): both loops unroll, the circle test folds away and every tap is
// one shared-memory load at an immediate offset and one constant-bank operand -- same taps in the same order
template <int DD>
__global__ void __launch_bounds__(BT_X * BT_Y) k_bilateral(const float* __restrict__ fused, int H, int W, BilatTaps bt,
                                                          const FuseState* __restrict__ fs, unsigned char* __restrict__ out8,
                                                          float* __restrict__ outf, int clip255)
{
    __shared__ float tile[BT_Y + FUSE_MAXD - 1][BT_X + FUSE_MAXD - 1];
    __shared__ float lut[LUT_BINS + 2];
    const int b = blockIdx.z;
    const FuseState* s = fs + b;
    const int d = DD ? DD : bt.d, r = d / 2, tw = BT_X + 2 * r, th = BT_Y + 2 * r;
    const int x0 = blockIdx.x * BT_X, y0 = blockIdx.y * BT_Y;
    const int tid = threadIdx.y * BT_X + threadIdx.x;
    const bool filter = s->filter != 0;
    const float* img = fused + (size_t)b * H * W;
    if (filter) {
        for (int i = tid; i < LUT_BINS + 2; i += BT_X * BT_Y) lut[i] = s->lut[i];
        for (int i = tid; i < tw * th; i += BT_X * BT_Y) {
            const int ty = i / tw, tx = i - ty * tw;
            tile[ty][tx] = img[(size_t)reflect101(y0 - r + ty, H) * W + reflect101(x0 - r + tx, W)];
        }
    }
    __syncthreads();
    const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
    if (x >= W || y >= H) return;
    float res;
    if (filter) {
        const float c = tile[threadIdx.y + r][threadIdx.x + r];
        const float scale = s->scale;
        float sum = 0.f, wsum = 0.f;
#pragma unroll
        for (int i = 0; i < (DD ? DD : d); ++i)
#pragma unroll
            for (int j = 0; j < (DD ? DD : d); ++j) {
                if (DD && (i - DD / 2) * (i - DD / 2) + (j - DD / 2) * (j - DD / 2) > (DD / 2) * (DD / 2)) continue;   // make_bilat_taps' circle
                const float sw = bt.sw[i * d + j];
                if (!DD && sw < 0.f) continue;
                const float v = tile[threadIdx.y + i][threadIdx.x + j];
                float al = __fmul_rn(fabsf(__fsub_rn(v, c)), scale);
                const int idx = (int)al;
                al = __fsub_rn(al, (float)idx);
                const float l0 = lut[idx], l1 = lut[idx + 1];
                const float w = __fmul_rn(sw, __fadd_rn(l0, __fmul_rn(al, __fsub_rn(l1, l0))));
                sum = __fmaf_rn(v, w, sum);
                wsum = __fadd_rn(wsum, w);
            }
        res = __fdiv_rn(sum, wsum);
    } else {
        res = img[(size_t)y * W + x];
    }
    if (clip255) res = fminf(fmaxf(res, 0.0f), 255.0f);
    const size_t o = ((size_t)b * H + y) * W + x;
    if (out8) out8[o] = (unsigned char)(int)res;
    if (outf) outf[o] = res;
}

extern "C" size_t svb_fuse_scratch_bytes(int B, int H, int W)
{
    return (size_t)2 * B * H * W * sizeof(float) + (size_t)B * sizeof(FuseState) + 256;
}

extern "C" int svb_fuse_depth_maps_dev(const float* stereo_dev, const float* conf_dev, const float* midas_dev,
                                       const float* flow_dev, int B, int H, int W, const svb_fuse_params* p,
                                       uint8_t* out_u8_dev, float* out_f32_dev, void* scratch_dev, size_t scratch_bytes,
                                       void* stream)
{
    if (!p || !out_u8_dev || !scratch_dev || B < 1 || H < 1 || W < 1 || (!stereo_dev && !midas_dev && !flow_dev)) {
        svb_set_error("svb_fuse_depth_maps_dev: bad argument (at least one of stereo/midas/flow is needed)");
        return SVB_ERR_PARAM;
    }
    if (p->blend_ksize < 1 || p->blend_ksize > FUSE_MAXK || !(p->blend_ksize & 1) || p->bilateral_d < 1 ||
        p->bilateral_d > FUSE_MAXD || !(p->bilateral_d & 1)) {
        svb_set_error("svb_fuse_depth_maps_dev: blend_ksize must be odd <= %d, bilateral_d odd <= %d", FUSE_MAXK, FUSE_MAXD);
        return SVB_ERR_UNSUPPORTED;
    }
    if (scratch_bytes < svb_fuse_scratch_bytes(B, H, W)) {
        svb_set_error("svb_fuse_depth_maps_dev: scratch too small (%zu < %zu)", scratch_bytes, svb_fuse_scratch_bytes(B, H, W));
        return SVB_ERR_PARAM;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n = (size_t)B * H * W;
    float* tmp = (float*)scratch_dev;
    float* fused = tmp + n;
    FuseState* fs = (FuseState*)(((uintptr_t)(fused + n) + 127) & ~(uintptr_t)127);

    // cv2.getGaussianKernel(k, sigma <= 0): sigma = 0.3*((k-1)*0.5 - 1) + 0.8, normalised, as float
    GaussTaps g;
    g.k = p->blend_ksize;
    {
        const double sigma = 0.3 * ((g.k - 1) * 0.5 - 1) + 0.8;
        double kd[FUSE_MAXK], sum = 0;
        for (int i = 0; i < g.k; ++i) {
            const double x = i - (g.k - 1) * 0.5;
            kd[i] = exp(-x * x / (2 * sigma * sigma));
            sum += kd[i];
        }
        for (int i = 0; i < g.k; ++i) g.w[i] = (float)(kd[i] / sum);
    }
    BilatTaps bt;
    bt.d = p->bilateral_d;
    {
        const double sc = p->bilateral_sigma_color > 0 ? p->bilateral_sigma_color : 1.0;
        const double ss = p->bilateral_sigma_space > 0 ? p->bilateral_sigma_space : 1.0;
        bt.gc = (float)(-0.5 / (sc * sc));
        const double gs = -0.5 / (ss * ss);
        const int r = bt.d / 2;
        for (int i = -r; i <= r; ++i)
            for (int j = -r; j <= r; ++j) {
                const double rr = sqrt((double)i * i + (double)j * j);
                bt.sw[(i + r) * bt.d + (j + r)] = rr > r ? -1.f : (float)exp(rr * rr * gs);
            }
    }
    BlendArgs a;
    a.stereo = stereo_dev; a.conf = conf_dev; a.midas = midas_dev; a.flow = flow_dev; a.tmp = tmp;
    a.ws = (float)p->stereo_weight; a.thr_conf = (float)p->stereo_conf_threshold; a.hole = (float)p->flow_hole_threshold;
    a.wf = (float)p->flow_fill_weight; a.wf1 = (float)(1.0 - p->flow_fill_weight);    // NumPy: array * (1 - w), w a Python float
    a.base = stereo_dev ? 0 : (midas_dev ? 1 : 2);
    a.midas_fill = stereo_dev && midas_dev && p->use_midas_fill;
    a.flow_fill = flow_dev && p->use_flow_fill && a.base != 2;

    k_fuse_init<<<svb_ceil_div(B, 128), 128, 0, s>>>(fs, B);
    dim3 grid(svb_ceil_div(W, 128), H, B);
    if (a.midas_fill) {
        const float* c = conf_dev;
        if (c && g.k == 15) k_fill_rows<15><<<grid, FR_T, 0, s>>>(c, H, W, (float)p->midas_fill_weight, g, tmp);
        else if (c && g.k == 31) k_fill_rows<31><<<grid, FR_T, 0, s>>>(c, H, W, (float)p->midas_fill_weight, g, tmp);
        else if (c) k_fill_rows<0><<<grid, FR_T, 0, s>>>(c, H, W, (float)p->midas_fill_weight, g, tmp);
        else SVB_CUDA(cudaMemsetAsync(tmp, 0, n * sizeof(float), s));      // conf == 1 everywhere -> weight 0
    }
    if (g.k == 15) k_blend<15><<<grid, 128, 0, s>>>(a, H, W, g, fused, fs);
    else if (g.k == 31) k_blend<31><<<grid, 128, 0, s>>>(a, H, W, g, fused, fs);
    else k_blend<0><<<grid, 128, 0, s>>>(a, H, W, g, fused, fs);
    k_bilat_lut<<<B, 256, 0, s>>>(fs, bt.gc, 10.0f);
    dim3 bgrid(svb_ceil_div(W, BT_X), svb_ceil_div(H, BT_Y), B);
    if (bt.d == 9) k_bilateral<9><<<bgrid, dim3(BT_X, BT_Y), 0, s>>>(fused, H, W, bt, fs, out_u8_dev, out_f32_dev, 1);
    else k_bilateral<0><<<bgrid, dim3(BT_X, BT_Y), 0, s>>>(fused, H, W, bt, fs, out_u8_dev, out_f32_dev, 1);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------------------------------
// a17: the five image statistics of detect_camera_occlusion (fused_depth_map.py:170-240).
//   stats[b] = { mean block std, low-variance block ratio, global std, histogram entropy, mean }
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_occ_blocks(const unsigned char* __restrict__ gray, int H, int W, int block, int bh,
                                                    int bw, double* __restrict__ stds)
{
    const int b = blockIdx.z, bi = blockIdx.y, bj = blockIdx.x;
    const int y1 = bi * block, y2 = min((bi + 1) * block, H), x1 = bj * block, x2 = min((bj + 1) * block, W);
    const int bwid = x2 - x1, cnt = (y2 - y1) * bwid;
    const unsigned char* img = gray + (size_t)b * H * W;
    unsigned long long s1 = 0, s2 = 0;
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
        const int yy = i / bwid, xx = i - yy * bwid;
        const unsigned v = img[(size_t)(y1 + yy) * W + x1 + xx];
        s1 += v;
        s2 += v * v;
    }
    __shared__ unsigned long long r1[8], r2[8];
    for (int o = 16; o; o >>= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    if ((threadIdx.x & 31) == 0) { r1[threadIdx.x >> 5] = s1; r2[threadIdx.x >> 5] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        s1 = s2 = 0;
        for (int i = 0; i < 8; ++i) { s1 += r1[i]; s2 += r2[i]; }
        // var = (n*S2 - S1^2) / n^2, exact integer numerator
        const double num = (double)((unsigned long long)cnt * s2 - s1 * s1);
        stds[((size_t)b * bh + bi) * bw + bj] = cnt ? sqrt(num / ((double)cnt * (double)cnt)) : 0.0;
    }
}

__global__ void __launch_bounds__(256) k_occ_hist(const unsigned char* __restrict__ gray, size_t n, unsigned* __restrict__ hist)
{
    __shared__ unsigned sh[8][256];                          // one histogram per warp: an eighth of the same-bin collisions
    const int b = blockIdx.y, wp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 8 * 256; i += 256) (&sh[0][0])[i] = 0;
    __syncthreads();
    const unsigned char* p = gray + (size_t)b * n;
    const size_t n16 = (reinterpret_cast<uintptr_t>(p) & 15) == 0 ? n / 16 : 0, stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = t0; i < n16; i += stride) {              // 16 pixels per load
        const uint4 v = reinterpret_cast<const uint4*>(p)[i];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            atomicAdd(&sh[wp][w[q] & 0xffu], 1u);
            atomicAdd(&sh[wp][(w[q] >> 8) & 0xffu], 1u);
            atomicAdd(&sh[wp][(w[q] >> 16) & 0xffu], 1u);
            atomicAdd(&sh[wp][w[q] >> 24], 1u);
        }
    }
    for (size_t i = 16 * n16 + t0; i < n; i += stride) atomicAdd(&sh[wp][p[i]], 1u);
    __syncthreads();
    unsigned c = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) c += sh[q][threadIdx.x];
    if (c) atomicAdd(&hist[b * 256 + threadIdx.x], c);
}

__device__ double block_sum_256(double v, double* red)
{
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0;
    for (int i = 0; i < 8; ++i) t += red[i];
    return t;
}

__global__ void __launch_bounds__(256) k_occ_final(const double* __restrict__ stds, int nblk, const unsigned* __restrict__ hist,
                                                   double n, double* __restrict__ stats)
{
    __shared__ double red[8];
    const int b = blockIdx.x;
    double a = 0, l = 0;
    for (int i = threadIdx.x; i < nblk; i += 256) {
        const double s = stds[(size_t)b * nblk + i];
        a += s;
        l += s < 12.0;
    }
    a = block_sum_256(a, red);
    l = block_sum_256(l, red);
    const unsigned h = hist[b * 256 + threadIdx.x];
    const double v = (double)threadIdx.x;
    const double mean = block_sum_256(v * h, red) / n;
    const double var = block_sum_256((v - mean) * (v - mean) * h, red) / n;
    // hist (float32, as cv2.calcHist returns) + 1e-10, normalised, Shannon entropy with log2(p + 1e-10)
    const float hf = __fadd_rn((float)h, 1e-10f);
    const float tot = (float)block_sum_256((double)hf, red);
    const float pr = __fdiv_rn(hf, tot);
    const double ent = -block_sum_256((double)__fmul_rn(pr, log2f(__fadd_rn(pr, 1e-10f))), red);
    if (threadIdx.x == 0) {
        double* o = stats + 5 * b;
        o[0] = nblk ? a / nblk : 0.0;
        o[1] = l / (nblk > 0 ? nblk : 1);
        o[2] = sqrt(var);
        o[3] = ent;
        o[4] = mean;
    }
}

extern "C" size_t svb_occlusion_scratch_bytes(int B, int H, int W, int block)
{
    if (block < 1) block = 48;
    const int bh = H / block > 1 ? H / block : 1, bw = W / block > 1 ? W / block : 1;
    return (size_t)B * bh * bw * sizeof(double) + (size_t)B * 256 * sizeof(unsigned) + 64;
}

extern "C" int svb_occlusion_stats_u8_dev(const uint8_t* gray_dev, int B, int H, int W, int block, double* stats_dev,
                                          void* scratch_dev, size_t scratch_bytes, void* stream)
{
    if (!gray_dev || !stats_dev || !scratch_dev || B < 1 || H < 1 || W < 1 || block < 1) {
        svb_set_error("svb_occlusion_stats_u8_dev: bad argument");
        return SVB_ERR_PARAM;
    }
    if (scratch_bytes < svb_occlusion_scratch_bytes(B, H, W, block)) {
        svb_set_error("svb_occlusion_stats_u8_dev: scratch too small");
        return SVB_ERR_PARAM;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int bh = H / block > 1 ? H / block : 1, bw = W / block > 1 ? W / block : 1;
    double* stds = (double*)scratch_dev;
    unsigned* hist = (unsigned*)(stds + (size_t)B * bh * bw);
    SVB_CUDA(cudaMemsetAsync(hist, 0, (size_t)B * 256 * sizeof(unsigned), s));
    k_occ_blocks<<<dim3(bw, bh, B), 256, 0, s>>>(gray_dev, H, W, block, bh, bw, stds);
    const size_t n = (size_t)H * W;
    const int nb = (int)min((size_t)svb_ceil_div(148 * 8, B), (n + 4095) / 4096);
    k_occ_hist<<<dim3(max(nb, 1), B), 256, 0, s>>>(gray_dev, n, hist);
    k_occ_final<<<B, 256, 0, s>>>(stds, bh * bw, hist, (double)n, stats_dev);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// ------------------------------------------------------------------------------------------
// the data-parallel parts around the mono estimators.
//   MiDaS post-processing (fused_depth_map.py:1126-1158): bicubic upsample of the network output
//   (torch.nn.functional.interpolate, mode="bicubic", align_corners=False), min-max normalisation to
//   0..255, contrast confidence = cv2.normalize(absdiff(gray, GaussianBlur(gray, 5x5)), 0, 1, MINMAX)
//   flow depth (fused_depth_map.py:1386-1403, 1492-1498): expected flow of the ego-motion homography,
//   residual magnitude, 1 / (r + 0.5), exponential average, bilateral 9/75/75
// ------------------------------------------------------------------------------------------
static void make_bilat_taps(BilatTaps& bt, int d, double sigma_color, double sigma_space)
{
    bt.d = d;
    const double sc = sigma_color > 0 ? sigma_color : 1.0, ss = sigma_space > 0 ? sigma_space : 1.0;
    bt.gc = (float)(-0.5 / (sc * sc));
    const double gs = -0.5 / (ss * ss);
    const int r = d / 2;
    for (int i = -r; i <= r; ++i)
        for (int j = -r; j <= r; ++j) {
            const double rr = sqrt((double)i * i + (double)j * j);
            bt.sw[(i + r) * d + (j + r)] = rr > r ? -1.f : (float)exp(rr * rr * gs);
        }
}

__global__ void k_minmax_f32(const float* __restrict__ src, size_t n, FuseState* __restrict__ fs)
{
    const int b = blockIdx.y;
    const float* p = src + (size_t)b * n;
    unsigned kmin = 0xffffffffu, kmax = 0;
    const size_t n4 = (reinterpret_cast<uintptr_t>(p) & 15) == 0 ? n / 4 : 0, stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = t0; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(p)[i];
        const unsigned k0 = f2key(a.x), k1 = f2key(a.y), k2 = f2key(a.z), k3 = f2key(a.w);
        kmin = min(min(kmin, k0), min(k1, min(k2, k3)));
        kmax = max(max(kmax, k0), max(k1, max(k2, k3)));
    }
    for (size_t i = 4 * n4 + t0; i < n; i += stride) {
        const unsigned k = f2key(p[i]);
        kmin = min(kmin, k);
        kmax = max(kmax, k);
    }
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    if ((threadIdx.x & 31) == 0) {
        if (kmin != 0xffffffffu && kmin < __ldcg(&fs[b].kmin)) atomicMin(&fs[b].kmin, kmin);
        if (kmax && kmax > __ldcg(&fs[b].kmax)) atomicMax(&fs[b].kmax, kmax);
    }
}

extern "C" size_t svb_bilateral_scratch_bytes(int B) { return (size_t)B * sizeof(FuseState) + 256; }

/* cv2.bilateralFilter(src f32, d, sigmaColor, sigmaSpace) (fused_depth_map.py:1400, 1690) */
extern "C" int svb_bilateral_f32_dev(const float* src_dev, int B, int H, int W, int d, double sigma_color,
                                     double sigma_space, float* dst_dev, void* scratch_dev, size_t scratch_bytes,
                                     void* stream)
{
    if (!src_dev || !dst_dev || !scratch_dev || B < 1 || H < 1 || W < 1) {
        svb_set_error("svb_bilateral_f32_dev: bad argument");
        return SVB_ERR_PARAM;
    }
    if (d < 1 || d > FUSE_MAXD || !(d & 1)) {
        svb_set_error("svb_bilateral_f32_dev: d must be odd and <= %d", FUSE_MAXD);
        return SVB_ERR_UNSUPPORTED;
    }
    if (scratch_bytes < svb_bilateral_scratch_bytes(B)) { svb_set_error("svb_bilateral_f32_dev: scratch too small"); return SVB_ERR_PARAM; }
    cudaStream_t s = (cudaStream_t)stream;
    FuseState* fs = (FuseState*)(((uintptr_t)scratch_dev + 127) & ~(uintptr_t)127);
    BilatTaps bt;
    make_bilat_taps(bt, d, sigma_color, sigma_space);
    const size_t n = (size_t)H * W;
    k_fuse_init<<<svb_ceil_div(B, 128), 128, 0, s>>>(fs, B);
    const int nb = (int)min((size_t)svb_ceil_div(148 * 8, B), (n + 1023) / 1024);
    k_minmax_f32<<<dim3(max(nb, 1), B), 256, 0, s>>>(src_dev, n, fs);
    k_bilat_lut<<<B, 256, 0, s>>>(fs, bt.gc, -FLT_MAX);
    const dim3 bgrid(svb_ceil_div(W, BT_X), svb_ceil_div(H, BT_Y), B);
    if (d == 9) k_bilateral<9><<<bgrid, dim3(BT_X, BT_Y), 0, s>>>(src_dev, H, W, bt, fs, nullptr, dst_dev, 0);
    else k_bilateral<0><<<bgrid, dim3(BT_X, BT_Y), 0, s>>>(src_dev, H, W, bt, fs, nullptr, dst_dev, 0);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// torch's upsample_bicubic2d (A = -0.75, align_corners = False): source index (dst + 0.5) * scale - 0.5,
// taps clamped to the image, x taps accumulated first, then y
__device__ __forceinline__ void cubic_coeffs(float t, float (&w)[4])
{
    const float A = -0.75f;
    float x = t + 1.0f;
    w[0] = ((A * x - 5.0f * A) * x + 8.0f * A) * x - 4.0f * A;
    w[1] = ((A + 2.0f) * t - (A + 3.0f)) * t * t + 1.0f;
    x = 1.0f - t;
    w[2] = ((A + 2.0f) * x - (A + 3.0f)) * x * x + 1.0f;
    x = 2.0f - t;
    w[3] = ((A * x - 5.0f * A) * x + 8.0f * A) * x - 4.0f * A;
}

__global__ void __launch_bounds__(256) k_resize_bicubic(const float* __restrict__ src, int Hi, int Wi, float* __restrict__ dst,
                                                        int Ho, int Wo, float sy, float sx, FuseState* __restrict__ fs)
{
    // A block writes 256 columns of one output row: its taps come from four source rows and, when upsampling, a few dozen source
    // columns -- staged (already clamped) in shared memory, so that the sixteen taps of a pixel are shared-memory loads instead
    // of dependent gathers from L2.  Wider source ranges (downsampling) keep the direct path.
    constexpr int TW = 320;
    __shared__ float tile[4][TW];
    const int x0 = blockIdx.x * blockDim.x, x = x0 + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    const float* img = src + (size_t)b * Hi * Wi;
    const float fy = __fsub_rn(__fmul_rn((float)y + 0.5f, sy), 0.5f);
    const float fly = floorf(fy);
    const int iy = (int)fly;
    const int xl = min(x0 + (int)blockDim.x, Wo) - 1;            // last output column of the block
    const int first = (int)floorf(__fsub_rn(__fmul_rn((float)x0 + 0.5f, sx), 0.5f)) - 1;
    const int ncols = (int)floorf(__fsub_rn(__fmul_rn((float)xl + 0.5f, sx), 0.5f)) + 2 - first + 1;
    const bool staged = ncols <= TW;
    if (staged) {
        for (int i = threadIdx.x; i < 4 * ncols; i += blockDim.x) {
            const int a = i / ncols, j = i - a * ncols;
            tile[a][j] = img[(size_t)min(max(iy - 1 + a, 0), Hi - 1) * Wi + min(max(first + j, 0), Wi - 1)];
        }
        __syncthreads();
    }
    unsigned kmin = 0xffffffffu, kmax = 0;
    if (x < Wo) {
        const float fx = __fsub_rn(__fmul_rn((float)x + 0.5f, sx), 0.5f);
        const float flx = floorf(fx);
        const int ix = (int)flx;
        float wy[4], wx[4];
        cubic_coeffs(fy - fly, wy);
        cubic_coeffs(fx - flx, wx);
        float acc = 0.f;
        if (staged) {
            const int j0 = ix - 1 - first;
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                float r = 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) r += tile[a][j0 + c] * wx[c];
                acc += r * wy[a];
            }
        } else {
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const float* row = img + (size_t)min(max(iy - 1 + a, 0), Hi - 1) * Wi;
                float r = 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) r += row[min(max(ix - 1 + c, 0), Wi - 1)] * wx[c];
                acc += r * wy[a];
            }
        }
        dst[((size_t)b * Ho + y) * Wo + x] = acc;
        kmin = kmax = f2key(acc);
    }
    if (fs) block_minmax_update(kmin, kmax, &fs[b].kmin, &fs[b].kmax);
}

// (x - min) / (max - min + 1e-8) * 255 in float32, zeros when max - min < 1e-6 (fused_depth_map.py:1138-1145)
__global__ void k_minmax_scale255(float* __restrict__ data, size_t n, const FuseState* __restrict__ fs)
{
    const int b = blockIdx.y;
    const float mn = key2f(fs[b].kmin), mx = key2f(fs[b].kmax);
    const float len = __fsub_rn(mx, mn);
    const bool flat = len < 1e-6f;
    const float den = __fadd_rn(len, 1e-8f);
    float* p = data + (size_t)b * n;
    auto sc = [&](float v) { return flat ? 0.0f : __fmul_rn(__fdiv_rn(__fsub_rn(v, mn), den), 255.0f); };
    const size_t n4 = (reinterpret_cast<uintptr_t>(p) & 15) == 0 ? n / 4 : 0, stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = t0; i < n4; i += stride) {
        const float4 a = reinterpret_cast<float4*>(p)[i];
        reinterpret_cast<float4*>(p)[i] = make_float4(sc(a.x), sc(a.y), sc(a.z), sc(a.w));
    }
    for (size_t i = 4 * n4 + t0; i < n; i += stride) p[i] = sc(p[i]);
}

extern "C" size_t svb_midas_scratch_bytes(int B) { return (size_t)B * sizeof(FuseState) + 256; }

extern "C" int svb_midas_upsample_normalize_dev(const float* pred_dev, int B, int Hi, int Wi, float* out_dev, int Ho, int Wo,
                                                void* scratch_dev, size_t scratch_bytes, void* stream)
{
    if (!pred_dev || !out_dev || !scratch_dev || B < 1 || Hi < 1 || Wi < 1 || Ho < 1 || Wo < 1) {
        svb_set_error("svb_midas_upsample_normalize_dev: bad argument");
        return SVB_ERR_PARAM;
    }
    if (scratch_bytes < svb_midas_scratch_bytes(B)) { svb_set_error("svb_midas_upsample_normalize_dev: scratch too small"); return SVB_ERR_PARAM; }
    cudaStream_t s = (cudaStream_t)stream;
    FuseState* fs = (FuseState*)(((uintptr_t)scratch_dev + 127) & ~(uintptr_t)127);
    k_fuse_init<<<svb_ceil_div(B, 128), 128, 0, s>>>(fs, B);
    k_resize_bicubic<<<dim3(svb_ceil_div(Wo, 256), Ho, B), 256, 0, s>>>(pred_dev, Hi, Wi, out_dev, Ho, Wo,
                                                                        (float)Hi / (float)Ho, (float)Wi / (float)Wo, fs);
    const size_t n = (size_t)Ho * Wo;
    const int nb = (int)min((size_t)svb_ceil_div(148 * 8, B), (n + 1023) / 1024);
    k_minmax_scale255<<<dim3(max(nb, 1), B), 256, 0, s>>>(out_dev, n, fs);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// contrast confidence (fused_depth_map.py:1150-1156): cv2.GaussianBlur(u8, 5x5, sigma 0) is the fixed-point
// binomial filter (sum w_i w_j p + 128) >> 8 with w = {1,4,6,4,1} and BORDER_REFLECT_101
// Four pixels per thread: the tile starts four columns left of the block so that every thread reads its 12-byte window as three
// aligned words; the horizontal taps of a row are two dp4a per pixel on byte windows cut out with PRMT, the contrast bytes leave
// as one 32-bit store.  Interior blocks fill the tile with 32-bit loads, border blocks byte by byte through the reflection.
constexpr int CC_X = 128, CC_Y = 8, CC_TW = CC_X / 4 + 2;     // tile: CC_Y + 4 rows of CC_TW words (pixels x0 - 4 .. x0 + CC_X + 3)
__global__ void __launch_bounds__(CC_X / 4 * CC_Y) k_contrast_u8(const unsigned char* __restrict__ gray, int H, int W,
                                                                 unsigned char* __restrict__ contrast, FuseState* __restrict__ fs)
{
    __shared__ uint32_t tile[CC_Y + 4][CC_TW];
    const int b = blockIdx.z, x0 = blockIdx.x * CC_X, y0 = blockIdx.y * CC_Y;
    const unsigned char* img = gray + (size_t)b * H * W;
    const int tid = threadIdx.y * (CC_X / 4) + threadIdx.x;
    const bool words = (W & 3) == 0 && (reinterpret_cast<uintptr_t>(img) & 3) == 0 && x0 >= 4 && x0 + CC_X + 4 <= W &&
                       y0 >= 2 && y0 + CC_Y + 2 <= H;
    for (int i = tid; i < (CC_Y + 4) * CC_TW; i += CC_X / 4 * CC_Y) {
        const int ty = i / CC_TW, tw = i - ty * CC_TW;
        uint32_t v;
        if (words) {
            v = *reinterpret_cast<const uint32_t*>(img + (size_t)(y0 + ty - 2) * W + x0 - 4 + 4 * tw);
        } else {
            const unsigned char* row = img + (size_t)reflect101(y0 + ty - 2, H) * W;
            v = 0;
#pragma unroll
            for (int q = 0; q < 4; ++q) v |= (uint32_t)row[reflect101(min(x0 - 4 + 4 * tw + q, W + 1), W)] << (8 * q);
        }
        tile[ty][tw] = v;
    }
    __syncthreads();
    const int x = x0 + 4 * threadIdx.x, y = y0 + threadIdx.y;
    unsigned cmin = 255, cmax = 0;
    if (x < W && y < H) {
        unsigned s[4] = {0u, 0u, 0u, 0u};
        uint32_t centre = 0;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const uint32_t a = tile[threadIdx.y + i][threadIdx.x], m = tile[threadIdx.y + i][threadIdx.x + 1],
                           c = tile[threadIdx.y + i][threadIdx.x + 2];
            // byte k of win[j] = pixel x + j - 2 + k
            const uint32_t win[5] = {__byte_perm(a, m, 0x5432), __byte_perm(a, m, 0x6543), m, __byte_perm(m, c, 0x4321),
                                     __byte_perm(m, c, 0x5432)};
            const unsigned wv = i == 0 || i == 4 ? 1u : i == 2 ? 6u : 4u;
#pragma unroll
            for (int j = 0; j < 4; ++j) s[j] += wv * __dp4a(win[j + 1], 0x01000000u, __dp4a(win[j], 0x04060401u, 0u));
            if (i == 2) centre = m;
        }
        uint32_t out = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int blur = (int)((s[j] + 128u) >> 8), g = (int)((centre >> (8 * j)) & 0xffu);
            const unsigned c = (unsigned)abs(g - blur);
            out |= c << (8 * j);
            if (x + j < W) { cmin = min(cmin, c); cmax = max(cmax, c); }
        }
        unsigned char* dst = contrast + ((size_t)b * H + y) * W + x;
        if (x + 3 < W && (reinterpret_cast<uintptr_t>(dst) & 3) == 0) *reinterpret_cast<uint32_t*>(dst) = out;
        else
            for (int j = 0; j < 4 && x + j < W; ++j) dst[j] = (unsigned char)(out >> (8 * j));
    }
    block_minmax_update(cmin, cmax, &fs[b].kmin, &fs[b].kmax);
}

// cv2.normalize(c, None, 0, 1, NORM_MINMAX) on CV_8U: scale = 1 / (max - min) (0 for a flat image),
// shift = -min * scale, result = saturate_u8(rint(c * scale + shift)) evaluated in float32
__global__ void k_contrast_norm(unsigned char* __restrict__ data, size_t n, const FuseState* __restrict__ fs)
{
    const int b = blockIdx.y;
    const int mn = (int)fs[b].kmin, mx = (int)fs[b].kmax;
    const double scale = mx > mn ? 1.0 / (double)(mx - mn) : 0.0;
    const float fsc = (float)scale, fsh = (float)(0.0 - (double)mn * scale);
    __shared__ unsigned char lut[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x)
        lut[i] = (unsigned char)min(max(__float2int_rn(__fadd_rn(__fmul_rn((float)i, fsc), fsh)), 0), 255);
    __syncthreads();
    unsigned char* p = data + (size_t)b * n;
    const size_t n16 = (reinterpret_cast<uintptr_t>(p) & 15) == 0 ? n / 16 : 0, stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = t0; i < n16; i += stride) {             // 16 pixels per thread and access
        uint4 v = reinterpret_cast<uint4*>(p)[i];
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; ++q)
            w[q] = (uint32_t)lut[w[q] & 0xffu] | ((uint32_t)lut[(w[q] >> 8) & 0xffu] << 8) |
                   ((uint32_t)lut[(w[q] >> 16) & 0xffu] << 16) | ((uint32_t)lut[w[q] >> 24] << 24);
        reinterpret_cast<uint4*>(p)[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    for (size_t i = 16 * n16 + t0; i < n; i += stride) p[i] = lut[p[i]];
}

extern "C" int svb_contrast_confidence_u8_dev(const uint8_t* gray_dev, int B, int H, int W, uint8_t* conf_dev, void* scratch_dev,
                                              size_t scratch_bytes, void* stream)
{
    if (!gray_dev || !conf_dev || !scratch_dev || B < 1 || H < 1 || W < 1) {
        svb_set_error("svb_contrast_confidence_u8_dev: bad argument");
        return SVB_ERR_PARAM;
    }
    if (scratch_bytes < svb_midas_scratch_bytes(B)) { svb_set_error("svb_contrast_confidence_u8_dev: scratch too small"); return SVB_ERR_PARAM; }
    cudaStream_t s = (cudaStream_t)stream;
    FuseState* fs = (FuseState*)(((uintptr_t)scratch_dev + 127) & ~(uintptr_t)127);
    k_fuse_init<<<svb_ceil_div(B, 128), 128, 0, s>>>(fs, B);
    k_contrast_u8<<<dim3(svb_ceil_div(W, CC_X), svb_ceil_div(H, CC_Y), B), dim3(CC_X / 4, CC_Y), 0, s>>>(gray_dev, H, W, conf_dev, fs);
    const size_t n = (size_t)H * W;
    const int nb = (int)min((size_t)svb_ceil_div(148 * 8, B), (n + 1023) / 1024);
    k_contrast_norm<<<dim3(max(nb, 1), B), 256, 0, s>>>(conf_dev, n, fs);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// flow depth (fused_depth_map.py:1386-1398 with the expected flow of :1491-1497): cv2.perspectiveTransform
// evaluates the homography in double on float32 coordinates and rounds the warped point to float32
struct Homography { double m[9]; };
__global__ void k_flow_residual_depth(const float2* __restrict__ flow, int H, int W, Homography hm, float* __restrict__ current,
                                      float* __restrict__ stable, int stable_valid, float a_keep, float a_new)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const double dx = (double)x, dy = (double)y;
    const double* m = hm.m;
    double w = __dadd_rn(__dadd_rn(__dmul_rn(dx, m[6]), __dmul_rn(dy, m[7])), m[8]);
    float ex = 0.f, ey = 0.f;
    if (fabs(w) > 2.220446049250313e-16) {
        w = __ddiv_rn(1.0, w);
        ex = (float)__dmul_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, m[0]), __dmul_rn(dy, m[1])), m[2]), w);
        ey = (float)__dmul_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, m[3]), __dmul_rn(dy, m[4])), m[5]), w);
    }
    ex = __fsub_rn(ex, (float)x);
    ey = __fsub_rn(ey, (float)y);
    const size_t o = (size_t)y * W + x;
    const float2 f = flow[o];
    const float rx = __fsub_rn(f.x, ex), ry = __fsub_rn(f.y, ey);
    const float mag = (float)__dsqrt_rn(__dadd_rn(__dmul_rn((double)rx, (double)rx), __dmul_rn((double)ry, (double)ry)));
    const float cur = __fdiv_rn(1.0f, __fadd_rn(mag, 0.5f));
    current[o] = cur;
    if (stable) stable[o] = stable_valid ? __fadd_rn(__fmul_rn(a_keep, stable[o]), __fmul_rn(a_new, cur)) : cur;
}

extern "C" int svb_flow_residual_depth_dev(const float* flow_dev, int B, int H, int W, const double* homography9, float* current_dev,
                                           float* stable_dev, int stable_valid, double alpha, void* stream)
{
    if (!flow_dev || !homography9 || !current_dev || B < 1 || H < 1 || W < 1) {
        svb_set_error("svb_flow_residual_depth_dev: bad argument");
        return SVB_ERR_PARAM;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n = (size_t)H * W;
    for (int b = 0; b < B; ++b) {
        Homography hm;
        for (int i = 0; i < 9; ++i) hm.m[i] = homography9[b * 9 + i];
        k_flow_residual_depth<<<dim3(svb_ceil_div(W, 256), H), 256, 0, s>>>(
            (const float2*)flow_dev + b * n, H, W, hm, current_dev + b * n, stable_dev ? stable_dev + b * n : nullptr, stable_valid,
            (float)alpha, (float)(1.0 - alpha));
    }
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
