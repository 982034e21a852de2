// bm.cu -- StereoBM (cv2.StereoBM_create(numDisparities, blockSize).compute; BASELINE.json
// config 2) for sm_100a, bit-exact on int16 output.
//
//   k_bm_prefilter  X-Sobel prefilter (cap-clipped), right image also as paired u16x2 planes
//   k_bm_tex_h/_v   texture test + ROI as a byte mask of the pixels that need a decision
//   k_bm2<NP>       a CTA walks down a strip of 64 window columns: vertical SAD sums in registers (entering row
//                   added, leaving row recomputed and subtracted), one barrier per row, horizontal box by
//                   sliding along each warp's run of output columns, argmin / uniqueness in place, subpixel
//                   division batched 32 pixels at a time.  The SAD volume never exists in HBM.
//                   (k_bm v1 -- one warp per row sliding along x, two block barriers per column -- ran at
//                   5 968 pairs/s on BASELINE config 2; this one at 12 492.)
//   speckle filter  (imgproc.cu) when speckleWindowSize > 0
#include "svb_common.cuh"
#include <algorithm>
#include <new>
#include <thread>
#include <vector>
#include <string.h>

struct BmGeo {
    int H, W, D, NP, minD, bs, r, cap, tex, uniq, FILTERED;
    int lofs, rofs, width1, Wh;
    int pft, pfs;              // prefilter type (0 normalized response, 1 x-Sobel) and window
};

// PREFILTER_NORMALIZED_RESPONSE, pass 1: vertical sums of preFilterSize rows (replicate borders) of both images
__global__ void __launch_bounds__(256) k_bm_vsum(const uint8_t* __restrict__ left, const uint8_t* __restrict__ right,
                                                 BmGeo g, uint16_t* __restrict__ vs)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    const int b = blockIdx.z >> 1, side = blockIdx.z & 1;
    if (x >= g.W) return;
    const size_t io = (size_t)b * g.H * g.W;
    const uint8_t* img = (side ? right : left) + io;
    const int w2 = g.pfs / 2;
    int s = 0;
    for (int j = -w2; j <= w2; ++j) s += img[(size_t)min(max(y + j, 0), g.H - 1) * g.W + x];
    vs[((size_t)side * gridDim.z / 2) * g.H * g.W + io + (size_t)y * g.W + x] = (uint16_t)s;   // <= 255 * 255
}

__global__ void __launch_bounds__(256) k_bm_prefilter(const uint8_t* __restrict__ left, const uint8_t* __restrict__ right,
                                                      BmGeo g, uint8_t* __restrict__ lp, uint32_t* __restrict__ rdesc,
                                                      const uint16_t* __restrict__ vs)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    const int H = g.H, W = g.W, cap = g.cap;
    if (x >= W) return;
    // stereobm.cpp prefilterNorm: ((4c + l + r + u + d) * scale_g - boxsum * scale_s) >> 10, clipped to [0, 2 cap]
    int scale_g = g.pfs * g.pfs / 8;
    const int scale_s = (1024 + scale_g) / (scale_g * 2);
    scale_g *= scale_s;
    const size_t io = (size_t)b * H * W;
    auto pf = [&](const uint8_t* img, int xx) -> int {
        if (g.pft == 0) {
            const uint16_t* v = vs + (img == left + io ? 0 : (size_t)gridDim.z * H * W) + io + (size_t)y * W;
            const int w2 = g.pfs / 2;
            int sum = 0;
            for (int i = -w2; i <= w2; ++i) sum += v[min(max(xx + i, 0), W - 1)];
            const uint8_t* cr = img + (size_t)y * W;
            const int up = img[(size_t)max(y - 1, 0) * W + xx], dn = img[(size_t)min(y + 1, H - 1) * W + xx];
            const int val = ((cr[xx] * 4 + cr[max(xx - 1, 0)] + cr[min(xx + 1, W - 1)] + up + dn) * scale_g - sum * scale_s) >> 10;
            return val < -cap ? 0 : (val > cap ? 2 * cap : val + cap);
        }
        if (xx <= 0 || xx >= W - 1) return cap;
        if ((H & 1) && y == H - 1) return cap;
        const int ya = y > 0 ? y - 1 : (H > 1 ? 1 : 0), yb = y < H - 1 ? y + 1 : (H > 1 ? H - 2 : 0);
        const uint8_t *r0 = img + (size_t)ya * W, *r1 = img + (size_t)y * W, *r2 = img + (size_t)yb * W;
        const int gx = ((int)r0[xx + 1] - r0[xx - 1]) + 2 * ((int)r1[xx + 1] - r1[xx - 1]) + ((int)r2[xx + 1] - r2[xx - 1]);
        return max(-cap, min(cap, gx)) + cap;
    };
    lp[io + (size_t)y * W + x] = (uint8_t)pf(left + io, x);
    const int v0 = pf(right + io, x), v1 = pf(right + io, max(x - 1, 0));
    uint32_t* base = rdesc + (size_t)(b * H + y) * 4 * g.Wh + (size_t)(x & 1) * g.Wh + (x >> 1);
    base[0] = (uint32_t)v0 | ((uint32_t)v1 << 16);
    base[2 * g.Wh] = (uint32_t)((-v0) & 0xffff) | ((uint32_t)((-v1) & 0xffff) << 16);
}

// |u - v| per half: max(u - v, v - u) with the negated operands precomputed
__device__ __forceinline__ uint32_t bm_absdiff(uint32_t U2, uint32_t nU2, uint32_t V2, uint32_t nV2)
{
    return __vmaxs2(__vadd2(U2, nV2), __vadd2(V2, nU2));
}

// ------------------------------------------------------------------------------------------
// k_bm_texture: T(y, x) = sum over the bs x bs window of |L - cap| (A.6 texture test), columns clamped like
// the SAD window's left columns.  Two tiny passes (horizontal box -> u16, vertical box -> s32).
// ------------------------------------------------------------------------------------------
// Horizontal pass: a block stages |L - cap| (<= 63) of TEX_P * 256 columns + the window's reach as bytes in shared memory; a thread
// sums the window of its first column and slides it over the next TEX_P - 1 (21 byte loads per four columns instead of 60 clamped
// global ones).
#define TEX_P 4
__global__ void __launch_bounds__(256) k_bm_tex_h(const uint8_t* __restrict__ lp, BmGeo g, uint16_t* __restrict__ th)
{
    __shared__ uint8_t a[256 * TEX_P + 32];
    const int x0 = blockIdx.x * 256 * TEX_P, y = blockIdx.y, b = blockIdx.z, r = g.r;
    const uint8_t* row = lp + ((size_t)b * g.H + y) * g.W;
    for (int i = threadIdx.x; i < 256 * TEX_P + 2 * r; i += 256)
        a[i] = (uint8_t)abs((int)row[min(max(x0 - r + i, 0), g.W - 1)] - g.cap);
    __syncthreads();
    const int xl = threadIdx.x * TEX_P, x = x0 + xl;
    if (x >= g.W) return;
    int s = 0;
    for (int j = 0; j <= 2 * r; ++j) s += a[xl + j];
    uint16_t o[TEX_P];
    o[0] = (uint16_t)s;                                      // <= 31 * 63
#pragma unroll
    for (int q = 1; q < TEX_P; ++q) {
        s += (int)a[xl + q + 2 * r] - (int)a[xl + q - 1];
        o[q] = (uint16_t)s;
    }
    uint16_t* dst = th + ((size_t)b * g.H + y) * g.W + x;
    if (x + TEX_P <= g.W && (reinterpret_cast<uintptr_t>(dst) & 7) == 0)
        *reinterpret_cast<uint2*>(dst) = make_uint2((uint32_t)o[0] | ((uint32_t)o[1] << 16), (uint32_t)o[2] | ((uint32_t)o[3] << 16));
    else
        for (int q = 0; q < TEX_P && x + q < g.W; ++q) dst[q] = o[q];
}
// ok(y, x) = 1 where k_bm2 has to decide a pixel: inside the rows / columns cv2 evaluates and textured enough.
// all_cols: disp12MaxDiff >= 0 -- every column of the width1 domain is evaluated (cv2 lets the columns outside
// the valid ROI vote in validateDisparity and blanks them afterwards).
// A thread walks TEXV_R rows of its column and slides the vertical window (2 loads per row after the first).
#define TEXV_R 8
__global__ void __launch_bounds__(256) k_bm_tex_v(const uint16_t* __restrict__ th, BmGeo g, int all_cols,
                                                  uint8_t* __restrict__ ok)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y0 = blockIdx.y * TEXV_R, b = blockIdx.z, r = g.r;
    if (x >= g.W) return;
    bool oc = x >= g.lofs && x < g.lofs + g.width1;
    if (!all_cols) oc = oc && x >= g.lofs + r && x < g.W - r;
    const uint16_t* col = th + (size_t)b * g.H * g.W + x;
    uint8_t* oo = ok + (size_t)b * g.H * g.W + x;
    int s = 0;
    bool have = false;                                       // s holds the window sum of the previous row
    for (int y = y0; y < min(y0 + TEXV_R, g.H); ++y) {
        bool o = oc && y >= r && y < g.H - r;
        if (o) {
            if (have) s += (int)col[(size_t)(y + r) * g.W] - (int)col[(size_t)(y - r - 1) * g.W];
            else {
                s = 0;
                for (int i = -r; i <= r; ++i) s += col[(size_t)(y + i) * g.W];
                have = true;
            }
            o = s >= g.tex;
        } else have = false;
        oo[(size_t)y * g.W] = o ? 1 : 0;
    }
}

// ------------------------------------------------------------------------------------------
// k_bm2<NP>: a CTA walks DOWN a strip of BM2_PC window columns of one image (a row segment of it).
//   vertical phase: warp w owns 4 window columns; their vertical SAD sums V(c, d) over the last bs rows live in
//     registers: per row  V += |L - R|(row v) - |L - R|(row v - bs)  (the leaving row is recomputed: three
//     packed ops and two cached loads beat a 100 KB shared-memory ring), then V goes to a double-buffered
//     shared row buffer;
//   ONE block barrier per row;
//   horizontal phase: warp w owns a run of output columns: S(oc) = sum of V over 2r+1 columns, slid along the
//     run; texture test from k_bm_texture's map; argmin / uniqueness / subpixel in place.
// The SAD volume never exists in HBM; no per-column barriers (the old k_bm had two per column step).
// ------------------------------------------------------------------------------------------
#define BM2_WARPS 16
#define BM2_CPW 4                          // window columns per warp
#define BM2_PC (BM2_WARPS * BM2_CPW)       // window columns per strip

// subpixel disparity of one parked pixel (A.6), written by the lane that holds it
__device__ __forceinline__ void bm2_finish(uint32_t key, uint32_t pn, uint32_t off, int D, int minD,
                                           int16_t* __restrict__ disp, int16_t* __restrict__ cost)
{
    const int minsad = (int)(key >> 16), mind = (int)(key & 0xffffu);
    const int p = (int)(pn & 0xffffu), n = (int)(pn >> 16);
    const int dd = p + n - 2 * minsad + abs(p - n);
    const int val = ((D - mind - 1 + minD) * 256 + (dd != 0 ? (p - n) * 256 / dd : 0) + 15) >> 4;
    disp[off] = (int16_t)val;
    if (cost) cost[off] = (int16_t)minsad;
}

template <int NP>
__global__ void __launch_bounds__(BM2_WARPS * 32) k_bm2(const uint8_t* __restrict__ lp, const uint32_t* __restrict__ rdesc,
                                                        const uint8_t* __restrict__ okmap, BmGeo g, int OC, int nstrips,
                                                        int rseg, int16_t* __restrict__ disp, int16_t* __restrict__ cost)
{
    constexpr int VW = NP * 32;
    extern __shared__ uint32_t smem[];
    uint32_t* vbuf = smem;                                   // [2][BM2_PC][VW] vertical sums of the current row
    uint32_t* scr = smem + 2 * BM2_PC * VW;                  // [BM2_WARPS][VW] per-warp scratch for the subpixel taps
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int strip = blockIdx.x % nstrips, seg = blockIdx.x / nstrips, b = blockIdx.z;
    const int r = g.r, bs = 2 * r + 1, D = g.D, W = g.W, H = g.H;
    const int ya = r + seg * rseg, yb = min(ya + rseg, H - r);   // output rows of this CTA
    if (ya >= yb) return;
    const int xs = strip * OC;                               // first output column (width1 domain)

    // row-invariant addressing of this warp's window columns (A.6: left column and right start clamp separately)
    int lc[BM2_CPW], pl[BM2_CPW], m[BM2_CPW][NP];
#pragma unroll
    for (int q = 0; q < BM2_CPW; ++q) {
        const int j = xs - r + BM2_CPW * warp + q;
        lc[q] = min(max(j + g.lofs, 0), W - 1);
        const int rs = min(max(j + g.rofs, 0), W - D);
        const int xr0 = rs + D - 1;                          // right column of disparity index d' = 0
        pl[q] = (xr0 & 1) * g.Wh;
#pragma unroll
        for (int k = 0; k < NP; ++k) m[q][k] = max((xr0 >> 1) - (32 * k + lane), 0);
    }
    uint32_t V[BM2_CPW][NP];
#pragma unroll
    for (int q = 0; q < BM2_CPW; ++q)
#pragma unroll
        for (int k = 0; k < NP; ++k) V[q][k] = 0;

    // per-lane constants of the decision: BM indices kk of this lane's two disparities, pad flags
    uint32_t kkpk[NP], padlo[NP], padhi[NP];
#pragma unroll
    for (int k = 0; k < NP; ++k) {
        const int d0 = 2 * (32 * k + lane);
        kkpk[k] = (uint32_t)((D - 1 - d0) & 0xffff) | ((uint32_t)((D - 2 - d0) & 0xffff) << 16);
        padlo[k] = d0 < D ? 0u : 0xffffffffu;
        padhi[k] = d0 + 1 < D ? 0u : 0xffffffffu;
    }
    uint32_t rkey = 0, rpn = 0, roff = 0;                    // parked pixel of this lane
    int nrec = 0;
    const int R = (OC + BM2_WARPS - 1) / BM2_WARPS;          // output columns per warp (<= BM2_CPW)
    const int oc_lo = warp * R, oc_hi = min(oc_lo + R, OC);
    const uint8_t* limg = lp + (size_t)b * H * W;
    const uint32_t* rimg = rdesc + (size_t)b * H * 4 * g.Wh;

    const int v0 = ya - r, v1 = yb - 1 + r;                  // input rows: all inside the image
    for (int v = v0; v <= v1; ++v) {
        const int it = v - v0, buf = it & 1;
        // ---- vertical phase
        {
            const uint8_t* lrow = limg + (size_t)v * W;
            const uint32_t* rrow = rimg + (size_t)v * 4 * g.Wh;
#pragma unroll
            for (int q = 0; q < BM2_CPW; ++q) {
                const int u = lrow[lc[q]];
                const uint32_t U2 = pk_dup((uint32_t)u), nU2 = pk_dup((uint32_t)(-u));
                const uint32_t* rp = rrow + pl[q];
#pragma unroll
                for (int k = 0; k < NP; ++k) V[q][k] += bm_absdiff(U2, nU2, __ldg(rp + m[q][k]), __ldg(rp + 2 * g.Wh + m[q][k]));
            }
            if (it >= bs) {
                const uint8_t* lold = lrow - (size_t)bs * W;
                const uint32_t* rold = rrow - (size_t)bs * 4 * g.Wh;
#pragma unroll
                for (int q = 0; q < BM2_CPW; ++q) {
                    const int u = lold[lc[q]];
                    const uint32_t U2 = pk_dup((uint32_t)u), nU2 = pk_dup((uint32_t)(-u));
                    const uint32_t* rp = rold + pl[q];
#pragma unroll
                    for (int k = 0; k < NP; ++k) V[q][k] -= bm_absdiff(U2, nU2, __ldg(rp + m[q][k]), __ldg(rp + 2 * g.Wh + m[q][k]));
                }
            }
            uint32_t* vb = vbuf + ((size_t)buf * BM2_PC + BM2_CPW * warp) * VW;
#pragma unroll
            for (int q = 0; q < BM2_CPW; ++q)
#pragma unroll
                for (int k = 0; k < NP; ++k) vb[q * VW + 32 * k + lane] = V[q][k];
        }
        __syncthreads();
        if (it < bs - 1) continue;                           // window not complete yet
        // ---- horizontal phase: output row y, this warp's run of output columns
        const int y = v - r;
        const uint32_t* vb = vbuf + (size_t)buf * BM2_PC * VW;
        const uint8_t* okrow = okmap + ((size_t)b * H + y) * W;
        uint32_t s[NP];
        for (int oc = oc_lo; oc < oc_hi; ++oc) {
            if (oc == oc_lo) {
#pragma unroll
                for (int k = 0; k < NP; ++k) s[k] = 0;
                for (int c = oc; c < oc + bs; ++c)
#pragma unroll
                    for (int k = 0; k < NP; ++k) s[k] += vb[c * VW + 32 * k + lane];
            } else {
#pragma unroll
                for (int k = 0; k < NP; ++k) s[k] += vb[(oc + bs - 1) * VW + 32 * k + lane] - vb[(oc - 1) * VW + 32 * k + lane];
            }
            const int x = xs + oc + g.lofs;
            if (x >= W) break;
            if (!okrow[x]) continue;
            // argmin over BM index kk = D-1-d' (lowest kk = largest disparity wins ties): keys (sad << 16 | kk)
            uint32_t key = 0xffffffffu;
#pragma unroll
            for (int k = 0; k < NP; ++k) {
                const uint32_t klo = __byte_perm(s[k], kkpk[k], 0x1054), khi = __byte_perm(s[k], kkpk[k], 0x3276);
                key = min(key, min(klo | padlo[k], khi | padhi[k]));
            }
            key = __reduce_min_sync(0xffffffffu, key);
            const int minsad = (int)(key >> 16), mind = (int)(key & 0xffffu);
            // the SADs at mind-1 / mind+1 (replicated at the ends of the range): subpixel taps, and the two entries
            // the uniqueness test leaves out
            uint32_t* sv = scr + (size_t)warp * VW;
#pragma unroll
            for (int k = 0; k < NP; ++k) sv[32 * k + lane] = s[k];
            __syncwarp();
            const uint16_t* sd = reinterpret_cast<const uint16_t*>(sv);         // indexed by d' = D-1-kk
            const bool hasp = mind + 1 < D, hasn = mind >= 1;
            const int p = sd[hasp ? D - 2 - mind : 1], n = sd[hasn ? D - mind : D - 2];   // sad[mind+1], sad[mind-1]
            __syncwarp();
            if (g.uniq > 0) {
                // cv2: reject iff some kk outside mind-1..mind+1 has sad <= thresh; count all sad <= thresh instead
                const int thresh = minsad + (minsad * g.uniq / 100);
                int cnt = 0;
#pragma unroll
                for (int k = 0; k < NP; ++k)
                    cnt += ((int)(s[k] & 0xffffu) <= thresh && !padlo[k]) + ((int)(s[k] >> 16) <= thresh && !padhi[k]);
                cnt = __reduce_add_sync(0xffffffffu, cnt);
                if (cnt > 1 + (hasp && p <= thresh) + (hasn && n <= thresh)) continue;
            }
            // park the pixel in lane nrec; 32 parked pixels are finished together (one division per lane)
            if (lane == nrec) { rkey = key; rpn = (uint32_t)p | ((uint32_t)n << 16); roff = (uint32_t)(y * W + x); }
            if (++nrec == 32) {
                bm2_finish(rkey, rpn, roff, D, g.minD, disp + (size_t)b * H * W, cost ? cost + (size_t)b * H * W : nullptr);
                nrec = 0;
            }
        }
    }
    if (lane < nrec)
        bm2_finish(rkey, rpn, roff, D, g.minD, disp + (size_t)b * H * W, cost ? cost + (size_t)b * H * W : nullptr);
}

// ------------------------------------------------------------------------------------------
// cv2.validateDisparity (calib3d; StereoBM calls it when disp12MaxDiff >= 0): per row, the right-view map
// is a scatter-min of the cost at x2 = x - round(d/16) (lowest x wins ties: key = cost | x | d), then a
// pixel is dropped iff BOTH roundings of its disparity hit right-view entries that differ by more than
// disp12MaxDiff*16.  Afterwards the columns outside [keep_lo, keep_hi) are blanked (StereoBM's ROI).
// ------------------------------------------------------------------------------------------
template <typename CT>
__global__ void __launch_bounds__(256) k_validate_scatter(const int16_t* __restrict__ disp, const CT* __restrict__ cost,
                                                          int H, int W, int minX1, int maxX1, int inv,
                                                          unsigned long long* __restrict__ keys)
{
    const int x = minX1 + blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= maxX1) return;
    const size_t row = ((size_t)blockIdx.z * H + blockIdx.y) * W;
    const int d = disp[row + x];
    if (d == inv) return;
    const int x2 = x - ((d + 8) >> 4);
    if (x2 < 0 || x2 >= W) return;                           // cannot happen for d inside [minD, maxD)
    const unsigned long long key = ((unsigned long long)((uint32_t)(int)cost[row + x] ^ 0x80000000u) << 32) |
                                   ((unsigned long long)(uint32_t)x << 16) | (uint16_t)d;
    atomicMin(keys + row + x2, key);
}

__global__ void __launch_bounds__(256) k_validate_check(const int16_t* __restrict__ src, int16_t* __restrict__ dst,
                                                        int H, int W, int minX1, int maxX1, int inv, int d12,
                                                        int keep_lo, int keep_hi,
                                                        const unsigned long long* __restrict__ keys)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= W) return;
    const size_t row = ((size_t)blockIdx.z * H + blockIdx.y) * W;
    int d = src[row + x];
    if (x >= minX1 && x < maxX1 && d != inv) {
        const int x0 = x - (d >> 4), x1 = x - ((d + 15) >> 4);
        auto bad = [&](int xx) {
            if (xx < 0 || xx >= W) return false;
            const unsigned long long k = keys[row + xx];
            if (k == ~0ull) return false;                    // nothing landed there
            const int d2 = (int16_t)(uint16_t)k;
            return d2 > inv && abs(d2 - d) > d12;
        };
        if (bad(x0) && bad(x1)) d = inv;
    }
    if (x < keep_lo || x >= keep_hi) d = inv;
    dst[row + x] = (int16_t)d;
}

// disp s16 [B,H,W] in place; cost s16 (cost_is_s32 == 0) or s32 [B,H,W]; scratch: B*H*W*8 bytes (keys) + B*H*W*2
// bytes (a copy of the input, so that the check does not read what it overwrites).
extern "C" int svb_validate_disparity_dev(int16_t* disp, const void* cost, int cost_is_s32, int B, int H, int W,
                                          int minDisparity, int numberOfDisparities, int disp12MaxDiff,
                                          void* scratch, void* stream)
{
    if (!disp || !cost || !scratch) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    if (B <= 0 || H <= 0 || W <= 0 || W > 65535) { svb_set_error("bad shape"); return SVB_ERR_PARAM; }
    cudaStream_t st = (cudaStream_t)stream;
    const int maxD = minDisparity + numberOfDisparities;
    const int minX1 = std::max(maxD, 0), maxX1 = W + std::min(minDisparity, 0);
    const int inv = (minDisparity - 1) * 16;
    const size_t npix = (size_t)B * H * W;
    unsigned long long* keys = (unsigned long long*)scratch;
    int16_t* copy = (int16_t*)(keys + npix);
    SVB_CUDA(cudaMemsetAsync(keys, 0xff, npix * 8, st));
    SVB_CUDA(cudaMemcpyAsync(copy, disp, npix * 2, cudaMemcpyDeviceToDevice, st));
    if (maxX1 > minX1) {
        dim3 grid(svb_ceil_div(maxX1 - minX1, 256), H, B);
        if (cost_is_s32) k_validate_scatter<int><<<grid, 256, 0, st>>>(copy, (const int*)cost, H, W, minX1, maxX1, inv, keys);
        else k_validate_scatter<int16_t><<<grid, 256, 0, st>>>(copy, (const int16_t*)cost, H, W, minX1, maxX1, inv, keys);
        SVB_LAUNCH_CHECK();
    }
    k_validate_check<<<dim3(svb_ceil_div(W, 256), H, B), 256, 0, st>>>(copy, disp, H, W, minX1, maxX1, inv,
                                                                       disp12MaxDiff * 16, 0, W, keys);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

// minDisparity > 0: the last minDisparity positions of cv2's width1 domain lie past the row end (left column
// lofs + xp >= W).  cv2 evaluates them like any other position (window columns clamp to the image) and its store
// lands in the first minDisparity columns of the NEXT row; inside the ROI rows the column blanking wipes that again,
// but the spill of the last ROI row (y = H-1-r) into row H-r survives (oracle/bm_oracle.c; probed on cv2 4.13).
// One warp per (image, spilled position): brute-force SAD over the window, lanes over the disparity index.
__global__ void __launch_bounds__(32) k_bm_spill(const uint8_t* __restrict__ lp, const uint32_t* __restrict__ rdesc,
                                                 BmGeo g, int16_t* __restrict__ disp)
{
    __shared__ int sad[512 + 2];
    const int lane = threadIdx.x, b = blockIdx.y;
    const int xp = g.W - g.lofs + blockIdx.x;                // lofs + xp = W + blockIdx.x
    const int y = g.H - 1 - g.r, D = g.D, r = g.r;
    const uint8_t* L = lp + (size_t)b * g.H * g.W;
    auto rv = [&](int yy, int x) -> int {                    // prefiltered right pixel: low half of its paired plane entry
        return (int)(rdesc[(size_t)(b * g.H + yy) * 4 * g.Wh + (size_t)(x & 1) * g.Wh + (x >> 1)] & 0xffffu);
    };
    int ts = 0;
    for (int k = lane; k < D; k += 32) sad[k + 1] = 0;
    for (int j = -r; j <= r; ++j) {
        const int lc = min(max(xp + j + g.lofs, 0), g.W - 1);
        const int rs = min(max(xp + j + g.rofs, 0), g.W - D);
        for (int i = -r; i <= r; ++i) {
            const int yy = min(max(y + i, 0), g.H - 1);
            const int lv = L[(size_t)yy * g.W + lc];
            ts += abs(lv - g.cap);
            for (int k = lane; k < D; k += 32) sad[k + 1] += abs(lv - rv(yy, rs + k));
        }
    }
    __syncwarp();
    if (ts < g.tex) return;
    uint32_t key = 0xffffffffu;
    for (int k = lane; k < D; k += 32) key = min(key, ((uint32_t)sad[k + 1] << 16) | (uint32_t)k);
    key = __reduce_min_sync(0xffffffffu, key);
    const int minsad = (int)(key >> 16), mind = (int)(key & 0xffffu);
    if (g.uniq > 0) {
        const int thresh = minsad + minsad * g.uniq / 100;
        bool rej = false;
        for (int k = lane; k < D; k += 32) rej |= (k < mind - 1 || k > mind + 1) && sad[k + 1] <= thresh;
        if (__any_sync(0xffffffffu, rej)) return;
    }
    if (lane == 0) {
        sad[0] = sad[2]; sad[D + 1] = sad[D - 1];
        const int pp = sad[mind + 2], n = sad[mind], dd = pp + n - 2 * minsad + abs(pp - n);
        disp[(size_t)b * g.H * g.W + (size_t)(y + 1) * g.W + blockIdx.x] =
            (int16_t)(((D - mind - 1 + g.minD) * 256 + (dd != 0 ? (pp - n) * 256 / dd : 0) + 15) >> 4);
    }
}

__global__ void k_bm_fill(int16_t* p, size_t n, int16_t v)      // 8 values (one 128-bit store) per thread
{
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i + 8 <= n && (reinterpret_cast<uintptr_t>(p + i) & 15) == 0) {
        const uint32_t w = (uint32_t)(uint16_t)v * 0x10001u;
        *reinterpret_cast<uint4*>(p + i) = make_uint4(w, w, w, w);
    } else {
        for (size_t j = i; j < n && j < i + 8; ++j) p[j] = v;
    }
}

// ------------------------------------------------------------------------------------------
struct svb_bm {
    svb_bm_params prm;
    SvbBuf ws, stage_in, stage_out;
    int launches = 0;
    int num_sms = 0;
    // *_host pipeline: upload of sub-batch i+1 and download of sub-batch i-1 overlap the kernels of sub-batch i
    // (every BM kernel is per image, so sub-batches cost nothing on the device)
    cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[4] = {}, ev_done[4] = {}, ev_mark[4] = {};
    SvbPinBuf pin_in, pin_out;      // pinned staging when the caller's host buffers are pageable (NumPy arrays)
    // successive computes of one handle share its workspace: a compute that arrives on another stream waits here
    cudaEvent_t ev_last = nullptr;
};

static int bm_geo(const svb_bm_params& p, int H, int W, BmGeo& g, bool& trivial)
{
    // what cv2 4.13 rejects (stereobm.cpp: CV_Error StsOutOfRange / StsBadArg)
    if (p.numDisparities <= 0 || p.numDisparities % 16 != 0) {
        svb_set_error("numDisparities must be positive and divisible by 16"); return SVB_ERR_PARAM;
    }
    if (p.blockSize < 5 || p.blockSize > 255 || p.blockSize % 2 == 0) {
        svb_set_error("SADWindowSize must be odd, be within 5..255"); return SVB_ERR_PARAM;
    }
    if (H > 0 && W > 0 && (p.blockSize >= W || p.blockSize >= H)) {
        svb_set_error("SADWindowSize must not be larger than image width or height"); return SVB_ERR_PARAM;
    }
    if (p.preFilterType != 0 && p.preFilterType != 1) {
        svb_set_error("preFilterType must be PREFILTER_XSOBEL or PREFILTER_NORMALIZED_RESPONSE"); return SVB_ERR_PARAM;
    }
    if (p.preFilterSize < 5 || p.preFilterSize > 255 || p.preFilterSize % 2 == 0) {
        svb_set_error("preFilterSize must be odd and be within 5..255"); return SVB_ERR_PARAM;
    }
    if (p.preFilterCap < 1 || p.preFilterCap > 63) { svb_set_error("preFilterCap must be within 1..63"); return SVB_ERR_PARAM; }
    if (p.textureThreshold < 0) { svb_set_error("texture threshold must be non-negative"); return SVB_ERR_PARAM; }
    if (p.uniquenessRatio < 0) { svb_set_error("uniqueness ratio must be non-negative"); return SVB_ERR_PARAM; }
    if (p.blockSize > 31) { svb_set_error("blockSize %d > 31 unsupported by the BM kernel", p.blockSize); return SVB_ERR_UNSUPPORTED; }
    if (p.numDisparities > 512) { svb_set_error("numDisparities > 512 unsupported"); return SVB_ERR_UNSUPPORTED; }
    if ((long long)p.blockSize * p.blockSize * 2 * p.preFilterCap > 65535) {
        svb_set_error("blockSize^2 * 2*preFilterCap exceeds the 16-bit SAD range of the BM kernel"); return SVB_ERR_UNSUPPORTED;
    }
    if (p.disp12MaxDiff >= 0 && (long long)p.blockSize * p.blockSize * 2 * p.preFilterCap > 32767) {
        svb_set_error("disp12MaxDiff >= 0: blockSize^2 * 2*preFilterCap exceeds the int16 cost map cv2 validates with");
        return SVB_ERR_UNSUPPORTED;
    }
    g.H = H; g.W = W; g.D = p.numDisparities; g.NP = (g.D + 63) / 64; g.minD = p.minDisparity;
    g.bs = p.blockSize; g.r = p.blockSize / 2; g.cap = p.preFilterCap; g.tex = p.textureThreshold;
    g.uniq = p.uniquenessRatio; g.FILTERED = (g.minD - 1) * 16;
    g.pft = p.preFilterType; g.pfs = p.preFilterSize;
    const int s = g.D - 1 + g.minD;
    g.lofs = std::max(s, 0); g.rofs = -std::min(s, 0);
    g.width1 = W - g.rofs - g.D + 1;
    g.Wh = (W + 1) / 2 + 4;
    trivial = g.lofs >= W || g.rofs >= W || g.width1 < 1;   // cv2 returns an all-FILTERED map
    return SVB_OK;
}

extern "C" int svb_bm_create(const svb_bm_params* p, svb_bm_t** out)
{
    if (!p || !out) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    BmGeo g; bool trivial;
    int rc = bm_geo(*p, 0, 0, g, trivial);
    if (rc != SVB_OK) return rc;
    svb_bm* h = new (std::nothrow) svb_bm();
    if (!h) return SVB_ERR_NOMEM;
    h->prm = *p;
    *out = h;
    return SVB_OK;
}
extern "C" void svb_bm_destroy(svb_bm_t* h)
{
    if (!h) return;
    for (int i = 0; i < 4; ++i) {
        if (h->ev_in[i]) cudaEventDestroy(h->ev_in[i]);
        if (h->ev_done[i]) cudaEventDestroy(h->ev_done[i]);
        if (h->ev_mark[i]) cudaEventDestroy(h->ev_mark[i]);
    }
    if (h->s_in) cudaStreamDestroy(h->s_in);
    if (h->s_run) cudaStreamDestroy(h->s_run);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    if (h->ev_last) cudaEventDestroy(h->ev_last);
    h->ws.release(); h->stage_in.release(); h->stage_out.release();
    h->pin_in.release(); h->pin_out.release();
    delete h;
}
extern "C" int svb_bm_last_launches(const svb_bm_t* h) { return h ? h->launches : 0; }

// ------------------------------------------------------------------------------------------
// k_bm2g: k_bm2 for numDisparities <= 64, where a vector is 32 words and k_bm2's warp-per-vector passes spend ~190 of
// their ~200 instructions per pixel on bookkeeping around three packed SAD operations.  A warp pass covers FOUR window /
// output columns: a group of 8 lanes per column, 4 words (8 disparities) per lane, so that the vertical sums, the row
// buffer, the horizontal box, the argmin keys and the uniqueness count are 4 columns wide per instruction; minima and
// counts are taken inside the groups by xor butterflies.  Strips are 128 window columns wide (k_bm2: 64): with a
// 15-wide window 114 of them are outputs instead of 50 of 64.  Same arithmetic as k_bm2 (tests: both against the oracle).
// ------------------------------------------------------------------------------------------
#ifndef BM_GROUPED
#define BM_GROUPED 1             // numDisparities <= 64: k_bm2g instead of k_bm2<1>
#endif
#define BMG_WARPS 16
#ifndef BMG_PC
#define BMG_PC 128                         // window columns per strip: BMG_NP passes of 4 per warp
#endif
#define BMG_NP (BMG_PC / (4 * BMG_WARPS))
#ifndef BMG_MAXSEG
#define BMG_MAXSEG 16            // most row segments per image
#endif
#ifndef BMG_STAGE
#define BMG_STAGE 1              // right descriptors of the strip staged in a shared-memory ring of blockSize + 2 rows
#endif
// words per staged plane: the strip's columns span <= (BMG_PC - 1) / 2 + 1 words, a vector 32 more; == 2 (mod 4) so that the
// four columns of a warp pass (two parities x two adjacent words) fall on different banks
#define BMG_NWP (((BMG_PC / 2 + 32 + 1) & ~3) + 2)
template <bool COST>
__global__ void __launch_bounds__(BMG_WARPS * 32) k_bm2g(const uint8_t* __restrict__ lp, const uint32_t* __restrict__ rdesc,
                                                         const uint8_t* __restrict__ okmap, BmGeo g, int OC, int nstrips,
                                                         int rseg, int16_t* __restrict__ disp, int16_t* __restrict__ cost)
{
    constexpr int VW = 32, WD = 8;
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* vbuf = smem;                                   // [2][BMG_PC][VW] vertical sums of the current row
    uint32_t* scr = smem + 2 * BMG_PC * VW;                  // [BMG_WARPS][4][VW] per-group scratch for the subpixel taps
    uint32_t* rring = scr + 4 * BMG_WARPS * VW;              // [bs + 2][4][BMG_NWP] right descriptors of the last rows (BMG_STAGE)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, grp = lane / WD, l = lane % WD;
    const int strip = blockIdx.x % nstrips, seg = blockIdx.x / nstrips, b = blockIdx.z;
    const int r = g.r, bs = 2 * r + 1, D = g.D, W = g.W, H = g.H;
    const int ya = r + seg * rseg, yb = min(ya + rseg, H - r);   // output rows of this CTA
    if (ya >= yb) return;
    const int xs = strip * OC;                               // first output column (width1 domain)

    // row-invariant addressing of this group's two window columns (A.6: left column and right start clamp separately)
    int lc[BMG_NP], pl[BMG_NP], m[BMG_NP][4];
    // BMG_STAGE: first descriptor word the strip needs (its first column's last disparity; negative where numDisparities < 64
    // leaves pad lanes -- the staged row holds word max(w, 0) there, what the clamp in m[][] reads).  Word k of a lane is then
    // mb[p] - 4k bytes into the staged row with no clamp of its own: two registers instead of eight
    const int wlo = ((min(max(xs - r + g.rofs, 0), W - D) + D - 1) >> 1) - 31;
    int mb[BMG_NP];
#pragma unroll
    for (int p = 0; p < BMG_NP; ++p) {
        const int j = xs - r + 4 * BMG_NP * warp + 4 * p + grp;
        lc[p] = min(max(j + g.lofs, 0), W - 1);
        const int rs = min(max(j + g.rofs, 0), W - D);
        const int xr0 = rs + D - 1;                          // right column of disparity index d' = 0
        pl[p] = (xr0 & 1) * g.Wh;
#pragma unroll
        for (int k = 0; k < 4; ++k) m[p][k] = max((xr0 >> 1) - (4 * l + k), 0);
        mb[p] = ((xr0 & 1) * BMG_NWP + (xr0 >> 1) - 4 * l - wlo) * 4;
    }
    uint32_t V[BMG_NP][4];
#pragma unroll
    for (int p = 0; p < BMG_NP; ++p)
#pragma unroll
        for (int k = 0; k < 4; ++k) V[p][k] = 0;
    // per-lane constants of the decision: BM indices kk of this lane's eight disparities, pad flags
    uint32_t kkpk[4], padlo[4], padhi[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int d0 = 2 * (4 * l + k);
        kkpk[k] = (uint32_t)((D - 1 - d0) & 0xffff) | ((uint32_t)((D - 2 - d0) & 0xffff) << 16);
        padlo[k] = d0 < D ? 0u : 0xffffffffu;
        padhi[k] = d0 + 1 < D ? 0u : 0xffffffffu;
    }
    uint32_t rkey = 0, rpn = 0, roff = 0;                    // parked pixel of this lane (one of its group's up to 8)
    int nrec = 0;                                            // pixels parked in this lane's group
    const int R = (OC + 4 * BMG_WARPS - 1) / (4 * BMG_WARPS);    // output columns per group
    const int oc_lo = (4 * warp + grp) * R, oc_hi = min(oc_lo + R, OC);
    const uint8_t* limg = lp + (size_t)b * H * W;
    const uint32_t* rimg = rdesc + (size_t)b * H * 4 * g.Wh;
    int16_t* dimg = disp + (size_t)b * H * W;
    int16_t* cimg = COST ? cost + (size_t)b * H * W : nullptr;
    uint32_t* sv = scr + (size_t)(4 * warp + grp) * VW;

    const int v0 = ya - r, v1 = yb - 1 + r;                  // input rows: all inside the image
    // BMG_STAGE: row v of the four descriptor planes, words wlo .. wlo + BMG_NWP - 1, into ring slot (v - v0) % (bs + 2); the
    // row is staged one iteration ahead, and the slot it takes was last read (as row v - bs) two barriers ago
    const int nslot = bs + 2;
    auto stage_row = [&](int v, int slot) {
        if (threadIdx.x < 4 * BMG_NWP) {
            const int q = threadIdx.x / BMG_NWP, i = threadIdx.x - q * BMG_NWP;
            rring[(slot * 4 + q) * BMG_NWP + i] = __ldg(rimg + ((size_t)v * 4 + q) * g.Wh + min(max(wlo + i, 0), g.Wh - 1));
        }
    };
    int snew = 0;
    if (BMG_STAGE) {
        stage_row(v0, 0);
        __syncthreads();
    }
    for (int v = v0; v <= v1; ++v) {
        const int it = v - v0, buf = it & 1;
        // ---- vertical phase: V += |L - R|(row v) - |L - R|(row v - bs), two passes of four columns
        if (BMG_STAGE) {
            const int sn = snew, so = sn + 2 >= nslot ? sn + 2 - nslot : sn + 2, sx = sn + 1 == nslot ? 0 : sn + 1;
            snew = sx;
            if (v + 1 <= v1) stage_row(v + 1, sx);
            const uint8_t* lrow = limg + (size_t)v * W;
            const char* rn = reinterpret_cast<const char*>(rring + sn * 4 * BMG_NWP);
            const char* ro = reinterpret_cast<const char*>(rring + so * 4 * BMG_NWP);
            auto rd = [](const char* row, int ofs) { return *reinterpret_cast<const uint32_t*>(row + ofs); };
            uint32_t* vb = vbuf + ((size_t)buf * BMG_PC + 4 * BMG_NP * warp + grp) * VW + 4 * l;
#pragma unroll
            for (int p = 0; p < BMG_NP; ++p) {
                {
                    const int u = lrow[lc[p]];
                    const uint32_t U2 = pk_dup((uint32_t)u), nU2 = pk_dup((uint32_t)(-u));
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        V[p][k] += bm_absdiff(U2, nU2, rd(rn, mb[p] - 4 * k), rd(rn, mb[p] - 4 * k + 2 * BMG_NWP * 4));
                }
                if (it >= bs) {
                    const int u = (lrow - (size_t)bs * W)[lc[p]];
                    const uint32_t U2 = pk_dup((uint32_t)u), nU2 = pk_dup((uint32_t)(-u));
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        V[p][k] -= bm_absdiff(U2, nU2, rd(ro, mb[p] - 4 * k), rd(ro, mb[p] - 4 * k + 2 * BMG_NWP * 4));
                }
                *reinterpret_cast<uint4*>(vb + 4 * p * VW) = make_uint4(V[p][0], V[p][1], V[p][2], V[p][3]);
            }
        } else {
            const uint8_t* lrow = limg + (size_t)v * W;
            const uint32_t* rrow = rimg + (size_t)v * 4 * g.Wh;
            uint32_t* vb = vbuf + ((size_t)buf * BMG_PC + 4 * BMG_NP * warp + grp) * VW + 4 * l;
#pragma unroll
            for (int p = 0; p < BMG_NP; ++p) {
                {
                    const int u = lrow[lc[p]];
                    const uint32_t U2 = pk_dup((uint32_t)u), nU2 = pk_dup((uint32_t)(-u));
                    const uint32_t* rp = rrow + pl[p];
#pragma unroll
                    for (int k = 0; k < 4; ++k) V[p][k] += bm_absdiff(U2, nU2, __ldg(rp + m[p][k]), __ldg(rp + 2 * g.Wh + m[p][k]));
                }
                if (it >= bs) {
                    const int u = (lrow - (size_t)bs * W)[lc[p]];
                    const uint32_t U2 = pk_dup((uint32_t)u), nU2 = pk_dup((uint32_t)(-u));
                    const uint32_t* rp = rrow - (size_t)bs * 4 * g.Wh + pl[p];
#pragma unroll
                    for (int k = 0; k < 4; ++k) V[p][k] -= bm_absdiff(U2, nU2, __ldg(rp + m[p][k]), __ldg(rp + 2 * g.Wh + m[p][k]));
                }
                *reinterpret_cast<uint4*>(vb + 4 * p * VW) = make_uint4(V[p][0], V[p][1], V[p][2], V[p][3]);
            }
        }
        __syncthreads();
        if (it < bs - 1) continue;                           // window not complete yet
        // ---- horizontal phase: output row y; group (warp, grp) owns the run of output columns oc_lo .. oc_hi - 1
        const int y = v - r;
        const uint32_t* vb = vbuf + (size_t)buf * BMG_PC * VW + 4 * l;
        const uint8_t* okrow = okmap + ((size_t)b * H + y) * W;
        uint32_t s[4];
        for (int jo = 0; jo < R; ++jo) {
            const int oc = oc_lo + jo;
            const int occ = min(oc, BMG_PC - bs);            // (an idle group re-reads valid columns, decides nothing)
            if (jo == 0) {
                uint4 t = *reinterpret_cast<const uint4*>(vb + occ * VW);
                s[0] = t.x; s[1] = t.y; s[2] = t.z; s[3] = t.w;
                for (int c = 1; c < bs; ++c) {
                    t = *reinterpret_cast<const uint4*>(vb + (occ + c) * VW);
                    s[0] += t.x; s[1] += t.y; s[2] += t.z; s[3] += t.w;
                }
            } else {
                const uint4 a = *reinterpret_cast<const uint4*>(vb + (occ + bs - 1) * VW);
                const uint4 e = *reinterpret_cast<const uint4*>(vb + (occ - 1) * VW);
                s[0] += a.x - e.x; s[1] += a.y - e.y; s[2] += a.z - e.z; s[3] += a.w - e.w;
            }
            const int x = xs + oc + g.lofs;
            bool ok = oc < oc_hi && x < W;
            if (ok) ok = okrow[x] != 0;
            if (!__any_sync(0xffffffffu, ok)) continue;
            // argmin over BM index kk = D-1-d' (lowest kk = largest disparity wins ties): keys (sad << 16 | kk)
            uint32_t key = 0xffffffffu;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t klo = __byte_perm(s[k], kkpk[k], 0x1054), khi = __byte_perm(s[k], kkpk[k], 0x3276);
                key = min(key, min(klo | padlo[k], khi | padhi[k]));
            }
#pragma unroll
            for (int o = WD / 2; o; o >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, o));
            const int minsad = (int)(key >> 16), mind = (int)(key & 0xffffu);
            // the SADs at mind-1 / mind+1 (replicated at the ends of the range): subpixel taps, and the two entries
            // the uniqueness test leaves out
            *reinterpret_cast<uint4*>(sv + 4 * l) = make_uint4(s[0], s[1], s[2], s[3]);
            __syncwarp();
            const uint16_t* sd = reinterpret_cast<const uint16_t*>(sv);         // indexed by d' = D-1-kk
            const bool hasp = mind + 1 < D, hasn = mind >= 1;
            const int p = sd[hasp ? D - 2 - mind : 1], n = sd[hasn ? D - mind : D - 2];   // sad[mind+1], sad[mind-1]
            __syncwarp();
            if (g.uniq > 0) {
                // cv2: reject iff some kk outside mind-1..mind+1 has sad <= thresh; count all sad <= thresh instead
                const int thresh = minsad + (minsad * g.uniq / 100);
                int cnt = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    cnt += ((int)(s[k] & 0xffffu) <= thresh && !padlo[k]) + ((int)(s[k] >> 16) <= thresh && !padhi[k]);
#pragma unroll
                for (int o = WD / 2; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
                if (cnt > 1 + (hasp && p <= thresh) + (hasn && n <= thresh)) ok = false;
            }
            // park the pixel in lane nrec of its group; 8 parked pixels are finished together (one division per lane)
            if (ok) {
                if (l == nrec) { rkey = key; rpn = (uint32_t)p | ((uint32_t)n << 16); roff = (uint32_t)(y * W + x); }
                if (++nrec == WD) {
                    bm2_finish(rkey, rpn, roff, D, g.minD, dimg, cimg);
                    nrec = 0;
                }
            }
        }
    }
    if (l < nrec) bm2_finish(rkey, rpn, roff, D, g.minD, dimg, cimg);
}
static int bm2g_launch(const BmGeo& g, const uint8_t* lp, const uint32_t* rdesc, const uint8_t* tex, int nb, int num_sms,
                       int16_t* disp, int16_t* cost, cudaStream_t st)
{
    // output columns per strip: all of them, or a multiple of the group count when that wastes little
    const int ocmax = BMG_PC - 2 * g.r, groups = 4 * BMG_WARPS, ocr = ocmax / groups * groups;
    const int OC = ocr >= groups && ocmax - ocr <= groups / 8 ? ocr : ocmax;
    const int bs = 2 * g.r + 1, nstrips = svb_ceil_div(g.width1, OC), rows = g.H - 2 * g.r;
    // row segments by waves x rows per CTA (two CTAs of 512 threads per SM), as for the SGBM cost kernel
    int nsegs = 1, rseg = rows;
    {
        const long long slots = 2LL * num_sms, base = (long long)nstrips * nb;
        const int maxseg = std::max(1, std::min(BMG_MAXSEG, rows / (4 * bs)));
        long long best = 0;
        for (int n = 1; n <= maxseg; ++n) {
            const int rs = svb_ceil_div(rows, n), nsg = svb_ceil_div(rows, rs);
            const long long cost = ((base * nsg + slots - 1) / slots) * (rs + bs - 1 + 6);
            if (n == 1 || cost < best) { best = cost; rseg = rs; nsegs = nsg; }
        }
    }
    const size_t smem = ((size_t)2 * BMG_PC + 4 * BMG_WARPS) * 32 * 4 + (BMG_STAGE ? (size_t)(bs + 2) * 4 * BMG_NWP * 4 : 0);
    if (cost) {
        SVB_CUDA(cudaFuncSetAttribute(k_bm2g<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_bm2g<true><<<dim3(nstrips * nsegs, 1, nb), BMG_WARPS * 32, smem, st>>>(lp, rdesc, tex, g, OC, nstrips, rseg, disp, cost);
    } else {
        SVB_CUDA(cudaFuncSetAttribute(k_bm2g<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_bm2g<false><<<dim3(nstrips * nsegs, 1, nb), BMG_WARPS * 32, smem, st>>>(lp, rdesc, tex, g, OC, nstrips, rseg, disp, cost);
    }
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

template <int NP>
static int bm2_launch(const BmGeo& g, const uint8_t* lp, const uint32_t* rdesc, const uint8_t* tex, int nb, int num_sms,
                      int16_t* disp, int16_t* cost, cudaStream_t st)
{
    // output columns per strip: a multiple of the warp count when that costs little (every warp gets the same run)
    // (phase 1 costs ~36 instructions per window column, phase 2 ~100 per output column of the longest run)
    const int ocmax = BM2_PC - 2 * g.r, ocr = ocmax / BM2_WARPS * BM2_WARPS;
    auto est = [](int oc) { return (36.0 * BM2_PC + 100.0 * BM2_WARPS * ((oc + BM2_WARPS - 1) / BM2_WARPS)) / oc; };
    const int OC = ocr >= BM2_WARPS && est(ocr) < est(ocmax) ? ocr : ocmax;
    const int bs = 2 * g.r + 1, nstrips = svb_ceil_div(g.width1, OC), rows = g.H - 2 * g.r;
    // row segments: enough CTAs to fill the SMs, few enough to keep the bs-1 warm-up rows cheap
    int nsegs = 1;
    while ((long long)nstrips * nb * nsegs < 4LL * num_sms && rows / (nsegs * 2) >= 4 * bs) nsegs *= 2;
    const int rseg = svb_ceil_div(rows, nsegs);
    nsegs = svb_ceil_div(rows, rseg);
    const size_t smem = ((size_t)2 * BM2_PC + BM2_WARPS) * NP * 32 * 4;
    SVB_CUDA(cudaFuncSetAttribute(k_bm2<NP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_bm2<NP><<<dim3(nstrips * nsegs, 1, nb), BM2_WARPS * 32, smem, st>>>(lp, rdesc, tex, g, OC, nstrips, rseg, disp, cost);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}

extern "C" int svb_bm_compute_dev(svb_bm_t* h, const uint8_t* left, const uint8_t* right, int B, int H, int W,
                                  int16_t* disp, void* stream)
{
    if (!h || !left || !right || !disp) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    if (B <= 0 || H <= 0 || W <= 0) { svb_set_error("bad shape"); return SVB_ERR_PARAM; }
    BmGeo g; bool trivial;
    int rc = bm_geo(h->prm, H, W, g, trivial);
    if (rc != SVB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    h->launches = 0;
    if (!h->ev_last) SVB_CUDA(cudaEventCreateWithFlags(&h->ev_last, cudaEventDisableTiming));
    else SVB_CUDA(cudaStreamWaitEvent(st, h->ev_last, 0));
    struct Rec {                                             // whatever path returns: the next compute waits for this one
        svb_bm* h; cudaStream_t st;
        ~Rec() { cudaEventRecord(h->ev_last, st); }
    } rec{h, st};
    const size_t npix = (size_t)B * H * W;
    k_bm_fill<<<svb_ceil_div(npix, 256 * 8), 256, 0, st>>>(disp, npix, (int16_t)g.FILTERED);
    SVB_LAUNCH_CHECK(); h->launches++;
    if (trivial) return SVB_OK;
    const size_t lp_bytes = (npix + 255) & ~(size_t)255;
    const size_t rd_bytes = ((size_t)B * H * 4 * g.Wh * 4 + 255) & ~(size_t)255;
    const bool validate = h->prm.disp12MaxDiff >= 0;
    // scratch shared by validateDisparity (keys u64 + input copy s16 + cost s16) and the speckle filter (2 x s32)
    const size_t sc_bytes = validate ? npix * 12 : (h->prm.speckleWindowSize > 0 ? npix * 8 : 0);
    const size_t vs_bytes = g.pft == 0 ? npix * 4 : 0;       // vertical sums of both images (u16), dead after the prefilter
    const size_t sv_bytes = (std::max(sc_bytes, vs_bytes) + 255) & ~(size_t)255;
    const size_t th_bytes = (npix * 2 + 255) & ~(size_t)255;     // texture map: horizontal pass u16, then s32
    rc = h->ws.reserve(lp_bytes + rd_bytes + sv_bytes + th_bytes + npix * 4);
    if (rc != SVB_OK) return rc;
    if (!h->num_sms) {
        int dev = 0;
        SVB_CUDA(cudaGetDevice(&dev));
        SVB_CUDA(cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    uint8_t* lp = (uint8_t*)h->ws.p;
    uint32_t* rdesc = (uint32_t*)((char*)h->ws.p + lp_bytes);
    uint16_t* vs = nullptr;
    if (g.pft == 0) {
        vs = (uint16_t*)((char*)h->ws.p + lp_bytes + rd_bytes);
        k_bm_vsum<<<dim3(svb_ceil_div(W, 256), H, 2 * B), 256, 0, st>>>(left, right, g, vs);
        SVB_LAUNCH_CHECK(); h->launches++;
    }
    k_bm_prefilter<<<dim3(svb_ceil_div(W, 256), H, B), 256, 0, st>>>(left, right, g, lp, rdesc, vs);
    SVB_LAUNCH_CHECK(); h->launches++;
    char* scratch = (char*)h->ws.p + lp_bytes + rd_bytes;
    int16_t* cost = validate ? (int16_t*)(scratch + npix * 10) : nullptr;
    uint16_t* th = (uint16_t*)(scratch + sv_bytes);
    uint8_t* tex = (uint8_t*)(scratch + sv_bytes + th_bytes);
    k_bm_tex_h<<<dim3(svb_ceil_div(W, 256 * TEX_P), H, B), 256, 0, st>>>(lp, g, th);
    SVB_LAUNCH_CHECK(); h->launches++;
    k_bm_tex_v<<<dim3(svb_ceil_div(W, 256), svb_ceil_div(H, TEXV_R), B), 256, 0, st>>>(th, g, validate ? 1 : 0, tex);
    SVB_LAUNCH_CHECK(); h->launches++;
    if (BM_GROUPED && g.NP == 1) rc = bm2g_launch(g, lp, rdesc, tex, B, h->num_sms, disp, cost, st);
    else switch (g.NP) {
#define CASE(N) case N: rc = bm2_launch<N>(g, lp, rdesc, tex, B, h->num_sms, disp, cost, st); break;
        CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
        default: svb_set_error("bad NP"); rc = SVB_ERR_UNSUPPORTED;
    }
    if (rc != SVB_OK) return rc;
    h->launches++;
    if (validate) {
        // the winning SAD fits int16 (checked in bm_geo); columns outside the valid ROI are blanked by the check
        unsigned long long* keys = (unsigned long long*)scratch;
        int16_t* copy = (int16_t*)(keys + npix);
        const int maxD = g.minD + g.D, minX1 = std::max(maxD, 0), maxX1 = W + std::min(g.minD, 0);
        SVB_CUDA(cudaMemsetAsync(keys, 0xff, npix * 8, st));
        SVB_CUDA(cudaMemcpyAsync(copy, disp, npix * 2, cudaMemcpyDeviceToDevice, st));
        if (maxX1 > minX1) {
            k_validate_scatter<int16_t><<<dim3(svb_ceil_div(maxX1 - minX1, 256), H, B), 256, 0, st>>>(
                copy, cost, H, W, minX1, maxX1, g.FILTERED, keys);
            SVB_LAUNCH_CHECK(); h->launches++;
        }
        k_validate_check<<<dim3(svb_ceil_div(W, 256), H, B), 256, 0, st>>>(
            copy, disp, H, W, minX1, maxX1, g.FILTERED, h->prm.disp12MaxDiff * 16, g.lofs + g.r, W - g.r, keys);
        SVB_LAUNCH_CHECK(); h->launches++;
    }
    if (g.minD > 0 && g.lofs + g.width1 > W) {
        k_bm_spill<<<dim3(g.lofs + g.width1 - W, B), 32, 0, st>>>(lp, rdesc, g, disp);
        SVB_LAUNCH_CHECK(); h->launches++;
    }
    if (h->prm.speckleWindowSize > 0) {
        // stereobm.cpp passes speckleRange unscaled (SGBM scales it by 16) -- probed on cv2 4.13
        rc = svb_filter_speckles_s16_dev(disp, B, H, W, g.FILTERED, h->prm.speckleWindowSize, h->prm.speckleRange,
                                         (int32_t*)scratch, st);
        if (rc != SVB_OK) return rc;
        h->launches += 4;
    }
    return SVB_OK;
}

extern "C" int svb_bm_compute_host(svb_bm_t* h, const uint8_t* left, const uint8_t* right, int B, int H, int W,
                                   int16_t* disp)
{
    if (!h || !left || !right || !disp) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    if (B <= 0 || H <= 0 || W <= 0) { svb_set_error("bad shape"); return SVB_ERR_PARAM; }
    const size_t n = (size_t)B * H * W;
    int rc = h->stage_in.reserve(2 * n);
    if (rc != SVB_OK) return rc;
    rc = h->stage_out.reserve(2 * n);
    if (rc != SVB_OK) return rc;
    uint8_t* dl = (uint8_t*)h->stage_in.p;
    uint8_t* dr = dl + n;
    int16_t* dout = (int16_t*)h->stage_out.p;
    if (!h->s_in) {
        SVB_CUDA(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
        SVB_CUDA(cudaStreamCreateWithFlags(&h->s_run, cudaStreamNonBlocking));
        SVB_CUDA(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
        for (int i = 0; i < 4; ++i) {
            SVB_CUDA(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
            SVB_CUDA(cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming));
            SVB_CUDA(cudaEventCreateWithFlags(&h->ev_mark[i], cudaEventDisableTiming));
        }
    }
    const int nsub = std::min(B, 4), per = (B + nsub - 1) / nsub;
    const size_t hw = (size_t)H * W;
    // Pageable caller buffers (NumPy arrays) of a batch: staged through pinned memory by the host pool, image by image
    // on the way in and per downloaded sub-range on the way out (see svb_sgbm_compute_cn_host)
    const bool staged = B > 1 && 4 * n >= ((size_t)8 << 20) && svb_host_pageable(left) && svb_host_pageable(right) &&
                        svb_host_pageable(disp);
    std::vector<std::atomic<int>> flags(staged ? B : 0);
    SvbTaskGroup gin, gout;
    int16_t* user_out = disp;
    if (staged) {
        rc = h->pin_in.reserve(2 * n);
        if (rc != SVB_OK) return rc;
        rc = h->pin_out.reserve(2 * n);
        if (rc != SVB_OK) return rc;
        uint8_t* pl = (uint8_t*)h->pin_in.p;
        uint8_t* pr = pl + n;
        for (auto& f : flags) f.store(0, std::memory_order_relaxed);
        const uint8_t *ul = left, *ur = right;
        for (int b = 0; b < B; ++b)
            svb_pool_submit(&gin, [=, &flags] {
                memcpy(pl + b * hw, ul + b * hw, hw);
                memcpy(pr + b * hw, ur + b * hw, hw);
                flags[b].store(1, std::memory_order_release);
            });
        left = pl; right = pr; disp = (int16_t*)h->pin_out.p;
    }
    struct Waiter { SvbTaskGroup *a, *b; ~Waiter() { svb_pool_wait(a); svb_pool_wait(b); } } waiter{&gin, &gout};   // on every return path
    int launches = 0, i = 0;
    for (int b0 = 0; b0 < B; b0 += per, ++i) {
        const int nb = std::min(per, B - b0);
        if (staged)
            for (int b = b0; b < b0 + nb; ++b)
                while (!flags[b].load(std::memory_order_acquire)) std::this_thread::yield();
        SVB_CUDA(cudaMemcpyAsync(dl + b0 * hw, left + b0 * hw, nb * hw, cudaMemcpyHostToDevice, h->s_in));
        SVB_CUDA(cudaMemcpyAsync(dr + b0 * hw, right + b0 * hw, nb * hw, cudaMemcpyHostToDevice, h->s_in));
        SVB_CUDA(cudaEventRecord(h->ev_in[i], h->s_in));
        SVB_CUDA(cudaStreamWaitEvent(h->s_run, h->ev_in[i], 0));
        rc = svb_bm_compute_dev(h, dl + b0 * hw, dr + b0 * hw, nb, H, W, dout + b0 * hw, h->s_run);
        if (rc != SVB_OK) { cudaDeviceSynchronize(); return rc; }
        launches += h->launches;
        SVB_CUDA(cudaEventRecord(h->ev_done[i], h->s_run));
        SVB_CUDA(cudaStreamWaitEvent(h->s_out, h->ev_done[i], 0));
        SVB_CUDA(cudaMemcpyAsync(disp + b0 * hw, dout + b0 * hw, nb * hw * 2, cudaMemcpyDeviceToHost, h->s_out));
        if (staged) SVB_CUDA(cudaEventRecord(h->ev_mark[i], h->s_out));
    }
    if (staged) {
        const int parts = svb_pool_threads();
        i = 0;
        for (int b0 = 0; b0 < B; b0 += per, ++i) {
            const int nb = std::min(per, B - b0);
            SVB_CUDA(cudaEventSynchronize(h->ev_mark[i]));
            const size_t bytes = (size_t)nb * hw * 2, chunk = (bytes / parts + 4095) & ~(size_t)4095;
            const char* src = (const char*)(disp + b0 * hw);
            char* dst = (char*)(user_out + b0 * hw);
            for (size_t o = 0; o < bytes; o += chunk)
                svb_pool_submit(&gout, [=] { memcpy(dst + o, src + o, std::min(chunk, bytes - o)); });
        }
        svb_pool_wait(&gout);
    }
    h->launches = launches;
    SVB_CUDA(cudaStreamSynchronize(h->s_out));
    SVB_CUDA(cudaStreamSynchronize(h->s_run));
    return SVB_OK;
}
