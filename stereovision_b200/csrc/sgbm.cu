// sgbm.cu -- semi-global block matching for sm_100a, bit-exact with
// cv2.StereoSGBM_create(...).compute (reference call sites depth_map.py:894-909,
// fused_depth_map.py:988-1004; arithmetic per SURVEY.md Appendix A.1-A.5 as
// corrected in oracle/sgbm_oracle.c).
//
// Pipeline (one CUDA stream, no host sync):
//   k_prefilter   x-Sobel + raw channel, Birchfield-Tomasi interval descriptors
//   k_cost        BT pixel cost (2 channels) + blockSize^2 box sum -> C volume (u16)
//   k_path<NP>    one launch per path direction: L_r recurrence, S += L_r (saturating)
//   k_wta<NP>     argmin / uniqueness / subpixel / right-view scatter-min
//   k_lrcheck     left-right consistency
//   median3 + speckle filter (imgproc.cu)
//
// Data layout in HBM: volumes are [job][row][xc][Dp] u16 with Dp = 64*NP >= D; the
// disparity axis is innermost so that one warp owns one pixel's whole disparity
// vector as NP packed u16x2 registers per lane (lane l holds d = 2*NP*l ... +2*NP-1).
#include "svb_common.cuh"
#include <algorithm>
#include <new>
#include <vector>
#include <string.h>

#define WALL 0x7fffu           // cv2's MAX_COST for the d=-1 / d=D neighbours and pad lanes
#define WALL2 0x7fff7fffu

struct SgbmGeo {
    int H, W, D, Dp, NP, minD, minX1, width1, bs, r, ftzero;
    int P1, P2, uniq, d12, INV, mode, nstripes, rowsMax;
    int Wh;                    // words per parity plane of the right descriptors
};

__host__ __device__ inline void job_rows(const SgbmGeo& g, int n, int& ys, int& ye, int& keep)
{
    if (g.nstripes == 1) { ys = 0; ye = g.H; keep = 0; return; }
    // 4 fixed stripes with overlap
    int ssz = (g.H + g.nstripes - 1) / g.nstripes;
    int overlap = (g.bs / 2 + 1) + (ssz + 9) / 10;
    ys = n * ssz - overlap; ys = ys < 0 ? 0 : (ys > g.H ? g.H : ys);
    ye = (n + 1) * ssz; ye = ye > g.H ? g.H : ye;
    keep = n * ssz;
    if (ye < ys) ye = ys;
}

// ------------------------------------------------------------------------------------------
// k_prefilter: per image row, the two BT channels (A.1 step 1-2) and their half-pixel
// interval bounds (A.1 step 3).  Left: one 32-byte record per pixel holding
// {u, -u, u0, -u1} x 2 channels, each duplicated into both halves of a u32.
// Right: 8 operand planes {v, -v, v0, -v1} x 2 channels; entry(xr) = a[xr] | a[xr-1]<<16,
// split by parity of xr so that a warp reading xr, xr-2, xr-4 ... is contiguous.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void bt_channels(const uint8_t* img, int H, int W, int y, int x, int ftzero,
                                            int& p0, int& p1)
{
    if (x <= 0 || x >= W - 1) { p0 = p1 = ftzero; return; }
    int yn = y > 0 ? y - 1 : y, ys = y < H - 1 ? y + 1 : y;
    const uint8_t *r = img + (size_t)y * W, *rn = img + (size_t)yn * W, *rs = img + (size_t)ys * W;
    int g = 2 * ((int)r[x + 1] - r[x - 1]) + ((int)rn[x + 1] - rn[x - 1]) + ((int)rs[x + 1] - rs[x - 1]);
    g = max(-ftzero, min(ftzero, g));
    p0 = g + ftzero; p1 = r[x];
}

__global__ void __launch_bounds__(256) k_prefilter(const uint8_t* __restrict__ left, const uint8_t* __restrict__ right,
                                                   int nimg, SgbmGeo g, uint32_t* __restrict__ ldesc,
                                                   uint32_t* __restrict__ rdesc)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y, b = blockIdx.z;
    if (x >= g.W) return;
    const int H = g.H, W = g.W;
    {
        const uint8_t* img = left + (size_t)b * H * W;
        int c0[3], c1[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) bt_channels(img, H, W, y, min(max(x + k - 1, 0), W - 1), g.ftzero, c0[k], c1[k]);
        uint32_t* rec = ldesc + ((size_t)(b * H + y) * W + x) * 8;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int* p = c ? c1 : c0;
            int u = p[1];
            int ua = x > 0 ? (u + p[0]) / 2 : u;
            int ub = x < W - 1 ? (u + p[2]) / 2 : u;
            int u0 = min(min(ua, ub), u), u1 = max(max(ua, ub), u);
            rec[c * 4 + 0] = pk_dup((uint32_t)u);
            rec[c * 4 + 1] = pk_dup((uint32_t)(-u));
            rec[c * 4 + 2] = pk_dup((uint32_t)u0);
            rec[c * 4 + 3] = pk_dup((uint32_t)(-u1));
        }
    }
    {
        const uint8_t* img = right + (size_t)b * H * W;
        int c0[4], c1[4];   // columns x-2 .. x+1
#pragma unroll
        for (int k = 0; k < 4; ++k) bt_channels(img, H, W, y, min(max(x + k - 2, 0), W - 1), g.ftzero, c0[k], c1[k]);
        uint32_t* base = rdesc + (size_t)(b * H + y) * 16 * g.Wh + (size_t)(x & 1) * g.Wh + (x >> 1);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int* p = c ? c1 : c0;
            int v[2], v0[2], v1[2];
#pragma unroll
            for (int s = 0; s < 2; ++s) {           // s=0: pixel x, s=1: pixel x-1
                int xx = x - s;
                int vv = p[2 - s];
                int va = xx < W - 1 ? (vv + p[3 - s]) / 2 : vv;
                int vb = xx > 0 ? (vv + p[1 - s]) / 2 : vv;
                v[s] = vv; v0[s] = min(min(va, vb), vv); v1[s] = max(max(va, vb), vv);
            }
            uint32_t ops[4] = {
                (uint32_t)(v[0] & 0xffff) | ((uint32_t)(v[1] & 0xffff) << 16),
                (uint32_t)((-v[0]) & 0xffff) | ((uint32_t)((-v[1]) & 0xffff) << 16),
                (uint32_t)(v0[0] & 0xffff) | ((uint32_t)(v0[1] & 0xffff) << 16),
                (uint32_t)((-v1[0]) & 0xffff) | ((uint32_t)((-v1[1]) & 0xffff) << 16)};
#pragma unroll
            for (int o = 0; o < 4; ++o) base[(size_t)(c * 4 + o) * 2 * g.Wh] = ops[o];
        }
    }
}

// ------------------------------------------------------------------------------------------
// k_cost: one block = TY output rows x one x-segment of one job; one warp per input row
// (TY + 2r of them, rows clamped into [ys, H-1] = replicate border, A.2).  Each warp slides
// along x keeping the horizontal box sum of the BT pixel cost (A.1 step 4) in registers
// (ring of the last bs pixel-cost vectors in shared memory); per column the warps meet in
// shared memory and the first TY warps add bs rows -> C(y, xc, :) -> HBM.
// Lane mapping here is "interleaved" (reg k, lane l <-> disparity pair e = 32k + l) so that
// right-image descriptor loads are contiguous; memory order of the volume is natural.
// ------------------------------------------------------------------------------------------
template <int NP>
__global__ void __launch_bounds__(1024) k_cost(const uint32_t* __restrict__ ldesc, const uint32_t* __restrict__ rdesc,
                                               SgbmGeo g, int TY, int XSEG, uint16_t* __restrict__ vol)
{
    extern __shared__ uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, NW = blockDim.x >> 5;
    const int job = blockIdx.z, b = job / g.nstripes, n = job % g.nstripes;
    int ys, ye, keep;
    job_rows(g, n, ys, ye, keep);
    const int y0 = ys + blockIdx.y * TY;
    if (y0 >= ye) return;
    const int xs = blockIdx.x * XSEG, xe = min(xs + XSEG, g.width1);
    const int r = g.r, bs = 2 * r + 1;
    uint32_t* ring = smem + (size_t)warp * bs * NP * 32;            // [slot][k][lane]
    uint32_t* hsb = smem + (size_t)NW * bs * NP * 32;               // [warp][k][lane]
    const int yy = min(max(y0 - r + warp, ys), g.H - 1);
    const uint32_t* lrow = ldesc + (size_t)(b * g.H + yy) * g.W * 8;
    const uint32_t* rrow = rdesc + (size_t)(b * g.H + yy) * 16 * g.Wh;
    for (int i = lane; i < bs * NP * 32; i += 32) ring[i] = 0;
    __syncwarp();
    uint32_t hs[NP];
#pragma unroll
    for (int k = 0; k < NP; ++k) hs[k] = 0;
    uint16_t* volrow = vol + ((size_t)job * g.rowsMax + (y0 - ys + warp)) * g.width1 * g.Dp;
    const bool out_warp = warp < TY && (y0 + warp) < ye;

    for (int j = xs - r; j < xe + r; ++j) {
        const int q = min(max(j, 0), g.width1 - 1);
        const int x = q + g.minX1;
        const uint4 l0 = *reinterpret_cast<const uint4*>(lrow + (size_t)x * 8);
        const uint4 l1 = *reinterpret_cast<const uint4*>(lrow + (size_t)x * 8 + 4);
        const int xr0 = x - g.minD;                  // right column of disparity index 0
        const uint32_t* rp = rrow + (size_t)(xr0 & 1) * g.Wh;
        const int slot = (j - (xs - r)) % bs;
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            const int m = max((xr0 >> 1) - (32 * k + lane), 0);
            uint32_t o[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) o[t] = __ldg(rp + (size_t)t * 2 * g.Wh + m);
            // channel 0 (prefiltered): c0 = max(0, u - v1, v0 - u); c1 = max(0, v - u1, u0 - v)
            uint32_t a = __vadd2(l0.x, o[3]);
            uint32_t c0 = __viaddmax_s16x2_relu(o[2], l0.y, a);
            uint32_t bb = __vadd2(o[0], l0.w);
            uint32_t c1 = __viaddmax_s16x2_relu(l0.z, o[1], bb);
            uint32_t pix = __vminu2(c0, c1);
            // channel 1 (raw intensity), cost >> 2
            a = __vadd2(l1.x, o[7]);
            c0 = __viaddmax_s16x2_relu(o[6], l1.y, a);
            bb = __vadd2(o[4], l1.w);
            c1 = __viaddmax_s16x2_relu(l1.z, o[5], bb);
            pix += (__vminu2(c0, c1) >> 2) & 0x3fff3fffu;
            uint32_t* rs = ring + (slot * NP + k) * 32 + lane;
            hs[k] += pix - *rs;
            *rs = pix;
        }
        if (j - r >= xs) {
            const int xc = j - r;
#pragma unroll
            for (int k = 0; k < NP; ++k) hsb[(warp * NP + k) * 32 + lane] = hs[k];
            __syncthreads();
            if (out_warp) {
                uint32_t* dst = reinterpret_cast<uint32_t*>(volrow + (size_t)xc * g.Dp);
#pragma unroll
                for (int k = 0; k < NP; ++k) {
                    uint32_t c = 0;
                    for (int t = 0; t < bs; ++t) c += hsb[((warp + t) * NP + k) * 32 + lane];
                    if (2 * (32 * k + lane) >= g.D) c = WALL2;
                    dst[32 * k + lane] = c;
                }
            }
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------------------
// Chunked per-lane view of one pixel's disparity vector: lane l, reg k <-> pair e = NP*l + k.
// ------------------------------------------------------------------------------------------
template <int NP>
__device__ __forceinline__ void load_vec(const uint16_t* p, int lane, uint32_t (&v)[NP])
{
    const uint32_t* w = reinterpret_cast<const uint32_t*>(p) + NP * lane;
    if constexpr (NP % 4 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 4; ++i) {
            uint4 t = reinterpret_cast<const uint4*>(w)[i];
            v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
        }
    } else if constexpr (NP % 2 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 2; ++i) {
            uint2 t = reinterpret_cast<const uint2*>(w)[i];
            v[2 * i] = t.x; v[2 * i + 1] = t.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < NP; ++i) v[i] = w[i];
    }
}
template <int NP>
__device__ __forceinline__ void store_vec(uint16_t* p, int lane, const uint32_t (&v)[NP])
{
    uint32_t* w = reinterpret_cast<uint32_t*>(p) + NP * lane;
    if constexpr (NP % 4 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 4; ++i)
            reinterpret_cast<uint4*>(w)[i] = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    } else if constexpr (NP % 2 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 2; ++i) reinterpret_cast<uint2*>(w)[i] = make_uint2(v[2 * i], v[2 * i + 1]);
    } else {
#pragma unroll
        for (int i = 0; i < NP; ++i) w[i] = v[i];
    }
}

// One step of the SGM recurrence (A.3) for a whole disparity vector held by a warp.
//   L(d) = C(d) + min(Lq(d), Lq(d-1)+P1, Lq(d+1)+P1, minLq+P2) - minLq
// Lq is updated in place; returns the warp-wide min of the new L (minLq for the next step).
template <int NP>
__device__ __forceinline__ uint32_t sgm_step(uint32_t (&Lq)[NP], const uint32_t (&C)[NP], uint32_t minq,
                                             uint32_t P1P1, uint32_t P2, int lane, int npairs_valid)
{
    // neighbours across lane boundaries; the vector ends see the wall
    uint32_t prev_last = __shfl_up_sync(0xffffffffu, Lq[NP - 1], 1);
    uint32_t next_first = __shfl_down_sync(0xffffffffu, Lq[0], 1);
    if (lane == 0) prev_last = WALL2;
    if (lane == 31) next_first = WALL2;
    const uint32_t mq2 = pk_dup(minq);
    const uint32_t mqP2 = pk_dup(minq + P2);
    uint32_t up = pk_shift_in(prev_last, Lq[0]);            // (Lq[d-1], Lq[d]) for reg 0
    uint32_t mloc = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < NP; ++k) {
        const uint32_t nxt = (k + 1 < NP) ? Lq[k + 1] : next_first;
        const uint32_t dn = pk_shift_in(Lq[k], nxt);        // (Lq[d+1], Lq[d+2])
        uint32_t t = pk_min(up, dn);
        t = pk_addmin(t, P1P1, Lq[k]);
        t = pk_min(t, mqP2);
        uint32_t L = C[k] + t - mq2;                        // halves stay in [0, 65535]: no carries
        if (NP * lane + k >= npairs_valid) L = WALL2;       // pad lanes keep the wall
        up = dn;
        Lq[k] = L;
        mloc = pk_min(mloc, L);
    }
    return __reduce_min_sync(0xffffffffu, pk_hmin(mloc));
}

// ------------------------------------------------------------------------------------------
// k_path: one warp follows one path line of direction (dx,dy) through the job's rows.
// dy == 0: line = one row, x runs 0..w1-1 (dx>0) or back (dx<0).
// dy != 0: line = start column x0; at step t the warp is at column (x0 + dx*t) mod w1 and
// restarts (Lq = 0, minLq = 0: "outside" predecessor) whenever it wraps around an image edge,
// so every warp does the same number of steps.
// ------------------------------------------------------------------------------------------
template <int NP>
__global__ void __launch_bounds__(256) k_path(const uint16_t* __restrict__ Cvol, uint16_t* __restrict__ Svol,
                                              SgbmGeo g, int dx, int dy, int first, int njobs)
{
    const int lane = threadIdx.x & 31;
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int w1 = g.width1;
    const int npv = g.D >> 1;
    const uint32_t P1P1 = pk_dup(g.P1);
    int job, ys, ye, keep;
    uint32_t Lq[NP], C[NP], S[NP];
#pragma unroll
    for (int k = 0; k < NP; ++k) Lq[k] = 0;
    uint32_t minq = 0;
    if (dy == 0) {
        job = (int)(gw / g.rowsMax);
        if (job >= njobs) return;
        job_rows(g, job % g.nstripes, ys, ye, keep);
        const int row = (int)(gw % g.rowsMax);
        if (row >= ye - ys) return;
        const size_t base = ((size_t)job * g.rowsMax + row) * w1 * g.Dp;
        for (int t = 0; t < w1; ++t) {
            const int xc = dx > 0 ? t : w1 - 1 - t;
            const size_t off = base + (size_t)xc * g.Dp;
            load_vec<NP>(Cvol + off, lane, C);
            if (!first) load_vec<NP>(Svol + off, lane, S);
            minq = sgm_step<NP>(Lq, C, minq, P1P1, g.P2, lane, npv);
#pragma unroll
            for (int k = 0; k < NP; ++k) S[k] = first ? pk_min(Lq[k], WALL2) : pk_addmin(S[k], Lq[k], WALL2);
            store_vec<NP>(Svol + off, lane, S);
        }
    } else {
        job = (int)(gw / w1);
        if (job >= njobs) return;
        job_rows(g, job % g.nstripes, ys, ye, keep);
        const int x0 = (int)(gw % w1);
        const int nrows = ye - ys;
        int xc = x0;
        for (int t = 0; t < nrows; ++t) {
            const int row = dy > 0 ? t : nrows - 1 - t;
            if (t > 0 && dx != 0) {
                xc += dx;
                bool wrap = false;
                if (xc >= w1) { xc = 0; wrap = true; }
                if (xc < 0) { xc = w1 - 1; wrap = true; }
                if (wrap) {
#pragma unroll
                    for (int k = 0; k < NP; ++k) Lq[k] = 0;
                    minq = 0;
                }
            }
            const size_t off = (((size_t)job * g.rowsMax + row) * w1 + xc) * g.Dp;
            load_vec<NP>(Cvol + off, lane, C);
            if (!first) load_vec<NP>(Svol + off, lane, S);
            minq = sgm_step<NP>(Lq, C, minq, P1P1, g.P2, lane, npv);
#pragma unroll
            for (int k = 0; k < NP; ++k) S[k] = first ? pk_min(Lq[k], WALL2) : pk_addmin(S[k], Lq[k], WALL2);
            store_vec<NP>(Svol + off, lane, S);
        }
    }
}

// ------------------------------------------------------------------------------------------
// k_wta: one warp per cost-domain pixel of a kept row (A.4): argmin (lowest d on ties,
// fully saturated column -> invalid), uniqueness, subpixel parabola, right-view map by
// atomicMin on key = minS<<16 | (0xffff - x): smallest cost wins, then the largest x, which is
// the pixel cv2's right-to-left loop would have kept.
// ------------------------------------------------------------------------------------------
template <int NP>
__global__ void __launch_bounds__(256) k_wta(const uint16_t* __restrict__ Svol, SgbmGeo g, int njobs, int img0,
                                             int16_t* __restrict__ raw, uint32_t* __restrict__ disp2key)
{
    const int lane = threadIdx.x & 31;
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int w1 = g.width1;
    const long long per_job = (long long)g.rowsMax * w1;
    const int job = (int)(gw / per_job);
    if (job >= njobs) return;
    int ys, ye, keep;
    job_rows(g, job % g.nstripes, ys, ye, keep);
    const int rem = (int)(gw % per_job);
    const int row = rem / w1, xc = rem % w1;
    const int y = ys + row;
    if (y >= ye || y < keep) return;
    const int b = job / g.nstripes;
    uint32_t S[NP];
    load_vec<NP>(Svol + (((size_t)job * g.rowsMax + row) * w1 + xc) * g.Dp, lane, S);
    const int dbase = 2 * NP * lane;
    uint32_t key = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < NP; ++k) {
        const int d = dbase + 2 * k;
        uint32_t lo = S[k] & 0xffffu, hi = S[k] >> 16;
        if (d < g.D) key = min(key, (lo << 16) | (uint32_t)d);
        if (d + 1 < g.D) key = min(key, (hi << 16) | (uint32_t)(d + 1));
    }
    key = __reduce_min_sync(0xffffffffu, key);
    const int minS = (int)(key >> 16), best = (int)(key & 0xffffu);
    const size_t pix = ((size_t)(img0 + b) * g.H + y) * g.W + (xc + g.minX1);
    if (minS >= (int)WALL) { if (lane == 0) raw[pix] = (int16_t)g.INV; return; }
    bool rej = false;
    if (g.mode == SVB_MODE_SGBM_3WAY) {
        // cv2's 3WAY vector loop: Sp[d] < (short)(100*minS/(100-u) + 1)  (see oracle/sgbm_oracle.c)
        if (g.uniq > 0) {
            const int th = (int)(int16_t)((100 * minS) / (100 - g.uniq) + 1);
#pragma unroll
            for (int k = 0; k < NP; ++k) {
                const int d = dbase + 2 * k;
                const int lo = (int)(int16_t)(S[k] & 0xffffu), hi = (int)(int16_t)(S[k] >> 16);
                rej |= (d < g.D && abs(d - best) > 1 && lo < th);
                rej |= (d + 1 < g.D && abs(d + 1 - best) > 1 && hi < th);
            }
        }
    } else {
        const int lim = minS * 100, f = 100 - g.uniq;
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            const int d = dbase + 2 * k;
            const int lo = (int)(S[k] & 0xffffu), hi = (int)(S[k] >> 16);
            rej |= (d < g.D && abs(d - best) > 1 && lo * f < lim);
            rej |= (d + 1 < g.D && abs(d + 1 - best) > 1 && hi * f < lim);
        }
    }
    rej = __any_sync(0xffffffffu, rej);
    if (rej) { if (lane == 0) raw[pix] = (int16_t)g.INV; return; }
    // neighbours of best for the parabola
    const uint32_t prev_last = __shfl_up_sync(0xffffffffu, S[NP - 1], 1) >> 16;
    const uint32_t next_first = __shfl_down_sync(0xffffffffu, S[0], 1) & 0xffffu;
    if (best >= dbase && best < dbase + 2 * NP) {
        int sm1 = 0, sp1 = 0;
#pragma unroll
        for (int j = 0; j < 2 * NP; ++j) {
            if (dbase + j == best) {
                sm1 = j == 0 ? (int)prev_last : (int)((j & 1) ? (S[j >> 1] & 0xffffu) : (S[(j - 1) >> 1] >> 16));
                sp1 = j == 2 * NP - 1 ? (int)next_first : (int)((j & 1) ? (S[(j + 1) >> 1] & 0xffffu) : (S[j >> 1] >> 16));
            }
        }
        int dsub = best * 16;
        if (best > 0 && best < g.D - 1) {
            const int den = max(sm1 + sp1 - 2 * minS, 1);
            dsub += ((sm1 - sp1) * 16 + den) / (2 * den);      // C division: truncates toward zero
        }
        raw[pix] = (int16_t)(dsub + g.minD * 16);
        const int x = xc + g.minX1;
        const int x2 = x - best - g.minD;
        atomicMin(disp2key + ((size_t)(img0 + b) * g.H + y) * g.W + x2, ((uint32_t)minS << 16) | (uint32_t)(0xffff - x));
    }
}

// LR consistency (A.4): invalidate iff both roundings of d disagree with the right-view map.
__global__ void __launch_bounds__(256) k_lrcheck(int16_t* __restrict__ raw, const uint32_t* __restrict__ disp2key,
                                                 SgbmGeo g, int nimg)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x + g.minX1;
    const int y = blockIdx.y, b = blockIdx.z;
    if (x >= g.minX1 + g.width1) return;
    const size_t rowoff = ((size_t)b * g.H + y) * g.W;
    const int d1 = raw[rowoff + x];
    if (d1 == g.INV) return;
    const int dl = d1 >> 4, dh = (d1 + 15) >> 4;
    const int xl = x - dl, xh = x - dh;
    bool f1 = false, f2 = false;
    if (xl >= 0 && xl < g.W) {
        const uint32_t k = disp2key[rowoff + xl];
        const int d2 = k == 0xffffffffu ? g.INV : (0xffff - (int)(k & 0xffffu)) - xl;
        f1 = d2 >= g.minD && abs(d2 - dl) > g.d12;
    }
    if (xh >= 0 && xh < g.W) {
        const uint32_t k = disp2key[rowoff + xh];
        const int d2 = k == 0xffffffffu ? g.INV : (0xffff - (int)(k & 0xffffu)) - xh;
        f2 = d2 >= g.minD && abs(d2 - dh) > g.d12;
    }
    if (f1 && f2) raw[rowoff + x] = (int16_t)g.INV;
}

__global__ void k_fill_s16(int16_t* p, size_t n, int16_t v)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------
struct svb_sgbm {
    svb_sgbm_params prm;
    size_t ws_limit = (size_t)24 << 30;
    SvbBuf ws;
    SvbBuf stage_in, stage_out;     // device staging for the *_host entry point
    int launches = 0;
    // optional per-stage timing with CUDA events on the launching stream (bench.py's roofline)
    bool profiling = false;
    struct Ev { int cat; cudaEvent_t a, b; };
    std::vector<Ev> evs;
    // last compute (for debug_stages)
    SgbmGeo last_geo{};
    size_t off_C = 0, off_S = 0, off_raw = 0;
    bool have_last = false;
};

static int make_geo(const svb_sgbm_params& p, int H, int W, SgbmGeo& g)
{
    if (H <= 0 || W <= 0) { svb_set_error("empty image (%dx%d)", W, H); return SVB_ERR_PARAM; }
    if (p.numDisparities <= 0) { svb_set_error("numDisparities must be > 0"); return SVB_ERR_PARAM; }
    if (p.blockSize < 1) { svb_set_error("blockSize must be >= 1"); return SVB_ERR_PARAM; }
    if (p.numDisparities % 16 != 0) {
        svb_set_error("numDisparities=%d: this engine requires a multiple of 16", p.numDisparities);
        return SVB_ERR_UNSUPPORTED;
    }
    if (p.mode == SVB_MODE_HH4) { svb_set_error("MODE_HH4 is not implemented"); return SVB_ERR_UNSUPPORTED; }
    if (p.mode < 0 || p.mode > 3) { svb_set_error("bad mode %d", p.mode); return SVB_ERR_PARAM; }
    g.H = H; g.W = W; g.D = p.numDisparities; g.minD = p.minDisparity;
    g.NP = (g.D + 63) / 64; g.Dp = 64 * g.NP;
    if (g.NP > 8) { svb_set_error("numDisparities=%d > 512 unsupported", g.D); return SVB_ERR_UNSUPPORTED; }
    const int maxD = g.minD + g.D;
    g.minX1 = std::max(maxD, 0);
    const int maxX1 = W + std::min(g.minD, 0);
    g.width1 = maxX1 - g.minX1;
    g.bs = p.blockSize; g.r = p.blockSize / 2;
    if (g.width1 <= g.r || g.width1 <= 0) {  // cv2 assert: width1 > SW2 (stereosgbm.cpp:511)
        svb_set_error("image too narrow: width - (minDisparity+numDisparities) must exceed blockSize/2");
        return SVB_ERR_PARAM;
    }
    if (W > 65535) { svb_set_error("width > 65535 unsupported"); return SVB_ERR_UNSUPPORTED; }
    g.ftzero = std::max(p.preFilterCap, 15) | 1;
    g.P1 = p.P1 > 0 ? p.P1 : 2;
    g.P2 = std::max(p.P2 > 0 ? p.P2 : 5, g.P1 + 1);
    g.uniq = p.uniquenessRatio >= 0 ? p.uniquenessRatio : 10;
    g.d12 = p.disp12MaxDiff > 0 ? p.disp12MaxDiff : 1;
    g.INV = (g.minD - 1) * 16;
    g.mode = p.mode;
    g.nstripes = p.mode == SVB_MODE_SGBM_3WAY ? 4 : 1;
    const int win = 2 * g.r + 1;
    const long long cmax = (long long)win * win * (2 * g.ftzero + 63);
    // L <= C + P2 must stay below cv2's MAX_COST (its int16 C+P2 buffer has the same limit)
    if (g.ftzero > 127 || cmax + g.P2 > 32766 || g.P1 > 32767) {
        svb_set_error("blockSize/preFilterCap/P2 overflow the 16-bit cost range "
                      "(max C %lld + P2 > 32766); cv2's int16 buffers wrap there too", cmax);
        return SVB_ERR_UNSUPPORTED;
    }
    if (g.uniq >= 100) { svb_set_error("uniquenessRatio must be < 100"); return SVB_ERR_PARAM; }
    g.rowsMax = 0;
    for (int n = 0; n < g.nstripes; ++n) {
        int ys, ye, keep; job_rows(g, n, ys, ye, keep);
        g.rowsMax = std::max(g.rowsMax, ye - ys);
    }
    g.Wh = (W + 1) / 2 + 4;
    return SVB_OK;
}

struct SgbmLayout { size_t ldesc, rdesc, C, S, raw, key, scratch, total; };
static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
static SgbmLayout make_layout(const SgbmGeo& g, int nb)
{
    SgbmLayout L; size_t o = 0;
    L.ldesc = o; o += al256((size_t)nb * g.H * g.W * 32);
    L.rdesc = o; o += al256((size_t)nb * g.H * 16 * g.Wh * 4);
    const size_t vol = (size_t)nb * g.nstripes * g.rowsMax * g.width1 * g.Dp * 2;
    L.C = o; o += al256(vol);
    L.S = o; o += al256(vol);
    L.raw = o; o += al256((size_t)nb * g.H * g.W * 2);
    L.key = o; o += al256((size_t)nb * g.H * g.W * 4);
    L.scratch = o; o += al256((size_t)nb * g.H * g.W * 8);
    L.total = o;
    return L;
}
static int pick_chunk(const svb_sgbm* h, const SgbmGeo& g, int B)
{
    const size_t per = make_layout(g, 1).total;
    size_t nb = h->ws_limit / per;
    if (nb < 1) nb = 1;
    return (int)std::min<size_t>(nb, (size_t)B);
}

extern "C" int svb_sgbm_create(const svb_sgbm_params* p, svb_sgbm_t** out)
{
    if (!p || !out) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    SgbmGeo g;
    // validate what does not depend on the image size with a generously wide dummy image
    int rc = make_geo(*p, 16, 65535, g);
    if (rc != SVB_OK) return rc;
    svb_sgbm* h = new (std::nothrow) svb_sgbm();
    if (!h) return SVB_ERR_NOMEM;
    h->prm = *p;
    *out = h;
    return SVB_OK;
}
extern "C" void svb_sgbm_destroy(svb_sgbm_t* h)
{
    if (!h) return;
    h->ws.release(); h->stage_in.release(); h->stage_out.release();
    delete h;
}
extern "C" size_t svb_sgbm_workspace_bytes(const svb_sgbm_t* h, int B, int H, int W)
{
    SgbmGeo g;
    if (!h || make_geo(h->prm, H, W, g) != SVB_OK) return 0;
    return make_layout(g, pick_chunk(h, g, B)).total;
}
extern "C" int svb_sgbm_set_workspace_limit(svb_sgbm_t* h, size_t bytes)
{
    if (!h) return SVB_ERR_PARAM;
    h->ws_limit = bytes ? bytes : ((size_t)24 << 30);
    return SVB_OK;
}
extern "C" int svb_sgbm_last_launches(const svb_sgbm_t* h) { return h ? h->launches : 0; }

enum { CAT_PREFILTER = 0, CAT_COST, CAT_PATH, CAT_WTA, CAT_LR, CAT_MEDIAN, CAT_SPECKLE, CAT_COUNT };
static const char* const kCatNames[CAT_COUNT] = {"k_prefilter", "k_cost", "k_path", "k_wta", "k_lrcheck", "k_median3_s16", "k_ccl_*"};
struct ProfScope {
    svb_sgbm* h; cudaStream_t st; int idx = -1;
    ProfScope(svb_sgbm* h_, int cat, cudaStream_t st_) : h(h_), st(st_)
    {
        if (!h->profiling) return;
        svb_sgbm::Ev e; e.cat = cat;
        cudaEventCreate(&e.a); cudaEventCreate(&e.b);
        cudaEventRecord(e.a, st);
        h->evs.push_back(e); idx = (int)h->evs.size() - 1;
    }
    ~ProfScope() { if (idx >= 0) cudaEventRecord(h->evs[idx].b, st); }
};

template <int NP>
static int run_chunk(svb_sgbm* h, const SgbmGeo& g, const SgbmLayout& L, const uint8_t* left, const uint8_t* right,
                     int nb, int16_t* disp, cudaStream_t st)
{
    char* ws = (char*)h->ws.p;
    uint32_t* ldesc = (uint32_t*)(ws + L.ldesc);
    uint32_t* rdesc = (uint32_t*)(ws + L.rdesc);
    uint16_t* Cv = (uint16_t*)(ws + L.C);
    uint16_t* Sv = (uint16_t*)(ws + L.S);
    int16_t* raw = (int16_t*)(ws + L.raw);
    uint32_t* key = (uint32_t*)(ws + L.key);
    int32_t* scratch = (int32_t*)(ws + L.scratch);
    const int njobs = nb * g.nstripes;

    {
        ProfScope ps(h, CAT_PREFILTER, st);
        k_prefilter<<<dim3(svb_ceil_div(g.W, 256), g.H, nb), 256, 0, st>>>(left, right, nb, g, ldesc, rdesc);
    }
    SVB_LAUNCH_CHECK(); h->launches++;

    // k_cost tiling: NW = TY + 2r warps, smem = (NW*bs + NW) * NP * 128 bytes
    {
        const int bs = 2 * g.r + 1;
        int TY = 16;
        auto smem_of = [&](int ty) { return (size_t)(ty + 2 * g.r) * (bs + 1) * NP * 128; };
        while (TY > 1 && (TY + 2 * g.r > 32 || smem_of(TY) > 200 * 1024)) --TY;
        if (TY + 2 * g.r > 32 || smem_of(TY) > 200 * 1024) {
            svb_set_error("blockSize %d too large for the cost kernel at numDisparities %d", g.bs, g.D);
            return SVB_ERR_UNSUPPORTED;
        }
        const int XSEG = 128;
        const size_t smem = smem_of(TY);
        SVB_CUDA(cudaFuncSetAttribute(k_cost<NP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dim3 grid(svb_ceil_div(g.width1, XSEG), svb_ceil_div(g.rowsMax, TY), njobs);
        ProfScope ps(h, CAT_COST, st);
        k_cost<NP><<<grid, 32 * (TY + 2 * g.r), smem, st>>>(ldesc, rdesc, g, TY, XSEG, Cv);
        SVB_LAUNCH_CHECK(); h->launches++;
    }

    // path directions (A.3): 3WAY {->, down, <-}; SGBM 5; HH 8
    static const int dirs[8][2] = {{1, 0}, {1, 1}, {0, 1}, {-1, 1}, {-1, 0}, {-1, -1}, {0, -1}, {1, -1}};
    static const int dirs3[3][2] = {{1, 0}, {0, 1}, {-1, 0}};
    const int npaths = g.mode == SVB_MODE_SGBM_3WAY ? 3 : (g.mode == SVB_MODE_HH ? 8 : 5);
    for (int k = 0; k < npaths; ++k) {
        const int dx = g.mode == SVB_MODE_SGBM_3WAY ? dirs3[k][0] : dirs[k][0];
        const int dy = g.mode == SVB_MODE_SGBM_3WAY ? dirs3[k][1] : dirs[k][1];
        const long long lines = (long long)njobs * (dy == 0 ? g.rowsMax : g.width1);
        ProfScope ps(h, CAT_PATH, st);
        k_path<NP><<<svb_ceil_div(lines * 32, 256), 256, 0, st>>>(Cv, Sv, g, dx, dy, k == 0, njobs);
        SVB_LAUNCH_CHECK(); h->launches++;
    }

    const size_t npix = (size_t)nb * g.H * g.W;
    k_fill_s16<<<svb_ceil_div(npix, 256), 256, 0, st>>>(raw, npix, (int16_t)g.INV);
    SVB_LAUNCH_CHECK(); h->launches++;
    SVB_CUDA(cudaMemsetAsync(key, 0xff, npix * 4, st));
    {
        const long long warps = (long long)njobs * g.rowsMax * g.width1;
        ProfScope ps(h, CAT_WTA, st);
        k_wta<NP><<<svb_ceil_div(warps * 32, 256), 256, 0, st>>>(Sv, g, njobs, 0, raw, key);
        SVB_LAUNCH_CHECK(); h->launches++;
    }
    {
        ProfScope ps(h, CAT_LR, st);
        k_lrcheck<<<dim3(svb_ceil_div(g.width1, 256), g.H, nb), 256, 0, st>>>(raw, key, g, nb);
    }
    SVB_LAUNCH_CHECK(); h->launches++;

    int rc;
    {
        ProfScope ps(h, CAT_MEDIAN, st);
        rc = svb_median3_s16_dev(raw, nb, g.H, g.W, disp, st);
    }
    if (rc != SVB_OK) return rc;
    h->launches++;
    if (h->prm.speckleWindowSize > 0) {
        ProfScope ps(h, CAT_SPECKLE, st);
        rc = svb_filter_speckles_s16_dev(disp, nb, g.H, g.W, g.INV, h->prm.speckleWindowSize,
                                         16 * h->prm.speckleRange, scratch, st);
        if (rc != SVB_OK) return rc;
        h->launches += 4;
    }
    return SVB_OK;
}

extern "C" int svb_sgbm_compute_dev(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right,
                                    int B, int H, int W, int16_t* disp, void* stream)
{
    if (!h || !left || !right || !disp) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    if (B <= 0) { svb_set_error("batch must be >= 1"); return SVB_ERR_PARAM; }
    SgbmGeo g;
    int rc = make_geo(h->prm, H, W, g);
    if (rc != SVB_OK) return rc;
    const int nb = pick_chunk(h, g, B);
    const SgbmLayout L = make_layout(g, nb);
    rc = h->ws.reserve(L.total);
    if (rc != SVB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    h->launches = 0;
    for (int b0 = 0; b0 < B; b0 += nb) {
        const int n = std::min(nb, B - b0);
        const size_t off = (size_t)b0 * H * W;
        switch (g.NP) {
#define CASE(N) case N: rc = run_chunk<N>(h, g, L, left + off, right + off, n, disp + off, st); break;
            CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
            default: svb_set_error("bad NP"); rc = SVB_ERR_UNSUPPORTED;
        }
        if (rc != SVB_OK) return rc;
    }
    h->last_geo = g; h->off_C = L.C; h->off_S = L.S; h->off_raw = L.raw; h->have_last = true;
    return SVB_OK;
}

extern "C" int svb_sgbm_compute_host(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right,
                                     int B, int H, int W, int16_t* disp)
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
    SVB_CUDA(cudaMemcpyAsync(dl, left, n, cudaMemcpyHostToDevice, 0));
    SVB_CUDA(cudaMemcpyAsync(dr, right, n, cudaMemcpyHostToDevice, 0));
    rc = svb_sgbm_compute_dev(h, dl, dr, B, H, W, (int16_t*)h->stage_out.p, 0);
    if (rc != SVB_OK) return rc;
    SVB_CUDA(cudaMemcpyAsync(disp, h->stage_out.p, 2 * n, cudaMemcpyDeviceToHost, 0));
    SVB_CUDA(cudaStreamSynchronize(0));
    return SVB_OK;
}

extern "C" int svb_sgbm_debug_stages(svb_sgbm_t* h, uint16_t* C_host, uint16_t* S_host, int16_t* raw_host)
{
    if (!h || !h->have_last) { svb_set_error("no previous compute"); return SVB_ERR_PARAM; }
    const SgbmGeo& g = h->last_geo;
    SVB_CUDA(cudaDeviceSynchronize());
    const size_t rowp = (size_t)g.width1 * g.Dp;
    uint16_t* tmp = (uint16_t*)malloc(rowp * 2);
    if (!tmp) return SVB_ERR_NOMEM;
    for (int which = 0; which < 2; ++which) {
        uint16_t* dst = which ? S_host : C_host;
        if (!dst) continue;
        const char* src = (const char*)h->ws.p + (which ? h->off_S : h->off_C);
        for (int n = 0; n < g.nstripes; ++n) {
            int ys, ye, keep; job_rows(g, n, ys, ye, keep);
            for (int y = std::max(ys, keep); y < ye; ++y) {
                cudaError_t e = cudaMemcpy(tmp, src + (((size_t)n * g.rowsMax + (y - ys)) * rowp) * 2, rowp * 2,
                                           cudaMemcpyDeviceToHost);
                if (e != cudaSuccess) { free(tmp); svb_set_error("memcpy: %s", cudaGetErrorString(e)); return SVB_ERR_CUDA; }
                for (int x = 0; x < g.width1; ++x)
                    memcpy(dst + ((size_t)y * g.width1 + x) * g.D, tmp + (size_t)x * g.Dp, (size_t)g.D * 2);
            }
        }
    }
    free(tmp);
    if (raw_host)
        SVB_CUDA(cudaMemcpy(raw_host, (const char*)h->ws.p + h->off_raw, (size_t)g.H * g.W * 2, cudaMemcpyDeviceToHost));
    return SVB_OK;
}

// ---- profiling hooks (bench.py): per-stage device time of all computes since enable/reset
extern "C" int svb_sgbm_set_profiling(svb_sgbm_t* h, int enable)
{
    if (!h) return SVB_ERR_PARAM;
    for (auto& e : h->evs) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    h->evs.clear();
    h->profiling = enable != 0;
    return SVB_OK;
}
extern "C" int svb_sgbm_get_profile(svb_sgbm_t* h, int max_stages, char* names /* max_stages x 32 */,
                                    float* ms, int* launches)
{
    if (!h || !names || !ms || !launches) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    SVB_CUDA(cudaDeviceSynchronize());
    const int n = std::min(max_stages, (int)CAT_COUNT);
    for (int i = 0; i < n; ++i) { ms[i] = 0.f; launches[i] = 0; strncpy(names + 32 * i, kCatNames[i], 31); names[32 * i + 31] = 0; }
    for (auto& e : h->evs) {
        float t = 0.f;
        SVB_CUDA(cudaEventElapsedTime(&t, e.a, e.b));
        if (e.cat < n) { ms[e.cat] += t; launches[e.cat] += e.cat == CAT_SPECKLE ? 4 : 1; }
    }
    return n;
}
