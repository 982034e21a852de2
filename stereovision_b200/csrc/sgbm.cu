// sgbm.cu -- semi-global block matching for sm_100a, bit-exact with
// cv2.StereoSGBM_create(...).compute (reference call sites depth_map.py:894-909,
// fused_depth_map.py:988-1004; arithmetic per SURVEY.md Appendix A.1-A.5 as
// corrected in oracle/sgbm_oracle.c).
//
// Pipeline (one CUDA stream, no host sync):
//   k_prefilter   x-Sobel + raw channel, Birchfield-Tomasi interval descriptors
//   k_cost        BT pixel cost (2 channels) + blockSize^2 box sum -> C volume (u16), one write
//   k_sweep<..,0> sweep 1 (top-down, left-to-right): up to 4 paths (-> \ | /) per pixel in ONE pass
//                 over C, partial sum S1 written once
//   k_sweep<..,1> sweep 2 (bottom-up / right-to-left): remaining paths, S = S1 + sum, then
//                 argmin / uniqueness / subpixel / right-view scatter-min fused (S never stored)
//   k_lrcheck     left-right consistency
//   median3 + speckle filter (imgproc.cu)
//
// HBM traffic per pair: C written once, read by each sweep; S1 written once, read once
// (5 volume passes; the direction-decomposed v1 made 25).
//
// Data layout in HBM: volumes are [job][row][xc][Dp] u16 with Dp = 64*NP >= D.  One pixel's
// disparity vector is HALF = Dp/2 u32 words; word w packs disparity index w (low half) and
// w + HALF (high half), so the d-1 / d+1 neighbours of BOTH halves sit in the adjacent words
// and the SGM recurrence needs no byte permutes.  A warp owns one pixel: lane l holds words
// NP*l .. NP*l+NP-1 in registers.
//
// k_sweep is a persistent wavefront kernel: a CTA takes a band of SW_R consecutive rows (one
// warp per row, each row at least two columns behind the row above so that its diagonal predecessors are ready), the
// L_r vectors of the row above travel through shared-memory rings, and the last row of a band
// hands its L_r to the next band (another CTA) through an L2-resident carry buffer guarded by
// per-band progress counters (release/acquire).  Bands are handed out by an atomic ticket in
// dependency order, so a waiting CTA only ever waits on a CTA that is already running.
#include "svb_common.cuh"
#include <thread>
#include <string.h>
#include <algorithm>
#include <type_traits>
#include <new>
#include <vector>
#include <string.h>

#define WALL 0x7fffu           // cv2's MAX_COST for the d=-1 / d=D neighbours and pad lanes
#define WALL2 0x7fff7fffu
#define FULLMASK 0xffffffffu

#ifndef SW_ROWS4
#define SW_ROWS4 18
#endif
// rows (= warps) per band
#ifndef SW_ROWS_SMALL
#define SW_ROWS_SMALL 30           // rows per band for numDisparities <= 128 (32 warps of <= 64 registers; 24 -> 30: 480p / D64 first sweep 6.3 -> 5.4 ms)
#endif
__host__ __device__ constexpr int sw_rows(int NP) { return NP <= 2 ? SW_ROWS_SMALL : (NP == 3 ? 22 : (NP == 4 ? SW_ROWS4 : (NP == 5 ? 16 : 8))); }
#ifndef SW_EXPERIMENT
#define SW_EXPERIMENT 0        // timing experiments only (results are wrong): 3 = no C/S1 loads, 4 = no S1 stores, 5 = both, 6 = no row hand-off waits
#endif
#ifndef SW_PLAIN_FLAG
// 1: a row publishes its progress counter with a plain volatile shared store after __syncwarp() instead of
// st.release.cta.  The data it guards are shared-memory stores of the same warp, issued earlier in program
// order; an SM performs a warp's shared-memory instructions in issue order, and the reader uses ld.acquire.
// The release fence also waited for the row's outstanding cp.async / global stores (15.3 -> 14.4 ms per 32
// pairs in the first sweep).  0 restores the fence.
#define SW_PLAIN_FLAG 1
#endif
#ifndef SW_FMA_ADDS
#define SW_FMA_ADDS 1          // P1 and sum additions of the recurrence as IMADs (FMA pipe) instead of ALU-pipe adds
#endif
#ifndef SW_SEAM_FMA
#define SW_SEAM_FMA 0          // 1: seam fix-ups of the disparity neighbours as IMAD / IMAD.HI instead of PRMTs (measured: 14.1 -> 15.8 ms, the multiply sits on the recurrence's critical path)
#endif
#ifndef SW_WAVE
#define SW_WAVE 1              // 0: k_sweep for every first sweep (A/B timing of k_wave)
#endif
#ifndef SW_UNROLL2
#define SW_UNROLL2 0           // 1: the rows' column loop unrolled by two (experiment)
#endif
#ifndef SW_HLAST
#define SW_HLAST 1             // 0: k_sweep for every last sweep (A/B timing of k_hlast)
#endif
#ifndef SW_SLEEP_NS
#define SW_SLEEP_NS 40         // back-off while a row waits for its neighbour's progress counter
#endif
// ring slots per path between two rows (>= 4; every extra slot is a column of slack between rows)
// Measured at 1080p / D256 / HH, 32 pairs (ms per launch): the last sweep with (ring, staging slots) = (4,4) 23.7,
// (4,3) 22.6, (5,3) 22.2, (6,2) 21.7: with 4 slots a row has 2 columns of slack to its neighbours and polls on every
// column (ncu: 4 polls per column on either side); the staging ring of C and S1 gives its shared memory to the
// inter-row rings.  The first sweep (it also writes S1) wants the deeper staging: (6,4) 14.2, (6,3) 14.6, (6,2) 14.9.
#ifndef SW_RING_FIRST
#define SW_RING_FIRST 6
#endif
#ifndef SW_RING_LAST
#define SW_RING_LAST 6
#endif
// NP <= 2 (numDisparities <= 128) leaves most of the shared memory unused: deeper rings and staging there
#ifndef SW_RING_SMALL
#define SW_RING_SMALL 6
#define SW_RING_SMALL_LAST 4
#define SW_STAGE_SMALL 4
#endif
__host__ __device__ constexpr int sw_ring(int NP, bool last)
{
    return NP <= 2 ? (last ? SW_RING_SMALL_LAST : SW_RING_SMALL) : (NP <= 4 ? (last ? SW_RING_LAST : SW_RING_FIRST) : 4);
}
// slots of the per-row cp.async staging ring for C (and S1 in the last sweep); prefetch distance = slots - 1
#ifndef SW_STAGE_LAST
#define SW_STAGE_LAST 2
#endif
#ifndef SW_STAGE_FIRST
#define SW_STAGE_FIRST 4
#endif
__host__ __device__ constexpr int sw_stage(int NP, bool last) { return NP <= 2 ? SW_STAGE_SMALL : (NP <= 4 ? (last ? SW_STAGE_LAST : SW_STAGE_FIRST) : (last ? 3 : 4)); }
__host__ __device__ constexpr int sw_roww(int NP) { return NP * 32 + 4; }     // ring row: the vector, its minimum, pad to 16 B
__host__ __device__ constexpr size_t sw_smem_words(int NP, bool last)
{
    return (size_t)(sw_rows(NP) + 1) * 3 * sw_ring(NP, last) * sw_roww(NP) +
           (size_t)sw_rows(NP) * sw_stage(NP, last) * NP * 32 * (last ? 2 : 1) + (last ? (size_t)sw_rows(NP) * NP * 32 : 0);
}
#define SW_FLAG_EVERY 8        // carry progress is published every this many columns
#define SW_SPIN_LIMIT (1 << 24)
#ifndef SVB_HOST_PIPE
#define SVB_HOST_PIPE 4        // sub-ranges per chunk in the *_host entry points (upload / per-image kernels / download overlap)
#endif

struct SgbmGeo {
    int H, W, D, Dp, NP, HALF, minD, minX1, width1, bs, r, ftzero;
    int P1, P2, uniq, d12, INV, mode, nstripes, rowsMax;
    int cn;                    // image channels (1, or 3: interleaved colour -- cv2 sums the BT cost of every channel)
    int Wp;                    // words per plane row of the right descriptors
    unsigned thMul;            // floor(2^32 / (100 - uniq)) + 1: exact division of the uniqueness threshold (0: divisor <= 1)
    long long rowPitch, imgPitch;   // bytes between image rows / between images of the INPUT batch (dense: W*cn, H*W*cn)
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
// interval bounds (A.1 step 3).
// Left: one 32-byte record per pixel {u, -u, u0, 1 - u1} x 2 channels, each duplicated into both
// halves of a u32.
// Right: 6 planes per row {-v1, v0, -v} x 2 channels; entry(xr) = a[xr] | a[xr - HALF] << 16,
// i.e. the operands of disparity index w (low) and w + HALF (high) for left pixel x = xr + w.
// ------------------------------------------------------------------------------------------
// img points at channel ch of pixel (0,0); cn = pixel stride in bytes
__device__ __forceinline__ void bt_channels(const uint8_t* img, long long pitch, int H, int W, int cn, int y, int x,
                                            int ftzero, int& p0, int& p1)
{
    if (x <= 0 || x >= W - 1) { p0 = p1 = ftzero; return; }
    int yn = y > 0 ? y - 1 : y, ys = y < H - 1 ? y + 1 : y;
    const uint8_t *r = img + y * pitch, *rn = img + yn * pitch, *rs = img + ys * pitch;
    const int e = (x + 1) * cn, w = (x - 1) * cn;
    int g = 2 * ((int)r[e] - r[w]) + ((int)rn[e] - rn[w]) + ((int)rs[e] - rs[w]);
    g = max(-ftzero, min(ftzero, g));
    p0 = g + ftzero; p1 = r[x * cn];
}

__global__ void __launch_bounds__(256) k_prefilter(const uint8_t* __restrict__ left, const uint8_t* __restrict__ right,
                                                   int nimg, int ch, SgbmGeo g, uint32_t* __restrict__ ldesc,
                                                   uint32_t* __restrict__ rdesc)
{
    // A block handles 256 consecutive left pixels / right entries of one image row.  The two BT channels of every pixel
    // the block needs -- the left window, the right window and the right window HALF pixels further left, one halo
    // pixel on either side -- are computed ONCE into shared memory (three passes of bt_channels per thread instead of
    // nine: the kernel was bound by its byte loads, not by its 60 bytes of output per pixel), then every thread builds
    // its record and its entry from three neighbouring values.
    __shared__ short sv[3][2][258];                          // [left | right | right - HALF][channel][x0 - 1 + i, clamped]
    const int x0 = blockIdx.x * blockDim.x, t = threadIdx.x, x = x0 + t;
    const int y = blockIdx.y, b = blockIdx.z;
    const int H = g.H, W = g.W;
    const uint8_t* imgL = left + b * g.imgPitch + ch;
    const uint8_t* imgR = right + b * g.imgPitch + ch;
    for (int i = t; i < 258; i += 256) {
#pragma unroll
        for (int w = 0; w < 3; ++w) {
            const int px = x0 - 1 + i - (w == 2 ? g.HALF : 0);
            int p0, p1;
            bt_channels(w == 0 ? imgL : imgR, g.rowPitch, H, W, g.cn, y, min(max(px, 0), W - 1), g.ftzero, p0, p1);
            sv[w][0][i] = (short)p0;
            sv[w][1][i] = (short)p1;
        }
    }
    __syncthreads();
    if (x < W) {
        uint32_t rec[8];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int p[3] = {sv[0][c][t], sv[0][c][t + 1], sv[0][c][t + 2]};
            int u = p[1];
            int ua = x > 0 ? (u + p[0]) / 2 : u;
            int ub = x < W - 1 ? (u + p[2]) / 2 : u;
            int u0 = min(min(ua, ub), u), u1 = max(max(ua, ub), u);
            rec[c * 4 + 0] = pk_dup((uint32_t)u);
            rec[c * 4 + 1] = pk_dup((uint32_t)(-u));
            rec[c * 4 + 2] = pk_dup((uint32_t)u0);
            rec[c * 4 + 3] = pk_dup((uint32_t)(1 - u1));     // v - u1 = ~(-v) + (1 - u1)
        }
        uint4* dst = reinterpret_cast<uint4*>(ldesc + ((size_t)(b * H + y) * W + x) * 8);   // 32-byte records: two 128-bit stores
        dst[0] = make_uint4(rec[0], rec[1], rec[2], rec[3]);
        dst[1] = make_uint4(rec[4], rec[5], rec[6], rec[7]);
    }
    if (x < g.Wp) {
        // entry x of the six planes = a[x] | a[x - HALF] << 16 (ONE coalesced 32-bit store per plane and entry);
        // halves whose pixel lies outside the image are 0.
        uint32_t ent[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int xr = half ? x - g.HALF : x;
            if (xr < 0 || xr >= W) continue;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int p[3] = {sv[1 + half][c][t], sv[1 + half][c][t + 1], sv[1 + half][c][t + 2]};
                int v = p[1];
                int va = xr < W - 1 ? (v + p[2]) / 2 : v;
                int vb = xr > 0 ? (v + p[0]) / 2 : v;
                int v0 = min(min(va, vb), v), v1 = max(max(va, vb), v);
                const uint32_t ops[3] = {(uint32_t)(-v1) & 0xffffu, (uint32_t)v0 & 0xffffu, (uint32_t)(-v) & 0xffffu};
#pragma unroll
                for (int o = 0; o < 3; ++o) ent[c * 3 + o] |= ops[o] << (16 * half);
            }
        }
        uint32_t* base = rdesc + (size_t)(b * H + y) * 6 * g.Wp + x;
#pragma unroll
        for (int q = 0; q < 6; ++q) base[(size_t)q * g.Wp] = ent[q];
    }
}

// ------------------------------------------------------------------------------------------
// cp.async helpers (LDGSTS)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
// one lane's NP-word slice of a pixel vector (16-byte copies when NP is a multiple of 4)
template <int NP>
__device__ __forceinline__ void cp_async_vec(uint32_t* smem_dst, const uint32_t* gsrc)
{
    if constexpr (NP % 4 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 4; ++i) cp_async16(smem_dst + 4 * i, gsrc + 4 * i);
    } else {
#pragma unroll
        for (int i = 0; i < NP; ++i) cp_async4(smem_dst + i, gsrc + i);
    }
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// ------------------------------------------------------------------------------------------
// Vector helpers for the sweep: lane l <-> words NP*l .. NP*l+NP-1 of a pixel's vector.
// ------------------------------------------------------------------------------------------
template <int NP, typename LD>
__device__ __forceinline__ void vec_ld(const uint32_t* p, int lane, uint32_t (&v)[NP], LD ld4)
{
    const uint32_t* w = p + NP * lane;
    if constexpr (NP % 4 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 4; ++i) {
            uint4 t = ld4(reinterpret_cast<const uint4*>(w) + i);
            v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
        }
    } else if constexpr (NP % 2 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 2; ++i) {
            uint2 t = ld4(reinterpret_cast<const uint2*>(w) + i);
            v[2 * i] = t.x; v[2 * i + 1] = t.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < NP; ++i) v[i] = ld4(w + i);
    }
}
struct LdPlain {
    __device__ __forceinline__ uint4 operator()(const uint4* p) const { return *p; }
    __device__ __forceinline__ uint2 operator()(const uint2* p) const { return *p; }
    __device__ __forceinline__ uint32_t operator()(const uint32_t* p) const { return *p; }
};
struct LdNc {     // read-only data written by an earlier kernel
    __device__ __forceinline__ uint4 operator()(const uint4* p) const { return __ldg(p); }
    __device__ __forceinline__ uint2 operator()(const uint2* p) const { return __ldg(p); }
    __device__ __forceinline__ uint32_t operator()(const uint32_t* p) const { return __ldg(p); }
};
struct LdCg {     // data written by another CTA of this kernel: bypass L1
    __device__ __forceinline__ uint4 operator()(const uint4* p) const { return __ldcg(p); }
    __device__ __forceinline__ uint2 operator()(const uint2* p) const { return __ldcg(p); }
    __device__ __forceinline__ uint32_t operator()(const uint32_t* p) const { return __ldcg(p); }
};
template <int NP>
__device__ __forceinline__ void vec_st(uint32_t* p, int lane, const uint32_t (&v)[NP])
{
    uint32_t* w = p + NP * lane;
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

// Shared-memory vector accesses by 32-bit shared-space address (no generic-address arithmetic in the rows' loop).
// volatile: kept in order with the progress-flag accesses (also volatile asm); the immediate offset is folded.
template <int NP, int OFS, int I = 0>
__device__ __forceinline__ void lds_vec(uint32_t sa, uint32_t (&v)[NP])
{
    if constexpr (NP % 4 == 0) {
        if constexpr (I < NP / 4) {
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4 + %5];"
                         : "=r"(v[4 * I]), "=r"(v[4 * I + 1]), "=r"(v[4 * I + 2]), "=r"(v[4 * I + 3]) : "r"(sa), "n"(OFS + 16 * I));
            lds_vec<NP, OFS, I + 1>(sa, v);
        }
    } else if constexpr (NP % 2 == 0) {
        if constexpr (I < NP / 2) {
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2 + %3];" : "=r"(v[2 * I]), "=r"(v[2 * I + 1]) : "r"(sa), "n"(OFS + 8 * I));
            lds_vec<NP, OFS, I + 1>(sa, v);
        }
    } else if constexpr (I < NP) {
        asm volatile("ld.shared.u32 %0, [%1 + %2];" : "=r"(v[I]) : "r"(sa), "n"(OFS + 4 * I));
        lds_vec<NP, OFS, I + 1>(sa, v);
    }
}
template <int NP, int OFS, int I = 0>
__device__ __forceinline__ void sts_vec(uint32_t sa, const uint32_t (&v)[NP])
{
    if constexpr (NP % 4 == 0) {
        if constexpr (I < NP / 4) {
            asm volatile("st.shared.v4.u32 [%0 + %1], {%2, %3, %4, %5};" ::"r"(sa), "n"(OFS + 16 * I), "r"(v[4 * I]),
                         "r"(v[4 * I + 1]), "r"(v[4 * I + 2]), "r"(v[4 * I + 3]));
            sts_vec<NP, OFS, I + 1>(sa, v);
        }
    } else if constexpr (NP % 2 == 0) {
        if constexpr (I < NP / 2) {
            asm volatile("st.shared.v2.u32 [%0 + %1], {%2, %3};" ::"r"(sa), "n"(OFS + 8 * I), "r"(v[2 * I]), "r"(v[2 * I + 1]));
            sts_vec<NP, OFS, I + 1>(sa, v);
        }
    } else if constexpr (I < NP) {
        asm volatile("st.shared.u32 [%0 + %1], %2;" ::"r"(sa), "n"(OFS + 4 * I), "r"(v[I]));
        sts_vec<NP, OFS, I + 1>(sa, v);
    }
}
template <int OFS>
__device__ __forceinline__ uint32_t lds_u32(uint32_t sa)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1 + %2];" : "=r"(v) : "r"(sa), "n"(OFS));
    return v;
}
template <int OFS>
__device__ __forceinline__ void sts_u32(uint32_t sa, uint32_t v)
{
    asm volatile("st.shared.u32 [%0 + %1], %2;" ::"r"(sa), "n"(OFS), "r"(v));
}

// ------------------------------------------------------------------------------------------
// k_cost: one CTA walks DOWN a strip of columns of one job (a row segment of it).  Per image
// row one elected thread asks the TMA for two tiles (cp.async.bulk.tensor, one row ahead,
// completion on an mbarrier): the 6 right-descriptor planes over the strip's disparity window
// and the strip's left records.  Every warp then computes the BT pixel-cost vectors (A.1
// step 4) of two strip columns and parks them in a shared buffer; after ONE block barrier every
// warp sums 2r+1 neighbouring columns (horizontal box, A.2), pushes the sum into a per-column
// ring of the last bs rows and keeps the vertical box sum C(y) = C(y-1) + hs(y+r) - hs(y-r-1)
// in registers -> HBM.  Replicate borders: columns clamp in the cost domain, rows clamp into
// [ys, H-1].
// Lane mappings: the pixel-cost phase uses (reg k, lane l) <-> word 32k + l (descriptor loads
// are contiguous across the warp); the box-sum phase is element-wise and uses the sweep's
// mapping (lane l <-> words NP*l..) so that shared and global accesses are 128-bit.
// ------------------------------------------------------------------------------------------
#define CK_WARPS 16
#ifndef CK_GROUPED
#define CK_GROUPED 1           // numDisparities <= 64: k_costg (four columns per warp pass) instead of k_cost
#endif
#ifndef CK_BUDGET_KB
#define CK_BUDGET_KB 200       // shared memory per CTA (experiment knob: 109 = two CTAs per SM at numDisparities 256 / blockSize 5)
#endif
#ifndef CK_SHARE
#define CK_SHARE 4             // k_cost: from this many words per lane on, a warp's two adjacent columns share one load of the right
                               // descriptor planes (1080p/D256 13.87 -> 13.77 ms, 480p/D320 4.11 -> 3.80; slower at D128: 8.06 -> 8.76)
#endif
#ifndef CK_UNROLL2_NP
#define CK_UNROLL2_NP 4        // k_cost: row loop unrolled by two (constant staging buffer index) up to this many words per lane
#endif
#ifndef CK_BSCONST
#define CK_BSCONST 1           // k_cost / k_costg instances with blockSize 3 / 5 / 7 / 9 as a compile-time constant
#endif
#ifndef CK_MAXSEG
#define CK_MAXSEG 16           // most row segments per job (1 = never cut a job's rows)
#endif
#define CK_PC 32               // pixel-cost columns per strip (2 per warp)
// staged entries per plane: the disparity window of the strip + 8 so that the window can start on a
// 16-byte boundary (TMA needs the inner coordinate's byte offset 16-byte aligned -- measured, tools/tma_test.cu)
__host__ __device__ constexpr int ck_se(int NP) { return 32 * NP + CK_PC + 8; }
__host__ __device__ constexpr int ck_nblk(int NP) { return ck_se(NP) > 256 ? 2 : 1; }   // TMA boxes are <= 256 wide
__host__ __device__ constexpr int ck_blkw(int NP) { return (6 * (ck_se(NP) / ck_nblk(NP)) + 31) & ~31; }   // words per staged box, 128-byte padded

// Birchfield-Tomasi pixel cost of one packed word (two disparities) from the left record {u, -u, u0, 1 - u1} x 2 channels
// and the six right planes {-v1, v0, -v} x 2 channels (A.1 step 4):
// channel 0: c0 = max(0, u - v1, v0 - u); c1 = max(0, v - u1, u0 - v); channel 1 (raw intensity) the same, cost >> 2.
// Operands are stored negated where that saves the subtraction (a packed negation would cost two more instructions).
__device__ __forceinline__ uint32_t bt_pix(const uint4& l0, const uint4& l1, uint32_t v1a, uint32_t v0a, uint32_t nva,
                                           uint32_t v1b, uint32_t v0b, uint32_t nvb)
{
    uint32_t a = __vadd2(l0.x, v1a);                         // u - v1
    uint32_t c0 = __viaddmax_s16x2_relu(v0a, l0.y, a);
    uint32_t bb = __vadd2(~nva, l0.w);                       // v - u1
    uint32_t c1 = __viaddmax_s16x2_relu(l0.z, nva, bb);
    uint32_t pix = __vminu2(c0, c1);
    a = __vadd2(l1.x, v1b);
    c0 = __viaddmax_s16x2_relu(v0b, l1.y, a);
    bb = __vadd2(~nvb, l1.w);
    c1 = __viaddmax_s16x2_relu(l1.z, nvb, bb);
    return pix + ((__vminu2(c0, c1) >> 2) & 0x3fff3fffu);
}

template <int V> struct CkInt { __device__ constexpr operator int() const { return V; } };   // a compile-time int as a lambda argument

struct CostArgs {
    int OC;                    // output columns per strip = CK_PC - 2r (or less when the ring is large)
    int nstrips, rseg, nsegs;  // strips per row, rows per segment, segments per job
    int accum;                 // add to the stored C (colour input: one pass per channel)
    int job0;                  // first job of this launch (the per-image kernels run on sub-ranges of a chunk)
};

// XTRA: the rare variants (colour accumulation passes, MODE_HH4's zero rows) live in their own instance so that
// the plain kernel keeps its code and register allocation (13.8 vs 14.5 ms per 32 pairs at 1080p/D256 otherwise)
// BS: blockSize as a compile-time constant (0 = read it from the geometry): the box loop unrolls, its loop control and
// register copies go away (k_cost runs ~75 % issue-bound)
template <int NP, bool FULL, bool XTRA, int BS>
__global__ void __launch_bounds__(CK_WARPS * 32) k_cost(const __grid_constant__ CUtensorMap tmR,
                                                        const __grid_constant__ CUtensorMap tmL,
                                                        SgbmGeo g, CostArgs ca, uint32_t* __restrict__ vol)
{
    constexpr int VW = NP * 32, SE = ck_se(NP), NBLK = ck_nblk(NP), BW = SE / NBLK, BLKW = ck_blkw(NP);
    extern __shared__ __align__(128) uint32_t smem[];
    __shared__ __align__(8) uint64_t mbar[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r = BS ? BS / 2 : g.r, bs = BS ? BS : 2 * r + 1, OC = ca.OC, PC = OC + 2 * r;
    uint32_t* stage = smem;                                  // [2][NBLK][6][BW] (each box padded to BLKW words)
    uint32_t* lstage = stage + 2 * NBLK * BLKW;                   // [2][CK_PC][8]
    uint32_t* pixbuf = lstage + 2 * CK_PC * 8;               // [2][CK_PC][VW]
    uint32_t* ring = pixbuf + 2 * CK_PC * VW;                // [OC][bs][VW]

    const int strip = blockIdx.x % ca.nstrips, seg = blockIdx.x / ca.nstrips;
    const int job = blockIdx.y + ca.job0, b = job / g.nstripes, n = job % g.nstripes;
    int ys, ye, keep;
    job_rows(g, n, ys, ye, keep);
    const int ya = ys + seg * ca.rseg, yb = min(ya + ca.rseg, ye);   // output rows of this CTA
    if (ya >= yb) return;
    const int xs = strip * OC;                               // first output column (cost domain)
    const int w1 = g.width1;

    // pix column c (0..PC-1) <-> cost column clamp(xs - r + c); image column = + minX1.
    // The staged window is NOT clamped (TMA zero-fills outside the tensor); indices into it are.
    // right entries cover xr in [xbase, xbase + SE): xbase = (xs - r + minX1 - minD) - (HALF - 1)
    const int xbase0 = xs - r + g.minX1 - g.minD - (g.HALF - 1);
    const int xbase = xbase0 & ~3;                           // floor to a multiple of 4 entries (16 bytes)
    const int xl0 = xs - r + g.minX1;                        // image column of left record 0

    auto issue_row = [&](int v, int buf) {                   // one thread
        const int yy = min(max(v, ys), g.H - 1);
        mbar_expect_tx(&mbar[buf], (6 * SE + CK_PC * 8) * 4);
#pragma unroll
        for (int q = 0; q < NBLK; ++q)
            tma_load_2d(stage + (buf * NBLK + q) * BLKW, &tmR, xbase + q * BW, (b * g.H + yy) * 6, &mbar[buf]);
        tma_load_2d(lstage + buf * CK_PC * 8, &tmL, xl0 * 8, b * g.H + yy, &mbar[buf]);
    };

    if (threadIdx.x == 0) {
        mbar_init(&mbar[0], 1);
        mbar_init(&mbar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const int v0 = ya - r, v1 = yb - 1 + r;                  // virtual input rows (clamped on use)
    if (threadIdx.x == 0) {
        issue_row(v0, 0);
        if (v0 + 1 <= v1) issue_row(v0 + 1, 1);
    }

    // pad masks: pix phase (w = 32k + lane) is not needed (masking happens at the store);
    // box phase mapping w = NP*lane + k
    uint32_t padm[NP];
#pragma unroll
    for (int k = 0; k < NP; ++k) {
        const int w = NP * lane + k;
        padm[k] = (w >= g.D ? 0xffffu : 0u) | (w + g.HALF >= g.D ? 0xffff0000u : 0u);
    }
    uint32_t acc[2][NP];
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int k = 0; k < NP; ++k) acc[c][k] = 0;

    // this warp's two pix columns: clamped column, index into the staged windows
    int eoff[2], loff[2];
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
        const int c = 2 * warp + cc;
        const int xq = min(max(xs - r + c, 0), w1 - 1);
        const int cq = xq - (xs - r);                        // column inside the window
        eoff[cc] = cq + g.HALF - 1 - lane + (xbase0 - xbase);   // entry of word w = lane; word 32k+lane: - 32k
        loff[cc] = cq * 8;
    }
    const int oc0 = 2 * warp;
    const bool two = oc0 + 1 < OC;
    const bool pair = 2 * warp + 1 < PC && eoff[1] == eoff[0] + 1;   // not at a clamped image border
    int slot = 0;
    // volume vector of this warp's first output column in the row being emitted (the first one is ya, at it == bs - 1)
    uint32_t* drow = vol + (((size_t)job * g.rowsMax + (ya - ys)) * w1 + xs + oc0) * VW;

    // one input row.  Up to 4 words per lane the row loop is unrolled by two and the staging buffer index is a constant
    // in each copy (720p/D128 7.27 -> 6.75 ms); with more words the doubled loop body no longer pays (480p/D320 3.47 -> 3.77)
    auto do_row = [&](const int v, auto bufc) {
        const int buf = bufc;
        const int it = v - v0;
        const uint32_t* st = stage + buf * NBLK * BLKW;
        const uint32_t* ls = lstage + buf * CK_PC * 8;
        uint32_t* pb = pixbuf + buf * CK_PC * VW;
        mbar_wait(&mbar[buf], (it >> 1) & 1);
        // ---- pixel costs of this warp's two strip columns
        if (CK_SHARE && NP >= CK_SHARE && pair) {
            // entry(c + 1, w + 1) == entry(c, w): the lane that holds the planes of (c, 32k + lane) uses them again for
            // word 32k + lane + 1 of the next column; only word 0 of that column (lane 31, last pass) needs its own load
            const int c = 2 * warp;
            const uint4 a0 = *reinterpret_cast<const uint4*>(ls + loff[0]);
            const uint4 a1 = *reinterpret_cast<const uint4*>(ls + loff[0] + 4);
            const uint4 b0 = *reinterpret_cast<const uint4*>(ls + loff[1]);
            const uint4 b1 = *reinterpret_cast<const uint4*>(ls + loff[1] + 4);
#pragma unroll
            for (int k = 0; k < NP; ++k) {
                int e = eoff[0] - 32 * k;
                const uint32_t* sp = st;
                if (NBLK > 1) { if (e >= BW) { e -= BW; sp += BLKW; } }
                uint32_t v1a = sp[0 * BW + e], v0a = sp[1 * BW + e], nva = sp[2 * BW + e];
                uint32_t v1b = sp[3 * BW + e], v0b = sp[4 * BW + e], nvb = sp[5 * BW + e];
                pb[c * VW + 32 * k + lane] = bt_pix(a0, a1, v1a, v0a, nva, v1b, v0b, nvb);
                if (k == NP - 1) {
                    int e2 = lane == 31 ? eoff[0] + 32 : eoff[0] - 32 * k;
                    sp = st;
                    if (NBLK > 1) { if (e2 >= BW) { e2 -= BW; sp += BLKW; } }
                    v1a = sp[0 * BW + e2]; v0a = sp[1 * BW + e2]; nva = sp[2 * BW + e2];
                    v1b = sp[3 * BW + e2]; v0b = sp[4 * BW + e2]; nvb = sp[5 * BW + e2];
                }
                const int wb = 32 * k + lane + 1;
                pb[(c + 1) * VW + (wb == VW ? 0 : wb)] = bt_pix(b0, b1, v1a, v0a, nva, v1b, v0b, nvb);
            }
        } else
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int c = 2 * warp + cc;
            if (c < PC) {
                const uint4 l0 = *reinterpret_cast<const uint4*>(ls + loff[cc]);
                const uint4 l1 = *reinterpret_cast<const uint4*>(ls + loff[cc] + 4);
#pragma unroll
                for (int k = 0; k < NP; ++k) {
                    int e = eoff[cc] - 32 * k;
                    const uint32_t* sp = st;
                    if (NBLK > 1) { if (e >= BW) { e -= BW; sp += BLKW; } }
                    const uint32_t v1a = sp[0 * BW + e], v0a = sp[1 * BW + e], nva = sp[2 * BW + e];
                    const uint32_t v1b = sp[3 * BW + e], v0b = sp[4 * BW + e], nvb = sp[5 * BW + e];
                    pb[c * VW + 32 * k + lane] = bt_pix(l0, l1, v1a, v0a, nva, v1b, v0b, nvb);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0 && v + 2 <= v1) issue_row(v + 2, buf);
        // ---- horizontal box over 2r+1 columns, vertical running sum, store
        if (oc0 < OC) {
            uint32_t hs0[NP], hs1[NP], t[NP];
            const uint32_t* pcol = pb + oc0 * VW;
            vec_ld<NP>(pcol, lane, hs0, LdPlain());
            uint32_t first[NP];
#pragma unroll
            for (int k = 0; k < NP; ++k) first[k] = hs0[k];
#pragma unroll
            for (int j = 1; j < (BS ? BS : bs); ++j) {
                vec_ld<NP>(pcol + j * VW, lane, t, LdPlain());
#pragma unroll
                for (int k = 0; k < NP; ++k) hs0[k] += t[k];
            }
            if (two) {
                vec_ld<NP>(pcol + bs * VW, lane, t, LdPlain());
#pragma unroll
                for (int k = 0; k < NP; ++k) hs1[k] = hs0[k] - first[k] + t[k];
            }
            const bool full = it >= bs;                       // ring holds hs(v - bs) in `slot`
            const int y = v - r;                              // output row once the window is complete
            const bool emit = it >= bs - 1;
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                if (cc == 1 && !two) break;
                const int oc = oc0 + cc;
                uint32_t* rg = ring + ((size_t)oc * bs + slot) * VW;
                const int xc = xs + oc;
                if (full) {
                    vec_ld<NP>(rg, lane, t, LdPlain());
#pragma unroll
                    for (int k = 0; k < NP; ++k) acc[cc][k] -= t[k];
                }
                if (cc == 0) {
                    vec_st<NP>(rg, lane, hs0);
#pragma unroll
                    for (int k = 0; k < NP; ++k) acc[0][k] += hs0[k];
                } else {
                    vec_st<NP>(rg, lane, hs1);
#pragma unroll
                    for (int k = 0; k < NP; ++k) acc[1][k] += hs1[k];
                }
                if (emit && xc < w1) {
                    uint32_t* dst = drow + cc * VW;
                    uint32_t o[NP];
                    if constexpr (XTRA) {
                        // MODE_HH4: cv2 leaves the rows whose window reaches below the image at cost 0 (oracle/sgbm_oracle.c)
                        const bool zrow = g.mode == SVB_MODE_HH4 && y + r > g.H - 1 && y > 0;
                        if (ca.accum) vec_ld<NP>(dst, lane, t, LdPlain());
#pragma unroll
                        for (int k = 0; k < NP; ++k) {
                            const uint32_t cv = zrow ? 0u : (ca.accum ? acc[cc][k] + t[k] : acc[cc][k]);
                            o[k] = FULL ? cv : ((cv & ~padm[k]) | (WALL2 & padm[k]));
                        }
                    } else {
#pragma unroll
                        for (int k = 0; k < NP; ++k) o[k] = FULL ? acc[cc][k] : ((acc[cc][k] & ~padm[k]) | (WALL2 & padm[k]));
                    }
                    vec_st<NP>(dst, lane, o);
                }
            }
        }
        if (it >= bs - 1) drow += (size_t)w1 * VW;
        slot = slot + 1 == bs ? 0 : slot + 1;
    };
    if constexpr (NP <= CK_UNROLL2_NP) {
        for (int v = v0; v <= v1; v += 2) {
            do_row(v, CkInt<0>());
            if (v + 1 <= v1) do_row(v + 1, CkInt<1>());
        }
    } else {
        for (int v = v0; v <= v1; ++v) do_row(v, (v - v0) & 1);
    }
}

// ------------------------------------------------------------------------------------------
// k_costg<NP>: k_cost for numDisparities <= 64 (NP 1; the NP 2 form works but is not faster, see run_chunk).  There a vector
// is 32 words, and k_cost's warp-per-vector passes run every packed instruction and every 32-bit shared-memory access for a
// quarter of the data a 256-disparity vector gives them.  Here a warp pass covers G = 4 / NP ADJACENT COLUMNS: a group of WD = 32 / G lanes per column, 4
// words (8 disparities) per lane, so that pixel costs, box sums, ring and volume accesses are all 128-bit per lane and
// one STG.128 of the warp writes G whole vectors (512 contiguous bytes).  Same strip walk, TMA staging, ring and
// arithmetic as k_cost (the tests compare the C volume of both with the oracle's); a strip is CK_WARPS * G columns wide.
// ------------------------------------------------------------------------------------------
__host__ __device__ constexpr int ckg_pc(int NP) { return CK_WARPS * (4 / NP); }             // pixel-cost columns per strip
__host__ __device__ constexpr int ckg_se(int NP) { return 32 * NP + ckg_pc(NP) + 8; }        // staged entries per plane (<= 256)
__host__ __device__ constexpr int ckg_blkw(int NP) { return (6 * ckg_se(NP) + 31) & ~31; }   // words per staged box, 128-byte padded
template <int NP, bool FULL, bool XTRA, int BS>
__global__ void __launch_bounds__(CK_WARPS * 32) k_costg(const __grid_constant__ CUtensorMap tmR,
                                                         const __grid_constant__ CUtensorMap tmL,
                                                         SgbmGeo g, CostArgs ca, uint32_t* __restrict__ vol)
{
    static_assert(NP == 1 || NP == 2, "k_costg: numDisparities <= 128");
    constexpr int G = 4 / NP, WD = 32 / G, VW = NP * 32, PCG = ckg_pc(NP), SE = ckg_se(NP), BLKW = ckg_blkw(NP);
    extern __shared__ __align__(128) uint32_t smem[];
    __shared__ __align__(8) uint64_t mbar[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, grp = lane / WD, l = lane % WD;
    const int r = BS ? BS / 2 : g.r, bs = BS ? BS : 2 * r + 1, OC = ca.OC, PC = OC + 2 * r;
    uint32_t* stage = smem;                                  // [2][6][SE] (each box padded to BLKW words)
    uint32_t* lstage = stage + 2 * BLKW;                     // [2][PCG][8]
    uint32_t* pixbuf = lstage + 2 * PCG * 8;                 // [2][PCG][VW]
    uint32_t* ring = pixbuf + 2 * PCG * VW;                  // [OC][bs][VW]

    const int strip = blockIdx.x % ca.nstrips, seg = blockIdx.x / ca.nstrips;
    const int job = blockIdx.y + ca.job0, b = job / g.nstripes, n = job % g.nstripes;
    int ys, ye, keep;
    job_rows(g, n, ys, ye, keep);
    const int ya = ys + seg * ca.rseg, yb = min(ya + ca.rseg, ye);   // output rows of this CTA
    if (ya >= yb) return;
    const int xs = strip * OC;                               // first output column (cost domain)
    const int w1 = g.width1;
    const int xbase0 = xs - r + g.minX1 - g.minD - (g.HALF - 1);
    const int xbase = xbase0 & ~3;                           // floor to a multiple of 4 entries (16 bytes)
    const int xl0 = xs - r + g.minX1;                        // image column of left record 0

    auto issue_row = [&](int v, int buf) {                   // one thread
        const int yy = min(max(v, ys), g.H - 1);
        mbar_expect_tx(&mbar[buf], (6 * SE + PCG * 8) * 4);
        tma_load_2d(stage + buf * BLKW, &tmR, xbase, (b * g.H + yy) * 6, &mbar[buf]);
#pragma unroll
        for (int q = 0; q < PCG / 32; ++q)                   // left records: boxes of 32 pixels (256 words)
            tma_load_2d(lstage + (buf * PCG + 32 * q) * 8, &tmL, (xl0 + 32 * q) * 8, b * g.H + yy, &mbar[buf]);
    };
    if (threadIdx.x == 0) {
        mbar_init(&mbar[0], 1);
        mbar_init(&mbar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const int v0 = ya - r, v1 = yb - 1 + r;                  // virtual input rows (clamped on use)
    if (threadIdx.x == 0) {
        issue_row(v0, 0);
        if (v0 + 1 <= v1) issue_row(v0 + 1, 1);
    }
    uint32_t padm[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int w = 4 * l + k;
        padm[k] = (w >= g.D ? 0xffffu : 0u) | (w + g.HALF >= g.D ? 0xffff0000u : 0u);
    }
    uint32_t acc[4] = {0u, 0u, 0u, 0u};
    // this group's strip column (pixel costs) / output column (box sums): the same index
    const int c = G * warp + grp;
    const int xq = min(max(xs - r + c, 0), w1 - 1);          // clamped cost column
    const int cq = xq - (xs - r);                            // ... inside the staged window
    // pixel-cost phase: register k of lane l <-> word 4l + k (NP 1: the four columns' entries are 32 consecutive words, and
    // the result leaves as one 128-bit store) or word 16k + l (NP 2: 4l + k would put lanes l and l + 8 on one bank)
    const int ebase = cq + g.HALF - 1 + (xbase0 - xbase);    // staged entry of word 0 (word w: ebase - w)
    const int loff = cq * 8;
    const bool has_pix = c < PC, has_out = c < OC;
    const int xc = xs + c;                                   // output column in the cost domain
    int slot = 0;
    // this lane's words of the volume vector being emitted (the first row is ya, at it == bs - 1)
    uint32_t* drow = vol + (((size_t)job * g.rowsMax + (ya - ys)) * w1 + xc) * VW + 4 * l;

    auto do_row = [&](const int v, auto bufc) {              // as in k_cost: row loop unrolled by two, constant buffer index
        const int buf = bufc;
        const int it = v - v0;
        const uint32_t* st = stage + buf * BLKW;
        const uint32_t* ls = lstage + buf * PCG * 8;
        uint32_t* pb = pixbuf + buf * PCG * VW;
        mbar_wait(&mbar[buf], (it >> 1) & 1);
        // ---- pixel costs of this group's strip column, 4 words per lane
        if (has_pix) {
            const uint4 l0 = *reinterpret_cast<const uint4*>(ls + loff);
            const uint4 l1 = *reinterpret_cast<const uint4*>(ls + loff + 4);
            uint32_t pix[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int e = ebase - (NP == 1 ? 4 * l + k : WD * k + l);
                const uint32_t v1a = st[0 * SE + e], v0a = st[1 * SE + e], nva = st[2 * SE + e];
                const uint32_t v1b = st[3 * SE + e], v0b = st[4 * SE + e], nvb = st[5 * SE + e];
                // as in k_cost: c0 = max(0, u - v1, v0 - u); c1 = max(0, v - u1, u0 - v); planes {-v1, v0, -v}
                uint32_t a = __vadd2(l0.x, v1a);
                uint32_t c0 = __viaddmax_s16x2_relu(v0a, l0.y, a);
                uint32_t bb = __vadd2(~nva, l0.w);
                uint32_t c1 = __viaddmax_s16x2_relu(l0.z, nva, bb);
                uint32_t px = __vminu2(c0, c1);
                a = __vadd2(l1.x, v1b);
                c0 = __viaddmax_s16x2_relu(v0b, l1.y, a);
                bb = __vadd2(~nvb, l1.w);
                c1 = __viaddmax_s16x2_relu(l1.z, nvb, bb);
                pix[k] = px + ((__vminu2(c0, c1) >> 2) & 0x3fff3fffu);
            }
            if constexpr (NP == 1) *reinterpret_cast<uint4*>(pb + c * VW + 4 * l) = make_uint4(pix[0], pix[1], pix[2], pix[3]);
            else {
#pragma unroll
                for (int k = 0; k < 4; ++k) pb[c * VW + WD * k + l] = pix[k];
            }
        }
        __syncthreads();
        if (threadIdx.x == 0 && v + 2 <= v1) issue_row(v + 2, buf);
        // ---- horizontal box over 2r+1 columns, vertical running sum, store
        if (has_out) {
            const uint32_t* pcol = pb + c * VW + 4 * l;
            uint4 hs = *reinterpret_cast<const uint4*>(pcol);
#pragma unroll
            for (int j = 1; j < (BS ? BS : bs); ++j) {
                const uint4 t = *reinterpret_cast<const uint4*>(pcol + j * VW);
                hs.x += t.x; hs.y += t.y; hs.z += t.z; hs.w += t.w;
            }
            uint32_t* rg = ring + ((size_t)c * bs + slot) * VW + 4 * l;
            if (it >= bs) {                                  // ring holds hs(v - bs) in `slot`
                const uint4 t = *reinterpret_cast<const uint4*>(rg);
                acc[0] -= t.x; acc[1] -= t.y; acc[2] -= t.z; acc[3] -= t.w;
            }
            *reinterpret_cast<uint4*>(rg) = hs;
            acc[0] += hs.x; acc[1] += hs.y; acc[2] += hs.z; acc[3] += hs.w;
            const int y = v - r;                             // output row once the window is complete
            if (it >= bs - 1 && xc < w1) {
                uint32_t* dst = drow;
                uint32_t o[4];
                if constexpr (XTRA) {
                    // MODE_HH4: cv2 leaves the rows whose window reaches below the image at cost 0 (oracle/sgbm_oracle.c)
                    const bool zrow = g.mode == SVB_MODE_HH4 && y + r > g.H - 1 && y > 0;
                    uint4 t = make_uint4(0u, 0u, 0u, 0u);
                    if (ca.accum) t = *reinterpret_cast<const uint4*>(dst);
                    const uint32_t tt[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t cv = zrow ? 0u : acc[k] + tt[k];
                        o[k] = FULL ? cv : ((cv & ~padm[k]) | (WALL2 & padm[k]));
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k) o[k] = FULL ? acc[k] : ((acc[k] & ~padm[k]) | (WALL2 & padm[k]));
                }
                *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
        if (it >= bs - 1) drow += (size_t)w1 * VW;
        slot = slot + 1 == bs ? 0 : slot + 1;
    };
    for (int v = v0; v <= v1; v += 2) {
        do_row(v, CkInt<0>());
        if (v + 1 <= v1) do_row(v + 1, CkInt<1>());
    }
}

// One step of the SGM recurrence (A.3) for a whole disparity vector held by a warp.
//   L(d) = C(d) + min(Lq(d), Lq(d-1)+P1, Lq(d+1)+P1, minLq+P2) - minLq
// Lq is updated in place; returns the warp-wide min of the new L (minLq for the next step).
// a + b on the FMA pipe (IMAD with a run-time 1) -- the integer ALU pipe is the busy one in the sweeps
__device__ __forceinline__ uint32_t add_fma(uint32_t a, uint32_t one, uint32_t b)
{
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(one), "r"(b));
    return r;
}

template <int NP, bool FULL>
__device__ __forceinline__ uint32_t sgm_step(uint32_t (&Lq)[NP], const uint32_t (&C)[NP], uint32_t mq2,
                                             uint32_t P1P1, uint32_t P2P2, int lane, const uint32_t (&padm)[NP],
                                             uint32_t one)
{
    // mq2 = minLq in both halves; returns the warp-wide min of the new L, again in both halves.
    // neighbours across lane boundaries; both halves of a word move together, except at the
    // two seams: word 0's high half continues word HALF-1's low half, and the ends see the wall
    uint32_t prev = __shfl_sync(FULLMASK, Lq[NP - 1], (lane + 31) & 31);
    uint32_t next = __shfl_sync(FULLMASK, Lq[0], (lane + 1) & 31);
#if SW_SEAM_FMA
    // the two seam fix-ups as integer multiply-adds (FMA pipe; the ALU pipe is the busy one and PRMT is half rate
    // there): x * 2^16 + wall = (lo = wall, hi = x.lo);  hi32(x * 2^16) + (wall << 16) = (lo = x.hi, hi = wall)
    {
        const uint32_t sh = one << 16;                       // 65536, opaque: the compiler would turn it back into a shift
        if (lane == 0) asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(prev) : "r"(prev), "r"(sh), "r"(WALL));
        if (lane == 31) asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(next) : "r"(next), "r"(sh), "r"(WALL << 16));
    }
#else
    if (lane == 0) prev = __byte_perm(prev, WALL2, 0x1054);     // (lo = wall, hi = lane31.lo)
    if (lane == 31) next = __byte_perm(next, WALL2, 0x5432);    // (lo = lane0.hi, hi = wall)
#endif
    const uint32_t mqP2 = mq2 + P2P2;                        // halves < 65536: no carry
    uint32_t up = prev;
    uint32_t mloc = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < NP; ++k) {
        const uint32_t dn = (k + 1 < NP) ? Lq[k + 1] : next;
#if SW_FMA_ADDS
        uint32_t t = add_fma(pk_min(up, dn), one, P1P1);     // halves stay below 65536 (wall + P1)
        t = pk_min(pk_min(t, Lq[k]), mqP2);
#else
        uint32_t t = pk_min(up, dn);
        t = pk_addmin(t, P1P1, Lq[k]);
        t = pk_min(t, mqP2);
#endif
        uint32_t L = C[k] + t - mq2;                         // halves stay in [0, 65535]: no carries
        if (!FULL) L = (L & ~padm[k]) | (WALL2 & padm[k]);   // pad disparities keep the wall
        up = Lq[k];
        Lq[k] = L;
        mloc = pk_min(mloc, L);
    }
    // min of the two halves in both halves; a packed pair of equal halves is monotone in its value
    return __reduce_min_sync(FULLMASK, pk_min(mloc, __byte_perm(mloc, 0, 0x1032)));
}

__device__ __forceinline__ int ld_acquire_gpu(const int* p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ int lds_acquire_cta(const int* p)
{
    int v;
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"((uint32_t)__cvta_generic_to_shared(p)) : "memory");
    return v;
}
__device__ __forceinline__ void sts_release_cta(int* p, int v)
{
    asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}


// one lane's NP-word slice, destination given as a shared-space address
template <int NP>
__device__ __forceinline__ void cp_async_vec_s(uint32_t sdst, const uint32_t* gsrc)
{
    if constexpr (NP % 4 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 4; ++i)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst + 16 * i), "l"(gsrc + 4 * i));
    } else if constexpr (NP % 2 == 0) {
#pragma unroll
        for (int i = 0; i < NP / 2; ++i)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sdst + 8 * i), "l"(gsrc + 2 * i));
    } else {
#pragma unroll
        for (int i = 0; i < NP; ++i)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst + 4 * i), "l"(gsrc + i));
    }
}

// Uniqueness threshold of A.4: a disparity d counts against the winner iff S[d] < th.
template <bool MAY3WAY>
__device__ __forceinline__ int wta_threshold(const SgbmGeo& g, int minS)
{
    const int f = 100 - g.uniq;                              // >= 1 (make_geo rejects uniquenessRatio >= 100)
    int th;
    if (MAY3WAY && g.mode == SVB_MODE_SGBM_3WAY) {
        // cv2's 3WAY vector loop: Sp[d] < (short)(100*minS/(100-u) + 1) (oracle/sgbm_oracle.c)
        const unsigned x = 100u * (unsigned)minS;
        th = g.uniq > 0 ? (int)(int16_t)((g.thMul ? __umulhi(x, g.thMul) : x) + 1u) : 0;
        th = max(th, 0);
    } else {
        const unsigned x = (unsigned)minS * 100u + (unsigned)f - 1u;   // S*f < minS*100  <=>  S < ceil(minS*100/f)
        th = (int)(g.thMul ? __umulhi(x, g.thMul) : x);
    }
    return min(th, 0x8000);
}

struct SweepArgs {
    const uint32_t* C;        // [job][rowsMax][w1][VW]
    uint32_t* S;              // same shape: sweep 1 writes S1, sweep 2 reads it
    uint32_t* carry;          // [job][2][w1][3][VW]   L_r of a band's last row (per path)
    uint32_t* cmin;           // [job][2][w1][4]       their minima
    int* prog;                // [job][bandsPerJob]    columns of the band's last row published
    int* ticket;              // band dispenser
    int* err;                 // set when a wait ran into the spin limit
    int16_t* raw;             // [img][H][W]   (sweep 2)
    uint32_t* key;            // [img][H][W]   right-view scatter-min keys (sweep 2)
    int njobs, bandsPerJob;
    int bandRows;             // rows per band of this launch (<= sw_rows(NP); fewer when there are not enough bands for the SMs)
    int dir;                  // +1: rows top-down, columns left-to-right; -1: the opposite
    int storeS;               // debug: sweep 2 writes the final S back
};

// VAR 0: D < Dp (pad disparities), saturating sums; 1: D == Dp, saturating sums;
// 2: D == Dp and the per-sweep sum of L_r cannot reach 32768 -> plain adds.
// MASK: paths of this sweep: 1 horizontal, 2 diagonal from behind, 4 vertical, 8 diagonal from ahead.
//
// Shared memory: every row (warp) of a band owns an INPUT ring per inter-row path: RING slots of
// ROWW = VW + 4 words: the L_r vector of one column of the previous row; the four tail words of the
// path-0 row of a slot hold that column's three path minima (both halves).  The row above writes it,
// the progress counters s_done[] order the accesses: row k may start column i once row k-1 has
// published i + 2 (column i + 1 is the diagonal-from-ahead predecessor) and row k-1 may overwrite a
// slot once row k has published the column that last read it.  Virtual columns -1 and w1
// (predecessors outside the image, A.3) are zero rows of the ring.
// All rows run the same lean loop: two helper warps play "the row above" for row 0 (the loader
// streams the carry of the previous band from L2 into row 0's ring with cp.async) and "the row
// below" for the last row (the drain copies that row's output ring into the carry for the next
// band and publishes the band's progress).  The pace of the whole wavefront is one warp's latency per column,
// so nothing but the recurrence itself may sit in the rows' loop.
template <int NP, int VAR, int MASK, bool LAST>
__global__ void __launch_bounds__((sw_rows(NP) + 2) * 32, 1) k_sweep(const SweepArgs a, const SgbmGeo g)
{
    constexpr bool FULL = VAR != 0, FAST = VAR == 2;
    constexpr bool VERT = (MASK & 14) != 0;                  // any path that crosses rows
    constexpr int VW = NP * 32, ROWW = sw_roww(NP);
    constexpr int SW_R = sw_rows(NP), RING = sw_ring(NP, LAST), NS = sw_stage(NP, LAST), PFD = NS - 1;
    constexpr int PATHW = RING * ROWW, RINGW = 3 * PATHW;
    extern __shared__ __align__(128) uint32_t smem[];
    uint32_t* ring = smem;                                   // [SW_R + 1][3][RING][ROWW] (+1: what the last row writes)
    uint32_t* cstage = ring + (SW_R + 1) * RINGW;            // [SW_R][NS][VW] cp.async ring of C columns
    uint32_t* sstage = cstage + SW_R * NS * VW;              // [SW_R][NS][VW] (last sweep) ... and of S1 columns
    uint32_t* wsc = sstage + (LAST ? SW_R * NS * VW : 0);    // [SW_R][VW] (last sweep) S of the current column for the subpixel taps
    __shared__ int s_ticket;
    __shared__ int s_done[SW_R + 2];                         // [k]: columns published by row k (or by the drain for k = nact); [SW_R + 1]: by the loader
    int* const s_load = &s_done[SW_R + 1];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool helper = warp >= SW_R;                        // SW_R: carry loader, SW_R + 1: carry drain
    const int w1 = g.width1;
    const uint32_t P1P1 = pk_dup(g.P1), P2P2 = pk_dup(g.P2);
    const int dir = a.dir;
    const uint32_t one = a.njobs > 0;                        // 1, opaque to the compiler (add_fma)
    const uint32_t himask = 0u - (one << 16);                // 0xffff0000 in a register (keeps the WTA keys' LOP3 a single op)
    uint32_t padm[NP];
#pragma unroll
    for (int k = 0; k < NP; ++k) {
        const int w = NP * lane + k;
        padm[k] = (w >= g.D ? 0xffffu : 0u) | (w + g.HALF >= g.D ? 0xffff0000u : 0u);
    }
    const int rw = helper ? 0 : warp;
    uint32_t* ring_in = ring + (size_t)rw * RINGW;
    uint32_t* ring_out = ring_in + RINGW;                    // input ring of warp + 1
    const uint32_t mysc_s = smem_u32(wsc + rw * VW);         // (last sweep) this row's S vector of the current column
    const uint32_t cst_s = smem_u32(cstage + (size_t)rw * NS * VW + NP * lane);   // this lane's slice, slot 0
    const uint32_t sst_s = smem_u32(sstage + (size_t)rw * NS * VW + NP * lane);

    auto zero_col = [&](uint32_t* rbase, int p, uint32_t slotoff) {   // one column of one path: vector + its minimum
        const uint32_t z[NP] = {};
        vec_st<NP>(rbase + p * PATHW + slotoff, lane, z);
        if (lane == 0) rbase[slotoff + VW + p] = 0;
    };

    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_ticket = atomicAdd(a.ticket, 1);
        __syncthreads();
        const int t = s_ticket;
        if (t >= a.njobs * a.bandsPerJob) break;
        const int job = t % a.njobs, band = t / a.njobs;
        int ys, ye, keep;
        job_rows(g, job % g.nstripes, ys, ye, keep);
        const int nrows = ye - ys, r0 = band * a.bandRows;
        const int nact = min(a.bandRows, nrows - r0);
        if (nact <= 0) continue;

        const bool active = !helper && warp < nact;
        const int sr = r0 + rw;                              // row index in sweep order
        const int row = dir > 0 ? sr : nrows - 1 - sr;       // storage row within the job
        const int y = ys + row;
        const bool rowSkip = LAST && !VERT && y < keep;      // 3WAY overlap rows: nothing depends on them
        const size_t rowbase = (((size_t)job * g.rowsMax + row) * w1) * VW;
        const int b = job / g.nstripes;
        const int nL = VERT && band > 0 ? w1 : 0;            // carry columns the loader brings in
        const int nD = VERT && (r0 + nact < nrows) ? w1 : 0; // carry columns the drain takes out

        if (active) {
            if (lane == 0) s_done[warp] = 0;
            if (VERT) {
                if (warp == 0 && band == 0) {                // first row of the job: no previous row at all
                    for (int q = lane; q < RINGW; q += 32) ring_in[q] = 0;
                } else {
                    zero_col(ring_in, 0, (RING - 1) * ROWW); // virtual column -1 of the diagonal-from-behind path
                }
            }
        } else if (warp == SW_R && lane == 0) {
            *s_load = nL ? 0 : w1 + 8;                       // no carry to wait for: row 0 runs free
            s_done[nact] = nD ? 0 : w1 + 64;                 // nobody below the last row: it never waits
        }
        __syncthreads();

        if (helper) {
            const int* prog_pred = a.prog + (size_t)job * a.bandsPerJob + band - 1;
            int* prog_mine = a.prog + (size_t)job * a.bandsPerJob + band;
            if (warp == SW_R && nL) {
                // ---- loader: the previous band's carry -> row 0's ring, DEPTH columns between request and publication
                const uint32_t* carry_in = a.carry + (((size_t)job * 2 + ((band + 1) & 1)) * w1) * 3 * VW;
                const uint32_t* cmin_in = a.cmin + (((size_t)job * 2 + ((band + 1) & 1)) * w1) * 4;
                uint32_t* ring0 = ring;
                const uint32_t ring0_s = smem_u32(ring0);
                constexpr int DEPTH = RING - 2;
                int seen0 = 0, seenG = 0, cl = 0, cp = 0;    // next column to request / to publish
                bool dead = false;
                long long spins = 0;
                while (cp < nL) {
                    // request as far ahead as row 0's ring allows ...
                    bool can = cl < nL && cl - cp < DEPTH;
                    if (can && seen0 < cl - RING + 2) { seen0 = lds_acquire_cta(&s_done[0]); can = seen0 >= cl - RING + 2; }
                    if (can && seenG < cl + 1) { seenG = ld_acquire_gpu(prog_pred); can = seenG >= cl + 1; }
                    if (can) {
                        const uint32_t so = (uint32_t)(cl % RING) * ROWW;
#pragma unroll
                        for (int p = 0; p < 3; ++p) {
                            if (!(MASK & (2 << p))) continue;
                            const uint32_t* src = carry_in + ((size_t)cl * 3 + p) * VW;
                            // 16-byte cp.async.cg whatever the vector width (L2 only: the carry was written by another
                            // CTA of this launch; the 4- and 8-byte forms would cache it in L1).  A vector is VW / 4
                            // 16-byte pieces: fewer than 32 for numDisparities <= 192, where a synchronous copy through
                            // registers (one dependent L2 round trip per column) used to pace the whole wavefront.
#pragma unroll
                            for (int q = 0; q < (VW / 4 + 31) / 32; ++q) {
                                const int piece = lane + 32 * q;
                                if (VW / 4 % 32 == 0 || piece < VW / 4)
                                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring0_s + (p * PATHW + so + 4 * piece) * 4),
                                                 "l"(src + 4 * piece));
                            }
                        }
                        if (lane == 0)
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring0_s + (so + VW) * 4),
                                         "l"(cmin_in + (size_t)cl * 4));
                        asm volatile("cp.async.commit_group;");
                        ++cl;
                        spins = 0;
                    } else if (cp < cl) {
                        // ... and whenever that is not possible, hand over the oldest requested column once it has landed
                        switch (cl - cp - 1) {
                            case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
                            case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
                            case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
                            default: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
                        }
                        __syncwarp();
                        ++cp;
                        if (lane == 0) sts_release_cta(s_load, cp);
                        spins = 0;
                    } else {
                        if (++spins > SW_SPIN_LIMIT) { dead = true; break; }
                        __nanosleep(SW_SLEEP_NS);
                    }
                }
                if (dead) {
                    if (lane == 0) { atomicExch(a.err, 1); sts_release_cta(s_load, w1 + 8); }   // let the rows run out
                } else {
                    // virtual column w1 for row 0, then the final count
                    while (seen0 < w1 - RING + 2) {
                        seen0 = lds_acquire_cta(&s_done[0]);
                        if (seen0 < w1 - RING + 2) __nanosleep(SW_SLEEP_NS);
                    }
                    zero_col(ring0, 2, (uint32_t)(w1 % RING) * ROWW);
                    __syncwarp();
                    if (lane == 0) sts_release_cta(s_load, w1 + 8);
                }
                asm volatile("cp.async.wait_all;" ::: "memory");
            }
            if (warp == SW_R + 1 && nD) {
                // ---- drain: the last row's output ring -> the carry for the next band, progress every SW_FLAG_EVERY columns
                uint32_t* carry_out = a.carry + (((size_t)job * 2 + (band & 1)) * w1) * 3 * VW;
                uint32_t* cmin_out = a.cmin + (((size_t)job * 2 + (band & 1)) * w1) * 4;
                const uint32_t* ringD = ring + (size_t)nact * RINGW;
                int seenL = 0;
                for (int cd = 0; cd < nD; ++cd) {
                    while (seenL < cd + 1) {
                        seenL = lds_acquire_cta(&s_done[nact - 1]);
                        if (seenL < cd + 1) __nanosleep(SW_SLEEP_NS);
                    }
                    const uint32_t so = (uint32_t)(cd % RING) * ROWW;
#pragma unroll
                    for (int p = 0; p < 3; ++p) {
                        if (!(MASK & (2 << p))) continue;
                        uint32_t v[NP];
                        vec_ld<NP>(ringD + p * PATHW + so, lane, v, LdPlain());
                        vec_st<NP>(carry_out + ((size_t)cd * 3 + p) * VW, lane, v);
                    }
                    if (lane < 4) cmin_out[(size_t)cd * 4 + lane] = ringD[so + VW + lane];
                    __syncwarp();
                    if (lane == 0) {
                        sts_release_cta(&s_done[nact], cd + 2);          // the slot of column cd is free again
                        if ((cd & (SW_FLAG_EVERY - 1)) == SW_FLAG_EVERY - 1 || cd == w1 - 1) st_release_gpu(prog_mine, cd + 1);
                    }
                }
            }
            continue;
        }

        uint32_t Lh[NP];
#pragma unroll
        for (int k = 0; k < NP; ++k) Lh[k] = 0;
        uint32_t minh2 = 0;                                  // min of the horizontal L_r, in both halves
        const int x0 = dir > 0 ? 0 : w1 - 1;
        const ptrdiff_t cstep = (ptrdiff_t)dir * VW;
        const uint32_t* pC0 = a.C + rowbase + (size_t)x0 * VW + NP * lane;
        const uint32_t* pS0 = a.S + rowbase + (size_t)x0 * VW + NP * lane;
        const int nloop = active && !rowSkip ? w1 : 0;

        // C (and S1) columns are staged PFD columns ahead through a per-row cp.async ring: every lane copies
        // and later reads back only its own slice, so a per-thread wait_group is all it takes.  The last PFD
        // requests of a row run past its end into the neighbouring row (valid memory, never read back).
        if (nloop) {
#pragma unroll
            for (int c = 0; c < PFD; ++c) {
                cp_async_vec_s<NP>(cst_s + c * VW * 4, pC0 + (ptrdiff_t)c * cstep);
                if (LAST) cp_async_vec_s<NP>(sst_s + c * VW * 4, pS0 + (ptrdiff_t)c * cstep);
                asm volatile("cp.async.commit_group;");
            }
        }
        int seenP = 0, seenS = 0;
        // Shared-space byte addresses of this lane's slice in the input ring slots of columns i-1, i, i+1 (path 0;
        // path p is PATHW words further, the output ring RINGW words further) -- rotated, never recomputed
        const uint32_t rin_s = smem_u32(ring_in + NP * lane), rin_end = rin_s + PATHW * 4;
        uint32_t a_m1 = rin_s + (RING - 1) * ROWW * 4, a_0 = rin_s, a_p1 = rin_s + ROWW * 4;
        const uint32_t tail = (uint32_t)(VW - NP * lane) * 4;      // from a lane's slice to the slot's minima
        uint32_t cs_cur = 0, cs_prev = (NS - 1) * VW * 4;          // stage offsets (bytes) of columns i, i-1
        // columns <= lim may run without looking at the neighbours' counters again
        int lim = VERT && SW_EXPERIMENT != 6 ? -1 : 0x7fffffff;
        const uint32_t* gC = pC0 + (ptrdiff_t)PFD * cstep;         // this lane's slice of column i + PFD
        const uint32_t* gS = pS0 + (ptrdiff_t)PFD * cstep;
        uint32_t* pS = a.S + rowbase + (size_t)x0 * VW + NP * lane;
        uint32_t rkey = 0, rcnt = 0, rtap = 0;                     // WTA record of column (i & ~31) + lane
        const int* done_pred = warp > 0 ? &s_done[warp - 1] : s_load;
        const int* done_succ = &s_done[warp + 1];

#if SW_UNROLL2
#pragma unroll 2
#endif
        for (int i = 0; i < nloop; ++i) {
            // column i + 1 of the row above is only needed by the diagonal-from-ahead path: without it (3WAY, HH4) a row
            // follows the row above one column behind instead of two (a single frame's dependent chain gets shorter)
            constexpr int NEED = (MASK & 8) ? 2 : 1;
            if (i > lim) {                                   // one compare on the common path
                while ((seenP = lds_acquire_cta(done_pred)) < i + NEED) __nanosleep(SW_SLEEP_NS);
                while ((seenS = lds_acquire_cta(done_succ)) < i - RING + 2) __nanosleep(SW_SLEEP_NS);
                lim = min(seenP - NEED, seenS + RING - 2);
            }
            // column i + PFD -> the stage slot column i - 1 has just left
            if (SW_EXPERIMENT != 3 && SW_EXPERIMENT != 5) {
                cp_async_vec_s<NP>(cst_s + cs_prev, gC);
                if (LAST) cp_async_vec_s<NP>(sst_s + cs_prev, gS);
            }
            asm volatile("cp.async.commit_group;");
            asm volatile("cp.async.wait_group %0;" ::"n"(PFD) : "memory");
            uint32_t C[NP], S[NP];
            lds_vec<NP, 0>(cst_s + cs_cur, C);
            if (LAST) lds_vec<NP, 0>(sst_s + cs_cur, S);
            else {
#pragma unroll
                for (int k = 0; k < NP; ++k) S[k] = 0u;
            }
            // predecessors on the previous row of the sweep
            uint32_t Lt[3][NP], mq2[3];
            if (VERT) {
#pragma unroll
                if (MASK & 2) { lds_vec<NP, 0>(a_m1, Lt[0]); mq2[0] = lds_u32<0>(a_m1 + tail); }
                if (MASK & 4) { lds_vec<NP, PATHW * 4>(a_0, Lt[1]); mq2[1] = lds_u32<4>(a_0 + tail); }
                if (MASK & 8) { lds_vec<NP, 2 * PATHW * 4>(a_p1, Lt[2]); mq2[2] = lds_u32<8>(a_p1 + tail); }
            }
            uint32_t sum[NP];                                // FAST: plain sum of this sweep's L_r
#pragma unroll
            for (int k = 0; k < NP; ++k) sum[k] = 0;
            if (MASK & 1) {                                  // horizontal: predecessor is the previous column
                minh2 = sgm_step<NP, FULL>(Lh, C, minh2, P1P1, P2P2, lane, padm, one);
#pragma unroll
                for (int k = 0; k < NP; ++k) {
                    if (FAST) sum[k] = Lh[k];                 // the first term of the sum
                    else S[k] = pk_addmin(S[k], Lh[k], WALL2);
                }
            }
            if (VERT) {
#pragma unroll
                for (int p = 0; p < 3; ++p) {
                    if (!(MASK & (2 << p))) continue;
                    const uint32_t m2 = sgm_step<NP, FULL>(Lt[p], C, mq2[p], P1P1, P2P2, lane, padm, one);
#pragma unroll
                    for (int k = 0; k < NP; ++k) {
                        if (FAST) sum[k] = SW_FMA_ADDS ? add_fma(Lt[p][k], one, sum[k]) : sum[k] + Lt[p][k];
                        else S[k] = pk_addmin(S[k], Lt[p][k], WALL2);
                    }
                    // the output ring is the next row's input ring: RINGW words further; lane 0's slice starts the slot
                    if (p == 0) { sts_vec<NP, RINGW * 4>(a_0, Lt[0]); if (lane == 0) sts_u32<(RINGW + VW) * 4>(a_0, m2); }
                    if (p == 1) { sts_vec<NP, (RINGW + PATHW) * 4>(a_0, Lt[1]); if (lane == 0) sts_u32<(RINGW + VW + 1) * 4>(a_0, m2); }
                    if (p == 2) { sts_vec<NP, (RINGW + 2 * PATHW) * 4>(a_0, Lt[2]); if (lane == 0) sts_u32<(RINGW + VW + 2) * 4>(a_0, m2); }
                }
            }
            if (FAST) {
#pragma unroll
                for (int k = 0; k < NP; ++k) S[k] = LAST ? pk_addmin(S[k], sum[k], WALL2) : sum[k];
            }
            if (LAST) {
                if (a.storeS) vec_st<NP>(pS - NP * lane, lane, S);
                if (y >= keep) {
                    // ---- WTA (A.4): argmin with lowest-d tie break via keys (S << 16 | d)
                    // (the lane's own 2 NP keys carry compile-time local indices k / HALF + k; the lane's first
                    //  disparity NP * lane is added once, before the warp-wide minimum)
                    uint32_t kmin = 0xffffffffu;
#pragma unroll
                    for (int k = 0; k < NP; ++k) {
                        const uint32_t klo = S[k] * 65536u + (uint32_t)k;                  // one IMAD (FMA pipe)
                        const uint32_t khi = (S[k] & himask) | (uint32_t)(32 * NP + k);     // one LOP3
                        kmin = min(kmin, min(klo, khi));
                    }
                    kmin = __reduce_min_sync(FULLMASK, kmin + (uint32_t)(NP * lane));
                    if constexpr (MASK == 1) {
                        // cv2's 3WAY argmin runs 8 int16 lanes over d (lane = d mod 8): within a lane a later d with an
                        // EQUAL cost replaces the earlier one, across lanes the lowest recorded d wins (oracle/sgbm_oracle.c).
                        // Only visible without a uniqueness test (any tie further than 1 apart is rejected otherwise).
                        if (g.mode == SVB_MODE_SGBM_3WAY && g.uniq == 0) {
                            const uint32_t m2 = pk_dup(kmin >> 16);
                            uint32_t ties = 0;
#pragma unroll
                            for (int k = 0; k < NP; ++k) {
                                const uint32_t x = S[k] ^ m2;
                                ties += ((x & 0xffffu) == 0u) + ((x >> 16) == 0u);
                            }
                            if (__reduce_add_sync(FULLMASK, ties) > 1u) {
                                int bestd = 0x7fff;
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    int v = -1;
#pragma unroll
                                    for (int k = 0; k < NP; ++k) {
                                        const int dlo = NP * lane + k;
                                        if ((dlo & 7) != j) continue;
                                        const uint32_t x = S[k] ^ m2;
                                        if ((x >> 16) == 0u) v = max(v, dlo + g.HALF);
                                        else if ((x & 0xffffu) == 0u) v = max(v, dlo);
                                    }
                                    v = __reduce_max_sync(FULLMASK, v);
                                    if (v >= 0) bestd = min(bestd, v);
                                }
                                kmin = (kmin & 0xffff0000u) | (uint32_t)bestd;
                            }
                        }
                    }
                    const int th = wta_threshold<MASK == 1>(g, (int)(kmin >> 16));
                    // uniqueness: count disparities with S < th; the ones at best-1..best+1 are taken out later
                    const uint32_t th2 = pk_dup((uint32_t)th);
                    uint32_t lt[NP];
#pragma unroll
                    for (int k = 0; k < NP; ++k) {
                        lt[k] = __vsub2(S[k], th2);                        // S, th in [0, 0x8000]: bit 15 = (S < th)
                        if (!FULL) lt[k] &= ~padm[k];
                    }
                    uint32_t bits = 0;
                    if constexpr (NP == 4) {                               // the sign bytes of four words in two permutes
                        bits = (__byte_perm(lt[0], lt[1], 0x7351) & 0x80808080u) |
                               ((__byte_perm(lt[2], lt[3], 0x7351) & 0x80808080u) >> 1);
                    } else if constexpr (NP == 2) {
                        bits = __byte_perm(lt[0], lt[1], 0x7351) & 0x80808080u;
                    } else {
#pragma unroll
                        for (int k = 0; k < NP; ++k) bits |= (lt[k] >> k) & (0x80008000u >> k);
                    }
                    const uint32_t cnt = __reduce_add_sync(FULLMASK, (uint32_t)__popc(bits));
                    // subpixel taps S[best - 1], S[best + 1] through the row's scratch vector (shared-space addresses;
                    // disparity d sits in half d / HALF of word d % HALF)
                    sts_vec<NP, 0>(mysc_s + NP * 4 * lane, S);
                    __syncwarp();
                    const int best = (int)(kmin & 0xffffu);
                    auto tap = [&](int d) {
                        uint32_t v;
                        const uint32_t ad = mysc_s + 4u * (uint32_t)d - (d >= 32 * NP ? (uint32_t)(4 * 32 * NP - 2) : 0u);
                        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(ad));
                        return v;
                    };
                    const uint32_t tm1 = tap(max(best - 1, 0)), tp1 = tap(min(best + 1, g.D - 1));
                    const uint32_t sm1v = best > 0 ? tm1 : 0u;
                    const uint32_t sp1v = best < g.D - 1 ? tp1 : 0u;
                    if (lane == (i & 31)) { rkey = kmin; rcnt = cnt; rtap = sm1v | (sp1v << 16); }
                    if ((i & 31) == 31 || i == w1 - 1) {
                        // 32 columns decided at once: lane j finishes column (i & ~31) + j
                        const int ci = (i & ~31) + lane;
                        if (ci <= i) {
                            const int xc = dir > 0 ? ci : w1 - 1 - ci;
                            const int x = xc + g.minX1;
                            const size_t rowpix = ((size_t)b * g.H + y) * g.W;
                            const int minS = (int)(rkey >> 16), bst = (int)(rkey & 0xffffu);
                            const int thr = wta_threshold<MASK == 1>(g, minS);
                            const int s1 = (int)(rtap & 0xffffu), s2 = (int)(rtap >> 16);
                            int out = g.INV;
                            if (minS < (int)WALL) {
                                const int excl = (minS < thr) + (bst > 0 && s1 < thr) + (bst < g.D - 1 && s2 < thr);
                                if ((int)rcnt <= excl) {
                                    int dsub = bst * 16;
                                    if (bst > 0 && bst < g.D - 1) {
                                        const int den = max(s1 + s2 - 2 * minS, 1);
                                        dsub += ((s1 - s2) * 16 + den) / (2 * den);   // C division: truncates toward zero
                                    }
                                    out = dsub + g.minD * 16;
                                    const int x2 = x - bst - g.minD;
                                    // smallest cost wins, then the largest x: the pixel cv2's right-to-left loop keeps
                                    atomicMin(a.key + rowpix + x2, ((uint32_t)minS << 16) | (uint32_t)(0xffff - x));
                                }
                            }
                            a.raw[rowpix + x] = (int16_t)out;
                        }
                    }
                }
            }
            if (VERT || LAST) __syncwarp();
#if SW_PLAIN_FLAG
            if (VERT && lane == 0) asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(&s_done[warp])), "r"(i + 1) : "memory");
#else
            if (VERT && lane == 0) sts_release_cta(&s_done[warp], i + 1);
#endif
            // (after the publication: its fence then has no global store of this column to wait for)
            if (!LAST && SW_EXPERIMENT != 4 && SW_EXPERIMENT != 5) vec_st<NP>(pS - NP * lane, lane, S);
            gC += cstep; gS += cstep; pS += cstep;
            a_m1 = a_0; a_0 = a_p1; a_p1 = a_p1 + ROWW * 4 == rin_end ? rin_s : a_p1 + ROWW * 4;
            cs_prev = cs_cur; cs_cur = cs_cur + VW * 4 == NS * VW * 4 ? 0u : cs_cur + VW * 4;
        }
        if (VERT && nloop) {
            // virtual column w1 for the row below (its diagonal-from-ahead predecessor at the last column),
            // then a final count no column index reaches
            while (seenS < w1 - RING + 2) {
                seenS = lds_acquire_cta(done_succ);
                if (seenS < w1 - RING + 2) __nanosleep(SW_SLEEP_NS);
            }
            zero_col(ring_out, 2, (uint32_t)(w1 % RING) * ROWW);
            __syncwarp();
            if (lane == 0) sts_release_cta(&s_done[warp], w1 + 8);
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
    }
}

// ------------------------------------------------------------------------------------------
// k_wave<NP, MASK>: the FIRST sweep (paths -> and the ones from the previous row, S1 out) for numDisparities <= 128,
// where k_sweep's warp-per-pixel loop spends ~100 of its ~140 instructions per column on per-vector overhead (shuffles,
// seams, reductions, ring and flag bookkeeping, polling) whatever the vector holds.  Here a warp carries G = 4 / NP
// CONSECUTIVE ROWS in lockstep, each in a group of WD = 32 / G lanes with 4 words (8 disparities) per lane -- the
// instruction stream of the 256-disparity case serving G pixels.  Group g runs SKEW columns behind group g - 1 (the
// predecessors (y-1, x-1..x+1) are then complete), so inside a warp the rows hand their L_r over through the same
// shared-memory rings as in k_sweep but need no flags: program order and __syncwarp do it.  Between warps (last group
// of warp w -> first group of warp w + 1), towards the previous / next band (loader / drain helper warps, L2 carry) and
// for the band dispenser everything is k_sweep's protocol.  Predecessors outside the image (column -1, column w1) are
// zeroed by the READER, so the rings hold real columns only.  Plain sums (VAR 0 / 2 with sums that cannot reach 32768) only.
// ------------------------------------------------------------------------------------------
#ifndef WV_WARPS1
#define WV_WARPS1 15           // row warps per CTA at numDisparities <= 64 (4 rows each)
#endif
#ifndef WV_WARPS2
#define WV_WARPS2 18           // ... at numDisparities <= 128 with 2 rows each (15 / 18 warps: first sweep 720p / D128 9.41 / 9.13 ms)
#endif
#ifndef WV_WARPS2X
#define WV_WARPS2X 12          // ... at numDisparities <= 128 with 4 rows each (8 words per lane)
#endif
#ifndef WV_G2
#define WV_G2 2                // rows per warp at numDisparities <= 128: 2, or 4 (8 words per lane, 4-slot rings, 12 warps: 720p / D128 9.1 -> 11.1 ms)
#endif
__host__ __device__ constexpr int wv_g(int NP) { return NP == 1 ? 4 : WV_G2; }
__host__ __device__ constexpr int wv_warps(int NP) { return NP == 1 ? WV_WARPS1 : (WV_G2 == 4 ? WV_WARPS2X : WV_WARPS2); }
__host__ __device__ constexpr int wv_rows(int NP) { return wv_warps(NP) * wv_g(NP); }
__host__ __device__ constexpr int wv_ring(int NP) { return NP * wv_g(NP) == 8 ? 4 : 6; }   // 8 words per lane: shallower rings so that enough rows fit
__host__ __device__ constexpr size_t wv_smem_words(int NP)
{
    return (size_t)(wv_rows(NP) + 1) * 3 * wv_ring(NP) * (NP * 32 + 4) + (size_t)wv_rows(NP) * 4 * NP * 32 + (NP * 32 + 4);
}
template <int NP, int MASK, bool FULL>
__global__ void __launch_bounds__((wv_warps(NP) + 2) * 32, 1) k_wave(const SweepArgs a, const SgbmGeo g)
{
    static_assert(NP == 1 || NP == 2, "k_wave: numDisparities <= 128");
    constexpr int G = wv_g(NP), WD = 32 / G, WPL = NP * G;   // rows per warp, lanes per row, words per lane (4 or 8)
    constexpr int WV_WARPS = wv_warps(NP);
    constexpr int VW = NP * 32, ROWW = VW + 4, ROWS = WV_WARPS * G;
    constexpr int RING = wv_ring(NP), NS = 4, PFD = NS - 1;
    constexpr int PATHW = RING * ROWW, RINGW = 3 * PATHW;
    constexpr int SKEW = (MASK & 8) ? 2 : 1;                 // columns a row runs behind the row above
    extern __shared__ __align__(128) uint32_t smem[];
    uint32_t* ring = smem;                                   // [ROWS + 1][3][RING][ROWW] (+1: what the last row writes)
    uint32_t* cstage = ring + (ROWS + 1) * RINGW;            // [ROWS][NS][VW] cp.async ring of C columns
    uint32_t* zrow = cstage + ROWS * NS * VW;                // [ROWW] zeros: the predecessor vector (and minimum) outside the image
    __shared__ int s_ticket;
    __shared__ int s_done[ROWS + 2];                         // [k]: columns published by row k (or the drain for k = nact); [ROWS + 1]: by the loader
    int* const s_load = &s_done[ROWS + 1];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool helper = warp >= WV_WARPS;                    // WV_WARPS: carry loader, WV_WARPS + 1: carry drain
    const int grp = lane / WD, l = lane % WD;
    if (threadIdx.x < ROWW) zrow[threadIdx.x] = 0;           // (the band loop starts with a block barrier)
    const int w1 = g.width1, dir = a.dir;
    const uint32_t P1P1 = pk_dup(g.P1), P2P2 = pk_dup(g.P2);
    const uint32_t one = a.njobs > 0;                        // 1, opaque to the compiler (add_fma)
    const int rw = helper ? 0 : warp * G + grp;              // this lane's row within the band
    uint32_t* ring_in = ring + (size_t)rw * RINGW;
    const int srcPrev = (lane & ~(WD - 1)) | ((l + WD - 1) & (WD - 1)), srcNext = (lane & ~(WD - 1)) | ((l + 1) & (WD - 1));
    const uint32_t selPrev = l == 0 ? 0x1054u : 0x3210u, selNext = l == WD - 1 ? 0x5432u : 0x3210u;
    const uint32_t cst_s = smem_u32(cstage + (size_t)rw * NS * VW + WPL * l);    // this lane's slice, stage slot 0
    uint32_t padm[WPL];                                      // !FULL (D < Dp): pad disparities keep the wall
#pragma unroll
    for (int k = 0; k < WPL; ++k) {
        const int w = WPL * l + k;
        padm[k] = (w >= g.D ? 0xffffu : 0u) | (w + g.HALF >= g.D ? 0xffff0000u : 0u);
    }
    // minimum over a row's group of lanes for TWO paths at once: (a, a) and (b, b) travel as (a, b)
    auto gmin2 = [&](uint32_t ma, uint32_t mb, uint32_t& ra, uint32_t& rb) {
        uint32_t v = __byte_perm(ma, mb, 0x7610);
#pragma unroll
        for (int o = WD / 2; o; o >>= 1) v = pk_min(v, __shfl_xor_sync(FULLMASK, v, o));
        ra = __byte_perm(v, v, 0x1010);
        rb = __byte_perm(v, v, 0x3232);
    };
    // one path's recurrence step for this lane's 4 words; returns the lane's minimum of the new L in both halves
    auto step = [&](uint32_t (&Lq)[WPL], const uint32_t (&C)[WPL], uint32_t mq2) {
        uint32_t prev = __shfl_sync(FULLMASK, Lq[WPL - 1], srcPrev);
        uint32_t next = __shfl_sync(FULLMASK, Lq[0], srcNext);
        prev = __byte_perm(prev, WALL2, selPrev);            // group lane 0: (lo = wall, hi = last lane's lo)
        next = __byte_perm(next, WALL2, selNext);            // last group lane: (lo = lane 0's hi, hi = wall)
        const uint32_t mqP2 = mq2 + P2P2;
        uint32_t up = prev, mloc = 0xffffffffu;
#pragma unroll
        for (int k = 0; k < WPL; ++k) {
            const uint32_t dn = k < WPL - 1 ? Lq[k + 1] : next;
            uint32_t t = add_fma(pk_min(up, dn), one, P1P1);
            t = pk_min(pk_min(t, Lq[k]), mqP2);
            uint32_t L = C[k] + t - mq2;
            if (!FULL) L = (L & ~padm[k]) | (WALL2 & padm[k]);
            up = Lq[k];
            Lq[k] = L;
            mloc = pk_min(mloc, L);
        }
        return pk_min(mloc, __byte_perm(mloc, 0, 0x1032));
    };

    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_ticket = atomicAdd(a.ticket, 1);
        __syncthreads();
        const int t = s_ticket;
        if (t >= a.njobs * a.bandsPerJob) break;
        const int job = t % a.njobs, band = t / a.njobs;
        int ys, ye, keep;
        job_rows(g, job % g.nstripes, ys, ye, keep);
        const int nrows = ye - ys, r0 = band * ROWS;
        const int nact = min(ROWS, nrows - r0);
        if (nact <= 0) continue;
        const bool rowAct = !helper && rw < nact;
        const int gLast = min(G - 1, nact - 1 - warp * G);   // last active group of this warp (< 0: none)
        const int nL = band > 0 ? w1 : 0;                    // carry columns the loader brings in
        const int nD = (r0 + nact < nrows) ? w1 : 0;         // carry columns the drain takes out

        if (rowAct) {
            if (l == 0) s_done[rw] = 0;
            if (rw == 0 && band == 0)                        // first row of the job: no previous row at all
                for (int q = l; q < RINGW; q += WD) ring_in[q] = 0;
        } else if (warp == WV_WARPS && lane == 0) {
            *s_load = nL ? 0 : w1 + 64;                      // no carry to wait for: row 0 runs free
            s_done[nact] = nD ? 0 : w1 + 64;                 // nobody below the last row: it never waits
        }
        __syncthreads();

        if (helper) {
            const int* prog_pred = a.prog + (size_t)job * a.bandsPerJob + band - 1;
            int* prog_mine = a.prog + (size_t)job * a.bandsPerJob + band;
            if (warp == WV_WARPS && nL) {
                // ---- loader: the previous band's carry -> row 0's ring (as in k_sweep)
                const uint32_t* carry_in = a.carry + (((size_t)job * 2 + ((band + 1) & 1)) * w1) * 3 * VW;
                const uint32_t* cmin_in = a.cmin + (((size_t)job * 2 + ((band + 1) & 1)) * w1) * 4;
                const uint32_t ring0_s = smem_u32(ring);
                constexpr int DEPTH = RING - 2;
                int seen0 = 0, seenG = 0, cl = 0, cp = 0;    // next column to request / to publish
                bool dead = false;
                long long spins = 0;
                while (cp < nL) {
                    bool can = cl < nL && cl - cp < DEPTH;
                    if (can && seen0 < cl - RING + 2) { seen0 = lds_acquire_cta(&s_done[0]); can = seen0 >= cl - RING + 2; }
                    if (can && seenG < cl + 1) { seenG = ld_acquire_gpu(prog_pred); can = seenG >= cl + 1; }
                    if (can) {
                        const uint32_t so = (uint32_t)(cl % RING) * ROWW;
#pragma unroll
                        for (int p = 0; p < 3; ++p) {
                            if (!(MASK & (2 << p))) continue;
                            const uint32_t* src = carry_in + ((size_t)cl * 3 + p) * VW;
                            if (lane < VW / 4)
                                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring0_s + (p * PATHW + so + 4 * lane) * 4),
                                             "l"(src + 4 * lane));
                        }
                        if (lane == 0)
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring0_s + (so + VW) * 4),
                                         "l"(cmin_in + (size_t)cl * 4));
                        asm volatile("cp.async.commit_group;");
                        ++cl;
                        spins = 0;
                    } else if (cp < cl) {
                        switch (cl - cp - 1) {
                            case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
                            case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
                            case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
                            default: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
                        }
                        __syncwarp();
                        ++cp;
                        if (lane == 0) sts_release_cta(s_load, cp);
                        spins = 0;
                    } else {
                        if (++spins > SW_SPIN_LIMIT) { dead = true; break; }
                        __nanosleep(SW_SLEEP_NS);
                    }
                }
                if (lane == 0) {
                    if (dead) atomicExch(a.err, 1);
                    sts_release_cta(s_load, w1 + 64);        // final count (or: let the rows run out)
                }
                asm volatile("cp.async.wait_all;" ::: "memory");
            }
            if (warp == WV_WARPS + 1 && nD) {
                // ---- drain: the last row's output ring -> the carry for the next band
                uint32_t* carry_out = a.carry + (((size_t)job * 2 + (band & 1)) * w1) * 3 * VW;
                uint32_t* cmin_out = a.cmin + (((size_t)job * 2 + (band & 1)) * w1) * 4;
                const uint32_t* ringD = ring + (size_t)nact * RINGW;
                int seenL = 0;
                long long spins = 0;
                for (int cd = 0; cd < nD; ++cd) {
                    while (seenL < cd + 1) {
                        seenL = lds_acquire_cta(&s_done[nact - 1]);
                        if (seenL < cd + 1) {
                            if (++spins > SW_SPIN_LIMIT) { if (lane == 0) atomicExch(a.err, 1); seenL = w1 + 64; }
                            __nanosleep(SW_SLEEP_NS);
                        }
                    }
                    const uint32_t so = (uint32_t)(cd % RING) * ROWW;
#pragma unroll
                    for (int p = 0; p < 3; ++p) {
                        if (!(MASK & (2 << p))) continue;
                        if (lane < VW / 4)
                            *reinterpret_cast<uint4*>(carry_out + ((size_t)cd * 3 + p) * VW + 4 * lane) =
                                *reinterpret_cast<const uint4*>(ringD + p * PATHW + so + 4 * lane);
                    }
                    if (lane < 4) cmin_out[(size_t)cd * 4 + lane] = ringD[so + VW + lane];
                    __syncwarp();
                    if (lane == 0) {
                        sts_release_cta(&s_done[nact], cd + 2);          // the slot of column cd is free again
                        if ((cd & (SW_FLAG_EVERY - 1)) == SW_FLAG_EVERY - 1 || cd == w1 - 1) st_release_gpu(prog_mine, cd + 1);
                    }
                }
            }
            continue;
        }

        if (gLast < 0) continue;                             // no active row in this warp
        // ---- the rows: group grp handles row r0 + rw, column c = step - SKEW * grp
        const int sr = min(r0 + rw, nrows - 1);              // (an idle group re-reads a valid row, stores nothing)
        const int row = dir > 0 ? sr : nrows - 1 - sr;       // storage row within the job
        const size_t rowbase = (((size_t)job * g.rowsMax + row) * w1) * VW;
        const int x0 = dir > 0 ? 0 : w1 - 1;
        const ptrdiff_t cstep = (ptrdiff_t)dir * VW;
        const uint32_t* pC0 = a.C + rowbase + (size_t)x0 * VW + WPL * l;
        uint32_t* pS0 = a.S + rowbase + (size_t)x0 * VW + WPL * l;
        auto ccol = [&](int c) { return pC0 + (ptrdiff_t)min(max(c, 0), w1 - 1) * cstep; };   // this lane's slice of column c (clamped)
        const int nsteps = w1 + SKEW * gLast;
#pragma unroll
        for (int s = 0; s < PFD; ++s) {
            cp_async_vec_s<WPL>(cst_s + s * VW * 4, ccol(s - SKEW * grp));
            asm volatile("cp.async.commit_group;");
        }
        uint32_t Lh[WPL];
#pragma unroll
        for (int k = 0; k < WPL; ++k) Lh[k] = 0u;
        uint32_t minh2 = 0;
        int seenP = 0, seenS = 0;
        // shared-space byte addresses of this lane's slice in the input ring slots of columns c-1, c, c+1 (path 0)
        const uint32_t rin_s = smem_u32(ring_in + WPL * l), rin_end = rin_s + PATHW * 4;
        auto slot_of = [&](int c) { return rin_s + (uint32_t)(((c % RING) + RING) % RING) * (ROWW * 4); };
        uint32_t a_m1 = slot_of(-SKEW * grp - 1), a_0 = slot_of(-SKEW * grp), a_p1 = slot_of(-SKEW * grp + 1);
        const uint32_t tail = (uint32_t)(VW - WPL * l) * 4;   // from a lane's slice to the slot's minima
        uint32_t cs_cur = 0, cs_prev = (NS - 1) * VW * 4;
        const int* done_pred = warp > 0 ? &s_done[warp * G - 1] : s_load;
        const int* done_succ = &s_done[warp * G + gLast + 1];
        const uint32_t done_mine = smem_u32(&s_done[rw]);
        long long spins = 0;
        const uint32_t z_vec = smem_u32(zrow + WPL * l), z_min = smem_u32(zrow + VW);
        const uint32_t* gC = ccol(PFD - SKEW * grp);          // this lane's slice of the column requested at this step
        uint32_t* pS = pS0;                                   // ... of S1 at this group's column (from column 0 on)
        int lim = -1;                                         // steps <= lim may run without looking at the neighbours' counters

        for (int st = 0; st < nsteps; ++st) {
            const int c = st - SKEW * grp;                   // this group's column
            const bool incol = (unsigned)c < (unsigned)w1, valid = rowAct && incol;
            // group 0 needs the row above (previous warp / loader) SKEW columns ahead; the last active group needs its
            // ring slot back from the row below (next warp / drain)
            if (st > lim) {
                const int needP = min(st + SKEW, w1), needS = st - SKEW * gLast - RING + 2;
                while (seenP < needP || seenS < needS) {
                    seenP = lds_acquire_cta(done_pred);
                    seenS = lds_acquire_cta(done_succ);
                    if (seenP < needP || seenS < needS) {
                        if (++spins > SW_SPIN_LIMIT) { if (lane == 0) atomicExch(a.err, 1); seenP = seenS = w1 + 64; }
                        __nanosleep(SW_SLEEP_NS);
                    }
                }
                lim = min(seenP >= w1 ? 0x3fffffff : seenP - SKEW, seenS + RING - 2 + SKEW * gLast);
            }
            cp_async_vec_s<WPL>(cst_s + cs_prev, gC);
            if ((unsigned)(c + PFD) < (unsigned)(w1 - 1)) gC += cstep;      // next request: column clamp(c + 1 + PFD)
            asm volatile("cp.async.commit_group;");
            asm volatile("cp.async.wait_group %0;" ::"n"(PFD) : "memory");
            uint32_t C[WPL];
            lds_vec<WPL, 0>(cst_s + cs_cur, C);
            uint32_t Lt[3][WPL], mq2[3];
            // (predecessors outside the image -- column -1 for the diagonal from behind, column w1 for the one from
            //  ahead -- are read from the row of zeros)
            if (MASK & 2) {
                const bool out = c <= 0;
                lds_vec<WPL, 0>(out ? z_vec : a_m1, Lt[0]);
                mq2[0] = lds_u32<0>(out ? z_min : a_m1 + tail);
            }
            if (MASK & 4) { lds_vec<WPL, PATHW * 4>(a_0, Lt[1]); mq2[1] = lds_u32<4>(a_0 + tail); }
            if (MASK & 8) {
                const bool out = c == w1 - 1;
                lds_vec<WPL, 2 * PATHW * 4>(out ? z_vec - 2 * PATHW * 4 : a_p1, Lt[2]);
                mq2[2] = lds_u32<8>(out ? z_min - 8 : a_p1 + tail);
            }
            if (c <= 0) {                                    // column 0: nothing to the left
#pragma unroll
                for (int k = 0; k < WPL; ++k) Lh[k] = 0u;
                minh2 = 0u;
            }
            uint32_t sum[WPL], mh, mv[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu};
            mh = step(Lh, C, minh2);
#pragma unroll
            for (int k = 0; k < WPL; ++k) sum[k] = Lh[k];
#pragma unroll
            for (int p = 0; p < 3; ++p) {
                if (!(MASK & (2 << p))) continue;
                mv[p] = step(Lt[p], C, mq2[p]);
#pragma unroll
                for (int k = 0; k < WPL; ++k) sum[k] = add_fma(Lt[p][k], one, sum[k]);
            }
            // minima over the row: two paths per butterfly
            uint32_t m0, m1, m2;
            if constexpr (MASK == 15) {
                gmin2(mh, mv[0], minh2, m0);
                gmin2(mv[1], mv[2], m1, m2);
            } else {                                         // MASK 5: -> and the vertical path
                gmin2(mh, mv[1], minh2, m1);
                m0 = m2 = 0u;
            }
            if (valid) {
                // the output ring is the next row's input ring: RINGW words further; group lane 0's slice starts the slot
                if (MASK & 2) { sts_vec<WPL, RINGW * 4>(a_0, Lt[0]); if (l == 0) sts_u32<(RINGW + VW) * 4>(a_0, m0); }
                if (MASK & 4) { sts_vec<WPL, (RINGW + PATHW) * 4>(a_0, Lt[1]); if (l == 0) sts_u32<(RINGW + VW + 1) * 4>(a_0, m1); }
                if (MASK & 8) { sts_vec<WPL, (RINGW + 2 * PATHW) * 4>(a_0, Lt[2]); if (l == 0) sts_u32<(RINGW + VW + 2) * 4>(a_0, m2); }
            }
            __syncwarp();
            if (valid && l == 0) asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(done_mine), "r"(c + 1) : "memory");
            if (!FULL) {                                     // pad disparities: the saturated sum of walls, as k_sweep<VAR 0> stores it
#pragma unroll
                for (int k = 0; k < WPL; ++k) sum[k] = (sum[k] & ~padm[k]) | (WALL2 & padm[k]);
            }
            if (valid) {
#pragma unroll
                for (int q = 0; q < WPL / 4; ++q)
                    reinterpret_cast<uint4*>(pS)[q] = make_uint4(sum[4 * q], sum[4 * q + 1], sum[4 * q + 2], sum[4 * q + 3]);
            }
            if (incol) pS += cstep;
            a_m1 = a_0; a_0 = a_p1; a_p1 = a_p1 + ROWW * 4 == rin_end ? rin_s : a_p1 + ROWW * 4;
            cs_prev = cs_cur; cs_cur = cs_cur + VW * 4 == NS * VW * 4 ? 0u : cs_cur + VW * 4;
        }
        __syncwarp();
        if (rowAct && l == 0) sts_release_cta(&s_done[rw], w1 + 64);   // a final count no column index reaches
        asm volatile("cp.async.wait_all;" ::: "memory");
    }
}

// ------------------------------------------------------------------------------------------
// k_hlast<G>: the last sweep of MODE_SGBM / SGBM_3WAY (one horizontal path + S = S1 + L + WTA, rows independent of
// each other) for numDisparities 64 and 128, where k_sweep's warp-per-pixel layout leaves three quarters / half of
// every packed instruction's lanes' worth of work on the table: a warp carries G = 4 / 2 ROWS at once, each in a group
// of WD = 32 / G lanes with 4 words (8 disparities) per lane -- the instruction stream of the 256-disparity case
// serving G pixels.  No rings, no flags, no helper warps: neighbours by group-relative shuffles, reductions by xor
// butterflies inside the group, C and S1 straight from global memory two columns ahead in registers.
// Same arithmetic as k_sweep<NP, VAR >= 1, MASK 1, LAST> (the tests compare both against the oracle).
// ------------------------------------------------------------------------------------------
#define HL_WARPS 8
// G rows per warp in groups of WD = 32 / G lanes, WPL = NP * G words per lane.  <1,4> and <2,2> are the multi-row forms
// (4 words per lane); <NP,1> is one row per warp for any width (what k_sweep<NP, ., 1, true> computes, without its
// ring / ticket / staging bookkeeping and with several CTAs per SM).
template <int NP, int G, bool FULL>
__global__ void __launch_bounds__(HL_WARPS * 32) k_hlast(const SweepArgs a, const SgbmGeo g)
{
    constexpr int VW = 32 * NP, WD = 32 / G, HF = 32 * NP, WPL = NP * G;     // HF = g.HALF
    extern __shared__ __align__(16) uint32_t smem[];                          // [HL_WARPS][G][VW]: S of the current column
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gl = lane & (WD - 1), grp = lane / WD;
    const int w1 = g.width1, dir = a.dir;
    const uint32_t P1P1 = pk_dup(g.P1), P2P2 = pk_dup(g.P2);
    const uint32_t one = a.njobs > 0, himask = 0u - (one << 16);
    const uint32_t sc_s = smem_u32(smem + ((size_t)warp * G + grp) * VW);    // this group's scratch vector
    const int blocksPerJob = (g.rowsMax + G - 1) / G;
    const long long units = (long long)a.njobs * blocksPerJob;
    const uint32_t selPrev = gl == 0 ? 0x1054u : 0x3210u, selNext = gl == WD - 1 ? 0x5432u : 0x3210u;
    uint32_t padm[WPL];                                      // !FULL (D < Dp): pad disparities keep the wall and never count
#pragma unroll
    for (int k = 0; k < WPL; ++k) {
        const int w = WPL * gl + k;
        padm[k] = (w >= g.D ? 0xffffu : 0u) | (w + g.HALF >= g.D ? 0xffff0000u : 0u);
    }
    auto gmin = [&](uint32_t v) {                                            // min over the group, in every lane of it
        if constexpr (G == 1) return __reduce_min_sync(FULLMASK, v);
        else {
#pragma unroll
            for (int o = WD / 2; o; o >>= 1) v = min(v, __shfl_xor_sync(FULLMASK, v, o));
            return v;
        }
    };
    auto gsum = [&](uint32_t v) {
        if constexpr (G == 1) return __reduce_add_sync(FULLMASK, v);
        else {
#pragma unroll
            for (int o = WD / 2; o; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
            return v;
        }
    };
    for (long long u = (long long)blockIdx.x * HL_WARPS + warp; u < units; u += (long long)gridDim.x * HL_WARPS) {
        const int job = (int)(u / blocksPerJob), rb = (int)(u % blocksPerJob);
        int ys, ye, keep;
        job_rows(g, job % g.nstripes, ys, ye, keep);
        const int nrows = ye - ys;
        const int row0 = rb * G + grp;
        const bool valid = row0 < nrows && ys + row0 >= keep;               // 3WAY overlap rows: nothing depends on them
        if (!__any_sync(FULLMASK, valid)) continue;
        const int row = min(row0, nrows - 1);                               // an idle group re-reads a valid row, stores nothing
        const int y = ys + row, b = job / g.nstripes;
        const size_t rowbase = (((size_t)job * g.rowsMax + row) * w1) * VW;
        const int x0 = dir > 0 ? 0 : w1 - 1;
        const ptrdiff_t cstep = (ptrdiff_t)dir * VW;
        const uint32_t* pC = a.C + rowbase + (size_t)x0 * VW;
        const uint32_t* pS = a.S + rowbase + (size_t)x0 * VW;
        // two columns ahead in registers; the loop is unrolled by two so that each buffer is refilled in place (no
        // register rotation), pointers advance incrementally and stop at the row's last column
        uint32_t cb[2][WPL], sb[2][WPL];
        vec_ld<WPL>(pC, gl, cb[0], LdNc()); vec_ld<WPL>(pS, gl, sb[0], LdNc());
        const uint32_t *qC = pC + (w1 > 1 ? cstep : 0), *qS = pS + (w1 > 1 ? cstep : 0);
        vec_ld<WPL>(qC, gl, cb[1], LdNc()); vec_ld<WPL>(qS, gl, sb[1], LdNc());
        int left = w1 - 2;                                   // columns beyond the two in flight
        uint32_t Lh[WPL];
#pragma unroll
        for (int k = 0; k < WPL; ++k) Lh[k] = 0;
        uint32_t minh2 = 0;
        uint32_t rkey = 0, rcnt = 0, rtap = 0;
        const size_t rowpix = ((size_t)b * g.H + y) * g.W;
        auto column = [&](const int i, uint32_t (&cbuf)[WPL], uint32_t (&sbuf)[WPL]) {
            uint32_t C[WPL], S[WPL];
#pragma unroll
            for (int k = 0; k < WPL; ++k) { C[k] = cbuf[k]; S[k] = sbuf[k]; }
            if (left > 0) { qC += cstep; qS += cstep; }
            --left;
            vec_ld<WPL>(qC, gl, cbuf, LdNc()); vec_ld<WPL>(qS, gl, sbuf, LdNc());     // column i + 2 (or the last one again)
            // ---- horizontal path (A.3), predecessor = the previous column
            uint32_t prev = __shfl_sync(FULLMASK, Lh[WPL - 1], (gl + WD - 1) & (WD - 1), WD);
            uint32_t next = __shfl_sync(FULLMASK, Lh[0], (gl + 1) & (WD - 1), WD);
            prev = __byte_perm(prev, WALL2, selPrev);        // group lane 0: (lo = wall, hi = last lane's lo); others: unchanged
            next = __byte_perm(next, WALL2, selNext);        // last group lane: (lo = lane 0's hi, hi = wall)
            const uint32_t mqP2 = minh2 + P2P2;
            uint32_t up = prev, mloc = 0xffffffffu;
#pragma unroll
            for (int k = 0; k < WPL; ++k) {
                const uint32_t dn = k + 1 < WPL ? Lh[k + 1] : next;
                uint32_t t = add_fma(pk_min(up, dn), one, P1P1);
                t = pk_min(pk_min(t, Lh[k]), mqP2);
                uint32_t L = C[k] + t - minh2;
                if (!FULL) L = (L & ~padm[k]) | (WALL2 & padm[k]);
                up = Lh[k];
                Lh[k] = L;
                mloc = pk_min(mloc, L);
            }
            minh2 = gmin(pk_min(mloc, __byte_perm(mloc, 0, 0x1032)));        // equal halves: u32 min = packed min
#pragma unroll
            for (int k = 0; k < WPL; ++k) S[k] = pk_addmin(S[k], Lh[k], WALL2);
            // ---- WTA (A.4), as in k_sweep: keys (S << 16 | d), lowest d wins ties
            uint32_t kmin = 0xffffffffu;
#pragma unroll
            for (int k = 0; k < WPL; ++k) {
                const uint32_t klo = S[k] * 65536u + (uint32_t)k;
                const uint32_t khi = (S[k] & himask) | (uint32_t)(HF + k);
                kmin = min(kmin, min(klo, khi));
            }
            kmin = gmin(kmin + (uint32_t)(WPL * gl));
            const int th = wta_threshold<true>(g, (int)(kmin >> 16));
            const uint32_t th2 = pk_dup((uint32_t)th);
            uint32_t lt[WPL];
#pragma unroll
            for (int k = 0; k < WPL; ++k) { lt[k] = __vsub2(S[k], th2); if (!FULL) lt[k] &= ~padm[k]; }
            uint32_t bits = 0;
            if constexpr (WPL == 4) {
                bits = (__byte_perm(lt[0], lt[1], 0x7351) & 0x80808080u) | ((__byte_perm(lt[2], lt[3], 0x7351) & 0x80808080u) >> 1);
            } else if constexpr (WPL == 2) {
                bits = __byte_perm(lt[0], lt[1], 0x7351) & 0x80808080u;
            } else {
#pragma unroll
                for (int k = 0; k < WPL; ++k) bits |= (lt[k] >> k) & (0x80008000u >> k);
            }
            const uint32_t cnt = gsum((uint32_t)__popc(bits));
            sts_vec<WPL, 0>(sc_s + 4 * WPL * gl, S);
            __syncwarp();
            const int best = (int)(kmin & 0xffffu);
            auto tap = [&](int d) {
                uint32_t v;
                const uint32_t ad = sc_s + 4u * (uint32_t)d - (d >= HF ? (uint32_t)(4 * HF - 2) : 0u);
                asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(ad));
                return v;
            };
            const uint32_t tm1 = tap(max(best - 1, 0)), tp1 = tap(min(best + 1, g.D - 1));
            const uint32_t sm1v = best > 0 ? tm1 : 0u, sp1v = best < g.D - 1 ? tp1 : 0u;
            if (gl == (i & (WD - 1))) { rkey = kmin; rcnt = cnt; rtap = sm1v | (sp1v << 16); }
            if ((i & (WD - 1)) == WD - 1 || i == w1 - 1) {
                // WD columns of this row decided at once: lane j of the group finishes column (i & ~(WD-1)) + j
                const int ci = (i & ~(WD - 1)) + gl;
                if (ci <= i && valid) {
                    const int xc = dir > 0 ? ci : w1 - 1 - ci;
                    const int x = xc + g.minX1;
                    const int minS = (int)(rkey >> 16), bst = (int)(rkey & 0xffffu);
                    const int thr = wta_threshold<true>(g, minS);
                    const int t1 = (int)(rtap & 0xffffu), t2 = (int)(rtap >> 16);
                    int out = g.INV;
                    if (minS < (int)WALL) {
                        const int excl = (minS < thr) + (bst > 0 && t1 < thr) + (bst < g.D - 1 && t2 < thr);
                        if ((int)rcnt <= excl) {
                            int dsub = bst * 16;
                            if (bst > 0 && bst < g.D - 1) {
                                const int den = max(t1 + t2 - 2 * minS, 1);
                                dsub += ((t1 - t2) * 16 + den) / (2 * den);
                            }
                            out = dsub + g.minD * 16;
                            atomicMin(a.key + rowpix + (x - bst - g.minD), ((uint32_t)minS << 16) | (uint32_t)(0xffff - x));
                        }
                    }
                    a.raw[rowpix + x] = (int16_t)out;
                }
            }
            __syncwarp();                                    // the scratch vector is rewritten by the next column
        };
        int i = 0;
        for (; i + 1 < w1; i += 2) { column(i, cb[0], sb[0]); column(i + 1, cb[1], sb[1]); }
        if (i < w1) column(i, cb[0], sb[0]);
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

// raw[] = INV: eight values per thread (one 128-bit store; `p` is a 256-byte aligned workspace slice), the tail singly
__global__ void __launch_bounds__(256) k_fill_s16(int16_t* p, size_t n, int16_t v)
{
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i + 8 <= n) {
        const uint32_t w = (uint32_t)(uint16_t)v * 0x10001u;
        *reinterpret_cast<uint4*>(p + i) = make_uint4(w, w, w, w);
    } else {
        for (size_t j = i; j < n; ++j) p[j] = v;
    }
}

// ------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------
struct svb_sgbm {
    svb_sgbm_params prm;
    size_t ws_limit = 0;            // 0: 60 % of the free device memory at first use
    SvbBuf ws;
    SvbBuf stage_in, stage_out;     // device staging for the *_host entry point
    SvbPinBuf pin_in, pin_out;      // pinned staging when the caller's host buffers are pageable (NumPy arrays)
    std::vector<cudaEvent_t> ev_mark;   // one per downloaded sub-range of such a call
    // *_host pipeline: H2D of sub-batch i+1 and D2H of sub-batch i-1 overlap the kernels of sub-batch i
    cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
    std::vector<cudaEvent_t> ev_in, ev_done;
    int launches = 0;
    int num_sms = 0;
    bool debug = false;             // keep the final S volume (debug_stages)
    int force_wave = -1;            // test hook: 1 / 0 = k_wave / k_sweep for the first sweep whatever the batch size
    // optional per-stage timing with CUDA events on the launching stream (bench.py's roofline)
    bool profiling = false;
    struct Ev { int cat; cudaEvent_t a, b; };
    std::vector<Ev> evs;
    // last compute (for debug_stages)
    SgbmGeo last_geo{};
    size_t off_C = 0, off_S = 0, off_raw = 0;
    bool have_last = false;
    // The wavefront's error flag (a wait that ran into its spin limit) lives in its own small allocation and is
    // STICKY: kernels only ever set it.  Every compute copies it to a pinned host word behind its kernels and
    // records ev_last there; the next compute on the handle (or svb_sgbm_check) reports a tripped flag.
    // ev_last also orders successive computes of the handle that arrive on DIFFERENT streams: they share one
    // workspace (ticket counter, progress flags, carry, C and S volumes) and must not overlap on the device.
    int* d_err = nullptr;
    int* h_err = nullptr;           // pinned
    cudaEvent_t ev_last = nullptr;
    bool pending = false;           // a compute has been enqueued whose flag has not been looked at yet
};

static int make_geo(const svb_sgbm_params& p, int H, int W, int cn, SgbmGeo& g)
{
    if (H <= 0 || W <= 0) { svb_set_error("empty image (%dx%d)", W, H); return SVB_ERR_PARAM; }
    if (p.numDisparities <= 0) { svb_set_error("numDisparities must be > 0"); return SVB_ERR_PARAM; }
    if (p.blockSize < 1) { svb_set_error("blockSize must be >= 1"); return SVB_ERR_PARAM; }
    if (p.numDisparities % 16 != 0) {
        svb_set_error("numDisparities=%d: this engine requires a multiple of 16", p.numDisparities);
        return SVB_ERR_UNSUPPORTED;
    }
    if (p.mode < 0 || p.mode > 3) { svb_set_error("bad mode %d", p.mode); return SVB_ERR_PARAM; }
    if (cn != 1 && cn != 3) { svb_set_error("images must have 1 or 3 channels (got %d)", cn); return SVB_ERR_PARAM; }
    g.H = H; g.W = W; g.cn = cn; g.D = p.numDisparities; g.minD = p.minDisparity;
    g.NP = (g.D + 63) / 64; g.Dp = 64 * g.NP; g.HALF = 32 * g.NP;
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
    if (g.bs > 25) { svb_set_error("blockSize %d > 25 unsupported", g.bs); return SVB_ERR_UNSUPPORTED; }
    g.ftzero = std::max(p.preFilterCap, 15) | 1;
    g.P1 = p.P1 > 0 ? p.P1 : 2;
    g.P2 = std::max(p.P2 > 0 ? p.P2 : 5, g.P1 + 1);
    g.uniq = p.uniquenessRatio >= 0 ? p.uniquenessRatio : 10;
    g.d12 = p.disp12MaxDiff > 0 ? p.disp12MaxDiff : 1;
    g.thMul = (100 - g.uniq) > 1 ? 0xffffffffu / (unsigned)(100 - g.uniq) + 1u : 0u;
    g.INV = (g.minD - 1) * 16;
    g.mode = p.mode;
    g.nstripes = p.mode == SVB_MODE_SGBM_3WAY ? 4 : 1;
    const int win = 2 * g.r + 1;
    const long long cmax = (long long)cn * win * win * (2 * g.ftzero + 63);
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
    g.Wp = ((W + g.HALF + 3) & ~3) + 4;
    g.rowPitch = (long long)W * cn; g.imgPitch = (long long)H * W * cn;
    return SVB_OK;
}

struct SgbmLayout { size_t ldesc, rdesc, C, S, carry, cmin, sync, raw, key, scratch, total; int bands; };
static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
static SgbmLayout make_layout(const SgbmGeo& g, int nb)
{
    SgbmLayout L; size_t o = 0;
    const size_t njobs = (size_t)nb * g.nstripes;
    // progress flags: one per band of the SHORTEST bands a sweep kernel of this width uses
    const int brmin = g.NP <= 2 ? std::min(sw_rows(g.NP), wv_rows(g.NP)) : sw_rows(g.NP);
    L.bands = (g.rowsMax + brmin - 1) / brmin;
    L.ldesc = o; o += al256((size_t)nb * g.H * g.W * 32);
    L.rdesc = o; o += al256((size_t)nb * g.H * 6 * g.Wp * 4);
    const size_t vol = njobs * g.rowsMax * g.width1 * g.Dp * 2;
    L.C = o; o += al256(vol);
    L.S = o; o += al256(vol);
    L.carry = o; o += al256(njobs * 2 * g.width1 * 3 * g.Dp * 2);
    L.cmin = o; o += al256(njobs * 2 * g.width1 * 4 * 4);
    L.sync = o; o += al256((njobs * g.rowsMax + 64) * 4);    // [0] ticket, [1] err, [16..] prog (one per band; bands <= rows)
    L.raw = o; o += al256((size_t)nb * g.H * g.W * 2);
    L.key = o; o += al256((size_t)nb * g.H * g.W * 4);
    L.scratch = o; o += al256((size_t)nb * g.H * g.W * 8);
    L.total = o;
    return L;
}
static size_t ws_limit_of(const svb_sgbm* h)
{
    if (h->ws_limit) return h->ws_limit;
    size_t fr = 0, tot = 0;
    if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) { cudaGetLastError(); return (size_t)24 << 30; }
    // what this handle already holds is "free" for its own purposes
    return (size_t)(0.6 * (double)(fr + h->ws.cap));
}
static int pick_chunk(const svb_sgbm* h, const SgbmGeo& g, int B)
{
    // the workspace of an earlier call already holds the whole batch: no cudaMemGetInfo (it costs ~0.1 ms,
    // which is most of a single 640x480 frame)
    if (!h->ws_limit && h->ws.cap && make_layout(g, B).total <= h->ws.cap) return B;
    const size_t per = make_layout(g, 1).total;
    size_t nb = ws_limit_of(h) / per;
    if (nb < 1) nb = 1;
    return (int)std::min<size_t>(nb, (size_t)B);
}

extern "C" int svb_sgbm_create(const svb_sgbm_params* p, svb_sgbm_t** out)
{
    if (!p || !out) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    SgbmGeo g;
    // validate what does not depend on the image size with a generously wide dummy image
    int rc = make_geo(*p, 16, 65535, 1, g);
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
    for (auto& e : h->evs) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    for (auto& e : h->ev_in) cudaEventDestroy(e);
    for (auto& e : h->ev_done) cudaEventDestroy(e);
    if (h->s_in) cudaStreamDestroy(h->s_in);
    if (h->s_run) cudaStreamDestroy(h->s_run);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    if (h->ev_last) cudaEventDestroy(h->ev_last);
    if (h->d_err) cudaFree(h->d_err);
    if (h->h_err) cudaFreeHost(h->h_err);
    for (auto& e : h->ev_mark) cudaEventDestroy(e);
    h->ws.release(); h->stage_in.release(); h->stage_out.release();
    h->pin_in.release(); h->pin_out.release();
    delete h;
}
extern "C" size_t svb_sgbm_workspace_bytes(const svb_sgbm_t* h, int B, int H, int W)
{
    SgbmGeo g;
    if (!h || make_geo(h->prm, H, W, 1, g) != SVB_OK) return 0;
    return make_layout(g, pick_chunk(h, g, B)).total;
}
extern "C" int svb_sgbm_set_workspace_limit(svb_sgbm_t* h, size_t bytes)
{
    if (!h) return SVB_ERR_PARAM;
    h->ws_limit = bytes;
    return SVB_OK;
}
extern "C" int svb_sgbm_set_debug(svb_sgbm_t* h, int enable)
{
    if (!h) return SVB_ERR_PARAM;
    h->debug = enable != 0;
    return SVB_OK;
}
extern "C" int svb_sgbm_last_launches(const svb_sgbm_t* h) { return h ? h->launches : 0; }

enum { CAT_PREFILTER = 0, CAT_COST, CAT_SWEEP1, CAT_SWEEP2, CAT_LR, CAT_MEDIAN, CAT_SPECKLE, CAT_COUNT };
static const char* const kCatNames[CAT_COUNT] = {"k_prefilter", "k_cost", "k_sweep<first>", "k_sweep<last+wta>",
                                                 "k_lrcheck", "k_median3_s16", "k_ccl_*"};
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

template <int NP, int VAR, int MASK, bool LAST>
static int launch_sweep(svb_sgbm* h, const SweepArgs& a, const SgbmGeo& g, cudaStream_t st)
{
    constexpr int SW_R = sw_rows(NP);
    const size_t smem = sw_smem_words(NP, LAST) * 4;
    auto kfn = k_sweep<NP, VAR, MASK, LAST>;
    SVB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    SVB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, (SW_R + 2) * 32, smem));
    if (per_sm < 1) { svb_set_error("k_sweep does not fit on an SM (smem %zu)", smem); return SVB_ERR_UNSUPPORTED; }
    const long long tickets = (long long)a.njobs * a.bandsPerJob;
    const int grid = (int)std::min<long long>(tickets, (long long)h->num_sms * per_sm);
    // ticket = 0, err untouched, prog = 0
    SVB_CUDA(cudaMemsetAsync(a.ticket, 0, 4, st));
    SVB_CUDA(cudaMemsetAsync(a.prog, 0, (size_t)tickets * 4, st));
    kfn<<<grid, (SW_R + 2) * 32, smem, st>>>(a, g);   // SW_R rows + the two carry helper warps
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
template <int NP, int MASK, bool FULL>
static int launch_wave(svb_sgbm* h, const SweepArgs& a, const SgbmGeo& g, cudaStream_t st)
{
    auto kfn = k_wave<NP, MASK, FULL>;
    const size_t smem = wv_smem_words(NP) * 4;
    SVB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    SVB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, (wv_warps(NP) + 2) * 32, smem));
    if (per_sm < 1) { svb_set_error("k_wave does not fit on an SM (smem %zu)", smem); return SVB_ERR_UNSUPPORTED; }
    const long long tickets = (long long)a.njobs * a.bandsPerJob;
    const int grid = (int)std::min<long long>(tickets, (long long)h->num_sms * per_sm);
    SVB_CUDA(cudaMemsetAsync(a.ticket, 0, 4, st));
    SVB_CUDA(cudaMemsetAsync(a.prog, 0, (size_t)tickets * 4, st));
    kfn<<<grid, (wv_warps(NP) + 2) * 32, smem, st>>>(a, g);
    SVB_LAUNCH_CHECK();
    return SVB_OK;
}
template <int NP, int MASK, bool LAST>
static int launch_sweep_var(svb_sgbm* h, const SweepArgs& a, const SgbmGeo& g, int var, cudaStream_t st)
{
    switch (var) {
        case 0: return launch_sweep<NP, 0, MASK, LAST>(h, a, g, st);
        case 1: return launch_sweep<NP, 1, MASK, LAST>(h, a, g, st);
        default: return launch_sweep<NP, 2, MASK, LAST>(h, a, g, st);
    }
}

// The *_host entry points hand run_chunk their host buffers: the per-image stages in front of the sweeps
// (prefilter, cost) and behind them (LR check, median, speckle) run on sub-ranges of the chunk, so that the
// upload of sub-range i+1 and the download of sub-range i-1 overlap them.  The wavefront sweeps always see the
// whole chunk (their ramp-up/-down is what makes small launches slow).
struct HostPipe {
    const uint8_t *hl, *hr;            // host images of the whole batch
    int16_t* hout;                     // host disparities of the whole batch
    size_t pin, pout;                  // bytes per image in, elements per image out
    int nsub;                          // sub-ranges per chunk
    // pageable caller buffers: hl / hr / hout are pinned staging; staged[i] turns 1 once image i of both inputs has
    // been copied there by the host pool, marks collects (event after the download, first image, count)
    std::atomic<int>* staged = nullptr;
    struct Mark { cudaEvent_t ev; int first, count; };
    std::vector<Mark>* marks = nullptr;
};

template <int NP>
static int run_chunk(svb_sgbm* h, const SgbmGeo& g, const SgbmLayout& L, const uint8_t* left, const uint8_t* right,
                     int nb, int16_t* disp, cudaStream_t st, const HostPipe* hp, int base)
{
    char* ws = (char*)h->ws.p;
    uint32_t* ldesc = (uint32_t*)(ws + L.ldesc);
    uint32_t* rdesc = (uint32_t*)(ws + L.rdesc);
    uint32_t* Cv = (uint32_t*)(ws + L.C);
    uint32_t* Sv = (uint32_t*)(ws + L.S);
    int* sync = (int*)(ws + L.sync);
    int16_t* raw = (int16_t*)(ws + L.raw);
    uint32_t* key = (uint32_t*)(ws + L.key);
    int32_t* scratch = (int32_t*)(ws + L.scratch);
    const int njobs = nb * g.nstripes;

    const size_t hw = (size_t)g.H * g.W;
    const int nsub = hp ? std::min(hp->nsub, nb) : 1, per = (nb + nsub - 1) / nsub;
    int evi = 0;
    for (int s0 = 0; s0 < nb; s0 += per) {
    const int ns = std::min(per, nb - s0);
    if (hp) {                                               // upload this sub-range while the previous one computes
        if (hp->staged)
            for (int i = base + s0; i < base + s0 + ns; ++i)
                while (!hp->staged[i].load(std::memory_order_acquire)) std::this_thread::yield();
        uint8_t* dl = const_cast<uint8_t*>(left) + s0 * hp->pin;
        uint8_t* dr = const_cast<uint8_t*>(right) + s0 * hp->pin;
        SVB_CUDA(cudaMemcpyAsync(dl, hp->hl + (base + s0) * hp->pin, ns * hp->pin, cudaMemcpyHostToDevice, h->s_in));
        SVB_CUDA(cudaMemcpyAsync(dr, hp->hr + (base + s0) * hp->pin, ns * hp->pin, cudaMemcpyHostToDevice, h->s_in));
        SVB_CUDA(cudaEventRecord(h->ev_in[evi], h->s_in));
        SVB_CUDA(cudaStreamWaitEvent(st, h->ev_in[evi], 0));
        ++evi;
    }
    // colour input: one prefilter + cost pass per channel, the box sums add up in C (A.1: cv2 sums the BT cost
    // of all 2*cn channels before the box filter, and the box filter is linear)
    for (int ch = 0; ch < g.cn; ++ch) {
    {
        ProfScope ps(h, CAT_PREFILTER, st);
        k_prefilter<<<dim3(svb_ceil_div(std::max(g.W, g.Wp), 256), g.H, ns), 256, 0, st>>>(
            left + s0 * g.imgPitch, right + s0 * g.imgPitch, ns, ch, g, ldesc + s0 * hw * 8,
            rdesc + (size_t)s0 * g.H * 6 * g.Wp);
    }
    SVB_LAUNCH_CHECK(); h->launches++;

    {
        // k_cost tiling: ring of bs rows per output column in shared memory
        const int bs = 2 * g.r + 1, VW = NP * 32;
        CostArgs ca;
        ca.accum = ch > 0;
        ca.job0 = s0 * g.nstripes;
        const int njobs = ns * g.nstripes;                   // jobs of this launch
        // numDisparities <= 64: k_costg (4 adjacent columns per warp pass, strips of 64 columns).  At numDisparities 128 (2
        // columns per pass) it was measured slower than k_cost: 8.06 -> 8.44 ms at 720p -- the pixel phase executes the same
        // number of loads there, only the box phase gets shorter.
        constexpr bool GROUPED = CK_GROUPED && NP == 1;
        constexpr int PCW = GROUPED ? CK_WARPS * (4 / (NP <= 2 ? NP : 1)) : CK_PC;   // pixel-cost columns per strip
        constexpr int SEW = GROUPED ? 32 * NP + PCW + 8 : ck_se(NP);
        constexpr int NBW = GROUPED ? 1 : ck_nblk(NP);
        constexpr int BLW = GROUPED ? ((6 * SEW + 31) & ~31) : ck_blkw(NP);
        const size_t fixed = ((size_t)2 * NBW * BLW + 2 * PCW * 8 + (size_t)2 * PCW * VW) * 4;
        const size_t budget = (size_t)CK_BUDGET_KB * 1024;
        int OC = PCW - 2 * g.r;
        while (OC > 1 && fixed + (size_t)OC * bs * VW * 4 > budget) --OC;
        if (OC < 1 || fixed + (size_t)OC * bs * VW * 4 > budget) {
            svb_set_error("blockSize %d too large for the cost kernel at numDisparities %d", g.bs, g.D);
            return SVB_ERR_UNSUPPORTED;
        }
        // A strip whose ring leaves room for only ONE CTA per SM (numDisparities 320 and up) may do better narrower: two
        // resident CTAs hide each other's row barrier.  Measured: 18 instead of 28 output columns at numDisparities 320 /
        // blockSize 5 -> k_cost 4.72 -> 4.11 ms (480p, 64 pairs), 11.75 -> 10.59 ms (1080p, 16 pairs); 12 instead of 26-28
        // columns (blockSize 7, or numDisparities 384) -> equal or slower.  Hence: only if >= 60 % of the columns remain.
        if (CK_BUDGET_KB == 200 && fixed + (size_t)OC * bs * VW * 4 > (size_t)113 * 1024) {
            int OC2 = OC;
            while (OC2 > 1 && fixed + (size_t)OC2 * bs * VW * 4 > (size_t)109 * 1024) --OC2;
            if (fixed + (size_t)OC2 * bs * VW * 4 <= (size_t)109 * 1024 && 10 * OC2 >= 6 * OC) OC = OC2;
        }
        ca.OC = OC;
        ca.nstrips = svb_ceil_div(g.width1, OC);
        // Row segments: every CTA walks rseg rows (+ bs-1 warm-up rows, + a few rows' worth of start-up).  The CTAs of a launch
        // run in waves of (resident CTAs per SM) x SMs; take the segment count that minimises waves x rows per CTA.  Measured:
        // a host-pipe sub-range of 8 pairs at 1080p/D256 (480 strips on 296 slots) needs the cut (e2e 568 -> 601 pairs/s);
        // 720p/D128, 32 pairs: k_cost 8.06 -> 7.81 ms; 32 pairs at 1080p/D256 (6.5 waves) gain nothing from it -- an SM left
        // with one CTA in the last wave runs it at nearly twice the speed, the kernel is bound per SM, not per CTA.
        auto plan_segs = [&](int per_sm) {
            const long long slots = (long long)std::max(per_sm, 1) * h->num_sms, base = (long long)ca.nstrips * njobs;
            const int maxseg = std::max(1, std::min(CK_MAXSEG, g.rowsMax / (16 * bs)));
            long long best = 0;
            for (int n = 1; n <= maxseg; ++n) {
                const int rs = svb_ceil_div(g.rowsMax, n), nsg = svb_ceil_div(g.rowsMax, rs);
                const long long cost = ((base * nsg + slots - 1) / slots) * (rs + bs - 1 + 6);
                if (n == 1 || cost < best) { best = cost; ca.rseg = rs; ca.nsegs = nsg; }
            }
        };
        const size_t smem = fixed + (size_t)OC * bs * VW * 4;
        // TMA views: right planes [nb*H*6][Wp] u32 (box 6 x BW), left records [nb*H][W*8] u32 (box 1 x 256)
        CUtensorMap tmR, tmL;
        int rc = svb_make_tmap_2d_u32(&tmR, rdesc, (uint64_t)g.Wp, (uint64_t)nb * g.H * 6, (uint64_t)g.Wp * 4,
                                      SEW / NBW, 6);
        if (rc != SVB_OK) return rc;
        rc = svb_make_tmap_2d_u32(&tmL, ldesc, (uint64_t)g.W * 8, (uint64_t)nb * g.H, (uint64_t)g.W * 32, CK_PC * 8, 1);
        if (rc != SVB_OK) return rc;
        ProfScope ps(h, CAT_COST, st);
        const bool xtra = ca.accum || g.mode == SVB_MODE_HH4;
        int per_sm = 0;
#define LAUNCH_COST_K(K) do { \
            SVB_CUDA(cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            SVB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, K, CK_WARPS * 32, smem)); \
            plan_segs(per_sm); \
            K<<<dim3(ca.nstrips * ca.nsegs, njobs), CK_WARPS * 32, smem, st>>>(tmR, tmL, g, ca, Cv); } while (0)
        // blockSize 3 / 5 / 7 / 9 (the reference's scripts use 7, and 5 at half scale) have instances with the size built in
#define LAUNCH_COST_BS(KERN, F, X) do { \
            const int cbs = (CK_BSCONST && !(X)) ? g.bs : 0; \
            if (cbs == 5) LAUNCH_COST_K((KERN<NP, F, false, 5>)); \
            else if (cbs == 7) LAUNCH_COST_K((KERN<NP, F, false, 7>)); \
            else if (cbs == 3) LAUNCH_COST_K((KERN<NP, F, false, 3>)); \
            else if (cbs == 9) LAUNCH_COST_K((KERN<NP, F, false, 9>)); \
            else LAUNCH_COST_K((KERN<NP, F, X, 0>)); } while (0)
#define LAUNCH_COST(F, X) LAUNCH_COST_BS(k_cost, F, X)
#define LAUNCH_COSTG(F, X) LAUNCH_COST_BS(k_costg, F, X)
        if constexpr (GROUPED) {
            if (g.D == g.Dp) { if (xtra) LAUNCH_COSTG(true, true); else LAUNCH_COSTG(true, false); }
            else             { if (xtra) LAUNCH_COSTG(false, true); else LAUNCH_COSTG(false, false); }
        } else {
            if (g.D == g.Dp) { if (xtra) LAUNCH_COST(true, true); else LAUNCH_COST(true, false); }
            else             { if (xtra) LAUNCH_COST(false, true); else LAUNCH_COST(false, false); }
        }
#undef LAUNCH_COSTG
#undef LAUNCH_COST
#undef LAUNCH_COST_BS
#undef LAUNCH_COST_K
        SVB_LAUNCH_CHECK(); h->launches++;
    }
    }
    }

    const size_t npix = (size_t)nb * g.H * g.W;
    k_fill_s16<<<svb_ceil_div(npix, 256 * 8), 256, 0, st>>>(raw, npix, (int16_t)g.INV);
    SVB_LAUNCH_CHECK(); h->launches++;
    SVB_CUDA(cudaMemsetAsync(key, 0xff, npix * 4, st));

    // path sets (A.3): 3WAY {->, down | <-}; SGBM {->, \, |, / | <-}; HH {->, \, |, / | <-, \, |, /};
    // HH4 {->, down | <-, up}
    SweepArgs a;
    a.C = Cv; a.S = Sv;
    a.carry = (uint32_t*)(ws + L.carry); a.cmin = (uint32_t*)(ws + L.cmin);
    a.ticket = sync; a.err = h->d_err; a.prog = sync + 16;
    a.raw = raw; a.key = key;
    // Rows per band: sw_rows(NP).  A sweep WITHOUT inter-row paths (the last sweep of MODE_SGBM / 3WAY: <- only) has
    // independent rows, so a small batch (a single live frame) gets shorter bands there and spreads over all SMs
    // (640x480 frame: 0.32 -> 0.22 ms).  Sweeps with vertical paths keep full bands: their latency is the dependent
    // chain (columns + 2 x rows) x one warp's column latency (~0.6 us), and every extra band boundary adds ~16 columns
    // to it (measured for the first sweep of that frame: 18 / 12 / 9 / 6 / 4 rows per band -> 0.51 / 0.51 / 0.53 /
    // 0.57 / 0.64 ms).
    const int rows_full = sw_rows(NP);
    int rows_free = rows_full;
    if ((long long)njobs * L.bands < h->num_sms)
        rows_free = std::min(rows_full, std::max(4, (int)(((long long)njobs * g.rowsMax + h->num_sms - 1) / h->num_sms)));
    auto set_band_rows = [&](int br) { a.bandRows = br; a.bandsPerJob = (g.rowsMax + br - 1) / br; };
    set_band_rows(rows_full);
    a.njobs = njobs;
    a.storeS = h->debug ? 1 : 0;
    const int win = 2 * g.r + 1;
    const long long lmax = (long long)g.cn * win * win * (2 * g.ftzero + 63) + g.P2;
    int var = g.D == g.Dp ? 1 : 0;
    if (var == 1 && 4 * lmax <= 32767) var = 2;
    int rc;
    {
        ProfScope ps(h, CAT_SWEEP1, st);
        a.dir = 1;
        const bool two = g.mode == SVB_MODE_SGBM_3WAY || g.mode == SVB_MODE_HH4;
        bool wave = false;
        if constexpr (NP <= 2) {
            // numDisparities <= 128 with plain sums: several rows per warp (k_wave)
            // ... when the launch fills the GPU, or without the diagonal paths.  One live frame in MODE_SGBM is a dependent
            // chain of columns + 2 x rows steps, and a k_wave step (4 / 2 rows) is longer than a k_sweep step: measured
            // 480p / D128 1.59 -> 1.82 ms, 720p / D128 2.90 -> 3.29 ms for ONE frame, while 3WAY / HH4 (one-column skew)
            // are faster with k_wave even then (480p / D64 0.80 -> 0.77 ms)
            const long long wtickets = (long long)njobs * ((g.rowsMax + wv_rows(NP) - 1) / wv_rows(NP));
            wave = SW_WAVE && (var == 2 || (var == 0 && 4 * lmax <= 32767)) && (h->force_wave < 0 ? (two || wtickets >= h->num_sms) : h->force_wave != 0);
            if (wave) {
                set_band_rows(wv_rows(NP));
                if (g.D == g.Dp) rc = two ? launch_wave<NP, 5, true>(h, a, g, st) : launch_wave<NP, 15, true>(h, a, g, st);
                else rc = two ? launch_wave<NP, 5, false>(h, a, g, st) : launch_wave<NP, 15, false>(h, a, g, st);
                set_band_rows(rows_full);
            }
        }
        if (!wave) rc = two ? launch_sweep_var<NP, 5, false>(h, a, g, var, st) : launch_sweep_var<NP, 15, false>(h, a, g, var, st);
        if (rc != SVB_OK) return rc;
        h->launches++;
    }
    {
        ProfScope ps(h, CAT_SWEEP2, st);
        a.dir = -1;
        if (g.mode != SVB_MODE_HH && g.mode != SVB_MODE_HH4) set_band_rows(rows_free);
        // numDisparities 64 / 128 with a horizontal-only last sweep: several rows per warp (k_hlast); not for the debug
        // S dump, nor for cv2's 3WAY tie rule without a uniqueness test (k_sweep has that special case)
        const bool hlast = SW_HLAST && g.mode != SVB_MODE_HH && g.mode != SVB_MODE_HH4 && !h->debug &&
                           !(g.mode == SVB_MODE_SGBM_3WAY && g.uniq == 0);
        if (hlast) {
            constexpr int G = NP == 1 ? 4 : (NP == 2 ? 2 : 1);
            const long long units = (long long)njobs * ((g.rowsMax + G - 1) / G);
            const size_t smem = (size_t)HL_WARPS * G * 32 * NP * 4;
            int per_sm = 0;
            auto kfn = g.D == g.Dp ? k_hlast<NP, G, true> : k_hlast<NP, G, false>;
            SVB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, HL_WARPS * 32, smem));
            const int grid = (int)std::min<long long>((units + HL_WARPS - 1) / HL_WARPS, (long long)h->num_sms * std::max(per_sm, 1));
            kfn<<<grid, HL_WARPS * 32, smem, st>>>(a, g);
            SVB_LAUNCH_CHECK();
            rc = SVB_OK;
        } else
        rc = g.mode == SVB_MODE_HH    ? launch_sweep_var<NP, 15, true>(h, a, g, var, st)
             : g.mode == SVB_MODE_HH4 ? launch_sweep_var<NP, 5, true>(h, a, g, var, st)
                                      : launch_sweep_var<NP, 1, true>(h, a, g, var, st);
        if (rc != SVB_OK) return rc;
        h->launches++;
    }
    evi = 0;
    for (int s0 = 0; s0 < nb; s0 += per) {
        const int ns = std::min(per, nb - s0);
        {
            ProfScope ps(h, CAT_LR, st);
            k_lrcheck<<<dim3(svb_ceil_div(g.width1, 256), g.H, ns), 256, 0, st>>>(raw + s0 * hw, key + s0 * hw, g, ns);
        }
        SVB_LAUNCH_CHECK(); h->launches++;
        {
            ProfScope ps(h, CAT_MEDIAN, st);
            rc = svb_median3_s16_dev(raw + s0 * hw, ns, g.H, g.W, disp + s0 * hw, st);
        }
        if (rc != SVB_OK) return rc;
        h->launches++;
        if (h->prm.speckleWindowSize > 0) {
            ProfScope ps(h, CAT_SPECKLE, st);
            rc = svb_filter_speckles_s16_dev(disp + s0 * hw, ns, g.H, g.W, g.INV, h->prm.speckleWindowSize,
                                             16 * h->prm.speckleRange, scratch + 2 * s0 * hw, st);
            if (rc != SVB_OK) return rc;
            h->launches += 4;
        }
        if (hp) {                                           // download this sub-range while the next one finishes
            SVB_CUDA(cudaEventRecord(h->ev_done[evi], st));
            SVB_CUDA(cudaStreamWaitEvent(h->s_out, h->ev_done[evi], 0));
            SVB_CUDA(cudaMemcpyAsync(hp->hout + (base + s0) * hp->pout, disp + s0 * hw, ns * hw * 2,
                                     cudaMemcpyDeviceToHost, h->s_out));
            if (hp->marks) {
                const size_t k = hp->marks->size();
                while (h->ev_mark.size() <= k) {
                    cudaEvent_t e;
                    SVB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                    h->ev_mark.push_back(e);
                }
                SVB_CUDA(cudaEventRecord(h->ev_mark[k], h->s_out));
                hp->marks->push_back({h->ev_mark[k], base + s0, ns});
            }
            ++evi;
        }
    }
    return SVB_OK;
}

extern "C" int svb_sgbm_compute_dev(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right,
                                    int B, int H, int W, int16_t* disp, void* stream)
{
    return svb_sgbm_compute_cn_dev(h, left, right, B, H, W, 1, disp, stream);
}

static int sgbm_compute_impl(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right,
                             int B, int H, int W, int cn, int16_t* disp, void* stream, const HostPipe* hp,
                             long long row_pitch = 0, long long img_pitch = 0);

// the last compute has finished: look at the flag it copied out, clear a tripped one on both sides
static int take_sweep_err(svb_sgbm* h)
{
    h->pending = false;
    if (!*h->h_err) return SVB_OK;
    *h->h_err = 0;
    SVB_CUDA(cudaMemset(h->d_err, 0, sizeof(int)));
    svb_set_error("k_sweep: an inter-band wait exceeded its spin limit; the disparities of the computes on this "
                  "handle since the last check are invalid");
    return SVB_ERR_CUDA;
}

extern "C" int svb_sgbm_compute_cn_dev(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right,
                                       int B, int H, int W, int cn, int16_t* disp, void* stream)
{
    return sgbm_compute_impl(h, left, right, B, H, W, cn, disp, stream, nullptr);
}
// The same on a pitched input batch: row_pitch / img_pitch in BYTES between rows / images of both
// inputs (a crop of a larger frame, a padded capture buffer); 0 = dense.  The output stays dense.
extern "C" int svb_sgbm_compute_pitched_dev(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right, int B, int H, int W,
                                            int cn, ptrdiff_t row_pitch, ptrdiff_t img_pitch, int16_t* disp, void* stream)
{
    if (row_pitch < 0 || img_pitch < 0) { svb_set_error("negative pitch"); return SVB_ERR_PARAM; }
    return sgbm_compute_impl(h, left, right, B, H, W, cn, disp, stream, nullptr, (long long)row_pitch, (long long)img_pitch);
}

static int sgbm_compute_impl(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right,
                             int B, int H, int W, int cn, int16_t* disp, void* stream, const HostPipe* hp,
                             long long row_pitch, long long img_pitch)
{
    if (!h || !left || !right || !disp) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    if (B <= 0) { svb_set_error("batch must be >= 1"); return SVB_ERR_PARAM; }
    SgbmGeo g;
    int rc = make_geo(h->prm, H, W, cn, g);
    if (rc != SVB_OK) return rc;
    if (row_pitch || img_pitch) {                            // pitched input batch (bytes); 0 = dense
        const long long rp = row_pitch ? row_pitch : (long long)W * cn, ip = img_pitch ? img_pitch : rp * H;
        if (rp < (long long)W * cn || (B > 1 && ip < rp * (H - 1) + (long long)W * cn)) {
            svb_set_error("pitches too small: row_pitch %lld < W*cn or images overlap (img_pitch %lld)", rp, ip);
            return SVB_ERR_PARAM;
        }
        g.rowPitch = rp;
        g.imgPitch = ip;
    }
    if (!h->num_sms) {
        int dev = 0;
        SVB_CUDA(cudaGetDevice(&dev));
        SVB_CUDA(cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const int nb = pick_chunk(h, g, B);
    const SgbmLayout L = make_layout(g, nb);
    cudaStream_t st = (cudaStream_t)stream;
    if (!h->ev_last) {
        SVB_CUDA(cudaEventCreateWithFlags(&h->ev_last, cudaEventDisableTiming));
        SVB_CUDA(cudaMalloc((void**)&h->d_err, 256));
        SVB_CUDA(cudaMemset(h->d_err, 0, 256));
        SVB_CUDA(cudaHostAlloc((void**)&h->h_err, sizeof(int), cudaHostAllocDefault));
        *h->h_err = 0;
    } else {
        // the previous compute of this handle: report its flag if it has finished (no synchronisation here) ...
        if (h->pending && cudaEventQuery(h->ev_last) == cudaSuccess) {
            rc = take_sweep_err(h);
            if (rc != SVB_OK) return rc;
        }
        (void)cudaGetLastError();                            // cudaErrorNotReady is not an error
        // ... and keep this one behind it, whatever stream it arrives on (one workspace per handle)
        SVB_CUDA(cudaStreamWaitEvent(st, h->ev_last, 0));
    }
    if (L.total > h->ws.cap) {
        rc = h->ws.reserve(L.total);
        if (rc != SVB_OK) return rc;
        // right-descriptor planes have never-written halves (xr < HALF): keep them defined
        SVB_CUDA(cudaMemsetAsync(h->ws.p, 0, L.total, st));
    } else {
        SVB_CUDA(cudaMemsetAsync((char*)h->ws.p + L.sync, 0, 64, st));
    }
    h->launches = 0;
    for (int b0 = 0; b0 < B; b0 += nb) {
        const int n = std::min(nb, B - b0);
        const size_t off = (size_t)b0 * H * W;
        switch (g.NP) {
#define CASE(N) case N: rc = run_chunk<N>(h, g, L, left + b0 * g.imgPitch, right + b0 * g.imgPitch, n, disp + off, st, hp, b0); break;
            CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
            default: svb_set_error("bad NP"); rc = SVB_ERR_UNSUPPORTED;
        }
        if (rc != SVB_OK) return rc;
    }
    h->last_geo = g; h->off_C = L.C; h->off_S = L.S; h->off_raw = L.raw; h->have_last = true;
    SVB_CUDA(cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
    SVB_CUDA(cudaEventRecord(h->ev_last, st));
    h->pending = true;
    return SVB_OK;
}

// Wait for the handle's last compute and report a wavefront wait that hit its spin limit (cannot happen
// unless the device was reset underneath us or two processes share the workspace; the results of the
// computes since the last check are invalid).  This is what a *_dev caller polls at its own sync points;
// the *_host entry points call it before they return.
extern "C" int svb_sgbm_check(svb_sgbm_t* h)
{
    if (!h) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    if (!h->ev_last || !h->pending) return SVB_OK;
    SVB_CUDA(cudaEventSynchronize(h->ev_last));
    return take_sweep_err(h);
}
static int check_sweep_err(svb_sgbm* h) { return svb_sgbm_check(h); }
// Test hook: raise the flag as a timed-out wait would, so that the reporting path can be exercised.
extern "C" int svb_sgbm_debug_force_wave(svb_sgbm_t* h, int mode)
{
    if (!h) { svb_set_error("null handle"); return SVB_ERR_PARAM; }
    h->force_wave = mode < 0 ? -1 : (mode != 0);
    return SVB_OK;
}

extern "C" int svb_sgbm_debug_trip_error(svb_sgbm_t* h)
{
    if (!h || !h->d_err) { svb_set_error("no previous compute"); return SVB_ERR_PARAM; }
    const int one = 1;
    SVB_CUDA(cudaMemcpy(h->d_err, &one, sizeof(int), cudaMemcpyHostToDevice));
    return SVB_OK;
}

extern "C" int svb_sgbm_compute_host(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right,
                                     int B, int H, int W, int16_t* disp)
{
    return svb_sgbm_compute_cn_host(h, left, right, B, H, W, 1, disp);
}

extern "C" int svb_sgbm_compute_cn_host(svb_sgbm_t* h, const uint8_t* left, const uint8_t* right,
                                        int B, int H, int W, int cn, int16_t* disp)
{
    if (!h || !left || !right || !disp) { svb_set_error("null argument"); return SVB_ERR_PARAM; }
    if (B <= 0 || H <= 0 || W <= 0 || (cn != 1 && cn != 3)) { svb_set_error("bad shape"); return SVB_ERR_PARAM; }
    const size_t n = (size_t)B * H * W, nin = n * cn;
    int rc = h->stage_in.reserve(2 * nin);
    if (rc != SVB_OK) return rc;
    rc = h->stage_out.reserve(2 * n);
    if (rc != SVB_OK) return rc;
    uint8_t* dl = (uint8_t*)h->stage_in.p;
    uint8_t* dr = dl + nin;
    int16_t* dout = (int16_t*)h->stage_out.p;
    if (!h->s_in) {
        SVB_CUDA(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
        SVB_CUDA(cudaStreamCreateWithFlags(&h->s_run, cudaStreamNonBlocking));
        SVB_CUDA(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
    }
    // SVB_HOST_PIPE sub-ranges per chunk: only the first upload and the last download are exposed
    const size_t need = (size_t)SVB_HOST_PIPE;
    while (h->ev_in.size() < need) {
        cudaEvent_t e1, e2;
        SVB_CUDA(cudaEventCreateWithFlags(&e1, cudaEventDisableTiming));
        SVB_CUDA(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
        h->ev_in.push_back(e1); h->ev_done.push_back(e2);
    }
    HostPipe hp;
    hp.hl = left; hp.hr = right; hp.hout = disp;
    hp.pin = (size_t)H * W * cn; hp.pout = (size_t)H * W;
    hp.nsub = SVB_HOST_PIPE;
    // Pageable caller buffers (the reference's call site: NumPy arrays, depth_map.py:909) of a batch: the host pool
    // copies them image by image into pinned staging while the first sub-ranges already upload and compute, and copies
    // each downloaded sub-range out while the later ones are still on the device.
    const bool staged = B > 1 && 2 * nin + 2 * n >= ((size_t)8 << 20) && svb_host_pageable(left) && svb_host_pageable(right) &&
                        svb_host_pageable(disp);
    if (!staged) {
        rc = sgbm_compute_impl(h, dl, dr, B, H, W, cn, dout, h->s_run, &hp);
        if (rc != SVB_OK) { cudaDeviceSynchronize(); return rc; }
        SVB_CUDA(cudaStreamSynchronize(h->s_out));
        SVB_CUDA(cudaStreamSynchronize(h->s_run));
        return check_sweep_err(h);
    }
    rc = h->pin_in.reserve(2 * nin);
    if (rc != SVB_OK) return rc;
    rc = h->pin_out.reserve(2 * n);
    if (rc != SVB_OK) return rc;
    uint8_t* pl = (uint8_t*)h->pin_in.p;
    uint8_t* pr = pl + nin;
    int16_t* po = (int16_t*)h->pin_out.p;
    std::vector<std::atomic<int>> flags(B);
    for (auto& f : flags) f.store(0, std::memory_order_relaxed);
    std::vector<HostPipe::Mark> marks;
    SvbTaskGroup gin, gout;
    struct Waiter { SvbTaskGroup *a, *b; ~Waiter() { svb_pool_wait(a); svb_pool_wait(b); } } waiter{&gin, &gout};   // on every return path
    const size_t pin = hp.pin, pout = hp.pout;
    for (int i = 0; i < B; ++i)
        svb_pool_submit(&gin, [=, &flags] {
            memcpy(pl + i * pin, left + i * pin, pin);
            memcpy(pr + i * pin, right + i * pin, pin);
            flags[i].store(1, std::memory_order_release);
        });
    hp.hl = pl; hp.hr = pr; hp.hout = po;
    hp.staged = flags.data();
    hp.marks = &marks;
    rc = sgbm_compute_impl(h, dl, dr, B, H, W, cn, dout, h->s_run, &hp);
    svb_pool_wait(&gin);                                     // (the tasks read the caller's arrays and `flags`)
    if (rc != SVB_OK) { cudaDeviceSynchronize(); return rc; }
    const int parts = svb_pool_threads();
    for (const auto& m : marks) {
        if (cudaEventSynchronize(m.ev) != cudaSuccess) {     // (tasks in flight write the caller's array: wait first)
            svb_pool_wait(&gout);
            SVB_CUDA(cudaGetLastError());
            svb_set_error("download of a disparity sub-range failed");
            return SVB_ERR_CUDA;
        }
        const size_t bytes = (size_t)m.count * pout * 2, per = (bytes / parts + 4095) & ~(size_t)4095;
        const char* src = (const char*)(po + (size_t)m.first * pout);
        char* dst = (char*)(disp + (size_t)m.first * pout);
        for (size_t o = 0; o < bytes; o += per)
            svb_pool_submit(&gout, [=] { memcpy(dst + o, src + o, std::min(per, bytes - o)); });
    }
    svb_pool_wait(&gout);
    SVB_CUDA(cudaStreamSynchronize(h->s_out));
    SVB_CUDA(cudaStreamSynchronize(h->s_run));
    return check_sweep_err(h);
}

// Unpack one pixel's vector (HALF words, word w = d | (d+HALF) << 16) into D u16 values.
static void unpack_vec(const uint32_t* src, int HALF, int D, uint16_t* dst)
{
    for (int d = 0; d < D; ++d) {
        const uint32_t w = src[d < HALF ? d : d - HALF];
        dst[d] = (uint16_t)(d < HALF ? (w & 0xffffu) : (w >> 16));
    }
}

extern "C" int svb_sgbm_debug_stages(svb_sgbm_t* h, uint16_t* C_host, uint16_t* S_host, int16_t* raw_host)
{
    if (!h || !h->have_last) { svb_set_error("no previous compute"); return SVB_ERR_PARAM; }
    if (S_host && !h->debug) { svb_set_error("S is only kept after svb_sgbm_set_debug(h, 1)"); return SVB_ERR_PARAM; }
    const SgbmGeo& g = h->last_geo;
    SVB_CUDA(cudaDeviceSynchronize());
    const size_t roww = (size_t)g.width1 * g.HALF;           // u32 words per volume row
    uint32_t* tmp = (uint32_t*)malloc(roww * 4);
    if (!tmp) return SVB_ERR_NOMEM;
    for (int which = 0; which < 2; ++which) {
        uint16_t* dst = which ? S_host : C_host;
        if (!dst) continue;
        const char* src = (const char*)h->ws.p + (which ? h->off_S : h->off_C);
        for (int n = 0; n < g.nstripes; ++n) {
            int ys, ye, keep; job_rows(g, n, ys, ye, keep);
            for (int y = std::max(ys, keep); y < ye; ++y) {
                cudaError_t e = cudaMemcpy(tmp, src + (((size_t)n * g.rowsMax + (y - ys)) * roww) * 4, roww * 4,
                                           cudaMemcpyDeviceToHost);
                if (e != cudaSuccess) { free(tmp); svb_set_error("memcpy: %s", cudaGetErrorString(e)); return SVB_ERR_CUDA; }
                for (int x = 0; x < g.width1; ++x)
                    unpack_vec(tmp + (size_t)x * g.HALF, g.HALF, g.D, dst + ((size_t)y * g.width1 + x) * g.D);
            }
        }
    }
    free(tmp);
    if (raw_host)
        SVB_CUDA(cudaMemcpy(raw_host, (const char*)h->ws.p + h->off_raw, (size_t)g.H * g.W * 2, cudaMemcpyDeviceToHost));
    return check_sweep_err(h);
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
