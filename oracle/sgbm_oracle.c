/*
 * oracle/sgbm_oracle.c -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Plain-C CPU restatement of the semi-global block matcher that the reference
 * drives through cv2.StereoSGBM_create(...).compute(left, right)
 *   (reference call sites: depth_map.py:894-909, fused_depth_map.py:988-1004).
 * The arithmetic itself lives in the un-vendored third-party dependency
 * OpenCV (opencv-python-headless 4.13.0.92, modules/calib3d/src/stereosgbm.cpp);
 * this file restates its published algorithm as distilled in SURVEY.md
 * Appendix A.1-A.5.  Parity is pinned by the golden vectors under tests/golden/
 * that cv2 4.13.0 produced (tests/golden/make_golden.py -> tests/test_oracle_golden.py),
 * by the SHA-256 of cv2's output at the BASELINE sizes (make_golden_fullsize.py ->
 * tests/test_fullsize_oracle.py: 1080p / D256 / HH, 720p / D128, 480p / D320 / 3WAY), and
 * by live-cv2 fuzz runs in the authoring container (counts in DESIGN.md section 2).
 *
 * Written for clarity, not speed: the whole cost volume C and the aggregated
 * volume S are materialised so that the stages can be handed to the tests.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_MODE_SGBM 0
#define ORC_MODE_HH 1
#define ORC_MODE_SGBM_3WAY 2
#define ORC_MODE_HH4 3

typedef struct {
    int minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff;
    int preFilterCap, uniquenessRatio, speckleWindowSize, speckleRange, mode;
} orc_sgbm_params;

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* x-Sobel prefiltered channel (p0) and raw channel (p1) of
 * one image row.  Columns 0 and W-1 of BOTH channels are ftzero. */
static void prefilter_row(const uint8_t* img, int H, int W, int cn, int ch, int y, int ftzero,
                          int* p0, int* p1)
{
    int yn = y > 0 ? y - 1 : y, ys = y < H - 1 ? y + 1 : y;
    const uint8_t *r = img + (size_t)y * W * cn + ch, *rn = img + (size_t)yn * W * cn + ch,
                  *rs = img + (size_t)ys * W * cn + ch;
    for (int x = 1; x < W - 1; ++x) {
        int e = (x + 1) * cn, w = (x - 1) * cn;
        int g = 2 * (r[e] - r[w]) + (rn[e] - rn[w]) + (rs[e] - rs[w]);
        p0[x] = iclamp(g, -ftzero, ftzero) + ftzero;
        p1[x] = r[x * cn];
    }
    p0[0] = p1[0] = ftzero;
    if (W > 1) p0[W - 1] = p1[W - 1] = ftzero;
}

/* Birchfield-Tomasi cost of one row into pix[xc*D + d]
 * (accumulated over the two channels; channel 1 is >>2). */
static void pixel_cost_row(const uint8_t* L, const uint8_t* R, int H, int W, int cn, int y,
                           int minD, int D, int ftzero, int minX1, int maxX1,
                           int* scratch /* 10*W ints */, int* pix)
{
    int *l0 = scratch, *l1 = l0 + W, *r0 = l1 + W, *r1 = r0 + W;
    int *v0 = r1 + W, *v1 = v0 + W, *u0 = v1 + W, *u1 = u0 + W;
    int width1 = maxX1 - minX1;
    memset(pix, 0, sizeof(int) * (size_t)width1 * D);
    /* 3-channel input (cv2 calcPixelCostBT, cn = 3): every colour channel contributes its own
     * gradient channel (unscaled) and raw channel (>>2); the costs add. */
    for (int ch = 0; ch < cn; ++ch) {
    prefilter_row(L, H, W, cn, ch, y, ftzero, l0, l1);
    prefilter_row(R, H, W, cn, ch, y, ftzero, r0, r1);
    for (int c = 0; c < 2; ++c) {
        const int* pl = c ? l1 : l0;
        const int* pr = c ? r1 : r0;
        int shift = c ? 2 : 0;
        for (int x = 0; x < W; ++x) {
            int v = pr[x];
            int va = x < W - 1 ? (v + pr[x + 1]) / 2 : v;
            int vb = x > 0 ? (v + pr[x - 1]) / 2 : v;
            v0[x] = imin(imin(va, vb), v);
            v1[x] = imax(imax(va, vb), v);
            int u = pl[x];
            int ua = x > 0 ? (u + pl[x - 1]) / 2 : u;
            int ub = x < W - 1 ? (u + pl[x + 1]) / 2 : u;
            u0[x] = imin(imin(ua, ub), u);
            u1[x] = imax(imax(ua, ub), u);
        }
        for (int x = minX1; x < maxX1; ++x) {
            int u = pl[x];
            int* out = pix + (size_t)(x - minX1) * D;
            for (int d = minD; d < minD + D; ++d) {
                int xr = x - d;
                int v = pr[xr];
                int c0 = imax(0, imax(u - v1[xr], v0[xr] - u));
                int c1 = imax(0, imax(v - u1[x], u0[x] - v));
                out[d - minD] += imin(c0, c1) >> shift;
            }
        }
    }
    }
}

/* C[y,xc,d] = box(blockSize) over pix, replicate borders in the cost
 * domain; rows clamp to [ylo, H-1].  (cv2 stores C+P2 and subtracts minLr+P2 in
 * the recurrence; the two cancel, so plain C is kept here -- probed: SURVEY A.3's
 * "+P2 per path" has the sign backwards, cv2's S is the textbook sum.) */
static void box_cost_volume(const uint8_t* L, const uint8_t* R, int H, int W, int cn,
                            int minD, int D, int ftzero, int bs, int P2,
                            int minX1, int maxX1, int ylo, int y0, int y1, int hh4,
                            int16_t* C /* rows y0..y1-1 stored at (y-y0) */)
{
    int width1 = maxX1 - minX1, r = bs / 2;
    size_t rowsz = (size_t)width1 * D;
    int nrows = H;
    int* scratch = (int*)malloc(sizeof(int) * 10 * (size_t)W);
    int* pix = (int*)malloc(sizeof(int) * rowsz);
    int32_t* hs = (int32_t*)calloc((size_t)nrows * rowsz, sizeof(int32_t));
    /* horizontal sums for every row that any window can touch */
    int ya = imax(ylo, y0 - r), yb = imin(H - 1, y1 - 1 + r);
    for (int y = ya; y <= yb; ++y) {
        pixel_cost_row(L, R, H, W, cn, y, minD, D, ftzero, minX1, maxX1, scratch, pix);
        int32_t* h = hs + (size_t)y * rowsz;
        for (int xc = 0; xc < width1; ++xc)
            for (int k = -r; k <= r; ++k) {
                const int* p = pix + (size_t)iclamp(xc + k, 0, width1 - 1) * D;
                for (int d = 0; d < D; ++d) h[(size_t)xc * D + d] += p[d];
            }
    }
    for (int y = y0; y < y1; ++y) {
        int16_t* c = C + (size_t)(y - y0) * rowsz;
        for (size_t i = 0; i < rowsz; ++i) {
            int s = 0;
            for (int k = -r; k <= r; ++k) s += hs[(size_t)iclamp(y + k, ylo, H - 1) * rowsz + i];
            /* MODE_HH4 (probed on cv2 4.13, 60/60 random cases bit-exact): rows whose box window
             * reaches below the image (y + r > H-1, y > 0) carry cost 0 for every disparity --
             * cv2's HH4 cost pass only updates C while the incoming row k = y + r exists. */
            if (hh4 && y + r > H - 1 && y > 0) s = 0;
            c[i] = (int16_t)s;
        }
    }
    free(hs); free(pix); free(scratch);
}

/* one path direction (dx,dy) over rows [y0,y1), accumulated into S
 * with int16 saturation.  Lr outside the domain is 0 with minLr 0; the d=-1 and
 * d=D neighbours are MAXC. */
static void aggregate_path(const int16_t* C, uint16_t* S, int y0, int y1, int width1,
                           int D, int P1, int P2, int dx, int dy)
{
    const int MAXC = 32767;
    size_t rowsz = (size_t)width1 * D;
    int* Lprev = (int*)calloc((size_t)(width1 + 2) * D, sizeof(int));
    int* Lcur = (int*)calloc((size_t)(width1 + 2) * D, sizeof(int));
    int* mprev = (int*)calloc(width1 + 2, sizeof(int));
    int* mcur = (int*)calloc(width1 + 2, sizeof(int));
    int forward_y = dy >= 0;
    int xfwd = dy != 0 ? 1 : (dx > 0);
    for (int yi = 0; yi < y1 - y0; ++yi) {
        int y = forward_y ? y0 + yi : y1 - 1 - yi;
        memset(Lcur, 0, sizeof(int) * (size_t)(width1 + 2) * D);
        memset(mcur, 0, sizeof(int) * (width1 + 2));
        for (int xi = 0; xi < width1; ++xi) {
            int xc = xfwd ? xi : width1 - 1 - xi;
            /* predecessor q = p - r, slot index = xc+1 (slots 0 and width1+1 are "outside") */
            const int* Lq; int mq;
            if (dy == 0) { Lq = Lcur + (size_t)(xc - dx + 1) * D; mq = mcur[xc - dx + 1]; }
            else         { Lq = Lprev + (size_t)(xc - dx + 1) * D; mq = mprev[xc - dx + 1]; }
            int* Lp = Lcur + (size_t)(xc + 1) * D;
            const int16_t* c = C + (size_t)(y - y0) * rowsz + (size_t)xc * D;
            uint16_t* s = S + (size_t)(y - y0) * rowsz + (size_t)xc * D;
            int m = MAXC;
            for (int d = 0; d < D; ++d) {
                int a = Lq[d];
                int b = (d > 0 ? Lq[d - 1] : MAXC) + P1;
                int e = (d < D - 1 ? Lq[d + 1] : MAXC) + P1;
                int L = c[d] + imin(imin(a, b), imin(e, mq + P2)) - mq;
                Lp[d] = L;
                m = imin(m, L);
                s[d] = (uint16_t)imin(s[d] + L, MAXC);
            }
            mcur[xc + 1] = m;
        }
        /* "outside" slots stay zero */
        memset(Lcur, 0, sizeof(int) * D); memset(Lcur + (size_t)(width1 + 1) * D, 0, sizeof(int) * D);
        mcur[0] = mcur[width1 + 1] = 0;
        int* t = Lprev; Lprev = Lcur; Lcur = t;
        t = mprev; mprev = mcur; mcur = t;
    }
    free(Lprev); free(Lcur); free(mprev); free(mcur);
}

/* WTA + uniqueness + subpixel + right-view map + LR check, one row. */
static void wta_row(const uint16_t* Srow, int W, int width1, int D, int minD, int minX1,
                    int uniq, int d12, int three_way, int16_t* disp_row, int* disp2, int* disp2cost)
{
    const int MAXC = 32767, INV = (minD - 1) * 16;
    for (int x = 0; x < W; ++x) { disp_row[x] = (int16_t)INV; disp2[x] = INV; disp2cost[x] = MAXC; }
    for (int xc = width1 - 1; xc >= 0; --xc) {
        const uint16_t* s = Srow + (size_t)xc * D;
        /* cv2 starts from minS=MAXC, best=-1 with a strict "<": a fully saturated
         * column keeps best=-1 and comes out as (minD-1)*16 = INV (probed). */
        int best = -1, minS = MAXC;
        for (int d = 0; d < D; ++d) if (s[d] < minS) { minS = s[d]; best = d; }
        if (best < 0) continue;
        if (three_way) {
            /* cv2's 3WAY argmin runs 8 int16 lanes over d (lane = d mod 8): within a lane a later d with an
             * EQUAL cost replaces the earlier one, across lanes the lowest recorded d wins.  Only visible with
             * uniquenessRatio = 0 (any tie further than 1 apart is rejected otherwise).  Probed on cv2 4.13. */
            int top[8];
            for (int j = 0; j < 8; ++j) top[j] = -1;
            for (int d = 0; d < D; ++d) if (s[d] == minS) top[d & 7] = d;
            best = D;
            for (int j = 0; j < 8; ++j) if (top[j] >= 0 && top[j] < best) best = top[j];
        }
        int reject = 0;
        if (three_way) {
            /* cv2's 3WAY vector loop tests Sp[d] < (short)(thresh+1), thresh = 100*minS/(100-u)
             * (integer division), i.e. "<=" where the other modes have "<"; probed on cv2 4.13. */
            if (uniq > 0) {
                int16_t th = (int16_t)((100 * minS) / (100 - uniq) + 1);
                for (int d = 0; d < D && !reject; ++d)
                    if (abs(d - best) > 1 && (int16_t)s[d] < th) reject = 1;
            }
        } else {
            for (int d = 0; d < D && !reject; ++d)
                if (abs(d - best) > 1 && (int)s[d] * (100 - uniq) < minS * 100) reject = 1;
        }
        if (reject) continue;
        int x2 = xc + minX1 - best - minD;
        if (disp2cost[x2] > minS) { disp2cost[x2] = minS; disp2[x2] = best + minD; }
        int dsub;
        if (best > 0 && best < D - 1) {
            int den = imax((int)s[best - 1] + s[best + 1] - 2 * (int)s[best], 1);
            dsub = best * 16 + (((int)s[best - 1] - (int)s[best + 1]) * 16 + den) / (2 * den);
        } else dsub = best * 16;
        disp_row[xc + minX1] = (int16_t)(dsub + minD * 16);
    }
    for (int x = minX1; x < minX1 + width1; ++x) {
        int d1 = disp_row[x];
        if (d1 == INV) continue;
        int dl = d1 >> 4, dh = (d1 + 15) >> 4;
        int xl = x - dl, xh = x - dh;
        int f1 = xl >= 0 && xl < W && disp2[xl] >= minD && abs(disp2[xl] - dl) > d12;
        int f2 = xh >= 0 && xh < W && disp2[xh] >= minD && abs(disp2[xh] - dh) > d12;
        if (f1 && f2) disp_row[x] = (int16_t)INV;
    }
}

void orc_median3_s16(const int16_t* src, int H, int W, int16_t* dst);
void orc_filter_speckles_s16(int16_t* img, int H, int W, int newVal, int maxSpeckleSize, int maxDiff);

/* Returns 0 on success, <0 on parameter errors (the checks cv2 4.13 makes).
 * Optional stage outputs (may be NULL): C_out/S_out [H*width1*D] (3WAY: kept rows of each stripe),
 * raw_out [H*W] = disparity before median3/speckle. */
int orc_sgbm_compute_cn(const uint8_t* left, const uint8_t* right, int H, int W, int cn,
                        const orc_sgbm_params* p, int16_t* disp,
                        int16_t* C_out, uint16_t* S_out, int16_t* raw_out);
int orc_sgbm_compute(const uint8_t* left, const uint8_t* right, int H, int W,
                     const orc_sgbm_params* p, int16_t* disp,
                     int16_t* C_out, uint16_t* S_out, int16_t* raw_out)
{
    return orc_sgbm_compute_cn(left, right, H, W, 1, p, disp, C_out, S_out, raw_out);
}

/* left/right: u8 [H,W] (cn = 1) or interleaved [H,W,3] (cn = 3). */
int orc_sgbm_compute_cn(const uint8_t* left, const uint8_t* right, int H, int W, int cn,
                        const orc_sgbm_params* p, int16_t* disp,
                        int16_t* C_out, uint16_t* S_out, int16_t* raw_out)
{
    if (cn != 1 && cn != 3) return -4;
    int minD = p->minDisparity, D = p->numDisparities, bs = p->blockSize;
    if (D <= 0 || bs < 1) return -1;
    int maxD = minD + D;
    int minX1 = imax(maxD, 0), maxX1 = W + imin(minD, 0), width1 = maxX1 - minX1;
    int INV = (minD - 1) * 16;
    if (width1 <= bs / 2) return -2;                /* cv2 assert, stereosgbm.cpp:511 */
    int ftzero = imax(p->preFilterCap, 15) | 1;
    int P1 = p->P1 > 0 ? p->P1 : 2;
    int P2 = imax(p->P2 > 0 ? p->P2 : 5, P1 + 1);
    int uniq = p->uniquenessRatio >= 0 ? p->uniquenessRatio : 10;
    int d12 = p->disp12MaxDiff > 0 ? p->disp12MaxDiff : 1;
    size_t rowsz = (size_t)width1 * D;
    int16_t* raw = (int16_t*)malloc(sizeof(int16_t) * (size_t)H * W);
    int* disp2 = (int*)malloc(sizeof(int) * 2 * (size_t)W);

    if (p->mode == ORC_MODE_SGBM_3WAY) {
        /* 4 fixed stripes, each an independent top-down sweep. */
        int nstripes = 4, ssz = (H + nstripes - 1) / nstripes;
        int overlap = (bs / 2 + 1) + (int)((ssz + 9) / 10); /* ceil(0.1*ssz) */
        for (int n = 0; n < nstripes; ++n) {
            int ys = iclamp(n * ssz - overlap, 0, H), ye = imin((n + 1) * ssz, H);
            int keep = n * ssz;
            if (ye <= ys) continue;
            int16_t* C = (int16_t*)malloc(sizeof(int16_t) * (size_t)(ye - ys) * rowsz);
            uint16_t* S = (uint16_t*)calloc((size_t)(ye - ys) * rowsz, sizeof(uint16_t));
            box_cost_volume(left, right, H, W, cn, minD, D, ftzero, bs, P2, minX1, maxX1, ys, ys, ye, 0, C);
            aggregate_path(C, S, ys, ye, width1, D, P1, P2, 1, 0);
            aggregate_path(C, S, ys, ye, width1, D, P1, P2, 0, 1);
            aggregate_path(C, S, ys, ye, width1, D, P1, P2, -1, 0);
            for (int y = imax(keep, ys); y < ye; ++y) {
                if (C_out) memcpy(C_out + (size_t)y * rowsz, C + (size_t)(y - ys) * rowsz, sizeof(int16_t) * rowsz);
                if (S_out) memcpy(S_out + (size_t)y * rowsz, S + (size_t)(y - ys) * rowsz, sizeof(uint16_t) * rowsz);
            }
            for (int y = imax(keep, ys); y < ye; ++y)
                wta_row(S + (size_t)(y - ys) * rowsz, W, width1, D, minD, minX1, uniq, d12, 1,
                        raw + (size_t)y * W, disp2, disp2 + W);
            free(C); free(S);
        }
    } else {
        int16_t* C = (int16_t*)malloc(sizeof(int16_t) * (size_t)H * rowsz);
        uint16_t* S = (uint16_t*)calloc((size_t)H * rowsz, sizeof(uint16_t));
        box_cost_volume(left, right, H, W, cn, minD, D, ftzero, bs, P2, minX1, maxX1, 0, 0, H, p->mode == ORC_MODE_HH4, C);
        static const int dirs[8][2] = {{1,0},{1,1},{0,1},{-1,1},{-1,0},{-1,-1},{0,-1},{1,-1}};
        int npaths = p->mode == ORC_MODE_SGBM ? 5 : 8;
        /* MODE_HH4 (cv2 computeDisparitySGBM_HH4): the four axis-aligned paths only. */
        int step = p->mode == ORC_MODE_HH4 ? 2 : 1;
        for (int k = 0; k < npaths; k += step)
            aggregate_path(C, S, 0, H, width1, D, P1, P2, dirs[k][0], dirs[k][1]);
        for (int y = 0; y < H; ++y)
            wta_row(S + (size_t)y * rowsz, W, width1, D, minD, minX1, uniq, d12, 0,
                    raw + (size_t)y * W, disp2, disp2 + W);
        if (C_out) memcpy(C_out, C, sizeof(int16_t) * (size_t)H * rowsz);
        if (S_out) memcpy(S_out, S, sizeof(uint16_t) * (size_t)H * rowsz);
        free(C); free(S);
    }
    if (raw_out) memcpy(raw_out, raw, sizeof(int16_t) * (size_t)H * W);
    orc_median3_s16(raw, H, W, disp);
    if (p->speckleWindowSize > 0)
        orc_filter_speckles_s16(disp, H, W, INV, p->speckleWindowSize, 16 * p->speckleRange);
    free(raw); free(disp2);
    return 0;
}
