/*
 * oracle/bm_oracle.c -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Plain-C CPU restatement of cv2.StereoBM_create(numDisparities, blockSize).compute
 * (BASELINE.json config 2; the reference itself never calls StereoBM -- SURVEY 0).
 * The arithmetic lives in OpenCV 4.13.0 (modules/calib3d/src/stereobm.cpp, not vendored);
 * this restates SURVEY.md Appendix A.6.  Pinned against live cv2 and tests/golden/.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int minDisparity, numDisparities, blockSize, preFilterCap, textureThreshold;
    int uniquenessRatio, speckleWindowSize, speckleRange, disp12MaxDiff;
    int preFilterType, preFilterSize;   /* cv2.STEREO_BM_PREFILTER_NORMALIZED_RESPONSE = 0, _XSOBEL = 1 (default) */
} orc_bm_params;

void orc_filter_speckles_s16(int16_t* img, int H, int W, int newVal, int maxSpeckleSize, int maxDiff);

static inline int bclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* A.6 prefilter (X-Sobel): rows are produced in pairs; the vertical neighbours of the
 * first/last row are reflected (row -1 -> 1, row H -> H-2); columns 0 and W-1 = cap;
 * for odd H the last row is all cap. */
void orc_bm_prefilter_xsobel(const uint8_t* src, int H, int W, int cap, uint8_t* dst)
{
    for (int y = 0; y < H; ++y) {
        uint8_t* d = dst + (size_t)y * W;
        if ((H & 1) && y == H - 1) { memset(d, cap, W); continue; }
        int ya = y > 0 ? y - 1 : (H > 1 ? 1 : 0);
        int yb = y < H - 1 ? y + 1 : (H > 1 ? H - 2 : 0);
        const uint8_t *r0 = src + (size_t)ya * W, *r1 = src + (size_t)y * W, *r2 = src + (size_t)yb * W;
        d[0] = (uint8_t)cap;
        if (W > 1) d[W - 1] = (uint8_t)cap;
        for (int x = 1; x < W - 1; ++x) {
            int g = (r0[x + 1] - r0[x - 1]) + 2 * (r1[x + 1] - r1[x - 1]) + (r2[x + 1] - r2[x - 1]);
            d[x] = (uint8_t)(bclamp(g, -cap, cap) + cap);
        }
    }
}

/* stereobm.cpp prefilterNorm (PREFILTER_NORMALIZED_RESPONSE): centre-weighted cross minus the mean of a
 * preFilterSize^2 box (replicate borders in both), in 10-bit fixed point, clipped to [0, 2*cap]. */
void orc_bm_prefilter_norm(const uint8_t* src, int H, int W, int winsize, int cap, uint8_t* dst)
{
    const int w2 = winsize / 2;
    int scale_g = winsize * winsize / 8;
    const int scale_s = (1024 + scale_g) / (scale_g * 2);
    scale_g *= scale_s;
    for (int y = 0; y < H; ++y) {
        const uint8_t* prev = src + (size_t)(y > 0 ? y - 1 : 0) * W;
        const uint8_t* curr = src + (size_t)y * W;
        const uint8_t* next = src + (size_t)(y < H - 1 ? y + 1 : H - 1) * W;
        for (int x = 0; x < W; ++x) {
            int sum = 0;
            for (int j = -w2; j <= w2; ++j) {
                const uint8_t* row = src + (size_t)bclamp(y + j, 0, H - 1) * W;
                for (int i = -w2; i <= w2; ++i) sum += row[bclamp(x + i, 0, W - 1)];
            }
            const int l = curr[x > 0 ? x - 1 : x], r = curr[x < W - 1 ? x + 1 : x];
            const int val = ((curr[x] * 4 + l + r + prev[x] + next[x]) * scale_g - sum * scale_s) >> 10;
            dst[(size_t)y * W + x] = (uint8_t)(val < -cap ? 0 : val > cap ? 2 * cap : val + cap);
        }
    }
}

/* cv2.validateDisparity (calib3d; StereoBM runs it when disp12MaxDiff >= 0): right-view map by scatter-min
 * of the cost at x - round(d/16) (first x wins ties), then a pixel is dropped iff BOTH roundings of its
 * disparity land on right-view entries that disagree by more than disp12MaxDiff*16.  Pinned bit-exact
 * against the public cv2.validateDisparity. */
void orc_validate_disparity(int16_t* disp, const int* cost, int H, int W, int minD, int D, int disp12MaxDiff)
{
    const int FILTERED = (minD - 1) * 16;
    const int maxD = minD + D, minX1 = maxD > 0 ? maxD : 0, maxX1 = W + (minD < 0 ? minD : 0);
    const int d12 = disp12MaxDiff * 16;
    int* d2 = (int*)malloc(sizeof(int) * 2 * (size_t)W);
    int* c2 = d2 + W;
    for (int y = 0; y < H; ++y) {
        int16_t* dp = disp + (size_t)y * W;
        for (int x = 0; x < W; ++x) { d2[x] = FILTERED; c2[x] = 0x7fffffff; }
        for (int x = minX1; x < maxX1; ++x) {
            const int d = dp[x], c = cost[(size_t)y * W + x];
            if (d == FILTERED) continue;
            const int x2 = x - ((d + 8) >> 4);
            if (x2 < 0 || x2 >= W) continue;
            if (c2[x2] > c) { c2[x2] = c; d2[x2] = d; }
        }
        for (int x = minX1; x < maxX1; ++x) {
            const int d = dp[x];
            if (d == FILTERED) continue;
            const int x0 = x - (d >> 4), x1 = x - ((d + 15) >> 4);
            if (0 <= x0 && x0 < W && d2[x0] > FILTERED && abs(d2[x0] - d) > d12 &&
                0 <= x1 && x1 < W && d2[x1] > FILTERED && abs(d2[x1] - d) > d12)
                dp[x] = (int16_t)FILTERED;
        }
    }
    free(d2);
}

/* returns 0, or <0 for what cv2 rejects */
int orc_bm_compute(const uint8_t* left, const uint8_t* right, int H, int W, const orc_bm_params* p, int16_t* disp)
{
    const int D = p->numDisparities, minD = p->minDisparity, bs = p->blockSize, r = bs / 2;
    const int cap = p->preFilterCap;
    if (D <= 0 || D % 16 != 0) return -1;
    if (bs < 5 || bs > 255 || bs % 2 == 0 || bs >= W || bs >= H) return -1;
    if (cap < 1 || cap > 63) return -1;
    if (p->preFilterType != 0 && p->preFilterType != 1) return -1;
    if (p->preFilterType == 0 && (p->preFilterSize < 5 || p->preFilterSize > 255 || p->preFilterSize % 2 == 0)) return -1;
    if (p->textureThreshold < 0 || p->uniquenessRatio < 0) return -1;
    const int FILTERED = (minD - 1) * 16;
    for (size_t i = 0; i < (size_t)H * W; ++i) disp[i] = (int16_t)FILTERED;
    const int lofs = D - 1 + minD > 0 ? D - 1 + minD : 0;
    const int rofs = D - 1 + minD < 0 ? -(D - 1 + minD) : 0;
    const int width1 = W - rofs - D + 1;
    if (lofs >= W || rofs >= W || width1 < 1) return 0;

    uint8_t* L = (uint8_t*)malloc((size_t)H * W);
    uint8_t* R = (uint8_t*)malloc((size_t)H * W);
    if (p->preFilterType == 0) {
        orc_bm_prefilter_norm(left, H, W, p->preFilterSize, cap, L);
        orc_bm_prefilter_norm(right, H, W, p->preFilterSize, cap, R);
    } else {
        orc_bm_prefilter_xsobel(left, H, W, cap, L);
        orc_bm_prefilter_xsobel(right, H, W, cap, R);
    }
    int* sad = (int*)malloc(sizeof(int) * (D + 2));
    int* cost = (int*)calloc((size_t)H * W, sizeof(int));
    const int x_lo = lofs + r, x_hi = W - r;              /* valid ROI columns */
    for (int y = r; y < H - r; ++y)
        for (int xp = 0; xp < width1; ++xp) {
            const int x = xp + lofs;
            /* cv2 evaluates every column of the width1 domain (windows clamp at the image sides); the
             * columns outside the valid ROI only matter as candidates of validateDisparity below and
             * are blanked afterwards */
            /* minDisparity > 0: the last minDisparity positions of the width1 domain lie past the row end
             * (x >= W).  cv2 computes them all the same (windows clamp to the image) and its store lands in
             * the first minDisparity columns of the NEXT row.  Inside the ROI rows the column blanking wipes
             * that again; the spill of the last ROI row (y = H-1-r) into row H-r survives -- that row was
             * blanked before the matcher ran (stereobm.cpp FindStereoCorrespInvoker).  Probed on cv2 4.13
             * (single stripe; with several stripes cv2 races with itself at the stripe seams). */
            const int spill = x >= W;
            if (spill && y != H - 1 - r) continue;
            if (!spill && (x < x_lo || x >= x_hi) && p->disp12MaxDiff < 0) continue;
            int16_t* out = spill ? disp + (size_t)(y + 1) * W + (x - W) : disp + (size_t)y * W + x;
            int ts = 0;
            for (int k = 0; k < D; ++k) sad[k + 1] = 0;
            for (int j = -r; j <= r; ++j) {
                const int lc = bclamp(xp + j + lofs, 0, W - 1);
                const int rs = bclamp(xp + j + rofs, 0, W - D);
                for (int i = -r; i <= r; ++i) {
                    const int yy = bclamp(y + i, 0, H - 1);
                    const int lv = L[(size_t)yy * W + lc];
                    ts += abs(lv - cap);
                    const uint8_t* rr = R + (size_t)yy * W + rs;
                    for (int k = 0; k < D; ++k) sad[k + 1] += abs(lv - rr[k]);
                }
            }
            if (ts < p->textureThreshold) continue;
            int mind = 0, minsad = sad[1];
            for (int k = 1; k < D; ++k) if (sad[k + 1] < minsad) { minsad = sad[k + 1]; mind = k; }
            if (p->uniquenessRatio > 0) {
                const int thresh = minsad + (minsad * p->uniquenessRatio / 100);
                int k;
                for (k = 0; k < D; ++k)
                    if ((k < mind - 1 || k > mind + 1) && sad[k + 1] <= thresh) break;
                if (k < D) continue;
            }
            sad[0] = sad[2]; sad[D + 1] = sad[D - 1];
            if (!spill) cost[(size_t)y * W + x] = sad[mind + 1];
            const int pp = sad[mind + 2], n = sad[mind], dd = pp + n - 2 * sad[mind + 1] + abs(pp - n);
            *out = (int16_t)((((D - mind - 1 + minD) * 256 + (dd != 0 ? (pp - n) * 256 / dd : 0) + 15) >> 4));
        }
    /* stereobm.cpp `if (disp12MaxDiff >= 0) validateDisparity(...)` with the winning SAD as cost; the
     * columns outside the valid ROI took part as candidates (probed, 40/40 random cv2 cases) and are blanked
     * afterwards. */
    if (p->disp12MaxDiff >= 0) {
        orc_validate_disparity(disp, cost, H, W, minD, D, p->disp12MaxDiff);
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
                if ((x < x_lo || x >= x_hi) && !(y == H - r && x < minD)) disp[(size_t)y * W + x] = (int16_t)FILTERED;
    }
    free(L); free(R); free(sad); free(cost);
    if (p->speckleWindowSize > 0)
        orc_filter_speckles_s16(disp, H, W, FILTERED, p->speckleWindowSize, p->speckleRange);
    return 0;
}
