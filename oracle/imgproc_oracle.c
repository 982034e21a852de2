/*
 * oracle/imgproc_oracle.c -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Plain-C CPU restatements of the OpenCV imgproc/calib3d helpers on the
 * reference's per-frame path (third-party dependency opencv-python-headless
 * 4.13.0.92, not vendored under /root/reference):
 *   cv2.remap INTER_LINEAR, CV_16SC2+CV_16UC1 maps   depth_map.py:815-826, fused_depth_map.py:480-491
 *   cv2.cvtColor BGR2GRAY                            depth_map.py:871-879, fused_depth_map.py:979-984
 *   cv2.medianBlur 3 (s16, inside StereoSGBM)        depth_map.py:909
 *   cv2.medianBlur 5 (f32)                           depth_map.py:912, fused_depth_map.py:1007
 *   cv2.filterSpeckles (inside StereoSGBM)           depth_map.py:909
 *   cv2.reprojectImageTo3D                           named by BASELINE.json (Q from depth_map.py:625)
 * Algorithms follow SURVEY.md Appendix A.4, A.7, A.8.  Parity is pinned against
 * live cv2 and the cv2-made golden vectors in tests/golden/.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline int iclampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* the four s16 weights are (32-fy)(32-fx)*32 etc. -- exact products
 * that always sum to 32768, so the table correction of cv2 never changes them. */
void orc_remap_u8(const uint8_t* src, int Hs, int Ws, int C,
                  const int16_t* map1, const uint16_t* map2, int Ho, int Wo, uint8_t* dst)
{
    for (int y = 0; y < Ho; ++y)
        for (int x = 0; x < Wo; ++x) {
            int sx = map1[2 * ((size_t)y * Wo + x)], sy = map1[2 * ((size_t)y * Wo + x) + 1];
            int f = map2[(size_t)y * Wo + x] & 1023, fx = f & 31, fy = f >> 5;
            int w[4] = {(32 - fy) * (32 - fx) * 32, (32 - fy) * fx * 32, fy * (32 - fx) * 32, fy * fx * 32};
            for (int c = 0; c < C; ++c) {
                int acc = 0;
                for (int k = 0; k < 4; ++k) {
                    int xx = sx + (k & 1), yy = sy + (k >> 1);
                    if (xx >= 0 && xx < Ws && yy >= 0 && yy < Hs)
                        acc += w[k] * src[((size_t)yy * Ws + xx) * C + c];
                }
                dst[((size_t)y * Wo + x) * C + c] = (uint8_t)iclampi((acc + 16384) >> 15, 0, 255);
            }
        }
}

/* SURVEY A.8 */
void orc_bgr2gray(const uint8_t* bgr, size_t n, uint8_t* gray)
{
    for (size_t i = 0; i < n; ++i)
        gray[i] = (uint8_t)((3735 * bgr[3 * i] + 19235 * bgr[3 * i + 1] + 9798 * bgr[3 * i + 2] + 16384) >> 15);
}

static int cmp_s16(const void* a, const void* b) { return (int)*(const int16_t*)a - (int)*(const int16_t*)b; }
static int cmp_f32(const void* a, const void* b)
{
    float x = *(const float*)a, y = *(const float*)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

/* exact 3x3 median, replicate border; invalid values are ordinary numbers (A.4) */
void orc_median3_s16(const int16_t* src, int H, int W, int16_t* dst)
{
    int16_t v[9];
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            int n = 0;
            for (int j = -1; j <= 1; ++j)
                for (int i = -1; i <= 1; ++i)
                    v[n++] = src[(size_t)iclampi(y + j, 0, H - 1) * W + iclampi(x + i, 0, W - 1)];
            qsort(v, 9, sizeof(int16_t), cmp_s16);
            dst[(size_t)y * W + x] = v[4];
        }
}

/* exact 5x5 median on float32, replicate border */
void orc_median5_f32(const float* src, int H, int W, float* dst)
{
    float v[25];
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            int n = 0;
            for (int j = -2; j <= 2; ++j)
                for (int i = -2; i <= 2; ++i)
                    v[n++] = src[(size_t)iclampi(y + j, 0, H - 1) * W + iclampi(x + i, 0, W - 1)];
            qsort(v, 25, sizeof(float), cmp_f32);
            dst[(size_t)y * W + x] = v[12];
        }
}

/* 4-connected components under |a-b|<=maxDiff among
 * pixels != newVal; components of size <= maxSpeckleSize are set to newVal. */
void orc_filter_speckles_s16(int16_t* img, int H, int W, int newVal, int maxSpeckleSize, int maxDiff)
{
    size_t n = (size_t)H * W;
    int32_t* label = (int32_t*)calloc(n, sizeof(int32_t));
    int32_t* stack = (int32_t*)malloc(n * sizeof(int32_t));
    uint8_t* small = (uint8_t*)malloc(n + 1);
    int cur = 0;
    for (size_t s = 0; s < n; ++s) {
        if (img[s] == newVal) continue;
        if (label[s]) { if (small[label[s]]) img[s] = (int16_t)newVal; continue; }
        int top = 0, count = 0;
        stack[top++] = (int32_t)s; label[s] = ++cur;
        while (top) {
            int32_t p = stack[--top]; ++count;
            int py = p / W, px = p % W, v = img[p];
            const int nb[4][2] = {{0, 1}, {0, -1}, {1, 0}, {-1, 0}};
            for (int k = 0; k < 4; ++k) {
                int qy = py + nb[k][0], qx = px + nb[k][1];
                if (qy < 0 || qy >= H || qx < 0 || qx >= W) continue;
                int32_t q = qy * W + qx;
                if (label[q] || img[q] == newVal || abs(v - img[q]) > maxDiff) continue;
                label[q] = cur; stack[top++] = q;
            }
        }
        small[cur] = count <= maxSpeckleSize;
        if (small[cur]) img[s] = (int16_t)newVal;
    }
    free(label); free(stack); free(small);
}

/* cv2.reprojectImageTo3D for CV_32F or CV_16S disparity (no /16 for int16, A.8):
 * [X Y Z W]^T = Q [x y d 1]^T, out = (X/W, Y/W, Z/W) computed in double, stored f32. */
void orc_reproject3d(const float* disp, int H, int W, const double* Q /* 4x4 row-major */,
                     int handleMissing, float* out)
{
    double minv = 0;
    if (handleMissing) {
        minv = disp[0];
        for (size_t i = 1; i < (size_t)H * W; ++i) if (disp[i] < minv) minv = disp[i];
    }
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            double d = disp[(size_t)y * W + x];
            double X = Q[0] * x + Q[1] * y + Q[2] * d + Q[3];
            double Y = Q[4] * x + Q[5] * y + Q[6] * d + Q[7];
            double Z = Q[8] * x + Q[9] * y + Q[10] * d + Q[11];
            double Wd = Q[12] * x + Q[13] * y + Q[14] * d + Q[15];
            float* o = out + 3 * ((size_t)y * W + x);
            double iw = 1.0 / Wd;
            o[0] = (float)(X * iw); o[1] = (float)(Y * iw); o[2] = (float)(Z * iw);
            if (handleMissing && d <= minv) o[2] = 10000.f;
        }
}

/* ---------------------------------------------------------------------------------------
 * cv2.resize(src, dsize, interpolation=INTER_LINEAR) -- the reference resizes both camera frames on every
 * processed frame (fused_depth_map.py:473-476, PROCESSING_SCALE = 0.33; depth_map.py:776, 809-811) and a mono
 * map to the stereo shape (fused_depth_map.py:1214-1217, float32).  OpenCV 4.13 imgproc/src/resize.cpp restated;
 * pinned against live cv2 (300/300 random u8 cases bit-exact, up- and down-scale, 1 and 3 channels).
 *
 * Coordinates: f = (float)((d + 0.5) * scale - 0.5), scale = 1.0 / ((double)dsize / ssize); s = floor(f); f -= s.
 * Horizontally s is clamped to the image with the fraction reset to 0 (s < 0 -> s = 0, f = 0; s >= ssize-1 ->
 * s = ssize-1, f = 0).  VERTICALLY the fraction is NOT reset: the two tap rows are clipped into the image and keep
 * their split weights (that changes the rounding in the first / last output rows when upscaling -- probed).
 * u8: weights are round-half-even(w * 2048) as int16, the horizontal pass keeps 32-bit sums, the vertical pass is
 * (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2.
 * ------------------------------------------------------------------------------------- */
#include <math.h>
static void resize_axis(int ssize, int dsize, int clamp_fraction, int* ofs, float* w /* 2 per entry */)
{
    const double scale = 1.0 / ((double)dsize / ssize);
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (clamp_fraction) {
            if (s < 0) { f = 0.f; s = 0; }
            if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
        }
        ofs[d] = s; w[2 * d] = 1.f - f; w[2 * d + 1] = f;
    }
}
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

void orc_resize_u8(const uint8_t* src, int Hs, int Ws, int C, uint8_t* dst, int Hd, int Wd)
{
    int* xo = (int*)malloc(sizeof(int) * (size_t)(Wd + Hd));
    int* yo = xo + Wd;
    float* xw = (float*)malloc(sizeof(float) * 2 * (size_t)(Wd + Hd));
    float* yw = xw + 2 * Wd;
    resize_axis(Ws, Wd, 1, xo, xw);
    resize_axis(Hs, Hd, 0, yo, yw);
    for (int y = 0; y < Hd; ++y) {
        const int y0 = clampi(yo[y], 0, Hs - 1), y1 = clampi(yo[y] + 1, 0, Hs - 1);
        const int b0 = (int)lrintf(yw[2 * y] * 2048.f), b1 = (int)lrintf(yw[2 * y + 1] * 2048.f);
        for (int x = 0; x < Wd; ++x) {
            const int x0 = xo[x], x1 = x0 + 1 < Ws ? x0 + 1 : Ws - 1;
            const int a0 = (int)lrintf(xw[2 * x] * 2048.f), a1 = (int)lrintf(xw[2 * x + 1] * 2048.f);
            for (int c = 0; c < C; ++c) {
                const int S0 = src[((size_t)y0 * Ws + x0) * C + c] * a0 + src[((size_t)y0 * Ws + x1) * C + c] * a1;
                const int S1 = src[((size_t)y1 * Ws + x0) * C + c] * a0 + src[((size_t)y1 * Ws + x1) * C + c] * a1;
                dst[((size_t)y * Wd + x) * C + c] =
                    (uint8_t)((((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2);
            }
        }
    }
    free(xo); free(xw);
}

/* float32, one channel (fused_depth_map.py:1214-1217).  This build of cv2 (IPP) computes the source coordinate
 * and its fraction in double and rounds the weight to float once (probed: 1.2e-7 relative, i.e. summation order,
 * against 1e-5 for the float coordinates of the u8 path); same border rules.  Tolerance 1e-6 relative. */
void orc_resize_f32(const float* src, int Hs, int Ws, float* dst, int Hd, int Wd)
{
    const double sx = 1.0 / ((double)Wd / Ws), sy = 1.0 / ((double)Hd / Hs);
    for (int y = 0; y < Hd; ++y) {
        const double fy = (y + 0.5) * sy - 0.5;
        const int ys = (int)floor(fy);
        const float wy1 = (float)(fy - ys), wy0 = 1.f - wy1;
        const int y0 = clampi(ys, 0, Hs - 1), y1 = clampi(ys + 1, 0, Hs - 1);
        for (int x = 0; x < Wd; ++x) {
            double fx = (x + 0.5) * sx - 0.5;
            int xs = (int)floor(fx);
            double fr = fx - xs;
            if (xs < 0) { fr = 0; xs = 0; }
            if (xs >= Ws - 1) { fr = 0; xs = Ws - 1; }
            const int x1 = xs + 1 < Ws ? xs + 1 : Ws - 1;
            const float wx1 = (float)fr, wx0 = 1.f - wx1;
            const float r0 = src[(size_t)y0 * Ws + xs] * wx0 + src[(size_t)y0 * Ws + x1] * wx1;
            const float r1 = src[(size_t)y1 * Ws + xs] * wx0 + src[(size_t)y1 * Ws + x1] * wx1;
            dst[(size_t)y * Wd + x] = r0 * wy0 + r1 * wy1;
        }
    }
}
