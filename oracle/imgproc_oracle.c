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
