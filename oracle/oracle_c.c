/* CPU oracle, C restatement  --  TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Plain-C (OpenMP over rows) restatement of the reference's per-frame geometry path.  It exists
 * (1) as the checker for parity tests at sizes where the numpy oracle (oracle/oracle.py) is slow
 * and (2) as the *timed* CPU baseline for the two stages the reference itself runs as CUDA
 * kernels and therefore has no CPU code for (preprocess, reproject).  Only tests/,
 * __graft_entry__.smoke() and bench.py's CPU-baseline legs may load it; the product never does.
 *
 * Build: oracle/build_oracle.py  ->  oracle/_build/liboracle_c.so
 *        gcc -O2 -fopenmp -ffp-contract=off (no -ffast-math): every product/sum rounds once,
 *        exactly like the numpy oracle it is cross-checked against in tests/test_oracle_c.py.
 *
 * Reference lines followed (paths relative to the reference tree):
 *   oc_init_maps   : OpenCV calib3d initUndistortRectifyMap(CV_16SC2) as called at
 *                    FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:139-140
 *   oc_remap       : OpenCV imgproc remap(INTER_LINEAR, BORDER_CONSTANT 0), same file :142-143
 *   oc_preprocess  : RAFTStereo/src/stereo_preprocess.cu:4-39, CREStereo/src/CREStereo_preprocess.cu:4-39,
 *                    FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:4-39, HitNet/src/HitNet_preprocess.cu:4-53
 *   oc_reproject   : RAFTStereo/src/stereo_preprocess.cu:41-68 and the three *_reprojectImageTo3D_kernel bodies
 *   oc_heatmap     : FastACVNet_plus/test/main.cpp:7-24 (minMaxLoc, subtract, divide, convertScaleAbs, applyColorMap)
 *   oc_resize      : OpenCV imgproc resize(8U, INTER_LINEAR) downscale (extension oracle)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { OC_RAFT = 0, OC_HITNET = 1, OC_CRE = 2, OC_FASTACV = 3 };

static void inv3(const double a[9], double t[9])
{
    double d = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
    if (d == 0.0) { memset(t, 0, 9 * sizeof(double)); return; }
    d = 1.0 / d;
    t[0] = (a[4] * a[8] - a[5] * a[7]) * d; t[1] = (a[2] * a[7] - a[1] * a[8]) * d; t[2] = (a[1] * a[5] - a[2] * a[4]) * d;
    t[3] = (a[5] * a[6] - a[3] * a[8]) * d; t[4] = (a[0] * a[8] - a[2] * a[6]) * d; t[5] = (a[2] * a[3] - a[0] * a[5]) * d;
    t[6] = (a[3] * a[7] - a[4] * a[6]) * d; t[7] = (a[1] * a[6] - a[0] * a[7]) * d; t[8] = (a[0] * a[4] - a[1] * a[3]) * d;
}

static int cv_round_i32(double v)
{
    double r = nearbyint(v);                 /* default rounding mode: half to even */
    if (!(r >= -2147483648.0 && r <= 2147483647.0)) return INT32_MIN;
    return (int)r;
}
static int16_t sat16(int v) { return (int16_t)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); }

/* K 3x3, D nD coefficients, R 3x3, P 3x4 row-major (ldP = 4) or 3x3 (ldP = 3) */
void oc_init_maps(const double* K, const double* D, int nD, const double* R, const double* P, int ldP,
                  int W, int H, int16_t* map1, uint16_t* map2)
{
    double d[14] = {0}, A[9], iR[9];
    for (int i = 0; i < nD && i < 14; i++) d[i] = D[i];
    const double k1 = d[0], k2 = d[1], p1 = d[2], p2 = d[3], k3 = d[4], k4 = d[5], k5 = d[6], k6 = d[7];
    const double s1 = d[8], s2 = d[9], s3 = d[10], s4 = d[11];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            A[i * 3 + j] = P[i * ldP + 0] * R[0 * 3 + j] + P[i * ldP + 1] * R[1 * 3 + j] + P[i * ldP + 2] * R[2 * 3 + j];
    inv3(A, iR);
    const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
#pragma omp parallel for schedule(static)
    for (int i = 0; i < H; i++) {
        for (int j = 0; j < W; j++) {
            double X = iR[0] * j + iR[1] * i + iR[2];
            double Y = iR[3] * j + iR[4] * i + iR[5];
            double Wd = iR[6] * j + iR[7] * i + iR[8];
            double w = 1.0 / Wd, x = X * w, y = Y * w;
            double x2 = x * x, y2 = y * y, r2 = x2 + y2, _2xy = 2 * x * y;
            double kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1 + ((k6 * r2 + k5) * r2 + k4) * r2);
            double xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2) + s1 * r2 + s2 * r2 * r2;
            double yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy + s3 * r2 + s4 * r2 * r2;
            double u = fx * xd + cx, v = fy * yd + cy;
            int iu = cv_round_i32(u * 32), iv = cv_round_i32(v * 32);
            size_t o = (size_t)i * W + j;
            map1[2 * o] = sat16(iu >> 5);
            map1[2 * o + 1] = sat16(iv >> 5);
            map2[o] = (uint16_t)((iv & 31) * 32 + (iu & 31));
        }
    }
}

/* src [Hs][Ws][3] u8 -> dst [H][W][3] u8 */
void oc_remap(const uint8_t* src, int Hs, int Ws, const int16_t* map1, const uint16_t* map2, int H, int W, uint8_t* dst)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < H; i++) {
        for (int j = 0; j < W; j++) {
            size_t o = (size_t)i * W + j;
            int sx = map1[2 * o], sy = map1[2 * o + 1];
            int fx = map2[o] & 31, fy = (map2[o] >> 5) & 31;
            int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
            int x0 = sx >= 0 && sx < Ws, x1 = sx + 1 >= 0 && sx + 1 < Ws, y0 = sy >= 0 && sy < Hs, y1 = sy + 1 >= 0 && sy + 1 < Hs;
            for (int c = 0; c < 3; c++) {
                int p00 = (x0 && y0) ? src[((size_t)sy * Ws + sx) * 3 + c] : 0;
                int p01 = (x1 && y0) ? src[((size_t)sy * Ws + sx + 1) * 3 + c] : 0;
                int p10 = (x0 && y1) ? src[((size_t)(sy + 1) * Ws + sx) * 3 + c] : 0;
                int p11 = (x1 && y1) ? src[((size_t)(sy + 1) * Ws + sx + 1) * 3 + c] : 0;
                dst[o * 3 + c] = (uint8_t)((p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11 + 16384) >> 15);
            }
        }
    }
}

/* one eye: img [H][W][3] BGR u8 -> out [3][H][W] f32 RGB planes, kernel-body arithmetic */
void oc_preprocess(int model, const uint8_t* img, int H, int W, float* out)
{
    size_t area = (size_t)H * W;
#pragma omp parallel for schedule(static)
    for (int dy = 0; dy < H; dy++) {
        for (int dx = 0; dx < W; dx++) {
            const uint8_t* v = img + ((size_t)dy * W + dx) * 3;
            float c0 = v[0], c1 = v[1], c2 = v[2];
            if (model == OC_HITNET) { c0 = (float)v[0] / 255.0; c1 = (float)v[1] / 255.0; c2 = (float)v[2] / 255.0; }
            float t = c2; c2 = c0; c0 = t;                       /* bgr -> rgb */
            if (model == OC_FASTACV) {
                c0 = (c0 / 255.0f - 0.485) / 0.229;
                c1 = (c1 / 255.0f - 0.456) / 0.224;
                c2 = (c2 / 255.0f - 0.406) / 0.225;
            }
            float* p = out + (size_t)dy * W + dx;
            p[0] = c0; p[area] = c1; p[2 * area] = c2;
        }
    }
}

/* float -> IEEE half, round to nearest even (what __float2half_rn / numpy astype(float16) do), integer arithmetic only */
static uint16_t f32_to_f16(float f)
{
    union { float f; uint32_t u; } v, magic;
    v.f = f;
    magic.u = ((127u - 15u) + (23u - 10u) + 1u) << 23;
    const uint32_t sign = v.u & 0x80000000u;
    uint16_t o;
    v.u ^= sign;
    if (v.u >= ((127u + 16u) << 23)) o = (v.u > (255u << 23)) ? 0x7e00 : 0x7c00;       /* NaN : overflow -> inf */
    else if (v.u < (113u << 23)) { v.f += magic.f; o = (uint16_t)(v.u - magic.u); }      /* half subnormal / zero */
    else { const uint32_t odd = (v.u >> 13) & 1u; v.u += ((15u - 127u) << 23) + 0xfffu; v.u += odd; o = (uint16_t)(v.u >> 13); }
    return (uint16_t)(o | (sign >> 16));
}

/* extension a10: the same kernel-body arithmetic, each value then rounded once to half (out: [3][H][W] as uint16 bit patterns) */
void oc_preprocess_f16(int model, const uint8_t* img, int H, int W, uint16_t* out)
{
    size_t area = (size_t)H * W;
#pragma omp parallel for schedule(static)
    for (int dy = 0; dy < H; dy++) {
        for (int dx = 0; dx < W; dx++) {
            const uint8_t* v = img + ((size_t)dy * W + dx) * 3;
            float c0 = v[0], c1 = v[1], c2 = v[2];
            if (model == OC_HITNET) { c0 = (float)v[0] / 255.0; c1 = (float)v[1] / 255.0; c2 = (float)v[2] / 255.0; }
            float t = c2; c2 = c0; c0 = t;                       /* bgr -> rgb */
            if (model == OC_FASTACV) {
                c0 = (c0 / 255.0f - 0.485) / 0.229;
                c1 = (c1 / 255.0f - 0.456) / 0.224;
                c2 = (c2 / 255.0f - 0.406) / 0.225;
            }
            uint16_t* p = out + (size_t)dy * W + dx;
            p[0] = f32_to_f16(c0); p[area] = f32_to_f16(c1); p[2 * area] = f32_to_f16(c2);
        }
    }
}

/* disparity [H][W] f32 (negated in place for RAFT), left [H][W][3] BGR u8, Q[16] f32 -> points [H*W][6] */
void oc_reproject(int model, float* disparity, const uint8_t* left, const float* Q, int H, int W, float* points)
{
#pragma omp parallel for schedule(static)
    for (int row = 0; row < H; row++) {
        for (int col = 0; col < W; col++) {
            size_t o = (size_t)row * W + col;
            float dv = disparity[o];
            if (model == OC_RAFT) { dv = -dv; disparity[o] = dv; }
            float w = Q[14] * dv;
            for (int i = 0; i < 3; i++)
                points[o * 6 + i] = (Q[i * 4] * col + Q[i * 4 + 1] * row + Q[i * 4 + 3] * 1) / w;
            points[o * 6 + 3] = (float)left[o * 3 + 2];
            points[o * 6 + 4] = (float)left[o * 3 + 1];
            points[o * 6 + 5] = (float)left[o * 3 + 0];
        }
    }
}

/* stable row-major compaction with predicate isfinite(d) && d > 0 on the disparity as used */
long oc_compact(const float* disparity_used, const float* dense, long n, float* out)
{
    long k = 0;
    for (long i = 0; i < n; i++) {
        float d = disparity_used[i];
        if (isfinite(d) && d > 0.0f) { memcpy(out + k * 6, dense + i * 6, 6 * sizeof(float)); k++; }
    }
    return k;
}

/* jet [256][3] BGR; out [H][W][3] */
void oc_heatmap(const float* disp, int H, int W, const uint8_t* jet, uint8_t* out)
{
    size_t n = (size_t)H * W;
    float mnf = disp[0], mxf = disp[0];
    for (size_t i = 1; i < n; i++) { if (disp[i] < mnf) mnf = disp[i]; if (disp[i] > mxf) mxf = disp[i]; }
    double mn = mnf, mx = mxf;
    float s = (float)((mx - mn) / 255);
    float m = (float)mn;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) {
        float t = disp[i] - m;
        float q = t / s;
        float a = fabsf(q);
        int idx;
        if (!(a == a)) idx = 0;                         /* NaN -> cvRound gives INT_MIN -> saturates to 0 */
        else { double r = nearbyint((double)a); idx = r > 255.0 ? 255 : (int)r; }
        out[i * 3 + 0] = jet[idx * 3 + 0]; out[i * 3 + 1] = jet[idx * 3 + 1]; out[i * 3 + 2] = jet[idx * 3 + 2];
    }
}

static void resize_tab(int D, int S, int* s0, int* s1, int* c0, int* c1)
{
    double scale = (double)S / D;
    for (int d = 0; d < D; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= S - 1) { s = S - 1; f = 0.f; }
        s0[d] = s; s1[d] = s + 1 < S ? s + 1 : S - 1;
        c0[d] = (int)nearbyintf((1.f - f) * 2048.f);
        c1[d] = (int)nearbyintf(f * 2048.f);
    }
}

/* src [H][W][3] -> dst [h][w][3], downscale only */
void oc_resize(const uint8_t* src, int H, int W, uint8_t* dst, int h, int w)
{
    if (h == H && w == W) { memcpy(dst, src, (size_t)H * W * 3); return; }
    if (W == 2 * w && H == 2 * h) {                       /* OpenCV routes exact 2x to INTER_AREA */
#pragma omp parallel for schedule(static)
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                for (int c = 0; c < 3; c++) {
                    const uint8_t* p = src + ((size_t)(2 * y) * W + 2 * x) * 3 + c;
                    dst[((size_t)y * w + x) * 3 + c] = (uint8_t)((p[0] + p[3] + p[(size_t)W * 3] + p[(size_t)W * 3 + 3] + 2) >> 2);
                }
        return;
    }
    int* tab = (int*)malloc(sizeof(int) * 4 * (size_t)(w + h));
    int *x0 = tab, *x1 = tab + w, *a0 = tab + 2 * w, *a1 = tab + 3 * w;
    int *y0 = tab + 4 * w, *y1 = y0 + h, *b0 = y0 + 2 * h, *b1 = y0 + 3 * h;
    resize_tab(w, W, x0, x1, a0, a1);
    resize_tab(h, H, y0, y1, b0, b1);
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; y++) {
        const uint8_t* r0 = src + (size_t)y0[y] * W * 3;
        const uint8_t* r1 = src + (size_t)y1[y] * W * 3;
        for (int x = 0; x < w; x++)
            for (int c = 0; c < 3; c++) {
                int T0 = r0[x0[x] * 3 + c] * a0[x] + r0[x1[x] * 3 + c] * a1[x];
                int T1 = r1[x0[x] * 3 + c] * a0[x] + r1[x1[x] * 3 + c] * a1[x];
                int v = (((b0[y] * (T0 >> 4)) >> 16) + ((b1[y] * (T1 >> 4)) >> 16) + 2) >> 2;
                dst[((size_t)y * w + x) * 3 + c] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
            }
    }
    free(tab);
}

int oc_max_threads(void);
#ifdef _OPENMP
#include <omp.h>
int oc_max_threads(void) { return omp_get_max_threads(); }
void oc_set_threads(int n) { omp_set_num_threads(n); }
#else
int oc_max_threads(void) { return 1; }
void oc_set_threads(int n) { (void)n; }
#endif
