/*
 * farneback_oracle.c -- TEST INFRASTRUCTURE ONLY (the parity checker).
 *
 * Plain-C CPU restatement of the dense Farneback optical flow that the
 * reference reaches through cv2.calcOpticalFlowFarneback (call sites:
 * optical_flow_seg.py:158,203,494; optical_flow_ob.py:225,270,616;
 * optical_flow_prediction.py:161,206,570; optical_flow_yolo.py:199,244,898).
 * The arithmetic lives in a third-party dependency that is NOT under
 * /root/reference: opencv-python==4.9.0 (requirements.txt:1),
 * modules/video/src/optflowgf.cpp, FarnebackOpticalFlowImpl::calc.  This file
 * restates that published algorithm and is PINNED
 * against cv2 4.13.0 outputs generated in the build container
 * (tests/golden/make_golden.py -> tests/golden/ npz files): flags = 0 with the three
 * parameter sets the reference ships, OPTFLOW_FARNEBACK_GAUSSIAN (farneback_cv2.npz),
 * OPTFLOW_USE_INITIAL_FLOW (farneback_initflow_cv2.npz; its INTER_AREA downscale is
 * bit-exact against cv2.resize, tests/test_oracle_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product path never does.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (see oracle/Makefile).
 * -ffp-contract=off matters: OpenCV's x86-64 baseline build of optflowgf.cpp
 * has no FMA, so every float multiply and add below rounds separately.
 *
 * Layouts: images row-major; R and M are AoS, 5 floats per pixel, exactly as
 * OpenCV keeps them; flow is 2 floats per pixel (dx, dy).
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#define NSOF_ORACLE_API __attribute__((visibility("default")))

static int cv_round(double v) { return (int)nearbyint(v); } /* round-half-even, like cvRound */
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

static int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

/* ---- A.2: cv::getGaussianKernel(ksize, sigma, CV_32F) ------------------- */
NSOF_ORACLE_API void nsof_oracle_gaussian_kernel(int ksize, double sigma, float *k)
{
    static const float small3[] = {0.25f, 0.5f, 0.25f};
    static const float small5[] = {0.0625f, 0.25f, 0.375f, 0.25f, 0.0625f};
    static const float small7[] = {0.03125f, 0.109375f, 0.21875f, 0.28125f,
                                   0.21875f, 0.109375f, 0.03125f};
    if (sigma <= 0 && (ksize == 1 || ksize == 3 || ksize == 5 || ksize == 7)) {
        if (ksize == 1) { k[0] = 1.f; return; }
        const float *t = ksize == 3 ? small3 : ksize == 5 ? small5 : small7;
        memcpy(k, t, sizeof(float) * ksize);
        return;
    }
    double sx = sigma > 0 ? sigma : ((ksize - 1) * 0.5 - 1) * 0.3 + 0.8;
    double scale2 = -0.5 / (sx * sx);
    double tmp[64];
    double sum = 0;
    for (int i = 0; i < ksize; i++) {
        double x = i - (ksize - 1) * 0.5;
        tmp[i] = exp(scale2 * x * x);
        sum += tmp[i];
    }
    sum = 1. / sum;
    for (int i = 0; i < ksize; i++) k[i] = (float)(tmp[i] * sum);
}

/* ---- A.2 + A.3: I_k = resize(GaussianBlur(f32(img)), (wk,hk), LINEAR) ---- */
/* f32 != 0: `img` holds float32 samples (pitch in bytes) -- cv2 accepts CV_32F frames, convertTo(CV_32F) is then the
 * identity; the reference's call sites pass uint8. */
NSOF_ORACLE_API int nsof_oracle_pyr_level_ex(const uint8_t *img, int h, int w, ptrdiff_t pitch, int f32,
                                             int ksize, double sigma, int hk, int wk, float *out);
NSOF_ORACLE_API int nsof_oracle_pyr_level(const uint8_t *img, int h, int w, ptrdiff_t pitch,
                                          int ksize, double sigma, int hk, int wk, float *out)
{
    return nsof_oracle_pyr_level_ex(img, h, w, pitch, 0, ksize, sigma, hk, wk, out);
}
NSOF_ORACLE_API int nsof_oracle_pyr_level_ex(const uint8_t *img, int h, int w, ptrdiff_t pitch, int f32,
                                             int ksize, double sigma, int hk, int wk, float *out)
{
    if (ksize > 63 || !(ksize & 1)) return -1;
    float kern[64];
    nsof_oracle_gaussian_kernel(ksize, sigma, kern);
    int r = ksize / 2;
    float *tmp = (float *)malloc(sizeof(float) * (size_t)h * w);
    float *blur = (float *)malloc(sizeof(float) * (size_t)h * w);
    if (!tmp || !blur) { free(tmp); free(blur); return -2; }
    /* row pass (symmetric pairs, float accumulate) */
    for (int y = 0; y < h; y++) {
        const uint8_t *s = img + (ptrdiff_t)y * pitch;
        const float *sf = (const float *)s;
#define SRC(x_) (f32 ? sf[x_] : (float)s[x_])
        for (int x = 0; x < w; x++) {
            float acc = kern[r] * SRC(x);
            for (int i = 1; i <= r; i++) {
                float a = SRC(reflect101(x - i, w));
                float b = SRC(reflect101(x + i, w));
                acc = acc + kern[r + i] * (a + b);
            }
            tmp[(size_t)y * w + x] = acc;
        }
#undef SRC
    }
    /* column pass */
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            float acc = kern[r] * tmp[(size_t)y * w + x];
            for (int i = 1; i <= r; i++) {
                float a = tmp[(size_t)reflect101(y - i, h) * w + x];
                float b = tmp[(size_t)reflect101(y + i, h) * w + x];
                acc = acc + kern[r + i] * (a + b);
            }
            blur[(size_t)y * w + x] = acc;
        }
    }
    free(tmp);
    if (hk == h && wk == w) {
        memcpy(out, blur, sizeof(float) * (size_t)h * w);
    } else if (h == 2 * hk && w == 2 * wk) {
        /* cv::resize switches INTER_LINEAR to the INTER_AREA fast path at exact 2x */
        for (int y = 0; y < hk; y++)
            for (int x = 0; x < wk; x++) {
                const float *s0 = blur + (size_t)(2 * y) * w + 2 * x;
                const float *s1 = s0 + w;
                out[(size_t)y * wk + x] = (s0[0] + s0[1] + s1[0] + s1[1]) * 0.25f;
            }
    } else {
        double sxr = (double)w / wk, syr = (double)h / hk;
        for (int y = 0; y < hk; y++) {
            float fy = (float)((y + 0.5) * syr - 0.5);
            int y0 = (int)floorf(fy);
            fy -= y0;
            if (y0 < 0) { y0 = 0; fy = 0.f; }
            if (y0 >= h - 1) { y0 = h - 1; fy = 0.f; }
            int y1 = imin(y0 + 1, h - 1);
            float b0 = 1.f - fy, b1 = fy;
            for (int x = 0; x < wk; x++) {
                float fx = (float)((x + 0.5) * sxr - 0.5);
                int x0 = (int)floorf(fx);
                fx -= x0;
                if (x0 < 0) { x0 = 0; fx = 0.f; }
                if (x0 >= w - 1) { x0 = w - 1; fx = 0.f; }
                int x1 = imin(x0 + 1, w - 1);
                float a0 = 1.f - fx, a1 = fx;
                float top = blur[(size_t)y0 * w + x0] * a0 + blur[(size_t)y0 * w + x1] * a1;
                float bot = blur[(size_t)y1 * w + x0] * a0 + blur[(size_t)y1 * w + x1] * a1;
                out[(size_t)y * wk + x] = top * b0 + bot * b1;
            }
        }
    }
    free(blur);
    return 0;
}

/* ---- A.4: FarnebackPrepareGaussian -------------------------------------- */
static void invert6(double a[6][6], double inv[6][6])
{
    double m[6][12];
    for (int i = 0; i < 6; i++)
        for (int j = 0; j < 6; j++) { m[i][j] = a[i][j]; m[i][j + 6] = (i == j); }
    for (int c = 0; c < 6; c++) {
        int p = c;
        for (int i = c + 1; i < 6; i++) if (fabs(m[i][c]) > fabs(m[p][c])) p = i;
        if (p != c) for (int j = 0; j < 12; j++) { double t = m[c][j]; m[c][j] = m[p][j]; m[p][j] = t; }
        double d = 1.0 / m[c][c];
        for (int j = 0; j < 12; j++) m[c][j] *= d;
        for (int i = 0; i < 6; i++) if (i != c) {
            double f = m[i][c];
            if (f != 0) for (int j = 0; j < 12; j++) m[i][j] -= f * m[c][j];
        }
    }
    for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) inv[i][j] = m[i][j + 6];
}

/* g, xg, xxg: arrays of 2n+1 floats, index [x+n] for x in -n..n; ig = {ig11, ig03, ig33, ig55} */
NSOF_ORACLE_API void nsof_oracle_prepare_gaussian(int n, double sigma, float *g, float *xg,
                                                  float *xxg, double *ig)
{
    if (sigma < FLT_EPSILON) sigma = n * 0.3;
    double s = 0.;
    for (int x = -n; x <= n; x++) {
        g[x + n] = (float)exp(-x * x / (2 * sigma * sigma));
        s += g[x + n];
    }
    s = 1. / s;
    for (int x = -n; x <= n; x++) {
        g[x + n] = (float)(g[x + n] * s);
        xg[x + n] = (float)(x * g[x + n]);
        xxg[x + n] = (float)(x * x * g[x + n]);
    }
    double G[6][6], inv[6][6];
    memset(G, 0, sizeof(G));
    for (int y = -n; y <= n; y++)
        for (int x = -n; x <= n; x++) {
            /* float products, accumulated in double (C++ evaluation order) */
            float gg = g[y + n] * g[x + n];
            G[0][0] += gg;
            G[1][1] += gg * x * x;
            G[3][3] += gg * x * x * x * x;
            G[5][5] += gg * x * x * y * y;
        }
    G[2][2] = G[0][3] = G[0][4] = G[3][0] = G[4][0] = G[1][1];
    G[4][4] = G[3][3];
    G[3][4] = G[4][3] = G[5][5];
    invert6(G, inv);
    ig[0] = inv[1][1];
    ig[1] = inv[0][3];
    ig[2] = inv[3][3];
    ig[3] = inv[5][5];
}

/* ---- A.5: FarnebackPolyExp  I (h x w float) -> R (h x w x 5 float) ------- */
NSOF_ORACLE_API int nsof_oracle_polyexp(const float *I, int h, int w, int n, double sigma, float *R)
{
    if (n < 1 || n > 31) return -1;
    float gb[64], xgb[64], xxgb[64];
    double igv[4];
    nsof_oracle_prepare_gaussian(n, sigma, gb, xgb, xxgb, igv);
    const float *g = gb + n, *xg = xgb + n, *xxg = xxgb + n;
    double ig11 = igv[0], ig03 = igv[1], ig33 = igv[2], ig55 = igv[3];
    float *rowbuf = (float *)malloc(sizeof(float) * (size_t)(w + 2 * n) * 3);
    if (!rowbuf) return -2;
    float *row = rowbuf + n * 3;
    for (int y = 0; y < h; y++) {
        const float *s0 = I + (size_t)y * w;
        float *d = R + (size_t)y * w * 5;
        float g0 = g[0];
        for (int x = 0; x < w; x++) {
            row[x * 3] = s0[x] * g0;
            row[x * 3 + 1] = row[x * 3 + 2] = 0.f;
        }
        for (int k = 1; k <= n; k++) {
            float gk = g[k], g1 = xg[k], g2 = xxg[k];
            const float *a = I + (size_t)imax(y - k, 0) * w;
            const float *b = I + (size_t)imin(y + k, h - 1) * w;
            for (int x = 0; x < w; x++) {
                float p = a[x] + b[x];
                float t0 = row[x * 3] + gk * p;
                float t1 = row[x * 3 + 1] + g1 * (b[x] - a[x]);
                float t2 = row[x * 3 + 2] + g2 * p;
                row[x * 3] = t0; row[x * 3 + 1] = t1; row[x * 3 + 2] = t2;
            }
        }
        for (int x = 0; x < n * 3; x++) {
            row[-1 - x] = row[2 - x];
            row[w * 3 + x] = row[w * 3 + x - 3];
        }
        for (int x = 0; x < w; x++) {
            g0 = g[0];
            double b1 = row[x * 3] * g0, b2 = 0, b3 = row[x * 3 + 1] * g0,
                   b4 = 0, b5 = row[x * 3 + 2] * g0, b6 = 0;
            for (int k = 1; k <= n; k++) {
                double tg = row[(x + k) * 3] + row[(x - k) * 3];
                g0 = g[k];
                b1 += tg * g0;
                b4 += tg * xxg[k];
                b2 += (row[(x + k) * 3] - row[(x - k) * 3]) * xg[k];
                b3 += (row[(x + k) * 3 + 1] + row[(x - k) * 3 + 1]) * g0;
                b6 += (row[(x + k) * 3 + 1] - row[(x - k) * 3 + 1]) * xg[k];
                b5 += (row[(x + k) * 3 + 2] + row[(x - k) * 3 + 2]) * g0;
            }
            d[x * 5 + 1] = (float)(b2 * ig11);
            d[x * 5] = (float)(b3 * ig11);
            d[x * 5 + 3] = (float)(b1 * ig03 + b4 * ig33);
            d[x * 5 + 2] = (float)(b1 * ig03 + b5 * ig33);
            d[x * 5 + 4] = (float)(b6 * ig55);
        }
    }
    free(rowbuf);
    return 0;
}

/* ---- A.6: FarnebackUpdateMatrices --------------------------------------- */
NSOF_ORACLE_API void nsof_oracle_update_matrices(const float *R0, const float *R1, const float *flow,
                                                 int h, int w, float *M)
{
    static const float border[5] = {0.14f, 0.14f, 0.4472f, 0.4472f, 0.4472f};
    const int BORDER = 5;
    size_t step1 = (size_t)w * 5;
    for (int y = 0; y < h; y++) {
        const float *fl = flow + (size_t)y * w * 2;
        const float *r0 = R0 + (size_t)y * w * 5;
        float *m = M + (size_t)y * w * 5;
        for (int x = 0; x < w; x++) {
            float dx = fl[x * 2], dy = fl[x * 2 + 1];
            float fx = x + dx, fy = y + dy;
            int x1 = (int)floorf(fx), y1 = (int)floorf(fy);
            float r2, r3, r4, r5, r6;
            fx -= x1; fy -= y1;
            if ((unsigned)x1 < (unsigned)(w - 1) && (unsigned)y1 < (unsigned)(h - 1)) {
                const float *p = R1 + (size_t)y1 * step1 + (size_t)x1 * 5;
                float a00 = (1.f - fx) * (1.f - fy), a01 = fx * (1.f - fy),
                      a10 = (1.f - fx) * fy, a11 = fx * fy;
                r2 = a00 * p[0] + a01 * p[5] + a10 * p[step1] + a11 * p[step1 + 5];
                r3 = a00 * p[1] + a01 * p[6] + a10 * p[step1 + 1] + a11 * p[step1 + 6];
                r4 = a00 * p[2] + a01 * p[7] + a10 * p[step1 + 2] + a11 * p[step1 + 7];
                r5 = a00 * p[3] + a01 * p[8] + a10 * p[step1 + 3] + a11 * p[step1 + 8];
                r6 = a00 * p[4] + a01 * p[9] + a10 * p[step1 + 4] + a11 * p[step1 + 9];
                r4 = (r0[x * 5 + 2] + r4) * 0.5f;
                r5 = (r0[x * 5 + 3] + r5) * 0.5f;
                r6 = (r0[x * 5 + 4] + r6) * 0.25f;
            } else {
                r2 = r3 = 0.f;
                r4 = r0[x * 5 + 2];
                r5 = r0[x * 5 + 3];
                r6 = r0[x * 5 + 4] * 0.5f;
            }
            r2 = (r0[x * 5] - r2) * 0.5f;
            r3 = (r0[x * 5 + 1] - r3) * 0.5f;
            r2 += r4 * dy + r6 * dx;
            r3 += r6 * dy + r5 * dx;
            if ((unsigned)(x - BORDER) >= (unsigned)(w - BORDER * 2) ||
                (unsigned)(y - BORDER) >= (unsigned)(h - BORDER * 2)) {
                float scale = (x < BORDER ? border[x] : 1.f) *
                              (x >= w - BORDER ? border[w - x - 1] : 1.f) *
                              (y < BORDER ? border[y] : 1.f) *
                              (y >= h - BORDER ? border[h - y - 1] : 1.f);
                r2 *= scale; r3 *= scale; r4 *= scale; r5 *= scale; r6 *= scale;
            }
            m[x * 5] = r4 * r4 + r6 * r6;
            m[x * 5 + 1] = (r4 + r5) * r6;
            m[x * 5 + 2] = r5 * r5 + r6 * r6;
            m[x * 5 + 3] = r4 * r2 + r6 * r3;
            m[x * 5 + 4] = r6 * r2 + r5 * r3;
        }
    }
}

/* ---- A.7: FarnebackUpdateFlow_Blur (box window, double sums) ------------- */
/* Window = rows/cols [-m, +m] with replicate borders, m = winsize/2; direct double
 * sums (OpenCV keeps running double sums; both are exact to ~1e-16 relative).
 * Quirk kept for winsize 1 (m = 0): OpenCV initialises its running sums with (m+2) copies of
 * row 0 / column 0, which is one copy too many when m = 0, and that copy is never subtracted:
 * the "window" sum becomes M(y,x) + M(0,x) + M(y,0) + M(0,0).  (libnsof rejects winsize 1.) */
NSOF_ORACLE_API int nsof_oracle_update_flow_blur(const float *M, int h, int w, int winsize, float *flow)
{
    int m = winsize / 2;
    double scale = 1. / (winsize * winsize);
    double *vsum = (double *)malloc(sizeof(double) * (size_t)w * 5);
    if (!vsum) return -2;
    for (int y = 0; y < h; y++) {
        for (int i = 0; i < w * 5; i++) vsum[i] = 0.;
        for (int dy = -m; dy <= m; dy++) {
            const float *r = M + (size_t)imin(imax(y + dy, 0), h - 1) * w * 5;
            for (int i = 0; i < w * 5; i++) vsum[i] += r[i];
        }
        if (m == 0)
            for (int i = 0; i < w * 5; i++) vsum[i] += M[i];
        for (int x = 0; x < w; x++) {
            double s[5] = {0, 0, 0, 0, 0};
            for (int dx = -m; dx <= m; dx++) {
                const double *v = vsum + (size_t)imin(imax(x + dx, 0), w - 1) * 5;
                for (int c = 0; c < 5; c++) s[c] += v[c];
            }
            if (m == 0)
                for (int c = 0; c < 5; c++) s[c] += vsum[c];
            double g11 = s[0] * scale, g12 = s[1] * scale, g22 = s[2] * scale,
                   h1 = s[3] * scale, h2 = s[4] * scale;
            double idet = 1. / (g11 * g22 - g12 * g12 + 1e-3);
            flow[((size_t)y * w + x) * 2] = (float)((g11 * h2 - g12 * h1) * idet);
            flow[((size_t)y * w + x) * 2 + 1] = (float)((g22 * h1 - g12 * h2) * idet);
        }
    }
    free(vsum);
    return 0;
}

/* ---- A.7': FarnebackUpdateFlow_GaussianBlur (OPTFLOW_FARNEBACK_GAUSSIAN = 256) ---- */
/* OpenCV optflowgf.cpp, FarnebackUpdateFlow_GaussianBlur: m = winsize/2, sigma = 0.3 m, float kernel
 * k[0] = 1, k[i] = f32(exp(-i^2 / (2 sigma^2))), normalised by s = 1 / (1 + 2 sum k[i]) in double and stored
 * float.  Vertical pass (float): v = M(y)*k[0]; v += (M(y+i) + M(y-i)) * k[i], rows clamped.  Horizontal pass
 * (float): s = v(x)*k[0]; s += k[i] * (v(x-i) + v(x+i)), columns clamped.  Solve: the products g11*g22 etc.
 * are FLOAT products (the g's are float variables), only "+ 1e-3", the reciprocal and the final product are
 * double.  Never selected by the reference's scripts ('flags': 0 everywhere); pinned against cv2 with flags=256. */
NSOF_ORACLE_API int nsof_oracle_gaussian_window(int winsize, float *k /* m+1 floats */)
{
    int m = winsize / 2;
    double sigma = m * 0.3, s = 1.;
    k[0] = 1.f;
    for (int i = 1; i <= m; i++) {
        float t = (float)exp(-i * i / (2 * sigma * sigma));
        k[i] = t;
        s += t * 2;
    }
    s = 1. / s;
    for (int i = 0; i <= m; i++) k[i] = (float)(k[i] * s);
    return m;
}

NSOF_ORACLE_API int nsof_oracle_update_flow_gaussian(const float *M, int h, int w, int winsize, float *flow)
{
    int m = winsize / 2;
    float *k = (float *)malloc(sizeof(float) * (size_t)(m + 1));
    float *vsum = (float *)malloc(sizeof(float) * (size_t)w * 5);
    if (!k || !vsum) return -2;
    nsof_oracle_gaussian_window(winsize, k);
    for (int y = 0; y < h; y++) {
        const float *r0 = M + (size_t)y * w * 5;
        for (int i = 0; i < w * 5; i++) vsum[i] = r0[i] * k[0];
        for (int j = 1; j <= m; j++) {
            const float *ra = M + (size_t)imax(y - j, 0) * w * 5;
            const float *rb = M + (size_t)imin(y + j, h - 1) * w * 5;
            for (int i = 0; i < w * 5; i++) vsum[i] += (rb[i] + ra[i]) * k[j];
        }
        for (int x = 0; x < w; x++) {
            float s[5];
            for (int c = 0; c < 5; c++) s[c] = vsum[(size_t)x * 5 + c] * k[0];
            for (int j = 1; j <= m; j++) {
                const float *va = vsum + (size_t)imax(x - j, 0) * 5, *vb = vsum + (size_t)imin(x + j, w - 1) * 5;
                for (int c = 0; c < 5; c++) s[c] += k[j] * (va[c] + vb[c]);
            }
            float g11 = s[0], g12 = s[1], g22 = s[2], h1 = s[3], h2 = s[4];
            double idet = 1. / (g11 * g22 - g12 * g12 + 1e-3);
            flow[((size_t)y * w + x) * 2] = (float)((g11 * h2 - g12 * h1) * idet);
            flow[((size_t)y * w + x) * 2 + 1] = (float)((g22 * h1 - g12 * h2) * idet);
        }
    }
    free(vsum);
    free(k);
    return 0;
}

/* ---- A.1/A.3: inter-level flow upsample: resize(LINEAR) * (1/pyr_scale) --- */
NSOF_ORACLE_API void nsof_oracle_upsample_flow(const float *src, int hs, int ws, int hd, int wd,
                                               double pyr_scale, float *dst)
{
    float mul = (float)(1. / pyr_scale);
    double sxr = (double)ws / wd, syr = (double)hs / hd;
    for (int y = 0; y < hd; y++) {
        float fy = (float)((y + 0.5) * syr - 0.5);
        int y0 = (int)floorf(fy);
        fy -= y0;
        if (y0 < 0) { y0 = 0; fy = 0.f; }
        if (y0 >= hs - 1) { y0 = hs - 1; fy = 0.f; }
        int y1 = imin(y0 + 1, hs - 1);
        float b0 = 1.f - fy, b1 = fy;
        for (int x = 0; x < wd; x++) {
            float fx = (float)((x + 0.5) * sxr - 0.5);
            int x0 = (int)floorf(fx);
            fx -= x0;
            if (x0 < 0) { x0 = 0; fx = 0.f; }
            if (x0 >= ws - 1) { x0 = ws - 1; fx = 0.f; }
            int x1 = imin(x0 + 1, ws - 1);
            float a0 = 1.f - fx, a1 = fx;
            for (int c = 0; c < 2; c++) {
                float top = src[((size_t)y0 * ws + x0) * 2 + c] * a0 + src[((size_t)y0 * ws + x1) * 2 + c] * a1;
                float bot = src[((size_t)y1 * ws + x0) * 2 + c] * a0 + src[((size_t)y1 * ws + x1) * 2 + c] * a1;
                dst[((size_t)y * wd + x) * 2 + c] = (top * b0 + bot * b1) * mul;
            }
        }
    }
}

/* ---- A.8: OPTFLOW_USE_INITIAL_FLOW (4): cv::resize(flow0, INTER_AREA) of a 2-channel float image ---- */
/* Restated from OpenCV imgproc/resize.cpp and pinned bit for bit against cv2.resize (tests/test_oracle_golden.py).
 * scale = src/dst per axis (>= 1 here).  Both scales integer ("area fast"): plain sum of the ix*iy source pixels in
 * row-major order, accumulated four at a time as sum += ((a+b)+c)+d (ResizeAreaFastNoVec's unrolled loop), times
 * f32(1/area).  Otherwise the DecimateAlpha tables (computeResizeAreaTab): per destination cell the partially covered
 * first source pixel, the fully covered ones (weight 1/cellWidth) and the partially covered last one; rows are combined
 * as sum = beta_0*buf_0 + beta_1*buf_1 + ... with buf = horizontal weighted sum, all float. */
static int area_tab(int ssize, int dsize, int dx, int *si, float *alpha /* room for (int)scale + 3 */)
{
    double scale = (double)ssize / dsize;
    double fsx1 = dx * scale, fsx2 = fsx1 + scale;
    double cell = scale < ssize - fsx1 ? scale : ssize - fsx1;
    int sx1 = (int)ceil(fsx1), sx2 = (int)floor(fsx2), k = 0;
    sx2 = imin(sx2, ssize - 1);
    sx1 = imin(sx1, sx2);
    if (sx1 - fsx1 > 1e-3) { si[k] = sx1 - 1; alpha[k++] = (float)((sx1 - fsx1) / cell); }
    for (int sx = sx1; sx < sx2; sx++) { si[k] = sx; alpha[k++] = (float)(1.0 / cell); }
    if (fsx2 - sx2 > 1e-3) {
        double t = fsx2 - sx2;
        if (t > 1.) t = 1.;
        if (t > cell) t = cell;
        si[k] = sx2;
        alpha[k++] = (float)(t / cell);
    }
    return k;
}

NSOF_ORACLE_API int nsof_oracle_resize_area_c2(const float *src, int hs, int ws, int hd, int wd, float *dst)
{
    if (hd > hs || wd > ws || hd < 1 || wd < 1) return -1;
    if (hs % hd == 0 && ws % wd == 0) {
        int iy = hs / hd, ix = ws / wd, area = ix * iy;
        float sc = 1.f / area;
        for (int dy = 0; dy < hd; dy++)
            for (int dx = 0; dx < wd; dx++)
                for (int c = 0; c < 2; c++) {
#define AV(k) src[(((size_t)dy * iy + (k) / ix) * ws + (size_t)dx * ix + (k) % ix) * 2 + c]
                    float sum = 0.f;
                    int k = 0;
                    for (; k <= area - 4; k += 4) sum += AV(k) + AV(k + 1) + AV(k + 2) + AV(k + 3);
                    for (; k < area; k++) sum += AV(k);
#undef AV
                    dst[((size_t)dy * wd + dx) * 2 + c] = sum * sc;
                }
        return 0;
    }
    int nx = ws / wd + 3, ny = hs / hd + 3;
    int *xi = (int *)malloc(sizeof(int) * (size_t)(nx + ny));
    float *xa = (float *)malloc(sizeof(float) * (size_t)(nx + ny));
    if (!xi || !xa) return -2;
    int *yi = xi + nx;
    float *ya = xa + nx;
    for (int dy = 0; dy < hd; dy++) {
        int ky = area_tab(hs, hd, dy, yi, ya);
        for (int dx = 0; dx < wd; dx++) {
            int kx = area_tab(ws, wd, dx, xi, xa);
            for (int c = 0; c < 2; c++) {
                float sum = 0.f;
                for (int j = 0; j < ky; j++) {
                    float buf = 0.f;
                    for (int i = 0; i < kx; i++) buf = buf + src[((size_t)yi[j] * ws + xi[i]) * 2 + c] * xa[i];
                    sum = j == 0 ? ya[j] * buf : sum + ya[j] * buf;
                }
                dst[((size_t)dy * wd + dx) * 2 + c] = sum;
            }
        }
    }
    free(xi);
    free(xa);
    return 0;
}

/* ---- A.1: level geometry (shared by tests to address stage dumps) -------- */
/* Returns the number of scales (levels_eff + 1); fills hk/wk/ksize/sigma for k = 0..levels_eff. */
NSOF_ORACLE_API int nsof_oracle_level_geometry(int h, int w, double pyr_scale, int levels,
                                               int *hk, int *wk, int *ksize, double *sigma)
{
    double scale = 1;
    int cropped = 0;
    for (; cropped < levels; cropped++) {
        scale *= pyr_scale;
        if (w * scale < 32 || h * scale < 32) break;
    }
    for (int k = 0; k <= cropped; k++) {
        scale = 1;
        for (int i = 0; i < k; i++) scale *= pyr_scale;
        double sg = (1. / scale - 1) * 0.5;
        int ks = cv_round(sg * 5) | 1;
        ks = imax(ks, 3);
        if (hk) hk[k] = cv_round(h * scale);
        if (wk) wk[k] = cv_round(w * scale);
        if (ksize) ksize[k] = ks;
        if (sigma) sigma[k] = sg;
    }
    return cropped + 1;
}

/* ---- A.1: the whole call, flags = 0 or OPTFLOW_FARNEBACK_GAUSSIAN (256), flow = None ---- */
/* stage_dump (optional, may be NULL): called after every stage with a tag so tests can
 * compare device intermediates; tag: 0=I(img0) 1=I(img1) 2=R0 3=R1 4=M(after first
 * UpdateMatrices) 5=flow(after iteration `it`), level k. */
typedef void (*nsof_oracle_dump_fn)(void *user, int tag, int k, int it, const float *data, int h, int w, int ch);

NSOF_ORACLE_API int nsof_oracle_farneback_dump(const uint8_t *prev, const uint8_t *next, int h, int w,
                                               ptrdiff_t pitch0, ptrdiff_t pitch1, double pyr_scale,
                                               int levels, int winsize, int iterations, int poly_n,
                                               double poly_sigma, int flags, float *flow_out,
                                               nsof_oracle_dump_fn dump, void *user)
{
    if (!prev || !next || h < 1 || w < 1) return -1;
    if (!(pyr_scale < 1.) || pyr_scale <= 0 || levels < 0 || winsize < 1 || iterations < 0 || poly_n < 1) return -1;
    if (flags & ~(256 | 4 | 0x10000)) return -3;  /* 0x10000: float32 frames (libnsof's NSOF_FB_SRC_F32) */
    /* flags & 4 (OPTFLOW_USE_INITIAL_FLOW): flow_out holds the initial flow on entry (no call site of the reference
     * passes one; restated for the drop-in signature and pinned against cv2) */
    int hk[64], wk[64], ks[64];
    double sg[64];
    if (levels > 62) return -1;
    int nscales = nsof_oracle_level_geometry(h, w, pyr_scale, levels, hk, wk, ks, sg);
    float *prev_flow = NULL;
    int ph = 0, pw = 0;
    for (int k = nscales - 1; k >= 0; k--) {
        int H = hk[k], W = wk[k];
        size_t N = (size_t)H * W;
        float *flow = (k == 0) ? flow_out : (float *)malloc(sizeof(float) * N * 2);
        float *I = (float *)malloc(sizeof(float) * N);
        float *R0 = (float *)malloc(sizeof(float) * N * 5);
        float *R1 = (float *)malloc(sizeof(float) * N * 5);
        float *M = (float *)malloc(sizeof(float) * N * 5);
        if (!flow || !I || !R0 || !R1 || !M) return -2;
        if (!prev_flow && (flags & 4)) {
            /* resize(flow0, INTER_AREA) then flow *= scale (Mat *= double: f32(v * f32(scale))) */
            double sc = 1;
            for (int i = 0; i < k; i++) sc *= pyr_scale;
            if (k > 0) {
                int rc = nsof_oracle_resize_area_c2(flow_out, h, w, H, W, flow);
                if (rc) return rc;
            }
            for (size_t i = 0; i < N * 2; i++) flow[i] = flow[i] * (float)sc;
        } else if (!prev_flow) memset(flow, 0, sizeof(float) * N * 2);
        else nsof_oracle_upsample_flow(prev_flow, ph, pw, H, W, pyr_scale, flow);
        for (int i = 0; i < 2; i++) {
            int rc = nsof_oracle_pyr_level_ex(i ? next : prev, h, w, i ? pitch1 : pitch0, (flags & 0x10000) != 0, ks[k], sg[k], H, W, I);
            if (rc) return rc;
            if (dump) dump(user, i, k, 0, I, H, W, 1);
            rc = nsof_oracle_polyexp(I, H, W, poly_n, poly_sigma, i ? R1 : R0);
            if (rc) return rc;
            if (dump) dump(user, 2 + i, k, 0, i ? R1 : R0, H, W, 5);
        }
        nsof_oracle_update_matrices(R0, R1, flow, H, W, M);
        if (dump) dump(user, 4, k, 0, M, H, W, 5);
        for (int it = 0; it < iterations; it++) {
            int rc = (flags & 256) ? nsof_oracle_update_flow_gaussian(M, H, W, winsize, flow)
                                   : nsof_oracle_update_flow_blur(M, H, W, winsize, flow);
            if (rc) return rc;
            if (dump) dump(user, 5, k, it, flow, H, W, 2);
            if (it < iterations - 1) nsof_oracle_update_matrices(R0, R1, flow, H, W, M);
        }
        free(I); free(R0); free(R1); free(M);
        free(prev_flow);
        prev_flow = (k == 0) ? NULL : flow;
        ph = H; pw = W;
    }
    return 0;
}

NSOF_ORACLE_API int nsof_oracle_farneback(const uint8_t *prev, const uint8_t *next, int h, int w,
                                          ptrdiff_t pitch0, ptrdiff_t pitch1, double pyr_scale,
                                          int levels, int winsize, int iterations, int poly_n,
                                          double poly_sigma, int flags, float *flow_out)
{
    return nsof_oracle_farneback_dump(prev, next, h, w, pitch0, pitch1, pyr_scale, levels, winsize,
                                      iterations, poly_n, poly_sigma, flags, flow_out, NULL, NULL);
}
