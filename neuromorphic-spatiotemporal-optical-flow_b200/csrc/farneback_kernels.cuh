// farneback_kernels.cuh -- stage 2 (dense Farneback) device code for sm_100a.
//
// Replaces, behind the reference's call `cv2.calcOpticalFlowFarneback(prev, next, None,
// **farneback_params)` (optical_flow_seg.py:158,203,494 and the copies in _ob/_prediction/_yolo),
// the arithmetic of OpenCV's FarnebackOpticalFlowImpl::calc (opencv-python 4.9.0,
// modules/video/src/optflowgf.cpp -- not vendored in the reference; specification in
// SURVEY.md Appendix A, CPU restatement in oracle/farneback_oracle.c).
//
// Data layout in HBM (all private workspace, per crop "item", per pyramid level):
//   I   : 2 planes  float  [img][h][pitch]            blurred+resized image (level input)
//   R   : 10 planes float  [img*5 + c][h][pitch]      polynomial expansion, c = (b_y,b_x,A_yy,A_xx,A_xy)
//   M   : 5 planes  float  [c][h][pitch]  (two buffers, ping-pong across iterations)
//   flow: float2 interleaved [h][fpitch]              (dx, dy) as cv2 returns it
// pitch is a multiple of 32 floats (128 B) so every row starts on a cache line and float4
// accesses at x % 4 == 0 are aligned.  Planar (SoA) instead of OpenCV's AoS-5 keeps every
// warp access a run of consecutive 4-byte words; byte counts are identical.
//
// All kernels take a device table of per-item jobs and run as grid (max_tiles, n_items[, 2]):
// blocks past an item's tile count exit at once, which is how ragged ROI crops batch.
#pragma once
#include "nsof_rt.h"

#define NSOF_MAX_POLY_N 31

struct FbJob {
    int h, w, pitch;       // this level
    int ch, cw, cpitch;    // coarser level (flow source), ch == 0 -> zero initial flow; cpitch in float2
    int sh, sw;            // full-resolution source
    int spitch0, spitch1;  // source row pitches, bytes
    int fpitch;            // pitch of `flow`, float2 units
    int resize_mode;       // 0 identity, 1 exact 2x (2x2 mean), 2 bilinear
    double sx, sy;         // src/dst ratio of the pyramid resize
    double fsx, fsy;       // coarse/this ratio of the flow upsample
    const uint8_t *src0, *src1;
    float *I;
    float *R;
    float *Ma, *Mb;
    const float2 *cflow;
    float2 *flow;
};

struct PolyConst {
    int n;
    float g[NSOF_MAX_POLY_N + 1], xg[NSOF_MAX_POLY_N + 1], xxg[NSOF_MAX_POLY_N + 1];
    float ig11, ig03, ig33, ig55;
};

// ---------------------------------------------------------------------------------------------
// S2-b: I_k = resize(GaussianBlur(f32(img), ksize, sigma), (w_k, h_k), INTER_LINEAR)   (A.2, A.3)
// Always evaluated from the full-resolution uint8 source, only at the blurred samples the resize
// touches.  gk[0] is the centre tap, gk[i] the tap at distance i (symmetric kernel, radius r).
// ---------------------------------------------------------------------------------------------
NSOF_DEV float fb_blur_row(const uint8_t *NSOF_RESTRICT row, int X, int w, const float *NSOF_RESTRICT gk, int r)
{
    float acc = gk[0] * (float)NSOF_LDG(row + X);
    if (X - r >= 0 && X + r < w) {
        for (int i = 1; i <= r; i++)
            acc = acc + gk[i] * ((float)NSOF_LDG(row + X - i) + (float)NSOF_LDG(row + X + i));
    } else {
        for (int i = 1; i <= r; i++)
            acc = acc + gk[i] * ((float)NSOF_LDG(row + nsof_reflect101(X - i, w)) +
                                 (float)NSOF_LDG(row + nsof_reflect101(X + i, w)));
    }
    return acc;
}

NSOF_DEV float fb_blur_at(const uint8_t *NSOF_RESTRICT src, int sp, int sh, int sw, int X, int Y,
                          const float *NSOF_RESTRICT gk, int r)
{
    float acc = gk[0] * fb_blur_row(src + (size_t)Y * sp, X, sw, gk, r);
    for (int i = 1; i <= r; i++) {
        float a = fb_blur_row(src + (size_t)nsof_reflect101(Y - i, sh) * sp, X, sw, gk, r);
        float b = fb_blur_row(src + (size_t)nsof_reflect101(Y + i, sh) * sp, X, sw, gk, r);
        acc = acc + gk[i] * (a + b);
    }
    return acc;
}

// cv::resize INTER_LINEAR source coordinate: fx = (float)((d + 0.5) * ratio - 0.5), clamped (A.3)
NSOF_DEV void fb_lin_coord(int d, double ratio, int len, int *i0, int *i1, float *a)
{
    float f = (float)(((double)d + 0.5) * ratio - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= len - 1) { s = len - 1; f = 0.f; }
    *i0 = s;
    *i1 = nsof_imin(s + 1, len - 1);
    *a = f;
}

#define FB_PYR_TW 32
#define FB_PYR_TH 8

NSOF_GLOBAL k_fb_pyr(const FbJob *NSOF_RESTRICT jobs, const float *NSOF_RESTRICT gk, int r)
{
    const FbJob J = jobs[NSOF_BY];
    const int tiles_x = (J.w + FB_PYR_TW - 1) / FB_PYR_TW, tiles_y = (J.h + FB_PYR_TH - 1) / FB_PYR_TH;
    if (NSOF_BX >= tiles_x * tiles_y) return;
    const int img = NSOF_BZ;
    const uint8_t *src = img ? J.src1 : J.src0;
    const int sp = img ? J.spitch1 : J.spitch0;
    float *dst = J.I + (size_t)img * J.pitch * J.h;
    const int tx0 = (NSOF_BX % tiles_x) * FB_PYR_TW, ty0 = (NSOF_BX / tiles_x) * FB_PYR_TH;
    for (int i = NSOF_TID; i < FB_PYR_TW * FB_PYR_TH; i += NSOF_NT) {
        const int x = tx0 + i % FB_PYR_TW, y = ty0 + i / FB_PYR_TW;
        if (x >= J.w || y >= J.h) continue;
        float v;
        if (J.resize_mode == 0) {
            v = fb_blur_at(src, sp, J.sh, J.sw, x, y, gk, r);
        } else if (J.resize_mode == 1) {
            float s00 = fb_blur_at(src, sp, J.sh, J.sw, 2 * x, 2 * y, gk, r);
            float s01 = fb_blur_at(src, sp, J.sh, J.sw, 2 * x + 1, 2 * y, gk, r);
            float s10 = fb_blur_at(src, sp, J.sh, J.sw, 2 * x, 2 * y + 1, gk, r);
            float s11 = fb_blur_at(src, sp, J.sh, J.sw, 2 * x + 1, 2 * y + 1, gk, r);
            v = (s00 + s01 + s10 + s11) * 0.25f;
        } else {
            int x0, x1, y0, y1;
            float ax, ay;
            fb_lin_coord(x, J.sx, J.sw, &x0, &x1, &ax);
            fb_lin_coord(y, J.sy, J.sh, &y0, &y1, &ay);
            float s00 = fb_blur_at(src, sp, J.sh, J.sw, x0, y0, gk, r);
            float s01 = fb_blur_at(src, sp, J.sh, J.sw, x1, y0, gk, r);
            float s10 = fb_blur_at(src, sp, J.sh, J.sw, x0, y1, gk, r);
            float s11 = fb_blur_at(src, sp, J.sh, J.sw, x1, y1, gk, r);
            float top = s00 * (1.f - ax) + s01 * ax;
            float bot = s10 * (1.f - ax) + s11 * ax;
            v = top * (1.f - ay) + bot * ay;
        }
        dst[(size_t)y * J.pitch + x] = v;
    }
}

// ---------------------------------------------------------------------------------------------
// S2-c: FarnebackPolyExp (A.5).  Tile TW x TH of R from an I tile with an n-pixel replicate halo
// staged in shared memory: vertical pass -> 3 row moments in shared memory -> horizontal pass.
// dynamic smem: (TH+2n)*(TW+2n) + 3*TH*(TW+2n) floats.
// ---------------------------------------------------------------------------------------------
#define FB_PE_TW 32
#define FB_PE_TH 32

NSOF_GLOBAL k_fb_polyexp(const FbJob *NSOF_RESTRICT jobs, PolyConst pc)
{
    const FbJob J = jobs[NSOF_BY];
    const int tiles_x = (J.w + FB_PE_TW - 1) / FB_PE_TW, tiles_y = (J.h + FB_PE_TH - 1) / FB_PE_TH;
    if (NSOF_BX >= tiles_x * tiles_y) return;
    const int img = NSOF_BZ;
    const size_t plane = (size_t)J.pitch * J.h;
    const float *I = J.I + img * plane;
    float *R = J.R + (size_t)img * 5 * plane;
    const int tx0 = (NSOF_BX % tiles_x) * FB_PE_TW, ty0 = (NSOF_BX / tiles_x) * FB_PE_TH;
    const int n = pc.n, IW = FB_PE_TW + 2 * n, IH = FB_PE_TH + 2 * n;
    NSOF_DYN_SMEM(smem_raw);
    float *sI = (float *)smem_raw;
    float *sr = sI + IW * IH;
    const int SRP = FB_PE_TH * IW;  // one moment plane

    for (int i = NSOF_TID; i < IW * IH; i += NSOF_NT) {
        const int lx = i % IW, ly = i / IW;
        const int gx = nsof_clampi(tx0 - n + lx, 0, J.w - 1), gy = nsof_clampi(ty0 - n + ly, 0, J.h - 1);
        sI[i] = NSOF_LDG(I + (size_t)gy * J.pitch + gx);
    }
    NSOF_SYNC();
    for (int i = NSOF_TID; i < IW * FB_PE_TH; i += NSOF_NT) {
        const int lx = i % IW, ly = i / IW;
        const float *c = sI + (ly + n) * IW + lx;
        float r0 = c[0] * pc.g[0], r1 = 0.f, r2 = 0.f;
        for (int k = 1; k <= n; k++) {
            const float a = c[-k * IW], b = c[k * IW];
            const float p = a + b;
            r0 = r0 + pc.g[k] * p;
            r1 = r1 + pc.xg[k] * (b - a);
            r2 = r2 + pc.xxg[k] * p;
        }
        sr[i] = r0;
        sr[SRP + i] = r1;
        sr[2 * SRP + i] = r2;
    }
    NSOF_SYNC();
    for (int i = NSOF_TID; i < FB_PE_TW * FB_PE_TH; i += NSOF_NT) {
        const int lx = i % FB_PE_TW, ly = i / FB_PE_TW;
        const int x = tx0 + lx, y = ty0 + ly;
        if (x >= J.w || y >= J.h) continue;
        const float *q0 = sr + ly * IW + lx + n, *q1 = q0 + SRP, *q2 = q1 + SRP;
        const float g0 = pc.g[0];
        float b1 = q0[0] * g0, b2 = 0.f, b3 = q1[0] * g0, b4 = 0.f, b5 = q2[0] * g0, b6 = 0.f;
        for (int k = 1; k <= n; k++) {
            const float gk = pc.g[k], xgk = pc.xg[k], xxgk = pc.xxg[k];
            const float p0 = q0[k], m0 = q0[-k], p1 = q1[k], m1 = q1[-k], p2 = q2[k], m2 = q2[-k];
            const float tg = p0 + m0;
            b1 += tg * gk;
            b4 += tg * xxgk;
            b2 += (p0 - m0) * xgk;
            b3 += (p1 + m1) * gk;
            b6 += (p1 - m1) * xgk;
            b5 += (p2 + m2) * gk;
        }
        const size_t o = (size_t)y * J.pitch + x;
        R[o] = b3 * pc.ig11;
        R[plane + o] = b2 * pc.ig11;
        R[2 * plane + o] = b1 * pc.ig03 + b5 * pc.ig33;
        R[3 * plane + o] = b1 * pc.ig03 + b4 * pc.ig33;
        R[4 * plane + o] = b6 * pc.ig55;
    }
}

// ---------------------------------------------------------------------------------------------
// S2-d: FarnebackUpdateMatrices for one pixel (A.6).  r0v = R0 at (x,y); R1 is gathered
// bilinearly at (x+dx, y+dy) through the read-only path (data-dependent: no TMA tiling).
// ---------------------------------------------------------------------------------------------
NSOF_DEV void fb_um_pixel(const float *NSOF_RESTRICT R1, size_t plane, int pitch, int h, int w, int x, int y,
                          float dx, float dy, const float r0v[5], float out[5])
{
    float fx = (float)x + dx, fy = (float)y + dy;
    const int x1 = (int)floorf(fx), y1 = (int)floorf(fy);
    float r2, r3, r4, r5, r6;
    fx -= (float)x1;
    fy -= (float)y1;
    if ((unsigned)x1 < (unsigned)(w - 1) && (unsigned)y1 < (unsigned)(h - 1)) {
        const float a00 = (1.f - fx) * (1.f - fy), a01 = fx * (1.f - fy), a10 = (1.f - fx) * fy, a11 = fx * fy;
        const float *p = R1 + (size_t)y1 * pitch + x1;
        r2 = a00 * NSOF_LDG(p) + a01 * NSOF_LDG(p + 1) + a10 * NSOF_LDG(p + pitch) + a11 * NSOF_LDG(p + pitch + 1);
        p += plane;
        r3 = a00 * NSOF_LDG(p) + a01 * NSOF_LDG(p + 1) + a10 * NSOF_LDG(p + pitch) + a11 * NSOF_LDG(p + pitch + 1);
        p += plane;
        r4 = a00 * NSOF_LDG(p) + a01 * NSOF_LDG(p + 1) + a10 * NSOF_LDG(p + pitch) + a11 * NSOF_LDG(p + pitch + 1);
        p += plane;
        r5 = a00 * NSOF_LDG(p) + a01 * NSOF_LDG(p + 1) + a10 * NSOF_LDG(p + pitch) + a11 * NSOF_LDG(p + pitch + 1);
        p += plane;
        r6 = a00 * NSOF_LDG(p) + a01 * NSOF_LDG(p + 1) + a10 * NSOF_LDG(p + pitch) + a11 * NSOF_LDG(p + pitch + 1);
        r4 = (r0v[2] + r4) * 0.5f;
        r5 = (r0v[3] + r5) * 0.5f;
        r6 = (r0v[4] + r6) * 0.25f;
    } else {
        r2 = r3 = 0.f;
        r4 = r0v[2];
        r5 = r0v[3];
        r6 = r0v[4] * 0.5f;
    }
    r2 = (r0v[0] - r2) * 0.5f;
    r3 = (r0v[1] - r3) * 0.5f;
    r2 += r4 * dy + r6 * dx;
    r3 += r6 * dy + r5 * dx;
    if ((unsigned)(x - 5) >= (unsigned)(w - 10) || (unsigned)(y - 5) >= (unsigned)(h - 10)) {
        // border attenuation table of OpenCV: 0.14, 0.14, 0.4472, 0.4472, 0.4472
        const int xr = w - 1 - x, yb = h - 1 - y;
        float scale = (x < 2 ? 0.14f : (x < 5 ? 0.4472f : 1.f));
        scale *= (xr < 2 ? 0.14f : (xr < 5 ? 0.4472f : 1.f));
        scale *= (y < 2 ? 0.14f : (y < 5 ? 0.4472f : 1.f));
        scale *= (yb < 2 ? 0.14f : (yb < 5 ? 0.4472f : 1.f));
        r2 *= scale; r3 *= scale; r4 *= scale; r5 *= scale; r6 *= scale;
    }
    out[0] = r4 * r4 + r6 * r6;
    out[1] = (r4 + r5) * r6;
    out[2] = r5 * r5 + r6 * r6;
    out[3] = r4 * r2 + r6 * r3;
    out[4] = r6 * r2 + r5 * r3;
}

// First UpdateMatrices of a level; the initial flow is zero (coarsest level) or the coarser
// level's flow upsampled on the fly (S2-f: resize INTER_LINEAR, times f32(1/pyr_scale)); the
// upsampled field itself is never written (UpdateFlow_Blur overwrites flow entirely).
#define FB_UM_TW 64
#define FB_UM_TH 4

NSOF_GLOBAL k_fb_um_first(const FbJob *NSOF_RESTRICT jobs, float flow_mul)
{
    const FbJob J = jobs[NSOF_BY];
    const int tiles_x = (J.w + FB_UM_TW - 1) / FB_UM_TW, tiles_y = (J.h + FB_UM_TH - 1) / FB_UM_TH;
    if (NSOF_BX >= tiles_x * tiles_y) return;
    const size_t plane = (size_t)J.pitch * J.h;
    const int tx0 = (NSOF_BX % tiles_x) * FB_UM_TW, ty0 = (NSOF_BX / tiles_x) * FB_UM_TH;
    for (int i = NSOF_TID; i < FB_UM_TW * FB_UM_TH; i += NSOF_NT) {
        const int x = tx0 + i % FB_UM_TW, y = ty0 + i / FB_UM_TW;
        if (x >= J.w || y >= J.h) continue;
        float dx = 0.f, dy = 0.f;
        if (J.ch > 0) {
            int x0, x1, y0, y1;
            float ax, ay;
            fb_lin_coord(x, J.fsx, J.cw, &x0, &x1, &ax);
            fb_lin_coord(y, J.fsy, J.ch, &y0, &y1, &ay);
            const float2 f00 = NSOF_LDG(J.cflow + (size_t)y0 * J.cpitch + x0), f01 = NSOF_LDG(J.cflow + (size_t)y0 * J.cpitch + x1);
            const float2 f10 = NSOF_LDG(J.cflow + (size_t)y1 * J.cpitch + x0), f11 = NSOF_LDG(J.cflow + (size_t)y1 * J.cpitch + x1);
            const float a0 = 1.f - ax, b0 = 1.f - ay;
            dx = ((f00.x * a0 + f01.x * ax) * b0 + (f10.x * a0 + f11.x * ax) * ay) * flow_mul;
            dy = ((f00.y * a0 + f01.y * ax) * b0 + (f10.y * a0 + f11.y * ax) * ay) * flow_mul;
        }
        const size_t o = (size_t)y * J.pitch + x;
        float r0v[5], m[5];
        for (int c = 0; c < 5; c++) r0v[c] = NSOF_LDG(J.R + c * plane + o);
        fb_um_pixel(J.R + 5 * plane, plane, J.pitch, J.h, J.w, x, y, dx, dy, r0v, m);
        for (int c = 0; c < 5; c++) J.Ma[c * plane + o] = m[c];
    }
}

// Level with zero iterations: flow is just the (scaled) upsampled coarse flow, or zero.
NSOF_GLOBAL k_fb_flow_init(const FbJob *NSOF_RESTRICT jobs, float flow_mul)
{
    const FbJob J = jobs[NSOF_BY];
    const int tiles_x = (J.w + FB_UM_TW - 1) / FB_UM_TW, tiles_y = (J.h + FB_UM_TH - 1) / FB_UM_TH;
    if (NSOF_BX >= tiles_x * tiles_y) return;
    const int tx0 = (NSOF_BX % tiles_x) * FB_UM_TW, ty0 = (NSOF_BX / tiles_x) * FB_UM_TH;
    for (int i = NSOF_TID; i < FB_UM_TW * FB_UM_TH; i += NSOF_NT) {
        const int x = tx0 + i % FB_UM_TW, y = ty0 + i / FB_UM_TW;
        if (x >= J.w || y >= J.h) continue;
        float dx = 0.f, dy = 0.f;
        if (J.ch > 0) {
            int x0, x1, y0, y1;
            float ax, ay;
            fb_lin_coord(x, J.fsx, J.cw, &x0, &x1, &ax);
            fb_lin_coord(y, J.fsy, J.ch, &y0, &y1, &ay);
            const float2 f00 = J.cflow[(size_t)y0 * J.cpitch + x0], f01 = J.cflow[(size_t)y0 * J.cpitch + x1];
            const float2 f10 = J.cflow[(size_t)y1 * J.cpitch + x0], f11 = J.cflow[(size_t)y1 * J.cpitch + x1];
            const float a0 = 1.f - ax, b0 = 1.f - ay;
            dx = ((f00.x * a0 + f01.x * ax) * b0 + (f10.x * a0 + f11.x * ax) * ay) * flow_mul;
            dy = ((f00.y * a0 + f01.y * ax) * b0 + (f10.y * a0 + f11.y * ax) * ay) * flow_mul;
        }
        J.flow[(size_t)y * J.fpitch + x] = make_float2(dx, dy);
    }
}

// ---------------------------------------------------------------------------------------------
// S2-e (+ S2-d fused): FarnebackUpdateFlow_Blur, flags = 0 (A.7).
//   phase 1  M tile (5 planes) with an m = winsize/2 replicate halo -> shared memory
//   phase 2  vertical (2m+1)-row window sums, one thread per (plane, column), sliding sum kept in
//            double (exact, like OpenCV's running double sums), written back in place as float
//   phase 3  4 pixels per work item: horizontal (2m+1) sums, 2x2 solve, then either
//              last iteration : write flow
//              otherwise      : UpdateMatrices with the new flow in registers -> M_out
//            so flow never touches HBM between iterations.
// dynamic smem: 5 * (TH+2m) * PW floats, PW = roundup4(TW+2m).   TW % 4 == 0, m >= 1.
// ---------------------------------------------------------------------------------------------
NSOF_GLOBAL k_fb_blur_solve(const FbJob *NSOF_RESTRICT jobs, int TW, int TH, int m, float scale, int src_is_a,
                            int last)
{
    const FbJob J = jobs[NSOF_BY];
    const int tiles_x = (J.w + TW - 1) / TW, tiles_y = (J.h + TH - 1) / TH;
    if (NSOF_BX >= tiles_x * tiles_y) return;
    const size_t plane = (size_t)J.pitch * J.h;
    const float *Min = src_is_a ? J.Ma : J.Mb;
    float *Mout = src_is_a ? J.Mb : J.Ma;
    const int tx0 = (NSOF_BX % tiles_x) * TW, ty0 = (NSOF_BX / tiles_x) * TH;
    const int LW = TW + 2 * m, LH = TH + 2 * m, PW = (LW + 3) & ~3;
    const int SP = LH * PW;  // smem plane
    NSOF_DYN_SMEM(smem_raw);
    float *sM = (float *)smem_raw;

    for (int i = NSOF_TID; i < 5 * LH * LW; i += NSOF_NT) {
        const int lx = i % LW, t = i / LW, ly = t % LH, c = t / LH;
        const int gx = nsof_clampi(tx0 - m + lx, 0, J.w - 1), gy = nsof_clampi(ty0 - m + ly, 0, J.h - 1);
        sM[c * SP + ly * PW + lx] = NSOF_LDG(Min + c * plane + (size_t)gy * J.pitch + gx);
    }
    NSOF_SYNC();
    for (int i = NSOF_TID; i < 5 * LW; i += NSOF_NT) {
        float *col = sM + (i / LW) * SP + (i % LW);
        double s = 0.;
        for (int j = 0; j <= 2 * m; j++) s += (double)col[j * PW];
        for (int y = 0; y < TH; y++) {
            const float old = col[y * PW];
            col[y * PW] = (float)s;
            if (y + 1 < TH) s += (double)col[(y + 2 * m + 1) * PW] - (double)old;
        }
    }
    NSOF_SYNC();
    const int gw = TW / 4;
    for (int i = NSOF_TID; i < gw * TH; i += NSOF_NT) {
        const int lx = (i % gw) * 4, ly = i / gw;
        const int x = tx0 + lx, y = ty0 + ly;
        if (x >= J.w || y >= J.h) continue;
        float s[5][4];
        for (int c = 0; c < 5; c++) {
            const float *v = sM + c * SP + ly * PW + lx;
            // common part: taps 3 .. 2m (shared by the four windows; empty when m == 1), then 3 private taps each
            float common = 0.f;
            for (int j = 3; j <= 2 * m; j++) common += v[j];
            const float v0 = v[0], v1 = v[1], v2 = v[2];
            const float w1 = v[2 * m + 1], w2 = v[2 * m + 2], w3 = v[2 * m + 3];
            s[c][0] = common + ((v0 + v1) + v2);
            s[c][1] = common + ((v1 + v2) + w1);
            s[c][2] = common + ((v2 + w1) + w2);
            s[c][3] = common + ((w1 + w2) + w3);
        }
        float fxv[4], fyv[4];
        for (int q = 0; q < 4; q++) {
            const float g11 = s[0][q] * scale, g12 = s[1][q] * scale, g22 = s[2][q] * scale;
            const float h1 = s[3][q] * scale, h2 = s[4][q] * scale;
            const float idet = 1.f / (g11 * g22 - g12 * g12 + 1e-3f);
            fxv[q] = (g11 * h2 - g12 * h1) * idet;
            fyv[q] = (g22 * h1 - g12 * h2) * idet;
        }
        if (last) {
            for (int q = 0; q < 4; q++)
                if (x + q < J.w) J.flow[(size_t)y * J.fpitch + x + q] = make_float2(fxv[q], fyv[q]);
        } else {
            const size_t o = (size_t)y * J.pitch + x;
            float mo[5][4];
            for (int q = 0; q < 4; q++) {
                float r0v[5], mm[5];
                if (x + q < J.w) {
                    for (int c = 0; c < 5; c++) r0v[c] = NSOF_LDG(J.R + c * plane + o + q);
                    fb_um_pixel(J.R + 5 * plane, plane, J.pitch, J.h, J.w, x + q, y, fxv[q], fyv[q], r0v, mm);
                } else {
                    for (int c = 0; c < 5; c++) mm[c] = 0.f;
                }
                for (int c = 0; c < 5; c++) mo[c][q] = mm[c];
            }
            for (int c = 0; c < 5; c++)
                *(float4 *)(Mout + c * plane + o) = make_float4(mo[c][0], mo[c][1], mo[c][2], mo[c][3]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// S2-e (+ S2-d fused), tuned variant for a compile-time window radius M and tile TW x TH.
// Same arithmetic, same order of operations per pixel as k_fb_blur_solve (results are bit-identical);
// what changes is how the work is laid out:
//   phase 1  the M tile + halo (5 planes x (TH+2M) rows) arrives by bulk asynchronous row copies
//            (cp.async.bulk global->shared on an mbarrier, SASS UBLKCP): no per-element index arithmetic,
//            no register staging, every row in flight at once.  Rows are 16-byte aligned runs starting at
//            tx0 - ML (ML = M rounded up to 4); replicate borders are patched in shared memory afterwards
//            (border tiles only).
//   phase 2  vertical running sums in double, one thread per (plane, column)
//   phase 3  4 pixels per thread: aligned float4 shared-memory reads of the 2M+4 window columns,
//            float4 loads of R0, float4 stores of M'
// dynamic smem: 16 B (mbarrier) + 5 * (TH+2M) * (TW+2ML) floats.
// ---------------------------------------------------------------------------------------------
template <int M, int TW, int TH>
NSOF_GLOBAL NSOF_LAUNCH_BOUNDS(256, 3) k_fb_blur_solve_t(const FbJob *NSOF_RESTRICT jobs, float scale, int src_is_a, int last)
{
    constexpr int ML = (M + 3) & ~3, PW = TW + 2 * ML, LH = TH + 2 * M, LW = TW + 2 * M, SP = LH * PW;
    constexpr int NV = (2 * M + 4 + (ML - M) + 3) / 4;  // float4s covering the window columns of 4 outputs
    const FbJob J = jobs[NSOF_BY];
    const int tiles_x = (J.w + TW - 1) / TW, tiles_y = (J.h + TH - 1) / TH;
    if (NSOF_BX >= tiles_x * tiles_y) return;
    const size_t plane = (size_t)J.pitch * J.h;
    const float *Min = src_is_a ? J.Ma : J.Mb;
    float *Mout = src_is_a ? J.Mb : J.Ma;
    const int tx0 = (NSOF_BX % tiles_x) * TW, ty0 = (NSOF_BX / tiles_x) * TH;
    NSOF_DYN_SMEM(smem_raw);
    nsof_mbar_t *bar = (nsof_mbar_t *)smem_raw;
    float *sM = (float *)(smem_raw + 16);

    // ---- phase 1
    const int gx0 = tx0 - ML;                                         // global x of shared column 0
    const int cx0 = nsof_imax(gx0, 0), cx1 = nsof_imin(tx0 + TW + ML, J.pitch);  // copied run, multiples of 4
    const unsigned row_bytes = (unsigned)(cx1 - cx0) * 4u;
    if (NSOF_TID == 0) nsof_mbar_init(bar, 1);
    NSOF_SYNC();
    if (NSOF_TID < NSOF_WARP) {
        if (NSOF_TID == 0) nsof_mbar_expect_tx(bar, row_bytes * (unsigned)(5 * LH));
        NSOF_SYNCWARP();
        for (int r = NSOF_TID; r < 5 * LH; r += NSOF_WARP) {
            const int c = r / LH, ly = r - c * LH;
            const int gy = nsof_clampi(ty0 - M + ly, 0, J.h - 1);
            nsof_bulk_g2s(sM + c * SP + ly * PW + (cx0 - gx0), Min + c * plane + (size_t)gy * J.pitch + cx0, row_bytes, bar);
        }
    }
    nsof_mbar_wait(bar, 0);
    if (tx0 - M < 0 || tx0 + TW + M > J.w) {  // replicate the first / last image column into the halo
        const int lane = NSOF_LANE, warp = NSOF_TID / NSOF_WARP, nw = nsof_imax(NSOF_NT / NSOF_WARP, 1);
        const int s_first = -gx0, s_last = J.w - 1 - gx0;  // shared columns of image x = 0 and x = w - 1
        for (int r = warp; r < 5 * LH; r += nw) {
            float *row = sM + (r / LH) * SP + (r % LH) * PW;
            if (tx0 - M < 0) {
                const float v = row[s_first];
                for (int sx = ML - M + lane; sx < s_first; sx += NSOF_WARP) row[sx] = v;
            }
            if (tx0 + TW + M > J.w) {
                const float v = row[s_last];
                NSOF_SYNCWARP();
                for (int sx = s_last + 1 + lane; sx < ML + TW + M; sx += NSOF_WARP) row[sx] = v;
            }
        }
    }
    NSOF_SYNC();
    // ---- phase 2
    for (int col = NSOF_TID; col < 5 * LW; col += NSOF_NT) {
        const int c = col / LW;
        float *p = sM + c * SP + (ML - M) + (col - c * LW);
        double s = 0.;
#pragma unroll
        for (int j = 0; j <= 2 * M; j++) s += (double)p[j * PW];
#pragma unroll 4
        for (int y = 0; y < TH; y++) {
            const float old = p[y * PW];
            p[y * PW] = (float)s;
            if (y + 1 < TH) s += (double)p[(y + 2 * M + 1) * PW] - (double)old;
        }
    }
    NSOF_SYNC();
    // ---- phase 3
    constexpr int GW = TW / 4;
    for (int i = NSOF_TID; i < GW * TH; i += NSOF_NT) {
        const int lx = (i % GW) * 4, ly = i / GW;
        const int x = tx0 + lx, y = ty0 + ly;
        if (x >= J.w || y >= J.h) continue;
        float s[5][4];
#pragma unroll
        for (int c = 0; c < 5; c++) {
            float v[NV * 4];
            const float4 *q = (const float4 *)(sM + c * SP + ly * PW + lx);
#pragma unroll
            for (int k = 0; k < NV; k++) {
                const float4 t = q[k];
                v[4 * k] = t.x; v[4 * k + 1] = t.y; v[4 * k + 2] = t.z; v[4 * k + 3] = t.w;
            }
            constexpr int B = ML - M;  // window of output 0 starts at v[B]
            float common = 0.f;
#pragma unroll
            for (int j = 3; j <= 2 * M; j++) common += v[B + j];
            const float v0 = v[B], v1 = v[B + 1], v2 = v[B + 2];
            const float w1 = v[B + 2 * M + 1], w2 = v[B + 2 * M + 2], w3 = v[B + 2 * M + 3];
            s[c][0] = common + ((v0 + v1) + v2);
            s[c][1] = common + ((v1 + v2) + w1);
            s[c][2] = common + ((v2 + w1) + w2);
            s[c][3] = common + ((w1 + w2) + w3);
        }
        float fxv[4], fyv[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float g11 = s[0][q] * scale, g12 = s[1][q] * scale, g22 = s[2][q] * scale;
            const float h1 = s[3][q] * scale, h2 = s[4][q] * scale;
            const float idet = 1.f / (g11 * g22 - g12 * g12 + 1e-3f);
            fxv[q] = (g11 * h2 - g12 * h1) * idet;
            fyv[q] = (g22 * h1 - g12 * h2) * idet;
        }
        if (last) {
            float2 *fo = J.flow + (size_t)y * J.fpitch + x;
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (x + q < J.w) fo[q] = make_float2(fxv[q], fyv[q]);
        } else {
            const size_t o = (size_t)y * J.pitch + x;
            float r0[5][4];
#pragma unroll
            for (int c = 0; c < 5; c++) {
                const float4 t = NSOF_LDG((const float4 *)(J.R + c * plane + o));
                r0[c][0] = t.x; r0[c][1] = t.y; r0[c][2] = t.z; r0[c][3] = t.w;
            }
            float mo[5][4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                float r0v[5], mm[5];
                if (x + q < J.w) {
#pragma unroll
                    for (int c = 0; c < 5; c++) r0v[c] = r0[c][q];
                    fb_um_pixel(J.R + 5 * plane, plane, J.pitch, J.h, J.w, x + q, y, fxv[q], fyv[q], r0v, mm);
                } else {
#pragma unroll
                    for (int c = 0; c < 5; c++) mm[c] = 0.f;
                }
#pragma unroll
                for (int c = 0; c < 5; c++) mo[c][q] = mm[c];
            }
#pragma unroll
            for (int c = 0; c < 5; c++)
                *(float4 *)(Mout + c * plane + o) = make_float4(mo[c][0], mo[c][1], mo[c][2], mo[c][3]);
        }
    }
}

#define FB_BS7_TW 64
#define FB_BS7_TH 32
// a plain function-pointer name: template commas do not survive nested launch macros
static void (*const FB_BLUR_SOLVE_M7)(const FbJob *, float, int, int) = k_fb_blur_solve_t<7, FB_BS7_TW, FB_BS7_TH>;
#define FB_BS7_SMEM (16 + 5 * (FB_BS7_TH + 14) * (FB_BS7_TW + 16) * sizeof(float))
