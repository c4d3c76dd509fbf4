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
// All kernels take a device table of per-item jobs and a FLAT TILE LIST (one uint2 = (item, tile of that item) per
// CTA, built on the device by k_fb_build_tiles from per-item runs) and run as grid (n_tiles_of_all_items[, 1, 2]):
// ragged ROI crops batch into exactly as many CTAs as there are tiles.  (Round 1 ran grid (max tiles, items) with
// early exits: one whole-frame item among 250 small crops made 90 % of the CTAs of every launch empty.)
#pragma once
#include <type_traits>
#include "nsof_rt.h"
#include <algorithm>

#define NSOF_MAX_POLY_N 31
#define FB_BS7_TW 64
#ifndef FB_BS7_TH
#define FB_BS7_TH 32
#endif


struct alignas(16) FbJob {
    int h, w, pitch;       // this level
    int ch, cw, cpitch;    // coarser level (flow source), ch == 0 -> zero initial flow; cpitch in float2
    int sh, sw;            // full-resolution source
    int spitch0, spitch1;  // source row pitches, bytes
    int fpitch;            // pitch of `flow`, float2 units
    int resize_mode;       // 0 identity, 1 exact 2x (2x2 mean), 2 bilinear
    int fast;              // 0: generic k_fb_pyr + k_fb_polyexp; 1 / 2: fused k_fb_pyr_polyexp_t<MODE 0 / 1>
    int share0;            // R[0] is another item's R[1] (shared frame): only image 1 of this item is expanded
    // Row window (spatial tiling across GPUs, nsof_fb_tiled_*): the kernels compute rows [y_lo, y_hi) of this level
    // (the generic pyramid kernels rows [iy_lo, iy_hi) of I) but keep addressing, borders and resize geometry of
    // the whole h x w level; rows outside the window are read where a stencil / gather reaches them (exchanged
    // halo rows).  Untiled: [0, h).
    int y_lo, y_hi, iy_lo, iy_hi;
    int gy_lo, gy_hi;      // rows of R1 that are valid on this rank (own + received halo); checked when guard != 0
    int *guard;            // device flag set when an UpdateMatrices gather leaves [gy_lo, gy_hi) (NULL: unchecked)
    double sx, sy;         // src/dst ratio of the pyramid resize
    double fsx, fsy;       // coarse/this ratio of the flow upsample
    const uint8_t *src0, *src1;
    float *I;
    // polynomial expansions of the two images, 5 planes each.  R[0] of an item whose `prev` crop IS another item's
    // `next` crop (consecutive pairs of a clip under an unchanged ROI rectangle, or whole frames) points at that item's
    // R[1]: the expansion of a shared frame is computed once per batch (the tile lists carry (item, image) entries).
    float *R[2];
    float *Ma, *Mb;
    const float2 *cflow;
    float2 *flow;
    // OPTFLOW_USE_INITIAL_FLOW, coarsest level of the item only.  init: 0 no; 1 the level's flow buffer holds
    // resize(flow0, INTER_AREA) * init_mul (k_fb_area_init); 2 level 0 is the coarsest: the caller's flow, in place.
    // cflow / cpitch then name the level's OWN flow buffer, read as it is.
    int init;
    float init_mul;
    const float2 *flow0;  // the caller's full-resolution flow (sh x sw, f0pitch in float2)
    int f0pitch;
};

// tile-list entry .x = item index (low 24 bits) | image index << 24 (pyramid / PolyExp lists; 0 elsewhere)
#define FB_TL_ITEM 0xffffffu
#define FB_TL_IMG_SHIFT 24

// One run of a flat tile list: `count` consecutive entries (item, 0), (item, 1), ... starting at entry `off`.
struct FbTileRun {
    unsigned off, item, count, pad;
};

NSOF_GLOBAL k_fb_build_tiles(const FbTileRun *NSOF_RESTRICT runs, uint2 *NSOF_RESTRICT tiles)
{
    const FbTileRun r = runs[NSOF_BX];
    for (unsigned t = (unsigned)NSOF_TID; t < r.count; t += (unsigned)NSOF_NT) tiles[r.off + t] = make_uint2(r.item, t);
}

struct PolyConst {
    int n;
    float g[NSOF_MAX_POLY_N + 1], xg[NSOF_MAX_POLY_N + 1], xxg[NSOF_MAX_POLY_N + 1];
    float ig11, ig03, ig33, ig55;
    // taps 0 .. 7 once more as (t, t) pairs: operands of the packed FP32 instructions of k_fb_pyr_polyexp_t
    float2 g2[8], xg2[8], xxg2[8];
};

// ---------------------------------------------------------------------------------------------
// S2-b: I_k = resize(GaussianBlur(f32(img), ksize, sigma), (w_k, h_k), INTER_LINEAR)   (A.2, A.3)
// Always evaluated from the full-resolution uint8 source, only at the blurred samples the resize
// touches.  gk[0] is the centre tap, gk[i] the tap at distance i (symmetric kernel, radius r).
// ---------------------------------------------------------------------------------------------
// T = uint8_t (the reference's grey frames) or float (cv2 accepts CV_32F frames too: convertTo(CV_32F) is then the identity)
template <typename T>
NSOF_DEV float fb_blur_row(const T *NSOF_RESTRICT row, int X, int w, const float *NSOF_RESTRICT gk, int r)
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

// sp: row pitch in BYTES
template <typename T>
NSOF_DEV float fb_blur_at(const T *NSOF_RESTRICT src, int sp, int sh, int sw, int X, int Y,
                          const float *NSOF_RESTRICT gk, int r)
{
#define FB_ROW(y_) ((const T *)((const unsigned char *)src + (size_t)(y_) * sp))
    float acc = gk[0] * fb_blur_row(FB_ROW(Y), X, sw, gk, r);
    for (int i = 1; i <= r; i++) {
        float a = fb_blur_row(FB_ROW(nsof_reflect101(Y - i, sh)), X, sw, gk, r);
        float b = fb_blur_row(FB_ROW(nsof_reflect101(Y + i, sh)), X, sw, gk, r);
        acc = acc + gk[i] * (a + b);
    }
#undef FB_ROW
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

template <typename T>
NSOF_GLOBAL k_fb_pyr_g(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl, const float *NSOF_RESTRICT gk, int r)
{
    const uint2 te = tl[NSOF_BX];
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + (te.x & FB_TL_ITEM));
    const int tile = (int)te.y;
    const int tiles_x = (J.w + FB_PYR_TW - 1) / FB_PYR_TW;
    const int img = (int)(te.x >> FB_TL_IMG_SHIFT);
    const T *src = (const T *)(img ? J.src1 : J.src0);
    const int sp = img ? J.spitch1 : J.spitch0;
    float *dst = J.I + (size_t)img * J.pitch * J.h;
    const int tx0 = (tile % tiles_x) * FB_PYR_TW, ty0 = J.iy_lo + (tile / tiles_x) * FB_PYR_TH;
    for (int i = NSOF_TID; i < FB_PYR_TW * FB_PYR_TH; i += NSOF_NT) {
        const int x = tx0 + i % FB_PYR_TW, y = ty0 + i / FB_PYR_TW;
        if (x >= J.w || y >= J.iy_hi) continue;
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

static void (*const k_fb_pyr)(const FbJob *, const uint2 *, const float *, int) = k_fb_pyr_g<uint8_t>;
static void (*const k_fb_pyr_f32)(const FbJob *, const uint2 *, const float *, int) = k_fb_pyr_g<float>;

// ---------------------------------------------------------------------------------------------
// S2-c: FarnebackPolyExp (A.5).  Tile TW x TH of R from an I tile with an n-pixel replicate halo
// staged in shared memory: vertical pass -> 3 row moments in shared memory -> horizontal pass.
// dynamic smem: (TH+2n)*(TW+2n) + 3*TH*(TW+2n) floats.
// ---------------------------------------------------------------------------------------------
#define FB_PE_TW 32
#define FB_PE_TH 32

NSOF_GLOBAL k_fb_polyexp(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl, PolyConst pc)
{
    const uint2 te = tl[NSOF_BX];
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + (te.x & FB_TL_ITEM));
    const int tile = (int)te.y;
    const int tiles_x = (J.w + FB_PE_TW - 1) / FB_PE_TW;
    const int img = (int)(te.x >> FB_TL_IMG_SHIFT);
    const size_t plane = (size_t)J.pitch * J.h;
    const float *I = J.I + img * plane;
    float *R = J.R[img];
    const int tx0 = (tile % tiles_x) * FB_PE_TW, ty0 = J.y_lo + (tile / tiles_x) * FB_PE_TH;
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
        if (x >= J.w || y >= J.y_hi) continue;
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
// WIN = 1: R1 is a window of the plane staged in shared memory (see fb_um_pixel2_t): `pitch` floats per row, `plane`
// floats per plane, element (0, 0) = image pixel (wx0, wy0), (ww, wh) the largest in-window cell offsets.
#define FB_UM_LD(p) (WIN ? *(p) : NSOF_LDG(p))
template <bool WIN>
NSOF_DEV void fb_um_pixel_t(const float *NSOF_RESTRICT R1, size_t plane, int pitch, int wx0, int wy0, int ww, int wh, int h,
                            int w, int x, int y, float dx, float dy, const float r0v[5], float out[5])
{
    float fx = (float)x + dx, fy = (float)y + dy;
    const int x1 = (int)floorf(fx), y1 = (int)floorf(fy);
    float r2, r3, r4, r5, r6;
    fx -= (float)x1;
    fy -= (float)y1;
    // The 20 loads are issued unconditionally (from a clamped, always valid address) and the two cases are
    // selected afterwards: without a branch around them the loads of several pixels can be in flight at once.
    const bool inb = (unsigned)x1 < (unsigned)(w - 1) && (unsigned)y1 < (unsigned)(h - 1);
    {
        const float a00 = (1.f - fx) * (1.f - fy), a01 = fx * (1.f - fy), a10 = (1.f - fx) * fy, a11 = fx * fy;
        const float *p;
        if (WIN) {
            const int rx = x1 - wx0, ry = y1 - wy0;
            p = R1 + ((inb && (unsigned)rx <= (unsigned)ww && (unsigned)ry <= (unsigned)wh) ? ry * pitch + rx : 0);
        } else {
            p = R1 + (inb ? (size_t)y1 * pitch + x1 : (size_t)0);
        }
        const float q2 = a00 * FB_UM_LD(p) + a01 * FB_UM_LD(p + 1) + a10 * FB_UM_LD(p + pitch) + a11 * FB_UM_LD(p + pitch + 1);
        p += plane;
        const float q3 = a00 * FB_UM_LD(p) + a01 * FB_UM_LD(p + 1) + a10 * FB_UM_LD(p + pitch) + a11 * FB_UM_LD(p + pitch + 1);
        p += plane;
        const float q4 = a00 * FB_UM_LD(p) + a01 * FB_UM_LD(p + 1) + a10 * FB_UM_LD(p + pitch) + a11 * FB_UM_LD(p + pitch + 1);
        p += plane;
        const float q5 = a00 * FB_UM_LD(p) + a01 * FB_UM_LD(p + 1) + a10 * FB_UM_LD(p + pitch) + a11 * FB_UM_LD(p + pitch + 1);
        p += plane;
        const float q6 = a00 * FB_UM_LD(p) + a01 * FB_UM_LD(p + 1) + a10 * FB_UM_LD(p + pitch) + a11 * FB_UM_LD(p + pitch + 1);
        r2 = inb ? q2 : 0.f;
        r3 = inb ? q3 : 0.f;
        r4 = inb ? (r0v[2] + q4) * 0.5f : r0v[2];
        r5 = inb ? (r0v[3] + q5) * 0.5f : r0v[3];
        r6 = inb ? (r0v[4] + q6) * 0.25f : r0v[4] * 0.5f;
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

NSOF_DEV void fb_um_pixel(const float *NSOF_RESTRICT R1, size_t plane, int pitch, int h, int w, int x, int y,
                          float dx, float dy, const float r0v[5], float out[5])
{
    fb_um_pixel_t<false>(R1, plane, pitch, 0, 0, 0, 0, h, w, x, y, dx, dy, r0v, out);
}

// Two pixels at once with the packed FP32 instructions of Blackwell (FMUL2 / FFMA2 / FADD2): the same operations as
// fb_um_pixel in the same order, the two pixels in the two halves of every operand, so the ~60 FP32 instructions per
// pixel of UpdateMatrices take half the issue slots.  The 2 x 20 gathers stay scalar loads (data-dependent addresses).
NSOF_DEV float nsof_um_border_scale(int x, int y, int w, int h)
{
    if ((unsigned)(x - 5) < (unsigned)(w - 10) && (unsigned)(y - 5) < (unsigned)(h - 10)) return 1.f;
    const int xr = w - 1 - x, yb = h - 1 - y;
    float scale = (x < 2 ? 0.14f : (x < 5 ? 0.4472f : 1.f));
    scale *= (xr < 2 ? 0.14f : (xr < 5 ? 0.4472f : 1.f));
    scale *= (y < 2 ? 0.14f : (y < 5 ? 0.4472f : 1.f));
    scale *= (yb < 2 ? 0.14f : (yb < 5 ? 0.4472f : 1.f));
    return scale;
}

// WIN = 1: R1 is not the image plane but a WINDOW of it staged in shared memory -- `pitch` floats per row, `plane` floats
// per plane, element (0, 0) = image pixel (wx0, wy0), ww x wh the largest in-window (x1, y1) offsets whose right / lower
// neighbours are still inside.  The caller guarantees that every pixel with an in-image bilinear cell lies inside (the
// window is placed on the tile's flow range); the offset test below only keeps the address legal for pixels whose
// values are discarded anyway (NaN flow).  Same loads, same arithmetic, same results as WIN = 0.
template <bool WIN>
NSOF_DEV void fb_um_pixel2_t(const float *NSOF_RESTRICT R1, size_t plane, int pitch, int wx0, int wy0, int ww, int wh, int h,
                             int w, int xa, int ya, int xb, int yb, float2 dx, float2 dy, const float2 r0v[5], float2 out[5])
{
    float2 fx = nsof_add2(make_float2((float)xa, (float)xb), dx), fy = nsof_add2(make_float2((float)ya, (float)yb), dy);
    const int x1a = (int)floorf(fx.x), y1a = (int)floorf(fy.x), x1b = (int)floorf(fx.y), y1b = (int)floorf(fy.y);
    fx = nsof_add2(fx, make_float2(-(float)x1a, -(float)x1b));
    fy = nsof_add2(fy, make_float2(-(float)y1a, -(float)y1b));
    const bool ina = (unsigned)x1a < (unsigned)(w - 1) && (unsigned)y1a < (unsigned)(h - 1);
    const bool inb = (unsigned)x1b < (unsigned)(w - 1) && (unsigned)y1b < (unsigned)(h - 1);
    const float2 one = make_float2(1.f, 1.f), mone = make_float2(-1.f, -1.f);
    const float2 ox = nsof_fma2(fx, mone, one), oy = nsof_fma2(fy, mone, one);  // 1 - fx, 1 - fy (one rounding each)
    const float2 a00 = nsof_mul2(ox, oy), a01 = nsof_mul2(fx, oy), a10 = nsof_mul2(ox, fy), a11 = nsof_mul2(fx, fy);
    const float *pa, *pb;
    if (WIN) {
        const int rxa = x1a - wx0, rya = y1a - wy0, rxb = x1b - wx0, ryb = y1b - wy0;
        const bool wa = ina && (unsigned)rxa <= (unsigned)ww && (unsigned)rya <= (unsigned)wh;
        const bool wb = inb && (unsigned)rxb <= (unsigned)ww && (unsigned)ryb <= (unsigned)wh;
        pa = R1 + (wa ? rya * pitch + rxa : 0);
        pb = R1 + (wb ? ryb * pitch + rxb : 0);
    } else {
        pa = R1 + (ina ? (size_t)y1a * pitch + x1a : (size_t)0);
        pb = R1 + (inb ? (size_t)y1b * pitch + x1b : (size_t)0);
    }
    float2 q[5];
#pragma unroll
    for (int c = 0; c < 5; c++) {
        float2 v00, v01, v10, v11;
        if (WIN) {  // shared memory: plain loads
            v00 = make_float2(pa[0], pb[0]); v01 = make_float2(pa[1], pb[1]);
            v10 = make_float2(pa[pitch], pb[pitch]); v11 = make_float2(pa[pitch + 1], pb[pitch + 1]);
        } else {
            v00 = make_float2(NSOF_LDG(pa), NSOF_LDG(pb)); v01 = make_float2(NSOF_LDG(pa + 1), NSOF_LDG(pb + 1));
            v10 = make_float2(NSOF_LDG(pa + pitch), NSOF_LDG(pb + pitch));
            v11 = make_float2(NSOF_LDG(pa + pitch + 1), NSOF_LDG(pb + pitch + 1));
        }
        q[c] = nsof_fma2(a11, v11, nsof_fma2(a10, v10, nsof_fma2(a01, v01, nsof_mul2(a00, v00))));
        pa += plane;
        pb += plane;
    }
    const float2 half = make_float2(0.5f, 0.5f), mhalf = make_float2(-0.5f, -0.5f), quarter = make_float2(0.25f, 0.25f);
    const float2 t4 = nsof_mul2(nsof_add2(r0v[2], q[2]), half), t5 = nsof_mul2(nsof_add2(r0v[3], q[3]), half);
    const float2 t6 = nsof_mul2(nsof_add2(r0v[4], q[4]), quarter), u6 = nsof_mul2(r0v[4], half);
    float2 r2 = make_float2(ina ? q[0].x : 0.f, inb ? q[0].y : 0.f);
    float2 r3 = make_float2(ina ? q[1].x : 0.f, inb ? q[1].y : 0.f);
    float2 r4 = make_float2(ina ? t4.x : r0v[2].x, inb ? t4.y : r0v[2].y);
    float2 r5 = make_float2(ina ? t5.x : r0v[3].x, inb ? t5.y : r0v[3].y);
    float2 r6 = make_float2(ina ? t6.x : u6.x, inb ? t6.y : u6.y);
    // (r0 - r) * 0.5 == fma(r, -0.5, r0 * 0.5): scaling by a power of two commutes with the rounding
    r2 = nsof_fma2(r2, mhalf, nsof_mul2(r0v[0], half));
    r3 = nsof_fma2(r3, mhalf, nsof_mul2(r0v[1], half));
    r2 = nsof_add2(r2, nsof_fma2(r6, dx, nsof_mul2(r4, dy)));
    r3 = nsof_add2(r3, nsof_fma2(r5, dx, nsof_mul2(r6, dy)));
    const bool bord = (unsigned)(xa - 5) >= (unsigned)(w - 10) || (unsigned)(ya - 5) >= (unsigned)(h - 10) ||
                      (unsigned)(xb - 5) >= (unsigned)(w - 10) || (unsigned)(yb - 5) >= (unsigned)(h - 10);
    if (bord) {
        const float2 sc = make_float2(nsof_um_border_scale(xa, ya, w, h), nsof_um_border_scale(xb, yb, w, h));
        r2 = nsof_mul2(r2, sc); r3 = nsof_mul2(r3, sc); r4 = nsof_mul2(r4, sc); r5 = nsof_mul2(r5, sc); r6 = nsof_mul2(r6, sc);
    }
    const float2 r66 = nsof_mul2(r6, r6);
    out[0] = nsof_fma2(r4, r4, r66);
    out[1] = nsof_mul2(nsof_add2(r4, r5), r6);
    out[2] = nsof_fma2(r5, r5, r66);
    out[3] = nsof_fma2(r4, r2, nsof_mul2(r6, r3));
    out[4] = nsof_fma2(r6, r2, nsof_mul2(r5, r3));
}

NSOF_DEV void fb_um_pixel2(const float *NSOF_RESTRICT R1, size_t plane, int pitch, int h, int w, int xa, int ya, int xb,
                           int yb, float2 dx, float2 dy, const float2 r0v[5], float2 out[5])
{
    fb_um_pixel2_t<false>(R1, plane, pitch, 0, 0, 0, 0, h, w, xa, ya, xb, yb, dx, dy, r0v, out);
}

// First UpdateMatrices of a level; the initial flow is zero (coarsest level) or the coarser
// level's flow upsampled on the fly (S2-f: resize INTER_LINEAR, times f32(1/pyr_scale)); the
// upsampled field itself is never written (UpdateFlow_Blur overwrites flow entirely).
#define FB_UM_TW 64   // (128 x 8 before the window kernel: its window needs rows, not width)
#ifndef FB_UM_TH
#define FB_UM_TH 16
#endif

// tiled runs: flag a gather that reaches R1 rows this rank neither owns nor received
NSOF_DEV void fb_um_guard(const FbJob &J, int y, float dy)
{
    const int y1 = (int)floorf((float)y + dy);
    if ((unsigned)y1 < (unsigned)(J.h - 1) && (y1 < J.gy_lo || y1 + 1 >= J.gy_hi)) *J.guard = 1;
}

NSOF_DEV float2 fb_upsampled_flow(const FbJob &J, int x, int y, float flow_mul)
{
    int x0, x1, y0, y1;
    float ax, ay;
    fb_lin_coord(x, J.fsx, J.cw, &x0, &x1, &ax);
    fb_lin_coord(y, J.fsy, J.ch, &y0, &y1, &ay);
    const float2 f00 = NSOF_LDG(J.cflow + (size_t)y0 * J.cpitch + x0), f01 = NSOF_LDG(J.cflow + (size_t)y0 * J.cpitch + x1);
    const float2 f10 = NSOF_LDG(J.cflow + (size_t)y1 * J.cpitch + x0), f11 = NSOF_LDG(J.cflow + (size_t)y1 * J.cpitch + x1);
    const float a0 = 1.f - ax, b0 = 1.f - ay;
    return make_float2(((f00.x * a0 + f01.x * ax) * b0 + (f10.x * a0 + f11.x * ax) * ay) * flow_mul,
                       ((f00.y * a0 + f01.y * ax) * b0 + (f10.y * a0 + f11.y * ax) * ay) * flow_mul);
}

// 4 pixels per thread: float4 loads of R0, float4 stores of M, 80 independent R1 gathers in flight per thread.
// (no launch bounds: 64 registers / 4 CTAs per SM as ptxas picks them; __launch_bounds__(256, 1) measured 2.35 -> 2.95 ms
// per full-frame step, (256, 5) with its 60-byte spill 2.51 ms)
// INIT = 1: the batch carries OPTFLOW_USE_INITIAL_FLOW items (J.init); a separate instantiation because the extra
// branch costs the hot path 6 registers and with them the fourth CTA per SM (2.35 -> 2.62 ms per full-frame step)
// Flat tile lists (round 2): 72 registers without a bound, 64 + a 12-byte spill with (256, 4): the fourth CTA per SM is kept
#ifndef FB_UM_MINB
#define FB_UM_MINB 4
#endif
template <int INIT>
NSOF_GLOBAL NSOF_LAUNCH_BOUNDS(256, FB_UM_MINB) k_fb_um_first_g(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl, float flow_mul)
{
    const uint2 te = tl[NSOF_BX];
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + te.x);
    const int tile = (int)te.y;
    const int tiles_x = (J.w + FB_UM_TW - 1) / FB_UM_TW;
    const size_t plane = (size_t)J.pitch * J.h;
    const int tx0 = (tile % tiles_x) * FB_UM_TW, ty0 = J.y_lo + (tile / tiles_x) * FB_UM_TH;
    // The R1 gather addresses depend on the coarse-flow loads: start pulling the tile's R1 rows (+- 4 rows for the
    // displacement) into L2 now, so that the second, dependent round of loads does not pay DRAM latency again.
    {
        const int py0 = nsof_imax(ty0 - 4, 0), py1 = nsof_imin(ty0 + FB_UM_TH + 4, J.h);
        const int px1 = nsof_imin(tx0 + FB_UM_TW, J.pitch);
        for (int r = NSOF_TID; r < 5 * (py1 - py0); r += NSOF_NT) {
            const int c = r / (py1 - py0), gy = py0 + r - c * (py1 - py0);
            nsof_prefetch_l2(J.R[1] + c * plane + (size_t)gy * J.pitch + tx0, (unsigned)(px1 - tx0) * 4u);
        }
    }
    // (one pixel per lane, as in phase 3b of the blur+solve kernel, was measured here and is SLOWER: 2.35 -> 2.60 ms per
    // full-frame step; this kernel has no shared-memory phase to overlap with, and 80 independent gathers per thread
    // hide more latency than the narrower L1 footprint saves)
    constexpr int GW = FB_UM_TW / 4;
    for (int i = NSOF_TID; i < GW * FB_UM_TH; i += NSOF_NT) {
        const int x = tx0 + (i % GW) * 4, y = ty0 + i / GW;
        if (x >= J.w || y >= J.y_hi) continue;
        const size_t o = (size_t)y * J.pitch + x;
        float r0[5][4];
#pragma unroll
        for (int c = 0; c < 5; c++) {
            const float4 t = NSOF_LDG((const float4 *)(J.R[0] + c * plane + o));
            r0[c][0] = t.x; r0[c][1] = t.y; r0[c][2] = t.z; r0[c][3] = t.w;
        }
        float mo[5][4];
        // (pixel pairs through fb_um_pixel2, as in phase 3b of the blur+solve kernel, need 90 registers here: 2.36 -> 2.97 ms
        // per full-frame step, 2.71 / 2.63 ms when held to 80 / 64 registers by launch bounds; left scalar)
#pragma unroll
        for (int q = 0; q < 4; q++) {
            float r0v[5], mm[5];
            if (x + q < J.w) {
                float2 d = make_float2(0.f, 0.f);
                if (INIT && J.init) d = NSOF_LDG(J.cflow + (size_t)y * J.cpitch + x + q);
                else if (J.ch > 0) d = fb_upsampled_flow(J, x + q, y, flow_mul);
#pragma unroll
                for (int c = 0; c < 5; c++) r0v[c] = r0[c][q];
                if (J.guard) fb_um_guard(J, y, d.y);
                fb_um_pixel(J.R[1], plane, J.pitch, J.h, J.w, x + q, y, d.x, d.y, r0v, mm);
            } else {
#pragma unroll
                for (int c = 0; c < 5; c++) mm[c] = 0.f;
            }
#pragma unroll
            for (int c = 0; c < 5; c++) mo[c][q] = mm[c];
        }
#pragma unroll
        for (int c = 0; c < 5; c++)
            *(float4 *)(J.Ma + c * plane + o) = make_float4(mo[c][0], mo[c][1], mo[c][2], mo[c][3]);
    }
}

static void (*const k_fb_um_first)(const FbJob *, const uint2 *, float) = k_fb_um_first_g<0>;
static void (*const k_fb_um_first_init)(const FbJob *, const uint2 *, float) = k_fb_um_first_g<1>;

// The same step with the R1 taps in shared memory (round 2; the kernel above stays as the window-off path).  ncu of the
// kernel above at level 0: L1 the busiest unit (87 %: with 4 pixels per thread a gather touches 4+ lines), top stall
// long_scoreboard 5.4, 297 instructions per pixel.  Here every thread owns one column of the tile (one pixel per lane, FB_UM_TH
// / (NT / TW) rows), computes its pixels' initial flow first (registers), the block reduces the range of the bilinear
// cells they touch, and when the range fits (TH+3) x (TW+8) pixels -- the initial flow is the upsampled coarser flow: smooth
// -- the five R1 planes of that window arrive by bulk copies and the 20 taps per pixel are LDS with immediate offsets.
// Tiles whose range does not fit gather from L2 as above (one pixel per lane).  Arithmetic per pixel = fb_um_pixel.
// dynamic smem: 32 B (mbarrier, 4 range integers) + 5 * FB_UMW_RH * FB_UMW_RS floats + (TW + TH) x 16 B of coordinates.
struct alignas(16) FbLin {
    int i0, i1;
    float a;
    int pad;
};
#define FB_UMW_RS (FB_UM_TW + 8)
#define FB_UMW_RH (FB_UM_TH + 3)
#define FB_UMW_SMEM (32 + 5 * FB_UMW_RH * FB_UMW_RS * sizeof(float) + (FB_UM_TW + FB_UM_TH) * 16)
// five CTAs per SM (48 registers, an 8-byte spill): the kernel waits on DRAM twice per tile (coarse flow, then window + R0),
// a fifth resident tile hides more of it than the spill costs -- 8.49 -> 8.33 ms per gated step, 12.55 -> 12.40 ms per
// full-frame step against four CTAs; six (40 registers, 72-byte spill): 8.45 / 12.65 ms
#ifndef FB_UMW_MINB
#define FB_UMW_MINB 5
#endif
template <int INIT>
NSOF_GLOBAL NSOF_LAUNCH_BOUNDS(256, FB_UMW_MINB) k_fb_um_first_w(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl, float flow_mul,
                                                                int flags, unsigned *stats)
{
    constexpr int TW = FB_UM_TW, TH = FB_UM_TH, NT = 256, RS = FB_UMW_RS, RH = FB_UMW_RH, SP = RH * RS;
    constexpr int ROWS = NT / TW, NPX = TH / ROWS;   // rows per pass, pixels per thread
    static_assert(NT % TW == 0 && TH % ROWS == 0 && RS % 4 == 0, "one column per thread");
    const uint2 te = tl[NSOF_BX];
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + te.x);
    const int tile = (int)te.y;
    const int tiles_x = (J.w + TW - 1) / TW;
    const size_t plane = (size_t)J.pitch * J.h;
    const int tx0 = (tile % tiles_x) * TW, ty0 = J.y_lo + (tile / tiles_x) * TH;
    NSOF_DYN_SMEM(smem_raw);
    nsof_mbar_t *bar = (nsof_mbar_t *)smem_raw;
    int *s_rng = (int *)(smem_raw + 16);
    float *sW = (float *)(smem_raw + 32);
    // source coordinates of the flow upsample (fb_lin_coord: double arithmetic) once per tile column / row instead of once
    // per pixel: {i0, i1, a} of the TW columns, then of the TH rows (the first profile of this kernel spent a third of its
    // instructions on them)
    FbLin *s_lin = (FbLin *)(sW + 5 * SP);
    if (NSOF_TID == 0) {
        nsof_mbar_init(bar, 1);
        s_rng[0] = s_rng[2] = 0x7fffffff;
        s_rng[1] = s_rng[3] = -0x7fffffff - 1;
    }
    if (J.ch > 0 && !(INIT && J.init))
        for (int i = NSOF_TID; i < TW + TH; i += NSOF_NT) {
            FbLin v;
            v.pad = 0;
            if (i < TW) fb_lin_coord(tx0 + i, J.fsx, J.cw, &v.i0, &v.i1, &v.a);
            else fb_lin_coord(ty0 + i - TW, J.fsy, J.ch, &v.i0, &v.i1, &v.a);
            s_lin[i] = v;
        }
    NSOF_SYNC();
    const float *R0 = J.R[0], *R1 = J.R[1];
    // (emulator: the one thread of a block walks the NT thread slots and keeps all their flows)
#ifdef NSOF_EMU
    float2 dv[NPX * NT];
#else
    float2 dv[NPX];
#endif
    float tmnx = 3.0e38f, tmxx = -3.0e38f, tmny = 3.0e38f, tmxy = -3.0e38f;
    // ---- phase A: the initial flow of this thread's pixels, and the range of fl(x + dx), fl(y + dy)
    NSOF_FOR_SLOTS(t, ts, NT) {
        const int lx = t % TW, x = tx0 + lx;
#pragma unroll
        for (int k = 0; k < NPX; k++) {
            const int y = ty0 + t / TW + k * ROWS;
            float2 d = make_float2(0.f, 0.f);
            if (x < J.w && y < J.y_hi) {
                if (INIT && J.init) d = NSOF_LDG(J.cflow + (size_t)y * J.cpitch + x);
                else if (J.ch > 0) {  // = fb_upsampled_flow(J, x, y, flow_mul) with the tabulated coordinates
                    const FbLin cx = s_lin[lx], cy = s_lin[TW + (y - ty0)];
                    const float ax = cx.a, ay = cy.a;
                    const float2 *r0p = J.cflow + (size_t)cy.i0 * J.cpitch, *r1p = J.cflow + (size_t)cy.i1 * J.cpitch;
                    const float2 f00 = NSOF_LDG(r0p + cx.i0), f01 = NSOF_LDG(r0p + cx.i1);
                    const float2 f10 = NSOF_LDG(r1p + cx.i0), f11 = NSOF_LDG(r1p + cx.i1);
                    const float a0 = 1.f - ax, b0 = 1.f - ay;
                    d = make_float2(((f00.x * a0 + f01.x * ax) * b0 + (f10.x * a0 + f11.x * ax) * ay) * flow_mul,
                                    ((f00.y * a0 + f01.y * ax) * b0 + (f10.y * a0 + f11.y * ax) * ay) * flow_mul);
                }
                const float px = (float)x + d.x, py = (float)y + d.y;
                tmnx = fminf(tmnx, px); tmxx = fmaxf(tmxx, px);
                tmny = fminf(tmny, py); tmxy = fmaxf(tmxy, py);
            }
            dv[ts * NPX + k] = d;
        }
    }
    {
        const int a = nsof_imax((int)floorf(fminf(fmaxf(tmnx, -1.f), 1.0e9f)), 0);
        const int b = nsof_imin((int)floorf(fmaxf(fminf(tmxx, 1.0e9f), -2.f)), J.w - 2);
        const int c = nsof_imax((int)floorf(fminf(fmaxf(tmny, -1.f), 1.0e9f)), 0);
        const int d = nsof_imin((int)floorf(fmaxf(fminf(tmxy, 1.0e9f), -2.f)), J.h - 2);
        nsof_block_minmax4(s_rng, a, b, c, d);
    }
    NSOF_SYNC();
    const int wx0 = s_rng[0] & ~3, wy0 = s_rng[2];
    const bool win = !(flags & 2) && s_rng[0] <= s_rng[1] && s_rng[2] <= s_rng[3] && s_rng[1] - wx0 <= RS - 2 &&
                     s_rng[3] - wy0 <= RH - 2;
    if (stats && NSOF_TID == 0) nsof_count(stats + (win ? 0 : 1));
    if (win) {
        const int ncol = nsof_imin(wx0 + RS, J.pitch) - wx0, nrow = s_rng[3] + 2 - wy0;
        if (NSOF_TID == 0) nsof_mbar_expect_tx(bar, (unsigned)(ncol * 4) * (unsigned)(5 * nrow));
        for (int r = NSOF_TID; r < 5 * nrow; r += NSOF_NT) {
            const int c = r / nrow, ly = r - c * nrow;
            nsof_bulk_g2s(sW + c * SP + ly * RS, R1 + c * plane + (size_t)(wy0 + ly) * J.pitch + wx0, (unsigned)(ncol * 4), bar);
        }
    }
    // ---- phase C: UpdateMatrices, one pixel per lane
    auto phase_c = [&](auto win_c) {
        constexpr bool WIN = decltype(win_c)::value;
        bool waited = !WIN;
        NSOF_FOR_SLOTS(t, ts, NT) {
            const int lx = t % TW, x = tx0 + lx;
            float r0[NPX][5];
            // R0 of all the thread's pixels first (independent of the window): in flight while the copy lands
#pragma unroll
            for (int k = 0; k < NPX; k++) {
                const int y = ty0 + t / TW + k * ROWS;
                if (x < J.w && y < J.y_hi) {
                    const size_t o = (size_t)y * J.pitch + x;
#pragma unroll
                    for (int c = 0; c < 5; c++) r0[k][c] = NSOF_LDG(R0 + c * plane + o);
                }
            }
            if (!waited) {
                if (NSOF_TID < NSOF_WARP) nsof_mbar_wait(bar, 0);
                NSOF_SYNC();
                waited = true;
            }
#pragma unroll
            for (int k = 0; k < NPX; k++) {
                const int y = ty0 + t / TW + k * ROWS;
                if (x < J.w && y < J.y_hi) {
                    const float2 d = dv[ts * NPX + k];
                    float mm[5];
                    if (J.guard) fb_um_guard(J, y, d.y);
                    if (WIN) fb_um_pixel_t<true>(sW, SP, RS, wx0, wy0, RS - 2, RH - 2, J.h, J.w, x, y, d.x, d.y, r0[k], mm);
                    else fb_um_pixel(R1, plane, J.pitch, J.h, J.w, x, y, d.x, d.y, r0[k], mm);
                    const size_t o = (size_t)y * J.pitch + x;
#pragma unroll
                    for (int c = 0; c < 5; c++) J.Ma[c * plane + o] = mm[c];
                }
            }
        }
    };
    if (win) phase_c(std::true_type{});
    else phase_c(std::false_type{});
}

static void (*const k_fb_um_first_win)(const FbJob *, const uint2 *, float, int, unsigned *) = k_fb_um_first_w<0>;
static void (*const k_fb_um_first_win_init)(const FbJob *, const uint2 *, float, int, unsigned *) = k_fb_um_first_w<1>;

// ---------------------------------------------------------------------------------------------
// OPTFLOW_USE_INITIAL_FLOW (A.8): flow_coarsest = resize(flow0, (w_k, h_k), INTER_AREA) * f32(scale_k), two-channel
// float.  OpenCV's two INTER_AREA paths, restated in oracle/farneback_oracle.c (bit-exact against cv2.resize there):
// both ratios integer -> sum of the ix*iy source pixels, row-major, accumulated four at a time as
// sum += ((a+b)+c)+d, times f32(1/area); otherwise the DecimateAlpha taps per axis (partially covered first pixel,
// fully covered ones at 1/cellWidth, partially covered last one), rows combined as sum = b0*buf0 + b1*buf1 + ...
// One thread per output pixel: the coarsest level is at most a few thousand pixels.
// ---------------------------------------------------------------------------------------------
struct FbAreaAxis {
    int s1, s2, has0, has1;
    float a0, am, a1;
};
NSOF_DEV FbAreaAxis fb_area_axis(int ssize, int dsize, int d)
{
    const double scale = (double)ssize / dsize;
    const double f1 = d * scale, f2 = f1 + scale;
    const double cell = fmin(scale, ssize - f1);
    int s1 = (int)ceil(f1), s2 = (int)floor(f2);
    s2 = nsof_imin(s2, ssize - 1);
    s1 = nsof_imin(s1, s2);
    FbAreaAxis A;
    A.s1 = s1; A.s2 = s2;
    A.has0 = (s1 - f1 > 1e-3) ? 1 : 0;
    A.has1 = (f2 - s2 > 1e-3) ? 1 : 0;
    A.a0 = (float)((s1 - f1) / cell);
    A.am = (float)(1.0 / cell);
    A.a1 = (float)(fmin(fmin(f2 - s2, 1.), cell) / cell);
    return A;
}
// tap i of an axis: source index and weight
NSOF_DEV void fb_area_tap(const FbAreaAxis &A, int i, int *si, float *a)
{
    const int j = i - A.has0;
    if (j < 0) { *si = A.s1 - 1; *a = A.a0; }
    else if (j < A.s2 - A.s1) { *si = A.s1 + j; *a = A.am; }
    else { *si = A.s2; *a = A.a1; }
}

NSOF_GLOBAL k_fb_area_init(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl)
{
    // runs on the UM tile list (initial flow is refused in the row-tiled mode, so [y_lo, y_hi) is the whole level)
    const uint2 te = tl[NSOF_BX];
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + te.x);
    if (J.init != 1) return;
    const int tile = (int)te.y;
    const int tiles_x = (J.w + FB_UM_TW - 1) / FB_UM_TW;
    const int tx0 = (tile % tiles_x) * FB_UM_TW, ty0 = (tile / tiles_x) * FB_UM_TH;
    const bool fastp = J.sh % J.h == 0 && J.sw % J.w == 0;
    for (int i = NSOF_TID; i < FB_UM_TW * FB_UM_TH; i += NSOF_NT) {
        const int x = tx0 + i % FB_UM_TW, y = ty0 + i / FB_UM_TW;
        if (x >= J.w || y >= J.h) continue;
        float sx = 0.f, sy = 0.f;
        if (fastp) {
            const int iy = J.sh / J.h, ix = J.sw / J.w, area = ix * iy;
            const float2 *base = J.flow0 + (size_t)y * iy * J.f0pitch + (size_t)x * ix;
#define FB_AV(k) NSOF_LDG(base + (size_t)((k) / ix) * J.f0pitch + (k) % ix)
            int k = 0;
            for (; k <= area - 4; k += 4) {
                const float2 a = FB_AV(k), b = FB_AV(k + 1), c = FB_AV(k + 2), d = FB_AV(k + 3);
                sx += ((a.x + b.x) + c.x) + d.x;
                sy += ((a.y + b.y) + c.y) + d.y;
            }
            for (; k < area; k++) {
                const float2 a = FB_AV(k);
                sx += a.x;
                sy += a.y;
            }
#undef FB_AV
            const float sc = 1.f / (float)area;
            sx *= sc;
            sy *= sc;
        } else {
            const FbAreaAxis AX = fb_area_axis(J.sw, J.w, x), AY = fb_area_axis(J.sh, J.h, y);
            const int nx = AX.has0 + (AX.s2 - AX.s1) + AX.has1, ny = AY.has0 + (AY.s2 - AY.s1) + AY.has1;
            for (int j = 0; j < ny; j++) {
                int yy;
                float beta;
                fb_area_tap(AY, j, &yy, &beta);
                const float2 *row = J.flow0 + (size_t)yy * J.f0pitch;
                float bx = 0.f, by = 0.f;
                for (int t = 0; t < nx; t++) {
                    int xx;
                    float al;
                    fb_area_tap(AX, t, &xx, &al);
                    const float2 v = NSOF_LDG(row + xx);
                    bx = bx + v.x * al;
                    by = by + v.y * al;
                }
                sx = j == 0 ? beta * bx : sx + beta * bx;
                sy = j == 0 ? beta * by : sy + beta * by;
            }
        }
        J.flow[(size_t)y * J.fpitch + x] = make_float2(sx * J.init_mul, sy * J.init_mul);
    }
}

// Level with zero iterations: flow is just the (scaled) upsampled coarse flow, or zero.
NSOF_GLOBAL k_fb_flow_init(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl, float flow_mul)
{
    const uint2 te = tl[NSOF_BX];
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + te.x);
    if (J.init) return;  // the initial flow is already in place
    const int tile = (int)te.y;
    const int tiles_x = (J.w + FB_UM_TW - 1) / FB_UM_TW;
    const int tx0 = (tile % tiles_x) * FB_UM_TW, ty0 = J.y_lo + (tile / tiles_x) * FB_UM_TH;
    for (int i = NSOF_TID; i < FB_UM_TW * FB_UM_TH; i += NSOF_NT) {
        const int x = tx0 + i % FB_UM_TW, y = ty0 + i / FB_UM_TW;
        if (x >= J.w || y >= J.y_hi) continue;
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
//
// S2-e' (OPTFLOW_FARNEBACK_GAUSSIAN, FarnebackUpdateFlow_GaussianBlur): same tile, gwin.on != 0.  The window is the
// separable float kernel gwin.k[0..m] (sigma = 0.3 m, normalised), both passes accumulate in float in OpenCV's order
// (centre tap first, then symmetric pairs outwards) and the sums are used unscaled.
// ---------------------------------------------------------------------------------------------
#define FB_MAX_WIN_M 63
struct FbGaussWin {
    int on;
    float k[FB_MAX_WIN_M + 1];
};

template <int GAUSS>  // compile-time: the box instantiation must not pay for the indexed weight table
NSOF_GLOBAL k_fb_blur_solve_g(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl, int TW, int TH, int m, float scale,
                              int src_is_a, int last, const FbGaussWin gwin)
{
    const uint2 te = tl[NSOF_BX];
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + te.x);
    const int tile = (int)te.y;
    const int tiles_x = (J.w + TW - 1) / TW;
    const size_t plane = (size_t)J.pitch * J.h;
    const float *Min = src_is_a ? J.Ma : J.Mb;
    float *Mout = src_is_a ? J.Mb : J.Ma;
    const int tx0 = (tile % tiles_x) * TW, ty0 = J.y_lo + (tile / tiles_x) * TH;
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
    if (GAUSS) {
        // in place: output row y overwrites input row y, the oldest row of its own window [y, y + 2m]
        for (int i = NSOF_TID; i < 5 * LW; i += NSOF_NT) {
            float *col = sM + (i / LW) * SP + (i % LW);
            for (int y = 0; y < TH; y++) {
                const float *c0 = col + (y + m) * PW;
                float s = c0[0] * gwin.k[0];
                for (int j = 1; j <= m; j++) s += (c0[j * PW] + c0[-j * PW]) * gwin.k[j];
                col[y * PW] = s;
            }
        }
    } else
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
        if (x >= J.w || y >= J.y_hi) continue;
        float s[5][4];
        if (GAUSS) {
            for (int c = 0; c < 5; c++) {
                const float *v = sM + c * SP + ly * PW + lx + m;
                for (int q = 0; q < 4; q++) {
                    float a = v[q] * gwin.k[0];
                    for (int j = 1; j <= m; j++) a += gwin.k[j] * (v[q - j] + v[q + j]);
                    s[c][q] = a;
                }
            }
        } else
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
                    for (int c = 0; c < 5; c++) r0v[c] = NSOF_LDG(J.R[0] + c * plane + o + q);
                    if (J.guard) fb_um_guard(J, y, fyv[q]);
                    fb_um_pixel(J.R[1], plane, J.pitch, J.h, J.w, x + q, y, fxv[q], fyv[q], r0v, mm);
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
// S2-e (+ S2-d fused), tuned variant for a compile-time window radius M, tile TW x TH and NT threads.
// Same arithmetic as k_fb_blur_solve except for the vertical window sums (float tree, see phase 2);
// what changes is how the work is laid out:
//   phase 1  the M tile + halo (5 planes x (TH+2M) rows) arrives by bulk asynchronous row copies
//            (cp.async.bulk global->shared on an mbarrier, SASS UBLKCP): no per-element index arithmetic,
//            no register staging, every row in flight at once.  Rows are 16-byte aligned runs starting at
//            tx0 - ML (ML = M rounded up to 4); replicate borders are patched in shared memory afterwards
//            (border tiles only).
//   phase 2  vertical window sums (float tree), one thread per (plane, column pair), written in place over rows
//            0 .. TH-1 of the tile; rows TH .. TH+2M-1 of every plane are dead from here on
//   phase 3a 4 pixels per thread: aligned float4 shared-memory reads of the 2M+4 window columns, 2x2 solve.
//            Last iteration: flow -> HBM.  Otherwise the flow tile goes to shared memory (into the dead rows)
//   phase 3b UpdateMatrices with ONE PIXEL PER LANE, consecutive lanes on consecutive x: the 20 bilinear R1
//            gathers of a warp instruction then fall into one or two 128-byte lines (with 4 pixels per thread
//            the lanes of one load sat 16 B apart: 4+ lines per instruction, and L1 was the busiest unit of the
//            kernel at 68 %); R0 loads and M' stores are 128 B per warp instruction.
//            The 20 gathers per pixel do not go to L2 any more when the tile's flow is coherent (round 2): phase 3a
//            also reduces the range of the bilinear cells (x1, y1) the tile will touch; if it fits a window of
//            RH x RS = (TH+3) x (TW+8) pixels, the five R1 planes of that window arrive by bulk row copies in the dead
//            rows 0 .. TH-1 of the five M planes (second phase of the same mbarrier) and the gathers are LDS.  The
//            window FOLLOWS the flow (its origin is the smallest x1 / y1), so the limit is the variation of the flow
//            inside a tile (3 rows, 4 - 7 columns), not its magnitude; tiles over a motion boundary keep the L2 path.
// dynamic smem: 32 B (mbarrier, 4 range integers) + 5 * (TH+2M) * (TW+2ML) floats.
// ---------------------------------------------------------------------------------------------
template <int M, int TW, int TH, int NT, int MINB>
NSOF_GLOBAL NSOF_LAUNCH_BOUNDS(NT, MINB) k_fb_blur_solve_t(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl, float scale,
                                                           int src_is_a, int flags, unsigned *stats)
{
    // flags: bit 0 last iteration of the level (flow -> HBM, no UpdateMatrices); bit 1 keep the R1 gathers on the L2
    // path (nsof_debug_r1_window(h, 0): A/B runs and the bit-identity test); bits 2 / 3 no L2 prefetch of the tile's R0 /
    // R1 rows in phase 1 (nsof_debug_bs_prefetch).  stats (NULL: off): tiles counted by path.
    const int last = flags & 1;
    constexpr int ML = (M + 3) & ~3, PW = TW + 2 * ML, LH = TH + 2 * M, SP = LH * PW;
    constexpr int NV = (2 * M + 4 + (ML - M) + 3) / 4;  // float4s covering the window columns of 4 outputs
    constexpr int RPP = (2 * M * PW) / (2 * TW);        // flow rows (TW float2) that fit the dead rows of one plane
    // small radii (winsize 3 / 4: M = 1 / 2): the dead halo rows cannot hold the flow tile; it gets its own region behind the
    // five planes, and the whole plane (not only rows 0 .. TH-1) is free for the R1 window of phase 3b
    constexpr bool FSEP = 5 * RPP < TH;
    constexpr int RPQ = FSEP ? 1 : RPP;
    static_assert((TH * PW) % 4 == 0 && (2 * TW) % 4 == 0 && SP % 4 == 0, "flow rows must stay 16-byte aligned");
    const auto flow_at = [](int ly) { return FSEP ? 5 * SP + ly * (2 * TW) : (ly / RPQ) * SP + TH * PW + (ly % RPQ) * (2 * TW); };
    const uint2 te = tl[NSOF_BX];
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + te.x);
    const int tile = (int)te.y;
    const int tiles_x = (J.w + TW - 1) / TW;
    const size_t plane = (size_t)J.pitch * J.h;
    const float *Min = src_is_a ? J.Ma : J.Mb;
    float *Mout = src_is_a ? J.Mb : J.Ma;
    const int tx0 = (tile % tiles_x) * TW, ty0 = J.y_lo + (tile / tiles_x) * TH;
    NSOF_DYN_SMEM(smem_raw);
    nsof_mbar_t *bar = (nsof_mbar_t *)smem_raw;
    int *s_rng = (int *)(smem_raw + 16);     // {min x1, max x1, min y1, max y1} of the tile's R1 gathers
    float *sM = (float *)(smem_raw + 32);
    // R1 window of phase 3b: RH rows of RS floats per plane, in rows 0 .. TH-1 of the plane's M tile
    constexpr int RS = TW + 8, RH = ((FSEP ? LH : TH) * PW) / RS;
    static_assert(RS % 4 == 0 && RH >= TH + 2, "R1 window: 16-byte rows, at least one row of slack");

    // ---- phase 1
    const int gx0 = tx0 - ML;                                         // global x of shared column 0
    const int cx0 = nsof_imax(gx0, 0), cx1 = nsof_imin(tx0 + TW + ML, J.pitch);  // copied run, multiples of 4
    const unsigned row_bytes = (unsigned)(cx1 - cx0) * 4u;
    if (NSOF_TID == 0) {
        nsof_mbar_init(bar, 1);
        nsof_mbar_expect_tx(bar, row_bytes * (unsigned)(5 * LH));
        s_rng[0] = s_rng[2] = 0x7fffffff;
        s_rng[1] = s_rng[3] = -0x7fffffff - 1;
    }
    NSOF_SYNC();
    // every thread issues at most a few row copies: the issue cost is spread over all warps
    for (int r = NSOF_TID; r < 5 * LH; r += NSOF_NT) {
        const int c = r / LH, ly = r - c * LH;
        const int gy = nsof_clampi(ty0 - M + ly, 0, J.h - 1);
        nsof_bulk_g2s(sM + c * SP + ly * PW + (cx0 - gx0), Min + c * plane + (size_t)gy * J.pitch + cx0, row_bytes, bar);
    }
    // phase 3b reads R0 and gathers R1 around this tile: start pulling those rows into L2 now (bulk prefetch, one
    // instruction per row; R1 with a margin of ML columns / M rows for the displacement), so that the dependent
    // loads two phases later find them there instead of paying DRAM latency at 35 % occupancy
    if (!last && (flags & 12) != 12) {
        const int px0 = nsof_imax(tx0 - ML, 0), px1 = nsof_imin(tx0 + TW + ML, J.pitch);
        const int py0 = nsof_imax(ty0 - M, 0), py1 = nsof_imin(ty0 + TH + M, J.h);
        const int t0 = nsof_imin(tx0 + TW, J.pitch);
        const bool pf0 = !(flags & 4), pf1 = !(flags & 8);   // A/B switches (nsof_debug_bs_prefetch)
        for (int r = NSOF_TID; r < 5 * (py1 - py0); r += NSOF_NT) {
            const int c = r / (py1 - py0), gy = py0 + r - c * (py1 - py0);
            if (pf1) nsof_prefetch_l2(J.R[1] + c * plane + (size_t)gy * J.pitch + px0, (unsigned)(px1 - px0) * 4u);
            if (pf0 && gy >= ty0 && gy < ty0 + TH) nsof_prefetch_l2(J.R[0] + c * plane + (size_t)gy * J.pitch + tx0, (unsigned)(t0 - tx0) * 4u);
        }
    }
    // one warp polls the mbarrier; the others sleep in the block barrier (a try_wait spin in every warp
    // steals issue slots from the co-resident CTAs: 32 % of all executed instructions in the first profile)
    if (NSOF_TID < NSOF_WARP) nsof_mbar_wait(bar, 0);
    NSOF_SYNC();
    if (tx0 - M < 0 || tx0 + TW + M > J.w) {  // replicate the first / last image column into the halo
        // one element per thread; on the right only the M columns past the last image column are ever read by a
        // valid output (a tile that is mostly outside the image used to fill all of them: 10 % of the kernel's
        // instructions on ROI-sized crops)
        const int s_first = -gx0, s_last = J.w - 1 - gx0;  // shared columns of image x = 0 and x = w - 1
        const int n_l = tx0 - M < 0 ? s_first - (ML - M) : 0;
        const int n_r = tx0 + TW + M > J.w ? nsof_imin(M, ML + TW + M - (s_last + 1)) : 0;
        const int per_row = n_l + n_r;
        for (int i = NSOF_TID; i < 5 * LH * per_row; i += NSOF_NT) {
            const int r = i / per_row, k = i - r * per_row;
            float *row = sM + (r / LH) * SP + (r % LH) * PW;
            if (k < n_l) row[ML - M + k] = row[s_first];
            else row[s_last + 1 + (k - n_l)] = row[s_last];
        }
        NSOF_SYNC();
    }

    // ---- phase 2: vertical 15-row window sums, one thread per (plane, pair of columns), all LH rows unrolled.
    // Each window is summed as a fixed tree  (v[y..y+7] + v[y+8..y+11]) + (v[y+12..y+13] + v[y+14])  built from
    // shared pair / quad / octet sums: 7 float additions per output instead of 14, depth 4, and no loop-carried
    // dependency.  (OpenCV keeps running sums in double; a running float sum would let rounding of large
    // neighbours leak into small windows, a direct tree does not: relative error <= 4 ulp of sum |v|.  The double
    // version cost 2 x 46 dependent FP64 additions per column and a float<->double conversion per element on
    // the quarter-rate XU pipe, which ncu showed as the busiest pipe of the kernel.)
    // Small radii (M <= 3: winsize 3 .. 7): the 2M+1 rows are added left to right as they stream by.
    static_assert(M == 7 || M <= 3, "window sums: the 15-row tree or the short direct sum");
    for (int it = NSOF_TID; it < 5 * (PW / 2); it += NSOF_NT) {
        const int c = it / (PW / 2);
        float2 *p = (float2 *)(sM + c * SP) + (it - c * (PW / 2));
        if constexpr (M == 7) {
            float2 v[LH], p2[LH], p4[LH], p8[LH];
#pragma unroll
            for (int j = 0; j < LH; j++) {
                v[j] = p[j * (PW / 2)];
                // the column pair is a natural operand of the packed FP32 add (FADD2): half the issue slots of this phase
                if (j >= 1) p2[j - 1] = nsof_add2(v[j - 1], v[j]);
                if (j >= 3) p4[j - 3] = nsof_add2(p2[j - 3], p2[j - 1]);
                if (j >= 7) p8[j - 7] = nsof_add2(p4[j - 7], p4[j - 3]);
                if (j >= 14) {
                    const int y = j - 14;  // row y of the column pair is dead (already loaded)
                    p[y * (PW / 2)] = nsof_add2(nsof_add2(p8[y], p4[y + 8]), nsof_add2(p2[y + 12], v[j]));
                }
            }
        } else {
            float2 v[LH];
#pragma unroll
            for (int j = 0; j < LH; j++) {
                v[j] = p[j * (PW / 2)];
                if (j >= 2 * M) {
                    const int y = j - 2 * M;  // row y is dead (already loaded)
                    float2 a = v[y];
#pragma unroll
                    for (int k = 1; k <= 2 * M; k++) a = nsof_add2(a, v[y + k]);
                    p[y * (PW / 2)] = a;
                }
            }
        }
    }
    NSOF_SYNC();
    // ---- phase 3a
    constexpr int GW = TW / 4;
#ifndef FB_BS_WINDOW
#define FB_BS_WINDOW 1
#endif
    // range of  fl(x + dx), fl(y + dy)  over the tile's pixels, in float (floor is monotonic: one conversion per thread)
    float tmnx = 3.0e38f, tmxx = -3.0e38f, tmny = 3.0e38f, tmxy = -3.0e38f;
    for (int i = NSOF_TID; i < GW * TH; i += NSOF_NT) {
        const int lx = (i % GW) * 4, ly = i / GW;
        const int x = tx0 + lx, y = ty0 + ly;
        if (x >= J.w || y >= J.y_hi) continue;
        float s[5][4];
#ifndef FB_BS_PACK3A
#define FB_BS_PACK3A 1
#endif
        float fxv[4], fyv[4];
        if constexpr (FB_BS_PACK3A && M == 7) {
        static_assert(M != 7 || (ML - M == 1 && NV == 5), "the packed window sums are written for columns v[1..18] of 20");
#pragma unroll
        for (int c = 0; c < 5; c++) {
            // window of output q = v[1+q .. 15+q].  v[4..15] is common to the four windows: it arrives as six aligned
            // pairs (halves of the float4 reads) and is summed as a tree of packed adds (5 FADD2 + 1 FADD instead of 11
            // FADD); the three private columns on either side are added with two shared partial sums.
            const float4 *q = (const float4 *)(sM + c * SP + ly * PW + lx);
            const float4 t0 = q[0], t1 = q[1], t2 = q[2], t3 = q[3], t4 = q[4];
            const float2 c2 = nsof_add2(nsof_add2(nsof_add2(make_float2(t1.x, t1.y), make_float2(t1.z, t1.w)),
                                                  nsof_add2(make_float2(t2.x, t2.y), make_float2(t2.z, t2.w))),
                                        nsof_add2(make_float2(t3.x, t3.y), make_float2(t3.z, t3.w)));
            const float common = c2.x + c2.y;
            const float a23 = t0.z + t0.w, b12 = t4.x + t4.y;  // v2 + v3, v16 + v17
            s[c][0] = common + (t0.y + a23);
            s[c][1] = common + (a23 + t4.x);
            s[c][2] = common + (t0.w + b12);
            s[c][3] = common + (b12 + t4.z);
        }
        {
            // 2x2 solve of two pixels per packed instruction (same operations as the scalar form below)
            const float2 sc = make_float2(scale, scale), nsc = make_float2(-scale, -scale), eps = make_float2(1e-3f, 1e-3f);
#pragma unroll
            for (int q = 0; q < 4; q += 2) {
                const float2 g11 = nsof_mul2(make_float2(s[0][q], s[0][q + 1]), sc), g12 = nsof_mul2(make_float2(s[1][q], s[1][q + 1]), sc);
                const float2 n12 = nsof_mul2(make_float2(s[1][q], s[1][q + 1]), nsc), g22 = nsof_mul2(make_float2(s[2][q], s[2][q + 1]), sc);
                const float2 h1 = nsof_mul2(make_float2(s[3][q], s[3][q + 1]), sc), h2 = nsof_mul2(make_float2(s[4][q], s[4][q + 1]), sc);
                const float2 det = nsof_fma2(n12, g12, nsof_fma2(g11, g22, eps));
                const float2 idet = make_float2(1.f / det.x, 1.f / det.y);
                const float2 fx2 = nsof_mul2(nsof_fma2(n12, h1, nsof_mul2(g11, h2)), idet);
                const float2 fy2 = nsof_mul2(nsof_fma2(n12, h2, nsof_mul2(g22, h1)), idet);
                fxv[q] = fx2.x; fxv[q + 1] = fx2.y;
                fyv[q] = fy2.x; fyv[q + 1] = fy2.y;
            }
        }
        } else {
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
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float g11 = s[0][q] * scale, g12 = s[1][q] * scale, g22 = s[2][q] * scale;
            const float h1 = s[3][q] * scale, h2 = s[4][q] * scale;
            const float idet = 1.f / (g11 * g22 - g12 * g12 + 1e-3f);
            fxv[q] = (g11 * h2 - g12 * h1) * idet;
            fyv[q] = (g22 * h1 - g12 * h2) * idet;
        }
        }
        if (last) {
            float2 *fo = J.flow + (size_t)y * J.fpitch + x;
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (x + q < J.w) fo[q] = make_float2(fxv[q], fyv[q]);
        } else {
            float4 *fo = (float4 *)(sM + flow_at(ly) + 2 * lx);
            fo[0] = make_float4(fxv[0], fyv[0], fxv[1], fyv[1]);
            fo[1] = make_float4(fxv[2], fyv[2], fxv[3], fyv[3]);
#if FB_BS_WINDOW
            // the same sums phase 3b forms (fb_um_pixel2_t: one rounded addition each); fminf / fmaxf drop a NaN
            const float fy = (float)y;
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (x + q < J.w) {
                    const float px = (float)(x + q) + fxv[q], py = fy + fyv[q];
                    tmnx = fminf(tmnx, px); tmxx = fmaxf(tmxx, px);
                    tmny = fminf(tmny, py); tmxy = fmaxf(tmxy, py);
                }
#endif
        }
    }
    if (last) return;
#if FB_BS_WINDOW
    {
        // cells outside the image are not gathered (the out-of-image branch of UpdateMatrices): clamp to the cells that are
        // (a thread without a valid pixel contributes the neutral pair 1e9 / -2)
        const int a = nsof_imax((int)floorf(fminf(fmaxf(tmnx, -1.f), 1.0e9f)), 0);
        const int b = nsof_imin((int)floorf(fmaxf(fminf(tmxx, 1.0e9f), -2.f)), J.w - 2);
        const int c = nsof_imax((int)floorf(fminf(fmaxf(tmny, -1.f), 1.0e9f)), 0);
        const int d = nsof_imin((int)floorf(fmaxf(fminf(tmxy, 1.0e9f), -2.f)), J.h - 2);
        nsof_block_minmax4(s_rng, a, b, c, d);
    }
#endif
    NSOF_SYNC();
    // ---- phase 3b
    const float *R0 = J.R[0], *R1 = J.R[1];
#if FB_BS_WINDOW
    const int wx0 = s_rng[0] & ~3, wy0 = s_rng[2];
    const bool win = !(flags & 2) && s_rng[0] <= s_rng[1] && s_rng[2] <= s_rng[3] && s_rng[1] - wx0 <= RS - 2 &&
                     s_rng[3] - wy0 <= RH - 2;
    if (stats && NSOF_TID == 0) nsof_count(stats + (win ? 0 : 1));
    if (!win && (flags & 8)) {
        // a tile that keeps the L2 gathers while phase 1 skipped the R1 prefetch (the default with the window on: for the
        // 99 % of tiles that take the window the 390 prefetch instructions were pure issue cost): start it now
        const int px0 = nsof_imax(tx0 - ML, 0), px1 = nsof_imin(tx0 + TW + ML, J.pitch);
        const int py0 = nsof_imax(ty0 - M, 0), py1 = nsof_imin(ty0 + TH + M, J.h);
        for (int r = NSOF_TID; r < 5 * (py1 - py0); r += NSOF_NT) {
            const int c = r / (py1 - py0), gy = py0 + r - c * (py1 - py0);
            nsof_prefetch_l2(R1 + c * plane + (size_t)gy * J.pitch + px0, (unsigned)(px1 - px0) * 4u);
        }
    }
    if (win) {
        // the window's rows of the five R1 planes -> rows 0 .. TH-1 of the M planes (dead since phase 3a)
        // (only the rows between the smallest y1 and the largest y1 + 1: 33 of the 35 for a uniform flow)
        const int ncol = nsof_imin(wx0 + RS, J.pitch) - wx0, nrow = s_rng[3] + 2 - wy0;
        if (NSOF_TID == 0) nsof_mbar_expect_tx(bar, (unsigned)(ncol * 4) * (unsigned)(5 * nrow));
        nsof_fence_proxy_async();
        for (int r = NSOF_TID; r < 5 * nrow; r += NSOF_NT) {
            const int c = r / nrow, ly = r - c * nrow;
            nsof_bulk_g2s(sM + c * SP + ly * RS, R1 + c * plane + (size_t)(wy0 + ly) * J.pitch + wx0, (unsigned)(ncol * 4), bar);
        }
        // (the wait is in phase3b, behind the first R0 loads)
    }
#else
    const int wx0 = 0, wy0 = 0;
    const bool win = false;
#endif
#ifndef FB_BS_UNROLL
#define FB_BS_UNROLL 2
#endif
#ifndef FB_BS_PAIR
#define FB_BS_PAIR 1
#endif
#if FB_BS_PAIR
    // two pixels per thread and iteration (i and i + NT: same column, NT / TW rows apart when NT is a multiple of TW), in
    // the two halves of packed FP32 operands (fb_um_pixel2)
    // (one instantiation per gather path: the block-uniform choice is made once, outside the pixel loop)
#ifndef FB_BS_R0_AHEAD
#define FB_BS_R0_AHEAD 1
#endif
    auto phase3b = [&](auto win_c) {
    constexpr bool WIN = decltype(win_c)::value;
    // window path: with the gathers in shared memory the R0 loads are the only L2 round trip left in an iteration; those
    // of the pixel pair of iteration k + 1 are issued before the arithmetic of iteration k (the first ones before the wait
    // for the window copy)
    constexpr bool AHEAD = WIN && FB_BS_R0_AHEAD;
    float2 r0n[5];
    auto r0_pair = [&](int i, float2 *r) {
        const int ib = i + NSOF_NT;
        const int xa = tx0 + i % TW, ya = ty0 + i / TW, xb = tx0 + ib % TW, yb = ty0 + ib / TW;
        if (ib < TW * TH && xa < J.w && ya < J.y_hi && xb < J.w && yb < J.y_hi) {
            const size_t oa = (size_t)ya * J.pitch + xa, ob = (size_t)yb * J.pitch + xb;
#pragma unroll
            for (int c = 0; c < 5; c++) r[c] = make_float2(NSOF_LDG(R0 + c * plane + oa), NSOF_LDG(R0 + c * plane + ob));
        }
    };
    if (AHEAD && NSOF_TID < TW * TH) r0_pair(NSOF_TID, r0n);
    if (WIN) {
        if (NSOF_TID < NSOF_WARP) nsof_mbar_wait(bar, 1);
        NSOF_SYNC();
    }
    for (int i = NSOF_TID; i < TW * TH; i += 2 * NSOF_NT) {
        const int ib = i + NSOF_NT;
        const int lxa = i % TW, lya = i / TW, lxb = ib % TW, lyb = ib / TW;
        const int xa = tx0 + lxa, ya = ty0 + lya, xb = tx0 + lxb, yb = ty0 + lyb;
        const bool va = xa < J.w && ya < J.y_hi, vb = ib < TW * TH && xb < J.w && yb < J.y_hi;
        float2 r0a[5];
        if (AHEAD) {  // every iteration hands over its prefetch and issues the next one, whichever branch it takes
#pragma unroll
            for (int c = 0; c < 5; c++) r0a[c] = r0n[c];
            if (i + 2 * NSOF_NT < TW * TH) r0_pair(i + 2 * NSOF_NT, r0n);
        }
        if (va && vb) {
            const float2 fa = *(const float2 *)(sM + flow_at(lya) + 2 * lxa);
            const float2 fb = *(const float2 *)(sM + flow_at(lyb) + 2 * lxb);
            const size_t oa = (size_t)ya * J.pitch + xa, ob = (size_t)yb * J.pitch + xb;
            float2 r0v[5], mm[5];
            if (AHEAD) {
#pragma unroll
                for (int c = 0; c < 5; c++) r0v[c] = r0a[c];
            } else {
#pragma unroll
                for (int c = 0; c < 5; c++) r0v[c] = make_float2(NSOF_LDG(R0 + c * plane + oa), NSOF_LDG(R0 + c * plane + ob));
            }
            if (J.guard) { fb_um_guard(J, ya, fa.y); fb_um_guard(J, yb, fb.y); }
            if (WIN)
                fb_um_pixel2_t<true>(sM, SP, RS, wx0, wy0, RS - 2, RH - 2, J.h, J.w, xa, ya, xb, yb, make_float2(fa.x, fb.x),
                                     make_float2(fa.y, fb.y), r0v, mm);
            else
                fb_um_pixel2(R1, plane, J.pitch, J.h, J.w, xa, ya, xb, yb, make_float2(fa.x, fb.x), make_float2(fa.y, fb.y), r0v, mm);
#pragma unroll
            for (int c = 0; c < 5; c++) { Mout[c * plane + oa] = mm[c].x; Mout[c * plane + ob] = mm[c].y; }
        } else if (va || vb) {
            const int lx = va ? lxa : lxb, ly = va ? lya : lyb, x = va ? xa : xb, y = va ? ya : yb;
            const float2 f = *(const float2 *)(sM + flow_at(ly) + 2 * lx);
            const size_t o = (size_t)y * J.pitch + x;
            // a pixel without a partner goes through the SAME packed routine (both halves = this pixel): which pixels
            // lose their partner depends on where the tiles start, and the scalar routine's compiler-chosen FMA
            // contraction differs from the packed one in the last bit -- a row-tiled run (tiles start at the band's
            // first row) must reproduce the untiled flow bit for bit
            float2 r0v[5], mm[5];
#pragma unroll
            for (int c = 0; c < 5; c++) { const float t = NSOF_LDG(R0 + c * plane + o); r0v[c] = make_float2(t, t); }
            if (J.guard) fb_um_guard(J, y, f.y);
            if (WIN)
                fb_um_pixel2_t<true>(sM, SP, RS, wx0, wy0, RS - 2, RH - 2, J.h, J.w, x, y, x, y, make_float2(f.x, f.x),
                                     make_float2(f.y, f.y), r0v, mm);
            else
                fb_um_pixel2(R1, plane, J.pitch, J.h, J.w, x, y, x, y, make_float2(f.x, f.x), make_float2(f.y, f.y), r0v, mm);
#pragma unroll
            for (int c = 0; c < 5; c++) Mout[c * plane + o] = mm[c].x;
        }
    }
    };
    if (win) phase3b(std::true_type{});
    else phase3b(std::false_type{});
#else
    constexpr int UNR3B = FB_BS_UNROLL;
#pragma unroll UNR3B
    for (int i = NSOF_TID; i < TW * TH; i += NSOF_NT) {
        const int lx = i % TW, ly = i / TW;
        const int x = tx0 + lx, y = ty0 + ly;
        if (x >= J.w || y >= J.y_hi) continue;
        const float2 f = *(const float2 *)(sM + flow_at(ly) + 2 * lx);
        const size_t o = (size_t)y * J.pitch + x;
        float r0v[5], mm[5];
#pragma unroll
        for (int c = 0; c < 5; c++) r0v[c] = NSOF_LDG(R0 + c * plane + o);
        if (J.guard) fb_um_guard(J, y, f.y);
        fb_um_pixel(R1, plane, J.pitch, J.h, J.w, x, y, f.x, f.y, r0v, mm);
#pragma unroll
        for (int c = 0; c < 5; c++) Mout[c * plane + o] = mm[c];
    }
#endif
}

static void (*const k_fb_blur_solve)(const FbJob *, const uint2 *, int, int, int, float, int, int, FbGaussWin) = k_fb_blur_solve_g<0>;
static void (*const k_fb_blur_solve_gauss)(const FbJob *, const uint2 *, int, int, int, float, int, int, FbGaussWin) = k_fb_blur_solve_g<1>;
// a plain function-pointer name: template commas do not survive nested launch macros
#ifndef FB_BS7_MINB
#define FB_BS7_MINB 3
#endif
#ifndef FB_BS7_NT
#define FB_BS7_NT 256
#endif
#define FB_BS7S_TW 32   // small tile: coarse levels / small batches are latency-bound, more and shorter CTAs
#define FB_BS7S_TH 16
#define FB_BS7S_NT 128
static void (*const FB_BLUR_SOLVE_M7)(const FbJob *, const uint2 *, float, int, int, unsigned *) = k_fb_blur_solve_t<7, FB_BS7_TW, FB_BS7_TH, FB_BS7_NT, FB_BS7_MINB>;
static void (*const FB_BLUR_SOLVE_M7S)(const FbJob *, const uint2 *, float, int, int, unsigned *) = k_fb_blur_solve_t<7, FB_BS7S_TW, FB_BS7S_TH, FB_BS7S_NT, 6>;
#define FB_BS7_SMEM (32 + 5 * (FB_BS7_TH + 14) * (FB_BS7_TW + 16) * sizeof(float))
// radii 1 and 2 (winsize 3 / 4, the two other parameter sets the reference ships): same kernel, flow tile in its own region
static void (*const FB_BLUR_SOLVE_M1)(const FbJob *, const uint2 *, float, int, int, unsigned *) = k_fb_blur_solve_t<1, FB_BS7_TW, FB_BS7_TH, FB_BS7_NT, FB_BS7_MINB>;
static void (*const FB_BLUR_SOLVE_M2)(const FbJob *, const uint2 *, float, int, int, unsigned *) = k_fb_blur_solve_t<2, FB_BS7_TW, FB_BS7_TH, FB_BS7_NT, FB_BS7_MINB>;
#define FB_BSM_SMEM(m) (32 + 5 * (FB_BS7_TH + 2 * (m)) * (FB_BS7_TW + 8) * sizeof(float) + FB_BS7_TH * FB_BS7_TW * 8)
// ... and their small tiles for latency-bound levels (a 161 x 161 uav frame is 18 tiles of 64 x 32)
static void (*const FB_BLUR_SOLVE_M1S)(const FbJob *, const uint2 *, float, int, int, unsigned *) = k_fb_blur_solve_t<1, FB_BS7S_TW, FB_BS7S_TH, FB_BS7S_NT, 6>;
static void (*const FB_BLUR_SOLVE_M2S)(const FbJob *, const uint2 *, float, int, int, unsigned *) = k_fb_blur_solve_t<2, FB_BS7S_TW, FB_BS7S_TH, FB_BS7S_NT, 6>;
#define FB_BSMS_SMEM(m) (32 + 5 * (FB_BS7S_TH + 2 * (m)) * (FB_BS7S_TW + 8) * sizeof(float) + FB_BS7S_TH * FB_BS7S_TW * 8)
#define FB_BS7S_SMEM (32 + 5 * (FB_BS7S_TH + 14) * (FB_BS7S_TW + 16) * sizeof(float))

// ---------------------------------------------------------------------------------------------
// S2-b + S2-c fused for the two pyramid shapes every BASELINE.json config spends its time in:
//   MODE 0  level 0        : I = GaussianBlur3x3(src)                       (no resize)
//   MODE 1  exact 2x level : I = 2x2 mean of GaussianBlur3x3(src)           (cv::resize at exactly 1/2)
// followed by FarnebackPolyExp with compile-time radius N.  The blurred image I never exists in HBM: the
// uint8 source window of the tile (+ N halo, + blur taps) is staged in shared memory, I is built there,
// the vertical moment pass writes three planes to shared memory and the horizontal pass writes R with
// float4 stores.  Per-pixel arithmetic (order of operations) is that of k_fb_pyr + k_fb_polyexp.
// grid (tiles, items, 2 images); items whose FbJob.fast != MODE + 1 exit (they take the generic kernels).
// static smem: max(3*TH*IP*4, SH*SW) + IH*IP*4 bytes  (IP = roundup4(TW+2N)).
// ---------------------------------------------------------------------------------------------
// FB_PP_PACKED = 1 builds the moment passes (phases C / D) in packed FP32 on row pairs.  Measured on B200 (round 2,
// profiles/r2_pyr_polyexp_packed_ab.txt): 25 % fewer executed instructions (FFMA2 takes its tap pairs straight from
// uniform registers, no MOVs) and 6 % MORE time (full-frame level 0: 35.6 vs 33.6 us per pair) -- the kernel is bound by
// its block barriers (stall `barrier` 4.1 cycles per issue: the wait for the bulk copy and the phase imbalance), not by
// issue slots or the FMA pipe, and the plane-at-a-time horizontal pass has less memory-level parallelism.  Off.
#ifndef FB_PP_PACKED
#define FB_PP_PACKED 0
#endif
template <int MODE, int N, int TW, int TH>
// 4 CTAs per SM (64 registers, no spill) hide the block barriers between the phases better than 3 (76 registers):
// 2.89 -> 2.60 ms per full-frame step; 5 and 6 CTAs (48 / 40 registers, spilling) measured 2.71 / 2.58 ms
#ifndef FB_PP_MINB
#define FB_PP_MINB 4
#endif
NSOF_GLOBAL NSOF_LAUNCH_BOUNDS(256, FB_PP_MINB) k_fb_pyr_polyexp_t(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl,
                                                         const float *NSOF_RESTRICT gk, PolyConst pc)
{
    constexpr int IW = TW + 2 * N, IH = TH + 2 * N, IP = (IW + 3) & ~3;
    // MODE 2: I comes from HBM (written by the generic pyramid kernel): no source window at all
    constexpr int SWN = MODE == 2 ? 4 : (MODE ? 2 * IW : IW) + 2, SHN = MODE == 2 ? 4 : (MODE ? 2 * IH : IH) + 2;  // source window
    constexpr int SW = (SWN + 3) & ~3;                                              // pitch of the float window
    constexpr int RP = (SWN + 32 + 15) & ~15;                                       // pitch of the raw uint8 window
    constexpr int A_BYTES = (3 * TH * IP * 4 > SHN * SW * 4 ? 3 * TH * IP * 4 : SHN * SW * 4);  // S (float) | sr
#if FB_PP_PACKED
    // Packed FP32 (FADD2 / FMUL2 / FFMA2) for the two moment passes: tile rows y and y + TH/2 travel in the two halves of
    // every operand.  For the halves to be an aligned register pair after a plain shared-memory load, I and the three
    // moment planes are stored ROW-PAIR INTERLEAVED: element (row r, column x) of I sits in float2 [r][x].x for
    // r < PR and in [r - HALF][x].y for r >= HALF (rows HALF .. PR-1 twice: they belong to the windows of both halves);
    // moment planes: float2 [y][x] = (row y, row y + HALF), y < HALF.
    constexpr int HALF = TH / 2, PR = IH - HALF;
    static_assert(TH % 8 == 0 && PR >= HALF && N <= 7, "row-pair layout");
    constexpr int SI_BYTES = PR * IP * 8;
#else
    constexpr int SI_BYTES = IH * IP * 4;
#endif
    constexpr int B_BYTES = (SI_BYTES > SHN * RP ? SI_BYTES : SHN * RP);                        // raw u8 | sI
#ifndef FB_PP_BSTRIP
#define FB_PP_BSTRIP 14
#endif
#ifndef FB_PP_STRIP
#define FB_PP_STRIP 8
#endif
    constexpr int BSTRIP = FB_PP_BSTRIP, BNS = (IH + BSTRIP - 1) / BSTRIP;
    static_assert(A_BYTES % 16 == 0 && B_BYTES % 16 == 0, "shared regions must stay 16-byte aligned");
    const uint2 te = tl[NSOF_BX];  // the list holds (item, image) entries of items with fast == MODE + 1 only
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + (te.x & FB_TL_ITEM));
    const int tile = (int)te.y;
    const int tiles_x = (J.w + TW - 1) / TW;
    const int img = (int)(te.x >> FB_TL_IMG_SHIFT);
    const uint8_t *src = img ? J.src1 : J.src0;
    const int sp = img ? J.spitch1 : J.spitch0;
    const size_t plane = (size_t)J.pitch * J.h;
    float *R = J.R[img];
    const int tx0 = (tile % tiles_x) * TW, ty0 = J.y_lo + (tile / tiles_x) * TH;
    NSOF_DYN_SMEM(smem_raw);
    nsof_mbar_t *bar = (nsof_mbar_t *)smem_raw;
    float *sS = (float *)(smem_raw + 16);                 // phases A2/B
    float *sr = (float *)(smem_raw + 16);                 // phases C/D (aliases sS)
    uint8_t *sraw = smem_raw + 16 + A_BYTES;              // phase A
    float *sI = (float *)(smem_raw + 16 + A_BYTES);       // phases B/C (aliases sraw)
    const int lane = NSOF_LANE, warp = NSOF_TID / NSOF_WARP, nw = nsof_imax(NSOF_NT / NSOF_WARP, 1);
#if FB_PP_PACKED
#define FB_SI_ST(ly, lx, v)                                                          \
    {                                                                                \
        const float v_ = (v);                                                        \
        if ((ly) < PR) sI[(((ly)) * IP + (lx)) * 2] = v_;                            \
        if ((ly) >= HALF) sI[(((ly) - HALF) * IP + (lx)) * 2 + 1] = v_;              \
    }
#define FB_SI_LD(ly, lx) ((ly) < PR ? sI[((ly) * IP + (lx)) * 2] : sI[(((ly) - HALF) * IP + (lx)) * 2 + 1])
#else
#define FB_SI_ST(ly, lx, v) sI[(ly) * IP + (lx)] = (v)
#define FB_SI_LD(ly, lx) sI[(ly) * IP + (lx)]
#endif

    // I positions of the tile + halo that lie inside the level image: lx in [lx_lo, lx_hi), ly in [ly_lo, ly_hi).
    // Positions outside are PolyExp's replicate border and are filled from these after phase B.
    const int lx_lo = nsof_imax(0, N - tx0), lx_hi = nsof_imin(IW, J.w - (tx0 - N));
    const int ly_lo = nsof_imax(0, N - ty0), ly_hi = nsof_imin(IH, J.h - (ty0 - N));
    const int cx_lo = tx0 - N + lx_lo, cy_lo = ty0 - N + ly_lo;
    if constexpr (MODE == 2) {
        // ---- A/B replaced: the in-image part of the I tile + halo straight from the level image in HBM
        const float *Iin = J.I + (size_t)img * J.pitch * J.h;
        const int wn = lx_hi - lx_lo, hn = ly_hi - ly_lo;
        for (int i = NSOF_TID; i < wn * hn; i += NSOF_NT) {
            const int ry = i / wn, rx = i - ry * wn;
            FB_SI_ST(ly_lo + ry, lx_lo + rx, NSOF_LDG(Iin + (size_t)(cy_lo + ry) * J.pitch + cx_lo + rx));
        }
        (void)src; (void)sp; (void)bar; (void)sS; (void)sraw; (void)gk;
    } else {
    // source window: shared (sy, sx) <-> source (Ylo + sy, Xlo + sx), BORDER_REFLECT_101 outside the source
    const int Xlo = (MODE ? 2 * cx_lo : cx_lo) - 1, Ylo = (MODE ? 2 * cy_lo : cy_lo) - 1;

    // ---- A: stage the window.  The in-image part arrives by one bulk asynchronous copy per row (16-byte aligned
    // superset of the row; the last source row only when the superset stays inside it), BORDER_REFLECT_101 is applied
    // to the raw bytes in shared memory (single bounce: the blur radius is 1), then uint8 -> float once per source
    // pixel.  Degenerate images (a side < 2) or an unaligned pitch take per-pixel loads.
    const int xs0 = nsof_imax(Xlo, 0), xs1 = nsof_imin(Xlo + SWN, J.sw);
    const int ys0 = nsof_imax(Ylo, 0), ys1 = nsof_imin(Ylo + SHN, J.sh);
    const bool bulk = J.sw > 1 && J.sh > 1 && (sp & 15) == 0 && xs1 > xs0 && ys1 > ys0;
    if (bulk) {
        const uint8_t *a0 = src + (size_t)ys0 * sp + xs0;
        const int mis = (int)((uintptr_t)a0 & 15);
        const int shift = (mis - (xs0 - Xlo)) & 15;              // window column sx lives at byte shift + sx of its row
        const unsigned row_bytes = (unsigned)((mis + (xs1 - xs0) + 15) & ~15);
        const bool last_safe = xs0 - mis + (int)row_bytes <= J.sw;  // the over-read of the last source row stays in it
        const int yb1 = (ys1 == J.sh && !last_safe) ? ys1 - 1 : ys1;
        if (NSOF_TID == 0) {
            nsof_mbar_init(bar, 1);
            nsof_mbar_expect_tx(bar, row_bytes * (unsigned)(yb1 - ys0));
        }
        NSOF_SYNC();
        for (int y = ys0 + NSOF_TID; y < yb1; y += NSOF_NT)
            nsof_bulk_g2s(sraw + (y - Ylo) * RP + shift + (xs0 - Xlo) - mis, a0 - mis + (size_t)(y - ys0) * sp, row_bytes, bar);
        if (yb1 < ys1) {
            const uint8_t *row = src + (size_t)yb1 * sp;
            for (int x = xs0 + NSOF_TID; x < xs1; x += NSOF_NT) sraw[(yb1 - Ylo) * RP + shift + (x - Xlo)] = NSOF_LDG(row + x);
        }
        if (NSOF_TID < NSOF_WARP) nsof_mbar_wait(bar, 0);  // one warp polls, the others sleep in the block barrier
        NSOF_SYNC();
        if (xs0 > Xlo || xs1 < Xlo + SWN) {  // left / right reflection, in-image rows
            const int nl = xs0 - Xlo, nr = Xlo + SWN - xs1;
            for (int i = NSOF_TID; i < (ys1 - ys0) * (nl + nr); i += NSOF_NT) {
                const int y = ys0 + i / (nl + nr), q = i % (nl + nr);
                const int X = q < nl ? Xlo + q : xs1 + (q - nl);
                uint8_t *row = sraw + (y - Ylo) * RP + shift;
                row[X - Xlo] = row[nsof_reflect101(X, J.sw) - Xlo];
            }
            NSOF_SYNC();
        }
        if (ys0 > Ylo || ys1 < Ylo + SHN) {  // top / bottom reflection, whole window rows
            const int nt = ys0 - Ylo, nb = Ylo + SHN - ys1;
            for (int i = NSOF_TID; i < (nt + nb) * SWN; i += NSOF_NT) {
                const int q = i / SWN, sx = i % SWN;
                const int Y = q < nt ? Ylo + q : ys1 + (q - nt);
                sraw[(Y - Ylo) * RP + shift + sx] = sraw[(nsof_reflect101(Y, J.sh) - Ylo) * RP + shift + sx];
            }
            NSOF_SYNC();
        }
        // A2: 4 source pixels per thread: two aligned words, funnel shift by the misalignment, 4 conversions
        constexpr int G = SW / 4;
        for (int i = NSOF_TID; i < SHN * G; i += NSOF_NT) {
            const int sy = i / G, g = i - sy * G;
            const uint32_t *wp = (const uint32_t *)(sraw + sy * RP) + g + (shift >> 2);
            const uint32_t w0 = wp[0], w1 = wp[1];
            const int sh8 = (shift & 3) * 8;
            const uint32_t u = sh8 ? ((w0 >> sh8) | (w1 << (32 - sh8))) : w0;
            *(float4 *)(sS + sy * SW + 4 * g) =
                make_float4(nsof_u8_to_f(u, 0), nsof_u8_to_f(u, 1), nsof_u8_to_f(u, 2), nsof_u8_to_f(u, 3));
        }
    } else {
        for (int sy = warp; sy < SHN; sy += nw) {
            const uint8_t *row = src + (size_t)nsof_reflect101(Ylo + sy, J.sh) * sp;
            for (int sx = lane; sx < SWN; sx += NSOF_WARP) sS[sy * SW + sx] = (float)NSOF_LDG(row + nsof_reflect101(Xlo + sx, J.sw));
        }
    }
    NSOF_SYNC();
    // ---- B: I tile with halo; one thread per (column, strip of BSTRIP rows), row-blurred samples slide down
    {
        const float g0 = gk[0], g1 = gk[1];
        for (int i = NSOF_TID; i < IW * BNS; i += NSOF_NT) {
            const int st = i / IW, lx = i - st * IW;
            if (lx < lx_lo || lx >= lx_hi) continue;
            const int y_a = nsof_imax(st * BSTRIP, ly_lo), y_b = nsof_imin(nsof_imin((st + 1) * BSTRIP, IH), ly_hi);
            if (y_a >= y_b) continue;
            if (MODE == 0) {
                const float *c = sS + (y_a - ly_lo) * SW + (lx - lx_lo) + 1;  // row above the first centre row
                float rbm, rb0;
                { float acc = g0 * c[0]; acc = acc + g1 * (c[-1] + c[1]); rbm = acc; }
                c += SW;
                { float acc = g0 * c[0]; acc = acc + g1 * (c[-1] + c[1]); rb0 = acc; }
                for (int ly = y_a; ly < y_b; ly++) {
                    c += SW;
                    float rbp = g0 * c[0];
                    rbp = rbp + g1 * (c[-1] + c[1]);
                    float acc = g0 * rb0;
                    acc = acc + g1 * (rbm + rbp);
                    FB_SI_ST(ly, lx, acc);
                    rbm = rb0;
                    rb0 = rbp;
                }
            } else {
                // I(ly) needs source rows 2(ly-ly_lo) .. +3 and columns 2(lx-lx_lo) .. +3 of the window
                const float *c = sS + 2 * (y_a - ly_lo) * SW + 2 * (lx - lx_lo);
                float h0[2], h1[2], h2[2], h3[2];
#define FB_HB(dst, row)                                                                  \
    {                                                                                    \
        const float2 pa = *(const float2 *)(row), pb = *(const float2 *)((row) + 2);     \
        float a0 = g0 * pa.y; a0 = a0 + g1 * (pa.x + pb.x);                              \
        float a1 = g0 * pb.x; a1 = a1 + g1 * (pa.y + pb.y);                              \
        dst[0] = a0; dst[1] = a1;                                                        \
    }
                FB_HB(h0, c);
                FB_HB(h1, c + SW);
                for (int ly = y_a; ly < y_b; ly++) {
                    FB_HB(h2, c + 2 * SW);
                    FB_HB(h3, c + 3 * SW);
                    float s00 = g0 * h1[0]; s00 = s00 + g1 * (h0[0] + h2[0]);
                    float s01 = g0 * h1[1]; s01 = s01 + g1 * (h0[1] + h2[1]);
                    float s10 = g0 * h2[0]; s10 = s10 + g1 * (h1[0] + h3[0]);
                    float s11 = g0 * h2[1]; s11 = s11 + g1 * (h1[1] + h3[1]);
                    FB_SI_ST(ly, lx, (s00 + s01 + s10 + s11) * 0.25f);
                    h0[0] = h2[0]; h0[1] = h2[1]; h1[0] = h3[0]; h1[1] = h3[1];
                    c += 2 * SW;
                }
#undef FB_HB
            }
        }
    }
    }  // MODE != 2
    NSOF_SYNC();
    if (lx_lo > 0 || lx_hi < IW || ly_lo > 0 || ly_hi < IH) {  // replicate border of the level image
        for (int ly = warp; ly < IH; ly += nw) {
            const int sy = nsof_clampi(ly, ly_lo, ly_hi - 1);
            for (int lx = lane; lx < IW; lx += NSOF_WARP) {
                const int sx = nsof_clampi(lx, lx_lo, lx_hi - 1);
                if (sx != lx || sy != ly) FB_SI_ST(ly, lx, FB_SI_LD(sy, sx));
            }
        }
        NSOF_SYNC();
    }
#if FB_PP_PACKED
    // ---- C (packed): vertical moments of rows y and y + HALF at once, one thread per (column, strip of 4 row pairs)
    {
        constexpr int STRIP = 4, NS = HALF / STRIP, PLANE = HALF * IP * 2;
        const float2 mone = make_float2(-1.f, -1.f);
        for (int i = NSOF_TID; i < IW * NS; i += NSOF_NT) {
            const int lx = i % IW, st = i / IW;
            float2 col[STRIP + 2 * N];
#pragma unroll
            for (int j = 0; j < STRIP + 2 * N; j++) col[j] = *(const float2 *)(sI + ((st * STRIP + j) * IP + lx) * 2);
#pragma unroll
            for (int y = 0; y < STRIP; y++) {
                float2 r0 = nsof_mul2(col[y + N], pc.g2[0]), r1 = make_float2(0.f, 0.f), r2 = make_float2(0.f, 0.f);
#pragma unroll
                for (int k = 1; k <= N; k++) {
                    const float2 a = col[y + N - k], b = col[y + N + k];
                    const float2 p = nsof_add2(a, b), d = nsof_fma2(a, mone, b);  // b - a, one rounding
                    r0 = nsof_fma2(pc.g2[k], p, r0);
                    r1 = nsof_fma2(pc.xg2[k], d, r1);
                    r2 = nsof_fma2(pc.xxg2[k], p, r2);
                }
                const int o = ((st * STRIP + y) * IP + lx) * 2;
                *(float2 *)(sr + o) = r0;
                *(float2 *)(sr + PLANE + o) = r1;
                *(float2 *)(sr + 2 * PLANE + o) = r2;
            }
        }
    }
    NSOF_SYNC();
    // ---- D (packed): horizontal pass, 2 columns x (row y, row y + HALF) per thread, one moment plane at a time (the
    // window of a plane is 2 + 2N float2 registers); the five outputs are finished in scalar arithmetic so that they
    // land in consecutive registers for the 8-byte stores
    {
        constexpr int CP = TW / 2, NQ = 2 + 2 * N, PLANE = HALF * IP * 2;
        static_assert(NQ % 2 == 0, "window loads are float4 = two float2");
        const float2 mone = make_float2(-1.f, -1.f);
        for (int i = NSOF_TID; i < CP * HALF; i += NSOF_NT) {
            const int lx = (i % CP) * 2, ly = i / CP;
            const int x = tx0 + lx, ya = ty0 + ly, yb = ya + HALF;
            if (x >= J.w || ya >= J.y_hi) continue;
            float2 q[NQ], b1[2], b2[2], b3[2], b4[2], b5[2], b6[2];
#define FB_PP_LOADQ(p)                                                                        \
    {                                                                                         \
        const float4 *s4 = (const float4 *)(sr + (p) * PLANE + (ly * IP + lx) * 2);           \
        _Pragma("unroll") for (int k = 0; k < NQ / 2; k++) {                                  \
            const float4 t = s4[k];                                                           \
            q[2 * k] = make_float2(t.x, t.y);                                                 \
            q[2 * k + 1] = make_float2(t.z, t.w);                                             \
        }                                                                                     \
    }
            FB_PP_LOADQ(0);
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int ctr = e + N;
                b1[e] = nsof_mul2(q[ctr], pc.g2[0]);
                b2[e] = make_float2(0.f, 0.f);
                b4[e] = make_float2(0.f, 0.f);
#pragma unroll
                for (int k = 1; k <= N; k++) {
                    const float2 tg = nsof_add2(q[ctr + k], q[ctr - k]), d = nsof_fma2(q[ctr - k], mone, q[ctr + k]);
                    b1[e] = nsof_fma2(tg, pc.g2[k], b1[e]);
                    b4[e] = nsof_fma2(tg, pc.xxg2[k], b4[e]);
                    b2[e] = nsof_fma2(d, pc.xg2[k], b2[e]);
                }
            }
            FB_PP_LOADQ(1);
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int ctr = e + N;
                b3[e] = nsof_mul2(q[ctr], pc.g2[0]);
                b6[e] = make_float2(0.f, 0.f);
#pragma unroll
                for (int k = 1; k <= N; k++) {
                    b3[e] = nsof_fma2(nsof_add2(q[ctr + k], q[ctr - k]), pc.g2[k], b3[e]);
                    b6[e] = nsof_fma2(nsof_fma2(q[ctr - k], mone, q[ctr + k]), pc.xg2[k], b6[e]);
                }
            }
            FB_PP_LOADQ(2);
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int ctr = e + N;
                b5[e] = nsof_mul2(q[ctr], pc.g2[0]);
#pragma unroll
                for (int k = 1; k <= N; k++) b5[e] = nsof_fma2(nsof_add2(q[ctr + k], q[ctr - k]), pc.g2[k], b5[e]);
            }
#undef FB_PP_LOADQ
            {
                const size_t o = (size_t)ya * J.pitch + x;
                *(float2 *)(R + o) = make_float2(b3[0].x * pc.ig11, b3[1].x * pc.ig11);
                *(float2 *)(R + plane + o) = make_float2(b2[0].x * pc.ig11, b2[1].x * pc.ig11);
                *(float2 *)(R + 2 * plane + o) = make_float2(b1[0].x * pc.ig03 + b5[0].x * pc.ig33, b1[1].x * pc.ig03 + b5[1].x * pc.ig33);
                *(float2 *)(R + 3 * plane + o) = make_float2(b1[0].x * pc.ig03 + b4[0].x * pc.ig33, b1[1].x * pc.ig03 + b4[1].x * pc.ig33);
                *(float2 *)(R + 4 * plane + o) = make_float2(b6[0].x * pc.ig55, b6[1].x * pc.ig55);
            }
            if (yb < J.y_hi) {
                const size_t o = (size_t)yb * J.pitch + x;
                *(float2 *)(R + o) = make_float2(b3[0].y * pc.ig11, b3[1].y * pc.ig11);
                *(float2 *)(R + plane + o) = make_float2(b2[0].y * pc.ig11, b2[1].y * pc.ig11);
                *(float2 *)(R + 2 * plane + o) = make_float2(b1[0].y * pc.ig03 + b5[0].y * pc.ig33, b1[1].y * pc.ig03 + b5[1].y * pc.ig33);
                *(float2 *)(R + 3 * plane + o) = make_float2(b1[0].y * pc.ig03 + b4[0].y * pc.ig33, b1[1].y * pc.ig03 + b4[1].y * pc.ig33);
                *(float2 *)(R + 4 * plane + o) = make_float2(b6[0].y * pc.ig55, b6[1].y * pc.ig55);
            }
        }
    }
#else
    // ---- C: vertical moments, one thread per (column, strip of 8 rows), the column window kept in registers
    {
        constexpr int STRIP = FB_PP_STRIP, NS = TH / STRIP;
        for (int i = NSOF_TID; i < IW * NS; i += NSOF_NT) {
            const int lx = i % IW, st = i / IW;
            float col[STRIP + 2 * N];
#pragma unroll
            for (int j = 0; j < STRIP + 2 * N; j++) col[j] = sI[(st * STRIP + j) * IP + lx];
#pragma unroll
            for (int y = 0; y < STRIP; y++) {
                float r0 = col[y + N] * pc.g[0], r1 = 0.f, r2 = 0.f;
#pragma unroll
                for (int k = 1; k <= N; k++) {
                    const float a = col[y + N - k], b = col[y + N + k];
                    const float p = a + b;
                    r0 = r0 + pc.g[k] * p;
                    r1 = r1 + pc.xg[k] * (b - a);
                    r2 = r2 + pc.xxg[k] * p;
                }
                const int o = (st * STRIP + y) * IP + lx;
                sr[o] = r0;
                sr[TH * IP + o] = r1;
                sr[2 * TH * IP + o] = r2;
            }
        }
    }
    NSOF_SYNC();
// FB_PP_D8 = 1: eight outputs per thread, one moment plane at a time (half the shared-memory loads per output, one
// round of 256 items per tile).  Measured SLOWER than the 4-output form (full-frame step: fb_pyr_polyexp 3.26 vs 2.90 ms,
// profiles/r2_pyr_polyexp_packed_ab.txt): longer dependent chains per thread and 2 x fewer independent items. Off.
#ifndef FB_PP_D8
#define FB_PP_D8 0
#endif
#if FB_PP_D8
    // ---- D: horizontal pass, EIGHT outputs per thread, one moment plane at a time (planes 0, 2, 1; the outputs that
    // depend on planes 0 and 2 are stored before plane 1 is loaded, so at most 32 partial sums are alive).  The kernel's
    // busiest unit is the shared-memory pipe (ncu: l1tex 72 %); a thread of the 4-output form loaded 15 float4 per 4
    // outputs, this one loads 15 per 8 -- and the 256 items of a tile are exactly one round of the 256 threads.
    {
        constexpr int GW = TW / 8, NV = (8 + 2 * N + 3) / 4;
        for (int i = NSOF_TID; i < GW * TH; i += NSOF_NT) {
            // 8 consecutive lanes = 4 column groups x 2 rows: their float4 loads (rows IP = 76 words apart) fall into
            // disjoint banks; 8 column groups side by side would be 32 bytes apart and conflict 2-way
            static_assert(GW == 8 && TH % 2 == 0, "lane mapping of phase D");
            const int j8 = i & 7, q8 = i >> 3;
            const int lx = ((((q8 & 1) << 2) | (j8 & 3))) * 8, ly = ((q8 >> 1) << 1) | (j8 >> 2);
            const int x = tx0 + lx, y = ty0 + ly;
            if (x >= J.w || y >= J.y_hi) continue;
            float q[NV * 4];
#define FB_PP_LOADQ(p)                                                                   \
    {                                                                                    \
        const float4 *s4 = (const float4 *)(sr + (p) * TH * IP + ly * IP + lx);          \
        _Pragma("unroll") for (int k = 0; k < NV; k++) {                                 \
            const float4 t = s4[k];                                                      \
            q[4 * k] = t.x; q[4 * k + 1] = t.y; q[4 * k + 2] = t.z; q[4 * k + 3] = t.w;  \
        }                                                                                \
    }
            const size_t o = (size_t)y * J.pitch + x;
            const float g0 = pc.g[0];
            float b1[8], b2[8], b4[8], b5[8];
            FB_PP_LOADQ(0);
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const int ctr = e + N;
                b1[e] = q[ctr] * g0; b2[e] = 0.f; b4[e] = 0.f;
#pragma unroll
                for (int k = 1; k <= N; k++) {
                    const float p0 = q[ctr + k], m0 = q[ctr - k];
                    const float tg = p0 + m0;
                    b1[e] += tg * pc.g[k];
                    b4[e] += tg * pc.xxg[k];
                    b2[e] += (p0 - m0) * pc.xg[k];
                }
            }
            FB_PP_LOADQ(2);
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const int ctr = e + N;
                b5[e] = q[ctr] * g0;
#pragma unroll
                for (int k = 1; k <= N; k++) b5[e] += (q[ctr + k] + q[ctr - k]) * pc.g[k];
            }
#pragma unroll
            for (int hq = 0; hq < 2; hq++) {
                const int e = 4 * hq;
                *(float4 *)(R + plane + o + e) = make_float4(b2[e] * pc.ig11, b2[e + 1] * pc.ig11, b2[e + 2] * pc.ig11, b2[e + 3] * pc.ig11);
                *(float4 *)(R + 2 * plane + o + e) = make_float4(b1[e] * pc.ig03 + b5[e] * pc.ig33, b1[e + 1] * pc.ig03 + b5[e + 1] * pc.ig33,
                                                                 b1[e + 2] * pc.ig03 + b5[e + 2] * pc.ig33, b1[e + 3] * pc.ig03 + b5[e + 3] * pc.ig33);
                *(float4 *)(R + 3 * plane + o + e) = make_float4(b1[e] * pc.ig03 + b4[e] * pc.ig33, b1[e + 1] * pc.ig03 + b4[e + 1] * pc.ig33,
                                                                 b1[e + 2] * pc.ig03 + b4[e + 2] * pc.ig33, b1[e + 3] * pc.ig03 + b4[e + 3] * pc.ig33);
            }
            FB_PP_LOADQ(1);
#undef FB_PP_LOADQ
            float b3[8], b6[8];
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const int ctr = e + N;
                b3[e] = q[ctr] * g0; b6[e] = 0.f;
#pragma unroll
                for (int k = 1; k <= N; k++) {
                    const float p1 = q[ctr + k], m1 = q[ctr - k];
                    b3[e] += (p1 + m1) * pc.g[k];
                    b6[e] += (p1 - m1) * pc.xg[k];
                }
            }
#pragma unroll
            for (int hq = 0; hq < 2; hq++) {
                const int e = 4 * hq;
                *(float4 *)(R + o + e) = make_float4(b3[e] * pc.ig11, b3[e + 1] * pc.ig11, b3[e + 2] * pc.ig11, b3[e + 3] * pc.ig11);
                *(float4 *)(R + 4 * plane + o + e) = make_float4(b6[e] * pc.ig55, b6[e + 1] * pc.ig55, b6[e + 2] * pc.ig55, b6[e + 3] * pc.ig55);
            }
        }
    }
#else
    // ---- D: horizontal pass, 4 outputs per thread
    {
        constexpr int GW = TW / 4, NV = (4 + 2 * N + 3) / 4;
        for (int i = NSOF_TID; i < GW * TH; i += NSOF_NT) {
            const int lx = (i % GW) * 4, ly = i / GW;
            const int x = tx0 + lx, y = ty0 + ly;
            if (x >= J.w || y >= J.y_hi) continue;
            float q[3][NV * 4];
#pragma unroll
            for (int p = 0; p < 3; p++) {
                const float4 *s4 = (const float4 *)(sr + p * TH * IP + ly * IP + lx);
#pragma unroll
                for (int k = 0; k < NV; k++) {
                    const float4 t = s4[k];
                    q[p][4 * k] = t.x; q[p][4 * k + 1] = t.y; q[p][4 * k + 2] = t.z; q[p][4 * k + 3] = t.w;
                }
            }
            float o0[4], o1[4], o2[4], o3[4], o4[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int ctr = e + N;
                const float g0 = pc.g[0];
                float b1 = q[0][ctr] * g0, b2 = 0.f, b3 = q[1][ctr] * g0, b4 = 0.f, b5 = q[2][ctr] * g0, b6 = 0.f;
#pragma unroll
                for (int k = 1; k <= N; k++) {
                    const float gkk = pc.g[k], xgk = pc.xg[k], xxgk = pc.xxg[k];
                    const float p0 = q[0][ctr + k], m0 = q[0][ctr - k], p1 = q[1][ctr + k], m1 = q[1][ctr - k];
                    const float p2 = q[2][ctr + k], m2 = q[2][ctr - k];
                    const float tg = p0 + m0;
                    b1 += tg * gkk;
                    b4 += tg * xxgk;
                    b2 += (p0 - m0) * xgk;
                    b3 += (p1 + m1) * gkk;
                    b6 += (p1 - m1) * xgk;
                    b5 += (p2 + m2) * gkk;
                }
                o0[e] = b3 * pc.ig11;
                o1[e] = b2 * pc.ig11;
                o2[e] = b1 * pc.ig03 + b5 * pc.ig33;
                o3[e] = b1 * pc.ig03 + b4 * pc.ig33;
                o4[e] = b6 * pc.ig55;
            }
            const size_t o = (size_t)y * J.pitch + x;
            *(float4 *)(R + o) = make_float4(o0[0], o0[1], o0[2], o0[3]);
            *(float4 *)(R + plane + o) = make_float4(o1[0], o1[1], o1[2], o1[3]);
            *(float4 *)(R + 2 * plane + o) = make_float4(o2[0], o2[1], o2[2], o2[3]);
            *(float4 *)(R + 3 * plane + o) = make_float4(o3[0], o3[1], o3[2], o3[3]);
            *(float4 *)(R + 4 * plane + o) = make_float4(o4[0], o4[1], o4[2], o4[3]);
        }
    }
#endif
#endif
#undef FB_SI_ST
#undef FB_SI_LD
}

#define FB_PP_TW 64
#define FB_PP_TH 32
#define FB_PP_N 5
static void (*const FB_PYR_POLYEXP_M0)(const FbJob *, const uint2 *, const float *, PolyConst) = k_fb_pyr_polyexp_t<0, FB_PP_N, FB_PP_TW, FB_PP_TH>;
static void (*const FB_PYR_POLYEXP_M1)(const FbJob *, const uint2 *, const float *, PolyConst) = k_fb_pyr_polyexp_t<1, FB_PP_N, FB_PP_TW, FB_PP_TH>;
// level 0 of the two other parameter sets the reference ships (poly_n 1 and 10; their pyr_scale 0.6 leaves only level 0
// on a fused shape)
static void (*const FB_PYR_POLYEXP_M0_N1)(const FbJob *, const uint2 *, const float *, PolyConst) = k_fb_pyr_polyexp_t<0, 1, FB_PP_TW, FB_PP_TH>;
static void (*const FB_PYR_POLYEXP_M0_N10)(const FbJob *, const uint2 *, const float *, PolyConst) = k_fb_pyr_polyexp_t<0, 10, FB_PP_TW, FB_PP_TH>;
// MODE 2: PolyExp alone with the same moment passes, I read from HBM -- the levels the generic pyramid kernel serves
// (ksize 9 / 19 of the coarse levels, bilinear ratios of pyr_scale 0.6) instead of the generic k_fb_polyexp
static void (*const FB_POLYEXP_T_N1)(const FbJob *, const uint2 *, const float *, PolyConst) = k_fb_pyr_polyexp_t<2, 1, FB_PP_TW, FB_PP_TH>;
static void (*const FB_POLYEXP_T_N10)(const FbJob *, const uint2 *, const float *, PolyConst) = k_fb_pyr_polyexp_t<2, 10, FB_PP_TW, FB_PP_TH>;
static inline size_t fb_pp_smem(int mode, int n = FB_PP_N)
{
    const int IW = FB_PP_TW + 2 * n, IH = FB_PP_TH + 2 * n, IP = (IW + 3) & ~3;
    const int SWN = mode == 2 ? 4 : (mode ? 2 * IW : IW) + 2, SHN = mode == 2 ? 4 : (mode ? 2 * IH : IH) + 2, SW = (SWN + 3) & ~3,
              RP = (SWN + 32 + 15) & ~15;
#if FB_PP_PACKED
    const int si_bytes = (IH - FB_PP_TH / 2) * IP * 8;
#else
    const int si_bytes = IH * IP * 4;
#endif
    const int a = std::max(3 * FB_PP_TH * IP * 4, SHN * SW * 4), b = std::max(si_bytes, SHN * RP);
    return 16 + (size_t)a + (size_t)b;
}

// ---------------------------------------------------------------------------------------------
// S2-b generic (any ksize, any resize ratio), separable through shared memory.  Same per-sample arithmetic as
// k_fb_pyr (horizontal taps first, then the vertical combination, then the resize formula) but each blurred
// sample the resize touches is computed once per tile instead of once per output pixel:
//   1  resize source coordinates of the tile's columns / rows                  -> cols[2*TW], rows[2*TH]
//   2  uint8 source window (reflect-101 applied while loading)                 -> sS
//   3  horizontal blur at the needed columns for every window row              -> sHb
//   4  vertical blur at the needed rows                                        -> sB  (2TH x 2TW)
//   5  resize combine, write I
// dynamic smem (host-sized from the largest ratio of the batch): see fb_ps_smem().
// ---------------------------------------------------------------------------------------------
#define FB_PS_TW 32
#define FB_PS_TH 16

NSOF_GLOBAL k_fb_pyr_sep(const FbJob *NSOF_RESTRICT jobs, const uint2 *NSOF_RESTRICT tl, const float *NSOF_RESTRICT gk, int r,
                         int SWmax, int SHmax)
{
    constexpr int TW = FB_PS_TW, TH = FB_PS_TH;
    const uint2 te = tl[NSOF_BX];  // the list holds (item, image) entries of the generic-path items (fast == 0) only
    NSOF_JOB_TO_SMEM(FbJob, J, jobs + (te.x & FB_TL_ITEM));
    const int tile = (int)te.y;
    const int tiles_x = (J.w + TW - 1) / TW;
    const int img = (int)(te.x >> FB_TL_IMG_SHIFT);
    const uint8_t *src = img ? J.src1 : J.src0;
    const int sp = img ? J.spitch1 : J.spitch0;
    float *dst = J.I + (size_t)img * J.pitch * J.h;
    const int tx0 = (tile % tiles_x) * TW, ty0 = J.iy_lo + (tile / tiles_x) * TH;
    const int twn = nsof_imin(TW, J.w - tx0), thn = nsof_imin(TH, J.iy_hi - ty0);
    const int SWp = (SWmax + 32 + 15) & ~15;
    NSOF_DYN_SMEM(smem_raw);
    nsof_mbar_t *bar = (nsof_mbar_t *)smem_raw;
    int *cols = (int *)(smem_raw + 16), *rows = cols + 2 * TW;
    float *ax = (float *)(rows + 2 * TH), *ay = ax + TW, *sg = ay + TH;  // sg: r + 1 taps
    float *sB = sg + ((r + 1 + 3) & ~3);
    float *sHb = sB + 2 * TH * 2 * TW;
    uint8_t *sS = (uint8_t *)(sHb + (size_t)SHmax * 2 * TW);
    const int lane = NSOF_LANE, warp = NSOF_TID / NSOF_WARP, nw = nsof_imax(NSOF_NT / NSOF_WARP, 1);

    for (int i = NSOF_TID; i < TW + TH + r + 1; i += NSOF_NT) {
        if (i < TW + TH) {
            const bool isx = i < TW;
            const int d = isx ? nsof_imin(tx0 + i, J.w - 1) : nsof_imin(ty0 + (i - TW), J.h - 1);
            int i0, i1;
            float a;
            if (J.resize_mode == 0) { i0 = i1 = d; a = 0.f; }
            else if (J.resize_mode == 1) { i0 = 2 * d; i1 = 2 * d + 1; a = 0.5f; }
            else fb_lin_coord(d, isx ? J.sx : J.sy, isx ? J.sw : J.sh, &i0, &i1, &a);
            if (isx) { cols[2 * i] = i0; cols[2 * i + 1] = i1; ax[i] = a; }
            else { rows[2 * (i - TW)] = i0; rows[2 * (i - TW) + 1] = i1; ay[i - TW] = a; }
        } else {
            sg[i - TW - TH] = gk[i - TW - TH];
        }
    }
    if (NSOF_TID == 0) nsof_mbar_init(bar, 1);
    NSOF_SYNC();
    const int Xmin = cols[0] - r, Ymin = rows[0] - r;
    const int swn = nsof_imin(cols[2 * twn - 1] + r - Xmin + 1, SWmax), shn = nsof_imin(rows[2 * thn - 1] + r - Ymin + 1, SHmax);
    // The in-image part of the window arrives by one bulk asynchronous copy per row (16-byte aligned superset of
    // the row; the last source row only when the superset stays inside it); BORDER_REFLECT_101 is then applied in
    // shared memory (single bounce: the mirrored sample is always inside the window).  Images narrower than the
    // kernel radius, or with an unaligned pitch, take per-pixel loads.
    const int xs0 = nsof_imax(Xmin, 0), xs1 = nsof_imin(Xmin + swn, J.sw);
    const int ys0 = nsof_imax(Ymin, 0), ys1 = nsof_imin(Ymin + shn, J.sh);
    const bool bulk = J.sw > r && J.sh > r && (sp & 15) == 0 && xs1 > xs0 && ys1 > ys0;
    int shift = 0;
    if (bulk) {
        const uint8_t *a0 = src + (size_t)ys0 * sp + xs0;
        const int mis = (int)((uintptr_t)a0 & 15);
        shift = (mis - (xs0 - Xmin)) & 15;                     // window column sx lives at byte shift + sx of its row
        const unsigned row_bytes = (unsigned)((mis + (xs1 - xs0) + 15) & ~15);
        const bool last_safe = xs0 - mis + (int)row_bytes <= J.sw;  // the over-read of the last source row stays in it
        const int yb1 = (ys1 == J.sh && !last_safe) ? ys1 - 1 : ys1;
        if (NSOF_TID == 0) nsof_mbar_expect_tx(bar, row_bytes * (unsigned)(yb1 - ys0));
        NSOF_SYNC();
        for (int y = ys0 + NSOF_TID; y < yb1; y += NSOF_NT)
            nsof_bulk_g2s(sS + (y - Ymin) * SWp + shift + (xs0 - Xmin) - mis, a0 - mis + (size_t)(y - ys0) * sp, row_bytes, bar);
        if (yb1 < ys1) {
            const uint8_t *row = src + (size_t)yb1 * sp;
            for (int x = xs0 + NSOF_TID; x < xs1; x += NSOF_NT) sS[(yb1 - Ymin) * SWp + shift + (x - Xmin)] = NSOF_LDG(row + x);
        }
        if (NSOF_TID < NSOF_WARP) nsof_mbar_wait(bar, 0);  // one warp polls, the others sleep in the block barrier
        NSOF_SYNC();
        if (xs0 > Xmin || xs1 < Xmin + swn) {  // left / right reflection, in-image rows
            const int nl = xs0 - Xmin, nr = Xmin + swn - xs1;
            for (int i = NSOF_TID; i < (ys1 - ys0) * (nl + nr); i += NSOF_NT) {
                const int y = ys0 + i / (nl + nr), q = i % (nl + nr);
                const int X = q < nl ? Xmin + q : xs1 + (q - nl);
                uint8_t *row = sS + (y - Ymin) * SWp + shift;
                row[X - Xmin] = row[nsof_reflect101(X, J.sw) - Xmin];
            }
            NSOF_SYNC();
        }
        if (ys0 > Ymin || ys1 < Ymin + shn) {  // top / bottom reflection, whole window rows
            const int nt = ys0 - Ymin, nb = Ymin + shn - ys1;
            for (int i = NSOF_TID; i < (nt + nb) * swn; i += NSOF_NT) {
                const int q = i / swn, sx = i % swn;
                const int Y = q < nt ? Ymin + q : ys1 + (q - nt);
                sS[(Y - Ymin) * SWp + shift + sx] = sS[(nsof_reflect101(Y, J.sh) - Ymin) * SWp + shift + sx];
            }
            NSOF_SYNC();
        }
    } else {
        for (int sy = warp; sy < shn; sy += nw) {
            const uint8_t *row = src + (size_t)nsof_reflect101(Ymin + sy, J.sh) * sp;
            for (int sx = lane; sx < swn; sx += NSOF_WARP) sS[sy * SWp + sx] = NSOF_LDG(row + nsof_reflect101(Xmin + sx, J.sw));
        }
        NSOF_SYNC();
    }
    const float g0 = sg[0];
    // Each output is a serial chain of r dependent FMAs (the order of the sum is OpenCV's), and a CTA runs at 8 warps:
    // every thread carries FOUR independent chains (2 rows x 2 columns) through the tap loop so the chains hide each
    // other's latency.  Outputs outside the tile are computed on clamped addresses and not stored.
    for (int sy = 2 * warp; sy < shn; sy += 2 * nw) {
        const bool rb = sy + 1 < shn;
        const uint8_t *row0 = sS + sy * SWp + shift, *row1 = sS + (rb ? sy + 1 : sy) * SWp + shift;
        for (int ci = lane; ci < 2 * twn; ci += 2 * NSOF_WARP) {
            const bool cb = ci + NSOF_WARP < 2 * twn;
            const int o0 = cols[ci] - Xmin, o1 = cols[cb ? ci + NSOF_WARP : ci] - Xmin;
            const uint8_t *c00 = row0 + o0, *c01 = row0 + o1, *c10 = row1 + o0, *c11 = row1 + o1;
            float a00 = g0 * nsof_small_int_to_f(c00[0]), a01 = g0 * nsof_small_int_to_f(c01[0]);
            float a10 = g0 * nsof_small_int_to_f(c10[0]), a11 = g0 * nsof_small_int_to_f(c11[0]);
            // the two taps at distance i are added as integers (exact) and converted once
            for (int i = 1; i <= r; i++) {
                const float g = sg[i];
                a00 = a00 + g * nsof_small_int_to_f((int)c00[-i] + (int)c00[i]);
                a01 = a01 + g * nsof_small_int_to_f((int)c01[-i] + (int)c01[i]);
                a10 = a10 + g * nsof_small_int_to_f((int)c10[-i] + (int)c10[i]);
                a11 = a11 + g * nsof_small_int_to_f((int)c11[-i] + (int)c11[i]);
            }
            sHb[sy * 2 * TW + ci] = a00;
            if (cb) sHb[sy * 2 * TW + ci + NSOF_WARP] = a01;
            if (rb) {
                sHb[(sy + 1) * 2 * TW + ci] = a10;
                if (cb) sHb[(sy + 1) * 2 * TW + ci + NSOF_WARP] = a11;
            }
        }
    }
    NSOF_SYNC();
    for (int ri = 2 * warp; ri < 2 * thn; ri += 2 * nw) {
        const bool rb = ri + 1 < 2 * thn;
        const float *rowc0 = sHb + (rows[ri] - Ymin) * 2 * TW, *rowc1 = sHb + (rows[rb ? ri + 1 : ri] - Ymin) * 2 * TW;
        for (int ci = lane; ci < 2 * twn; ci += 2 * NSOF_WARP) {
            const bool cb = ci + NSOF_WARP < 2 * twn;
            const int c1 = cb ? ci + NSOF_WARP : ci;
            const float *c00 = rowc0 + ci, *c01 = rowc0 + c1, *c10 = rowc1 + ci, *c11 = rowc1 + c1;
            float a00 = g0 * c00[0], a01 = g0 * c01[0], a10 = g0 * c10[0], a11 = g0 * c11[0];
            for (int i = 1; i <= r; i++) {
                const float g = sg[i];
                const int d = i * 2 * TW;
                a00 = a00 + g * (c00[-d] + c00[d]);
                a01 = a01 + g * (c01[-d] + c01[d]);
                a10 = a10 + g * (c10[-d] + c10[d]);
                a11 = a11 + g * (c11[-d] + c11[d]);
            }
            sB[ri * 2 * TW + ci] = a00;
            if (cb) sB[ri * 2 * TW + c1] = a01;
            if (rb) {
                sB[(ri + 1) * 2 * TW + ci] = a10;
                if (cb) sB[(ri + 1) * 2 * TW + c1] = a11;
            }
        }
    }
    NSOF_SYNC();
    for (int ly = warp; ly < thn; ly += nw) {
        for (int lx = lane; lx < twn; lx += NSOF_WARP) {
            const float s00 = sB[(2 * ly) * 2 * TW + 2 * lx], s01 = sB[(2 * ly) * 2 * TW + 2 * lx + 1];
            const float s10 = sB[(2 * ly + 1) * 2 * TW + 2 * lx], s11 = sB[(2 * ly + 1) * 2 * TW + 2 * lx + 1];
            float v;
            if (J.resize_mode == 0) v = s00;
            else if (J.resize_mode == 1) v = (s00 + s01 + s10 + s11) * 0.25f;
            else {
                const float a = ax[lx], b = ay[ly];
                const float top = s00 * (1.f - a) + s01 * a;
                const float bot = s10 * (1.f - a) + s11 * a;
                v = top * (1.f - b) + bot * b;
            }
            dst[(size_t)(ty0 + ly) * J.pitch + tx0 + lx] = v;
        }
    }
}

static inline size_t fb_ps_smem(int r, int SWmax, int SHmax)
{
    const size_t SWp = (SWmax + 32 + 15) & ~15;
    return 16 + (2 * FB_PS_TW + 2 * FB_PS_TH) * sizeof(int) + (FB_PS_TW + FB_PS_TH + ((r + 1 + 3) & ~3)) * sizeof(float) +
           (size_t)2 * FB_PS_TH * 2 * FB_PS_TW * sizeof(float) + (size_t)SHmax * 2 * FB_PS_TW * sizeof(float) + SWp * SHmax;
}
