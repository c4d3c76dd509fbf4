// frames.cu -- the frame-side neighbours of the hot path (include/nsof.h: nsof_rgb2gray_u8*, nsof_remap_flow_*).
//
// Front end: `cv2.cvtColor(frame, cv2.COLOR_RGB2GRAY)` on the decoded BGR frame (optical_flow_seg.py:442-443 and the
// same lines of the other scripts) -- OpenCV's 8-bit path: (c0 * 9798 + c1 * 19235 + c2 * 3735 + 16384) >> 15 with c0
// the FIRST channel of the array as stored (the scripts hand a BGR array to an RGB2GRAY conversion; so do we).
// Back end: the prediction task of optical_flow_prediction.py:281-352 -- flow_map = (x, y) + flow as float32, then
// `cv2.remap(next_frame[:, :, c], map_x, map_y, INTER_LINEAR, borderMode=BORDER_REPLICATE)` per channel.  OpenCV's
// 8-bit bilinear remap is fixed point: coordinates rounded to 1/32 px (cvRound(v * 32)), weights
// (32 - ax)(32 - ay) * 32 ... with 15 fractional bits, (sum + 2^14) >> 15, taps clamped to the image.
// Both restatements match cv2 4.13 bit for bit (tests/golden/frames_reference.npz).
#include "nsof_handle.h"

#include <algorithm>
#include <climits>
#include <cstring>

struct GreyJob {
    const uint8_t *src;
    uint8_t *dst;
    ptrdiff_t spitch, sstride, dpitch, dstride;
    int H, W;
};

// grid (chunks, frames): 4 pixels per thread
NSOF_GLOBAL k_rgb2gray_u8(GreyJob J)
{
    const uint8_t *src = J.src + (size_t)NSOF_BY * J.sstride;
    uint8_t *dst = J.dst + (size_t)NSOF_BY * J.dstride;
    const int qw = (J.W + 3) / 4, total = J.H * qw;
    for (int idx = NSOF_BX * NSOF_NT + NSOF_TID; idx < total; idx += nsof_grid_x() * NSOF_NT) {
        const int y = idx / qw, x = (idx - y * qw) * 4;
        const uint8_t *s = src + (size_t)y * J.spitch + 3 * (size_t)x;
        uint8_t *o = dst + (size_t)y * J.dpitch + x;
        const int nq = nsof_imin(4, J.W - x);
        uint32_t packed = 0;
        if (nq == 4 && (((uintptr_t)s) & 3) == 0) {  // 12 source bytes as three aligned words
            const uint32_t w0 = NSOF_LDG((const uint32_t *)s), w1 = NSOF_LDG((const uint32_t *)s + 1), w2 = NSOF_LDG((const uint32_t *)s + 2);
            const uint32_t b[12] = {w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u, w0 >> 24, w1 & 255u, (w1 >> 8) & 255u,
                                    (w1 >> 16) & 255u, w1 >> 24, w2 & 255u, (w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24};
#pragma unroll
            for (int q = 0; q < 4; q++)
                packed |= ((b[3 * q] * 9798u + b[3 * q + 1] * 19235u + b[3 * q + 2] * 3735u + 16384u) >> 15) << (8 * q);
        } else {
            for (int q = 0; q < nq; q++) {
                const uint32_t g = ((uint32_t)NSOF_LDG(s + 3 * q) * 9798u + (uint32_t)NSOF_LDG(s + 3 * q + 1) * 19235u +
                                    (uint32_t)NSOF_LDG(s + 3 * q + 2) * 3735u + 16384u) >> 15;
                packed |= g << (8 * q);
            }
        }
        if (nq == 4 && (((uintptr_t)o) & 3) == 0) *(uint32_t *)o = packed;
        else for (int q = 0; q < nq; q++) o[q] = (uint8_t)(packed >> (8 * q));
    }
}

struct WarpJob {
    const uint8_t *frame;
    const float *flow;
    uint8_t *out;
    ptrdiff_t fpitch_bytes, opitch;
    int flow_pitch;  // floats
    int H, W, ch, x0, y0, h, w;
};

// cvRound of a float on x86 (cvtss2si): round half to even, "integer indefinite" for NaN / out of range
NSOF_HD int frames_cv_round(float v)
{
    if (!(fabsf(v) < 2147483648.f)) return INT_MIN;
#if defined(__CUDA_ARCH__)
    return __float2int_rn(v);
#else
    return (int)lrintf(v);
#endif
}

// one output pixel: fixed-point coordinates, four clamped taps per channel
NSOF_DEV void remap_pixel(const WarpJob &J, int x, int y, float sign, uint8_t *px)
{
    const float2 f = *(const float2 *)(J.flow + (size_t)y * J.flow_pitch + 2 * x);
    // flow_map = (coords + flow).astype(float32): the sum is formed in double (int64 + float promote to float64)
    const float mx = (float)((double)(J.x0 + x) + (double)(sign * f.x));
    const float my = (float)((double)(J.y0 + y) + (double)(sign * f.y));
    const int sx = frames_cv_round(mx * 32.f), sy = frames_cv_round(my * 32.f);
    const int ax = sx & 31, ay = sy & 31;
    const int ix = nsof_clampi(sx >> 5, -32768, 32767), iy = nsof_clampi(sy >> 5, -32768, 32767);  // saturate_cast<short>
    const int xa = nsof_clampi(ix, 0, J.W - 1), xb = nsof_clampi(ix + 1, 0, J.W - 1);
    const int ya = nsof_clampi(iy, 0, J.H - 1), yb = nsof_clampi(iy + 1, 0, J.H - 1);
    const int w00 = (32 - ax) * (32 - ay) * 32, w01 = ax * (32 - ay) * 32, w10 = (32 - ax) * ay * 32, w11 = ax * ay * 32;
    const uint8_t *ra = J.frame + (size_t)ya * J.fpitch_bytes, *rb = J.frame + (size_t)yb * J.fpitch_bytes;
    for (int c = 0; c < J.ch; c++) {
        const int v = (int)NSOF_LDG(ra + xa * J.ch + c) * w00 + (int)NSOF_LDG(ra + xb * J.ch + c) * w01 +
                      (int)NSOF_LDG(rb + xa * J.ch + c) * w10 + (int)NSOF_LDG(rb + xb * J.ch + c) * w11;
        px[c] = (uint8_t)((v + 16384) >> 15);
    }
}

// grid (chunks, regions): four output pixels (all channels) per thread; for 3-channel images whose output rows are
// 4-byte aligned the 12 bytes leave as three words
NSOF_GLOBAL k_remap_flow(const WarpJob *NSOF_RESTRICT jobs, float sign)
{
    const WarpJob J = jobs[NSOF_BY];
    const int qw = (J.w + 3) / 4, total = J.h * qw;
    for (int idx = NSOF_BX * NSOF_NT + NSOF_TID; idx < total; idx += nsof_grid_x() * NSOF_NT) {
        const int y = idx / qw, x = (idx - y * qw) * 4;
        uint8_t *o = J.out + (size_t)y * J.opitch + (size_t)x * J.ch;
        const int nq = nsof_imin(4, J.w - x);
        if (J.ch == 3 && nq == 4 && (((uintptr_t)o) & 3) == 0) {
            uint8_t b[12];
#pragma unroll
            for (int q = 0; q < 4; q++) remap_pixel(J, x + q, y, sign, b + 3 * q);
            uint32_t *o4 = (uint32_t *)o;
#pragma unroll
            for (int k = 0; k < 3; k++)
                o4[k] = (uint32_t)b[4 * k] | ((uint32_t)b[4 * k + 1] << 8) | ((uint32_t)b[4 * k + 2] << 16) | ((uint32_t)b[4 * k + 3] << 24);
        } else {
            for (int q = 0; q < nq; q++) {
                uint8_t b[4];
                remap_pixel(J, x + q, y, sign, b);
                for (int c = 0; c < J.ch; c++) o[q * J.ch + c] = b[c];
            }
        }
    }
}

extern "C" int nsof_rgb2gray_u8(nsof_handle *h, const uint8_t *src, ptrdiff_t src_pitch, ptrdiff_t src_frame_stride, int n,
                                int H, int W, uint8_t *dst, ptrdiff_t dst_pitch, ptrdiff_t dst_frame_stride, void *cuda_stream)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "rgb2gray: NULL handle");
    if (n < 0 || H < 1 || W < 1 || (n > 0 && (!src || !dst))) NSOF_FAIL(NSOF_ERR_BAD_ARG, "rgb2gray: NULL / empty frames");
    if (src_pitch < (ptrdiff_t)W * 3 || dst_pitch < W) NSOF_FAIL(NSOF_ERR_BAD_ARG, "rgb2gray: pitch smaller than a row");
    if (n == 0) return NSOF_OK;
    if (n > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "rgb2gray: at most 65535 frames per call");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    GreyJob J = {src, dst, src_pitch, src_frame_stride, dst_pitch, dst_frame_stride, H, W};
    const int gx = std::max(1, std::min((H * ((W + 3) / 4) + 255) / 256, 2048));
    NSOF_KLAUNCH(h, NSOF_K_RGB2GRAY, 4. * n * H * W, k_rgb2gray_u8, gx, n, 1, 256, 0, (cudaStream_t)cuda_stream, J);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_rgb2gray_u8_host(nsof_handle *h, const uint8_t *src, ptrdiff_t src_pitch, int H, int W, uint8_t *dst,
                                     ptrdiff_t dst_pitch)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "rgb2gray_host: NULL handle");
    if (!src || !dst || H < 1 || W < 1 || src_pitch < (ptrdiff_t)W * 3 || dst_pitch < W)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "rgb2gray_host: NULL / empty frame or pitch smaller than a row");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t sb = (((size_t)W * 3 * H + 255) / 256) * 256;
    int rc = h->io.ensure(sb + (size_t)W * H + 16);
    if (rc) return rc;
    uint8_t *d_src = (uint8_t *)h->io.p, *d_dst = d_src + sb;
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(d_src, (size_t)W * 3, src, (size_t)src_pitch, (size_t)W * 3, H, cudaMemcpyHostToDevice, nullptr));
    rc = nsof_rgb2gray_u8(h, d_src, (ptrdiff_t)W * 3, 0, 1, H, W, d_dst, W, 0, nullptr);
    if (rc) return rc;
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)dst_pitch, d_dst, (size_t)W, (size_t)W, H, cudaMemcpyDeviceToHost, nullptr));
    NSOF_CUDA_TRY(cudaStreamSynchronize(nullptr));
    return NSOF_OK;
}

extern "C" int nsof_remap_flow_batch(nsof_handle *h, const nsof_warp_desc *items, int n, int flow_sign, void *cuda_stream)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "remap_flow: NULL handle");
    if (n < 0 || (n > 0 && !items)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "remap_flow: bad item list");
    if (flow_sign != 1 && flow_sign != -1) NSOF_FAIL(NSOF_ERR_BAD_ARG, "remap_flow: flow_sign must be +1 or -1");
    if (n == 0) return NSOF_OK;
    if (n > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "remap_flow: at most 65535 regions per call");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t bytes = (size_t)n * sizeof(WarpJob);
    NsofStageSlot *slot;
    int rc = nsof_stage_acquire(h, bytes, &slot);
    if (rc) return rc;
    rc = h->warp_tab.ensure(bytes + 16);
    if (rc) return rc;
    WarpJob *hj = (WarpJob *)slot->host;
    int max_px = 0;
    double px = 0.;
    for (int i = 0; i < n; i++) {
        const nsof_warp_desc &d = items[i];
        if (!d.frame || !d.flow || !d.out || d.H < 1 || d.W < 1 || d.h < 1 || d.w < 1 || d.channels < 1 || d.channels > 4 ||
            d.frame_pitch < (ptrdiff_t)d.W * d.channels || d.out_pitch < (ptrdiff_t)d.w * d.channels ||
            d.flow_pitch < (ptrdiff_t)d.w * 8 || d.flow_pitch % 8 || (uintptr_t)d.flow % 8)
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "remap_flow: region %d is malformed", i);
        if ((double)d.h * d.w > 2e9) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "remap_flow: region %d too large", i);
        WarpJob &J = hj[i];
        J.frame = d.frame; J.flow = d.flow; J.out = d.out;
        J.fpitch_bytes = d.frame_pitch; J.opitch = d.out_pitch; J.flow_pitch = (int)(d.flow_pitch / 4);
        J.H = d.H; J.W = d.W; J.ch = d.channels; J.x0 = d.x0; J.y0 = d.y0; J.h = d.h; J.w = d.w;
        max_px = std::max(max_px, d.h * d.w);
        px += (double)d.h * d.w * (8 + 2 * d.channels);
    }
    rc = nsof_table_upload(h, h->warp_tab.p, slot, bytes, stream);
    if (rc) return rc;
    const int gx = std::max(1, std::min((max_px / 4 + 255) / 256, 2048));
    NSOF_KLAUNCH(h, NSOF_K_REMAP, px, k_remap_flow, gx, n, 1, 256, 0, stream, (const WarpJob *)h->warp_tab.p, (float)flow_sign);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_remap_flow_host(nsof_handle *h, const uint8_t *frame, ptrdiff_t frame_pitch, int H, int W, int channels,
                                    const float *flow, ptrdiff_t flow_pitch, int x0, int y0, int height, int width,
                                    int flow_sign, uint8_t *out, ptrdiff_t out_pitch)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "remap_flow_host: NULL handle");
    if (!frame || !flow || !out || H < 1 || W < 1 || height < 1 || width < 1 || channels < 1 || channels > 4 ||
        frame_pitch < (ptrdiff_t)W * channels || flow_pitch < (ptrdiff_t)width * 8 || out_pitch < (ptrdiff_t)width * channels)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "remap_flow_host: NULL / empty arrays or pitch smaller than a row");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    auto up = [](size_t v) { return ((v + 255) / 256) * 256; };
    const size_t fb = up((size_t)W * channels * H), lb = up((size_t)width * 8 * height), ob = (size_t)width * channels * height;
    int rc = h->io.ensure(fb + lb + ob + 16);
    if (rc) return rc;
    uint8_t *d_frame = (uint8_t *)h->io.p, *d_out = d_frame + fb + lb;
    float *d_flow = (float *)(d_frame + fb);
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(d_frame, (size_t)W * channels, frame, (size_t)frame_pitch, (size_t)W * channels, H,
                                    cudaMemcpyHostToDevice, nullptr));
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(d_flow, (size_t)width * 8, flow, (size_t)flow_pitch, (size_t)width * 8, height,
                                    cudaMemcpyHostToDevice, nullptr));
    nsof_warp_desc d = {d_frame, (ptrdiff_t)W * channels, H, W, channels, d_flow, (ptrdiff_t)width * 8, x0, y0, height, width,
                        d_out, (ptrdiff_t)width * channels};
    rc = nsof_remap_flow_batch(h, &d, 1, flow_sign, nullptr);
    if (rc) return rc;
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(out, (size_t)out_pitch, d_out, (size_t)width * channels, (size_t)width * channels, height,
                                    cudaMemcpyDeviceToHost, nullptr));
    NSOF_CUDA_TRY(cudaStreamSynchronize(nullptr));
    return NSOF_OK;
}


// ---------------------------------------------------------------------------------------------------------------------
// SSIM of the prediction task (optical_flow_prediction.py:113-115): skimage.metrics.structural_similarity(true[:, :, 2],
// prediction[:, :, 2], data_range=255.0) with its defaults -- 7 x 7 uniform window, sample covariance (x 49 / 48),
// K1 = 0.01, K2 = 0.03, mean over the image cropped by 3 pixels on every side (so every window lies inside the
// image and scipy's boundary mode never matters).  scikit-image is not vendored in the reference (requirements.txt pins
// scikit-image); restated from its published algorithm, oracle/frames_oracle.py `ssim_u8`.
// Both inputs are uint8, so the five window sums (x, y, xx, yy, xy) are exact integers; the per-pixel S and the mean are
// float64.  Partial sums per CTA are reduced in a fixed order by a second launch: the result is deterministic.
// ---------------------------------------------------------------------------------------------------------------------
#define SSIM_TW 32
#define SSIM_TH 32
struct SsimJob {
    const uint8_t *a, *b;
    ptrdiff_t apitch, bpitch;
    int astride, bstride;  // bytes between pixels (3 for channel 2 of an interleaved BGR frame, 1 for a grey plane)
    int H, W;
    double C1, C2;
    double *partial;  // one per CTA
};

NSOF_GLOBAL k_ssim_tiles(SsimJob J)
{
    constexpr int LW = SSIM_TW + 6, LH = SSIM_TH + 6;
    const int iw = J.W - 6, ih = J.H - 6;                     // interior (cropped) size
    const int tiles_x = (iw + SSIM_TW - 1) / SSIM_TW;
    const int tx0 = (NSOF_BX % tiles_x) * SSIM_TW, ty0 = (NSOF_BX / tiles_x) * SSIM_TH;  // interior coordinates
    NSOF_DYN_SMEM(smem_raw);
    int *sa = (int *)smem_raw, *sb = sa + LW * LH;              // pixels
    int *hs = sb + LW * LH;                                     // horizontal 7-sums: 5 planes of LH x SSIM_TW
    for (int i = NSOF_TID; i < LW * LH; i += NSOF_NT) {
        const int lx = i % LW, ly = i / LW;
        const int x = nsof_imin(tx0 + lx, J.W - 1), y = nsof_imin(ty0 + ly, J.H - 1);  // clamped loads feed unused outputs only
        sa[i] = NSOF_LDG(J.a + (size_t)y * J.apitch + (size_t)x * J.astride);
        sb[i] = NSOF_LDG(J.b + (size_t)y * J.bpitch + (size_t)x * J.bstride);
    }
    NSOF_SYNC();
    for (int i = NSOF_TID; i < SSIM_TW * LH; i += NSOF_NT) {
        const int lx = i % SSIM_TW, ly = i / SSIM_TW;
        int s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0;
#pragma unroll
        for (int k = 0; k < 7; k++) {
            const int u = sa[ly * LW + lx + k], v = sb[ly * LW + lx + k];
            s0 += u; s1 += v; s2 += u * u; s3 += v * v; s4 += u * v;
        }
        int *o = hs + i;
        o[0] = s0; o[SSIM_TW * LH] = s1; o[2 * SSIM_TW * LH] = s2; o[3 * SSIM_TW * LH] = s3; o[4 * SSIM_TW * LH] = s4;
    }
    NSOF_SYNC();
    double acc = 0.;
    for (int i = NSOF_TID; i < SSIM_TW * SSIM_TH; i += NSOF_NT) {
        const int lx = i % SSIM_TW, ly = i / SSIM_TW;
        if (tx0 + lx >= iw || ty0 + ly >= ih) continue;
        int s[5] = {0, 0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < 7; k++)
#pragma unroll
            for (int c = 0; c < 5; c++) s[c] += hs[c * SSIM_TW * LH + (ly + k) * SSIM_TW + lx];
        // uniform_filter means, then skimage's formula (cov_norm = NP / (NP - 1), NP = 49)
        const double inv = 1. / 49., cn = 49. / 48.;
        const double ux = s[0] * inv, uy = s[1] * inv, uxx = s[2] * inv, uyy = s[3] * inv, uxy = s[4] * inv;
        const double vx = cn * (uxx - ux * ux), vy = cn * (uyy - uy * uy), vxy = cn * (uxy - ux * uy);
        const double A1 = 2 * ux * uy + J.C1, A2 = 2 * vxy + J.C2, B1 = ux * ux + uy * uy + J.C1, B2 = vx + vy + J.C2;
        acc += (A1 * A2) / (B1 * B2);
    }
    // fixed-order block reduction through shared memory (reuses sa: 8-byte aligned)
    double *red = (double *)smem_raw;
    NSOF_SYNC();
    red[NSOF_TID] = acc;
    NSOF_SYNC();
    if (NSOF_TID == 0) {
        double t = 0.;
        for (int i = 0; i < NSOF_NT; i++) t += red[i];
        J.partial[NSOF_BX] = t;
    }
}

NSOF_GLOBAL k_ssim_finish(const double *partial, int n, double count, double *out)
{
    if (NSOF_TID != 0 || NSOF_BX != 0) return;
    double t = 0.;
    for (int i = 0; i < n; i++) t += partial[i];
    *out = t / count;
}

extern "C" int nsof_ssim_u8(nsof_handle *h, const uint8_t *a, ptrdiff_t a_pitch, int a_pixel_stride, const uint8_t *b,
                            ptrdiff_t b_pitch, int b_pixel_stride, int height, int width, double data_range, double *out_dev,
                            void *cuda_stream)
{
    if (!h || !a || !b || !out_dev) NSOF_FAIL(NSOF_ERR_BAD_ARG, "ssim: NULL argument");
    if (height < 7 || width < 7) NSOF_FAIL(NSOF_ERR_BAD_ARG, "ssim: win_size 7 exceeds the image extent (skimage raises ValueError)");
    if (a_pixel_stride < 1 || b_pixel_stride < 1 || a_pitch < (ptrdiff_t)width * a_pixel_stride || b_pitch < (ptrdiff_t)width * b_pixel_stride)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "ssim: bad pitch / pixel stride");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    const int iw = width - 6, ih = height - 6;
    const int tiles = ((iw + SSIM_TW - 1) / SSIM_TW) * ((ih + SSIM_TH - 1) / SSIM_TH);
    int rc = h->warp_tab.ensure((size_t)tiles * sizeof(double) + 64);
    if (rc) return rc;
    SsimJob J = {a, b, a_pitch, b_pitch, a_pixel_stride, b_pixel_stride, height, width,
                 (0.01 * data_range) * (0.01 * data_range), (0.03 * data_range) * (0.03 * data_range), (double *)h->warp_tab.p};
    const size_t smem = (2 * (SSIM_TW + 6) * (SSIM_TH + 6) + 5 * SSIM_TW * (SSIM_TH + 6)) * sizeof(int);
    NSOF_KLAUNCH(h, NSOF_K_SSIM, 2. * height * width, k_ssim_tiles, tiles, 1, 1, 256, smem, stream, J);
    NSOF_KLAUNCH(h, NSOF_K_SSIM, 8. * tiles, k_ssim_finish, 1, 1, 1, 32, 0, stream, (const double *)h->warp_tab.p, tiles,
                 (double)iw * ih, out_dev);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

// host arrays in, the score out (synchronous)
extern "C" int nsof_ssim_u8_host(nsof_handle *h, const uint8_t *a, ptrdiff_t a_pitch, int a_pixel_stride, const uint8_t *b,
                                 ptrdiff_t b_pitch, int b_pixel_stride, int height, int width, double data_range, double *out)
{
    if (!h || !a || !b || !out) NSOF_FAIL(NSOF_ERR_BAD_ARG, "ssim: NULL argument");
    if (height < 1 || width < 1 || a_pixel_stride < 1 || b_pixel_stride < 1) NSOF_FAIL(NSOF_ERR_BAD_ARG, "ssim: bad geometry");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    // bytes of a row that hold pixels: the last pixel is ONE byte (a channel view of an interleaved frame starts at the
    // channel's offset, so width * stride bytes from there would run past the end of the caller's array)
    const size_t rowa = (size_t)(width - 1) * a_pixel_stride + 1, rowb = (size_t)(width - 1) * b_pixel_stride + 1;
    const size_t pa = (rowa + 255) & ~(size_t)255, pb = (rowb + 255) & ~(size_t)255;
    int rc = h->io.ensure(pa * height + pb * height + 256);
    if (rc) return rc;
    uint8_t *da = (uint8_t *)h->io.p, *db = da + pa * height;
    double *dout = (double *)(db + pb * height + (256 - (pb * height) % 256) % 256);
    cudaStream_t stream = nullptr;
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(da, pa, a, a_pitch, rowa, height, cudaMemcpyHostToDevice, stream));
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(db, pb, b, b_pitch, rowb, height, cudaMemcpyHostToDevice, stream));
    rc = nsof_ssim_u8(h, da, (ptrdiff_t)pa, a_pixel_stride, db, (ptrdiff_t)pb, b_pixel_stride, height, width, data_range, dout, stream);
    if (rc) return rc;
    NSOF_CUDA_TRY(cudaMemcpyAsync(out, dout, sizeof(double), cudaMemcpyDeviceToHost, stream));
    NSOF_CUDA_TRY(cudaStreamSynchronize(stream));
    return NSOF_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// nsof_ssim_u8_noalloc: the SSIM above for callers that may not allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The per-CTA partial sums live in a file-scope device array instead of the handle's growable scratch buffer, so the
// call allocates nothing (no cudaMalloc / cudaFree, no growable buffer).  Limits and rules:
//   * at most NSOF_SSIM_NOALLOC_MAX_TILES (4096) tiles of 32 x 32 interior pixels, e.g. 1080 x 1920 (2040 tiles) or
//     2048 x 2048 (4096); larger images are rejected with NSOF_ERR_UNSUPPORTED;
//   * the array (32 KB) is held by every GPU that loads the library until the process exits;
//   * every call on a device shares it: calls that overlap on the GPU (different streams) race -- serialise them;
//   * it keeps the last call's contents; nothing has to start at zero (k_ssim_tiles writes partial[0 .. tiles) before
//     k_ssim_finish reads exactly that range).
// In host code `g_ssim_scratch` is only a stand-in for the symbol: its device address is taken once per device, with
// that device current, through cudaGetSymbolAddress.
// ---------------------------------------------------------------------------------------------------------------------
#define NSOF_SSIM_NOALLOC_MAX_TILES 4096
#define NSOF_MAX_DEVICES 64
#ifdef NSOF_EMU
static double g_ssim_scratch[NSOF_SSIM_NOALLOC_MAX_TILES];
#else
static __device__ __align__(256) double g_ssim_scratch[NSOF_SSIM_NOALLOC_MAX_TILES];
#endif
static double *g_ssim_scratch_ptr[NSOF_MAX_DEVICES];  // per device; written once (benign if two threads race: same value)

static int ssim_scratch_address(int device, double **out)
{
    if (device < 0 || device >= NSOF_MAX_DEVICES) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "ssim_noalloc: device %d", device);
    if (!g_ssim_scratch_ptr[device]) {
#ifdef NSOF_EMU
        g_ssim_scratch_ptr[device] = g_ssim_scratch;
#else
        double *p = nullptr;
        NSOF_CUDA_TRY(cudaGetSymbolAddress((void **)&p, g_ssim_scratch));  // the device is current (caller set it)
        g_ssim_scratch_ptr[device] = p;
#endif
    }
    *out = g_ssim_scratch_ptr[device];
    return NSOF_OK;
}

extern "C" int nsof_ssim_u8_noalloc(nsof_handle *h, const uint8_t *a, ptrdiff_t a_pitch, int a_pixel_stride, const uint8_t *b,
                                    ptrdiff_t b_pitch, int b_pixel_stride, int height, int width, double data_range,
                                    double *out_dev, void *cuda_stream)
{
    if (!h || !a || !b || !out_dev) NSOF_FAIL(NSOF_ERR_BAD_ARG, "ssim_noalloc: NULL argument");
    if (height < 7 || width < 7) NSOF_FAIL(NSOF_ERR_BAD_ARG, "ssim_noalloc: win_size 7 exceeds the image extent");
    if (a_pixel_stride < 1 || b_pixel_stride < 1 || a_pitch < (ptrdiff_t)width * a_pixel_stride || b_pitch < (ptrdiff_t)width * b_pixel_stride)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "ssim_noalloc: bad pitch / pixel stride");
    const int iw = width - 6, ih = height - 6;
    const long long tiles = (long long)((iw + SSIM_TW - 1) / SSIM_TW) * ((ih + SSIM_TH - 1) / SSIM_TH);
    if (tiles > NSOF_SSIM_NOALLOC_MAX_TILES)
        NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "ssim_noalloc: %d x %d needs %lld tiles of 32 x 32, the static scratch holds %d (use nsof_ssim_u8)",
                  height, width, tiles, NSOF_SSIM_NOALLOC_MAX_TILES);
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    double *scratch = nullptr;
    int rc = ssim_scratch_address(h->device, &scratch);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    SsimJob J = {a, b, a_pitch, b_pitch, a_pixel_stride, b_pixel_stride, height, width,
                 (0.01 * data_range) * (0.01 * data_range), (0.03 * data_range) * (0.03 * data_range), scratch};
    const size_t smem = (2 * (SSIM_TW + 6) * (SSIM_TH + 6) + 5 * SSIM_TW * (SSIM_TH + 6)) * sizeof(int);
    NSOF_KLAUNCH(h, NSOF_K_SSIM, 2. * height * width, k_ssim_tiles, (int)tiles, 1, 1, 256, smem, stream, J);
    NSOF_KLAUNCH(h, NSOF_K_SSIM, 8. * tiles, k_ssim_finish, 1, 1, 1, 32, 0, stream, (const double *)scratch, (int)tiles,
                 (double)iw * ih, out_dev);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}
