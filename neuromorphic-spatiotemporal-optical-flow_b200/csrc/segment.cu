// segment.cu -- motion segmentation of flow regions on bit planes (include/nsof.h: nsof_flow_segment_*).
//
// Reference: task_results + process_flow_region (optical_flow_seg.py:255-320, 322-357; optical_flow_yolo.py has the
// same code).  Per region: mask = (|flow| > SEG_TH) with cv2.cartToPolar's magnitude (float32 arithmetic even for the
// reference's float64 flow array), then `iterations` x [dilate, erode] with a cv2 structuring element, then
// "> 1 -> 255".
//
// Layout: one bit per pixel, 32 pixels per word (bit b of word X = pixel 32 (X - 1) + b), every row carries one zero
// word on either side so a thread reads the words X - 1, X, X + 1 of a row without bounds tests.  A region of an ROI
// crop (560 x 420) is 29 KB per plane: all passes of a step run out of L2.  One thread per output word; the rows of
// the element that share one horizontal run are OR-ed vertically first, then the run is applied once (the 10 x 10
// ellipse: 4 groups, 27 shifts instead of 78).  erode(x) = ~dilate(~x) with out-of-image bits held at zero on both
// sides, which is exactly cv2's "pixels outside the image do not take part".
#include "nsof_handle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

struct SegJob {
    const float *flow;
    uint8_t *mask;
    uint32_t *A, *B;  // bit planes, h x pw words each
    int fpitch;       // floats per flow row
    int mpitch;       // bytes per mask row
    int h, w, pw;
};

#define SEG_MAX_GROUPS 32
struct SegElement {
    int n_groups, ay;               // anchor row
    int lo[SEG_MAX_GROUPS], hi[SEG_MAX_GROUPS];  // horizontal run of the group, offsets relative to the anchor column
    uint32_t rows[SEG_MAX_GROUPS];  // bit i = element row i belongs to the group
    int kh;                         // element rows
    signed char grp_of_row[32];     // group of element row i, -1 when the row is empty
    int nested;                     // groups are sorted by length and every run contains the previous one
};

// bits of word X (1-based inside the padded row) that are pixels of the image
NSOF_HD uint32_t seg_valid(int X, int w)
{
    const int x0 = (X - 1) * 32;
    if (X < 1 || x0 >= w) return 0u;
    return w - x0 >= 32 ? 0xffffffffu : ((1u << (w - x0)) - 1u);
}

// cv2.cartToPolar converts float64 input to float32 and computes mag = sqrtf(fmaf(x, x, y * y)) (pinned by the
// golden masks: tests/golden/segment_reference.npz, case edge_values).  mag > th  <=>  fmaf(x, x, y * y) > s_star with
// s_star the largest float whose correctly rounded square root is still <= th (host-computed): no sqrt per pixel.
NSOF_DEV void seg_pack_words(const SegJob &J, uint32_t *NSOF_RESTRICT dst, float s_star, int first, int stride)
{
    const int total = J.h * J.pw;
    for (int idx = first; idx < total; idx += stride) {
        const int y = idx / J.pw, X = idx - y * J.pw;
        uint32_t word = 0;
        const int x0 = (X - 1) * 32;
        if (X >= 1 && x0 < J.w) {
            const float2 *row = (const float2 *)(J.flow + (size_t)y * J.fpitch) + x0;
            const int nb = nsof_imin(32, J.w - x0);
            for (int b = 0; b < nb; b++) {
                const float2 f = row[b];
#if defined(__CUDA_ARCH__)
                const float s = __fmaf_rn(f.x, f.x, __fmul_rn(f.y, f.y));
#else
                volatile float yy = f.y * f.y;  // rounded product, then one fused multiply-add
                const float s = fmaf(f.x, f.x, yy);
#endif
                word |= (s > s_star ? 1u : 0u) << b;
            }
        }
        dst[idx] = word;
        if (X == 0 || X == J.pw - 1) J.B[idx] = 0u;  // the tiled pass never writes the pad words of either plane
    }
}

// One morphological pass: dst = dilate(src) or erode(src).
NSOF_DEV void seg_morph_words(const SegJob &J, const uint32_t *NSOF_RESTRICT src, uint32_t *NSOF_RESTRICT dst,
                              const SegElement &E, int erode, int first, int stride)
{
    const int total = J.h * J.pw;
    for (int idx = first; idx < total; idx += stride) {
        const int y = idx / J.pw, X = idx - y * J.pw;
        const uint32_t vb = seg_valid(X, J.w);
        if (!vb) { dst[idx] = 0u; continue; }
        const uint32_t va = seg_valid(X - 1, J.w), vc = seg_valid(X + 1, J.w);
        uint32_t acc = 0u;
        for (int g = 0; g < E.n_groups; g++) {
            uint32_t a = 0u, b = 0u, c = 0u;
            for (uint32_t m = E.rows[g]; m; m &= m - 1) {
#if defined(__CUDA_ARCH__)
                const int i = __ffs((int)m) - 1;
#else
                const int i = __builtin_ctz(m);
#endif
                const int yy = y + i - E.ay;
                if (yy < 0 || yy >= J.h) continue;
                const uint32_t *r = src + (size_t)yy * J.pw + X;
                uint32_t ra = r[-1], rb = r[0], rc = r[1];
                if (erode) { ra = ~ra & va; rb = ~rb & vb; rc = ~rc & vc; }
                a |= ra; b |= rb; c |= rc;
            }
            if (!(a | b | c)) continue;
            // pixel x + d of the row: bit 32 + d of b:a for d < 0, bit d of c:b for d >= 0 -- two branch-free shift loops
            // (the first version selected the word pair per tap: 49 % of the kernel's instructions, ncu)
            const uint64_t lo64 = ((uint64_t)b << 32) | a, hi64 = ((uint64_t)c << 32) | b;
            const int d0 = E.lo[g], d1 = E.hi[g];
            for (int d = d0; d <= nsof_imin(d1, -1); d++) acc |= (uint32_t)(lo64 >> (32 + d));
            for (int d = nsof_imax(d0, 0); d <= d1; d++) acc |= (uint32_t)(hi64 >> d);
        }
        dst[idx] = (erode ? ~acc : acc) & vb;
    }
}

// bit plane -> uint8 0 / 255 (the final threshold(., 1, 255) of the reference), 4 pixels per thread
NSOF_DEV void seg_unpack_bytes(const SegJob &J, const uint32_t *NSOF_RESTRICT src, int first, int stride)
{
    const int qw = (J.w + 3) / 4, total = J.h * qw;
    for (int idx = first; idx < total; idx += stride) {
        const int y = idx / qw, x = (idx - y * qw) * 4;
        const uint32_t bits = (src[(size_t)y * J.pw + 1 + (x >> 5)] >> (x & 31)) & 0xfu;
        uint8_t *o = J.mask + (size_t)y * J.mpitch + x;
        if (x + 4 <= J.w && (((uintptr_t)o) & 3) == 0) {
            *(uint32_t *)o = (bits & 1u ? 0x000000ffu : 0u) | (bits & 2u ? 0x0000ff00u : 0u) | (bits & 4u ? 0x00ff0000u : 0u) |
                             (bits & 8u ? 0xff000000u : 0u);
        } else {
            for (int q = 0; q < 4 && x + q < J.w; q++) o[q] = (bits >> q) & 1u ? 255 : 0;
        }
    }
}

// grid (chunks, regions): regions of any size, planes in global memory (L2-resident).
// Measured alternative (B200, 30 ROI regions per launch): ONE launch per batch with both planes of a region in the
// shared memory of one CTA -- 0.50 ms per batch against 0.14 ms for these 12 launches: a pass is ~150 instructions
// per word and a region is 7 350 words, so ten passes on one SM lose to ten launches spread over all 148.
NSOF_GLOBAL k_seg_pack(const SegJob *NSOF_RESTRICT jobs, float s_star)
{
    const SegJob J = jobs[NSOF_BY];
#ifdef NSOF_EMU
    seg_pack_words(J, J.A, s_star, NSOF_BX * NSOF_NT + NSOF_TID, nsof_grid_x() * NSOF_NT);
#else
    // A warp produces 32 consecutive words: for word k the 32 lanes test its 32 pixels (one coalesced 256-byte load
    // per instruction) and a ballot assembles the word; lane k keeps it.  One thread walking its own 32 pixels put
    // the lanes of every load 256 bytes apart (32 lines per instruction, L1 thrash: 1.3 TB/s).
    const int total = J.h * J.pw, lane = NSOF_TID & 31;
    const int wid = (NSOF_BX * NSOF_NT + NSOF_TID) >> 5, nwarps = (nsof_grid_x() * NSOF_NT) >> 5;
    for (int w0 = wid * 32; w0 < total; w0 += nwarps * 32) {
        uint32_t mine = 0u;
        const int nk = nsof_imin(32, total - w0);
        for (int k = 0; k < nk; k++) {
            const int idx = w0 + k, y = idx / J.pw, X = idx - y * J.pw, x = (X - 1) * 32 + lane;
            bool p = false;
            if (X >= 1 && x < J.w) {
                const float2 f = *((const float2 *)(J.flow + (size_t)y * J.fpitch) + x);
                p = __fmaf_rn(f.x, f.x, __fmul_rn(f.y, f.y)) > s_star;
            }
            const uint32_t word = __ballot_sync(0xffffffffu, p);
            if (lane == k) mine = word;
        }
        if (lane < nk) {
            const int idx = w0 + lane, X = idx % J.pw;
            J.A[idx] = mine;
            if (X == 0 || X == J.pw - 1) J.B[idx] = 0u;
        }
    }
#endif
}
NSOF_GLOBAL k_seg_morph(const SegJob *NSOF_RESTRICT jobs, SegElement E, int erode, int src_is_b)
{
    const SegJob J = jobs[NSOF_BY];
    seg_morph_words(J, src_is_b ? J.B : J.A, src_is_b ? J.A : J.B, E, erode, NSOF_BX * NSOF_NT + NSOF_TID, nsof_grid_x() * NSOF_NT);
}
NSOF_GLOBAL k_seg_unpack(const SegJob *NSOF_RESTRICT jobs, int src_is_b)
{
    const SegJob J = jobs[NSOF_BY];
    seg_unpack_bytes(J, src_is_b ? J.B : J.A, NSOF_BX * NSOF_NT + NSOF_TID, nsof_grid_x() * NSOF_NT);
}

// The pass as a separable-style tile kernel (the one the driver takes; the per-word kernel above stays as the generic
// path and as its cross-check).  A CTA owns 32 rows x 8 words of the output:
//   1  the source rows the tile touches (+ 1 word either side) -> shared memory, inverted and masked for erode
//   2  per source word ONCE: the horizontal run-OR of every group; the runs of a cv2 element are nested (each contains
//      the previous one), so group g + 1 extends group g by the few offsets it adds (10 shifts for the whole 10 x 10
//      ellipse instead of 27 per OUTPUT word)
//   3  per output word: OR over the element rows of the plane of the row's group -- loads and ORs only
// 145 instead of ~700 instructions per output word (ncu: the per-word kernel is issue-bound at 80 %).
#define SEG_TWW 8
#define SEG_TROWS 32
NSOF_GLOBAL k_seg_morph_tiled(const SegJob *NSOF_RESTRICT jobs, SegElement E, int erode, int src_is_b)
{
    const SegJob J = jobs[NSOF_BY];
    const int nwords = J.pw - 2, tiles_x = (nwords + SEG_TWW - 1) / SEG_TWW, tiles_y = (J.h + SEG_TROWS - 1) / SEG_TROWS;
    if (NSOF_BX >= tiles_x * tiles_y) return;
    const uint32_t *NSOF_RESTRICT src = src_is_b ? J.B : J.A;
    uint32_t *NSOF_RESTRICT dst = src_is_b ? J.A : J.B;
    const int X0 = 1 + (NSOF_BX % tiles_x) * SEG_TWW, y0 = (NSOF_BX / tiles_x) * SEG_TROWS;  // first word / row of the tile
    const int nrows = SEG_TROWS + E.kh - 1;          // source rows y0 - ay .. y0 + TROWS - 1 + (kh - 1 - ay)
    NSOF_DYN_SMEM(smem_raw);
    uint32_t *sIn = (uint32_t *)smem_raw;            // [nrows][TWW + 2]
    uint32_t *sH = sIn + nrows * (SEG_TWW + 2);      // [n_groups][nrows][TWW]
    for (int i = NSOF_TID; i < nrows * (SEG_TWW + 2); i += NSOF_NT) {
        const int r = i / (SEG_TWW + 2), c = i - r * (SEG_TWW + 2);
        const int yy = y0 - E.ay + r, X = X0 - 1 + c;
        uint32_t v = 0u;
        if (yy >= 0 && yy < J.h && X < J.pw) {
            v = src[(size_t)yy * J.pw + X];
            if (erode) v = ~v & seg_valid(X, J.w);
        }
        sIn[i] = v;
    }
    NSOF_SYNC();
    for (int i = NSOF_TID; i < nrows * SEG_TWW; i += NSOF_NT) {
        const int r = i / SEG_TWW, c = i - r * SEG_TWW;
        const uint32_t *p = sIn + r * (SEG_TWW + 2) + c;
        const uint64_t lo64 = ((uint64_t)p[1] << 32) | p[0], hi64 = ((uint64_t)p[2] << 32) | p[1];
        uint32_t acc = 0u;
        int plo = 1, phi = 0;  // offsets already in acc: [plo, phi] (empty)
        for (int g = 0; g < E.n_groups; g++) {
            const int d0 = E.lo[g], d1 = E.hi[g];
            if (!E.nested || plo > phi) { acc = 0u; plo = d1 + 1; phi = d1; }  // from scratch: add [d0, d1]
            for (int d = d0; d < plo; d++) acc |= d < 0 ? (uint32_t)(lo64 >> (32 + d)) : (uint32_t)(hi64 >> d);
            for (int d = phi + 1; d <= d1; d++) acc |= d < 0 ? (uint32_t)(lo64 >> (32 + d)) : (uint32_t)(hi64 >> d);
            plo = d0; phi = d1;
            sH[(g * nrows + r) * SEG_TWW + c] = acc;
        }
    }
    NSOF_SYNC();
    for (int i = NSOF_TID; i < SEG_TROWS * SEG_TWW; i += NSOF_NT) {
        const int ly = i / SEG_TWW, c = i - ly * SEG_TWW;
        const int y = y0 + ly, X = X0 + c;
        if (y >= J.h || X > nwords) continue;
        uint32_t acc = 0u;
        for (int k = 0; k < E.kh; k++) {
            const int g = E.grp_of_row[k];
            if (g >= 0) acc |= sH[(g * nrows + ly + k) * SEG_TWW + c];  // source row y + k - ay = tile row ly + k
        }
        dst[(size_t)y * J.pw + X] = (erode ? ~acc : acc) & seg_valid(X, J.w);
    }
}

// cv::getStructuringElement (imgproc/src/morph.dispatch.cpp): rows are single runs [j1, j2)
static void seg_element_rows(int shape, int kw, int kh, int *j1, int *j2)
{
    const int r = kh / 2, c = kw / 2;
    const double inv_r2 = r ? 1.0 / ((double)r * r) : 0.0;
    for (int i = 0; i < kh; i++) {
        j1[i] = j2[i] = 0;
        if (shape == 0 || (shape == 1 && i == r)) { j1[i] = 0; j2[i] = kw; }
        else if (shape == 1) { j1[i] = c; j2[i] = c + 1; }
        else {
            const int dy = i - r;
            if (std::abs(dy) <= r) {
                const int dx = (int)lrint(c * std::sqrt((r * r - dy * dy) * inv_r2));  // cvRound
                j1[i] = std::max(c - dx, 0);
                j2[i] = std::min(c + dx + 1, kw);
            }
        }
    }
}

extern "C" int nsof_structuring_element(int shape, int ksize_w, int ksize_h, uint8_t *out)
{
    if (shape < 0 || shape > 2 || ksize_w < 1 || ksize_h < 1 || ksize_w > 32 || ksize_h > 32 || !out)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "structuring_element: shape 0..2, sizes 1..32, non-NULL output");
    int j1[32], j2[32];
    seg_element_rows(shape, ksize_w, ksize_h, j1, j2);
    for (int i = 0; i < ksize_h; i++)
        for (int j = 0; j < ksize_w; j++) out[i * ksize_w + j] = (j >= j1[i] && j < j2[i]) ? 1 : 0;
    return NSOF_OK;
}

static int seg_check_params(const nsof_seg_params *p)
{
    if (!p) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment: NULL params");
    if (p->shape < 0 || p->shape > 2) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment: shape must be 0 (rect), 1 (cross) or 2 (ellipse)");
    if (p->ksize_w < 1 || p->ksize_h < 1 || p->ksize_w > 32 || p->ksize_h > 32)
        NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "flow_segment: structuring element %d x %d outside 1..32", p->ksize_w, p->ksize_h);
    if (p->iterations < 0 || p->iterations > 64) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment: iterations outside 0..64");
    if (p->mag_thresh != p->mag_thresh) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment: mag_thresh is NaN");
    return NSOF_OK;
}

// largest float s with (double)sqrtf(s) <= th; th < 0 -> every (non-NaN) magnitude passes; th = inf -> none
static float seg_s_star(double th)
{
    if (th < 0) return -1.0f;
    if (std::isinf(th)) return INFINITY;
    const double t2 = th * th;
    float s = t2 >= 3.4028234663852886e38 ? 3.4028234663852886e38f : (float)t2;
    while (s > 0.f && (double)sqrtf(s) > th) s = nextafterf(s, -INFINITY);
    while (s < 3.4028234663852886e38f && (double)sqrtf(nextafterf(s, INFINITY)) <= th) s = nextafterf(s, INFINITY);
    return s;
}

extern "C" int nsof_flow_segment_batch(nsof_handle *h, const nsof_seg_desc *items, int n, const nsof_seg_params *params,
                                       void *cuda_stream)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment: NULL handle");
    int rc = seg_check_params(params);
    if (rc) return rc;
    if (n < 0 || (n > 0 && !items)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment: bad item list");
    if (n == 0) return NSOF_OK;
    if (n > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "flow_segment: at most 65535 regions per call");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));

    const size_t table_bytes = (((size_t)n * sizeof(SegJob) + 255) / 256) * 256;
    size_t words = 0;
    int max_words = 0, max_tiles = 0;
    double px = 0.;
    for (int i = 0; i < n; i++) {
        const nsof_seg_desc &d = items[i];
        if (!d.flow || !d.mask || d.h < 1 || d.w < 1 || d.flow_pitch < (ptrdiff_t)d.w * 8 || d.flow_pitch % 8 ||
            (uintptr_t)d.flow % 8 || d.mask_pitch < d.w)
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment: region %d is malformed (NULL, empty, pitch too small or flow not 8-byte aligned)", i);
        const int pw = (d.w + 31) / 32 + 2;
        if ((double)d.h * pw > 1e9) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "flow_segment: region %d too large", i);
        words += (size_t)d.h * pw;
        max_words = std::max(max_words, d.h * pw);
        max_tiles = std::max(max_tiles, ((pw - 2 + SEG_TWW - 1) / SEG_TWW) * ((d.h + SEG_TROWS - 1) / SEG_TROWS));
        px += (double)d.h * d.w;
    }
    NsofStageSlot *slot;
    rc = nsof_stage_acquire(h, table_bytes, &slot);
    if (rc) return rc;
    rc = h->seg.ensure(table_bytes + 2 * words * 4 + 16);
    if (rc) return rc;
    SegJob *hj = (SegJob *)slot->host;
    uint32_t *planes = (uint32_t *)((unsigned char *)h->seg.p + table_bytes);
    size_t off = 0;
    for (int i = 0; i < n; i++) {
        const nsof_seg_desc &d = items[i];
        SegJob &J = hj[i];
        J.flow = d.flow;
        J.mask = d.mask;
        J.fpitch = (int)(d.flow_pitch / 4);
        J.mpitch = (int)d.mask_pitch;
        J.h = d.h;
        J.w = d.w;
        J.pw = (d.w + 31) / 32 + 2;
        J.A = planes + off;
        J.B = planes + off + (size_t)d.h * J.pw;
        off += 2 * (size_t)d.h * J.pw;
    }
    rc = nsof_table_upload(h, h->seg.p, slot, (size_t)n * sizeof(SegJob), stream);
    if (rc) return rc;
    const SegJob *dj = (const SegJob *)h->seg.p;

    // element rows -> groups of rows that share one run
    SegElement E;
    memset(&E, 0, sizeof(E));
    int j1[32], j2[32];
    seg_element_rows(params->shape, params->ksize_w, params->ksize_h, j1, j2);
    const int ax = params->ksize_w / 2;
    E.ay = params->ksize_h / 2;
    for (int i = 0; i < params->ksize_h; i++) {
        if (j2[i] <= j1[i]) continue;
        const int lo = j1[i] - ax, hi = j2[i] - 1 - ax;
        int g = 0;
        while (g < E.n_groups && !(E.lo[g] == lo && E.hi[g] == hi)) g++;
        if (g == E.n_groups) { E.lo[g] = lo; E.hi[g] = hi; E.n_groups++; }
        E.rows[g] |= 1u << i;
    }
    // groups by increasing run length; nested = every run contains the previous one (true for every cv2 element)
    for (int a = 0; a < E.n_groups; a++)
        for (int b = a + 1; b < E.n_groups; b++)
            if (E.hi[b] - E.lo[b] < E.hi[a] - E.lo[a]) {
                std::swap(E.lo[a], E.lo[b]); std::swap(E.hi[a], E.hi[b]); std::swap(E.rows[a], E.rows[b]);
            }
    E.nested = 1;
    for (int g = 1; g < E.n_groups; g++)
        if (E.lo[g] > E.lo[g - 1] || E.hi[g] < E.hi[g - 1]) E.nested = 0;
    E.kh = params->ksize_h;
    for (int i = 0; i < 32; i++) {
        E.grp_of_row[i] = -1;
        for (int g = 0; g < E.n_groups; g++)
            if (i < E.kh && (E.rows[g] >> i) & 1u) E.grp_of_row[i] = (signed char)g;
    }

    const float s_star = seg_s_star(params->mag_thresh);
    const int gx_words = std::max(1, std::min((max_words + 255) / 256, 1024));
    NSOF_KLAUNCH(h, NSOF_K_SEG_PACK, 8. * px + words * 4., k_seg_pack, gx_words, n, 1, 256, 0, stream, dj, s_star);
    int src_is_b = 0;
    const int nrows = SEG_TROWS + params->ksize_h - 1;
    const size_t tiled_smem = ((size_t)nrows * (SEG_TWW + 2) + (size_t)E.n_groups * nrows * SEG_TWW) * sizeof(uint32_t);
    const bool tiled = !h->dbg_no_fuse && tiled_smem <= 48 * 1024;  // nsof_debug_fuse(h, 0): the per-word kernel (tests)
    for (int it = 0; it < 2 * params->iterations; it++) {
        if (tiled)
            NSOF_KLAUNCH(h, NSOF_K_SEG_MORPH, words * 8., k_seg_morph_tiled, max_tiles, n, 1, 256, tiled_smem, stream, dj, E, it & 1,
                         src_is_b);
        else
            NSOF_KLAUNCH(h, NSOF_K_SEG_MORPH, words * 8., k_seg_morph, gx_words, n, 1, 256, 0, stream, dj, E, it & 1, src_is_b);
        src_is_b ^= 1;
    }
    NSOF_KLAUNCH(h, NSOF_K_SEG_UNPACK, px + words * 4., k_seg_unpack, gx_words, n, 1, 256, 0, stream, dj, src_is_b);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

// zero the mask rectangles of `n` regions (the paste target `motion_binary = np.zeros(...)` of the next frame pair)
NSOF_GLOBAL k_seg_mask_clear(const SegJob *NSOF_RESTRICT jobs)
{
    const SegJob J = jobs[NSOF_BY];
    const int qw = (J.w + 3) / 4, total = J.h * qw;
    for (int idx = NSOF_BX * NSOF_NT + NSOF_TID; idx < total; idx += nsof_grid_x() * NSOF_NT) {
        const int y = idx / qw, x = (idx - y * qw) * 4;
        uint8_t *o = J.mask + (size_t)y * J.mpitch + x;
        if (x + 4 <= J.w && (((uintptr_t)o) & 3) == 0) *(uint32_t *)o = 0u;
        else for (int q = 0; q < 4 && x + q < J.w; q++) o[q] = 0;
    }
}

extern "C" int nsof_mask_clear(nsof_handle *h, const nsof_seg_desc *items, int n, void *cuda_stream)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "mask_clear: NULL handle");
    if (n < 0 || (n > 0 && !items)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "mask_clear: bad item list");
    if (n == 0) return NSOF_OK;
    if (n > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "mask_clear: at most 65535 regions per call");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t bytes = (size_t)n * sizeof(SegJob);
    NsofStageSlot *slot;
    int rc = nsof_stage_acquire(h, bytes, &slot);
    if (rc) return rc;
    rc = h->clear_tab.ensure(bytes + 16);
    if (rc) return rc;
    SegJob *hj = (SegJob *)slot->host;
    int max_q = 0;
    double px = 0.;
    for (int i = 0; i < n; i++) {
        const nsof_seg_desc &d = items[i];
        if (!d.mask || d.h < 1 || d.w < 1 || d.mask_pitch < d.w) NSOF_FAIL(NSOF_ERR_BAD_ARG, "mask_clear: region %d is malformed", i);
        memset(&hj[i], 0, sizeof(SegJob));
        hj[i].mask = d.mask; hj[i].mpitch = (int)d.mask_pitch; hj[i].h = d.h; hj[i].w = d.w;
        max_q = std::max(max_q, d.h * ((d.w + 3) / 4));
        px += (double)d.h * d.w;
    }
    rc = nsof_table_upload(h, h->clear_tab.p, slot, bytes, stream);
    if (rc) return rc;
    NSOF_KLAUNCH(h, NSOF_K_SEG_UNPACK, px, k_seg_mask_clear, std::max(1, std::min((max_q + 255) / 256, 1024)), n, 1, 256, 0, stream,
                 (const SegJob *)h->clear_tab.p);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_flow_segment_host(nsof_handle *h, const float *flow, ptrdiff_t flow_pitch, int height, int width,
                                      const nsof_seg_params *params, uint8_t *mask, ptrdiff_t mask_pitch)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment_host: NULL handle");
    if (!flow || !mask || height < 1 || width < 1 || flow_pitch < (ptrdiff_t)width * 8 || mask_pitch < width)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_segment_host: NULL / empty region or pitch smaller than a row");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t frow = (size_t)width * 8, fbytes = ((frow * height + 255) / 256) * 256;
    int rc = h->io.ensure(fbytes + (size_t)width * height + 16);
    if (rc) return rc;
    float *dflow = (float *)h->io.p;
    uint8_t *dmask = (uint8_t *)h->io.p + fbytes;
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(dflow, frow, flow, (size_t)flow_pitch, frow, height, cudaMemcpyHostToDevice, nullptr));
    nsof_seg_desc d = {dflow, (ptrdiff_t)frow, height, width, dmask, (ptrdiff_t)width};
    rc = nsof_flow_segment_batch(h, &d, 1, params, nullptr);
    if (rc) return rc;
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(mask, (size_t)mask_pitch, dmask, (size_t)width, (size_t)width, height, cudaMemcpyDeviceToHost, nullptr));
    NSOF_CUDA_TRY(cudaStreamSynchronize(nullptr));
    return NSOF_OK;
}
