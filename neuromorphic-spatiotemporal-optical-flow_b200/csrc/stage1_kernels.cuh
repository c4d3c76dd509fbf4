// stage1_kernels.cuh -- stage 1 (synaptic temporal memory -> ROI) device code for sm_100a.
//
// Event mode  : eventsim/event_mem_sim.py:40-63 (update_state, resistance_exp), :100-104
//               (bincount_2d), :204-228 (scheme-1 loop).
// Frame mode  : simulation/simulationcode_v4_transistor_vehicle.m:111-121 (compress_image),
//               :146-184 (difference matrix, func1-3, update_state), :187-236, :306-321.
// ROI tail    : optical_flow_seg.py:426-435 (current -> uint8 map), :115-121
//               (update_transition_pic), :223-224 (connectedComponentsWithStats, 4-connectivity),
//               :129-138 / :174-183 (rectangles).
#pragma once
#include "nsof_rt.h"
#include "../../include/nsof.h"

// float32 copies of the model constants, demoted the way numpy demotes Python floats that meet
// float32 arrays (event_mem_sim.py:42-56)
struct SynF32 {
    float voff, von, koff, kon, son, soff, bon, boff, alphaoff, alphaon, dt;
};

// update_state for one element, float32 (event_mem_sim.py:40-57)
NSOF_DEV float s1_update_state_f32(float w, float V, const SynF32 &p)
{
    float dwdt = 0.f;
    if (V < p.voff) {
        float a = V / p.voff - 1.f;
        if (p.alphaoff != 1.f) a = powf(a, p.alphaoff);
        dwdt = p.koff * a * powf(1.f - w * p.soff, p.boff);
    } else if (V > p.von) {
        float a = V / p.von - 1.f;
        if (p.alphaon != 1.f) a = powf(a, p.alphaon);
        dwdt = p.kon * a * powf(1.f - w * p.son, p.bon);
    }
    float wn = w + dwdt * p.dt;
    // np.clip(x, 0, 1): NaN propagates
    wn = wn < 0.f ? 0.f : wn;
    wn = wn > 1.f ? 1.f : wn;
    return wn;
}

NSOF_GLOBAL k_s1_synapse_step(float *NSOF_RESTRICT w, const float *NSOF_RESTRICT V, size_t n, SynF32 p, int per_block)
{
    const size_t base = (size_t)NSOF_BX * per_block;
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const size_t e = base + i;
        if (e < n) w[e] = s1_update_state_f32(w[e], V[e], p);
    }
}

// bincount_2d: counts must be zeroed by the caller (memset on the same stream)
NSOF_GLOBAL k_s1_event_hist(const int16_t *NSOF_RESTRICT x, const int16_t *NSOF_RESTRICT y, size_t n, int H, int W,
                            int32_t *counts, int per_block)
{
    const size_t base = (size_t)NSOF_BX * per_block;
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const size_t e = base + i;
        if (e >= n) continue;
        const int xx = x[e], yy = y[e];
        if ((unsigned)xx < (unsigned)W && (unsigned)yy < (unsigned)H) {
#ifdef NSOF_EMU
            counts[(size_t)yy * W + xx] += 1;
#else
            atomicAdd(counts + (size_t)yy * W + xx, 1);
#endif
        }
    }
}

// Scheme 1, chunk of <= 32 slices.  Pass 1: per-pixel hit mask (bit s = pixel reached theta events
// in slice s).  theta == 1: one atomicOr per event.  theta > 1 uses the counting variant below.
NSOF_GLOBAL k_s1_slice_bits(const int16_t *NSOF_RESTRICT x, const int16_t *NSOF_RESTRICT y,
                            const int64_t *NSOF_RESTRICT slice_begin, int n_slices, int H, int W, uint32_t *bits,
                            int per_block)
{
    const int64_t e0 = slice_begin[0], e1 = slice_begin[n_slices];
    const int64_t base = e0 + (int64_t)NSOF_BX * per_block;
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const int64_t e = base + i;
        if (e >= e1) continue;
        // slice of event e: last s with slice_begin[s] <= e
        int lo = 0, hi = n_slices - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (slice_begin[mid] <= e) lo = mid; else hi = mid - 1;
        }
        const int xx = x[e], yy = y[e];
        if ((unsigned)xx < (unsigned)W && (unsigned)yy < (unsigned)H) {
#ifdef NSOF_EMU
            bits[(size_t)yy * W + xx] |= 1u << lo;
#else
            atomicOr(bits + (size_t)yy * W + xx, 1u << lo);
#endif
        }
    }
}

// Pass 2: every pixel replays its <= 32 slices with w in a register.
NSOF_GLOBAL k_s1_replay_bits(float *NSOF_RESTRICT w, uint32_t *NSOF_RESTRICT bits, size_t n, int n_slices,
                             float active_v, float silent_v, SynF32 p, int per_block)
{
    const size_t base = (size_t)NSOF_BX * per_block;
    const bool silent_is_idle = !(silent_v < p.voff) && !(silent_v > p.von);
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const size_t e = base + i;
        if (e >= n) continue;
        uint32_t b = bits[e];
        bits[e] = 0;  // leave the scratch clean for the next chunk
        if (silent_is_idle && b == 0) continue;
        float wv = w[e];
        for (int s = 0; s < n_slices; s++) {
            const bool hit = (b >> s) & 1u;
            if (!hit && silent_is_idle) continue;
            wv = s1_update_state_f32(wv, hit ? active_v : silent_v, p);
        }
        w[e] = wv;
    }
}

// General theta: one slice at a time from an int32 histogram.
NSOF_GLOBAL k_s1_step_from_counts(float *NSOF_RESTRICT w, const int32_t *NSOF_RESTRICT counts, size_t n, int theta,
                                  float active_v, float silent_v, SynF32 p, int per_block)
{
    const size_t base = (size_t)NSOF_BX * per_block;
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const size_t e = base + i;
        if (e < n) w[e] = s1_update_state_f32(w[e], counts[e] >= theta ? active_v : silent_v, p);
    }
}

// resistance_exp: float32 w, float64 result (np.log returns a float64 scalar, which promotes)
NSOF_GLOBAL k_s1_resistance_exp(const float *NSOF_RESTRICT w, size_t n, double Ron, double lam, double *NSOF_RESTRICT R,
                                int per_block)
{
    const size_t base = (size_t)NSOF_BX * per_block;
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const size_t e = base + i;
        if (e < n) R[e] = Ron / exp(-lam * (double)(1.0f - w[e]));
    }
}

// ---------------------------------------------------------------------------------------------
// Frame mode.  Lanczos-3 antialiased imresize, dimension 1 (rows) first as MATLAB does for equal
// scales.  Weight/index tables come from the host (double); P taps per output.
// pass 1: tmp[n][Hc][W] = sum_k wy[cy][k] * (src[iy[cy][k]][x] / 255)
// ---------------------------------------------------------------------------------------------
NSOF_GLOBAL k_s1_lanczos_rows(const uint8_t *NSOF_RESTRICT frames, ptrdiff_t pitch, ptrdiff_t frame_stride, int W,
                              int Hc, const double *NSOF_RESTRICT wy, const int *NSOF_RESTRICT iy, int P,
                              double *NSOF_RESTRICT tmp, int cols_per_block)
{
    const int cy = NSOF_BY, f = NSOF_BZ;
    const uint8_t *src = frames + (size_t)f * frame_stride;
    const int x0 = NSOF_BX * cols_per_block;
    for (int i = NSOF_TID; i < cols_per_block; i += NSOF_NT) {
        const int x = x0 + i;
        if (x >= W) continue;
        double acc = 0.;
        for (int k = 0; k < P; k++) {
            const double wgt = wy[cy * P + k];
            if (wgt != 0.) acc += wgt * ((double)NSOF_LDG(src + (size_t)iy[cy * P + k] * pitch + x) / 255.0);
        }
        tmp[((size_t)f * Hc + cy) * W + x] = acc;
    }
}

// pass 2: cells[n][Hc][Wc] = sum_k wx[cx][k] * tmp[cy][ix[cx][k]]
NSOF_GLOBAL k_s1_lanczos_cols(const double *NSOF_RESTRICT tmp, int W, int Hc, int Wc, const double *NSOF_RESTRICT wx,
                              const int *NSOF_RESTRICT ix, int P, double *NSOF_RESTRICT cells)
{
    const int f = NSOF_BX;
    for (int i = NSOF_TID; i < Hc * Wc; i += NSOF_NT) {
        const int cx = i % Wc, cy = i / Wc;
        const double *row = tmp + ((size_t)f * Hc + cy) * W;
        double acc = 0.;
        for (int k = 0; k < P; k++) {
            const double wgt = wx[cx * P + k];
            if (wgt != 0.) acc += wgt * row[ix[cx * P + k]];
        }
        cells[(size_t)f * Hc * Wc + i] = acc;
    }
}

struct FrameSimConst {
    double th1, th2, dt_sub, vds;
    int nsub;
    double alphaoff, alphaon, voff, von, koff, kon, son, soff, bon, boff, Ron, lam;
};

// calculate_difference_matrix + modulatefunc + nSubSteps x update_state + resistance -> current
NSOF_GLOBAL k_s1_frame_step(const double *NSOF_RESTRICT a, const double *NSOF_RESTRICT b, double *NSOF_RESTRICT w,
                            double *NSOF_RESTRICT current, size_t n, FrameSimConst c, int per_block)
{
    const size_t base = (size_t)NSOF_BX * per_block;
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const size_t e = base + i;
        if (e >= n) continue;
        const double d = fabs(a[e] * 256. - b[e] * 256.);
        // assignment order of the .m file: (d > th2) func3, then (th1 < d <= th2) func2, then (d <= th1) func1
        double V = 0.;
        if (d > c.th2) V = (d + 4.) * 0.75;
        if (d > c.th1 && d <= c.th2) V = (d + 4.) * 0.75;
        if (d <= c.th1) V = (d - 5.5) * 0.6;
        // modulatefunc: a=0.3,b=0 for V>0 ; c=3,d=-3 for V<0 ; (V == 0 leaves v_mod undefined in MATLAB: 0 here)
        double vm = 0.;
        if (V > 0.) vm = 0.3 * V + 0.;
        else if (V < 0.) vm = 3. * V + -3.;
        vm = -vm;
        double wv = w[e];
        if (vm < c.voff) {
            double g = c.koff * pow(vm / c.voff - 1., c.alphaoff);
            for (int s = 0; s < c.nsub; s++) {
                const double dwdt = g * pow(1. - wv * c.soff, c.boff);
                wv = wv + dwdt * c.dt_sub;
                wv = fmax(0., fmin(wv, 1.));
            }
        } else if (vm > c.von) {
            double g = c.kon * pow(vm / c.von - 1., c.alphaon);
            for (int s = 0; s < c.nsub; s++) {
                const double dwdt = g * pow(1. - wv * c.son, c.bon);
                wv = wv + dwdt * c.dt_sub;
                wv = fmax(0., fmin(wv, 1.));
            }
        }
        w[e] = wv;
        if (current) current[e] = c.vds / (c.Ron / exp(-c.lam * (1. - wv)));
    }
}

// ---------------------------------------------------------------------------------------------
// ROI tail: one block per array state.  dynamic smem: Hc*Wc int32 labels + (1 + 5*NSOF_MAX_RECTS) ints.
// ---------------------------------------------------------------------------------------------
NSOF_GLOBAL k_s1_roi(const double *NSOF_RESTRICT currents, nsof_roi_params p, const double *NSOF_RESTRICT thr,
                     uint8_t *memimg, int32_t *labels_out, int32_t *n_components, int32_t *n_rects, nsof_rect *rects,
                     int32_t *stats)
{
    const int s = NSOF_BX, Hc = p.cells_h, Wc = p.cells_w, N = Hc * Wc;
    const double *I = currents + (size_t)s * N;
    NSOF_DYN_SMEM(smem_raw);
    int *lab = (int *)smem_raw;
    int *flag = lab + N;
    int *st = flag + 1;  // [NSOF_MAX_RECTS][5] left, top, right(exclusive), bottom(exclusive), area

    // (1) current -> uint8 map by monotone threshold table; mask = (u8 >= THRES)
    for (int i = NSOF_TID; i < N; i += NSOF_NT) {
        const double v = I[i];
        int u = 0;
        if (v > 0. && v < 1.) {
            int lo = 0, hi = 255;  // largest u in [0,255] with thr[u] <= v (thr[0] = -inf conceptually)
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (thr[mid] <= v) lo = mid; else hi = mid - 1;
            }
            u = lo;
        }
        if (memimg) memimg[(size_t)s * N + i] = (uint8_t)u;
        lab[i] = (u >= p.thres) ? i : -1;
    }
    NSOF_SYNC();
    // (2) 4-connected min-label propagation to a fixed point
    for (int iter = 0; iter < N + 1; iter++) {
        if (NSOF_TID == 0) *flag = 0;
        NSOF_SYNC();
        for (int i = NSOF_TID; i < N; i += NSOF_NT) {
            int l = lab[i];
            if (l < 0) continue;
            const int x = i % Wc, y = i / Wc;
            int best = l;
            if (x > 0 && lab[i - 1] >= 0) best = nsof_imin(best, lab[i - 1]);
            if (x + 1 < Wc && lab[i + 1] >= 0) best = nsof_imin(best, lab[i + 1]);
            if (y > 0 && lab[i - Wc] >= 0) best = nsof_imin(best, lab[i - Wc]);
            if (y + 1 < Hc && lab[i + Wc] >= 0) best = nsof_imin(best, lab[i + Wc]);
            // pointer jump
            if (best >= 0 && lab[best] >= 0) best = nsof_imin(best, lab[best]);
            if (best < l) { lab[i] = best; *flag = 1; }
        }
        NSOF_SYNC();
        const int changed = *flag;
        NSOF_SYNC();
        if (!changed) break;
    }
    // (3) cv2 numbering (raster order of each component's first cell) + stats + rectangles: tiny, serial
    if (NSOF_TID == 0) {
        int ncomp = 0;
        int uxmin = Wc, uymin = Hc, uxmax = 0, uymax = 0;
        for (int i = 0; i < N; i++) {
            const int l = lab[i];
            if (l < 0) { if (labels_out) labels_out[(size_t)s * N + i] = 0; continue; }
            const int x = i % Wc, y = i / Wc;
            int id;
            if (l == i) {
                id = ++ncomp;  // root: first cell of its component in raster order
                lab[i] = -(id + 1);  // encode the final id as -(id+1) so later cells can look it up
                if (id <= NSOF_MAX_RECTS) { int *q = st + (id - 1) * 5; q[0] = x; q[1] = y; q[2] = x + 1; q[3] = y + 1; q[4] = 0; }
            } else {
                id = -lab[l] - 1;  // root precedes i in raster order, already renumbered
            }
            if (id <= NSOF_MAX_RECTS) {
                int *q = st + (id - 1) * 5;
                q[0] = nsof_imin(q[0], x); q[1] = nsof_imin(q[1], y);
                q[2] = nsof_imax(q[2], x + 1); q[3] = nsof_imax(q[3], y + 1); q[4] += 1;
            }
            uxmin = nsof_imin(uxmin, x); uymin = nsof_imin(uymin, y);
            uxmax = nsof_imax(uxmax, x + 1); uymax = nsof_imax(uymax, y + 1);
            if (labels_out) labels_out[(size_t)s * N + i] = id;
        }
        const int nst = nsof_imin(ncomp, NSOF_MAX_RECTS);
        if (stats)
            for (int c = 0; c < nst; c++) {
                const int *q = st + c * 5;
                int32_t *o = stats + ((size_t)s * NSOF_MAX_RECTS + c) * 5;
                o[0] = q[0]; o[1] = q[1]; o[2] = q[2] - q[0]; o[3] = q[3] - q[1]; o[4] = q[4];
            }
        int nr = 0;
        if (ncomp > 0) {
            if (p.flag == 2) {
                nsof_rect r;
                r.x0 = nsof_imax(uxmin * p.pixel_w - p.ext_left, 0);
                r.y0 = nsof_imax(uymin * p.pixel_h - p.ext_upper, 0);
                r.x1 = nsof_imin(uxmax * p.pixel_w + p.ext_right, p.frame_w);
                r.y1 = nsof_imin(uymax * p.pixel_h + p.ext_lower, p.frame_h);
                if (rects) rects[(size_t)s * NSOF_MAX_RECTS] = r;
                nr = 1;
            } else {
                for (int c = 0; c < nst; c++) {
                    const int *q = st + c * 5;
                    nsof_rect r;
                    r.x0 = nsof_imax(q[0] * p.pixel_w - p.ext_left, 0);
                    r.y0 = nsof_imax(q[1] * p.pixel_h - p.ext_upper, 0);
                    r.x1 = nsof_imin(q[2] * p.pixel_w + p.ext_right, p.frame_w);
                    r.y1 = nsof_imin(q[3] * p.pixel_h + p.ext_lower, p.frame_h);
                    if (rects) rects[(size_t)s * NSOF_MAX_RECTS + c] = r;
                }
                nr = ncomp;  // > NSOF_MAX_RECTS tells the host wrapper that the list was truncated
            }
        }
        if (n_rects) n_rects[s] = nr;
        if (n_components) n_components[s] = ncomp;
    }
}
