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

// ---------------------------------------------------------------------------------------------
// Scheme 2 (eventsim/event_mem_sim.py:231-286): DC bias everywhere, bias + active_v on the pixels that saw a
// matching event in the slice AND whose refractory stamp next_ok <= t_us[sl.start]; those pixels get
// next_ok = t_us[sl.stop - 1] + REFRACTORY_US.  'split': ON events (p == 1) drive array A, OFF events (p == 0)
// array B (any other polarity value is ignored, as in the reference -- its synthetic generator writes -1 for
// OFF); 'magnitude': every event drives array A.  Same two passes as scheme 1, chunks of <= 32 slices.
// ---------------------------------------------------------------------------------------------
NSOF_GLOBAL k_s2_slice_bits(const int16_t *NSOF_RESTRICT x, const int16_t *NSOF_RESTRICT y, const int8_t *NSOF_RESTRICT pol,
                            const int64_t *NSOF_RESTRICT slice_begin, int n_slices, int H, int W, int magnitude,
                            uint32_t *bits_a, uint32_t *bits_b, int per_block)
{
    const int64_t e0 = slice_begin[0], e1 = slice_begin[n_slices];
    const int64_t base = e0 + (int64_t)NSOF_BX * per_block;
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const int64_t e = base + i;
        if (e >= e1) continue;
        int lo = 0, hi = n_slices - 1;  // slice of event e: last s with slice_begin[s] <= e
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (slice_begin[mid] <= e) lo = mid; else hi = mid - 1;
        }
        const int xx = x[e], yy = y[e], pp = pol[e];
        uint32_t *bits = magnitude ? bits_a : (pp == 1 ? bits_a : (pp == 0 ? bits_b : nullptr));
        if (bits && (unsigned)xx < (unsigned)W && (unsigned)yy < (unsigned)H) {
#ifdef NSOF_EMU
            bits[(size_t)yy * W + xx] |= 1u << lo;
#else
            atomicOr(bits + (size_t)yy * W + xx, 1u << lo);
#endif
        }
    }
}

// first / last event time of every slice of the chunk (t_first is what every pixel's refractory test reads)
NSOF_GLOBAL k_s2_slice_times(const int64_t *NSOF_RESTRICT t, const int64_t *NSOF_RESTRICT slice_begin, int n_slices,
                             int64_t *NSOF_RESTRICT t_first, int64_t *NSOF_RESTRICT t_last)
{
    for (int s = NSOF_TID; s < n_slices; s += NSOF_NT) {
        const int64_t a = slice_begin[s], b = slice_begin[s + 1];
        t_first[s] = b > a ? t[a] : 0;
        t_last[s] = b > a ? t[b - 1] : 0;
    }
}

NSOF_GLOBAL k_s2_replay_bits(float *NSOF_RESTRICT w, uint32_t *NSOF_RESTRICT bits, int64_t *NSOF_RESTRICT next_ok, size_t n,
                             int n_slices, const int64_t *NSOF_RESTRICT t_first, const int64_t *NSOF_RESTRICT t_last,
                             int64_t refractory, float active_v, float bias_v, SynF32 p, int per_block)
{
    const size_t base = (size_t)NSOF_BX * per_block;
    const bool bias_is_idle = !(bias_v < p.voff) && !(bias_v > p.von);
    const float v_on = bias_v + active_v;  // Va[ys_ok, xs_ok] += active_v on a float32 map
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const size_t e = base + i;
        if (e >= n) continue;
        const uint32_t b = bits[e];
        bits[e] = 0;  // leave the scratch clean for the next chunk
        if (bias_is_idle && b == 0) continue;
        float wv = w[e];
        int64_t nk = next_ok[e];
        for (int s = 0; s < n_slices; s++) {
            bool hit = (b >> s) & 1u;
            if (hit) {
                hit = nk <= t_first[s];
                if (hit) nk = t_last[s] + refractory;
            }
            if (!hit && bias_is_idle) continue;
            wv = s1_update_state_f32(wv, hit ? v_on : bias_v, p);
        }
        w[e] = wv;
        next_ok[e] = nk;
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

// Event-mode ROI glue (builder-defined, SURVEY.md 8(d) config 3: the reference never thresholds the eventsim
// state): one current per MEMSIZE x MEMSIZE cell = the largest pixel current 1 / resistance_exp(w) inside it.
// The map w -> current is increasing, so the cell current is the current of the cell's largest w: a float max
// (exact) followed by the same double arithmetic numpy applies to that element.  One block per cell.
NSOF_GLOBAL k_s1_cell_current_from_w(const float *NSOF_RESTRICT w, int W, int memsize_h, int memsize_w, int Wc, double Ron,
                                     double lam, double *NSOF_RESTRICT currents)
{
    const int cell = NSOF_BX, cx = cell % Wc, cy = cell / Wc;
    NSOF_DYN_SMEM(smem_raw);
    float *red = (float *)smem_raw;
    float m = -1.f;
    for (int i = NSOF_TID; i < memsize_h * memsize_w; i += NSOF_NT) {
        const int ly = i / memsize_w, lx = i - ly * memsize_w;
        m = fmaxf(m, w[(size_t)(cy * memsize_h + ly) * W + cx * memsize_w + lx]);
    }
    red[NSOF_TID] = m;
    NSOF_SYNC();
    if (NSOF_TID == 0) {
        for (int i = 1; i < NSOF_NT; i++) m = fmaxf(m, red[i]);
        currents[cell] = 1.0 / (Ron / exp(-lam * (double)(1.0f - m)));
    }
}

// ---------------------------------------------------------------------------------------------
// Frame mode.  compress_image = im2double + imresize(.,'lanczos3') with antialiasing is a separable
// linear map: cells = Wy * (img / 255) * Wx^T with DENSE per-axis matrices Wy [Hc][H], Wx [Wc][W]
// built on the host in double from MATLAB's contribution tables (mirrored taps that land on the same
// source index are merged).  Only a band of each row of Wy / Wx is non-zero (about 6/scale taps).
//
// pass 1 (k_s1_lanczos_rows): every source byte is read ONCE (uchar4 per thread, 128 B per warp).
//   grid (ceil(W/4/NT), S row segments, n frames).  A segment of seg_rows source rows touches at most
//   LZ_MAXC consecutive cell rows starting at cbase[seg]; the thread keeps LZ_MAXC x 4 double
//   accumulators of sum_y Wy * u8 in registers and writes them to part[f][seg][j][x] (im2double's 1/255 is
//   applied in pass 2).  Weights of the segment are staged
//   in shared memory: sw[j][r] = Wy[cbase + j][y0 + r].   dynamic smem: LZ_MAXC * seg_rows doubles.
// pass 2 (k_s1_lanczos_cols): block per (cell row, frame): tmp[x] = sum over segments in fixed order
//   (deterministic), then one warp per cell column reduces its tap band with a fixed-order shuffle tree.
// ---------------------------------------------------------------------------------------------
#define LZ_MAXC 8
#define LZ_TILE_W 512   // source columns per block: 128 threads x uchar4
#define LZ_STAGES 3     // the segment's rows arrive in 3 groups, each on its own mbarrier

// dynamic smem: 64 B (mbarriers) + LZ_MAXC * seg_rows doubles + seg_rows * LZ_TILE_W bytes.
// The block's seg_rows x LZ_TILE_W source window is fetched by bulk asynchronous row copies (cp.async.bulk, SASS
// UBLKCP): ~48 KB in flight per block with no register staging (per-thread loads left the kernel latency-bound at
// 0.4 TB/s); the FP64 accumulation starts as soon as the first third of the rows has landed.
NSOF_GLOBAL k_s1_lanczos_rows(const uint8_t *NSOF_RESTRICT frames, ptrdiff_t pitch, ptrdiff_t frame_stride, int H, int W,
                              int Hc, const double *NSOF_RESTRICT Wy, const int *NSOF_RESTRICT cbase, int seg_rows,
                              int n_seg, double *NSOF_RESTRICT part)
{
    const int seg = NSOF_BY, f = NSOF_BZ;
    const int y0 = seg * seg_rows, y1 = nsof_imin(y0 + seg_rows, H), rows = y1 - y0;
    const int cb = cbase[seg];
    const int x0 = NSOF_BX * LZ_TILE_W, tw = nsof_imin(LZ_TILE_W, W - x0);
    NSOF_DYN_SMEM(smem_raw);
    nsof_mbar_t *bar = (nsof_mbar_t *)smem_raw;
    double *sw = (double *)(smem_raw + 64);
    uint8_t *tile = smem_raw + 64 + (size_t)LZ_MAXC * seg_rows * sizeof(double);
    const uint8_t *src = frames + (size_t)f * frame_stride + (size_t)y0 * pitch + x0;
    const bool bulk = ((uintptr_t)src & 15) == 0 && (pitch & 15) == 0 && (tw & 15) == 0;
    const int per_stage = (rows + LZ_STAGES - 1) / LZ_STAGES;
    if (bulk && NSOF_TID == 0) {
        for (int sg = 0; sg < LZ_STAGES; sg++) {
            const int r0 = sg * per_stage, r1 = nsof_imin(rows, r0 + per_stage);
            nsof_mbar_init(bar + sg, 1);
            nsof_mbar_expect_tx(bar + sg, (unsigned)(nsof_imax(r1 - r0, 0) * tw));
        }
    }
    for (int i = NSOF_TID; i < LZ_MAXC * seg_rows; i += NSOF_NT) {
        const int j = i / seg_rows, r = i % seg_rows;
        const int cy = cb + j, y = y0 + r;
        sw[i] = (cy < Hc && y < y1) ? Wy[(size_t)cy * H + y] : 0.;
    }
    NSOF_SYNC();
    if (bulk)
        for (int r = NSOF_TID; r < rows; r += NSOF_NT)
            nsof_bulk_g2s(tile + (size_t)r * LZ_TILE_W, src + (size_t)r * pitch, (unsigned)tw, bar + r / per_stage);
    const int ngroups = (tw + 3) / 4;
    for (int g = NSOF_TID; g < ngroups; g += NSOF_NT) {
        const int x = x0 + g * 4;
        double acc[LZ_MAXC][4];
#pragma unroll
        for (int j = 0; j < LZ_MAXC; j++) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.;
        for (int sg = 0; sg < LZ_STAGES; sg++) {
            const int r0 = sg * per_stage, r1 = nsof_imin(rows, r0 + per_stage);
            if (bulk) nsof_mbar_wait(bar + sg, 0);
#pragma unroll 4
            for (int r = r0; r < r1; r++) {
                double v[4];
                if (bulk) {
                    const uint32_t u = *(const uint32_t *)(tile + (size_t)r * LZ_TILE_W + g * 4);
                    v[0] = (double)(u & 255u); v[1] = (double)((u >> 8) & 255u);
                    v[2] = (double)((u >> 16) & 255u); v[3] = (double)(u >> 24);
                } else {
                    const uint8_t *row = src + (size_t)r * pitch + g * 4;
                    for (int q = 0; q < 4; q++) v[q] = x + q < W ? (double)NSOF_LDG(row + q) : 0.;
                }
                const double *wr = sw + r;
#pragma unroll
                for (int j = 0; j < LZ_MAXC; j++) {
                    const double wj = wr[j * seg_rows];
                    acc[j][0] += wj * v[0]; acc[j][1] += wj * v[1]; acc[j][2] += wj * v[2]; acc[j][3] += wj * v[3];
                }
            }
        }
        for (int j = 0; j < LZ_MAXC; j++) {
            double *o = part + (((size_t)f * n_seg + seg) * LZ_MAXC + j) * W + x;
            for (int q = 0; q < 4; q++)
                if (x + q < W) o[q] = acc[j][q];
        }
    }
}

// dynamic smem: W doubles
NSOF_GLOBAL k_s1_lanczos_cols(const double *NSOF_RESTRICT part, int W, int Hc, int Wc, const double *NSOF_RESTRICT Wx,
                              const int *NSOF_RESTRICT xband, const int *NSOF_RESTRICT cbase, int n_seg,
                              double *NSOF_RESTRICT cells)
{
    const int cy = NSOF_BX, f = NSOF_BY;
    NSOF_DYN_SMEM(smem_raw);
    double *tmp = (double *)smem_raw;
    for (int x = NSOF_TID; x < W; x += NSOF_NT) {
        double s = 0.;
        for (int seg = 0; seg < n_seg; seg++) {
            const int j = cy - cbase[seg];
            if (j >= 0 && j < LZ_MAXC) s += part[(((size_t)f * n_seg + seg) * LZ_MAXC + j) * W + x];
        }
        tmp[x] = s / 255.0;  // im2double, applied once to the row-pass sum (linear map)
    }
    NSOF_SYNC();
    const int lane = NSOF_TID & 31, warp = NSOF_TID >> 5, nwarps = NSOF_NT >> 5;
    for (int cx = warp; cx < Wc; cx += (nwarps > 0 ? nwarps : 1)) {
        const int xa = xband[2 * cx], xb = xband[2 * cx + 1];
        double s = 0.;
        for (int x = xa + lane; x < xb; x += NSOF_WARP) s += Wx[(size_t)cx * W + x] * tmp[x];
        s = nsof_warp_sum_f64(s);
        if (lane == 0) cells[((size_t)f * Hc + cy) * Wc + cx] = s;
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

// simulate_memristor_array over a whole clip in one launch: thread = (clip, cell), loop over the frames
// inside; cells[t] is [n_cells] at stride `stride_t` doubles; current_out[t] = state after the pair (t, t+1).
//
// The nsub Euler sub-steps of a frame are a chain of dependent x^b evaluations: pure latency (a few thousand
// cells, one warp per block).  The chain is shortened without changing the recurrence
//     w <- clip(w + c * (1 - s w)^b, 0, 1),   c = g * dt_sub:
//  * the state is carried as u = 1 - s w (in [1 - s, 1]) and z = u^(b-1).  One Euler step is then
//        u' = u (1 + d),  d = -s c z          (because u' - u = -s (w' - w) = -s c u^b = u * (-s c z))
//    and the clip of w to [0, 1] is the clip of u' to [1 - s, 1], where z is a constant (1, or (1-s)^(b-1));
//  * z is carried forward multiplicatively, z' = z (1 + d)^(b-1), by a 6-term binomial series (|d| <= 2e-3:
//    truncation ~1e-16, Estrin form) instead of an exp(log) pair (~45 dependent operations);
//  * z is re-evaluated exactly, exp((b-1) log u), every 64 sub-steps and whenever |d| is larger, which bounds the
//    accumulated rounding to ~1e-14 relative.
NSOF_DEV double s1_pow_pos(double x, double b) { return x > 0. ? exp(b * log(x)) : pow(x, b); }

struct S1Binom { double a1, a2, a3, a4, a5, a6; };
NSOF_DEV S1Binom s1_binom(double b)
{
    S1Binom c;
    c.a1 = b;
    c.a2 = c.a1 * (b - 1.) / 2.;
    c.a3 = c.a2 * (b - 2.) / 3.;
    c.a4 = c.a3 * (b - 3.) / 4.;
    c.a5 = c.a4 * (b - 4.) / 5.;
    c.a6 = c.a5 * (b - 5.) / 6.;
    return c;
}

// nsub sub-steps of w <- clip(w + c * (1 - s w)^b, 0, 1) with c = g * dt_sub, 0 < s < 1.
// The loop carries d = -s c z itself (d' = d (1 + d)^(b-1)): the dependency is d -> series -> d, three FP64 levels.
// The series is evaluated unconditionally and ONE predicate (no clip, |d| small, not a re-evaluation step), ready
// before the series is, guards the rare path: a branch per test would serialise the in-order pipeline on each.
NSOF_DEV double s1_euler_chain(double wv, double c, double s, double b, int nsub)
{
    if (c == 0.) return wv;
    const double e = b - 1.;
    const S1Binom k = s1_binom(e);
    const double kap = -s * c;
    const double u_lo = 1. - s, z_lo = s1_pow_pos(u_lo, e);  // w = 1
    double u = 1. - wv * s;
    double d = kap * s1_pow_pos(u, e);
    for (int i = 0; i < nsub; i++) {
        // d (1 + a1 d + ... + a6 d^6) = (d + d^2 (a1 + a2 d)) + d^4 ((a3 + a4 d) + d^2 (a5 + a6 d)): three levels
        const double d2 = d * d, d4 = d2 * d2;
        double dn = (d + d2 * (k.a1 + k.a2 * d)) + d4 * ((k.a3 + k.a4 * d) + d2 * (k.a5 + k.a6 * d));
        double un = u + u * d;
        const bool fast = (un < 1.) & (un > u_lo) & (fabs(d) <= 2e-3) & ((i & 63) != 63);
        if (!fast) {
            if (un >= 1.) { un = 1.; dn = kap; }                 // w clipped at 0: z = 1
            else if (un <= u_lo) { un = u_lo; dn = kap * z_lo; }  // w clipped at 1
            else dn = kap * s1_pow_pos(un, e);                   // exact re-evaluation
        }
        u = un;
        d = dn;
    }
    if (u >= 1.) return 0.;
    if (u <= u_lo) return 1.;
    return fmax(0., fmin((1. - u) / s, 1.));
}

NSOF_GLOBAL k_s1_frame_run(const double *NSOF_RESTRICT cells, size_t stride_t, int n_frames, double *NSOF_RESTRICT w,
                           double *NSOF_RESTRICT current, size_t n, FrameSimConst c, int per_block)
{
    const size_t base = (size_t)NSOF_BX * per_block;
    for (int i = NSOF_TID; i < per_block; i += NSOF_NT) {
        const size_t e = base + i;
        if (e >= n) continue;
        double wv = w[e];
        double a = cells[e];
        for (int t = 0; t + 1 < n_frames; t++) {
            const double b = cells[(size_t)(t + 1) * stride_t + e];
            const double d = fabs(a * 256. - b * 256.);
            double V = 0.;
            if (d > c.th2) V = (d + 4.) * 0.75;
            if (d > c.th1 && d <= c.th2) V = (d + 4.) * 0.75;
            if (d <= c.th1) V = (d - 5.5) * 0.6;
            double vm = 0.;
            if (V > 0.) vm = 0.3 * V + 0.;
            else if (V < 0.) vm = 3. * V + -3.;
            vm = -vm;
            // one call for both branches (per-lane constants) so that a warp never runs the chain twice
            const bool off = vm < c.voff, on = !off && vm > c.von;
            const double gc = off ? c.koff * pow(vm / c.voff - 1., c.alphaoff) : on ? c.kon * pow(vm / c.von - 1., c.alphaon) : 0.;
            wv = s1_euler_chain(wv, gc * c.dt_sub, off ? c.soff : c.son, off ? c.boff : c.bon, c.nsub);
            if (current) current[(size_t)t * stride_t + e] = c.vds / (c.Ron / exp(-c.lam * (1. - wv)));
            a = b;
        }
        w[e] = wv;
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

// ---------------------------------------------------------------------------------------------
// S1-a on the tensor cores.  compress_image is a dense contraction, cells = Wy . img . Wx^T, and img is uint8: the
// COLUMN pass (contraction over x, the contiguous axis) is an exact integer GEMM once the double weights are cut
// into int8 slices (Ozaki splitting):  Wx[c][x] = 2^-E * sum_{s<7} q_s[c][x] * 2^{-7(s+1)},  |q_s| <= 127  (49 bits of
// every weight).  Every slice product sum_x q_s[c][x] * img[y][x] is an int32 dot product the
// int8 tensor-core path computes exactly (mma.sync m16n8k32 u8 x s8 -> s32: 16 rows x 32 columns of pixels against 32
// columns x 8 cells of one weight slice); the eight results are combined in double with power-of-two scales.  The row
// pass over the (H x cells) intermediate is then 0.1 % of the work and stays in FP64 (k_s1_lz_rows_f64).
// The FP64 kernels above cost 8 DFMA per source pixel and ran at 46 % of the FP64 pipe; this path issues one tensor-core
// instruction per 512 pixels and slice, reads pixels straight from global memory as fragment registers (no shared memory)
// and is bound by HBM.  Result vs the numpy restatement: the integer sums are exact, so it is the MORE accurate side
// (differences ~1e-16; tests hold 1e-13).  Cells are handled in windows of 8 (the N of the instruction); a window loops
// over the 32-column chunks its cells' taps touch.
// grid (ceil(H / 128), windows, frames), 256 threads: a warp owns 16 rows.   tmpx: [frame][H][windows * 8] doubles.
// ---------------------------------------------------------------------------------------------
#define LZ_SLICES 7   // 49 bits of every weight: residual < 2^-54 per tap relative to the largest weight (<= 1.5e-14 on a cell)
struct LzMmaConst {
    double scale[LZ_SLICES];  // 2^-(E + 7 (s + 1))
};

// W8 is stored in FRAGMENT ORDER: [window][chunk][slice][half][lane] 32-bit words, so that a warp's B operand is one
// contiguous 256-byte read (the first version read weights as [cell][x] rows: every load instruction touched 8 cache
// lines, 160 L1 wavefronts per 8 tensor-core instructions, and ran slower than the FP64 kernels).  The contraction index
// is permuted so that a lane's two A words of a row are adjacent too: lane t of a group covers physical columns
// 8t .. 8t+7 of the chunk (logical k = 4t + j for the first word, 16 + 4t + j for the second), one 8-byte load per row.
NSOF_HD size_t lz_w8_word(int nch, int win, int ch, int s, int half, int lane)
{
    return ((((size_t)win * nch + ch) * LZ_SLICES + s) * 32 + lane) * 2 + half;  // a lane's two words are one 8-byte load
}

NSOF_GLOBAL NSOF_LAUNCH_BOUNDS(256, 3) k_s1_lz_cols_mma(const uint8_t *NSOF_RESTRICT frames, ptrdiff_t pitch, ptrdiff_t frame_stride, int H, int W, int nch,
                             const unsigned *NSOF_RESTRICT W8, const int *NSOF_RESTRICT winx, LzMmaConst K, double *NSOF_RESTRICT tmpx)
{
    const int win = NSOF_BY, f = NSOF_BZ, nwin = nsof_grid_y();
    const int c0 = winx[2 * win], nc = winx[2 * win + 1];
    const uint8_t *img = frames + (size_t)f * frame_stride;
    const int NC = nwin * 8;
#ifdef NSOF_EMU
    for (int r = 0; r < 128; r++) {
        const int y = NSOF_BX * 128 + r;
        if (y >= H) break;
        for (int n = 0; n < 8; n++) {
            double v = 0.;
            for (int s = 0; s < LZ_SLICES; s++) {
                long long acc = 0;
                for (int x = c0 * 32; x < (c0 + nc) * 32 && x < W; x++) {
                    const int ch = x / 32, xo = x % 32, t = xo / 8, j8 = xo % 8;
                    const unsigned word = W8[lz_w8_word(nch, win, ch, s, j8 / 4, n * 4 + t)];
                    acc += (long long)(int8_t)((word >> (8 * (j8 % 4))) & 255u) * img[(size_t)y * pitch + x];
                }
                v += (double)acc * K.scale[s];
            }
            tmpx[((size_t)f * H + y) * NC + win * 8 + n] = v;
        }
    }
#else
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int y0 = blockIdx.x * 128 + warp * 16;
    if (y0 >= H) return;
    const uint8_t *rowA = img + (size_t)nsof_imin(y0 + g, H - 1) * pitch;
    const uint8_t *rowB = img + (size_t)nsof_imin(y0 + g + 8, H - 1) * pitch;
    const int xmax = W - 8;  // W % 8 == 0: an 8-byte group beyond it lies wholly in columns whose weights are zero
    int acc[LZ_SLICES][4];
#pragma unroll
    for (int s = 0; s < LZ_SLICES; s++) acc[s][0] = acc[s][1] = acc[s][2] = acc[s][3] = 0;
    // software pipeline: the pixels of chunk ch + 1 and ALL sixteen weight words of chunk ch are in flight before the
    // first tensor-core instruction of chunk ch issues (the first version used two register pairs for the weights and
    // every IMMA waited for its own load: ~1200 cycles per chunk)
    const int x0 = nsof_imin(c0 * 32 + 8 * t, xmax);
    uint2 pa = __ldg((const uint2 *)(rowA + x0)), pb = __ldg((const uint2 *)(rowB + x0));
    for (int ch = c0; ch < c0 + nc; ch++) {
        const unsigned *wp = W8 + lz_w8_word(nch, win, ch, 0, 0, lane);
        unsigned b[LZ_SLICES][2];
#pragma unroll
        for (int s = 0; s < LZ_SLICES; s++) { const uint2 w2 = __ldg((const uint2 *)(wp + s * 64)); b[s][0] = w2.x; b[s][1] = w2.y; }
        const int xn = nsof_imin(nsof_imin(ch + 1, c0 + nc - 1) * 32 + 8 * t, xmax);
        const uint2 na = __ldg((const uint2 *)(rowA + xn)), nb = __ldg((const uint2 *)(rowB + xn));
#pragma unroll
        for (int s = 0; s < LZ_SLICES; s++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(acc[s][0]), "+r"(acc[s][1]), "+r"(acc[s][2]), "+r"(acc[s][3])
                         : "r"(pa.x), "r"(pb.x), "r"(pa.y), "r"(pb.y), "r"(b[s][0]), "r"(b[s][1]));
        pa = na;
        pb = nb;
    }
    double v[4] = {0., 0., 0., 0.};
#pragma unroll
    for (int s = 0; s < LZ_SLICES; s++)
#pragma unroll
        for (int i = 0; i < 4; i++) v[i] += (double)acc[s][i] * K.scale[s];
    const int ya = y0 + g, yb = y0 + g + 8, col = win * 8 + 2 * t;
    if (ya < H) { double *o = tmpx + ((size_t)f * H + ya) * NC + col; o[0] = v[0]; o[1] = v[1]; }
    if (yb < H) { double *o = tmpx + ((size_t)f * H + yb) * NC + col; o[0] = v[2]; o[1] = v[3]; }
#endif
}

// Row pass over the intermediate: cells[f][r][c] = (1 / 255) * sum_y Wy[r][y] * tmpx[f][y][c], fixed summation order.
// grid (Hc, frames), 256 threads = 8 parts of the tap band x 32 cell columns.
NSOF_GLOBAL k_s1_lz_rows_f64(const double *NSOF_RESTRICT tmpx, int H, int NC, int Hc, int Wc, const double *NSOF_RESTRICT Wy,
                             const int *NSOF_RESTRICT yband, double *NSOF_RESTRICT cells)
{
    const int r = NSOF_BX, f = NSOF_BY;
    const int ya = yband[2 * r], yb = yband[2 * r + 1];
    NSOF_DYN_SMEM(smem_raw);
    double *part = (double *)smem_raw;  // [8][32]
    const int parts = 8, per = (yb - ya + parts - 1) / parts;
    for (int cb = 0; cb < Wc; cb += 32) {
        for (int i = NSOF_TID; i < parts * 32; i += NSOF_NT) {
            const int p = i / 32, c = cb + i % 32;
            double s = 0.;
            if (c < Wc) {
                const int y0 = ya + p * per, y1 = nsof_imin(yb, y0 + per);
                for (int y = y0; y < y1; y++) s += Wy[(size_t)r * H + y] * tmpx[((size_t)f * H + y) * NC + c];
            }
            part[i] = s;
        }
        NSOF_SYNC();
        for (int i = NSOF_TID; i < 32; i += NSOF_NT) {
            const int c = cb + i;
            if (c < Wc) {
                double s = 0.;
                for (int p = 0; p < parts; p++) s += part[p * 32 + i];
                cells[((size_t)f * Hc + r) * Wc + c] = s / 255.0;
            }
        }
        NSOF_SYNC();
    }
}
