// stage1.cu -- host driver + C ABI of stage 1 (synaptic temporal memory -> ROI), see include/nsof.h.
#include "nsof_handle.h"
#include "stage1_kernels.cuh"
#include <algorithm>
#include <cmath>

static const int S1_NT = 256;
static const int S1_PER_BLOCK = 1024;

static SynF32 syn_f32(const nsof_synapse_params *p, double dt)
{
    SynF32 s;
    s.voff = (float)p->voff; s.von = (float)p->von; s.koff = (float)p->koff; s.kon = (float)p->kon;
    s.son = (float)p->son; s.soff = (float)p->soff; s.bon = (float)p->bon; s.boff = (float)p->boff;
    s.alphaoff = (float)p->alphaoff; s.alphaon = (float)p->alphaon; s.dt = (float)dt;
    return s;
}

static inline int blocks_for(size_t n, int per_block) { return (int)((n + per_block - 1) / per_block); }

extern "C" int nsof_event_hist(nsof_handle *h, const int16_t *x, const int16_t *y, size_t n_events, int H, int W,
                               int32_t *counts, void *cuda_stream)
{
    if (!h || !counts || H < 1 || W < 1 || (n_events && (!x || !y))) NSOF_FAIL(NSOF_ERR_BAD_ARG, "event_hist: bad argument");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    NSOF_CUDA_TRY(cudaMemsetAsync(counts, 0, (size_t)H * W * sizeof(int32_t), st));
    if (n_events) {
        NSOF_KLAUNCH(h, NSOF_K_EVENT_HIST, 4. * n_events + 4. * H * W, k_s1_event_hist, blocks_for(n_events, S1_PER_BLOCK), 1, 1, S1_NT, 0, st, x, y, n_events, H, W, counts,
                    S1_PER_BLOCK);
        NSOF_CUDA_TRY(cudaGetLastError());
    }
    return NSOF_OK;
}

extern "C" int nsof_synapse_step_f32(nsof_handle *h, float *w, const float *V, size_t n, const nsof_synapse_params *p,
                                     double dt, void *cuda_stream)
{
    if (!h || !w || !V || !p) NSOF_FAIL(NSOF_ERR_BAD_ARG, "synapse_step: bad argument");
    if (n == 0) return NSOF_OK;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    NSOF_KLAUNCH(h, NSOF_K_SYNAPSE_STEP, 12. * n, k_s1_synapse_step, blocks_for(n, S1_PER_BLOCK), 1, 1, S1_NT, 0, (cudaStream_t)cuda_stream, w, V, n,
                syn_f32(p, dt), S1_PER_BLOCK);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_event_scheme1_f32(nsof_handle *h, const int16_t *x, const int16_t *y, const int64_t *slice_begin,
                                      int n_slices, int H, int W, int theta, float active_v, float silent_v,
                                      const nsof_synapse_params *p, double dt, float *w, void *cuda_stream)
{
    if (!h || !x || !y || !slice_begin || !p || !w || n_slices < 0 || H < 1 || W < 1)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "event_scheme1: bad argument");
    if (n_slices == 0) return NSOF_OK;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t npx = (size_t)H * W;
    const bool first_use = h->s1.bytes < npx * 4;
    int rc = h->s1.ensure(npx * 4);
    if (rc) return rc;
    const SynF32 sp = syn_f32(p, dt);
    // slice_begin lives on the device; the host needs the event counts per chunk to size grids.
    std::vector<int64_t> sb(n_slices + 1);
    NSOF_CUDA_TRY(cudaMemcpyAsync(sb.data(), slice_begin, sizeof(int64_t) * (n_slices + 1), cudaMemcpyDeviceToHost, st));
    NSOF_CUDA_TRY(cudaStreamSynchronize(st));
    if (theta <= 1) {
        uint32_t *bits = (uint32_t *)h->s1.p;
        if (first_use || true) NSOF_CUDA_TRY(cudaMemsetAsync(bits, 0, npx * 4, st));
        for (int s0 = 0; s0 < n_slices; s0 += 32) {
            const int ns = std::min(32, n_slices - s0);
            const int64_t ne = sb[s0 + ns] - sb[s0];
            if (ne > 0) {
                NSOF_KLAUNCH(h, NSOF_K_EVENT_HIST, 4. * ne, k_s1_slice_bits, blocks_for((size_t)ne, S1_PER_BLOCK), 1, 1, S1_NT, 0, st, x, y, slice_begin + s0,
                            ns, H, W, bits, S1_PER_BLOCK);
            }
            const bool silent_is_idle = !(silent_v < sp.voff) && !(silent_v > sp.von);
            if (ne > 0 || !silent_is_idle) {
                NSOF_KLAUNCH(h, NSOF_K_EVENT_REPLAY, 8. * npx, k_s1_replay_bits, blocks_for(npx, S1_PER_BLOCK), 1, 1, S1_NT, 0, st, w, bits, npx, ns, active_v,
                            silent_v, sp, S1_PER_BLOCK);
            }
        }
    } else {
        int32_t *counts = (int32_t *)h->s1.p;
        for (int s = 0; s < n_slices; s++) {
            const int64_t ne = sb[s + 1] - sb[s];
            NSOF_CUDA_TRY(cudaMemsetAsync(counts, 0, npx * 4, st));
            if (ne > 0)
                NSOF_KLAUNCH(h, NSOF_K_EVENT_HIST, 4. * ne, k_s1_event_hist, blocks_for((size_t)ne, S1_PER_BLOCK), 1, 1, S1_NT, 0, st, x + sb[s], y + sb[s],
                            (size_t)ne, H, W, counts, S1_PER_BLOCK);
            NSOF_KLAUNCH(h, NSOF_K_EVENT_REPLAY, 12. * npx, k_s1_step_from_counts, blocks_for(npx, S1_PER_BLOCK), 1, 1, S1_NT, 0, st, w, counts, npx, theta,
                        active_v, silent_v, sp, S1_PER_BLOCK);
        }
    }
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_event_scheme2_f32(nsof_handle *h, const int16_t *x, const int16_t *y, const int8_t *pol, const int64_t *t,
                                      const int64_t *slice_begin, int n_slices, int H, int W, int magnitude, float active_v,
                                      float bias_v, long long refractory_us, const nsof_synapse_params *p, double dt, float *w_a,
                                      float *w_b, int64_t *next_ok_a, int64_t *next_ok_b, void *cuda_stream)
{
    if (!h || !x || !y || !pol || !t || !slice_begin || !p || !w_a || !next_ok_a || n_slices < 0 || H < 1 || W < 1)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "event_scheme2: bad argument");
    if (!magnitude && (!w_b || !next_ok_b)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "event_scheme2: split polarity needs the B arrays");
    if (n_slices == 0) return NSOF_OK;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t npx = (size_t)H * W;
    // scratch: two hit-mask planes + first/last event time of the 32 slices of a chunk
    int rc = h->s1.ensure(npx * 8 + 64 * sizeof(int64_t));
    if (rc) return rc;
    uint32_t *bits_a = (uint32_t *)h->s1.p, *bits_b = bits_a + npx;
    int64_t *t_first = (int64_t *)(bits_b + npx), *t_last = t_first + 32;
    const SynF32 sp = syn_f32(p, dt);
    std::vector<int64_t> sb(n_slices + 1);
    NSOF_CUDA_TRY(cudaMemcpyAsync(sb.data(), slice_begin, sizeof(int64_t) * (n_slices + 1), cudaMemcpyDeviceToHost, st));
    NSOF_CUDA_TRY(cudaStreamSynchronize(st));
    NSOF_CUDA_TRY(cudaMemsetAsync(bits_a, 0, npx * 8, st));
    const bool bias_is_idle = !(bias_v < sp.voff) && !(bias_v > sp.von);
    for (int s0 = 0; s0 < n_slices; s0 += 32) {
        const int ns = std::min(32, n_slices - s0);
        const int64_t ne = sb[s0 + ns] - sb[s0];
        if (ne == 0 && bias_is_idle) continue;
        if (ne > 0) {
            NSOF_KLAUNCH(h, NSOF_K_EVENT_HIST, 5. * ne, k_s2_slice_bits, blocks_for((size_t)ne, S1_PER_BLOCK), 1, 1, S1_NT, 0, st, x, y, pol,
                         slice_begin + s0, ns, H, W, magnitude, bits_a, bits_b, S1_PER_BLOCK);
            NSOF_KLAUNCH(h, NSOF_K_EVENT_HIST, 16. * ns, k_s2_slice_times, 1, 1, 1, 32, 0, st, t, slice_begin + s0, ns, t_first, t_last);
        }
        NSOF_KLAUNCH(h, NSOF_K_EVENT_REPLAY, 24. * npx, k_s2_replay_bits, blocks_for(npx, S1_PER_BLOCK), 1, 1, S1_NT, 0, st, w_a, bits_a,
                     next_ok_a, npx, ns, t_first, t_last, (int64_t)refractory_us, active_v, bias_v, sp, S1_PER_BLOCK);
        if (!magnitude)
            NSOF_KLAUNCH(h, NSOF_K_EVENT_REPLAY, 24. * npx, k_s2_replay_bits, blocks_for(npx, S1_PER_BLOCK), 1, 1, S1_NT, 0, st, w_b, bits_b,
                         next_ok_b, npx, ns, t_first, t_last, (int64_t)refractory_us, active_v, bias_v, sp, S1_PER_BLOCK);
    }
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_resistance_exp_f32(nsof_handle *h, const float *w, size_t n, const nsof_synapse_params *p, double *R,
                                       void *cuda_stream)
{
    if (!h || !w || !R || !p) NSOF_FAIL(NSOF_ERR_BAD_ARG, "resistance_exp: bad argument");
    if (n == 0) return NSOF_OK;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const double lam = log(p->Roff / p->Ron);
    NSOF_KLAUNCH(h, NSOF_K_RESISTANCE, 12. * n, k_s1_resistance_exp, blocks_for(n, S1_PER_BLOCK), 1, 1, S1_NT, 0, (cudaStream_t)cuda_stream, w, n, p->Ron,
                lam, R, S1_PER_BLOCK);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_event_cell_currents_f32(nsof_handle *h, const float *w, int H, int W, int memsize_h, int memsize_w,
                                            const nsof_synapse_params *p, double *currents, void *cuda_stream)
{
    if (!h || !w || !p || !currents || H < 1 || W < 1 || memsize_h < 1 || memsize_w < 1 || memsize_h > H || memsize_w > W)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "event_cell_currents: bad argument");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const int Hc = H / memsize_h, Wc = W / memsize_w;
    const double lam = log(p->Roff / p->Ron);
    NSOF_KLAUNCH(h, NSOF_K_RESISTANCE, 4. * Hc * memsize_h * Wc * memsize_w, k_s1_cell_current_from_w, Hc * Wc, 1, 1, 128,
                 128 * sizeof(float), (cudaStream_t)cuda_stream, w, W, memsize_h, memsize_w, Wc, p->Ron, lam, currents);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

// ------------------------------------------------------------------------------ frame mode
// MATLAB imresize 'lanczos3' contributions for one dimension (antialiased when shrinking):
//   scale = out/in; kernel width 6 (/scale when scale < 1); u = x/scale + 0.5*(1 - 1/scale), x = 1..out;
//   left = floor(u - kw/2); P = ceil(kw) + 2 taps; weights h(u - idx) normalised to sum 1;
//   indices mirrored at the borders (aux = [1:in, in:-1:1]).
static double lanczos3(double x)
{
    const double pi = 3.14159265358979323846;
    const double eps = 2.220446049250313e-16;
    double f = (sin(pi * x) * sin(pi * x / 3.) + eps) / ((pi * pi * x * x / 3.) + eps);
    return fabs(x) < 3. ? f : 0.;
}
static int lanczos_contrib(int in, int out, std::vector<double> &wgt, std::vector<int> &idx)
{
    const double scale = (double)out / in;
    double kw = 6.;
    const bool aa = scale < 1.;
    if (aa) kw = kw / scale;
    const int P = (int)ceil(kw) + 2;
    wgt.assign((size_t)out * P, 0.);
    idx.assign((size_t)out * P, 0);
    for (int o = 0; o < out; o++) {
        const double x = o + 1;
        const double u = x / scale + 0.5 * (1. - 1. / scale);
        const double left = floor(u - kw / 2.);
        double sum = 0.;
        for (int k = 0; k < P; k++) {
            const double ind = left + k;
            const double v = aa ? scale * lanczos3(scale * (u - ind)) : lanczos3(u - ind);
            wgt[(size_t)o * P + k] = v;
            sum += v;
            // mirror: aux = [1:in, in:-1:1]; aux(mod(ind-1, 2*in) + 1)
            long long m = ((long long)ind - 1) % (2LL * in);
            if (m < 0) m += 2LL * in;
            const int one_based = m < in ? (int)m + 1 : (int)(2LL * in - m);
            idx[(size_t)o * P + k] = one_based - 1;
        }
        for (int k = 0; k < P; k++) wgt[(size_t)o * P + k] /= sum;
    }
    return P;
}

// Dense per-axis matrix D[out][in] of the resize (mirrored taps merged) and, per output, its non-zero band.
static void lanczos_dense(int in, int out, std::vector<double> &D, std::vector<int> &band)
{
    std::vector<double> wgt;
    std::vector<int> idx;
    const int P = lanczos_contrib(in, out, wgt, idx);
    D.assign((size_t)out * in, 0.);
    band.assign((size_t)out * 2, 0);
    for (int o = 0; o < out; o++) {
        int lo = in, hi = 0;
        for (int k = 0; k < P; k++) {
            const double v = wgt[(size_t)o * P + k];
            if (v == 0.) continue;
            const int i = idx[(size_t)o * P + k];
            D[(size_t)o * in + i] += v;
            lo = std::min(lo, i);
            hi = std::max(hi, i + 1);
        }
        if (lo >= hi) { lo = 0; hi = 0; }
        band[2 * o] = lo;
        band[2 * o + 1] = hi;
    }
}

extern "C" int nsof_lanczos3_compress_u8(nsof_handle *h, const uint8_t *frames, int n, int H, int W, ptrdiff_t pitch,
                                         ptrdiff_t frame_stride, int Hc, int Wc, double *cells, void *cuda_stream)
{
    if (!h || !frames || !cells || n < 0 || H < 1 || W < 1 || Hc < 1 || Wc < 1 || pitch < W)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "lanczos3_compress: bad argument");
    if (n == 0) return NSOF_OK;
    if (n > 65535 || Hc > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "lanczos3_compress: too many frames/cells per call");
    if ((size_t)W * sizeof(double) > 200 * 1024) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "lanczos3_compress: frame wider than 25600 px");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    int rc;
    if (h->lz_key[0] != H || h->lz_key[1] != W || h->lz_key[2] != Hc || h->lz_key[3] != Wc) {
        std::vector<double> Dy, Dx;
        std::vector<int> yband, xband;
        lanczos_dense(H, Hc, Dy, yband);
        lanczos_dense(W, Wc, Dx, xband);
        // row segments: the largest seg_rows (<= 96) whose segments all touch at most LZ_MAXC consecutive cell rows
        int seg_rows = std::min(96, H), n_seg = 0;
        std::vector<int> cbase;
        for (;; seg_rows = std::max(1, seg_rows / 2)) {
            n_seg = (H + seg_rows - 1) / seg_rows;
            cbase.assign(n_seg, 0);
            bool ok = true;
            for (int sgm = 0; sgm < n_seg && ok; sgm++) {
                const int y0 = sgm * seg_rows, y1 = std::min(H, y0 + seg_rows);
                int lo = Hc, hi = -1;
                for (int cy = 0; cy < Hc; cy++)
                    if (yband[2 * cy] < y1 && yband[2 * cy + 1] > y0) { lo = std::min(lo, cy); hi = std::max(hi, cy); }
                if (hi < lo) { lo = 0; hi = 0; }
                cbase[sgm] = lo;
                if (hi - lo + 1 > LZ_MAXC) ok = false;
            }
            if (ok) break;
            if (seg_rows == 1) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "lanczos3_compress: a source row feeds more than %d cell rows", LZ_MAXC);
        }
        // tensor-core column pass (k_s1_lz_cols_mma): windows of 8 cells, their 32-column chunk ranges, and the weights of
        // every window cut into 8 int8 slices of 7 bits: Wx = 2^-E sum_s q_s 2^{-7 (s + 1)}
        const int nwin = (Wc + 7) / 8, Wp = (W + 31) & ~31;
        std::vector<int> winx(2 * (size_t)nwin, 0);
        const int nch = Wp / 32;
        std::vector<unsigned> W8((size_t)nwin * nch * LZ_SLICES * 64, 0u);  // fragment order, see lz_w8_word
        double maxw = 0.;
        for (double v : Dx) maxw = std::max(maxw, fabs(v));
        int e2 = 0;
        frexp(maxw, &e2);  // maxw = m 2^e2, m in [0.5, 1)  ->  |w| 2^-e2 < 1
        for (int win = 0; win < nwin; win++) {
            int lo = W, hi = 0;
            for (int c = win * 8; c < std::min(Wc, win * 8 + 8); c++) { lo = std::min(lo, xband[2 * c]); hi = std::max(hi, xband[2 * c + 1]); }
            if (lo >= hi) { lo = 0; hi = 1; }
            winx[2 * win] = lo / 32;
            winx[2 * win + 1] = (hi + 31) / 32 - lo / 32;
            for (int c = win * 8; c < std::min(Wc, win * 8 + 8); c++)
                for (int x = 0; x < W; x++) {
                    double r = ldexp(Dx[(size_t)c * W + x], -e2);
                    const int ch = x / 32, xo = x % 32, t = xo / 8, j8 = xo % 8;
                    for (int sl = 0; sl < LZ_SLICES; sl++) {
                        r *= 128.;
                        const double q = trunc(r);
                        r -= q;
                        const unsigned byte = (unsigned)(uint8_t)(int8_t)q;
                        W8[lz_w8_word(nch, win, ch, sl, j8 / 4, (c - win * 8) * 4 + t)] |= byte << (8 * (j8 % 4));
                    }
                }
        }
        h->lz_mma[0] = nwin; h->lz_mma[1] = Wp; h->lz_mma[2] = e2;
        const size_t off_yband = (Dy.size() + Dx.size()) * sizeof(double) + (xband.size() + cbase.size()) * sizeof(int);
        const size_t off_winx = off_yband + yband.size() * sizeof(int);
        const size_t off_w8 = (off_winx + winx.size() * sizeof(int) + 15) & ~(size_t)15;
        h->lz_off[0] = off_yband; h->lz_off[1] = off_winx; h->lz_off[2] = off_w8;
        const size_t bytes = off_w8 + W8.size() * sizeof(unsigned) + 64;
        rc = h->lz.ensure(bytes);
        if (rc) return rc;
        // tables are small and change only with the geometry: plain synchronous upload
        char *d = (char *)h->lz.p;
        NSOF_CUDA_TRY(cudaStreamSynchronize(st));
        NSOF_CUDA_TRY(cudaMemcpyAsync(d, Dy.data(), Dy.size() * 8, cudaMemcpyHostToDevice, st)); d += Dy.size() * 8;
        NSOF_CUDA_TRY(cudaMemcpyAsync(d, Dx.data(), Dx.size() * 8, cudaMemcpyHostToDevice, st)); d += Dx.size() * 8;
        NSOF_CUDA_TRY(cudaMemcpyAsync(d, xband.data(), xband.size() * 4, cudaMemcpyHostToDevice, st)); d += xband.size() * 4;
        NSOF_CUDA_TRY(cudaMemcpyAsync(d, cbase.data(), cbase.size() * 4, cudaMemcpyHostToDevice, st));
        NSOF_CUDA_TRY(cudaMemcpyAsync((char *)h->lz.p + off_yband, yband.data(), yband.size() * 4, cudaMemcpyHostToDevice, st));
        NSOF_CUDA_TRY(cudaMemcpyAsync((char *)h->lz.p + off_winx, winx.data(), winx.size() * 4, cudaMemcpyHostToDevice, st));
        NSOF_CUDA_TRY(cudaMemcpyAsync((char *)h->lz.p + off_w8, W8.data(), W8.size() * sizeof(unsigned), cudaMemcpyHostToDevice, st));
        NSOF_CUDA_TRY(cudaStreamSynchronize(st));
        h->lz_key[0] = H; h->lz_key[1] = W; h->lz_key[2] = Hc; h->lz_key[3] = Wc;
        h->lz_P[0] = seg_rows; h->lz_P[1] = n_seg;
    }
    const int seg_rows = h->lz_P[0], n_seg = h->lz_P[1];
    const double *d_Dy = (const double *)h->lz.p;
    const double *d_Dx = d_Dy + (size_t)Hc * H;
    const int *d_xband = (const int *)(d_Dx + (size_t)Wc * W);
    const int *d_cbase = d_xband + 2 * (size_t)Wc;
    // tensor-core path: word-aligned rows (the fragment registers are loaded straight from global memory)
    const bool mma_ok = !h->dbg_lz_f64 && W >= 32 && W % 8 == 0 && pitch % 8 == 0 && frame_stride % 8 == 0 && ((uintptr_t)frames & 7) == 0;
    if (mma_ok) {
        const int nwin = h->lz_mma[0], Wp = h->lz_mma[1], e2 = h->lz_mma[2], NC = nwin * 8;
        rc = h->s1.ensure((size_t)n * H * NC * sizeof(double));
        if (rc) return rc;
        double *tmpx = (double *)h->s1.p;
        LzMmaConst K;
        for (int sl = 0; sl < LZ_SLICES; sl++) K.scale[sl] = ldexp(1., e2 - 7 * (sl + 1));
        const char *tb = (const char *)h->lz.p;
        NSOF_KLAUNCH(h, NSOF_K_LANCZOS, (double)n * H * W, k_s1_lz_cols_mma, (H + 127) / 128, nwin, n, 256, 0, st, frames, pitch,
                     frame_stride, H, W, Wp / 32, (const unsigned *)(tb + h->lz_off[2]), (const int *)(tb + h->lz_off[1]), K, tmpx);
        h->prof_level = 1;  // the per-launch profile shows the two passes as s1_lanczos (column pass) and s1_lanczos.L1 (row pass)
        NSOF_KLAUNCH(h, NSOF_K_LANCZOS, 0., k_s1_lz_rows_f64, Hc, n, 1, 256, 8 * 32 * sizeof(double), st, (const double *)tmpx, H, NC, Hc,
                     Wc, d_Dy, (const int *)(tb + h->lz_off[0]), cells);
        h->prof_level = -1;
        NSOF_CUDA_TRY(cudaGetLastError());
        return NSOF_OK;
    }
    rc = h->s1.ensure((size_t)n * n_seg * LZ_MAXC * W * sizeof(double));
    if (rc) return rc;
    double *part = (double *)h->s1.p;
    if (!h->lz_smem_set) {
        NSOF_CUDA_TRY(cudaFuncSetAttribute(k_s1_lanczos_cols, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        NSOF_CUDA_TRY(cudaFuncSetAttribute(k_s1_lanczos_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        h->lz_smem_set = true;
    }
    const size_t rows_smem = 64 + (size_t)LZ_MAXC * seg_rows * sizeof(double) + (size_t)seg_rows * LZ_TILE_W;
    NSOF_KLAUNCH(h, NSOF_K_LANCZOS, (double)n * H * W, k_s1_lanczos_rows, (W + LZ_TILE_W - 1) / LZ_TILE_W, n_seg, n,
                 LZ_TILE_W / 4, rows_smem, st, frames, pitch, frame_stride, H, W, Hc, d_Dy, d_cbase, seg_rows, n_seg, part);
    NSOF_KLAUNCH(h, NSOF_K_LANCZOS, 0., k_s1_lanczos_cols, Hc, n, 1, S1_NT, (size_t)W * sizeof(double), st, part, W, Hc, Wc,
                 d_Dx, d_xband, d_cbase, n_seg, cells);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

// 1: take the FP64 Lanczos kernels even where the tensor-core column pass applies (tests, A/B)
extern "C" int nsof_debug_lanczos_f64(nsof_handle *h, int enable)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->dbg_lz_f64 = enable != 0;
    return NSOF_OK;
}

static FrameSimConst frame_sim_const(const nsof_synapse_params *p, const nsof_frame_sim_params *fp)
{
    FrameSimConst c;
    c.th1 = fp->th1; c.th2 = fp->th2; c.dt_sub = fp->dt / fp->nsubsteps; c.vds = fp->vds; c.nsub = fp->nsubsteps;
    c.alphaoff = p->alphaoff; c.alphaon = p->alphaon; c.voff = p->voff; c.von = p->von; c.koff = p->koff; c.kon = p->kon;
    c.son = p->son; c.soff = p->soff; c.bon = p->bon; c.boff = p->boff; c.Ron = p->Ron; c.lam = log(p->Roff / p->Ron);
    return c;
}

extern "C" int nsof_frame_sim_run_f64(nsof_handle *h, const double *cells, int n_frames, int n, int Hc, int Wc, double *w,
                                      double *current_out, const nsof_synapse_params *p, const nsof_frame_sim_params *fp,
                                      void *cuda_stream)
{
    if (!h || !cells || !w || !p || !fp || n < 0 || n_frames < 1 || Hc < 1 || Wc < 1 || fp->nsubsteps < 1)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "frame_sim_run: bad argument");
    if (n == 0 || n_frames < 2) return NSOF_OK;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const FrameSimConst c = frame_sim_const(p, fp);
    const size_t total = (size_t)n * Hc * Wc;
    const int per_block = 32;  // few cells, long dependent chains: one warp per block spreads the cells over the SMs
    NSOF_KLAUNCH(h, NSOF_K_FRAME_SIM, 8. * total * (2. * n_frames - 1), k_s1_frame_run, blocks_for(total, per_block), 1, 1, 32,
                 0, (cudaStream_t)cuda_stream, cells, total, n_frames, w, current_out, total, c, per_block);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_frame_sim_step_f64(nsof_handle *h, const double *cells_prev, const double *cells_next, double *w,
                                       double *current_out, int n, int Hc, int Wc, const nsof_synapse_params *p,
                                       const nsof_frame_sim_params *fp, void *cuda_stream)
{
    if (!h || !cells_prev || !cells_next || !w || !p || !fp || n < 0 || Hc < 1 || Wc < 1 || fp->nsubsteps < 1)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "frame_sim_step: bad argument");
    if (n == 0) return NSOF_OK;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const FrameSimConst c = frame_sim_const(p, fp);
    const size_t total = (size_t)n * Hc * Wc;
    // few cells, long dependent chains: small blocks so the cells spread over many SMs
    const int per_block = 32;
    NSOF_KLAUNCH(h, NSOF_K_FRAME_SIM, 40. * total, k_s1_frame_step, blocks_for(total, per_block), 1, 1, 32, 0, (cudaStream_t)cuda_stream, cells_prev, cells_next,
                w, current_out, total, c, per_block);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

// ------------------------------------------------------------------------------ ROI tail
static int validate_roi(const nsof_roi_params *p)
{
    if (!p) NSOF_FAIL(NSOF_ERR_BAD_ARG, "roi: params is NULL");
    if (p->cells_h < 1 || p->cells_w < 1 || p->frame_h < 1 || p->frame_w < 1 || p->pixel_h < 1 || p->pixel_w < 1)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "roi: bad geometry");
    if ((size_t)p->cells_h * p->cells_w > 40000) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "roi: more than 40000 cells per state");
    if (p->flag != 1 && p->flag != 2) NSOF_FAIL(NSOF_ERR_BAD_ARG, "roi: flag must be 1 (separate) or 2 (merged)");
    return NSOF_OK;
}

extern "C" int nsof_roi_from_state_f64(nsof_handle *h, const double *currents, int n, const nsof_roi_params *p,
                                       const double *thr_table, uint8_t *memimg, int32_t *labels, int32_t *n_components,
                                       int32_t *n_rects, nsof_rect *rects, int32_t *stats, void *cuda_stream)
{
    if (!h || !currents || !thr_table || n < 0) NSOF_FAIL(NSOF_ERR_BAD_ARG, "roi: bad argument");
    int rc = validate_roi(p);
    if (rc) return rc;
    if (n == 0) return NSOF_OK;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t smem = ((size_t)p->cells_h * p->cells_w + 1 + 5 * NSOF_MAX_RECTS) * sizeof(int);
    if (!h->roi_smem_set) {  // the attribute is per device: kept per handle, like lz_smem_set
        NSOF_CUDA_TRY(cudaFuncSetAttribute(k_s1_roi, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        h->roi_smem_set = true;
    }
    NSOF_KLAUNCH(h, NSOF_K_ROI, (double)n * p->cells_h * p->cells_w * 13., k_s1_roi, n, 1, 1, 128, smem, (cudaStream_t)cuda_stream, currents, *p, thr_table, memimg, labels,
                n_components, n_rects, rects, stats);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_roi_from_state_f64_host(nsof_handle *h, const double *currents, int n, const nsof_roi_params *p,
                                            const double *thr_table, uint8_t *memimg, int32_t *labels,
                                            int32_t *n_components, int32_t *n_rects, nsof_rect *rects, int32_t *stats)
{
    if (!h || !currents || !thr_table || n < 0) NSOF_FAIL(NSOF_ERR_BAD_ARG, "roi: bad argument");
    int rc = validate_roi(p);
    if (rc) return rc;
    if (n == 0) return NSOF_OK;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t N = (size_t)p->cells_h * p->cells_w;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t o_cur = 0, o_thr = o_cur + up(n * N * 8), o_mem = o_thr + up(256 * 8), o_lab = o_mem + up(n * N),
                 o_nc = o_lab + up(n * N * 4), o_nr = o_nc + up((size_t)n * 4), o_rc = o_nr + up((size_t)n * 4),
                 o_st = o_rc + up((size_t)n * NSOF_MAX_RECTS * sizeof(nsof_rect)),
                 total = o_st + up((size_t)n * NSOF_MAX_RECTS * 5 * 4);
    rc = h->io.ensure(total);
    if (rc) return rc;
    char *d = (char *)h->io.p;
    cudaStream_t st = nullptr;
    NSOF_CUDA_TRY(cudaMemcpyAsync(d + o_cur, currents, n * N * 8, cudaMemcpyHostToDevice, st));
    NSOF_CUDA_TRY(cudaMemcpyAsync(d + o_thr, thr_table, 256 * 8, cudaMemcpyHostToDevice, st));
    rc = nsof_roi_from_state_f64(h, (const double *)(d + o_cur), n, p, (const double *)(d + o_thr), (uint8_t *)(d + o_mem),
                                 (int32_t *)(d + o_lab), (int32_t *)(d + o_nc), (int32_t *)(d + o_nr), (nsof_rect *)(d + o_rc),
                                 (int32_t *)(d + o_st), st);
    if (rc) return rc;
    if (memimg) NSOF_CUDA_TRY(cudaMemcpyAsync(memimg, d + o_mem, n * N, cudaMemcpyDeviceToHost, st));
    if (labels) NSOF_CUDA_TRY(cudaMemcpyAsync(labels, d + o_lab, n * N * 4, cudaMemcpyDeviceToHost, st));
    if (n_components) NSOF_CUDA_TRY(cudaMemcpyAsync(n_components, d + o_nc, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    if (n_rects) NSOF_CUDA_TRY(cudaMemcpyAsync(n_rects, d + o_nr, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    if (rects) NSOF_CUDA_TRY(cudaMemcpyAsync(rects, d + o_rc, (size_t)n * NSOF_MAX_RECTS * sizeof(nsof_rect), cudaMemcpyDeviceToHost, st));
    if (stats) NSOF_CUDA_TRY(cudaMemcpyAsync(stats, d + o_st, (size_t)n * NSOF_MAX_RECTS * 5 * 4, cudaMemcpyDeviceToHost, st));
    NSOF_CUDA_TRY(cudaStreamSynchronize(st));
    return NSOF_OK;
}
