// farneback.cu -- host driver + C ABI of stage 2 (dense Farneback), see include/nsof.h.
//
// Follows the level loop of OpenCV's FarnebackOpticalFlowImpl::calc, restated for
// a batch of independent crop pairs: every kernel launch covers one stage of one pyramid level of
// ALL crops of the batch (grid = max tiles x items), so small ROI crops still fill 148 SMs.
#include "nsof_handle.h"
#include "farneback_kernels.cuh"
#include <float.h>
#include <algorithm>
#include <unordered_map>

// ------------------------------------------------------------------------------ error + handle
static thread_local char g_err[512] = "";
void nsof_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char *nsof_last_error(void) { return g_err; }
extern "C" int nsof_version(void) { return 100; }

int NsofDevBuf::ensure(size_t need)
{
    if (need <= bytes) return NSOF_OK;
    if (p) {
        NSOF_CUDA_TRY(cudaDeviceSynchronize());
        NSOF_CUDA_TRY(cudaFree(p));
        p = nullptr;
        bytes = 0;
    }
    size_t want = need + need / 8;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
        p = nullptr;
        (void)cudaGetLastError();
        NSOF_FAIL(NSOF_ERR_NOMEM, "cudaMalloc of %zu bytes failed: %s", want, cudaGetErrorString(e));
    }
    bytes = want;
    return NSOF_OK;
}
void NsofDevBuf::release()
{
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
}

int nsof_stage_acquire(nsof_handle *h, size_t bytes, NsofStageSlot **out)
{
    NsofStageSlot &s = h->stage[h->stage_next];
    h->stage_next = (h->stage_next + 1) % 4;
    if (s.pending) {
        NSOF_CUDA_TRY(cudaEventSynchronize(s.done));
        s.pending = false;
    }
    if (!s.done) NSOF_CUDA_TRY(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    if (s.bytes < bytes + 16) {  // uploads read whole 16-byte words
        if (s.host) NSOF_CUDA_TRY(cudaFreeHost(s.host));
        s.host = nullptr;
        s.bytes = 0;
        NSOF_CUDA_TRY(cudaMallocHost(&s.host, bytes + bytes / 4 + 4096));
        s.bytes = bytes + bytes / 4 + 4096;
    }
    *out = &s;
    return NSOF_OK;
}

// Job tables reach the device through a kernel that reads the pinned staging slot over PCIe, not through
// cudaMemcpyAsync: a host-to-device copy queues on the copy engine behind every frame copy already submitted (5 ms
// of them per step in the host pipeline) and the batch's first kernel would sit behind all of them.
NSOF_GLOBAL k_table_upload(float4 *NSOF_RESTRICT dst, const float4 *NSOF_RESTRICT src, size_t n16)
{
    for (size_t i = (size_t)NSOF_BX * NSOF_NT + NSOF_TID; i < n16; i += (size_t)nsof_grid_x() * NSOF_NT) dst[i] = src[i];
}

// `dst` (NsofDevBuf memory) and the slot are both allocated with >= 16 bytes of slack past `bytes`.
int nsof_table_upload(nsof_handle *h, void *dst, NsofStageSlot *slot, size_t bytes, cudaStream_t stream)
{
    const size_t n16 = (bytes + 15) / 16;
    const int grid = (int)std::min<size_t>(32, std::max<size_t>(1, n16 / 256));
    NSOF_KLAUNCH(h, NSOF_K_TABLE_UPLOAD, 2.0 * (double)bytes, k_table_upload, grid, 1, 1, 256, 0, stream, (float4 *)dst,
                 (const float4 *)slot->host, n16);
    NSOF_CUDA_TRY(cudaGetLastError());
    NSOF_CUDA_TRY(cudaEventRecord(slot->done, stream));
    slot->pending = true;
    return NSOF_OK;
}

struct FbItemPlan {
    int scales;
    int hk[64], wk[64], pitch[64];
    size_t off_I[64], off_R[64];                  // per level: the levels' pyramids are built concurrently
    size_t off_Ma, off_Mb, off_fA, off_fB;        // shared by the levels (serial chain); floats from arena base
    int r0_from;                                  // item whose R[1] serves as this item's R[0] (shared frame), or -1
    nsof_crop_desc d;
};

static void fb_ctx_free(nsof_handle *h);  // defined with FbBatchCtx
static void fb_link_close(nsof_handle *h);  // defined with nsof_fb_tiled_run

extern "C" int nsof_create(int device, nsof_handle **out)
{
    if (!out) NSOF_FAIL(NSOF_ERR_BAD_ARG, "nsof_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0) {
        (void)cudaGetLastError();
        NSOF_FAIL(NSOF_ERR_CUDA, "nsof_create: no CUDA device (%s); libnsof has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (device < 0 || device >= count) NSOF_FAIL(NSOF_ERR_BAD_ARG, "nsof_create: device %d of %d", device, count);
    NSOF_CUDA_TRY(cudaSetDevice(device));
    nsof_handle *h = new nsof_handle();
    h->device = device;
    h->last_plan = new std::vector<FbItemPlan>();
    if (const char *e = getenv("NSOF_SHARE_FRAMES")) h->dbg_no_dedup = e[0] == '0';  // A/B measurements (default: on)
    if (const char *e = getenv("NSOF_R1_WINDOW")) {  // A/B measurements (default: on; 2 = blur+solve only)
        h->dbg_no_r1_window = e[0] == '0';
        h->dbg_um_round1 = e[0] == '0' || e[0] == '2';
    }
    if (const char *e = getenv("NSOF_GENERIC_BLUR")) h->dbg_generic_blur = e[0] == '1';   // A/B: winsize 3 .. 5 on the generic kernel
    if (const char *e = getenv("NSOF_BS_NO_PREFETCH")) h->dbg_bs_no_prefetch = atoi(e) < 0 ? -1 : (atoi(e) & 3);
    if (const char *e = getenv("NSOF_LANCZOS_F64")) h->dbg_lz_f64 = e[0] == '1';     // A/B: FP64 Lanczos kernels
    *out = h;
    return NSOF_OK;
}

extern "C" int nsof_destroy(nsof_handle *h)
{
    if (!h) return NSOF_OK;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    h->arena.release();
    h->tile_buf.release();
    nsof_jpeg_release(h);
    h->jpeg_buf.release();
    fb_link_close(h);
    h->link_flags.release();
    h->link_tab.release();
    h->tables.release();
    h->clear_tab.release();
    h->guard_buf.release();
    h->r1_stats_buf.release();
    fb_ctx_free(h);
    h->io.release();
    h->s1.release();
    h->seg.release();
    h->warp_tab.release();
    h->lz.release();
    for (auto &s : h->stage) {
        if (s.host) cudaFreeHost(s.host);
        if (s.done) cudaEventDestroy(s.done);
    }
    for (auto e : h->prof_pool) cudaEventDestroy(e);
    for (auto e : h->ev_level) cudaEventDestroy(e);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    for (auto &a : h->aux)
        if (a) cudaStreamDestroy(a);
    delete h->last_plan;
    delete h;
    return NSOF_OK;
}

extern "C" size_t nsof_workspace_bytes(const nsof_handle *h)
{
    return h ? h->arena.bytes + h->tables.bytes + h->io.bytes + h->s1.bytes : 0;
}

// ------------------------------------------------------------------------------ launch accounting
const char *const nsof_kernel_class_names[NSOF_K_CLASSES] = {
    "fb_pyr", "fb_polyexp", "fb_pyr_polyexp", "fb_um_first", "fb_blur_um", "fb_blur_flow", "fb_flow_init", "s1_lanczos", "s1_frame_sim",
    "s1_roi", "s1_event_hist", "s1_event_replay", "s1_synapse_step", "s1_resistance", "fb_flow_clear", "halo_copy", "table_upload",
    "seg_pack", "seg_morph", "seg_unpack", "rgb2gray", "remap_flow", "fb_tile_list", "ssim"};

extern "C" unsigned long long nsof_launch_count(const nsof_handle *h) { return h ? h->launches : 0; }

extern "C" int nsof_profile(nsof_handle *h, int enable)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "nsof_profile: NULL handle");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    NSOF_CUDA_TRY(cudaDeviceSynchronize());
    h->prof_on = enable != 0;
    if (enable) {
        h->prof_used = 0;
        h->prof_recs.clear();
    }
    return NSOF_OK;
}

extern "C" int nsof_profile_read(nsof_handle *h, nsof_kernel_stat *out, int max_out, int *n_out)
{
    if (!h || !out || !n_out) NSOF_FAIL(NSOF_ERR_BAD_ARG, "nsof_profile_read: NULL argument");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    NSOF_CUDA_TRY(cudaDeviceSynchronize());
    const int LV = 9;  // levels 0..7 and "no level"
    std::vector<nsof_kernel_stat> acc((size_t)NSOF_K_CLASSES * LV);
    memset(acc.data(), 0, acc.size() * sizeof(nsof_kernel_stat));
    for (const NsofProfRec &r : h->prof_recs) {
        float ms = 0.f;
        NSOF_CUDA_TRY(cudaEventElapsedTime(&ms, h->prof_pool[r.ev], h->prof_pool[r.ev + 1]));
        const int lv = (r.level >= 0 && r.level < LV - 1) ? r.level : LV - 1;
        nsof_kernel_stat &a = acc[(size_t)r.cls * LV + lv];
        a.launches += 1;
        a.ms += ms;
        a.algorithmic_bytes += r.bytes;
    }
    int n = 0;
    for (int c = 0; c < NSOF_K_CLASSES; c++)
        for (int lv = 0; lv < LV && n < max_out; lv++) {
            nsof_kernel_stat &a = acc[(size_t)c * LV + lv];
            if (!a.launches) continue;
            if (lv == LV - 1) snprintf(a.name, sizeof(a.name), "%s", nsof_kernel_class_names[c]);
            else snprintf(a.name, sizeof(a.name), "%s.L%d", nsof_kernel_class_names[c], lv);
            out[n++] = a;
        }
    *n_out = n;
    return NSOF_OK;
}

// ------------------------------------------------------------------------------ geometry (A.1)
static int cv_round(double v) { return (int)nearbyint(v); }  // round-half-even like cvRound

extern "C" int nsof_fb_level_geometry(int h, int w, double pyr_scale, int levels, int *hk, int *wk, int *ksize,
                                      double *sigma)
{
    if (levels > 62) levels = 62;
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
        ks = std::max(ks, 3);
        if (hk) hk[k] = cv_round(h * scale);
        if (wk) wk[k] = cv_round(w * scale);
        if (ksize) ksize[k] = ks;
        if (sigma) sigma[k] = sg;
    }
    return cropped + 1;
}

// cv::getGaussianKernel(ksize, sigma, CV_32F) -> half kernel: gk[0] centre, gk[i] distance i (A.2)
static void gaussian_half_kernel(int ksize, double sigma, float *gk)
{
    const int r = ksize / 2;
    if (sigma <= 0 && ksize <= 7) {
        static const float t1[] = {1.f};
        static const float t3[] = {0.5f, 0.25f};
        static const float t5[] = {0.375f, 0.25f, 0.0625f};
        static const float t7[] = {0.28125f, 0.21875f, 0.109375f, 0.03125f};
        const float *t = ksize == 1 ? t1 : ksize == 3 ? t3 : ksize == 5 ? t5 : t7;
        for (int i = 0; i <= r; i++) gk[i] = t[i];
        return;
    }
    const double sx = sigma > 0 ? sigma : ((ksize - 1) * 0.5 - 1) * 0.3 + 0.8;
    const double scale2 = -0.5 / (sx * sx);
    std::vector<double> tmp(ksize);
    double sum = 0;
    for (int i = 0; i < ksize; i++) {
        double x = i - (ksize - 1) * 0.5;
        tmp[i] = exp(scale2 * x * x);
        sum += tmp[i];
    }
    sum = 1. / sum;
    for (int i = 0; i <= r; i++) gk[i] = (float)(tmp[r + i] * sum);
}

// 6x6 inverse by Gauss-Jordan with partial pivoting (double)
static void invert6(double a[6][6], double inv[6][6])
{
    double m[6][12];
    for (int i = 0; i < 6; i++)
        for (int j = 0; j < 6; j++) {
            m[i][j] = a[i][j];
            m[i][j + 6] = (i == j);
        }
    for (int c = 0; c < 6; c++) {
        int p = c;
        for (int i = c + 1; i < 6; i++)
            if (fabs(m[i][c]) > fabs(m[p][c])) p = i;
        if (p != c)
            for (int j = 0; j < 12; j++) std::swap(m[c][j], m[p][j]);
        const double d = 1.0 / m[c][c];
        for (int j = 0; j < 12; j++) m[c][j] *= d;
        for (int i = 0; i < 6; i++)
            if (i != c) {
                const double f = m[i][c];
                if (f != 0)
                    for (int j = 0; j < 12; j++) m[i][j] -= f * m[c][j];
            }
    }
    for (int i = 0; i < 6; i++)
        for (int j = 0; j < 6; j++) inv[i][j] = m[i][j + 6];
}

// FarnebackPrepareGaussian (A.4)
static void prepare_poly(int n, double sigma, PolyConst *pc)
{
    if (sigma < FLT_EPSILON) sigma = n * 0.3;
    std::vector<float> g(2 * n + 1), xg(2 * n + 1), xxg(2 * n + 1);
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
    double G[6][6] = {}, inv[6][6];
    for (int y = -n; y <= n; y++)
        for (int x = -n; x <= n; x++) {
            const float gg = g[y + n] * g[x + n];
            G[0][0] += gg;
            G[1][1] += gg * x * x;
            G[3][3] += gg * x * x * x * x;
            G[5][5] += gg * x * x * y * y;
        }
    G[2][2] = G[0][3] = G[0][4] = G[3][0] = G[4][0] = G[1][1];
    G[4][4] = G[3][3];
    G[3][4] = G[4][3] = G[5][5];
    invert6(G, inv);
    memset(pc, 0, sizeof(*pc));
    pc->n = n;
    for (int k = 0; k <= n; k++) {
        pc->g[k] = g[n + k];
        pc->xg[k] = xg[n + k];
        pc->xxg[k] = xxg[n + k];
    }
    for (int k = 0; k <= n && k < 8; k++) {
        pc->g2[k] = make_float2(pc->g[k], pc->g[k]);
        pc->xg2[k] = make_float2(pc->xg[k], pc->xg[k]);
        pc->xxg2[k] = make_float2(pc->xxg[k], pc->xxg[k]);
    }
    pc->ig11 = (float)inv[1][1];
    pc->ig03 = (float)inv[0][3];
    pc->ig33 = (float)inv[3][3];
    pc->ig55 = (float)inv[5][5];
}

static inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int validate_params(const nsof_fb_params *p)
{
    if (!p) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: params is NULL");
    if (!(p->pyr_scale < 1.) || !(p->pyr_scale > 0.))
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: pyr_scale must be in (0, 1) (cv2: assertion pyrScale_ < 1)");
    if (p->levels < 0 || p->winsize < 1 || p->iterations < 0)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: levels >= 0, winsize >= 1, iterations >= 0 required");
    if (p->winsize < 2)
        NSOF_FAIL(NSOF_ERR_UNSUPPORTED,
                  "farneback: winsize 1 is not implemented (OpenCV's running-sum initialisation adds row 0 / "
                  "column 0 to every 1x1 window there; see oracle/farneback_oracle.c A.7)");
    if (p->poly_n < 1 || p->poly_n > NSOF_MAX_POLY_N)
        NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "farneback: poly_n must be in 1..%d", NSOF_MAX_POLY_N);
    if (p->flags & ~(NSOF_FB_GAUSSIAN | NSOF_FB_USE_INITIAL_FLOW | NSOF_FB_SRC_F32))
        NSOF_FAIL(NSOF_ERR_UNSUPPORTED,
                  "farneback: flags=%d; OPTFLOW_USE_INITIAL_FLOW (4) and OPTFLOW_FARNEBACK_GAUSSIAN (256) are the flags "
                  "cv2 defines", p->flags);
    if ((p->flags & NSOF_FB_GAUSSIAN) && p->winsize / 2 > FB_MAX_WIN_M)
        NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "farneback: Gaussian window radius %d > %d", p->winsize / 2, FB_MAX_WIN_M);
    return NSOF_OK;
}

// opt-in limit of dynamic shared memory per CTA: 227 KB minus the static part (the job descriptor, NSOF_JOB_TO_SMEM)
#define FB_MAX_DYN_SMEM (227 * 1024 - 512)

// Tile of k_fb_blur_solve for a window radius m: largest candidate that keeps >= 3 CTAs per SM,
// else anything that fits the 227 KB opt-in limit.
static bool pick_blur_tile(int m, int *TW, int *TH, size_t *smem)
{
    static const int cand[][2] = {{64, 32}, {64, 16}, {32, 32}, {32, 16}, {16, 16}, {16, 8}, {8, 8}, {4, 4}};
    for (int pass = 0; pass < 2; pass++) {
        const size_t limit = pass == 0 ? 75 * 1024 : FB_MAX_DYN_SMEM;
        for (auto &c : cand) {
            const size_t pw = round_up(c[0] + 2 * m, 4);
            const size_t b = 5 * (size_t)(c[1] + 2 * m) * pw * sizeof(float);
            if (b <= limit) {
                *TW = c[0];
                *TH = c[1];
                *smem = b;
                return true;
            }
        }
    }
    return false;
}

extern "C" double nsof_fb_algorithmic_bytes(int h, int w, const nsof_fb_params *p)
{
    // DESIGN.md byte model == SURVEY.md 8(d): per level k (N_k pixels, N_0 source pixels, T iterations), with I
    // never in HBM:   2*N_0 (uint8 sources) + 40*N_k (write R0,R1) + 60*N_k + 8*N_{k+1} (first UpdateMatrices)
    //                 + 80*N_k*(T-1) (read M,R0,R1; write M) + 28*N_k (read M, write flow)
    if (!p || h < 1 || w < 1) return 0.;
    int hk[64], wk[64];
    const int S = nsof_fb_level_geometry(h, w, p->pyr_scale, p->levels, hk, wk, nullptr, nullptr);
    const double N0 = (double)h * w;
    double bytes = 0;
    for (int k = 0; k < S; k++) {
        const double Nk = (double)hk[k] * wk[k];
        const double Nc = k + 1 < S ? (double)hk[k + 1] * wk[k + 1] : 0.;
        bytes += 2 * N0 + 40 * Nk;
        if (p->iterations > 0) bytes += 60 * Nk + 8 * Nc + 80 * Nk * (p->iterations - 1) + 28 * Nk;
        else bytes += 8 * Nc + 8 * Nk;
    }
    return bytes;
}

// ------------------------------------------------------------------------------ batch driver
static int configure_smem(nsof_handle *h)
{
    if (h->smem_configured) return NSOF_OK;
    NSOF_CUDA_TRY(cudaFuncSetAttribute(k_fb_blur_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_MAX_DYN_SMEM));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(k_fb_blur_solve_gauss, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_MAX_DYN_SMEM));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_BLUR_SOLVE_M7, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FB_BS7_SMEM));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_BLUR_SOLVE_M7S, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FB_BS7S_SMEM));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_BLUR_SOLVE_M1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FB_BSM_SMEM(1)));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_BLUR_SOLVE_M2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FB_BSM_SMEM(2)));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_PYR_POLYEXP_M0, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fb_pp_smem(0)));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_PYR_POLYEXP_M1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fb_pp_smem(1)));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_PYR_POLYEXP_M0_N1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fb_pp_smem(0, 1)));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_PYR_POLYEXP_M0_N10, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fb_pp_smem(0, 10)));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_POLYEXP_T_N1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fb_pp_smem(2, 1)));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(FB_POLYEXP_T_N10, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fb_pp_smem(2, 10)));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(k_fb_pyr_sep, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_MAX_DYN_SMEM));
    NSOF_CUDA_TRY(cudaFuncSetAttribute(k_fb_polyexp, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_MAX_DYN_SMEM));
    h->smem_configured = true;
    return NSOF_OK;
}

static int plan_item(const nsof_crop_desc &d, const nsof_fb_params *p, FbItemPlan *pl, int index)
{
    if (!d.prev || !d.next || !d.flow) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: item %d has a NULL pointer", index);
    if (d.h < 1 || d.w < 1)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: item %d is empty (%dx%d) (cv2: assertion !_src.empty())", index, d.h, d.w);
    const ptrdiff_t esz = (p->flags & NSOF_FB_SRC_F32) ? 4 : 1;
    if (d.prev_pitch < d.w * esz || d.next_pitch < d.w * esz || d.flow_pitch < (ptrdiff_t)d.w * 8 || d.flow_pitch % 8 ||
        (esz == 4 && ((d.prev_pitch | d.next_pitch) % 4 || ((uintptr_t)d.prev | (uintptr_t)d.next) % 4)))
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: item %d has a bad pitch", index);
    if ((uintptr_t)d.flow % 8) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: item %d flow pointer is not 8-byte aligned", index);
    pl->d = d;
    pl->scales = nsof_fb_level_geometry(d.h, d.w, p->pyr_scale, p->levels, pl->hk, pl->wk, nullptr, nullptr);
    for (int k = 0; k < pl->scales; k++) {
        if (pl->hk[k] < 1 || pl->wk[k] < 1) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: item %d level %d is empty", index, k);
        pl->pitch[k] = (int)round_up(pl->wk[k], 32);
    }
    return NSOF_OK;
}

extern "C" int nsof_reserve(nsof_handle *h, size_t total_crop_pixels, int max_items)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "nsof_reserve: NULL handle");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    // what fb_prepare's plan asks for, per level-0 pixel: (I 2 + R 10) on every level (x 4/3 for pyr_scale 0.5),
    // Ma 5 + Mb 5, coarse flow 2 x 2 / 4  ->  27 floats (28 with slack); rows are padded to 32 floats: <= 31 floats on
    // each of ~4/3 h rows per item (budgeted for crops up to ~1150 rows).  A hint only: fb_prepare sizes every batch
    // exactly and grows the arena when a batch needs more (other pyr_scale, taller crops).
    const size_t floats = total_crop_pixels * 28 + (size_t)max_items * 31 * 1536 * 28;
    int rc = h->arena.ensure(floats * sizeof(float));
    if (rc) return rc;
    return h->tables.ensure((size_t)max_items * 8 * sizeof(FbJob) + 65536);
}

// Everything one batch needs between planning and the last launch.  nsof_farneback_batch fills it and runs all
// stages; the spatially tiled entry points (nsof_fb_tiled_*) fill it once and run one stage per call, with the
// caller exchanging halo rows in between.
struct FbTiling {
    int rank, world;   // this rank owns rows [rank * h_k / world, (rank + 1) * h_k / world) of every level k
    int halo_r1;       // R1 rows received from each neighbour per level (reach of the UpdateMatrices gather)
    int *guard;        // device flag, see FbJob::guard
};

// Tile lists of a level: which items they hold and the tile geometry
//   PP0 / PP1  fused pyramid + PolyExp, items with fast == 1 / 2       64 x 32 over rows [y_lo, y_hi)
//   PYR        generic pyramid, items with fast == 0                    32 x 16 (separable) or 32 x 8 (direct) over I rows
//   PE         generic PolyExp, items with fast == 0                    32 x 32
//   UM         first UpdateMatrices / flow initialisation, all items    128 x 8
//   BS         blur + solve, all items                                  BTW x BTH, or 32 x 16 on latency-bound levels
enum { FB_L_PP0 = 0, FB_L_PP1, FB_L_PYR, FB_L_PE, FB_L_UM, FB_L_BS, FB_NLISTS };

struct FbBatchCtx {
    bool valid = false;
    int n = 0, max_scales = 0, stop_level = 0, m = 0;
    nsof_fb_params params = {};
    bool tiled = false;
    FbTiling tiling = {0, 1, 0, nullptr};
    int ksize[64];
    // tiled batches keep 4 job tables per level: 0 the rank's band, 1 / 2 its top / bottom `edge` rows (what the row
    // neighbours wait for), 3 the rows in between; every per-level vector below except gk_off is indexed
    // [variant * max_scales + level]
    int n_variants = 1, edge = 0;
    std::vector<size_t> gk_off, job_off;
    std::vector<int> n_active, path_n, ps_sw, ps_sh;
    std::vector<long long> sum_tiles_bs;
    // flat tile lists (one CTA per entry), per [kk * FB_NLISTS + list]: first entry and entry count in h->tile_buf
    std::vector<unsigned> list_off, list_n;
    std::vector<char> bs_small, pyr_sep;  // per kk: geometry of FB_L_BS (32x16 instead of 64x32) / FB_L_PYR (separable)
    const uint2 *dtiles = nullptr;
    std::vector<int> n_init;  // per level: items whose coarsest level it is and that take an initial flow (k > 0)
    std::vector<double> sumN, sumNc, path_N, path_N0;
    int BTW = 0, BTH = 0;
    size_t bs_smem = 0, pe_smem = 0;
    bool bs_m7 = false;
    bool pe_tuned = false;   // PolyExp of the generic-pyramid items: k_fb_pyr_polyexp_t<2, poly_n> (I from HBM) instead of k_fb_polyexp
    int bs_msmall = 0;   // 1 / 2: winsize 3 / 4 .. 5 with the box window on the tuned kernel's small-radius instantiations
    PolyConst pc;
    float flow_mul = 0.f, bscale = 0.f;
    FbGaussWin gwin;  // S2-e': weights of the Gaussian window (flags & NSOF_FB_GAUSSIAN)
    unsigned char *ds = nullptr;  // device tables
};

static void fb_ctx_free(nsof_handle *h)
{
    delete h->ctx;
    h->ctx = nullptr;
}

// rows [lo, hi) of a level of height hk owned by `rank` of `world`
static inline void band_rows(int hk, int rank, int world, int *lo, int *hi)
{
    *lo = (int)((long long)rank * hk / world);
    *hi = (int)((long long)(rank + 1) * hk / world);
}

static int fb_prepare(nsof_handle *h, const nsof_crop_desc *items, int n, const nsof_fb_params *params,
                      const FbTiling *tiling, cudaStream_t stream)
{
    if (!h->ctx) h->ctx = new FbBatchCtx();
    FbBatchCtx &X = *h->ctx;
    X.valid = false;
    int rc = configure_smem(h);
    if (rc) return rc;
    X.n = n;
    X.params = *params;
    X.tiled = tiling != nullptr;
    if (tiling) X.tiling = *tiling;

    // ---- plan: geometry + arena offsets
    std::vector<FbItemPlan> &plan = *h->last_plan;
    plan.resize(n);
    size_t arena_floats = 0;
    int max_scales = 0;
    for (int i = 0; i < n; i++) {
        rc = plan_item(items[i], params, &plan[i], i);
        if (rc) return rc;
        FbItemPlan &pl = plan[i];
        const size_t n0 = (size_t)pl.pitch[0] * pl.hk[0];
        const size_t n1 = pl.scales > 1 ? (size_t)pl.pitch[1] * pl.hk[1] : 0;
        for (int k = 0; k < pl.scales; k++) {
            const size_t nk = (size_t)pl.pitch[k] * pl.hk[k];
            pl.off_I[k] = arena_floats;  arena_floats += 2 * nk;
            pl.off_R[k] = arena_floats;  arena_floats += 10 * nk;
        }
        pl.off_Ma = arena_floats; arena_floats += 5 * n0;
        pl.off_Mb = arena_floats; arena_floats += 5 * n0;
        pl.off_fA = arena_floats; arena_floats += 2 * n1;
        pl.off_fB = arena_floats; arena_floats += 2 * n1;
        arena_floats = round_up(arena_floats, 64);
        max_scales = std::max(max_scales, pl.scales);
    }
    // Shared frames: item j's `prev` crop that IS item i's `next` crop (same address, size and pitch: consecutive pairs
    // of a clip under an unchanged ROI rectangle, or whole frames) has the same pyramid and polynomial expansion at
    // every level -- R[0] of j is R[1] of i, computed once.  (OpenCV expands both images of every call.)
    {
        std::unordered_map<const uint8_t *, int> by_next;
        if (!h->dbg_no_dedup)
            for (int i = 0; i < n; i++) by_next.emplace(plan[i].d.next, i);
        for (int j = 0; j < n; j++) {
            plan[j].r0_from = -1;
            auto it = by_next.find(plan[j].d.prev);
            if (it == by_next.end() || it->second == j) continue;
            const nsof_crop_desc &a = plan[it->second].d, &b = plan[j].d;
            if (a.h == b.h && a.w == b.w && a.next_pitch == b.prev_pitch) plan[j].r0_from = it->second;
        }
    }
    rc = h->arena.ensure(arena_floats * sizeof(float));
    if (rc) return rc;
    float *base = (float *)h->arena.p;
    X.max_scales = max_scales;

    double sigma[64];
    {
        // ksize/sigma depend on the level index only; take them from a geometry with all levels present
        double scale = 1;
        for (int k = 0; k < max_scales; k++) {
            sigma[k] = (1. / scale - 1) * 0.5;
            X.ksize[k] = std::max(cv_round(sigma[k] * 5) | 1, 3);
            scale *= params->pyr_scale;
        }
    }
    X.stop_level = (h->dbg_level >= 0 && h->dbg_level < max_scales) ? h->dbg_level : 0;
    h->last_stop_level = X.stop_level;
    h->last_params = *params;

    // ---- job tables for every level + Gaussian taps, one upload
    size_t gk_floats = 0;
    const int NV = tiling ? 4 : 1, LV = NV * max_scales;
    X.n_variants = NV;
    X.edge = tiling ? tiling->halo_r1 : 0;
    X.gk_off.assign(max_scales, 0);
    X.job_off.assign(LV, 0);
    X.n_active.assign(LV, 0);
    for (int k = 0; k < max_scales; k++) {
        if (X.ksize[k] / 2 > 4096) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "farneback: pyramid blur ksize %d too large", X.ksize[k]);
        X.gk_off[k] = gk_floats;
        gk_floats += round_up(X.ksize[k] / 2 + 1, 4);
        for (int i = 0; i < n; i++) X.n_active[k] += plan[i].scales > k;
        for (int v = 1; v < NV; v++) X.n_active[v * max_scales + k] = X.n_active[k];
    }
    size_t table_bytes = round_up(gk_floats * sizeof(float), 16);
    for (int kk = 0; kk < LV; kk++) {
        X.job_off[kk] = table_bytes;
        table_bytes += round_up((size_t)X.n_active[kk] * sizeof(FbJob), 16);
    }
    // runs of the flat tile lists: at most one per (level, list, item); only the used part is uploaded
    const size_t runs_off = table_bytes;
    {
        size_t max_runs = 0;
        for (int kk = 0; kk < LV; kk++) max_runs += (size_t)X.n_active[kk] * (FB_NLISTS + 4);  // pyramid lists: 2 images
        table_bytes += round_up(max_runs * sizeof(FbTileRun), 16);
    }
    NsofStageSlot *slot;
    rc = nsof_stage_acquire(h, table_bytes, &slot);
    if (rc) return rc;
    rc = h->tables.ensure(table_bytes + 16);
    if (rc) return rc;
    unsigned char *hs = (unsigned char *)slot->host;
    X.ds = (unsigned char *)h->tables.p;
    X.sumN.assign(LV, 0.);
    X.sumNc.assign(LV, 0.);
    // per level and pyramid path (0 generic, 1 fused level-0 shape, 2 fused exact-2x shape): items, pixels, source pixels
    X.path_n.assign(3 * LV, 0);
    X.path_N.assign(3 * LV, 0.);
    X.path_N0.assign(3 * LV, 0.);
    for (auto *v : {&X.ps_sw, &X.ps_sh}) v->assign(LV, 0);
    X.sum_tiles_bs.assign(LV, 0);
    X.n_init.assign(LV, 0);
    // fused pyramid + PolyExp: poly_n 5 on the level-0 and exact-2x shapes, poly_n 1 / 10 (the other shipped sets) on level 0
    const bool may_fuse = !h->dbg_no_fuse && (params->poly_n == FB_PP_N || params->poly_n == 1 || params->poly_n == 10);
    const int m = params->winsize / 2;
    X.m = m;
    X.BTW = X.BTH = 0;
    X.bs_smem = 0;
    // winsize 14/15 with the box window: the tuned kernel (every BASELINE.json config)
    X.bs_m7 = (m == 7) && !(params->flags & NSOF_FB_GAUSSIAN);
    X.bs_msmall = ((m == 1 || m == 2) && !(params->flags & NSOF_FB_GAUSSIAN) && !h->dbg_generic_blur) ? m : 0;
    if (X.bs_m7) { X.BTW = FB_BS7_TW; X.BTH = FB_BS7_TH; X.bs_smem = FB_BS7_SMEM; }
    else if (X.bs_msmall) { X.BTW = FB_BS7_TW; X.BTH = FB_BS7_TH; X.bs_smem = FB_BSM_SMEM(m); }
    else if (params->iterations > 0 && !pick_blur_tile(m, &X.BTW, &X.BTH, &X.bs_smem))
        NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "farneback: winsize %d needs more than 227 KB of shared memory", params->winsize);
    // poly_n 1 / 10 only: on the grasp set (poly_n 5) the tuned moment passes at the coarse levels, whose multiply-adds the
    // compiler contracts differently, moved the flow of ill-conditioned crops away from the C oracle (99.9 % quantile of the
    // EPE on a smooth synthetic background 1e-3 -> 4.6e-3 px, mean 8e-6 -> 1.5e-5) for 1 % of the step: not taken
    X.pe_tuned = may_fuse && params->poly_n != FB_PP_N;
    const int pn = params->poly_n;
    for (int kk = 0; kk < LV; kk++) {
        const int k = kk % max_scales, variant = kk / max_scales;
        if (variant == 0) gaussian_half_kernel(X.ksize[k], sigma[k], (float *)hs + X.gk_off[k]);
        FbJob *jobs = (FbJob *)(hs + X.job_off[kk]);
        int a = 0;
        for (int i = 0; i < n; i++) {
            const FbItemPlan &pl = plan[i];
            if (pl.scales <= k) continue;
            FbJob J;
            memset(&J, 0, sizeof(J));
            J.h = pl.hk[k]; J.w = pl.wk[k]; J.pitch = pl.pitch[k];
            const bool coarsest = (k == pl.scales - 1);
            if (!coarsest) {
                J.ch = pl.hk[k + 1]; J.cw = pl.wk[k + 1]; J.cpitch = pl.pitch[k + 1];
                J.fsx = (double)J.cw / J.w; J.fsy = (double)J.ch / J.h;
            }
            J.sh = pl.d.h; J.sw = pl.d.w;
            J.spitch0 = (int)pl.d.prev_pitch; J.spitch1 = (int)pl.d.next_pitch;
            J.src0 = pl.d.prev; J.src1 = pl.d.next;
            if (J.h == J.sh && J.w == J.sw) J.resize_mode = 0;
            else if (J.sh == 2 * J.h && J.sw == 2 * J.w) J.resize_mode = 1;  // cv::resize takes INTER_AREA at exact 2x
            else J.resize_mode = 2;
            J.sx = (double)J.sw / J.w; J.sy = (double)J.sh / J.h;
            J.fast = (may_fuse && X.ksize[k] == 3 && J.resize_mode < 2) ? J.resize_mode + 1 : 0;
            if (J.fast == 2 && params->poly_n != FB_PP_N) J.fast = 0;
            if (params->flags & NSOF_FB_SRC_F32) J.fast = 0;  // float32 frames: the direct pyramid kernel (typed loads)
            // row window
            J.y_lo = 0; J.y_hi = J.h; J.gy_lo = 0; J.gy_hi = J.h; J.guard = nullptr;
            if (tiling) {
                band_rows(J.h, tiling->rank, tiling->world, &J.y_lo, &J.y_hi);
                J.gy_lo = std::max(0, J.y_lo - tiling->halo_r1);
                J.gy_hi = std::min(J.h, J.y_hi + tiling->halo_r1);
                J.guard = tiling->guard;
                // sub-windows of the band (empty when the band is shorter than two edges)
                const int e = std::min(X.edge, (J.y_hi - J.y_lo) / 2);
                if (variant == 1) J.y_hi = J.y_lo + e;
                else if (variant == 2) J.y_lo = J.y_hi - e;
                else if (variant == 3) { J.y_lo += e; J.y_hi -= e; }
            }
            J.iy_lo = std::max(0, J.y_lo - pn); J.iy_hi = std::min(J.h, J.y_hi + pn);
            if (J.y_hi <= J.y_lo) J.iy_hi = J.iy_lo;
            const int rows = J.y_hi - J.y_lo, irows = J.iy_hi - J.iy_lo;
            const double frac = (double)rows / J.h;  // share of the level this rank computes (1 when untiled)
            // (the byte accounting of the pyramid launches counts the images actually expanded: 1 or 2 per item)
            const double n_img = pl.r0_from >= 0 ? 0.5 : 1.0;
            X.path_n[3 * kk + J.fast] += 1;
            X.path_N[3 * kk + J.fast] += (double)rows * J.w * n_img;
            X.path_N0[3 * kk + J.fast] += (double)J.sh * J.sw * frac * n_img;
            J.I = base + pl.off_I[k]; J.Ma = base + pl.off_Ma; J.Mb = base + pl.off_Mb;
            J.R[0] = pl.r0_from >= 0 ? base + plan[pl.r0_from].off_R[k] + 5 * (size_t)J.pitch * J.h : base + pl.off_R[k];
            J.R[1] = base + pl.off_R[k] + 5 * (size_t)J.pitch * J.h;
            J.share0 = pl.r0_from >= 0;
            // coarse flow ping-pong: level k writes A when k is odd, B when even; level 0 writes the caller's buffer
            float *fl_this = (k & 1) ? base + pl.off_fA : base + pl.off_fB;
            float *fl_coarse = (k & 1) ? base + pl.off_fB : base + pl.off_fA;
            if (k == 0) { J.flow = (float2 *)pl.d.flow; J.fpitch = (int)(pl.d.flow_pitch / 8); }
            else { J.flow = (float2 *)fl_this; J.fpitch = J.pitch; }
            J.cflow = (const float2 *)fl_coarse;
            if (coarsest && (params->flags & NSOF_FB_USE_INITIAL_FLOW)) {
                double sc = 1;  // pyr_scale^k by repeated multiplication, as the level geometry
                for (int i2 = 0; i2 < k; i2++) sc *= params->pyr_scale;
                J.init = k == 0 ? 2 : 1;
                J.init_mul = (float)sc;
                J.flow0 = (const float2 *)pl.d.flow;
                J.f0pitch = (int)(pl.d.flow_pitch / 8);
                J.cflow = J.flow;  // the level's own buffer, read as it is by the first UpdateMatrices
                J.cpitch = J.fpitch;
                if (k > 0) X.n_init[kk] += 1;
            }
            jobs[a++] = J;
            X.sumN[kk] += (double)rows * J.w;
            if (!coarsest) X.sumNc[kk] += (double)J.ch * J.cw * frac;
            if (!J.fast) {
                X.ps_sw[kk] = std::max(X.ps_sw[kk], (int)ceil(FB_PS_TW * J.sx) + 2 * (X.ksize[k] / 2) + 4);
                X.ps_sh[kk] = std::max(X.ps_sh[kk], (int)ceil(FB_PS_TH * J.sy) + 2 * (X.ksize[k] / 2) + 4);
            }
            if (X.bs_m7 || X.bs_msmall) X.sum_tiles_bs[kk] += (long long)((J.w + X.BTW - 1) / X.BTW) * ((rows + X.BTH - 1) / X.BTH);
            (void)irows;
        }
    }
    // ---- flat tile lists: per (level, list) the runs (item, tile count) of the items it holds, in item order
    X.list_off.assign((size_t)LV * FB_NLISTS, 0);
    X.list_n.assign((size_t)LV * FB_NLISTS, 0);
    X.bs_small.assign(LV, 0);
    X.pyr_sep.assign(LV, 0);
    FbTileRun *runs = (FbTileRun *)(hs + runs_off);
    size_t n_runs = 0;
    unsigned n_tiles = 0;
    for (int kk = 0; kk < LV; kk++) {
        const int k = kk % max_scales;
        // fewer 64x32 tiles than two waves of the 148 SMs (3 CTAs each): the level is latency-bound, take the 32x16
        // tile (4x the CTAs, each a quarter of the serial work)
        X.bs_small[kk] = (X.bs_m7 || X.bs_msmall) && X.sum_tiles_bs[kk] < 2 * 148 * 3 && !h->dbg_big_tiles;
        // two CTAs per SM; larger windows (extreme resize ratios) take the direct pyramid kernel
        X.pyr_sep[kk] = fb_ps_smem(X.ksize[k] / 2, X.ps_sw[kk], X.ps_sh[kk]) <= 100 * 1024 && !(params->flags & NSOF_FB_SRC_F32);
        const FbJob *jobs = (const FbJob *)(hs + X.job_off[kk]);
        for (int l = 0; l < FB_NLISTS; l++) {
            X.list_off[(size_t)kk * FB_NLISTS + l] = n_tiles;
            for (int a = 0; a < X.n_active[kk]; a++) {
                const FbJob &J = jobs[a];
                const int rows = J.y_hi - J.y_lo, irows = J.iy_hi - J.iy_lo;
                int tw = 0, th = 0, nr = rows;
                switch (l) {
                case FB_L_PP0: if (J.fast == 1) { tw = FB_PP_TW; th = FB_PP_TH; } break;
                case FB_L_PP1: if (J.fast == 2) { tw = FB_PP_TW; th = FB_PP_TH; } break;
                case FB_L_PYR:
                    if (!J.fast) { tw = X.pyr_sep[kk] ? FB_PS_TW : FB_PYR_TW; th = X.pyr_sep[kk] ? FB_PS_TH : FB_PYR_TH; nr = irows; }
                    break;
                case FB_L_PE: if (!J.fast) { tw = X.pe_tuned ? FB_PP_TW : FB_PE_TW; th = X.pe_tuned ? FB_PP_TH : FB_PE_TH; } break;
                case FB_L_UM: tw = FB_UM_TW; th = FB_UM_TH; break;
                case FB_L_BS:
                    if (X.BTW) { tw = X.bs_small[kk] ? FB_BS7S_TW : X.BTW; th = X.bs_small[kk] ? FB_BS7S_TH : X.BTH; }
                    break;
                }
                if (!tw || nr <= 0) continue;
                const unsigned cnt = (unsigned)(((J.w + tw - 1) / tw) * ((nr + th - 1) / th));
                if (!cnt) continue;
                const bool per_image = l <= FB_L_PE;  // pyramid / PolyExp lists: one run per image to expand
                for (int img = per_image && J.share0 ? 1 : 0; img < (per_image ? 2 : 1); img++) {
                    FbTileRun r = {n_tiles, (unsigned)a | ((unsigned)img << FB_TL_IMG_SHIFT), cnt, 0u};
                    runs[n_runs++] = r;
                    n_tiles += cnt;
                }
            }
            X.list_n[(size_t)kk * FB_NLISTS + l] = n_tiles - X.list_off[(size_t)kk * FB_NLISTS + l];
        }
    }
    rc = h->tile_buf.ensure((size_t)std::max(n_tiles, 1u) * sizeof(uint2));
    if (rc) return rc;
    X.dtiles = (const uint2 *)h->tile_buf.p;
    { int rc_up = nsof_table_upload(h, X.ds, slot, runs_off + n_runs * sizeof(FbTileRun), stream); if (rc_up) return rc_up; }
    if (n_runs) {
        NSOF_KLAUNCH(h, NSOF_K_TILE_LIST, (double)n_runs * sizeof(FbTileRun) + (double)n_tiles * sizeof(uint2), k_fb_build_tiles,
                     (int)n_runs, 1, 1, 128, 0, stream, (const FbTileRun *)(X.ds + runs_off), (uint2 *)h->tile_buf.p);
        NSOF_CUDA_TRY(cudaGetLastError());
    }

    prepare_poly(params->poly_n, params->poly_sigma, &X.pc);
    X.pe_smem = ((size_t)(FB_PE_TH + 2 * pn) * (FB_PE_TW + 2 * pn) + 3 * (size_t)FB_PE_TH * (FB_PE_TW + 2 * pn)) * sizeof(float);
    X.flow_mul = (float)(1. / params->pyr_scale);
    X.bscale = (float)(1. / ((double)params->winsize * params->winsize));
    memset(&X.gwin, 0, sizeof(X.gwin));
    if (params->flags & NSOF_FB_GAUSSIAN) {
        // FarnebackUpdateFlow_GaussianBlur: sigma = 0.3 m; k[0] = 1, k[i] = f32(exp(-i^2 / 2 sigma^2)); normalised in double
        const double gs = m * 0.3;
        double sum = 1.;
        X.gwin.on = 1;
        X.gwin.k[0] = 1.f;
        for (int i = 1; i <= m; i++) {
            const float t = (float)exp(-i * i / (2 * gs * gs));
            X.gwin.k[i] = t;
            sum += t * 2;
        }
        sum = 1. / sum;
        for (int i = 0; i <= m; i++) X.gwin.k[i] = (float)(X.gwin.k[i] * sum);
        X.bscale = 1.f;  // the Gaussian sums are used as they are
    }
    X.valid = true;
    return NSOF_OK;
}

// S2-b + S2-c of level k on `stream` (independent of the flow: needs the sources and the tables only)
static int fb_launch_pyramid(nsof_handle *h, int k, cudaStream_t stream, int variant = 0)
{
    FbBatchCtx &X = *h->ctx;
    const int kk = variant * X.max_scales + k;  // job table / geometry of the band (0) or one of its sub-windows
    const FbJob *jobs = (const FbJob *)(X.ds + X.job_off[kk]);
    const float *gk = (const float *)X.ds + X.gk_off[k];
    const int na = X.n_active[kk];
    const PolyConst &pc = X.pc;
    h->prof_level = k;
    const auto ln = [&](int l) { return (int)X.list_n[(size_t)kk * FB_NLISTS + l]; };
    const auto lp = [&](int l) { return X.dtiles + X.list_off[(size_t)kk * FB_NLISTS + l]; };
    (void)na;
    if (ln(FB_L_PP0))
        NSOF_KLAUNCH(h, NSOF_K_PYR_POLYEXP, 2 * X.path_N0[3 * kk + 1] + 40 * X.path_N[3 * kk + 1],
                     (pc.n == 1 ? FB_PYR_POLYEXP_M0_N1 : pc.n == 10 ? FB_PYR_POLYEXP_M0_N10 : FB_PYR_POLYEXP_M0),
                     ln(FB_L_PP0), 1, 1, 256, fb_pp_smem(0, pc.n), stream, jobs, lp(FB_L_PP0), gk, pc);
    if (ln(FB_L_PP1))
        NSOF_KLAUNCH(h, NSOF_K_PYR_POLYEXP, 2 * X.path_N0[3 * kk + 2] + 40 * X.path_N[3 * kk + 2], FB_PYR_POLYEXP_M1,
                     ln(FB_L_PP1), 1, 1, 256, fb_pp_smem(1), stream, jobs, lp(FB_L_PP1), gk, pc);
    if (ln(FB_L_PYR) && ln(FB_L_PE)) {
        const int r = X.ksize[k] / 2;
        if (X.pyr_sep[kk])
            NSOF_KLAUNCH(h, NSOF_K_PYR, 2 * X.path_N0[3 * kk], k_fb_pyr_sep, ln(FB_L_PYR), 1, 1, 256,
                         fb_ps_smem(r, X.ps_sw[kk], X.ps_sh[kk]), stream, jobs, lp(FB_L_PYR), gk, r, X.ps_sw[kk], X.ps_sh[kk]);
        else if (X.params.flags & NSOF_FB_SRC_F32)
            NSOF_KLAUNCH(h, NSOF_K_PYR, 8 * X.path_N0[3 * kk], k_fb_pyr_f32, ln(FB_L_PYR), 1, 1, 256, 0, stream, jobs, lp(FB_L_PYR), gk, r);
        else
            NSOF_KLAUNCH(h, NSOF_K_PYR, 2 * X.path_N0[3 * kk], k_fb_pyr, ln(FB_L_PYR), 1, 1, 256, 0, stream, jobs, lp(FB_L_PYR), gk, r);
        if (X.pe_tuned)
            NSOF_KLAUNCH(h, NSOF_K_POLYEXP, 40 * X.path_N[3 * kk],
                         (pc.n == 1 ? FB_POLYEXP_T_N1 : FB_POLYEXP_T_N10), ln(FB_L_PE), 1, 1, 256,
                         fb_pp_smem(2, pc.n), stream, jobs, lp(FB_L_PE), gk, pc);
        else
            NSOF_KLAUNCH(h, NSOF_K_POLYEXP, 40 * X.path_N[3 * kk], k_fb_polyexp, ln(FB_L_PE), 1, 1, 256, X.pe_smem, stream,
                         jobs, lp(FB_L_PE), pc);
    }
    NSOF_CUDA_TRY(cudaGetLastError());
    h->prof_level = -1;
    return NSOF_OK;
}

// S2-f + first S2-d of level k (or the flow initialisation of a level with zero iterations)
static int fb_launch_um_first(nsof_handle *h, int k, cudaStream_t stream, int variant = 0)
{
    FbBatchCtx &X = *h->ctx;
    const int kk = variant * X.max_scales + k;
    const FbJob *jobs = (const FbJob *)(X.ds + X.job_off[kk]);
    const int na = X.n_active[kk];
    h->prof_level = k;
    const int n_um = (int)X.list_n[(size_t)kk * FB_NLISTS + FB_L_UM];
    const uint2 *tl_um = X.dtiles + X.list_off[(size_t)kk * FB_NLISTS + FB_L_UM];
    (void)na;
    if (!n_um) { h->prof_level = -1; return NSOF_OK; }  // empty sub-window
    if (X.n_init[kk])  // OPTFLOW_USE_INITIAL_FLOW: items whose coarsest level this is
        NSOF_KLAUNCH(h, NSOF_K_FLOW_INIT, 8 * X.sumN[kk], k_fb_area_init, n_um, 1, 1, 256, 0, stream, jobs, tl_um);
    if (X.params.iterations == 0)
        NSOF_KLAUNCH(h, NSOF_K_FLOW_INIT, 8 * X.sumNc[kk] + 8 * X.sumN[kk], k_fb_flow_init, n_um, 1, 1, 256, 0, stream,
                     jobs, tl_um, X.flow_mul);
    else if (!h->dbg_um_round1) {  // R1 taps from a shared-memory window wherever the tile's initial flow allows it
        unsigned *um_stats = h->dbg_r1_stats ? (unsigned *)h->r1_stats_buf.p + 2 : nullptr;
        if (X.params.flags & NSOF_FB_USE_INITIAL_FLOW)
            NSOF_KLAUNCH(h, NSOF_K_UM_FIRST, 60 * X.sumN[kk] + 8 * X.sumNc[kk], k_fb_um_first_win_init, n_um, 1, 1, 256,
                         FB_UMW_SMEM, stream, jobs, tl_um, X.flow_mul, 0, um_stats);
        else
            NSOF_KLAUNCH(h, NSOF_K_UM_FIRST, 60 * X.sumN[kk] + 8 * X.sumNc[kk], k_fb_um_first_win, n_um, 1, 1, 256,
                         FB_UMW_SMEM, stream, jobs, tl_um, X.flow_mul, 0, um_stats);
    } else if (X.params.flags & NSOF_FB_USE_INITIAL_FLOW)
        NSOF_KLAUNCH(h, NSOF_K_UM_FIRST, 60 * X.sumN[kk] + 8 * X.sumNc[kk], k_fb_um_first_init, n_um, 1, 1, 256, 0, stream,
                     jobs, tl_um, X.flow_mul);
    else
        NSOF_KLAUNCH(h, NSOF_K_UM_FIRST, 60 * X.sumN[kk] + 8 * X.sumNc[kk], k_fb_um_first, n_um, 1, 1, 256, 0, stream,
                     jobs, tl_um, X.flow_mul);
    NSOF_CUDA_TRY(cudaGetLastError());
    h->prof_level = -1;
    return NSOF_OK;
}

// iteration `it` of S2-e on level k: blur + solve, then UpdateMatrices (it < iters - 1) or the flow write
static int fb_launch_blur(nsof_handle *h, int k, int it, int iters, cudaStream_t stream, int variant = 0)
{
    FbBatchCtx &X = *h->ctx;
    const int kk = variant * X.max_scales + k;
    const FbJob *jobs = (const FbJob *)(X.ds + X.job_off[kk]);
    const int na = X.n_active[kk];
    const bool last = it == iters - 1;
    const double Nk = X.sumN[kk];
    h->prof_level = k;
    const int n_bs = (int)X.list_n[(size_t)kk * FB_NLISTS + FB_L_BS];
    const uint2 *tl_bs = X.dtiles + X.list_off[(size_t)kk * FB_NLISTS + FB_L_BS];
    (void)na;
    if (!n_bs) { h->prof_level = -1; return NSOF_OK; }  // empty sub-window
    // L2 prefetch of the tile's R rows in phase 1: only with the window off (with it, measured on the bench workload:
    // 8.80 -> 8.50 ms per step without the prefetches -- the window copy replaces the R1 gathers and the R0 loads run one
    // iteration ahead, so the 390 prefetch instructions per tile were issue cost only)
    const int pf_off = h->dbg_bs_no_prefetch >= 0 ? (h->dbg_bs_no_prefetch & 3) : (h->dbg_no_r1_window ? 0 : 3);
    const int bs_flags = (last ? 1 : 0) | (h->dbg_no_r1_window ? 2 : 0) | (pf_off << 2);
    unsigned *bs_stats = h->dbg_r1_stats ? (unsigned *)h->r1_stats_buf.p : nullptr;
    if (X.bs_m7 && X.bs_small[kk])  // latency-bound level: the 32x16 tile (see fb_prepare)
        NSOF_KLAUNCH(h, last ? NSOF_K_BLUR_FLOW : NSOF_K_BLUR_UM, (last ? 28 : 80) * Nk, FB_BLUR_SOLVE_M7S, n_bs, 1, 1,
                     FB_BS7S_NT, FB_BS7S_SMEM, stream, jobs, tl_bs, X.bscale, (it & 1) == 0, bs_flags, bs_stats);
    else if (X.bs_m7)
        NSOF_KLAUNCH(h, last ? NSOF_K_BLUR_FLOW : NSOF_K_BLUR_UM, (last ? 28 : 80) * Nk, FB_BLUR_SOLVE_M7, n_bs, 1, 1,
                     FB_BS7_NT, X.bs_smem, stream, jobs, tl_bs, X.bscale, (it & 1) == 0, bs_flags, bs_stats);
    else if (X.bs_msmall && X.bs_small[kk])
        NSOF_KLAUNCH(h, last ? NSOF_K_BLUR_FLOW : NSOF_K_BLUR_UM, (last ? 28 : 80) * Nk,
                     (X.bs_msmall == 1 ? FB_BLUR_SOLVE_M1S : FB_BLUR_SOLVE_M2S), n_bs, 1, 1, FB_BS7S_NT,
                     (X.bs_msmall == 1 ? FB_BSMS_SMEM(1) : FB_BSMS_SMEM(2)), stream, jobs, tl_bs, X.bscale, (it & 1) == 0, bs_flags,
                     bs_stats);
    else if (X.bs_msmall == 1)
        NSOF_KLAUNCH(h, last ? NSOF_K_BLUR_FLOW : NSOF_K_BLUR_UM, (last ? 28 : 80) * Nk, FB_BLUR_SOLVE_M1, n_bs, 1, 1,
                     FB_BS7_NT, X.bs_smem, stream, jobs, tl_bs, X.bscale, (it & 1) == 0, bs_flags, bs_stats);
    else if (X.bs_msmall == 2)
        NSOF_KLAUNCH(h, last ? NSOF_K_BLUR_FLOW : NSOF_K_BLUR_UM, (last ? 28 : 80) * Nk, FB_BLUR_SOLVE_M2, n_bs, 1, 1,
                     FB_BS7_NT, X.bs_smem, stream, jobs, tl_bs, X.bscale, (it & 1) == 0, bs_flags, bs_stats);
    else if (X.gwin.on)
        NSOF_KLAUNCH(h, last ? NSOF_K_BLUR_FLOW : NSOF_K_BLUR_UM, (last ? 28 : 80) * Nk, k_fb_blur_solve_gauss, n_bs, 1, 1,
                     256, X.bs_smem, stream, jobs, tl_bs, X.BTW, X.BTH, X.m, X.bscale, (it & 1) == 0, last, X.gwin);
    else
        NSOF_KLAUNCH(h, last ? NSOF_K_BLUR_FLOW : NSOF_K_BLUR_UM, (last ? 28 : 80) * Nk, k_fb_blur_solve, n_bs, 1, 1,
                     256, X.bs_smem, stream, jobs, tl_bs, X.BTW, X.BTH, X.m, X.bscale, (it & 1) == 0, last, X.gwin);
    NSOF_CUDA_TRY(cudaGetLastError());
    h->prof_level = -1;
    return NSOF_OK;
}

static int fb_check_batch_args(nsof_handle *h, const nsof_crop_desc *items, int n, const nsof_fb_params *params)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: NULL handle");
    if (n < 0 || (n > 0 && !items)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: bad item list");
    int rc = validate_params(params);
    if (rc) return rc;
    if (n > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "farneback: at most 65535 items per batch");
    return NSOF_OK;
}

extern "C" int nsof_farneback_batch(nsof_handle *h, const nsof_crop_desc *items, int n, const nsof_fb_params *params,
                                    void *cuda_stream)
{
    int rc = fb_check_batch_args(h, items, n, params);
    if (rc) return rc;
    if (n == 0) return NSOF_OK;
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    rc = fb_prepare(h, items, n, params, nullptr, stream);
    if (rc) return rc;
    FbBatchCtx &X = *h->ctx;
    const int max_scales = X.max_scales, stop_level = X.stop_level;

    // fork: every level but the coarsest builds its R0, R1 on an internal stream while the caller's stream iterates
    // on the coarser levels (small, latency-bound launches); joined per level by an event before its first
    // UpdateMatrices.  Serial when profiling (one event pair per launch must not overlap another).
    const int top = max_scales - 1;
    const bool fork = !h->prof_on && !h->no_fork && top > stop_level;
    if (fork) {
        for (auto &a : h->aux)
            if (!a) NSOF_CUDA_TRY(cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking));
        if (!h->ev_fork) NSOF_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
        while ((int)h->ev_level.size() < max_scales) {
            cudaEvent_t e = nullptr;
            NSOF_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            h->ev_level.push_back(e);
        }
        NSOF_CUDA_TRY(cudaEventRecord(h->ev_fork, stream));  // tables uploaded, previous batch's readers of R done
        rc = fb_launch_pyramid(h, top, stream);
        if (rc) return rc;
        for (int k = top - 1; k >= stop_level; k--) {
            cudaStream_t a = h->aux[k & 1];
            NSOF_CUDA_TRY(cudaStreamWaitEvent(a, h->ev_fork, 0));
            rc = fb_launch_pyramid(h, k, a);
            if (rc) return rc;
            NSOF_CUDA_TRY(cudaEventRecord(h->ev_level[k], a));
        }
    }

    // ---- level loop, coarse to fine (A.1)
    for (int k = max_scales - 1; k >= stop_level; k--) {
        int iters = params->iterations;
        if (k == stop_level && h->dbg_level >= 0 && h->dbg_iters >= 0) iters = std::min(iters, h->dbg_iters);
        if (!fork) {
            rc = fb_launch_pyramid(h, k, stream);
            if (rc) return rc;
        } else if (k < top) {
            NSOF_CUDA_TRY(cudaStreamWaitEvent(stream, h->ev_level[k], 0));
        }
        rc = fb_launch_um_first(h, k, stream);
        if (rc) return rc;
        if (params->iterations > 0)
            for (int it = 0; it < iters; it++) {
                rc = fb_launch_blur(h, k, it, iters, stream);
                if (rc) return rc;
            }
    }
    return NSOF_OK;
}

// ------------------------------------------------------------------------------ spatial tiling
// One frame pair (or a batch of pairs) split by rows over `world` GPUs.  Every rank plans the WHOLE problem (full-size
// workspace: addressing, borders and resize geometry stay those of the whole frame) but computes only its band of
// rows of every level; rows a stencil or gather reads beyond the band are halo rows the caller exchanges between
// the stage calls (NCCL send/recv between row neighbours, parallel.py).  Buffers are reported by nsof_fb_tiled_query.
extern "C" int nsof_fb_tiled_begin(nsof_handle *h, const nsof_crop_desc *items, int n, const nsof_fb_params *params, int rank,
                                   int world, int halo_r1, void *cuda_stream)
{
    int rc = fb_check_batch_args(h, items, n, params);
    if (rc) return rc;
    if (n < 1) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_begin: no items");
    if (world < 1 || rank < 0 || rank >= world || halo_r1 < 1) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_begin: bad rank/world/halo");
    if (params->iterations < 1) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "fb_tiled_begin: iterations >= 1 required");
    if (params->flags & NSOF_FB_USE_INITIAL_FLOW)
        NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "fb_tiled_begin: OPTFLOW_USE_INITIAL_FLOW is not available in the row-tiled mode");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    rc = h->guard_buf.ensure(sizeof(int));
    if (rc) return rc;
    NSOF_CUDA_TRY(cudaMemsetAsync(h->guard_buf.p, 0, sizeof(int), stream));
    FbTiling t = {rank, world, halo_r1, (int *)h->guard_buf.p};
    rc = fb_prepare(h, items, n, params, &t, stream);
    if (rc) return rc;
    // every item must have every level, with a band of at least the blur radius on every rank
    const std::vector<FbItemPlan> &plan = *h->last_plan;
    for (int i = 0; i < n; i++) {
        if (plan[i].scales != h->ctx->max_scales) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "fb_tiled_begin: items must share the level count");
        const int hk = plan[i].hk[plan[i].scales - 1];
        if (hk / world < params->winsize / 2 + 1 || hk / world < halo_r1)
            NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "fb_tiled_begin: the coarsest level (%d rows) is too small for %d ranks", hk, world);
    }
    return NSOF_OK;
}

extern "C" int nsof_fb_tiled_levels(const nsof_handle *h)
{
    return (h && h->ctx && h->ctx->valid && h->ctx->tiled) ? h->ctx->max_scales : 0;
}

// stage: 0 pyramid + PolyExp, 1 first UpdateMatrices, 2 blur + solve iteration `it`
extern "C" int nsof_fb_tiled_stage(nsof_handle *h, int level, int stage, int it, void *cuda_stream)
{
    return nsof_fb_tiled_stage_rows(h, level, stage, it, 0, cuda_stream);
}

// The same stage on a sub-window of the band: part 0 the whole band, 1 / 2 its top / bottom `halo_r1` rows (everything
// a row neighbour waits for), 3 the rows in between.  Parts 1 + 2 + 3 write exactly what part 0 writes, so a caller
// can start the halo exchange after the edges and compute the interior while the rows travel.
extern "C" int nsof_fb_tiled_stage_rows(nsof_handle *h, int level, int stage, int it, int part, void *cuda_stream)
{
    if (!h || !h->ctx || !h->ctx->valid || !h->ctx->tiled) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_stage: no tiled batch (call nsof_fb_tiled_begin)");
    FbBatchCtx &X = *h->ctx;
    if (level < 0 || level >= X.max_scales) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_stage: level %d", level);
    if (part < 0 || part >= X.n_variants) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_stage: part %d", part);
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    switch (stage) {
    case 0: return fb_launch_pyramid(h, level, stream, part);
    case 1: return fb_launch_um_first(h, level, stream, part);
    case 2:
        if (it < 0 || it >= X.params.iterations) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_stage: iteration %d", it);
        return fb_launch_blur(h, level, it, X.params.iterations, stream, part);
    default: NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_stage: stage %d", stage);
    }
}

// Buffer `what` of item `item` on `level`: 0 R1 (5 planes), 1 Ma, 2 Mb (5 planes), 3 the level's flow (float2 rows).
// M written by UpdateMatrices number u of a level (u = 0 first, then one per non-last iteration) is Ma for even u.
extern "C" int nsof_fb_tiled_query(nsof_handle *h, int item, int level, int what, nsof_tiled_buf *out)
{
    if (!h || !out || !h->ctx || !h->ctx->valid || !h->ctx->tiled) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_query: no tiled batch");
    FbBatchCtx &X = *h->ctx;
    const std::vector<FbItemPlan> &plan = *h->last_plan;
    if (item < 0 || item >= X.n || level < 0 || level >= X.max_scales) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_query: item/level");
    const FbItemPlan &pl = plan[item];
    float *base = (float *)h->arena.p;
    const size_t planef = (size_t)pl.pitch[level] * pl.hk[level];
    memset(out, 0, sizeof(*out));
    out->h = pl.hk[level];
    out->w = pl.wk[level];
    band_rows(out->h, X.tiling.rank, X.tiling.world, &out->y_lo, &out->y_hi);
    switch (what) {
    case 0: out->ptr = base + pl.off_R[level] + 5 * planef; out->planes = 5; out->row_floats = pl.pitch[level]; out->plane_floats = planef; break;
    case 1: out->ptr = base + pl.off_Ma; out->planes = 5; out->row_floats = pl.pitch[level]; out->plane_floats = planef; break;
    case 2: out->ptr = base + pl.off_Mb; out->planes = 5; out->row_floats = pl.pitch[level]; out->plane_floats = planef; break;
    case 3:
        out->planes = 1;
        if (level == 0) { out->ptr = pl.d.flow; out->row_floats = (size_t)(pl.d.flow_pitch / 4); }
        else { out->ptr = (level & 1) ? base + pl.off_fA : base + pl.off_fB; out->row_floats = 2 * (size_t)pl.pitch[level]; }
        out->plane_floats = out->row_floats * out->h;
        break;
    default: NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_query: what=%d", what);
    }
    return NSOF_OK;
}

// ---- row tiling over peer memory (NVLink): the whole level loop in ONE host call -------------------------------------
// Every rank plans the whole frame with the same descriptors, so a row of any workspace buffer has the SAME offset
// from the arena base on every rank.  With the row neighbours' arenas mapped into this process (CUDA IPC, or plain
// pointers when the ranks are devices / threads of one process), an exchange is one kernel: k_fb_halo_push stores the
// band's edge rows of all pairs and planes straight into the neighbour's halo rows at the same offsets and then
// raises the neighbour's flag to the exchange's sequence number; the consumer's stream carries a one-thread kernel
// (k_fb_halo_wait) that spins until BOTH of its own flags have reached that number.  No pack / unpack kernels, no
// message buffers, no library collective, no host round trip between stages.  Producers never wait on consumers, and
// every overwrite of a halo row is separated from its last reader by a flag wait (tiled_begin demands >= 1 iteration),
// so ranks cannot deadlock or race however far one runs ahead.  The ranks MUST run on different devices (or, in the
// emulator, different threads): a spinning kernel and the kernel it waits for cannot share one GPU.
struct FbPushBlock {
    void *dst;
    const void *src;
    size_t bytes;
};

NSOF_DEV void nsof_store_release_sys(unsigned long long *p, unsigned long long v)
{
#ifdef NSOF_EMU
    __atomic_store_n(p, v, __ATOMIC_RELEASE);
#else
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
#endif
}
NSOF_DEV unsigned long long nsof_load_acquire_sys(const unsigned long long *p)
{
#ifdef NSOF_EMU
    return __atomic_load_n(p, __ATOMIC_ACQUIRE);
#else
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
#endif
}

// grid (chunks, blocks).  `done` counts finished CTAs (local memory, reset by the last one); the last CTA publishes.
NSOF_GLOBAL k_fb_halo_push(const FbPushBlock *NSOF_RESTRICT blocks, unsigned total_ctas, unsigned *done,
                           unsigned long long *flag_up, unsigned long long *flag_dn, unsigned long long seq)
{
    const FbPushBlock B = blocks[NSOF_BY];
    const size_t stride = (size_t)nsof_grid_x() * NSOF_NT;
    const float4 *s4 = (const float4 *)B.src;   // rows of pitch-padded planes: 16-byte aligned multiples of 16 bytes
    float4 *d4 = (float4 *)B.dst;
    for (size_t i = (size_t)NSOF_BX * NSOF_NT + NSOF_TID; i < B.bytes / 16; i += stride) d4[i] = s4[i];
#ifdef NSOF_EMU
    if (++*done == total_ctas) {
        *done = 0;
        if (flag_up) nsof_store_release_sys(flag_up, seq);
        if (flag_dn) nsof_store_release_sys(flag_dn, seq);
    }
#else
    __threadfence_system();  // this thread's peer stores are performed system-wide before the CTA reports
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned prev = atomicAdd(done, 1u);
        if (prev + 1 == total_ctas) {
            *done = 0;
            __threadfence_system();
            if (flag_up) nsof_store_release_sys(flag_up, seq);
            if (flag_dn) nsof_store_release_sys(flag_dn, seq);
        }
    }
#endif
}

NSOF_GLOBAL k_fb_halo_wait(const unsigned long long *from_up, const unsigned long long *from_dn, unsigned long long seq)
{
    if (NSOF_TID != 0) return;
    if (from_up) while (nsof_load_acquire_sys(from_up) < seq) { }
    if (from_dn) while (nsof_load_acquire_sys(from_dn) < seq) { }
}

// flags buffer of a handle (peer-writable): [0] raised by the upper neighbour, [1] by the lower one, [2] CTA counter
static int fb_link_flags(nsof_handle *h)
{
    if (h->link_flags.p) return NSOF_OK;
    int rc = h->link_flags.ensure(256);
    if (rc) return rc;
    NSOF_CUDA_TRY(cudaMemset(h->link_flags.p, 0, 256));
    return NSOF_OK;
}

static void fb_link_close(nsof_handle *h)
{
#ifndef NSOF_EMU
    for (void *p : h->link_ipc) cudaIpcCloseMemHandle(p);
#endif
    h->link_ipc.clear();
    h->link_up_arena = h->link_dn_arena = nullptr;
    h->link_up_flags = h->link_dn_flags = nullptr;
    h->link_arena_seen = nullptr;
}

// Export: 64-byte CUDA IPC handles of this rank's arena and flag buffer (call after nsof_fb_tiled_begin: the arena is
// sized by the plan; `arena_bytes` lets the caller detect a re-allocation), plus their raw addresses for ranks that
// share a process.
extern "C" int nsof_fb_tiled_link_export(nsof_handle *h, void *arena_ipc64, void *flags_ipc64, void **arena_ptr, void **flags_ptr,
                                         size_t *arena_bytes)
{
    if (!h || !h->ctx || !h->ctx->valid || !h->ctx->tiled) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_link_export: no tiled batch");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    int rc = fb_link_flags(h);
    if (rc) return rc;
#ifndef NSOF_EMU
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handles travel as 64 bytes");
    if (arena_ipc64) NSOF_CUDA_TRY(cudaIpcGetMemHandle((cudaIpcMemHandle_t *)arena_ipc64, h->arena.p));
    if (flags_ipc64) NSOF_CUDA_TRY(cudaIpcGetMemHandle((cudaIpcMemHandle_t *)flags_ipc64, h->link_flags.p));
#else
    if (arena_ipc64) memset(arena_ipc64, 0, 64);
    if (flags_ipc64) memset(flags_ipc64, 0, 64);
#endif
    if (arena_ptr) *arena_ptr = h->arena.p;
    if (flags_ptr) *flags_ptr = h->link_flags.p;
    if (arena_bytes) *arena_bytes = h->arena.bytes;
    return NSOF_OK;
}

// Attach the row neighbours.  use_ipc != 0: the four arguments are 64-byte IPC handles of OTHER processes (NULL = no
// such neighbour) and are opened here; use_ipc == 0: they are device pointers valid in this process.
extern "C" int nsof_fb_tiled_link_attach(nsof_handle *h, const void *up_arena, const void *up_flags, const void *dn_arena,
                                         const void *dn_flags, int use_ipc)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_link_attach: NULL handle");
    if ((up_arena == nullptr) != (up_flags == nullptr) || (dn_arena == nullptr) != (dn_flags == nullptr))
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_link_attach: a neighbour needs both its arena and its flags");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    int rc = fb_link_flags(h);
    if (rc) return rc;
    fb_link_close(h);
    void *p[4] = {nullptr, nullptr, nullptr, nullptr};
    const void *in[4] = {up_arena, up_flags, dn_arena, dn_flags};
    for (int i = 0; i < 4; i++) {
        if (!in[i]) continue;
        if (use_ipc) {
#ifndef NSOF_EMU
            cudaIpcMemHandle_t hd;
            memcpy(&hd, in[i], sizeof(hd));
            NSOF_CUDA_TRY(cudaIpcOpenMemHandle(&p[i], hd, cudaIpcMemLazyEnablePeerAccess));
            h->link_ipc.push_back(p[i]);
#else
            NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "fb_tiled_link_attach: no CUDA IPC in the emulator build");
#endif
        } else {
            p[i] = const_cast<void *>(in[i]);
        }
    }
    h->link_up_arena = (float *)p[0];
    h->link_up_flags = (unsigned long long *)p[1];
    h->link_dn_arena = (float *)p[2];
    h->link_dn_flags = (unsigned long long *)p[3];
    h->link_arena_seen = h->arena.p;
    return NSOF_OK;
}

// Edge rows of buffer `what` (nsof_fb_tiled_query numbering) of level k, all items and planes, as push blocks
static void fb_push_blocks(nsof_handle *h, int k, int what, int rows, std::vector<FbPushBlock> *out, size_t *biggest)
{
    FbBatchCtx &X = *h->ctx;
    const std::vector<FbItemPlan> &plan = *h->last_plan;
    float *base = (float *)h->arena.p;
    for (int i = 0; i < X.n; i++) {
        const FbItemPlan &pl = plan[i];
        const size_t planef = (size_t)pl.pitch[k] * pl.hk[k];
        size_t off, row_floats = pl.pitch[k], plane_floats = planef;
        int planes = 5;
        switch (what) {
        case 0: off = pl.off_R[k] + 5 * planef; break;
        case 1: off = pl.off_Ma; break;
        case 2: off = pl.off_Mb; break;
        default: off = (k & 1) ? pl.off_fA : pl.off_fB; planes = 1; row_floats = 2 * (size_t)pl.pitch[k]; plane_floats = row_floats * pl.hk[k]; break;
        }
        int y_lo, y_hi;
        band_rows(pl.hk[k], X.tiling.rank, X.tiling.world, &y_lo, &y_hi);
        const int r = std::min(rows, y_hi - y_lo);
        const size_t bytes = (size_t)r * row_floats * sizeof(float);
        if (!bytes) continue;
        for (int c = 0; c < planes; c++) {
            const size_t top = off + c * plane_floats + (size_t)y_lo * row_floats;
            const size_t bot = off + c * plane_floats + (size_t)(y_hi - r) * row_floats;
            if (h->link_up_arena) out->push_back({h->link_up_arena + top, base + top, bytes});
            if (h->link_dn_arena) out->push_back({h->link_dn_arena + bot, base + bot, bytes});
            *biggest = std::max(*biggest, bytes);
        }
    }
}

// The level loop of the batch planned by nsof_fb_tiled_begin, every stage followed by its halo push and flag wait.
// `halo_flow` = rows of the coarser flow the bilinear upsampling reaches beyond the band (4).  Asynchronous on the
// stream; check nsof_fb_tiled_guard afterwards.  `first_seq` = this batch's first exchange number: every rank passes
// the same, strictly increasing value (the driver keeps the running count and gets the next one back).
extern "C" int nsof_fb_tiled_run(nsof_handle *h, int halo_flow, unsigned long long first_seq, unsigned long long *next_seq,
                                 void *cuda_stream)
{
    if (!h || !h->ctx || !h->ctx->valid || !h->ctx->tiled) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_run: no tiled batch (call nsof_fb_tiled_begin)");
    FbBatchCtx &X = *h->ctx;
    const int world = X.tiling.world, rank = X.tiling.rank;
    if (world > 1) {
        if ((rank > 0 && !h->link_up_arena) || (rank < world - 1 && !h->link_dn_arena))
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_run: row neighbours are not attached (nsof_fb_tiled_link_attach)");
        if (h->link_arena_seen != h->arena.p)
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_run: the workspace was re-allocated after the neighbours were attached; export and attach again");
    }
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    int rc = fb_link_flags(h);
    if (rc) return rc;
    const int T = X.params.iterations, m = X.m;
    // ---- push tables of every exchange of the batch, one upload
    struct Ex { int k, what, rows; size_t first, count, biggest; };
    std::vector<Ex> exs;
    std::vector<FbPushBlock> blocks;
    auto add = [&](int k, int what, int rows) {
        Ex e = {k, what, rows, blocks.size(), 0, 0};
        if (world > 1) fb_push_blocks(h, k, what, rows, &blocks, &e.biggest);
        e.count = blocks.size() - e.first;
        exs.push_back(e);
    };
    for (int k = X.max_scales - 1; k >= X.stop_level; k--) {
        add(k, 0, X.tiling.halo_r1);
        add(k, 1, m);
        for (int it = 0; it < T; it++) {
            if (it < T - 1) add(k, (it % 2 == 0) ? 2 : 1, m);
            else add(k, 3, k > 0 ? halo_flow : 0);
        }
    }
    const FbPushBlock *dblocks = nullptr;
    if (!blocks.empty()) {
        const size_t bytes = blocks.size() * sizeof(FbPushBlock);
        NsofStageSlot *slot;
        rc = nsof_stage_acquire(h, bytes, &slot);
        if (rc) return rc;
        rc = h->link_tab.ensure(bytes + 16);
        if (rc) return rc;
        memcpy(slot->host, blocks.data(), bytes);
        rc = nsof_table_upload(h, h->link_tab.p, slot, bytes, stream);
        if (rc) return rc;
        dblocks = (const FbPushBlock *)h->link_tab.p;
    }
    unsigned long long *fl = (unsigned long long *)h->link_flags.p;
    unsigned long long seq = first_seq;
    size_t ei = 0;
    // an exchange = push (store the edge rows into the neighbours' halo rows, raise their flags) + wait (own flags)
    auto push = [&]() -> int {
        const Ex &e = exs[ei];
        seq++;
        if (world == 1 || e.rows <= 0) return NSOF_OK;  // (a skipped exchange still consumes its number on every rank)
        // peer flags: the UPPER neighbour's "raised by the lower one" slot [1], the LOWER neighbour's slot [0]
        unsigned long long *f_up = h->link_up_arena ? h->link_up_flags + 1 : nullptr;
        unsigned long long *f_dn = h->link_dn_arena ? h->link_dn_flags + 0 : nullptr;
        if (e.count) {
            double total = 0.;
            for (size_t b = 0; b < e.count; b++) total += (double)blocks[e.first + b].bytes;
            const int chunks = (int)std::min<size_t>(32, std::max<size_t>(1, e.biggest / (16 * 256 * 4)));
            NSOF_KLAUNCH(h, NSOF_K_HALO_COPY, 2 * total, k_fb_halo_push, chunks, (int)e.count, 1, 256, 0, stream, dblocks + e.first,
                         (unsigned)(chunks * e.count), (unsigned *)(fl + 2), f_up, f_dn, seq);
        }
        return NSOF_OK;
    };
    auto wait = [&]() -> int {
        const Ex &e = exs[ei++];
        if (world == 1 || e.rows <= 0) return NSOF_OK;
        NSOF_KLAUNCH(h, NSOF_K_HALO_COPY, 0., k_fb_halo_wait, 1, 1, 1, 32, 0, stream,
                     (const unsigned long long *)(h->link_up_arena ? fl + 0 : nullptr),
                     (const unsigned long long *)(h->link_dn_arena ? fl + 1 : nullptr), seq);
        NSOF_CUDA_TRY(cudaGetLastError());
        return NSOF_OK;
    };
    // A stage of a level whose bands are at least four edges high runs as  top + bottom edge rows -> push -> rows in
    // between -> wait  (the sub-window job tables of nsof_fb_tiled_stage_rows): the rows travel, and the neighbours catch
    // up, while the bulk of the band is computed.  Opt-in (NSOF_TILED_SPLIT=1): measured SLOWER on 2 B200 (8K, 8 pairs:
    // 527 - 530 vs 549 pairs/s) -- three launches and two more partially filled tile rows per stage cost more than the
    // ~20 us of flag latency they hide.
    const char *env_split = getenv("NSOF_TILED_SPLIT");
    const bool want_split = world > 1 && env_split && env_split[0] == '1';
    auto stage = [&](int k, int what, int it) -> int {
        bool split = want_split && exs[ei].rows > 0;
        if (split) {
            const std::vector<FbItemPlan> &plan = *h->last_plan;
            for (int i = 0; i < X.n && split; i++) {
                int lo, hi;
                band_rows(plan[i].hk[k], rank, world, &lo, &hi);
                split = hi - lo >= 4 * X.edge;
            }
        }
        auto run = [&](int variant) -> int {
            return what == 0 ? fb_launch_pyramid(h, k, stream, variant) : what == 1 ? fb_launch_um_first(h, k, stream, variant)
                                                                                   : fb_launch_blur(h, k, it, T, stream, variant);
        };
        int r2;
        if (!split) {
            if ((r2 = run(0))) return r2;
            if ((r2 = push())) return r2;
            return wait();
        }
        if ((r2 = run(1))) return r2;
        if ((r2 = run(2))) return r2;
        if ((r2 = push())) return r2;
        if ((r2 = run(3))) return r2;
        return wait();
    };
    for (int k = X.max_scales - 1; k >= X.stop_level; k--) {
        if ((rc = stage(k, 0, 0))) return rc;
        if ((rc = stage(k, 1, 0))) return rc;
        for (int it = 0; it < T; it++)
            if ((rc = stage(k, 2, it))) return rc;
    }
    if (next_seq) *next_seq = seq;
    return NSOF_OK;
}

// 1 when an UpdateMatrices gather of the tiled batch reached R1 rows outside the received halo (the result is then
// invalid: rerun with a larger halo_r1).  Synchronises the stream.
extern "C" int nsof_fb_tiled_guard(nsof_handle *h, int *tripped, void *cuda_stream)
{
    if (!h || !tripped || !h->guard_buf.p) NSOF_FAIL(NSOF_ERR_BAD_ARG, "fb_tiled_guard: no tiled batch");
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    int v = 0;
    NSOF_CUDA_TRY(cudaMemcpyAsync(&v, h->guard_buf.p, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)cuda_stream));
    // read and clear
    NSOF_CUDA_TRY(cudaMemsetAsync(h->guard_buf.p, 0, sizeof(int), (cudaStream_t)cuda_stream));
    NSOF_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)cuda_stream));
    *tripped = v;
    return NSOF_OK;
}

extern "C" int nsof_farneback_u8(nsof_handle *h, const uint8_t *prev, const uint8_t *next, int height, int width,
                                 ptrdiff_t prev_pitch, ptrdiff_t next_pitch, const nsof_fb_params *params, float *flow,
                                 ptrdiff_t flow_pitch, void *cuda_stream)
{
    nsof_crop_desc d = {prev, next, height, width, prev_pitch, next_pitch, flow, flow_pitch};
    return nsof_farneback_batch(h, &d, 1, params, cuda_stream);
}

// HOST-pointer entry: stage crops into h->io, run, copy flow back, synchronise.
extern "C" int nsof_farneback_batch_host(nsof_handle *h, const nsof_crop_desc *items, int n, const nsof_fb_params *params)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: NULL handle");
    if (n < 0 || (n > 0 && !items)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: bad item list");
    int rc = validate_params(params);
    if (rc) return rc;
    if (n == 0) return NSOF_OK;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    std::vector<nsof_crop_desc> dev(n);
    const ptrdiff_t esz = (params->flags & NSOF_FB_SRC_F32) ? 4 : 1;  // float32 frames (cv2 accepts them)
    size_t bytes = 0;
    std::vector<size_t> off_img(n), off_flow(n), ipitch(n);
    for (int i = 0; i < n; i++) {
        const nsof_crop_desc &d = items[i];
        if (!d.prev || !d.next || !d.flow) NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: item %d has a NULL pointer", i);
        if (d.h < 1 || d.w < 1)
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: item %d is empty (%dx%d) (cv2: assertion !_src.empty())", i, d.h, d.w);
        if (d.prev_pitch < d.w * esz || d.next_pitch < d.w * esz || d.flow_pitch < (ptrdiff_t)d.w * 8)
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "farneback: item %d has a bad pitch", i);
        ipitch[i] = round_up((size_t)d.w * esz, 128);
        off_img[i] = bytes;
        bytes += 2 * ipitch[i] * d.h;
        bytes = round_up(bytes, 256);
        off_flow[i] = bytes;
        bytes += (size_t)d.w * d.h * 8;
        bytes = round_up(bytes, 256);
    }
    rc = h->io.ensure(bytes);
    if (rc) return rc;
    unsigned char *io = (unsigned char *)h->io.p;
    cudaStream_t stream = nullptr;
    for (int i = 0; i < n; i++) {
        const nsof_crop_desc &d = items[i];
        uint8_t *p0 = io + off_img[i], *p1 = p0 + ipitch[i] * d.h;
        NSOF_CUDA_TRY(cudaMemcpy2DAsync(p0, ipitch[i], d.prev, d.prev_pitch, (size_t)d.w * esz, d.h, cudaMemcpyHostToDevice, stream));
        NSOF_CUDA_TRY(cudaMemcpy2DAsync(p1, ipitch[i], d.next, d.next_pitch, (size_t)d.w * esz, d.h, cudaMemcpyHostToDevice, stream));
        dev[i].prev = p0; dev[i].next = p1; dev[i].h = d.h; dev[i].w = d.w;
        dev[i].prev_pitch = dev[i].next_pitch = (ptrdiff_t)ipitch[i];
        dev[i].flow = (float *)(io + off_flow[i]);
        dev[i].flow_pitch = (ptrdiff_t)d.w * 8;
        if (params->flags & NSOF_FB_USE_INITIAL_FLOW)  // flow is in/out, as cv2's InputOutputArray
            NSOF_CUDA_TRY(cudaMemcpy2DAsync(dev[i].flow, dev[i].flow_pitch, d.flow, d.flow_pitch, (size_t)d.w * 8, d.h,
                                            cudaMemcpyHostToDevice, stream));
    }
    rc = nsof_farneback_batch(h, dev.data(), n, params, stream);
    if (rc) return rc;
    for (int i = 0; i < n; i++) {
        const nsof_crop_desc &d = items[i];
        NSOF_CUDA_TRY(cudaMemcpy2DAsync(d.flow, d.flow_pitch, dev[i].flow, dev[i].flow_pitch, (size_t)d.w * 8, d.h,
                                        cudaMemcpyDeviceToHost, stream));
    }
    NSOF_CUDA_TRY(cudaStreamSynchronize(stream));
    return NSOF_OK;
}

extern "C" int nsof_farneback_u8_host(nsof_handle *h, const uint8_t *prev, const uint8_t *next, int height, int width,
                                      ptrdiff_t prev_pitch, ptrdiff_t next_pitch, const nsof_fb_params *params,
                                      float *flow, ptrdiff_t flow_pitch)
{
    nsof_crop_desc d = {prev, next, height, width, prev_pitch, next_pitch, flow, flow_pitch};
    return nsof_farneback_batch_host(h, &d, 1, params);
}

// ------------------------------------------------------------------------------ S2-g helpers
struct FbClearJob {
    float2 *p;
    int h, w, fpitch;  // fpitch in float2
};

// grid (row groups, items): 8 rows per block, one warp per row, float2 stores (crops are only 8-byte aligned)
NSOF_GLOBAL k_fb_flow_clear(const FbClearJob *NSOF_RESTRICT jobs)
{
    const FbClearJob J = jobs[NSOF_BY];
    const int nw = nsof_imax(NSOF_NT / NSOF_WARP, 1), warp = NSOF_TID / NSOF_WARP, lane = NSOF_LANE;
    for (int y = NSOF_BX * nw + warp; y < J.h; y += nsof_grid_x() * nw) {
        float2 *row = J.p + (size_t)y * J.fpitch;
        for (int x = lane; x < J.w; x += NSOF_WARP) row[x] = make_float2(0.f, 0.f);
    }
}

extern "C" int nsof_flow_clear(nsof_handle *h, const nsof_crop_desc *items, int n, void *cuda_stream)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_clear: NULL handle");
    if (n < 0 || (n > 0 && !items)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_clear: bad item list");
    if (n == 0) return NSOF_OK;
    if (n > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "flow_clear: at most 65535 items per call");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t bytes = (size_t)n * sizeof(FbClearJob);
    NsofStageSlot *slot;
    int rc = nsof_stage_acquire(h, bytes, &slot);
    if (rc) return rc;
    rc = h->clear_tab.ensure(bytes + 16);
    if (rc) return rc;
    FbClearJob *hj = (FbClearJob *)slot->host;
    int max_h = 0;
    double px = 0.;
    for (int i = 0; i < n; i++) {
        const nsof_crop_desc &d = items[i];
        if (!d.flow || d.h < 1 || d.w < 1 || d.flow_pitch < (ptrdiff_t)d.w * 8 || d.flow_pitch % 8 || (uintptr_t)d.flow % 8)
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "flow_clear: item %d is malformed", i);
        hj[i].p = (float2 *)d.flow;
        hj[i].h = d.h;
        hj[i].w = d.w;
        hj[i].fpitch = (int)(d.flow_pitch / 8);
        max_h = std::max(max_h, d.h);
        px += (double)d.h * d.w;
    }
    rc = nsof_table_upload(h, h->clear_tab.p, slot, bytes, stream);
    if (rc) return rc;
    NSOF_KLAUNCH(h, NSOF_K_FLOW_CLEAR, 8 * px, k_fb_flow_clear, std::min((max_h + 7) / 8, 64), n, 1, 256, 0, stream,
                 (const FbClearJob *)h->clear_tab.p);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

// grid (chunks, blocks): every block of bytes is copied by up to 64 CTAs, 16 bytes per thread where aligned
NSOF_GLOBAL k_copy_blocks(const nsof_copy_block *NSOF_RESTRICT blocks)
{
    const nsof_copy_block B = blocks[NSOF_BY];
    const size_t stride = (size_t)nsof_grid_x() * NSOF_NT;
    if ((((uintptr_t)B.dst | (uintptr_t)B.src | B.bytes) & 15) == 0) {
        const float4 *s4 = (const float4 *)B.src;
        float4 *d4 = (float4 *)B.dst;
        for (size_t i = (size_t)NSOF_BX * NSOF_NT + NSOF_TID; i < B.bytes / 16; i += stride) d4[i] = s4[i];
    } else {
        const unsigned char *s1 = (const unsigned char *)B.src;
        unsigned char *d1 = (unsigned char *)B.dst;
        for (size_t i = (size_t)NSOF_BX * NSOF_NT + NSOF_TID; i < B.bytes; i += stride) d1[i] = s1[i];
    }
}

extern "C" int nsof_copy_blocks(nsof_handle *h, const nsof_copy_block *blocks, int n, void *cuda_stream)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "copy_blocks: NULL handle");
    if (n < 0 || (n > 0 && !blocks)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "copy_blocks: bad block list");
    if (n == 0) return NSOF_OK;
    if (n > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "copy_blocks: at most 65535 blocks per call");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const size_t bytes = (size_t)n * sizeof(nsof_copy_block);
    NsofStageSlot *slot;
    int rc = nsof_stage_acquire(h, bytes, &slot);
    if (rc) return rc;
    rc = h->clear_tab.ensure(bytes + 16);
    if (rc) return rc;
    size_t biggest = 0;
    double total = 0.;
    for (int i = 0; i < n; i++) {
        if (!blocks[i].dst || !blocks[i].src) NSOF_FAIL(NSOF_ERR_BAD_ARG, "copy_blocks: block %d has a NULL pointer", i);
        biggest = std::max(biggest, blocks[i].bytes);
        total += (double)blocks[i].bytes;
    }
    memcpy(slot->host, blocks, bytes);
    rc = nsof_table_upload(h, h->clear_tab.p, slot, bytes, stream);
    if (rc) return rc;
    const int chunks = (int)std::min<size_t>(64, std::max<size_t>(1, biggest / (16 * 256 * 4)));
    NSOF_KLAUNCH(h, NSOF_K_HALO_COPY, 2 * total, k_copy_blocks, chunks, n, 1, 256, 0, stream, (const nsof_copy_block *)h->clear_tab.p);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

// The same copy with the block table already RESIDENT in device memory (uploaded once by the caller, tiled.py keeps one
// per exchange plan): no staging slot and no upload kernel per call.
extern "C" int nsof_copy_blocks_resident(nsof_handle *h, const nsof_copy_block *device_blocks, int n, size_t biggest_block,
                                         double total_bytes, void *cuda_stream)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "copy_blocks_resident: NULL handle");
    if (n < 0 || (n > 0 && !device_blocks)) NSOF_FAIL(NSOF_ERR_BAD_ARG, "copy_blocks_resident: bad block list");
    if (n == 0) return NSOF_OK;
    if (n > 65535) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "copy_blocks_resident: at most 65535 blocks per call");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    const int chunks = (int)std::min<size_t>(64, std::max<size_t>(1, biggest_block / (16 * 256 * 4)));
    NSOF_KLAUNCH(h, NSOF_K_HALO_COPY, 2 * total_bytes, k_copy_blocks, chunks, n, 1, 256, 0, stream, device_blocks);
    NSOF_CUDA_TRY(cudaGetLastError());
    return NSOF_OK;
}

extern "C" int nsof_copy_2d_async(void *dst, ptrdiff_t dst_pitch, const void *src, ptrdiff_t src_pitch, size_t row_bytes,
                                  size_t rows, int kind, void *cuda_stream)
{
    if (!dst || !src) NSOF_FAIL(NSOF_ERR_BAD_ARG, "copy_2d: NULL pointer");
    if (kind != 1 && kind != 2) NSOF_FAIL(NSOF_ERR_BAD_ARG, "copy_2d: kind must be 1 (H2D) or 2 (D2H)");
    if (dst_pitch < (ptrdiff_t)row_bytes || src_pitch < (ptrdiff_t)row_bytes) NSOF_FAIL(NSOF_ERR_BAD_ARG, "copy_2d: bad pitch");
    if (!row_bytes || !rows) return NSOF_OK;
    NSOF_CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)dst_pitch, src, (size_t)src_pitch, row_bytes, rows,
                                    kind == 1 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, (cudaStream_t)cuda_stream));
    return NSOF_OK;
}

// ------------------------------------------------------------------------------ test hooks
extern "C" int nsof_debug_truncate(nsof_handle *h, int stop_level, int stop_iters)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->dbg_level = stop_level;
    h->dbg_iters = stop_iters;
    return NSOF_OK;
}

extern "C" int nsof_debug_big_tiles(nsof_handle *h, int enable)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->dbg_big_tiles = enable != 0;
    return NSOF_OK;
}

// 1: winsize 3 .. 5 take the generic blur+solve kernel (k_fb_blur_solve_g) instead of the tuned kernel's small-radius
// instantiations (tests: the two must agree within the cv2 tolerance; A/B)
extern "C" int nsof_debug_generic_blur(nsof_handle *h, int enable)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->dbg_generic_blur = enable != 0;
    return NSOF_OK;
}

// 0: the UpdateMatrices half of the tuned blur+solve kernel gathers R1 from L2 in every tile (A/B, bit-identity test)
extern "C" int nsof_debug_r1_window(nsof_handle *h, int enable)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->dbg_no_r1_window = enable == 0;
    h->dbg_um_round1 = enable == 0 || enable == 2;   // 2: window in blur+solve only (A/B of the two kernels)
    return NSOF_OK;
}

// mask bit 0 / 1: the tuned blur+solve kernel does not prefetch the tile's R0 / R1 rows into L2 in phase 1 (A/B);
// -1: the default (no prefetch while the R1 window is on, both prefetches with the window off)
extern "C" int nsof_debug_bs_prefetch(nsof_handle *h, int off_mask)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->dbg_bs_no_prefetch = off_mask < 0 ? -1 : (off_mask & 3);
    return NSOF_OK;
}

// out2[0..3] (FOUR counters) = {blur+solve tiles whose R1 gathers came from the shared-memory window, blur+solve tiles on
// the L2 path, the same two for the first UpdateMatrices of a level} since the previous call, then zeroed; the first call (or out2 == NULL) switches the counting on.  Synchronises `cuda_stream`.
extern "C" int nsof_debug_r1_window_stats(nsof_handle *h, unsigned long long *out2, void *cuda_stream)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (!h->dbg_r1_stats) {
        int rc = h->r1_stats_buf.ensure(4 * sizeof(unsigned));
        if (rc != NSOF_OK) return rc;
        NSOF_CUDA_TRY(cudaMemsetAsync(h->r1_stats_buf.p, 0, 4 * sizeof(unsigned), st));
        NSOF_CUDA_TRY(cudaStreamSynchronize(st));
        h->dbg_r1_stats = true;
        if (out2) out2[0] = out2[1] = out2[2] = out2[3] = 0;
        return NSOF_OK;
    }
    unsigned v[4] = {0, 0, 0, 0};
    NSOF_CUDA_TRY(cudaMemcpyAsync(v, h->r1_stats_buf.p, sizeof(v), cudaMemcpyDeviceToHost, st));
    NSOF_CUDA_TRY(cudaMemsetAsync(h->r1_stats_buf.p, 0, sizeof(v), st));
    NSOF_CUDA_TRY(cudaStreamSynchronize(st));
    if (out2) { out2[0] = v[0]; out2[1] = v[1]; out2[2] = v[2]; out2[3] = v[3]; }
    return NSOF_OK;
}

// 0: expand both images of every item even when a frame is shared between two items (tests, A/B)
extern "C" int nsof_debug_share_frames(nsof_handle *h, int enable)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->dbg_no_dedup = !enable;
    return NSOF_OK;
}

extern "C" int nsof_debug_fork(nsof_handle *h, int enable)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->no_fork = !enable;
    return NSOF_OK;
}

extern "C" int nsof_debug_fuse(nsof_handle *h, int enable)
{
    if (!h) NSOF_FAIL(NSOF_ERR_BAD_ARG, "NULL handle");
    h->dbg_no_fuse = !enable;
    return NSOF_OK;
}

extern "C" int nsof_debug_fetch(nsof_handle *h, int item, int what, float *out, int *out_h, int *out_w)
{
    if (!h || !out) NSOF_FAIL(NSOF_ERR_BAD_ARG, "debug_fetch: NULL argument");
    std::vector<FbItemPlan> &plan = *h->last_plan;
    if (item < 0 || item >= (int)plan.size()) NSOF_FAIL(NSOF_ERR_BAD_ARG, "debug_fetch: no item %d in the last batch", item);
    const FbItemPlan &pl = plan[item];
    const int k = h->last_stop_level;
    if (k >= pl.scales) NSOF_FAIL(NSOF_ERR_BAD_ARG, "debug_fetch: item %d has no level %d", item, k);
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    NSOF_CUDA_TRY(cudaDeviceSynchronize());
    const int hh = pl.hk[k], ww = pl.wk[k], pitch = pl.pitch[k];
    const size_t plane = (size_t)pitch * hh;
    float *base = (float *)h->arena.p;
    if (out_h) *out_h = hh;
    if (out_w) *out_w = ww;
    const float *src;
    int ch;
    if (what <= 1 && !h->dbg_no_fuse)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "debug_fetch: I is not materialised by the fused pyramid kernels; call nsof_debug_fuse(h, 0) first");
    switch (what) {
    case 0: src = base + pl.off_I[k]; ch = 1; break;
    case 1: src = base + pl.off_I[k] + plane; ch = 1; break;
    case 2: src = pl.r0_from >= 0 ? base + plan[pl.r0_from].off_R[k] + 5 * plane : base + pl.off_R[k]; ch = 5; break;
    case 3: src = base + pl.off_R[k] + 5 * plane; ch = 5; break;
    case 4: {
        // M currently valid: Ma after um_first, then the iterations alternate a->b->a ...
        int iters = h->last_params.iterations;
        if (h->dbg_level >= 0 && h->dbg_iters >= 0) iters = std::min(iters, h->dbg_iters);
        const int writes = iters > 0 ? iters - 1 : 0;  // the last iteration writes flow, not M
        src = base + ((writes & 1) ? pl.off_Mb : pl.off_Ma);
        ch = 5;
        break;
    }
    case 5: {
        if (k == 0) {
            NSOF_CUDA_TRY(cudaMemcpy2DAsync(out, (size_t)ww * 8, pl.d.flow, pl.d.flow_pitch, (size_t)ww * 8, hh,
                                            cudaMemcpyDeviceToHost, nullptr));
        } else {
            const float *f = (k & 1) ? base + pl.off_fA : base + pl.off_fB;
            NSOF_CUDA_TRY(cudaMemcpy2DAsync(out, (size_t)ww * 8, f, (size_t)pitch * 8, (size_t)ww * 8, hh,
                                            cudaMemcpyDeviceToHost, nullptr));
        }
        NSOF_CUDA_TRY(cudaDeviceSynchronize());
        return NSOF_OK;
    }
    default: NSOF_FAIL(NSOF_ERR_BAD_ARG, "debug_fetch: what=%d", what);
    }
    for (int c = 0; c < ch; c++)
        NSOF_CUDA_TRY(cudaMemcpy2DAsync(out + (size_t)c * hh * ww, (size_t)ww * 4, src + c * plane, (size_t)pitch * 4,
                                        (size_t)ww * 4, hh, cudaMemcpyDeviceToHost, nullptr));
    NSOF_CUDA_TRY(cudaDeviceSynchronize());
    return NSOF_OK;
}

#ifdef NSOF_EMU
namespace nsof_emu {
thread_local int bx, by, bz, gdx, gdy;
thread_local unsigned char *smem;
}
#endif
