// nsof_handle.h -- the per-(GPU, stream) context behind `nsof_handle` and the error plumbing
// shared by farneback.cu and stage1.cu.
#pragma once
#include "../../include/nsof.h"
#include "nsof_rt.h"
#ifdef NSOF_EMU
#include "emu_cuda.h"
#endif
#include <stdarg.h>
#include <stdio.h>
#include <vector>

void nsof_set_error(const char *fmt, ...);

#define NSOF_CUDA_TRY(expr)                                                                      \
    do {                                                                                         \
        cudaError_t e_ = (expr);                                                                 \
        if (e_ != cudaSuccess) {                                                                 \
            nsof_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return NSOF_ERR_CUDA;                                                                \
        }                                                                                        \
    } while (0)

#define NSOF_FAIL(code, ...)         \
    do {                             \
        nsof_set_error(__VA_ARGS__); \
        return (code);               \
    } while (0)

// Growable device buffer (grow = free + malloc after a device sync; contents are scratch).
struct NsofDevBuf {
    void *p = nullptr;
    size_t bytes = 0;
    int ensure(size_t need);
    void release();
};

// Pinned host staging ring for job tables: an upload may still be in flight when the next
// batch is being described, so each slot carries the event of its last copy.
struct NsofStageSlot {
    void *host = nullptr;
    size_t bytes = 0;
    cudaEvent_t done = nullptr;
    bool pending = false;
};

struct FbItemPlan;  // farneback.cu
struct FbBatchCtx;  // farneback.cu

// Kernel classes for the launch counter and the optional per-launch event timing (nsof_profile*).
enum NsofKernelClass {
    NSOF_K_PYR = 0, NSOF_K_POLYEXP, NSOF_K_PYR_POLYEXP, NSOF_K_UM_FIRST, NSOF_K_BLUR_UM, NSOF_K_BLUR_FLOW, NSOF_K_FLOW_INIT,
    NSOF_K_LANCZOS, NSOF_K_FRAME_SIM, NSOF_K_ROI, NSOF_K_EVENT_HIST, NSOF_K_EVENT_REPLAY, NSOF_K_SYNAPSE_STEP,
    NSOF_K_RESISTANCE, NSOF_K_FLOW_CLEAR, NSOF_K_HALO_COPY, NSOF_K_TABLE_UPLOAD, NSOF_K_SEG_PACK, NSOF_K_SEG_MORPH,
    NSOF_K_SEG_UNPACK, NSOF_K_RGB2GRAY, NSOF_K_REMAP, NSOF_K_TILE_LIST, NSOF_K_SSIM, NSOF_K_CLASSES
};
extern const char *const nsof_kernel_class_names[NSOF_K_CLASSES];

struct NsofProfRec {
    int cls, level;  // pyramid level of a stage-2 launch, -1 otherwise
    size_t ev;  // index of the start event in the pool; stop event is ev + 1
    double bytes;
};

struct nsof_handle {
    int device = 0;
    NsofDevBuf arena;      // I, R, M, coarse flow of every item of the current batch
    NsofDevBuf tables;     // job tables + Gaussian taps on the device
    NsofDevBuf tile_buf;   // flat tile lists of the current batch (uint2 = item, tile), built by k_fb_build_tiles
    // row tiling over peer memory (nsof_fb_tiled_link_* / nsof_fb_tiled_run)
    NsofDevBuf link_flags;  // [0] raised by the upper neighbour, [1] by the lower one, [2] CTA counter of k_fb_halo_push
    NsofDevBuf link_tab;    // push-block tables of the current batch
    float *link_up_arena = nullptr, *link_dn_arena = nullptr;  // the neighbours' arenas, mapped here
    unsigned long long *link_up_flags = nullptr, *link_dn_flags = nullptr;
    void *link_arena_seen = nullptr;  // this rank's arena address when the neighbours were attached
    std::vector<void *> link_ipc;     // IPC mappings to close
    void *jpeg = nullptr;   // nvJPEG handle + state (decode.cu), created on first use
    NsofDevBuf jpeg_buf;    // decoded BGR frames of the current nsof_jpeg_decode_gray call
    NsofDevBuf clear_tab;  // rectangle table of nsof_flow_clear
    NsofDevBuf guard_buf;  // halo guard flag of a tiled batch
    FbBatchCtx *ctx = nullptr;  // plan, tables and launch geometry of the current batch
    NsofDevBuf io;         // device staging of the *_host entry points (crops in, flow out)
    NsofDevBuf s1;         // stage-1 scratch
    NsofDevBuf warp_tab;   // region table of nsof_remap_flow_batch (frames.cu)
    NsofDevBuf seg;        // flow segmentation: job table + two bit planes per item (segment.cu)
    NsofDevBuf lz;         // Lanczos-3 weight/index tables (device), cached for lz_key
    int lz_key[4] = {0, 0, 0, 0};
    int lz_P[2] = {0, 0};  // seg_rows, n_seg of the cached Lanczos geometry
    bool lz_smem_set = false;
    int lz_mma[3] = {0, 0, 0};        // windows of 8 cells, padded width, weight exponent of the tensor-core column pass
    size_t lz_off[3] = {0, 0, 0};     // byte offsets of yband / window table / int8 weight slices in `lz`
    bool dbg_lz_f64 = false;          // force the FP64 Lanczos kernels (tests, A/B)
    bool roi_smem_set = false;  // k_s1_roi opted into > 48 KB of dynamic shared memory on this handle's device
    NsofStageSlot stage[4];
    int stage_next = 0;
    // debug truncation (tests only)
    int dbg_level = -1, dbg_iters = -1;
    bool dbg_no_dedup = false;  // expand shared frames once per item (nsof_debug_share_frames(h, 0))
    bool dbg_no_fuse = false;  // force the generic pyramid + PolyExp kernels (so that I exists and can be fetched)
    // geometry of the last batch, kept for nsof_debug_fetch
    std::vector<FbItemPlan> *last_plan = nullptr;
    int last_stop_level = 0;
    nsof_fb_params last_params = {};
    bool smem_configured = false;
    // fork-join inside nsof_farneback_batch: the pyramid + PolyExp launches of every level but the coarsest do not
    // depend on the flow, so they run on internal streams beside the (latency-bound) coarse-level iterations
    cudaStream_t aux[2] = {nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr;
    std::vector<cudaEvent_t> ev_level;
    bool dbg_big_tiles = false;  // keep the 64x32 blur tile on every level (tests)
    bool dbg_generic_blur = false;  // winsize 3 .. 5 on the generic blur+solve kernel (nsof_debug_generic_blur)
    bool dbg_um_round1 = false;     // first UpdateMatrices of a level: the round-1 kernel (L2 gathers) instead of the window kernel
    bool dbg_no_r1_window = false;  // blur+solve: gather R1 from L2 in every tile (nsof_debug_r1_window(h, 0))
    int dbg_bs_no_prefetch = -1;    // bit 0 / 1: no L2 prefetch of R0 / R1 rows in blur+solve phase 1; -1: default (see the blur+solve launch in farneback.cu)
    bool dbg_r1_stats = false;      // count the tiles of either path (nsof_debug_r1_window_stats)
    NsofDevBuf r1_stats_buf;        // 2 x unsigned: tiles through the shared-memory window, tiles on the L2 path
    bool no_fork = false;  // nsof_debug_fork(h, 0): everything on the caller's stream
    // launch accounting
    unsigned long long launches = 0;
    bool prof_on = false;
    int prof_level = -1;
    std::vector<cudaEvent_t> prof_pool;
    size_t prof_used = 0;
    std::vector<NsofProfRec> prof_recs;
};

// Bracket one kernel launch: counts it, and when profiling is on records a CUDA event pair on `st`.
static inline void nsof_prof_pre(nsof_handle *h, int cls, double bytes, cudaStream_t st)
{
    h->launches++;
    if (!h->prof_on) return;
    while (h->prof_pool.size() < h->prof_used + 2) {
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        h->prof_pool.push_back(e);
    }
    NsofProfRec r = {cls, h->prof_level, h->prof_used, bytes};
    h->prof_recs.push_back(r);
    cudaEventRecord(h->prof_pool[h->prof_used], st);
}
static inline void nsof_prof_post(nsof_handle *h, cudaStream_t st)
{
    if (!h->prof_on) return;
    cudaEventRecord(h->prof_pool[h->prof_used + 1], st);
    h->prof_used += 2;
}
#define NSOF_KLAUNCH(h, cls, bytes, kern, gx, gy, gz, nt, smem_bytes, stream, ...)  \
    do {                                                                            \
        nsof_prof_pre((h), (cls), (bytes), (stream));                               \
        NSOF_LAUNCH(kern, gx, gy, gz, nt, smem_bytes, stream, __VA_ARGS__);         \
        nsof_prof_post((h), (stream));                                              \
    } while (0)

void nsof_jpeg_release(nsof_handle *h);  // decode.cu
int nsof_stage_acquire(nsof_handle *h, size_t bytes, NsofStageSlot **slot);
// device <- pinned staging slot by a kernel (no copy engine); both sides allocated with >= 16 bytes of slack
int nsof_table_upload(nsof_handle *h, void *dst, NsofStageSlot *slot, size_t bytes, cudaStream_t stream);
