/*
 * nsof.h -- C ABI of libnsof: the B200 (sm_100a) implementation of the ROI-gated Farneback
 * hot path of RTCartist/Neuromorphic-Spatiotemporal-Optical-Flow.
 *
 * Every entry point is `extern "C"`, takes plain pointers and sizes, returns an int status
 * (0 = ok, negative = error, text via nsof_last_error()) and never throws.  Citations
 * (file:line) are relative to the reference repository and name the interface each entry point
 * replaces.  There is no CPU fallback: without a CUDA device nsof_create() fails.
 *
 * Threading: one handle per (GPU, stream).  Calls on one handle are serialised by the caller.
 * "_host" entry points take host pointers, copy in/out and synchronise; the others take device
 * pointers, enqueue on `cuda_stream` (a cudaStream_t passed as void*) and return.
 */
#ifndef NSOF_H
#define NSOF_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define NSOF_API __declspec(dllexport)
#else
#define NSOF_API __attribute__((visibility("default")))
#endif

#define NSOF_OK 0
#define NSOF_ERR_BAD_ARG (-1)     /* what cv2 reports as cv2.error -215 (empty input, size mismatch, pyr_scale >= 1) */
#define NSOF_ERR_UNSUPPORTED (-2) /* legal for OpenCV but outside this library (winsize 1, poly_n > 31, ...) */
#define NSOF_ERR_CUDA (-3)
#define NSOF_ERR_NOMEM (-4)

typedef struct nsof_handle nsof_handle;

/* cv2.OPTFLOW_FARNEBACK_GAUSSIAN: Gaussian winsize x winsize window instead of the box (SURVEY.md 8(a) S2-e') */
#define NSOF_FB_GAUSSIAN 256
/* cv2.OPTFLOW_USE_INITIAL_FLOW: `flow` is in/out; the coarsest level starts from resize(flow, INTER_AREA) * scale
 * instead of zero (no call site of the reference passes a flow; implemented for the drop-in signature) */
#define NSOF_FB_USE_INITIAL_FLOW 4
/* Not a cv2 flag: `prev` / `next` point to FLOAT32 frames (pitches in bytes).  cv2.calcOpticalFlowFarneback accepts
 * CV_32F images as well as CV_8U (its convertTo(CV_32F) is then the identity); no reference call site passes them.
 * They take the direct pyramid kernel with typed loads (the tuned kernels stage uint8 source windows). */
#define NSOF_FB_SRC_F32 0x10000

/* The kwargs of `farneback_params` (optical_flow_seg.py:73-81), same names and meaning as
 * cv2.calcOpticalFlowFarneback(prev, next, flow, pyr_scale, levels, winsize, iterations, poly_n,
 * poly_sigma, flags).  poly_n is the radius of the (2*poly_n+1)^2 fit window, as in OpenCV. */
typedef struct nsof_fb_params {
    double pyr_scale;
    int levels;
    int winsize;
    int iterations;
    int poly_n;
    double poly_sigma;
    int flags; /* 0 (every call site of the reference), NSOF_FB_GAUSSIAN, NSOF_FB_USE_INITIAL_FLOW or both */
} nsof_fb_params;

/* One crop pair = one independent Farneback problem (the `prev_region` / `next_region` views of
 * optical_flow_seg.py:186-187).  Pointers address the first pixel of the crop inside its frame;
 * pitches are the frame's row strides in bytes, so a crop is described without copying.
 * flow: (dx, dy) float32 interleaved, the layout cv2 returns; flow_pitch in bytes (multiple of 8). */
typedef struct nsof_crop_desc {
    const uint8_t *prev;
    const uint8_t *next;
    int h, w;
    ptrdiff_t prev_pitch, next_pitch;
    float *flow;
    ptrdiff_t flow_pitch;
} nsof_crop_desc;

NSOF_API const char *nsof_last_error(void);
NSOF_API int nsof_version(void);

/* Creates the per-(GPU, stream) context that owns the workspace (I, R0, R1, M ping-pong, coarse
 * flow, job tables).  Fails with NSOF_ERR_CUDA when `device` is not a usable CUDA device. */
NSOF_API int nsof_create(int device, nsof_handle **out);
NSOF_API int nsof_destroy(nsof_handle *h);
/* Optional: size the workspace once for batches of up to `max_items` crops totalling
 * `total_crop_pixels` pixels (otherwise it grows on demand, which synchronises the device). */
NSOF_API int nsof_reserve(nsof_handle *h, size_t total_crop_pixels, int max_items);
/* Bytes of device workspace currently held. */
NSOF_API size_t nsof_workspace_bytes(const nsof_handle *h);

/* ---- launch accounting ------------------------------------------------------------------- */
/* Kernels launched on this handle since nsof_create (bench.py's gpu_launches). */
NSOF_API unsigned long long nsof_launch_count(const nsof_handle *h);
/* Optional per-launch timing: nsof_profile(h, 1) synchronises, clears the records and brackets every
 * following kernel launch with a CUDA event pair on its stream; nsof_profile(h, 0) stops;
 * nsof_profile_read() synchronises and returns one row per (kernel class, pyramid level) that ran
 * (name "fb_blur_um.L0", ...): launches, summed
 * device time, summed algorithmic bytes (the DESIGN.md byte model).  Used by bench.py's roofline leg. */
typedef struct nsof_kernel_stat {
    char name[32];
    int launches;
    double ms;
    double algorithmic_bytes;
} nsof_kernel_stat;
NSOF_API int nsof_profile(nsof_handle *h, int enable);
NSOF_API int nsof_profile_read(nsof_handle *h, nsof_kernel_stat *out, int max_out, int *n_out);

/* ---- stage 2: dense Farneback ----------------------------------------------------------- */
/* Replaces cv2.calcOpticalFlowFarneback(prev_region, next_region, None, **farneback_params)
 * (optical_flow_seg.py:158, 203, 494; optical_flow_ob.py:225, 270, 616;
 *  optical_flow_prediction.py:161, 206, 570; optical_flow_yolo.py:199, 244, 898)
 * for `n` crop pairs at once; device pointers; asynchronous on `cuda_stream`. */
NSOF_API int nsof_farneback_batch(nsof_handle *h, const nsof_crop_desc *items, int n,
                                  const nsof_fb_params *params, void *cuda_stream);
/* Single pair, device pointers. */
NSOF_API int nsof_farneback_u8(nsof_handle *h, const uint8_t *prev, const uint8_t *next, int height, int width,
                               ptrdiff_t prev_pitch, ptrdiff_t next_pitch, const nsof_fb_params *params,
                               float *flow, ptrdiff_t flow_pitch, void *cuda_stream);
/* Same with HOST pointers: H2D of the crops, kernels, D2H of the flow, synchronous -- the call
 * the Python `calcOpticalFlowFarneback` drop-in makes for numpy arrays. */
NSOF_API int nsof_farneback_batch_host(nsof_handle *h, const nsof_crop_desc *items, int n,
                                       const nsof_fb_params *params);
NSOF_API int nsof_farneback_u8_host(nsof_handle *h, const uint8_t *prev, const uint8_t *next, int height,
                                    int width, ptrdiff_t prev_pitch, ptrdiff_t next_pitch,
                                    const nsof_fb_params *params, float *flow, ptrdiff_t flow_pitch);

/* S2-g, the reference's paste into `flow = np.zeros((h, w, 2))` (optical_flow_seg.py:215, 204): zero the flow
 * rectangles of `n` crops (only .h, .w, .flow, .flow_pitch of each item are read) in ONE launch.  A sequence
 * driver that owns its flow frames clears the previous step's rectangles with this instead of re-zeroing whole
 * frames: outside the rectangles the frames are still zero. */
NSOF_API int nsof_flow_clear(nsof_handle *h, const nsof_crop_desc *items, int n, void *cuda_stream);
/* Strided host<->device copy of `rows` runs of `row_bytes` (cudaMemcpy2DAsync; kind 1 = host to device,
 * 2 = device to host), used by the host-buffer sequence path to land a clip's frames in the time-major device
 * layout.  Pinned host memory keeps it asynchronous. */
NSOF_API int nsof_copy_2d_async(void *dst, ptrdiff_t dst_pitch, const void *src, ptrdiff_t src_pitch,
                                size_t row_bytes, size_t rows, int kind, void *cuda_stream);

/* ---- stage 2 split by rows over several GPUs (BASELINE.json configs[4]: 8K frames, halo-row exchange) -------
 * The reference has no multi-GPU path; this is the B200 design for frames too large to be worth one GPU each.
 * Every rank calls nsof_fb_tiled_begin with the SAME full-frame crop descriptors (its own device pointers; only
 * the source rows its band's stencils reach are read, and only its band of `flow` is written) and then drives the
 * level loop of FarnebackOpticalFlowImpl::calc (A.1) stage by stage:
 *     for level = levels-1 .. 0:
 *         stage 0 (pyramid + PolyExp of the band)      -> exchange `halo_r1` rows of R1 (what = 0) with both
 *                                                          neighbours, and (level < top) 4 rows of the coarser
 *                                                          level's flow (what = 3, level + 1)
 *         stage 1 (first UpdateMatrices)               -> exchange winsize/2 rows of Ma (what = 1)
 *         stage 2, it = 0 .. iterations-1              -> after every iteration but the last: exchange winsize/2
 *                                                          rows of the M buffer just written (Mb, Ma, Mb, ...)
 * Rank r owns rows [r*h_k/world, (r+1)*h_k/world) of level k (nsof_tiled_buf.y_lo / y_hi).  An exchange sends the
 * first / last rows of the band to the upper / lower neighbour and receives theirs into the rows just outside the
 * band, in place (the workspace is full size on every rank).  parallel.TiledFarneback does this over NCCL. */
typedef struct nsof_tiled_buf {
    void *ptr;             /* device pointer of row 0 of plane 0 (rows of the WHOLE level, not of the band) */
    size_t row_floats;     /* floats per row */
    size_t plane_floats;   /* floats between planes */
    int planes;
    int h, w;              /* the level's size */
    int y_lo, y_hi;        /* this rank's band of rows */
} nsof_tiled_buf;
NSOF_API int nsof_fb_tiled_begin(nsof_handle *h, const nsof_crop_desc *items, int n, const nsof_fb_params *params,
                                 int rank, int world, int halo_r1, void *cuda_stream);
NSOF_API int nsof_fb_tiled_levels(const nsof_handle *h);
/* stage: 0 pyramid + PolyExp, 1 first UpdateMatrices, 2 blur + solve iteration `it` (ignored otherwise). */
NSOF_API int nsof_fb_tiled_stage(nsof_handle *h, int level, int stage, int it, void *cuda_stream);
/* The same stage on a sub-window of the band: part 0 = the whole band, 1 / 2 = its top / bottom `halo_r1` rows (all a
 * row neighbour waits for), 3 = the rows in between.  Parts 1 + 2 + 3 write exactly what part 0 writes: start the halo
 * exchange after the edges, compute the interior while the rows travel (tiled.py, overlap=True). */
NSOF_API int nsof_fb_tiled_stage_rows(nsof_handle *h, int level, int stage, int it, int part, void *cuda_stream);
/* what: 0 R1, 1 Ma, 2 Mb, 3 flow of `level` (level 0: the caller's flow buffer). */
NSOF_API int nsof_fb_tiled_query(nsof_handle *h, int item, int level, int what, nsof_tiled_buf *out);
/* *tripped = 1 when an UpdateMatrices gather reached R1 rows beyond the received halo (result invalid: rerun with
 * a larger halo_r1).  Synchronises the stream. */
NSOF_API int nsof_fb_tiled_guard(nsof_handle *h, int *tripped, void *cuda_stream);

/* ---- the same row tiling over PEER MEMORY (NVLink), the whole level loop in one host call -----------------------
 * Every rank plans the whole frame from the same descriptors, so a row of any workspace buffer has the same offset
 * from the workspace base on every rank.  After nsof_fb_tiled_begin a rank exports its workspace and its flag buffer
 * (nsof_fb_tiled_link_export: CUDA IPC handles for ranks in other processes, raw addresses for ranks in this one),
 * the driver hands every rank its row neighbours' exports (nsof_fb_tiled_link_attach), and nsof_fb_tiled_run
 * enqueues every stage of the batch with, after each, ONE kernel that stores the band's edge rows straight into
 * the neighbours' halo rows and raises their flags to the exchange's sequence number, and a one-thread kernel
 * that waits for this rank's own two flags.  No message buffers, no pack / unpack kernels, no library collective and
 * no host round trip between stages.  Ranks must sit on different devices (a waiting kernel and the kernel it waits
 * for cannot share a GPU).  Re-export and re-attach when `arena_bytes` changes (a larger batch re-allocated the
 * workspace).  `first_seq`: the running exchange count, identical on every rank (start at 0, pass *next_seq on). */
NSOF_API int nsof_fb_tiled_link_export(nsof_handle *h, void *arena_ipc64, void *flags_ipc64, void **arena_ptr,
                                       void **flags_ptr, size_t *arena_bytes);
NSOF_API int nsof_fb_tiled_link_attach(nsof_handle *h, const void *up_arena, const void *up_flags, const void *dn_arena,
                                       const void *dn_flags, int use_ipc);
NSOF_API int nsof_fb_tiled_run(nsof_handle *h, int halo_flow, unsigned long long first_seq, unsigned long long *next_seq,
                               void *cuda_stream);

/* Device-to-device copy of `n` byte blocks in ONE launch (halo rows <-> contiguous message buffers of the tiled
 * path: one NCCL message per neighbour instead of one per plane and pair). */
typedef struct nsof_copy_block {
    void *dst;
    const void *src;
    size_t bytes;
} nsof_copy_block;
NSOF_API int nsof_copy_blocks(nsof_handle *h, const nsof_copy_block *blocks, int n, void *cuda_stream);
/* The same with the block table already resident in DEVICE memory (uploaded once by the caller; `biggest_block` and
 * `total_bytes` describe it): no staging slot and no upload kernel per call. */
NSOF_API int nsof_copy_blocks_resident(nsof_handle *h, const nsof_copy_block *device_blocks, int n, size_t biggest_block,
                                       double total_bytes, void *cuda_stream);

/* ---- SSIM of the prediction task (optical_flow_prediction.py:113-115) --------------------
 * skimage.metrics.structural_similarity(a, b, data_range=...) with its defaults for 2-D uint8 images: 7 x 7 uniform
 * window, sample covariance, K1 = 0.01, K2 = 0.03, mean over the image cropped by 3 pixels.  `*_pixel_stride` = bytes
 * between pixels (3 selects one channel of an interleaved BGR frame, as the script's `[:, :, 2]` does).  The score is a
 * float64 written to DEVICE memory (`_host`: to host memory, synchronous). */
NSOF_API int nsof_ssim_u8(nsof_handle *h, const uint8_t *a, ptrdiff_t a_pitch, int a_pixel_stride, const uint8_t *b,
                          ptrdiff_t b_pitch, int b_pixel_stride, int height, int width, double data_range, double *out_dev,
                          void *cuda_stream);
NSOF_API int nsof_ssim_u8_host(nsof_handle *h, const uint8_t *a, ptrdiff_t a_pitch, int a_pixel_stride, const uint8_t *b,
                               ptrdiff_t b_pitch, int b_pixel_stride, int height, int width, double data_range, double *out);

/* ---- JPEG frames straight to grey on the device (optical_flow_seg.py:439-443) --------------
 * `cv2.imread(path)` + `cv2.cvtColor(frame, COLOR_RGB2GRAY)` for `n` JPEG streams in HOST memory, all `height` x `width`:
 * grey frame i lands in DEVICE memory at gray + i * frame_stride (rows `pitch` bytes apart), so a frame crosses PCIe as
 * its JPEG bytes (~0.2 MB at 1080p) instead of decoded pixels.  Entropy decode + IDCT are nvJPEG's (library code, loaded
 * with dlopen on first use: NSOF_ERR_UNSUPPORTED when libnvjpeg is absent); the BGR -> grey step is the exact integer
 * formula of nsof_rgb2gray_u8.  The decoder is not cv2's libjpeg-turbo: grey values agree to a few levels (tests state the
 * bound), not bit for bit.  nsof_jpeg_info: size of one stream (host-side parse). */
NSOF_API int nsof_jpeg_info(nsof_handle *h, const uint8_t *jpeg, size_t bytes, int *height, int *width);
NSOF_API int nsof_jpeg_decode_gray(nsof_handle *h, const uint8_t *const *jpegs, const size_t *bytes, int n, int height,
                                   int width, uint8_t *gray, ptrdiff_t pitch, ptrdiff_t frame_stride, void *cuda_stream);

/* nsof_ssim_u8 for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments and
 * result as nsof_ssim_u8; the per-CTA partial sums live in a file-scope `static __device__` array, so the call makes
 * no cudaMalloc / cudaFree and touches no growable buffer.  Limit: at most 4096 tiles of 32 x 32 interior pixels
 * (1080 x 1920 = 2040 tiles, 2048 x 2048 = 4096); larger images return NSOF_ERR_UNSUPPORTED.  The 32 KB array is held
 * by every GPU that loads the library until exit and is shared by all calls on a device: calls that overlap on the GPU
 * race (serialise them on one stream); it keeps the last call's contents and needs no clearing. */
NSOF_API int nsof_ssim_u8_noalloc(nsof_handle *h, const uint8_t *a, ptrdiff_t a_pitch, int a_pixel_stride, const uint8_t *b,
                                  ptrdiff_t b_pitch, int b_pixel_stride, int height, int width, double data_range,
                                  double *out_dev, void *cuda_stream);

/* Pyramid geometry exactly as the driver computes it (A.1): returns the number of scales and
 * fills hk/wk/ksize/sigma[0 .. scales-1] (any pointer may be NULL).  Pure host arithmetic. */
NSOF_API int nsof_fb_level_geometry(int height, int width, double pyr_scale, int levels, int *hk, int *wk,
                                    int *ksize, double *sigma);
/* Algorithmic HBM bytes of one crop pair under the per-level fusion this library implements
 * (DESIGN.md "byte model"); used by bench.py for the roofline. */
NSOF_API double nsof_fb_algorithmic_bytes(int height, int width, const nsof_fb_params *params);

/* Test hooks (used by tests/ only): truncate the pyramid after `stop_level` with `stop_iters`
 * UpdateFlow iterations on that level (-1, -1 = off), then read back a stage of item `item`:
 * what = 0 I(prev) 1 I(next) 2 R0 3 R1 4 M (current) 5 flow of that level.  `out` is a HOST
 * buffer of ch*h*w floats, planar [ch][h][w] (flow: interleaved [h][w][2]). */
NSOF_API int nsof_debug_truncate(nsof_handle *h, int stop_level, int stop_iters);
/* enable = 0 forces the unfused pyramid + PolyExp kernels, which write I to the workspace (what = 0, 1). */
NSOF_API int nsof_debug_fuse(nsof_handle *h, int enable);
/* enable = 0 keeps every launch of nsof_farneback_batch on the caller's stream (default: the pyramid + PolyExp
 * launches of the finer levels run on internal streams beside the coarse-level iterations). */
NSOF_API int nsof_debug_share_frames(nsof_handle *h, int enable);
NSOF_API int nsof_debug_lanczos_f64(nsof_handle *h, int enable);
NSOF_API int nsof_debug_fork(nsof_handle *h, int enable);
/* enable = 1 keeps the large blur+solve tile on every level (default: latency-bound levels take the small one). */
NSOF_API int nsof_debug_big_tiles(nsof_handle *h, int enable);
/* enable = 1: winsize 3 .. 5 (box window) run the generic blur+solve kernel instead of the tuned kernel's small-radius
 * instantiations (default). */
NSOF_API int nsof_debug_generic_blur(nsof_handle *h, int enable);
/* enable = 2: window in the blur+solve kernel only.  enable = 0: the tuned blur+solve kernel gathers R1 from L2 in every tile (default: from a shared-memory window placed
 * on the tile's flow range whenever the range fits; same values either way), and the first UpdateMatrices of a level runs
 * the round-1 kernel.  nsof_debug_r1_window_stats: out2[0..3] (FOUR counters) = {blur+solve tiles through the window,
 * blur+solve tiles on the L2 path, the same two for the first UpdateMatrices} since the previous call (the first call
 * switches the counting on). */
NSOF_API int nsof_debug_r1_window(nsof_handle *h, int enable);
/* off_mask bit 0 / 1: no bulk L2 prefetch of the tile's R0 / R1 rows in phase 1 of the tuned blur+solve kernel (A/B);
 * -1 restores the default (no prefetch while the R1 window is on). */
NSOF_API int nsof_debug_bs_prefetch(nsof_handle *h, int off_mask);
NSOF_API int nsof_debug_r1_window_stats(nsof_handle *h, unsigned long long *out2, void *cuda_stream);
NSOF_API int nsof_debug_fetch(nsof_handle *h, int item, int what, float *out, int *out_h, int *out_w);

/* ---- stage 1: synaptic temporal memory -> ROI ------------------------------------------- */
/* Rectangle in frame pixels, half-open: rows [y0, y1), cols [x0, x1)  (regions_info tuples of
 * optical_flow_seg.py:138, 180-183 are (x_start, y_start, x_end, y_end)). */
typedef struct nsof_rect {
    int x0, y0, x1, y1;
} nsof_rect;

typedef struct nsof_roi_params {
    int cells_h, cells_w;  /* Hc x Wc: shape of one slice of constructed3DMatrix */
    int frame_h, frame_w;  /* H x W of the grey frames */
    int pixel_h, pixel_w;  /* MEMSIZE (optical_flow_seg.py:455-458 passes it twice) */
    int ext_left, ext_right, ext_upper, ext_lower; /* EXTEND_* (optical_flow_seg.py:38-41) */
    int flag;              /* FLAG: 1 separate regions, 2 merged region (optical_flow_seg.py:31) */
    int thres;             /* THRES on the uint8 map (optical_flow_seg.py:36) */
    double map_a, map_b;   /* u8 = uint8(clip(map_a / log10(I) + map_b, 0, 255)); reference: -3366, -306 */
} nsof_roi_params;

#define NSOF_MAX_RECTS 256

/* Replaces, for `n` independent array states at once, the stage-1 tail of the reference:
 * current -> uint8 map (optical_flow_seg.py:426-435), update_transition_pic (:115-121),
 * cv2.connectedComponentsWithStats(connectivity=4) (:223-224) and the bbox arithmetic of
 * process_merged_region / process_separate_regions (:129-138, :174-183).
 * currents: n slices of cells_h x cells_w doubles (device).  thr_table: 256 doubles (device),
 * thr_table[v] = smallest current whose uint8 map value is >= v, found by the host wrapper by
 * bisection with numpy's own log10 so that the device comparison is bit-exact by construction
 * (thr_table[0] unused).  Outputs (device): memimg n*Hc*Wc uint8 map; labels n*Hc*Wc int32 (cv2
 * numbering: raster order of each component's first cell, 0 = background); n_components[n]
 * (= cv2's num_labels - 1); n_rects[n] (flag 2: 0 or 1; flag 1: n_components, of which at most
 * NSOF_MAX_RECTS are stored); rects[n][NSOF_MAX_RECTS]; stats[n][NSOF_MAX_RECTS][5] = cv2
 * CC_STAT_(LEFT,TOP,WIDTH,HEIGHT,AREA) of components 1.. (row 0 = component 1).
 * Any output pointer may be NULL. */
NSOF_API int nsof_roi_from_state_f64(nsof_handle *h, const double *currents, int n, const nsof_roi_params *p,
                                     const double *thr_table, uint8_t *memimg, int32_t *labels,
                                     int32_t *n_components, int32_t *n_rects, nsof_rect *rects, int32_t *stats,
                                     void *cuda_stream);
/* HOST-pointer variant (copies in/out, synchronous). */
NSOF_API int nsof_roi_from_state_f64_host(nsof_handle *h, const double *currents, int n,
                                          const nsof_roi_params *p, const double *thr_table, uint8_t *memimg,
                                          int32_t *labels, int32_t *n_components, int32_t *n_rects,
                                          nsof_rect *rects, int32_t *stats);

/* Memristor model constants: PARAMS of eventsim/event_mem_sim.py:20-27 == simulation/
 * simulationcode_v4_transistor_vehicle.m:26-33. */
typedef struct nsof_synapse_params {
    double alphaoff, alphaon, voff, von, koff, kon, son, soff, bon, boff, Ron, Roff, won, woff, wini;
} nsof_synapse_params;

/* bincount_2d (eventsim/event_mem_sim.py:100-104): per-slice event histogram, int32, bit-exact.
 * x, y: device int16 arrays of n_events; counts: device int32 H*W, overwritten. */
NSOF_API int nsof_event_hist(nsof_handle *h, const int16_t *x, const int16_t *y, size_t n_events, int H, int W,
                             int32_t *counts, void *cuda_stream);
/* update_state (eventsim/event_mem_sim.py:40-57), float32 arrays of n elements, in place on w:
 * w <- clip(w + dwdt(w, V) * dt, 0, 1) with the float32-demoted constants numpy uses. */
NSOF_API int nsof_synapse_step_f32(nsof_handle *h, float *w, const float *V, size_t n,
                                   const nsof_synapse_params *p, double dt, void *cuda_stream);
/* Scheme 1 of simulate() (eventsim/event_mem_sim.py:204-228) for a run of `n_slices` consecutive
 * time windows, fused: events are time-sorted, slice s owns events [slice_begin[s], slice_begin[s+1]).
 * For each slice: counts = bincount_2d; V = silent_v, V[counts >= theta] = active_v; w = update_state(w, V).
 * w stays in registers across the slices of a chunk; only the final state is written.
 * slice_begin: device int64[n_slices + 1]. */
NSOF_API int nsof_event_scheme1_f32(nsof_handle *h, const int16_t *x, const int16_t *y,
                                    const int64_t *slice_begin, int n_slices, int H, int W, int theta,
                                    float active_v, float silent_v, const nsof_synapse_params *p, double dt,
                                    float *w, void *cuda_stream);
/* Scheme 2 of simulate() (eventsim/event_mem_sim.py:231-286): DC bias + event overlay with a per-pixel refractory
 * period, for a run of `n_slices` consecutive time windows.  Per slice and array: V = bias_v; pixels with a matching
 * event in the slice whose next_ok <= t[first event of the slice] get V = bias_v + active_v and
 * next_ok = t[last event of the slice] + refractory_us (REFRACTORY_US = 800, :34); w = update_state(w, V).
 * magnitude = 0 ('split'): ON events (pol == 1) drive (w_a, next_ok_a), OFF events (pol == 0) drive (w_b, next_ok_b);
 * magnitude = 1: every event drives array A and the B pointers may be NULL.  pol: int8 per event; t: int64 us;
 * next_ok_*: device int64[H*W], zero before the first call, carried between calls. */
NSOF_API int nsof_event_scheme2_f32(nsof_handle *h, const int16_t *x, const int16_t *y, const int8_t *pol,
                                    const int64_t *t, const int64_t *slice_begin, int n_slices, int H, int W,
                                    int magnitude, float active_v, float bias_v, long long refractory_us,
                                    const nsof_synapse_params *p, double dt, float *w_a, float *w_b,
                                    int64_t *next_ok_a, int64_t *next_ok_b, void *cuda_stream);
/* resistance_exp (eventsim/event_mem_sim.py:60-63): R = Ron / exp(-ln(Roff/Ron) * (1 - w)) for float32 w.
 * The result is float64, as in numpy (np.log returns a float64 scalar, which promotes the array). */
NSOF_API int nsof_resistance_exp_f32(nsof_handle *h, const float *w, size_t n, const nsof_synapse_params *p,
                                     double *R, void *cuda_stream);

/* Event-mode ROI glue (builder-defined; the reference never thresholds the eventsim state):
 * currents[cy][cx] = max over the memsize_h x memsize_w pixels of cell (cy, cx) of 1 / resistance_exp(w)
 * (Hc = H / memsize_h, Wc = W / memsize_w, floor), double, ready for nsof_roi_from_state_f64. */
NSOF_API int nsof_event_cell_currents_f32(nsof_handle *h, const float *w, int H, int W, int memsize_h, int memsize_w,
                                          const nsof_synapse_params *p, double *currents, void *cuda_stream);

/* Frame-mode simulator (simulation/simulationcode_v4_transistor_vehicle.m:111-121, 146-236, 306-321):
 * one step of simulate_memristor_array for `n` clips at once.
 *   cells_prev/cells_next: device double [n][Hc][Wc], the Lanczos-3 compressed frames in [0,1]
 *   w: device double [n][Hc][Wc] state, updated in place with nsubsteps Euler sub-steps of dt/nsubsteps
 *   current_out (may be NULL): device double [n][Hc][Wc] = vds / R(w)   (fixture-compatible current)
 */
typedef struct nsof_frame_sim_params {
    double th1, th2; /* vehicle.m:40-41 */
    double dt;       /* vehicle.m:55 */
    int nsubsteps;   /* vehicle.m:56-59 */
    double vds;      /* vehicle.m:36 */
} nsof_frame_sim_params;
NSOF_API int nsof_frame_sim_step_f64(nsof_handle *h, const double *cells_prev, const double *cells_next,
                                     double *w, double *current_out, int n, int Hc, int Wc,
                                     const nsof_synapse_params *p, const nsof_frame_sim_params *fp,
                                     void *cuda_stream);
/* simulate_memristor_array (vehicle.m:187-227) over a whole clip in ONE launch: cells is device double
 * [n_frames][n][Hc][Wc] (time-major, so that time slice t of all n clips is contiguous); w [n][Hc][Wc] is
 * the state, updated in place through all n_frames - 1 pairs; current_out (may be NULL) is
 * [n_frames - 1][n][Hc][Wc], slice t = vds / R(w) after the pair (t, t + 1). */
NSOF_API int nsof_frame_sim_run_f64(nsof_handle *h, const double *cells, int n_frames, int n, int Hc, int Wc,
                                    double *w, double *current_out, const nsof_synapse_params *p,
                                    const nsof_frame_sim_params *fp, void *cuda_stream);
/* compress_image (vehicle.m:111-121): im2double + imresize(., [floor(H/n), floor(W/m)], 'lanczos3')
 * with MATLAB's antialiasing, for `n` uint8 frames (device) of H x W with row pitch `pitch` bytes and
 * frame stride `frame_stride` bytes -> device double [n][Hc][Wc]. */
NSOF_API int nsof_lanczos3_compress_u8(nsof_handle *h, const uint8_t *frames, int n, int H, int W,
                                       ptrdiff_t pitch, ptrdiff_t frame_stride, int Hc, int Wc, double *cells,
                                       void *cuda_stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Next row of the path: the motion-segmentation post-process that consumes the flow.
 *
 * Replaces, per region, `mag, ang = cv2.cartToPolar(flow_region[..., 0], flow_region[..., 1])` followed by
 * `process_flow_region(mag, ang)` (optical_flow_seg.py:285-287, 308-311, 322-357; same code in optical_flow_yolo.py)
 * and the paste `motion_binary[y_min:y_max, x_min:x_max] = ...` (:287).  What that function returns depends only on
 *     motion_mask = 255 where mag > SEG_TH            (:345-346; the HSV / BGR / grey / `binary` images are dead code)
 *     5 x [ dilate, erode ] with getStructuringElement(MORPH_ELLIPSE, (10, 10))   (:349-352)
 *     threshold(., 1, 255, THRESH_BINARY)             (:355)
 * The reference's flow is float64 there (pasted into np.zeros, negated, :215, :461) but holds float32 values, and
 * cv2.cartToPolar computes in float32 regardless: mag = sqrtf(fmaf(x, x, y*y)), compared with mag_thresh in double
 * (pinned by the golden masks of the reference's own function).  Morphology follows cv2: anchor = kernel
 * centre (ksize / 2), pixels outside the image do not take part (dilate ignores them, erode ignores them).
 * Masks are bit-exact.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct nsof_seg_params {
    double mag_thresh;  /* SEG_TH (optical_flow_seg.py:49) */
    int shape;          /* cv2.MORPH_RECT 0, MORPH_CROSS 1, MORPH_ELLIPSE 2 (getStructuringElement) */
    int ksize_w, ksize_h; /* 1..32 each; (10, 10) in the reference */
    int iterations;     /* dilate + erode rounds, 5 in the reference; 0 = threshold only */
} nsof_seg_params;
typedef struct nsof_seg_desc {
    const float *flow;    /* device, h x w x 2 float32 (dx, dy), row pitch flow_pitch BYTES */
    ptrdiff_t flow_pitch;
    int h, w;
    uint8_t *mask;        /* device, h x w uint8 (0 / 255), row pitch mask_pitch BYTES: the region of motion_binary */
    ptrdiff_t mask_pitch;
} nsof_seg_desc;
/* All `n` regions in 2 + 2 * iterations launches (threshold + pack to bit planes, one launch per morphological pass
 * over all regions, unpack).  Asynchronous on `cuda_stream`. */
NSOF_API int nsof_flow_segment_batch(nsof_handle *h, const nsof_seg_desc *items, int n, const nsof_seg_params *params,
                                     void *cuda_stream);
/* `motion_binary = np.zeros((h, w))` of the next frame pair (optical_flow_seg.py:262) for a driver that owns its mask
 * frames: zero the mask rectangles of `n` regions (only .mask, .mask_pitch, .h, .w are read) in one launch. */
NSOF_API int nsof_mask_clear(nsof_handle *h, const nsof_seg_desc *items, int n, void *cuda_stream);
/* One region with HOST pointers (flow in, mask out), synchronous: the call a `process_flow_region` drop-in makes
 * for numpy arrays. */
NSOF_API int nsof_flow_segment_host(nsof_handle *h, const float *flow, ptrdiff_t flow_pitch, int height, int width,
                                    const nsof_seg_params *params, uint8_t *mask, ptrdiff_t mask_pitch);
/* cv2.getStructuringElement(shape, (ksize_w, ksize_h)) as uint8 [ksize_h][ksize_w] (host), for tests and callers
 * that want to look at the element the passes use. */
NSOF_API int nsof_structuring_element(int shape, int ksize_w, int ksize_h, uint8_t *out);

/* ---------------------------------------------------------------------------------------------------------------
 * Frame-side neighbours of the path.
 *
 * nsof_rgb2gray_u8 replaces `cv2.cvtColor(frame, cv2.COLOR_RGB2GRAY)` on the decoded 3-channel frame
 * (optical_flow_seg.py:442-443; same lines in _ob / _prediction / _yolo): OpenCV's 8-bit formula
 * (c0 * 9798 + c1 * 19235 + c2 * 3735 + 16384) >> 15 on the channels AS STORED (the scripts pass a BGR array).
 * `n` frames of H x W x 3 (row pitch / frame stride in bytes) -> n frames of H x W; device pointers, asynchronous.
 * Decoding (cv2.imread) stays on the host.  Bit-exact.
 * ------------------------------------------------------------------------------------------------------------- */
NSOF_API int nsof_rgb2gray_u8(nsof_handle *h, const uint8_t *src, ptrdiff_t src_pitch, ptrdiff_t src_frame_stride, int n,
                              int H, int W, uint8_t *dst, ptrdiff_t dst_pitch, ptrdiff_t dst_frame_stride, void *cuda_stream);
NSOF_API int nsof_rgb2gray_u8_host(nsof_handle *h, const uint8_t *src, ptrdiff_t src_pitch, int H, int W, uint8_t *dst,
                                   ptrdiff_t dst_pitch);
/* The prediction task (optical_flow_prediction.py:281-352, 257-361): per region
 *     flow_map = (x, y) + flow_region  (formed in double, stored float32)
 *     new_frame[y0:y1, x0:x1, c] = cv2.remap(next_frame[:, :, c], map_x, map_y, INTER_LINEAR, borderMode=BORDER_REPLICATE)
 * with OpenCV's 8-bit fixed-point bilinear (coordinates rounded to 1/32 px, 15-bit weights, taps clamped).
 * `frame` is the whole source image (H x W x channels, interleaved), `flow` the region's flow (h x w x 2 float32),
 * `out` the region of the destination image; flow_sign -1 applies the reference's `flow = -flow` first.  Bit-exact. */
typedef struct nsof_warp_desc {
    const uint8_t *frame; ptrdiff_t frame_pitch; int H, W, channels;
    const float *flow; ptrdiff_t flow_pitch;
    int x0, y0, h, w;
    uint8_t *out; ptrdiff_t out_pitch;
} nsof_warp_desc;
NSOF_API int nsof_remap_flow_batch(nsof_handle *h, const nsof_warp_desc *items, int n, int flow_sign, void *cuda_stream);
NSOF_API int nsof_remap_flow_host(nsof_handle *h, const uint8_t *frame, ptrdiff_t frame_pitch, int H, int W, int channels,
                                  const float *flow, ptrdiff_t flow_pitch, int x0, int y0, int height, int width,
                                  int flow_sign, uint8_t *out, ptrdiff_t out_pitch);

#ifdef __cplusplus
}
#endif
#endif /* NSOF_H */
