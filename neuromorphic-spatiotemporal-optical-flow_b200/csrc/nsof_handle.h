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

struct nsof_handle {
    int device = 0;
    NsofDevBuf arena;      // I, R, M, coarse flow of every item of the current batch
    NsofDevBuf tables;     // job tables + Gaussian taps on the device
    NsofDevBuf io;         // device staging of the *_host entry points (crops in, flow out)
    NsofDevBuf s1;         // stage-1 scratch
    NsofDevBuf lz;         // Lanczos-3 weight/index tables (device), cached for lz_key
    int lz_key[4] = {0, 0, 0, 0};
    int lz_P[2] = {0, 0};
    NsofStageSlot stage[4];
    int stage_next = 0;
    // debug truncation (tests only)
    int dbg_level = -1, dbg_iters = -1;
    // geometry of the last batch, kept for nsof_debug_fetch
    std::vector<FbItemPlan> *last_plan = nullptr;
    int last_stop_level = 0;
    nsof_fb_params last_params = {};
    bool smem_configured = false;
};

int nsof_stage_acquire(nsof_handle *h, size_t bytes, NsofStageSlot **slot);
