// nsof_rt.h -- kernel/launch vocabulary shared by every .cu file of libnsof.
//
// The product build (nvcc, sm_100a) maps these macros 1:1 onto CUDA.  A second,
// TEST-ONLY build (g++ -DNSOF_EMU, see tests/emu/) maps them onto a sequential
// "one logical thread per block" executor so that the *same kernel source* can be
// checked for indexing / border / tiling logic against the oracle on a box with
// no GPU.  That build is never loaded by the package; it exists for
// `pytest -m "not gpu"` only.  To make it possible, kernels are written in the
// block-stride style:
//     for (int i = NSOF_TID; i < n_work_items; i += NSOF_NT) { ... }
//     NSOF_SYNC();
// and never assume a particular thread count for correctness.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <math.h>

#ifdef NSOF_EMU
// ------------------------------------------------------------------ emulator
#include <stdlib.h>
#include <string.h>
#include <vector>
struct float2 { float x, y; };
struct float4 { float x, y, z, w; };
struct uint2 { unsigned x, y; };
static inline uint2 make_uint2(unsigned x, unsigned y) { uint2 r = {x, y}; return r; }
static inline float2 make_float2(float x, float y) { float2 r = {x, y}; return r; }
static inline float4 make_float4(float x, float y, float z, float w) { float4 r = {x, y, z, w}; return r; }
namespace nsof_emu {
extern thread_local int bx, by, bz, gdx, gdy;
extern thread_local unsigned char *smem;
template <class F> static inline void launch(int gx, int gy, int gz, size_t smem_bytes, F f)
{
    std::vector<unsigned char> buf(smem_bytes + 64);
    smem = buf.data();
    gdx = gx;
    gdy = gy;
    for (bz = 0; bz < gz; bz++)
        for (by = 0; by < gy; by++)
            for (bx = 0; bx < gx; bx++) f();
    smem = nullptr;
}
}  // namespace nsof_emu
#define NSOF_GLOBAL static void
#define NSOF_DEV static inline
#define NSOF_HD static inline
#define NSOF_RESTRICT
#define NSOF_BX (nsof_emu::bx)
#define NSOF_BY (nsof_emu::by)
#define NSOF_BZ (nsof_emu::bz)
#define NSOF_TID 0
#define NSOF_NT 1
#define NSOF_WARP 1
static inline int nsof_grid_x() { return nsof_emu::gdx; }
static inline int nsof_grid_y() { return nsof_emu::gdy; }
static inline double nsof_warp_sum_f64(double v) { return v; }
#define NSOF_SYNC() ((void)0)
#define NSOF_DYN_SMEM(name) unsigned char *name = nsof_emu::smem
#define NSOF_LDG(p) (*(p))
#define NSOF_LAUNCH(kern, gx, gy, gz, nt, smem_bytes, stream, ...) \
    nsof_emu::launch((gx), (gy), (gz), (smem_bytes), [&] { kern(__VA_ARGS__); })
#define NSOF_LAUNCH_BOUNDS(nt, minb)
#else
// ---------------------------------------------------------------- CUDA build
#include <cuda_runtime.h>
#define NSOF_GLOBAL __global__ void
#define NSOF_DEV __device__ __forceinline__
#define NSOF_HD __host__ __device__ __forceinline__
#define NSOF_RESTRICT __restrict__
#define NSOF_BX ((int)blockIdx.x)
#define NSOF_BY ((int)blockIdx.y)
#define NSOF_BZ ((int)blockIdx.z)
#define NSOF_TID ((int)threadIdx.x)
#define NSOF_NT ((int)blockDim.x)
#define NSOF_WARP 32
__device__ __forceinline__ int nsof_grid_x() { return (int)gridDim.x; }
__device__ __forceinline__ int nsof_grid_y() { return (int)gridDim.y; }
// fixed-order butterfly: deterministic for a given lane assignment
__device__ __forceinline__ double nsof_warp_sum_f64(double v)
{
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
#define NSOF_SYNC() __syncthreads()
#define NSOF_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#define NSOF_LDG(p) __ldg(p)
#define NSOF_LAUNCH(kern, gx, gy, gz, nt, smem_bytes, stream, ...) \
    kern<<<dim3((gx), (gy), (gz)), dim3(nt), (smem_bytes), (stream)>>>(__VA_ARGS__)
#define NSOF_LAUNCH_BOUNDS(nt, minb) __launch_bounds__(nt, minb)
#endif

// The CTA's job descriptor, staged once in shared memory.  With flat tile lists the descriptor's address depends on a
// loaded index (not on blockIdx), so a by-value copy read lazily from global memory pins the address and the live fields
// in registers for the whole kernel (k_fb_um_first: 64 -> 72 registers, its fourth CTA per SM); shared-memory fields
// have a constant address and can be re-read wherever they are needed.  T: 16-byte aligned, sizeof(T) % 16 == 0.
#ifdef NSOF_EMU
#define NSOF_JOB_TO_SMEM(T, J, src) \
    const T J##_copy = *(src);      \
    const T &J = J##_copy
#else
#define NSOF_JOB_TO_SMEM(T, J, src)                                                                             \
    __shared__ __align__(16) T J##_s;                                                                           \
    static_assert(sizeof(T) % 16 == 0, "job descriptors are copied as 16-byte words");                          \
    for (int i_ = (int)threadIdx.x; i_ < (int)(sizeof(T) / 16); i_ += (int)blockDim.x)                          \
        ((int4 *)&J##_s)[i_] = __ldg((const int4 *)(src) + i_);                                                 \
    __syncthreads();                                                                                            \
    const T &J = J##_s
#endif

// Small non-negative integer (< 2^23) -> float without the conversion pipe: the integer is planted in the mantissa of
// 2^23 and 2^23 is subtracted (exact).  I2F runs on the quarter-rate XU pipe; this is one integer add / byte permute and
// one FADD.  nsof_u8_to_f(word, k): byte k of a packed word.
#ifdef NSOF_EMU
static inline float nsof_small_int_to_f(int v) { return (float)v; }
static inline float nsof_u8_to_f(uint32_t word, int k) { return (float)((word >> (8 * k)) & 255u); }
#else
__device__ __forceinline__ float nsof_small_int_to_f(int v) { return __int_as_float(0x4B000000 + v) - 8388608.f; }
__device__ __forceinline__ float nsof_u8_to_f(uint32_t word, int k)
{
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7540u + (unsigned)k)) - 8388608.f;
}
#endif

NSOF_HD int nsof_imin(int a, int b) { return a < b ? a : b; }
NSOF_HD int nsof_imax(int a, int b) { return a > b ? a : b; }
NSOF_HD int nsof_clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
// cv::BORDER_REFLECT_101 (gfedcb|abcdefgh|gfedcba)
NSOF_HD int nsof_reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// ---------------------------------------------------------------------------------------------
// Bulk asynchronous row copies global -> shared (cp.async.bulk, the non-tensor TMA path: SASS UBLKCP)
// completing on an mbarrier.  Rows of a halo tile are contiguous runs of 16-byte multiples, so no
// tensor map is needed; border replication is fixed up in shared memory afterwards.
// The emulator build copies synchronously.
// ---------------------------------------------------------------------------------------------
#ifdef NSOF_EMU
typedef unsigned long long nsof_mbar_t;
static inline void nsof_mbar_init(nsof_mbar_t *, int) {}
static inline void nsof_mbar_expect_tx(nsof_mbar_t *, unsigned) {}
static inline void nsof_bulk_g2s(void *dst, const void *src, unsigned bytes, nsof_mbar_t *) { memcpy(dst, src, bytes); }
static inline void nsof_mbar_wait(nsof_mbar_t *, unsigned) {}
static inline void nsof_prefetch_l2(const void *, unsigned) {}
static inline void nsof_fence_proxy_async() {}
// A kernel whose threads keep per-thread state across a block barrier: the one emulated thread of a block walks the NT
// thread slots (slot index ts for its state arrays); on the device every thread is its own slot 0.
#define NSOF_FOR_SLOTS(t, ts, NT) for (int t = 0, ts = 0; t < (NT); t++, ts++)
static inline void nsof_count(unsigned *p) { ++*p; }
// block-wide minimum / maximum into four shared integers {min a, max a, min b, max b} (one emulated thread per block)
static inline void nsof_block_minmax4(int *s, int mna, int mxa, int mnb, int mxb)
{
    if (mna < s[0]) s[0] = mna;
    if (mxa > s[1]) s[1] = mxa;
    if (mnb < s[2]) s[2] = mnb;
    if (mxb > s[3]) s[3] = mxb;
}
#define NSOF_LANE 0
#define NSOF_SYNCWARP() ((void)0)
static inline float2 nsof_add2(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
static inline float2 nsof_mul2(float2 a, float2 b) { return make_float2(a.x * b.x, a.y * b.y); }
static inline float2 nsof_fma2(float2 a, float2 b, float2 c) { return make_float2(a.x * b.x + c.x, a.y * b.y + c.y); }
#else
typedef unsigned long long nsof_mbar_t;
__device__ __forceinline__ unsigned nsof_smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void nsof_mbar_init(nsof_mbar_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(nsof_smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void nsof_mbar_expect_tx(nsof_mbar_t *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(nsof_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void nsof_bulk_g2s(void *dst, const void *src, unsigned bytes, nsof_mbar_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     nsof_smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(nsof_smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void nsof_mbar_wait(nsof_mbar_t *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "NSOF_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra NSOF_MBAR_DONE;\n"
        "bra NSOF_MBAR_WAIT;\n"
        "NSOF_MBAR_DONE:\n"
        "}\n" ::"r"(nsof_smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// bulk L2 prefetch of a 16-byte aligned run (cp.async.bulk.prefetch: no register, no LSU slot, no completion)
__device__ __forceinline__ void nsof_prefetch_l2(const void *src, unsigned bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
// generic-proxy accesses of shared memory (before the last barrier) -> bulk async copies into the same bytes
__device__ __forceinline__ void nsof_fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void nsof_count(unsigned *p) { atomicAdd(p, 1u); }
#define NSOF_FOR_SLOTS(t, ts, NT) for (int t = (int)threadIdx.x, ts = 0; ts < 1; ts++)
#define NSOF_LANE ((int)(threadIdx.x & 31))
#define NSOF_SYNCWARP() __syncwarp()
// block-wide minimum / maximum into four shared integers {min a, max a, min b, max b}: REDUX per warp, then one
// shared-memory atomic per warp and value.  Every thread of the block calls it (converged).
__device__ __forceinline__ void nsof_block_minmax4(int *s, int mna, int mxa, int mnb, int mxb)
{
    mna = __reduce_min_sync(0xffffffffu, mna);
    mxa = __reduce_max_sync(0xffffffffu, mxa);
    mnb = __reduce_min_sync(0xffffffffu, mnb);
    mxb = __reduce_max_sync(0xffffffffu, mxb);
    if ((threadIdx.x & 31) == 0) {
        atomicMin(s + 0, mna);
        atomicMax(s + 1, mxa);
        atomicMin(s + 2, mnb);
        atomicMax(s + 3, mxb);
    }
}
// Packed FP32 add of Blackwell (add.rn.f32x2 -> SASS FADD2): two IEEE round-to-nearest additions in one issue slot,
// per component identical to a scalar FADD.  For kernels that are bound by instruction issue, not by the FP32 pipe.
#ifndef NSOF_PACKED_F32
#define NSOF_PACKED_F32 1
#endif
__device__ __forceinline__ float2 nsof_add2(float2 a, float2 b)
{
#if NSOF_PACKED_F32
    return __fadd2_rn(a, b);
#else
    return make_float2(a.x + b.x, a.y + b.y);
#endif
}
// FMUL2 / FFMA2: per component the same as __fmul_rn / __fmaf_rn
__device__ __forceinline__ float2 nsof_mul2(float2 a, float2 b)
{
#if NSOF_PACKED_F32
    return __fmul2_rn(a, b);
#else
    return make_float2(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y));
#endif
}
__device__ __forceinline__ float2 nsof_fma2(float2 a, float2 b, float2 c)
{
#if NSOF_PACKED_F32
    return __ffma2_rn(a, b, c);
#else
    return make_float2(__fmaf_rn(a.x, b.x, c.x), __fmaf_rn(a.y, b.y, c.y));
#endif
}
#endif
