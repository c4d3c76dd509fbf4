// decode.cu -- JPEG frames straight to grey on the device (optical_flow_seg.py:439-443:
// `cv2.imread` -> `cv2.cvtColor(frame, COLOR_RGB2GRAY)` on the BGR array).
//
// The entropy decode + IDCT is NVIDIA's nvJPEG (library code, like calling cuBLAS: it is not a kernel of this path), loaded
// at run time with dlopen so that libnsof has no hard dependency on it; what is ours is the boundary (host JPEG bytes in,
// pitched grey frames in HBM out) and the grey conversion, which is the exact integer formula of frames.cu on nvJPEG's
// interleaved BGR output.  A frame then crosses PCIe as ~0.2 MB of JPEG instead of 2 MB of grey pixels.
// PARITY: the grey formula is bit-exact; the JPEG decoder is NOT the reference's (cv2.imread = libjpeg-turbo: islow IDCT,
// fancy chroma upsampling) -- decoders agree to a few grey levels, the test states the measured bound.
#include "nsof_handle.h"

#include <cstring>
#include <mutex>

#ifndef NSOF_EMU
#include <dlfcn.h>
#include <nvjpeg.h>

struct NvJpegApi {
    void *so = nullptr;
    nvjpegStatus_t (*CreateSimple)(nvjpegHandle_t *) = nullptr;
    nvjpegStatus_t (*Destroy)(nvjpegHandle_t) = nullptr;
    nvjpegStatus_t (*StateCreate)(nvjpegHandle_t, nvjpegJpegState_t *) = nullptr;
    nvjpegStatus_t (*StateDestroy)(nvjpegJpegState_t) = nullptr;
    nvjpegStatus_t (*GetImageInfo)(nvjpegHandle_t, const unsigned char *, size_t, int *, nvjpegChromaSubsampling_t *, int *, int *) = nullptr;
    nvjpegStatus_t (*Decode)(nvjpegHandle_t, nvjpegJpegState_t, const unsigned char *, size_t, nvjpegOutputFormat_t, nvjpegImage_t *,
                             cudaStream_t) = nullptr;
    bool ok = false;
};
static NvJpegApi g_nvjpeg;
static std::once_flag g_nvjpeg_once;

static void nvjpeg_load()
{
    const char *names[] = {"libnvjpeg.so.12", "libnvjpeg.so", "/usr/local/cuda/lib64/libnvjpeg.so.12", "/usr/local/cuda/lib64/libnvjpeg.so"};
    for (const char *n : names) {
        g_nvjpeg.so = dlopen(n, RTLD_NOW | RTLD_LOCAL);
        if (g_nvjpeg.so) break;
    }
    if (!g_nvjpeg.so) return;
#define NSOF_SYM(field, name) *(void **)(&g_nvjpeg.field) = dlsym(g_nvjpeg.so, name)
    NSOF_SYM(CreateSimple, "nvjpegCreateSimple");
    NSOF_SYM(Destroy, "nvjpegDestroy");
    NSOF_SYM(StateCreate, "nvjpegJpegStateCreate");
    NSOF_SYM(StateDestroy, "nvjpegJpegStateDestroy");
    NSOF_SYM(GetImageInfo, "nvjpegGetImageInfo");
    NSOF_SYM(Decode, "nvjpegDecode");
#undef NSOF_SYM
    g_nvjpeg.ok = g_nvjpeg.CreateSimple && g_nvjpeg.Destroy && g_nvjpeg.StateCreate && g_nvjpeg.StateDestroy &&
                  g_nvjpeg.GetImageInfo && g_nvjpeg.Decode;
}

// nvjpegDecode is asynchronous on the stream and a decoder state owns the pinned staging of its image: a state may
// only be reused once the work queued with it has finished.  Two states alternate, each with an event recorded behind
// its decode; the host waits for that event before the state's next use (the entropy decode of frame i + 1 then runs
// under the device part of frame i).
#define NSOF_JPEG_STATES 2
struct NsofJpeg {
    nvjpegHandle_t handle = nullptr;
    nvjpegJpegState_t state[NSOF_JPEG_STATES] = {nullptr, nullptr};
    cudaEvent_t done[NSOF_JPEG_STATES] = {nullptr, nullptr};
    bool busy[NSOF_JPEG_STATES] = {false, false};
    int next = 0;
};
#endif


void nsof_jpeg_release(nsof_handle *h)
{
#ifndef NSOF_EMU
    NsofJpeg *j = (NsofJpeg *)h->jpeg;
    if (!j) return;
    for (int k = 0; k < NSOF_JPEG_STATES; k++) {
        if (j->done[k]) { cudaEventSynchronize(j->done[k]); cudaEventDestroy(j->done[k]); }
        if (j->state[k]) g_nvjpeg.StateDestroy(j->state[k]);
    }
    if (j->handle) g_nvjpeg.Destroy(j->handle);
    delete j;
#endif
    h->jpeg = nullptr;
}

#ifndef NSOF_EMU
static int jpeg_state(nsof_handle *h, NsofJpeg **out)
{
    std::call_once(g_nvjpeg_once, nvjpeg_load);
    if (!g_nvjpeg.ok) NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "jpeg: libnvjpeg could not be loaded (%s)", g_nvjpeg.so ? "missing symbols" : "dlopen failed");
    if (!h->jpeg) {
        NsofJpeg *j = new NsofJpeg();
        bool ok = g_nvjpeg.CreateSimple(&j->handle) == NVJPEG_STATUS_SUCCESS;
        for (int k = 0; ok && k < NSOF_JPEG_STATES; k++)
            ok = g_nvjpeg.StateCreate(j->handle, &j->state[k]) == NVJPEG_STATUS_SUCCESS &&
                 cudaEventCreateWithFlags(&j->done[k], cudaEventDisableTiming) == cudaSuccess;
        if (!ok) {
            h->jpeg = j;
            nsof_jpeg_release(h);
            NSOF_FAIL(NSOF_ERR_CUDA, "jpeg: nvjpegCreateSimple / nvjpegJpegStateCreate failed");
        }
        h->jpeg = j;
    }
    *out = (NsofJpeg *)h->jpeg;
    return NSOF_OK;
}
#endif

// height / width of a JPEG (host bytes).  Host arithmetic inside nvJPEG; no device work.
extern "C" int nsof_jpeg_info(nsof_handle *h, const uint8_t *jpeg, size_t bytes, int *height, int *width)
{
    if (!h || !jpeg || !bytes || !height || !width) NSOF_FAIL(NSOF_ERR_BAD_ARG, "jpeg_info: bad argument");
#ifdef NSOF_EMU
    NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "jpeg: no nvJPEG in the emulator build");
#else
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    NsofJpeg *j;
    int rc = jpeg_state(h, &j);
    if (rc) return rc;
    int comps = 0, hs[NVJPEG_MAX_COMPONENT], ws[NVJPEG_MAX_COMPONENT];
    nvjpegChromaSubsampling_t ss;
    if (g_nvjpeg.GetImageInfo(j->handle, jpeg, bytes, &comps, &ss, ws, hs) != NVJPEG_STATUS_SUCCESS)
        NSOF_FAIL(NSOF_ERR_BAD_ARG, "jpeg_info: not a JPEG stream nvJPEG can parse (cv2.imread would return None)");
    *height = hs[0];
    *width = ws[0];
    return NSOF_OK;
#endif
}

// `n` JPEG streams in HOST memory (all height x width) -> grey frames in DEVICE memory, frame i at
// gray + i * frame_stride, rows `pitch` bytes apart: cv2.imread + cv2.cvtColor(COLOR_RGB2GRAY) of the scripts.
// Asynchronous on the stream after the host-side entropy decode of each frame.
extern "C" int nsof_jpeg_decode_gray(nsof_handle *h, const uint8_t *const *jpegs, const size_t *bytes, int n, int height, int width,
                                     uint8_t *gray, ptrdiff_t pitch, ptrdiff_t frame_stride, void *cuda_stream)
{
    if (!h || n < 0 || (n > 0 && (!jpegs || !bytes || !gray))) NSOF_FAIL(NSOF_ERR_BAD_ARG, "jpeg_decode: bad argument");
    if (height < 1 || width < 1 || pitch < width) NSOF_FAIL(NSOF_ERR_BAD_ARG, "jpeg_decode: bad geometry");
    if (n == 0) return NSOF_OK;
#ifdef NSOF_EMU
    NSOF_FAIL(NSOF_ERR_UNSUPPORTED, "jpeg: no nvJPEG in the emulator build");
#else
    NSOF_CUDA_TRY(cudaSetDevice(h->device));
    NsofJpeg *j;
    int rc = jpeg_state(h, &j);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    const size_t bpitch = ((size_t)width * 3 + 255) & ~(size_t)255, bframe = bpitch * height;
    rc = h->jpeg_buf.ensure(bframe * n);
    if (rc) return rc;
    for (int i = 0; i < n; i++) {
        int comps = 0, hs[NVJPEG_MAX_COMPONENT], ws[NVJPEG_MAX_COMPONENT];
        nvjpegChromaSubsampling_t ss;
        if (!jpegs[i] || g_nvjpeg.GetImageInfo(j->handle, jpegs[i], bytes[i], &comps, &ss, ws, hs) != NVJPEG_STATUS_SUCCESS)
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "jpeg_decode: stream %d is not a JPEG nvJPEG can parse", i);
        if (hs[0] != height || ws[0] != width)
            NSOF_FAIL(NSOF_ERR_BAD_ARG, "jpeg_decode: stream %d is %d x %d, expected %d x %d", i, hs[0], ws[0], height, width);
        nvjpegImage_t out;
        memset(&out, 0, sizeof(out));
        out.channel[0] = (unsigned char *)h->jpeg_buf.p + bframe * i;
        out.pitch[0] = bpitch;
        // BGR interleaved = the array cv2.imread returns (a grey JPEG is replicated into the three channels, as imread does)
        const int k = j->next;
        j->next = (k + 1) % NSOF_JPEG_STATES;
        if (j->busy[k]) NSOF_CUDA_TRY(cudaEventSynchronize(j->done[k]));
        const nvjpegStatus_t st = g_nvjpeg.Decode(j->handle, j->state[k], jpegs[i], bytes[i], NVJPEG_OUTPUT_BGRI, &out, stream);
        if (st != NVJPEG_STATUS_SUCCESS) NSOF_FAIL(NSOF_ERR_CUDA, "jpeg_decode: nvjpegDecode failed on stream %d (status %d)", i, (int)st);
        NSOF_CUDA_TRY(cudaEventRecord(j->done[k], stream));
        j->busy[k] = true;
    }
    return nsof_rgb2gray_u8(h, (const uint8_t *)h->jpeg_buf.p, (ptrdiff_t)bpitch, (ptrdiff_t)bframe, n, height, width, gray, pitch,
                            frame_stride, cuda_stream);
#endif
}
