// Device-side JPEG decode of the camera's side-by-side capture frames.
//
// The reference's camera delivers MJPG, 1280x480, left | right in one frame (usb_test.py:7-11); cv2.VideoCapture decodes every
// frame on the CPU and FastACVNet_plus_Algorithm.cpp:116-131 then splits it with two clone()s.  Here the compressed frames go
// to the GPU as they are and nvJPEG decodes them into the handle's capture buffer (BGR, interleaved, 6W bytes per row), which
// stage A reads in place: left eye = the buffer, right eye = the buffer + 3W bytes, both with the doubled pitch.
//
// nvJPEG ships with the CUDA toolkit.  It is loaded with dlopen on first use, so that libstereo_geom.so itself depends on
// nothing but the driver: a box without libnvjpeg loses this entry point (clear error), not the library.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>

#include <mutex>
#include <thread>
#include <vector>

#include <nvjpeg.h>

#include "common.h"

namespace sa {

namespace {

struct NvJpegApi {
    void* dl = nullptr;
    nvjpegStatus_t (*CreateEx)(nvjpegBackend_t, nvjpegDevAllocator_t*, nvjpegPinnedAllocator_t*, unsigned int, nvjpegHandle_t*) = nullptr;
    nvjpegStatus_t (*CreateSimple)(nvjpegHandle_t*) = nullptr;
    nvjpegStatus_t (*Destroy)(nvjpegHandle_t) = nullptr;
    nvjpegStatus_t (*StateCreate)(nvjpegHandle_t, nvjpegJpegState_t*) = nullptr;
    nvjpegStatus_t (*StateDestroy)(nvjpegJpegState_t) = nullptr;
    nvjpegStatus_t (*GetImageInfo)(nvjpegHandle_t, const unsigned char*, size_t, int*, nvjpegChromaSubsampling_t*, int*, int*) = nullptr;
    nvjpegStatus_t (*Decode)(nvjpegHandle_t, nvjpegJpegState_t, const unsigned char*, size_t, nvjpegOutputFormat_t, nvjpegImage_t*, cudaStream_t) = nullptr;
    nvjpegStatus_t (*BatchedInit)(nvjpegHandle_t, nvjpegJpegState_t, int, int, nvjpegOutputFormat_t) = nullptr;
    nvjpegStatus_t (*Batched)(nvjpegHandle_t, nvjpegJpegState_t, const unsigned char* const*, const size_t*, nvjpegImage_t*, cudaStream_t) = nullptr;
    std::string why;   // why it could not be loaded
};

// one per process; never unloaded (handles of several GPUs share it)
const NvJpegApi& api()
{
    static NvJpegApi a;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnvjpeg.so.12", "libnvjpeg.so", "/usr/local/cuda/lib64/libnvjpeg.so.12", "/usr/local/cuda/lib64/libnvjpeg.so",
                               "/usr/local/cuda/targets/x86_64-linux/lib/libnvjpeg.so.12"};
        for (const char* n : names) {
            a.dl = dlopen(n, RTLD_NOW | RTLD_LOCAL);
            if (a.dl) break;
        }
        if (!a.dl) {
            a.why = "libnvjpeg.so.12 not found (dlopen)";
            return;
        }
        bool ok = true;
        auto sym = [&](const char* name) {
            void* p = dlsym(a.dl, name);
            if (!p) { ok = false; a.why = std::string("libnvjpeg lacks ") + name; }
            return p;
        };
        a.CreateEx = reinterpret_cast<decltype(a.CreateEx)>(sym("nvjpegCreateEx"));
        a.CreateSimple = reinterpret_cast<decltype(a.CreateSimple)>(sym("nvjpegCreateSimple"));
        a.Destroy = reinterpret_cast<decltype(a.Destroy)>(sym("nvjpegDestroy"));
        a.StateCreate = reinterpret_cast<decltype(a.StateCreate)>(sym("nvjpegJpegStateCreate"));
        a.StateDestroy = reinterpret_cast<decltype(a.StateDestroy)>(sym("nvjpegJpegStateDestroy"));
        a.GetImageInfo = reinterpret_cast<decltype(a.GetImageInfo)>(sym("nvjpegGetImageInfo"));
        a.Decode = reinterpret_cast<decltype(a.Decode)>(sym("nvjpegDecode"));
        a.BatchedInit = reinterpret_cast<decltype(a.BatchedInit)>(sym("nvjpegDecodeBatchedInitialize"));
        a.Batched = reinterpret_cast<decltype(a.Batched)>(sym("nvjpegDecodeBatched"));
        if (!ok) { dlclose(a.dl); a.dl = nullptr; }
    });
    return a;
}

const char* status_text(nvjpegStatus_t st)
{
    switch (st) {
        case NVJPEG_STATUS_SUCCESS: return "success";
        case NVJPEG_STATUS_NOT_INITIALIZED: return "not initialized";
        case NVJPEG_STATUS_INVALID_PARAMETER: return "invalid parameter";
        case NVJPEG_STATUS_BAD_JPEG: return "bad JPEG";
        case NVJPEG_STATUS_JPEG_NOT_SUPPORTED: return "JPEG not supported";
        case NVJPEG_STATUS_ALLOCATOR_FAILURE: return "allocator failure";
        case NVJPEG_STATUS_EXECUTION_FAILED: return "execution failed";
        case NVJPEG_STATUS_ARCH_MISMATCH: return "architecture mismatch";
        case NVJPEG_STATUS_INTERNAL_ERROR: return "internal error";
        case NVJPEG_STATUS_IMPLEMENTATION_NOT_SUPPORTED: return "implementation not supported";
        case NVJPEG_STATUS_INCOMPLETE_BITSTREAM: return "incomplete bitstream";
        default: return "unknown status";
    }
}

}  // namespace

struct JpegDecoder {
    nvjpegHandle_t lib = nullptr;
    nvjpegJpegState_t state = nullptr;
    int batch = 0;          // size the batched decoder is initialised for (0 = not)
    bool batched_ok = true; // cleared when the batched decoder refused a stream this handle was given: single-image calls from then on
};

void jpeg_release(sa_handle* h)
{
    if (!h || !h->jpeg) return;
    const NvJpegApi& a = api();
    if (a.dl) {
        if (h->jpeg->state) a.StateDestroy(h->jpeg->state);
        if (h->jpeg->lib) a.Destroy(h->jpeg->lib);
    }
    delete h->jpeg;
    h->jpeg = nullptr;
}

// n JPEG streams of `width` x `height` pixels -> BGR interleaved frames at dst (device), `pitch` bytes per row, `frame` bytes apart; queued on s
int jpeg_decode_batch(sa_handle* h, int n, const void* const* jpeg, const size_t* bytes, uint8_t* dst, size_t pitch, size_t frame, int width, int height,
                      cudaStream_t s)
{
    const NvJpegApi& a = api();
    if (!a.dl) return fail(h, SA_ERR_IO, "JPEG decode unavailable: %s", a.why.c_str());
    if (!h->jpeg) {
        JpegDecoder* d = new JpegDecoder;
        // GPU-assisted Huffman decoding for large batches of baseline streams (what a camera's MJPG is); plain hybrid if the build lacks it
        // (SA_NVJPEG_BACKEND = hardware | hybrid | default overrides, for experiments)
        nvjpegBackend_t be = NVJPEG_BACKEND_GPU_HYBRID;
        if (const char* e = getenv("SA_NVJPEG_BACKEND")) {
            if (!strcmp(e, "hardware")) be = NVJPEG_BACKEND_HARDWARE;
            else if (!strcmp(e, "hybrid")) be = NVJPEG_BACKEND_HYBRID;
            else if (!strcmp(e, "default")) be = NVJPEG_BACKEND_DEFAULT;
        }
        nvjpegStatus_t st = a.CreateEx(be, nullptr, nullptr, 0, &d->lib);
        if (st != NVJPEG_STATUS_SUCCESS) st = a.CreateSimple(&d->lib);
        if (st == NVJPEG_STATUS_SUCCESS) st = a.StateCreate(d->lib, &d->state);
        if (st != NVJPEG_STATUS_SUCCESS) {
            if (d->lib) a.Destroy(d->lib);
            delete d;
            return fail(h, SA_ERR_CUDA, "nvJPEG initialisation failed: %s", status_text(st));
        }
        h->jpeg = d;
    }
    JpegDecoder* d = h->jpeg;
    for (int i = 0; i < n; i++) {
        if (!jpeg[i] || bytes[i] == 0) return fail(h, SA_ERR_INVALID_ARG, "JPEG frame %d: null or empty", i);
        int comps = 0, ws[NVJPEG_MAX_COMPONENT] = {0}, hs[NVJPEG_MAX_COMPONENT] = {0};
        nvjpegChromaSubsampling_t sub;
        const nvjpegStatus_t st = a.GetImageInfo(d->lib, static_cast<const unsigned char*>(jpeg[i]), bytes[i], &comps, &sub, ws, hs);
        if (st != NVJPEG_STATUS_SUCCESS) return fail(h, SA_ERR_PARSE, "JPEG frame %d: %s", i, status_text(st));
        if (ws[0] != width || hs[0] != height)
            return fail(h, SA_ERR_INVALID_ARG, "JPEG frame %d is %dx%d, the handle expects %dx%d (two eyes side by side)", i, ws[0], hs[0], width, height);
    }
    std::vector<nvjpegImage_t> out((size_t)n);
    for (int i = 0; i < n; i++) {
        for (int c = 0; c < NVJPEG_MAX_COMPONENT; c++) { out[i].channel[c] = nullptr; out[i].pitch[c] = 0; }
        out[i].channel[0] = dst + (size_t)i * frame;
        out[i].pitch[0] = pitch;
    }
    if (d->batched_ok && n > 1) {
        nvjpegStatus_t st = NVJPEG_STATUS_SUCCESS;
        if (d->batch != n) {
            const unsigned hw = std::thread::hardware_concurrency();
            st = a.BatchedInit(d->lib, d->state, n, (int)std::max(1u, std::min(hw ? hw : 1u, (unsigned)n)), NVJPEG_OUTPUT_BGRI);
            d->batch = st == NVJPEG_STATUS_SUCCESS ? n : 0;
        }
        if (st == NVJPEG_STATUS_SUCCESS) {
            std::vector<const unsigned char*> data((size_t)n);
            for (int i = 0; i < n; i++) data[i] = static_cast<const unsigned char*>(jpeg[i]);
            st = a.Batched(d->lib, d->state, data.data(), bytes, out.data(), s);
            if (st == NVJPEG_STATUS_SUCCESS) return SA_OK;
        }
        // e.g. progressive streams: decode one by one from now on
        d->batched_ok = false;
        d->batch = 0;
    }
    for (int i = 0; i < n; i++) {
        const nvjpegStatus_t st = a.Decode(d->lib, d->state, static_cast<const unsigned char*>(jpeg[i]), bytes[i], NVJPEG_OUTPUT_BGRI, &out[i], s);
        if (st != NVJPEG_STATUS_SUCCESS)
            return fail(h, st == NVJPEG_STATUS_BAD_JPEG || st == NVJPEG_STATUS_INCOMPLETE_BITSTREAM ? SA_ERR_PARSE : SA_ERR_CUDA, "JPEG frame %d: %s", i, status_text(st));
    }
    return SA_OK;
}

}  // namespace sa
