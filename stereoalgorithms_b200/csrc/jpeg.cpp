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
    // encoder
    nvjpegStatus_t (*EncStateCreate)(nvjpegHandle_t, nvjpegEncoderState_t*, cudaStream_t) = nullptr;
    nvjpegStatus_t (*EncStateDestroy)(nvjpegEncoderState_t) = nullptr;
    nvjpegStatus_t (*EncParamsCreate)(nvjpegHandle_t, nvjpegEncoderParams_t*, cudaStream_t) = nullptr;
    nvjpegStatus_t (*EncParamsDestroy)(nvjpegEncoderParams_t) = nullptr;
    nvjpegStatus_t (*EncSetQuality)(nvjpegEncoderParams_t, const int, cudaStream_t) = nullptr;
    nvjpegStatus_t (*EncSetSampling)(nvjpegEncoderParams_t, const nvjpegChromaSubsampling_t, cudaStream_t) = nullptr;
    nvjpegStatus_t (*EncSetHuffman)(nvjpegEncoderParams_t, const int, cudaStream_t) = nullptr;
    nvjpegStatus_t (*EncodeImage)(nvjpegHandle_t, nvjpegEncoderState_t, const nvjpegEncoderParams_t, const nvjpegImage_t*, nvjpegInputFormat_t, int, int, cudaStream_t) = nullptr;
    nvjpegStatus_t (*EncodeYUV)(nvjpegHandle_t, nvjpegEncoderState_t, const nvjpegEncoderParams_t, const nvjpegImage_t*, nvjpegChromaSubsampling_t, int, int, cudaStream_t) = nullptr;
    nvjpegStatus_t (*EncRetrieve)(nvjpegHandle_t, nvjpegEncoderState_t, unsigned char*, size_t*, cudaStream_t) = nullptr;
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
        a.EncStateCreate = reinterpret_cast<decltype(a.EncStateCreate)>(sym("nvjpegEncoderStateCreate"));
        a.EncStateDestroy = reinterpret_cast<decltype(a.EncStateDestroy)>(sym("nvjpegEncoderStateDestroy"));
        a.EncParamsCreate = reinterpret_cast<decltype(a.EncParamsCreate)>(sym("nvjpegEncoderParamsCreate"));
        a.EncParamsDestroy = reinterpret_cast<decltype(a.EncParamsDestroy)>(sym("nvjpegEncoderParamsDestroy"));
        a.EncSetQuality = reinterpret_cast<decltype(a.EncSetQuality)>(sym("nvjpegEncoderParamsSetQuality"));
        a.EncSetSampling = reinterpret_cast<decltype(a.EncSetSampling)>(sym("nvjpegEncoderParamsSetSamplingFactors"));
        a.EncSetHuffman = reinterpret_cast<decltype(a.EncSetHuffman)>(sym("nvjpegEncoderParamsSetOptimizedHuffman"));
        a.EncodeImage = reinterpret_cast<decltype(a.EncodeImage)>(sym("nvjpegEncodeImage"));
        a.EncodeYUV = reinterpret_cast<decltype(a.EncodeYUV)>(sym("nvjpegEncodeYUV"));
        a.EncRetrieve = reinterpret_cast<decltype(a.EncRetrieve)>(sym("nvjpegEncodeRetrieveBitstream"));
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
    nvjpegEncoderState_t enc = nullptr;
    nvjpegEncoderParams_t enc_params = nullptr;
    int batch = 0;          // size the batched decoder is initialised for (0 = not)
    bool batched_ok = true; // cleared when the batched decoder refused a stream this handle was given: single-image calls from then on
};

int jpeg_open(sa_handle* h);
int jpeg_decode_checked(sa_handle* h, int n, const void* const* jpeg, const size_t* bytes, uint8_t* dst, size_t pitch, size_t frame, cudaStream_t s);

void jpeg_release(sa_handle* h)
{
    if (!h || !h->jpeg) return;
    const NvJpegApi& a = api();
    if (a.dl) {
        if (h->jpeg->enc_params) a.EncParamsDestroy(h->jpeg->enc_params);
        if (h->jpeg->enc) a.EncStateDestroy(h->jpeg->enc);
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
    if (const int rc = jpeg_open(h)) return rc;
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
    return jpeg_decode_checked(h, n, jpeg, bytes, dst, pitch, frame, s);
}

// the handle's nvJPEG state, created on first use
int jpeg_open(sa_handle* h)
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
    return SA_OK;
}

int jpeg_decode_checked(sa_handle* h, int n, const void* const* jpeg, const size_t* bytes, uint8_t* dst, size_t pitch, size_t frame, cudaStream_t s)
{
    const NvJpegApi& a = api();
    JpegDecoder* d = h->jpeg;
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

// Size of a JPEG stream (first image component) without decoding it; SA_ERR_PARSE if it is not one.
int jpeg_image_size(sa_handle* h, const void* jpeg, size_t bytes, int* width, int* height)
{
    const NvJpegApi& a = api();
    if (const int rc = jpeg_open(h)) return rc;
    int comps = 0, ws[NVJPEG_MAX_COMPONENT] = {0}, hs[NVJPEG_MAX_COMPONENT] = {0};
    nvjpegChromaSubsampling_t sub;
    const nvjpegStatus_t st = a.GetImageInfo(h->jpeg->lib, static_cast<const unsigned char*>(jpeg), bytes, &comps, &sub, ws, hs);
    if (st != NVJPEG_STATUS_SUCCESS) return fail(h, SA_ERR_PARSE, "JPEG stream: %s", status_text(st));
    *width = ws[0];
    *height = hs[0];
    return SA_OK;
}

// One 8-bit image on the device (channels = 1: grey, 3: BGR interleaved) -> baseline JPEG in `out` (host).  Colour is stored 4:2:0 like
// cv::imwrite's default; the Huffman tables are the standard ones.  Synchronises s.
int jpeg_encode(sa_handle* h, const uint8_t* dev_image, int width, int height, int channels, size_t pitch, int quality, std::vector<unsigned char>& out,
                cudaStream_t s)
{
    const NvJpegApi& a = api();
    if (const int rc = jpeg_open(h)) return rc;
    JpegDecoder* d = h->jpeg;
    nvjpegStatus_t st = NVJPEG_STATUS_SUCCESS;
    if (!d->enc) st = a.EncStateCreate(d->lib, &d->enc, s);
    if (st == NVJPEG_STATUS_SUCCESS && !d->enc_params) st = a.EncParamsCreate(d->lib, &d->enc_params, s);
    if (st == NVJPEG_STATUS_SUCCESS) st = a.EncSetQuality(d->enc_params, quality, s);
    if (st == NVJPEG_STATUS_SUCCESS) st = a.EncSetHuffman(d->enc_params, 0, s);
    if (st == NVJPEG_STATUS_SUCCESS) st = a.EncSetSampling(d->enc_params, channels == 1 ? NVJPEG_CSS_GRAY : NVJPEG_CSS_420, s);
    if (st != NVJPEG_STATUS_SUCCESS) return fail(h, SA_ERR_CUDA, "nvJPEG encoder set-up failed: %s", status_text(st));
    nvjpegImage_t src;
    for (int c = 0; c < NVJPEG_MAX_COMPONENT; c++) { src.channel[c] = nullptr; src.pitch[c] = 0; }
    src.channel[0] = const_cast<unsigned char*>(dev_image);
    src.pitch[0] = pitch;
    if (channels == 1) st = a.EncodeYUV(d->lib, d->enc, d->enc_params, &src, NVJPEG_CSS_GRAY, width, height, s);
    else st = a.EncodeImage(d->lib, d->enc, d->enc_params, &src, NVJPEG_INPUT_BGRI, width, height, s);
    if (st != NVJPEG_STATUS_SUCCESS) return fail(h, SA_ERR_CUDA, "nvJPEG encode failed: %s", status_text(st));
    size_t len = 0;
    st = a.EncRetrieve(d->lib, d->enc, nullptr, &len, s);
    if (st != NVJPEG_STATUS_SUCCESS) return fail(h, SA_ERR_CUDA, "nvJPEG bitstream size: %s", status_text(st));
    if (cudaStreamSynchronize(s) != cudaSuccess) return fail(h, SA_ERR_CUDA, "cudaStreamSynchronize after the JPEG encode failed");
    out.resize(len);
    st = a.EncRetrieve(d->lib, d->enc, out.data(), &len, s);
    if (st != NVJPEG_STATUS_SUCCESS) return fail(h, SA_ERR_CUDA, "nvJPEG bitstream retrieval: %s", status_text(st));
    if (cudaStreamSynchronize(s) != cudaSuccess) return fail(h, SA_ERR_CUDA, "cudaStreamSynchronize after the JPEG encode failed");
    out.resize(len);
    return SA_OK;
}

}  // namespace sa
