// The reference's Initialize / Run / Version / Release semantics on top of the geometry library
// (include/stereo_algorithms_compat.h).  Mirrors class FastACVNet_plus_Algorithm
// (FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:6-96) and its three siblings, minus OpenCV and TensorRT.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/stereo_algorithms_compat.h"

namespace {

struct Algorithm {
    int model = 0;
    int gpu_id = 0;
    std::string model_path, calib_path;
    bool calib_exists = false;
    int initflag = -1;  // FastACVNet_plus_Algorithm::initflag (:23)
    sa_handle* geom = nullptr;
    int W = 0, H = 0;
    sa_engine_fn engine = nullptr;
    void* engine_user = nullptr;
    unsigned stub_seed = 1;
    std::string err;
    std::vector<uint8_t> tmp;
    // page-locking of a point-cloud buffer the caller reuses (sa_ref_SetHostRegister)
    bool host_register = false;    // opt-in: see sa_ref_SetHostRegister
    float* pc_seen = nullptr;      // pointer of the previous Run
    float* pc_reg = nullptr;       // registered pointer ...
    size_t pc_reg_bytes = 0;       // ... and its size
    float* pc_failed = nullptr;    // registration of this pointer failed: do not retry on every call
};

void drop_registration(Algorithm* a)
{
    if (a->pc_reg) sa_host_unregister(a->pc_reg);
    a->pc_reg = nullptr;
    a->pc_reg_bytes = 0;
}

void track_pointcloud_buffer(Algorithm* a, float* pointcloud, size_t bytes)
{
    if (!a->host_register) return;
    if (pointcloud == a->pc_reg && bytes <= a->pc_reg_bytes) { a->pc_seen = pointcloud; return; }
    if (a->pc_reg) drop_registration(a);                             // another buffer (or a larger image): forget the old one
    if (pointcloud == a->pc_seen && pointcloud != a->pc_failed) {   // second call in a row with this buffer: worth pinning
        if (sa_host_register(pointcloud, bytes) == SA_OK) { a->pc_reg = pointcloud; a->pc_reg_bytes = bytes; }
        else a->pc_failed = pointcloud;                              // e.g. the caller registered it already - which is just as good
    }
    a->pc_seen = pointcloud;
}

bool file_exists(const char* p)
{
    if (!p) return false;
    FILE* f = std::fopen(p, "rb");
    if (!f) return false;
    std::fclose(f);
    return true;
}

// (re)create the geometry handle for a W x H camera; the reference hard-codes 640x480
// (FastACVNet_plus/src/TRTFastACVNet_plus.cpp:13-14) and asserts on anything else (:121-122).
int make_geometry(Algorithm* a, int W, int H)
{
    if (a->geom) { sa_destroy(a->geom); a->geom = nullptr; }
    sa_config cfg;
    std::memset(&cfg, 0, sizeof cfg);
    cfg.struct_size = (int)sizeof cfg;
    cfg.gpu_id = a->gpu_id;
    cfg.width = W; cfg.height = H;
    cfg.model = a->model;
    cfg.tensor_dtype = SA_F32;   // the reference's bindings are float (TRTFastACVNet_plus.h:68)
    cfg.max_batch = 1;           // #define ..._BATCH_SIZE 1 (TRTFastACVNet_plus.h:15)
    int rc = sa_create(&cfg, &a->geom);
    if (rc != SA_OK) { a->err = sa_last_error(nullptr); a->geom = nullptr; return rc; }
    a->W = W; a->H = H;
    if (a->calib_exists) {
        rc = sa_load_calibration_yaml(a->geom, a->calib_path.c_str());
        if (rc != SA_OK) { a->err = sa_last_error(a->geom); return rc; }
    }
    if (a->engine) sa_set_engine(a->geom, a->engine, a->engine_user);
    return SA_OK;
}

}  // namespace

extern "C" {

void* sa_ref_Initialize(int model, const char* model_path, int gpu_id, const char* calibration_path)
{
    Algorithm* a = new Algorithm();   // the reference always returns the object, even on failure (:152-157)
    a->model = model;
    a->gpu_id = gpu_id;
    a->model_path = model_path ? model_path : "";
    a->calib_path = calibration_path ? calibration_path : "";
    a->calib_exists = file_exists(calibration_path);
    if (!a->calib_exists) a->err = "calibration_path does not exist!!!";   // the reference prints this and goes on (:42-45)
    if (model < SA_MODEL_RAFTSTEREO || model > SA_MODEL_FASTACVNET_PLUS) { a->err = "unknown model id"; return a; }
    const bool warned = !a->calib_exists;
    if (make_geometry(a, 640, 480) != SA_OK) { a->initflag = -1; return a; }   // "init failed!!" (:47-51)
    a->initflag = 0;
    if (!warned) a->err.clear();
    return a;
}

int sa_ref_Run(void* p, int rectify, sa_image* left, sa_image* right, float* pointcloud, float* disparity)
{
    Algorithm* a = static_cast<Algorithm*>(p);
    if (!a) return -1;
    if (a->initflag != 0) { if (a->err.empty()) a->err = "init failed,please check!!"; return -1; }     // (:61-65)
    if (!left || !right || !left->data || !right->data || left->rows <= 0 || left->cols <= 0 || right->rows <= 0 || right->cols <= 0) {
        a->err = "Image_src is empty!!!";                                                               // (:66-70)
        return -1;
    }
    if (left->rows != right->rows || left->cols != right->cols) { a->err = "left and right image sizes differ"; return -1; }
    if (!pointcloud) { a->err = "pointcloud is null"; return -1; }
    const int W = left->cols, H = left->rows;
    if (W != a->W || H != a->H) {
        if (make_geometry(a, W, H) != SA_OK) { a->initflag = -1; return -1; }
    }
    sa_handle* g = a->geom;
    const size_t row = (size_t)W * 3;
    const size_t lstep = left->step ? left->step : row, rstep = right->step ? right->step : row;
    int rc;
    if (lstep == rstep) {
        rc = sa_rectify_preprocess(g, 1, left->data, right->data, lstep, SA_HOST, rectify, nullptr);
    } else {   // different pitches: repack the right image
        a->tmp.resize(row * H);
        for (int y = 0; y < H; y++) std::memcpy(&a->tmp[y * row], right->data + (size_t)y * rstep, row);
        std::vector<uint8_t> ltmp(row * H);
        for (int y = 0; y < H; y++) std::memcpy(&ltmp[y * row], left->data + (size_t)y * lstep, row);
        rc = sa_rectify_preprocess(g, 1, ltmp.data(), a->tmp.data(), row, SA_HOST, rectify, nullptr);
        if (rc == SA_OK) rc = sa_synchronize(g);
    }
    if (rc != SA_OK) { a->err = sa_last_error(g); return -1; }
    // network slot: context->enqueueV3(stream) (TRTFastACVNet_plus.cpp:152).  The callback gets the stream stage A was queued
    // on and stage B will be queued on (the handle's own, never NULL), so an engine enqueued there is ordered after the
    // tensors are written and before the disparity is read - without the cudaStreamSynchronize of the reference (:159).
    if (a->engine) {
        rc = a->engine(a->engine_user, sa_get_stream(g), 0, 1, sa_input_tensor(g, SA_LEFT), sa_input_tensor(g, SA_RIGHT), sa_disparity_buffer(g));
        if (rc != 0) { a->err = "engine callback failed"; sa_synchronize(g); return -1; }
    } else if (a->stub_seed != 0) {
        rc = sa_stub_disparity(g, 1, a->stub_seed, nullptr);
        if (rc != SA_OK) { a->err = sa_last_error(g); return -1; }
    }
    track_pointcloud_buffer(a, pointcloud, (size_t)W * H * 6 * sizeof(float));
    rc = sa_reproject(g, 1, SA_DENSE, nullptr, nullptr, SA_DEVICE, nullptr);           // cloud stays in the internal buffer ...
    if (rc != SA_OK) { a->err = sa_last_error(g); sa_synchronize(g); return -1; }
    // ... and every read-back of the call is queued behind it and waited for once (the reference: two blocking cudaMemcpy,
    // :163-164).  cv::remap(img, img, ...) leaves the rectified image in the caller's Mat (FastACVNet_plus_Algorithm.cpp:142-143):
    // copied straight into its rows, whatever their step.
    rc = sa_download(g, 1, pointcloud, disparity, rectify ? left->data : nullptr, lstep, rectify ? right->data : nullptr, rstep, nullptr, nullptr);
    if (rc != SA_OK) { a->err = sa_last_error(g); return -1; }
    return 0;
}

const char* sa_ref_Version(void* p)
{
    Algorithm* a = static_cast<Algorithm*>(p);
    switch (a ? a->model : -1) {
        case SA_MODEL_RAFTSTEREO: return "RAFTStereoAlgorithm_V1.0";             // RAFTStereo/src/RAFTStereoAlgorithm.cpp:154
        case SA_MODEL_HITNET: return "HitNetAlgorithm_V1.0";                     // HitNet/src/HitNetAlgorithm.cpp:132
        case SA_MODEL_CRESTEREO: return "CREStereoAlgorithm_V1.0";               // CREStereo/src/CREStereoAlgorithm.cpp:180
        case SA_MODEL_FASTACVNET_PLUS: return "FastACVNet_plus_Algorithm_V1.0";  // FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:180
        default: return "unknown";
    }
}

int sa_ref_Release(void* p)
{
    Algorithm* a = static_cast<Algorithm*>(p);
    if (!a) return -1;
    drop_registration(a);
    if (a->geom) sa_destroy(a->geom);
    delete a;
    return 0;
}

int sa_ref_SetHostRegister(void* p, int on)
{
    Algorithm* a = static_cast<Algorithm*>(p);
    if (!a) return -1;
    a->host_register = on != 0;
    if (!on) drop_registration(a);
    return 0;
}

int sa_ref_SetEngine(void* p, sa_engine_fn fn, void* user)
{
    Algorithm* a = static_cast<Algorithm*>(p);
    if (!a) return -1;
    a->engine = fn; a->engine_user = user;
    if (a->geom) sa_set_engine(a->geom, fn, user);
    return 0;
}

int sa_ref_SetStubSeed(void* p, unsigned seed)
{
    Algorithm* a = static_cast<Algorithm*>(p);
    if (!a) return -1;
    a->stub_seed = seed;
    return 0;
}

sa_handle* sa_ref_Geometry(void* p) { return p ? static_cast<Algorithm*>(p)->geom : nullptr; }

const char* sa_ref_LastError(void* p) { return p ? static_cast<Algorithm*>(p)->err.c_str() : "null handle"; }

int sa_ref_Heatmap(void* p, uint8_t* out_bgr)
{
    Algorithm* a = static_cast<Algorithm*>(p);
    if (!a || !a->geom || !out_bgr) return -1;
    int rc = sa_heatmap(a->geom, 1, out_bgr, SA_HOST, nullptr);
    if (rc != SA_OK) { a->err = sa_last_error(a->geom); return -1; }
    return 0;
}

int sa_write_pointcloud_txt(const char* path, const float* points, size_t n_points)
{
    if (!path || !points) return SA_ERR_INVALID_ARG;
    FILE* f = std::fopen(path, "w");
    if (!f) return SA_ERR_IO;
    // std::ostream << float with default flags == printf("%g") (6 significant digits)
    for (size_t i = 0; i < n_points; i++) {
        const float* q = points + i * 6;
        std::fprintf(f, "%g %g %g %g %g %g\n", q[0], q[1], q[2], q[3], q[4], q[5]);
    }
    return std::fclose(f) == 0 ? SA_OK : SA_ERR_IO;
}

int sa_write_pointcloud_ply(const char* path, const float* points, size_t n_points, int binary, int skip_non_finite)
{
    if (!path || !points) return SA_ERR_INVALID_ARG;
    auto finite = [](float v) { return v == v && v - v == 0.0f; };
    size_t kept = n_points;
    if (skip_non_finite) {
        kept = 0;
        for (size_t i = 0; i < n_points; i++) kept += finite(points[i * 6]) && finite(points[i * 6 + 1]) && finite(points[i * 6 + 2]);
    }
    FILE* f = std::fopen(path, "wb");
    if (!f) return SA_ERR_IO;
    std::fprintf(f, "ply\nformat %s 1.0\ncomment StereoAlgorithms geometry path\nelement vertex %zu\n"
                    "property float x\nproperty float y\nproperty float z\nproperty uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n",
                 binary ? "binary_little_endian" : "ascii", kept);
    for (size_t i = 0; i < n_points; i++) {
        const float* q = points + i * 6;
        if (skip_non_finite && !(finite(q[0]) && finite(q[1]) && finite(q[2]))) continue;
        const unsigned char c[3] = {(unsigned char)q[3], (unsigned char)q[4], (unsigned char)q[5]};
        if (binary) { std::fwrite(q, sizeof(float), 3, f); std::fwrite(c, 1, 3, f); }
        else std::fprintf(f, "%g %g %g %u %u %u\n", q[0], q[1], q[2], c[0], c[1], c[2]);
    }
    return std::fclose(f) == 0 ? SA_OK : SA_ERR_IO;
}

int sa_write_pointcloud_pcd(const char* path, const float* points, size_t n_points, int skip_non_finite)
{
    if (!path || !points) return SA_ERR_INVALID_ARG;
    auto finite = [](float v) { return v == v && v - v == 0.0f; };
    size_t kept = n_points;
    if (skip_non_finite) {
        kept = 0;
        for (size_t i = 0; i < n_points; i++) kept += finite(points[i * 6]) && finite(points[i * 6 + 1]) && finite(points[i * 6 + 2]);
    }
    FILE* f = std::fopen(path, "wb");
    if (!f) return SA_ERR_IO;
    std::fprintf(f, "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z rgb\nSIZE 4 4 4 4\nTYPE F F F U\nCOUNT 1 1 1 1\n"
                    "WIDTH %zu\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %zu\nDATA binary\n", kept, kept);
    bool ok = true;
    for (size_t i = 0; i < n_points && ok; i++) {
        const float* q = points + i * 6;
        if (skip_non_finite && !(finite(q[0]) && finite(q[1]) && finite(q[2]))) continue;
        const uint32_t rgb = ((uint32_t)(unsigned char)q[3] << 16) | ((uint32_t)(unsigned char)q[4] << 8) | (uint32_t)(unsigned char)q[5];
        ok = std::fwrite(q, sizeof(float), 3, f) == 3 && std::fwrite(&rgb, 4, 1, f) == 1;
    }
    return (std::fclose(f) == 0 && ok) ? SA_OK : SA_ERR_IO;
}

int sa_write_pgm(const char* path, const uint8_t* grey, int w, int h, size_t step)
{
    if (!path || !grey || w <= 0 || h <= 0 || (step != 0 && step < (size_t)w)) return SA_ERR_INVALID_ARG;
    FILE* f = std::fopen(path, "wb");
    if (!f) return SA_ERR_IO;
    std::fprintf(f, "P5\n%d %d\n255\n", w, h);
    const size_t st = step ? step : (size_t)w;
    bool ok = true;
    for (int y = 0; y < h && ok; y++) ok = std::fwrite(grey + (size_t)y * st, 1, (size_t)w, f) == (size_t)w;
    return (std::fclose(f) == 0 && ok) ? SA_OK : SA_ERR_IO;
}

int sa_write_ppm_bgr(const char* path, const uint8_t* bgr, int w, int h, size_t step)
{
    if (!path || !bgr || w <= 0 || h <= 0 || (step != 0 && step < (size_t)w * 3)) return SA_ERR_INVALID_ARG;
    FILE* f = std::fopen(path, "wb");
    if (!f) return SA_ERR_IO;
    std::fprintf(f, "P6\n%d %d\n255\n", w, h);
    const size_t st = step ? step : (size_t)w * 3;
    std::vector<uint8_t> row((size_t)w * 3);
    bool ok = true;
    for (int y = 0; y < h && ok; y++) {
        const uint8_t* p = bgr + (size_t)y * st;
        for (int x = 0; x < w; x++) { row[3 * x] = p[3 * x + 2]; row[3 * x + 1] = p[3 * x + 1]; row[3 * x + 2] = p[3 * x]; }
        ok = std::fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    return (std::fclose(f) == 0 && ok) ? SA_OK : SA_ERR_IO;
}

}  // extern "C"
