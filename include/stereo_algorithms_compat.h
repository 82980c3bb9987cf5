/* stereo_algorithms_compat.h - the reference's per-model entry points as a true C ABI.
 *
 * The reference exports, from one shared library per model, the unmangled names
 *     void* Initialize(char* model_path, int gpu_id, char* calibration_path);
 *     int   Run<Model>[_RectifyImage](void* p, cv::Mat& left, cv::Mat& right, float* pointcloud, cv::Mat& disparity);
 *     const char* Version(void* p);   int Release(void* p);
 * (FastACVNet_plus/include/FastACVNet_plus_Algorithm.h:23-43, RAFTStereo/include/RAFTStereoAlgorithm.h:24-39,
 *  HitNet/include/HitNetAlgorithm.h:24-39, CREStereo/include/CREStereoAlgorithm.h:24-44).
 * Their `extern "C"` signatures carry cv::Mat&, so they are bound to the caller's OpenCV C++ ABI.  The functions
 * below are the same four operations with every cv::Mat replaced by (data, rows, cols, step); the header-only
 * shims in shims/ put the reference names and cv::Mat signatures back on top of them for C++ callers.
 *
 * Behaviour kept from the reference: Initialize always returns a non-null handle and latches
 * failure; Run returns -1 after a failed init or for empty images, 0 otherwise; the *_RectifyImage flavour (and
 * RunRAFTStereo / RunHitNet, which always rectify) overwrites left/right with the rectified images; pointcloud is
 * a caller-allocated float[rows*cols*6]; disparity comes back as rows*cols floats (for RAFTStereo already
 * sign-flipped).  Differences: nothing is printed or asserted (sa_ref_LastError), Release also frees the object,
 * any image size is accepted (the reference asserts 640x480), and the TensorRT engine between the two halves is
 * not part of this library: it is a callback (sa_ref_SetEngine) or, when none is set, a synthetic disparity.
 */
#ifndef STEREO_ALGORITHMS_COMPAT_H_
#define STEREO_ALGORITHMS_COMPAT_H_

#include "stereo_geom.h"

#ifdef __cplusplus
extern "C" {
#endif

/* A cv::Mat of type CV_8UC3 seen from C: Mat::data, Mat::rows, Mat::cols, Mat::step[0]. */
typedef struct sa_image {
    uint8_t* data;
    int rows, cols;
    size_t step; /* bytes between rows; 0 = cols*3 */
} sa_image;

/* Initialize (FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:152-157 -> :35-56).  model: enum sa_model. */
SA_API void* sa_ref_Initialize(int model, const char* model_path, int gpu_id, const char* calibration_path);
/* Run<Model> (rectify = 0; CREStereo, FastACVNet_plus only in the reference, :77-91) and Run<Model>_RectifyImage /
 * RunRAFTStereo / RunHitNet (rectify = 1, :59-75).  disparity: rows*cols floats, may be NULL. */
SA_API int sa_ref_Run(void* p, int rectify, sa_image* left, sa_image* right, float* pointcloud, float* disparity);
/* Version (:178-181): "<Model>Algorithm_V1.0" exactly as the reference spells it per model. */
SA_API const char* sa_ref_Version(void* p);
/* Release (:186-190). */
SA_API int sa_ref_Release(void* p);

/* The network slot: a callback with the device tensors and the handle's stream (sa_engine_fn in stereo_geom.h: the stream is
 * the one both halves run on, never NULL - enqueue on it, do not synchronise), or (default) the synthetic stub with this seed;
 * seed 0 = leave the disparity buffer as the caller prepared it through sa_ref_Geometry + sa_set_disparity. */
SA_API int sa_ref_SetEngine(void* p, sa_engine_fn fn, void* user);
SA_API int sa_ref_SetStubSeed(void* p, unsigned seed);
/* Not in the reference.  The demos allocate `pointcloud` once and pass it to every Run (FastACVNet_plus/test/main.cpp:40-50).
 * With on = 1, when sa_ref_Run sees the same pointcloud pointer on two consecutive calls it page-locks that buffer
 * (sa_host_register), which takes the 7.4 MB device-to-host copy of a 640x480 cloud from ~0.6 ms (pageable) to ~0.15 ms.  The
 * registration is dropped when another pointer shows up, when this is switched off, and in sa_ref_Release.
 * OFF by default, because it puts an obligation on the caller the reference does not have: a page-locked buffer must not be
 * freed behind the library's back (free it after Release, after SetHostRegister(p, 0), or after a Run with another buffer) -
 * otherwise a later allocation that lands at the same address would be treated as still page-locked and the copy would go to
 * the old pages. */
SA_API int sa_ref_SetHostRegister(void* p, int on);
/* The underlying geometry handle (created lazily for the first image size seen; 640x480 after Initialize). */
SA_API sa_handle* sa_ref_Geometry(void* p);
SA_API const char* sa_ref_LastError(void* p);
/* heatmap() of the demos (FastACVNet_plus/test/main.cpp:7-24) on the disparity of the last Run: rows*cols*3 BGR. */
SA_API int sa_ref_Heatmap(void* p, uint8_t* out_bgr);
/* The demo's pointcloud.txt writer (FastACVNet_plus/test/main.cpp:62-69): one "x y z r g b" line per point. */
SA_API int sa_write_pointcloud_txt(const char* path, const float* points, size_t n_points);
/* Same points as a PLY file (the text dump is 50+ bytes per point): x y z as float, r g b as uchar,
 * binary little-endian (binary != 0, 15 bytes per point) or ASCII.  skip_non_finite != 0 drops points whose x, y or z is
 * inf/NaN (the dense cloud has them wherever the disparity is 0). */
SA_API int sa_write_pointcloud_ply(const char* path, const float* points, size_t n_points, int binary, int skip_non_finite);
/* ... and as a PCD v0.7 file (Point Cloud Library): FIELDS x y z rgb, rgb packed 0x00RRGGBB in a 32-bit unsigned, DATA binary. */
SA_API int sa_write_pointcloud_pcd(const char* path, const float* points, size_t n_points, int skip_non_finite);
/* Lossless stand-ins for the demos' cv::imwrite calls (FastACVNet_plus/test/main.cpp:55,59; no JPEG codec here):
 * binary Netpbm.  sa_write_pgm: 8-bit grey (P5), e.g. the output of sa_disparity_u8.  sa_write_ppm_bgr: a BGR image
 * as OpenCV holds it (rectified frame, heat map) written as RGB (P6).  step = bytes per row, 0 = packed. */
SA_API int sa_write_pgm(const char* path, const uint8_t* grey, int w, int h, size_t step);
SA_API int sa_write_ppm_bgr(const char* path, const uint8_t* bgr, int w, int h, size_t step);

#ifdef __cplusplus
}
#endif
#endif
