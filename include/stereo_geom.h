/* stereo_geom.h - C ABI of libstereo_geom.so, the B200 (sm_100a) geometry library for the
 * per-frame path that wraps every network of StereoAlgorithms:
 *
 *     rectify -> preprocess -> [network, untouched] -> reproject (point cloud) -> heatmap
 *
 * Plain C: opaque handle, raw pointers, sizes and pitches.  No OpenCV, TensorRT or torch types.
 * Every function returns SA_OK (0) or a negative sa_status; nothing asserts or prints.  The
 * message for the last failure on a handle is sa_last_error(h).  There is no CPU fallback: if no
 * CUDA device is usable, sa_create fails with SA_ERR_CUDA.
 *
 * Each declaration cites the reference interface it replaces (paths relative to the reference
 * tree).  The reference-named entry points (Initialize / Run<Model>[_RectifyImage] / Version /
 * Release) are in stereo_algorithms_compat.h, built on top of this header.
 *
 * Layouts (N = number of stereo pairs in the call, W x H = camera image, w x h = network size,
 * equal to W x H unless sa_config.net_width/net_height say otherwise):
 *   raw / rectified image   uint8  [N][H][W][3]   BGR, rows `pitch` bytes apart on input
 *   network tensor          f32|f16 [N][3][h][w]  RGB planes, one tensor per eye
 *                           (HitNet: a single [N][6][h][w] tensor, L_r L_g L_b R_r R_g R_b)
 *   disparity               float  [N][planes][h][w]  planes = 2 for CREStereo (plane 0 is used), else 1
 *   point cloud, dense      float  [N][h*w][6]    x y z r g b, row-major pixels, nothing filtered
 *   point cloud, compact    float  [N][h*w][6]    first counts[i] rows of frame i are valid
 *   heatmap                 uint8  [N][h][w][3]   BGR (JET)
 */
#ifndef STEREO_GEOM_H_
#define STEREO_GEOM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SA_API __attribute__((visibility("default")))

typedef struct sa_handle sa_handle;
typedef void* sa_stream; /* a cudaStream_t, or NULL for the handle's own stream */

enum sa_status {
    SA_OK = 0,
    SA_ERR_INVALID_ARG = -1,
    SA_ERR_NOT_READY = -2, /* calibration / maps / disparity missing for the requested stage */
    SA_ERR_CUDA = -3,
    SA_ERR_IO = -4,
    SA_ERR_PARSE = -5,
    SA_ERR_TOO_LARGE = -6, /* n > max_batch, or a size beyond a documented limit */
    SA_ERR_NOMEM = -7
};

/* Which network the tensors are prepared for.  Selects the normalisation of the preprocess
 * stage and the disparity convention of the reproject stage:
 *   RAFTSTEREO       f32(c), no scaling          RAFTStereo/src/stereo_preprocess.cu:4-39; reproject negates the
 *                                                disparity and writes it back (:54-55)
 *   HITNET           f32(f64(c)/255.0), 6-ch     HitNet/src/HitNet_preprocess.cu:4-53
 *   CRESTEREO        f32(c), no scaling          CREStereo/src/CREStereo_preprocess.cu:4-39; disparity = plane 0 of 2
 *                                                (CREStereo/src/TRTCREStereo.cpp:15,133)
 *   FASTACVNET_PLUS  (c/255.0f - mean)/std with double literals   FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:27-29 */
enum sa_model { SA_MODEL_RAFTSTEREO = 0, SA_MODEL_HITNET = 1, SA_MODEL_CRESTEREO = 2, SA_MODEL_FASTACVNET_PLUS = 3 };
enum sa_dtype { SA_F32 = 0, SA_F16 = 1 };
enum sa_where { SA_HOST = 0, SA_DEVICE = 1 };
enum sa_eye { SA_LEFT = 0, SA_RIGHT = 1 };
enum sa_reproject_mode { SA_DENSE = 0, SA_COMPACT = 1 };
enum sa_stage { SA_STAGE_RECTIFY = 0, SA_STAGE_RESIZE = 1, SA_STAGE_REPROJECT = 2, SA_STAGE_HEATMAP = 3, SA_STAGE_STUB = 4,
                SA_STAGE_MAPS = 5, SA_STAGE_COUNT_ = 6 };

enum sa_flags {
    SA_FLAG_TIMING = 1u << 0,      /* record CUDA events around every stage (sa_stage_time) */
    SA_FLAG_NO_HEATMAP_BUF = 1u << 1, /* do not allocate the internal heatmap buffer */
    SA_FLAG_NO_POINT_BUF = 1u << 2    /* do not allocate the internal point-cloud buffer (caller passes device `out`) */
};

/* Replaces the constants the reference hard-codes (INPUT_H/INPUT_W/batch 1,
 * FastACVNet_plus/src/TRTFastACVNet_plus.cpp:13-18, FastACVNet_plus/include/TRTFastACVNet_plus.h:15). */
typedef struct sa_config {
    int struct_size; /* = sizeof(sa_config) */
    int gpu_id;      /* cudaSetDevice(gpu_id), TRTFastACVNet_plus.cpp:45 - re-applied on every call */
    int width, height;         /* camera image size */
    int net_width, net_height; /* network input size; 0 = same as the image (the reference asserts equality, :121-122) */
    int model;                 /* enum sa_model */
    int tensor_dtype;          /* enum sa_dtype (the reference is f32 only) */
    int max_batch;             /* buffers are sized for this many stereo pairs (the reference: 1) */
    unsigned flags;            /* enum sa_flags */
} sa_config;

/* Buffer allocation of FastACVNet_plus::Initialize (TRTFastACVNet_plus.cpp:91-115) without the TensorRT part. */
SA_API int sa_create(const sa_config* cfg, sa_handle** out);
/* FastACVNet_plus::Release (TRTFastACVNet_plus.cpp:171-192); unlike the reference the handle itself is freed too. */
SA_API int sa_destroy(sa_handle* h);
SA_API const char* sa_version(void);
/* Replaces CUDA_CHECK's stderr + assert(0) (FastACVNet_plus/include/TRTFastACVNet_plus.h:18-26).  h may be NULL
 * (message of the last failed sa_create on this thread). */
SA_API const char* sa_last_error(const sa_handle* h);

/* ReadObjectYml (FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:98-114): the 11 !!opencv-matrix nodes of a
 * StereoCalibration.yml.  Also computes the rectification maps for both eyes on the device - once, not per frame
 * as the reference does (:139-140). */
SA_API int sa_load_calibration_yaml(sa_handle* h, const char* path);
/* Same, from memory (struct CalibrationParam, FastACVNet_plus/include/TRTFastACVNet_plus.h:30-43).  Row-major
 * doubles: K 3x3, D nD (4,5,8,12 or 14) coefficients, R 3x3, P 3x4, Q 4x4. */
SA_API int sa_set_calibration(sa_handle* h, const double* K_l, const double* D_l, int nD_l, const double* R_L, const double* P1,
                              const double* K_r, const double* D_r, int nD_r, const double* R_R, const double* P2, const double* Q);
/* Calibrationparam_Q[i] = (float)Q[i] (TRTFastACVNet_plus.cpp:107-113); use after a resize changed the geometry. */
SA_API int sa_set_Q(sa_handle* h, const double* Q16);
SA_API int sa_get_Q(sa_handle* h, double* Q16);
/* Maps in cv::initUndistortRectifyMap's CV_16SC2 + CV_16UC1 layout (host pointers): map1 int16 [H][W][2] = (x,y),
 * map2 uint16 [H][W] = (fy<<5)|fx.  sa_set_maps injects caller-computed maps instead of the library's. */
SA_API int sa_set_maps(sa_handle* h, int eye, const int16_t* map1, const uint16_t* map2);
SA_API int sa_get_maps(sa_handle* h, int eye, int16_t* map1, uint16_t* map2);

/* Stage A.  RectifyImage (FastACVNet_plus_Algorithm.cpp:132-145: 2x remap INTER_LINEAR, bit-exact) fused with
 * <Model>_preprocess (TRTFastACVNet_plus.cpp:130-134) for n pairs in one launch.  left/right: [n][H] rows of `pitch`
 * bytes (0 = 3*W), host or device.  rectify = 0 is the reference's plain Run<Model> (images already rectified).
 * Results stay on the device: sa_input_tensor, sa_copy_rectified. */
SA_API int sa_rectify_preprocess(sa_handle* h, int n, const uint8_t* left, const uint8_t* right, size_t pitch, int where,
                                 int rectify, sa_stream stream);
/* buffers[inputIndex1/2] and buffers[outputIndex] of the reference (TRTFastACVNet_plus.cpp:93-95): the device
 * pointers a TensorRT context is bound to.  HitNet: one 6-channel tensor, `eye` is ignored. */
SA_API void* sa_input_tensor(sa_handle* h, int eye);
SA_API float* sa_disparity_buffer(sa_handle* h);
SA_API int sa_disparity_planes(const sa_handle* h);
/* Stand-in for context->enqueueV3(stream) (TRTFastACVNet_plus.cpp:152) when no engine is available: fills the
 * disparity buffer with a smooth synthetic field, ~5 % zeros and ~1 % negatives (RAFTStereo: sign-flipped, as
 * that network emits it). */
SA_API int sa_stub_disparity(sa_handle* h, int n, unsigned seed, sa_stream stream);
/* Upload a network output [n][planes][h][w] (tests, or an engine that ran elsewhere). */
SA_API int sa_set_disparity(sa_handle* h, int n, const float* disparity, int where, sa_stream stream);

/* Stage B.  <Model>_reprojectImageTo3D (FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:41-63,75-80) for n frames.
 * mode SA_DENSE: every pixel, like the reference.  SA_COMPACT (extension): only pixels with a finite disparity
 * > 0, order preserved, counts[i] rows valid in frame i's slot.  out: NULL (keep in the internal buffer),
 * a device buffer (written directly) or a host buffer (copied; the call then synchronises).  counts: host or device
 * like `out`, may be NULL for SA_DENSE. */
SA_API int sa_reproject(sa_handle* h, int n, int mode, float* out, int* counts, int where, sa_stream stream);
SA_API float* sa_pointcloud_buffer(sa_handle* h);

/* Stage C.  heatmap() of the demo (FastACVNet_plus/test/main.cpp:7-24), per frame, on the disparity as returned
 * (after RAFTStereo's sign flip).  out as for sa_reproject. */
SA_API int sa_heatmap(sa_handle* h, int n, uint8_t* out, int where, sa_stream stream);

/* The disparity as the 8-bit image the demos write: cv::imwrite("disparity.jpg", disparity) on a CV_32FC1 Mat
 * (FastACVNet_plus/test/main.cpp:55, RAFTStereo/test/main.cpp:31) first does convertTo(CV_8U), i.e.
 * saturate_cast<uchar>(cvRound(d)): round half to even, clamp to 0..255, NaN / +-inf / |d| >= 2^31 -> 0.
 * out: uint8[n][h][w] (plane 0 of the disparity), device or host like sa_heatmap.  The JPEG encoding itself is not
 * reproduced; sa_write_pgm (stereo_algorithms_compat.h) stores the bytes losslessly. */
SA_API int sa_disparity_u8(sa_handle* h, int n, uint8_t* out, int where, sa_stream stream);

/* Read-backs (the cudaMemcpy D2H + clone of TRTFastACVNet_plus.cpp:163-168; rectified images are what the
 * reference leaves in the caller's cv::Mat). */
SA_API int sa_copy_rectified(sa_handle* h, int n, int eye, uint8_t* out, int where, sa_stream stream);

/* Compressed capture frames.  The reference's camera delivers MJPG frames, 2W x H with the left and the
 * right eye side by side (usb_test.py:7-11: 1280x480); cv2.VideoCapture decodes each on the CPU and the three-argument
 * RectifyImage overload splits it with two clone()s (FastACVNet_plus_Algorithm.cpp:116-131).  Here the n compressed frames
 * (host pointers, baseline or progressive JPEG, any chroma subsampling, each exactly 2W x H) are decoded on the GPU by nvJPEG
 * into the handle's capture buffer - u8 [n][H][2W][3], BGR, allocated on the first call - in stream order, and stage A reads
 * the two eyes out of it in place (byte offset 3W, row pitch 6W).  nvJPEG is loaded from the CUDA toolkit's libnvjpeg.so.12 on
 * first use: SA_ERR_IO if it is not installed; SA_ERR_PARSE for a damaged stream, SA_ERR_INVALID_ARG for one of another size.
 * The Huffman stage of small batches runs on the calling thread, so the call returns when the GPU part is queued, not earlier.
 * Decoded pixels are nvJPEG's, not libjpeg's (what cv2 uses): measured on 1280x480 frames they lie half a grey level lower on
 * average (nvJPEG truncates where libjpeg rounds), 1 % of the pixels differ by more than 2 and single pixels at chroma edges of
 * 4:2:0 streams by up to 10 (tests/test_gpu_jpeg.py).  Everything after the decode is exact as everywhere else. */
SA_API int sa_decode_capture_jpeg(sa_handle* h, int n, const void* const* jpeg, const size_t* jpeg_bytes, sa_stream stream);
SA_API uint8_t* sa_capture_buffer(sa_handle* h); /* device; NULL before the first decode */
/* The first n decoded frames, dense u8 [n][H][2W][3], to host (synchronises) or device memory. */
SA_API int sa_copy_capture(sa_handle* h, int n, uint8_t* out, int where, sa_stream stream);
/* sa_decode_capture_jpeg + sa_rectify_preprocess(h, n, capture, capture + 3W, 6W, SA_DEVICE, rectify, stream). */
SA_API int sa_rectify_preprocess_jpeg(sa_handle* h, int n, const void* const* jpeg, const size_t* jpeg_bytes, int rectify, sa_stream stream);

/* JPEG files of the 8-bit outputs.  The demos store the disparity and the heat map with
 * cv::imwrite("disparity.jpg", ...) / cv::imwrite("heatmap.jpg", ...) (FastACVNet_plus/test/main.cpp:55,59): baseline JPEG,
 * quality 95, colour 4:2:0.  sa_encode_jpeg does that on the GPU (nvJPEG, same run-time dependency as above) for one image -
 * channels = 1 (grey: sa_disparity_u8's output) or 3 (BGR interleaved: sa_heatmap's, sa_copy_rectified's), host or device memory,
 * `pitch` bytes per row (0 = dense).  The stream is returned in `out` (host; out = NULL: only *out_bytes, the size needed) -
 * nvJPEG's bytes, not libjpeg's: same format and quality setting, not the same file.  Synchronises the stream.
 * sa_write_jpeg = encode + fwrite.  sa_decode_jpeg is the counterpart for ONE image of any size (BGR interleaved out, host or
 * device; out = NULL: only the size), e.g. to read a stored frame back. */
SA_API int sa_encode_jpeg(sa_handle* h, const uint8_t* image, int where, int width, int height, int channels, size_t pitch, int quality, void* out,
                          size_t out_capacity, size_t* out_bytes, sa_stream stream);
SA_API int sa_write_jpeg(sa_handle* h, const char* path, const uint8_t* image, int where, int width, int height, int channels, size_t pitch, int quality);
SA_API int sa_decode_jpeg(sa_handle* h, const void* jpeg, size_t jpeg_bytes, uint8_t* out, int where, size_t pitch, int* width, int* height, sa_stream stream);
SA_API int sa_copy_net_image(sa_handle* h, int n, int eye, uint8_t* out, int where, sa_stream stream); /* resized u8, if resizing */
SA_API int sa_copy_tensor(sa_handle* h, int n, int eye, void* out, int where, sa_stream stream);
SA_API int sa_copy_disparity(sa_handle* h, int n, float* out, int where, sa_stream stream); /* plane 0 -> [n][h][w] */
SA_API int sa_synchronize(sa_handle* h);
/* All read-backs of one Run in one call: the two cudaMemcpy D2H of TRTFastACVNet_plus.cpp:163-164 plus the rectified images
 * the reference leaves in the caller's cv::Mat (FastACVNet_plus_Algorithm.cpp:142-143).  Every requested copy is queued on
 * one stream and the call synchronises ONCE at the end.  HOST destinations; any of them may be NULL.  pointcloud: the dense
 * cloud of the last sa_reproject that kept its result in the internal buffer (out = NULL); disparity: plane 0, [n][h][w];
 * rect_*: rows `*_pitch` bytes apart (0 = 3*W, e.g. cv::Mat::step); heatmap: the result of the last sa_heatmap with out = NULL. */
SA_API int sa_download(sa_handle* h, int n, float* pointcloud, float* disparity, uint8_t* rect_left, size_t rect_left_pitch, uint8_t* rect_right,
                       size_t rect_right_pitch, uint8_t* heatmap, sa_stream stream);
/* The handle's own compute stream (a cudaStream_t; the reference creates one per object with cudaStreamCreate,
 * TRTFastACVNet_plus.cpp:91).  It is what every call uses when its `stream` argument is NULL, and it is created
 * cudaStreamNonBlocking: work a caller queues on the legacy default stream is NOT ordered against it.  Never NULL for a
 * live handle. */
SA_API sa_stream sa_get_stream(sa_handle* h);

/* The reference's whole Run<Model>[_RectifyImage] data flow for n pairs with HOST buffers (TRTFastACVNet_plus.cpp:
 * 119-170): upload, stage A, network slot, stage B, download - pipelined in chunks over copy/compute streams.
 * Any output pointer may be NULL to skip that download.  The network slot runs the engine callback if one is set,
 * else the synthetic stub when stub_seed != 0, else uses the disparity buffer as it is. */
/* The callback replaces context->enqueueV3(stream) (TRTFastACVNet_plus.cpp:152).  `stream` is the stream stage A was queued on
 * and stage B will be queued on - always a real stream, never NULL: enqueue the network on it (or make it wait for an event
 * of your own stream and signal back) and return without synchronising.  Return 0 on success. */
typedef int (*sa_engine_fn)(void* user, sa_stream stream, int first_frame, int n, const void* left_tensor,
                            const void* right_tensor, float* disparity);
SA_API int sa_set_engine(sa_handle* h, sa_engine_fn fn, void* user);
SA_API int sa_run_batch(sa_handle* h, int n, const uint8_t* left, const uint8_t* right, int rectify, unsigned stub_seed,
                        float* pointcloud, float* disparity, uint8_t* rect_left, uint8_t* rect_right, uint8_t* heatmap);

/* One host batch over several GPUs from one process.  The reference has no multi-GPU path (one object, one device:
 * cudaSetDevice in Initialize only, TRTFastACVNet_plus.cpp:45); frames share nothing, so the batch is cut frame-wise:
 * handle g (created on GPU g with max_batch >= ceil(n / n_handles), same sizes / model / tensor type as handle 0, same
 * calibration loaded) gets the contiguous block sa_shard_range(n, g, n_handles) of the n pairs and runs sa_run_batch on it,
 * on its own host thread, reading and writing its slice of the caller's HOST buffers (layouts and NULL rules as for
 * sa_run_batch; use pinned memory - sa_host_alloc - so that the copies of all GPUs really overlap).  Returns when every
 * shard is home; on failure the status of the first failing shard, its index in *failed_handle (may be NULL) and its
 * message in sa_last_error(handles[*failed_handle]).  No collective, no peer traffic.  The synthetic stub numbers frames
 * across shards, so the result equals sa_run_batch of the whole batch on one handle bit for bit. */
SA_API int sa_run_batch_multi(sa_handle* const* handles, int n_handles, int n, const uint8_t* left, const uint8_t* right, int rectify, unsigned stub_seed,
                              float* pointcloud, float* disparity, uint8_t* rect_left, uint8_t* rect_right, uint8_t* heatmap, int* failed_handle);
/* The partition sa_run_batch_multi uses (and bench.py's ranks, stereoalgorithms_b200/sharding.py): blocks of
 * ceil(n_total / world) pairs, the last ones shorter or empty.  Host only. */
SA_API int sa_shard_range(int n_total, int rank, int world, int* begin, int* end);

/* Host-only helpers: usable without a CUDA device.
 * sa_parse_calibration_yaml is the file-reading half of ReadObjectYml (FastACVNet_plus_Algorithm.cpp:98-114): all 11
 * matrices of struct CalibrationParam (TRTFastACVNet_plus.h:30-43), row-major.  err (optional) receives the message.
 * sa_norm_lut gives the 3x256 normalisation table (R,G,B planes) that stage A applies for `model`; sa_resize_table the
 * source index / weight table (s0 s1 c0 c1, each [dst]) of one axis of the 8-bit INTER_LINEAR resize. */
typedef struct sa_calibration {
    double K_l[9], D_l[14], R_L[9], P1[12];
    double K_r[9], D_r[14], R_R[9], P2[12];
    double Q[16], R[9], T[3];
    int nD_l, nD_r;
} sa_calibration;
SA_API int sa_parse_calibration_yaml(const char* path, sa_calibration* out, char* err, size_t err_len);
/* The writing half of the calibration format (Stereo_Calibration/Stereo_Calibration.cpp:165-179 writes it with
 * cv::FileStorage): "%YAML:1.0", "---", then one !!opencv-matrix node per matrix, doubles printed so that they read back
 * bit-identically.  sa_scale_calibration derives the calibration of a resized / cropped camera from an existing one
 * (done by hand in the reference's harnesses): rows 0-1 of K_l, K_r, P1, P2 are multiplied by
 * `scale`, (crop_x, crop_y) pixels are subtracted from the principal points, and Q is rebuilt to match
 * (Q[3] = -cx', Q[7] = -cy', Q[11] = f', Q[15] *= scale; Q[14] = -1/Tx is unchanged).  in == out is allowed. */
SA_API int sa_write_calibration_yaml(const char* path, const sa_calibration* c);
SA_API int sa_scale_calibration(const sa_calibration* in, double scale, double crop_x, double crop_y, sa_calibration* out);
/* Re-rectification: cv::stereoRectify as the reference's calibration tool calls it (Stereo_Calibration/Stereo_Calibration.cpp:162,
 * flags = CALIB_ZERO_DISPARITY, alpha = 0, new size = image size), restated in f64 host code.  From the cameras (K, D with
 * 0..14 coefficients), the image size and the pose R (3x3), T of the right camera it produces R_L, R_R (3x3), P1, P2 (3x4),
 * Q (4x4, may be NULL) and the valid ROIs {x, y, w, h} (may be NULL) for an output of new_width x new_height (0 x 0 = image
 * size).  alpha < 0: keep the focal length; 0: only valid pixels; 1: all source pixels.  sa_rectify_calibration applies it
 * to a parsed calibration in place (uses K_*, D_*, R, T; overwrites R_L, R_R, P1, P2, Q).  Parity with OpenCV 4.13: see
 * csrc/stereo_rectify.cpp. */
#define SA_CALIB_ZERO_DISPARITY 1024 /* = cv::CALIB_ZERO_DISPARITY */
SA_API int sa_stereo_rectify(const double* K_l, const double* D_l, int nD_l, const double* K_r, const double* D_r, int nD_r, int width, int height,
                             const double* R, const double* T, int flags, double alpha, int new_width, int new_height, double* R_L, double* R_R,
                             double* P1, double* P2, double* Q, int* roi_l, int* roi_r);
SA_API int sa_rectify_calibration(sa_calibration* c, int width, int height, int flags, double alpha, int new_width, int new_height, int* roi_l,
                                  int* roi_r);
SA_API int sa_norm_lut(int model, float* lut768);
SA_API int sa_resize_table(int dst, int src, int* table);

/* No counterpart in the reference.  Compacting reprojection (the SA_COMPACT flavour of sa_reproject) for callers that are
 * not allowed to allocate device memory at the point where they need it - a TensorRT plugin's enqueue(), a camera / ROS
 * callback under a "no allocation after start-up" rule, a process whose own pool already owns the whole GPU.  It takes no
 * handle and calls no cudaMalloc / cudaFree or any other allocating routine: the per-CTA counts and offsets the
 * compaction needs live in a 512 KiB device array built into the library (present on every GPU that loads it, until
 * the process exits).
 *   Limit: ceil(w*h / 1024) * n <= 65536 scratch slots per call (218 frames of 640x480, 32 of 1920x1080, 8 of
 *   3840x2160) and n <= 65535 frames; beyond that the call returns SA_ERR_TOO_LARGE and explains itself in `err`.
 *   `points` must be 8-byte aligned (the rows are written as pairs of floats), `disparity` and `counts` 4-byte aligned;
 *   otherwise SA_ERR_INVALID_ARG.
 *   All image/point pointers are DEVICE pointers on the CURRENT device (there is no handle to select one); Q16 is 16
 *   host floats, row-major.  Strides of 0 mean "contiguous" (left_pitch = 3*w, frame strides = one frame).
 *   disparity [n][..] float (sign-flipped in place when negate != 0, like RAFTStereo's kernel), left_bgr the rectified
 *   left images, points [n][w*h][6] with the first counts[i] rows of frame i valid, counts [n] ints on the device.
 *   Work is queued on `stream` (NULL = the default stream); nothing is synchronised.
 *   The first call on each device also looks up the scratch array's device address; make that first call before any
 *   stream capture is started.
 *   All calls share the one scratch array: calls that can overlap on the GPU (different streams, different threads)
 *   race on it and corrupt each other's point counts - serialise them.  The array keeps whatever the previous call
 *   wrote; that is harmless because every slot a call reads is written by that same call first.
 *   err (optional, err_len bytes) receives the failure text. */
SA_API int sa_reproject_compact_noalloc(float* disparity, size_t disparity_frame_stride, const uint8_t* left_bgr, size_t left_pitch,
                                        size_t left_frame_stride, const float* Q16, int w, int h, int n, int negate, float* points, int* counts,
                                        sa_stream stream, char* err, size_t err_len);

/* Pinned host memory for the buffers above (cudaMallocHost of TRTFastACVNet_plus.cpp:98-99). */
SA_API int sa_host_alloc(void** p, size_t bytes);
SA_API int sa_host_free(void* p);
/* Page-lock a buffer the caller already owns (cudaHostRegister), so that copies into it run at PCIe speed instead of through
 * the driver's bounce buffer; unregister before the buffer is freed. */
SA_API int sa_host_register(void* p, size_t bytes);
SA_API int sa_host_unregister(void* p);

/* Observability (the reference only prints a wall-clock "inference time", TRTFastACVNet_plus.cpp:165-166). */
SA_API int sa_stage_time(sa_handle* h, int stage, double* total_ms, long* launches); /* needs SA_FLAG_TIMING; synchronises */
SA_API int sa_reset_timing(sa_handle* h);
SA_API long sa_launch_count(const sa_handle* h); /* kernels launched by this handle so far */
SA_API int sa_set_option(sa_handle* h, const char* key, long value); /* tuning knobs, see DESIGN.md */

#ifdef __cplusplus
}
#endif
#endif /* STEREO_GEOM_H_ */
