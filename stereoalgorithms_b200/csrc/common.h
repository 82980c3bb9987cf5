// Internal declarations shared by the translation units of libstereo_geom.so.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/stereo_geom.h"

namespace sa {

// ---- tiling of the fused rectify kernel -----------------------------------------------------
constexpr int kTileW = 64;          // output pixels per tile row (one warp covers 32 of them)
constexpr int kTileH = 16;          // output rows per tile
constexpr int kRectThreads = 256;   // 4 output pixels per thread
constexpr int kFootprintWords = 3072;  // staged source pixels per tile (BGRx words, 12 KB)

// Bounding box, in source-image pixels, of all bilinear taps of one output tile.  Coordinates are
// "extended": they may reach 2 pixels outside the image, where the staging writes zeros
// (BORDER_CONSTANT 0).  w == 0 marks a tile whose footprint does not fit kFootprintWords: it is
// gathered directly from global memory.
struct TileInfo {
    short x0, y0;      // top-left (x0 is a multiple of 4)
    short w, h;        // extent (w is a multiple of 4)
    uint32_t inv_gw;   // ceil(2^20 / (w/4)): exact division of a group index < 768 by the groups per row
    uint32_t flags;    // kTileInterior: the whole footprint lies inside the image
};
constexpr uint32_t kTileInterior = 1u;

struct MapParams {
    double iR[9];
    double k[12];  // k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4
    double fx, fy, cx, cy;
};

struct StageTimer {
    std::vector<cudaEvent_t> begin[SA_STAGE_COUNT_], end[SA_STAGE_COUNT_];
    size_t used[SA_STAGE_COUNT_] = {0};
};

}  // namespace sa

namespace sa {
struct JpegDecoder;   // nvJPEG state of one handle (csrc/jpeg.cpp)
}

struct sa_handle {
    sa_config cfg{};
    int W = 0, H = 0, w = 0, h = 0;  // image / network size
    bool resize = false;
    int planes = 1;
    int max_n = 0;
    size_t tensor_elt = 4;

    cudaStream_t stream = nullptr, s_h2d = nullptr, s_d2h = nullptr;
    std::vector<cudaEvent_t> ev_pool;

    // calibration
    bool have_calib = false, have_q = false;
    bool have_maps[2] = {false, false};
    double K[2][9]{}, D[2][14]{}, R[2][9]{}, P[2][12]{}, Q[16]{};
    int nD[2] = {0, 0};
    float Qf[16]{};

    // device buffers
    uint8_t* raw[2] = {nullptr, nullptr};     // uploaded camera frames   [N][H][W][3]
    uint8_t* capture = nullptr;               // decoded side-by-side capture frames [N][H][2W][3] (sa_decode_capture_jpeg; allocated on first use)
    sa::JpegDecoder* jpeg = nullptr;
    uint8_t* jpeg_scratch = nullptr;          // device staging for sa_encode_jpeg of host images / sa_decode_jpeg to host memory (grown on demand)
    size_t jpeg_scratch_bytes = 0;
    uint8_t* rect[2] = {nullptr, nullptr};    // rectified frames         [N][H][W][3]
    uint8_t* net_u8[2] = {nullptr, nullptr};  // resized frames (resize only) [N][h][w][3]
    void* tensor[2] = {nullptr, nullptr};     // network inputs; HitNet: tensor[1] = tensor[0] + 3 planes
    float* disp = nullptr;                    // [N][planes][h][w]
    float* points = nullptr;                  // [N][h*w][6]
    uint8_t* heat = nullptr;                  // [N][h][w][3]
    int* counts = nullptr;                    // [N]
    int* map1[2] = {nullptr, nullptr};        // (x | y << 16) int16 pairs, [H][W]
    uint16_t* map2[2] = {nullptr, nullptr};
    sa::TileInfo* tiles[2] = {nullptr, nullptr};
    uint32_t* tmap[2] = {nullptr, nullptr};   // derived per-pixel table of the fast rectify path, [H][W]
    size_t raw_pitch = 0;                     // bytes per row of the uploaded frames: 3 W rounded up to 16
    size_t rect_pitch = 0, rect_frame = 0;    // rectified rows: 3 W bytes when W % 4 == 0, else 3 * roundup(W, 4) rounded up to 16 bytes, so
                                              // that every row starts 16-byte aligned and holds whole 4-pixel (12-byte) groups
    int* fp_max = nullptr;                    // device: {max footprint width, height} over staged full tiles of both eyes
    int box_w = 0, box_h = 0;                 // host copy of fp_max
    int unstaged_tiles = 0;                   // tiles (both eyes) whose footprint does not fit the staging buffers
    float* lut32 = nullptr;                   // [3][256]
    __half* lut16 = nullptr;
    unsigned* minmax = nullptr;               // [N][2] ordered-uint encodings (min, ~max)
    int* blk_counts = nullptr;                // [N][blocks per frame]  (compaction scratch)
    unsigned* blk_minmax = nullptr;           // [N][blocks per frame][2]  (min/max partials written by reproject)
    int* blk_offsets = nullptr;
    unsigned long long* blk_status = nullptr; // [N][blocks per frame] + 1: look-back words of the single-pass compaction, then the tile ticket
    unsigned cmp_epoch = 0;                   // launch counter: a look-back word is valid only if it carries the current launch's epoch
    long opt_single_pass_compact = 0;         // 1 = one-launch decoupled look-back compaction instead of count -> scan -> compact (measured 0.27 against 0.15 ms)
    int* rs_x = nullptr;                      // resize tables: x0 x1 a0 a1 [w], y0 y1 b0 b1 [h]
    int* rs_y = nullptr;
    int* rs_nx = nullptr;                     // fused resize: first network column / row owned by each tile column / row ([tiles + 1])
    int* rs_ny = nullptr;
    int rs_max_net = 0;
    int net_right_frames = 0;                 // frames of net_u8[1] that are current (the fused kernel writes only the left image)

    // state
    const uint8_t* color_src[2] = {nullptr, nullptr};  // what reproject takes colours from / copy_rectified returns
    size_t color_pitch = 0, color_frame = 0;
    int minmax_valid = 0;  // number of frames whose final min/max is current
    int partial_valid = 0; // number of frames whose per-CTA min/max partials are current
    int n_last = 0;
    sa_engine_fn engine = nullptr;
    void* engine_user = nullptr;

    // options
    long opt_frame_fastest = 1;
    long opt_chunk = 8;
    long opt_force_direct = 0;
    long opt_force_generic = 0;
    long opt_exact_div = 0;
    long opt_persistent = 0;          // CTAs per SM of the persistent rectify kernel (0 = one CTA per tile, the default: measured faster)
    long opt_batch_ctas = 0;           // CTAs per SM of the batch kernel; 0 = automatic (launch_rectify_t)
    long opt_force_unaligned = 0;
    long opt_scalar_preprocess = 0;   // preprocess_kernel (one pixel per thread) instead of preprocess4_kernel
    long opt_no_fixed_size = 0;
    long opt_chunk_taper = 50;        // see RectifyArgs::chunk_taper (measured: 1-2 % on the 640x480 and 1242x375 workloads, neutral at 1080p)
    long opt_l2_prefetch = -1;        // see RectifyArgs::l2_prefetch; -1 = automatic
    long opt_bulk = 1;                // 1 = rectify_ws_kernel (tensor copies + warp specialisation) when the rows are 16-byte aligned and W % 16 == 0; 2 = any even W;
                                      // 0 = same as no_align16
    long opt_batch_loop = 1;          // rectify_batch_kernel (default); 0 = one CTA per (tile, eye, frame)
    long opt_separate_preprocess = 0; // resize path: 1 = resize_kernel then preprocess4_kernel (round 1) instead of resize_norm_kernel
    long opt_fused_resize = 0;        // 1 = one-launch rectify + resize + normalise (rectify_ws_kernel<..., RESIZE>) where it applies; 0 = three launches
                                      // (measured on 64 pairs 1080p -> 640x480: 0.79 ms fused against 0.75 ms - the fused kernel is issue-bound)
    long opt_no_align16 = 0;          // 1 = keep the 4-byte-aligned frame loop even where the 16-byte one applies
    long opt_frame_base = 0;          // global index of this handle's frame 0 (sa_run_batch_multi: the stub numbers frames across shards)
    int sm_count = 148;

    long launches = 0;
    sa::StageTimer timer;
    std::string err;
};

namespace sa {

int fail(sa_handle* h, int code, const char* fmt, ...);
// csrc/jpeg.cpp: n JPEG streams of width x height -> BGR frames at dst (device; `pitch` bytes per row, `frame` bytes apart), queued on s
int jpeg_decode_batch(sa_handle* h, int n, const void* const* jpeg, const size_t* bytes, uint8_t* dst, size_t pitch, size_t frame, int width, int height, cudaStream_t s);
void jpeg_release(sa_handle* h);
int jpeg_image_size(sa_handle* h, const void* jpeg, size_t bytes, int* width, int* height);
int jpeg_decode_checked(sa_handle* h, int n, const void* const* jpeg, const size_t* bytes, uint8_t* dst, size_t pitch, size_t frame, cudaStream_t s);
// one 8-bit device image (1 channel grey / 3 channels BGR interleaved) -> baseline JPEG bytes; synchronises s
int jpeg_encode(sa_handle* h, const uint8_t* dev_image, int width, int height, int channels, size_t pitch, int quality, std::vector<unsigned char>& out, cudaStream_t s);
void set_global_error(const char* msg);

#define SA_CUDA(h, expr)                                                                               \
    do {                                                                                               \
        cudaError_t e_ = (expr);                                                                       \
        if (e_ != cudaSuccess) return sa::fail((h), SA_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct StageScope {  // records timing events around a stage when SA_FLAG_TIMING is set
    sa_handle* h;
    int stage;
    cudaStream_t s;
    StageScope(sa_handle* h, int stage, cudaStream_t s);
    ~StageScope();
};

// calib_yaml.cpp
struct YamlMat {
    std::string name;
    int rows = 0, cols = 0;
    std::vector<double> data;
};
int parse_opencv_yaml(const char* path, std::vector<YamlMat>& out, std::string& err);

// kernels (each returns a cudaError_t from the launch)
cudaError_t launch_init_maps(const MapParams& p, int W, int H, int* map1, uint16_t* map2, cudaStream_t s);
cudaError_t launch_tile_info(const int* map1, const uint16_t* map2, int W, int H, TileInfo* tiles, uint32_t* tmap, int* fp_max, int tiles_x, int tiles_y,
                             int force_direct, int halo, cudaStream_t s);

constexpr int kMaxChunks = 32;

struct RectifyArgs {
    const uint8_t* src[2];
    uint8_t* rect[2];
    void* tensor[2];
    const int* map1[2];
    const uint16_t* map2[2];
    const TileInfo* tiles[2];
    const uint32_t* tmap[2];  // per pixel: staged-footprint word offset | fx << 12 | fy << 17 (0xffffffff in unstaged tiles)
    const void* lut;          // float[3][256] or __half[3][256]
    size_t src_pitch, src_frame;       // bytes
    size_t tensor_frame;               // elements between consecutive frames of one eye
    int W, H, n;
    int model, dtype;
    int write_tensor;                  // 0 when a resize follows
    int frame_fastest;
    int force_generic;                 // testing: never take the fast tile path
    int persistent_ctas;               // > 0: persistent pipelined kernel with this many CTAs (a multiple of the SM count)
    int batch_loop;                    // > 0 (= SM count): one CTA per (tile, eye) walks a run of frames (rectify_batch_kernel)
    int bulk_copy;                     // option bulk: 0 = never the tensor-copy kernel (arrives as no_align16), 1 = where W % 16 == 0, 2 = any even width with 16-byte-aligned rows
    int no_fixed_size;                 // batch kernel: do not use the 640x480 instantiation (tests, tuning)
    int chunk_taper;                   // batch kernel: percent by which the first / last frame chunk is longer / shorter than the average (0 = equal chunks)
    int chunk_tab;                     // 1: chunk c covers frames [chunk_begin[c], chunk_begin[c + 1]) (filled in by launch_rectify_t)
    int chunk_begin[kMaxChunks + 1];
    size_t rect_pitch, rect_frame;     // bytes per row / per frame of the rectified output (rows padded, see sa_handle::rect_pitch)
    int force_unaligned;               // batch kernel: take the 2-byte-aligned code path even for 4-byte aligned frames (tests, tuning)
    int l2_prefetch;                   // batch kernel: prefetch.global.L2 of the frame this many iterations ahead (0 = off, < 0 = automatic)
    int box_w, box_h;                  // largest staged footprint of the calibration (pixels, rows), both eyes
    int unstaged_tiles;                // tiles whose footprint is not staged (bounds-checked path)
    int batch_ctas_per_sm;             // target number of CTAs per SM for the frame chunking of the batch kernel
    int no_align16;                    // batch kernel: do not take the 16-byte-aligned frame loop (tests, tuning)
    // fused resize (rectify at W x H, tensors at rs_w x rs_h; launch_rectify reports through *fused whether it took this path)
    int rs_w, rs_h, rs_area2x;         // network size (0 = no resize); 1 = exact 2x decimation (rounded 2x2 mean)
    const int* rs_tab_x;               // x0 x1 a0 a1, each [rs_w]   (build_resize_table)
    const int* rs_tab_y;               // y0 y1 b0 b1, each [rs_h]
    const int* rs_nx;                  // [tiles_x + 1]: first network column whose left tap x0 lies in tile column i or beyond
    const int* rs_ny;                  // [tiles_y + 1]: same for rows
    int rs_max_net;                    // most network pixels any tile owns
    uint8_t* rs_net_left;              // resized LEFT image [n][rs_h][rs_w][3] (stage B's colours); the right one is not written
    int fused_resize;                  // 1: take the one-launch rectify + resize + normalise kernel where it applies (default: three launches, measured faster)
};
cudaError_t launch_rectify(const RectifyArgs& a, cudaStream_t s, long* launches, int* fused = nullptr);

struct PreprocArgs {  // rectify = 0 path, and the second half of the resize path
    const uint8_t* src[2];
    void* tensor[2];
    const void* lut;
    size_t src_pitch, src_frame, tensor_frame;
    int W, H, n, model, dtype;
    int force_scalar;   // 1: one pixel per thread even where the four-pixel kernel applies (tests)
};
cudaError_t launch_preprocess(const PreprocArgs& a, cudaStream_t s, long* launches);

struct ResizeArgs {
    const uint8_t* src[2];   // rectified [n][H][W][3]
    uint8_t* dst[2];         // [n][h][w][3]
    const int* tab_x;        // x0 x1 a0 a1, each [w]
    const int* tab_y;        // y0 y1 b0 b1, each [h]
    size_t src_pitch, src_frame;
    int W, H, w, h, n, area2x;
    int one_eye;             // 1: src[0] / dst[0] only (n frames), not both eyes
    // optional fused normalisation (resize_norm_kernel): tensors written straight from the resized values, u8 image only for eye 0
    void* tensor[2];         // null: plain u8 resize of both eyes
    size_t tensor_frame;     // elements between consecutive frames of one eye
    const void* lut;         // float[3][256] or __half[3][256]
    int dtype, ident;
};
cudaError_t launch_resize(const ResizeArgs& a, cudaStream_t s, long* launches);

struct ReprojectArgs {
    float* disp;              // frame f at disp + f*disp_frame (plane 0)
    size_t disp_frame;        // floats
    const uint8_t* color;     // BGR
    size_t color_pitch, color_frame;
    float* points;            // frame f at points + f*w*h*6
    float q[16];
    int w, h, n;
    int negate;               // RAFTStereo
    int compact;
    int exact_div;            // 1: IEEE divisions (bit-identical to the reference arithmetic), 0: shared-reciprocal quotients
    unsigned* blk_minmax;     // scratch [n][nblk][2]: per-CTA (min, ~max) of the disparity as used; may be null
    int* blk_counts;          // scratch [n][nblk]
    int* blk_offsets;         // scratch [n][nblk]
    int* counts;              // [n] device
    unsigned long long* status;   // single-pass compaction: [n][nblk] look-back words (null: three-launch compaction)
    unsigned long long* ticket;   // ... the tile ticket (zeroed by the launcher)
    unsigned epoch;               // ... and this launch's epoch (> 0)
};
cudaError_t launch_reproject(const ReprojectArgs& a, cudaStream_t s, long* launches);
int reproject_blocks_per_frame(int w, int h);
cudaError_t launch_reduce_minmax(const unsigned* blk_minmax, int nblk, int n, unsigned* minmax, cudaStream_t s, long* launches);

cudaError_t launch_minmax(const float* disp, size_t disp_frame, int npix, int n, unsigned* minmax, cudaStream_t s, long* launches);
constexpr int kHeatFoldMaxBlocks = 1024;   // up to this many stage B partials per frame every heat map CTA folds them itself (640x480: 300)
cudaError_t launch_heatmap(const float* disp, size_t disp_frame, int npix, int n, unsigned* minmax, const unsigned* blk_minmax, int nblk, uint8_t* out, cudaStream_t s,
                           long* launches);
cudaError_t launch_stub(float* disp, size_t disp_frame, int planes, int w, int h, int first_frame, int n, unsigned seed, int negate,
                        cudaStream_t s, long* launches);
cudaError_t launch_disparity_u8(const float* disp, size_t disp_frame, size_t npix, int n, uint8_t* out, cudaStream_t s, long* launches);
cudaError_t launch_copy_plane0(const float* disp, size_t disp_frame, size_t npix, int n, float* out, cudaStream_t s, long* launches);

void build_norm_lut(int model, float* lut /*[3][256]*/);
void build_resize_table(int D, int S, int* s0, int* s1, int* c0, int* c1);

}  // namespace sa
