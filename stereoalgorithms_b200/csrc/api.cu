// C ABI of libstereo_geom.so (include/stereo_geom.h): handle, calibration, buffers, stage dispatch, copies.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>

#include "common.h"

namespace sa {

static thread_local std::string g_err;

void set_global_error(const char* msg) { g_err = msg ? msg : ""; }

int fail(sa_handle* h, int code, const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf; else g_err = buf;
    return code;
}

StageScope::StageScope(sa_handle* h_, int stage_, cudaStream_t s_) : h(h_), stage(stage_), s(s_)
{
    if (!(h->cfg.flags & SA_FLAG_TIMING)) return;
    StageTimer& t = h->timer;
    size_t i = t.used[stage];
    if (i >= t.begin[stage].size()) {
        cudaEvent_t b = nullptr, e = nullptr;
        if (cudaEventCreate(&b) != cudaSuccess || cudaEventCreate(&e) != cudaSuccess) return;
        t.begin[stage].push_back(b);
        t.end[stage].push_back(e);
    }
    cudaEventRecord(t.begin[stage][i], s);
}
StageScope::~StageScope()
{
    if (!(h->cfg.flags & SA_FLAG_TIMING)) return;
    StageTimer& t = h->timer;
    size_t i = t.used[stage];
    if (i < t.end[stage].size()) { cudaEventRecord(t.end[stage][i], s); t.used[stage] = i + 1; }
}

// Normalisation tables, built with the exact type promotions of the reference kernel bodies:
//   RAFTStereo / CREStereo: float c = v            (RAFTStereo/src/stereo_preprocess.cu:17-19, normalise lines commented out)
//   HitNet:                 (float)v / 255.0       (HitNet/src/HitNet_preprocess.cu:19-25; the literal is a double)
//   FastACVNet_plus:        (c/255.0f - mean)/std  (FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:27-29; mean/std are doubles)
// Output channel order is R, G, B.
void build_norm_lut(int model, float* lut)
{
    static const double mean[3] = {0.485, 0.456, 0.406}, stdv[3] = {0.229, 0.224, 0.225};
    for (int ch = 0; ch < 3; ch++)
        for (int v = 0; v < 256; v++) {
            float c = (float)v, o;
            switch (model) {
                case SA_MODEL_HITNET: o = (float)((double)c / 255.0); break;
                case SA_MODEL_FASTACVNET_PLUS: { float t = c / 255.0f; o = (float)(((double)t - mean[ch]) / stdv[ch]); break; }
                default: o = c;
            }
            lut[ch * 256 + v] = o;
        }
}

// cv::resize's 8-bit INTER_LINEAR coefficient tables
void build_resize_table(int D, int S, int* s0, int* s1, int* c0, int* c1)
{
    const double scale = (double)S / D;
    for (int d = 0; d < D; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= S - 1) { s = S - 1; f = 0.f; }
        s0[d] = s;
        s1[d] = s + 1 < S ? s + 1 : S - 1;
        c0[d] = (int)nearbyintf((1.f - f) * 2048.f);
        c1[d] = (int)nearbyintf(f * 2048.f);
    }
}

static void inv3(const double a[9], double t[9])
{
    double d = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
    if (d == 0.0) { for (int i = 0; i < 9; i++) t[i] = 0.0; return; }
    d = 1.0 / d;
    t[0] = (a[4] * a[8] - a[5] * a[7]) * d; t[1] = (a[2] * a[7] - a[1] * a[8]) * d; t[2] = (a[1] * a[5] - a[2] * a[4]) * d;
    t[3] = (a[5] * a[6] - a[3] * a[8]) * d; t[4] = (a[0] * a[8] - a[2] * a[6]) * d; t[5] = (a[2] * a[3] - a[0] * a[5]) * d;
    t[6] = (a[3] * a[7] - a[4] * a[6]) * d; t[7] = (a[1] * a[6] - a[0] * a[7]) * d; t[8] = (a[0] * a[4] - a[1] * a[3]) * d;
}

static int tiles_x(const sa_handle* h) { return (h->W + kTileW - 1) / kTileW; }
static int tiles_y(const sa_handle* h) { return (h->H + kTileH - 1) / kTileH; }

// Largest staged footprint over both eyes: the box of the TMA rectify path.  Recomputed from the tile tables that exist.
static int refresh_box(sa_handle* h)
{
    int mx[3] = {0, 0, 0};
    SA_CUDA(h, cudaMemcpy(mx, h->fp_max, sizeof mx, cudaMemcpyDeviceToHost));
    h->box_w = mx[0];
    h->box_h = mx[1];
    h->unstaged_tiles = mx[2];
    return SA_OK;
}

static int compute_maps(sa_handle* h)
{
    SA_CUDA(h, cudaMemsetAsync(h->fp_max, 0, 3 * sizeof(int), h->stream));
    for (int eye = 0; eye < 2; eye++) {
        MapParams p{};
        double A[9];
        const double* P = h->P[eye];
        const double* R = h->R[eye];
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) A[i * 3 + j] = P[i * 4 + 0] * R[0 * 3 + j] + P[i * 4 + 1] * R[1 * 3 + j] + P[i * 4 + 2] * R[2 * 3 + j];
        inv3(A, p.iR);
        for (int i = 0; i < 12; i++) p.k[i] = h->D[eye][i];
        p.fx = h->K[eye][0]; p.fy = h->K[eye][4]; p.cx = h->K[eye][2]; p.cy = h->K[eye][5];
        {
            StageScope t(h, SA_STAGE_MAPS, h->stream);
            SA_CUDA(h, launch_init_maps(p, h->W, h->H, h->map1[eye], h->map2[eye], h->stream));
            SA_CUDA(h, launch_tile_info(h->map1[eye], h->map2[eye], h->W, h->H, h->tiles[eye], h->tmap[eye], h->fp_max, tiles_x(h), tiles_y(h), (int)h->opt_force_direct, h->resize ? 1 : 0, h->stream));
            h->launches += 2;
        }
        h->have_maps[eye] = true;
    }
    SA_CUDA(h, cudaStreamSynchronize(h->stream));
    return refresh_box(h);
}

static int set_q(sa_handle* h, const double* Q)
{
    for (int i = 0; i < 16; i++) { h->Q[i] = Q[i]; h->Qf[i] = (float)Q[i]; }
    h->have_q = true;
    return SA_OK;
}

template <typename T>
static cudaError_t dev_alloc(T** p, size_t count) { return cudaMalloc((void**)p, count * sizeof(T) > 0 ? count * sizeof(T) : 16); }

static cudaStream_t pick(sa_handle* h, sa_stream s) { return s ? (cudaStream_t)s : h->stream; }

// Row pitch of host frames on the device.  Dense - one linear copy per chunk - unless padding the rows to 16 bytes is what lets
// the tensor-copy rectify kernel run: widths that are not a multiple of 16 with option bulk = 2 (a 2-D copy of 3726-byte rows
// costs the 1242x375 end-to-end path 20 % of its PCIe rate, so it is not done for nothing).
static size_t upload_pitch(const sa_handle* h, int rectify)
{
    const size_t row = (size_t)h->W * 3;
    return (rectify && h->opt_bulk >= 2 && h->W % 2 == 0) ? h->raw_pitch : row;
}

// where == host: async copy + (for device->host results) a synchronise so the caller may read at once.
static int copy_out(sa_handle* h, void* dst, const void* src, size_t bytes, int where, cudaStream_t s)
{
    if (where == SA_HOST) {
        SA_CUDA(h, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s));
        SA_CUDA(h, cudaStreamSynchronize(s));
    } else {
        SA_CUDA(h, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, s));
    }
    return SA_OK;
}

#define SA_ENTER(h)                                                        \
    if (!(h)) return SA_ERR_INVALID_ARG;                                   \
    {                                                                      \
        cudaError_t e0_ = cudaSetDevice((h)->cfg.gpu_id);                  \
        if (e0_ != cudaSuccess) return sa::fail((h), SA_ERR_CUDA, "cudaSetDevice(%d): %s", (h)->cfg.gpu_id, cudaGetErrorString(e0_)); \
    }

#define SA_CHECK_N(h, n)                                                                                    \
    if ((n) < 0) return sa::fail((h), SA_ERR_INVALID_ARG, "n = %d is negative", (n));                          \
    if ((n) > (h)->max_n) return sa::fail((h), SA_ERR_TOO_LARGE, "n = %d exceeds max_batch = %d", (n), (h)->max_n);

// ---- stage bodies over a frame range [f0, f0 + n), shared by the public calls and sa_run_batch ---------------
static int stage_rectify(sa_handle* h, int f0, int n, const uint8_t* src_l, const uint8_t* src_r, size_t pitch, size_t frame_bytes, int rectify,
                         cudaStream_t s)
{
    // src_l / src_r point at frame f0 already
    const size_t rect_frame = h->rect_frame;
    const size_t tplane = (size_t)h->w * h->h;
    const bool hit = h->cfg.model == SA_MODEL_HITNET;
    const size_t tframe = hit ? 6 * tplane : 3 * tplane;
    void* t_l = (char*)h->tensor[0] + (size_t)f0 * tframe * h->tensor_elt;
    void* t_r = (char*)h->tensor[1] + (size_t)f0 * tframe * h->tensor_elt;
    const void* lut = h->cfg.tensor_dtype == SA_F16 ? (const void*)h->lut16 : (const void*)h->lut32;

    const uint8_t* net_src[2];
    size_t net_pitch, net_frame;
    int fused = 0;                    // 1: the rectify launch also resized and wrote the tensors (+ the resized left image)
    if (rectify) {
        if (!h->have_maps[0] || !h->have_maps[1]) return fail(h, SA_ERR_NOT_READY, "no calibration / maps loaded (sa_load_calibration_yaml, sa_set_calibration or sa_set_maps)");
        RectifyArgs a{};
        a.src[0] = src_l; a.src[1] = src_r;
        a.rect[0] = h->rect[0] + (size_t)f0 * rect_frame; a.rect[1] = h->rect[1] + (size_t)f0 * rect_frame;
        a.tensor[0] = t_l; a.tensor[1] = t_r;
        for (int e = 0; e < 2; e++) { a.map1[e] = h->map1[e]; a.map2[e] = h->map2[e]; a.tiles[e] = h->tiles[e]; a.tmap[e] = h->tmap[e]; }
        a.lut = lut;
        a.src_pitch = pitch; a.src_frame = frame_bytes; a.tensor_frame = tframe;
        a.rect_pitch = h->rect_pitch; a.rect_frame = h->rect_frame;
        a.W = h->W; a.H = h->H; a.n = n; a.model = h->cfg.model; a.dtype = h->cfg.tensor_dtype;
        a.write_tensor = h->resize ? 0 : 1;
        a.frame_fastest = (int)h->opt_frame_fastest;
        a.force_generic = (int)h->opt_force_generic;
        a.persistent_ctas = (int)(h->opt_persistent > 0 ? h->opt_persistent * h->sm_count : 0);
        a.batch_loop = h->opt_batch_loop ? h->sm_count : 0;
        a.batch_ctas_per_sm = (int)h->opt_batch_ctas;
        a.bulk_copy = (int)h->opt_bulk;
        a.l2_prefetch = (int)h->opt_l2_prefetch;
        a.chunk_taper = (int)h->opt_chunk_taper;
        a.force_unaligned = (int)h->opt_force_unaligned;
        a.no_fixed_size = (int)h->opt_no_fixed_size;
        a.no_align16 = (h->opt_no_align16 || !h->opt_bulk) ? 1 : 0;
        a.box_w = h->box_w; a.box_h = h->box_h;
        a.unstaged_tiles = h->unstaged_tiles;
        if (h->resize) {
            a.rs_w = h->w; a.rs_h = h->h;
            a.rs_area2x = (h->W == 2 * h->w && h->H == 2 * h->h) ? 1 : 0;
            a.rs_tab_x = h->rs_x; a.rs_tab_y = h->rs_y; a.rs_nx = h->rs_nx; a.rs_ny = h->rs_ny;
            a.rs_max_net = h->rs_max_net;
            a.rs_net_left = h->net_u8[0] + (size_t)f0 * tplane * 3;
            a.fused_resize = (int)h->opt_fused_resize;
        }
        {
            StageScope t(h, SA_STAGE_RECTIFY, s);
            SA_CUDA(h, launch_rectify(a, s, &h->launches, &fused));
        }
        net_src[0] = a.rect[0]; net_src[1] = a.rect[1];
        net_pitch = h->rect_pitch; net_frame = rect_frame;
    } else {
        net_src[0] = src_l; net_src[1] = src_r;
        net_pitch = pitch; net_frame = frame_bytes;
    }
    if (f0 == 0) {  // what reproject colours come from, and what sa_copy_rectified returns
        h->color_src[0] = rectify ? h->rect[0] : src_l;
        h->color_src[1] = rectify ? h->rect[1] : src_r;
        h->color_pitch = net_pitch; h->color_frame = net_frame;
    }
    if (h->resize && fused) {
        const size_t nf = tplane * 3;
        if (f0 == 0) { h->color_src[0] = h->net_u8[0]; h->color_src[1] = nullptr; h->color_pitch = (size_t)h->w * 3; h->color_frame = nf; h->net_right_frames = 0; }
    } else if (h->resize) {
        const size_t nf = tplane * 3;
        if (f0 == 0) h->net_right_frames = 0;
        h->net_right_frames = std::max(h->net_right_frames, f0 + n);
        ResizeArgs r{};
        r.src[0] = net_src[0]; r.src[1] = net_src[1];
        r.dst[0] = h->net_u8[0] + (size_t)f0 * nf; r.dst[1] = h->net_u8[1] + (size_t)f0 * nf;
        r.tab_x = h->rs_x; r.tab_y = h->rs_y;
        r.src_pitch = net_pitch; r.src_frame = net_frame;
        r.W = h->W; r.H = h->H; r.w = h->w; r.h = h->h; r.n = n;
        r.area2x = (h->W == 2 * h->w && h->H == 2 * h->h) ? 1 : 0;
        PreprocArgs p{};
        p.src[0] = r.dst[0]; p.src[1] = r.dst[1];
        p.tensor[0] = t_l; p.tensor[1] = t_r; p.lut = lut;
        p.src_pitch = (size_t)h->w * 3; p.src_frame = nf; p.tensor_frame = tframe;
        p.W = h->w; p.H = h->h; p.n = n; p.model = h->cfg.model; p.dtype = h->cfg.tensor_dtype;
        p.force_scalar = (int)h->opt_scalar_preprocess;
        StageScope t(h, SA_STAGE_RESIZE, s);
        if (h->opt_separate_preprocess) {
            SA_CUDA(h, launch_resize(r, s, &h->launches));
            SA_CUDA(h, launch_preprocess(p, s, &h->launches));
            if (f0 == 0) { h->color_src[0] = h->net_u8[0]; h->color_src[1] = h->net_u8[1]; h->color_pitch = (size_t)h->w * 3; h->color_frame = nf; }
        } else {
            // resize + normalise in one launch: tensors straight from the resized values, the u8 image for the left eye only
            r.tensor[0] = t_l; r.tensor[1] = t_r; r.tensor_frame = tframe; r.lut = lut;
            r.dtype = h->cfg.tensor_dtype;
            r.ident = (h->cfg.model == SA_MODEL_RAFTSTEREO || h->cfg.model == SA_MODEL_CRESTEREO) ? 1 : 0;
            SA_CUDA(h, launch_resize(r, s, &h->launches));
            h->net_right_frames = 0;                                   // the right resized image is made on request (sa_copy_net_image)
            if (f0 == 0) { h->color_src[0] = h->net_u8[0]; h->color_src[1] = nullptr; h->color_pitch = (size_t)h->w * 3; h->color_frame = nf; }
        }
    } else if (!rectify) {
        PreprocArgs p{};
        p.src[0] = src_l; p.src[1] = src_r; p.tensor[0] = t_l; p.tensor[1] = t_r; p.lut = lut;
        p.src_pitch = pitch; p.src_frame = frame_bytes; p.tensor_frame = tframe;
        p.W = h->W; p.H = h->H; p.n = n; p.model = h->cfg.model; p.dtype = h->cfg.tensor_dtype;
        p.force_scalar = (int)h->opt_scalar_preprocess;
        StageScope t(h, SA_STAGE_RECTIFY, s);
        SA_CUDA(h, launch_preprocess(p, s, &h->launches));
    }
    return SA_OK;
}

static size_t disp_frame_floats(const sa_handle* h) { return (size_t)h->planes * h->w * h->h; }

static int stage_reproject(sa_handle* h, int f0, int n, int mode, float* dev_out /* frame f0 */, int* dev_counts /* frame f0 */, cudaStream_t s)
{
    if (!h->have_q) return fail(h, SA_ERR_NOT_READY, "no Q matrix (load a calibration or call sa_set_Q)");
    if (!h->color_src[0]) return fail(h, SA_ERR_NOT_READY, "no left image on the device (run sa_rectify_preprocess first)");
    if (!dev_out) return fail(h, SA_ERR_INVALID_ARG, "reproject: no destination for the points (handle created with SA_FLAG_NO_POINT_BUF and no device `out`)");
    const size_t npix = (size_t)h->w * h->h;
    ReprojectArgs a{};
    a.disp = h->disp + (size_t)f0 * disp_frame_floats(h);
    a.disp_frame = disp_frame_floats(h);
    a.color = h->color_src[0] + (size_t)f0 * h->color_frame;
    a.color_pitch = h->color_pitch; a.color_frame = h->color_frame;
    a.points = dev_out;
    for (int i = 0; i < 16; i++) a.q[i] = h->Qf[i];
    a.w = h->w; a.h = h->h; a.n = n;
    a.negate = h->cfg.model == SA_MODEL_RAFTSTEREO;
    a.compact = mode == SA_COMPACT;
    a.exact_div = (int)h->opt_exact_div;
    const int nblk = reproject_blocks_per_frame(h->w, h->h);
    a.blk_minmax = h->blk_minmax + 2 * (size_t)f0 * nblk;
    a.blk_counts = h->blk_counts + (size_t)f0 * nblk;
    a.blk_offsets = h->blk_offsets + (size_t)f0 * nblk;
    if (mode == SA_COMPACT && h->opt_single_pass_compact) {
        a.status = h->blk_status + (size_t)f0 * nblk;
        a.ticket = h->blk_status + (size_t)h->max_n * nblk;
        a.epoch = ++h->cmp_epoch;
        if (h->cmp_epoch >= (1u << 30)) h->cmp_epoch = 0;  // 30 bits in the word; the wrap needs a clean slate
        if (h->cmp_epoch == 0) { SA_CUDA(h, cudaMemsetAsync(h->blk_status, 0, ((size_t)h->max_n * nblk + 1) * sizeof(unsigned long long), s)); a.epoch = h->cmp_epoch = 1; }
    }
    a.counts = dev_counts;
    (void)npix;
    {
        StageScope t(h, SA_STAGE_REPROJECT, s);
        SA_CUDA(h, launch_reproject(a, s, &h->launches));
    }
    return SA_OK;
}

// minmax_state: 2 = the final per-frame min/max is current, 1 = only reproject's per-CTA partials are, 0 = neither
static int stage_heatmap(sa_handle* h, int f0, int n, uint8_t* dev_out /* frame f0 */, int minmax_state, cudaStream_t s)
{
    const int npix = h->w * h->h;
    const float* d = h->disp + (size_t)f0 * disp_frame_floats(h);
    unsigned* mm = h->minmax + 2 * (size_t)f0;
    StageScope t(h, SA_STAGE_HEATMAP, s);
    const int nblk = reproject_blocks_per_frame(h->w, h->h);
    const unsigned* partials = nullptr;
    if (minmax_state == 1 && nblk <= kHeatFoldMaxBlocks) {
        partials = h->blk_minmax + 2 * (size_t)f0 * nblk;              // folded inside the heat map kernel: one launch, not two
    } else if (minmax_state == 1) {
        SA_CUDA(h, launch_reduce_minmax(h->blk_minmax + 2 * (size_t)f0 * nblk, nblk, n, mm, s, &h->launches));
    } else if (minmax_state == 0) {
        SA_CUDA(h, cudaMemsetAsync(mm, 0xff, sizeof(unsigned) * 2 * (size_t)n, s));
        SA_CUDA(h, launch_minmax(d, disp_frame_floats(h), npix, n, mm, s, &h->launches));
    }
    SA_CUDA(h, launch_heatmap(d, disp_frame_floats(h), npix, n, mm, partials, nblk, dev_out, s, &h->launches));
    return SA_OK;
}

}  // namespace sa

using namespace sa;

// =============================================================================================================
extern "C" {

const char* sa_version(void) { return "stereo_geom_b200 0.1 (sm_100a)"; }

const char* sa_last_error(const sa_handle* h) { return h ? h->err.c_str() : sa::g_err.c_str(); }

int sa_create(const sa_config* cfg, sa_handle** out)
{
    if (!cfg || !out) return fail(nullptr, SA_ERR_INVALID_ARG, "sa_create: null argument");
    *out = nullptr;
    if (cfg->struct_size != (int)sizeof(sa_config)) return fail(nullptr, SA_ERR_INVALID_ARG, "sa_create: struct_size %d != %d", cfg->struct_size, (int)sizeof(sa_config));
    if (cfg->width <= 0 || cfg->height <= 0 || cfg->width > 16384 || cfg->height > 16384)
        return fail(nullptr, SA_ERR_INVALID_ARG, "sa_create: image size %dx%d out of range (1..16384)", cfg->width, cfg->height);
    if (cfg->model < 0 || cfg->model > 3) return fail(nullptr, SA_ERR_INVALID_ARG, "sa_create: unknown model %d", cfg->model);
    if (cfg->tensor_dtype != SA_F32 && cfg->tensor_dtype != SA_F16) return fail(nullptr, SA_ERR_INVALID_ARG, "sa_create: unknown tensor dtype %d", cfg->tensor_dtype);
    if (cfg->max_batch <= 0 || cfg->max_batch > 16384) return fail(nullptr, SA_ERR_INVALID_ARG, "sa_create: max_batch %d out of range (1..16384)", cfg->max_batch);
    const int w = cfg->net_width > 0 ? cfg->net_width : cfg->width, hh = cfg->net_height > 0 ? cfg->net_height : cfg->height;
    if (w > cfg->width || hh > cfg->height) return fail(nullptr, SA_ERR_INVALID_ARG, "sa_create: network size %dx%d larger than the image (upscaling is not supported)", w, hh);

    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(nullptr, SA_ERR_CUDA, "sa_create: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    if (cfg->gpu_id < 0 || cfg->gpu_id >= ndev) return fail(nullptr, SA_ERR_INVALID_ARG, "sa_create: gpu_id %d but %d device(s)", cfg->gpu_id, ndev);
    if ((e = cudaSetDevice(cfg->gpu_id)) != cudaSuccess) return fail(nullptr, SA_ERR_CUDA, "cudaSetDevice(%d): %s", cfg->gpu_id, cudaGetErrorString(e));

    sa_handle* h = new (std::nothrow) sa_handle();
    if (!h) return fail(nullptr, SA_ERR_NOMEM, "sa_create: out of host memory");
    h->cfg = *cfg;
    h->W = cfg->width; h->H = cfg->height; h->w = w; h->h = hh;
    h->resize = (w != h->W || hh != h->H);
    h->planes = cfg->model == SA_MODEL_CRESTEREO ? 2 : 1;
    h->max_n = cfg->max_batch;
    h->tensor_elt = cfg->tensor_dtype == SA_F16 ? 2 : 4;
    { int sms = 0; if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cfg->gpu_id) == cudaSuccess && sms > 0) h->sm_count = sms; }

    const size_t N = (size_t)h->max_n, img = (size_t)h->W * h->H * 3, npix = (size_t)h->w * h->h, mpix = (size_t)h->W * h->H;
    const int ntiles = tiles_x(h) * tiles_y(h);
    const int nblk = reproject_blocks_per_frame(h->w, h->h);
    bool ok = true;
    auto A = [&](cudaError_t r) { if (r != cudaSuccess && ok) { ok = false; e = r; } };
    A(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    A(cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking));
    A(cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking));
    // rectified rows: dense when W is a multiple of 4 (12-byte pixel groups then tile every row); otherwise padded so that
    // every row starts 16-byte aligned and ends on a whole group - the kernels' packed word stores need that, and byte
    // stores into 2-byte-aligned rows cost 1242-wide frames 20 % of stage A.  Readers get the pitch (colour source of stage
    // B, resize, sa_copy_rectified).
    h->rect_pitch = h->W % 4 == 0 ? (size_t)h->W * 3 : ((size_t)((h->W + 3) / 4) * 12 + 15) / 16 * 16;
    h->rect_frame = h->rect_pitch * h->H;
    // uploaded frames: rows padded to 16 bytes, so that host input of any even width (1242 ...) takes the tensor-copy kernel
    h->raw_pitch = ((size_t)h->W * 3 + 15) / 16 * 16;
    for (int eye = 0; eye < 2 && ok; eye++) {
        A(dev_alloc(&h->raw[eye], N * h->raw_pitch * h->H));
        A(dev_alloc(&h->rect[eye], N * h->rect_frame));
        if (h->resize) A(dev_alloc(&h->net_u8[eye], N * npix * 3));
        A(dev_alloc(&h->map1[eye], mpix));
        A(dev_alloc(&h->map2[eye], mpix));
        A(dev_alloc(&h->tiles[eye], (size_t)ntiles));
        A(dev_alloc(&h->tmap[eye], mpix));
    }
    if (ok) {
        // one allocation for both eyes' tensors: [left N x 3 planes][right N x 3 planes], HitNet: N x 6 planes interleaved per frame
        char* t = nullptr;
        A(cudaMalloc((void**)&t, N * 6 * npix * h->tensor_elt));
        h->tensor[0] = t;
        h->tensor[1] = t ? (cfg->model == SA_MODEL_HITNET ? t + 3 * npix * h->tensor_elt : t + N * 3 * npix * h->tensor_elt) : nullptr;
    }
    A(dev_alloc(&h->disp, N * h->planes * npix));
    if (!(cfg->flags & SA_FLAG_NO_POINT_BUF)) A(dev_alloc(&h->points, N * npix * 6));
    if (!(cfg->flags & SA_FLAG_NO_HEATMAP_BUF)) A(dev_alloc(&h->heat, N * npix * 3));
    A(dev_alloc(&h->counts, N));
    A(dev_alloc(&h->fp_max, 4));
    if (ok) A(cudaMemset(h->fp_max, 0, 4 * sizeof(int)));
    A(dev_alloc(&h->minmax, N * 2));
    A(dev_alloc(&h->blk_counts, N * nblk));
    A(dev_alloc(&h->blk_offsets, N * nblk));
    A(dev_alloc(&h->blk_status, N * nblk + 1));
    if (ok) A(cudaMemset(h->blk_status, 0, (N * nblk + 1) * sizeof(unsigned long long)));
    A(dev_alloc(&h->blk_minmax, N * nblk * 2));
    A(dev_alloc(&h->lut32, 768));
    A(dev_alloc(&h->lut16, 768));
    if (h->resize) {
        A(dev_alloc(&h->rs_x, 4 * (size_t)h->w)); A(dev_alloc(&h->rs_y, 4 * (size_t)h->h));
        A(dev_alloc(&h->rs_nx, (size_t)tiles_x(h) + 1)); A(dev_alloc(&h->rs_ny, (size_t)tiles_y(h) + 1));
    }
    if (ok) {
        float lut[768];
        __half lut_h[768];
        build_norm_lut(cfg->model, lut);
        for (int i = 0; i < 768; i++) lut_h[i] = __float2half_rn(lut[i]);
        A(cudaMemcpy(h->lut32, lut, sizeof lut, cudaMemcpyHostToDevice));
        A(cudaMemcpy(h->lut16, lut_h, sizeof lut_h, cudaMemcpyHostToDevice));
        A(cudaMemset(h->disp, 0, N * h->planes * npix * sizeof(float)));
        if (h->resize) {
            std::vector<int> tx(4 * (size_t)h->w), ty(4 * (size_t)h->h);
            build_resize_table(h->w, h->W, &tx[0], &tx[h->w], &tx[2 * h->w], &tx[3 * h->w]);
            build_resize_table(h->h, h->H, &ty[0], &ty[h->h], &ty[2 * h->h], &ty[3 * h->h]);
            A(cudaMemcpy(h->rs_x, tx.data(), tx.size() * sizeof(int), cudaMemcpyHostToDevice));
            A(cudaMemcpy(h->rs_y, ty.data(), ty.size() * sizeof(int), cudaMemcpyHostToDevice));
            // which network pixels a rectified tile owns (fused rectify + resize): those whose top-left tap (x0, y0) lies in it
            auto owners = [](const int* s0, int D, int tiles, int tile, std::vector<int>& first) {
                first.assign((size_t)tiles + 1, D);
                int d = 0;
                for (int t = 0; t <= tiles; t++) {
                    while (d < D && s0[d] < t * tile) d++;                 // s0 is non-decreasing
                    first[t] = d;
                }
                first[tiles] = D;
                int most = 0;
                for (int t = 0; t < tiles; t++) most = std::max(most, first[t + 1] - first[t]);
                return most;
            };
            std::vector<int> nx, ny;
            const int mx = owners(&tx[0], h->w, tiles_x(h), kTileW, nx), my = owners(&ty[0], h->h, tiles_y(h), kTileH, ny);
            h->rs_max_net = mx * my;
            A(cudaMemcpy(h->rs_nx, nx.data(), nx.size() * sizeof(int), cudaMemcpyHostToDevice));
            A(cudaMemcpy(h->rs_ny, ny.data(), ny.size() * sizeof(int), cudaMemcpyHostToDevice));
        }
    }
    if (!ok) {
        int code = (e == cudaErrorMemoryAllocation) ? SA_ERR_NOMEM : SA_ERR_CUDA;
        fail(nullptr, code, "sa_create: %s (max_batch %d at %dx%d needs about %.1f MiB of device memory)", cudaGetErrorString(e), h->max_n, h->W, h->H,
             (double)(N * (4 * img + 6 * npix * h->tensor_elt + npix * (4 * h->planes + 24 + 3))) / 1048576.0);
        cudaGetLastError();
        sa_destroy(h);
        return code;
    }
    *out = h;
    return SA_OK;
}

int sa_destroy(sa_handle* h)
{
    if (!h) return SA_OK;
    cudaSetDevice(h->cfg.gpu_id);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->s_h2d) cudaStreamSynchronize(h->s_h2d);
    if (h->s_d2h) cudaStreamSynchronize(h->s_d2h);
    for (int e = 0; e < 2; e++) {
        cudaFree(h->raw[e]); cudaFree(h->rect[e]); cudaFree(h->net_u8[e]); cudaFree(h->map1[e]); cudaFree(h->map2[e]); cudaFree(h->tiles[e]); cudaFree(h->tmap[e]);
    }
    jpeg_release(h);
    cudaFree(h->capture);
    cudaFree(h->jpeg_scratch);
    cudaFree(h->tensor[0]); cudaFree(h->disp); cudaFree(h->points); cudaFree(h->heat); cudaFree(h->counts); cudaFree(h->minmax);
    cudaFree(h->fp_max); cudaFree(h->blk_counts); cudaFree(h->blk_offsets); cudaFree(h->blk_status); cudaFree(h->blk_minmax); cudaFree(h->lut32); cudaFree(h->lut16); cudaFree(h->rs_x); cudaFree(h->rs_y); cudaFree(h->rs_nx); cudaFree(h->rs_ny);
    for (int s = 0; s < SA_STAGE_COUNT_; s++) {
        for (auto ev : h->timer.begin[s]) cudaEventDestroy(ev);
        for (auto ev : h->timer.end[s]) cudaEventDestroy(ev);
    }
    for (auto ev : h->ev_pool) cudaEventDestroy(ev);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->s_h2d) cudaStreamDestroy(h->s_h2d);
    if (h->s_d2h) cudaStreamDestroy(h->s_d2h);
    delete h;
    return SA_OK;
}

int sa_set_calibration(sa_handle* h, const double* K_l, const double* D_l, int nD_l, const double* R_L, const double* P1, const double* K_r,
                       const double* D_r, int nD_r, const double* R_R, const double* P2, const double* Q)
{
    SA_ENTER(h);
    if (!K_l || !R_L || !P1 || !K_r || !R_R || !P2) return fail(h, SA_ERR_INVALID_ARG, "sa_set_calibration: null matrix");
    if (nD_l < 0 || nD_l > 14 || nD_r < 0 || nD_r > 14 || (nD_l && !D_l) || (nD_r && !D_r)) return fail(h, SA_ERR_INVALID_ARG, "sa_set_calibration: distortion vectors must hold 0..14 coefficients");
    const double* Ks[2] = {K_l, K_r};
    const double* Ds[2] = {D_l, D_r};
    const int nDs[2] = {nD_l, nD_r};
    const double* Rs[2] = {R_L, R_R};
    const double* Ps[2] = {P1, P2};
    for (int e = 0; e < 2; e++) {
        std::memcpy(h->K[e], Ks[e], 9 * sizeof(double));
        std::memset(h->D[e], 0, sizeof h->D[e]);
        for (int i = 0; i < nDs[e]; i++) h->D[e][i] = Ds[e][i];
        h->nD[e] = nDs[e];
        std::memcpy(h->R[e], Rs[e], 9 * sizeof(double));
        std::memcpy(h->P[e], Ps[e], 12 * sizeof(double));
        if (nDs[e] > 12 && (h->D[e][12] != 0.0 || h->D[e][13] != 0.0)) return fail(h, SA_ERR_INVALID_ARG, "sa_set_calibration: tilted sensor model (tauX, tauY) is not supported");
    }
    if (Q) set_q(h, Q);
    h->have_calib = true;
    return compute_maps(h);
}

int sa_parse_calibration_yaml(const char* path, sa_calibration* out, char* err, size_t err_len)
{
    auto bail = [&](int code, const std::string& m) {
        if (err && err_len) { std::snprintf(err, err_len, "%s", m.c_str()); }
        return code;
    };
    if (!path || !out) return bail(SA_ERR_INVALID_ARG, "null argument");
    std::vector<YamlMat> mats;
    std::string perr;
    int rc = parse_opencv_yaml(path, mats, perr);
    if (rc != SA_OK) return bail(rc, perr);
    std::memset(out, 0, sizeof *out);
    struct Want { const char* name; double* dst; int rows, cols; bool required; };
    const Want want[] = {{"intrinsic_left", out->K_l, 3, 3, true}, {"intrinsic_right", out->K_r, 3, 3, true}, {"R_L", out->R_L, 3, 3, true},
                         {"R_R", out->R_R, 3, 3, true},           {"P1", out->P1, 3, 4, true},             {"P2", out->P2, 3, 4, true},
                         {"Q", out->Q, 4, 4, true},               {"R", out->R, 3, 3, false},              {"T", out->T, 3, 1, false}};
    for (const Want& w : want) {
        const YamlMat* m = nullptr;
        for (auto& c : mats) if (c.name == w.name) m = &c;
        if (!m) { if (w.required) return bail(SA_ERR_PARSE, std::string("missing node '") + w.name + "'"); continue; }
        if ((int)m->data.size() != w.rows * w.cols) return bail(SA_ERR_PARSE, std::string("node '") + w.name + "' has the wrong shape");
        std::memcpy(w.dst, m->data.data(), sizeof(double) * m->data.size());
    }
    const char* dn[2] = {"distCoeffs_left", "distCoeffs_right"};
    double* dd[2] = {out->D_l, out->D_r};
    int* dc[2] = {&out->nD_l, &out->nD_r};
    for (int e = 0; e < 2; e++) {
        const YamlMat* m = nullptr;
        for (auto& c : mats) if (c.name == dn[e]) m = &c;
        if (!m) return bail(SA_ERR_PARSE, std::string("missing node '") + dn[e] + "'");
        if (m->data.size() > 14) return bail(SA_ERR_PARSE, std::string("node '") + dn[e] + "' has more than 14 coefficients");
        std::memcpy(dd[e], m->data.data(), sizeof(double) * m->data.size());
        *dc[e] = (int)m->data.size();
    }
    return SA_OK;
}

int sa_write_calibration_yaml(const char* path, const sa_calibration* c)
{
    if (!path || !c) return SA_ERR_INVALID_ARG;
    if (c->nD_l < 0 || c->nD_l > 14 || c->nD_r < 0 || c->nD_r > 14) return SA_ERR_INVALID_ARG;
    FILE* f = std::fopen(path, "w");
    if (!f) return SA_ERR_IO;
    std::fprintf(f, "%%YAML:1.0\n---\n");
    auto node = [&](const char* name, int rows, int cols, const double* d) {
        std::fprintf(f, "%s: !!opencv-matrix\n   rows: %d\n   cols: %d\n   dt: d\n   data: [ ", name, rows, cols);
        int col = 11;
        for (int i = 0; i < rows * cols; i++) {
            char buf[40];
            int len;
            if (d[i] == (double)(long long)d[i] && std::fabs(d[i]) < 1e15) len = std::snprintf(buf, sizeof buf, "%lld.", (long long)d[i]);   // "0." "1." like OpenCV
            else len = std::snprintf(buf, sizeof buf, "%.17g", d[i]);
            if (col + len + 2 > 78) { std::fprintf(f, "\n       "); col = 7; }
            std::fprintf(f, "%s%s", buf, i + 1 < rows * cols ? ", " : " ]\n");
            col += len + 2;
        }
    };
    node("intrinsic_left", 3, 3, c->K_l);
    node("distCoeffs_left", 1, c->nD_l, c->D_l);
    node("intrinsic_right", 3, 3, c->K_r);
    node("distCoeffs_right", 1, c->nD_r, c->D_r);
    node("R", 3, 3, c->R);
    node("T", 3, 1, c->T);
    node("R_L", 3, 3, c->R_L);
    node("R_R", 3, 3, c->R_R);
    node("P1", 3, 4, c->P1);
    node("P2", 3, 4, c->P2);
    node("Q", 4, 4, c->Q);
    return std::fclose(f) == 0 ? SA_OK : SA_ERR_IO;
}

int sa_scale_calibration(const sa_calibration* in, double scale, double crop_x, double crop_y, sa_calibration* out)
{
    if (!in || !out || !(scale > 0.0)) return SA_ERR_INVALID_ARG;
    sa_calibration c = *in;
    double* k33[2] = {c.K_l, c.K_r};
    for (double* K : k33) {
        for (int i = 0; i < 6; i++) K[i] *= scale;       // rows 0 and 1
        K[2] -= crop_x;
        K[5] -= crop_y;
    }
    double* p34[2] = {c.P1, c.P2};
    for (double* P : p34) {
        for (int i = 0; i < 8; i++) P[i] *= scale;
        P[2] -= crop_x;
        P[6] -= crop_y;
    }
    c.Q[3] = -c.P1[2];
    c.Q[7] = -c.P1[6];
    c.Q[11] = c.P1[0];
    c.Q[15] *= scale;
    *out = c;
    return SA_OK;
}

int sa_norm_lut(int model, float* lut768)
{
    if (!lut768 || model < 0 || model > 3) return SA_ERR_INVALID_ARG;
    build_norm_lut(model, lut768);
    return SA_OK;
}

int sa_resize_table(int dst, int src, int* table)
{
    if (!table || dst <= 0 || src <= 0) return SA_ERR_INVALID_ARG;
    build_resize_table(dst, src, table, table + dst, table + 2 * dst, table + 3 * dst);
    return SA_OK;
}

int sa_load_calibration_yaml(sa_handle* h, const char* path)
{
    SA_ENTER(h);
    if (!path) return fail(h, SA_ERR_INVALID_ARG, "sa_load_calibration_yaml: null path");
    sa_calibration c;
    char msg[512] = {0};
    int rc = sa_parse_calibration_yaml(path, &c, msg, sizeof msg);
    if (rc != SA_OK) return fail(h, rc, "%s: %s", path, msg);
    return sa_set_calibration(h, c.K_l, c.D_l, c.nD_l, c.R_L, c.P1, c.K_r, c.D_r, c.nD_r, c.R_R, c.P2, c.Q);
}

int sa_set_Q(sa_handle* h, const double* Q16)
{
    SA_ENTER(h);
    if (!Q16) return fail(h, SA_ERR_INVALID_ARG, "sa_set_Q: null");
    return set_q(h, Q16);
}

int sa_get_Q(sa_handle* h, double* Q16)
{
    SA_ENTER(h);
    if (!Q16) return fail(h, SA_ERR_INVALID_ARG, "sa_get_Q: null");
    if (!h->have_q) return fail(h, SA_ERR_NOT_READY, "no Q matrix set");
    std::memcpy(Q16, h->Q, sizeof h->Q);
    return SA_OK;
}

int sa_set_maps(sa_handle* h, int eye, const int16_t* map1, const uint16_t* map2)
{
    SA_ENTER(h);
    if ((eye != 0 && eye != 1) || !map1 || !map2) return fail(h, SA_ERR_INVALID_ARG, "sa_set_maps: bad argument");
    const size_t mpix = (size_t)h->W * h->H;
    SA_CUDA(h, cudaMemcpyAsync(h->map1[eye], map1, mpix * 4, cudaMemcpyHostToDevice, h->stream));
    SA_CUDA(h, cudaMemcpyAsync(h->map2[eye], map2, mpix * 2, cudaMemcpyHostToDevice, h->stream));
    SA_CUDA(h, launch_tile_info(h->map1[eye], h->map2[eye], h->W, h->H, h->tiles[eye], h->tmap[eye], h->fp_max, tiles_x(h), tiles_y(h), (int)h->opt_force_direct, h->resize ? 1 : 0, h->stream));
    h->launches += 1;
    SA_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_maps[eye] = true;
    return refresh_box(h);      // fp_max only grows here (the other eye's table may be older): a larger box is always valid
}

int sa_get_maps(sa_handle* h, int eye, int16_t* map1, uint16_t* map2)
{
    SA_ENTER(h);
    if (eye != 0 && eye != 1) return fail(h, SA_ERR_INVALID_ARG, "sa_get_maps: eye must be 0 or 1");
    if (!h->have_maps[eye]) return fail(h, SA_ERR_NOT_READY, "no maps for eye %d", eye);
    const size_t mpix = (size_t)h->W * h->H;
    if (map1) SA_CUDA(h, cudaMemcpyAsync(map1, h->map1[eye], mpix * 4, cudaMemcpyDeviceToHost, h->stream));
    if (map2) SA_CUDA(h, cudaMemcpyAsync(map2, h->map2[eye], mpix * 2, cudaMemcpyDeviceToHost, h->stream));
    SA_CUDA(h, cudaStreamSynchronize(h->stream));
    return SA_OK;
}

int sa_rectify_preprocess(sa_handle* h, int n, const uint8_t* left, const uint8_t* right, size_t pitch, int where, int rectify, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (n == 0) return SA_OK;
    if (!left || !right) return fail(h, SA_ERR_INVALID_ARG, "sa_rectify_preprocess: null image pointer");
    const size_t row = (size_t)h->W * 3;
    if (pitch == 0) pitch = row;
    if (pitch < row) return fail(h, SA_ERR_INVALID_ARG, "sa_rectify_preprocess: pitch %zu < 3*width = %zu", pitch, row);
    if (where != SA_HOST && where != SA_DEVICE) return fail(h, SA_ERR_INVALID_ARG, "sa_rectify_preprocess: bad `where`");
    cudaStream_t s = pick(h, stream);
    const uint8_t* src[2] = {left, right};
    size_t src_pitch = pitch;
    const size_t up_pitch = upload_pitch(h, rectify);
    if (where == SA_HOST) {
        for (int e = 0; e < 2; e++) {
            if (pitch == row && up_pitch == row) SA_CUDA(h, cudaMemcpyAsync(h->raw[e], src[e], (size_t)n * h->H * row, cudaMemcpyHostToDevice, s));
            else SA_CUDA(h, cudaMemcpy2DAsync(h->raw[e], up_pitch, src[e], pitch, row, (size_t)n * h->H, cudaMemcpyHostToDevice, s));
            src[e] = h->raw[e];
        }
        src_pitch = up_pitch;
    }
    h->n_last = n;
    h->minmax_valid = 0;
    h->partial_valid = 0;
    return stage_rectify(h, 0, n, src[0], src[1], src_pitch, src_pitch * h->H, rectify, s);
}

// ---- f3: compressed side-by-side capture frames ----------------------------------------------------------------------------
int sa_decode_capture_jpeg(sa_handle* h, int n, const void* const* jpeg, const size_t* jpeg_bytes, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (n == 0) return SA_OK;
    if (!jpeg || !jpeg_bytes) return fail(h, SA_ERR_INVALID_ARG, "sa_decode_capture_jpeg: null argument");
    const size_t pitch = (size_t)h->W * 6, frame = pitch * h->H;
    if (!h->capture) SA_CUDA(h, cudaMalloc(&h->capture, frame * (size_t)h->max_n));
    return jpeg_decode_batch(h, n, jpeg, jpeg_bytes, h->capture, pitch, frame, 2 * h->W, h->H, pick(h, stream));
}

uint8_t* sa_capture_buffer(sa_handle* h) { return h ? h->capture : nullptr; }

int sa_copy_capture(sa_handle* h, int n, uint8_t* out, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (!out || (where != SA_HOST && where != SA_DEVICE)) return fail(h, SA_ERR_INVALID_ARG, "sa_copy_capture: bad argument");
    if (!h->capture) return fail(h, SA_ERR_NOT_READY, "sa_copy_capture: no capture frame has been decoded yet");
    cudaStream_t s = pick(h, stream);
    SA_CUDA(h, cudaMemcpyAsync(out, h->capture, (size_t)n * h->H * h->W * 6, where == SA_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, s));
    if (where == SA_HOST) SA_CUDA(h, cudaStreamSynchronize(s));
    return SA_OK;
}

int sa_rectify_preprocess_jpeg(sa_handle* h, int n, const void* const* jpeg, const size_t* jpeg_bytes, int rectify, sa_stream stream)
{
    const int rc = sa_decode_capture_jpeg(h, n, jpeg, jpeg_bytes, stream);
    if (rc != SA_OK || n == 0) return rc;
    // the two eyes are the same buffer 3W bytes apart, rows 6W bytes apart: no split, no copy (the reference clones both halves)
    return sa_rectify_preprocess(h, n, h->capture, h->capture + (size_t)h->W * 3, (size_t)h->W * 6, SA_DEVICE, rectify, stream);
}

// ---- f4: JPEG files of the demos' 8-bit outputs -------------------------------------------------------------------------------
static int jpeg_scratch(sa_handle* h, size_t bytes)
{
    if (h->jpeg_scratch_bytes >= bytes) return SA_OK;
    cudaFree(h->jpeg_scratch);
    h->jpeg_scratch = nullptr;
    h->jpeg_scratch_bytes = 0;
    SA_CUDA(h, cudaMalloc(&h->jpeg_scratch, bytes));
    h->jpeg_scratch_bytes = bytes;
    return SA_OK;
}

int sa_encode_jpeg(sa_handle* h, const uint8_t* image, int where, int width, int height, int channels, size_t pitch, int quality, void* out, size_t out_capacity,
                   size_t* out_bytes, sa_stream stream)
{
    SA_ENTER(h);
    if (!image || !out_bytes || width <= 0 || height <= 0 || (channels != 1 && channels != 3) || (where != SA_HOST && where != SA_DEVICE))
        return fail(h, SA_ERR_INVALID_ARG, "sa_encode_jpeg: bad argument");
    if (width > 65535 || height > 65535) return fail(h, SA_ERR_TOO_LARGE, "sa_encode_jpeg: %dx%d exceeds JPEG's 65535 x 65535", width, height);
    const size_t row = (size_t)width * channels;
    if (pitch == 0) pitch = row;
    if (pitch < row) return fail(h, SA_ERR_INVALID_ARG, "sa_encode_jpeg: pitch %zu < row bytes %zu", pitch, row);
    if (quality < 1 || quality > 100) return fail(h, SA_ERR_INVALID_ARG, "sa_encode_jpeg: quality %d outside 1..100", quality);
    cudaStream_t s = pick(h, stream);
    const uint8_t* dev = image;
    if (where == SA_HOST) {
        if (const int rc = jpeg_scratch(h, row * height)) return rc;
        SA_CUDA(h, cudaMemcpy2DAsync(h->jpeg_scratch, row, image, pitch, row, (size_t)height, cudaMemcpyHostToDevice, s));
        dev = h->jpeg_scratch;
        pitch = row;
    }
    std::vector<unsigned char> bytes;
    if (const int rc = jpeg_encode(h, dev, width, height, channels, pitch, quality, bytes, s)) return rc;
    *out_bytes = bytes.size();
    if (!out) return SA_OK;                                            // size query
    if (out_capacity < bytes.size()) return fail(h, SA_ERR_TOO_LARGE, "sa_encode_jpeg: the stream has %zu bytes, the buffer %zu", bytes.size(), out_capacity);
    std::memcpy(out, bytes.data(), bytes.size());
    return SA_OK;
}

int sa_write_jpeg(sa_handle* h, const char* path, const uint8_t* image, int where, int width, int height, int channels, size_t pitch, int quality)
{
    SA_ENTER(h);
    if (!path) return fail(h, SA_ERR_INVALID_ARG, "sa_write_jpeg: null path");
    size_t n = 0;
    if (const int rc = sa_encode_jpeg(h, image, where, width, height, channels, pitch, quality, nullptr, 0, &n, nullptr)) return rc;
    std::vector<unsigned char> buf(n);
    if (const int rc = sa_encode_jpeg(h, image, where, width, height, channels, pitch, quality, buf.data(), buf.size(), &n, nullptr)) return rc;
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(h, SA_ERR_IO, "sa_write_jpeg: cannot open %s", path);
    const bool ok = std::fwrite(buf.data(), 1, n, f) == n;
    if (std::fclose(f) != 0 || !ok) return fail(h, SA_ERR_IO, "sa_write_jpeg: short write to %s", path);
    return SA_OK;
}

int sa_decode_jpeg(sa_handle* h, const void* jpeg, size_t jpeg_bytes, uint8_t* out, int where, size_t pitch, int* width, int* height, sa_stream stream)
{
    SA_ENTER(h);
    if (!jpeg || jpeg_bytes == 0 || !width || !height || (where != SA_HOST && where != SA_DEVICE)) return fail(h, SA_ERR_INVALID_ARG, "sa_decode_jpeg: bad argument");
    int w = 0, hh = 0;
    if (const int rc = jpeg_image_size(h, jpeg, jpeg_bytes, &w, &hh)) return rc;
    *width = w;
    *height = hh;
    if (!out) return SA_OK;                                            // size query
    const size_t row = (size_t)w * 3;
    if (pitch == 0) pitch = row;
    if (pitch < row) return fail(h, SA_ERR_INVALID_ARG, "sa_decode_jpeg: pitch %zu < 3*width = %zu", pitch, row);
    cudaStream_t s = pick(h, stream);
    const void* one[1] = {jpeg};
    const size_t len[1] = {jpeg_bytes};
    if (where == SA_DEVICE) return jpeg_decode_checked(h, 1, one, len, out, pitch, 0, s);
    if (const int rc = jpeg_scratch(h, row * hh)) return rc;
    if (const int rc = jpeg_decode_checked(h, 1, one, len, h->jpeg_scratch, row, 0, s)) return rc;
    SA_CUDA(h, cudaMemcpy2DAsync(out, pitch, h->jpeg_scratch, row, row, (size_t)hh, cudaMemcpyDeviceToHost, s));
    SA_CUDA(h, cudaStreamSynchronize(s));
    return SA_OK;
}

void* sa_input_tensor(sa_handle* h, int eye) { return h ? h->tensor[(eye == 1 && h->cfg.model != SA_MODEL_HITNET) ? 1 : 0] : nullptr; }
float* sa_disparity_buffer(sa_handle* h) { return h ? h->disp : nullptr; }
int sa_disparity_planes(const sa_handle* h) { return h ? h->planes : 0; }
float* sa_pointcloud_buffer(sa_handle* h) { return h ? h->points : nullptr; }

int sa_stub_disparity(sa_handle* h, int n, unsigned seed, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    cudaStream_t s = pick(h, stream);
    h->minmax_valid = 0;
    h->partial_valid = 0;
    StageScope t(h, SA_STAGE_STUB, s);
    SA_CUDA(h, launch_stub(h->disp, disp_frame_floats(h), h->planes, h->w, h->h, 0, n, seed, h->cfg.model == SA_MODEL_RAFTSTEREO, s, &h->launches));
    return SA_OK;
}

int sa_set_disparity(sa_handle* h, int n, const float* disparity, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (!disparity) return fail(h, SA_ERR_INVALID_ARG, "sa_set_disparity: null");
    cudaStream_t s = pick(h, stream);
    h->minmax_valid = 0;
    h->partial_valid = 0;
    SA_CUDA(h, cudaMemcpyAsync(h->disp, disparity, (size_t)n * disp_frame_floats(h) * sizeof(float),
                               where == SA_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, s));
    if (where == SA_HOST) SA_CUDA(h, cudaStreamSynchronize(s));  // the caller may reuse its buffer
    return SA_OK;
}

int sa_reproject(sa_handle* h, int n, int mode, float* out, int* counts, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (n == 0) return SA_OK;
    if (mode != SA_DENSE && mode != SA_COMPACT) return fail(h, SA_ERR_INVALID_ARG, "sa_reproject: unknown mode %d", mode);
    if (mode == SA_COMPACT && !counts) return fail(h, SA_ERR_INVALID_ARG, "sa_reproject: SA_COMPACT needs a counts array");
    cudaStream_t s = pick(h, stream);
    const size_t npix = (size_t)h->w * h->h;
    float* dev = (out && where == SA_DEVICE) ? out : h->points;
    if (!dev) return fail(h, SA_ERR_INVALID_ARG, "sa_reproject: no internal point buffer (SA_FLAG_NO_POINT_BUF) and no device `out`");
    int* dev_counts = (counts && where == SA_DEVICE) ? counts : (counts ? h->counts : nullptr);
    int rc = stage_reproject(h, 0, n, mode, dev, dev_counts, s);
    if (rc != SA_OK) return rc;
    h->minmax_valid = 0;
    h->partial_valid = n;
    if (where == SA_HOST && (out || counts)) {
        if (mode == SA_DENSE) {
            if (counts) for (int i = 0; i < n; i++) counts[i] = (int)npix;
            if (out) return copy_out(h, out, dev, (size_t)n * npix * 6 * sizeof(float), SA_HOST, s);
            SA_CUDA(h, cudaStreamSynchronize(s));
        } else {
            SA_CUDA(h, cudaMemcpyAsync(counts, h->counts, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, s));
            SA_CUDA(h, cudaStreamSynchronize(s));
            if (out) {
                for (int i = 0; i < n; i++)
                    if (counts[i] > 0)
                        SA_CUDA(h, cudaMemcpyAsync(out + (size_t)i * npix * 6, dev + (size_t)i * npix * 6, (size_t)counts[i] * 6 * sizeof(float), cudaMemcpyDeviceToHost, s));
                SA_CUDA(h, cudaStreamSynchronize(s));
            }
        }
    }
    return SA_OK;
}

int sa_heatmap(sa_handle* h, int n, uint8_t* out, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (n == 0) return SA_OK;
    cudaStream_t s = pick(h, stream);
    uint8_t* dev = (out && where == SA_DEVICE) ? out : h->heat;
    if (!dev) return fail(h, SA_ERR_INVALID_ARG, "sa_heatmap: no internal heatmap buffer (SA_FLAG_NO_HEATMAP_BUF) and no device `out`");
    int rc = stage_heatmap(h, 0, n, dev, h->minmax_valid >= n ? 2 : (h->partial_valid >= n ? 1 : 0), s);
    if (rc != SA_OK) return rc;
    h->minmax_valid = n;
    if (out && where == SA_HOST) return copy_out(h, out, dev, (size_t)n * h->w * h->h * 3, SA_HOST, s);
    return SA_OK;
}

int sa_disparity_u8(sa_handle* h, int n, uint8_t* out, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (!out) return fail(h, SA_ERR_INVALID_ARG, "sa_disparity_u8: null");
    if (n == 0) return SA_OK;
    cudaStream_t s = pick(h, stream);
    uint8_t* dev = where == SA_DEVICE ? out : h->heat;               // host output: staged in the heatmap buffer (3x as large)
    if (!dev) return fail(h, SA_ERR_INVALID_ARG, "sa_disparity_u8: no internal staging buffer (SA_FLAG_NO_HEATMAP_BUF); pass a device `out`");
    const size_t npix = (size_t)h->w * h->h;
    SA_CUDA(h, launch_disparity_u8(h->disp, disp_frame_floats(h), npix, n, dev, s, &h->launches));
    if (where == SA_HOST) return copy_out(h, out, dev, (size_t)n * npix, SA_HOST, s);
    return SA_OK;
}

int sa_copy_rectified(sa_handle* h, int n, int eye, uint8_t* out, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if ((eye != 0 && eye != 1) || !out) return fail(h, SA_ERR_INVALID_ARG, "sa_copy_rectified: bad argument");
    // the full-resolution image the network path started from: the rectified buffer, or the caller's own image when rectify = 0
    const uint8_t* src = h->resize ? h->rect[eye] : h->color_src[eye];
    if (!src) return fail(h, SA_ERR_NOT_READY, "sa_copy_rectified: nothing processed yet");
    cudaStream_t s = pick(h, stream);
    const size_t row = (size_t)h->W * 3;
    const size_t pitch = h->resize ? h->rect_pitch : h->color_pitch;
    SA_CUDA(h, cudaMemcpy2DAsync(out, row, src, pitch, row, (size_t)n * h->H, where == SA_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, s));
    if (where == SA_HOST) SA_CUDA(h, cudaStreamSynchronize(s));
    return SA_OK;
}

int sa_copy_net_image(sa_handle* h, int n, int eye, uint8_t* out, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if ((eye != 0 && eye != 1) || !out) return fail(h, SA_ERR_INVALID_ARG, "sa_copy_net_image: bad argument");
    if (!h->resize) return sa_copy_rectified(h, n, eye, out, where, stream);
    cudaStream_t s = pick(h, stream);
    if (eye == 1 && h->net_right_frames < n) {
        // the fused rectify + resize kernel writes only the LEFT resized image (stage B's colours); the right one is made on
        // request from the rectified right image, with the stand-alone resize kernel
        if (n > h->n_last) return fail(h, SA_ERR_NOT_READY, "sa_copy_net_image: only %d frame(s) processed", h->n_last);
        ResizeArgs r{};
        r.src[0] = r.src[1] = h->rect[1];
        r.dst[0] = r.dst[1] = h->net_u8[1];
        r.tab_x = h->rs_x; r.tab_y = h->rs_y;
        r.src_pitch = h->rect_pitch; r.src_frame = h->rect_frame;
        r.W = h->W; r.H = h->H; r.w = h->w; r.h = h->h; r.n = n;
        r.area2x = (h->W == 2 * h->w && h->H == 2 * h->h) ? 1 : 0;
        r.one_eye = 1;
        SA_CUDA(h, launch_resize(r, s, &h->launches));
        h->net_right_frames = n;
    }
    return copy_out(h, out, h->net_u8[eye], (size_t)n * h->w * h->h * 3, where, s);
}

int sa_copy_tensor(sa_handle* h, int n, int eye, void* out, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (!out) return fail(h, SA_ERR_INVALID_ARG, "sa_copy_tensor: null");
    const size_t npix = (size_t)h->w * h->h;
    const bool hit = h->cfg.model == SA_MODEL_HITNET;
    const size_t bytes = (size_t)n * (hit ? 6 : 3) * npix * h->tensor_elt;
    return copy_out(h, out, sa_input_tensor(h, eye), bytes, where, pick(h, stream));
}

int sa_copy_disparity(sa_handle* h, int n, float* out, int where, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (!out) return fail(h, SA_ERR_INVALID_ARG, "sa_copy_disparity: null");
    cudaStream_t s = pick(h, stream);
    const size_t npix = (size_t)h->w * h->h;
    // plane 0 of every frame (CREStereo has two planes, CREStereo/src/TRTCREStereo.cpp:15,137)
    SA_CUDA(h, cudaMemcpy2DAsync(out, npix * sizeof(float), h->disp, disp_frame_floats(h) * sizeof(float), npix * sizeof(float), (size_t)n,
                                 where == SA_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, s));
    if (where == SA_HOST) SA_CUDA(h, cudaStreamSynchronize(s));
    return SA_OK;
}

int sa_synchronize(sa_handle* h)
{
    SA_ENTER(h);
    SA_CUDA(h, cudaStreamSynchronize(h->stream));
    SA_CUDA(h, cudaStreamSynchronize(h->s_h2d));
    SA_CUDA(h, cudaStreamSynchronize(h->s_d2h));
    return SA_OK;
}

sa_stream sa_get_stream(sa_handle* h) { return h ? (sa_stream)h->stream : nullptr; }

int sa_download(sa_handle* h, int n, float* pointcloud, float* disparity, uint8_t* rect_left, size_t rect_left_pitch, uint8_t* rect_right,
                size_t rect_right_pitch, uint8_t* heatmap, sa_stream stream)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (n == 0) return SA_OK;
    cudaStream_t s = pick(h, stream);
    const size_t npix = (size_t)h->w * h->h, row = (size_t)h->W * 3;
    if (pointcloud) {
        if (!h->points) return fail(h, SA_ERR_INVALID_ARG, "sa_download: handle was created without a point buffer");
        SA_CUDA(h, cudaMemcpyAsync(pointcloud, h->points, (size_t)n * npix * 6 * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    if (disparity)
        SA_CUDA(h, cudaMemcpy2DAsync(disparity, npix * sizeof(float), h->disp, disp_frame_floats(h) * sizeof(float), npix * sizeof(float), (size_t)n,
                                     cudaMemcpyDeviceToHost, s));
    uint8_t* rect_out[2] = {rect_left, rect_right};
    const size_t rect_out_pitch[2] = {rect_left_pitch ? rect_left_pitch : row, rect_right_pitch ? rect_right_pitch : row};
    for (int e = 0; e < 2; e++) {
        if (!rect_out[e]) continue;
        if (rect_out_pitch[e] < row) return fail(h, SA_ERR_INVALID_ARG, "sa_download: pitch %zu < 3*width = %zu", rect_out_pitch[e], row);
        const uint8_t* src = h->resize ? h->rect[e] : h->color_src[e];
        if (!src) return fail(h, SA_ERR_NOT_READY, "sa_download: nothing processed yet");
        const size_t pitch = h->resize ? h->rect_pitch : h->color_pitch;
        SA_CUDA(h, cudaMemcpy2DAsync(rect_out[e], rect_out_pitch[e], src, pitch, row, (size_t)n * h->H, cudaMemcpyDeviceToHost, s));
    }
    if (heatmap) {
        if (!h->heat) return fail(h, SA_ERR_INVALID_ARG, "sa_download: handle was created without a heatmap buffer");
        SA_CUDA(h, cudaMemcpyAsync(heatmap, h->heat, (size_t)n * npix * 3, cudaMemcpyDeviceToHost, s));
    }
    SA_CUDA(h, cudaStreamSynchronize(s));
    return SA_OK;
}

int sa_set_engine(sa_handle* h, sa_engine_fn fn, void* user)
{
    if (!h) return SA_ERR_INVALID_ARG;
    h->engine = fn; h->engine_user = user;
    return SA_OK;
}

static cudaEvent_t pool_event(sa_handle* h, size_t i)
{
    while (h->ev_pool.size() <= i) {
        cudaEvent_t e = nullptr;
        if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        h->ev_pool.push_back(e);
    }
    return h->ev_pool[i];
}

// The chunk loop of sa_run_batch.  Returns at the first failure with work possibly still queued on the three streams;
// the caller drains them before the error reaches the user.
static int run_batch_chunks(sa_handle* h, int n, const uint8_t* left, const uint8_t* right, int rectify, unsigned stub_seed, float* pointcloud,
                            float* disparity, uint8_t* rect_left, uint8_t* rect_right, uint8_t* heatmap)
{
    const size_t img = (size_t)h->W * h->H * 3, npix = (size_t)h->w * h->h, dfl = disp_frame_floats(h);
    const size_t up_pitch = upload_pitch(h, rectify);
    const size_t rawf = up_pitch * h->H;                              // one uploaded frame
    const int chunk = (int)(h->opt_chunk > 0 ? h->opt_chunk : 8);
    const int nchunks = (n + chunk - 1) / chunk;
    uint8_t* rect_out[2] = {rect_left, rect_right};
    const uint8_t* host_in[2] = {left, right};
    for (int c = 0; c < nchunks; c++) {
        const int f0 = c * chunk, m = (f0 + chunk <= n) ? chunk : n - f0;
        cudaEvent_t up = pool_event(h, 2 * (size_t)c), done = pool_event(h, 2 * (size_t)c + 1);
        if (!up || !done) return fail(h, SA_ERR_CUDA, "sa_run_batch: cannot create events");
        // upload (cudaMemcpyAsync H2D x2, TRTFastACVNet_plus.cpp:130-131)
        for (int e = 0; e < 2; e++) {
            if (up_pitch == (size_t)h->W * 3) SA_CUDA(h, cudaMemcpyAsync(h->raw[e] + (size_t)f0 * img, host_in[e] + (size_t)f0 * img, (size_t)m * img, cudaMemcpyHostToDevice, h->s_h2d));
            else SA_CUDA(h, cudaMemcpy2DAsync(h->raw[e] + (size_t)f0 * rawf, up_pitch, host_in[e] + (size_t)f0 * img, (size_t)h->W * 3, (size_t)h->W * 3, (size_t)m * h->H, cudaMemcpyHostToDevice, h->s_h2d));
        }
        SA_CUDA(h, cudaEventRecord(up, h->s_h2d));
        SA_CUDA(h, cudaStreamWaitEvent(h->stream, up, 0));
        // stage A
        int rc = stage_rectify(h, f0, m, h->raw[0] + (size_t)f0 * rawf, h->raw[1] + (size_t)f0 * rawf, up_pitch, rawf, rectify, h->stream);
        if (rc != SA_OK) return rc;
        // network slot (enqueueV3, TRTFastACVNet_plus.cpp:152)
        const size_t tframe = (h->cfg.model == SA_MODEL_HITNET ? 6 : 3) * npix * h->tensor_elt;
        if (h->engine) {
            rc = h->engine(h->engine_user, (sa_stream)h->stream, f0, m, (char*)h->tensor[0] + (size_t)f0 * tframe, (char*)h->tensor[1] + (size_t)f0 * tframe,
                           h->disp + (size_t)f0 * dfl);
            if (rc != 0) return fail(h, SA_ERR_INVALID_ARG, "sa_run_batch: engine callback returned %d", rc);
        } else if (stub_seed != 0) {
            StageScope t(h, SA_STAGE_STUB, h->stream);
            SA_CUDA(h, launch_stub(h->disp + (size_t)f0 * dfl, dfl, h->planes, h->w, h->h, f0 + (int)h->opt_frame_base, m, stub_seed, h->cfg.model == SA_MODEL_RAFTSTEREO, h->stream, &h->launches));
        }
        // stage B (+ C)
        rc = stage_reproject(h, f0, m, SA_DENSE, h->points + (size_t)f0 * npix * 6, nullptr, h->stream);
        if (rc != SA_OK) return rc;
        if (heatmap) { rc = stage_heatmap(h, f0, m, h->heat + (size_t)f0 * npix * 3, 1, h->stream); if (rc != SA_OK) return rc; }
        SA_CUDA(h, cudaEventRecord(done, h->stream));
        SA_CUDA(h, cudaStreamWaitEvent(h->s_d2h, done, 0));
        // download (cudaMemcpy D2H x2, TRTFastACVNet_plus.cpp:163-164; plus the rectified images the reference leaves in left/right).
        // One linear copy per output wherever the device layout is dense (single-plane disparity, unpadded rectified rows):
        // a 2-D copy of h rows costs the copy engine a descriptor per row.
        if (pointcloud) SA_CUDA(h, cudaMemcpyAsync(pointcloud + (size_t)f0 * npix * 6, h->points + (size_t)f0 * npix * 6, (size_t)m * npix * 6 * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
        if (disparity) {
            if (h->planes == 1) SA_CUDA(h, cudaMemcpyAsync(disparity + (size_t)f0 * npix, h->disp + (size_t)f0 * dfl, (size_t)m * npix * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
            else SA_CUDA(h, cudaMemcpy2DAsync(disparity + (size_t)f0 * npix, npix * sizeof(float), h->disp + (size_t)f0 * dfl, dfl * sizeof(float), npix * sizeof(float), (size_t)m, cudaMemcpyDeviceToHost, h->s_d2h));
        }
        if (rectify)
            for (int e = 0; e < 2; e++) {
                if (!rect_out[e]) continue;
                if (h->rect_pitch == (size_t)h->W * 3) SA_CUDA(h, cudaMemcpyAsync(rect_out[e] + (size_t)f0 * img, h->rect[e] + (size_t)f0 * h->rect_frame, (size_t)m * img, cudaMemcpyDeviceToHost, h->s_d2h));
                else SA_CUDA(h, cudaMemcpy2DAsync(rect_out[e] + (size_t)f0 * img, (size_t)h->W * 3, h->rect[e] + (size_t)f0 * h->rect_frame, h->rect_pitch, (size_t)h->W * 3, (size_t)m * h->H, cudaMemcpyDeviceToHost, h->s_d2h));
            }
        if (heatmap) SA_CUDA(h, cudaMemcpyAsync(heatmap + (size_t)f0 * npix * 3, h->heat + (size_t)f0 * npix * 3, (size_t)m * npix * 3, cudaMemcpyDeviceToHost, h->s_d2h));
    }
    return SA_OK;
}

int sa_run_batch(sa_handle* h, int n, const uint8_t* left, const uint8_t* right, int rectify, unsigned stub_seed, float* pointcloud, float* disparity,
                 uint8_t* rect_left, uint8_t* rect_right, uint8_t* heatmap)
{
    SA_ENTER(h);
    SA_CHECK_N(h, n);
    if (n == 0) return SA_OK;
    if (!left || !right) return fail(h, SA_ERR_INVALID_ARG, "sa_run_batch: null image pointer");
    // stage B always writes the cloud (its min/max partials feed the heat map), so the internal buffer is needed even when
    // the caller wants only the disparity or the heat map back
    if (!h->points) return fail(h, SA_ERR_INVALID_ARG, "sa_run_batch: handle was created without a point buffer (SA_FLAG_NO_POINT_BUF)");
    if (!h->heat && heatmap) return fail(h, SA_ERR_INVALID_ARG, "sa_run_batch: handle was created without a heatmap buffer");
    h->n_last = n;
    h->minmax_valid = 0;
    h->partial_valid = 0;
    const int rc = run_batch_chunks(h, n, left, right, rectify, stub_seed, pointcloud, disparity, rect_left, rect_right, heatmap);
    if (rc != SA_OK) {
        // Earlier chunks may still be copying into the caller's buffers: nothing of this call is in flight once it returns,
        // also on failure.  (The message of the first failure is kept.)
        const std::string first = h->err;
        cudaStreamSynchronize(h->s_h2d);
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->s_d2h);
        cudaGetLastError();
        h->err = first;
        return rc;
    }
    SA_CUDA(h, cudaStreamSynchronize(h->s_d2h));
    SA_CUDA(h, cudaStreamSynchronize(h->stream));
    h->partial_valid = n;
    h->minmax_valid = heatmap ? n : 0;
    return SA_OK;
}

int sa_host_alloc(void** p, size_t bytes)
{
    if (!p) return SA_ERR_INVALID_ARG;
    cudaError_t e = cudaMallocHost(p, bytes ? bytes : 1);
    if (e != cudaSuccess) return fail(nullptr, e == cudaErrorMemoryAllocation ? SA_ERR_NOMEM : SA_ERR_CUDA, "sa_host_alloc(%zu): %s", bytes, cudaGetErrorString(e));
    return SA_OK;
}

int sa_host_register(void* p, size_t bytes)
{
    if (!p || bytes == 0) return SA_ERR_INVALID_ARG;
    const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterDefault);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(nullptr, SA_ERR_CUDA, "sa_host_register(%zu): %s", bytes, cudaGetErrorString(e)); }
    return SA_OK;
}

int sa_host_unregister(void* p)
{
    if (!p) return SA_OK;
    const cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(nullptr, SA_ERR_CUDA, "sa_host_unregister: %s", cudaGetErrorString(e)); }
    return SA_OK;
}

int sa_host_free(void* p)
{
    if (!p) return SA_OK;
    cudaError_t e = cudaFreeHost(p);
    if (e != cudaSuccess) return fail(nullptr, SA_ERR_CUDA, "sa_host_free: %s", cudaGetErrorString(e));
    return SA_OK;
}

int sa_stage_time(sa_handle* h, int stage, double* total_ms, long* launches)
{
    SA_ENTER(h);
    if (stage < 0 || stage >= SA_STAGE_COUNT_) return fail(h, SA_ERR_INVALID_ARG, "sa_stage_time: bad stage %d", stage);
    if (!(h->cfg.flags & SA_FLAG_TIMING)) return fail(h, SA_ERR_NOT_READY, "sa_stage_time: handle was created without SA_FLAG_TIMING");
    SA_CUDA(h, cudaDeviceSynchronize());
    double tot = 0.0;
    for (size_t i = 0; i < h->timer.used[stage]; i++) {
        float ms = 0.f;
        SA_CUDA(h, cudaEventElapsedTime(&ms, h->timer.begin[stage][i], h->timer.end[stage][i]));
        tot += ms;
    }
    if (total_ms) *total_ms = tot;
    if (launches) *launches = (long)h->timer.used[stage];
    return SA_OK;
}

int sa_reset_timing(sa_handle* h)
{
    SA_ENTER(h);
    SA_CUDA(h, cudaDeviceSynchronize());
    for (int s = 0; s < SA_STAGE_COUNT_; s++) h->timer.used[s] = 0;
    return SA_OK;
}

long sa_launch_count(const sa_handle* h) { return h ? h->launches : 0; }

int sa_set_option(sa_handle* h, const char* key, long value)
{
    if (!h || !key) return SA_ERR_INVALID_ARG;
    if (!std::strcmp(key, "frame_fastest")) h->opt_frame_fastest = value;
    else if (!std::strcmp(key, "chunk")) h->opt_chunk = value;
    else if (!std::strcmp(key, "force_direct")) h->opt_force_direct = value;   // takes effect at the next map computation
    else if (!std::strcmp(key, "force_generic")) h->opt_force_generic = value;
    else if (!std::strcmp(key, "exact_div")) h->opt_exact_div = value;
    else if (!std::strcmp(key, "persistent")) h->opt_persistent = value;
    else if (!std::strcmp(key, "batch_loop")) h->opt_batch_loop = value;
    else if (!std::strcmp(key, "batch_ctas")) h->opt_batch_ctas = value;
    else if (!std::strcmp(key, "bulk")) h->opt_bulk = value;
    else if (!std::strcmp(key, "force_unaligned")) h->opt_force_unaligned = value;
    else if (!std::strcmp(key, "no_fixed_size")) h->opt_no_fixed_size = value;
    else if (!std::strcmp(key, "scalar_preprocess")) h->opt_scalar_preprocess = value;
    else if (!std::strcmp(key, "l2_prefetch")) h->opt_l2_prefetch = value;
    else if (!std::strcmp(key, "frame_base")) h->opt_frame_base = value;
    else if (!std::strcmp(key, "single_pass_compact")) h->opt_single_pass_compact = value;
    else if (!std::strcmp(key, "separate_preprocess")) h->opt_separate_preprocess = value;
    else if (!std::strcmp(key, "no_align16")) h->opt_no_align16 = value;
    else if (!std::strcmp(key, "fused_resize")) h->opt_fused_resize = value;
    else if (!std::strcmp(key, "chunk_taper")) h->opt_chunk_taper = value < 0 ? 0 : (value > 95 ? 95 : value);
    else return fail(h, SA_ERR_INVALID_ARG, "sa_set_option: unknown key '%s'", key);
    return SA_OK;
}

}  // extern "C"
