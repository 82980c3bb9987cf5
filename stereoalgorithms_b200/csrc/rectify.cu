// Stage A: rectification fused with the network preprocess, one HBM pass per eye-frame.
//
// Replaces, per stereo pair, the reference's  cv::remap x2 on the CPU
// (FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:142-143)  +  H2D  +  <Model>_preprocess_kernel x2
// (FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:4-39 and the three siblings).
//
// One CTA (256 threads) produces a 64x16 tile of the rectified image:
//   1. each thread reads the per-pixel table entry (footprint offset | fx | fy, derived from the maps per calibration)
//      of its 4 pixels - once per batch in the default kernel, which keeps them in registers while it walks the frames;
//   2. the CTA stages the tile's source footprint - a bounding box precomputed per tile at calibration time -
//      from global memory into shared memory as one BGRx word per pixel, zero outside the image, so that
//      BORDER_CONSTANT needs no per-tap test afterwards;
//   3. each thread blends its pixels with OpenCV's exact fixed-point arithmetic
//         dst = (p00*w00 + p01*w01 + p10*w10 + p11*w11 + 2^14) >> 15,  w = (32-fx|fx)*(32-fy|fy)*32
//      evaluated as ((top*(32-fy) + bot*fy + 512) >> 10), which is the same integer because nothing is rounded
//      in between: the horizontal pass on the byte dot-product unit (IDP.4A), the vertical one as a 16-bit dot product
//      with the rounding constant as its accumulator (IDP.2A);
//   4. the normalised RGB planes go straight to global memory (a warp writes 32 consecutive floats/halves of a
//      plane row); the rectified BGR bytes of four neighbouring lanes are packed into three words with one
//      shuffle + byte-permute and leave as 96-byte coalesced row segments (the rectified buffer is internal and its
//      rows are padded to whole 4-pixel groups, so this works for every width).
//
// Kernels in this file (DESIGN.md section 3 has the measurements):
//   rectify_batch_kernel       default: one CTA per (tile, eye) walks a run of frames; A4 / !A4 / PART / 640x480 instantiations
//   rectify_tma_kernel         option bulk=1: same decomposition, footprint by cp.async.bulk.tensor + mbarrier ring (slower)
//   rectify_kernel             one CTA per (tile, frame); batch_loop=0 and the bounds-checked tile_generic path
//   rectify_persistent_kernel  option persistent>0: software-pipelined over arbitrary tiles (slower)
//   preprocess4_kernel / preprocess_kernel   rectify = 0 and the tail of the resize path: normalise + planarise only
#include <cuda.h>   // CUtensorMap and its enums only; the one driver function used is fetched with cudaGetDriverEntryPoint

#include <type_traits>

#include "common.h"

#ifndef SA_BLEND_DP4A
#define SA_BLEND_DP4A 1
#endif

namespace sa {

namespace {

__device__ __forceinline__ uint32_t fetch_px(const uint8_t* __restrict__ src, size_t pitch, int W, int H, int x, int y)
{
    if ((unsigned)x < (unsigned)W && (unsigned)y < (unsigned)H) {
        const uint8_t* p = src + (size_t)y * pitch + 3 * x;
        return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
    }
    return 0u;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <typename V>
__device__ __forceinline__ void copy_rows(uint8_t* __restrict__ g, size_t gpitch, const uint8_t* s, int rows, int row_bytes)
{
    constexpr int kRow = kTileW * 3;
    constexpr int per_row = kRow / (int)sizeof(V);
    const int nfull = row_bytes / (int)sizeof(V);
    for (int i = threadIdx.x; i < rows * per_row; i += kRectThreads) {
        int r = i / per_row, v = i - r * per_row;
        if (v < nfull) reinterpret_cast<V*>(g + (size_t)r * gpitch)[v] = reinterpret_cast<const V*>(s + r * kRow)[v];
    }
    const int tail0 = nfull * (int)sizeof(V), ntail = row_bytes - tail0;
    if (ntail > 0)
        for (int i = threadIdx.x; i < rows * ntail; i += kRectThreads) {
            int r = i / ntail, b = tail0 + (i - r * ntail);
            g[(size_t)r * gpitch + b] = s[r * kRow + b];
        }
}

template <int DTYPE>
struct TensorT { using type = float; };
template <>
struct TensorT<SA_F16> { using type = __half; };

// OpenCV's fixed-point bilinear blend of one output pixel from its four BGRx taps.
//   (p00*w00 + p01*w01 + p10*w10 + p11*w11 + 2^14) >> 15  with  w = (32-fx|fx)*(32-fy|fy)*32
// == ((top*(32-fy) + bot*fy + 512) >> 10)  with  top/bot = p_left*(32-fx) + p_right*fx   (nothing rounded in between).
// B and R travel in the two 16-bit lanes of one word through the horizontal pass (each lane <= 255*32).
__device__ __forceinline__ void blend(uint32_t p00, uint32_t p01, uint32_t p10, uint32_t p11, uint32_t fx, uint32_t fy, uint32_t& vB, uint32_t& vG,
                                      uint32_t& vR)
{
#if SA_BLEND_DP4A
    // horizontal pass on the byte dot-product unit: (B_l, B_r, G_l, G_r) . (ax, fx, 0, 0) etc.
    const uint32_t wx = (32u - fx) | (fx << 8), wxh = wx << 16, ay = 32u - fy;
    const uint32_t tBG = __byte_perm(p00, p01, 0x5140), tRR = __byte_perm(p00, p01, 0x6262);
    const uint32_t bBG = __byte_perm(p10, p11, 0x5140), bRR = __byte_perm(p10, p11, 0x6262);
    const uint32_t tB = __dp4a(tBG, wx, 0u), tG = __dp4a(tBG, wxh, 0u), tR = __dp4a(tRR, wx, 0u);
    const uint32_t bB = __dp4a(bBG, wx, 0u), bG = __dp4a(bBG, wxh, 0u), bR = __dp4a(bRR, wx, 0u);
    vB = (tB * ay + bB * fy + 512u) >> 10;
    vG = (tG * ay + bG * fy + 512u) >> 10;
    vR = (tR * ay + bR * fy + 512u) >> 10;
#else
    const uint32_t ax = 32u - fx, ay = 32u - fy;
    const uint32_t tBR = (p00 & 0x00ff00ffu) * ax + (p01 & 0x00ff00ffu) * fx;
    const uint32_t bBR = (p10 & 0x00ff00ffu) * ax + (p11 & 0x00ff00ffu) * fx;
    const uint32_t tG = __byte_perm(p00, 0, 0x4441) * ax + __byte_perm(p01, 0, 0x4441) * fx;
    const uint32_t bG = __byte_perm(p10, 0, 0x4441) * ax + __byte_perm(p11, 0, 0x4441) * fx;
    vB = ((tBR & 0xffffu) * ay + (bBR & 0xffffu) * fy + 512u) >> 10;
    vR = ((tBR >> 16) * ay + (bBR >> 16) * fy + 512u) >> 10;
    vG = (tG * ay + bG * fy + 512u) >> 10;
#endif
}

// First word a lane stores of the rectified row `row` (bytes per row = pitch, a multiple of 4): lanes m = 0..2 of every
// group of four pixels write the three words of the group's 12 bytes.
__device__ __forceinline__ uint32_t* rect_words(uint8_t* frame_base, size_t pitch, int row, int x, int m)
{
    return reinterpret_cast<uint32_t*>(frame_base + (size_t)row * pitch) + (3 * ((x - m) >> 2) + m);
}

template <int DTYPE, bool IDENT, typename T>
__device__ __forceinline__ void normalise(const T* s_lut, uint32_t vB, uint32_t vG, uint32_t vR, T& r, T& g, T& b)
{
    if (IDENT) {
        if (DTYPE == SA_F16) { r = (T)__ushort2half_rn((unsigned short)vR); g = (T)__ushort2half_rn((unsigned short)vG); b = (T)__ushort2half_rn((unsigned short)vB); }
        else {  // (float)v for v < 2^23 without the conversion unit: 2^23 + v, exactly, minus 2^23
            r = (T)(__uint_as_float(0x4b000000u | vR) - 8388608.0f);
            g = (T)(__uint_as_float(0x4b000000u | vG) - 8388608.0f);
            b = (T)(__uint_as_float(0x4b000000u | vB) - 8388608.0f);
        }
    } else {
        r = s_lut[vR]; g = s_lut[256 + vG]; b = s_lut[512 + vB];
    }
}

// The warp-specialised batch kernel keeps FOUR interleaved copies of each normalisation table in shared memory (entry e of copy j at word
// 4 e + j, one 32-bit word per value): the lanes of a warp are spread over four disjoint sets of banks by lane & 3, which cuts
// the bank conflicts of 32 data-dependent look-ups from ~3.5 to ~2.3 wavefronts.  3 x 256 x 4 words = 12 KB.
constexpr int kLut4Words = 3 * 256 * 4;
template <int DTYPE>
__device__ __forceinline__ void fill_lut4(uint32_t* s_lut4, const void* lut, int nthreads)
{
    using T = typename TensorT<DTYPE>::type;
    const T* l = static_cast<const T*>(lut);
    for (int i = threadIdx.x; i < 768; i += nthreads) {
        uint32_t bits;
        if (DTYPE == SA_F16) bits = __half_as_ushort(*reinterpret_cast<const __half*>(l + i));
        else bits = __float_as_uint(*reinterpret_cast<const float*>(l + i));
        *reinterpret_cast<uint4*>(&s_lut4[4 * i]) = make_uint4(bits, bits, bits, bits);
    }
}
// lut_lane = (const char*)s_lut4 + 4 * (lane & 3)
template <typename T>
__device__ __forceinline__ void normalise_lut4(const char* lut_lane, uint32_t vB, uint32_t vG, uint32_t vR, T& r, T& g, T& b)
{
    r = *reinterpret_cast<const T*>(lut_lane + vR * 16u);
    g = *reinterpret_cast<const T*>(lut_lane + 4096u + vG * 16u);
    b = *reinterpret_cast<const T*>(lut_lane + 8192u + vB * 16u);
}

// Stage a footprint row group (4 source pixels, 12 bytes -> four BGRx words) with per-pixel bounds checks.
template <bool ALIGNED4>
__device__ __forceinline__ uint4 stage_checked(const uint8_t* __restrict__ src, size_t pitch, int W, int H, int xx, int yy)
{
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if ((unsigned)yy < (unsigned)H) {
        if (ALIGNED4 && xx >= 0 && xx + 3 < W) {
            const uint32_t* p = reinterpret_cast<const uint32_t*>(src + (size_t)yy * pitch + 3 * xx);
            const uint32_t u0 = __ldg(p), u1 = __ldg(p + 1), u2 = __ldg(p + 2);
            o.x = u0;  // byte 3 of a staged word is never read
            o.y = __byte_perm(u0, u1, 0x4543);
            o.z = __byte_perm(u1, u2, 0x4432);
            o.w = u2 >> 8;
        } else {
            o.x = fetch_px(src, pitch, W, H, xx, yy);
            o.y = fetch_px(src, pitch, W, H, xx + 1, yy);
            o.z = fetch_px(src, pitch, W, H, xx + 2, yy);
            o.w = fetch_px(src, pitch, W, H, xx + 3, yy);
        }
    }
    return o;
}

// ---- fast path: a full 64x16 tile, staged footprint, 4-byte aligned rectified rows ----------------------------
// No bounds tests on the output side, 32-bit offsets, and the rectified BGR bytes leave straight from registers:
// lanes 4j..4j+3 hold 4 consecutive pixels = 12 bytes = 3 words; one shuffle + one byte-permute per lane builds
// word m in lane 4j+m (m < 3), so a warp row segment (32 pixels) is one 96-byte coalesced store.
template <int DTYPE, bool IDENT, bool WRITE_TENSOR, bool ALIGNED4>
__device__ __forceinline__ void tile_fast(const RectifyArgs& a, const TileInfo ti, int eye, int frame, int tx0, int ty0, uint32_t* s_src,
                                          const typename TensorT<DTYPE>::type* s_lut)
{
    using T = typename TensorT<DTYPE>::type;
    const int tid = threadIdx.x;
    const int W = a.W;
    const int pix0 = (ty0 + (tid >> 6)) * W + tx0 + (tid & (kTileW - 1));
    const int rstep = 4 * W;

    // per-pixel staged-footprint offset + sub-pixel weights, precomputed with the tile table (maps.cu)
    const uint32_t* __restrict__ tmap = a.tmap[eye] + pix0;
    uint32_t tm[4];
#pragma unroll
    for (int k = 0; k < 4; k++) tm[k] = __ldg(tmap + k * rstep);

    const uint8_t* __restrict__ src = a.src[eye] + (size_t)frame * a.src_frame;
    const int fw = ti.w;
    {
        const int gw = fw >> 2, total = gw * ti.h;
        const uint32_t inv = ti.inv_gw;                        // ceil(2^20 / gw): g / gw == (g * inv) >> 20 for g < 768
        if (ALIGNED4 && (ti.flags & kTileInterior)) {
            const int pitch = (int)a.src_pitch;
            const uint8_t* base = src + ti.y0 * pitch + 3 * ti.x0;   // frame < 2 GB (checked on the host)
            for (int g = tid; g < total; g += kRectThreads) {
                const int r = (int)(((uint32_t)g * inv) >> 20), c = g - r * gw;
                const uint32_t* p = reinterpret_cast<const uint32_t*>(base + r * pitch + 12 * c);
                const uint32_t u0 = __ldg(p), u1 = __ldg(p + 1), u2 = __ldg(p + 2);
                uint4 o;
                o.x = u0;
                o.y = __byte_perm(u0, u1, 0x4543);
                o.z = __byte_perm(u1, u2, 0x4432);
                o.w = u2 >> 8;
                *reinterpret_cast<uint4*>(&s_src[g << 2]) = o;   // row stride == fw words, so word index = 4*g
            }
        } else {
            const int X0 = ti.x0, Y0 = ti.y0;
            for (int g = tid; g < total; g += kRectThreads) {
                const int r = (int)(((uint32_t)g * inv) >> 20), c4 = (g - r * gw) << 2;
                *reinterpret_cast<uint4*>(&s_src[g << 2]) = stage_checked<ALIGNED4>(src, a.src_pitch, W, a.H, X0 + c4, Y0 + r);
            }
        }
    }
    __syncthreads();

    const int plane = W * a.H;
    const int m = tid & 3;
    const uint32_t sel = m == 0 ? 0x4210u : (m == 1 ? 0x5421u : 0x6542u);
    uint32_t* __restrict__ rp = rect_words(a.rect[eye] + (size_t)frame * a.rect_frame, a.rect_pitch, ty0 + (tid >> 6), tx0 + (tid & (kTileW - 1)), m);
    T* __restrict__ tp = WRITE_TENSOR ? static_cast<T*>(a.tensor[eye]) + (size_t)frame * a.tensor_frame + pix0 : nullptr;
    const int wstep = (int)a.rect_pitch;                             // words per 4 rows
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t t = tm[k];
        const uint32_t* p = s_src + (t & 0xfffu);
        uint32_t vB, vG, vR;
        blend(p[0], p[1], p[fw], p[fw + 1], (t >> 12) & 31u, (t >> 17) & 31u, vB, vG, vR);
        const uint32_t v = vB | (vG << 8) | (vR << 16);
        const uint32_t nb = __shfl_down_sync(0xffffffffu, v, 1);
        if (m < 3) rp[k * wstep] = __byte_perm(v, nb, sel);
        if (WRITE_TENSOR) {
            T r, g, b;
            normalise<DTYPE, IDENT, T>(s_lut, vB, vG, vR, r, g, b);
            tp[k * rstep] = r; tp[k * rstep + plane] = g; tp[k * rstep + 2 * plane] = b;
        }
    }
}

// ---- generic path: partial tiles at the right/bottom edge, footprints too large to stage, odd alignments ---------
// CTA-wide barrier, or (NAMED) barrier 1 over the first kRectThreads threads only - the consumer threads of the warp-specialised
// kernel, whose producer warps have left by then.
template <bool NAMED>
__device__ __forceinline__ void tile_sync()
{
    if (NAMED) asm volatile("bar.sync 1, %0;" ::"n"(kRectThreads) : "memory");
    else __syncthreads();
}

template <int DTYPE, bool IDENT, bool WRITE_TENSOR, bool ALIGNED4, bool NAMED_BAR = false>
__device__ __forceinline__ void tile_generic(const RectifyArgs& a, const TileInfo ti, int eye, int frame, int tx0, int ty0, int out_align,
                                             uint32_t* s_src, uint32_t* s_out, const typename TensorT<DTYPE>::type* s_lut)
{
    using T = typename TensorT<DTYPE>::type;
    const int tid = threadIdx.x;
    const int W = a.W, H = a.H;
    const uint8_t* __restrict__ src = (eye ? a.src[1] : a.src[0]) + (size_t)frame * a.src_frame;
    const size_t pitch = a.src_pitch;

    const int lx = tid & (kTileW - 1), ly0 = tid >> 6;
    const int x = tx0 + lx;
    const int* __restrict__ map1 = eye ? a.map1[1] : a.map1[0];
    const uint16_t* __restrict__ map2 = eye ? a.map2[1] : a.map2[0];

    const int fw = ti.w, X0 = ti.x0, Y0 = ti.y0;
    const bool staged = fw > 0;
    if (staged) {
        const int gw = fw >> 2, total = gw * ti.h;
        const uint32_t inv = ti.inv_gw;
        for (int g = tid; g < total; g += kRectThreads) {
            const int r = (int)(((uint32_t)g * inv) >> 20), c4 = (g - r * gw) << 2;
            *reinterpret_cast<uint4*>(&s_src[g << 2]) = stage_checked<ALIGNED4>(src, pitch, W, H, X0 + c4, Y0 + r);
        }
    }
    tile_sync<NAMED_BAR>();

    uint8_t* s_out_b = reinterpret_cast<uint8_t*>(s_out);
    const size_t plane = (size_t)W * H;
    T* __restrict__ tbase = WRITE_TENSOR ? static_cast<T*>(eye ? a.tensor[1] : a.tensor[0]) + (size_t)frame * a.tensor_frame : nullptr;
#pragma unroll 1
    for (int k = 0; k < 4; k++) {
        const int ly = ly0 + 4 * k, y = ty0 + ly;
        if (x < W && y < H) {
            const size_t mo = (size_t)y * W + x;
            const int m1 = __ldg(map1 + mo);
            const uint32_t m2 = __ldg(map2 + mo);
            int sx = (int)(short)(m1 & 0xffff), sy = m1 >> 16;
            uint32_t p00, p01, p10, p11;
            if (staged) {
                sx = max(-2, min(W, sx));
                sy = max(-2, min(H, sy));
                const uint32_t* p = s_src + (sy - Y0) * fw + (sx - X0);
                p00 = p[0]; p01 = p[1]; p10 = p[fw]; p11 = p[fw + 1];
            } else {
                p00 = fetch_px(src, pitch, W, H, sx, sy);
                p01 = fetch_px(src, pitch, W, H, sx + 1, sy);
                p10 = fetch_px(src, pitch, W, H, sx, sy + 1);
                p11 = fetch_px(src, pitch, W, H, sx + 1, sy + 1);
            }
            uint32_t vB, vG, vR;
            blend(p00, p01, p10, p11, m2 & 31u, (m2 >> 5) & 31u, vB, vG, vR);
            uint8_t* so = s_out_b + (ly * kTileW + lx) * 3;
            so[0] = (uint8_t)vB; so[1] = (uint8_t)vG; so[2] = (uint8_t)vR;
            if (WRITE_TENSOR) {
                T* t = tbase + (size_t)y * W + x;
                T r, g, b;
                normalise<DTYPE, IDENT, T>(s_lut, vB, vG, vR, r, g, b);
                t[0] = r; t[plane] = g; t[2 * plane] = b;
            }
        }
    }
    tile_sync<NAMED_BAR>();

    // rectified BGR tile -> global, widest stores the alignment allows
    const int rows = min(kTileH, H - ty0), row_bytes = min(kTileW, W - tx0) * 3;
    uint8_t* g = (eye ? a.rect[1] : a.rect[0]) + (size_t)frame * a.rect_frame + (size_t)ty0 * a.rect_pitch + (size_t)tx0 * 3;
    const size_t gpitch = a.rect_pitch;
    if (out_align >= 16) copy_rows<uint4>(g, gpitch, s_out_b, rows, row_bytes);
    else if (out_align >= 4) copy_rows<uint32_t>(g, gpitch, s_out_b, rows, row_bytes);
    else if (out_align >= 2) copy_rows<uint16_t>(g, gpitch, s_out_b, rows, row_bytes);
    else copy_rows<uint8_t>(g, gpitch, s_out_b, rows, row_bytes);
}

// grid: frame_fastest ? (2n, tiles_x, tiles_y) : (tiles_x, tiles_y, 2n)   [eye = fe & 1, frame = fe >> 1]
template <int DTYPE, bool IDENT, bool WRITE_TENSOR, bool ALIGNED4>
__global__ void __launch_bounds__(kRectThreads, 6) rectify_kernel(const RectifyArgs a, int tiles_x, int out_align, int fast_ok)
{
    using T = typename TensorT<DTYPE>::type;
    __shared__ __align__(16) uint32_t s_src[kFootprintWords];
    __shared__ __align__(16) uint32_t s_out[kTileH * kTileW * 3 / 4];
    __shared__ T s_lut[IDENT ? 1 : 768];

    int txi, tyi, fe;
    if (a.frame_fastest) { fe = blockIdx.x; txi = blockIdx.y; tyi = blockIdx.z; }
    else { txi = blockIdx.x; tyi = blockIdx.y; fe = blockIdx.z; }
    const int eye = fe & 1, frame = fe >> 1;
    const int tx0 = txi * kTileW, ty0 = tyi * kTileH;
    const TileInfo ti = (eye ? a.tiles[1] : a.tiles[0])[tyi * tiles_x + txi];

    if (!IDENT) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = threadIdx.x; i < 768; i += kRectThreads) s_lut[i] = lut[i];
    }
    const bool full = tx0 + kTileW <= a.W && ty0 + kTileH <= a.H;
    if (fast_ok && full && ti.w > 0) tile_fast<DTYPE, IDENT, WRITE_TENSOR, ALIGNED4>(a, ti, eye, frame, tx0, ty0, s_src, s_lut);
    else tile_generic<DTYPE, IDENT, WRITE_TENSOR, ALIGNED4>(a, ti, eye, frame, tx0, ty0, out_align, s_src, s_out, s_lut);
}

// ---- persistent, software-pipelined variant of the fast path ------------------------------------------------------------
// One CTA per SM slot walks a contiguous run of tiles (in frame-fastest order consecutive tiles are the SAME tile
// position in successive eye-frames, so the tile table entry and the per-pixel table stay L1-resident).  While tile i is
// blended out of shared memory, the raw source words and table entries of tile i+1 are already in flight into
// registers; they are expanded to BGRx and stored to shared memory after the barrier that ends tile i.  Global-load
// latency therefore overlaps the arithmetic of the previous tile instead of stalling every CTA at its start.
// Tiles that cannot take the fast path (partial tiles at the image edge, unstaged footprints) run tile_generic in place.
template <int DTYPE, bool IDENT, bool WRITE_TENSOR>
__global__ void __launch_bounds__(kRectThreads, 4) rectify_persistent_kernel(const RectifyArgs a, int tiles_x, int tiles_y, int out_align)
{
    using T = typename TensorT<DTYPE>::type;
    __shared__ __align__(16) uint32_t s_src[kFootprintWords];
    __shared__ __align__(16) uint32_t s_out[kTileH * kTileW * 3 / 4];
    __shared__ T s_lut[IDENT ? 1 : 768];

    const int tid = threadIdx.x;
    const int W = a.W, H = a.H, nfe = 2 * a.n;
    const int tpf = tiles_x * tiles_y;
    const long long total = (long long)tpf * nfe;
    int it = (int)(total * blockIdx.x / gridDim.x);
    const int end = (int)(total * (blockIdx.x + 1) / gridDim.x);
    if (it >= end) return;
    if (!IDENT) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = tid; i < 768; i += kRectThreads) s_lut[i] = lut[i];
    }

    // position of tile `it`
    int fe, txi, tyi;
    if (a.frame_fastest) { fe = it % nfe; const int tile = it / nfe; tyi = tile / tiles_x; txi = tile - tyi * tiles_x; }
    else { const int tile = it % tpf; fe = it / tpf; tyi = tile / tiles_x; txi = tile - tyi * tiles_x; }
    auto advance = [&](int& f, int& x, int& y) {
        if (a.frame_fastest) { if (++f == nfe) { f = 0; if (++x == tiles_x) { x = 0; ++y; } } }
        else { if (++x == tiles_x) { x = 0; if (++y == tiles_y) { y = 0; ++f; } } }
    };
    auto tile_fast_ok = [&](const TileInfo& t, int x, int y) { return t.w > 0 && x * kTileW + kTileW <= W && y * kTileH + kTileH <= H; };

    const int pitch = (int)a.src_pitch;
    const int rstep = 4 * W, wstep = (int)a.rect_pitch, plane = W * H;
    const int m = tid & 3;
    const uint32_t sel = m == 0 ? 0x4210u : (m == 1 ? 0x5421u : 0x6542u);
    const int lpix = (tid >> 6) * W + (tid & (kTileW - 1));     // pixel offset of this thread inside a tile

    // ---- pipeline registers: tile i+1 (table entries + raw source words), tile info of i+1 and i+2 -------------------
    uint32_t tm_n[4], u[3][3];
    TileInfo ti_c, ti_n;
    int fe_n = fe, tx_n = txi, ty_n = tyi;

    auto issue_loads = [&](const TileInfo& t, int f, int x, int y) {   // all loads of one fast tile, nothing consumed here
        const int eye = f & 1, frame = f >> 1;
        const uint32_t* __restrict__ tmap = a.tmap[eye] + (y * kTileH * W + x * kTileW + lpix);
#pragma unroll
        for (int k = 0; k < 4; k++) tm_n[k] = __ldg(tmap + k * rstep);
        const uint8_t* __restrict__ src = a.src[eye] + (size_t)frame * a.src_frame;
        const int gw = t.w >> 2, ngroups = gw * t.h;
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const int g = tid + j * kRectThreads;
            u[j][0] = u[j][1] = u[j][2] = 0u;
            if (g < ngroups) {
                const int r = (int)(((uint32_t)g * t.inv_gw) >> 20), c = g - r * gw;
                const int yy = t.y0 + r, xx = t.x0 + 4 * c;           // x0 and W are multiples of 4: a group is all in or all out
                if ((unsigned)yy < (unsigned)H && (unsigned)xx < (unsigned)W) {
                    const uint32_t* p = reinterpret_cast<const uint32_t*>(src + yy * pitch + 3 * xx);
                    u[j][0] = __ldg(p); u[j][1] = __ldg(p + 1); u[j][2] = __ldg(p + 2);
                }
            }
        }
    };

    ti_c = a.tiles[fe & 1][tyi * tiles_x + txi];
    bool fast_c = tile_fast_ok(ti_c, txi, tyi);
    if (fast_c) issue_loads(ti_c, fe, txi, tyi);
    advance(fe_n, tx_n, ty_n);
    if (it + 1 < end) ti_n = a.tiles[fe_n & 1][ty_n * tiles_x + tx_n];

    for (; it < end; ++it) {
        const int eye = fe & 1, frame = fe >> 1;
        const int tx0 = txi * kTileW, ty0 = tyi * kTileH;
        const bool has_next = it + 1 < end;
        const bool fast_n = has_next && tile_fast_ok(ti_n, tx_n, ty_n);
        if (fast_c) {
            // expand the prefetched raw words of THIS tile into the staged footprint
            const int ngroups = (ti_c.w >> 2) * ti_c.h;
#pragma unroll
            for (int j = 0; j < 3; j++) {
                const int g = tid + j * kRectThreads;
                if (g < ngroups) {
                    uint4 o;
                    o.x = u[j][0];
                    o.y = __byte_perm(u[j][0], u[j][1], 0x4543);
                    o.z = __byte_perm(u[j][1], u[j][2], 0x4432);
                    o.w = u[j][2] >> 8;
                    *reinterpret_cast<uint4*>(&s_src[g << 2]) = o;
                }
            }
            uint32_t tm[4];
#pragma unroll
            for (int k = 0; k < 4; k++) tm[k] = tm_n[k];
            __syncthreads();
            if (fast_n) issue_loads(ti_n, fe_n, tx_n, ty_n);            // tile i+1 in flight during the blend of tile i

            const int fw = ti_c.w;
            const int pix0 = ty0 * W + tx0 + lpix;
            uint32_t* __restrict__ rp = rect_words(a.rect[eye] + (size_t)frame * a.rect_frame, a.rect_pitch, ty0 + (tid >> 6), tx0 + (tid & (kTileW - 1)), m);
            T* __restrict__ tp = WRITE_TENSOR ? static_cast<T*>(a.tensor[eye]) + (size_t)frame * a.tensor_frame + pix0 : nullptr;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t t = tm[k];
                const uint32_t* p = s_src + (t & 0xfffu);
                uint32_t vB, vG, vR;
                blend(p[0], p[1], p[fw], p[fw + 1], (t >> 12) & 31u, (t >> 17) & 31u, vB, vG, vR);
                const uint32_t v = vB | (vG << 8) | (vR << 16);
                const uint32_t nb = __shfl_down_sync(0xffffffffu, v, 1);
                if (m < 3) rp[k * wstep] = __byte_perm(v, nb, sel);
                if (WRITE_TENSOR) {
                    T r, g, b;
                    normalise<DTYPE, IDENT, T>(s_lut, vB, vG, vR, r, g, b);
                    tp[k * rstep] = r; tp[k * rstep + plane] = g; tp[k * rstep + 2 * plane] = b;
                }
            }
            __syncthreads();                                            // everyone is done reading s_src
        } else {
            __syncthreads();                                            // (also orders the LUT fill before its first use)
            tile_generic<DTYPE, IDENT, WRITE_TENSOR, true>(a, ti_c, eye, frame, tx0, ty0, out_align, s_src, s_out, s_lut);
            __syncthreads();
            if (fast_n) issue_loads(ti_n, fe_n, tx_n, ty_n);
        }
        // rotate the pipeline
        fe = fe_n; txi = tx_n; tyi = ty_n;
        ti_c = ti_n; fast_c = fast_n;
        advance(fe_n, tx_n, ty_n);
        if (it + 2 < end) ti_n = a.tiles[fe_n & 1][ty_n * tiles_x + tx_n];
    }
}

// ---- batch-loop variant: the tile is fixed, the CTA walks the frames --------------------------------------------------
// Everything that depends only on the calibration - the tile's footprint box, each thread's staging addresses and
// bounds predicates, the per-pixel footprint offsets and blend weights, the output offsets - is the same for every frame
// of the batch.  So a CTA takes one (tile, eye) and a run of frames: it sets those values up once, keeps them in
// registers, and per frame only (a) stores the previously fetched raw source words into the staged footprint,
// (b) issues the loads of the NEXT frame's words, (c) blends and stores its four pixels.  Per-frame work shrinks to the
// arithmetic that really depends on the pixels, and the global loads of frame f+1 are in flight during the blend of f.
// grid = (tiles_x, tiles_y, 2 * chunks), blockIdx.z = eye + 2 * chunk, chunk c covers frames [c*F, min(n, (c+1)*F)).
#ifndef SA_BATCH_MINB
#define SA_BATCH_MINB 4
#endif
#ifndef SA_BATCH_MINB_FIXED
#define SA_BATCH_MINB_FIXED 5
#endif

// prefetch.global.L2 of `p` if off >= 0, as ONE predicated instruction (a branch around it costs more than the prefetch)
__device__ __forceinline__ void l2_prefetch_if(const void* p, int off)
{
    asm volatile("{\n.reg .pred q;\nsetp.ge.s32 q, %1, 0;\n@q prefetch.global.L2 [%0];\n}\n" ::"l"(p), "r"(off));
}

// A4: every input row and every rectified row starts on a 4-byte boundary and W is a multiple of 4 (640, 1920, 3840 ...).
// !A4 (e.g. 1242-wide frames, whose rows and frames are only 2-byte aligned): source words are loaded from the enclosing
// aligned words and funnel-shifted by the address's low bits, rectified bytes leave as byte stores.
// PART: the tile reaches past the right/bottom image edge; its out-of-image pixels are masked.  Kept out of the common
// instantiation: the extra predicates cost ~15 % there.
// FW, FH: compile-time image size (0 = run-time).  With the size fixed every per-pixel output offset - 4 rows x 3 planes of
// tensor stores, 4 rows of rectified words - becomes an immediate of ONE base register that advances once per frame,
// instead of a 64-bit multiply-add per store: ~30 of 257 instructions per warp and frame.  Instantiated for 640x480, the
// only size the reference accepts (TRTFastACVNet_plus.cpp:121-122).
template <int DTYPE, bool IDENT, bool WRITE_TENSOR, bool A4, bool PART, int FW = 0, int FH = 0>
__device__ __forceinline__ void rectify_batch_body(const RectifyArgs& a, int tiles_x, int out_align, int frames_per_cta, uint32_t (*s_src2)[kFootprintWords],
                                                   typename TensorT<DTYPE>::type* s_lut)
{
    using T = typename TensorT<DTYPE>::type;
    uint32_t* const s_src = s_src2[0];
    uint32_t* const s_out = s_src2[1];                               // only the generic path needs s_out; it never overlaps the loop below

    const int tid = threadIdx.x;
    const int W = FW ? FW : a.W, H = FH ? FH : a.H;
    const int eye = blockIdx.z & 1, chunk = blockIdx.z >> 1;
    // equal chunks of frames_per_cta frames, or (chunk_taper) the tapered schedule of launch_rectify_t
    const int f0 = a.chunk_tab ? a.chunk_begin[chunk] : chunk * frames_per_cta;
    int f1 = a.chunk_tab ? a.chunk_begin[chunk + 1] : min(a.n, f0 + frames_per_cta);
    const int tx0 = blockIdx.x * kTileW, ty0 = blockIdx.y * kTileH;
    const TileInfo ti = a.tiles[eye][blockIdx.y * tiles_x + blockIdx.x];
    const bool fast = ti.w > 0;                                        // partial tiles at the right/bottom edge stay on this path, with their
                                                                       // out-of-image pixels masked (pvalid below)
    // !A4: a group of 4 source pixels that straddles x = W is loaded whole, i.e. up to 9 bytes past the end of its row.
    // In the last image row that runs into the next frame - harmless - except in the LAST frame of the batch, where it
    // would leave the caller's buffer.  Tiles whose footprint holds that group (the bottom-right one or two) therefore
    // take the bounds-checked path for that one frame only.  (An earlier version sent them down the generic path for
    // every frame: ~50x the work per frame, and with 85 frames per CTA the tail of the whole launch.)
    if (!A4 && fast && ti.y0 + ti.h >= H && ti.x0 + ti.w >= W - 4 && f1 == a.n) {
        f1--;
        __syncthreads();
        tile_generic<DTYPE, IDENT, WRITE_TENSOR, A4>(a, ti, eye, f1, tx0, ty0, out_align, s_src, s_out, s_lut);
        __syncthreads();
        if (f0 >= f1) return;
    }
    if (!fast) {                                                      // edge tile / unstaged footprint: generic path, frame by frame
        for (int f = f0; f < f1; f++) {
            __syncthreads();
            tile_generic<DTYPE, IDENT, WRITE_TENSOR, A4>(a, ti, eye, f, tx0, ty0, out_align, s_src, s_out, s_lut);
        }
        return;
    }

    // ---- per-calibration state of this thread (frame independent) ----------------------------------------------
    const int pix0 = (ty0 + (tid >> 6)) * W + tx0 + (tid & (kTileW - 1));
    const int rstep = 4 * W, plane = W * H;
    const size_t rect_pitch = FW ? (size_t)3 * FW : a.rect_pitch, rect_frame = FW ? (size_t)3 * FW * FH : a.rect_frame;
    const int wstep = (int)rect_pitch;                                // words per 4 rows of the rectified image
    uint32_t pvalid = 0xfu;             // bit k: pixel k of this thread lies inside the image
    if (PART) {
        pvalid = 0;
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (tx0 + (tid & (kTileW - 1)) < W && ty0 + (tid >> 6) + 4 * k < H) pvalid |= 1u << k;
    }
    uint32_t toff[4], twt[4];           // byte offset of the top-left tap in s_src; (32-fx, fx, 32-fy, fy) as four bytes
    {
        const uint32_t* __restrict__ tmap = a.tmap[eye] + pix0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t t = (!PART || ((pvalid >> k) & 1u)) ? __ldg(tmap + k * rstep) : 0u;
            const uint32_t fx = (t >> 12) & 31u, fy = (t >> 17) & 31u;
            toff[k] = (t & 0xfffu) << 2;
            twt[k] = (32u - fx) | (fx << 8) | ((32u - fy) << 16) | (fy << 24);
        }
    }
    const int fw4 = ti.w << 2;          // bytes per staged row
    const int ngroups = (ti.w >> 2) * ti.h;
    const int pitch = (int)a.src_pitch;
    int soff[3];                        // byte offset of this thread's staging groups inside a frame; -1 = nothing to load
    uint32_t gcut = 0;                  // !A4: 3 bits per slot, number of pixels of the group inside the image if it straddles x = W (else 0)
#pragma unroll
    for (int j = 0; j < 3; j++) {
        const int g = tid + j * kRectThreads;
        soff[j] = -1;
        if (g < ngroups) {
            const int gw = ti.w >> 2;
            const int r = (int)(((uint32_t)g * ti.inv_gw) >> 20), c = g - r * gw;
            const int yy = ti.y0 + r, xx = ti.x0 + 4 * c;             // A4: x0 and W are multiples of 4, a group is all in or all out
            if ((unsigned)yy < (unsigned)H && (unsigned)xx < (unsigned)W) {
                soff[j] = yy * pitch + 3 * xx;
                if (!A4 && xx + 4 > W) gcut |= (uint32_t)(W - xx) << (3 * j);
            }
        }
    }
    const int m = tid & 3;
    uint32_t sel = m == 0 ? 0x4210u : (m == 1 ? 0x5421u : 0x6542u);
    asm volatile("" : "+r"(sel));       // keep it in a register: recomputing the selects inside the frame loop costs 5 instructions per pixel
    const uint8_t* __restrict__ src = a.src[eye] + (size_t)f0 * a.src_frame;
    // The rectified buffer is internal and its rows are padded to 4-pixel groups on 4-byte boundaries (sa_handle::rect_pitch),
    // so the packed word stores below work for every width; lanes of a group that straddles x = W write their neighbour's
    // (unused) bytes into the row padding.
    uint32_t* __restrict__ rp = rect_words(a.rect[eye] + (size_t)f0 * rect_frame, rect_pitch, ty0 + (tid >> 6), tx0 + (tid & (kTileW - 1)), m);
    T* __restrict__ tp = WRITE_TENSOR ? static_cast<T*>(a.tensor[eye]) + (size_t)f0 * a.tensor_frame + pix0 : nullptr;
    const size_t rect_frame_words = rect_frame / 4;

    // Raw source words of the next frame.  Slots that have nothing to load (outside the image, or beyond the footprint)
    // stay zero for the whole loop: they are cleared once here and never written again.
    uint32_t u[3][A4 ? 3 : 4];
    uint32_t ualign = 0;                                              // !A4: low address bits of each slot's group, 2 bits per slot
#pragma unroll
    for (int j = 0; j < 3; j++) { u[j][0] = u[j][1] = u[j][2] = 0u; if (!A4) u[j][A4 ? 0 : 3] = 0u; }
    const bool third = ngroups > 2 * kRectThreads;                   // CTA-uniform: most footprints fit two slots per thread
    auto fetch_slot = [&](const uint8_t* __restrict__ base, int j) {
        if (soff[j] >= 0) {
            if (A4) {
                const uint32_t* p = reinterpret_cast<const uint32_t*>(base + soff[j]);
                u[j][0] = __ldg(p); u[j][1] = __ldg(p + 1); u[j][2] = __ldg(p + 2);
            } else {
                const uintptr_t A = reinterpret_cast<uintptr_t>(base + soff[j]);
                const uint32_t al = (uint32_t)A & 3u;
                const uint32_t* p = reinterpret_cast<const uint32_t*>(A - al);
                u[j][0] = __ldg(p); u[j][1] = __ldg(p + 1); u[j][2] = __ldg(p + 2);
                u[j][A4 ? 0 : 3] = al ? __ldg(p + 3) : 0u;
                ualign = (ualign & ~(3u << (2 * j))) | (al << (2 * j));
            }
        }
    };
    auto fetch = [&](const uint8_t* __restrict__ base) {
        fetch_slot(base, 0);
        fetch_slot(base, 1);
        if (third) fetch_slot(base, 2);
    };
    auto store_slot = [&](uint32_t* buf, int j) {
        const int g = tid + j * kRectThreads;
        if (g < ngroups) {
            uint32_t w0 = u[j][0], w1 = u[j][1], w2 = u[j][2];
            if (!A4) {
                const uint32_t sh = ((ualign >> (2 * j)) & 3u) << 3;
                w0 = __funnelshift_r(u[j][0], u[j][1], sh);
                w1 = __funnelshift_r(u[j][1], u[j][2], sh);
                w2 = __funnelshift_r(u[j][2], u[j][A4 ? 0 : 3], sh);
            }
            uint4 o;
            o.x = w0;
            o.y = __byte_perm(w0, w1, 0x4543);
            o.z = __byte_perm(w1, w2, 0x4432);
            o.w = w2 >> 8;
            if (!A4) {                                                // pixels of a straddling group beyond x = W are border (zero)
                const uint32_t nv = (gcut >> (3 * j)) & 7u;
                if (nv) { if (nv < 2) o.y = 0u; if (nv < 3) o.z = 0u; o.w = 0u; }
            }
            *reinterpret_cast<uint4*>(&buf[g << 2]) = o;
        }
    };
    fetch(src);
    // frames ahead (0 = off); automatic: 2, but off on the 2-byte-aligned path (measured 3 % slower there, 9 % faster elsewhere)
    const int pfd = a.l2_prefetch >= 0 ? a.l2_prefetch : (A4 ? 2 : 0);
    const uint8_t* pf = src + (size_t)pfd * a.src_frame;              // frame f + pfd, advanced with the loop

    for (int f = f0; f < f1; f++) {
        uint32_t* const buf = s_src2[(f - f0) & 1];
        store_slot(buf, 0);
        store_slot(buf, 1);
        if (third) store_slot(buf, 2);
        __syncthreads();      // the only barrier of the iteration: the other buffer was last read before the previous one
        if (f + 1 < f1) { src += a.src_frame; fetch(src); }            // next frame's words fly during this frame's blend
        if (pfd && f + pfd < f1) {                                      // and a later frame is pulled into L2, so that its fetch
            l2_prefetch_if(pf + soff[0], soff[0]);                      // is an L2 hit, not a DRAM round trip
            l2_prefetch_if(pf + soff[1], soff[1]);
            if (third) l2_prefetch_if(pf + soff[2], soff[2]);
        }
        pf += a.src_frame;

        const char* sb = reinterpret_cast<const char*>(buf);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const char* p = sb + toff[k];
            const uint32_t p00 = *reinterpret_cast<const uint32_t*>(p), p01 = *reinterpret_cast<const uint32_t*>(p + 4);
            const uint32_t p10 = *reinterpret_cast<const uint32_t*>(p + fw4), p11 = *reinterpret_cast<const uint32_t*>(p + fw4 + 4);
            const uint32_t wt = twt[k];
            const uint32_t wx = wt & 0xffffu, wxh = wt << 16;
            const uint32_t tBG = __byte_perm(p00, p01, 0x5140), tRR = __byte_perm(p00, p01, 0x6262);
            const uint32_t bBG = __byte_perm(p10, p11, 0x5140), bRR = __byte_perm(p10, p11, 0x6262);
            const uint32_t tB = __dp4a(tBG, wx, 0u), tG = __dp4a(tBG, wxh, 0u), tR = __dp4a(tRR, wx, 0u);
            const uint32_t bB = __dp4a(bBG, wx, 0u), bG = __dp4a(bBG, wxh, 0u), bR = __dp4a(bRR, wx, 0u);
            // vertical pass: (t, b) as two 16-bit lanes . (32-fy, fy) = bytes 2,3 of wt, plus the rounding constant
            const uint32_t vB = __dp2a_hi(__byte_perm(tB, bB, 0x5410), wt, 512u) >> 10;
            const uint32_t vG = __dp2a_hi(__byte_perm(tG, bG, 0x5410), wt, 512u) >> 10;
            const uint32_t vR = __dp2a_hi(__byte_perm(tR, bR, 0x5410), wt, 512u) >> 10;
            const bool pv = !PART || ((pvalid >> k) & 1u);
            {
                const uint32_t v = vB | (vG << 8) | (vR << 16);
                const uint32_t nb = __shfl_down_sync(0xffffffffu, v, 1);
                if (m < 3 && pv) rp[k * wstep] = __byte_perm(v, nb, sel);
            }
            if (WRITE_TENSOR && pv) {
                T r, g, b;
                normalise<DTYPE, IDENT, T>(s_lut, vB, vG, vR, r, g, b);
                tp[k * rstep] = r; tp[k * rstep + plane] = g; tp[k * rstep + 2 * plane] = b;
            }
        }
        rp += rect_frame_words;
        if (WRITE_TENSOR) tp += a.tensor_frame;
    }
}

// ---- warp-specialised flavour of the batch loop (default for 16-byte-aligned input) --------------------------------------------
// Same work decomposition as rectify_batch_kernel - one CTA per (tile, eye) walks a run of frames - but the two halves of a
// frame's work no longer share the threads or a CTA-wide barrier:
//   * 2 PRODUCER warps stage the footprints: a staging unit is 16 source pixels = 48 bytes = three 16-byte loads by one lane,
//     expanded to sixteen BGRx words and stored as four 16-byte shared-memory stores.  A footprint has at most 6 x 32 units
//     (typically ~100), so a lane owns two, rarely three.  They fill a ring of three BGRx stages, run up to two frames
//     ahead of the blend, and pull the frame after that into L2 (two prefetches per unit: a 48-byte unit lies in exactly two
//     32-byte sectors).  Their load latency is exposed only to themselves.
//   * 8 CONSUMER warps (256 threads, the pixel mapping of the kernel above) do nothing but blend and store: they wait on the
//     stage's "full" mbarrier, read their taps with LDS [reg + imm] (the staged rows have a compile-time stride), and release the
//     stage with one "empty" arrive per warp.  No __syncthreads, no loads, no staging arithmetic in their loop.
// In the kernel above every thread does both jobs and every frame ends at a barrier the slowest warp decides; ncu shows it
// issue-/latency-bound (issue slots 71 % busy, 1.5 warps per issue stalled at the barrier, 2.1 on the loads' scoreboard) with 83 of
// 231 instructions per thread and frame spent on staging and bookkeeping (profiles/r02_stage_a.md).
// Other changes that ride along:
//   * staged row stride 100 words and units assigned to lanes column-major: the eight lanes of a quarter-warp store to eight
//     different 16-byte bank groups (a 96-word stride with row-major units gave 4-way conflicts on every store);
//   * units that lie outside the image are zeroed once in all stages, not re-stored every frame;
//   * RAFT / CRE f16 tensors: the vertical dot product starts from 512 + (0x6400 << 10), so that after the shift the low 16 bits
//     are the half-precision bit pattern of 1024 + v; one half subtraction gives (half)v - no integer-to-float conversion
//     (I2F.F16 issues at a quarter of the ALU rate).
// Needs: image rows, frames and bases 16-byte aligned, W % 16 == 0 (a unit is then all inside or all outside the image), and
// every footprint of the calibration at most 96 pixels wide and 32 rows high (launch_rectify_t checks the calibration's largest
// footprint; a single larger tile falls back to the bounds-checked path).
constexpr int kWsStride = 100;                          // words per staged footprint row
constexpr int kWsStrideB = kWsStride * 4;
constexpr int kWsMaxW = 96, kWsMaxH = 32;               // largest footprint (pixels, rows)
constexpr int kWsStages = 2;                            // BGRx stages: one being blended, one being filled
constexpr int kWsRawStages = 3;                         // raw (packed BGR) stages: two tensor copies in flight while one is expanded
constexpr int kWsProducers = 128;                       // four warps = one warpgroup (setmaxnreg works on whole warpgroups)
constexpr int kWsThreads = kRectThreads + kWsProducers;

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"      // (a suspend-time hint makes ptxas wrap the wait in a
        "@p bra DONE_%=;\n"                                              // NANOSLEEP back-off loop: measured 2 % slower)
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// 12 source bytes -> four BGRx words (byte 3 of a staged word is never read), one 16-byte shared-memory store
__device__ __forceinline__ void put4(uint32_t addr, uint32_t w0, uint32_t w1, uint32_t w2)
{
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w0), "r"(__byte_perm(w0, w1, 0x4543)), "r"(__byte_perm(w1, w2, 0x4432)), "r"(w2 >> 8)
                 : "memory");
}
__device__ __forceinline__ void put_unit(uint32_t addr, const uint4& u0, const uint4& u1, const uint4& u2)
{
    put4(addr, u0.x, u0.y, u0.z);
    put4(addr + 16u, u0.w, u1.x, u1.y);
    put4(addr + 32u, u1.z, u1.w, u2.x);
    put4(addr + 48u, u2.y, u2.z, u2.w);
}

// ---- fused rectify -> resize -> normalise (extension a8) ---------------------------------------------------------------------------
// Same rings and roles as above, plus: the consumer warps keep the tile they have just rectified in shared memory (one BGRx word
// per pixel, 65 x 17: the tile plus the pixel column right of it and the pixel row below it, which are blended a second time
// here because the bilinear taps of the resize straddle tile edges), and the PRODUCER warpgroup - idle most of the frame - resizes
// out of it two frames later: OpenCV's 8-bit INTER_LINEAR (cv::resize, 11-bit coefficients; exact 2x decimation = rounded 2x2
// mean), normalise, store the network planes.  The rectified image is written to global memory once and never read back; the
// resized u8 image exists in registers only, except the LEFT one, which stage B needs for the point colours.  A tile owns the
// network pixels whose top-left tap lies in it; their constants (tap offset, coefficients, output index) sit in a shared-memory
// table built once per CTA.  No extra barrier is needed: when the consumers release BGRx stage (f mod 2) after blending frame f
// they have also finished tile buffer (f mod 2), and they do not touch that buffer again before the producers - who resize frame
// f - 2 right after they have handed over frame f, out of a ring of three tile buffers - say so.
constexpr int kRtStride = 68;                                    // words per row of the shared rectified tile
constexpr int kRtWords = kRtStride * (kTileH + 1);
constexpr int kRtB = kRtWords * 4;
constexpr int kRtBuffers = 3;                                    // tile of frame f in buffer f mod 3: written (f), waiting (f - 1), being resized (f - 2)

struct WsResize {
    uint32_t rt0;                 // shared-memory byte address of the two rectified-tile buffers
    uint4* rz;                    // per owned network pixel: {tap byte offsets left | right << 16, a0 | a1 << 16, b0 | b1 << 16, y * w + x | dy << 31}
    int nnet;                     // network pixels this tile owns
};

// Table of the network pixels tile (bx, by) owns; called by the 128 producer threads (p = 0..127).
__device__ __forceinline__ int ws_resize_setup(const RectifyArgs& a, uint4* rz, int tx0, int ty0, int p)
{
    const int nx_lo = __ldg(a.rs_nx + blockIdx.x), nx_n = __ldg(a.rs_nx + blockIdx.x + 1) - nx_lo;
    const int ny_lo = __ldg(a.rs_ny + blockIdx.y), ny_n = __ldg(a.rs_ny + blockIdx.y + 1) - ny_lo;
    const int nnet = nx_n * ny_n;
    const int w = a.rs_w, h = a.rs_h;
    for (int q = p; q < nnet; q += kWsProducers) {
        const int qy = q / nx_n, qx = q - qy * nx_n;
        const int X = nx_lo + qx, Y = ny_lo + qy;
        const int x0 = __ldg(a.rs_tab_x + X), x1 = __ldg(a.rs_tab_x + w + X), a0 = __ldg(a.rs_tab_x + 2 * w + X), a1 = __ldg(a.rs_tab_x + 3 * w + X);
        const int y0 = __ldg(a.rs_tab_y + Y), y1 = __ldg(a.rs_tab_y + h + Y), b0 = __ldg(a.rs_tab_y + 2 * h + Y), b1 = __ldg(a.rs_tab_y + 3 * h + Y);
        uint4 c;
        const uint32_t o00 = (uint32_t)((y0 - ty0) * kRtStride + (x0 - tx0)) * 4u;          // byte offset of the top-left tap in a tile buffer
        c.x = o00 | ((o00 + (uint32_t)(x1 - x0) * 4u) << 16);                                // ... and of the top-right one
        c.y = (uint32_t)a0 | ((uint32_t)a1 << 16);
        c.z = (uint32_t)b0 | ((uint32_t)b1 << 16);
        c.w = (uint32_t)(Y * w + X) | ((uint32_t)(y1 - y0) << 31);                           // output index; bit 31: the bottom taps are one row down
        rz[q] = c;
    }
    return nnet;
}

// Resize + normalise + store the owned network pixels of one frame out of tile buffer b; 128 producer threads.
// tf / uf: this frame's tensor (plane 0) and resized left image (null for the right eye).
template <int DTYPE, bool IDENT>
__device__ __forceinline__ void ws_resize_frame(const WsResize& rs, typename TensorT<DTYPE>::type* __restrict__ tf, uint8_t* __restrict__ uf, size_t nplane, bool area,
                                                uint32_t b, const typename TensorT<DTYPE>::type* s_lut, int p)
{
    using T = typename TensorT<DTYPE>::type;
    const uint32_t rt = rs.rt0 + b * (uint32_t)kRtB;
#ifdef SA_EXP_NO_RESIZE_BODY
    if (nplane == 1)
#endif
#pragma unroll 1      // (unrolled by two, the three warps that own one pixel per lane execute the second, empty half as well)
    for (int q = p; q < rs.nnet; q += kWsProducers) {
        const uint4 c = rs.rz[q];
        const uint32_t a00 = rt + (c.x & 0xffffu), a01 = rt + (c.x >> 16), dyb = (c.w >> 31) * (uint32_t)(kRtStride * 4), idx = c.w & 0x7fffffffu;
        uint32_t p00, p01, p10, p11;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(p00) : "r"(a00));
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(p01) : "r"(a01));
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(p10) : "r"(a00 + dyb));
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(p11) : "r"(a01 + dyb));
        uint32_t v[3];
        if (area) {                                                // rounded mean of the 2x2 block, per channel
            const uint32_t tA = __byte_perm(p00, p01, 0x5140), tB = __byte_perm(p10, p11, 0x5140);       // B0 B1 G0 G1
            const uint32_t tC = __byte_perm(p00, p01, 0x6262), tD = __byte_perm(p10, p11, 0x6262);       // R0 R1 R0 R1
            v[0] = __dp4a(__byte_perm(tA, tB, 0x5410), 0x01010101u, 2u) >> 2;
            v[1] = __dp4a(__byte_perm(tA, tB, 0x7632), 0x01010101u, 2u) >> 2;
            v[2] = __dp4a(__byte_perm(tC, tD, 0x5410), 0x01010101u, 2u) >> 2;
        } else {
            const uint32_t b0s = c.z << 16, b1s = c.z & 0xffff0000u;
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                const uint32_t s = ch == 0 ? 0x7740u : (ch == 1 ? 0x7751u : 0x7762u);    // (p_left.ch, p_right.ch, 0, 0): byte 3 of a tile word is 0
                const uint32_t T0 = __dp2a_lo(c.y, __byte_perm(p00, p01, s), 0u);           // p_left * a0 + p_right * a1
                const uint32_t T1 = __dp2a_lo(c.y, __byte_perm(p10, p11, s), 0u);
                // ((b0 * (T0 >> 4)) >> 16) + ((b1 * (T1 >> 4)) >> 16) + 2) >> 2 as OpenCV's VResizeLinear; at most 255: the coefficient
                // pairs sum to at most 2049, so the two terms sum to at most 1021
                v[ch] = (__umulhi(T0 >> 4, b0s) + __umulhi(T1 >> 4, b1s) + 2u) >> 2;
            }
        }
        T r, g, bb;
        normalise<DTYPE, IDENT, T>(s_lut, v[0], v[1], v[2], r, g, bb);
        T* const t = tf + idx;
        t[0] = r; t[nplane] = g; t[2 * nplane] = bb;
        if (uf) { uint8_t* const u = uf + idx * 3u; u[0] = (uint8_t)v[0]; u[1] = (uint8_t)v[1]; u[2] = (uint8_t)v[2]; }
    }
}

__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* tm, int c0, int c1, int c2, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)),
                 "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_prefetch_3d(const CUtensorMap* tm, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(tm), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

// Producer warps: bring frames [f0, f1) of one (tile, eye) into the BGRx ring.
// The packed source bytes of a frame's footprint arrive as ONE tensor copy (cp.async.bulk.tensor.3d: box = the largest footprint of
// the calibration, bw words x bh rows x 1 frame of a 3-D map over 4-byte words; whatever lies outside the image - negative
// coordinates included - is zero-filled by the copy engine, which is BORDER_CONSTANT 0) into one of two raw stages, completing
// on that stage's mbarrier; the copy of frame f+1 is in flight while frame f is expanded, and frame f+3 is pulled into L2 by a
// tensor prefetch.  No thread spends an instruction on global addresses, bounds or load latency.  The 64 producer lanes then
// expand the box, 16 pixels (three 16-byte shared loads -> four 16-byte shared stores) per lane and round, into the BGRx stage
// the consumers read.
template <bool RESIZE, int DTYPE, bool IDENT>
__device__ __forceinline__ void ws_produce(const RectifyArgs& a, const TileInfo ti, const CUtensorMap* tm, int eye, int bw, int bh, int f0, int f1, int pfd, uint32_t raw0,
                                           uint32_t raw_stage_b, uint32_t stage0, uint32_t stage_b, uint64_t* s_rawfull, uint64_t* s_full, uint64_t* s_empty,
                                           WsResize rs, const typename TensorT<DTYPE>::type* s_lut)
{
    const int p = threadIdx.x - kRectThreads;                          // 0..127
    const int cx = 6 * (ti.x0 >> 2), cy = ti.y0;                       // box origin: column in 16-bit elements (x0 is a multiple of 16 pixels = 48 bytes), row
    const uint32_t rowb = (uint32_t)bw * 4u, box_bytes = rowb * (uint32_t)bh;
    // 16-pixel units to expand: those of THIS tile's footprint (the copy brings the calibration's largest box; most tiles need fewer
    // rows and some fewer columns of it - one producer warp in four then has nothing to do).  The fused resize kernel expands the
    // whole box: its halo pixels tap outside the tile's own footprint.
    const int ucols = RESIZE ? bw / 12 : min(bw / 12, (ti.w + 15) >> 4), urows = RESIZE ? bh : min(bh, (int)ti.h);
    const int nunits = ucols * urows;
    if (p == 0) {
#pragma unroll
        for (int j = 0; j < kWsRawStages; j++)
            if (f0 + j < f1) {
                mbar_expect_tx(&s_rawfull[j], box_bytes);
                tma_load_3d(reinterpret_cast<void*>(__cvta_shared_to_generic(raw0 + j * raw_stage_b)), tm, cx, cy, f0 + j, &s_rawfull[j]);
            }
        for (int j = kWsRawStages; j < kWsRawStages + pfd; j++)
            if (f0 + j < f1) tma_prefetch_3d(tm, cx, cy, f0 + j);
    }
    // this lane's units (at most two: a box has at most 192): byte offset inside a raw stage / inside a BGRx stage
    uint32_t uraw[2], udst[2];
#pragma unroll
    for (int j = 0; j < 2; j++) {
        const int u = p + j * kWsProducers;
        const int c = u / urows, r = u - c * urows;                    // column-major: the lanes of a quarter-warp hit different bank groups
        uraw[j] = u < nunits ? (uint32_t)r * rowb + 48u * (uint32_t)c : 0xffffffffu;       // 0xffffffff = none
        udst[j] = (uint32_t)(r * kWsStride + 16 * c) * 4u;
    }
    const bool second = nunits > kWsProducers;                         // CTA-uniform: most boxes need one unit per lane
    auto expand = [](uint32_t ra, uint32_t da) {                       // 48 packed bytes -> sixteen BGRx words
        uint4 u0, u1, u2;
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(u0.x), "=r"(u0.y), "=r"(u0.z), "=r"(u0.w) : "r"(ra));
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4+16];" : "=r"(u1.x), "=r"(u1.y), "=r"(u1.z), "=r"(u1.w) : "r"(ra));
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4+32];" : "=r"(u2.x), "=r"(u2.y), "=r"(u2.z), "=r"(u2.w) : "r"(ra));
        put_unit(da, u0, u1, u2);
    };
    if (RESIZE) rs.nnet = ws_resize_setup(a, rs.rz, blockIdx.x * kTileW, blockIdx.y * kTileH, p);   // (each thread reads back only what it wrote)
    using TT = typename TensorT<DTYPE>::type;
    const size_t nplane = RESIZE ? (size_t)a.rs_w * a.rs_h : 0;
    TT* tf = RESIZE ? static_cast<TT*>(a.tensor[eye]) + (size_t)f0 * a.tensor_frame : nullptr;      // tensor / resized left image of the next frame to resize
    uint8_t* uf = (RESIZE && eye == 0) ? a.rs_net_left + (size_t)f0 * nplane * 3 : nullptr;
    const bool area = RESIZE && a.rs_area2x != 0;
    auto resize_next = [&](uint32_t b) {
        ws_resize_frame<DTYPE, IDENT>(rs, tf, uf, nplane, area, b, s_lut, p);
        tf += a.tensor_frame;
        if (uf) uf += nplane * 3;
    };
    uint32_t rs_i = 0, rround = 0;                                     // raw ring position
    uint32_t rzb = 0;                                                  // tile buffer of the next frame to resize
    for (int f = f0; f < f1; f++) {
        const uint32_t i = (uint32_t)(f - f0), st = i & 1u, round = i >> 1;
        mbar_wait(&s_rawfull[rs_i], rround & 1u);                       // the packed bytes of frame f have landed
        if (round) mbar_wait(&s_empty[st], (round - 1) & 1u);          // the consumers are done with the BGRx stage's previous frame (f - 2)
                                                                        // ... and, fused resize, have left its rectified tile behind
        const uint32_t rsrc = raw0 + rs_i * raw_stage_b, rdst = stage0 + st * stage_b;
        if (uraw[0] != 0xffffffffu) expand(rsrc + uraw[0], rdst + udst[0]);
        if (second && uraw[1] != 0xffffffffu) expand(rsrc + uraw[1], rdst + udst[1]);
        mbar_arrive(&s_full[st]);                                       // release: this lane's stores are visible to whoever acquires the barrier
        // the raw stage is free once both producer warps have read it: refill it with frame f + kWsRawStages
        asm volatile("bar.sync 2, %0;" ::"n"(kWsProducers) : "memory");
        if (p == 0) {
            if (f + kWsRawStages < f1) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy reads of the stage before the async-proxy write
                mbar_expect_tx(&s_rawfull[rs_i], box_bytes);
                tma_load_3d(reinterpret_cast<void*>(__cvta_shared_to_generic(rsrc)), tm, cx, cy, f + kWsRawStages, &s_rawfull[rs_i]);
            }
            if (pfd && f + kWsRawStages + pfd < f1) tma_prefetch_3d(tm, cx, cy, f + kWsRawStages + pfd);
        }
        // fused resize: frame f is on its way to the consumers; now resize frame f - 2 out of its tile buffer (the consumers
        // will not write that buffer again before frame f + 1 is handed over)
        if (RESIZE && round) { resize_next(rzb); if (++rzb == kRtBuffers) rzb = 0; }
        if (++rs_i == kWsRawStages) { rs_i = 0; ++rround; }
    }
    if (RESIZE) {                                                       // the last two frames: wait for the consumers, then resize
        const int nf = f1 - f0;
        for (int j = max(nf - 2, 0); j < nf; j++) {
            mbar_wait(&s_empty[j & 1], (uint32_t)(j >> 1) & 1u);
            resize_next(rzb);
            if (++rzb == kRtBuffers) rzb = 0;
        }
    }
}

// Consumer warps: blend frames [f0, f1) of one (tile, eye) out of the ring and store rectified bytes + tensors.
template <int DTYPE, bool IDENT, bool WRITE_TENSOR, bool PART, int FW, int FH>
__device__ __forceinline__ void ws_consume(const RectifyArgs& a, const TileInfo ti, int eye, int W, int H, int f0, int f1, int tx0, int ty0, uint32_t stage0,
                                           uint32_t stage_b, uint64_t* s_full, uint64_t* s_empty, const uint32_t* s_lut)
{
    using T = typename TensorT<DTYPE>::type;
    constexpr bool HBIAS = IDENT && WRITE_TENSOR && DTYPE == SA_F16;   // blend results carry the 0x6400 half-precision bias
    const int tid = threadIdx.x;
    const int lx = tid & (kTileW - 1), ly = tid >> 6;
    const int pix0 = (ty0 + ly) * W + tx0 + lx;
    const int rstep = 4 * W, plane = W * H;
    const size_t rect_pitch = FW ? (size_t)3 * FW : a.rect_pitch, rect_frame = FW ? (size_t)3 * FW * FH : a.rect_frame;
    const int wstep = (int)rect_pitch;                                // words per 4 rows of the rectified image
    uint32_t pvalid = 0xfu;                                           // bit k: pixel k of this thread lies inside the image
    if (PART) {
        pvalid = 0;
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (tx0 + lx < W && ty0 + ly + 4 * k < H) pvalid |= 1u << k;
    }
    uint32_t toff[4], twt[4];           // shared-memory byte address of the top-left tap in stage 0; (32-fx, fx, 32-fy, fy) as four bytes
    {
        const uint32_t* __restrict__ tmap = a.tmap[eye] + pix0;
        const uint32_t fw = (uint32_t)ti.w;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t t = (!PART || ((pvalid >> k) & 1u)) ? __ldg(tmap + k * rstep) : 0u;
            const uint32_t fx = (t >> 12) & 31u, fy = (t >> 17) & 31u, off = t & 0xfffu;
            const uint32_t ry = ((off >> 2) * ti.inv_gw) >> 20, rx = off - ry * fw;   // off / fw: the table is laid out with the tile's own row length
            toff[k] = stage0 + (ry * kWsStride + rx) * 4u;
            twt[k] = (32u - fx) | (fx << 8) | ((32u - fy) << 16) | (fy << 24);
        }
    }
    const int m = tid & 3;
    uint32_t sel = m == 0 ? 0x4210u : (m == 1 ? 0x5421u : 0x6542u);
    asm volatile("" : "+r"(sel));       // keep it in a register: recomputing the selects inside the frame loop costs 5 instructions per pixel
    uint32_t* __restrict__ rp = rect_words(a.rect[eye] + (size_t)f0 * rect_frame, rect_pitch, ty0 + ly, tx0 + lx, m);
    T* __restrict__ tp = WRITE_TENSOR ? static_cast<T*>(a.tensor[eye]) + (size_t)f0 * a.tensor_frame + pix0 : nullptr;
    const size_t rect_frame_words = rect_frame / 4;
    constexpr uint32_t kAcc = HBIAS ? (512u + (0x6400u << 10)) : 512u;
    const int lane = tid & 31;
    const char* const lut_r = reinterpret_cast<const char*>(s_lut) + 4 * (lane & 3);      // this lane's copy of the tables

    // one frame out of BGRx stage `stg` (at byte offset `so` from stage 0); the frame loop below names the two stages explicitly,
    // so that the stage index, its offset and the barrier addresses are loop constants instead of per-frame arithmetic
    auto one_frame = [&](const uint32_t so, const int stg, const uint32_t parity) {
        mbar_wait(&s_full[stg], parity);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            uint32_t p00, p01, p10, p11;
            const uint32_t ad = toff[k] + so;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(p00) : "r"(ad));
            asm volatile("ld.shared.b32 %0, [%1+4];" : "=r"(p01) : "r"(ad));
            asm volatile("ld.shared.b32 %0, [%1+%2];" : "=r"(p10) : "r"(ad), "n"(kWsStrideB));
            asm volatile("ld.shared.b32 %0, [%1+%2];" : "=r"(p11) : "r"(ad), "n"(kWsStrideB + 4));
            const uint32_t wt = twt[k];
            const uint32_t wx = wt & 0xffffu, wxh = wt << 16;
            const uint32_t tBG = __byte_perm(p00, p01, 0x5140), tRR = __byte_perm(p00, p01, 0x6262);
            const uint32_t bBG = __byte_perm(p10, p11, 0x5140), bRR = __byte_perm(p10, p11, 0x6262);
            const uint32_t tB = __dp4a(tBG, wx, 0u), tG = __dp4a(tBG, wxh, 0u), tR = __dp4a(tRR, wx, 0u);
            const uint32_t bB = __dp4a(bBG, wx, 0u), bG = __dp4a(bBG, wxh, 0u), bR = __dp4a(bRR, wx, 0u);
            // vertical pass: (t, b) as two 16-bit lanes . (32-fy, fy) = bytes 2,3 of wt, plus the rounding constant (and the half bias)
            const uint32_t vB = __dp2a_hi(__byte_perm(tB, bB, 0x5410), wt, kAcc) >> 10;
            const uint32_t vG = __dp2a_hi(__byte_perm(tG, bG, 0x5410), wt, kAcc) >> 10;
            const uint32_t vR = __dp2a_hi(__byte_perm(tR, bR, 0x5410), wt, kAcc) >> 10;
            const bool pv = !PART || ((pvalid >> k) & 1u);
            {
                // B | G << 8 | R << 16 from the low bytes (with the half bias the upper bytes are not clean)
                const uint32_t v = HBIAS ? __byte_perm(__byte_perm(vB, vG, 0x7740), vR, 0x7410) : (vB | (vG << 8) | (vR << 16));
                const uint32_t nb = __shfl_down_sync(0xffffffffu, v, 1);
                if (m < 3 && pv) rp[k * wstep] = __byte_perm(v, nb, sel);
            }
            if (WRITE_TENSOR && pv) {
                T r, g, b;
                if (HBIAS) {
                    // (half)(1024 + v) - 1024, one scalar subtraction per value (as a vector op the compiler pays for it in
                    // byte-permutes that pack and unpack the pairs)
                    unsigned short hr, hg, hb;
                    asm("sub.rn.f16 %0, %1, %2;" : "=h"(hr) : "h"((unsigned short)vR), "h"((unsigned short)0x6400));
                    asm("sub.rn.f16 %0, %1, %2;" : "=h"(hg) : "h"((unsigned short)vG), "h"((unsigned short)0x6400));
                    asm("sub.rn.f16 %0, %1, %2;" : "=h"(hb) : "h"((unsigned short)vB), "h"((unsigned short)0x6400));
                    r = (T)__ushort_as_half(hr); g = (T)__ushort_as_half(hg); b = (T)__ushort_as_half(hb);
                } else if (IDENT) {
                    normalise<DTYPE, IDENT, T>(nullptr, vB, vG, vR, r, g, b);
                } else {
                    normalise_lut4<T>(lut_r, vB, vG, vR, r, g, b);
                }
                tp[k * rstep] = r; tp[k * rstep + plane] = g; tp[k * rstep + 2 * plane] = b;
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[stg]);                      // this warp has read everything it needs from the stage
        rp += rect_frame_words;
        if (WRITE_TENSOR) tp += a.tensor_frame;
    };
    static_assert(kWsStages == 2, "the frame loop alternates between two stages");
    uint32_t parity = 0;
    for (int f = f0; f < f1; f += 2) {
        one_frame(0u, 0, parity);
        if (f + 1 < f1) one_frame(stage_b, 1, parity);
        parity ^= 1u;
    }
}

// Consumer warps of the fused resize kernel: blend frames [f0, f1), store the rectified bytes, and leave every frame's tile
// (+ halo column / row) as BGRx words in tile buffer (f - f0) mod 2 for the producers' resize (ws_resize_frame).
template <bool PART, int FW, int FH>
__device__ __forceinline__ void ws_consume_rt(const RectifyArgs& a, const TileInfo ti, int eye, int W, int H, int f0, int f1, int tx0, int ty0, uint32_t stage0,
                                              uint32_t stage_b, uint64_t* s_full, uint64_t* s_empty, uint32_t rt0)
{
    const int tid = threadIdx.x;
    const int lx = tid & (kTileW - 1), ly = tid >> 6;
    const int pix0 = (ty0 + ly) * W + tx0 + lx;
    const int rstep = 4 * W;
    const size_t rect_pitch = FW ? (size_t)3 * FW : a.rect_pitch, rect_frame = FW ? (size_t)3 * FW * FH : a.rect_frame;
    const int wstep = (int)rect_pitch;
    uint32_t pvalid = 0xfu;
    if (PART) {
        pvalid = 0;
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (tx0 + lx < W && ty0 + ly + 4 * k < H) pvalid |= 1u << k;
    }
    // own pixels (k = 0..3) and, for the first 81 threads, one halo pixel (k = 4): row 16 of the tile, then column 64, then the corner
    uint32_t toff[5], twt[5];
    {
        const uint32_t* __restrict__ tmap = a.tmap[eye] + pix0;
        const uint32_t fw = (uint32_t)ti.w;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t t = (!PART || ((pvalid >> k) & 1u)) ? __ldg(tmap + k * rstep) : 0u;
            const uint32_t fx = (t >> 12) & 31u, fy = (t >> 17) & 31u, off = t & 0xfffu;
            const uint32_t ry = ((off >> 2) * ti.inv_gw) >> 20, rx = off - ry * fw;
            toff[k] = stage0 + (ry * kWsStride + rx) * 4u;
            twt[k] = (32u - fx) | (fx << 8) | ((32u - fy) << 16) | (fy << 24);
        }
    }
    int hx = -1, hy = -1;                                            // halo pixel of this thread, tile coordinates
    if (tid < kTileW) { hx = tid; hy = kTileH; }
    else if (tid < kTileW + kTileH) { hx = kTileW; hy = tid - kTileW; }
    else if (tid == kTileW + kTileH) { hx = kTileW; hy = kTileH; }
    bool halo = hx >= 0 && tx0 + hx < W && ty0 + hy < H;
    if (halo) {
        // Needed at all?  Only the LAST network column / row a tile owns can have its left / top tap in the tile's last pixel
        // column / row (x0, y0 are non-decreasing), and only if its right / bottom tap carries weight.
        const int nx_hi = __ldg(a.rs_nx + blockIdx.x + 1), ny_hi = __ldg(a.rs_ny + blockIdx.y + 1);
        bool need_col = false, need_row = false;
        if (nx_hi > __ldg(a.rs_nx + blockIdx.x)) need_col = __ldg(a.rs_tab_x + nx_hi - 1) == tx0 + kTileW - 1 && __ldg(a.rs_tab_x + 3 * a.rs_w + nx_hi - 1) != 0;
        if (ny_hi > __ldg(a.rs_ny + blockIdx.y)) need_row = __ldg(a.rs_tab_y + ny_hi - 1) == ty0 + kTileH - 1 && __ldg(a.rs_tab_y + 3 * a.rs_h + ny_hi - 1) != 0;
        if (a.rs_area2x) need_col = need_row = false;                // 2 x 2 blocks start on even pixels: never across a tile edge
        halo = (hx == kTileW ? need_col : true) && (hy == kTileH ? need_row : true);
    }
#ifdef SA_EXP_NO_HALO
    halo = false;
#endif
    toff[4] = stage0;
    twt[4] = 0;
    if (halo) {
        // the halo pixel belongs to a neighbouring tile: its taps, relative to THIS tile's box, from the maps themselves
        const size_t mo = (size_t)(ty0 + hy) * W + tx0 + hx;
        const int m1 = __ldg(a.map1[eye] + mo);
        const uint32_t m2 = __ldg(a.map2[eye] + mo);
        const int sx = max(-2, min(W, (int)(short)(m1 & 0xffff))), sy = max(-2, min(H, m1 >> 16));
        const uint32_t fx = m2 & 31u, fy = (m2 >> 5) & 31u;
        toff[4] = stage0 + (uint32_t)((sy - ti.y0) * kWsStride + (sx - ti.x0)) * 4u;
        twt[4] = (32u - fx) | (fx << 8) | ((32u - fy) << 16) | (fy << 24);
    }
    const int m = tid & 3;
    uint32_t sel = m == 0 ? 0x4210u : (m == 1 ? 0x5421u : 0x6542u);
    asm volatile("" : "+r"(sel));
    uint32_t* __restrict__ rp = rect_words(a.rect[eye] + (size_t)f0 * rect_frame, rect_pitch, ty0 + ly, tx0 + lx, m);
    const size_t rect_frame_words = rect_frame / 4;
    const uint32_t rt_mine = rt0 + (uint32_t)(ly * kRtStride + lx) * 4u;              // this thread's pixel k = 0 in tile buffer 0
    const uint32_t rt_halo = rt0 + (uint32_t)(max(hy, 0) * kRtStride + max(hx, 0)) * 4u;
    const int lane = tid & 31;

    auto blend1 = [&](uint32_t ad, uint32_t wt) -> uint32_t {        // one rectified pixel as a BGRx word
        uint32_t p00, p01, p10, p11;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(p00) : "r"(ad));
        asm volatile("ld.shared.b32 %0, [%1+4];" : "=r"(p01) : "r"(ad));
        asm volatile("ld.shared.b32 %0, [%1+%2];" : "=r"(p10) : "r"(ad), "n"(kWsStrideB));
        asm volatile("ld.shared.b32 %0, [%1+%2];" : "=r"(p11) : "r"(ad), "n"(kWsStrideB + 4));
        const uint32_t wx = wt & 0xffffu, wxh = wt << 16;
        const uint32_t tBG = __byte_perm(p00, p01, 0x5140), tRR = __byte_perm(p00, p01, 0x6262);
        const uint32_t bBG = __byte_perm(p10, p11, 0x5140), bRR = __byte_perm(p10, p11, 0x6262);
        const uint32_t tB = __dp4a(tBG, wx, 0u), tG = __dp4a(tBG, wxh, 0u), tR = __dp4a(tRR, wx, 0u);
        const uint32_t bB = __dp4a(bBG, wx, 0u), bG = __dp4a(bBG, wxh, 0u), bR = __dp4a(bRR, wx, 0u);
        const uint32_t vB = __dp2a_hi(__byte_perm(tB, bB, 0x5410), wt, 512u) >> 10;
        const uint32_t vG = __dp2a_hi(__byte_perm(tG, bG, 0x5410), wt, 512u) >> 10;
        const uint32_t vR = __dp2a_hi(__byte_perm(tR, bR, 0x5410), wt, 512u) >> 10;
        return vB | (vG << 8) | (vR << 16);
    };

    uint32_t st = 0, round = 0, so = 0, rb = 0;                        // rb: tile buffer of frame f = (f - f0) mod kRtBuffers
    for (int f = f0; f < f1; f++) {
        mbar_wait(&s_full[st], round & 1u);                             // frame f is staged - and tile buffer rb has been resized out of
        const uint32_t rtb = rb * (uint32_t)kRtB;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t v = blend1(toff[k] + so, twt[k]);
            const bool pv = !PART || ((pvalid >> k) & 1u);
            const uint32_t nb = __shfl_down_sync(0xffffffffu, v, 1);
            if (m < 3 && pv) rp[k * wstep] = __byte_perm(v, nb, sel);
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(rt_mine + rtb + (uint32_t)(4 * k * kRtStride * 4)), "r"(v) : "memory");
        }
        if (halo) {
            const uint32_t v = blend1(toff[4] + so, twt[4]);
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(rt_halo + rtb), "r"(v) : "memory");
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[st]);                       // BGRx stage read, tile buffer written: both free for the producers
        rp += rect_frame_words;
        so += stage_b;
        if (++st == kWsStages) { st = 0; so = 0; ++round; }
        if (++rb == kRtBuffers) rb = 0;
    }
}

#ifndef SA_WS_MINB
#define SA_WS_MINB 4
#endif
template <int DTYPE, bool IDENT, bool WRITE_TENSOR, int FW = 0, int FH = 0, bool RESIZE = false>
__global__ void __launch_bounds__(kWsThreads, SA_WS_MINB)
    rectify_ws_kernel(const RectifyArgs a, const __grid_constant__ CUtensorMap tm_left, const __grid_constant__ CUtensorMap tm_right, int tiles_x, int out_align,
                      int frames_per_cta, int bw, int bh)
{
    using T = typename TensorT<DTYPE>::type;
    // dynamic shared memory, sized by the calibration's largest footprint (launch_rectify_t): kWsRawStages raw boxes of
    // raw_b bytes (bw words x bh rows, rounded up to 128), then kWsStages BGRx stages of stage_b bytes (bh rows of kWsStride words)
    extern __shared__ __align__(128) uint8_t s_dyn[];
    __shared__ __align__(8) uint64_t s_rawfull[kWsRawStages], s_full[kWsStages], s_empty[kWsStages];
    __shared__ __align__(16) uint32_t s_lut[(IDENT || RESIZE) ? 4 : kLut4Words];   // [channel][entry][copy] (fill_lut4)
    __shared__ T s_lut_plain[(RESIZE && !IDENT) ? 768 : 1];            // fused resize: few look-ups per frame, plain [3][256] table
    const uint32_t raw_b = ((uint32_t)(bw * 4 * bh) + 127u) & ~127u, stage_b = (uint32_t)(kWsStrideB * bh);
    const int tid = threadIdx.x;
    if (!IDENT && !RESIZE) fill_lut4<DTYPE>(s_lut, a.lut, kWsThreads);
    if (!IDENT && RESIZE) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = tid; i < 768; i += kWsThreads) s_lut_plain[i] = lut[i];
    }
    const int W = FW ? FW : a.W, H = FH ? FH : a.H;
    const int eye = blockIdx.z & 1, chunk = blockIdx.z >> 1;
    const int f0 = a.chunk_tab ? a.chunk_begin[chunk] : chunk * frames_per_cta;
    const int f1 = a.chunk_tab ? a.chunk_begin[chunk + 1] : min(a.n, f0 + frames_per_cta);
    const int tx0 = blockIdx.x * kTileW, ty0 = blockIdx.y * kTileH;
    const TileInfo ti = a.tiles[eye][blockIdx.y * tiles_x + blockIdx.x];
    if (tid == 0) {
        for (int st = 0; st < kWsStages; st++) { mbar_init(&s_full[st], kWsProducers); mbar_init(&s_empty[st], kRectThreads / 32); }
        for (int st = 0; st < kWsRawStages; st++) mbar_init(&s_rawfull[st], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();                                                    // barriers initialised, LUT in place; the only CTA-wide barrier
    const bool staged = ti.w > 0 && ti.w * 3 <= bw * 4 && ti.h <= bh;    // the footprint lies inside the box of the tensor copies
    if (!staged && RESIZE) return;       // cannot happen: the host takes the fused path only when every tile's box fits (launch_rectify_t)
    if (!staged) {       // unstaged / oversized footprint: bounds-checked path, frame by frame, on the 256 consumer threads
        if (tid >= kRectThreads) return;
        for (int f = f0; f < f1; f++) {
            tile_sync<true>();
            tile_generic<DTYPE, IDENT, WRITE_TENSOR, true, true>(a, ti, eye, f, tx0, ty0, out_align, reinterpret_cast<uint32_t*>(s_dyn),
                                                                 reinterpret_cast<uint32_t*>(s_dyn) + kFootprintWords, static_cast<const T*>(a.lut));   // (the launch reserves at least this much)
        }
        return;
    }
    const uint32_t raw0 = smem_u32(s_dyn), stage0 = raw0 + kWsRawStages * raw_b;
    WsResize rsz;
    rsz.rt0 = stage0 + kWsStages * stage_b;
    rsz.rz = reinterpret_cast<uint4*>(s_dyn + (rsz.rt0 - raw0) + kRtBuffers * kRtB);
    rsz.nnet = 0;
    // Register split: the kernel is launched with 40 registers per thread (4 CTAs of 384 threads per SM); the producer
    // warpgroup gives back down to 24 - it holds one 48-byte unit and a few addresses - and the two consumer warpgroups take
    // 48 each: 128 x 24 + 256 x 48 = 384 x 40.  (Fused resize: the producers also resize, everybody stays at 40.)
    if (tid >= kRectThreads) {
        if (!RESIZE) asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
        ws_produce<RESIZE, DTYPE, IDENT>(a, ti, eye ? &tm_right : &tm_left, eye, bw, bh, f0, f1, a.l2_prefetch >= 0 ? a.l2_prefetch : 2, raw0, raw_b, stage0, stage_b,
                                         s_rawfull, s_full, s_empty, rsz, s_lut_plain);
        return;
    }
    constexpr bool whole = FW && FW % kTileW == 0 && FH % kTileH == 0;   // no partial tiles at all
    const bool full = whole || (tx0 + kTileW <= W && ty0 + kTileH <= H);
    if (RESIZE) {
        if (full) ws_consume_rt<false, FW, FH>(a, ti, eye, W, H, f0, f1, tx0, ty0, stage0, stage_b, s_full, s_empty, rsz.rt0);
        else ws_consume_rt<true, FW, FH>(a, ti, eye, W, H, f0, f1, tx0, ty0, stage0, stage_b, s_full, s_empty, rsz.rt0);
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 48;");
    if (full) ws_consume<DTYPE, IDENT, WRITE_TENSOR, false, FW, FH>(a, ti, eye, W, H, f0, f1, tx0, ty0, stage0, stage_b, s_full, s_empty, s_lut);
    else ws_consume<DTYPE, IDENT, WRITE_TENSOR, true, FW, FH>(a, ti, eye, W, H, f0, f1, tx0, ty0, stage0, stage_b, s_full, s_empty, s_lut);
}

template <int DTYPE, bool IDENT, bool WRITE_TENSOR, bool A4, int FW = 0, int FH = 0>
// 5 CTAs per SM for the fixed-size identity-normalisation instantiations (RAFT / CRE): without the address arithmetic they
// fit 48 registers without spilling (0.102 -> 0.099 ms); the LUT variants spill at 48 and lose 8 %, so they stay at 4.
__global__ void __launch_bounds__(kRectThreads, (FW && IDENT) ? SA_BATCH_MINB_FIXED : SA_BATCH_MINB) rectify_batch_kernel(const RectifyArgs a, int tiles_x, int out_align, int frames_per_cta)
{
    using T = typename TensorT<DTYPE>::type;
    __shared__ __align__(16) uint32_t s_src2[2][kFootprintWords];   // double buffered: one barrier per frame
    // (the four-copy table of the warp-specialised kernel measured 5 % SLOWER here on 1242x375: plain [3][256] table)
    __shared__ T s_lut[IDENT ? 1 : 768];
    if (!IDENT) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = threadIdx.x; i < 768; i += kRectThreads) s_lut[i] = lut[i];
    }
    if (FW) {                                                        // fixed sizes are multiples of the tile: no partial tiles
        static_assert(FW % kTileW == 0 && FH % kTileH == 0, "fixed-size instantiation needs whole tiles");
        rectify_batch_body<DTYPE, IDENT, WRITE_TENSOR, A4, false, FW, FH>(a, tiles_x, out_align, frames_per_cta, s_src2, s_lut);
        return;
    }
    const bool full = (int)(blockIdx.x + 1) * kTileW <= a.W && (int)(blockIdx.y + 1) * kTileH <= a.H;
    if (full) rectify_batch_body<DTYPE, IDENT, WRITE_TENSOR, A4, false>(a, tiles_x, out_align, frames_per_cta, s_src2, s_lut);
    else rectify_batch_body<DTYPE, IDENT, WRITE_TENSOR, A4, true>(a, tiles_x, out_align, frames_per_cta, s_src2, s_lut);
}

// Encodes the 3-D tensor map of one eye's source batch.  The driver entry point is looked up once (no link-time libcuda).
static bool encode_source_map(CUtensorMap* tm, const uint8_t* base, int W, int H, int n, size_t pitch, size_t frame, int bw, int bh)
{
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (encode_fn)p;
    }
    if (!fn) return false;
    // 16-bit elements: a row is 3W/2 of them for every even W (1242 included; 4-byte elements would need W % 4 == 0), a box row of
    // up to 96 pixels is 144 <= 256 of them, and bytes beyond 3W - the row padding, the next row - are outside the tensor: zero-filled
    const cuuint64_t dims[3] = {(cuuint64_t)(3 * W / 2), (cuuint64_t)H, (cuuint64_t)n};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frame};
    const cuuint32_t box[3] = {(cuuint32_t)(2 * bw), (cuuint32_t)bh, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    return fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, const_cast<uint8_t*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// rectify = 0 (images already rectified) and the tail of the resize path: BGR u8 -> normalised planes only.
template <int DTYPE, bool IDENT>
__global__ void __launch_bounds__(256) preprocess_kernel(const PreprocArgs a)
{
    using T = typename TensorT<DTYPE>::type;
    __shared__ T s_lut[IDENT ? 1 : 768];
    if (!IDENT) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = threadIdx.x; i < 768; i += 256) s_lut[i] = lut[i];
        __syncthreads();
    }
    const int eye = blockIdx.y & 1, frame = blockIdx.y >> 1;
    const size_t plane = (size_t)a.W * a.H;
    const size_t p = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= plane) return;
    const int y = (int)(p / a.W), x = (int)(p - (size_t)y * a.W);
    const uint8_t* v = (eye ? a.src[1] : a.src[0]) + (size_t)frame * a.src_frame + (size_t)y * a.src_pitch + 3 * x;
    const uint32_t vB = v[0], vG = v[1], vR = v[2];
    T* t = static_cast<T*>(eye ? a.tensor[1] : a.tensor[0]) + (size_t)frame * a.tensor_frame + p;
    T r, g, b;
    if (IDENT) {
        if (DTYPE == SA_F16) { r = (T)__ushort2half_rn((unsigned short)vR); g = (T)__ushort2half_rn((unsigned short)vG); b = (T)__ushort2half_rn((unsigned short)vB); }
        else { r = (T)(float)vR; g = (T)(float)vG; b = (T)(float)vB; }
    } else {
        r = s_lut[vR]; g = s_lut[256 + vG]; b = s_lut[512 + vB];
    }
    t[0] = r; t[plane] = g; t[2 * plane] = b;
}

// Same operation, four pixels per thread: for rows that start on 4-byte boundaries and widths that are a multiple of 4 (every
// 640 / 1280 / 1920-wide input; the resized network images) a thread reads its 12 bytes as three words and writes four
// consecutive elements of each plane with one 16-byte (f32) or 8-byte (f16) store.  A CTA covers 4096 pixels so that the LUT
// is staged once per 4096 look-ups; a thread's four groups are 256 groups apart (a warp reads 384 contiguous bytes).
// Measured on 64 pairs of 640x480 f32: 0.183 ms with one pixel per thread (byte loads, 4-byte stores) -> see DESIGN.md section 5.
constexpr int kPreGroups = 4;     // 4-pixel groups per thread

template <int DTYPE, bool IDENT>
__global__ void __launch_bounds__(256) preprocess4_kernel(const PreprocArgs a)
{
    using T = typename TensorT<DTYPE>::type;
    __shared__ T s_lut[IDENT ? 1 : 768];
    if (!IDENT) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = threadIdx.x; i < 768; i += 256) s_lut[i] = lut[i];
    }
    const int eye = blockIdx.y & 1, frame = blockIdx.y >> 1;
    const int gw = a.W >> 2;                                       // groups per row
    const int ngroups = gw * a.H;
    const uint8_t* src = (eye ? a.src[1] : a.src[0]) + (size_t)frame * a.src_frame;
    T* tbase = static_cast<T*>(eye ? a.tensor[1] : a.tensor[0]) + (size_t)frame * a.tensor_frame;
    const size_t plane = (size_t)a.W * a.H;
    uint32_t u[kPreGroups][3];
    int gi[kPreGroups];
#pragma unroll
    for (int j = 0; j < kPreGroups; j++) {                            // all loads first
        const int g = (blockIdx.x * kPreGroups + j) * 256 + threadIdx.x;
        gi[j] = g < ngroups ? g : -1;
        u[j][0] = u[j][1] = u[j][2] = 0u;
        if (g < ngroups) {
            const int y = g / gw, x4 = g - y * gw;
            const uint32_t* p = reinterpret_cast<const uint32_t*>(src + (size_t)y * a.src_pitch) + 3 * x4;
            u[j][0] = __ldg(p); u[j][1] = __ldg(p + 1); u[j][2] = __ldg(p + 2);
        }
    }
    if (!IDENT) __syncthreads();
#pragma unroll
    for (int j = 0; j < kPreGroups; j++) {
        if (gi[j] < 0) continue;
        // bytes: B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
        const uint32_t w0 = u[j][0], w1 = u[j][1], w2 = u[j][2];
        const uint32_t B[4] = {w0 & 0xffu, w0 >> 24, (w1 >> 16) & 0xffu, (w2 >> 8) & 0xffu};
        const uint32_t G[4] = {(w0 >> 8) & 0xffu, w1 & 0xffu, w1 >> 24, (w2 >> 16) & 0xffu};
        const uint32_t Rr[4] = {(w0 >> 16) & 0xffu, (w1 >> 8) & 0xffu, w2 & 0xffu, w2 >> 24};
        T r[4], g[4], b[4];
#pragma unroll
        for (int k = 0; k < 4; k++) normalise<DTYPE, IDENT, T>(s_lut, B[k], G[k], Rr[k], r[k], g[k], b[k]);
        T* t = tbase + (size_t)gi[j] * 4;
        if (DTYPE == SA_F16) {
            struct alignas(8) H4 { T v[4]; };
            *reinterpret_cast<H4*>(t) = H4{{r[0], r[1], r[2], r[3]}};
            *reinterpret_cast<H4*>(t + plane) = H4{{g[0], g[1], g[2], g[3]}};
            *reinterpret_cast<H4*>(t + 2 * plane) = H4{{b[0], b[1], b[2], b[3]}};
        } else {
            struct alignas(16) F4 { T v[4]; };
            *reinterpret_cast<F4*>(t) = F4{{r[0], r[1], r[2], r[3]}};
            *reinterpret_cast<F4*>(t + plane) = F4{{g[0], g[1], g[2], g[3]}};
            *reinterpret_cast<F4*>(t + 2 * plane) = F4{{b[0], b[1], b[2], b[3]}};
        }
    }
}

// Dense frames of ANY width (1242x375: W % 4 != 0, frames and planes only 2- or 8-byte aligned): with contiguous rows a frame is
// one linear run of pixels, so a thread again takes four of them - 12 bytes from the enclosing aligned words, funnel-shifted
// by the address's low bits - and writes four consecutive elements of each plane, in stores as wide as the alignment of the
// planes allows (SW bytes: 16 / 8 / 4 for f32, 8 / 4 / 2 for f16).  The npix % 4 pixels at the end of a frame are done by
// the first threads of the last CTA one at a time.
template <int DTYPE, bool IDENT, int SW>
__global__ void __launch_bounds__(256) preprocess_lin_kernel(const PreprocArgs a)
{
    using T = typename TensorT<DTYPE>::type;
    __shared__ T s_lut[IDENT ? 1 : 768];
    if (!IDENT) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = threadIdx.x; i < 768; i += 256) s_lut[i] = lut[i];
    }
    const int eye = blockIdx.y & 1, frame = blockIdx.y >> 1;
    const size_t plane = (size_t)a.W * a.H;
    const int ngroups = (int)(plane >> 2);
    const uint8_t* src = (eye ? a.src[1] : a.src[0]) + (size_t)frame * a.src_frame;
    T* tbase = static_cast<T*>(eye ? a.tensor[1] : a.tensor[0]) + (size_t)frame * a.tensor_frame;
    uint32_t u[kPreGroups][4];
    uint32_t sh[kPreGroups];
    int gi[kPreGroups];
#pragma unroll
    for (int j = 0; j < kPreGroups; j++) {                            // all loads first
        const int g = (blockIdx.x * kPreGroups + j) * 256 + threadIdx.x;
        gi[j] = g < ngroups ? g : -1;
        u[j][0] = u[j][1] = u[j][2] = u[j][3] = 0u;
        sh[j] = 0u;
        if (g < ngroups) {
            const uintptr_t A = reinterpret_cast<uintptr_t>(src + (size_t)g * 12);
            const uint32_t al = (uint32_t)A & 3u;
            const uint32_t* p = reinterpret_cast<const uint32_t*>(A - al);
            u[j][0] = __ldg(p); u[j][1] = __ldg(p + 1); u[j][2] = __ldg(p + 2);
            if (al) u[j][3] = __ldg(p + 3);                           // holds bytes of the group whenever it is needed: never past the data
            sh[j] = al << 3;
        }
    }
    if (!IDENT) __syncthreads();
#pragma unroll
    for (int j = 0; j < kPreGroups; j++) {
        if (gi[j] < 0) continue;
        const uint32_t w0 = __funnelshift_r(u[j][0], u[j][1], sh[j]), w1 = __funnelshift_r(u[j][1], u[j][2], sh[j]), w2 = __funnelshift_r(u[j][2], u[j][3], sh[j]);
        const uint32_t B[4] = {w0 & 0xffu, w0 >> 24, (w1 >> 16) & 0xffu, (w2 >> 8) & 0xffu};
        const uint32_t G[4] = {(w0 >> 8) & 0xffu, w1 & 0xffu, w1 >> 24, (w2 >> 16) & 0xffu};
        const uint32_t Rr[4] = {(w0 >> 16) & 0xffu, (w1 >> 8) & 0xffu, w2 & 0xffu, w2 >> 24};
        T v[3][4];
#pragma unroll
        for (int k = 0; k < 4; k++) normalise<DTYPE, IDENT, T>(s_lut, B[k], G[k], Rr[k], v[0][k], v[1][k], v[2][k]);
        T* t = tbase + (size_t)gi[j] * 4;
        constexpr int per = SW / (int)sizeof(T);                      // elements per store: 4, 2 or 1
#pragma unroll
        for (int c = 0; c < 3; c++) {
            T* q = t + (size_t)c * plane;
            if (per == 4) { struct alignas(SW) V { T e[4]; }; *reinterpret_cast<V*>(q) = V{{v[c][0], v[c][1], v[c][2], v[c][3]}}; }
            else if (per == 2) {
                struct alignas(SW) V { T e[2]; };
                *reinterpret_cast<V*>(q) = V{{v[c][0], v[c][1]}};
                *reinterpret_cast<V*>(q + 2) = V{{v[c][2], v[c][3]}};
            } else { q[0] = v[c][0]; q[1] = v[c][1]; q[2] = v[c][2]; q[3] = v[c][3]; }
        }
    }
    // the last npix % 4 pixels of the frame
    const int tail = (int)(plane & 3);
    if (blockIdx.x == gridDim.x - 1 && (int)threadIdx.x < tail) {
        const size_t p = (plane & ~(size_t)3) + threadIdx.x;
        const uint8_t* q = src + p * 3;
        T r, g, b;
        normalise<DTYPE, IDENT, T>(s_lut, q[0], q[1], q[2], r, g, b);
        tbase[p] = r; tbase[plane + p] = g; tbase[2 * plane + p] = b;
    }
}

inline bool lut_is_identity(int model) { return model == SA_MODEL_RAFTSTEREO || model == SA_MODEL_CRESTEREO; }

// want_fused != nullptr: launch only the fused rectify + resize kernel, if it applies (*want_fused = 1), else nothing
template <int DTYPE, bool IDENT, bool WT>
cudaError_t launch_rectify_t(const RectifyArgs& a, bool aligned4, int tiles_x, int out_align, int fast_ok, dim3 grid, cudaStream_t s, int* want_fused = nullptr)
{
    if (a.batch_loop && !a.force_generic && a.src_pitch * (size_t)a.H < (1u << 31) && (size_t)a.W * a.H * 3 < (1u << 30)) {
        const bool a4 = fast_ok && aligned4 && !a.force_unaligned;
        // the warp-specialised kernel does not need W % 4 == 0 (fast_ok does: the batch kernel's 12-byte groups)
        const bool fast_ok_any_w = out_align >= 4 && !a.force_generic;
        const int tiles_y = (a.H + kTileH - 1) / kTileH;
        // frames per CTA.  Many short CTAs balance better (partial tiles at the image edge take longer than full ones, and
        // the last wave of a launch is only as fast as its slowest CTA); long CTAs amortise the per-tile set-up.  Measured
        // (profiles/r01_workloads.md): ~128 CTAs per SM as long as each still walks >= 12 frames.  batch_ctas > 0 fixes
        // the CTAs per SM instead (tests use it to force one frame per CTA or the whole batch in one CTA).
        const long long tile_eyes = (long long)tiles_x * tiles_y * 2;
        const long long want = (long long)a.batch_loop * (a.batch_ctas_per_sm > 0 ? a.batch_ctas_per_sm : 128);   // a.batch_loop carries the SM count
        int fpc = (int)((tile_eyes * a.n + want - 1) / want);
        if (fpc < 1) fpc = 1;
        if (a.batch_ctas_per_sm <= 0) {
            if (fpc < 12) fpc = 12;
            const int whole = a.n / fpc > 0 ? a.n / fpc : 1;            // chunks of at least fpc frames ...
            fpc = (a.n + whole - 1) / whole;                            // ... of equal length
        }
        if (fpc > a.n) fpc = a.n;
        const int chunks = (a.n + fpc - 1) / fpc;
        // Tapered chunks: CTAs are dispatched in grid order (chunk index slowest), so the chunks that start last are made the
        // shortest - the launch then ends with many short CTAs instead of a few long ones.  Chunk c of C gets
        // avg * (1 + t * (1 - 2c/(C-1))) frames (t = chunk_taper / 100; the sizes still sum to n).
        RectifyArgs at = a;
        at.chunk_tab = 0;
        if (a.chunk_taper > 0 && chunks >= 2 && chunks <= kMaxChunks) {
            const double t = a.chunk_taper / 100.0, avg = (double)a.n / chunks;
            double acc = 0.0;
            int pos = 0;
            for (int c = 0; c < chunks; c++) {
                at.chunk_begin[c] = pos;
                acc += avg * (1.0 + t * (1.0 - 2.0 * c / (chunks - 1)));
                int next = (int)(acc + 0.5);
                if (next <= pos) next = pos + 1;                       // every chunk has at least one frame ...
                if (next > a.n - (chunks - 1 - c)) next = a.n - (chunks - 1 - c);   // ... and leaves one for each chunk after it
                pos = next;
            }
            at.chunk_begin[chunks] = a.n;
            at.chunk_tab = 1;
        }
        if (2 * chunks <= 65535) {
            const dim3 g((unsigned)tiles_x, (unsigned)tiles_y, 2u * (unsigned)chunks);
            // 16-byte-aligned input and footprints that fit a ring stage: the warp-specialised kernel
            // (any even width: what matters is that every ROW starts 16-byte aligned - dense 640 / 1920 / 3840-wide frames, or a
            // padded pitch such as cudaMallocPitch / cv::cuda::GpuMat give a 1242-wide one)
            // Widths that are not a multiple of 16 take it only on request (option bulk = 2): measured on 256 pairs of 1242x375 with
            // 3728-byte rows it is slower than the batch kernel (1.18 against 0.98 ms; ncu: warps parked on the copies' mbarriers).
            const bool a16 = fast_ok_any_w && !a.no_align16 && !a.force_unaligned && (a.W % 16 == 0 || (a.W % 2 == 0 && a.bulk_copy >= 2)) && a.src_pitch % 16 == 0 &&
                             a.src_frame % 16 == 0 &&
                             ((uintptr_t)a.src[0] | (uintptr_t)a.src[1]) % 16 == 0 && a.box_w > 0 && a.box_w % 16 == 0 && a.box_w <= kWsMaxW && a.box_h <= kWsMaxH;
            if (a16) {
                // box of the tensor copies = the largest footprint of the calibration (a multiple of 16 pixels wide), in 4-byte words
                const int bw = a.box_w * 3 / 4, bh = a.box_h;
                CUtensorMap tl, tr;
                if (encode_source_map(&tl, a.src[0], a.W, a.H, a.n, a.src_pitch, a.src_frame, bw, bh) &&
                    encode_source_map(&tr, a.src[1], a.W, a.H, a.n, a.src_pitch, a.src_frame, bw, bh)) {
                    size_t dyn = (size_t)kWsRawStages * (((size_t)bw * 4 * bh + 127) & ~(size_t)127) + (size_t)kWsStages * kWsStrideB * bh;
                    if (want_fused) {
                        // + two rectified-tile buffers and the table of the network pixels a tile owns
                        dyn += (size_t)kRtBuffers * kRtB + 16 * (size_t)a.rs_max_net;
                        if (dyn > 96 * 1024) return cudaSuccess;           // (a near-1:1 resize of a strongly distorted camera: three launches)
#define SA_WS_RESIZE(FWV, FHV)                                                                                                      \
    do {                                                                                                                            \
        auto kern = rectify_ws_kernel<DTYPE, IDENT, true, FWV, FHV, true>;                                                           \
        if (dyn > 30 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);                     \
        kern<<<g, kWsThreads, dyn, s>>>(at, tl, tr, tiles_x, out_align, fpc, bw, bh);                                                \
    } while (0)
                        if (a.W == 1920 && a.H == 1080 && !a.no_fixed_size) SA_WS_RESIZE(1920, 1080);
                        else SA_WS_RESIZE(0, 0);
#undef SA_WS_RESIZE
                        *want_fused = 1;
                        return cudaGetLastError();
                    }
                    const size_t generic_need = (size_t)(kFootprintWords + kTileH * kTileW * 3 / 4) * 4;   // scratch of the bounds-checked path
                    if (dyn < generic_need) dyn = generic_need;
#define SA_WS_LAUNCH(FWV, FHV)                                                                                                      \
    do {                                                                                                                            \
        auto kern = rectify_ws_kernel<DTYPE, IDENT, WT, FWV, FHV>;                                                                   \
        /* static (barriers, 12 KB of look-up tables) + dynamic beyond the 48 KB default needs the opt-in */                         \
        if (dyn > 30 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);                     \
        kern<<<g, kWsThreads, dyn, s>>>(at, tl, tr, tiles_x, out_align, fpc, bw, bh);                                                \
    } while (0)
                    if (a.W == 640 && a.H == 480 && !a.no_fixed_size) SA_WS_LAUNCH(640, 480);
                    else if (a.W == 1920 && a.H == 1080 && !a.no_fixed_size) SA_WS_LAUNCH(1920, 1080);
                    else SA_WS_LAUNCH(0, 0);
#undef SA_WS_LAUNCH
                    return cudaGetLastError();
                }
            }
            if (want_fused) return cudaSuccess;                        // the fused kernel does not apply: the caller takes the three-launch path
            if (a4 && a.W == 640 && a.H == 480 && !a.no_fixed_size) rectify_batch_kernel<DTYPE, IDENT, WT, true, 640, 480><<<g, kRectThreads, 0, s>>>(at, tiles_x, out_align, fpc);
            else if (a4) rectify_batch_kernel<DTYPE, IDENT, WT, true><<<g, kRectThreads, 0, s>>>(at, tiles_x, out_align, fpc);
            else rectify_batch_kernel<DTYPE, IDENT, WT, false><<<g, kRectThreads, 0, s>>>(at, tiles_x, out_align, fpc);
            return cudaGetLastError();
        }
    }
    if (want_fused) return cudaSuccess;
    if (fast_ok && aligned4 && a.persistent_ctas > 0) {
        const int tiles_y = (a.H + kTileH - 1) / kTileH;
        const long long total = (long long)tiles_x * tiles_y * 2 * a.n;
        const unsigned ctas = (unsigned)(total < a.persistent_ctas ? total : a.persistent_ctas);
        rectify_persistent_kernel<DTYPE, IDENT, WT><<<ctas, kRectThreads, 0, s>>>(a, tiles_x, tiles_y, out_align);
        return cudaGetLastError();
    }
    if (aligned4) rectify_kernel<DTYPE, IDENT, WT, true><<<grid, kRectThreads, 0, s>>>(a, tiles_x, out_align, fast_ok);
    else rectify_kernel<DTYPE, IDENT, WT, false><<<grid, kRectThreads, 0, s>>>(a, tiles_x, out_align, fast_ok);
    return cudaGetLastError();
}

}  // namespace

// largest power of two (capped at 256) dividing v
static int low_align(uintptr_t v) { if (v == 0) return 256; uintptr_t l = v & (~v + 1); return l > 256 ? 256 : (int)l; }
static int ptr_align(const void* p) { return low_align((uintptr_t)p); }
static int size_align(size_t v) { return low_align((uintptr_t)v); }
static int min3(int a, int b, int c) { return a < b ? (a < c ? a : c) : (b < c ? b : c); }

cudaError_t launch_rectify(const RectifyArgs& a, cudaStream_t s, long* launches, int* fused)
{
    if (fused) *fused = 0;
    if (a.n <= 0) return cudaSuccess;
    const int tiles_x = (a.W + kTileW - 1) / kTileW, tiles_y = (a.H + kTileH - 1) / kTileH;
    if (tiles_x > 65535 || tiles_y > 65535 || 2 * a.n > 65535) return cudaErrorInvalidConfiguration;
    const dim3 grid = a.frame_fastest ? dim3(2u * (unsigned)a.n, (unsigned)tiles_x, (unsigned)tiles_y)
                                      : dim3((unsigned)tiles_x, (unsigned)tiles_y, 2u * (unsigned)a.n);
    // input alignment: every row of every frame of both eyes starts on a 4-byte boundary
    const bool aligned4 = min3(ptr_align(a.src[0]), ptr_align(a.src[1]), min3(size_align(a.src_pitch), size_align(a.src_frame), 256)) >= 4;
    // output alignment of the rectified rows (tile x offsets are multiples of 192 bytes)
    const int out_align = min3(ptr_align(a.rect[0]), ptr_align(a.rect[1]), min3(size_align(a.rect_pitch), size_align(a.rect_frame), 256));
    // fast path: 4-byte aligned rectified rows and 32-bit offsets inside a frame
    const int fast_ok = (out_align >= 4 && a.W % 4 == 0 && a.src_pitch * (size_t)a.H < (1u << 31) && (size_t)a.W * a.H * 3 < (1u << 30) &&
                         !a.force_generic) ? 1 : 0;
    const bool ident = lut_is_identity(a.model);
    cudaError_t e;
    if (fused && a.rs_w > 0 && a.fused_resize && a.unstaged_tiles == 0 && a.rs_net_left && a.tensor[0]) {
        // rectify -> resize -> normalise in one launch (the rectified image never comes back from global memory)
        int did = 0;
#define SA_FUSED(DT, ID) e = launch_rectify_t<DT, ID, true>(a, aligned4, tiles_x, out_align, fast_ok, grid, s, &did)
        if (a.dtype == SA_F16) { if (ident) SA_FUSED(SA_F16, true); else SA_FUSED(SA_F16, false); }
        else { if (ident) SA_FUSED(SA_F32, true); else SA_FUSED(SA_F32, false); }
#undef SA_FUSED
        if (e != cudaSuccess) return e;
        if (did) { *fused = 1; if (launches) ++*launches; return cudaSuccess; }
    }
#define SA_RECT(DT, ID, WT) e = launch_rectify_t<DT, ID, WT>(a, aligned4, tiles_x, out_align, fast_ok, grid, s)
    if (!a.write_tensor) SA_RECT(SA_F32, true, false);
    else if (a.dtype == SA_F16) { if (ident) SA_RECT(SA_F16, true, true); else SA_RECT(SA_F16, false, true); }
    else { if (ident) SA_RECT(SA_F32, true, true); else SA_RECT(SA_F32, false, true); }
#undef SA_RECT
    if (launches) ++*launches;
    return e;
}

cudaError_t launch_preprocess(const PreprocArgs& a, cudaStream_t s, long* launches)
{
    if (a.n <= 0) return cudaSuccess;
    const size_t plane = (size_t)a.W * a.H;
    const bool ident = lut_is_identity(a.model);
    const size_t telt = a.dtype == SA_F16 ? 2 : 4;
    // four pixels per thread: word-aligned source rows, W % 4 == 0, and planes / frames of the tensor on 16-byte boundaries
    const bool vec = !a.force_scalar && a.W % 4 == 0 && ((uintptr_t)a.src[0] | (uintptr_t)a.src[1] | a.src_pitch | a.src_frame) % 4 == 0 &&
                     ((uintptr_t)a.tensor[0] | (uintptr_t)a.tensor[1] | (a.tensor_frame * telt) | (plane * telt)) % 16 == 0 && plane / 4 < (1u << 30);
    if (vec) {
        const unsigned per_cta = 256u * kPreGroups;
        dim3 grid((unsigned)((plane / 4 + per_cta - 1) / per_cta), 2u * (unsigned)a.n);
        if (a.dtype == SA_F16) {
            if (ident) preprocess4_kernel<SA_F16, true><<<grid, 256, 0, s>>>(a); else preprocess4_kernel<SA_F16, false><<<grid, 256, 0, s>>>(a);
        } else {
            if (ident) preprocess4_kernel<SA_F32, true><<<grid, 256, 0, s>>>(a); else preprocess4_kernel<SA_F32, false><<<grid, 256, 0, s>>>(a);
        }
    } else if (!a.force_scalar && a.src_pitch == (size_t)a.W * 3 && plane >= 4 && plane / 4 < (1u << 30) &&
               ((uintptr_t)a.tensor[0] | (uintptr_t)a.tensor[1]) % 16 == 0) {
        // dense frames of any width: linear four-pixel kernel, stores as wide as the planes' alignment allows
        const size_t al = (a.tensor_frame * telt) | (plane * telt);
        const int sw = al % 16 == 0 ? 16 : (al % 8 == 0 ? 8 : (al % 4 == 0 ? 4 : 2));
        const unsigned per_cta = 256u * kPreGroups;
        dim3 grid((unsigned)((plane / 4 + per_cta - 1) / per_cta), 2u * (unsigned)a.n);
#define SA_LIN(DT, SWB) do { if (ident) preprocess_lin_kernel<DT, true, SWB><<<grid, 256, 0, s>>>(a); else preprocess_lin_kernel<DT, false, SWB><<<grid, 256, 0, s>>>(a); } while (0)
        if (a.dtype == SA_F16) { if (sw >= 8) SA_LIN(SA_F16, 8); else if (sw == 4) SA_LIN(SA_F16, 4); else SA_LIN(SA_F16, 2); }
        else { if (sw == 16) SA_LIN(SA_F32, 16); else if (sw == 8) SA_LIN(SA_F32, 8); else SA_LIN(SA_F32, 4); }
#undef SA_LIN
    } else {
        dim3 grid((unsigned)((plane + 255) / 256), 2u * (unsigned)a.n);
        if (a.dtype == SA_F16) {
            if (ident) preprocess_kernel<SA_F16, true><<<grid, 256, 0, s>>>(a); else preprocess_kernel<SA_F16, false><<<grid, 256, 0, s>>>(a);
        } else {
            if (ident) preprocess_kernel<SA_F32, true><<<grid, 256, 0, s>>>(a); else preprocess_kernel<SA_F32, false><<<grid, 256, 0, s>>>(a);
        }
    }
    if (launches) ++*launches;
    return cudaGetLastError();
}

}  // namespace sa
