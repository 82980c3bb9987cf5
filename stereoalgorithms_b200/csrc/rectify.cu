// Stage A: rectification fused with the network preprocess, one HBM pass per eye-frame.
//
// Replaces, per stereo pair, the reference's  cv::remap x2 on the CPU
// (FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:142-143)  +  H2D  +  <Model>_preprocess_kernel x2
// (FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:4-39 and the three siblings).
//
// One CTA (256 threads) produces a 64x16 tile of the rectified image:
//   1. each thread reads the map entries of its 4 pixels (coalesced; maps are batch-shared and L2 resident);
//   2. the CTA stages the tile's source footprint - a bounding box precomputed per tile at calibration time -
//      from global memory into shared memory as one BGRx word per pixel, zero outside the image, so that
//      BORDER_CONSTANT needs no per-tap test afterwards;
//   3. each thread blends its pixels with OpenCV's exact fixed-point arithmetic
//         dst = (p00*w00 + p01*w01 + p10*w10 + p11*w11 + 2^14) >> 15,  w = (32-fx|fx)*(32-fy|fy)*32
//      evaluated as ((top*(32-fy) + bot*fy + 512) >> 10), which is the same integer because nothing is rounded
//      in between; B and R share one 32-bit multiply (two 16-bit lanes);
//   4. the normalised RGB planes go straight to global memory (a warp writes 32 consecutive floats/halves of a
//      plane row), the rectified BGR bytes go through shared memory and leave as 16-byte row segments.
#include "common.h"

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

template <int DTYPE, bool IDENT, bool WRITE_TENSOR, bool ALIGNED4>
__global__ void __launch_bounds__(kRectThreads, 6) rectify_kernel(const RectifyArgs a, int tiles_x, int tiles_per_frame, int out_align)
{
    using T = typename TensorT<DTYPE>::type;
    __shared__ __align__(16) uint32_t s_src[kFootprintWords];
    __shared__ __align__(16) uint32_t s_out[kTileH * kTileW * 3 / 4];
    __shared__ T s_lut[IDENT ? 1 : 768];

    const int tid = threadIdx.x;
    int tile, fe;
    if (a.frame_fastest) { fe = blockIdx.x % (2 * a.n); tile = blockIdx.x / (2 * a.n); }
    else { tile = blockIdx.x % tiles_per_frame; fe = blockIdx.x / tiles_per_frame; }
    const int eye = fe & 1, frame = fe >> 1;
    const int W = a.W, H = a.H;
    const int tyi = tile / tiles_x;
    const int tx0 = (tile - tyi * tiles_x) * kTileW, ty0 = tyi * kTileH;

    const uint8_t* __restrict__ src = a.src[eye] + (size_t)frame * a.src_frame;
    const size_t pitch = a.src_pitch;
    const TileInfo ti = a.tiles[eye][tile];

    if (!IDENT) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = tid; i < 768; i += kRectThreads) s_lut[i] = lut[i];
    }

    // ---- 1. map entries of this thread's pixels -------------------------------------------------
    const int lx = tid & (kTileW - 1), ly0 = tid >> 6;
    const int x = tx0 + lx;
    int m1[4];
    uint32_t m2[4];
    const int* __restrict__ map1 = a.map1[eye];
    const uint16_t* __restrict__ map2 = a.map2[eye];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int y = ty0 + ly0 + 4 * k;
        m1[k] = 0; m2[k] = 0;
        if (x < W && y < H) {
            const size_t o = (size_t)y * W + x;
            m1[k] = __ldg(map1 + o);
            m2[k] = __ldg(map2 + o);
        }
    }

    // ---- 2. stage the source footprint --------------------------------------------------------
    const int fw = ti.w, X0 = ti.x0, Y0 = ti.y0;
    const bool staged = fw > 0;
    if (staged) {
        const int gw = fw >> 2, total = gw * ti.h;
        for (int g = tid; g < total; g += kRectThreads) {
            const int r = g / gw, c4 = (g - r * gw) << 2;
            const int yy = Y0 + r, xx = X0 + c4;
            uint4 o = make_uint4(0u, 0u, 0u, 0u);
            if ((unsigned)yy < (unsigned)H) {
                const uint8_t* rowp = src + (size_t)yy * pitch;
                if (ALIGNED4 && xx >= 0 && xx + 3 < W) {
                    const uint32_t* p = reinterpret_cast<const uint32_t*>(rowp + 3 * xx);
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
            *reinterpret_cast<uint4*>(&s_src[r * fw + c4]) = o;
        }
    }
    __syncthreads();

    // ---- 3. blend, normalise, store ------------------------------------------------------------
    uint8_t* s_out_b = reinterpret_cast<uint8_t*>(s_out);
    const size_t plane = (size_t)W * H;
    T* __restrict__ tbase = WRITE_TENSOR ? static_cast<T*>(a.tensor[eye]) + (size_t)frame * a.tensor_frame : nullptr;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int ly = ly0 + 4 * k, y = ty0 + ly;
        if (x < W && y < H) {
            int sx = (int)(short)(m1[k] & 0xffff), sy = m1[k] >> 16;
            const uint32_t fx = m2[k] & 31u, fy = (m2[k] >> 5) & 31u;
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
            const uint32_t ax = 32u - fx, ay = 32u - fy;
            // horizontal pass: B and R in the two 16-bit lanes of one word (<= 255*32 each), G alone
            const uint32_t tBR = (p00 & 0x00ff00ffu) * ax + (p01 & 0x00ff00ffu) * fx;
            const uint32_t bBR = (p10 & 0x00ff00ffu) * ax + (p11 & 0x00ff00ffu) * fx;
            const uint32_t tG = ((p00 >> 8) & 0xffu) * ax + ((p01 >> 8) & 0xffu) * fx;
            const uint32_t bG = ((p10 >> 8) & 0xffu) * ax + ((p11 >> 8) & 0xffu) * fx;
            // vertical pass + rounding: (sum*32 + 2^14) >> 15 == (sum + 512) >> 10
            const uint32_t vB = ((tBR & 0xffffu) * ay + (bBR & 0xffffu) * fy + 512u) >> 10;
            const uint32_t vR = ((tBR >> 16) * ay + (bBR >> 16) * fy + 512u) >> 10;
            const uint32_t vG = (tG * ay + bG * fy + 512u) >> 10;

            uint8_t* so = s_out_b + (ly * kTileW + lx) * 3;
            so[0] = (uint8_t)vB; so[1] = (uint8_t)vG; so[2] = (uint8_t)vR;

            if (WRITE_TENSOR) {
                T* t = tbase + (size_t)y * W + x;
                T r, g, b;
                if (IDENT) {
                    if (DTYPE == SA_F16) { r = (T)__ushort2half_rn((unsigned short)vR); g = (T)__ushort2half_rn((unsigned short)vG); b = (T)__ushort2half_rn((unsigned short)vB); }
                    else { r = (T)(float)vR; g = (T)(float)vG; b = (T)(float)vB; }
                } else {
                    r = s_lut[vR]; g = s_lut[256 + vG]; b = s_lut[512 + vB];
                }
                t[0] = r; t[plane] = g; t[2 * plane] = b;
            }
        }
    }
    __syncthreads();

    // ---- 4. rectified BGR tile -> global, widest stores the alignment allows ---------------------
    const int rows = min(kTileH, H - ty0), row_bytes = min(kTileW, W - tx0) * 3;
    uint8_t* g = a.rect[eye] + (size_t)frame * plane * 3 + ((size_t)ty0 * W + tx0) * 3;
    const size_t gpitch = (size_t)W * 3;
    if (out_align >= 16) copy_rows<uint4>(g, gpitch, s_out_b, rows, row_bytes);
    else if (out_align >= 4) copy_rows<uint32_t>(g, gpitch, s_out_b, rows, row_bytes);
    else if (out_align >= 2) copy_rows<uint16_t>(g, gpitch, s_out_b, rows, row_bytes);
    else copy_rows<uint8_t>(g, gpitch, s_out_b, rows, row_bytes);
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

inline bool lut_is_identity(int model) { return model == SA_MODEL_RAFTSTEREO || model == SA_MODEL_CRESTEREO; }

template <int DTYPE, bool IDENT, bool WT>
cudaError_t launch_rectify_t(const RectifyArgs& a, bool aligned4, int tiles_x, int tpf, int out_align, unsigned grid, cudaStream_t s)
{
    if (aligned4) rectify_kernel<DTYPE, IDENT, WT, true><<<grid, kRectThreads, 0, s>>>(a, tiles_x, tpf, out_align);
    else rectify_kernel<DTYPE, IDENT, WT, false><<<grid, kRectThreads, 0, s>>>(a, tiles_x, tpf, out_align);
    return cudaGetLastError();
}

}  // namespace

// largest power of two (capped at 256) dividing v
static int low_align(uintptr_t v) { if (v == 0) return 256; uintptr_t l = v & (~v + 1); return l > 256 ? 256 : (int)l; }
static int ptr_align(const void* p) { return low_align((uintptr_t)p); }
static int size_align(size_t v) { return low_align((uintptr_t)v); }
static int min3(int a, int b, int c) { return a < b ? (a < c ? a : c) : (b < c ? b : c); }

cudaError_t launch_rectify(const RectifyArgs& a, cudaStream_t s, long* launches)
{
    if (a.n <= 0) return cudaSuccess;
    const int tiles_x = (a.W + kTileW - 1) / kTileW, tiles_y = (a.H + kTileH - 1) / kTileH;
    const int tpf = tiles_x * tiles_y;
    const unsigned grid = (unsigned)tpf * 2u * (unsigned)a.n;
    // input alignment: every row of every frame of both eyes starts on a 4-byte boundary
    const bool aligned4 = min3(ptr_align(a.src[0]), ptr_align(a.src[1]), min3(size_align(a.src_pitch), size_align(a.src_frame), 256)) >= 4;
    // output alignment of the rectified rows (tile x offsets are multiples of 192 bytes)
    const int out_align = min3(ptr_align(a.rect[0]), ptr_align(a.rect[1]), min3(size_align((size_t)a.W * 3), size_align((size_t)a.W * a.H * 3), 256));
    const bool ident = lut_is_identity(a.model);
    cudaError_t e;
#define SA_RECT(DT, ID, WT) e = launch_rectify_t<DT, ID, WT>(a, aligned4, tiles_x, tpf, out_align, grid, s)
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
    dim3 grid((unsigned)((plane + 255) / 256), 2u * (unsigned)a.n);
    const bool ident = lut_is_identity(a.model);
    if (a.dtype == SA_F16) {
        if (ident) preprocess_kernel<SA_F16, true><<<grid, 256, 0, s>>>(a); else preprocess_kernel<SA_F16, false><<<grid, 256, 0, s>>>(a);
    } else {
        if (ident) preprocess_kernel<SA_F32, true><<<grid, 256, 0, s>>>(a); else preprocess_kernel<SA_F32, false><<<grid, 256, 0, s>>>(a);
    }
    if (launches) ++*launches;
    return cudaGetLastError();
}

}  // namespace sa
