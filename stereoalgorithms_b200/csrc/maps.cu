// Rectification maps on the device, once per calibration (the reference recomputes them on the CPU
// for every frame: FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:139-140), plus the per-tile
// source footprint table the fused rectify kernel stages from.
//
// init_maps_kernel reproduces cv::initUndistortRectifyMap(K, D, R, P, size, CV_16SC2) bit for bit:
// float64 throughout, every product and sum rounded separately (no FMA contraction - hence the
// explicit __dmul_rn/__dadd_rn), the same operation order as OpenCV's per-pixel loop, cvRound's
// x86 behaviour for values that do not fit an int (INT_MIN), saturating int16 packing.
#include "common.h"

namespace sa {

__device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dvd(double a, double b) { return __ddiv_rn(a, b); }

__device__ __forceinline__ int cv_round_i32(double v)
{
    double r = rint(v);  // round half to even
    if (!(r >= -2147483648.0 && r <= 2147483647.0)) return INT_MIN;
    return (int)r;
}

__global__ void __launch_bounds__(256) init_maps_kernel(MapParams p, int W, int H, int* __restrict__ map1, uint16_t* __restrict__ map2)
{
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    int i = blockIdx.y;
    if (j >= W || i >= H) return;
    const double dj = (double)j, di = (double)i;
    double X = add(add(mul(p.iR[0], dj), mul(p.iR[1], di)), p.iR[2]);
    double Y = add(add(mul(p.iR[3], dj), mul(p.iR[4], di)), p.iR[5]);
    double Wd = add(add(mul(p.iR[6], dj), mul(p.iR[7], di)), p.iR[8]);
    double w = dvd(1.0, Wd), x = mul(X, w), y = mul(Y, w);
    double x2 = mul(x, x), y2 = mul(y, y), r2 = add(x2, y2), _2xy = mul(mul(2.0, x), y);
    const double k1 = p.k[0], k2 = p.k[1], p1 = p.k[2], p2 = p.k[3], k3 = p.k[4], k4 = p.k[5], k5 = p.k[6], k6 = p.k[7];
    const double s1 = p.k[8], s2 = p.k[9], s3 = p.k[10], s4 = p.k[11];
    double num = add(1.0, mul(add(mul(add(mul(k3, r2), k2), r2), k1), r2));
    double den = add(1.0, mul(add(mul(add(mul(k6, r2), k5), r2), k4), r2));
    double kr = dvd(num, den);
    double xd = add(add(add(add(mul(x, kr), mul(p1, _2xy)), mul(p2, add(r2, mul(2.0, x2)))), mul(s1, r2)), mul(mul(s2, r2), r2));
    double yd = add(add(add(add(mul(y, kr), mul(p1, add(r2, mul(2.0, y2)))), mul(p2, _2xy)), mul(s3, r2)), mul(mul(s4, r2), r2));
    double u = add(mul(p.fx, xd), p.cx), v = add(mul(p.fy, yd), p.cy);
    int iu = cv_round_i32(mul(u, 32.0)), iv = cv_round_i32(mul(v, 32.0));
    int sx = max(-32768, min(32767, iu >> 5)), sy = max(-32768, min(32767, iv >> 5));
    size_t o = (size_t)i * W + j;
    map1[o] = (sx & 0xffff) | (sy << 16);
    map2[o] = (uint16_t)(((iv & 31) << 5) | (iu & 31));
}

cudaError_t launch_init_maps(const MapParams& p, int W, int H, int* map1, uint16_t* map2, cudaStream_t s)
{
    dim3 grid((W + 255) / 256, H);
    init_maps_kernel<<<grid, 256, 0, s>>>(p, W, H, map1, map2);
    return cudaGetLastError();
}

// One CTA per output tile: min/max of the (clamped) integer source coordinates of its pixels.
__global__ void __launch_bounds__(256) tile_info_kernel(const int* __restrict__ map1, const uint16_t* __restrict__ map2, int W, int H,
                                                        TileInfo* __restrict__ tiles, uint32_t* __restrict__ tmap, int* __restrict__ fp_max, int tiles_x,
                                                        int force_direct, int halo)
{
    __shared__ TileInfo s_t;
    const int tile = blockIdx.x;
    const int tx0 = (tile % tiles_x) * kTileW, ty0 = (tile / tiles_x) * kTileH;
    int xmin = 1 << 20, xmax = -(1 << 20), ymin = 1 << 20, ymax = -(1 << 20);
    // halo = 1 (handles that resize): the box also covers the taps of the pixel column right of the tile and the pixel row
    // below it, which the fused rectify + resize kernel re-blends for the resize taps that straddle a tile edge
    const int bw = kTileW + halo, bh = kTileH + halo;
    for (int k = threadIdx.x; k < bw * bh; k += blockDim.x) {
        int x = tx0 + (k % bw), y = ty0 + (k / bw);
        if (x < W && y < H) {
            int m = map1[(size_t)y * W + x];
            int sx = (int)(short)(m & 0xffff), sy = m >> 16;
            sx = max(-2, min(W, sx));
            sy = max(-2, min(H, sy));
            xmin = min(xmin, sx); xmax = max(xmax, sx);
            ymin = min(ymin, sy); ymax = max(ymax, sy);
        }
    }
    __shared__ int red[4][8];
    for (int o = 16; o; o >>= 1) {
        xmin = min(xmin, __shfl_xor_sync(0xffffffffu, xmin, o));
        xmax = max(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
        ymin = min(ymin, __shfl_xor_sync(0xffffffffu, ymin, o));
        ymax = max(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
    }
    int warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { red[0][warp] = xmin; red[1][warp] = xmax; red[2][warp] = ymin; red[3][warp] = ymax; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; k++) {
            xmin = min(xmin, red[0][k]); xmax = max(xmax, red[1][k]);
            ymin = min(ymin, red[2][k]); ymax = max(ymax, red[3][k]);
        }
        TileInfo t;
        // Footprint columns start and end on a multiple of 4 pixels (12 bytes = 3 words: the staging unit of the 4-byte-aligned
        // paths) - or of 16 pixels (48 bytes = three 16-byte loads, the unit of the warp-specialised kernel) when the image
        // width allows the tensor-copy path at all.  Wider by 12 pixels on average; rows are fetched in 32-byte
        // sectors anyway.
        const int al = (W % 2 == 0) ? 15 : 3;    // (even widths can take the tensor-copy path if their rows are padded to 16 bytes)
        int x0 = xmin & ~al;                     // floor (also for negatives)
        int w = ((xmax + 2 - x0) + al) & ~al;    // taps reach sx+1
        int h = ymax + 2 - ymin;
        t.x0 = (short)x0; t.y0 = (short)ymin; t.w = (short)w; t.h = (short)h;
        t.inv_gw = 0; t.flags = 0;
        if (force_direct || w * h > kFootprintWords || w <= 0 || h <= 0) { t.w = 0; t.h = 0; if (fp_max) atomicAdd(fp_max + 2, 1); }   // count of unstaged tiles
        else {
            const uint32_t gw = (uint32_t)w >> 2;
            t.inv_gw = ((1u << 20) + gw - 1) / gw;
            if (x0 >= 0 && ymin >= 0 && x0 + w <= W && ymin + h <= H) t.flags |= kTileInterior;
            if (fp_max) { atomicMax(fp_max, w); atomicMax(fp_max + 1, h); }   // box of the tensor copies: the largest staged footprint, partial tiles included
        }
        tiles[tile] = t;
        s_t = t;
    }
    __syncthreads();
    // derived per-pixel table: word offset of the top-left tap inside the staged footprint, and the two 5-bit weights
    const TileInfo t = s_t;
    for (int k = threadIdx.x; k < kTileW * kTileH; k += blockDim.x) {
        int x = tx0 + (k % kTileW), y = ty0 + (k / kTileW);
        if (x < W && y < H) {
            const size_t o = (size_t)y * W + x;
            uint32_t v = 0xffffffffu;
            if (t.w > 0) {
                int m = map1[o];
                int sx = max(-2, min(W, (int)(short)(m & 0xffff))), sy = max(-2, min(H, m >> 16));
                const uint32_t off = (uint32_t)((sy - t.y0) * t.w + (sx - t.x0));
                const uint32_t f = map2[o];
                v = off | ((f & 31u) << 12) | (((f >> 5) & 31u) << 17);
            }
            tmap[o] = v;
        }
    }
}

cudaError_t launch_tile_info(const int* map1, const uint16_t* map2, int W, int H, TileInfo* tiles, uint32_t* tmap, int* fp_max, int tiles_x, int tiles_y,
                             int force_direct, int halo, cudaStream_t s)
{
    tile_info_kernel<<<tiles_x * tiles_y, 256, 0, s>>>(map1, map2, W, H, tiles, tmap, fp_max, tiles_x, force_direct, halo ? 1 : 0);
    return cudaGetLastError();
}

}  // namespace sa
