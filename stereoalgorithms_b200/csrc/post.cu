// Stage C (heatmap) and the small kernels around the network slot: per-frame disparity min/max, JET heatmap,
// synthetic-disparity stub, plane-0 extraction, and the u8 INTER_LINEAR resize of the optional resize path.
#include <cfloat>

#include "common.h"

namespace sa {

namespace {

// in global memory, not __constant__: each CTA copies it to shared memory with one coalesced load per thread (a
// constant-bank read indexed by threadIdx serialises 32 ways)
__device__ const uint32_t c_jet[256] = {
#include "jet_lut.inc"
};

__device__ __forceinline__ unsigned enc_ordered(float f)
{
    unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float dec_ordered(unsigned e)
{
    return __uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e);
}

// ---- min/max (standalone; normally fused into the reproject kernels) ---------------------------------------
__global__ void __launch_bounds__(256) minmax_kernel(const float* __restrict__ disp, size_t disp_frame, int npix, unsigned* __restrict__ minmax)
{
    const int frame = blockIdx.y;
    const float* d = disp + (size_t)frame * disp_frame;
    float mn = INFINITY, mx = -INFINITY;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
        const float v = d[i];
        mn = fminf(mn, v); mx = fmaxf(mx, v);
    }
    __shared__ unsigned s_mm[16];
    const unsigned a = __reduce_min_sync(0xffffffffu, enc_ordered(mn)), b = __reduce_min_sync(0xffffffffu, ~enc_ordered(mx));
    if ((threadIdx.x & 31) == 0) { s_mm[2 * (threadIdx.x >> 5)] = a; s_mm[2 * (threadIdx.x >> 5) + 1] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {      // one atomic pair per CTA (a few dozen per frame)
        unsigned x = s_mm[0], y = s_mm[1];
        for (int i = 1; i < 8; i++) { x = min(x, s_mm[2 * i]); y = min(y, s_mm[2 * i + 1]); }
        atomicMin(minmax + 2 * frame, x);
        atomicMin(minmax + 2 * frame + 1, y);
    }
}

// ---- heatmap: heatmap() of FastACVNet_plus/test/main.cpp:7-24 -------------------------------------------------
//   s = f32((max-min)/255)  [double arithmetic];  n = f32(f32(d - f32(min)) / s);  idx = sat_u8(round_half_even(|n|))
// One CTA colours kHeatIters x 1024 pixels: the JET table is staged once per CTA, and all of a thread's disparity loads are
// issued before the first one is used.
constexpr int kHeatIters = 4;

// FOLD: `minmax` holds the per-CTA partials stage B left behind (nblk pairs per frame) instead of the frame's final pair; every CTA
// folds them itself while its disparity loads are in flight (a few hundred 8-byte words out of L2), which saves the separate
// one-CTA-per-frame reduction launch in front of this kernel; the first CTA of a frame also stores the final pair to final_mm.
template <bool VEC, bool FOLD>
__global__ void __launch_bounds__(256) heatmap_kernel(const float* __restrict__ disp, size_t disp_frame, int npix, const unsigned* __restrict__ minmax, int nblk,
                                                      unsigned* __restrict__ final_mm, uint8_t* __restrict__ out)
{
    __shared__ uint32_t s_jet[256];
    __shared__ unsigned s_mm[16];
    __shared__ float s_par[3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    s_jet[tid] = c_jet[tid];
    const int frame = blockIdx.y;
    const float* d = disp + (size_t)frame * disp_frame;
    uint8_t* o = out + (size_t)frame * npix * 3;
    float v[kHeatIters][4];
#pragma unroll
    for (int it = 0; it < kHeatIters; it++) {
        const int p0 = ((blockIdx.x * kHeatIters + it) * 256 + tid) * 4;
        v[it][0] = v[it][1] = v[it][2] = v[it][3] = 0.f;
        if (VEC) { if (p0 < npix) { const float4 t = *reinterpret_cast<const float4*>(d + p0); v[it][0] = t.x; v[it][1] = t.y; v[it][2] = t.z; v[it][3] = t.w; } }
        else { for (int k = 0; k < 4; k++) if (p0 + k < npix) v[it][k] = d[p0 + k]; }
    }
    // frame constants: warp 0 alone finishes the min/max fold and does the double-precision division (a CTA colours only 4096 pixels:
    // done by every thread, this preamble was a third of all the instructions the kernel executed), the others pick them up
    if (FOLD) {
        const uint2* in = reinterpret_cast<const uint2*>(minmax) + (size_t)frame * nblk;
        unsigned a = 0xffffffffu, b = 0xffffffffu;
#pragma unroll 1
        for (int i = tid; i < nblk; i += 256) { const uint2 t = __ldg(in + i); a = min(a, t.x); b = min(b, t.y); }
        a = __reduce_min_sync(0xffffffffu, a);
        b = __reduce_min_sync(0xffffffffu, b);
        if (lane == 0) { s_mm[2 * warp] = a; s_mm[2 * warp + 1] = b; }
        __syncthreads();
    }
    if (warp == 0) {
        unsigned emn, emx;                                           // (min, ~max) in the order-preserving encoding
        if (FOLD) {
            emn = __reduce_min_sync(0xffffffffu, s_mm[2 * (lane & 7)]);
            emx = __reduce_min_sync(0xffffffffu, s_mm[2 * (lane & 7) + 1]);
            if (blockIdx.x == 0 && lane == 0) { final_mm[2 * frame] = emn; final_mm[2 * frame + 1] = emx; }
        } else {
            emn = minmax[2 * frame]; emx = minmax[2 * frame + 1];
        }
        const float mn0 = dec_ordered(emn), mx0 = dec_ordered(~emx);
        const float s0 = (float)(((double)mx0 - (double)mn0) / 255.0);
        if (lane == 0) { s_par[0] = mn0; s_par[1] = s0; s_par[2] = 1.0f / s0; }   // 1/s = inf for a constant frame: every pixel then takes the exact path
    }
    __syncthreads();                                                 // (also: the colour table is staged)
    const float mn = s_par[0], s = s_par[1], rs = s_par[2];
#pragma unroll
    for (int it = 0; it < kHeatIters; it++) {
        const int chunk = blockIdx.x * kHeatIters + it;
        const int p0 = (chunk * 256 + tid) * 4;
        // idx = sat_u8(round_half_even(|fl(x / s)|)), x = fl(d - min).  The IEEE division is a dozen instructions per pixel; the product
        // with the reciprocal is within 2 ulp of it (< 3.1e-5 below 256).  Adding 512 + 1/2 + 2/16384 rounds that product to 1/16384ths
        // and leaves them in the sum's mantissa: bits 14.. = floor(q + 1/2), bits 0..13 = how far q lies above the half-integer (+2).
        // Low bits 0..4 = within 2/16384 = 1.2e-4 of a half-integer: only there could the product and the quotient round to different
        // integers.  Those pixels (5 in 16384), NaN, inf, s = 0 and anything >= 500 take the division; everything else that is
        // >= 255.5 saturates either way.  One test per thread (four pixels), not per pixel.
        uint32_t off[4];
        bool ok = true;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const float x = __fsub_rn(v[it][k], mn);
            const float q = fabsf(__fmul_rn(x, rs));
            const uint32_t b = __float_as_uint(__fadd_rn(q, 512.5001220703125f));
            ok = ok && (q < 500.0f) && ((b & 16383u) >= 5u);
            off[k] = min((b >> 12) & 0x7fcu, 1020u);                    // byte offset into the colour table
        }
        if (!ok) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                int idx = __float2int_rn(fabsf(__fdiv_rn(__fsub_rn(v[it][k], mn), s)));     // NaN -> 0, like saturate_cast<uchar>(cvRound(NaN) = INT_MIN)
                off[k] = 4u * (uint32_t)min(255, max(0, idx));
            }
        }
        uint32_t px[4];
#pragma unroll
        for (int k = 0; k < 4; k++) px[k] = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(s_jet) + off[k]);
        if (VEC) {
            // 12 bytes per lane as three 4-byte stores, 12 bytes apart from lane to lane: the three of them fill every sector they
            // touch, and L2 merges them before anything goes to DRAM.  (A warp-private transpose through shared memory for unit-stride
            // stores costs more instructions than it saves L1 passes: this kernel is issue-bound, not L1-bound.)
            if (p0 < npix) {
                uint32_t* ow = reinterpret_cast<uint32_t*>(o) + (size_t)(p0 >> 2) * 3;
                ow[0] = __byte_perm(px[0], px[1], 0x4210);
                ow[1] = __byte_perm(px[1], px[2], 0x5421);
                ow[2] = __byte_perm(px[2], px[3], 0x6542);
            }
        } else {
            for (int k = 0; k < 4; k++)
                if (p0 + k < npix) {
                    uint8_t* q = o + (size_t)(p0 + k) * 3;
                    q[0] = (uint8_t)px[k]; q[1] = (uint8_t)(px[k] >> 8); q[2] = (uint8_t)(px[k] >> 16);
                }
        }
    }
}

// ---- synthetic disparity (stand-in for the TensorRT enqueue) ------------------------------------------------
__device__ __forceinline__ uint32_t hash32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

__global__ void __launch_bounds__(256) stub_kernel(float* __restrict__ disp, size_t disp_frame, int planes, int w, int h, int first_frame,
                                                   unsigned seed, int negate)
{
    const int frame = blockIdx.y;
    const size_t npix = (size_t)w * h;
    const size_t p = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= npix) return;
    const int r = (int)(p / w), c = (int)(p - (size_t)r * w);
    const int gframe = first_frame + frame;
    float v = 4.0f + 60.0f * (0.5f + 0.5f * __sinf(6.2831853f * 3.0f * c / w + 0.37f * gframe)) * (0.5f + 0.5f * __cosf(6.2831853f * 2.0f * r / h));
    const uint32_t u = hash32((uint32_t)p * 2654435761u ^ hash32(seed + 0x9e3779b9u * (uint32_t)gframe));
    const uint32_t pm = u % 1000u;
    if (pm < 50u) v = 0.0f;
    else if (pm < 60u) v = -1.0f;
    if (negate) v = -v;
    float* d = disp + (size_t)frame * disp_frame;
    d[p] = v;
    for (int k = 1; k < planes; k++) d[(size_t)k * npix + p] = 0.0f;
}

__global__ void __launch_bounds__(256) copy_plane0_kernel(const float* __restrict__ disp, size_t disp_frame, size_t npix, float* __restrict__ out)
{
    const int frame = blockIdx.y;
    const size_t p = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (p < npix) out[(size_t)frame * npix + p] = disp[(size_t)frame * disp_frame + p];
}

// ---- disparity as an 8-bit image: what cv::imwrite("disparity.jpg", disparity) encodes (FastACVNet_plus/test/main.cpp:55) ----
// imwrite converts a CV_32FC1 Mat with convertTo(CV_8U): saturate_cast<uchar>(cvRound(d)).  cvRound is cvtss2si: round half
// to even, and 0x80000000 for NaN and for |d| >= 2^31 - which saturates to 0, not 255 (probed with cv2 4.13: +inf -> 0).
__device__ __forceinline__ uint32_t disparity_to_u8(float d)
{
    if (!(fabsf(d) < 2147483648.0f)) return 0u;
    return (uint32_t)min(255, max(0, __float2int_rn(d)));
}

template <bool VEC>
__global__ void __launch_bounds__(256) disparity_u8_kernel(const float* __restrict__ disp, size_t disp_frame, size_t npix, uint8_t* __restrict__ out)
{
    const int frame = blockIdx.y;
    const float* d = disp + (size_t)frame * disp_frame;
    uint8_t* o = out + (size_t)frame * npix;
    const size_t p = ((size_t)blockIdx.x * 256 + threadIdx.x) * 4;
    if (VEC) {
        if (p < npix) {
            const float4 t = *reinterpret_cast<const float4*>(d + p);
            *reinterpret_cast<uint32_t*>(o + p) = disparity_to_u8(t.x) | (disparity_to_u8(t.y) << 8) | (disparity_to_u8(t.z) << 16) | (disparity_to_u8(t.w) << 24);
        }
    } else {
        for (int k = 0; k < 4; k++)
            if (p + k < npix) o[p + k] = (uint8_t)disparity_to_u8(d[p + k]);
    }
}

// ---- u8 resize (extension a8): OpenCV's fixed-point INTER_LINEAR for downscales ----------------------------
__global__ void __launch_bounds__(256) resize_kernel(const ResizeArgs a)
{
    const int eye = a.one_eye ? 0 : (blockIdx.y & 1), frame = a.one_eye ? blockIdx.y : (blockIdx.y >> 1);
    const size_t npix = (size_t)a.w * a.h;
    const size_t p = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= npix) return;
    const int y = (int)(p / a.w), x = (int)(p - (size_t)y * a.w);
    const uint8_t* src = (eye ? a.src[1] : a.src[0]) + (size_t)frame * a.src_frame;
    uint8_t* dst = (eye ? a.dst[1] : a.dst[0]) + ((size_t)frame * npix + p) * 3;
    if (a.area2x) {  // exact 2x decimation is executed by OpenCV as INTER_AREA: rounded mean of the 2x2 block
        const uint8_t* r0 = src + (size_t)(2 * y) * a.src_pitch + 6 * x;
        const uint8_t* r1 = r0 + a.src_pitch;
#pragma unroll
        for (int c = 0; c < 3; c++) dst[c] = (uint8_t)((r0[c] + r0[3 + c] + r1[c] + r1[3 + c] + 2) >> 2);
        return;
    }
    const int x0 = a.tab_x[x], x1 = a.tab_x[a.w + x], a0 = a.tab_x[2 * a.w + x], a1 = a.tab_x[3 * a.w + x];
    const int y0 = a.tab_y[y], y1 = a.tab_y[a.h + y], b0 = a.tab_y[2 * a.h + y], b1 = a.tab_y[3 * a.h + y];
    const uint8_t* r0 = src + (size_t)y0 * a.src_pitch;
    const uint8_t* r1 = src + (size_t)y1 * a.src_pitch;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const int T0 = r0[3 * x0 + c] * a0 + r0[3 * x1 + c] * a1;
        const int T1 = r1[3 * x0 + c] * a0 + r1[3 * x1 + c] * a1;
        const int v = (((b0 * (T0 >> 4)) >> 16) + ((b1 * (T1 >> 4)) >> 16) + 2) >> 2;
        dst[c] = (uint8_t)min(255, max(0, v));
    }
}

// The same gather with the network normalisation fused in: the resized value goes through the look-up table (or the identity)
// straight into the planes; the resized u8 image is written for the LEFT eye only (stage B's colours).  Replaces resize_kernel
// + preprocess4_kernel on the default resize path: the resized images are not written and read back, one launch less.
// (A tiled flavour - 64 x 16 network pixels per CTA, the source box staged in shared memory with coalesced 16-byte loads - measured
// 0.39 ms against 0.33 ms for this gather on 64 pairs 1080p -> 640x480, like round 1's cp.async attempt: L1 serves the gather well.)
template <typename T, bool IDENT>
__global__ void __launch_bounds__(256) resize_norm_kernel(const ResizeArgs a)
{
    __shared__ T s_lut[IDENT ? 1 : 768];
    if (!IDENT) {
        const T* lut = static_cast<const T*>(a.lut);
        for (int i = threadIdx.x; i < 768; i += 256) s_lut[i] = lut[i];
        __syncthreads();
    }
    const int eye = blockIdx.y & 1, frame = blockIdx.y >> 1;
    const size_t npix = (size_t)a.w * a.h;
    const size_t p = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= npix) return;
    const int y = (int)(p / a.w), x = (int)(p - (size_t)y * a.w);
    const uint8_t* src = (eye ? a.src[1] : a.src[0]) + (size_t)frame * a.src_frame;
    int v[3];
    if (a.area2x) {
        const uint8_t* r0 = src + (size_t)(2 * y) * a.src_pitch + 6 * x;
        const uint8_t* r1 = r0 + a.src_pitch;
#pragma unroll
        for (int c = 0; c < 3; c++) v[c] = (r0[c] + r0[3 + c] + r1[c] + r1[3 + c] + 2) >> 2;
    } else {
        const int x0 = __ldg(a.tab_x + x), x1 = __ldg(a.tab_x + a.w + x), a0 = __ldg(a.tab_x + 2 * a.w + x), a1 = __ldg(a.tab_x + 3 * a.w + x);
        const int y0 = __ldg(a.tab_y + y), y1 = __ldg(a.tab_y + a.h + y), b0 = __ldg(a.tab_y + 2 * a.h + y), b1 = __ldg(a.tab_y + 3 * a.h + y);
        const uint8_t* r0 = src + (size_t)y0 * a.src_pitch;
        const uint8_t* r1 = src + (size_t)y1 * a.src_pitch;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const int T0 = r0[3 * x0 + c] * a0 + r0[3 * x1 + c] * a1;
            const int T1 = r1[3 * x0 + c] * a0 + r1[3 * x1 + c] * a1;
            v[c] = min(255, max(0, (((b0 * (T0 >> 4)) >> 16) + ((b1 * (T1 >> 4)) >> 16) + 2) >> 2));
        }
    }
    T* t = static_cast<T*>(eye ? a.tensor[1] : a.tensor[0]) + (size_t)frame * a.tensor_frame + p;
    if (IDENT) { t[0] = (T)(float)v[2]; t[npix] = (T)(float)v[1]; t[2 * npix] = (T)(float)v[0]; }
    else { t[0] = s_lut[v[2]]; t[npix] = s_lut[256 + v[1]]; t[2 * npix] = s_lut[512 + v[0]]; }
    if (eye == 0) {
        uint8_t* dst = a.dst[0] + ((size_t)frame * npix + p) * 3;
        dst[0] = (uint8_t)v[0]; dst[1] = (uint8_t)v[1]; dst[2] = (uint8_t)v[2];
    }
}

}  // namespace

cudaError_t launch_minmax(const float* disp, size_t disp_frame, int npix, int n, unsigned* minmax, cudaStream_t s, long* launches)
{
    if (n <= 0) return cudaSuccess;
    int bx = (npix + 256 * 8 - 1) / (256 * 8);
    if (bx < 1) bx = 1;
    minmax_kernel<<<dim3((unsigned)bx, (unsigned)n), 256, 0, s>>>(disp, disp_frame, npix, minmax);
    if (launches) ++*launches;
    return cudaGetLastError();
}

// blk_minmax != nullptr: fold stage B's per-CTA partials (nblk pairs per frame) inside the kernel and leave the result in minmax
cudaError_t launch_heatmap(const float* disp, size_t disp_frame, int npix, int n, unsigned* minmax, const unsigned* blk_minmax, int nblk, uint8_t* out,
                           cudaStream_t s, long* launches)
{
    if (n <= 0) return cudaSuccess;
    const bool vec = (npix % 4 == 0) && (disp_frame % 4 == 0) && (((uintptr_t)disp) % 16 == 0) && (((uintptr_t)out) % 4 == 0);
    dim3 grid((unsigned)((npix + 1024 * kHeatIters - 1) / (1024 * kHeatIters)), (unsigned)n);
    if (blk_minmax) {
        if (vec) heatmap_kernel<true, true><<<grid, 256, 0, s>>>(disp, disp_frame, npix, blk_minmax, nblk, minmax, out);
        else heatmap_kernel<false, true><<<grid, 256, 0, s>>>(disp, disp_frame, npix, blk_minmax, nblk, minmax, out);
    } else if (vec) heatmap_kernel<true, false><<<grid, 256, 0, s>>>(disp, disp_frame, npix, minmax, 0, nullptr, out);
    else heatmap_kernel<false, false><<<grid, 256, 0, s>>>(disp, disp_frame, npix, minmax, 0, nullptr, out);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_stub(float* disp, size_t disp_frame, int planes, int w, int h, int first_frame, int n, unsigned seed, int negate, cudaStream_t s,
                        long* launches)
{
    if (n <= 0) return cudaSuccess;
    const size_t npix = (size_t)w * h;
    stub_kernel<<<dim3((unsigned)((npix + 255) / 256), (unsigned)n), 256, 0, s>>>(disp, disp_frame, planes, w, h, first_frame, seed, negate);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_copy_plane0(const float* disp, size_t disp_frame, size_t npix, int n, float* out, cudaStream_t s, long* launches)
{
    if (n <= 0) return cudaSuccess;
    copy_plane0_kernel<<<dim3((unsigned)((npix + 255) / 256), (unsigned)n), 256, 0, s>>>(disp, disp_frame, npix, out);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_disparity_u8(const float* disp, size_t disp_frame, size_t npix, int n, uint8_t* out, cudaStream_t s, long* launches)
{
    if (n <= 0) return cudaSuccess;
    const bool vec = npix % 4 == 0 && disp_frame % 4 == 0 && ((uintptr_t)disp) % 16 == 0 && ((uintptr_t)out) % 4 == 0;
    const dim3 grid((unsigned)((npix + 1023) / 1024), (unsigned)n);
    if (vec) disparity_u8_kernel<true><<<grid, 256, 0, s>>>(disp, disp_frame, npix, out);
    else disparity_u8_kernel<false><<<grid, 256, 0, s>>>(disp, disp_frame, npix, out);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_resize(const ResizeArgs& a, cudaStream_t s, long* launches)
{
    if (a.n <= 0) return cudaSuccess;
    const size_t npix = (size_t)a.w * a.h;
    if (a.tensor[0] && a.tensor[1]) {
        const dim3 grid((unsigned)((npix + 255) / 256), 2u * (unsigned)a.n);
        if (a.dtype == SA_F16) { if (a.ident) resize_norm_kernel<__half, true><<<grid, 256, 0, s>>>(a); else resize_norm_kernel<__half, false><<<grid, 256, 0, s>>>(a); }
        else { if (a.ident) resize_norm_kernel<float, true><<<grid, 256, 0, s>>>(a); else resize_norm_kernel<float, false><<<grid, 256, 0, s>>>(a); }
        if (launches) ++*launches;
        return cudaGetLastError();
    }
    resize_kernel<<<dim3((unsigned)((npix + 255) / 256), (a.one_eye ? 1u : 2u) * (unsigned)a.n), 256, 0, s>>>(a);
    if (launches) ++*launches;
    return cudaGetLastError();
}

}  // namespace sa
