// Stage B: disparity -> x y z r g b point cloud through Q, dense (reference behaviour) or compacted.
//
// Replaces <Model>_reprojectImageTo3D_kernel (FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:41-63,
// RAFTStereo/src/stereo_preprocess.cu:41-68 with its negate + write-back, CREStereo/..., HitNet/...), which
// runs one thread per pixel with six scalar stores at a 24-byte stride.
//
// Here one thread owns 4 consecutive pixels (one 16-byte disparity load, three 4-byte colour loads), a warp owns
// 128 pixels = 3072 contiguous output bytes.  The 24 floats of a thread go to a warp-private shared-memory
// region (16-byte slots, XOR-swizzled so neither the 96-byte-strided writes nor the consecutive reads conflict)
// and leave as six fully coalesced 16-byte stores per lane.  Arithmetic is the reference's, each product, sum
// and quotient rounded once (no FMA contraction, IEEE division):
//     w = Q[14]*d;  X_i = (Q[4i]*col + Q[4i+1]*row + Q[4i+3]) / w;  r,g,b = (float) of the rectified left BGR.
// The per-frame disparity min/max needed by the heatmap is reduced on the way (one atomic pair per warp).
//
// Compact mode (extension): keep pixels with finite disparity > 0, in row-major order.
// Three launches: per-block counts -> per-frame exclusive scan -> reproject with a warp-ballot style prefix
// (4-bit masks, shuffle scan) so that every warp writes one contiguous run of points.
#include <cfloat>

#include "common.h"

namespace sa {

namespace {

constexpr int kThreads = 256;
constexpr int kPxPerBlock = kThreads * 4;

__device__ __forceinline__ unsigned enc_ordered(float f)
{
    unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__device__ __forceinline__ bool valid_disp(float d) { return d > 0.0f && d <= FLT_MAX; }

struct Q10 { float q0, q1, q3, q4, q5, q7, q8, q9, q11, q14; };

__device__ __forceinline__ void point_xyz(const Q10& q, float d, int col, int row, float& X, float& Y, float& Z)
{
    const float fc = (float)col, fr = (float)row;
    const float w = __fmul_rn(q.q14, d);
    X = __fdiv_rn(__fadd_rn(__fadd_rn(__fmul_rn(q.q0, fc), __fmul_rn(q.q1, fr)), q.q3), w);
    Y = __fdiv_rn(__fadd_rn(__fadd_rn(__fmul_rn(q.q4, fc), __fmul_rn(q.q5, fr)), q.q7), w);
    Z = __fdiv_rn(__fadd_rn(__fadd_rn(__fmul_rn(q.q8, fc), __fmul_rn(q.q9, fr)), q.q11), w);
}

// Loads disparity + colour of the thread's 4 pixels.  VEC: everything 16/4-byte aligned and contiguous.
template <bool VEC>
__device__ __forceinline__ void load4(const ReprojectArgs& a, int frame, int p0, int npix, int row0, int col0, float d[4], uint32_t c[4])
{
    const float* disp = a.disp + (size_t)frame * a.disp_frame;
    const uint8_t* color = a.color + (size_t)frame * a.color_frame;
    if (VEC) {
        if (p0 < npix) {
            const float4 v = *reinterpret_cast<const float4*>(disp + p0);
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
            const uint32_t* cp = reinterpret_cast<const uint32_t*>(color + (size_t)p0 * 3);
            const uint32_t u0 = __ldg(cp), u1 = __ldg(cp + 1), u2 = __ldg(cp + 2);
            c[0] = u0; c[1] = __byte_perm(u0, u1, 0x4543); c[2] = __byte_perm(u1, u2, 0x4432); c[3] = u2 >> 8;
        } else {
            d[0] = d[1] = d[2] = d[3] = 0.f; c[0] = c[1] = c[2] = c[3] = 0u;
        }
    } else {
        int row = row0, col = col0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            d[k] = 0.f; c[k] = 0u;
            if (p0 + k < npix) {
                d[k] = disp[p0 + k];
                const uint8_t* v = color + (size_t)row * a.color_pitch + 3 * col;
                c[k] = (uint32_t)v[0] | ((uint32_t)v[1] << 8) | ((uint32_t)v[2] << 16);
            }
            if (++col == a.w) { col = 0; ++row; }
        }
    }
}

template <bool VEC>
__device__ __forceinline__ void store_disp4(const ReprojectArgs& a, int frame, int p0, int npix, const float d[4])
{
    float* disp = a.disp + (size_t)frame * a.disp_frame;
    if (VEC) { if (p0 < npix) *reinterpret_cast<float4*>(disp + p0) = make_float4(d[0], d[1], d[2], d[3]); }
    else {
#pragma unroll
        for (int k = 0; k < 4; k++) if (p0 + k < npix) disp[p0 + k] = d[k];
    }
}

__device__ __forceinline__ void warp_minmax(unsigned* mm, float mn, float mx)
{
    for (int o = 16; o; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) {
        if (mn <= FLT_MAX || mn > FLT_MAX) {  // not NaN
            atomicMin(mm, enc_ordered(mn));
            atomicMin(mm + 1, ~enc_ordered(mx));
        }
    }
}

template <bool VEC>
__global__ void __launch_bounds__(kThreads) reproject_dense_kernel(const ReprojectArgs a)
{
    __shared__ __align__(16) float4 s_pts[kThreads / 32][192];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int frame = blockIdx.y, npix = a.w * a.h;
    const int p0 = (blockIdx.x * kThreads + tid) * 4;
    const Q10 q = {a.q[0], a.q[1], a.q[3], a.q[4], a.q[5], a.q[7], a.q[8], a.q[9], a.q[11], a.q[14]};
    int row = p0 / a.w, col = p0 - row * a.w;

    float d[4];
    uint32_t c[4];
    load4<VEC>(a, frame, p0, npix, row, col, d, c);
    if (a.negate) {
#pragma unroll
        for (int k = 0; k < 4; k++) d[k] = -d[k];
        store_disp4<VEC>(a, frame, p0, npix, d);
    }

    float f[24];
    float mn = INFINITY, mx = -INFINITY;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        point_xyz(q, d[k], col, row, f[6 * k], f[6 * k + 1], f[6 * k + 2]);
        f[6 * k + 3] = (float)((c[k] >> 16) & 0xffu);
        f[6 * k + 4] = (float)((c[k] >> 8) & 0xffu);
        f[6 * k + 5] = (float)(c[k] & 0xffu);
        if (p0 + k < npix) { mn = fminf(mn, d[k]); mx = fmaxf(mx, d[k]); }
        if (++col == a.w) { col = 0; ++row; }
    }
    if (a.minmax) warp_minmax(a.minmax + 2 * frame, mn, mx);

    float4* sw = s_pts[warp];
    const int flip = (lane >> 2) & 1;
#pragma unroll
    for (int i = 0; i < 6; i++) sw[(lane * 6 + i) ^ flip] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
    __syncwarp();

    float* out = a.points + (size_t)frame * npix * 6;
    const size_t warp_px0 = (size_t)(blockIdx.x * kThreads + warp * 32) * 4;
    if (VEC) {
        float4* out4 = reinterpret_cast<float4*>(out) + warp_px0 * 6 / 4;
        const size_t lim = ((size_t)npix * 6 / 4 > warp_px0 * 6 / 4) ? (size_t)npix * 6 / 4 - warp_px0 * 6 / 4 : 0;
#pragma unroll
        for (int m = 0; m < 6; m++) {
            const int s = lane + 32 * m;
            if ((size_t)s < lim) out4[s] = sw[s ^ ((s / 24) & 1)];
        }
    } else {
        const float* sf = reinterpret_cast<const float*>(sw);
        float* o = out + warp_px0 * 6;
        const size_t lim = ((size_t)npix * 6 > warp_px0 * 6) ? (size_t)npix * 6 - warp_px0 * 6 : 0;
#pragma unroll 4
        for (int m = 0; m < 24; m++) {
            const int idx = lane + 32 * m, s = idx >> 2;
            if ((size_t)idx < lim) o[idx] = sf[((s ^ ((s / 24) & 1)) << 2) | (idx & 3)];
        }
    }
}

// ---- compaction ------------------------------------------------------------------------------------------
template <bool VEC>
__global__ void __launch_bounds__(kThreads) count_valid_kernel(const ReprojectArgs a, int nblk)
{
    __shared__ int s_w[kThreads / 32];
    const int tid = threadIdx.x, frame = blockIdx.y, npix = a.w * a.h;
    const int p0 = (blockIdx.x * kThreads + tid) * 4;
    const float* disp = a.disp + (size_t)frame * a.disp_frame;
    int cnt = 0;
    if (VEC) {
        if (p0 < npix) {
            float4 v = *reinterpret_cast<const float4*>(disp + p0);
            if (a.negate) { v.x = -v.x; v.y = -v.y; v.z = -v.z; v.w = -v.w; }
            cnt = (int)valid_disp(v.x) + (int)valid_disp(v.y) + (int)valid_disp(v.z) + (int)valid_disp(v.w);
        }
    } else {
        for (int k = 0; k < 4; k++)
            if (p0 + k < npix) { float v = disp[p0 + k]; if (a.negate) v = -v; cnt += (int)valid_disp(v); }
    }
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((tid & 31) == 0) s_w[tid >> 5] = cnt;
    __syncthreads();
    if (tid == 0) {
        int t = 0;
        for (int i = 0; i < kThreads / 32; i++) t += s_w[i];
        a.blk_counts[(size_t)frame * nblk + blockIdx.x] = t;
    }
}

// one CTA per frame: exclusive scan of its block counts, total -> counts[frame]
__global__ void __launch_bounds__(kThreads) scan_blocks_kernel(const int* __restrict__ blk_counts, int* __restrict__ blk_offsets, int* __restrict__ counts,
                                                               int nblk)
{
    __shared__ int s_part[kThreads];
    const int tid = threadIdx.x, frame = blockIdx.x;
    const int* in = blk_counts + (size_t)frame * nblk;
    int* out = blk_offsets + (size_t)frame * nblk;
    const int per = (nblk + kThreads - 1) / kThreads;
    const int b0 = tid * per, b1 = min(nblk, b0 + per);
    int sum = 0;
    for (int i = b0; i < b1; i++) sum += in[i];
    s_part[tid] = sum;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int i = 0; i < kThreads; i++) { int t = s_part[i]; s_part[i] = run; run += t; }
        counts[frame] = run;
    }
    __syncthreads();
    int run = s_part[tid];
    for (int i = b0; i < b1; i++) { out[i] = run; run += in[i]; }
}

template <bool VEC>
__global__ void __launch_bounds__(kThreads) reproject_compact_kernel(const ReprojectArgs a, int nblk)
{
    __shared__ __align__(16) float2 s_pts[kThreads / 32][128 * 3];
    __shared__ int s_wtot[kThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int frame = blockIdx.y, npix = a.w * a.h;
    const int p0 = (blockIdx.x * kThreads + tid) * 4;
    const Q10 q = {a.q[0], a.q[1], a.q[3], a.q[4], a.q[5], a.q[7], a.q[8], a.q[9], a.q[11], a.q[14]};
    int row = p0 / a.w, col = p0 - row * a.w;

    float d[4];
    uint32_t c[4];
    load4<VEC>(a, frame, p0, npix, row, col, d, c);
    if (a.negate) {
#pragma unroll
        for (int k = 0; k < 4; k++) d[k] = -d[k];
        store_disp4<VEC>(a, frame, p0, npix, d);
    }
    unsigned mask = 0;
    float mn = INFINITY, mx = -INFINITY;
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (p0 + k < npix) {
            mn = fminf(mn, d[k]); mx = fmaxf(mx, d[k]);
            if (valid_disp(d[k])) mask |= 1u << k;
        }
    if (a.minmax) warp_minmax(a.minmax + 2 * frame, mn, mx);

    const int cnt = __popc(mask);
    int incl = cnt;
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    const int wtotal = __shfl_sync(0xffffffffu, incl, 31);
    if (lane == 0) s_wtot[warp] = wtotal;
    int slot = incl - cnt;

    float2* sw = s_pts[warp];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (mask & (1u << k)) {
            float X, Y, Z;
            point_xyz(q, d[k], col, row, X, Y, Z);
            sw[slot * 3 + 0] = make_float2(X, Y);
            sw[slot * 3 + 1] = make_float2(Z, (float)((c[k] >> 16) & 0xffu));
            sw[slot * 3 + 2] = make_float2((float)((c[k] >> 8) & 0xffu), (float)(c[k] & 0xffu));
            ++slot;
        }
        if (++col == a.w) { col = 0; ++row; }
    }
    __syncthreads();
    int base = a.blk_offsets[(size_t)frame * nblk + blockIdx.x];
    for (int i = 0; i < warp; i++) base += s_wtot[i];
    float2* out = reinterpret_cast<float2*>(a.points + (size_t)frame * npix * 6) + (size_t)base * 3;
    for (int i = lane; i < wtotal * 3; i += 32) out[i] = sw[i];
}

__global__ void fill_int_kernel(int* p, int n, int v)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

}  // namespace

int reproject_blocks_per_frame(int w, int h) { return (int)(((size_t)w * h + kPxPerBlock - 1) / kPxPerBlock); }

cudaError_t launch_reproject(const ReprojectArgs& a, cudaStream_t s, long* launches)
{
    if (a.n <= 0) return cudaSuccess;
    const size_t npix = (size_t)a.w * a.h;
    const int nblk = reproject_blocks_per_frame(a.w, a.h);
    const bool vec = (npix % 4 == 0) && (a.disp_frame % 4 == 0) && (((uintptr_t)a.disp) % 16 == 0) && (a.color_pitch == (size_t)a.w * 3) &&
                     (a.color_frame % 4 == 0) && (((uintptr_t)a.color) % 4 == 0) && (((uintptr_t)a.points) % 16 == 0);
    dim3 grid((unsigned)nblk, (unsigned)a.n);
    if (!a.compact) {
        if (vec) reproject_dense_kernel<true><<<grid, kThreads, 0, s>>>(a); else reproject_dense_kernel<false><<<grid, kThreads, 0, s>>>(a);
        if (launches) ++*launches;
        if (a.counts) { fill_int_kernel<<<(a.n + 255) / 256, 256, 0, s>>>(a.counts, a.n, (int)npix); if (launches) ++*launches; }
        return cudaGetLastError();
    }
    if (vec) count_valid_kernel<true><<<grid, kThreads, 0, s>>>(a, nblk); else count_valid_kernel<false><<<grid, kThreads, 0, s>>>(a, nblk);
    scan_blocks_kernel<<<a.n, kThreads, 0, s>>>(a.blk_counts, a.blk_offsets, a.counts, nblk);
    if (vec) reproject_compact_kernel<true><<<grid, kThreads, 0, s>>>(a, nblk); else reproject_compact_kernel<false><<<grid, kThreads, 0, s>>>(a, nblk);
    if (launches) *launches += 3;
    return cudaGetLastError();
}

}  // namespace sa
