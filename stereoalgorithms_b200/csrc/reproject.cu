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
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdarg>

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

__device__ __forceinline__ float rcp_approx(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// The three quotients of one pixel share one reciprocal.
//   FAST: r = rcp(w) refined by one Newton step, then q = n*r corrected with the exact remainder (two FMAs): the result
//   is within 1 ulp of the IEEE quotient (the parity gate for XYZ is 1e-5 relative), without the ~60-instruction slow
//   path that a division by zero takes - and ~5 % of real disparities are zero.  w = 0, inf, NaN: n * rcp(w) is IEEE's
//   answer in kind and sign (n/0 = +-inf, 0/0 = NaN, n/inf = 0).  |w| outside [2^-60, 2^60] otherwise: exact division.
//   The host only selects FAST when every Q entry used is finite and small enough that the numerators stay below 2^36,
//   so inside that window a quotient is at most 2^96: n * r1 cannot overflow to inf (and then turn the correction step
//   into inf - inf = NaN where IEEE division answers +-inf), and r1 is never a flushed denormal.
//   !FAST: three IEEE divisions, each product and sum rounded once - bit-identical to the reference arithmetic without
//   FMA contraction.
template <bool FAST>
__device__ __forceinline__ void point_xyz(const Q10& q, float d, float fc, float fr, float& X, float& Y, float& Z)
{
    const float w = __fmul_rn(q.q14, d);
    const float nx = __fadd_rn(__fadd_rn(__fmul_rn(q.q0, fc), __fmul_rn(q.q1, fr)), q.q3);
    const float ny = __fadd_rn(__fadd_rn(__fmul_rn(q.q4, fc), __fmul_rn(q.q5, fr)), q.q7);
    const float nz = __fadd_rn(__fadd_rn(__fmul_rn(q.q8, fc), __fmul_rn(q.q9, fr)), q.q11);
    if (FAST) {
        const float aw = fabsf(w);
        float r = rcp_approx(w);
        const bool common = aw >= 0x1p-60f && aw <= 0x1p60f;
        const bool special = aw == 0.0f || !(aw < INFINITY);
        if (common || special) {
            const float e = fmaf(-w, r, 1.0f);
            const float r1 = fmaf(r, e, r);
            const float qx = nx * r1, qy = ny * r1, qz = nz * r1;
            const float fx = fmaf(fmaf(-qx, w, nx), r1, qx), fy = fmaf(fmaf(-qy, w, ny), r1, qy), fz = fmaf(fmaf(-qz, w, nz), r1, qz);
            X = common ? fx : nx * r;
            Y = common ? fy : ny * r;
            Z = common ? fz : nz * r;
            return;
        }
    }
    X = __fdiv_rn(nx, w);
    Y = __fdiv_rn(ny, w);
    Z = __fdiv_rn(nz, w);
}

// (float) of byte k of a packed BGRx word without the conversion unit: 2^23 + byte, exactly, minus 2^23
__device__ __forceinline__ float byte_to_float(uint32_t c, uint32_t sel) { return __uint_as_float(__byte_perm(c, 0x4b000000u, sel)) - 8388608.0f; }

// Pitched colour rows: may the thread's 4 pixels be fetched as 16-byte windows (one per row they touch)?  Each window must end
// inside the frame's pixel data - a caller's buffer has no slack after the last row.
__device__ __forceinline__ bool pitched_windows_ok(const ReprojectArgs& a, int p0, int npix, int row0, int col0)
{
    if (p0 + 4 > npix) return false;
    const size_t data_end = (size_t)(a.h - 1) * a.color_pitch + 3 * (size_t)a.w;
    if ((size_t)row0 * a.color_pitch + 3 * (size_t)col0 + 16 > data_end) return false;
    if (col0 + 4 > a.w && (size_t)(row0 + 1) * a.color_pitch + 16 > data_end) return false;   // also false when row0 is the last row
    return true;
}

// Loads disparity + colour of the thread's 4 pixels.
//   VEC: disparity 16-byte aligned, colour rows contiguous and 4-byte aligned -> one float4 + three words.
//   !VEC: scalar disparity loads; colour from the enclosing aligned words, funnel-shifted by the frame's misalignment
//   (e.g. 1242x375 frames start on 2-byte boundaries), as long as the rows are contiguous and the 16-byte window stays
//   inside the frame; the last pixels of a frame and pitched images fall back to byte loads.
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
    } else if (a.color_pitch == (size_t)a.w * 3 && p0 + 8 <= npix) {
        // contiguous rows: the 4 pixels are 12 consecutive bytes, even across a row end
#pragma unroll
        for (int k = 0; k < 4; k++) d[k] = disp[p0 + k];
        const uintptr_t A = reinterpret_cast<uintptr_t>(color + (size_t)p0 * 3);
        const uint32_t al = (uint32_t)A & 3u, sh = al << 3;
        const uint32_t* cp = reinterpret_cast<const uint32_t*>(A - al);
        const uint32_t r0 = __ldg(cp), r1 = __ldg(cp + 1), r2 = __ldg(cp + 2), r3 = __ldg(cp + 3);   // cp + 3 is inside the frame: p0 + 8 <= npix
        const uint32_t u0 = __funnelshift_r(r0, r1, sh), u1 = __funnelshift_r(r1, r2, sh), u2 = __funnelshift_r(r2, r3, sh);
        c[0] = u0; c[1] = __byte_perm(u0, u1, 0x4543); c[2] = __byte_perm(u1, u2, 0x4432); c[3] = u2 >> 8;
    } else if (a.color_pitch != (size_t)a.w * 3 && pitched_windows_ok(a, p0, npix, row0, col0)) {
        // pitched rows (the padded rectified buffer of widths that are not a multiple of 4, or a caller's ROI): one 16-byte
        // window at the group's first pixel and, if the group runs over the row end, a second one at the start of the next
        // row - both issued before either is used
#pragma unroll
        for (int k = 0; k < 4; k++) d[k] = disp[p0 + k];
        const int n0 = min(4, a.w - col0);                                  // pixels of the group in row0
        auto window = [&](size_t off, uint32_t o[4]) {
            const uintptr_t A = reinterpret_cast<uintptr_t>(color + off);
            const uint32_t al = (uint32_t)A & 3u, sh = al << 3;
            const uint32_t* cp = reinterpret_cast<const uint32_t*>(A - al);
            const uint32_t r0 = __ldg(cp), r1 = __ldg(cp + 1), r2 = __ldg(cp + 2), r3 = __ldg(cp + 3);
            const uint32_t u0 = __funnelshift_r(r0, r1, sh), u1 = __funnelshift_r(r1, r2, sh), u2 = __funnelshift_r(r2, r3, sh);
            o[0] = u0; o[1] = __byte_perm(u0, u1, 0x4543); o[2] = __byte_perm(u1, u2, 0x4432); o[3] = u2 >> 8;
        };
        uint32_t ca[4], cb[4] = {0u, 0u, 0u, 0u};
        window((size_t)row0 * a.color_pitch + 3 * (size_t)col0, ca);
        if (n0 < 4) window((size_t)(row0 + 1) * a.color_pitch, cb);
        c[0] = ca[0];
        c[1] = n0 > 1 ? ca[1] : cb[0];
        c[2] = n0 > 2 ? ca[2] : (n0 == 2 ? cb[0] : cb[1]);
        c[3] = n0 > 3 ? ca[3] : (n0 == 3 ? cb[0] : (n0 == 2 ? cb[1] : cb[2]));
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

// Per-CTA disparity min/max -> scratch, no atomics: thousands of warps of one frame hammering the same two words
// serialise in the L2 atomic unit (it showed up as ~half of this kernel's time).  One REDUX per warp on the
// order-preserving integer encoding, one shared-memory hop, one 8-byte store per CTA; reduce_minmax_kernel folds the
// partials of a frame when a heatmap is actually requested.  Must be called by all threads of the CTA.
__device__ __forceinline__ void block_minmax(unsigned* __restrict__ blk_mm, unsigned* s_mm, float mn, float mx)
{
    const unsigned emn = __reduce_min_sync(0xffffffffu, enc_ordered(mn));
    const unsigned emx = __reduce_max_sync(0xffffffffu, enc_ordered(mx));
    const int warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { s_mm[2 * warp] = emn; s_mm[2 * warp + 1] = emx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned a = s_mm[0], b = s_mm[1];
#pragma unroll
        for (int i = 1; i < kThreads / 32; i++) { a = min(a, s_mm[2 * i]); b = max(b, s_mm[2 * i + 1]); }
        *reinterpret_cast<uint2*>(blk_mm) = make_uint2(a, ~b);      // same (min, ~max) convention as the final array
    }
}

// one CTA per frame: fold the per-CTA partials
__global__ void __launch_bounds__(kThreads) reduce_minmax_kernel(const unsigned* __restrict__ blk_mm, int nblk, unsigned* __restrict__ minmax)
{
    __shared__ unsigned s_mm[2 * kThreads / 32];
    const int frame = blockIdx.x;
    const uint2* in = reinterpret_cast<const uint2*>(blk_mm) + (size_t)frame * nblk;
    unsigned a = 0xffffffffu, b = 0xffffffffu;
    for (int i = threadIdx.x; i < nblk; i += kThreads) { const uint2 v = in[i]; a = min(a, v.x); b = min(b, v.y); }
    a = __reduce_min_sync(0xffffffffu, a);
    b = __reduce_min_sync(0xffffffffu, b);
    const int warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { s_mm[2 * warp] = a; s_mm[2 * warp + 1] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 1; i < kThreads / 32; i++) { a = min(a, s_mm[2 * i]); b = min(b, s_mm[2 * i + 1]); }
        minmax[2 * frame] = a;
        minmax[2 * frame + 1] = b;
    }
}

// VEC: aligned inputs (see load4).  VOUT: every frame of the point buffer starts on a 16-byte boundary.
template <bool VEC, bool VOUT, bool FAST>
__global__ void __launch_bounds__(kThreads) reproject_dense_kernel(const ReprojectArgs a)
{
    __shared__ __align__(16) float4 s_pts[kThreads / 32][192];
    __shared__ unsigned s_mm[2 * kThreads / 32];
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
    float fc = (float)col, fr = (float)row;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        point_xyz<FAST>(q, d[k], fc, fr, f[6 * k], f[6 * k + 1], f[6 * k + 2]);
        f[6 * k + 3] = byte_to_float(c[k], 0x7542);
        f[6 * k + 4] = byte_to_float(c[k], 0x7541);
        f[6 * k + 5] = byte_to_float(c[k], 0x7540);
        if (p0 + k < npix) { mn = fminf(mn, d[k]); mx = fmaxf(mx, d[k]); }
        fc += 1.0f;                                     // exact: columns and rows are far below 2^24
        if (++col == a.w) { col = 0; fc = 0.0f; fr += 1.0f; }
    }
    float4* sw = s_pts[warp];
    const int flip = (lane >> 2) & 1;
#pragma unroll
    for (int i = 0; i < 6; i++) sw[(lane * 6 + i) ^ flip] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
    __syncwarp();

    float* out = a.points + (size_t)frame * npix * 6;
    const size_t warp_px0 = (size_t)(blockIdx.x * kThreads + warp * 32) * 4;
    if (VOUT) {
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
    if (a.blk_minmax) block_minmax(a.blk_minmax + 2 * ((size_t)frame * gridDim.x + blockIdx.x), s_mm, mn, mx);
}

// ---- compaction ------------------------------------------------------------------------------------------
// Valid pixels per 1024-pixel block (the unit reproject_compact_kernel works in).  One CTA counts kCountBlocks consecutive blocks:
// four independent 16-byte loads per thread in flight, one REDUX per block and warp - as one block per CTA (one load per thread,
// a five-step shuffle reduction) the kernel ran at 3.6 TB/s, bounded by CTA turnover, not by the 4 bytes per pixel it reads.
constexpr int kCountBlocks = 4;
template <bool VEC>
__global__ void __launch_bounds__(kThreads) count_valid_kernel(const ReprojectArgs a, int nblk)
{
    __shared__ int s_w[kCountBlocks][kThreads / 32];
    const int tid = threadIdx.x, frame = blockIdx.y, npix = a.w * a.h;
    const float* disp = a.disp + (size_t)frame * a.disp_frame;
    float4 v[kCountBlocks];
#pragma unroll
    for (int j = 0; j < kCountBlocks; j++) {
        const int p0 = ((blockIdx.x * kCountBlocks + j) * kThreads + tid) * 4;
        v[j] = make_float4(0.f, 0.f, 0.f, 0.f);                          // 0 is not a valid disparity
        if (VEC) { if (p0 < npix) v[j] = *reinterpret_cast<const float4*>(disp + p0); }
        else {
            if (p0 < npix) v[j].x = disp[p0];
            if (p0 + 1 < npix) v[j].y = disp[p0 + 1];
            if (p0 + 2 < npix) v[j].z = disp[p0 + 2];
            if (p0 + 3 < npix) v[j].w = disp[p0 + 3];
        }
    }
#pragma unroll
    for (int j = 0; j < kCountBlocks; j++) {
        float4 t = v[j];
        if (a.negate) { t.x = -t.x; t.y = -t.y; t.z = -t.z; t.w = -t.w; }
        const int cnt = (int)valid_disp(t.x) + (int)valid_disp(t.y) + (int)valid_disp(t.z) + (int)valid_disp(t.w);
        const int w = __reduce_add_sync(0xffffffffu, cnt);
        if ((tid & 31) == 0) s_w[j][tid >> 5] = w;
    }
    __syncthreads();
    if (tid < kCountBlocks) {
        const int blk = blockIdx.x * kCountBlocks + tid;
        if (blk < nblk) {
            int t = 0;
#pragma unroll
            for (int i = 0; i < kThreads / 32; i++) t += s_w[tid][i];
            a.blk_counts[(size_t)frame * nblk + blk] = t;
        }
    }
}

// one CTA per frame: exclusive scan of its block counts, total -> counts[frame]
__global__ void __launch_bounds__(kThreads) scan_blocks_kernel(const int* __restrict__ blk_counts, int* __restrict__ blk_offsets, int* __restrict__ counts,
                                                               int nblk)
{
    __shared__ int s_warp[kThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, frame = blockIdx.x;
    const int* in = blk_counts + (size_t)frame * nblk;
    int* out = blk_offsets + (size_t)frame * nblk;
    const int per = (nblk + kThreads - 1) / kThreads;
    const int b0 = tid * per, b1 = min(nblk, b0 + per);
    int sum = 0;
    for (int i = b0; i < b1; i++) sum += in[i];
    int incl = sum;                                                      // inclusive scan over the warp, then over the warp totals
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int run = incl - sum;
#pragma unroll
    for (int i = 0; i < kThreads / 32; i++) {
        const int t = s_warp[i];
        if (i < warp) run += t;
        if (tid == kThreads - 1 && i == kThreads / 32 - 1) counts[frame] = run + sum;
    }
    for (int i = b0; i < b1; i++) { out[i] = run; run += in[i]; }
}

// Look-back word of the single-pass compaction: value (32) | flag (2) << 32 | epoch (30) << 34.  A word counts only if it carries
// the current launch's epoch, so the array is never cleared between launches.
constexpr unsigned long long kFlagAggregate = 1ull << 32, kFlagPrefix = 2ull << 32;
__device__ __forceinline__ unsigned long long ld_acquire(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ONEPASS: validity filter, order-preserving compaction and reprojection in ONE launch ("decoupled look-back"): a CTA takes the
// next 1024-pixel tile from a ticket (so every tile before it is running or done), publishes its own count, and finds its output
// offset by adding up the published counts of the tiles before it in the frame, 32 at a time, back to the first one that already
// knows its own offset.  !ONEPASS (the default): the offsets come from count_valid_kernel + scan_blocks_kernel (three launches).
// Measured on 64 frames of 640x480: 0.274 ms single pass against 0.147 ms for the three launches - a 1024-pixel tile is ~5 us of
// work, about what one look-back costs, and the CTAs hold their SM slots while they poll (same time with blockIdx instead of the
// ticket).  Option single_pass_compact = 1 selects it.
template <bool VEC, bool FAST, bool ONEPASS = false>
__global__ void __launch_bounds__(kThreads) reproject_compact_kernel(const ReprojectArgs a, int nblk)
{
    __shared__ __align__(16) float2 s_pts[kThreads / 32][128 * 3 + 2];   // (+ 2: the staged run starts one float2 in when its destination does)
    __shared__ int s_wtot[kThreads / 32];
    __shared__ unsigned s_mm[2 * kThreads / 32];
    __shared__ int s_tile, s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int frame = blockIdx.y, bx = blockIdx.x;
    if (ONEPASS) {
        if (tid == 0) s_tile = (int)atomicAdd(a.ticket, 1ull);
        __syncthreads();
        frame = s_tile / nblk;
        bx = s_tile - frame * nblk;
    }
    const int npix = a.w * a.h;
    const int p0 = (bx * kThreads + tid) * 4;
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
    const int cnt = __popc(mask);
    int incl = cnt;
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    const int wtotal = __shfl_sync(0xffffffffu, incl, 31);
    if (lane == 0) s_wtot[warp] = wtotal;
    int slot = incl - cnt;
    __syncthreads();
    int base;
    if (ONEPASS) {
        if (warp == 0) {
            int total = 0;
#pragma unroll
            for (int i = 0; i < kThreads / 32; i++) total += s_wtot[i];
            unsigned long long* st = a.status + (size_t)frame * nblk;
            const unsigned long long tag = (unsigned long long)a.epoch << 34;
            int excl = 0;
            if (bx > 0) {
                if (lane == 0) st_release(st + bx, tag | kFlagAggregate | (unsigned)total);       // my count, for the tiles after me
                int j = bx - 1;                                                                     // look back, 32 tiles per step
                while (true) {
                    const int k = j - lane;
                    unsigned long long v = 0;
                    bool ready = k < 0;                                                           // before the frame's first tile: nothing to add
                    if (!ready) { v = ld_acquire(st + k); ready = (v >> 34) == a.epoch && (v & (kFlagAggregate | kFlagPrefix)) != 0; }
                    if (!__all_sync(0xffffffffu, ready)) continue;                                // some tile in the window has not published yet
                    const unsigned pref = __ballot_sync(0xffffffffu, k >= 0 && (v & kFlagPrefix));
                    const int stop = pref ? __ffs(pref) - 1 : 32;                                 // first lane (nearest tile) that knows its prefix
                    int add = (k >= 0 && lane <= stop) ? (int)(unsigned)v : 0;
                    for (int o = 16; o; o >>= 1) add += __shfl_xor_sync(0xffffffffu, add, o);
                    excl += add;
                    if (pref || j - 32 < 0) break;
                    j -= 32;
                }
            }
            if (lane == 0) {
                st_release(st + bx, tag | kFlagPrefix | (unsigned)(excl + total));                   // my inclusive prefix
                if (bx == nblk - 1) a.counts[frame] = excl + total;
                s_base = excl;
            }
        }
        __syncthreads();
        base = s_base;
    } else {
        base = a.blk_offsets[(size_t)frame * nblk + bx];
    }
    for (int i = 0; i < warp; i++) base += s_wtot[i];
    // The warp's points leave as one contiguous run of wtotal * 24 bytes.  They are staged in shared memory with the same 16-byte phase
    // as their destination (a run that starts on an odd 8-byte word is staged one float2 in), so that all but at most one float2 at
    // either end go out as 16-byte loads and stores.
    float2* out = reinterpret_cast<float2*>(a.points + (size_t)frame * npix * 6) + (size_t)base * 3;
    const int par = (int)((reinterpret_cast<uintptr_t>(out) >> 3) & 1u);
    float2* sw = s_pts[warp] + par;
    float fc = (float)col, fr = (float)row;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (mask & (1u << k)) {
            float X, Y, Z;
            point_xyz<FAST>(q, d[k], fc, fr, X, Y, Z);
            sw[slot * 3 + 0] = make_float2(X, Y);
            sw[slot * 3 + 1] = make_float2(Z, byte_to_float(c[k], 0x7542));
            sw[slot * 3 + 2] = make_float2(byte_to_float(c[k], 0x7541), byte_to_float(c[k], 0x7540));
            ++slot;
        }
        fc += 1.0f;
        if (++col == a.w) { col = 0; fc = 0.0f; fr += 1.0f; }
    }
    __syncwarp();
    const int n2 = wtotal * 3;
    if (par && lane == 0 && n2 > 0) out[0] = sw[0];
    const int n4 = (n2 - par) >> 1;                                      // (n2 = 0 with par = 1: -1 >> 1 = -1, no iteration)
    float4* out4 = reinterpret_cast<float4*>(out + par);
    const float4* sw4 = reinterpret_cast<const float4*>(sw + par);
    for (int i = lane; i < n4; i += 32) out4[i] = sw4[i];
    if (n2 > par && ((n2 - par) & 1) && lane == 0) out[n2 - 1] = sw[n2 - 1];
    if (a.blk_minmax) block_minmax(a.blk_minmax + 2 * ((size_t)frame * nblk + bx), s_mm, mn, mx);
}

__global__ void fill_int_kernel(int* p, int n, int v)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

}  // namespace

int reproject_blocks_per_frame(int w, int h) { return (int)(((size_t)w * h + kPxPerBlock - 1) / kPxPerBlock); }

cudaError_t launch_reduce_minmax(const unsigned* blk_minmax, int nblk, int n, unsigned* minmax, cudaStream_t s, long* launches)
{
    if (n <= 0) return cudaSuccess;
    reduce_minmax_kernel<<<n, kThreads, 0, s>>>(blk_minmax, nblk, minmax);
    if (launches) ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_reproject(const ReprojectArgs& a, cudaStream_t s, long* launches)
{
    if (a.n <= 0) return cudaSuccess;
    const size_t npix = (size_t)a.w * a.h;
    const int nblk = reproject_blocks_per_frame(a.w, a.h);
    const bool vec = (npix % 4 == 0) && (a.disp_frame % 4 == 0) && (((uintptr_t)a.disp) % 16 == 0) && (a.color_pitch == (size_t)a.w * 3) &&
                     (a.color_frame % 4 == 0) && (((uintptr_t)a.color) % 4 == 0) && (((uintptr_t)a.points) % 16 == 0);
    // the shared-reciprocal quotient needs bounded, finite numerators: |Q entry| <= 2^20 and image sides <= 2^14
    bool fast = !a.exact_div && a.w <= 16384 && a.h <= 16384;
    for (int i : {0, 1, 3, 4, 5, 7, 8, 9, 11}) fast = fast && (fabsf(a.q[i]) <= 1048576.0f);   // false for NaN too
    // point rows leave as 16-byte stores when every frame of the point buffer starts on a 16-byte boundary (even pixel count)
    const bool vout = (npix % 2 == 0) && (((uintptr_t)a.points) % 16 == 0);
    dim3 grid((unsigned)nblk, (unsigned)a.n);
#define SA_DISPATCH(KERNEL, ...)                                                                         \
    do {                                                                                                 \
        if (vec) { if (fast) KERNEL<true, true><<<grid, kThreads, 0, s>>>(__VA_ARGS__); else KERNEL<true, false><<<grid, kThreads, 0, s>>>(__VA_ARGS__); } \
        else { if (fast) KERNEL<false, true><<<grid, kThreads, 0, s>>>(__VA_ARGS__); else KERNEL<false, false><<<grid, kThreads, 0, s>>>(__VA_ARGS__); }   \
    } while (0)
    if (!a.compact) {
#define SA_DENSE(V, O, F) reproject_dense_kernel<V, O, F><<<grid, kThreads, 0, s>>>(a)
        if (vec) { if (fast) SA_DENSE(true, true, true); else SA_DENSE(true, true, false); }
        else if (vout) { if (fast) SA_DENSE(false, true, true); else SA_DENSE(false, true, false); }
        else { if (fast) SA_DENSE(false, false, true); else SA_DENSE(false, false, false); }
#undef SA_DENSE
        if (launches) ++*launches;
        if (a.counts) { fill_int_kernel<<<(a.n + 255) / 256, 256, 0, s>>>(a.counts, a.n, (int)npix); if (launches) ++*launches; }
        return cudaGetLastError();
    }
    if (a.status && a.ticket && a.epoch) {                            // single pass
        cudaError_t e = cudaMemsetAsync(a.ticket, 0, sizeof(unsigned long long), s);
        if (e != cudaSuccess) return e;
        if (vec) { if (fast) reproject_compact_kernel<true, true, true><<<grid, kThreads, 0, s>>>(a, nblk); else reproject_compact_kernel<true, false, true><<<grid, kThreads, 0, s>>>(a, nblk); }
        else { if (fast) reproject_compact_kernel<false, true, true><<<grid, kThreads, 0, s>>>(a, nblk); else reproject_compact_kernel<false, false, true><<<grid, kThreads, 0, s>>>(a, nblk); }
        if (launches) ++*launches;
        return cudaGetLastError();
    }
    const dim3 cgrid((unsigned)((nblk + kCountBlocks - 1) / kCountBlocks), (unsigned)a.n);
    if (vec) count_valid_kernel<true><<<cgrid, kThreads, 0, s>>>(a, nblk); else count_valid_kernel<false><<<cgrid, kThreads, 0, s>>>(a, nblk);
    scan_blocks_kernel<<<a.n, kThreads, 0, s>>>(a.blk_counts, a.blk_offsets, a.counts, nblk);
    SA_DISPATCH(reproject_compact_kernel, a, nblk);
#undef SA_DISPATCH
    if (launches) *launches += 3;
    return cudaGetLastError();
}

}  // namespace sa

// =============================================================================================================
// sa_reproject_compact_noalloc: compacting reprojection for callers that may not allocate device memory.
//
// The compaction kernels above need scratch memory beyond their inputs and outputs: one valid-pixel count and one
// prefix offset per CTA of 1024 pixels.  sa_reproject takes that scratch from buffers the handle cudaMalloc'ed in
// sa_create.  Some callers can never reach a point where allocating is allowed: a TensorRT plugin's enqueue(), a
// camera or ROS callback that runs under a "no allocation after start-up" rule, or a process whose own memory pool
// has already claimed the whole device, so that a later cudaMalloc simply fails.  For them the scratch lives in the
// library image itself: a file-scope device array, reserved on every GPU that loads the library, for as long as the
// process lives.  The price is a fixed ceiling (kNoAllocMaxBlocks CTAs per call) and no concurrency between calls.
namespace sa {
namespace {

constexpr int kNoAllocMaxBlocks = 65536;   // 1024-pixel CTAs over all frames of one call: 218 frames of 640x480, 32 of 1080p, 8 of 4K

// counts in [0, kNoAllocMaxBlocks), offsets in [kNoAllocMaxBlocks, 2 * kNoAllocMaxBlocks): 512 KiB on each GPU.
static __device__ __align__(256) int g_noalloc_scratch[2 * kNoAllocMaxBlocks];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not an address a kernel could use; the address is
// looked up once per device (with that device current) and remembered here.
constexpr int kNoAllocMaxDevices = 64;
static std::atomic<int*> g_noalloc_addr[kNoAllocMaxDevices];

int noalloc_fail(char* err, size_t err_len, int code, const char* fmt, ...)
{
    if (err && err_len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace
}  // namespace sa

extern "C" int sa_reproject_compact_noalloc(float* disparity, size_t disparity_frame_stride, const uint8_t* left_bgr, size_t left_pitch,
                                            size_t left_frame_stride, const float* Q16, int w, int h, int n, int negate, float* points, int* counts,
                                            sa_stream stream, char* err, size_t err_len)
{
    using namespace sa;
    if (err && err_len) err[0] = 0;
    if (!disparity || !left_bgr || !Q16 || !points || !counts)
        return noalloc_fail(err, err_len, SA_ERR_INVALID_ARG, "sa_reproject_compact_noalloc: null pointer argument");
    if (w <= 0 || h <= 0 || w > 16384 || h > 16384 || n <= 0)
        return noalloc_fail(err, err_len, SA_ERR_INVALID_ARG, "sa_reproject_compact_noalloc: bad size %dx%d x %d frames", w, h, n);
    if (left_pitch == 0) left_pitch = (size_t)w * 3;
    if (left_pitch < (size_t)w * 3) return noalloc_fail(err, err_len, SA_ERR_INVALID_ARG, "sa_reproject_compact_noalloc: left_pitch %zu < 3*w", left_pitch);
    if (disparity_frame_stride == 0) disparity_frame_stride = (size_t)w * h;
    if (left_frame_stride == 0) left_frame_stride = left_pitch * (size_t)h;
    if (reinterpret_cast<uintptr_t>(points) % 8 != 0 || reinterpret_cast<uintptr_t>(disparity) % 4 != 0 || reinterpret_cast<uintptr_t>(counts) % 4 != 0)
        return noalloc_fail(err, err_len, SA_ERR_INVALID_ARG,
                            "sa_reproject_compact_noalloc: points must be 8-byte aligned (rows leave as float2 stores), disparity and counts 4-byte aligned");
    const long long nblk = reproject_blocks_per_frame(w, h);
    if (n > 65535)   // one grid row per frame; small frames would otherwise pass the slot limit below and fail as an opaque launch error
        return noalloc_fail(err, err_len, SA_ERR_TOO_LARGE, "sa_reproject_compact_noalloc: %d frames in one call, the limit is 65535; split the call", n);
    if (nblk * n > kNoAllocMaxBlocks)
        return noalloc_fail(err, err_len, SA_ERR_TOO_LARGE,
                            "sa_reproject_compact_noalloc: %d frames of %dx%d need %lld scratch slots, the built-in scratch holds %d "
                            "(one slot per 1024 pixels per frame); split the call or use sa_reproject",
                            n, w, h, nblk * n, kNoAllocMaxBlocks);

    // set-up, once per device: resolve the device address of the scratch array for the CURRENT device
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail(err, err_len, SA_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kNoAllocMaxDevices) return noalloc_fail(err, err_len, SA_ERR_INVALID_ARG, "device ordinal %d not supported (max %d)", dev, kNoAllocMaxDevices - 1);
    int* scratch = g_noalloc_addr[dev].load(std::memory_order_acquire);
    if (!scratch) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return noalloc_fail(err, err_len, SA_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
        scratch = static_cast<int*>(p);
        g_noalloc_addr[dev].store(scratch, std::memory_order_release);   // racing first calls resolve the same address
    }

    ReprojectArgs a{};
    a.disp = disparity;
    a.disp_frame = disparity_frame_stride;
    a.color = left_bgr;
    a.color_pitch = left_pitch;
    a.color_frame = left_frame_stride;
    a.points = points;
    for (int i = 0; i < 16; i++) a.q[i] = Q16[i];
    a.w = w; a.h = h; a.n = n;
    a.negate = negate ? 1 : 0;
    a.compact = 1;
    a.exact_div = 0;
    a.blk_minmax = nullptr;
    // Every slot this call reads is written by this call first (count kernel -> scan kernel -> compaction), so nothing
    // has to be zeroed; whatever the previous call left behind in the rest of the array is never looked at.
    a.blk_counts = scratch;
    a.blk_offsets = scratch + kNoAllocMaxBlocks;
    a.counts = counts;
    e = launch_reproject(a, static_cast<cudaStream_t>(stream), nullptr);
    if (e != cudaSuccess) return noalloc_fail(err, err_len, SA_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
    return SA_OK;
}
