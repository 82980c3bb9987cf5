// Micro-test of the TMA staging used by rectify_tma_kernel: one 3-D tensor copy of a (bw x bh x 1) box of 32-bit words,
// negative / overhanging coordinates zero-filled.  The box origin must be 16-byte aligned in the innermost dimension
// (a word coordinate that is not a multiple of 4 raises 'illegal instruction'): measured with this program.  nvcc -gencode arch=compute_100a,code=sm_100a -o tma_test tma_test.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int VARIANT>
__global__ void k(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int which, int cx, int cy, int cz, int bw, int bh, uint32_t* out)
{
    __shared__ __align__(128) uint32_t buf[3072];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const CUtensorMap* tm = VARIANT == 0 ? &tmA : (which ? &tmB : &tmA);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bw * bh * 4) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(buf)),
                     "l"(tm), "r"(cx), "r"(cy), "r"(cz), "r"(smem_u32(&bar))
                     : "memory");
    }
    asm volatile(
        "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(smem_u32(&bar)), "r"(0)
        : "memory");
    for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = buf[i];
}

int main()
{
    const int W = 640, H = 480, n = 3, bw = 56, bh = 20;
    const size_t pitch = 3 * W, frame = pitch * H;
    std::vector<uint8_t> img(frame * n);
    for (size_t i = 0; i < img.size(); i++) img[i] = (uint8_t)(i * 2654435761u >> 13);
    uint8_t* d; uint32_t* o;
    cudaMalloc(&d, img.size()); cudaMalloc(&o, bw * bh * 4);
    cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice);
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    printf("entry point: %d %d %p\n", (int)e, (int)q, p);
    CUtensorMap tm;
    const cuuint64_t dims[3] = {3 * W / 4, H, n}; const cuuint64_t strides[2] = {pitch, frame};
    const cuuint32_t box[3] = {bw, bh, 1}; const cuuint32_t es[3] = {1, 1, 1};
    CUresult r = ((encode_fn)p)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    if (r != CUDA_SUCCESS) return 1;
    const int cases[][3] = {{12, 30, 1}, {-4, -2, 0}, {3 * W / 4 - 20, H - 5, 2}, {0, 0, 2}, {-12, 470, 1}};
    int bad_total = 0;
    for (int v = 0; v < 2; v++)
        for (auto& c : cases) {
            cudaMemset(o, 0xff, bw * bh * 4);
            if (v == 0) k<0><<<1, 256>>>(tm, tm, 1, c[0], c[1], c[2], bw, bh, o);
            else k<1><<<1, 256>>>(tm, tm, 1, c[0], c[1], c[2], bw, bh, o);
            cudaError_t se = cudaDeviceSynchronize();
            if (se != cudaSuccess) { printf("variant %d case (%d,%d,%d): %s\n", v, c[0], c[1], c[2], cudaGetErrorString(se)); return 2; }
            std::vector<uint32_t> got(bw * bh);
            cudaMemcpy(got.data(), o, bw * bh * 4, cudaMemcpyDeviceToHost);
            int bad = 0;
            for (int y = 0; y < bh; y++)
                for (int x = 0; x < bw; x++) {
                    const int gx = c[0] + x, gy = c[1] + y;
                    uint32_t want = 0;
                    if (gx >= 0 && gx < 3 * W / 4 && gy >= 0 && gy < H) memcpy(&want, &img[c[2] * frame + gy * pitch + 4 * (size_t)gx], 4);
                    bad += got[y * bw + x] != want;
                }
            printf("variant %d case (%d,%d,%d): bad %d\n", v, c[0], c[1], c[2], bad);
            bad_total += bad;
        }
    return bad_total ? 3 : 0;
}
