// Debug tool: one cp.async.bulk.tensor.3d tile load of a uint8 plane, checked against the source.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__global__ void k(const __grid_constant__ CUtensorMap map, int x, int y, int z, unsigned char* out, int bw, int bh) {
    extern __shared__ __align__(128) unsigned char tile[];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bw * bh) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(smem_u32(tile)), "l"(&map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(&bar)) : "memory");
    }
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = tile[i];
}
int main(int argc, char** argv) {
    const int W = 1241, H = 376, P = 1280, N = 3, BW = argc > 1 ? atoi(argv[1]) : 144, BH = argc > 2 ? atoi(argv[2]) : 22;
    std::vector<unsigned char> h((size_t)P * H * N);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (unsigned char)(i * 2654435761u >> 24);
    unsigned char *d, *o;
    cudaMalloc(&d, h.size()); cudaMalloc(&o, BW * BH);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
    typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    CUtensorMap map;
    cuuint64_t gd[3] = {W, H, N}, gs[2] = {P, (cuuint64_t)P * H};
    cuuint32_t box[3] = {(cuuint32_t)BW, (cuuint32_t)BH, 1}, es[3] = {1, 1, 1};
    CUresult r = ((Enc)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, gd, gs, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc %d (qr %d)\n", (int)r, (int)qr);
    const int xs[6] = {112, 8, -16, 1192, 4, -8}, ys[6] = {10, 10, -4, 360, -4, 5};
    for (int test = 0; test < 6; ++test) {
        const int x = xs[test], y = ys[test], z = test % 3;
        k<<<1, 128, BW * BH>>>(map, x, y, z, o, BW, BH);
        cudaError_t e = cudaDeviceSynchronize();
        printf("test %d (x %d y %d z %d): %s\n", test, x, y, z, cudaGetErrorString(e));
        if (e != cudaSuccess) { printf("   (stopping)\n"); return 1; }
        std::vector<unsigned char> t(BW * BH);
        cudaMemcpy(t.data(), o, t.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int r_ = 0; r_ < BH; ++r_)
            for (int c = 0; c < BW; ++c) {
                const int gx = x + c, gy = y + r_;
                const unsigned char want = (gx < 0 || gx >= W || gy < 0 || gy >= H) ? 0 : h[(size_t)z * P * H + (size_t)gy * P + gx];
                bad += t[r_ * BW + c] != want;
            }
        printf("   mismatches %d\n", bad);
    }
    return 0;
}
