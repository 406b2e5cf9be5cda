// Probe: do CUDA green contexts (SM partitioning) work on this box through the runtime + driver entry points?
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <set>
#include <vector>

#define DRV(name) decltype(&name) p_##name = nullptr; { cudaDriverEntryPointQueryResult qr; \
    if (cudaGetDriverEntryPoint(#name, (void**)&p_##name, cudaEnableDefault, &qr) != cudaSuccess || !p_##name) { printf("no %s\n", #name); return 1; } }

__global__ void smid_kernel(int* out, int spin) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    long long t0 = clock64();
    while (clock64() - t0 < spin) {}
    if (threadIdx.x == 0) out[blockIdx.x] = (int)smid;
}

int main() {
    cudaFree(0);
    DRV(cuDeviceGetDevResource) DRV(cuDevSmResourceSplitByCount) DRV(cuDevResourceGenerateDesc)
    DRV(cuGreenCtxCreate) DRV(cuGreenCtxStreamCreate) DRV(cuGreenCtxDestroy)
    CUdevResource sm;
    CUresult r = p_cuDeviceGetDevResource(0, &sm, CU_DEV_RESOURCE_TYPE_SM);
    printf("getDevResource %d smCount %u\n", (int)r, sm.sm.smCount);
    CUdevResource groups[2], rem;
    unsigned ng = 1;
    r = p_cuDevSmResourceSplitByCount(groups, &ng, &sm, &rem, 0, 24);
    printf("split %d ng %u first %u remaining %u\n", (int)r, ng, groups[0].sm.smCount, rem.sm.smCount);
    CUdevResourceDesc d0, d1;
    r = p_cuDevResourceGenerateDesc(&d0, &groups[0], 1); printf("desc0 %d\n", (int)r);
    r = p_cuDevResourceGenerateDesc(&d1, &rem, 1); printf("desc1 %d\n", (int)r);
    CUgreenCtx g0, g1;
    r = p_cuGreenCtxCreate(&g0, d0, 0, CU_GREEN_CTX_DEFAULT_STREAM); printf("create0 %d\n", (int)r);
    r = p_cuGreenCtxCreate(&g1, d1, 0, CU_GREEN_CTX_DEFAULT_STREAM); printf("create1 %d\n", (int)r);
    CUstream s0, s1;
    r = p_cuGreenCtxStreamCreate(&s0, g0, CU_STREAM_NON_BLOCKING, 0); printf("stream0 %d\n", (int)r);
    r = p_cuGreenCtxStreamCreate(&s1, g1, CU_STREAM_NON_BLOCKING, 0); printf("stream1 %d\n", (int)r);
    int* out;
    cudaMalloc(&out, 4096 * sizeof(int));
    std::vector<int> h(4096);
    for (int k = 0; k < 2; ++k) {
        smid_kernel<<<2048, 128, 0, (cudaStream_t)(k ? s1 : s0)>>>(out, 20000);
        cudaError_t e = cudaStreamSynchronize((cudaStream_t)(k ? s1 : s0));
        cudaMemcpy(h.data(), out, 2048 * sizeof(int), cudaMemcpyDeviceToHost);
        std::set<int> u(h.begin(), h.begin() + 2048);
        printf("stream %d: err %d distinct SMs %zu (min %d max %d)\n", k, (int)e, u.size(), *u.begin(), *u.rbegin());
    }
    // concurrency: long kernel on s1, short on s0 -- the short one must not wait
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    smid_kernel<<<124 * 16, 128, 0, (cudaStream_t)s1>>>(out, 4000000);
    cudaEventRecord(a, (cudaStream_t)s0);
    smid_kernel<<<24, 128, 0, (cudaStream_t)s0>>>(out + 2048, 2000);
    cudaEventRecord(b, (cudaStream_t)s0);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    printf("short kernel beside the long one: %.3f ms, long still running: %d\n", ms, cudaStreamQuery((cudaStream_t)s1) == cudaErrorNotReady);
    cudaDeviceSynchronize();
    return 0;
}
