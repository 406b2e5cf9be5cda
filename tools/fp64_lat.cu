// Debug tool: dependent-chain latency of FP64 / FP32 / integer ops on this GPU, full warp vs one active lane.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE> __device__ __forceinline__ double step(double x, double y) {
    if (MODE == 0) return x + y;
    if (MODE == 1) return x * y;
    if (MODE == 2) return fma(x, y, y);
    if (MODE == 3) return y / x;
    if (MODE == 4) return sqrt(x);
    return x;
}
template <int MODE> __global__ void lat(double* out, long long* cyc, double x0, double y0, int active) {
    if ((int)threadIdx.x >= active) return;
    double x = x0, y = y0;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 64; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) x = step<MODE>(x, y);
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void latf(float* out, long long* cyc, float x0, float y0) {
    float x = x0, y = y0;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 64; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) x = __fadd_rn(x, y);
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
    double* o; long long* c; cudaMalloc(&o, 32 * 8); cudaMalloc(&c, 8);
    const char* nm[5] = {"dadd", "dmul", "dfma", "ddiv", "dsqrt"};
    for (int active : {32, 1}) {
        for (int m = 0; m < 5; ++m) {
            for (int rep = 0; rep < 2; ++rep) {
                if (m == 0) lat<0><<<1, 32>>>(o, c, 1.0000001, 1e-9, active);
                if (m == 1) lat<1><<<1, 32>>>(o, c, 1.0000001, 1.0000000001, active);
                if (m == 2) lat<2><<<1, 32>>>(o, c, 0.5, 0.25, active);
                if (m == 3) lat<3><<<1, 32>>>(o, c, 1.7, 1.3, active);
                if (m == 4) lat<4><<<1, 32>>>(o, c, 1.7, 1.3, active);
            }
            long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
            printf("active lanes %2d  %-6s %6.1f cycles per dependent op\n", active, nm[m], h / 1024.0);
        }
    }
    float* of = (float*)o;
    latf<<<1, 32>>>(of, c, 1.f, 1e-3f); latf<<<1, 32>>>(of, c, 1.f, 1e-3f);
    long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("fadd   %6.1f cycles per dependent op\n", h / 1024.0);
    return 0;
}
