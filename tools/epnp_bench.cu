// Debug tool: phase timing (clock64) of one EPnP hypothesis as pnp_hyp_kernel runs it (one warp per hypothesis).
//   nvcc -O3 -std=c++17 --fmad=false -gencode arch=compute_100a,code=sm_100a -o tools/epnp_bench tools/epnp_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ long long g_tick[64][8];
__host__ __device__ inline void tick(int k) {
#if defined(__CUDA_ARCH__)
    if (threadIdx.x == ((k == 5 || k == 6) ? 64 : 0)) g_tick[blockIdx.x][k] = clock64();
#endif
}
#define SVO_EPNP_TICK(k) tick(k);
#include "../stereo_visual_odometry_b200/csrc/svo_epnp.cuh"
using namespace svo;
struct WarpRedD {
    static __device__ __forceinline__ void sync() { __syncwarp(); }
    static __device__ __forceinline__ void sync_all() { __syncthreads(); }
    static __device__ __forceinline__ bool any(bool v) { return __any_sync(0xffffffffu, v); }
    static __device__ __forceinline__ double sumd(double v) { return v; }
};
__global__ void k(const float* X, const float* uv, Intrinsics K, double* out) {
    __shared__ EpnpWork w;
    Pose P;
    const float* x = X + 15 * blockIdx.x;
    const float* u = uv + 10 * blockIdx.x;
    const bool ok = epnp5_solve<32, 3, WarpRedD>(w, K, x, u, threadIdx.x & 31, threadIdx.x >> 5, P);
    if (threadIdx.x == 0) {
        for (int i = 0; i < 3; ++i) out[blockIdx.x * 4 + i] = P.t[i];
        out[blockIdx.x * 4 + 3] = ok;
    }
}
__global__ void lat(double* out, long long* cyc, double x0) {
    double x = x0, y = x0 * 0.5;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; ++i) x = x + y;            // dependent DADD
    long long t1 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; ++i) x = x * y + 1.0;      // DMUL + DADD (fmad=false)
    long long t2 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; ++i) x = 1.0 / (x + 1.5);  // DADD + DDIV
    long long t3 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; ++i) x = sqrt(x + 1.5);    // DADD + DSQRT
    long long t4 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; }
}
int main() {
    {
        double* o; long long* c; cudaMalloc(&o, 32 * 8); cudaMalloc(&c, 4 * 8);
        lat<<<1, 32>>>(o, c, 1.0000001); lat<<<1, 32>>>(o, c, 1.0000001);
        long long h[4]; cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
        printf("FP64 dependent latencies (cycles / step, loop overhead included): dadd %.1f  dmul+dadd %.1f  dadd+ddiv %.1f  dadd+dsqrt %.1f\n",
               h[0] / 256.0, h[1] / 256.0, h[2] / 256.0, h[3] / 256.0);
    }
    const int N = 64;
    static float hX[N * 15], hU[N * 10];
    unsigned s = 12345;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (float)(s >> 8) / 16777216.f; };
    Intrinsics K{718.856, 718.856, 607.1928, 185.2157};
    for (int b = 0; b < N; ++b)
        for (int i = 0; i < 5; ++i) {
            const float x = -8 + 16 * rnd(), y = -1 + 3 * rnd(), z = 4 + 40 * rnd();
            hX[b * 15 + 3 * i] = x; hX[b * 15 + 3 * i + 1] = y; hX[b * 15 + 3 * i + 2] = z;
            hU[b * 10 + 2 * i] = (float)(K.fx * (x + 0.01) / (z - 0.05) + K.cx) + 0.2f * (rnd() - 0.5f);
            hU[b * 10 + 2 * i + 1] = (float)(K.fy * y / (z - 0.05) + K.cy) + 0.2f * (rnd() - 0.5f);
        }
    float *dX, *dU; double* dO;
    cudaMalloc(&dX, sizeof(hX)); cudaMalloc(&dU, sizeof(hU)); cudaMalloc(&dO, N * 4 * sizeof(double));
    cudaMemcpy(dX, hX, sizeof(hX), cudaMemcpyHostToDevice); cudaMemcpy(dU, hU, sizeof(hU), cudaMemcpyHostToDevice);
    for (int rep = 0; rep < 3; ++rep) k<<<N, 96>>>(dX, dU, K, dO);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<<<N, 96>>>(dX, dU, K, dO); cudaEventRecord(e1);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    static long long t[64][8]; cudaMemcpyFromSymbol(t, g_tick, sizeof(t));
    static double o[N * 4]; cudaMemcpy(o, dO, sizeof(o), cudaMemcpyDeviceToHost);
    const char* nm[7] = {"setup(PCA,alphas)", "M,MtM", "SVD12", "L", "approx solve (SVD6xk)", "gauss-newton", "R_and_t"};
    double acc[7] = {0}, mx[7] = {0};
    for (int b = 0; b < N; ++b)
        for (int p = 0; p < 7; ++p) { const double d = (double)(t[b][p + 1] - t[b][p]); acc[p] += d / N; if (d > mx[p]) mx[p] = d; }
    printf("kernel %.1f us for %d hypotheses (1 warp each)\n", ms * 1e3, N);
    double tot = 0;
    for (int p = 0; p < 7; ++p) { printf("  %-24s mean %9.0f cycles  max %9.0f\n", nm[p], acc[p], mx[p]); tot += acc[p]; }
    printf("  total mean %.0f cycles; pose[0] t = %.9f %.9f %.9f ok %g\n", tot, o[0], o[1], o[2], o[3]);
    return 0;
}
