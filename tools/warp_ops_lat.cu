// Debug tool: latency (one dependent chain) and throughput (8 independent chains per warp, 1..8 warps per SM
// sub-partition) of the warp-level ops the KLT reductions are built from: REDUX, SHFL, I2F, F2I, LDS, S2R.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE> __device__ __forceinline__ int step(int x, int lane, const int* sm) {
    if (MODE == 0) return __reduce_add_sync(0xffffffffu, x);
    if (MODE == 1) return __shfl_xor_sync(0xffffffffu, x, 1) + 1;
    if (MODE == 2) return (int)__int2float_rn(x) + 1;   // I2F + F2I
    if (MODE == 3) return sm[(x + lane) & 31];
    if (MODE == 4) { int t; asm volatile("mov.u32 %0, %%laneid;" : "=r"(t)); return x + t; }
    if (MODE == 5) return x * 3 + 1;
    if (MODE == 6) return __float_as_int(__fadd_rn(__int_as_float(x), 1.5f));
    return x;
}
template <int MODE, int ILP> __global__ void run(int* out, long long* cyc, int x0) {
    __shared__ int sm[32];
    const int lane = threadIdx.x & 31;
    if (threadIdx.x < 32) sm[lane] = (lane * 7) & 31;
    __syncthreads();
    int x[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) x[k] = x0 + k + lane;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 128; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < ILP; ++k) x[k] = step<MODE>(x[k], lane, sm);
    }
    long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
template <int MODE> void bench(const char* name, int* o, long long* c) {
    long long h;
    for (int rep = 0; rep < 2; ++rep) run<MODE, 1><<<1, 32>>>(o, c, 1);
    cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("%-10s latency %6.1f cycles", name, h / 1024.0);
    for (int warps : {4, 8, 16, 32}) {   // warps per block = per SM (4 sub-partitions)
        for (int rep = 0; rep < 2; ++rep) run<MODE, 8><<<1, 32 * warps>>>(o, c, 1);
        cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        // ops issued per SM sub-partition: warps / 4 * 8 * 1024 in h cycles
        printf("  | %2d warps: %5.2f cyc/op/SMSP", warps, h / (warps / 4.0 * 8 * 1024.0));
    }
    printf("\n");
}
int main() {
    int* o; long long* c; cudaMalloc(&o, 4096 * 4); cudaMalloc(&c, 8);
    bench<0>("REDUX.ADD", o, c);
    bench<1>("SHFL+IADD", o, c);
    bench<2>("I2F+F2I+1", o, c);
    bench<3>("LDS", o, c);
    bench<4>("S2R+IADD", o, c);
    bench<5>("IMAD", o, c);
    bench<6>("FADD", o, c);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
