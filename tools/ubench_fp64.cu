// FP64 issue rate and dependent-issue latency of this GPU (the non-minimal fit and the FP64 reference-order kernels are FP64 code):
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_fp64 tools/ubench_fp64.cu && tools/ubench_fp64
#include <cstdio>
#include <cuda_runtime.h>
template <int CHAINS>
__global__ void k_dfma(double* out, unsigned long long* cyc, int iters, double a, double b)
{
    double x[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; c++) x[c] = a + c + threadIdx.x;
    const unsigned long long t0 = clock64();
    for (int i = 0; i < iters; i++)
    {
#pragma unroll
        for (int r = 0; r < 16; r++)
#pragma unroll
            for (int c = 0; c < CHAINS; c++) x[c] = __fma_rn(x[c], b, a);
    }
    const unsigned long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_dadd_chain(double* out, unsigned long long* cyc, int iters, double a)
{
    double x = a + threadIdx.x;
    const unsigned long long t0 = clock64();
    for (int i = 0; i < iters; i++)
    {
#pragma unroll
        for (int r = 0; r < 16; r++) x = __dadd_rn(x, a);
    }
    const unsigned long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_ffma_chain(float* out, unsigned long long* cyc, int iters, float a, float b)
{
    float x = a + threadIdx.x;
    const unsigned long long t0 = clock64();
    for (int i = 0; i < iters; i++)
    {
#pragma unroll
        for (int r = 0; r < 16; r++) x = __fmaf_rn(x, b, a);
    }
    const unsigned long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main()
{
    int n_sm = 0, khz = 0;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double* out; unsigned long long* cyc; float* outf;
    cudaMalloc(&out, 8 * 1024 * n_sm); cudaMalloc(&outf, 4 * 1024); cudaMalloc(&cyc, 8 * n_sm);
    unsigned long long h[1024];
    const int iters = 2000;
    for (int rep = 0; rep < 2; rep++)
    {
        k_dfma<1><<<1, 32>>>(out, cyc, iters, 1.0, 0.999999);
        cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
        if (rep) std::printf("DFMA dependent chain, one warp:           %.1f cycles per instruction\n", double(h[0]) / (16.0 * iters));
        k_dadd_chain<<<1, 32>>>(out, cyc, iters, 1e-9);
        cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
        if (rep) std::printf("DADD dependent chain, one warp:           %.1f cycles per instruction\n", double(h[0]) / (16.0 * iters));
        k_dadd_chain<<<1, 1>>>(out, cyc, iters, 1e-9);
        cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
        if (rep) std::printf("DADD dependent chain, one thread:         %.1f cycles per instruction\n", double(h[0]) / (16.0 * iters));
        k_ffma_chain<<<1, 32>>>(outf, cyc, iters, 1.0f, 0.999f);
        cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
        if (rep) std::printf("FFMA dependent chain, one warp:           %.1f cycles per instruction\n", double(h[0]) / (16.0 * iters));
        k_dfma<4><<<1, 32>>>(out, cyc, iters, 1.0, 0.999999);
        cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
        if (rep) std::printf("DFMA 4 independent chains, one warp:      %.1f cycles per instruction\n", double(h[0]) / (64.0 * iters));
        k_dfma<4><<<n_sm, 1024>>>(out, cyc, iters, 1.0, 0.999999);
        cudaMemcpy(h, cyc, 8 * n_sm, cudaMemcpyDeviceToHost);
        double worst = 0; for (int i = 0; i < n_sm; i++) worst = h[i] > worst ? double(h[i]) : worst;
        const double per_sm = 32.0 * 64.0 * iters / worst; // warp instructions per clock per SM
        if (rep) std::printf("DFMA throughput, 32 warps x 4 chains / SM: %.3f warp-instr/clk/SM = %.1f lanes/clk/SM -> %.2f Tflop/s at %d MHz x %d SMs\n", per_sm, per_sm * 32,
                             per_sm * 32 * 2 * (khz * 1e3) * n_sm * 1e-12, khz / 1000, n_sm);
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { std::printf("CUDA error\n"); return 1; }
    return 0;
}
