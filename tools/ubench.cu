// Pipe-throughput microbenchmarks for the B200 (sm_100a) non-tensor datapath: the denominators of the
// hybrid-MSAC kernel's issue-bound roofline.  Prints warp-instructions per clock per SM for each op mix.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench.cu && ./ubench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048
#define NACC 8

template <int KIND>
__global__ void __launch_bounds__(256) k_bench(float* out, unsigned long long* cycles, float s0, float s1, float s2)
{
    float a[NACC];
    float b = s0 + threadIdx.x * 1e-9f, c = s1, d = s2;
    unsigned int ia[NACC];
    unsigned long long pa[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) { a[i] = s0 * i + threadIdx.x; ia[i] = threadIdx.x + i; pa[i] = threadIdx.x * 0x100000001ull + i; }
    __shared__ float4 sh[64];
    if (threadIdx.x < 64) sh[threadIdx.x] = make_float4(s0, s1, s2, s0);
    __syncthreads();
    const unsigned long long t0 = clock64();
    for (int it = 0; it < ITERS; it++)
    {
#pragma unroll
        for (int i = 0; i < NACC; i++)
        {
            if (KIND == 0) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(b), "f"(c));                       // FFMA, 2 distinct + acc
            if (KIND == 1) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));                       // FFMA chain form
            if (KIND == 2) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c));                                   // FADD
            if (KIND == 3) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c));                                   // FMUL
            if (KIND == 4) asm volatile("{.reg .pred p; setp.le.f32 p, %0, %1; @p add.s32 %2, %2, 1;}" : "+f"(a[i]) : "f"(c), "r"(ia[i])); // FSETP + pred IADD (not chained)
            if (KIND == 5) asm volatile("min.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c));                                      // FMNMX
            if (KIND == 6) asm volatile("add.s32 %0, %0, %1;" : "+r"(ia[i]) : "r"(ia[(i + 1) % NACC]));                    // IADD
            if (KIND == 7) asm volatile("lg2.approx.f32 %0, %0;" : "+f"(a[i]));                                            // MUFU.LG2
            if (KIND == 8) asm volatile("rcp.approx.f32 %0, %0;" : "+f"(a[i]));                                            // MUFU.RCP
            if (KIND == 9) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pa[i]) : "l"(pa[(i + 1) % NACC]), "l"(pa[(i + 2) % NACC])); // FFMA2
            if (KIND == 10) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));                     // 1 FFMA : 1 FMNMX
                              asm volatile("min.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b)); }
            if (KIND == 11) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));                     // 2 FFMA : 1 FMNMX
                              asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(d), "f"(c));
                              asm volatile("min.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b)); }
            if (KIND == 12) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));                     // 4 FFMA : 1 MUFU
                              asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(d), "f"(c));
                              asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));
                              asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(d), "f"(c));
                              asm volatile("lg2.approx.f32 %0, %0;" : "+f"(a[i])); }
            if (KIND == 13) { float4 v = sh[(it + i) & 63]; a[i] += v.x + v.y + v.z + v.w; }                                // LDS.128 broadcast + 4 FADD
            if (KIND == 14) { asm volatile("{.reg .pred p; setp.le.f32 p, %0, %1; @p fma.rn.f32 %2, %0, %0, %2;}" : "+f"(a[i]) : "f"(c), "f"(d)); } // FSETP + pred FFMA
            if (KIND == 15) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(ia[i]) : "r"(ia[(i + 1) % NACC]), "r"(ia[(i + 2) % NACC])); // LOP3
            if (KIND == 16) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pa[i]) : "l"(pa[(i + 1) % NACC]), "l"(pa[(i + 2) % NACC])); // FFMA2 : FMNMX 1:1
                              asm volatile("min.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b)); }
            if (KIND == 17) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pa[i]) : "l"(pa[(i + 1) % NACC]), "l"(pa[(i + 2) % NACC])); // FFMA2 : MUFU 2:1
                              asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pa[i]) : "l"(pa[(i + 2) % NACC]), "l"(pa[(i + 1) % NACC]));
                              asm volatile("lg2.approx.f32 %0, %0;" : "+f"(a[i])); }
        }
    }
    const unsigned long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += a[i] + (float)ia[i] + (float)(pa[i] & 0xffff);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int KIND>
void run(const char* name, int instr_per_slot, int blocks_per_sm, int nsm, float* out, unsigned long long* cyc)
{
    const int grid = nsm * blocks_per_sm;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_bench<KIND><<<grid, 256>>>(out, cyc, 1.0001f, 0.9999f, 1e-3f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k_bench<KIND><<<grid, 256>>>(out, cyc, 1.0001f, 0.9999f, 1e-3f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    unsigned long long* h = new unsigned long long[grid];
    cudaMemcpy(h, cyc, grid * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < grid; i++) avg += (double)h[i]; avg /= grid;
    // warp-instructions per SM: blocks_per_sm * 8 warps * ITERS * NACC * instr_per_slot, all resident concurrently
    const double winstr = (double)blocks_per_sm * 8.0 * ITERS * NACC * instr_per_slot;
    printf("%-34s bps=%d  cycles=%9.0f  warp-instr/clk/SM=%6.3f  lane-ops/clk/SM=%7.1f  kernel=%.3f ms  eff.clock=%.0f MHz\n", name,
           blocks_per_sm, avg, winstr / avg, 32.0 * winstr / avg, ms, avg / (ms * 1e3));
    delete[] h;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
}

int main()
{
    int nsm = 0; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("SMs=%d  clockRate attr=%d kHz\n", nsm, clk);
    float* out; unsigned long long* cyc;
    cudaMalloc(&out, sizeof(float) * nsm * 8 * 256);
    cudaMalloc(&cyc, sizeof(unsigned long long) * nsm * 8);
    for (int bps : {2, 4})
    {
        run<0>("FFMA acc=b*c+acc", 1, bps, nsm, out, cyc);
        run<1>("FFMA acc=acc*c+d", 1, bps, nsm, out, cyc);
        run<2>("FADD", 1, bps, nsm, out, cyc);
        run<3>("FMUL", 1, bps, nsm, out, cyc);
        run<4>("FSETP + @p IADD", 2, bps, nsm, out, cyc);
        run<5>("FMNMX", 1, bps, nsm, out, cyc);
        run<6>("IADD", 1, bps, nsm, out, cyc);
        run<15>("LOP3", 1, bps, nsm, out, cyc);
        run<7>("MUFU.LG2", 1, bps, nsm, out, cyc);
        run<8>("MUFU.RCP", 1, bps, nsm, out, cyc);
        run<9>("FFMA2 (f32x2)", 1, bps, nsm, out, cyc);
        run<10>("FFMA:FMNMX 1:1", 2, bps, nsm, out, cyc);
        run<11>("FFMA:FMNMX 2:1", 3, bps, nsm, out, cyc);
        run<12>("FFMA:MUFU 4:1", 5, bps, nsm, out, cyc);
        run<13>("LDS.128 bcast + 4 FADD", 5, bps, nsm, out, cyc);
        run<14>("FSETP + @p FFMA", 2, bps, nsm, out, cyc);
        run<16>("FFMA2:FMNMX 1:1", 2, bps, nsm, out, cyc);
        run<17>("FFMA2:MUFU 2:1", 3, bps, nsm, out, cyc);
    }
    return 0;
}
