// Issue-rate microbenchmarks of the B200 (sm_100a) non-tensor datapath: the measured denominators of the hybrid-MSAC scoring
// kernels' roofline (bench.py reads the JSON this prints; SURVEY 8d asks for the nominal FP32 peak AND an FMA measurement).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu && tools/ubench [--json profiles/x.json]
//
// Method: one CTA of 1024 threads per SM (8 warps per scheduler), every thread runs kIters iterations of kIlp INDEPENDENT dependency
// chains; every instruction is its own `asm volatile` with a read-write operand, so nothing can be folded, hoisted or removed.
// Cycles = SM clock (clock64) from the first warp's start to the last warp's end inside each CTA, averaged over the CTAs, so the
// rates are per SM clock and do not depend on the frequency the box happens to run at; the same launch is also timed with CUDA
// events (-> instructions per second at the clock the box actually sustained, which is what a roofline denominator is).
// A rate above 4 warp-instructions / clock / SM (the issue limit of 4 schedulers) would mean a broken measurement: asserted.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

constexpr int kIlp = 8;
constexpr int kThreads = 1024;

enum Kind
{
    K_FFMA, K_FFMA_ACC, K_FADD, K_FMUL, K_FFMA2, K_FSETP_PFFMA, K_FMNMX, K_LOP3, K_IADD, K_MUFU_LG2, K_MUFU_RCP, K_LDS128,
    K_MIX_SCALE, K_MIX_ORIENT, K_FFMA_MUFU_7_1, K_FFMA_FMNMX_1_1, K_COUNT
};

struct Desc { const char* name; int instr_per_step; int flops_per_step; const char* note; };
static const Desc kDesc[K_COUNT] = {
    {"FFMA  d = a*b + d (3 distinct registers)", 1, 2, "the FP32 peak denominator"},
    {"FFMA  d = d*b + c (chain form)", 1, 2, ""},
    {"FADD", 1, 1, ""},
    {"FMUL", 1, 1, ""},
    {"FFMA2 (fma.rn.f32x2: two FMAs per instruction)", 1, 4, "packed FP32 pair"},
    {"FSETP + predicated FFMA", 2, 3, "verdict + accumulate pattern"},
    {"FMNMX", 1, 1, "alu pipe"},
    {"LOP3", 1, 0, "alu pipe"},
    {"IADD3", 1, 0, "alu pipe"},
    {"MUFU.LG2", 1, 1, "xu pipe"},
    {"MUFU.RCP + FADD (the FADD keeps ptxas from folding rcp(rcp(x)))", 2, 2, "xu pipe: expect 2 x 0.5"},
    {"LDS.128 broadcast", 1, 0, "shared-memory feature read of the scoring kernels"},
    {"scale-evaluation mix: 2 FFMA + MUFU.LG2 + FADD + 2 FSETP + pred FFMA + 2 pred FADD", 9, 11, "one scale evaluation as issued (no loads)"},
    {"orientation-evaluation mix: 8 FFMA + 4 FMUL + 3 FADD + FMNMX + FMNMX3 + MUFU.RCP + 2 FSETP", 20, 30, "one orientation evaluation as issued (no loads)"},
    {"7 FFMA : 1 MUFU.LG2", 8, 14, "xu co-issue"},
    {"1 FFMA : 1 FMNMX", 2, 3, "fma + alu co-issue"},
};

template <int KIND>
__global__ void __launch_bounds__(kThreads, 1) k_bench(float* out, unsigned long long* cycles, int iters, float s0, float s1, float s2)
{
    __shared__ unsigned long long t_first, t_last;
    __shared__ float4 sh[256];
    if (threadIdx.x == 0) { t_first = ~0ull; t_last = 0ull; }
    if (threadIdx.x < 256) sh[threadIdx.x] = make_float4(s0, s1, s2, s0);
    float a[kIlp], acc[kIlp];
    unsigned int ia[kIlp];
    unsigned long long pa[kIlp];
    const float b = s0 + threadIdx.x * 1e-9f, c = s1, d = s2;
    const unsigned long long pb = (static_cast<unsigned long long>(__float_as_uint(s1)) << 32) | __float_as_uint(s1);
    const unsigned long long pc = (static_cast<unsigned long long>(__float_as_uint(s2)) << 32) | __float_as_uint(s2);
#pragma unroll
    for (int i = 0; i < kIlp; i++)
    {
        a[i] = 1.0f + 1e-3f * (i + threadIdx.x);
        acc[i] = 0.f;
        ia[i] = threadIdx.x * 17 + i;
        pa[i] = (static_cast<unsigned long long>(__float_as_uint(a[i])) << 32) | __float_as_uint(a[i]);
    }
    const unsigned sh_base = static_cast<unsigned>(__cvta_generic_to_shared(sh));
    __syncthreads();
    const unsigned long long t0 = clock64();
    for (int it = 0; it < iters; it++)
    {
#pragma unroll
        for (int i = 0; i < kIlp; i++)
        {
            if (KIND == K_FFMA) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(b), "f"(c));
            if (KIND == K_FFMA_ACC) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));
            if (KIND == K_FADD) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c));
            if (KIND == K_FMUL) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c));
            if (KIND == K_FFMA2) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pa[i]) : "l"(pb), "l"(pc));
            if (KIND == K_FSETP_PFFMA)
                // the compared value changes every step (ptxas optimises the PTX it is given: a loop-invariant setp would be hoisted)
                asm volatile("{.reg .pred p; setp.le.f32 p, %1, %2; @p fma.rn.f32 %1, %0, %3, %1;}" : "+f"(a[i]), "+f"(acc[i]) : "f"(1e30f), "f"(d));
            if (KIND == K_FMNMX) asm volatile("min.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c));
            if (KIND == K_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(ia[i]) : "r"(ia[(i + 1) % kIlp]), "r"(ia[(i + 2) % kIlp]));
            if (KIND == K_IADD) asm volatile("add.s32 %0, %0, %1;" : "+r"(ia[i]) : "r"(ia[(i + 1) % kIlp]));
            if (KIND == K_MUFU_LG2) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            if (KIND == K_MUFU_RCP) // (a chain of reciprocals alone is folded by ptxas: x -> 1/x -> x; the FADD keeps it, counted below)
            {
                asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
                asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c));
            }
            if (KIND == K_LDS128)
            {
                float x, y, z, w;
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(sh_base + 16u * ((it * kIlp + i) & 255)));
                asm volatile("" ::"f"(x), "f"(y), "f"(z), "f"(w));
                if (it == iters + 1) a[i] += x + y + z + w; // never true: keeps the destination registers live without extra instructions
            }
            if (KIND == K_MIX_SCALE)
            {
                // v = fma(h8, y, fma(h7, x, iq)); L = lg2 v; dd = A - L; in = |dd| <= Tin; unc = |dd| <= Tout; @in acc += dd*dd; @in cnt += 1
                float v, L;
                asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(v) : "f"(b), "f"(a[i]), "f"(c));
                asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(v) : "f"(d), "f"(a[i]));
                asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(L) : "f"(v));
                asm volatile("sub.rn.f32 %0, %1, %0;" : "+f"(L) : "f"(c));
                asm volatile("{.reg .pred p, q; .reg .f32 m; abs.f32 m, %2; setp.le.f32 p, m, %3; setp.le.f32 q, m, %4; @p fma.rn.f32 %0, %2, %2, %0; @p add.rn.f32 %1, %1, 0f3F800000; @q add.rn.f32 %1, %1, 0f00000000;}"
                             : "+f"(acc[i]), "+f"(a[i]) : "f"(L), "f"(d), "f"(b));
            }
            if (KIND == K_MIX_ORIENT)
            {
                float n, dd, c1, s1v, mn, mx, ri, t, lo, hi, z, p;
                asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(n) : "f"(a[i]), "f"(b), "f"(c));
                asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(dd) : "f"(a[i]), "f"(d), "f"(c));
                asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(c1) : "f"(dd), "f"(c));
                asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(c1) : "f"(n), "f"(d));
                asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(s1v) : "f"(dd), "f"(d));
                asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(s1v) : "f"(n), "f"(c));
                asm volatile("{.reg .f32 m, n; abs.f32 m, %2; abs.f32 n, %3; min.f32 %0, m, n; max.f32 %1, m, n;}" : "=f"(mn), "=f"(mx) : "f"(c1), "f"(s1v));
                asm volatile("max.f32 %0, %0, 0f0DA24260;" : "+f"(mx));
                asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(ri) : "f"(mx));
                asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(t) : "f"(mn), "f"(ri));
                asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(lo) : "f"(b), "f"(ri), "f"(c));
                asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(hi) : "f"(d), "f"(ri), "f"(c));
                asm volatile("mul.rn.f32 %0, %1, %1;" : "=f"(z) : "f"(t));
                asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(p) : "f"(z), "f"(c));
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(p) : "f"(z), "f"(d));
                asm volatile("{.reg .pred p, q; setp.le.f32 p, %2, %4; setp.le.f32 q, %2, %5; @p fma.rn.f32 %0, %3, %6, %0; @p add.rn.f32 %1, %1, 0f3F800000; @q add.rn.f32 %1, %1, 0f00000000;}"
                             : "+f"(acc[i]), "+f"(a[i]) : "f"(t), "f"(z), "f"(lo), "f"(hi), "f"(p));
            }
            if (KIND == K_FFMA_MUFU_7_1)
            {
#pragma unroll
                for (int r = 0; r < 7; r++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));
                asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(acc[i]));
            }
            if (KIND == K_FFMA_FMNMX_1_1)
            {
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));
                asm volatile("min.f32 %0, %0, %1;" : "+f"(acc[i]) : "f"(b));
            }
        }
    }
    const unsigned long long t1 = clock64();
    atomicMin(&t_first, t0);
    atomicMax(&t_last, t1);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kIlp; i++) s += a[i] + acc[i] + static_cast<float>(ia[i]) + static_cast<float>(pa[i] & 0xffff);
    if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t_last - t_first;
}

template <int KIND>
static void launch(int grid, float* out, unsigned long long* cyc, int iters, cudaStream_t s)
{
    k_bench<KIND><<<grid, kThreads, 0, s>>>(out, cyc, iters, 1.0000001f, 0.9999999f, 1e-7f);
}
typedef void (*launch_fn)(int, float*, unsigned long long*, int, cudaStream_t);
template <int K> struct Table { static void fill(launch_fn* t) { t[K] = launch<K>; Table<K + 1>::fill(t); } };
template <> struct Table<K_COUNT> { static void fill(launch_fn*) {} };

int main(int argc, char** argv)
{
    const char* json_path = nullptr;
    for (int i = 1; i + 1 < argc; i++) if (!std::strcmp(argv[i], "--json")) json_path = argv[i + 1];
    int dev = 0, n_sm = 0, clock_khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, dev);
    float* out; unsigned long long* cyc;
    cudaMalloc(&out, sizeof(float) * n_sm * kThreads);
    cudaMalloc(&cyc, sizeof(unsigned long long) * n_sm);
    launch_fn table[K_COUNT];
    Table<0>::fill(table);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::string json = "{\"n_sm\": " + std::to_string(n_sm) + ", \"max_clock_mhz\": " + std::to_string(clock_khz / 1000) + ", \"ops\": {";
    std::printf("SMs=%d  max SM clock=%d MHz  (%d threads/SM, %d independent chains per thread)\n", n_sm, clock_khz / 1000, kThreads, kIlp);
    double ffma_tflops_burst = 0, ffma_tflops_sustained = 0, ffma_rate = 0;
    bool broken = false;
    for (int k = 0; k < K_COUNT; k++)
    {
        const int iters = (k == K_MUFU_LG2 || k == K_MUFU_RCP) ? 4096 : 16384;
        table[k](n_sm, out, cyc, 256, 0); // warm-up (instruction cache, clocks)
        cudaDeviceSynchronize();
        // burst: one launch (~1 ms); sustained: back-to-back launches for ~0.5 s, the last one timed
        float ms_burst = 0.f, ms_sus = 0.f;
        cudaEventRecord(e0); table[k](n_sm, out, cyc, iters, 0); cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms_burst, e0, e1);
        std::vector<unsigned long long> h(n_sm);
        cudaMemcpy(h.data(), cyc, sizeof(unsigned long long) * n_sm, cudaMemcpyDeviceToHost);
        const int reps = std::max(4, static_cast<int>(500.0f / std::max(ms_burst, 0.05f)));
        for (int r = 0; r < reps; r++) table[k](n_sm, out, cyc, iters, 0);
        cudaEventRecord(e0); table[k](n_sm, out, cyc, iters, 0); cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms_sus, e0, e1);
        if (cudaGetLastError() != cudaSuccess) { std::printf("CUDA error\n"); return 1; }
        double cyc_avg = 0;
        for (auto v : h) cyc_avg += static_cast<double>(v);
        cyc_avg /= n_sm;
        const double instr_per_warp = static_cast<double>(iters) * kIlp * kDesc[k].instr_per_step;
        const double rate = (kThreads / 32) * instr_per_warp / cyc_avg; // warp-instructions / clock / SM
        const double total_instr = instr_per_warp * (kThreads / 32) * n_sm;
        const double flops = static_cast<double>(iters) * kIlp * kDesc[k].flops_per_step * kThreads * n_sm;
        const double tf_burst = flops / (ms_burst * 1e-3) / 1e12, tf_sus = flops / (ms_sus * 1e-3) / 1e12;
        const double eff_mhz = total_instr / n_sm / rate / (ms_burst * 1e-3) / 1e6;
        std::printf("%-64.64s  %6.3f warp-instr/clk/SM  %7.2f lane-ops/clk/SM  burst %6.3f ms (%6.1f Tflop/s, eff. clock %4.0f MHz)  sustained %6.3f ms (%6.1f Tflop/s)\n",
                    kDesc[k].name, rate, rate * 32.0, ms_burst, tf_burst, eff_mhz, ms_sus, tf_sus);
        if (rate > 4.05) { std::printf("BROKEN MEASUREMENT above: more than 4 warp-instructions per clock per SM\n"); broken = true; }
        if (k == K_FFMA) { ffma_tflops_burst = tf_burst; ffma_tflops_sustained = tf_sus; ffma_rate = rate; }
        char buf[512];
        std::snprintf(buf, sizeof(buf), "%s\"%d\": {\"name\": \"%s\", \"warp_instr_per_clk_per_sm\": %.4f, \"burst_ms\": %.4f, \"sustained_ms\": %.4f, "
                      "\"tflops_burst\": %.3f, \"tflops_sustained\": %.3f, \"effective_clock_mhz\": %.0f}", k ? ", " : "", k, kDesc[k].name, rate, ms_burst,
                      ms_sus, tf_burst, tf_sus, eff_mhz);
        json += buf;
    }
    char tail[512];
    std::snprintf(tail, sizeof(tail), "}, \"fp32_ffma_tflops_burst\": %.3f, \"fp32_ffma_tflops_sustained\": %.3f, \"ffma_warp_instr_per_clk_per_sm\": %.4f, "
                  "\"fp32_nominal_tflops\": %.3f}", ffma_tflops_burst, ffma_tflops_sustained, ffma_rate, n_sm * 128.0 * 2.0 * clock_khz * 1e3 / 1e12);
    json += tail;
    std::printf("FP32 FFMA measured: burst %.2f Tflop/s, sustained %.2f Tflop/s (nominal %d SM x 128 x 2 x %.3f GHz = %.2f)\n", ffma_tflops_burst,
                ffma_tflops_sustained, n_sm, clock_khz / 1e6, n_sm * 128.0 * 2.0 * clock_khz * 1e3 / 1e12);
    if (json_path && !broken)
    {
        FILE* f = std::fopen(json_path, "w");
        if (f) { std::fputs(json.c_str(), f); std::fputs("\n", f); std::fclose(f); }
    }
    return broken ? 2 : 0;
}
