// Phase timeline of one small fit in k_fit_rows (the finishing CTA's %globaltimer stamps):
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -DGCR_FIT_STAMPS -I graph-cut-ransac_b200/csrc -o tools/fit_stamps tools/fit_stamps.cu
//   tools/fit_stamps [n_scale n_orient]
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <unordered_map>
#include <vector>
#include "gcr_kernels_common.cuh"
#include "gcr_fit.cuh"
using namespace gcr;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
int main(int argc, char** argv)
{
    const int ns = argc > 1 ? std::atoi(argv[1]) : 14, no = argc > 2 ? std::atoi(argv[2]) : 14, N = 2000;
    std::vector<double> scale(3 * N), orient(3 * N);
    srand(5);
    auto u = [] { return rand() / double(RAND_MAX); };
    const double h7 = 2.3e-4, h8 = -1.7e-4, alpha = 0.62;
    for (int i = 0; i < N; i++)
    {
        const double x = 4000 * u(), y = 3000 * u(), w = 1.0 - h7 * x - h8 * y;
        scale[3 * i] = x; scale[3 * i + 1] = y; scale[3 * i + 2] = std::pow(w / alpha, 3.0) * (1.0 + 1e-3 * u());
        orient[3 * i] = 4000 * u(); orient[3 * i + 1] = 3000 * u(); orient[3 * i + 2] = 3.14159 * u();
    }
    std::vector<unsigned int> idx_s(ns), idx_o(no);
    for (int i = 0; i < ns; i++) idx_s[i] = rand() % N;
    for (int i = 0; i < no; i++) idx_o[i] = rand() % N;
    FitDesc D{};
    D.ns = ns; D.no = no; D.valid = 1;
    const long long rows = ns + 1ll * no * (no - 1) / 2;
    D.unit_rows = 1024;
    while (D.unit_rows < kFitUnitRows && (rows + D.unit_rows - 1) / D.unit_rows > kFitUnitsAim) D.unit_rows *= 2;
    const unsigned int units = static_cast<unsigned int>((rows + D.unit_rows - 1) / D.unit_rows);
    double *d_scale, *d_orient, *d_lines, *d_partial, *d_models;
    unsigned int *d_is, *d_io, *d_done, *d_nent; int* d_ok; FitDesc* d_desc; FitEntry* d_ent;
    CK(cudaMalloc(&d_scale, 24 * N)); CK(cudaMalloc(&d_orient, 24 * N)); CK(cudaMalloc(&d_lines, 24 * (no + 1))); CK(cudaMalloc(&d_partial, 144 * units));
    CK(cudaMalloc(&d_models, 32)); CK(cudaMalloc(&d_is, 4 * ns)); CK(cudaMalloc(&d_io, 4 * no)); CK(cudaMalloc(&d_done, 4)); CK(cudaMalloc(&d_nent, 4));
    CK(cudaMalloc(&d_ok, 4)); CK(cudaMalloc(&d_desc, sizeof(FitDesc))); CK(cudaMalloc(&d_ent, sizeof(FitEntry) * kPhiBins));
    CK(cudaMemcpy(d_scale, scale.data(), 24 * N, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_orient, orient.data(), 24 * N, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_is, idx_s.data(), 4 * ns, cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_io, idx_o.data(), 4 * no, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_desc, &D, sizeof(D), cudaMemcpyHostToDevice)); CK(cudaMemset(d_done, 0, 4));
    FitBuffers F{};
    F.scale = d_scale; F.orient = d_orient; F.desc = d_desc; F.idx_s = d_is; F.idx_o = d_io;
    F.lines = no <= kFitSmemLines ? nullptr : d_lines; F.done = d_done; F.partial = d_partial; F.models = d_models; F.ok = d_ok; F.n_ent = d_nent; F.ent = d_ent;
    F.n_fits = 1; F.n_units = units; F.n_lines = no; F.problem = 0;
    {
        std::unordered_map<int, double> probe;
        size_t last = probe.bucket_count();
        for (int i = 0; i < kPhiBins; ++i)
        {
            probe[i] += 1.0;
            if (probe.bucket_count() != last && F.n_rehash < 8) { last = probe.bucket_count(); F.rehash_at[F.n_rehash] = i + 1; F.rehash_to[F.n_rehash++] = int(last); }
        }
    }
    FitSlots S{nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 8; rep++)
    {
        const int passes = rep >= 6 ? 2 : 1; // the last two launches run the finish twice: the second pass finds its code in the instruction caches
        CK(cudaMemcpyToSymbol(g_fit_passes, &passes, sizeof(int)));
        if (F.lines) k_fit_lines<<<(no + kFitThreads - 1) / kFitThreads, kFitThreads>>>(F);
        cudaEventRecord(e0);
        k_fit_rows<<<units, kFitThreads>>>(F, S);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        unsigned long long st[32];
        CK(cudaMemcpyFromSymbol(st, g_fit_stamps, sizeof(st)));
        double m[4]; int ok;
        CK(cudaMemcpy(m, d_models, 32, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&ok, d_ok, 4, cudaMemcpyDeviceToHost));
        std::printf("rep %d: %llu SM cycles in %llu ns: effective SM clock %.0f MHz\n", rep, st[24] - st[16], st[8] - st[0], 1e3 * double(st[24] - st[16]) / double(st[8] - st[0]));
        std::printf("rep %d (finish x %d): cycles of the last pass: unit sums %llu combine %llu elimination %llu back-substitution %llu | solve %llu phi-hist %llu entries %llu mode %llu\n", rep,
                    passes, st[25] - st[19], st[26] - st[25], st[28] - st[26], st[27] - st[28], st[20] - st[19], st[21] - st[20], st[22] - st[21], st[23] - st[22]);
        std::printf("rep %d: active mask at the top of elimination step 0: %08llx, later steps: %08llx\n", rep, st[13], st[14]);
        std::printf("rep %d: kernel %.1f us (events) | ns from start: lines %llu rows %llu counter %llu solve %llu phi-hist %llu entries %llu mode %llu end %llu | ok %d model %.6g %.6g %.6g %.6g\n",
                    rep, ms * 1e3, st[1] - st[0], st[2] - st[0], st[3] - st[0], st[4] - st[0], st[5] - st[0], st[6] - st[0], st[7] - st[0], st[8] - st[0], ok, m[0], m[1], m[2], m[3]);
    }
    return 0;
}
