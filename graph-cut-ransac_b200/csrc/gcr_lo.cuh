// One round of the graph-cut local optimisation as ONE launch sequence (gcr_lo_round):
//   GCRANSAC::graphCutLocalOptimization, GCRANSAC.h:873-1062 (two classes) / :759-870 (one class), one pass of its while loop:
//     relabel every feature with the truncated threshold r^2 <= (1.5 t)^2                                   GCRANSAC.h:931-937
//     inner RANSAC: `n_trials` times draw min(7 x sample size, n) labelled inliers per class, fit, getScore   GCRANSAC.h:955-1046
//     keep the first trial that scores strictly higher than everything before it                             GCRANSAC.h:1034
// Launches: k_lo_label (labels of every feature -> mask words) | k_lo_draw (label words -> index lists, the
// trials' sub-samples and fit descriptors) | k_fit_rows (gcr_fit.cuh: fits, phi, candidate slots) | k_exact_items (scores, masks)
// | k_lo_select (the strict-< update, result into page-locked memory).  The host issued two synchronous calls per round before
// (labels back, sub-samples drawn on the host and copied over, fits + scores back); now it synchronises once and reads 100 bytes
// plus, when the round improved the model, its mask.
//
// The sub-samples are the HOST's, draw for draw: its sampler is SplitMix64, whose i-th output is a function of (state + i x golden
// ratio), so thread (trial, class) computes the draws that the host's sequential loop would hand to that trial and walks the same
// partial Fisher-Yates shuffle on a virtual copy of the list (only the <= 32 displaced positions are kept).
#pragma once
#include "gcr_fit.cuh"

namespace gcr
{

constexpr int kLoMaxSubset = 32;   // largest sub-sample per class this path draws (7 x minimal sample = 14 / 21 in the reference)
constexpr int kLoThreads = 1024;

struct LoRound // what the host knows before the round (kernel argument)
{
    Model4 model;                 // the so-far-best model: labels come from it
    double score;                 // its score: a trial must beat it
    unsigned long long rng_state; // SplitMix64 state before the round's first draw
    int n_trials;
    int lim[2];     // sub-sample size per class (inlierLimit, rectifying_homography_estimator.h:45-55)
    int min_inl[2]; // minimal sample per class: labels of a class no more numerous than this end the optimisation (sampling fails)
    int hybrid;     // two classes
};

struct LoResult // page-locked
{
    double model[4];
    double score;
    unsigned int n_inl[2];
    unsigned int updated;         // a trial beat the so-far-best: model / score / n_inl / the mask words behind this block are its
    unsigned int sampled;         // 0: too few labelled inliers, no trial was run
    unsigned int n_models_scored; // trials whose fit produced a model
    unsigned int winner_trial;
    unsigned long long draws;     // random numbers the round consumed
    unsigned int n_in[2];         // labelled inliers per class
};

struct LoDevice
{
    unsigned int* in_idx; // [ns + no] labelled inlier indices, ascending, scale class then orientation class
    unsigned int* info;   // [0..1]: labelled inliers per class, [2]: sampled, [3]: draws per trial
    FitDesc* desc;        // [n_trials]
    unsigned int* idx_s;  // [n_trials, lim[0]]
    unsigned int* idx_o;  // [n_trials, lim[1]]
    LoResult* out;
    unsigned int* out_words; // [W] page-locked
};

__device__ __forceinline__ unsigned long long splitmix_at(unsigned long long state, unsigned long long i) // output of the (i + 1)-th next()
{
    unsigned long long z = state + (i + 1ull) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// Labels of every feature against the so-far-best model (kernel argument): r^2 <= tsq, or < for the one-class problems, written as the
// packed mask words of candidate slot 0 (only the words: the labelling needs no sums).  Unit = 256 consecutive features of one class,
// a warp = one mask word, as in exact_items_phase.  First kernel of the round: plain launch.
__global__ void __launch_bounds__(kExactThreads) k_lo_label(Buffers B, Thresholds T, LoRound R)
{
    pdl_launch_dependents();
    pdl_wait();
    const unsigned int ups = (B.ns + kExactThreads - 1) / kExactThreads, upo = (B.no + kExactThreads - 1) / kExactThreads;
    const int lane = threadIdx.x & 31;
    for (unsigned int u = blockIdx.x; u < ups + upo; u += gridDim.x)
    {
        const int cls = u >= ups;
        const int f = static_cast<int>((cls ? u - ups : u) * kExactThreads) + threadIdx.x;
        const int n = cls ? B.no : B.ns;
        const double* F = cls ? B.orient : B.scale;
        bool in = false;
        if (f < n)
        {
            const double r = cls ? orientation_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], R.model)
                                 : scale_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], R.model, T.problem == 2);
            const double r2 = r * r;
            in = T.strict ? r2 < T.tsq[cls] : r2 <= T.tsq[cls];
        }
        __syncwarp();
        const unsigned int word = __ballot_sync(0xffffffffu, in);
        if (lane == 0 && (f - lane) < n) B.cand_masks[(cls ? B.ws : 0) + ((f - lane) >> 5)] = word;
    }
}

// Label words of candidate slot 0 -> ascending index lists (wordsToIndices of the host driver), then every trial's sub-samples and
// fit descriptor.  One CTA.
__global__ void __launch_bounds__(kLoThreads) k_lo_draw(Buffers B, LoRound R, LoDevice L)
{
    pdl_launch_dependents();
    pdl_wait();
    __shared__ unsigned int s_warp[kLoThreads / 32];
    __shared__ unsigned int s_base, s_n[2], s_sampled, s_draws[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int* words = B.cand_masks; // slot 0
    for (int cls = 0; cls < 2; cls++)
    {
        const int n_words = cls ? (B.no + 31) / 32 : B.ws;
        const unsigned int* w = words + (cls ? B.ws : 0);
        unsigned int* out = L.in_idx + (cls ? B.ns : 0);
        if (tid == 0) s_base = 0u;
        __syncthreads();
        for (int w0 = 0; w0 < n_words; w0 += kLoThreads)
        {
            const unsigned int bits = w0 + tid < n_words ? __ldcg(w + w0 + tid) : 0u;
            const unsigned int c = __popc(bits);
            unsigned int incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const unsigned int up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += up;
            }
            if (lane == 31) s_warp[warp] = incl;
            __syncthreads();
            if (warp == 0)
            {
                unsigned int v = s_warp[lane], iv = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1)
                {
                    const unsigned int up = __shfl_up_sync(0xffffffffu, iv, o);
                    if (lane >= o) iv += up;
                }
                s_warp[lane] = iv - v; // exclusive prefix of the warp totals
            }
            __syncthreads();
            unsigned int at = s_base + s_warp[warp] + incl - c;
            unsigned int b = bits;
            while (b)
            {
                const int bit = __ffs(b) - 1;
                b &= b - 1u;
                out[at++] = static_cast<unsigned int>(32 * (w0 + tid) + bit);
            }
            __syncthreads();
            if (tid == kLoThreads - 1) s_base = at; // (the last thread's end is the chunk's end: its own bits were added above)
            __syncthreads();
        }
        if (tid == 0) s_n[cls] = s_base;
        __syncthreads();
    }
    if (tid == 0)
    {
        const unsigned int n_s = s_n[0], n_o = R.hybrid ? s_n[1] : 0u;
        // GCRANSAC.h:955-996: the draw fails when a class has no more labelled inliers than its minimal sample
        const bool ok = !(n_s <= static_cast<unsigned int>(R.min_inl[0]) || (R.hybrid && n_o <= static_cast<unsigned int>(R.min_inl[1])));
        const unsigned int k_s = min(static_cast<unsigned int>(R.lim[0]), n_s), k_o = min(static_cast<unsigned int>(R.lim[1]), n_o);
        s_sampled = ok ? 1u : 0u;
        s_draws[0] = k_s < n_s ? k_s : 0u; // a class taken whole costs no random numbers
        s_draws[1] = k_o < n_o ? k_o : 0u;
        L.info[0] = n_s;
        L.info[1] = n_o;
        L.info[2] = s_sampled;
        L.info[3] = s_draws[0] + s_draws[1];
    }
    __syncthreads();
    // One thread per (trial, class).  (A warp per job with one record per lane and votes for the look-ups measured slower, 82 against
    // 77 us for the whole round at C1: four dependent collectives per draw.)
    const unsigned int per_trial = s_draws[0] + s_draws[1];
    for (int job = tid; job < 2 * R.n_trials; job += kLoThreads)
    {
        const int trial = job >> 1, cls = job & 1;
        const unsigned int n = cls ? (R.hybrid ? s_n[1] : 0u) : s_n[0];
        const unsigned int k = s_sampled ? min(static_cast<unsigned int>(R.lim[cls]), n) : 0u;
        const unsigned int* from = L.in_idx + (cls ? B.ns : 0);
        unsigned int* to = (cls ? L.idx_o : L.idx_s) + static_cast<size_t>(trial) * R.lim[cls];
        if (k < n)
        {
            // partial Fisher-Yates on a virtual copy of `from`: draw i swaps positions i and j >= i and emits the new entry at i; only the
            // positions written so far differ from `from` (position i itself is never looked at again).  The swap partners depend on the
            // generator only: every list entry the shuffle will read is prefetched into L1 before the first is needed -- one trip to L2
            // instead of 2k of them in a row.  (The list was written by this CTA in this launch and not read since: L1 holds no older copy.)
            unsigned int pos[kLoMaxSubset], val[kLoMaxSubset];
            const unsigned long long first = static_cast<unsigned long long>(trial) * per_trial + (cls ? s_draws[0] : 0u);
#pragma unroll 1
            for (unsigned int i = 0; i < k; i++)
            {
                const unsigned long long z = splitmix_at(R.rng_state, first + i);
                pos[i] = i + static_cast<unsigned int>(((z >> 32) * static_cast<unsigned long long>(n - i)) >> 32);
                asm volatile("prefetch.global.L1 [%0];" ::"l"(from + pos[i]));
                if ((i & 31u) == 0u) asm volatile("prefetch.global.L1 [%0];" ::"l"(from + i));
            }
#pragma unroll 1
            for (unsigned int i = 0; i < k; i++)
            {
                const unsigned int j = pos[i];
                unsigned int vi = from[i], vj = from[j];
                for (unsigned int q = 0; q < i; q++) // the latest record of a position wins
                {
                    if (pos[q] == i) vi = val[q];
                    if (pos[q] == j) vj = val[q];
                }
                val[i] = vi; // position j holds what position i held
                to[i] = j == i ? vi : vj;
            }
        }
        else
            for (unsigned int i = 0; i < k; i++) to[i] = __ldcg(from + i);
        if (cls == 0)
        {
            const unsigned int n_o = R.hybrid ? s_n[1] : 0u;
            FitDesc D;
            D.s_off = static_cast<long long>(trial) * R.lim[0];
            D.o_off = static_cast<long long>(trial) * R.lim[1];
            D.ns = static_cast<int>(k);
            D.no = s_sampled ? static_cast<int>(min(static_cast<unsigned int>(R.lim[1]), n_o)) : 0;
            D.unit_first = static_cast<unsigned int>(trial);
            D.ent_off = static_cast<unsigned int>(trial) * static_cast<unsigned int>(min(R.lim[1], kPhiBins));
            D.valid = !s_sampled ? 0 : (R.hybrid ? ((D.ns >= 2 && D.no >= 2 && !(D.ns == 2 && D.no == 2)) ? 1 : 0) : (D.ns > 3 ? 1 : 0));
            D.unit_rows = 1024;
            L.desc[trial] = D;
        }
    }
}

// The trials' scores -> the strict-< update in trial order (GCRANSAC.h:1034): the first trial with the highest score, if that beats the
// so-far-best.  One CTA; the result block and, if the model improved, its mask go to page-locked memory.
__global__ void __launch_bounds__(256) k_lo_select(Buffers B, Thresholds T, LoRound R, LoDevice L, const double* fit_models, const int* fit_ok)
{
    pdl_launch_dependents();
    pdl_wait();
    __shared__ double s_val[8];
    __shared__ int s_t[8];
    __shared__ unsigned int s_cnt[8];
    __shared__ int s_winner;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool sampled = __ldcg(L.info + 2) != 0u;
    double best = -1.0;
    int bt = -1;
    unsigned int n_ok = 0u;
    if (sampled)
        for (int t = tid; t < R.n_trials; t += 256)
        {
            if (!__ldcg(fit_ok + t)) continue;
            n_ok++;
            cand_value(B, T, static_cast<unsigned int>(t));
            const double v = B.cand_score[t];
            if (v > best) { best = v; bt = t; } // (ascending t: the first of equals stays)
        }
    __syncwarp();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int ot = __shfl_xor_sync(0xffffffffu, bt, o);
        n_ok += __shfl_xor_sync(0xffffffffu, n_ok, o);
        if (ot >= 0 && (ov > best || (ov == best && (bt < 0 || ot < bt)))) { best = ov; bt = ot; }
    }
    if (lane == 0)
    {
        s_val[warp] = best;
        s_t[warp] = bt;
        s_cnt[warp] = n_ok;
    }
    __syncthreads();
    if (tid == 0)
    {
        unsigned int total = 0u;
        for (int w = 0; w < 8; w++)
        {
            total += s_cnt[w];
            if (w > 0 && s_t[w] >= 0 && (s_val[w] > best || (s_val[w] == best && (bt < 0 || s_t[w] < bt)))) { best = s_val[w]; bt = s_t[w]; }
        }
        const bool updated = bt >= 0 && R.score < best;
        LoResult r;
        r.model[0] = R.model.h7; r.model[1] = R.model.h8; r.model[2] = R.model.alpha; r.model[3] = R.model.phi;
        r.score = R.score;
        r.n_inl[0] = r.n_inl[1] = 0u;
        if (updated)
        {
            for (int c = 0; c < 4; c++) r.model[c] = __ldcg(fit_models + 4 * bt + c);
            r.score = best;
            r.n_inl[0] = B.cand_cnt[2 * bt];
            r.n_inl[1] = B.cand_cnt[2 * bt + 1];
        }
        r.updated = updated ? 1u : 0u;
        r.sampled = sampled ? 1u : 0u;
        r.n_models_scored = total;
        r.winner_trial = updated ? static_cast<unsigned int>(bt) : 0xffffffffu;
        r.draws = sampled ? static_cast<unsigned long long>(R.n_trials) * __ldcg(L.info + 3) : 0ull;
        r.n_in[0] = __ldcg(L.info);
        r.n_in[1] = __ldcg(L.info + 1);
        *L.out = r;
        s_winner = updated ? bt : -1;
    }
    __syncthreads();
    if (s_winner >= 0)
        for (int w = tid; w < B.wmask; w += 256) L.out_words[w] = __ldcg(B.cand_masks + static_cast<size_t>(s_winner) * B.wmask + w);
}

} // namespace gcr
