// Feature front end on the device: what the reference's examples do on the host between
// cv2.SIFT and pygcransac (examples/utils.py:5-49), for keypoints that already live in device memory.
//   scale_features_from_sift        utils.py:5-26   one (x, y, size) row per unique (int(x), int(y)), the FIRST keypoint of
//                                                   each coordinate, rows in order of first appearance (dict order)
//   orientation_features_from_sift  utils.py:29-49  (x, y, deg2rad(angle)) for every keypoint with angle != -1, in input
//                                                   order, plus 0.5 * size per row
// Keypoints are rows (x, y, size, angle_degrees), float32 (what cv2.KeyPoint holds) or float64.
//
// De-duplication: an open-addressing table of keypoint indices; a slot is claimed by compare-and-swap and then lowered
// with atomicMin by every later keypoint of the same coordinate, so the survivor of a coordinate is its lowest index
// whatever the thread order.  Compaction is a stable two-level scan, so the row order is the reference's.
// Integer work: results are bit-identical to the Python functions (deg2rad is one FP64 multiplication by pi/180).
#pragma once
#include <cstdint>

namespace gcr
{

constexpr int kFrontThreads = 256;
constexpr unsigned int kFrontEmpty = 0xffffffffu;

struct FrontBuffers
{
    const void* kp;      // [n,4] rows (x, y, size, angle)
    int is_f32;
    unsigned int n;
    unsigned int* table; // [cap] keypoint index or kFrontEmpty
    unsigned int cap_mask;
    unsigned int* block_counts; // [2, n_blocks] then scanned in place to exclusive offsets; [2 * n_blocks .. +2] = totals
    unsigned int n_blocks;
    double* scale;  // out f64[<= n, 3]
    double* orient; // out f64[<= n, 3]
    double* sizes;  // out f64[<= n]
    unsigned int* votes; // [2, ceil(n / 32)]: per warp of 32 consecutive keypoints, who yields a scale row / an orientation row (written by the
                         // counting pass, so that the emitting pass does not walk the table again)
};
constexpr int kFrontScanThreads = 1024;

__device__ __forceinline__ void front_load(const FrontBuffers& F, unsigned int i, double v[4])
{
    if (F.is_f32)
    {
        const float4 r = reinterpret_cast<const float4*>(F.kp)[i];
        v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
    }
    else
    {
        const double2* p = reinterpret_cast<const double2*>(F.kp) + 2 * static_cast<size_t>(i);
        const double2 a = p[0], b = p[1];
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
    }
}

// (int(x), int(y)) of utils.py:19: truncation toward zero (saturating outside the int32 range)
__device__ __forceinline__ unsigned long long front_key(double x, double y)
{
    const int xi = __double2int_rz(x), yi = __double2int_rz(y);
    return (static_cast<unsigned long long>(static_cast<unsigned int>(xi)) << 32) | static_cast<unsigned int>(yi);
}

__device__ __forceinline__ unsigned int front_hash(unsigned long long k)
{
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdull;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ull;
    k ^= k >> 33;
    return static_cast<unsigned int>(k);
}

__global__ void __launch_bounds__(kFrontThreads) k_front_insert(FrontBuffers F)
{
    const unsigned int i = blockIdx.x * kFrontThreads + threadIdx.x;
    if (i >= F.n) return;
    double v[4];
    front_load(F, i, v);
    const unsigned long long key = front_key(v[0], v[1]);
    unsigned int slot = front_hash(key) & F.cap_mask;
    for (;;)
    {
        unsigned int cur = F.table[slot];
        if (cur == kFrontEmpty)
        {
            cur = atomicCAS(&F.table[slot], kFrontEmpty, i);
            if (cur == kFrontEmpty) return; // claimed
        }
        double w[4];
        front_load(F, cur, w); // every index a slot ever holds has the slot's coordinate
        if (front_key(w[0], w[1]) == key)
        {
            atomicMin(&F.table[slot], i);
            return;
        }
        slot = (slot + 1) & F.cap_mask;
    }
}

// first keypoint of its coordinate?
__device__ __forceinline__ bool front_is_first(const FrontBuffers& F, unsigned int i, unsigned long long key)
{
    unsigned int slot = front_hash(key) & F.cap_mask;
    for (;;)
    {
        const unsigned int cur = F.table[slot];
        if (cur == i) return true;
        if (cur == kFrontEmpty) return false; // cannot happen after k_front_insert
        double w[4];
        front_load(F, cur, w);
        if (front_key(w[0], w[1]) == key) return false;
        slot = (slot + 1) & F.cap_mask;
    }
}

// EMIT = false: per-block counts of the two outputs.  EMIT = true: rows written at block offset + rank inside the block.
template <bool EMIT>
__global__ void __launch_bounds__(kFrontThreads) k_front_compact(FrontBuffers F)
{
    __shared__ unsigned int s_warp[2][kFrontThreads / 32];
    const unsigned int i = blockIdx.x * kFrontThreads + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double v[4] = {0, 0, 0, 0};
    bool is_scale = false, is_orient = false;
    const unsigned int n_warps = (F.n + 31u) / 32u, gw = i >> 5;
    unsigned int bs, bo;
    if (!EMIT)
    {
        if (i < F.n)
        {
            front_load(F, i, v);
            is_scale = front_is_first(F, i, front_key(v[0], v[1]));
            is_orient = v[3] != -1.0; // utils.py:45
        }
        __syncwarp();
        bs = __ballot_sync(0xffffffffu, is_scale);
        bo = __ballot_sync(0xffffffffu, is_orient);
        if (lane == 0 && gw < n_warps)
        {
            F.votes[gw] = bs;
            F.votes[n_warps + gw] = bo;
        }
    }
    else
    {
        bs = gw < n_warps ? F.votes[gw] : 0u;
        bo = gw < n_warps ? F.votes[n_warps + gw] : 0u;
        is_scale = (bs >> lane) & 1u;
        is_orient = (bo >> lane) & 1u;
        if (is_scale || is_orient) front_load(F, i, v);
    }
    if (lane == 0)
    {
        s_warp[0][warp] = __popc(bs);
        s_warp[1][warp] = __popc(bo);
    }
    __syncthreads();
    unsigned int before_s = 0, before_o = 0, total_s = 0, total_o = 0;
#pragma unroll
    for (int w = 0; w < kFrontThreads / 32; w++)
    {
        if (w < warp) { before_s += s_warp[0][w]; before_o += s_warp[1][w]; }
        total_s += s_warp[0][w];
        total_o += s_warp[1][w];
    }
    if (!EMIT)
    {
        if (threadIdx.x == 0)
        {
            F.block_counts[blockIdx.x] = total_s;
            F.block_counts[F.n_blocks + blockIdx.x] = total_o;
        }
        return;
    }
    const unsigned int lower = (1u << lane) - 1u;
    if (is_scale)
    {
        const size_t r = F.block_counts[blockIdx.x] + before_s + __popc(bs & lower);
        F.scale[3 * r] = v[0];
        F.scale[3 * r + 1] = v[1];
        F.scale[3 * r + 2] = v[2];
    }
    if (is_orient)
    {
        const size_t r = F.block_counts[F.n_blocks + blockIdx.x] + before_o + __popc(bo & lower);
        F.orient[3 * r] = v[0];
        F.orient[3 * r + 1] = v[1];
        F.orient[3 * r + 2] = v[3] * (3.14159265358979323846 / 180.0); // np.deg2rad
        F.sizes[r] = 0.5 * v[2];
    }
}

// exclusive scan of the two count rows in place (one CTA), totals appended: chunks of 1 024 counts per row, coalesced, a shuffle scan
// per warp and one over the warp totals (the strided one-thread-per-segment form took 58 us for 16 384 blocks)
__global__ void __launch_bounds__(kFrontScanThreads) k_front_scan(FrontBuffers F)
{
    __shared__ unsigned int s_warp[2][kFrontScanThreads / 32];
    __shared__ unsigned int s_run[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 2) s_run[tid] = 0u;
    __syncthreads();
    for (unsigned int b0 = 0; b0 < F.n_blocks; b0 += kFrontScanThreads)
    {
        const unsigned int b = b0 + tid;
        unsigned int c[2], incl[2];
#pragma unroll
        for (int row = 0; row < 2; row++)
        {
            c[row] = b < F.n_blocks ? F.block_counts[static_cast<size_t>(row) * F.n_blocks + b] : 0u;
            incl[row] = c[row];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const unsigned int up = __shfl_up_sync(0xffffffffu, incl[row], o);
                if (lane >= o) incl[row] += up;
            }
            if (lane == 31) s_warp[row][warp] = incl[row];
        }
        __syncthreads();
        if (warp < 2)
        {
            const unsigned int v = s_warp[warp][lane];
            unsigned int iv = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const unsigned int up = __shfl_up_sync(0xffffffffu, iv, o);
                if (lane >= o) iv += up;
            }
            s_warp[warp][lane] = iv - v; // exclusive prefix of the warp totals of row `warp`
        }
        __syncthreads();
        unsigned int end[2];
#pragma unroll
        for (int row = 0; row < 2; row++)
        {
            end[row] = s_run[row] + s_warp[row][warp] + incl[row];
            if (b < F.n_blocks) F.block_counts[static_cast<size_t>(row) * F.n_blocks + b] = end[row] - c[row];
        }
        __syncthreads();
        if (tid == kFrontScanThreads - 1)
        {
            s_run[0] = end[0];
            s_run[1] = end[1];
        }
        __syncthreads();
    }
    if (tid < 2) F.block_counts[2 * static_cast<size_t>(F.n_blocks) + tid] = s_run[tid];
}

} // namespace gcr
