// ba_setup.cu — structure build of the window fast path ON THE DEVICE.
//
// What g2o does on the host in addVertex / addEdge / initializeOptimization (reference
// include/bundle_adjuster.h:122-166) — and what the first version of this backend did in a
// single host thread, costing 20x the LM loop itself — happens here in a few kernels over the
// raw edge arrays exactly as the caller passed them:
//   1. validate + key      edge -> 64-bit key (global point << 32 | global pose)
//   2. sort                stable radix sort of (key, edge index)            [cub, setup only]
//   3. gather + CSR        obs_cam / obs_uv in (point, pose) order, pt_ptr, duplicate detection
//   4. tiles               greedy cut of every problem's points into tiles (<= kWinPtCap points,
//                          <= kWinObsCap observations), one thread per problem
//   5. per tile (CTA)      keyframe-major slots (win_meta, win_uv, win_slot), pair counts per
//                          S block;  after a scan over the tiles: the pair lists themselves
// Everything is deterministic: ties are broken by the stable sort, the lists are ordered by
// point, and the only atomics are integer counters whose totals do not depend on order.
#include "ba_setup.h"

#include <algorithm>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace visoba {

namespace {

constexpr int kSetupThreads = 128;

__device__ __forceinline__ int find_segment(const int64_t* __restrict__ off, int n, int64_t x)
{ // largest b in [0, n) with off[b] <= x
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (off[mid] <= x) lo = mid;
        else hi = mid;
    }
    return lo;
}

// ---- 1. validate + key ----
// key = point << cam_bits | pose, cam_bits = the bits n_cams needs: the radix sort walks only the bits in use
// (C5: 35 bits = 5 passes of 8 instead of the 7 that point << 32 | pose costs)
__global__ void k_setup_keys(SetupIn in, int cam_bits, unsigned long long* __restrict__ keys, unsigned int* __restrict__ vals, int* __restrict__ err)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= in.n_edges) return;
    int pose = in.e_pose[e], point = in.e_point[e];
    const float2 uv = in.e_uv[e];
    int code = 0;
    if (in.local_indices) {
        const int b = find_segment(in.edge_off, in.n_problems, e);
        const int nc = in.prob_cam0[b + 1] - in.prob_cam0[b], np = in.prob_pt0[b + 1] - in.prob_pt0[b];
        if (pose < 0 || pose >= nc || point < 0 || point >= np) code = 2; // VISO_BA_ERR_INDEX
        pose += in.prob_cam0[b];
        point += in.prob_pt0[b];
    } else if (pose < 0 || pose >= in.n_cams || point < 0 || point >= in.n_points) {
        code = 2;
    }
    if (!isfinite(uv.x) || !isfinite(uv.y)) code = code ? code : 5; // VISO_BA_ERR_NONFINITE
    if (code) {
        atomicMax(err, code);
        atomicMin(reinterpret_cast<unsigned long long*>(err + 2), (unsigned long long)e);
        pose = 0, point = 0;
    }
    keys[e] = ((unsigned long long)(unsigned)point << cam_bits) | (unsigned)pose;
    vals[e] = (unsigned int)e;
}

// ---- 3. gather + CSR ----
__global__ void k_setup_gather(SetupIn in, int cam_bits, const unsigned long long* __restrict__ keys, const unsigned int* __restrict__ vals,
    int32_t* __restrict__ obs_cam, float2* __restrict__ obs_uv, int* __restrict__ err)
{
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= in.n_edges) return;
    const unsigned long long k = keys[s];
    obs_cam[s] = (int32_t)(k & ((1ull << cam_bits) - 1ull));
    obs_uv[s] = in.e_uv[vals[s]];
    if (s > 0 && keys[s - 1] == k) err[1] = 1; // duplicate (point, pose) edge: the general path handles it
}

__global__ void k_setup_ptptr(int64_t n_edges, int32_t n_points, int cam_bits, const unsigned long long* __restrict__ keys, int32_t* __restrict__ pt_ptr)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p > n_points) return;
    const unsigned long long want = (unsigned long long)(unsigned)p << cam_bits;
    int64_t lo = 0, hi = n_edges; // first s with keys[s] >= want
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (keys[mid] < want) lo = mid + 1;
        else hi = mid;
    }
    pt_ptr[p] = (int32_t)lo;
}

// ---- 4. tiles ----
// greedy cut, identical in the counting and the filling pass
template <bool kFill>
__device__ __forceinline__ int cut_tiles(const int32_t* __restrict__ pt_ptr, int q, int hi, WinTile proto, WinTile* out, int64_t part_stride, int64_t blkptr_stride)
{
    int n = 0;
    while (q < hi) {
        int e = q;
        const int o0 = pt_ptr[q];
        { // the largest e <= min(hi, q + kWinPtCap) with pt_ptr[e] - o0 <= kWinObsCap: pt_ptr is monotone, so the
          // greedy scan's end point is found by bisection (23 tiles x 7 probes instead of 2000 loads per window)
            int lo = q, up = min(hi, q + kWinPtCap);
            while (lo < up) {
                const int mid = (lo + up + 1) >> 1;
                if (pt_ptr[mid] - o0 <= kWinObsCap) lo = mid;
                else up = mid - 1;
            }
            e = lo;
        }
        if (e == q) ++e; // a point with more observations than a tile holds (only with duplicate edges,
                         // which send the handle to the general path anyway): keep the loop finite
        if (kFill) {
            WinTile t = proto;
            t.p0 = q, t.np = e - q, t.o0 = o0, t.no = pt_ptr[e] - o0;
            t.part_off = proto.part_off + n * part_stride;
            t.blkptr_off = proto.blkptr_off + n * blkptr_stride;
            t.list_off = 0;
            out[n] = t;
        }
        ++n;
        q = e;
    }
    return n;
}

// One CTA per problem: the problem's slice of pt_ptr is staged in shared memory (a window: 2001 entries), so the
// probes of the sequential cut cost a shared-memory round trip instead of a dependent L2 miss each; larger
// problems read pt_ptr in place.
constexpr int kCutStage = 4096;
constexpr int kCutThreads = 128;

__device__ __forceinline__ const int32_t* stage_pt_ptr(const int32_t* __restrict__ pt_ptr, int q0, int hi, int32_t* s_buf)
{
    const bool fits = hi - q0 + 1 <= kCutStage;
    if (fits)
        for (int i = threadIdx.x; i <= hi - q0; i += kCutThreads) s_buf[i] = pt_ptr[q0 + i];
    __syncthreads();
    return fits ? s_buf - q0 : pt_ptr;
}

__global__ void __launch_bounds__(kCutThreads) k_setup_tiles_count(SetupIn in, const int32_t* __restrict__ pt_ptr, int32_t* __restrict__ prob_ntiles)
{
    __shared__ int32_t s_buf[kCutStage];
    const int b = blockIdx.x;
    const int q0 = in.prob_pt0[b], hi = in.prob_pt0[b + 1];
    const int32_t* src = stage_pt_ptr(pt_ptr, q0, hi, s_buf);
    if (threadIdx.x) return;
    WinTile proto = {};
    prob_ntiles[b] = cut_tiles<false>(src, q0, hi, proto, nullptr, 0, 0);
}

// one thread: prefix sums over the problems (B is small next to the edge count)
__global__ void k_setup_prob_offsets(SetupIn in, const int32_t* __restrict__ prob_ntiles, const int32_t* __restrict__ prob_nfree,
    int32_t* __restrict__ prob_tile_ptr, int64_t* __restrict__ prob_part_off, int64_t* __restrict__ prob_blkptr_off, int64_t* __restrict__ totals)
{
    if (blockIdx.x || threadIdx.x) return;
    int32_t t = 0;
    int64_t part = 0, bp = 0;
    for (int b = 0; b < in.n_problems; ++b) {
        prob_tile_ptr[b] = t, prob_part_off[b] = part, prob_blkptr_off[b] = bp;
        const int64_t nblk = (int64_t)prob_nfree[b] * (prob_nfree[b] + 1) / 2;
        t += prob_ntiles[b];
        part += prob_ntiles[b] * nblk * kItemPart;
        bp += prob_ntiles[b] * win_blkptr_entries(nblk);
    }
    prob_tile_ptr[in.n_problems] = t, prob_part_off[in.n_problems] = part, prob_blkptr_off[in.n_problems] = bp;
    totals[0] = t, totals[1] = part, totals[2] = bp;
}

__global__ void __launch_bounds__(kCutThreads) k_setup_tiles_fill(SetupIn in, const int32_t* __restrict__ pt_ptr, const int32_t* __restrict__ prob_nfree,
    const int32_t* __restrict__ prob_tile_ptr, const int64_t* __restrict__ prob_part_off, const int64_t* __restrict__ prob_blkptr_off,
    WinTile* __restrict__ tiles)
{
    __shared__ int32_t s_buf[kCutStage];
    const int b = blockIdx.x;
    const int q0 = in.prob_pt0[b], hi = in.prob_pt0[b + 1];
    const int32_t* src = stage_pt_ptr(pt_ptr, q0, hi, s_buf);
    if (threadIdx.x) return;
    const int64_t nblk = (int64_t)prob_nfree[b] * (prob_nfree[b] + 1) / 2;
    WinTile proto = {};
    proto.prob = b, proto.cam0 = in.prob_cam0[b], proto.ncam = in.prob_cam0[b + 1] - in.prob_cam0[b], proto.nf = prob_nfree[b];
    proto.part_off = prob_part_off[b], proto.blkptr_off = prob_blkptr_off[b];
    cut_tiles<true>(src, q0, hi, proto, tiles + prob_tile_ptr[b], nblk * kItemPart, win_blkptr_entries(nblk));
}

// ---- 5a. per tile: keyframe-major slots and pair counts ----
__global__ void __launch_bounds__(kSetupThreads) k_setup_tile_slots(const WinTile* __restrict__ tiles, const int32_t* __restrict__ pt_ptr,
    const int32_t* __restrict__ obs_cam, const float2* __restrict__ obs_uv, const int32_t* __restrict__ cam_free,
    uint32_t* __restrict__ pt_mask, uint32_t* __restrict__ win_meta, float2* __restrict__ win_uv, uint16_t* __restrict__ win_slot,
    uint32_t* __restrict__ win_blkptr, int64_t* __restrict__ tile_pairs)
{
    __shared__ unsigned char s_cam[kWinObsCap];
    __shared__ unsigned short s_slot[kWinObsCap];
    __shared__ unsigned short s_obs0[kWinPtCap + 1];
    __shared__ int s_start[kWinMaxCams + 1];
    __shared__ signed char s_free[kWinMaxCams];
    __shared__ unsigned int s_cnt[kSmallMaxN / 6 * (kSmallMaxN / 6 + 1) / 2 + 1];
    __shared__ unsigned int s_mask[kWinPtCap];
    __shared__ unsigned char s_camof[kWinMaxCams];
    const WinTile t = tiles[blockIdx.x];
    const int tid = threadIdx.x;
    const int nblk = t.nf * (t.nf + 1) / 2;
    for (int o = tid; o < t.no; o += kSetupThreads) s_cam[o] = (unsigned char)(obs_cam[t.o0 + o] - t.cam0);
    for (int p = tid; p <= t.np; p += kSetupThreads) s_obs0[p] = (unsigned short)(pt_ptr[t.p0 + p] - t.o0);
    for (int c = tid; c < t.ncam; c += kSetupThreads) s_free[c] = (signed char)cam_free[t.cam0 + c];
    for (int k = tid; k <= nblk; k += kSetupThreads) s_cnt[k] = 0u;
    if (tid <= kWinMaxCams) s_start[tid] = 0;
    __syncthreads();
    for (int o = tid; o < t.no; o += kSetupThreads) atomicAdd(&s_start[s_cam[o] + 1], 1);
    __syncthreads();
    if (tid == 0)
        for (int c = 0; c < t.ncam; ++c) s_start[c + 1] += s_start[c];
    __syncthreads();
    // stable counting sort by keyframe: thread c walks the tile's observations in (point, pose) order
    if (tid < t.ncam) {
        int n = s_start[tid];
        for (int o = 0; o < t.no; ++o)
            if (s_cam[o] == tid) s_slot[o] = (unsigned short)n++;
    }
    __syncthreads();
    for (int p = tid; p < t.np; p += kSetupThreads) {
        const int a0 = s_obs0[p], a1 = s_obs0[p + 1];
        unsigned int m = 0u;
        for (int o = a0; o < a1; ++o) {
            const int c = s_cam[o], sl = s_slot[o];
            m |= 1u << c;
            win_meta[t.o0 + sl] = (unsigned)c | ((unsigned)p << 8);
            win_uv[t.o0 + sl] = obs_uv[t.o0 + o];
            win_slot[t.o0 + o] = (unsigned short)sl;
        }
        pt_mask[t.p0 + p] = m;
        s_mask[p] = m;
    }
    if (tid < t.ncam && s_free[tid] >= 0) s_camof[s_free[tid]] = (unsigned char)tid;
    __syncthreads();
    // pairs of S block (fa, fc) in this tile = points that both keyframes see (no duplicate edges on this path):
    // one thread per block counts them from the visibility masks
    for (int k = tid; k < nblk; k += kSetupThreads) {
        int fa = (int)((sqrtf(8.0f * (float)k + 1.0f) - 1.0f) * 0.5f);
        while ((fa + 1) * (fa + 2) / 2 <= k) ++fa;
        while (fa * (fa + 1) / 2 > k) --fa;
        const int fc = k - fa * (fa + 1) / 2;
        const unsigned int both = (1u << s_camof[fa]) | (1u << s_camof[fc]);
        unsigned int n = 0u;
        for (int p = 0; p < t.np; ++p) n += (s_mask[p] & both) == both;
        s_cnt[k + 1] = n;
    }
    __syncthreads();
    // counts -> global (entry k + 1 holds the count of block k; the scan happens in pass b)
    for (int k = tid; k <= nblk; k += kSetupThreads) win_blkptr[t.blkptr_off + k] = s_cnt[k];
    if (tid == 0) {
        int64_t n = 0;
        for (int k = 1; k <= nblk; ++k) n += s_cnt[k];
        tile_pairs[blockIdx.x] = n;
    }
}

// ---- 5b. per tile: block offsets and the pair lists, ordered by point ----
__global__ void __launch_bounds__(kSetupThreads) k_setup_tile_pairs(WinTile* __restrict__ tiles, const int64_t* __restrict__ tile_list_off,
    const int32_t* __restrict__ pt_ptr, const uint32_t* __restrict__ pt_mask, const uint16_t* __restrict__ win_slot,
    const int32_t* __restrict__ cam_free, uint32_t* __restrict__ win_blkptr, uint32_t* __restrict__ win_pairs)
{
    __shared__ unsigned short s_slot[kWinObsCap];
    __shared__ unsigned short s_obs0[kWinPtCap + 1];
    __shared__ unsigned int s_mask[kWinPtCap];
    __shared__ unsigned char s_fcam[kWinMaxCams];
    __shared__ unsigned int s_bp[kSmallMaxN / 6 * (kSmallMaxN / 6 + 1) / 2 + 1];
    WinTile t = tiles[blockIdx.x];
    const int64_t list_off = tile_list_off[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nblk = t.nf * (t.nf + 1) / 2;
    if (tid == 0) tiles[blockIdx.x].list_off = list_off;
    for (int o = tid; o < t.no; o += kSetupThreads) s_slot[o] = win_slot[t.o0 + o];
    for (int p = tid; p <= t.np; p += kSetupThreads) s_obs0[p] = (unsigned short)(pt_ptr[t.p0 + p] - t.o0);
    for (int p = tid; p < t.np; p += kSetupThreads) s_mask[p] = pt_mask[t.p0 + p];
    for (int c = tid; c < t.ncam; c += kSetupThreads) {
        const int f = cam_free[t.cam0 + c];
        if (f >= 0) s_fcam[f] = (unsigned char)c;
    }
    for (int k = tid; k <= nblk; k += kSetupThreads) s_bp[k] = win_blkptr[t.blkptr_off + k];
    __syncthreads();
    // the order the tile kernel takes the blocks in: by pair count, longest first (ties: lower index first).  Rank sort: the
    // rank of block k is the number of blocks that come before it.  Sixteen sub-warps take sixteen consecutive entries per pass
    // and the four of one warp wait for their longest list, so neighbours in this order should be of similar length.
    for (int k = tid; k < nblk; k += kSetupThreads) {
        const unsigned int ck = s_bp[k + 1];
        int rank = 0;
        for (int j = 0; j < nblk; ++j) {
            const unsigned int cj = s_bp[j + 1];
            rank += (cj > ck || (cj == ck && j < k)) ? 1 : 0;
        }
        win_blkptr[t.blkptr_off + nblk + 1 + rank] = (unsigned int)k;
    }
    __syncthreads();
    if (tid == 0) { // exclusive scan of the counts (nblk <= 378)
        unsigned int run = 0;
        for (int k = 0; k < nblk; ++k) {
            const unsigned int c = s_bp[k + 1];
            s_bp[k] = run;
            run += c;
        }
        s_bp[nblk] = run;
    }
    __syncthreads();
    for (int k = tid; k <= nblk; k += kSetupThreads) win_blkptr[t.blkptr_off + k] = s_bp[k];
    uint32_t* __restrict__ lst = win_pairs + list_off;
    for (int blk = warp; blk < nblk; blk += kSetupThreads / 32) { // one warp per S block
        int I = (int)((sqrtf(8.0f * (float)blk + 1.0f) - 1.0f) * 0.5f);
        while ((I + 1) * (I + 2) / 2 <= blk) ++I;
        while (I * (I + 1) / 2 > blk) --I;
        const int J = blk - I * (I + 1) / 2;
        const int cI = s_fcam[I], cJ = s_fcam[J];
        const unsigned int belowI = (1u << cI) - 1u, belowJ = (1u << cJ) - 1u;
        unsigned int pos = s_bp[blk];
        for (int base = 0; base < t.np; base += 32) {
            const int p = base + lane;
            const unsigned int m = p < t.np ? s_mask[p] : 0u;
            const bool both = ((m >> cI) & (m >> cJ) & 1u) != 0u;
            const unsigned int bal = __ballot_sync(0xffffffffu, both);
            if (both) {
                const int o0 = s_obs0[p];
                const unsigned int sa = s_slot[o0 + __popc(m & belowI)], sb = s_slot[o0 + __popc(m & belowJ)];
                lst[pos + __popc(bal & ((1u << lane) - 1u))] = sa | (sb << 16);
            }
            pos += __popc(bal);
        }
    }
}

__global__ void k_setup_fill_pt_prob(int32_t n_points, int32_t n_problems, const int32_t* __restrict__ prob_pt0, int32_t* __restrict__ pt_prob)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_points) return;
    int lo = 0, hi = n_problems; // largest b with prob_pt0[b] <= p
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (prob_pt0[mid] <= p) lo = mid;
        else hi = mid;
    }
    pt_prob[p] = lo;
}

inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

} // namespace

size_t setup_sort_temp_bytes(int64_t n_edges)
{
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
        (const unsigned int*)nullptr, (unsigned int*)nullptr, n_edges);
    size_t scan = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan, (const int64_t*)nullptr, (int64_t*)nullptr, (int)std::min<int64_t>(n_edges + 1, 0x7fffffff));
    return std::max(bytes, scan) + 256;
}

// bits of the pose index inside a sort key
static int key_cam_bits(const SetupIn& in)
{
    int bits = 1;
    while (bits < 31 && ((int64_t)1 << bits) < (int64_t)in.n_cams) ++bits;
    return bits;
}

int setup_stage0(const SetupIn& in, const SetupWork& w, cudaStream_t st)
{
    cudaMemsetAsync(w.err, 0, 4 * sizeof(int), st);
    cudaMemsetAsync(w.err + 2, 0xff, 2 * sizeof(int), st); // first offending edge: atomicMin target
    if (in.n_edges <= 0) return 0;
    k_setup_keys<<<cdiv(in.n_edges, 256), 256, 0, st>>>(in, key_cam_bits(in), w.keys, w.vals, w.err);
    return 1;
}

int setup_stage1(const SetupIn& in, const SetupWork& w, const SetupOut& o, cudaStream_t st)
{
    int launches = 0;
    const int cam_bits = key_cam_bits(in);
    if (in.n_edges > 0) {
        int end_bit = cam_bits + 1;
        while (end_bit < 64 && ((int64_t)1 << (end_bit - cam_bits)) < (int64_t)in.n_points) ++end_bit;
        size_t bytes = w.temp_bytes;
        cub::DeviceRadixSort::SortPairs(w.temp, bytes, w.keys, w.keys_sorted, w.vals, w.vals_sorted, in.n_edges, 0, end_bit, st);
        k_setup_gather<<<cdiv(in.n_edges, 256), 256, 0, st>>>(in, cam_bits, w.keys_sorted, w.vals_sorted, o.obs_cam, o.obs_uv, w.err);
        launches += 2;
    }
    k_setup_ptptr<<<cdiv((int64_t)in.n_points + 1, 256), 256, 0, st>>>(in.n_edges, in.n_points, cam_bits, w.keys_sorted, o.pt_ptr);
    k_setup_tiles_count<<<in.n_problems, kCutThreads, 0, st>>>(in, o.pt_ptr, w.prob_ntiles);
    k_setup_prob_offsets<<<1, 32, 0, st>>>(in, w.prob_ntiles, in.prob_nfree, o.prob_tile_ptr, o.prob_part_off, w.prob_blkptr_off, w.totals);
    k_setup_fill_pt_prob<<<cdiv(in.n_points, 256), 256, 0, st>>>(in.n_points, in.n_problems, in.prob_pt0, o.pt_prob);
    return launches + 4;
}

int setup_stage2(const SetupIn& in, const SetupWork& w, const SetupOut& o, int32_t n_tiles, cudaStream_t st)
{
    if (n_tiles <= 0) return 0;
    cudaMemsetAsync(w.tile_pairs, 0, sizeof(int64_t) * ((size_t)n_tiles + 1), st);
    k_setup_tiles_fill<<<in.n_problems, kCutThreads, 0, st>>>(in, o.pt_ptr, in.prob_nfree, o.prob_tile_ptr, o.prob_part_off, w.prob_blkptr_off, o.tiles);
    k_setup_tile_slots<<<n_tiles, kSetupThreads, 0, st>>>(o.tiles, o.pt_ptr, o.obs_cam, o.obs_uv, in.cam_free, o.pt_mask, o.win_meta, o.win_uv,
        o.win_slot, o.win_blkptr, w.tile_pairs);
    size_t bytes = w.temp_bytes;
    cub::DeviceScan::ExclusiveSum(w.temp, bytes, w.tile_pairs, w.tile_list_off, n_tiles + 1, st); // entry n_tiles = total
    return 3;
}

int setup_stage3(const SetupIn& in, const SetupWork& w, const SetupOut& o, int32_t n_tiles, cudaStream_t st)
{
    if (n_tiles <= 0) return 0;
    k_setup_tile_pairs<<<n_tiles, kSetupThreads, 0, st>>>(o.tiles, w.tile_list_off, o.pt_ptr, o.pt_mask, o.win_slot, in.cam_free, o.win_blkptr, o.win_pairs);
    return 1;
}

// ------------------------------------------------------------------------------------------
// general path (one large problem, B == 1): structure build on the device.  Same tables as the host build in viso_ba.cpp
// (build_general) leaves: observations in (point, pose) order, the pair list sorted by S block with the pairs of a block
// in point order (the radix sort is stable and the pairs are generated point by point), blk_ptr.
// ------------------------------------------------------------------------------------------
namespace {

__global__ void k_gen_present(int32_t n_points, const int32_t* __restrict__ pt_ptr, const int32_t* __restrict__ obs_cam,
    const int32_t* __restrict__ cam_free, uint8_t* __restrict__ present)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_points) return;
    const int s0 = pt_ptr[p], s1 = pt_ptr[p + 1];
    for (int a = s0; a < s1; ++a) {
        const int fa = cam_free[obs_cam[a]];
        if (fa < 0) continue;
        const size_t row = (size_t)fa * (fa + 1) / 2;
        for (int c = s0; c < a; ++c) {
            const int fc = cam_free[obs_cam[c]];
            if (fc >= 0 && !present[row + (size_t)fc]) present[row + (size_t)fc] = 1; // every writer writes 1
        }
    }
}

__global__ void k_gen_pair_counts(int32_t pt_begin, int32_t pt_end, const int32_t* __restrict__ pt_ptr, const int32_t* __restrict__ obs_cam,
    const int32_t* __restrict__ cam_free, int64_t* __restrict__ counts)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = pt_end - pt_begin;
    if (i > n) return;
    int64_t f = 0;
    if (i < n)
        for (int s = pt_ptr[pt_begin + i]; s < pt_ptr[pt_begin + i + 1]; ++s) f += cam_free[obs_cam[s]] >= 0 ? 1 : 0;
    counts[i] = f * (f + 1) / 2; // entry n: 0, becomes the total after the scan
}

__global__ void k_gen_pairs(int32_t pt_begin, int32_t pt_end, const int32_t* __restrict__ pt_ptr, const int32_t* __restrict__ obs_cam,
    const int32_t* __restrict__ cam_free, const int32_t* __restrict__ remap, const int64_t* __restrict__ offsets,
    unsigned int* __restrict__ keys, unsigned long long* __restrict__ vals)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pt_end - pt_begin) return;
    const int p = pt_begin + i;
    const int s0 = pt_ptr[p], s1 = pt_ptr[p + 1];
    int64_t q = offsets[i];
    for (int a = s0; a < s1; ++a) { // the order of build_general: a ascending, c ascending up to a
        const int fa = cam_free[obs_cam[a]];
        if (fa < 0) continue;
        for (int c = s0; c <= a; ++c) { // sorted by pose: fc <= fa
            const int fc = cam_free[obs_cam[c]];
            if (fc < 0) continue;
            const int tri = fa * (fa + 1) / 2 + fc;
            keys[q] = (unsigned int)(remap ? remap[tri] : tri);
            vals[q] = (unsigned long long)(unsigned int)a | ((unsigned long long)(unsigned int)c << 32);
            ++q;
        }
    }
}

__global__ void k_gen_blkptr(const unsigned int* __restrict__ keys, int64_t n_pairs, int32_t n_blocks, int64_t* __restrict__ blk_ptr)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n_blocks) return;
    int64_t lo = 0, hi = n_pairs; // first pair with key >= k
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (keys[mid] < (unsigned int)k) lo = mid + 1;
        else hi = mid;
    }
    blk_ptr[k] = lo;
}

} // namespace

int setup_sort_only(const SetupIn& in, const SetupWork& w, const SetupOut& o, cudaStream_t st)
{
    int launches = 0;
    const int cam_bits = key_cam_bits(in);
    if (in.n_edges > 0) {
        int end_bit = cam_bits + 1;
        while (end_bit < 64 && ((int64_t)1 << (end_bit - cam_bits)) < (int64_t)in.n_points) ++end_bit;
        size_t bytes = w.temp_bytes;
        cub::DeviceRadixSort::SortPairs(w.temp, bytes, w.keys, w.keys_sorted, w.vals, w.vals_sorted, in.n_edges, 0, end_bit, st);
        k_setup_gather<<<cdiv(in.n_edges, 256), 256, 0, st>>>(in, cam_bits, w.keys_sorted, w.vals_sorted, o.obs_cam, o.obs_uv, w.err);
        launches += 2;
    }
    k_setup_ptptr<<<cdiv((int64_t)in.n_points + 1, 256), 256, 0, st>>>(in.n_edges, in.n_points, cam_bits, w.keys_sorted, o.pt_ptr);
    k_setup_fill_pt_prob<<<cdiv(in.n_points, 256), 256, 0, st>>>(in.n_points, in.n_problems, in.prob_pt0, o.pt_prob);
    return launches + 2;
}

int gen_present(int32_t n_points, const int32_t* pt_ptr, const int32_t* obs_cam, const int32_t* cam_free, uint8_t* present, cudaStream_t st)
{
    if (n_points <= 0) return 0;
    k_gen_present<<<cdiv(n_points, 128), 128, 0, st>>>(n_points, pt_ptr, obs_cam, cam_free, present);
    return 1;
}

size_t gen_temp_bytes(int64_t n_points, int64_t n_pairs)
{
    size_t scan = 0, sort = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan, (const int64_t*)nullptr, (int64_t*)nullptr, (int)std::min<int64_t>(n_points + 1, 0x7fffffff));
    cub::DeviceRadixSort::SortPairs(nullptr, sort, (const unsigned int*)nullptr, (unsigned int*)nullptr, (const unsigned long long*)nullptr,
        (unsigned long long*)nullptr, std::max<int64_t>(n_pairs, 1));
    return std::max(scan, sort) + 256;
}

int gen_pair_offsets(int32_t pt_begin, int32_t pt_end, const int32_t* pt_ptr, const int32_t* obs_cam, const int32_t* cam_free,
                     int64_t* counts, void* temp, size_t temp_bytes, cudaStream_t st)
{
    const int n = pt_end - pt_begin;
    k_gen_pair_counts<<<cdiv((int64_t)n + 1, 256), 256, 0, st>>>(pt_begin, pt_end, pt_ptr, obs_cam, cam_free, counts);
    size_t bytes = temp_bytes;
    cub::DeviceScan::ExclusiveSum(temp, bytes, counts, counts, n + 1, st);
    return 2;
}

int gen_pairs(int32_t pt_begin, int32_t pt_end, const int32_t* pt_ptr, const int32_t* obs_cam, const int32_t* cam_free,
              const int32_t* remap, const int64_t* offsets, int64_t n_pairs, int32_t n_blocks, unsigned int* keys, unsigned int* keys_sorted,
              unsigned long long* vals, unsigned long long* vals_sorted, int64_t* blk_ptr, void* temp, size_t temp_bytes, cudaStream_t st)
{
    int launches = 0;
    const int n = pt_end - pt_begin;
    if (n > 0 && n_pairs > 0) {
        k_gen_pairs<<<cdiv(n, 128), 128, 0, st>>>(pt_begin, pt_end, pt_ptr, obs_cam, cam_free, remap, offsets, keys, vals);
        int end_bit = 1;
        while (end_bit < 32 && ((int64_t)1 << end_bit) < (int64_t)n_blocks) ++end_bit;
        size_t bytes = temp_bytes;
        cub::DeviceRadixSort::SortPairs(temp, bytes, keys, keys_sorted, vals, vals_sorted, n_pairs, 0, end_bit, st);
        launches += 2;
    }
    k_gen_blkptr<<<cdiv((int64_t)n_blocks + 1, 256), 256, 0, st>>>(keys_sorted, n_pairs, n_blocks, blk_ptr);
    return launches + 1;
}

} // namespace visoba
