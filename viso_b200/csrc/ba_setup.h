// ba_setup.h — device-side structure build of the window fast path (ba_setup.cu).
#ifndef VISO_BA_SETUP_H
#define VISO_BA_SETUP_H

#include "ba_device.h"

#include <stddef.h>

namespace visoba {

struct SetupIn { // everything is device memory
    int64_t n_edges;
    int32_t n_points, n_cams, n_problems;
    int32_t local_indices;          // 1: e_pose / e_point are local to the edge's problem (batched bulk call)
    const int32_t* e_pose;          // [Ne] caller order
    const int32_t* e_point;
    const float2* e_uv;
    const int64_t* edge_off;        // [B+1] (local_indices only)
    const int32_t* prob_cam0;       // [B+1]
    const int32_t* prob_pt0;        // [B+1]
    const int32_t* prob_nfree;      // [B]
    const int32_t* cam_free;        // [Nc]
};

struct SetupWork { // scratch, device memory
    unsigned long long *keys, *keys_sorted; // [Ne]
    unsigned int *vals, *vals_sorted;       // [Ne]
    void* temp;                             // cub
    size_t temp_bytes;
    int32_t* prob_ntiles;                   // [B]
    int64_t* prob_blkptr_off;               // [B+1]
    int64_t* tile_pairs;                    // [max_tiles + 1]
    int64_t* tile_list_off;                 // [max_tiles + 1]
    int64_t* totals;                        // [4] n_tiles, part doubles, blkptr entries, -
    int* err;                               // [4] status code, duplicate flag, first bad edge (64 bit)
};

struct SetupOut { // the structure the window kernels read (DevProblem fields)
    int32_t* pt_ptr;       // [Np+1]
    int32_t* pt_prob;      // [Np]
    int32_t* obs_cam;      // [Ne] (point, pose) order, global pose
    float2* obs_uv;        // [Ne]
    int32_t* prob_tile_ptr;// [B+1]
    int64_t* prob_part_off;// [B+1]
    WinTile* tiles;        // [n_tiles]
    uint32_t* pt_mask;     // [Np]
    uint32_t* win_meta;    // [Ne]
    float2* win_uv;        // [Ne]
    uint16_t* win_slot;    // [Ne]
    uint32_t* win_blkptr;  // [totals[2]]
    uint32_t* win_pairs;   // [n_pairs]
};

size_t setup_sort_temp_bytes(int64_t n_edges);
// stage 0: validate the raw edges and key them -> err (needs no gauge / free-pose information)
int setup_stage0(const SetupIn& in, const SetupWork& w, cudaStream_t st);
// stage 1: sort, CSR, tile counts  -> totals / err (read them back before stage 2)
int setup_stage1(const SetupIn& in, const SetupWork& w, const SetupOut& o, cudaStream_t st);
// stage 2: tiles, slots, pair counts, scan   -> tile_list_off[n_tiles] = number of pairs
int setup_stage2(const SetupIn& in, const SetupWork& w, const SetupOut& o, int32_t n_tiles, cudaStream_t st);
// stage 3: pair lists
int setup_stage3(const SetupIn& in, const SetupWork& w, const SetupOut& o, int32_t n_tiles, cudaStream_t st);

// ---- general path (one large problem): the same sort, then the pair lists by S block, on the device ----
// sort + CSR only (stage 1 without the tiles): obs_cam / obs_uv in (point, pose) order, pt_ptr, pt_prob
int setup_sort_only(const SetupIn& in, const SetupWork& w, const SetupOut& o, cudaStream_t st);
// present[I (I + 1) / 2 + J] = 1 for every pair of FREE poses I >= J some point sees (the structure of S; all points)
int gen_present(int32_t n_points, const int32_t* pt_ptr, const int32_t* obs_cam, const int32_t* cam_free, uint8_t* present, cudaStream_t st);
size_t gen_temp_bytes(int64_t n_points, int64_t n_pairs);
// counts[i] = pairs of point pt_begin + i (free observations f: f (f + 1) / 2), then the exclusive scan in place: counts[n] = total
int gen_pair_offsets(int32_t pt_begin, int32_t pt_end, const int32_t* pt_ptr, const int32_t* obs_cam, const int32_t* cam_free,
                     int64_t* counts, void* temp, size_t temp_bytes, cudaStream_t st);
// the pairs of the points [pt_begin, pt_end) in point order: key = S block (through remap when the storage is compact), value =
// observation a | observation c << 32; stable radix sort by block; blk_ptr[k] = first pair of block k
int gen_pairs(int32_t pt_begin, int32_t pt_end, const int32_t* pt_ptr, const int32_t* obs_cam, const int32_t* cam_free,
              const int32_t* remap, const int64_t* offsets, int64_t n_pairs, int32_t n_blocks, unsigned int* keys, unsigned int* keys_sorted,
              unsigned long long* vals, unsigned long long* vals_sorted, int64_t* blk_ptr, void* temp, size_t temp_bytes, cudaStream_t st);

} // namespace visoba
#endif
