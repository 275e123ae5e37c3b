// dnp_pairs.h -- internal interface of the all-pairs kernels (dnp_pairs.cu).
#pragma once

#include <atomic>
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "dipole.h"

namespace dnp {

// Fused epilogue description (K7).  Rows are addressed by r = tgt_idx ? tgt_idx[m] : m.
struct PairEpilogue {
    int mode = DNP_EPI_STORE;
    float* out = nullptr;
    int64_t ld_out = 3;
    float* normals = nullptr;
    int64_t ld_n = 6;
    const int64_t* tgt_idx = nullptr;
    const uint8_t* tgt_skip = nullptr;
    int32_t* nonfinite = nullptr;
};

// Sources chosen on the device (the device-driven greedy loop, dnp_drivers.cu): the source rows are
// src_idx[offsets[*patch] .. offsets[*patch + 1]); the S given to run_pairs is then only the upper bound the launch
// is planned for.  *patch < 0: no sources.
struct SourceSelect {
    const int32_t* patch = nullptr;
    const int64_t* offsets = nullptr;
    // the caller has already written the packed rows [0, plan.S_pad) at the start of the workspace
    // (pack_source_row below; null dipoles beyond the patch): run_pairs launches no pack kernel
    bool prepacked = false;
};

#ifdef __CUDACC__
// One source in the layout the all-pairs kernel streams through shared memory: 48 bytes,
// [x y z nx | ny nz -3nx -3ny | -3nz 0 0 0].
__device__ __forceinline__ void pack_source_row(const float v[6], float4* __restrict__ dst) {
    dst[0] = make_float4(v[0], v[1], v[2], v[3]);
    dst[1] = make_float4(v[4], v[5], v[3] * -3.f, v[4] * -3.f);
    dst[2] = make_float4(v[5] * -3.f, 0.f, 0.f, 0.f);
}
#endif

struct PairPlan {
    int tp = 2;            // target pairs per thread (2*tp targets)
    int nsplit = 1;        // source splits (grid.y)
    int tile_src = 256;    // sources per shared-memory tile (256, or 64 for calls too small to fill the machine)
    int tiles_total = 1;   // source tiles
    int m_tiles = 1;       // target tiles (grid.x)
    int64_t S_pad = 0;
    size_t packed_bytes = 0;
    size_t partial_bytes = 0;
};

// plan of the constant-window path (dnp_upairs.cu)
struct UPairPlan {
    int tp = 4;            // target pairs per thread (1024- or 256-target tiles)
    int lanes = 1;         // 1: everything on the caller's stream; 2 / 4: chunks rotate over that many lane streams
    int n_chunks = 1;      // launches
    int chunk = 4;         // sources per launch (multiple of 4, <= one window slot)
    int m_tiles = 1;       // target tiles
    int grid = 1;          // CTAs per launch
    int per = 4;           // (tile, source) iterations per CTA (multiple of 4)
    int ways = 2;          // partial row sets per lane
    bool whole_tiles = false;  // one target tile per CTA over the whole slot (many chunks x many tiles)
    int slot_floats = 8064;    // window floats per lane's slot (half or a quarter of the window)
    int64_t way_stride = 0;    // floats between row sets
    int64_t S_pad = 0;
    size_t packed_bytes = 0;
    size_t partial_bytes = 0;
};

int sm_count_cached();
PairPlan make_pair_plan(int64_t S, int64_t M, int sm_count, int force_tp, int force_nsplit);
size_t pairs_workspace_bytes(int64_t S, int64_t M);
size_t pairs_workspace_bytes_upper(int64_t S, int64_t M);   // covers every source count <= S

// potential == true: phi epilogue (STORE/ACCUM on a scalar per row).
// lattice_n > 0: targets are the nodes of the n^3 lattice (tgt ignored).
int run_pairs(bool potential, const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx,
              const float* tgt, int64_t M, int64_t ld_t, int lattice_n, float eps,
              const PairEpilogue& ep, void* workspace, size_t workspace_bytes, cudaStream_t stream,
              int force_tp = 0, int force_nsplit = 0, SourceSelect sel = SourceSelect());

// Sources through the device-resident constant window (dnp_upairs.cu): asynchronous, capturable.
bool use_ring_path(int64_t S, int64_t M, int lattice_n);
UPairPlan make_upair_plan(bool potential, int64_t S, int64_t M, int sm_count);
size_t upairs_workspace_bytes(int64_t S, int64_t M);
int run_upairs(bool potential, const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx, const float* tgt,
               int64_t M, int64_t ld_t, float eps, const PairEpilogue& ep, void* workspace, size_t workspace_bytes,
               cudaStream_t stream);
// One field slab per patch of a CSR (the slab schedule of the greedy drivers, dnp_slab.cu): see dnp_upairs.cu.
int upairs_slab_quantum();                 // patches are padded to a multiple of this many window rows
size_t upairs_slab_packed_bytes(int64_t total_pad);
int run_upairs_patch_slabs(const float* pts, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                           const int64_t* offsets_host, const int64_t* pad_off_dev, const int64_t* pad_off_host,
                           int64_t P, const int32_t* slot_host, const int64_t* first_target_host,
                           const int64_t* tgt_idx, int64_t M, float eps, float* packed, float* slabs,
                           int64_t slab_stride, cudaStream_t stream);
// Mean potential over the n^3 lattice with the sources in registers and the nodes as uniform operands (dnp_lattice.cu).
size_t lattice_mean_workspace_bytes();
int run_lattice_mean(const float* src, int64_t S, int64_t ld_s, int n, float* mean_out, void* workspace,
                     size_t workspace_bytes, cudaStream_t stream);
// fixed-order sum of `nsplit` partial row sets + the fused epilogue; set_stride = floats between row sets
// (0: M rows of 1 / 3 floats)
int run_split_reduce(bool potential, const float* partial, int nsplit, int64_t M, const PairEpilogue& ep,
                     cudaStream_t stream, int64_t set_stride = 0);
// records one event of the armed per-thread kernel timer (no-op otherwise)
void timer_mark(cudaStream_t stream);

// out[0] = scale * sum_i in[i*stride], single CTA, fixed order.
int run_sum(const float* in, int64_t n, int64_t stride, float scale, float* out, cudaStream_t stream);

void profile_begin();
int profile_end(float* ms_out, int capacity, int* n_out);

int64_t launch_count();
void count_launch(int n);

}  // namespace dnp
