// dnp_scores.h -- internal interface of the score / driver kernels
// (dnp_scores.cu, dnp_points.cu, dnp_microbench.cu).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace dnp {

int run_edge_interactions_batched(const float* pts, int64_t ld, const int64_t* perm, const int64_t* offsets,
                                  int64_t P, const int64_t* tgt_perm, const int64_t* tgt_offsets,
                                  const int64_t* edges, int64_t n_edges, float eps, float* out,
                                  cudaStream_t stream);

// Optional tail of the arg-max for the device-driven greedy loop (dnp_drivers.cu): cur = {patch, flip, previous
// patch} is advanced to the winner on the device, order / value (device, optional) record the trace at `step`.
struct PatchAdvance {
    int32_t* cur = nullptr;
    int32_t* order = nullptr;
    float* value = nullptr;
    int step = 0;
};

#ifdef __CUDACC__
// Block-wide first maximum of |vals[p]| over the active patches, as ONE unsigned key per candidate:
// (|v| bits << 32) | inverted index, so that the larger key is the larger |v| and, among equals, the smaller
// index (torch's abs().argmax()).  Returned to thread 0 (0 = nothing active); warp_best: 32 shared slots.
__device__ __forceinline__ unsigned long long block_first_absmax(const float* __restrict__ vals,
                                                                 const uint8_t* __restrict__ active, int64_t P,
                                                                 unsigned long long* warp_best) {
    unsigned long long best = 0ull;
    for (int64_t p = threadIdx.x; p < P; p += blockDim.x) {
        if (active && !active[p]) continue;
        const unsigned long long k =
            ((unsigned long long)__float_as_uint(fabsf(vals[p])) << 32) | (0xffffffffu - (uint32_t)p);
        if (k > best) best = k;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        if (other > best) best = other;
    }
    if ((threadIdx.x & 31) == 0) warp_best[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x < 32) {
        best = threadIdx.x < (blockDim.x >> 5) ? warp_best[threadIdx.x] : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            if (other > best) best = other;
        }
    }
    return best;
}
#endif

// Multi-GPU greedy loop (dnp_comm.cu): every patch is OWNED by one rank, which holds its rows as targets and computes
// its interaction; `owned` (device, P bytes) zeroes the others and `combine` sums the vectors of all ranks on the stream
// (x + 0 is exact, so every rank sees the owner's value bit for bit) before anything reads them.
struct ShardHook {
    const uint8_t* owned = nullptr;
    int (*combine)(void* ctx, float* vals, int64_t n, cudaStream_t stream) = nullptr;
    void* ctx = nullptr;
    int world = 1;
};

int run_patch_interactions(const float* E, int64_t ld_e, const float* normals, int64_t ld_n, const int64_t* perm,
                           const int64_t* offsets, int64_t P, const uint8_t* active, float* out,
                           int32_t* best_idx, float* best_val, cudaStream_t stream, PatchAdvance adv = PatchAdvance(),
                           const ShardHook* shard = nullptr);

// The whole greedy patch loop as one persistent cooperative kernel (dnp_loop.cu), for small clouds with disjoint
// patches covering the targets; *taken = 0 when the preconditions do not hold (the caller runs its own schedule).
size_t patch_loop_workspace_bytes(int64_t N, int64_t P);
int run_patch_loop_persistent(float* pts, int64_t N, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                              const int64_t* offsets_host, int64_t P, const int64_t* rest_perm,
                              const int64_t* rest_offsets_host, const int64_t* tgt_idx, int64_t n_tgt,
                              int64_t start_patch, float eps, int diffuse, float* E, int64_t ld_e, int32_t* order_host,
                              float* value_host, void* workspace, size_t workspace_bytes, cudaStream_t stream,
                              int* taken);

// disjoint patches (perm[0..T) distinct rows of [0, N)) and no rest row among them?  flags: N ints, counts: 64 ints of
// scratch; one 12-byte read on the stream.  *ok = 0: overlapping input (not an error); rows out of range: DNP_ERR_INVALID.
int check_patch_cover(const int64_t* perm, int64_t T, int64_t N, const int64_t* rest_rows, int64_t R, int* flags,
                      int* counts, cudaStream_t stream, int* ok);

// The greedy patch loop as P independent per-patch field launches + a loop on the P x P interaction matrix
// (dnp_slab.cu); *taken = 0 when the preconditions do not hold.
// world: ranks the slabs are split over (every rank holds ~P / world of them: the budget is per GPU)
bool slab_schedule_fits(int64_t P, int64_t T, int64_t max_patch, int world = 1);
bool slab_schedule_pays(int64_t P, int64_t T, int64_t max_patch, int world = 1);
size_t patch_slab_workspace_bytes(int64_t N, int64_t P, int64_t T, int64_t n_slabs = -1);
int run_patch_loop_slab(float* pts, int64_t N, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                        const int64_t* offsets_host, int64_t P, const int64_t* rest_perm,
                        const int64_t* rest_offsets_host, const int64_t* tgt_idx, int64_t n_tgt, int64_t start_patch,
                        float eps, int diffuse, float* E, int64_t ld_e, int32_t* order_host, float* value_host,
                        void* workspace, size_t workspace_bytes, cudaStream_t stream, const ShardHook* shard,
                        int* taken);

int run_patch_covariance(const float* pts, int64_t ld, const int64_t* perm, const int64_t* offsets, int64_t P,
                         double* cov, cudaStream_t stream);

int run_field_sign(const float* E, int64_t ld_e, float* normals, int64_t ld_n, const int64_t* idx, int64_t n,
                   int ge, cudaStream_t stream);

int64_t points_max_n();        // single-cluster capacity
int64_t points_grid_max_n();   // cooperative-grid capacity (one CTA per SM)
size_t points_workspace_bytes(int64_t N);
int run_points_propagate(float* pts, int64_t N, int64_t ld, int64_t start, float eps, int diffuse,
                         int32_t* order, uint8_t* flipped, void* workspace, size_t workspace_bytes,
                         cudaStream_t stream);

// greedy patch propagation loop (dnp_drivers.cu)
size_t patch_propagate_workspace_bytes(int64_t N, int64_t P, int64_t max_patch, int64_t n_tgt);
// the sharded loop: n_tgt = rows of this rank's patches, patch_rows = offsets[P], n_owned = patches this rank owns
size_t patch_propagate_sharded_workspace_bytes(int64_t N, int64_t P, int64_t max_patch, int64_t n_tgt,
                                               int64_t patch_rows, int64_t n_owned, int world);
int run_patch_propagate(float* pts, int64_t N, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                        const int64_t* offsets_host, int64_t P, const int64_t* rest_perm,
                        const int64_t* rest_offsets_host, const int64_t* tgt_idx, int64_t n_tgt, int64_t start_patch,
                        float eps, int diffuse, float* E, int64_t ld_e, int32_t* order_host, float* value_host,
                        void* workspace, size_t workspace_bytes, cudaStream_t stream, const ShardHook* shard = nullptr);

int64_t points_batched_max_n();
int run_points_propagate_batched(float* pts, int64_t ld, const int64_t* offsets, int64_t P, int64_t max_patch,
                                 float eps, int diffuse, cudaStream_t stream);
int run_patch_lattice_sign(float* pts, int64_t ld, const int64_t* offsets, int64_t P, int n_lat, float* mean_out,
                           int flip, cudaStream_t stream);

// "xie" reflected-normal interaction (dnp_xie.cu)
int run_xie_matrix(bool field, const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t,
                   float C, const double* rad2, float* out, int64_t ld_o, cudaStream_t stream);
int64_t xie_max_n(bool knn);
int run_knn_radius2(const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t, int k,
                    double* out, cudaStream_t stream);
size_t xie_workspace_bytes(int64_t N, int64_t T);
int run_xie_propagate(const float* pts, int64_t N, int64_t ld, const int64_t* order, int64_t T, float C,
                      const double* rad2, int diffuse, uint8_t* flips, float* inter, void* workspace,
                      size_t workspace_bytes, cudaStream_t stream);

// voxel partition (dnp_partition.cu)
size_t voxel_partition_workspace_bytes(int64_t N);
int run_voxel_partition(const float* pts, int64_t N, int64_t ld, int n_part, double range_lo, double range_hi,
                        int64_t* perm, int64_t* offsets, int32_t* ijk, int64_t* counts, void* workspace,
                        size_t workspace_bytes, cudaStream_t stream);

// patch-graph flip solve (dnp_flip.cu)
size_t flip_solve_workspace_bytes(int64_t n, int exact_max_n, int replicas);
int run_flip_solve(const double* A, const double* B, int64_t n, int64_t ld, int exact_max_n, int replicas, int sweeps,
                   uint64_t seed, uint8_t* x_out, double* obj_out, int32_t* status_out, void* workspace,
                   size_t workspace_bytes, cudaStream_t stream);

// text IO + wire format (dnp_textio.cu)
size_t xyz_parse_workspace_bytes(int64_t n_bytes, int64_t max_lines);
int run_xyz_parse(const char* txt, int64_t n_bytes, int64_t max_lines, float* out, int64_t max_rows, int64_t* counts,
                  void* workspace, size_t workspace_bytes, cudaStream_t stream);
size_t xyz_format_workspace_bytes(int64_t n_values);
int run_xyz_format(const float* pts, int64_t n_rows, int cols, int64_t ld_row, int64_t ld_col, char* out,
                   int64_t capacity, int64_t* n_bytes_out, void* workspace, size_t workspace_bytes, cudaStream_t stream);
int run_wire_decode(const double* xyz, int64_t n, float* out, int64_t ld, cudaStream_t stream);
int run_wire_encode(const float* pts, int64_t n, int64_t ld, double* out, cudaStream_t stream);

int run_microbench(int kind, int iters, float* ms_out, double* ops_out, cudaStream_t stream);

}  // namespace dnp
