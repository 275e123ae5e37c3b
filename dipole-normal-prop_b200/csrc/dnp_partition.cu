// dnp_partition.cu -- voxel partition of a point cloud into patches.
//
// Replaces util._divide_pc (util.py:110-150): the reference walks all (n_part + 1)^3 cells in Python and
// builds three boolean masks over the whole cloud per cell -- 74 088 passes over the points at
// `--number_parts 41`, minutes on a million points.  Here every point computes its cell once
// (cell_key_kernel), one stable LSD radix sort by cell key (cub::DeviceRadixSort, a library primitive)
// groups them, and the run heads give the CSR layout the patch kernels (K3 / K4) consume:
//   perm[offsets[c] .. offsets[c+1])  = rows of the c-th non-empty cell, ascending (stable sort),
//   ijk[c]                            = its grid coordinates, cells in lexicographic (i, j, k) order
// -- exactly the order of the reference's nested loops.  HBM bound: 12 B read per point for the keys, then
// 8 B per point per radix pass.
//
// Cell membership follows the reference to the bit: cell t of an axis holds l_t < x <= l_t + edge with
// l_t = edge * t + lo evaluated in double and rounded to float (torch casts the Python scalar to the
// tensor's dtype before comparing); the host builds both bound tables that way and the kernel only
// compares.  Points outside every cell (or NaN) belong to no patch, as in the reference.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"

namespace dnp {

namespace {

constexpr int kMaxParts = 255;                     // n_part + 1 <= 256 cells per axis

struct CellBounds {
    float lo[kMaxParts + 1];
    float hi[kMaxParts + 1];
    float origin, inv_edge;
    int cells;                                     // n_part + 1
};

__device__ __forceinline__ int axis_cell(const CellBounds& b, float x) {
    // estimate, then settle on the exact float bounds (first matching cell in ascending order)
    if (!(fabsf(x) <= 3.402823466e38f)) return -1;     // NaN / Inf: in no cell
    int t = (int)floorf((x - b.origin) * b.inv_edge);
    t = max(1, min(b.cells - 2, t));
#pragma unroll
    for (int d = -1; d <= 1; ++d) {
        const int c = t + d;
        if (c >= 0 && c < b.cells && b.lo[c] < x && x <= b.hi[c]) return c;
    }
    // estimate off by more than one only at the far ends of a huge range: fall back to a scan
    for (int c = 0; c < b.cells; ++c)
        if (b.lo[c] < x && x <= b.hi[c]) return c;
    return -1;
}

__global__ void cell_key_kernel(const float* __restrict__ pts, int64_t N, int64_t ld, const __grid_constant__ CellBounds b,
                                uint32_t sentinel, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float* p = pts + i * ld;
    const int cx = axis_cell(b, __ldg(p)), cy = axis_cell(b, __ldg(p + 1)), cz = axis_cell(b, __ldg(p + 2));
    keys[i] = (cx < 0 || cy < 0 || cz < 0) ? sentinel : (uint32_t)((cx * b.cells + cy) * b.cells + cz);
    vals[i] = (uint32_t)i;
}

__global__ void head_flags_kernel(const uint32_t* __restrict__ keys, int64_t N, uint32_t sentinel,
                                  uint8_t* __restrict__ flags) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const uint32_t k = keys[i];
    flags[i] = (k != sentinel && (i == 0 || keys[i - 1] != k)) ? 1 : 0;
}

// offsets[c] = heads[c], offsets[n_cells] = number of points inside the box; ijk[c] from the key
__global__ void finish_csr_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals, int64_t N,
                                  uint32_t sentinel, int cells, const int32_t* __restrict__ heads,
                                  const int32_t* __restrict__ n_cells_ptr, int64_t* __restrict__ perm,
                                  int64_t* __restrict__ offsets, int32_t* __restrict__ ijk, int64_t* __restrict__ counts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int n_cells = *n_cells_ptr;
    if (i < N) perm[i] = (int64_t)vals[i];
    if (i < n_cells) {
        const int32_t h = heads[i];
        offsets[i] = h;
        const uint32_t k = keys[h];
        ijk[3 * i] = (int32_t)(k / (uint32_t)(cells * cells));
        ijk[3 * i + 1] = (int32_t)((k / (uint32_t)cells) % (uint32_t)cells);
        ijk[3 * i + 2] = (int32_t)(k % (uint32_t)cells);
    }
    if (i == 0) {
        // points inside the box = first position holding the sentinel (keys are sorted): binary search
        int64_t lo = 0, hi = N;
        while (lo < hi) {
            const int64_t mid = (lo + hi) / 2;
            if (keys[mid] == sentinel) hi = mid; else lo = mid + 1;
        }
        offsets[n_cells] = lo;
        counts[0] = n_cells;
        counts[1] = lo;
    }
}

inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }

struct PartitionLayout {
    size_t keys_a, keys_b, vals_a, vals_b, flags, heads, n_sel, cub_temp, total;
    size_t cub_bytes;
};

PartitionLayout partition_layout(int64_t N) {
    PartitionLayout L;
    const size_t n = (size_t)(N < 1 ? 1 : N);
    size_t sort_bytes = 0, select_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 32);
    cub::DeviceSelect::Flagged(nullptr, select_bytes, thrust::counting_iterator<int32_t>(0), (const uint8_t*)nullptr,
                               (int32_t*)nullptr, (int32_t*)nullptr, (int)n);
    L.cub_bytes = sort_bytes > select_bytes ? sort_bytes : select_bytes;
    size_t off = 0;
    L.keys_a = off; off += up256(n * 4);
    L.keys_b = off; off += up256(n * 4);
    L.vals_a = off; off += up256(n * 4);
    L.vals_b = off; off += up256(n * 4);
    L.flags = off; off += up256(n);
    L.heads = off; off += up256(n * 4);
    L.n_sel = off; off += 256;
    L.cub_temp = off; off += up256(L.cub_bytes);
    L.total = off;
    return L;
}

}  // namespace

size_t voxel_partition_workspace_bytes(int64_t N) { return partition_layout(N).total; }

int run_voxel_partition(const float* pts, int64_t N, int64_t ld, int n_part, double range_lo, double range_hi,
                        int64_t* perm, int64_t* offsets, int32_t* ijk, int64_t* counts, void* workspace,
                        size_t workspace_bytes, cudaStream_t stream) {
    if (n_part < 1 || n_part > kMaxParts) {
        set_error("n_part must be in [1, %d] (got %d)", kMaxParts, n_part);
        return DNP_ERR_INVALID;
    }
    if (N >= (int64_t)0x7fffffff) {
        set_error("N too large for one partition call");
        return DNP_ERR_INVALID;
    }
    const PartitionLayout L = partition_layout(N);
    if (!workspace || workspace_bytes < L.total) {
        set_error("workspace too small: need %zu bytes (dnp_voxel_partition_workspace_bytes), got %zu", L.total,
                  workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    if (N == 0) {
        DNP_CHECK_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(int64_t), stream));
        DNP_CHECK_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int64_t), stream));
        return DNP_OK;
    }
    // bounds exactly as util.py:126-136 builds them: double arithmetic, compared in float
    CellBounds b;
    const double edge = (range_hi - range_lo) / (double)n_part;
    b.cells = n_part + 1;
    for (int t = 0; t < b.cells; ++t) {
        const double l = edge * (double)t + range_lo;
        b.lo[t] = (float)l;
        b.hi[t] = (float)(l + edge);
    }
    b.origin = (float)range_lo;
    b.inv_edge = (float)(1.0 / edge);
    const uint32_t sentinel = (uint32_t)b.cells * b.cells * b.cells;           // one past the last cell key
    int end_bit = 1;
    while ((1u << end_bit) <= sentinel && end_bit < 32) ++end_bit;

    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    uint32_t* keys_a = reinterpret_cast<uint32_t*>(ws + L.keys_a);
    uint32_t* keys_b = reinterpret_cast<uint32_t*>(ws + L.keys_b);
    uint32_t* vals_a = reinterpret_cast<uint32_t*>(ws + L.vals_a);
    uint32_t* vals_b = reinterpret_cast<uint32_t*>(ws + L.vals_b);
    uint8_t* flags = ws + L.flags;
    int32_t* heads = reinterpret_cast<int32_t*>(ws + L.heads);
    int32_t* n_sel = reinterpret_cast<int32_t*>(ws + L.n_sel);
    void* cub_temp = ws + L.cub_temp;
    size_t cub_bytes = L.cub_bytes;

    const int threads = 256;
    const unsigned blocks = (unsigned)((N + threads - 1) / threads);
    cell_key_kernel<<<blocks, threads, 0, stream>>>(pts, N, ld, b, sentinel, keys_a, vals_a);
    DNP_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(cub_temp, cub_bytes, keys_a, keys_b, vals_a, vals_b, (int)N, 0, end_bit,
                                                   stream));
    head_flags_kernel<<<blocks, threads, 0, stream>>>(keys_b, N, sentinel, flags);
    cub_bytes = L.cub_bytes;
    DNP_CHECK_CUDA(cub::DeviceSelect::Flagged(cub_temp, cub_bytes, thrust::counting_iterator<int32_t>(0), flags, heads,
                                              n_sel, (int)N, stream));
    finish_csr_kernel<<<blocks, threads, 0, stream>>>(keys_b, vals_b, N, sentinel, b.cells, heads, n_sel, perm, offsets,
                                                      ijk, counts);
    count_launch(3);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
