// dnp_lattice.cu -- the global-orientation test: mean dipole potential over the n^3 lattice of util.gen_grid
// (field_utils.py:7-9, util.py:26-36), for callers that need only the MEAN (every pipeline: they test `< 0`).
//
// The generic all-pairs kernel keeps TARGETS in registers and streams sources; with 1000 lattice nodes as targets
// that fills 4 CTAs' worth of threads and needs a source split + a reduction pass to use the machine (round 1:
// 0.68 ms per million sources, 0.375 of the FP32 roofline).  Here the roles are swapped:
//   * every thread keeps 2*TP SOURCES (position + moment) in registers, packed two per 64-bit register pair;
//   * the lattice node is the warp-uniform operand.  Nodes are (c[x], c[y], c[z]) with an n-entry coordinate table
//     computed on the host with the reference's float operations and passed by value (LDCU -> uniform registers);
//   * the three nested node loops let everything that depends on x only (rx, rx^2, nx rx) and on (x, y) only
//     (rx^2 + ry^2, nx rx + ny ry) be hoisted: the innermost z step costs 6 packed FP32 operations + 2 MUFU.RSQ
//     per (2 sources x 1 node) instead of 12 + 2, which makes the kernel SFU-bound (1 MUFU per pair: 4.65e12
//     pairs/s ceiling on B200) instead of FP32-bound (3.1e12);
//   * only the grand total is needed, so a thread simply keeps one running sum over all nodes and its sources:
//     no per-node reduction at all.  The (source tile, x) items are dealt to the CTAs in equal contiguous shares
//     (host-computed start table, see dnp_upairs.cu for why), one wave, then a fixed-order block reduction and a
//     fixed-order sum of the per-CTA partials: deterministic, no atomics.
// A source that sits exactly on a node contributes 0 (n.R = 0 exactly, clamped rsqrt), as in dnp_pairs.cu.
#include <stdlib.h>

#include <atomic>

#include "dnp_common.cuh"
#include "dnp_pairs.h"

namespace dnp {

namespace {

constexpr int kLBlock = 128;
constexpr int kLMaxGrid = 148 * 8 + 16;
// r2 is formed as rx * rx + kLTiny + ..., so that rsqrt stays finite (<= 1e12, cubed <= 1e36) and a source exactly on
// a node adds (n.R) u^3 = 0 * finite = 0 with no clamp instruction (dnp_pairmath.cuh: every non-FMA instruction is an
// issue cycle; without the two FMNMX the step is purely MUFU-bound)
constexpr float kLTiny = 1.0e-24f;

struct LatticeTable {
    float c[256];                                  // node coordinate per index, util.py:26-36
};
struct LatticeShares {
    int tile[kLMaxGrid];                           // first source tile of CTA b
    int x0[kLMaxGrid];                             // first lattice x index inside it
};

// N > 0: lattice side known at compile time (the z loop unrolls); N == 0: run-time side n.
template <int N, int TP>
__global__ void __launch_bounds__(kLBlock, (TP >= 4 ? 3 : 6))
lattice_sum_kernel(const __grid_constant__ LatticeTable tab, const __grid_constant__ LatticeShares shares, int per,
                   int m_tiles, int n_rt, const float* __restrict__ src, int64_t S, int64_t ld_s,
                   float* __restrict__ partial) {
    const int n = N > 0 ? N : n_rt;
    const int tid = threadIdx.x;
    const f32x2 zero2 = pack2(0.f, 0.f);
    const f32x2 tiny2 = pack2(tab.c[255], tab.c[255]);          // kLTiny, delivered through the constant bank
    int tile = shares.tile[blockIdx.x];
    int x = shares.x0[blockIdx.x];
    int left = per;                                // (tile, x) items this CTA still owns
    f32x2 total[TP];
#pragma unroll
    for (int j = 0; j < TP; ++j) total[j] = zero2;

    while (left > 0 && tile < m_tiles) {
        int x_end = x + left;
        if (x_end > n) x_end = n;
        // ---- this tile's sources: lane a of pair j is row tile_base + 2j*kLBlock + tid, lane b the next kLBlock
        const int64_t tile_base = (int64_t)tile * (kLBlock * 2 * TP);
        f32x2 px[TP], py[TP], pz[TP], mx[TP], my[TP], mz[TP];
#pragma unroll
        for (int j = 0; j < TP; ++j) {
            float v[2][6];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int64_t s = tile_base + (int64_t)(2 * j + h) * kLBlock + tid;
#pragma unroll
                for (int c = 0; c < 6; ++c) v[h][c] = 0.f;                   // past the end: a null dipole at the origin
                if (s < S) {
                    const float* p = src + s * ld_s;
#pragma unroll
                    for (int c = 0; c < 6; ++c) v[h][c] = __ldg(p + c);
                }
            }
            px[j] = pack2(v[0][0], v[1][0]); py[j] = pack2(v[0][1], v[1][1]); pz[j] = pack2(v[0][2], v[1][2]);
            mx[j] = pack2(v[0][3], v[1][3]); my[j] = pack2(v[0][4], v[1][4]); mz[j] = pack2(v[0][5], v[1][5]);
        }

        for (int xi = x; xi < x_end; ++xi) {
            const float qx = tab.c[xi];
            const f32x2 qx2 = pack2(qx, qx);
            f32x2 rx2[TP], dx[TP];
#pragma unroll
            for (int j = 0; j < TP; ++j) {
                const f32x2 rx = sub2(px[j], qx2);                           // R = source - node (field_utils.py:47)
                rx2[j] = fma2(rx, rx, tiny2);
                dx[j] = mul2(mx[j], rx);
            }
#pragma unroll 1
            for (int yi = 0; yi < n; ++yi) {
                const float qy = tab.c[yi];
                const f32x2 qy2 = pack2(qy, qy);
                f32x2 rxy[TP], dxy[TP];
#pragma unroll
                for (int j = 0; j < TP; ++j) {
                    const f32x2 ry = sub2(py[j], qy2);
                    rxy[j] = fma2(ry, ry, rx2[j]);
                    dxy[j] = fma2(my[j], ry, dx[j]);
                }
                f32x2 phi[TP];
#pragma unroll
                for (int j = 0; j < TP; ++j) phi[j] = zero2;
#pragma unroll(N > 0 ? N : 2)
                for (int zi = 0; zi < n; ++zi) {
                    const float qz = tab.c[zi];
                    const f32x2 qz2 = pack2(qz, qz);
#pragma unroll
                    for (int j = 0; j < TP; ++j) {
                        const f32x2 rz = sub2(pz[j], qz2);
                        const f32x2 r2 = fma2(rz, rz, rxy[j]);
                        const f32x2 d = fma2(mz[j], rz, dxy[j]);
                        float r2a, r2b;
                        unpack2(r2, r2a, r2b);
                        const f32x2 u = pack2(rsqrt_sfu(r2a), rsqrt_sfu(r2b));      // finite: r2 >= kLTiny
                        phi[j] = fma2(d, mul2(mul2(u, u), u), phi[j]);
                    }
                }
#pragma unroll
                for (int j = 0; j < TP; ++j) total[j] = add2(total[j], phi[j]);   // second-level sum per (x, y) row
            }
        }
        left -= x_end - x;
        ++tile;
        x = 0;
    }

    // ---- block total, fixed order: lanes of a pair, pairs, warp shuffle, warps
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float a, b;
        unpack2(total[j], a, b);
        acc += a + b;
    }
    // a non-finite thread total would poison the mean: the reference zeroes non-finite node potentials
    // (field_utils.py:53-54); coincident source/node pairs already contribute 0, so this only catches NaN / Inf INPUT
    if (!(fabsf(acc) <= 3.402823466e38f)) acc = 0.f;
    __shared__ float warp_part[kLBlock / 32];
    acc = warp_sum(acc);
    if ((tid & 31) == 0) warp_part[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kLBlock / 32; ++w) t += warp_part[w];
        partial[blockIdx.x] = t;
    }
}


// Scratch of dnp_lattice_mean_potential_noalloc_f32 (below): the per-CTA partial sums, one float per CTA of the largest
// grid the kernel is ever launched with.  File scope: every device that loads the library holds these 4.9 KB until
// the process exits.  In HOST code this name is a stand-in, not a device address -- the launcher gets the address of
// the current device's copy from cudaGetSymbolAddress.
constexpr size_t kStaticPartialFloats = kLMaxGrid;
static __device__ __align__(256) float g_static_partials[kStaticPartialFloats];

}  // namespace

size_t lattice_mean_workspace_bytes() { return (size_t)kLMaxGrid * sizeof(float); }

// mean_out[0] = (1 / n^3) sum_nodes sum_sources potential.  workspace: lattice_mean_workspace_bytes().
int run_lattice_mean(const float* src, int64_t S, int64_t ld_s, int n, float* mean_out, void* workspace,
                     size_t workspace_bytes, cudaStream_t stream) {
    if (n < 1 || n > 255) {
        set_error("lattice side out of range");
        return DNP_ERR_INVALID;
    }
    if (!workspace || workspace_bytes < lattice_mean_workspace_bytes()) {
        set_error("workspace too small: need %zu bytes, got %zu", lattice_mean_workspace_bytes(), workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    float* partial = reinterpret_cast<float*>(workspace);
    LatticeTable tab;
    for (int i = 0; i < 256; ++i) {
        // util.py:36: (index.float() / n - 0.5) * 2, float32 operations in that order
        volatile float q = (float)i / (float)n;
        volatile float h = q - 0.5f;
        tab.c[i] = i < n ? h * 2.f : 0.f;
    }
    tab.c[255] = kLTiny;                               // n <= 255 nodes per axis use the table; the last slot carries the constant
    // few sources: 2 per thread (256-source tiles) so that the tiles x n items cover the machine
    const int sms = sm_count_cached();
    const int tp = S >= (int64_t)sms * 3 * kLBlock * 8 ? 4 : 1;
    const int64_t tile_src = (int64_t)kLBlock * 2 * tp;
    const int64_t m_tiles = S > 0 ? (S + tile_src - 1) / tile_src : 0;
    const int64_t items = m_tiles * n;
    int64_t grid = (int64_t)sms * (tp == 4 ? 3 : 6);
    if (grid > kLMaxGrid) grid = kLMaxGrid;
    if (grid > items) grid = items;
    if (grid < 1) grid = 1;
    const int per = (int)((items + grid - 1) / grid);
    grid = items > 0 ? (items + per - 1) / per : 1;
    LatticeShares sh;
    for (int b = 0; b < grid; ++b) {
        const int64_t it = (int64_t)b * per;
        sh.tile[b] = (int)(it / n);
        sh.x0[b] = (int)(it % n);
    }
    const int per_arg = items > 0 ? per : 0;
    if (n == 10) {
        if (tp == 4) lattice_sum_kernel<10, 4><<<(unsigned)grid, kLBlock, 0, stream>>>(tab, sh, per_arg, (int)m_tiles, n, src, S, ld_s, partial);
        else lattice_sum_kernel<10, 1><<<(unsigned)grid, kLBlock, 0, stream>>>(tab, sh, per_arg, (int)m_tiles, n, src, S, ld_s, partial);
    } else {
        if (tp == 4) lattice_sum_kernel<0, 4><<<(unsigned)grid, kLBlock, 0, stream>>>(tab, sh, per_arg, (int)m_tiles, n, src, S, ld_s, partial);
        else lattice_sum_kernel<0, 1><<<(unsigned)grid, kLBlock, 0, stream>>>(tab, sh, per_arg, (int)m_tiles, n, src, S, ld_s, partial);
    }
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    const double nodes = (double)n * n * n;
    return run_sum(partial, grid, 1, (float)(1.0 / nodes), mean_out, stream);
}

}  // namespace dnp

// ------------------------------------------------------------------ allocation-free entry point
// No counterpart in the reference: added for callers that cannot allocate
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Same kernels as the workspace variant.
extern "C" int dnp_lattice_mean_potential_noalloc_f32(const float* src, int64_t S, int64_t ld_s, int n, float* mean_out,
                                                      void* stream) {
    using namespace dnp;
    if (S < 0 || S > DNP_NOALLOC_MAX_SOURCES) {
        set_error("dnp_lattice_mean_potential_noalloc_f32: %lld sources, the allocation-free variant takes at most %lld "
                  "(DNP_NOALLOC_MAX_SOURCES); use dnp_lattice_mean_potential_f32 with a workspace",
                  (long long)S, (long long)DNP_NOALLOC_MAX_SOURCES);
        return DNP_ERR_INVALID;
    }
    if (n < 1 || n > DNP_NOALLOC_MAX_LATTICE) {
        set_error("dnp_lattice_mean_potential_noalloc_f32: lattice side %d outside [1, %d] (DNP_NOALLOC_MAX_LATTICE)", n,
                  DNP_NOALLOC_MAX_LATTICE);
        return DNP_ERR_INVALID;
    }
    if (!mean_out || (S > 0 && !src) || ld_s < 6) {
        set_error("dnp_lattice_mean_potential_noalloc_f32: null pointer or row stride below 6");
        return DNP_ERR_INVALID;
    }
    // the device address of THIS device's copy of the array, looked up once per device with that device current
    static std::atomic<int> have[kMaxDevices];
    static float* address[kMaxDevices];
    int dev = 0;
    DNP_CHECK_CUDA(cudaGetDevice(&dev));
    float* partials = nullptr;
    if (dev >= 0 && dev < kMaxDevices && have[dev].load(std::memory_order_acquire)) {
        partials = address[dev];
    } else {
        DNP_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&partials), g_static_partials));
        if (dev >= 0 && dev < kMaxDevices) {
            address[dev] = partials;
            have[dev].store(1, std::memory_order_release);
        }
    }
    // nothing has to start at zero: every CTA of the launch stores its own slot and the final sum reads exactly the
    // slots of this launch, so the previous call's contents are never seen
    return run_lattice_mean(src, S, ld_s, n, mean_out, partials, kStaticPartialFloats * sizeof(float),
                            (cudaStream_t)stream);
}

