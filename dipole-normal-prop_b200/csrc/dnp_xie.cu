// dnp_xie.cu -- the "xie" reflected-normal interaction.
//
// Replaces, from the reference's field_utils.py:
//   xie_field        :431-469   out[t, s, :] = (n_s - C (n_s.Rh) Rh) / |R|^3,  R = p_s - q_t   ([M, S, 3])
//   xie_intersaction :509-519   inter[t, s]  = xie_field[t, s] . n_t, NaN / Inf -> 0           ([M, S])
//   xie_propagation_points_in_order :569-605   T sequential chains over an [N, N] interaction matrix
// A coincident pair keeps the un-normalised value n_s (R_unit = R = 0 and no division, :456-460), so the
// diagonal of the self-interaction matrix is |n|^2.  The optional kNN mask (:447-455, a source acts only
// on its k nearest targets) arrives as one squared radius per source.
//
// The reference materialises the [N, N] float matrix (and an [N, N, 3] temporary) and then runs N
// sequential steps of `sum(mat[idx] * weights)` per chain.  Here the matrix is only materialised when the
// caller asks for it (xie_matrix_kernel, HBM-write bound).  The propagation never forms it:
//   * xie_chain_kernel: one thread-block cluster per chain keeps every point (position, normal, weight)
//     in shared memory; a step evaluates the visited point's row on the fly against the points already
//     weighted (unvisited points have weight 0 and are skipped), reduces in double -- warp shuffle, CTA,
//     then a DSMEM exchange of one double per CTA and one hardware cluster barrier -- and the owner sets
//     the new weight.  The T chains are independent clusters of one launch.  Latency bound (~2 us/step).
//   * xie_matvec_kernel (the `diffuse` epilogue, :598-604): out[t, i] = sum_j inter(i, j) w[t, j] as an
//     all-pairs pass, each pair evaluated once for all T chains, double accumulators.
#include <stdlib.h>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"

namespace dnp {

namespace {

// inter(target i, source j) of field_utils.py:440-462 dotted with the target normal (:513).
// rad2 < 0: no kNN mask; otherwise the pair counts only if |R|^2 <= rad2 (double, exact differences).
struct XiePair {
    float vx, vy, vz;      // the field vector (xie_field)
    bool masked;
};

__device__ __forceinline__ XiePair xie_pair(float px, float py, float pz, float nx, float ny, float nz,
                                            float qx, float qy, float qz, float C, double rad2) {
    XiePair o;
    o.masked = false;
    const float rx = __fsub_rn(px, qx), ry = __fsub_rn(py, qy), rz = __fsub_rn(pz, qz);
    if (rad2 >= 0.0) {
        const double dx = (double)px - (double)qx, dy = (double)py - (double)qy, dz = (double)pz - (double)qz;
        const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
        if (d2 > rad2) { o.vx = o.vy = o.vz = 0.f; o.masked = true; return o; }
    }
    const float r2 = __fadd_rn(__fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry)), __fmul_rn(rz, rz));
    if (r2 == 0.f) { o.vx = nx; o.vy = ny; o.vz = nz; return o; }       // coincident: un-normalised n_s
    float u = rsqrt_sfu(r2);
    u = u * (1.5f - 0.5f * r2 * u * u);                                  // one Newton step: full FP32 accuracy
    const float ux = rx * u, uy = ry * u, uz = rz * u;
    const float cd = C * (nx * ux + ny * uy + nz * uz);
    const float u3 = u * u * u;
    o.vx = (nx - cd * ux) * u3;
    o.vy = (ny - cd * uy) * u3;
    o.vz = (nz - cd * uz) * u3;
    return o;
}

__device__ __forceinline__ float finite_or_zero(float v) { return (fabsf(v) <= 3.402823466e38f) ? v : 0.f; }

__device__ __forceinline__ float xie_inter(float px, float py, float pz, float nx, float ny, float nz,
                                           float qx, float qy, float qz, float tx, float ty, float tz,
                                           float C, double rad2) {
    const XiePair p = xie_pair(px, py, pz, nx, ny, nz, qx, qy, qz, C, rad2);
    return finite_or_zero(p.vx * tx + p.vy * ty + p.vz * tz);           // :517-518
}

// ------------------------------------------------------------------ materialised matrix / field
constexpr int kMatThreads = 128;      // sources (columns) per CTA
constexpr int kMatRows = 64;          // targets (rows) per CTA

template <bool kField>
__global__ void __launch_bounds__(kMatThreads)
xie_matrix_kernel(const float* __restrict__ src, int64_t S, int64_t ld_s, const float* __restrict__ tgt, int64_t M,
                  int64_t ld_t, float C, const double* __restrict__ rad2, float* __restrict__ out, int64_t ld_o) {
    __shared__ float tq[kMatRows][6];
    const int64_t s = (int64_t)blockIdx.x * kMatThreads + threadIdx.x;
    const int64_t t0 = (int64_t)blockIdx.y * kMatRows;
    const int rows = (int)min((int64_t)kMatRows, M - t0);
    for (int i = threadIdx.x; i < rows * 6; i += kMatThreads) {
        const int r = i / 6, c = i % 6;
        tq[r][c] = (kField && c >= 3) ? 0.f : tgt[(t0 + r) * ld_t + c];
    }
    __syncthreads();
    if (s >= S) return;
    const float* p = src + s * ld_s;
    const float px = p[0], py = p[1], pz = p[2], nx = p[3], ny = p[4], nz = p[5];
    const double r2max = rad2 ? rad2[s] : -1.0;
    for (int r = 0; r < rows; ++r) {
        if (kField) {
            const XiePair v = xie_pair(px, py, pz, nx, ny, nz, tq[r][0], tq[r][1], tq[r][2], C, r2max);
            float* o = out + (t0 + r) * ld_o + s * 3;
            o[0] = v.vx; o[1] = v.vy; o[2] = v.vz;
        } else {
            out[(t0 + r) * ld_o + s] = xie_inter(px, py, pz, nx, ny, nz, tq[r][0], tq[r][1], tq[r][2],
                                                 tq[r][3], tq[r][4], tq[r][5], C, r2max);
        }
    }
}

// ------------------------------------------------------------------ in-order propagation, one cluster per chain
constexpr int kXThreads = 512;
constexpr int kXWarps = kXThreads / 32;
constexpr int kXMaxCluster = 16;

__device__ __forceinline__ uint32_t x_cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t x_cluster_size() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void x_cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n"
                 "barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void x_dsmem_store_f64(double* local_smem, uint32_t rank, double v) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_smem)), "r"(rank));
    asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(remote), "d"(v) : "memory");
}
__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// shared memory of one CTA: the exchange slots first (8-byte aligned), then the SoA state of `cap` points
size_t xie_smem_bytes(int cap, bool knn) {
    return sizeof(double) * (2 * kXMaxCluster + kXWarps) + (size_t)cap * (7 * sizeof(float) + (knn ? sizeof(double) : 0)) + 16;
}

__global__ void __launch_bounds__(kXThreads, 1)
xie_chain_kernel(const float* __restrict__ pts, int64_t N, int64_t ld, int cap, const int64_t* __restrict__ order,
                 float C, const double* __restrict__ rad2, uint8_t* __restrict__ flips, float* __restrict__ inter,
                 float* __restrict__ weights) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* slot = reinterpret_cast<double*>(smem_raw);                 // [2][kXMaxCluster]
    double* warp_part = slot + 2 * kXMaxCluster;                         // [kXWarps]
    double* s_rad = warp_part + kXWarps;                                 // [cap] if rad2
    float* f = reinterpret_cast<float*>(s_rad + (rad2 ? cap : 0));
    float *sx = f, *sy = f + cap, *sz = f + 2 * cap, *snx = f + 3 * cap, *sny = f + 4 * cap, *snz = f + 5 * cap,
          *sw = f + 6 * cap;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rank = x_cluster_rank(), ncta = x_cluster_size();
    const int64_t chain = blockIdx.x / ncta;
    const int64_t base = (int64_t)rank * cap;
    const int n_local = (int)max((int64_t)0, min((int64_t)cap, N - base));
    const int64_t* ord = order + chain * N;

    for (int i = tid; i < n_local; i += kXThreads) {
        const float* p = pts + (base + i) * ld;
        sx[i] = p[0]; sy[i] = p[1]; sz[i] = p[2];
        snx[i] = p[3]; sny[i] = p[4]; snz[i] = p[5];
        sw[i] = 0.f;                                                     // unvisited: weight 0 (:587-588)
        if (rad2) s_rad[i] = rad2[base + i];
    }
    __syncthreads();

    int64_t idx = ord[0];
    for (int64_t step = 0; step < N; ++step) {
        const int64_t idx_next = step + 1 < N ? ord[step + 1] : 0;       // prefetch
        const float* q = pts + idx * ld;
        const float qx = __ldg(q), qy = __ldg(q + 1), qz = __ldg(q + 2);
        const float tx = __ldg(q + 3), ty = __ldg(q + 4), tz = __ldg(q + 5);
        // interactions[idx] = sum_j mat[idx, j] * weights[j]            (:594)
        double acc = 0.0;
        for (int j = tid; j < n_local; j += kXThreads) {
            const float w = sw[j];
            if (w != 0.f)
                acc += (double)(xie_inter(sx[j], sy[j], sz[j], snx[j], sny[j], snz[j], qx, qy, qz, tx, ty, tz, C,
                                          rad2 ? s_rad[j] : -1.0) * w);
        }
        acc = warp_sum_f64(acc);
        if (lane == 0) warp_part[warp] = acc;
        __syncthreads();
        double* sl = slot + (step & 1) * kXMaxCluster;
        if (warp == 0) {
            double v = lane < kXWarps ? warp_part[lane] : 0.0;
            v = warp_sum_f64(v);
            if ((uint32_t)lane < ncta) x_dsmem_store_f64(&sl[rank], (uint32_t)lane, v);
        }
        x_cluster_barrier();
        double total = 0.0;
        for (uint32_t r = 0; r < ncta; ++r) total += sl[r];              // fixed order: identical in every CTA
        const int64_t li = idx - base;
        if (li >= 0 && li < n_local && tid == (int)(li % kXThreads)) {
            const bool neg = total < 0.0;
            sw[li] = neg ? -1.f : 1.f;                                   // :596
            flips[chain * N + idx] = neg ? 1 : 0;
            inter[chain * N + idx] = (float)total;
        }
        idx = idx_next;
        // sw[li] is only ever read by the thread that wrote it; warp_part / slot reuse is ordered by the
        // __syncthreads and the cluster barrier of the next step (slots are double-buffered).
    }
    __syncthreads();
    if (weights)
        for (int i = tid; i < n_local; i += kXThreads) weights[chain * N + base + i] = sw[i];
    x_cluster_barrier();     // no CTA may exit while a peer can still address its shared memory
}

// ------------------------------------------------------------------ diffuse epilogue: out[t, i] = sum_j inter(i, j) w[t, j]
constexpr int kMvThreads = 128;
constexpr int kMvTile = 128;
constexpr int kMvChains = 8;          // chains per pass

__global__ void __launch_bounds__(kMvThreads)
xie_matvec_kernel(const float* __restrict__ pts, int64_t N, int64_t ld, float C, const double* __restrict__ rad2,
                  const float* __restrict__ weights, int64_t T, int64_t t0, uint8_t* __restrict__ flips,
                  float* __restrict__ inter) {
    __shared__ float sp[kMvTile][6];
    __shared__ float swt[kMvChains][kMvTile];
    __shared__ double srad[kMvTile];
    const int64_t i = (int64_t)blockIdx.x * kMvThreads + threadIdx.x;
    const int nt = (int)min((int64_t)kMvChains, T - t0);
    float qx = 0.f, qy = 0.f, qz = 0.f, tx = 0.f, ty = 0.f, tz = 0.f;
    if (i < N) {
        const float* q = pts + i * ld;
        qx = q[0]; qy = q[1]; qz = q[2]; tx = q[3]; ty = q[4]; tz = q[5];
    }
    double acc[kMvChains];
#pragma unroll
    for (int t = 0; t < kMvChains; ++t) acc[t] = 0.0;
    for (int64_t j0 = 0; j0 < N; j0 += kMvTile) {
        const int n = (int)min((int64_t)kMvTile, N - j0);
        __syncthreads();
        for (int k = threadIdx.x; k < n * 6; k += kMvThreads) sp[k / 6][k % 6] = pts[(j0 + k / 6) * ld + k % 6];
        for (int k = threadIdx.x; k < kMvChains * kMvTile; k += kMvThreads) {
            const int t = k / kMvTile, j = k % kMvTile;
            swt[t][j] = (t < nt && j < n) ? weights[(t0 + t) * N + j0 + j] : 0.f;
        }
        if (rad2) for (int k = threadIdx.x; k < n; k += kMvThreads) srad[k] = rad2[j0 + k];
        __syncthreads();
        if (i < N) {
            for (int j = 0; j < n; ++j) {
                const float v = xie_inter(sp[j][0], sp[j][1], sp[j][2], sp[j][3], sp[j][4], sp[j][5], qx, qy, qz,
                                          tx, ty, tz, C, rad2 ? srad[j] : -1.0);
#pragma unroll
                for (int t = 0; t < kMvChains; ++t) acc[t] += (double)(v * swt[t][j]);
            }
        }
    }
    if (i < N) {
        for (int t = 0; t < nt; ++t) {
            inter[(t0 + t) * N + i] = (float)acc[t];
            flips[(t0 + t) * N + i] = acc[t] < 0.0 ? 1 : 0;
        }
    }
}

}  // namespace

int run_xie_matrix(bool field, const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t,
                   float C, const double* rad2, float* out, int64_t ld_o, cudaStream_t stream) {
    if (S <= 0 || M <= 0) return DNP_OK;
    dim3 grid((unsigned)((S + kMatThreads - 1) / kMatThreads), (unsigned)((M + kMatRows - 1) / kMatRows));
    if (grid.y > 65535u) {
        set_error("too many targets for one xie matrix call (%lld rows)", (long long)M);
        return DNP_ERR_INVALID;
    }
    if (field) xie_matrix_kernel<true><<<grid, kMatThreads, 0, stream>>>(src, S, ld_s, tgt, M, ld_t, C, rad2, out, ld_o);
    else xie_matrix_kernel<false><<<grid, kMatThreads, 0, stream>>>(src, S, ld_s, tgt, M, ld_t, C, rad2, out, ld_o);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp

namespace dnp {

int64_t xie_max_n(bool knn) {
    const size_t per_point = 7 * sizeof(float) + (knn ? sizeof(double) : 0);
    const size_t avail = 227 * 1024 - xie_smem_bytes(0, knn);
    return (int64_t)(avail / per_point) * kXMaxCluster;
}

size_t xie_workspace_bytes(int64_t N, int64_t T) {
    if (N < 1) N = 1;
    if (T < 1) T = 1;
    return (size_t)((N * T * (int64_t)sizeof(float) + 255) / 256 * 256);
}

int run_xie_propagate(const float* pts, int64_t N, int64_t ld, const int64_t* order, int64_t T, float C,
                      const double* rad2, int diffuse, uint8_t* flips, float* inter, void* workspace,
                      size_t workspace_bytes, cudaStream_t stream) {
    if (N <= 0 || T <= 0) return DNP_OK;
    const bool knn = rad2 != nullptr;
    if (N > xie_max_n(knn)) {
        set_error("N=%lld exceeds the single-cluster capacity of dnp_xie_propagate_in_order_f32 (%lld points)",
                  (long long)N, (long long)xie_max_n(knn));
        return DNP_ERR_INVALID;
    }
    if (diffuse && (!workspace || workspace_bytes < xie_workspace_bytes(N, T))) {
        set_error("workspace too small: need %zu bytes (dnp_xie_workspace_bytes), got %zu", xie_workspace_bytes(N, T),
                  workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    static std::atomic<int> attr_done[kMaxDevices];
    DNP_ONCE_PER_DEVICE(attr_done, {
        DNP_CHECK_CUDA(cudaFuncSetAttribute(xie_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        DNP_CHECK_CUDA(cudaFuncSetAttribute(xie_chain_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    });
    // smallest power-of-two cluster that gives every thread <= 4 points and fits the shared memory
    int ncta = 1;
    while (ncta < kXMaxCluster && (int64_t)ncta * kXThreads * 4 < N) ncta *= 2;
    while (ncta < kXMaxCluster && xie_smem_bytes((int)((N + ncta - 1) / ncta), knn) > 227 * 1024) ncta *= 2;
    float* weights = diffuse ? reinterpret_cast<float*>(workspace) : nullptr;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.blockDim = dim3(kXThreads);
    cfg.stream = stream;
    int cap = 0;
    for (;;) {
        cap = (int)((N + ncta - 1) / ncta);
        cfg.gridDim = dim3((unsigned)(T * ncta));
        cfg.dynamicSmemBytes = xie_smem_bytes(cap, knn);
        attr[0].val.clusterDim.x = ncta;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        int max_clusters = 0;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&max_clusters, xie_chain_kernel, &cfg);
        if (e == cudaSuccess && max_clusters >= 1) break;
        (void)cudaGetLastError();
        if (ncta == 1 || xie_smem_bytes((int)((N + ncta / 2 - 1) / (ncta / 2)), knn) > 227 * 1024) {
            set_error("cannot schedule a %d-CTA cluster for N=%lld", ncta, (long long)N);
            return DNP_ERR_CUDA;
        }
        ncta /= 2;
    }
    DNP_CHECK_CUDA(cudaLaunchKernelEx(&cfg, xie_chain_kernel, pts, N, ld, cap, order, C, rad2, flips, inter, weights));
    count_launch(1);
    if (diffuse) {                                                       // :598-604
        const unsigned blocks = (unsigned)((N + kMvThreads - 1) / kMvThreads);
        for (int64_t t0 = 0; t0 < T; t0 += kMvChains) {
            xie_matvec_kernel<<<blocks, kMvThreads, 0, stream>>>(pts, N, ld, C, rad2, weights, T, t0, flips, inter);
            count_launch(1);
        }
        DNP_CHECK_CUDA(cudaGetLastError());
    }
    return DNP_OK;
}

// ---------------------------------------------------------------- k-th neighbour radius (the kNN mask, :447-455)
// out[s] = squared distance (double, exact differences of the float coordinates: the expression xie_pair tests) from
// source s to its k-th nearest target.  The reference builds a scipy KD-tree per call; round 1 used torch cdist +
// topk (S x M doubles materialised in chunks).  Here: one thread per source, targets staged through shared
// memory, the k smallest distances kept in a per-thread sorted array (insertion only when a candidate beats the
// current k-th, which becomes rare after the first few hundred targets).  k <= 64.
namespace {
constexpr int kKnnThreads = 128;
constexpr int kKnnTile = 1024;
constexpr int kKnnMax = 64;

__global__ void __launch_bounds__(kKnnThreads)
knn_radius2_kernel(const float* __restrict__ src, int64_t S, int64_t ld_s, const float* __restrict__ tgt, int64_t M,
                   int64_t ld_t, int k, double* __restrict__ out) {
    __shared__ float st[kKnnTile * 3];
    const int64_t s = (int64_t)blockIdx.x * kKnnThreads + threadIdx.x;
    double px = 0.0, py = 0.0, pz = 0.0;
    if (s < S) { px = src[s * ld_s]; py = src[s * ld_s + 1]; pz = src[s * ld_s + 2]; }
    double best[kKnnMax];                                 // ascending; best[k - 1] is the current k-th
#pragma unroll 1
    for (int i = 0; i < k; ++i) best[i] = 1.0e300;
    for (int64_t t0 = 0; t0 < M; t0 += kKnnTile) {
        const int n_here = (int)min((int64_t)kKnnTile, M - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < n_here * 3; i += kKnnThreads) st[i] = tgt[(t0 + i / 3) * ld_t + i % 3];
        __syncthreads();
        if (s >= S) continue;
        for (int j = 0; j < n_here; ++j) {
            const double dx = px - (double)st[3 * j], dy = py - (double)st[3 * j + 1], dz = pz - (double)st[3 * j + 2];
            const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
            if (d2 < best[k - 1]) {
                int i = k - 1;
                while (i > 0 && best[i - 1] > d2) { best[i] = best[i - 1]; --i; }
                best[i] = d2;
            }
        }
    }
    if (s < S) out[s] = best[k - 1];
}
}  // namespace

int run_knn_radius2(const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t, int k,
                    double* out, cudaStream_t stream) {
    if (S <= 0) return DNP_OK;
    if (k < 1 || k > kKnnMax || k > M) {
        set_error("k must be in [1, min(%d, targets)] (got %d)", kKnnMax, k);
        return DNP_ERR_INVALID;
    }
    knn_radius2_kernel<<<(unsigned)((S + kKnnThreads - 1) / kKnnThreads), kKnnThreads, 0, stream>>>(src, S, ld_s, tgt, M,
                                                                                               ld_t, k, out);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
