// dnp_scores.cu -- patch-pair flip scores (K3) and per-patch interactions (K4).
//
// K3 replaces the per-edge chain field_grad -> (E*n).sum(-1).sum() -> .cpu()
// of field_utils.field_edge_calculator (field_utils.py:145-160) for a whole
// edge list in one launch: one CTA per edge, source patch staged through shared
// memory, targets in registers, fixed-order block reduction, no atomics.
// K4 replaces the list comprehension at field_utils.py:244,316 (one tiny launch
// and one host sync per remaining patch per step) and the abs().argmax() at
// :247,319 by two launches and no host synchronisation.
#include "dnp_common.cuh"
#include "dnp_pairmath.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"

namespace dnp {

namespace {

constexpr int kEdgeThreads = 256;                 // the per-patch reduction kernels below
constexpr int kEdgeBlock = 128;                   // edge kernel: 128 threads x 2 targets = one 256-target slice per pass
constexpr int kEdgeChunk = 256;                   // sources staged per pass (48 B each: 12 KB)

__device__ __forceinline__ float block_sum_256(float v, float* warp_part) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    if (threadIdx.x < 32) {
        r = threadIdx.x < (blockDim.x >> 5) ? warp_part[threadIdx.x] : 0.f;
        r = warp_sum(r);
    }
    return r;   // valid in warp 0
}

// One CTA per edge.  The source patch is staged through shared memory in the all-pairs kernel's packed row layout
// ([x y z nx | ny nz mx my | mz . . .], m = -3 n) and read back with broadcast LDS.128; a thread owns a target PAIR
// (rows t and t + 128 of the current 256-target slice) and runs the packed 19-instruction pair arithmetic of
// dnp_pairmath.cuh.  Warps whose targets are all past the end of the patch skip the arithmetic (they still help
// staging), so a 1048-row patch costs 4.1 slices of work, not 5.  Fixed-order block reduction, no atomics.
__global__ void __launch_bounds__(kEdgeBlock)
edge_interactions_kernel(const float* __restrict__ pts, int64_t ld, const int64_t* __restrict__ perm,
                         const int64_t* __restrict__ offsets, const int64_t* __restrict__ tperm,
                         const int64_t* __restrict__ toffsets, const int64_t* __restrict__ edges, float eps,
                         float* __restrict__ out) {
    __shared__ float4 s_src[kEdgeChunk * 3];
    __shared__ float warp_part[32];
    const int64_t e = blockIdx.x;
    const int64_t ps = edges[2 * e], pt = edges[2 * e + 1];
    const int64_t s0 = offsets[ps], ns = offsets[ps + 1] - s0;
    const int64_t t0 = toffsets[pt], nt = toffsets[pt + 1] - t0;
    const f32x2 eps2 = pack2(eps, eps), zero2 = pack2(0.f, 0.f);
    float total = 0.f;
    for (int64_t tb = 0; tb < nt; tb += 2 * kEdgeBlock) {
        const int64_t ta = tb + threadIdx.x, tc = ta + kEdgeBlock;
        float q[2][6] = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
        if (ta < nt) {
            const float* p = pts + tperm[t0 + ta] * ld;
#pragma unroll
            for (int c = 0; c < 6; ++c) q[0][c] = p[c];
        }
        if (tc < nt) {
            const float* p = pts + tperm[t0 + tc] * ld;
#pragma unroll
            for (int c = 0; c < 6; ++c) q[1][c] = p[c];
        }
        // does this warp own any target of the slice?  (lane 0 holds the warp's smallest row)
        const bool warp_live = tb + (threadIdx.x & ~31) < nt;
        const f32x2 qx = pack2(q[0][0], q[1][0]), qy = pack2(q[0][1], q[1][1]), qz = pack2(q[0][2], q[1][2]);
        f32x2 ex = zero2, ey = zero2, ez = zero2;
        for (int64_t sb = 0; sb < ns; sb += kEdgeChunk) {
            const int n_here = (int)min((int64_t)kEdgeChunk, ns - sb);
            __syncthreads();
            for (int i = threadIdx.x; i < n_here; i += kEdgeBlock) {
                const float* p = pts + perm[s0 + sb + i] * ld;
                float v[6];
#pragma unroll
                for (int c = 0; c < 6; ++c) v[c] = p[c];
                pack_source_row(v, s_src + 3 * i);
            }
            __syncthreads();
            if (warp_live) {
                f32x2 ax = zero2, ay = zero2, az = zero2;
#pragma unroll 4
                for (int i = 0; i < n_here; ++i) {
                    const SrcX2 src(s_src[3 * i], s_src[3 * i + 1], s_src[3 * i + 2]);
                    pair_field_x2(src, qx, qy, qz, eps2, ax, ay, az);
                }
                ex = add2(ex, ax); ey = add2(ey, ay); ez = add2(ez, az);
            }
        }
        float fx[2], fy[2], fz[2];
        unpack2(ex, fx[0], fx[1]);
        unpack2(ey, fy[0], fy[1]);
        unpack2(ez, fz[0], fz[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            // non-finite field components are zeroed like the leaf does (field_utils.py:110-115): a NaN point makes
            // the field of every target it touches vanish instead of poisoning the score
            if (!(fabsf(fx[h]) <= 3.402823466e38f)) fx[h] = 0.f;
            if (!(fabsf(fy[h]) <= 3.402823466e38f)) fy[h] = 0.f;
            if (!(fabsf(fz[h]) <= 3.402823466e38f)) fz[h] = 0.f;
            if ((h == 0 ? ta : tc) < nt)
                total += __fadd_rn(__fadd_rn(__fmul_rn(fx[h], q[h][3]), __fmul_rn(fy[h], q[h][4])), __fmul_rn(fz[h], q[h][5]));
        }
    }
    __syncthreads();
    const float r = block_sum_256(total, warp_part);
    if (threadIdx.x == 0) out[e] = r;
}

// out[p] = sum_{t in patch p} E_t . n_t  (0 for inactive patches)
__global__ void __launch_bounds__(kEdgeThreads)
patch_interactions_kernel(const float* __restrict__ E, int64_t ld_e, const float* __restrict__ normals,
                          int64_t ld_n, const int64_t* __restrict__ perm, const int64_t* __restrict__ offsets,
                          const uint8_t* __restrict__ active, const uint8_t* __restrict__ owned,
                          float* __restrict__ out) {
    __shared__ float warp_part[32];
    const int64_t p = blockIdx.x;
    if ((active && !active[p]) || (owned && !owned[p])) {
        if (threadIdx.x == 0) out[p] = 0.f;
        return;
    }
    const int64_t t0 = offsets[p], nt = offsets[p + 1] - t0;
    float acc = 0.f;
    for (int64_t t = threadIdx.x; t < nt; t += kEdgeThreads) {
        const int64_t r = perm[t0 + t];
        const float* e = E + r * ld_e;
        const float* n = normals + r * ld_n;
        acc += __fadd_rn(__fadd_rn(__fmul_rn(e[0], n[0]), __fmul_rn(e[1], n[1])), __fmul_rn(e[2], n[2]));
    }
    const float r = block_sum_256(acc, warp_part);
    if (threadIdx.x == 0) out[p] = r;
}

// first maximal |out[p]| among active patches (torch abs().argmax() on the remaining list)
__global__ void __launch_bounds__(1024)
patch_argmax_kernel(const float* __restrict__ vals, const uint8_t* __restrict__ active, int64_t P,
                    int32_t* __restrict__ best_idx, float* __restrict__ best_val, PatchAdvance adv) {
    __shared__ unsigned long long warp_best[32];
    const unsigned long long best = block_first_absmax(vals, active, P, warp_best);
    if (threadIdx.x == 0) {
        int32_t idx = -1;
        float val = 0.f;
        if (best != 0ull) {
            idx = (int32_t)(0xffffffffu - (uint32_t)best);
            val = vals[idx];
        }
        *best_idx = idx;
        *best_val = val;
        if (adv.cur) {                           // device-driven loop: the winner becomes the next step's patch
            adv.cur[2] = adv.cur[0];
            adv.cur[0] = idx;
            adv.cur[1] = val < 0.f;              // field_utils.py:250,322
            if (adv.order) adv.order[adv.step] = idx;
            if (adv.value) adv.value[adv.step] = val;
        }
    }
}


// per-patch xyz covariance about the patch mean, divided by the patch size
// (util.pca_eigen_values, util.py:495-497), accumulated in double in a fixed
// order: one CTA per patch, two passes (mean, then second moments).
__device__ __forceinline__ double block_sum_f64(double v, double* scratch) {
    __syncthreads();
    scratch[threadIdx.x] = v;
    __syncthreads();
    for (int o = kEdgeThreads / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) scratch[threadIdx.x] += scratch[threadIdx.x + o];
        __syncthreads();
    }
    return scratch[0];
}

__global__ void __launch_bounds__(kEdgeThreads)
patch_covariance_kernel(const float* __restrict__ pts, int64_t ld, const int64_t* __restrict__ perm,
                        const int64_t* __restrict__ offsets, double* __restrict__ cov) {
    __shared__ double scratch[kEdgeThreads];
    const int64_t p = blockIdx.x;
    const int64_t t0 = offsets[p], nt = offsets[p + 1] - t0;
    double m[3] = {0.0, 0.0, 0.0};
    for (int64_t t = threadIdx.x; t < nt; t += kEdgeThreads) {
        const float* q = pts + perm[t0 + t] * ld;
        m[0] += q[0]; m[1] += q[1]; m[2] += q[2];
    }
    const double inv = nt > 0 ? 1.0 / (double)nt : 0.0;
#pragma unroll
    for (int c = 0; c < 3; ++c) m[c] = block_sum_f64(m[c], scratch) * inv;
    double c6[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};   // xx xy xz yy yz zz
    for (int64_t t = threadIdx.x; t < nt; t += kEdgeThreads) {
        const float* q = pts + perm[t0 + t] * ld;
        const double x = q[0] - m[0], y = q[1] - m[1], z = q[2] - m[2];
        c6[0] += x * x; c6[1] += x * y; c6[2] += x * z; c6[3] += y * y; c6[4] += y * z; c6[5] += z * z;
    }
#pragma unroll
    for (int c = 0; c < 6; ++c) c6[c] = block_sum_f64(c6[c], scratch) * inv;
    if (threadIdx.x == 0) {
        double* o = cov + 9 * p;
        o[0] = c6[0]; o[1] = c6[1]; o[2] = c6[2];
        o[3] = c6[1]; o[4] = c6[3]; o[5] = c6[4];
        o[6] = c6[2]; o[7] = c6[4]; o[8] = c6[5];
    }
}

// streamed per-point sign: n_r *= (E_r . n_r > 0 ? 1 : -1)  (>= for DNP_EPI_SIGN_GE)
// field_utils.py:267-271,337-342,382-385 (strict >) and :196-199 (>=).  36 bytes of HBM traffic per point.
__global__ void __launch_bounds__(256)
field_sign_kernel(const float* __restrict__ E, int64_t ld_e, float* __restrict__ normals, int64_t ld_n,
                  const int64_t* __restrict__ idx, int64_t n, int ge) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const int64_t r = idx ? idx[k] : k;
        const float* e = E + r * ld_e;
        float* nr = normals + r * ld_n;
        const float nx = nr[0], ny = nr[1], nz = nr[2];
        const float dot = __fadd_rn(__fadd_rn(__fmul_rn(e[0], nx), __fmul_rn(e[1], ny)), __fmul_rn(e[2], nz));
        const float sgn = (ge ? (dot >= 0.f) : (dot > 0.f)) ? 1.f : -1.f;
        nr[0] = nx * sgn; nr[1] = ny * sgn; nr[2] = nz * sgn;
    }
}

}  // namespace

int run_edge_interactions_batched(const float* pts, int64_t ld, const int64_t* perm, const int64_t* offsets,
                                  int64_t P, const int64_t* tgt_perm, const int64_t* tgt_offsets,
                                  const int64_t* edges, int64_t n_edges, float eps, float* out,
                                  cudaStream_t stream) {
    (void)P;
    if (n_edges <= 0) return DNP_OK;
    edge_interactions_kernel<<<(unsigned)n_edges, kEdgeBlock, 0, stream>>>(pts, ld, perm, offsets, tgt_perm,
                                                                             tgt_offsets, edges, eps, out);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

int run_patch_interactions(const float* E, int64_t ld_e, const float* normals, int64_t ld_n, const int64_t* perm,
                           const int64_t* offsets, int64_t P, const uint8_t* active, float* out,
                           int32_t* best_idx, float* best_val, cudaStream_t stream, PatchAdvance adv,
                           const ShardHook* shard) {
    if (P <= 0) return DNP_OK;
    patch_interactions_kernel<<<(unsigned)P, kEdgeThreads, 0, stream>>>(E, ld_e, normals, ld_n, perm, offsets, active,
                                                                       shard ? shard->owned : nullptr, out);
    count_launch(1);
    if (shard && shard->combine) {
        DNP_CHECK_CUDA(cudaGetLastError());
        const int rc = shard->combine(shard->ctx, out, P, stream);
        if (rc != DNP_OK) return rc;
    }
    if (best_idx && best_val) {
        patch_argmax_kernel<<<1, 1024, 0, stream>>>(out, active, P, best_idx, best_val, adv);
        count_launch(1);
    }
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

int run_patch_covariance(const float* pts, int64_t ld, const int64_t* perm, const int64_t* offsets, int64_t P,
                         double* cov, cudaStream_t stream) {
    if (P <= 0) return DNP_OK;
    patch_covariance_kernel<<<(unsigned)P, kEdgeThreads, 0, stream>>>(pts, ld, perm, offsets, cov);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

int run_field_sign(const float* E, int64_t ld_e, float* normals, int64_t ld_n, const int64_t* idx, int64_t n,
                   int ge, cudaStream_t stream) {
    if (n <= 0) return DNP_OK;
    const int64_t want = (n + 255) / 256;
    const int64_t cap = (int64_t)sm_count_cached() * 16;
    field_sign_kernel<<<(unsigned)(want < cap ? want : cap), 256, 0, stream>>>(E, ld_e, normals, ld_n, idx, n, ge);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
