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
#include "dnp_pairs.h"
#include "dnp_scores.h"

namespace dnp {

namespace {

constexpr int kEdgeThreads = 256;
constexpr int kEdgeChunk = 512;                   // sources staged per pass (12 KB)
constexpr float kRsqrtClamp = 1.152921504606847e18f;

// scalar twin of pair_field_x2 (dnp_pairs.cu): same formula, same SFU approximations
__device__ __forceinline__ void pair_field_scalar(float px, float py, float pz, float nx, float ny, float nz,
                                                  float qx, float qy, float qz, float eps,
                                                  float& ax, float& ay, float& az, float& bx, float& by, float& bz) {
    const float rx = px - qx, ry = py - qy, rz = pz - qz;
    const float r2 = fmaf(rz, rz, fmaf(ry, ry, rx * rx));
    const float u = rsqrt_sfu(r2);
    const float uc = fminf(u, kRsqrtClamp);
    const float den = fmaf(r2, r2 * u, eps);
    const float w = fmaxf(rcp_sfu(den), 0.f);
    const float d = fmaf(nz, rz, fmaf(ny, ry, nx * rx));
    const float c = (d * w) * (uc * uc);
    ax = fmaf(w, nx, ax); ay = fmaf(w, ny, ay); az = fmaf(w, nz, az);
    bx = fmaf(c, rx, bx); by = fmaf(c, ry, by); bz = fmaf(c, rz, bz);
}

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

__global__ void __launch_bounds__(kEdgeThreads)
edge_interactions_kernel(const float* __restrict__ pts, int64_t ld, const int64_t* __restrict__ perm,
                         const int64_t* __restrict__ offsets, const int64_t* __restrict__ tperm,
                         const int64_t* __restrict__ toffsets, const int64_t* __restrict__ edges, float eps,
                         float* __restrict__ out) {
    __shared__ float4 s_src[kEdgeChunk * 2];      // [x y z nx] [ny nz 0 0]
    __shared__ float warp_part[32];
    const int64_t e = blockIdx.x;
    const int64_t ps = edges[2 * e], pt = edges[2 * e + 1];
    const int64_t s0 = offsets[ps], ns = offsets[ps + 1] - s0;
    const int64_t t0 = toffsets[pt], nt = toffsets[pt + 1] - t0;
    float total = 0.f;
    for (int64_t tb = 0; tb < nt; tb += kEdgeThreads) {
        const int64_t t = tb + threadIdx.x;
        float qx = 0.f, qy = 0.f, qz = 0.f, tnx = 0.f, tny = 0.f, tnz = 0.f;
        if (t < nt) {
            const float* q = pts + tperm[t0 + t] * ld;
            qx = q[0]; qy = q[1]; qz = q[2]; tnx = q[3]; tny = q[4]; tnz = q[5];
        }
        float ex = 0.f, ey = 0.f, ez = 0.f;
        for (int64_t sb = 0; sb < ns; sb += kEdgeChunk) {
            const int n_here = (int)min((int64_t)kEdgeChunk, ns - sb);
            __syncthreads();
            for (int i = threadIdx.x; i < n_here; i += kEdgeThreads) {
                const float* p = pts + perm[s0 + sb + i] * ld;
                s_src[2 * i] = make_float4(p[0], p[1], p[2], p[3]);
                s_src[2 * i + 1] = make_float4(p[4], p[5], 0.f, 0.f);
            }
            __syncthreads();
            float ax = 0.f, ay = 0.f, az = 0.f, bx = 0.f, by = 0.f, bz = 0.f;
#pragma unroll 4
            for (int i = 0; i < n_here; ++i) {
                const float4 A = s_src[2 * i], B = s_src[2 * i + 1];
                pair_field_scalar(A.x, A.y, A.z, A.w, B.x, B.y, qx, qy, qz, eps, ax, ay, az, bx, by, bz);
            }
            ex += fmaf(-3.f, bx, ax); ey += fmaf(-3.f, by, ay); ez += fmaf(-3.f, bz, az);
        }
        // non-finite field components are zeroed like the leaf does (field_utils.py:110-115): a NaN point makes
        // the field of every target it touches vanish instead of poisoning the score
        if (!(fabsf(ex) <= 3.402823466e38f)) ex = 0.f;
        if (!(fabsf(ey) <= 3.402823466e38f)) ey = 0.f;
        if (!(fabsf(ez) <= 3.402823466e38f)) ez = 0.f;
        if (t < nt) total += __fadd_rn(__fadd_rn(__fmul_rn(ex, tnx), __fmul_rn(ey, tny)), __fmul_rn(ez, tnz));
    }
    __syncthreads();
    const float r = block_sum_256(total, warp_part);
    if (threadIdx.x == 0) out[e] = r;
}

// out[p] = sum_{t in patch p} E_t . n_t  (0 for inactive patches)
__global__ void __launch_bounds__(kEdgeThreads)
patch_interactions_kernel(const float* __restrict__ E, int64_t ld_e, const float* __restrict__ normals,
                          int64_t ld_n, const int64_t* __restrict__ perm, const int64_t* __restrict__ offsets,
                          const uint8_t* __restrict__ active, float* __restrict__ out) {
    __shared__ float warp_part[32];
    const int64_t p = blockIdx.x;
    if (active && !active[p]) {
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
    edge_interactions_kernel<<<(unsigned)n_edges, kEdgeThreads, 0, stream>>>(pts, ld, perm, offsets, tgt_perm,
                                                                             tgt_offsets, edges, eps, out);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

int run_patch_interactions(const float* E, int64_t ld_e, const float* normals, int64_t ld_n, const int64_t* perm,
                           const int64_t* offsets, int64_t P, const uint8_t* active, float* out,
                           int32_t* best_idx, float* best_val, cudaStream_t stream, PatchAdvance adv) {
    if (P <= 0) return DNP_OK;
    patch_interactions_kernel<<<(unsigned)P, kEdgeThreads, 0, stream>>>(E, ld_e, normals, ld_n, perm, offsets, active,
                                                                       out);
    count_launch(1);
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
