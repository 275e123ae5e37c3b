// dnp_points.cu -- per-point greedy dipole propagation as ONE persistent kernel (K5).
//
// Replaces field_utils.strongest_field_propagation_points (field_utils.py:353-388).
// The reference runs N-1 dependent steps, each a 1 x (N-1) field_grad call, a
// masked dot product, abs().argmax() and a conditional flip -- dozens of launches
// and >= 8 host synchronisations per step.  Here the whole loop is one launch of
// a single thread-block cluster (<= 16 CTAs):
//   * every point's position, normal, running field E and visited flag live in
//     the shared memory of the CTA that owns it for the whole run;
//   * one step = add the current source's field to all owned points, masked
//     |E.n| arg-max by warp REDUX -> CTA -> cluster: the CTA winner (key, position,
//     normal) is pushed into every CTA's shared memory through DSMEM
//     (st.shared::cluster) and one hardware cluster barrier ends the step;
//   * the arg-max key packs (|v| bits, inverted index, sign) into 64 bits, so
//     "first maximum in index order" (torch argmax) and the flip decision
//     (v < 0, :375) come out of one unsigned max.
// Arithmetic is IEEE round-to-nearest in the reference's evaluation order
// (sqrt.rn / div.rn, no FMA contraction): with a single source the leaf's sum
// over sources is the identity, so E and every decision match the reference's
// CPU float32 path bit for bit (tests/golden/points_prop.npz).
// This stage is latency-bound (about a thousand cycles per step), not
// FLOP-bound; the all-pairs throughput kernels are in dnp_pairs.cu.
#include <stdlib.h>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"

namespace dnp {

namespace {

constexpr int kPtThreads = 512;
constexpr int kMaxCluster = 16;
constexpr int kWarps = kPtThreads / 32;

struct __align__(16) Candidate {
    unsigned long long key;   // (|v| bits << 32) | ((0x7fffffff - idx) << 1) | sign(v)
    float qx, qy, qz;         // position of the candidate
    float nx, ny, nz;         // its normal BEFORE the flip decision
};
static_assert(sizeof(Candidate) == 32, "candidate must be two 16-byte vectors");

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_size() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n"
                 "barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// 16-byte store into the shared memory of CTA `rank` of this cluster (DSMEM)
__device__ __forceinline__ void dsmem_store16(void* local_smem, uint32_t rank, uint4 v) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_smem)), "r"(rank));
    asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(remote), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

__device__ __forceinline__ unsigned long long make_key(float v, int64_t idx) {
    const uint32_t a = __float_as_uint(fabsf(v));
    const uint32_t lo = ((0x7fffffffu - (uint32_t)idx) << 1) | (v < 0.f ? 1u : 0u);
    return ((unsigned long long)a << 32) | lo;
}

// 64-bit max over a warp with two 32-bit REDUX passes
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long k) {
    const uint32_t hi = (uint32_t)(k >> 32);
    const uint32_t hmax = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t lo = (hi == hmax) ? (uint32_t)k : 0u;
    const uint32_t lmax = __reduce_max_sync(0xffffffffu, lo);
    return ((unsigned long long)hmax << 32) | lmax;
}

// shared-memory carve-up: SoA state of `cap` points, then the exchange buffers
struct PointSmem {
    float *qx, *qy, *qz, *nx, *ny, *nz, *ex, *ey, *ez;
    uint8_t* visited;
    Candidate* warp_cand;   // [kWarps]
    Candidate* cand;        // [2][kMaxCluster]
};

__device__ __forceinline__ PointSmem carve(unsigned char* base, int cap) {
    PointSmem s;
    Candidate* c = reinterpret_cast<Candidate*>(base);
    s.cand = c;
    s.warp_cand = c + 2 * kMaxCluster;
    float* f = reinterpret_cast<float*>(c + 2 * kMaxCluster + kWarps);
    s.qx = f; s.qy = f + cap; s.qz = f + 2 * cap;
    s.nx = f + 3 * cap; s.ny = f + 4 * cap; s.nz = f + 5 * cap;
    s.ex = f + 6 * cap; s.ey = f + 7 * cap; s.ez = f + 8 * cap;
    s.visited = reinterpret_cast<uint8_t*>(f + 9 * cap);
    return s;
}

size_t point_smem_bytes(int cap) {
    return sizeof(Candidate) * (2 * kMaxCluster + kWarps) + size_t(cap) * 9 * sizeof(float) + size_t(cap) + 16;
}

// One step's sweep over the points this CTA owns (field_utils.py:379-380 with the leaf :96-116, then :369):
//   E[m] += field of the current source at m (m != cur, IEEE ops in the reference's order),
//   best = max over unvisited points of key(E[m].n[m], m).
__device__ __forceinline__ void sweep_points(const PointSmem& sm, int n_local, int64_t base, int64_t cur,
                                             float px, float py, float pz, float p3, float p4, float p5, float eps,
                                             unsigned long long& best, int& best_i) {
    for (int i = threadIdx.x; i < n_local; i += kPtThreads) {
        const int64_t m = base + i;
        float ex = sm.ex[i], ey = sm.ey[i], ez = sm.ez[i];
        if (m != cur) {
            const float rx = __fsub_rn(px, sm.qx[i]), ry = __fsub_rn(py, sm.qy[i]), rz = __fsub_rn(pz, sm.qz[i]);
            const float r = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry)), __fmul_rn(rz, rz)));
            if (r != 0.f) {
                const float ux = __fdiv_rn(rx, r), uy = __fdiv_rn(ry, r), uz = __fdiv_rn(rz, r);
                const float d3 = __fmul_rn(3.f, __fadd_rn(__fadd_rn(__fmul_rn(p3, ux), __fmul_rn(p4, uy)), __fmul_rn(p5, uz)));
                const float den = __fadd_rn(__fmul_rn(__fmul_rn(r, r), r), eps);
                ex = __fsub_rn(ex, __fdiv_rn(__fsub_rn(__fmul_rn(d3, ux), p3), den));
                ey = __fsub_rn(ey, __fdiv_rn(__fsub_rn(__fmul_rn(d3, uy), p4), den));
                ez = __fsub_rn(ez, __fdiv_rn(__fsub_rn(__fmul_rn(d3, uz), p5), den));
                sm.ex[i] = ex; sm.ey[i] = ey; sm.ez[i] = ez;
            }
        }
        if (!sm.visited[i]) {
            const float v = __fadd_rn(__fadd_rn(__fmul_rn(ex, sm.nx[i]), __fmul_rn(ey, sm.ny[i])), __fmul_rn(ez, sm.nz[i]));
            const unsigned long long k = make_key(v, m);
            if (best_i < 0 || k > best) { best = k; best_i = i; }
        }
    }
}

// The whole greedy loop for ONE cloud of N points on the calling thread-block cluster (every CTA of the cluster
// calls it with the same arguments).  N == 0 is allowed (the barriers still pair up).
__device__ __forceinline__ void propagate_on_cluster(unsigned char* smem_raw, float* __restrict__ pts, int64_t N,
                                                     int64_t ld, int cap, int64_t start, float eps, int diffuse,
                                                     int32_t* __restrict__ order, uint8_t* __restrict__ flipped) {
    const PointSmem sm = carve(smem_raw, cap);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rank = cluster_rank(), ncta = cluster_size();
    const int64_t base = (int64_t)rank * cap;
    const int n_local = (int)max((int64_t)0, min((int64_t)cap, N - base));

    for (int i = tid; i < n_local; i += kPtThreads) {
        const float* p = pts + (base + i) * ld;
        sm.qx[i] = p[0]; sm.qy[i] = p[1]; sm.qz[i] = p[2];
        sm.nx[i] = p[3]; sm.ny[i] = p[4]; sm.nz[i] = p[5];
        sm.ex[i] = 0.f; sm.ey[i] = 0.f; sm.ez[i] = 0.f;
        sm.visited[i] = (base + i == start);
    }
    // current source (field_utils.py:358-362): the starting point, never flipped
    int64_t cur = start;
    float px = 0.f, py = 0.f, pz = 0.f, p3 = 0.f, p4 = 0.f, p5 = 0.f;
    if (N > 0) {
        const float* p = pts + start * ld;
        px = p[0]; py = p[1]; pz = p[2]; p3 = p[3]; p4 = p[4]; p5 = p[5];
    }
    if (rank == 0 && tid == 0 && N > 0) {
        if (order) order[0] = (int32_t)start;
        if (flipped) flipped[0] = 0;
    }
    __syncthreads();

    for (int64_t step = 1; step <= N; ++step) {
        // ---- E[others] += field_grad(pts[cur], pts[others,:3], eps)   (:379-380, leaf :96-116)
        //      interaction = (E * n).sum(-1) over unvisited points      (:369)
        unsigned long long best = 0ull;
        int best_i = -1;
        sweep_points(sm, n_local, base, cur, px, py, pz, p3, p4, p5, eps, best, best_i);
        if (step == N) break;   // the last iteration only spreads the last source's field

        // ---- abs().argmax(): warp -> CTA -> cluster                                 (:372-373)
        // keys of real candidates are never 0 (the inverted index field is >= 1)
        const unsigned long long wbest = warp_max_u64(best);
        if (best_i >= 0 && best == wbest && wbest != 0ull) {
            Candidate c;
            c.key = best;
            c.qx = sm.qx[best_i]; c.qy = sm.qy[best_i]; c.qz = sm.qz[best_i];
            c.nx = sm.nx[best_i]; c.ny = sm.ny[best_i]; c.nz = sm.nz[best_i];
            sm.warp_cand[warp] = c;
        } else if (lane == 0 && wbest == 0ull) {
            sm.warp_cand[warp].key = 0ull;
        }
        __syncthreads();
        Candidate* slot = sm.cand + (step & 1) * kMaxCluster;
        if (warp == 0) {
            const unsigned long long k = lane < kWarps ? sm.warp_cand[lane].key : 0ull;
            const unsigned long long cbest = warp_max_u64(k);
            const uint32_t src_warp = __ffs(__ballot_sync(0xffffffffu, lane < kWarps && k == cbest)) - 1;
            // lane r pushes this CTA's winner into slot[rank] of CTA r
            if ((uint32_t)lane < ncta) {
                const uint4* w = reinterpret_cast<const uint4*>(&sm.warp_cand[src_warp]);
                uint4 v0 = w[0], v1 = w[1];
                if (cbest == 0ull) { v0.x = 0u; v0.y = 0u; }
                uint4* dst = reinterpret_cast<uint4*>(&slot[rank]);
                dsmem_store16(dst, (uint32_t)lane, v0);
                dsmem_store16(dst + 1, (uint32_t)lane, v1);
            }
        }
        cluster_barrier();

        // every thread scans the <= 16 CTA winners (shared-memory broadcast reads)
        unsigned long long gbest = 0ull;
        int gsrc = 0;
        for (uint32_t r = 0; r < ncta; ++r) {
            const unsigned long long k = slot[r].key;
            if (k > gbest) { gbest = k; gsrc = (int)r; }
        }
        const Candidate win = slot[gsrc];
        const int64_t idx = (int64_t)(0x7fffffffu - ((uint32_t)gbest >> 1));
        const bool flip = (gbest & 1ull) != 0ull;     // interaction < 0  (:375)
        const float sgn = flip ? -1.f : 1.f;
        cur = idx;
        px = win.qx; py = win.qy; pz = win.qz;
        p3 = win.nx * sgn; p4 = win.ny * sgn; p5 = win.nz * sgn;
        const int64_t li = idx - base;
        if (li >= 0 && li < n_local && tid == (int)(li % kPtThreads)) {
            sm.visited[li] = 1;
            sm.nx[li] = p3; sm.ny[li] = p4; sm.nz[li] = p5;      // pts[idx, 3:] *= -1  (:376)
            if (order) order[step] = (int32_t)idx;
            if (flipped) flipped[step] = flip ? 1 : 0;
        }
        // the owner thread is the only one that touches point li, so no barrier is needed
        // before the next sweep; the warp_cand slots are protected by the cluster barrier.
    }

    __syncthreads();
    for (int i = tid; i < n_local; i += kPtThreads) {
        float nx = sm.nx[i], ny = sm.ny[i], nz = sm.nz[i];
        if (diffuse) {                                           // :382-385, strict > 0
            const float v = __fadd_rn(__fadd_rn(__fmul_rn(sm.ex[i], nx), __fmul_rn(sm.ey[i], ny)), __fmul_rn(sm.ez[i], nz));
            const float s = v > 0.f ? 1.f : -1.f;
            nx *= s; ny *= s; nz *= s;
        }
        float* p = pts + (base + i) * ld;
        p[3] = nx; p[4] = ny; p[5] = nz;
    }
    cluster_barrier();   // no CTA may exit while a peer can still address its shared memory
}

__global__ void __launch_bounds__(kPtThreads, 1)
points_propagate_kernel(float* __restrict__ pts, int64_t N, int64_t ld, int cap, int64_t start, float eps,
                        int diffuse, int32_t* __restrict__ order, uint8_t* __restrict__ flipped) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    propagate_on_cluster(smem_raw, pts, N, ld, cap, start, eps, diffuse, order, flipped);
}

// Many independent clouds in ONE launch (the per-patch stage of the patch graph, util.py:185-196 with
// graph_dipole.st_propagation_points_file): cluster c owns patch c = rows offsets[c] .. offsets[c+1] of `pts`
// (patches stored contiguously), every patch starts at its first row.  Same code, hence the same bits, as the
// single-cloud kernel; the clusters never talk to each other.
__global__ void __launch_bounds__(kPtThreads, 1)
points_propagate_batched_kernel(float* __restrict__ pts, int64_t ld, const int64_t* __restrict__ offsets, int cap,
                                float eps, int diffuse) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int64_t patch = blockIdx.x / cluster_size();
    const int64_t o = offsets[patch];
    propagate_on_cluster(smem_raw, pts + o * ld, offsets[patch + 1] - o, ld, cap, 0, eps, diffuse, nullptr, nullptr);
}

// Lattice sign test per patch (field_utils.py:7-9 + util.py:26-36 on every patch of a CSR): one CTA per patch,
// thread t owns lattice nodes t, t + 512, ...; mean potential -> mean_out[patch]; with `flip`, the normals of a
// patch whose mean is negative are negated in place (graph_dipole.py:30-31).
__global__ void __launch_bounds__(kPtThreads)
patch_lattice_sign_kernel(float* __restrict__ pts, int64_t ld, const int64_t* __restrict__ offsets, int n_lat,
                          float* __restrict__ mean_out, int flip) {
    constexpr int kTile = 256;
    constexpr int kNodesPerThread = 4;                          // n_lat^3 <= 2048
    __shared__ float s_src[kTile][6];
    __shared__ float s_part[kWarps];
    __shared__ float s_mean;
    const int64_t o = offsets[blockIdx.x], N = offsets[blockIdx.x + 1] - o;
    float* P = pts + o * ld;
    const int M = n_lat * n_lat * n_lat;
    float qx[kNodesPerThread], qy[kNodesPerThread], qz[kNodesPerThread], phi[kNodesPerThread];
#pragma unroll
    for (int k = 0; k < kNodesPerThread; ++k) {
        const int node = threadIdx.x + k * kPtThreads;
        const int iz = node % n_lat, ixy = node / n_lat, iy = ixy % n_lat, ix = ixy / n_lat;
        qx[k] = __fmul_rn(__fsub_rn(__fdiv_rn((float)ix, (float)n_lat), 0.5f), 2.f);
        qy[k] = __fmul_rn(__fsub_rn(__fdiv_rn((float)iy, (float)n_lat), 0.5f), 2.f);
        qz[k] = __fmul_rn(__fsub_rn(__fdiv_rn((float)iz, (float)n_lat), 0.5f), 2.f);
        phi[k] = 0.f;
    }
    for (int64_t t0 = 0; t0 < N; t0 += kTile) {
        const int n_here = (int)min((int64_t)kTile, N - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < n_here * 6; i += kPtThreads) s_src[i / 6][i % 6] = P[(t0 + i / 6) * ld + i % 6];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kNodesPerThread; ++k) {
            if (threadIdx.x + k * kPtThreads >= M) continue;
            float acc = phi[k];
            for (int i = 0; i < n_here; ++i) {
                const float rx = s_src[i][0] - qx[k], ry = s_src[i][1] - qy[k], rz = s_src[i][2] - qz[k];
                const float r2 = fmaf(rz, rz, fmaf(ry, ry, rx * rx));
                const float r = sqrtf(r2);
                const float dot = fmaf(s_src[i][5], rz, fmaf(s_src[i][4], ry, s_src[i][3] * rx));
                acc += (r2 > 0.f) ? dot / (r2 * r) : 0.f;
            }
            phi[k] = acc;
        }
    }
    float total = 0.f;
#pragma unroll
    for (int k = 0; k < kNodesPerThread; ++k) {
        const bool live = threadIdx.x + k * kPtThreads < M;
        const float v = phi[k];
        total += (live && fabsf(v) <= 3.402823466e38f) ? v : 0.f;   // non-finite node totals -> 0 (field_utils.py:53-54)
    }
    total = warp_sum(total);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = total;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < kWarps; ++w) t += s_part[w];
        s_mean = t / (float)M;
        mean_out[blockIdx.x] = s_mean;
    }
    __syncthreads();
    if (flip && s_mean < 0.f) {
        for (int64_t i = threadIdx.x; i < N; i += kPtThreads) {
            float* q = P + i * ld;
            q[3] = -q[3]; q[4] = -q[4]; q[5] = -q[5];
        }
    }
}


// ------------------------------------------------------------------ N beyond one cluster: cooperative grid
// Same algorithm and the same arithmetic (sweep_points), but the point state is spread over the shared
// memory of up to one CTA per SM and a step ends with a grid-wide exchange through L2: every CTA
// publishes its winner in slots[step & 1][cta], arrives on a monotonically increasing counter, and warp 0
// of every CTA picks the global winner once all CTAs of the step have arrived.  The kernel is launched
// cooperatively, so all CTAs are co-resident and the spin cannot deadlock.
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(kPtThreads, 1)
points_propagate_grid_kernel(float* __restrict__ pts, int64_t N, int64_t ld, int cap, int64_t start, float eps,
                             int diffuse, int32_t* __restrict__ order, uint8_t* __restrict__ flipped,
                             Candidate* __restrict__ slots, unsigned int* __restrict__ counter) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PointSmem sm = carve(smem_raw, cap);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rank = blockIdx.x, ncta = gridDim.x;
    const int64_t base = (int64_t)rank * cap;
    const int n_local = (int)max((int64_t)0, min((int64_t)cap, N - base));

    for (int i = tid; i < n_local; i += kPtThreads) {
        const float* p = pts + (base + i) * ld;
        sm.qx[i] = p[0]; sm.qy[i] = p[1]; sm.qz[i] = p[2];
        sm.nx[i] = p[3]; sm.ny[i] = p[4]; sm.nz[i] = p[5];
        sm.ex[i] = 0.f; sm.ey[i] = 0.f; sm.ez[i] = 0.f;
        sm.visited[i] = (base + i == start);
    }
    int64_t cur = start;
    float px, py, pz, p3, p4, p5;
    {
        const float* p = pts + start * ld;
        px = p[0]; py = p[1]; pz = p[2]; p3 = p[3]; p4 = p[4]; p5 = p[5];
    }
    if (rank == 0 && tid == 0) {
        if (order) order[0] = (int32_t)start;
        if (flipped) flipped[0] = 0;
    }
    Candidate* win_smem = sm.cand;      // the cluster exchange buffer is free here: slot 0 holds the step's winner
    __syncthreads();

    for (int64_t step = 1; step <= N; ++step) {
        unsigned long long best = 0ull;
        int best_i = -1;
        sweep_points(sm, n_local, base, cur, px, py, pz, p3, p4, p5, eps, best, best_i);
        if (step == N) break;

        const unsigned long long wbest = warp_max_u64(best);
        if (best_i >= 0 && best == wbest && wbest != 0ull) {
            Candidate c;
            c.key = best;
            c.qx = sm.qx[best_i]; c.qy = sm.qy[best_i]; c.qz = sm.qz[best_i];
            c.nx = sm.nx[best_i]; c.ny = sm.ny[best_i]; c.nz = sm.nz[best_i];
            sm.warp_cand[warp] = c;
        } else if (lane == 0 && wbest == 0ull) {
            sm.warp_cand[warp].key = 0ull;
        }
        __syncthreads();
        Candidate* slot = slots + (size_t)(step & 1) * ncta;
        if (warp == 0) {
            const unsigned long long k = lane < kWarps ? sm.warp_cand[lane].key : 0ull;
            const unsigned long long cbest = warp_max_u64(k);
            const uint32_t src_warp = __ffs(__ballot_sync(0xffffffffu, lane < kWarps && k == cbest)) - 1;
            if (lane == 0) {
                const uint4* w = reinterpret_cast<const uint4*>(&sm.warp_cand[src_warp]);
                uint4 v0 = w[0], v1 = w[1];
                if (cbest == 0ull) { v0.x = 0u; v0.y = 0u; }
                uint4* dst = reinterpret_cast<uint4*>(&slot[rank]);
                __stcg(dst, v0);
                __stcg(dst + 1, v1);
                __threadfence();
                atomicAdd(counter, 1u);
                const unsigned int target = (unsigned int)step * ncta;       // < 2^32: checked by the launcher
                while (ld_acquire_u32(counter) < target) { }
            }
            __syncwarp();
            // the winners of all CTAs: lanes scan them in strides of 32 (L2 reads, never L1)
            unsigned long long gbest = 0ull;
            uint32_t gsrc = 0;
            for (uint32_t r = lane; r < ncta; r += 32) {
                const unsigned long long kk = __ldcg(&slot[r].key);
                if (kk > gbest) { gbest = kk; gsrc = r; }
            }
            const unsigned long long all = warp_max_u64(gbest);
            const uint32_t owner = __ffs(__ballot_sync(0xffffffffu, gbest == all)) - 1;
            if ((uint32_t)lane == owner) {
                const uint4* w = reinterpret_cast<const uint4*>(&slot[gsrc]);
                uint4* d = reinterpret_cast<uint4*>(win_smem);
                d[0] = __ldcg(w);
                d[1] = __ldcg(w + 1);
            }
        }
        __syncthreads();
        const Candidate win = *win_smem;
        const unsigned long long gbest = win.key;
        const int64_t idx = (int64_t)(0x7fffffffu - ((uint32_t)gbest >> 1));
        const bool flip = (gbest & 1ull) != 0ull;
        const float sgn = flip ? -1.f : 1.f;
        cur = idx;
        px = win.qx; py = win.qy; pz = win.qz;
        p3 = win.nx * sgn; p4 = win.ny * sgn; p5 = win.nz * sgn;
        const int64_t li = idx - base;
        if (li >= 0 && li < n_local && tid == (int)(li % kPtThreads)) {
            sm.visited[li] = 1;
            sm.nx[li] = p3; sm.ny[li] = p4; sm.nz[li] = p5;
            if (order) order[step] = (int32_t)idx;
            if (flipped) flipped[step] = flip ? 1 : 0;
        }
        // win_smem is rewritten only after the next step's first __syncthreads, which every thread
        // reaches after it has copied `win` above.
    }

    __syncthreads();
    for (int i = tid; i < n_local; i += kPtThreads) {
        float nx = sm.nx[i], ny = sm.ny[i], nz = sm.nz[i];
        if (diffuse) {
            const float v = __fadd_rn(__fadd_rn(__fmul_rn(sm.ex[i], nx), __fmul_rn(sm.ey[i], ny)), __fmul_rn(sm.ez[i], nz));
            const float sg = v > 0.f ? 1.f : -1.f;
            nx *= sg; ny *= sg; nz *= sg;
        }
        float* p = pts + (base + i) * ld;
        p[3] = nx; p[4] = ny; p[5] = nz;
    }
}

constexpr size_t kGridWsHeader = 256;      // the arrival counter, alone in its sectors

}  // namespace

int64_t points_max_n() {
    // one cluster: 227 KB of shared memory per CTA, 37 bytes of state per point, up to 16 CTAs
    const size_t avail = 227 * 1024 - point_smem_bytes(0);
    return (int64_t)(avail / 37) * kMaxCluster;
}

int64_t points_grid_max_n() {
    // cooperative grid: one CTA per SM, the same 37 bytes of shared memory per point
    const size_t avail = 227 * 1024 - point_smem_bytes(0);
    return (int64_t)(avail / 37) * sm_count_cached();
}

size_t points_workspace_bytes(int64_t N) {
    if (N <= points_max_n()) return 0;                       // the single-cluster kernel needs no scratch
    return kGridWsHeader + 2 * (size_t)sm_count_cached() * sizeof(Candidate);
}

static int run_points_propagate_grid(float* pts, int64_t N, int64_t ld, int64_t start, float eps, int diffuse,
                                     int32_t* order, uint8_t* flipped, void* workspace, size_t workspace_bytes,
                                     cudaStream_t stream) {
    if (N > points_grid_max_n()) {
        set_error("N=%lld exceeds the capacity of dnp_points_propagate_f32 (%lld points on this device)",
                  (long long)N, (long long)points_grid_max_n());
        return DNP_ERR_INVALID;
    }
    const size_t need = points_workspace_bytes(N);
    if (!workspace || workspace_bytes < need) {
        set_error("workspace too small: need %zu bytes (dnp_points_workspace_bytes), got %zu", need, workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    static std::atomic<int> attr_done[kMaxDevices];
    DNP_ONCE_PER_DEVICE(attr_done, DNP_CHECK_CUDA(cudaFuncSetAttribute(
        points_propagate_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024)));
    // about 2048 points per CTA (4 per thread) until every SM has one: fewer CTAs make the per-step exchange cheaper
    const int sms = sm_count_cached();
    int ncta = (int)((N + 2047) / 2048);
    if (ncta > sms) ncta = sms;
    while (point_smem_bytes((int)((N + ncta - 1) / ncta)) > 227 * 1024 && ncta < sms) ++ncta;
    const int cap = (int)((N + ncta - 1) / ncta);
    const size_t smem = point_smem_bytes(cap);
    int per_sm = 0;
    DNP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, points_propagate_grid_kernel, kPtThreads, smem));
    if (per_sm < 1 || (int64_t)per_sm * sms < ncta) {
        set_error("cannot co-schedule %d CTAs for N=%lld", ncta, (long long)N);
        return DNP_ERR_CUDA;
    }
    unsigned int* counter = reinterpret_cast<unsigned int*>(workspace);
    Candidate* slots = reinterpret_cast<Candidate*>(reinterpret_cast<unsigned char*>(workspace) + kGridWsHeader);
    DNP_CHECK_CUDA(cudaMemsetAsync(counter, 0, kGridWsHeader, stream));
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3(ncta);
    cfg.blockDim = dim3(kPtThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    DNP_CHECK_CUDA(cudaLaunchKernelEx(&cfg, points_propagate_grid_kernel, pts, N, ld, cap, start, eps, diffuse, order,
                                      flipped, slots, counter));
    count_launch(1);
    return DNP_OK;
}

int64_t points_batched_max_n() { return points_max_n(); }

int run_points_propagate_batched(float* pts, int64_t ld, const int64_t* offsets, int64_t P, int64_t max_patch,
                                 float eps, int diffuse, cudaStream_t stream) {
    if (P <= 0 || max_patch <= 0) return DNP_OK;
    if (max_patch > points_max_n()) {
        set_error("largest patch (%lld points) exceeds the single-cluster capacity (%lld)", (long long)max_patch,
                  (long long)points_max_n());
        return DNP_ERR_INVALID;
    }
    static std::atomic<int> attr_done[kMaxDevices];
    DNP_ONCE_PER_DEVICE(attr_done, {
        DNP_CHECK_CUDA(cudaFuncSetAttribute(points_propagate_batched_kernel,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        DNP_CHECK_CUDA(cudaFuncSetAttribute(points_propagate_batched_kernel,
                                            cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    });
    // <= 1024 points (2 per thread) per CTA; the cluster size is the same for every patch of the launch
    int ncta = 1;
    while (ncta < kMaxCluster && (int64_t)ncta * kPtThreads * 2 < max_patch) ncta *= 2;
    while (point_smem_bytes((int)((max_patch + ncta - 1) / ncta)) > 227 * 1024 && ncta < kMaxCluster) ncta *= 2;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.blockDim = dim3(kPtThreads);
    cfg.stream = stream;
    int cap = 0;
    for (;;) {
        cap = (int)((max_patch + ncta - 1) / ncta);
        cfg.gridDim = dim3((unsigned)(P * ncta));
        cfg.dynamicSmemBytes = point_smem_bytes(cap);
        attr[0].val.clusterDim.x = ncta;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        int max_clusters = 0;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&max_clusters, points_propagate_batched_kernel, &cfg);
        if (e == cudaSuccess && max_clusters >= 1) break;
        (void)cudaGetLastError();
        if (ncta == 1 || point_smem_bytes((int)((max_patch + ncta / 2 - 1) / (ncta / 2))) > 227 * 1024) {
            set_error("cannot schedule a %d-CTA cluster for patches of %lld points", ncta, (long long)max_patch);
            return DNP_ERR_CUDA;
        }
        ncta /= 2;
    }
    DNP_CHECK_CUDA(cudaLaunchKernelEx(&cfg, points_propagate_batched_kernel, pts, ld, offsets, cap, eps, diffuse));
    count_launch(1);
    return DNP_OK;
}

int run_patch_lattice_sign(float* pts, int64_t ld, const int64_t* offsets, int64_t P, int n_lat, float* mean_out,
                           int flip, cudaStream_t stream) {
    if (P <= 0) return DNP_OK;
    if (n_lat < 1 || n_lat * n_lat * n_lat > 4 * kPtThreads) {
        set_error("lattice size must satisfy n^3 <= %d (got n = %d)", 4 * kPtThreads, n_lat);
        return DNP_ERR_INVALID;
    }
    patch_lattice_sign_kernel<<<(unsigned)P, kPtThreads, 0, stream>>>(pts, ld, offsets, n_lat, mean_out, flip);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

int run_points_propagate(float* pts, int64_t N, int64_t ld, int64_t start, float eps, int diffuse,
                         int32_t* order, uint8_t* flipped, void* workspace, size_t workspace_bytes,
                         cudaStream_t stream) {
    if (N <= 0) return DNP_OK;
    const bool force_grid = getenv("DNP_POINTS_FORCE_GRID") != nullptr;           // tests: grid kernel at any N
    if (N > points_max_n() || (force_grid && workspace))
        return run_points_propagate_grid(pts, N, ld, start, eps, diffuse, order, flipped, workspace, workspace_bytes,
                                         stream);
    static std::atomic<int> attr_done[kMaxDevices];
    DNP_ONCE_PER_DEVICE(attr_done, {
        DNP_CHECK_CUDA(cudaFuncSetAttribute(points_propagate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            227 * 1024));
        DNP_CHECK_CUDA(cudaFuncSetAttribute(points_propagate_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    });
    // smallest power-of-two cluster that gives every thread <= 1 point, capped by what the device can co-schedule
    // (tools/points_probe.py, us per step: N = 2000: 2 CTAs 1.76, 4 CTAs 1.38, 8 CTAs 1.53; N = 5000: 8 CTAs 1.97,
    // 16 CTAs 1.85; N = 10 000: 8 CTAs 2.46, 16 CTAs 2.27 -- a second point per thread costs more than a wider
    // cluster exchange; 1024-thread CTAs were slower still, 2.85 at N = 10 000)
    int ncta = 1;
    while (ncta < kMaxCluster && (int64_t)ncta * kPtThreads < N) ncta *= 2;
    if (const char* forced = getenv("DNP_POINTS_CTAS")) {                       // tuning override (power of two)
        const int f = atoi(forced);
        if (f == 1 || f == 2 || f == 4 || f == 8 || f == 16) ncta = f;
    }
    for (;; ncta *= 2) {
        const int cap = (int)((N + ncta - 1) / ncta);
        if (point_smem_bytes(cap) <= 227 * 1024) break;
        if (ncta >= kMaxCluster) {
            set_error("N=%lld exceeds the single-cluster capacity of dnp_points_propagate_f32 (%lld points)",
                      (long long)N, (long long)points_max_n());
            return DNP_ERR_INVALID;
        }
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.blockDim = dim3(kPtThreads);
    cfg.stream = stream;
    int cap = 0;
    for (;;) {
        cap = (int)((N + ncta - 1) / ncta);
        cfg.gridDim = dim3(ncta);
        cfg.dynamicSmemBytes = point_smem_bytes(cap);
        attr[0].val.clusterDim.x = ncta;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        int max_clusters = 0;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&max_clusters, points_propagate_kernel, &cfg);
        if (e == cudaSuccess && max_clusters >= 1) break;
        (void)cudaGetLastError();
        // a 16-CTA cluster may not be schedulable on every GPC layout: fall back to fewer, fatter CTAs
        if (ncta == 1 || point_smem_bytes((int)((N + ncta / 2 - 1) / (ncta / 2))) > 227 * 1024) {
            set_error("cannot schedule a %d-CTA cluster for N=%lld", ncta, (long long)N);
            return DNP_ERR_CUDA;
        }
        ncta /= 2;
    }
    DNP_CHECK_CUDA(cudaLaunchKernelEx(&cfg, points_propagate_kernel, pts, N, ld, cap, start, eps, diffuse, order,
                                      flipped));
    count_launch(1);
    return DNP_OK;
}

}  // namespace dnp
