// dnp_loop.cu -- the greedy patch-propagation loop of a SMALL cloud as one persistent cooperative kernel
// (field_utils.py:239-265, :311-335).
//
// At fandisk size (72 patches of 100-268 rows, 11 031 points) one step of the loop is 3e6 pair interactions,
// two microseconds of arithmetic, and the multi-launch schedule of dnp_drivers.cu spends 27 us on it: three or
// four dependent launches of ~7 us each.  Here the whole loop is ONE launch of up to one CTA per SM:
//   * the rows of the patch CSR are dealt to the CTAs in chunks of 8, round robin, so that every CTA holds a
//     sample of the whole cloud and the rows still alive at a step are spread evenly however far the loop has got;
//     a CTA keeps its rows' positions, normals and running field E in shared memory for the whole loop (nothing
//     of the loop state lives in HBM),
//   * per step every CTA stages the chosen patch's rows (flipped if its interaction was negative) as packed sources
//     in shared memory, adds their field to its own rows (a thread owns a PAIR of rows and every L-th source,
//     through the packed f32x2 pair arithmetic of the all-pairs kernel, fixed-order reduction; the rows still
//     alive are compacted first, so visited rows cost nothing and free their lanes), and sums E.n over its rows of
//     every unvisited patch it holds a part of (one warp per run of rows, fixed order),
//   * ONE grid-wide barrier per step (a monotonically increasing arrival counter in L2), after which every CTA
//     adds up the partial interactions of every patch in chunk order and takes the first abs-arg-max itself --
//     no winner broadcast.
// Preconditions (checked by the launcher, otherwise the multi-launch schedule runs): the patches are disjoint and
// cover the target rows, the largest patch has <= 1024 rows, <= 2048 rows per CTA (303 000 rows), P <= 8192, and,
// when diffuse, <= 1.5e8 pairs per step.
// Deterministic: no float atomics, every sum in a fixed order.  The running field is summed in a different order
// than the all-pairs kernel's, so E differs from the multi-launch schedule in the last bits; the decisions
// (visit order, flips) are pinned by the reference's golden fixtures for both.
#include <stdio.h>
#include <stdlib.h>

#include <atomic>

#include "dnp_common.cuh"
#include "dnp_pairmath.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"
#include "dipole.h"

namespace dnp {
namespace {

constexpr int kLoopThreads = 1024;
constexpr int kLoopWarps = kLoopThreads / 32;
constexpr int64_t kLoopMaxRows = 2048;      // target rows of one CTA
constexpr int64_t kLoopMaxSrc = 1024;       // rows of the largest patch
constexpr int64_t kLoopMaxPatches = 8192;
// pairs per step beyond which the all-pairs kernel's throughput wins (tools/loop_size_probe.py): when diffuse every
// row is alive at every step; otherwise half of them on average, compacted away here but still walked by the launches
constexpr double kLoopMaxPairsDiffuse = 1.5e8;
constexpr double kLoopMaxPairs = 1.0e9;

__device__ __forceinline__ unsigned int loop_ld_acquire(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ float finite_or_zero(float v) { return fabsf(v) <= 3.402823466e38f ? v : 0.f; }

// every CTA arrives once per step; `target` = arrivals expected so far
__device__ __forceinline__ void grid_arrive_wait(unsigned int* counter, unsigned int target) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicAdd(counter, 1u);
        while (loop_ld_acquire(counter) < target) { }
        __threadfence();
    }
    __syncthreads();
}

struct LoopArgs {
    float* pts;                     // [N, ld] xyz + normals; normals of flipped patches are negated in place
    int64_t ld;
    const int64_t* perm;            // CSR rows of the patches
    const int64_t* offsets;         // device, P + 1
    int P;
    const int64_t* rest_perm;       // optional: rows flipped together with a patch
    const int64_t* rest_offsets;    // device
    int start;
    float eps;
    int diffuse;
    float* E;                       // [N, ld_e] final field of the patch rows
    int64_t ld_e;
    float* inter;                   // device, 2 * P * G: partial interactions [p][c], double-buffered by step parity
    unsigned int* counter;          // zeroed by the launcher
    int32_t* order;                 // optional device traces, P entries
    float* value;
    int cap_rows, cap_src;          // shared-memory capacities this launch was sized for
    int lanes;                      // > 0: forced threads per target row (DNP_LOOP_LANES)
    long long* prof;                // optional device array of 8 cycle counters (DNP_LOOP_DEBUG & 4), CTA 0
    int debug;                      // DNP_LOOP_DEBUG=4: print CTA 0's cycles per phase to stderr
};

// Rows are dealt to the CTAs in chunks of kChunk consecutive CSR rows, round robin: chunk q belongs to CTA q % G and
// is its local chunk q / G.  Every CTA then holds a sample of the whole cloud, so the rows still alive at a step
// (the unvisited patches) are spread evenly over the CTAs however far the loop has got.
constexpr int kChunk = 8;

__global__ void __launch_bounds__(kLoopThreads, 1) patch_loop_kernel(LoopArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* src = reinterpret_cast<float4*>(smem_raw);                               // cap_src * 3
    float* tq = reinterpret_cast<float*>(src + (size_t)a.cap_src * 3);               // cap_rows * 3 each
    float* tn = tq + (size_t)a.cap_rows * 3;
    float* tE = tn + (size_t)a.cap_rows * 3;
    float* dots = tE + (size_t)a.cap_rows * 3;                                       // cap_rows
    int32_t* tpatch = reinterpret_cast<int32_t*>(dots + a.cap_rows);                 // cap_rows, non-decreasing
    int32_t* seg_lb = tpatch + a.cap_rows;                                           // cap_rows: my runs of one patch
    int32_t* seg_le = seg_lb + a.cap_rows;
    uint16_t* live = reinterpret_cast<uint16_t*>(seg_le + a.cap_rows);               // cap_rows: the rows alive this step
    uint8_t* tstate = reinterpret_cast<uint8_t*>(live + a.cap_rows);                 // cap_rows: 1 = visited
    uint8_t* visited = tstate + a.cap_rows;                                          // P
    __shared__ float part[kLoopThreads * 6];                                          // per-thread partial fields (a row pair)
    __shared__ unsigned long long warp_best[32];
    __shared__ float warp_val[32];
    __shared__ int32_t sh_next[2];
    __shared__ int sh_nseg, sh_nlive[2];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int G = (int)gridDim.x, c = (int)blockIdx.x;
    const int64_t* __restrict__ off = a.offsets;
    const int64_t T = off[a.P];
    const int64_t NQ = (T + kChunk - 1) / kChunk;                                    // chunks of the CSR
    const int nk = NQ > c ? (int)((NQ - 1 - c) / G) + 1 : 0;                         // my chunks
    int R = nk * kChunk;                                                             // my rows: local index li <->
    if (nk > 0 && (NQ - 1) % G == c) R -= (int)(NQ * kChunk - T);                    // CSR row ((li/8) G + c) 8 + li%8
#define LOOP_GLOBAL_ROW(li) ((((int64_t)((li) / kChunk) * G + c) * kChunk) + ((li) % kChunk))

    if (tid == 0) { sh_nseg = 0; sh_nlive[0] = 0; sh_nlive[1] = 0; }
    for (int li = tid; li < R; li += kLoopThreads) {
        const int64_t i = LOOP_GLOBAL_ROW(li);
        const float* q = a.pts + a.perm[i] * a.ld;
        tq[3 * li] = q[0]; tq[3 * li + 1] = q[1]; tq[3 * li + 2] = q[2];
        tn[3 * li] = q[3]; tn[3 * li + 1] = q[4]; tn[3 * li + 2] = q[5];
        tE[3 * li] = 0.f; tE[3 * li + 1] = 0.f; tE[3 * li + 2] = 0.f;
        tstate[li] = 0;
        int lo = 0, hi = a.P;                                                        // the patch of row i:
        while (hi - lo > 1) {                                                        // last p with off[p] <= i
            const int mid = (lo + hi) >> 1;
            if (off[mid] <= i) lo = mid; else hi = mid;
        }
        tpatch[li] = lo;
    }
    for (int p = tid; p < a.P; p += kLoopThreads) visited[p] = 0;
    __syncthreads();
    // my runs of rows of one patch (local order follows CSR order, so a patch's rows here are one run)
    for (int li = tid; li < R; li += kLoopThreads) {
        if (li == 0 || tpatch[li] != tpatch[li - 1]) {
            const int p = tpatch[li];
            int lo = li, hi = R;                                                     // first index > li with another patch
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (tpatch[mid] == p) lo = mid; else hi = mid;
            }
            const int s = atomicAdd(&sh_nseg, 1);
            seg_lb[s] = li;
            seg_le[s] = hi;
        }
    }
    __syncthreads();
    const int nseg = sh_nseg;
    int LP = 32;                                                                     // lanes per patch in the arg-max:
    while (LP > 1 && (int64_t)a.P * LP > kLoopThreads) LP >>= 1;                     // one pass over the patches if possible

    const f32x2 eps2 = pack2(a.eps, a.eps);
    const f32x2 zero2 = pack2(0.f, 0.f);
    int j = a.start, flip = 0;
    const bool prof = a.prof && blockIdx.x == 0 && tid == 0;
    long long t0 = prof ? clock64() : 0, acc[6] = {0, 0, 0, 0, 0, 0};
#define LOOP_MARK(k) if (prof) { const long long t1 = clock64(); acc[k] += t1 - t0; t0 = t1; }
    for (int step = 0; step < a.P; ++step) {
        if (blockIdx.x == 0 && tid == 0) {
            if (a.order) a.order[step] = j;
            if (step == 0 && a.value) a.value[0] = 0.f;
        }
        if (tid == 0) visited[j] = 1;
        // ---- the chosen patch as packed sources (its commit: flipped if the interaction was negative, :250,322)
        const int64_t so = off[j];
        const int n = (int)(off[j + 1] - so);
        for (int i = tid; i < n; i += kLoopThreads) {
            float v[6];
            const float* q = a.pts + a.perm[so + i] * a.ld;
#pragma unroll
            for (int k = 0; k < 6; ++k) v[k] = q[k];
            if (flip) { v[3] = -v[3]; v[4] = -v[4]; v[5] = -v[5]; }
            pack_source_row(v, src + 3 * i);
        }
        // my rows of the chosen patch: the chunks qa .. qb of the patch that are mine, qm0, qm0 + G, ...
        const int qa = (int)(so / kChunk), qb = n > 0 ? (int)((so + n - 1) / kChunk) : qa - 1;
        const int qm0 = qa + (((c - qa) % G) + G) % G;
        const int nm = qm0 <= qb ? (qb - qm0) / G + 1 : 0;
        for (int idx = tid; idx < nm * kChunk; idx += kLoopThreads) {
            const int q = qm0 + (idx / kChunk) * G;
            const int64_t i = (int64_t)q * kChunk + idx % kChunk;
            if (i >= so && i < so + n) tstate[(q / G) * kChunk + idx % kChunk] = 1;
        }
        if (flip && a.rest_perm && n > 0 && qa % G == c) {                        // :250-252: the rest of the patch follows
            const int64_t ro = a.rest_offsets[j], rn = a.rest_offsets[j + 1] - ro;
            for (int64_t i = tid; i < rn; i += kLoopThreads) {
                float* q = a.pts + a.rest_perm[ro + i] * a.ld + 3;
                q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2];
            }
        }
        // the rows alive this step, compacted (so that dead rows cost nothing and free their lanes): not of the
        // chosen patch and, unless diffuse, not of a patch visited before
        for (int base = 0; base < R; base += kLoopThreads) {
            const int t = base + tid;
            const bool f = t < R && tpatch[t] != j && (a.diffuse || !tstate[t]);
            const unsigned m = __ballot_sync(0xffffffffu, f);
            int wbase = 0;
            if (lane == 0 && m) wbase = atomicAdd(&sh_nlive[step & 1], __popc(m));
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            if (f) live[wbase + __popc(m & ((1u << lane) - 1u))] = (uint16_t)t;
        }
        __syncthreads();
        LOOP_MARK(0)
        // ---- E += field of the patch on my rows (:261,265,330,335); visited rows are skipped, or, when diffuse,
        // only the source patch itself (:256-261,326-331).  A thread owns a PAIR of rows (the two halves of the
        // packed arithmetic) and every L-th source; consecutive threads are consecutive row pairs, so a warp reads
        // one or two sources at a time (broadcast shared-memory loads).  The L partial fields of a row are then
        // summed in lane order.
        const int nlive = sh_nlive[step & 1];
        if (tid == 0) sh_nlive[(step + 1) & 1] = 0;                               // next step's counter
        const int pairs = (nlive + 1) >> 1;
        int L = pairs > 0 ? kLoopThreads / pairs : 1;
        const int l_cap = (n + 3) >> 2;
        L = L > l_cap ? l_cap : L;
        L = L < 1 ? 1 : L;
        if (a.lanes > 0 && a.lanes * pairs <= kLoopThreads) L = a.lanes;
        for (int pbase = 0; pbase < pairs; pbase += kLoopThreads) {               // one pass unless L == 1
            const int span = pairs - pbase < kLoopThreads ? pairs - pbase : kLoopThreads;
            const int tp = tid % span, l = tid / span;
            const bool on = l < L && (L == 1 ? tid < span : true);
            f32x2 ax = zero2, ay = zero2, az = zero2;
            if (on) {
                const int m0 = 2 * (pbase + tp);
                const int t0r = live[m0], t1r = m0 + 1 < nlive ? live[m0 + 1] : t0r;
                const f32x2 qx = pack2(tq[3 * t0r], tq[3 * t1r]), qy = pack2(tq[3 * t0r + 1], tq[3 * t1r + 1]),
                            qz = pack2(tq[3 * t0r + 2], tq[3 * t1r + 2]);
#pragma unroll 2
                for (int s = l; s < n; s += L) {
                    const SrcX2 one(src[3 * s], src[3 * s + 1], src[3 * s + 2]);
                    pair_field_x2(one, qx, qy, qz, eps2, ax, ay, az);
                }
                float* o = part + 6 * tid;
                unpack2(ax, o[0], o[1]); unpack2(ay, o[2], o[3]); unpack2(az, o[4], o[5]);
            }
            __syncthreads();
            const int rows_here = (2 * (pbase + span) < nlive ? 2 * (pbase + span) : nlive) - 2 * pbase;
            for (int m = tid; m < rows_here; m += kLoopThreads) {                 // m: position among this pass's live rows
                const float* o = part + 6 * (m >> 1) + (m & 1);
                float ex = 0.f, ey = 0.f, ez = 0.f;
                for (int l2 = 0; l2 < L; ++l2) {                                  // the row's lanes, in lane order
                    const float* ol = o + (size_t)6 * span * l2;
                    ex += ol[0]; ey += ol[2]; ez += ol[4];
                }
                const int t = live[2 * pbase + m];
                const float Ex = tE[3 * t] + finite_or_zero(ex), Ey = tE[3 * t + 1] + finite_or_zero(ey),
                            Ez = tE[3 * t + 2] + finite_or_zero(ez);              // field_utils.py:114-115
                tE[3 * t] = Ex; tE[3 * t + 1] = Ey; tE[3 * t + 2] = Ez;
                // E.n of the row for the interactions below (only live rows belong to unvisited patches)
                dots[t] = __fadd_rn(__fadd_rn(__fmul_rn(Ex, tn[3 * t]), __fmul_rn(Ey, tn[3 * t + 1])),
                                    __fmul_rn(Ez, tn[3 * t + 2]));
            }
            __syncthreads();
        }
        LOOP_MARK(1)
        // ---- my share of sum E.n of every unvisited patch I hold rows of (:244-247,316-319): slot [p][c] of this
        // step's buffer
        float* part_inter = a.inter + (size_t)(step & 1) * a.P * G;
        if (step + 1 < a.P) {
            for (int s = warp; s < nseg; s += kLoopWarps) {
                const int b = seg_lb[s], e = seg_le[s], p = tpatch[b];
                if (visited[p]) continue;
                float sum = 0.f;
                for (int i = b + lane; i < e; i += 32) sum += dots[i];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                if (lane == 0) __stcg(part_inter + (size_t)p * G + c, sum);
            }
        }
        LOOP_MARK(2)
        grid_arrive_wait(a.counter, (unsigned int)(step + 1) * (unsigned int)G);
        LOOP_MARK(3)
        // every CTA is done reading the sources: commit the flip of my rows of the patch to the cloud
        if (flip) {
            for (int idx = tid; idx < nm * kChunk; idx += kLoopThreads) {
                const int64_t i = (int64_t)(qm0 + (idx / kChunk) * G) * kChunk + idx % kChunk;
                if (i >= so && i < so + n) {
                    float* nrm = a.pts + a.perm[i] * a.ld + 3;
                    nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2];
                }
            }
        }
        if (step + 1 == a.P) break;
        // ---- I_p = the partial sums of the CTAs that hold chunks of p, in chunk order, then the first abs-arg-max
        // over the unvisited patches -- by every CTA for itself, with identical arithmetic
        unsigned long long best = 0ull;
        float best_v = 0.f;
        for (int base = 0; base < a.P; base += kLoopThreads / LP) {                // LP lanes of a warp per patch
            const int p = base + tid / LP, lp = tid % LP;
            const bool on = p < a.P && !visited[p];
            float v = 0.f;
            if (on) {
                const int64_t b = off[p], e = off[p + 1];
                if (e > b) {
                    const int pa = (int)(b / kChunk), pq = (int)((e - 1) / kChunk) - pa + 1;
                    const int cnt = pq < G ? pq : G;
                    const int c0 = pa % G;
                    const float* row = part_inter + (size_t)p * G;
                    // chunk order, strided by lane; four loads in flight (an L2 round trip each)
                    for (int i = lp; i < cnt; i += 4 * LP) {
                        float w[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int ii = i + u * LP, cc = c0 + ii;
                            w[u] = ii < cnt ? __ldcg(row + (cc >= G ? cc - G : cc)) : 0.f;
                        }
                        v += w[0]; v += w[1]; v += w[2]; v += w[3];
                    }
                }
            }
            for (int o = LP >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (on && lp == 0) {
                const unsigned long long k =
                    ((unsigned long long)__float_as_uint(fabsf(v)) << 32) | (0xffffffffu - (uint32_t)p);
                if (k > best) { best = k; best_v = v; }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            const float other_v = __shfl_xor_sync(0xffffffffu, best_v, o);
            if (other > best) { best = other; best_v = other_v; }
        }
        if (lane == 0) { warp_best[warp] = best; warp_val[warp] = best_v; }
        __syncthreads();
        if (warp == 0) {
            best = lane < kLoopWarps ? warp_best[lane] : 0ull;
            best_v = lane < kLoopWarps ? warp_val[lane] : 0.f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
                const float other_v = __shfl_xor_sync(0xffffffffu, best_v, o);
                if (other > best) { best = other; best_v = other_v; }
            }
            if (lane == 0) {
                sh_next[0] = (int32_t)(0xffffffffu - (uint32_t)best);             // best != 0: a patch is left
                sh_next[1] = best_v < 0.f;
                if (blockIdx.x == 0 && a.value) a.value[step + 1] = best_v;
            }
        }
        __syncthreads();
        j = sh_next[0];
        flip = sh_next[1];
        LOOP_MARK(4)
    }
#undef LOOP_MARK
    if (prof)
        for (int k = 0; k < 6; ++k) a.prof[k] = acc[k];
    // ---- the final field of my rows
    for (int li = tid; li < R; li += kLoopThreads) {
        float* o = a.E + a.perm[LOOP_GLOBAL_ROW(li)] * a.ld_e;
        o[0] = tE[3 * li]; o[1] = tE[3 * li + 1]; o[2] = tE[3 * li + 2];
    }
#undef LOOP_GLOBAL_ROW
}

// distinct rows named by the CSR (disjoint patches <=> distinct == T) and rows out of range
__global__ void loop_cover_kernel(const int64_t* __restrict__ perm, int64_t T, int64_t N, int* __restrict__ flags,
                                  int* __restrict__ counts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T) return;
    const int64_t r = perm[i];
    if (r < 0 || r >= N) { atomicAdd(counts + 1, 1); return; }
    if (atomicExch(flags + r, 1) == 0) atomicAdd(counts, 1);
}

// rest rows that are ALSO patch (representative) rows: the loop caches every CTA's normals in shared memory once, and
// a "rest" flip only reaches the cloud in global memory -- with such an overlap a later patch would be scored from
// stale normals (the reference flips pts[rest_patch] at once, field_utils.py:250-252), so the call is refused
__global__ void loop_rest_overlap_kernel(const int64_t* __restrict__ rest_perm, int64_t R, int64_t N,
                                         const int* __restrict__ flags, int* __restrict__ counts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const int64_t r = rest_perm[i];
    if (r < 0 || r >= N) { atomicAdd(counts + 1, 1); return; }
    if (flags[r]) atomicAdd(counts + 2, 1);
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

size_t loop_smem_bytes(int cap_rows, int cap_src, int P) {
    return (size_t)cap_src * 48 + (size_t)cap_rows * (10 * 4 + 3 * 4 + 2 + 1) + (size_t)P + 64;
}

}  // namespace

size_t patch_loop_workspace_bytes(int64_t N, int64_t P) {
    return up256((size_t)N * 4) + 256 + up256((size_t)2 * P * sm_count_cached() * 4 + 64) + up256((size_t)P * 4) * 2;
}

int check_patch_cover(const int64_t* perm, int64_t T, int64_t N, const int64_t* rest_rows, int64_t R, int* flags,
                      int* counts, cudaStream_t stream, int* ok) {
    *ok = 0;
    DNP_CHECK_CUDA(cudaMemsetAsync(flags, 0, (size_t)N * 4, stream));
    DNP_CHECK_CUDA(cudaMemsetAsync(counts, 0, 256, stream));
    loop_cover_kernel<<<(unsigned)((T + 255) / 256), 256, 0, stream>>>(perm, T, N, flags, counts);
    count_launch(1);
    if (R > 0) {
        loop_rest_overlap_kernel<<<(unsigned)((R + 255) / 256), 256, 0, stream>>>(rest_rows, R, N, flags, counts);
        count_launch(1);
    }
    int host_counts[3] = {0, 0, 0};
    DNP_CHECK_CUDA(cudaMemcpyAsync(host_counts, counts, sizeof(host_counts), cudaMemcpyDeviceToHost, stream));
    DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
    if (host_counts[1] != 0) {
        set_error("patch rows out of range [0, %lld)", (long long)N);
        return DNP_ERR_INVALID;
    }
    *ok = host_counts[0] == T && host_counts[2] == 0;
    return DNP_OK;
}

// Returns DNP_OK and sets *taken = 1 when the loop ran here; *taken = 0 (nothing launched that matters) when the
// preconditions do not hold and the caller has to run the multi-launch schedule.
int run_patch_loop_persistent(float* pts, int64_t N, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                              const int64_t* offsets_host, int64_t P, const int64_t* rest_perm,
                              const int64_t* rest_offsets_host, const int64_t* tgt_idx, int64_t n_tgt,
                              int64_t start_patch, float eps, int diffuse, float* E, int64_t ld_e, int32_t* order_host,
                              float* value_host, void* workspace, size_t workspace_bytes, cudaStream_t stream,
                              int* taken) {
    *taken = 0;
    {   // the precondition check reads 8 bytes back: not on a capturing stream (the device-driven launches capture)
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) { cudaGetLastError(); return DNP_OK; }
        if (st != cudaStreamCaptureStatusNone) return DNP_OK;
    }
    const int64_t T = offsets_host[P] - offsets_host[0];
    if (P < 1 || P > kLoopMaxPatches || offsets_host[0] != 0 || T < 1) return DNP_OK;
    if (tgt_idx ? (tgt_idx != perm || n_tgt != T) : T != N) return DNP_OK;       // the patches must be the targets
    int64_t max_patch = 0;
    for (int64_t p = 0; p < P; ++p) {
        const int64_t n = offsets_host[p + 1] - offsets_host[p];
        if (n > max_patch) max_patch = n;
    }
    const char* cap_env = getenv("DNP_LOOP_MAX_PAIRS");                               // tuning override (tools/loop_probe.py)
    const double max_pairs = cap_env ? atof(cap_env) : (diffuse ? kLoopMaxPairsDiffuse : kLoopMaxPairs);
    if (max_patch > kLoopMaxSrc || (double)max_patch * (double)T > max_pairs) return DNP_OK;
    if (workspace_bytes < patch_loop_workspace_bytes(N, P)) return DNP_OK;
    // chunks of 8 CSR rows dealt round robin to the CTAs; at least 4 chunks each (fewer CTAs: a cheaper barrier)
    const int sms = sm_count_cached();
    const int64_t NQ = (T + kChunk - 1) / kChunk;
    int G = (int)(NQ / 4 < sms ? NQ / 4 : sms);
    if (G < 1) G = 1;
    const int64_t max_rows = (NQ + G - 1) / G * kChunk;
    if (max_rows > kLoopMaxRows) return DNP_OK;

    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    size_t off = 0;
    int* flags = reinterpret_cast<int*>(ws + off); off += up256((size_t)N * 4);
    int* counts = reinterpret_cast<int*>(ws + off);                                   // [0] distinct, [1] out of range
    unsigned int* counter = reinterpret_cast<unsigned int*>(ws + off + 128); off += 256;
    float* inter = reinterpret_cast<float*>(ws + off); off += up256((size_t)2 * P * sms * 4 + 64);
    int32_t* order_dev = reinterpret_cast<int32_t*>(ws + off); off += up256((size_t)P * 4);
    float* value_dev = reinterpret_cast<float*>(ws + off); off += up256((size_t)P * 4);

    // disjoint patches covering the targets?  (one 8-byte read per CALL)
    DNP_CHECK_CUDA(cudaMemsetAsync(flags, 0, (size_t)N * 4, stream));
    DNP_CHECK_CUDA(cudaMemsetAsync(counts, 0, 256, stream));
    loop_cover_kernel<<<(unsigned)((T + 255) / 256), 256, 0, stream>>>(perm, T, N, flags, counts);
    count_launch(1);
    const int64_t R = rest_perm ? rest_offsets_host[P] - rest_offsets_host[0] : 0;
    if (R > 0) {
        loop_rest_overlap_kernel<<<(unsigned)((R + 255) / 256), 256, 0, stream>>>(rest_perm + rest_offsets_host[0], R, N,
                                                                                  flags, counts);
        count_launch(1);
    }
    int host_counts[3] = {0, 0, 0};
    DNP_CHECK_CUDA(cudaMemcpyAsync(host_counts, counts, sizeof(host_counts), cudaMemcpyDeviceToHost, stream));
    DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
    if (host_counts[1] != 0) {
        set_error("patch rows out of range [0, %lld)", (long long)N);
        return DNP_ERR_INVALID;
    }
    if (host_counts[0] != T) return DNP_OK;                                           // overlapping patches
    if (host_counts[2] != 0) return DNP_OK;                                           // rest rows among the patch rows

    // rest offsets on the device (they arrive on the host): reuse the flags area, no longer needed
    int64_t* rest_off_dev = nullptr;
    if (rest_perm) {
        if ((size_t)(P + 1) * 8 > up256((size_t)N * 4)) return DNP_OK;
        rest_off_dev = reinterpret_cast<int64_t*>(flags);
        DNP_CHECK_CUDA(cudaMemcpyAsync(rest_off_dev, rest_offsets_host, (size_t)(P + 1) * 8, cudaMemcpyHostToDevice,
                                       stream));
    }

    LoopArgs a;
    a.pts = pts; a.ld = ld; a.perm = perm; a.offsets = offsets_dev; a.P = (int)P;
    a.rest_perm = rest_perm; a.rest_offsets = rest_off_dev;
    a.start = (int)start_patch; a.eps = eps; a.diffuse = diffuse; a.E = E; a.ld_e = ld_e;
    a.inter = inter; a.counter = counter;
    a.order = order_host ? order_dev : nullptr;
    a.value = value_host ? value_dev : nullptr;
    a.cap_rows = (int)((max_rows + 3) & ~(int64_t)3);
    a.cap_src = (int)((max_patch + 1) & ~(int64_t)1);
    if (a.cap_src < 2) a.cap_src = 2;
    a.lanes = 0;
    a.debug = getenv("DNP_LOOP_DEBUG") ? atoi(getenv("DNP_LOOP_DEBUG")) : 0;
    a.prof = (a.debug & 4) ? reinterpret_cast<long long*>(reinterpret_cast<unsigned char*>(inter) +
                                                           (((size_t)2 * P * sms * 4 + 7) & ~(size_t)7))
                           : nullptr;
    if (getenv("DNP_LOOP_LANES")) a.lanes = atoi(getenv("DNP_LOOP_LANES"));
    const size_t smem = loop_smem_bytes(a.cap_rows, a.cap_src, (int)P);
    static std::atomic<int> smem_opted[64];
    int dev = 0;
    DNP_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !smem_opted[dev].load()) {
        DNP_CHECK_CUDA(cudaFuncSetAttribute(patch_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        smem_opted[dev].store(1);
    }
    if (smem > 200 * 1024) return DNP_OK;
    int per_sm = 0;
    DNP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, patch_loop_kernel, kLoopThreads, smem));
    if (per_sm < 1 || (int64_t)per_sm * sms < G) return DNP_OK;

    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3((unsigned)G);
    cfg.blockDim = dim3(kLoopThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    DNP_CHECK_CUDA(cudaLaunchKernelEx(&cfg, patch_loop_kernel, a));
    count_launch(1);
    *taken = 1;
    if (a.prof) {
        long long h[6];
        DNP_CHECK_CUDA(cudaMemcpyAsync(h, a.prof, sizeof(h), cudaMemcpyDeviceToHost, stream));
        DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
        fprintf(stderr, "[patch_loop] cycles per step, CTA 0: stage %.0f field %.0f interactions %.0f barrier %.0f argmax %.0f (P=%lld G=%d rows/CTA<=%lld)\n",
                h[0] / (double)P, h[1] / (double)P, h[2] / (double)P, h[3] / (double)P, h[4] / (double)P, (long long)P, G, (long long)max_rows);
    }
    if (order_host)
        DNP_CHECK_CUDA(cudaMemcpyAsync(order_host, order_dev, (size_t)P * 4, cudaMemcpyDeviceToHost, stream));
    if (value_host)
        DNP_CHECK_CUDA(cudaMemcpyAsync(value_host, value_dev, (size_t)P * 4, cudaMemcpyDeviceToHost, stream));
    if (order_host || value_host) DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
