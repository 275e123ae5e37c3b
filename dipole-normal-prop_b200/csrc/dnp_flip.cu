// dnp_flip.cu -- patch-graph flip solve.
//
// Replaces graph.MIQP(A, B) (graph.py:444-470), which hands
//     maximise  sum_ij A_ij (1 - (x_i - x_j)^2) + B_ij (x_i - x_j)^2 ,  x in {0,1}^n     (graph.py:431-442)
// to Gurobi.  With s_i = 1 - 2 x_i the objective is  C0 + sum_{i<j} J_ij s_i s_j ,
//     J_ij = (A_ij - B_ij + A_ji - B_ji) / 2 ,  C0 = sum_{i!=j} (A_ij + B_ij)/2 + sum_i A_ii ,
// i.e. the ground state of an Ising model whose couplings are the patch flip scores.
//
//   * n <= exact_max_n (<= 40): exhaustive.  One spin is pinned (global flip symmetry); a thread owns one
//     pattern of the high spins and walks the 2^10 states of the 10 low spins in Gray-code order, fully
//     unrolled: the flipped bit, its new sign and the state of all lower bits at that moment are compile-time
//     constants, so a state costs one FADD for the energy, ~1 amortised FADD for the local fields and one FMNMX
//     for the running maximum; the position of the maximum is recovered afterwards by walking the winning
//     patterns once more (flip_resolve_kernel).  Runs of up to 64 high patterns are themselves visited in Gray
//     order with incremental updates.  Issue-bound, memory-free walk.
//   * larger n: simulated annealing, one warp per replica (spins and local fields in shared memory, exact
//     sequential Metropolis order inside a warp: the 32 lanes test 32 spins at once and the first accepted
//     one is applied before the rest are re-tested), then a zero-temperature descent.  Counter-based hash
//     RNG: the result is a pure function of (A, B, seed).
//   * flip_select_kernel re-evaluates the best candidates in double from A and B, picks the maximum
//     (ties: lowest candidate), canonicalises the global sign (fewer flipped patches; tie: x_0 = 0) and
//     reports the bound C0 + sum |J|: a solution that reaches it is optimal whatever found it.
#include <math.h>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"
#include "dipole.h"

namespace dnp {
namespace {

constexpr int kLow = 10;                 // spins walked by the unrolled Gray code
constexpr int kLowStates = 1 << kLow;
constexpr int kExactMax = 40;            // spins of the exhaustive path (incl. the pinned one)
constexpr int kExactMin = kLow + 1;      // smaller problems are padded with uncoupled spins
constexpr int kTopK = 32;                // candidates re-evaluated in double
constexpr int kAnnealWarps = 4;          // replicas per CTA
constexpr int kMaxN = 4096;

__device__ __forceinline__ constexpr int ctz_const(int t) {
    int b = 0;
    while (!((t >> b) & 1)) ++b;
    return b;
}

__device__ __forceinline__ double jd(const double* A, const double* B, int64_t ld, int i, int j) {
    return 0.5 * ((A[i * ld + j] - B[i * ld + j]) + (A[j * ld + i] - B[j * ld + i]));
}

// ------------------------------------------------------------------ couplings, constant term, bound, scale
// J (float, [np x np], zero diagonal, zero padding), info[0] = C0, info[1] = C0 + sum_{i<j} |J_ij|,
// scale[0] = max_i sum_j |J_ij|.  One CTA, fixed reduction order (deterministic).
__global__ void __launch_bounds__(1024) flip_prepare_kernel(const double* __restrict__ A, const double* __restrict__ B,
                                                            int n, int64_t ld, int np, float* __restrict__ J,
                                                            double* __restrict__ info, float* __restrict__ scale) {
    __shared__ double sh_c[32], sh_u[32];
    __shared__ float sh_r[32];
    double c0 = 0.0, ub = 0.0;
    float rowmax = 0.f;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int i = warp; i < np; i += nwarps) {                // one warp per row
        double rc = 0.0, ru = 0.0;
        for (int j = lane; j < np; j += 32) {
            float v = 0.f;
            if (i < n && j < n) {
                if (i == j) {
                    rc += A[i * ld + i];
                } else {
                    const double d = jd(A, B, ld, i, j);
                    rc += 0.5 * (A[i * ld + j] + B[i * ld + j]);
                    ru += fabs(d);
                    v = (float)d;
                }
            }
            J[(int64_t)i * np + j] = v;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            rc += __shfl_xor_sync(0xffffffffu, rc, o);
            ru += __shfl_xor_sync(0xffffffffu, ru, o);
        }
        c0 += rc;
        ub += 0.5 * ru;                                       // every pair is seen from both rows
        rowmax = fmaxf(rowmax, (float)ru);
    }
    if (lane == 0) { sh_c[warp] = c0; sh_u[warp] = ub; sh_r[warp] = rowmax; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double c = 0.0, u = 0.0;
        float r = 0.f;
        for (int w = 0; w < nwarps; ++w) { c += sh_c[w]; u += sh_u[w]; r = fmaxf(r, sh_r[w]); }
        info[0] = c;
        info[1] = c + u;
        scale[0] = r;
    }
}

// ------------------------------------------------------------------ exhaustive walk
// One Gray-code step: T is a template parameter, so the flipped bit, its new sign and the loop bounds are
// compile-time constants and D / J4 stay in registers.  kTrack = false keeps only the running maximum (one FMNMX
// on the ALU pipe): the search pass; kTrack = true also remembers where it was: the resolve pass.
template <int T, bool kTrack>
__device__ __forceinline__ void gray_step(float& E, float (&D)[kLow], const float (&J4)[kLow][kLow], float& tb,
                                          int& tbest) {
    constexpr int b = ctz_const(T);
    constexpr int gray = T ^ (T >> 1);
    constexpr bool minus = (gray >> b) & 1;                   // new value of spin b is -1
    E = minus ? E - D[b] : E + D[b];
#pragma unroll
    for (int c2 = 0; c2 < b; ++c2) D[c2] = minus ? D[c2] - J4[c2][b] : D[c2] + J4[c2][b];
    if (kTrack) {
        if (E > tb) { tb = E; tbest = T; }
    } else {
        tb = fmaxf(tb, E);
    }
}
template <int LO, int HI, bool kTrack>
__device__ __forceinline__ void gray_walk(float& E, float (&D)[kLow], const float (&J4)[kLow][kLow], float& tb,
                                          int& tbest) {
    if constexpr (HI - LO == 1) {
        gray_step<LO, kTrack>(E, D, J4, tb, tbest);
    } else {
        constexpr int MID = (LO + HI) / 2;
        gray_walk<LO, MID, kTrack>(E, D, J4, tb, tbest);
        gray_walk<MID, HI, kTrack>(E, D, J4, tb, tbest);
    }
}

// Spin layout: low bit b <-> spin b (b < kLow), high bit k <-> spin kLow + k, spin np-1 pinned to +1.
struct LowTables {
    float J4[kLow][kLow];      // 4 J among the low spins
    float d_static[kLow];      // static part of the doubled local field of bit b at the moment it flips
    float e_low;               // sum_{b<c} J_bc with every low spin +1
};

__device__ __forceinline__ void low_tables(const float* Js, int np, LowTables& t) {
#pragma unroll
    for (int a = 0; a < kLow; ++a)
#pragma unroll
        for (int b = 0; b < kLow; ++b) t.J4[a][b] = 4.f * Js[a * np + b];
    // when bit b flips, all lower bits are +1 except bit b-1 (Gray code); all higher low bits start at +1
    t.e_low = 0.f;
#pragma unroll
    for (int b = 0; b < kLow; ++b) {
        float lo = 0.f, hi = 0.f;
#pragma unroll
        for (int c = 0; c < kLow; ++c) {
            if (c < b) lo += (c == b - 1) ? -Js[b * np + c] : Js[b * np + c];
            if (c > b) hi += Js[b * np + c];
        }
        t.d_static[b] = 2.f * (lo + hi);
        t.e_low += hi;
    }
}

// energy among the high spins (+ the pinned one) and their field on every low spin, from scratch
__device__ __forceinline__ void high_setup(const float* Js, int np, int high_bits, unsigned long long hp, float& c,
                                           float (&g)[kLow]) {
    c = 0.f;
#pragma unroll
    for (int b = 0; b < kLow; ++b) g[b] = Js[b * np + (np - 1)];
    for (int k = 0; k < high_bits; ++k) {
        const float sk = ((hp >> k) & 1ull) ? -1.f : 1.f;
        const float* row = Js + (kLow + k) * np;
        float acc = row[np - 1];                              // pinned spin
        for (int m = k + 1; m < high_bits; ++m) acc += ((hp >> m) & 1ull) ? -row[kLow + m] : row[kLow + m];
        c = fmaf(sk, acc, c);
#pragma unroll
        for (int b = 0; b < kLow; ++b) g[b] = fmaf(sk, row[b], g[b]);
    }
}

// the 2^kLow states of the low spins under one high pattern: maximum (and, with kTrack, its Gray index)
template <bool kTrack>
__device__ __forceinline__ float walk_low(const LowTables& t, float c, const float (&g)[kLow], int& tbest) {
    float D[kLow];
    float E = c + t.e_low;
#pragma unroll
    for (int b = 0; b < kLow; ++b) {
        E += g[b];
        D[b] = fmaf(2.f, g[b], t.d_static[b]);
    }
    float tb = E;
    tbest = 0;
    gray_walk<1, kLowStates, kTrack>(E, D, t.J4, tb, tbest);
    return tb;
}

// Search pass.  A thread takes runs of 2^mid_bits consecutive high patterns and walks them in Gray order too: one
// high spin changes between neighbours, so c and g are updated in O(high_bits) instead of rebuilt in O(high_bits^2).
// Per CTA: the best energy (without C0) and the HIGH pattern that reached it.
__global__ void __launch_bounds__(256) flip_exhaustive_kernel(const float* __restrict__ Jg, int np, int high_bits,
                                                              int mid_bits, float* __restrict__ cand_e,
                                                              unsigned long long* __restrict__ cand_s) {
    extern __shared__ float Js[];                            // [np][np]
    for (int i = threadIdx.x; i < np * np; i += blockDim.x) Js[i] = Jg[i];
    __syncthreads();
    LowTables t;
    low_tables(Js, np, t);

    float best = -INFINITY;
    unsigned long long best_hp = 0;
    const unsigned long long runs = 1ull << (high_bits - mid_bits);
    const int per_run = 1 << mid_bits;
    for (unsigned long long run = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; run < runs;
         run += (unsigned long long)gridDim.x * blockDim.x) {
        unsigned long long hp = run << mid_bits;
        float c, g[kLow];
        high_setup(Js, np, high_bits, hp, c, g);
#pragma unroll 1
        for (int j = 0; j < per_run; ++j) {
            if (j > 0) {
                const int k = __ffs(j) - 1;                    // Gray code: bit k of the pattern flips
                hp ^= 1ull << k;
                const float two_s = ((hp >> k) & 1ull) ? -2.f : 2.f;
                const float* row = Js + (kLow + k) * np;
                float hh = row[np - 1];
                for (int m = 0; m < high_bits; ++m)
                    if (m != k) hh += ((hp >> m) & 1ull) ? -row[kLow + m] : row[kLow + m];
                c = fmaf(two_s, hh, c);
#pragma unroll
                for (int b = 0; b < kLow; ++b) g[b] = fmaf(two_s, row[b], g[b]);
            }
            int tbest;
            const float tb = walk_low<false>(t, c, g, tbest);
            if (tb > best) { best = tb; best_hp = hp; }
        }
    }
    // CTA maximum, ties -> smaller pattern
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oe = __shfl_xor_sync(0xffffffffu, best, o);
        const unsigned long long os = __shfl_xor_sync(0xffffffffu, best_hp, o);
        if (oe > best || (oe == best && os < best_hp)) { best = oe; best_hp = os; }
    }
    __shared__ float sh_e[8];
    __shared__ unsigned long long sh_s[8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sh_e[warp] = best; sh_s[warp] = best_hp; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
            if (sh_e[w] > best || (sh_e[w] == best && sh_s[w] < best_hp)) { best = sh_e[w]; best_hp = sh_s[w]; }
        cand_e[blockIdx.x] = best;
        cand_s[blockIdx.x] = best_hp;
    }
}

// Resolve pass: one thread per CTA candidate walks its pattern again, this time remembering the best low state,
// and replaces the pattern by the full 64-bit state (bit i = x_i).
__global__ void __launch_bounds__(64) flip_resolve_kernel(const float* __restrict__ Jg, int np, int high_bits,
                                                          int n_cand, float* __restrict__ cand_e,
                                                          unsigned long long* __restrict__ cand_s) {
    extern __shared__ float Js[];
    for (int i = threadIdx.x; i < np * np; i += blockDim.x) Js[i] = Jg[i];
    __syncthreads();
    const int cnd = blockIdx.x * blockDim.x + threadIdx.x;
    if (cnd >= n_cand) return;
    LowTables t;
    low_tables(Js, np, t);
    const unsigned long long hp = cand_s[cnd];
    float c, g[kLow];
    high_setup(Js, np, high_bits, hp, c, g);
    int tbest;
    const float tb = walk_low<true>(t, c, g, tbest);
    cand_e[cnd] = tb;
    cand_s[cnd] = (hp << kLow) | (unsigned long long)(tbest ^ (tbest >> 1));
}

// ------------------------------------------------------------------ spanning-forest orientation
// Breadth-first over the non-zero couplings: s_v = s_u * sign(J_uv) for the edge that discovers v.  On an
// unfrustrated graph (a consistently orientable surface) this IS the optimum; it seeds annealing replica 0.
// One CTA; queue order is fixed (discoveries of a pop are appended in thread/iteration order through a ballot
// scan), so the result is deterministic.  n pops x one row scan each.
__global__ void __launch_bounds__(1024) flip_tree_kernel(const float* __restrict__ J, int n, int np,
                                                         signed char* __restrict__ s_out) {
    extern __shared__ unsigned char tree_smem[];
    int* queue = reinterpret_cast<int*>(tree_smem);                       // [np]
    signed char* s = reinterpret_cast<signed char*>(queue + np);          // [np], 0 = not reached yet
    __shared__ int warp_cnt[32];
    __shared__ int q_tail;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < np; i += blockDim.x) s[i] = (i >= n) ? 1 : 0;
    if (threadIdx.x == 0) q_tail = 0;
    __syncthreads();
    int head = 0, next_root = 0;
    while (true) {
        __syncthreads();
        int tail = q_tail;
        if (head == tail) {                                               // new component: lowest unreached vertex
            while (next_root < n && s[next_root] != 0) ++next_root;
            if (next_root >= n) break;
            __syncthreads();
            if (threadIdx.x == 0) { queue[tail] = next_root; s[next_root] = 1; q_tail = tail + 1; }
            __syncthreads();
            tail = tail + 1;
        }
        const int u = queue[head++];
        const signed char su = s[u];
        const float* row = J + (int64_t)u * np;
        for (int base = 0; base < n; base += blockDim.x) {
            const int v = base + threadIdx.x;
            const float w = (v < n) ? row[v] : 0.f;
            const bool found = (v < n) && (w != 0.f) && (s[v] == 0);
            const uint32_t m = __ballot_sync(0xffffffffu, found);
            if (lane == 0) warp_cnt[warp] = __popc(m);
            __syncthreads();
            int before = 0, total = 0;
            for (int k = 0; k < 32; ++k) {
                const int c = warp_cnt[k];
                if (k < warp) before += c;
                total += c;
            }
            if (found) {
                queue[tail + before + __popc(m & ((1u << lane) - 1u))] = v;
                s[v] = (w > 0.f) ? su : (signed char)-su;
            }
            tail += total;
            __syncthreads();
        }
        if (threadIdx.x == 0) q_tail = tail;
    }
    for (int i = threadIdx.x; i < np; i += blockDim.x) s_out[i] = s[i];
}

// ------------------------------------------------------------------ annealing
__device__ __forceinline__ uint32_t mix32(uint32_t a, uint32_t b) {
    uint32_t h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u + (a << 6) + (a >> 2));
    h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
    return h;
}

// one warp = one replica.  Shared memory per warp: h[np] float, s[np] int8.
__global__ void __launch_bounds__(kAnnealWarps * 32) flip_anneal_kernel(const float* __restrict__ J, int n, int np,
                                                                        int sweeps, uint32_t seed_lo, uint32_t seed_hi,
                                                                        const float* __restrict__ scale,
                                                                        const signed char* __restrict__ s_tree,
                                                                        float* __restrict__ cand_e,
                                                                        uint32_t* __restrict__ cand_s, int words) {
    extern __shared__ unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* h = reinterpret_cast<float*>(smem_raw) + (size_t)warp * np;
    signed char* s = reinterpret_cast<signed char*>(smem_raw + (size_t)kAnnealWarps * np * sizeof(float)) + (size_t)warp * np;
    const int replica = blockIdx.x * kAnnealWarps + warp;
    const uint32_t key = mix32(seed_lo ^ (uint32_t)replica * 0x632BE5ABu, seed_hi + (uint32_t)replica);

    // replica 0 starts from the spanning-forest orientation, replica 1 from "no flip", the others from random states
    for (int i = lane; i < np; i += 32)
        s[i] = (replica == 0) ? s_tree[i] : ((replica == 1 || i >= n || !(mix32(key, 0x51ED27u + i) & 1u)) ? 1 : -1);
    __syncwarp();
    auto rebuild = [&]() {
        for (int i = lane; i < n; i += 32) {
            const float* row = J + (int64_t)i * np;
            float acc = 0.f;
            for (int j = 0; j < n; ++j) acc = fmaf(row[j], (float)s[j], acc);
            h[i] = acc;
        }
        __syncwarp();
    };
    rebuild();

    const float jscale = fmaxf(scale[0], 1e-30f);
    // geometric schedule; replicas differ in their starting temperature
    const float t_hot = jscale * exp2f(-(float)(replica & 3));
    const float t_cold = jscale * 1e-3f;
    const float ratio = (sweeps > 1) ? powf(t_cold / t_hot, 1.f / (float)(sweeps - 1)) : 1.f;
    float temp = t_hot;
    const int anneal_sweeps = (replica < 2) ? 0 : sweeps;    // replicas 0 and 1: pure descent from their start

    auto sweep = [&](uint32_t sweep_id, float inv_t, bool greedy, float tol) -> int {
        int flips = 0;
        for (int c = 0; c < n; c += 32) {
            const int i = c + lane;
            float u = 0.f;
            if (!greedy && i < n) u = (float)(mix32(key ^ sweep_id * 0x2545F491u, (uint32_t)i) >> 8) * (1.f / 16777216.f);
            int lo = 0;
            while (true) {
                bool acc = false;
                if (i < n && lane >= lo) {
                    const float gain = -2.f * (float)s[i] * h[i];
                    acc = greedy ? (gain > tol) : (gain >= 0.f || u < __expf(gain * inv_t));
                }
                const uint32_t mask = __ballot_sync(0xffffffffu, acc);
                if (!mask) break;
                const int b = __ffs(mask) - 1;
                const int f = c + b;
                const float two_s = -2.f * (float)s[f];       // 2 x new value
                __syncwarp();
                if (lane == b) s[f] = (signed char)-s[f];
                const float* row = J + (int64_t)f * np;
                for (int j = lane; j < n; j += 32) h[j] = fmaf(two_s, row[j], h[j]);
                __syncwarp();
                ++flips;
                lo = b + 1;
            }
        }
        return flips;
    };

    for (int sw = 0; sw < anneal_sweeps; ++sw) {
        sweep((uint32_t)sw, 1.f / temp, false, 0.f);
        temp *= ratio;
    }
    rebuild();                                               // drop the accumulated rounding of h
    const float tol = jscale * 1e-6f;
    for (int sw = 0; sw < 4 * n + 64; ++sw)
        if (sweep(0u, 0.f, true, tol) == 0) break;

    float e = 0.f;
    for (int i = lane; i < n; i += 32) e = fmaf((float)s[i], h[i], e);
    e = 0.5f * warp_sum(e);
    for (int w = lane; w < words; w += 32) {
        uint32_t bits = 0;
        for (int k = 0; k < 32; ++k) {
            const int i = w * 32 + k;
            if (i < n && s[i] < 0) bits |= 1u << k;
        }
        cand_s[(int64_t)replica * words + w] = bits;
    }
    if (lane == 0) cand_e[replica] = e;
}

// ------------------------------------------------------------------ selection (one CTA of kTopK warps)
// status: 1 = exhaustive (optimal), 2 = reaches the bound C0 + sum|J| (optimal), 0 = best found.
__global__ void __launch_bounds__(kTopK * 32) flip_select_kernel(const double* __restrict__ A, const double* __restrict__ B,
                                                                 int n, int64_t ld, const float* __restrict__ cand_e,
                                                                 const uint32_t* __restrict__ cand_s, int words,
                                                                 int n_cand, const double* __restrict__ info,
                                                                 int exhaustive, uint8_t* __restrict__ x_out,
                                                                 double* __restrict__ obj_out,
                                                                 int32_t* __restrict__ status_out) {
    __shared__ int top[kTopK];
    __shared__ double val[kTopK];
    __shared__ int winner, invert;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // warp 0: the kTopK largest float energies, ties -> lower candidate index (serial selection, tiny)
    if (warp == 0) {
        float prev_e = INFINITY;
        int prev_i = -1;
        for (int k = 0; k < kTopK; ++k) {
            float be = -INFINITY;
            int bi = 0x7fffffff;
            for (int c = lane; c < n_cand; c += 32) {
                const float e = cand_e[c];
                const bool after = (e < prev_e) || (e == prev_e && c > prev_i);   // not selected yet
                if (after && (e > be || (e == be && c < bi))) { be = e; bi = c; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float oe = __shfl_xor_sync(0xffffffffu, be, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (oe > be || (oe == be && oi < bi)) { be = oe; bi = oi; }
            }
            if (lane == 0) top[k] = (bi == 0x7fffffff) ? -1 : bi;
            prev_e = be;
            prev_i = bi;
        }
    }
    __syncthreads();
    // every warp: objective of its candidate in double, fixed order
    {
        const int c = top[warp];
        double acc = 0.0;
        if (c >= 0) {
            const uint32_t* st = cand_s + (int64_t)c * words;
            for (int i = 0; i < n; ++i) {
                const int si = (st[i >> 5] >> (i & 31)) & 1u;
                double row = 0.0;
                for (int j = i + 1 + lane; j < n; j += 32) {
                    const int sj = (st[j >> 5] >> (j & 31)) & 1u;
                    const double d = jd(A, B, ld, i, j);
                    row += (si == sj) ? d : -d;
                }
                acc += row;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        }
        if (lane == 0) val[warp] = (c >= 0) ? acc : -INFINITY;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int w = 0;
        for (int k = 1; k < kTopK; ++k)
            if (val[k] > val[w]) w = k;
        winner = top[w];
        const uint32_t* st = cand_s + (int64_t)winner * words;
        int ones = 0;
        for (int i = 0; i < n; ++i) ones += (st[i >> 5] >> (i & 31)) & 1u;
        invert = (2 * ones > n) || (2 * ones == n && (st[0] & 1u));
        const double obj = info[0] + val[w];
        obj_out[0] = obj;
        obj_out[1] = info[1];
        const double slack = 1e-9 * fmax(fabs(info[1]), 1.0);
        status_out[0] = exhaustive ? 1 : ((obj >= info[1] - slack) ? 2 : 0);
    }
    __syncthreads();
    const uint32_t* st = cand_s + (int64_t)winner * words;
    for (int i = threadIdx.x; i < n; i += blockDim.x) x_out[i] = (uint8_t)((((st[i >> 5] >> (i & 31)) & 1u) ^ (uint32_t)invert));
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

struct FlipLayout {
    int np, words, n_cand, blocks, mid_bits;
    bool exhaustive;
    size_t J, info, scale, tree, cand_e, cand_s, total;
};

FlipLayout flip_layout(int64_t n, int exact_max_n, int replicas) {
    FlipLayout L{};
    if (exact_max_n > kExactMax) exact_max_n = kExactMax;
    L.exhaustive = n <= exact_max_n;
    if (L.exhaustive) {
        L.np = (int)(n < kExactMin ? kExactMin : n);
        const int high_bits = L.np - 1 - kLow;
        // runs of 2^mid_bits patterns per thread visit, sized so that the grid still gets >= 2^22 runs to balance
        L.mid_bits = high_bits - 22 < 0 ? 0 : (high_bits - 22 > 6 ? 6 : high_bits - 22);
        const unsigned long long patterns = 1ull << (high_bits - L.mid_bits);
        unsigned long long blocks = (patterns + 255) / 256;
        if (blocks > 148ull * 8) blocks = 148ull * 8;
        L.blocks = (int)blocks;
        L.n_cand = L.blocks;
        L.words = 2;
    } else {
        L.np = (int)((n + 31) & ~31ll);
        if (replicas < kAnnealWarps) replicas = kAnnealWarps;
        L.blocks = (replicas + kAnnealWarps - 1) / kAnnealWarps;
        L.n_cand = L.blocks * kAnnealWarps;
        L.words = L.np / 32;
    }
    size_t off = 0;
    L.J = off; off += up256((size_t)L.np * L.np * sizeof(float));
    L.info = off; off += 256;
    L.scale = off; off += 256;
    L.tree = off; off += up256((size_t)L.np);
    L.cand_e = off; off += up256((size_t)L.n_cand * sizeof(float));
    L.cand_s = off; off += up256((size_t)L.n_cand * L.words * sizeof(uint32_t));
    L.total = off;
    return L;
}

}  // namespace

size_t flip_solve_workspace_bytes(int64_t n, int exact_max_n, int replicas) {
    if (n < 1 || n > kMaxN) return 0;
    return flip_layout(n, exact_max_n, replicas).total;
}

int run_flip_solve(const double* A, const double* B, int64_t n, int64_t ld, int exact_max_n, int replicas, int sweeps,
                   uint64_t seed, uint8_t* x_out, double* obj_out, int32_t* status_out, void* workspace,
                   size_t workspace_bytes, cudaStream_t stream) {
    if (n > kMaxN) {
        set_error("flip solve: at most %d patches (got %lld)", kMaxN, (long long)n);
        return DNP_ERR_INVALID;
    }
    if (sweeps < 1) sweeps = 1;
    const FlipLayout L = flip_layout(n, exact_max_n, replicas);
    if (!workspace || workspace_bytes < L.total) {
        set_error("workspace too small: need %zu bytes (dnp_flip_solve_workspace_bytes), got %zu", L.total,
                  workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    float* J = reinterpret_cast<float*>(ws + L.J);
    double* info = reinterpret_cast<double*>(ws + L.info);
    float* scale = reinterpret_cast<float*>(ws + L.scale);
    float* cand_e = reinterpret_cast<float*>(ws + L.cand_e);
    uint32_t* cand_s = reinterpret_cast<uint32_t*>(ws + L.cand_s);

    flip_prepare_kernel<<<1, 1024, 0, stream>>>(A, B, (int)n, ld, L.np, J, info, scale);
    if (L.exhaustive) {
        const size_t smem = (size_t)L.np * L.np * sizeof(float);
        unsigned long long* cand_s64 = reinterpret_cast<unsigned long long*>(cand_s);
        flip_exhaustive_kernel<<<L.blocks, 256, smem, stream>>>(J, L.np, L.np - 1 - kLow, L.mid_bits, cand_e, cand_s64);
        flip_resolve_kernel<<<(L.n_cand + 63) / 64, 64, smem, stream>>>(J, L.np, L.np - 1 - kLow, L.n_cand, cand_e,
                                                                        cand_s64);
        count_launch(1);
    } else {
        signed char* s_tree = reinterpret_cast<signed char*>(ws + L.tree);
        flip_tree_kernel<<<1, 1024, (size_t)L.np * 5, stream>>>(J, (int)n, L.np, s_tree);
        count_launch(1);
        const size_t smem = (size_t)kAnnealWarps * L.np * (sizeof(float) + 1);
        if (smem > 48 * 1024)
            DNP_CHECK_CUDA(cudaFuncSetAttribute(flip_anneal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        flip_anneal_kernel<<<L.blocks, kAnnealWarps * 32, smem, stream>>>(J, (int)n, L.np, sweeps, (uint32_t)seed,
                                                                          (uint32_t)(seed >> 32), scale, s_tree, cand_e, cand_s,
                                                                          L.words);
    }
    flip_select_kernel<<<1, kTopK * 32, 0, stream>>>(A, B, (int)n, ld, cand_e, cand_s, L.words, L.n_cand, info,
                                                     L.exhaustive ? 1 : 0, x_out, obj_out, status_out);
    count_launch(3);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
