// dnp_pairs.cu -- all-pairs dipole potential / field kernels (K1, K2, K6, K7).
//
// Replaces the leaf bodies of field_utils.potential (field_utils.py:46-55) and
// field_utils.field_grad (field_utils.py:96-116) and the recursive chunking
// around them (:23-44, :73-94): memory is O(S + M), nothing is materialised per
// pair.
//
// Design (B200, sm_100a):
//   * sources are re-packed once per call into 48-byte rows [x y z nx | ny nz mx my | mz 0 0 0] with
//     m = -3 n; the TMA unit streams 12 KB tiles of them into a 3-stage shared-memory ring
//     (cp.async.bulk + mbarrier, SASS UBLKCP) and every warp reads a source with broadcast LDS.128.
//   * every thread keeps 2*TP targets in registers.  Two targets share one 64-bit register pair and the
//     whole per-pair arithmetic runs on packed FFMA2/FMUL2/FADD2 (two FP32 ops per lane per instruction)
//     whose source-side operand is a 32-bit register broadcast to both halves (`R.F32`): 19 packed
//     instructions + 2 MUFU.RSQ + 2 MUFU.RCP + 4 FMNMX per (2 targets x 1 source),
//         E += w (n + t R),  t = (m.R) u^2,  u = rsqrt(r2),  w = 1 / (r2^2 u + eps).
//     Calls above ~2e8 pairs take the constant-window kernels of dnp_upairs.cu instead, where the source
//     operand is a uniform register and costs no register-file read at all.
//   * no tensor cores: the 1/(|R|^3+eps) weight is per pair, not a contraction.
//   * coincident pairs: rsqrt(0)=inf makes r2^2 u = 0*inf NaN, so the weight w becomes NaN and
//     max(w, 0) returns 0 (FMNMX drops the NaN operand); t uses the clamped rsqrt so it stays finite
//     (and is exactly 0 because m.R = 0).
//   * per-target sums are flushed into a second-level accumulator after every 256-source tile, and
//     (for nsplit > 1) source ranges are reduced by a second kernel in a fixed order: deterministic,
//     no float atomics.
#include <stdlib.h>

#include <vector>

#include "dnp_common.cuh"
#include "dnp_pairmath.cuh"
#include "dnp_epilogue.cuh"
#include "dnp_pairs.h"

namespace dnp {

namespace {

constexpr int kBlock = 128;                      // threads per CTA
constexpr int kTileSrc = 256;                    // sources per shared-memory stage
constexpr int kTileSmall = 64;                   // ... for calls too small to fill the machine with 256-source tiles
constexpr int kStages = 3;                       // TMA ring depth
constexpr int kSrcBytes = 48;                    // packed bytes per source
constexpr size_t smem_bytes(int tile_src) { return size_t(kStages) * tile_src * kSrcBytes + kStages * sizeof(uint64_t); }

// ------------------------------------------------------------------ source packing
__global__ void pack_sources_kernel(const float* __restrict__ src, int64_t S, int64_t ld_s,
                                    const int64_t* __restrict__ idx, int64_t S_pad,
                                    float4* __restrict__ packed) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S_pad) return;
    float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // padding rows: a null dipole
    if (s < S) {
        const float* p = src + (idx ? idx[s] : s) * ld_s;
#pragma unroll
        for (int c = 0; c < 6; ++c) v[c] = __ldg(p + c);
    }
    pack_source_row(v, packed + 3 * s);
}

// the same for a source range chosen on the device: rows idx[o .. o + n) of patch *sel.patch, null dipoles up to S_pad
__global__ void pack_selected_kernel(const float* __restrict__ src, int64_t ld_s, const int64_t* __restrict__ idx,
                                     SourceSelect sel, int64_t S_pad, float4* __restrict__ packed) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S_pad) return;
    const int pj = *sel.patch;
    const int64_t o = pj < 0 ? 0 : sel.offsets[pj], n = pj < 0 ? 0 : sel.offsets[pj + 1] - o;
    float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (s < n) {
        const float* p = src + idx[o + s] * ld_s;
#pragma unroll
        for (int c = 0; c < 6; ++c) v[c] = __ldg(p + c);
    }
    pack_source_row(v, packed + 3 * s);
}

// lattice node i of util.gen_grid (util.py:26-36), same float operations
__device__ __forceinline__ void lattice_node(int64_t i, int n, float& x, float& y, float& z) {
    const int zi = (int)(i % n), xy = (int)(i / n), yi = xy % n, xi = xy / n;
    const float fn = (float)n;
    x = __fmul_rn(__fsub_rn(__fdiv_rn((float)xi, fn), 0.5f), 2.f);
    y = __fmul_rn(__fsub_rn(__fdiv_rn((float)yi, fn), 0.5f), 2.f);
    z = __fmul_rn(__fsub_rn(__fdiv_rn((float)zi, fn), 0.5f), 2.f);
}

// ------------------------------------------------------------------ the all-pairs kernel
// grid.x = target tiles of kBlock*2*TP targets, grid.y = source splits.
template <int TP, bool kPotential, int TS>
__global__ void __launch_bounds__(kBlock, (TP >= 4 ? 3 : (TP == 2 ? 5 : 8)))
pairs_kernel(const float4* __restrict__ packed, int tiles_total, int nsplit,
             const float* __restrict__ tgt, int64_t M, int64_t ld_t, int lattice_n,
             float eps, float* __restrict__ partial, PairEpilogue ep, SourceSelect sel) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    if (sel.patch) {                                 // device-chosen source range: only its tiles are walked
        const int pj = *sel.patch;
        const int64_t n = pj < 0 ? 0 : sel.offsets[pj + 1] - sel.offsets[pj];
        tiles_total = (int)((n + TS - 1) / TS);
    }
    float4* stage_buf = reinterpret_cast<float4*>(smem_raw);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + size_t(kStages) * (uint32_t)(TS * kSrcBytes));

    const int tid = threadIdx.x;
    const int64_t tile_base = (int64_t)blockIdx.x * (kBlock * 2 * TP);
    const int split = blockIdx.y;
    const int t_begin = (int)((int64_t)tiles_total * split / nsplit);
    const int t_end = (int)((int64_t)tiles_total * (split + 1) / nsplit);
    const int n_tiles = t_end - t_begin;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) mbar_init(&bars[s], 1);
        mbar_init_fence();
    }
    __syncthreads();
    if (tid == 0) {
        for (int i = 0; i < kStages && i < n_tiles; ++i) {
            mbar_arrive_expect_tx(&bars[i], (uint32_t)(TS * kSrcBytes));
            tma_bulk_g2s(stage_buf + size_t(i) * TS * 3,
                         packed + size_t(t_begin + i) * TS * 3, (uint32_t)(TS * kSrcBytes), &bars[i]);
        }
    }

    // ---- targets: lane a of pair j is target tile_base + (2j)*kBlock + tid, lane b the next kBlock
    f32x2 qx[TP], qy[TP], qz[TP];
    int64_t row[2 * TP];
#pragma unroll
    for (int j = 0; j < 2 * TP; ++j) {
        const int64_t m = tile_base + (int64_t)j * kBlock + tid;
        row[j] = -1;
        if (m < M) {
            const int64_t r = ep.tgt_idx ? ep.tgt_idx[m] : m;
            if (!(ep.tgt_skip && ep.tgt_skip[r])) row[j] = r;
        }
    }
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float x[2] = {0.f, 0.f}, y[2] = {0.f, 0.f}, z[2] = {0.f, 0.f};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t r = row[2 * j + h];
            if (r >= 0) {
                if (lattice_n > 0) lattice_node(r, lattice_n, x[h], y[h], z[h]);
                else { const float* q = tgt + r * ld_t; x[h] = __ldg(q); y[h] = __ldg(q + 1); z[h] = __ldg(q + 2); }
            }
        }
        qx[j] = pack2(x[0], x[1]); qy[j] = pack2(y[0], y[1]); qz[j] = pack2(z[0], z[1]);
    }

    const f32x2 zero2 = pack2(0.f, 0.f);
    const f32x2 eps2 = pack2(eps, eps);
    // second-level totals (one per component per target pair)
    f32x2 tx[TP], ty[TP], tz[TP];
#pragma unroll
    for (int j = 0; j < TP; ++j) { tx[j] = zero2; ty[j] = zero2; tz[j] = zero2; }

    for (int it = 0; it < n_tiles; ++it) {
        const int st = it % kStages;
        mbar_wait(&bars[st], (uint32_t)((it / kStages) & 1));
        const float4* sp = stage_buf + size_t(st) * TS * 3;

        if (kPotential) {
            f32x2 phi[TP];
#pragma unroll
            for (int j = 0; j < TP; ++j) phi[j] = zero2;
#pragma unroll 2
            for (int s = 0; s < TS; ++s) {
                const SrcX2 src(sp[3 * s], sp[3 * s + 1], sp[3 * s + 2]);
#pragma unroll
                for (int j = 0; j < TP; ++j) pair_potential_x2(src, qx[j], qy[j], qz[j], phi[j]);
            }
#pragma unroll
            for (int j = 0; j < TP; ++j) tx[j] = add2(tx[j], phi[j]);
        } else {
            f32x2 ax[TP], ay[TP], az[TP];
#pragma unroll
            for (int j = 0; j < TP; ++j) { ax[j] = ay[j] = az[j] = zero2; }
#pragma unroll 2
            for (int s = 0; s < TS; ++s) {
                const SrcX2 src(sp[3 * s], sp[3 * s + 1], sp[3 * s + 2]);
#pragma unroll
                for (int j = 0; j < TP; ++j)
                    pair_field_x2<false>(src, qx[j], qy[j], qz[j], eps2, ax[j], ay[j], az[j]);
            }
            // clamp-free fast pass; a duplicate of a target in this tile shows up as NaN and the tile is redone
            bool bad = false;
#pragma unroll
            for (int j = 0; j < TP; ++j) bad |= sums_nonfinite(ax[j], ay[j], az[j]);
            if (__any_sync(0xffffffffu, bad)) {
#pragma unroll
                for (int j = 0; j < TP; ++j) { ax[j] = ay[j] = az[j] = zero2; }
#pragma unroll 2
                for (int s = 0; s < TS; ++s) {
                    const SrcX2 src(sp[3 * s], sp[3 * s + 1], sp[3 * s + 2]);
#pragma unroll
                    for (int j = 0; j < TP; ++j)
                        pair_field_x2<true>(src, qx[j], qy[j], qz[j], eps2, ax[j], ay[j], az[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < TP; ++j) {               // tile sums folded into the totals
                tx[j] = add2(tx[j], ax[j]);
                ty[j] = add2(ty[j], ay[j]);
                tz[j] = add2(tz[j], az[j]);
            }
        }

        __syncthreads();                                  // every warp is done with stage `st`
        if (tid == 0 && it + kStages < n_tiles) {
            mbar_arrive_expect_tx(&bars[st], (uint32_t)(TS * kSrcBytes));
            tma_bulk_g2s(stage_buf + size_t(st) * TS * 3,
                         packed + size_t(t_begin + it + kStages) * TS * 3, (uint32_t)(TS * kSrcBytes), &bars[st]);
        }
    }

    // ---- write-out: direct epilogue, or one partial row per source split
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float ex[2], ey[2], ez[2];
        unpack2(tx[j], ex[0], ex[1]);
        unpack2(ty[j], ey[0], ey[1]);
        unpack2(tz[j], ez[0], ez[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t r = row[2 * j + h];
            if (r < 0) continue;
            const int64_t m = tile_base + (int64_t)(2 * j + h) * kBlock + tid;
            if (kPotential) {
                if (partial) partial[(int64_t)split * M + m] = ex[h];
                else apply_potential_epilogue(ep, r, ex[h]);
            } else {
                if (partial) {
                    float* o = partial + ((int64_t)split * M + m) * 3;
                    o[0] = ex[h]; o[1] = ey[h]; o[2] = ez[h];
                } else {
                    apply_field_epilogue(ep, r, ex[h], ey[h], ez[h]);
                }
            }
        }
    }
}

// ------------------------------------------------------------------ fixed-order reduction over source splits
template <bool kPotential>
__global__ void split_reduce_kernel(const float* __restrict__ partial, int nsplit, int64_t M, int64_t set_stride,
                                    PairEpilogue ep) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const int64_t r = ep.tgt_idx ? ep.tgt_idx[m] : m;
    if (ep.tgt_skip && ep.tgt_skip[r]) return;
    if (kPotential) {
        float acc = 0.f;
        for (int k = 0; k < nsplit; ++k) acc += partial[(int64_t)k * set_stride + m];
        apply_potential_epilogue(ep, r, acc);
    } else {
        float ex = 0.f, ey = 0.f, ez = 0.f;
        for (int k = 0; k < nsplit; ++k) {
            const float* p = partial + (int64_t)k * set_stride + m * 3;
            ex += p[0]; ey += p[1]; ez += p[2];
        }
        apply_field_epilogue(ep, r, ex, ey, ez);
    }
}

// single-CTA deterministic sum (fixed strided order + fixed tree); scale applied last
__global__ void __launch_bounds__(1024) sum_kernel(const float* __restrict__ in, int64_t n, int64_t stride,
                                                   float scale, float* __restrict__ out) {
    __shared__ float warp_part[32];
    float acc = 0.f;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) acc += in[i * stride];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < (blockDim.x >> 5) ? warp_part[threadIdx.x] : 0.f;
        v = warp_sum(v);
        if (threadIdx.x == 0) *out = v * scale;
    }
}

std::atomic<int64_t> g_launches{0};

// per-thread kernel timer (dnp_profile_begin / dnp_profile_end): event pairs recorded on the launching
// stream around every all-pairs kernel launched by THIS host thread while it is armed.
struct KernelTimer {
    bool armed = false;
    std::vector<cudaEvent_t> events;   // start0, stop0, start1, stop1, ...
};
thread_local KernelTimer t_timer;

void timer_mark_impl(cudaStream_t stream) {
    if (!t_timer.armed) return;
    cudaEvent_t ev;
    if (cudaEventCreate(&ev) != cudaSuccess) return;
    cudaEventRecord(ev, stream);
    t_timer.events.push_back(ev);
}

// tuning overrides, read once (immutable afterwards, so the library stays re-entrant)
int env_int(const char* name) {
    const char* v = getenv(name);
    return v ? atoi(v) : 0;
}

inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

}  // namespace

int64_t launch_count() { return g_launches.load(); }

void timer_mark(cudaStream_t stream) { timer_mark_impl(stream); }

int run_split_reduce(bool potential, const float* partial, int nsplit, int64_t M, const PairEpilogue& ep,
                     cudaStream_t stream, int64_t set_stride) {
    const int threads = 256;
    const unsigned blocks = (unsigned)((M + threads - 1) / threads);
    if (set_stride <= 0) set_stride = M * (potential ? 1 : 3);
    if (potential) split_reduce_kernel<true><<<blocks, threads, 0, stream>>>(partial, nsplit, M, set_stride, ep);
    else split_reduce_kernel<false><<<blocks, threads, 0, stream>>>(partial, nsplit, M, set_stride, ep);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

void profile_begin() {
    for (cudaEvent_t ev : t_timer.events) cudaEventDestroy(ev);
    t_timer.events.clear();
    t_timer.armed = true;
}

int profile_end(float* ms_out, int capacity, int* n_out) {
    t_timer.armed = false;
    const int n = (int)(t_timer.events.size() / 2);
    int rc = DNP_OK;
    for (int i = 0; i < n; ++i) {
        float ms = 0.f;
        cudaError_t e = cudaEventSynchronize(t_timer.events[2 * i + 1]);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, t_timer.events[2 * i], t_timer.events[2 * i + 1]);
        if (e != cudaSuccess && rc == DNP_OK) rc = cuda_fail(e, "kernel timer");
        if (ms_out && i < capacity) ms_out[i] = ms;
    }
    for (cudaEvent_t ev : t_timer.events) cudaEventDestroy(ev);
    t_timer.events.clear();
    if (n_out) *n_out = n;
    return rc;
}
void count_launch(int n) { g_launches.fetch_add(n); }

// ------------------------------------------------------------------ planning
int sm_count_cached() {
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int v = cached[dev].load();
    if (v == 0) {
        cudaDeviceProp prop;
        v = (cudaGetDeviceProperties(&prop, dev) == cudaSuccess) ? prop.multiProcessorCount : 148;
        cached[dev].store(v);
    }
    return v;
}

PairPlan make_pair_plan(int64_t S, int64_t M, int sm_count, int force_tp, int force_nsplit) {
    PairPlan p;
    // (targets per thread, source tile, source splits) by a small cost model of the grid, in units of one full
    // wave of the candidate (resident CTAs on every SM): full waves cost 1 each; the last, partial wave costs the
    // deepest SM's share of it, but never less than what ONE CTA needs when it has an SM to itself (few warps
    // cannot hide the MUFU / shared-memory latency); a split adds the reduction pass over its partial sums.
    // Calls too small to fill the machine with 256-source tiles (a 150-row patch against 11 000 points is one
    // tile x 44 CTAs) take 64-source tiles, so that the sources can be split over more CTAs.  The constants are
    // from tools/plan_probe.py on B200 (profiles/README.md): before this the S = 500, M = 524288 step of the
    // representatives loop ran 512 TP4 CTAs on 444 slots, 1.15 waves at 1.13e12 pairs/s, where TP2 reaches 1.33e12.
    if (force_tp <= 0) force_tp = env_int("DNP_FORCE_TP");
    if (force_nsplit <= 0) force_nsplit = env_int("DNP_FORCE_NSPLIT");
    if (force_tp != 1 && force_tp != 2 && force_tp != 4) force_tp = 0;
    const int force_tile = env_int("DNP_FORCE_TILE");
    double best_cost = 0.0;
    p.tp = 0;
    for (int cand = 0; cand < 4; ++cand) {
        const int tp = cand == 0 ? 4 : (cand == 1 ? 2 : 1);
        const int tile = cand == 3 ? kTileSmall : kTileSrc;               // small tiles: TP1 only
        if (force_tp && tp != force_tp) continue;
        if (force_tile > 0 ? tile != force_tile : (force_tp && tile != kTileSrc)) continue;
        int tiles = (int)((S + tile - 1) / tile);
        if (tiles < 1) tiles = 1;
        const int max_split = tiles < 64 ? tiles : 64;
        const int want_split = force_nsplit > max_split ? max_split : force_nsplit;
        const int resident = tp >= 4 ? 3 : (tp == 2 ? 5 : 8);
        const double alone = tp >= 4 ? 0.55 : (tp == 2 ? 0.40 : 0.30);    // one CTA alone on an SM / a full wave
        const double rate = tp == 2 ? 1.0 : 0.97;                          // measured steady-state rate vs TP2
        const int64_t tile_tgts = (int64_t)kBlock * 2 * tp;
        int64_t m_tiles = (M + tile_tgts - 1) / tile_tgts;
        if (m_tiles < 1) m_tiles = 1;
        const int64_t wave = (int64_t)sm_count * resident;
        for (int k = 1; k <= max_split; ++k) {
            if (want_split > 0 && k != want_split) continue;
            const int64_t ctas = m_tiles * k;
            const int64_t full = ctas / wave, tail = ctas - full * wave;
            double waves = (double)full;
            if (tail > 0) {
                const double deepest = (double)((tail + sm_count - 1) / sm_count) / resident;
                waves += deepest > alone ? deepest : alone;
            }
            // sources one CTA walks (the longest split; 8 iterations' worth of barrier + fold per tile) + its
            // prologue / epilogue, in source iterations
            const double cta_src = (double)((tiles + k - 1) / k) * (tile + 8.0) + 32.0;
            double cost = waves * (double)wave * (double)tile_tgts * cta_src / rate;
            if (k > 1) cost += 3.5 * (double)(k + 1) * (double)M;          // partial sums written, read, reduced
            if (k > 1) cost += 3.0e6;                                       // ... by one more launch (~2 us of pairs)
            if (p.tp == 0 || cost < best_cost * 0.99) {                     // ties: fewer splits, more targets per thread
                best_cost = cost;
                p.tp = tp;
                p.tile_src = tile;
                p.tiles_total = tiles;
                p.nsplit = k;
                p.m_tiles = (int)m_tiles;
            }
            if (ctas >= 8 * wave) break;                                    // deep enough: stop splitting
        }
    }
    p.S_pad = (int64_t)p.tiles_total * p.tile_src;
    p.packed_bytes = (size_t)round_up(p.S_pad * kSrcBytes, 256);
    p.partial_bytes = p.nsplit > 1 ? (size_t)round_up((int64_t)p.nsplit * M * 3 * sizeof(float), 256) : 0;
    return p;
}

size_t pairs_workspace_bytes(int64_t S, int64_t M) {
    // exactly what run_pairs will carve (same plan), plus the per-target scratch
    // row of the DOT / lattice paths and one scalar slot
    if (S < 1) S = 1;
    if (M < 1) M = 1;
    const PairPlan p = make_pair_plan(S, M, sm_count_cached(), 0, 0);
    const size_t packed = (size_t)round_up(round_up(S, kTileSrc) * kSrcBytes, 256);   // either tile's padding
    size_t core = packed + p.partial_bytes;
    if (use_ring_path(S, M, 0)) {
        const size_t c = upairs_workspace_bytes(S, M);
        if (c > core) core = c;
    }
    return core + (size_t)round_up(M * 4, 256) + 1024;
}

// For callers that size ONE workspace for their largest source set and reuse it for smaller ones (the greedy
// drivers): a smaller set may want more source splits, so bound the split count of ANY plan with <= S sources
// instead of planning every size (the plan stops splitting at 8 waves of CTAs; run_pairs clamps in any case).
size_t pairs_workspace_bytes_upper(int64_t S, int64_t M) {
    if (S < 1) S = 1;
    if (M < 1) M = 1;
    const int sm = sm_count_cached();
    int64_t kmax = 1;
    for (int tp = 4; tp >= 1; tp >>= 1) {
        const int resident = tp >= 4 ? 3 : (tp == 2 ? 5 : 8);
        const int64_t tile_tgts = (int64_t)kBlock * 2 * tp;
        const int64_t m_tiles = (M + tile_tgts - 1) / tile_tgts;
        const int64_t k = (8 * (int64_t)sm * resident + m_tiles - 1) / m_tiles;
        if (k > kmax) kmax = k;
    }
    const int64_t tiles_small = (S + kTileSmall - 1) / kTileSmall;
    if (kmax > tiles_small) kmax = tiles_small;
    if (kmax > 64) kmax = 64;
    const size_t partial = kmax > 1 ? (size_t)round_up(kmax * M * 3 * (int64_t)sizeof(float), 256) : 0;
    const size_t exact = pairs_workspace_bytes(S, M);
    const size_t upper = (size_t)round_up(round_up(S, kTileSrc) * kSrcBytes, 256) + partial +
                         (size_t)round_up(M * 4, 256) + 1024;
    return upper > exact ? upper : exact;
}

// ------------------------------------------------------------------ launchers
template <int TP, bool kPot, int TS = kTileSrc>
static int launch_pairs_tp(const PairPlan& plan, const float4* packed, const float* tgt, int64_t M,
                           int64_t ld_t, int lattice_n, float eps, float* partial, const PairEpilogue& ep,
                           const SourceSelect& sel, cudaStream_t stream) {
    dim3 grid((unsigned)plan.m_tiles, (unsigned)plan.nsplit);
    timer_mark(stream);
    pairs_kernel<TP, kPot, TS><<<grid, kBlock, smem_bytes(TS), stream>>>(packed, plan.tiles_total, plan.nsplit, tgt,
                                                                         M, ld_t, lattice_n, eps, partial, ep, sel);
    timer_mark(stream);
    count_launch(1);
    return DNP_OK;
}

int run_pairs(bool potential, const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx,
              const float* tgt, int64_t M, int64_t ld_t, int lattice_n, float eps,
              const PairEpilogue& ep, void* workspace, size_t workspace_bytes, cudaStream_t stream,
              int force_tp, int force_nsplit, SourceSelect sel) {
    if (M <= 0) return DNP_OK;
    if (sel.patch && (!src_idx || !sel.offsets)) {
        set_error("device-selected sources need src_idx and offsets");
        return DNP_ERR_INVALID;
    }
    if (!sel.patch && force_tp <= 0 && force_nsplit <= 0 && use_ring_path(S, M, lattice_n))
        return run_upairs(potential, src, S, ld_s, src_idx, tgt, M, ld_t, eps, ep, workspace, workspace_bytes, stream);
    PairPlan plan = make_pair_plan(S, M, sm_count_cached(), force_tp, force_nsplit);
    if (workspace && plan.nsplit > 1 && workspace_bytes < plan.packed_bytes + plan.partial_bytes) {
        // a workspace sized for another source count: keep the splits that fit
        const size_t per_split = (size_t)M * 3 * sizeof(float);
        const size_t room = workspace_bytes > plan.packed_bytes ? workspace_bytes - plan.packed_bytes : 0;
        const int fit = (int)(room / per_split < 64 ? room / per_split : 64);
        plan.nsplit = fit >= 2 ? (fit < plan.nsplit ? fit : plan.nsplit) : 1;
        plan.partial_bytes = plan.nsplit > 1 ? (size_t)plan.nsplit * per_split : 0;
    }
    if (!workspace || workspace_bytes < plan.packed_bytes + plan.partial_bytes) {
        set_error("workspace too small: need %zu bytes, got %zu", plan.packed_bytes + plan.partial_bytes,
                  workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    float4* packed = reinterpret_cast<float4*>(workspace);
    float* partial = plan.nsplit > 1
                         ? reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + plan.packed_bytes)
                         : nullptr;
    {
        const int threads = 256;
        const unsigned blocks = (unsigned)((plan.S_pad + threads - 1) / threads);
        if (sel.patch && sel.prepacked) {
            // packed by the caller's own kernel (dnp_drivers.cu)
        } else {
            if (sel.patch)
                pack_selected_kernel<<<blocks, threads, 0, stream>>>(src, ld_s, src_idx, sel, plan.S_pad, packed);
            else
                pack_sources_kernel<<<blocks, threads, 0, stream>>>(src, S < 0 ? 0 : S, ld_s, src_idx, plan.S_pad,
                                                                    packed);
            count_launch(1);
        }
    }
    const float4* pk = packed;
    int rc;
#define DNP_LAUNCH(TP)                                                                                       \
    rc = potential ? launch_pairs_tp<TP, true>(plan, pk, tgt, M, ld_t, lattice_n, eps, partial, ep, sel, stream)  \
                   : launch_pairs_tp<TP, false>(plan, pk, tgt, M, ld_t, lattice_n, eps, partial, ep, sel, stream)
    if (plan.tile_src == kTileSmall) {
        rc = potential ? launch_pairs_tp<1, true, kTileSmall>(plan, pk, tgt, M, ld_t, lattice_n, eps, partial, ep, sel,
                                                              stream)
                       : launch_pairs_tp<1, false, kTileSmall>(plan, pk, tgt, M, ld_t, lattice_n, eps, partial, ep,
                                                               sel, stream);
    } else switch (plan.tp) {
        case 4: DNP_LAUNCH(4); break;
        case 2: DNP_LAUNCH(2); break;
        default: DNP_LAUNCH(1); break;
    }
#undef DNP_LAUNCH
    if (rc != DNP_OK) return rc;
    DNP_CHECK_CUDA(cudaGetLastError());
    if (plan.nsplit > 1) return run_split_reduce(potential, partial, plan.nsplit, M, ep, stream);
    return DNP_OK;
}

int run_sum(const float* in, int64_t n, int64_t stride, float scale, float* out, cudaStream_t stream) {
    sum_kernel<<<1, 1024, 0, stream>>>(in, n, stride, scale, out);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
