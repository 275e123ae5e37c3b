// dnp_upairs.cu -- all-pairs field / potential with the sources in UNIFORM registers, fed from a
// device-resident constant window.  field_utils.py:96-116 (field leaf) / :46-55 (potential leaf) and the
// recursive chunking around them (:23-44, :73-94).
//
// Why the constant bank (profiles/README.md, round 1): a packed FFMA2 issues every 2 cycles only if it reads
// at most two fresh 64-bit register pairs; a source is the same for all 32 lanes, so it belongs in a uniform
// register (`FFMA2 R, R.pair, UR.F32, R.pair` costs no register-file read), and ptxas fills uniform registers
// at speed only from a constant bank (LDCU) -- values produced by LDS / LDG / REDUX are moved back into vector
// registers.  Round 1 used the kernel-PARAMETER bank (sources by value): that needs the sources on the host,
// i.e. a device->host copy and a stream synchronisation per call.  This file uses the module's own constant
// bank instead:
//
//   * `c_window[2][8064]`: two 32 256-byte slots (896 field sources x 36 B or 1344 potential sources x 24 B).
//     A slot is filled ON THE DEVICE by an asynchronous device-to-device copy from the packed source array
//     (cudaMemcpyToSymbolAsync), ordered before the launch that reads it by the stream.  Nothing touches the
//     host: the call is asynchronous and capturable into a CUDA graph.
//   * one launch = one slot (chunk of sources) x all targets.  The (target tile, source) iteration space of a
//     launch is cut into EQUAL contiguous shares, one per resident CTA (stream-K style): a CTA walks up to a
//     few (tile, source range) segments, so a single launch is exactly one full wave whatever M is -- the
//     500 x 524 288 step of the representatives loop no longer runs 1.15 or 1.38 waves.  A tile cut between
//     CTAs lands in different partial row sets ("ways"); partial rows are accumulated by read-modify-write
//     in L2 (launches that share a row set are serialised by their stream) and summed in a fixed order by
//     split_reduce_kernel together with the fused epilogue: deterministic, no float atomics.
//   * calls with more than one chunk alternate over two lanes (streams), one slot each, so the copy into one
//     slot and the ramp of its launch overlap the other lane's launch.
//   * the window is a per-device resource: calls are queued on it in submission order (a mutex held while
//     enqueueing + an event the next user's stream waits for).  Large calls fill the machine anyway.
//
// Arithmetic per (2 targets x 1 source): 19 packed FP32 operations + 2 MUFU.RSQ + 2 MUFU.RCP (field),
// 12 + 2 MUFU.RSQ (potential); coincident pairs contribute exactly 0 (see dnp_pairs.cu).
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "dnp_common.cuh"
#include "dnp_epilogue.cuh"
#include "dnp_pairmath.cuh"
#include "dnp_pairs.h"

namespace dnp {

namespace {

constexpr int kUBlock = 128;                       // threads per CTA
constexpr int kWindowFloats = 16128;               // 64 512 B of the 64 KB constant bank
constexpr int kSlotFloats = 8064;                  // two slots of 32 256 B ...
constexpr int kQuarterFloats = 4032;               // ... or four of 16 128 B (more launches in flight for fewer targets)
constexpr int kMaxSlots = 4;
constexpr int kSlabQuantum = 12;                   // per-patch launches of the slab schedule: sources per loop iteration
constexpr int kFixTiles = 1 << 20;                 // target tiles a fused single-launch call may have (4 MB of counters)
constexpr float kRsqrtClampU = kRsqrtClamp;         // 2^60 (dnp_pairmath.cuh, with kTinyR4 / kTinyR2)

__constant__ float c_window[kWindowFloats];
// kTinyR4 / kTinyR2 from the constant bank: as uniform-register operands they cost no vector register (an immediate
// would be re-materialised with MOVs inside the loop)
__constant__ float c_tiny[2] = {kTinyR4, kTinyR2};

template <bool kPotential>
struct Win {
    static constexpr int kStride = kPotential ? 6 : 9;            // floats per source in the window
    static constexpr int kFold = kPotential ? 336 : 224;          // sources per first-level sum
};

struct UTargets {
    const float* tgt;
    int64_t M;
    int64_t ld_t;
    const int64_t* tgt_idx;
    const uint8_t* tgt_skip;
};

// sources -> dense window rows: field [x y z | nx ny nz | -3nx -3ny -3nz], potential [x y z | nx ny nz];
// rows >= S are null dipoles.  n_dev (optional): the live row count is read on the device.
template <bool kPotential>
__global__ void pack_window_kernel(const float* __restrict__ src, int64_t S, int64_t ld_s,
                                   const int64_t* __restrict__ idx, int64_t S_pad, float* __restrict__ packed) {
    constexpr int ST = Win<kPotential>::kStride;
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S_pad) return;
    float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (s < S) {
        const float* p = src + (idx ? idx[s] : s) * ld_s;
#pragma unroll
        for (int c = 0; c < 6; ++c) v[c] = __ldg(p + c);
    }
    float* o = packed + ST * s;
#pragma unroll
    for (int c = 0; c < 6; ++c) o[c] = v[c];
    if (!kPotential) { o[6] = v[3] * -3.f; o[7] = v[4] * -3.f; o[8] = v[5] * -3.f; }
}

// fill of a slot by a kernel writing through the window's global address (DNP_RING_FILL=2, experiment)
__global__ void fill_window_kernel(const float4* __restrict__ from, float4* __restrict__ to, int n4) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n4) to[i] = from[i];
}

// Field contributions of window sources [s0, s0 + n) (n a multiple of 4) to TP target pairs.
// kSafe = false: no clamp at all -- a coincident pair (r2 = 0, u = +inf) makes t = (m.R) u^2 = 0 * inf = NaN and
// poisons that lane's sums, which the caller detects and repairs by re-running the fold with kSafe = true (clamped
// u: the pair then adds exactly 0).  Coincidences are exact duplicates of a target among the sources: rare, and every
// FMNMX saved is an issue cycle of the ~45 a pair of pairs costs (dnp_pairmath.cuh).
template <int TP, bool kSafe, int kQuads = 1>
__device__ __forceinline__ void field_fold(const float* __restrict__ win, int s0, int n, const f32x2 (&qx)[TP],
                                           const f32x2 (&qy)[TP], const f32x2 (&qz)[TP], f32x2 eps2, f32x2 tiny4,
                                           f32x2 (&ax)[TP], f32x2 (&ay)[TP], f32x2 (&az)[TP]) {
    // kQuads > 1 (n a multiple of 4 kQuads): that many source quads per iteration, for callers whose n is a run-time
    // value -- ptxas unrolls a loop of constant length three quads deep by itself, a loop of unknown length not at all
    for (int sq = s0; sq < s0 + n; sq += 4 * kQuads)
#pragma unroll
    for (int quad = 0; quad < kQuads; ++quad) {
        const int s = sq + 4 * quad;
        const float4* wq = reinterpret_cast<const float4*>(win + 9 * s);
        float v[36];
#pragma unroll
        for (int i = 0; i < 9; ++i) { const float4 f = wq[i]; v[4 * i] = f.x; v[4 * i + 1] = f.y; v[4 * i + 2] = f.z; v[4 * i + 3] = f.w; }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float px = v[9 * k], py = v[9 * k + 1], pz = v[9 * k + 2];
            const float nx = v[9 * k + 3], ny = v[9 * k + 4], nz = v[9 * k + 5];
            const float mx = v[9 * k + 6], my = v[9 * k + 7], mz = v[9 * k + 8];
            const f32x2 px2 = pack2(px, px), py2 = pack2(py, py), pz2 = pack2(pz, pz);
            const f32x2 nx2 = pack2(nx, nx), ny2 = pack2(ny, ny), nz2 = pack2(nz, nz);
            const f32x2 mx2 = pack2(mx, mx), my2 = pack2(my, my), mz2 = pack2(mz, mz);
#pragma unroll
            for (int j = 0; j < TP; ++j) {
                const f32x2 rx = sub2(px2, qx[j]), ry = sub2(py2, qy[j]), rz = sub2(pz2, qz[j]);
                const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, mul2(rx, rx)));
                float r2a, r2b;
                unpack2(r2, r2a, r2b);
                const float ua = rsqrt_sfu(r2a), ub = rsqrt_sfu(r2b);              // +inf on a coincident pair
                const f32x2 u = pack2(ua, ub);
                const f32x2 uc = kSafe ? pack2(fminf(ua, kRsqrtClampU), fminf(ub, kRsqrtClampU)) : u;
                const f32x2 den = fma2(fma2(r2, r2, tiny4), u, eps2);              // |R|^3 + eps; +inf if R = 0
                float da, db;
                unpack2(den, da, db);
                const f32x2 w = pack2(rcp_sfu(da), rcp_sfu(db));                   // rcp(+inf) = +0
                const f32x2 d = fma2(mz2, rz, fma2(my2, ry, mul2(mx2, rx)));       // -3 n.R (exactly 0 when R = 0)
                const f32x2 t = mul2(d, mul2(uc, uc));
                const f32x2 vx = fma2(t, rx, nx2), vy = fma2(t, ry, ny2), vz = fma2(t, rz, nz2);
                ax[j] = fma2(w, vx, ax[j]); ay[j] = fma2(w, vy, ay[j]); az[j] = fma2(w, vz, az[j]);
            }
        }
    }
}

// true if any of the partial sums is NaN or Inf (x * 0 is NaN for both)
template <int TP>
__device__ __forceinline__ bool any_nonfinite(const f32x2 (&ax)[TP], const f32x2 (&ay)[TP], const f32x2 (&az)[TP]) {
    float chk = 0.f;
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float a, b;
        unpack2(ax[j], a, b); chk = fmaf(a, 0.f, chk); chk = fmaf(b, 0.f, chk);
        unpack2(ay[j], a, b); chk = fmaf(a, 0.f, chk); chk = fmaf(b, 0.f, chk);
        unpack2(az[j], a, b); chk = fmaf(a, 0.f, chk); chk = fmaf(b, 0.f, chk);
    }
    return chk != 0.f;                                                             // NaN != 0
}


// Where each CTA's share of the (tile, source) space starts.  Computed on the host and passed by value so that
// the kernel reads it with LDCU: everything the source index depends on has to stay in the uniform datapath
// (an in-kernel division would land in vector registers and turn every LDCU of the window into an LDC).
constexpr int kMaxGrid = 148 * 12 + 16;
struct ShareTable {
    int tile[kMaxGrid];          // first tile of CTA c
    // first source QUAD inside that tile (9 bits) | partial row set of that first segment << 9 (7 bits) |
    // CTAs sharing the tile the share BEGINS in << 16 (8 bits) | CTAs sharing the tile it ENDS in << 24 (8 bits).
    // One word on purpose: a third table read per CTA made ptxas give up the uniform datapath (see the note below).
    unsigned q_way[kMaxGrid];
};

// What a single-launch call does with a finished segment (fused != 0): a tile covered by ONE segment goes straight
// through the epilogue; a cut tile's segments store their partial rows, and whichever CTA arrives last at the
// tile's counter sums the rows of all ways IN WAY ORDER (so the result does not depend on who is last) and applies
// the epilogue.  Counters live in a per-device buffer that is all zero between calls (the last arriver resets).
struct FusedFix {
    int enabled;
    int* counters;
    PairEpilogue ep;
};

// One launch: the window slot at float offset win_off (4 n_q sources) x all target tiles.  CTA c owns per_q quads of the flattened
// (tile, source) space, starting where the table says.
//
// NOTE for maintainers: the share LENGTH is a plain kernel parameter on purpose.  Whenever the trip count of the
// segment loop depended on a value loaded from the table (per-CTA lengths, an end position, bits of q_way), ptxas
// 12.9 moved the whole index chain -- blockIdx included -- into vector registers and every LDCU.64 of the window
// became an LDC.64 (tests/test_abi.py::test_window_kernel_reads_sources_through_uniform_registers guards this).
template <bool kPotential, int TP>
__global__ void __launch_bounds__(kUBlock, (TP >= 4 ? 3 : (TP == 2 ? 5 : 8)))
upairs_kernel(const __grid_constant__ ShareTable tab, int win_q, int n_q, int m_tiles, int per_q, UTargets T,
              float eps, float* __restrict__ part, int64_t way_stride, const __grid_constant__ FusedFix fix) {
    constexpr int ST = Win<kPotential>::kStride;
    constexpr int kFold = Win<kPotential>::kFold;
    constexpr int NC = kPotential ? 1 : 3;
    const int tid = threadIdx.x;
    const float* __restrict__ win = c_window + 4 * win_q;     // slot offsets in units of 4 floats: 16-byte aligned reads
    const f32x2 zero2 = pack2(0.f, 0.f);
    const f32x2 eps2 = pack2(eps, eps);
    const float t4 = c_tiny[0], t2 = c_tiny[1];
    const f32x2 tiny4 = pack2(t4, t4), tiny2 = pack2(t2, t2);

    // all source positions in QUADS (4 sources; n_q and the table): lets ptxas prove the alignment of the
    // LDCU.64 window reads, and keeps shifts / divisions out of the kernel
    int tile = tab.tile[blockIdx.x];
    int q_first = (int)(tab.q_way[blockIdx.x] & 0x1ffu);
    // which partial row set: the number of CTA boundaries between the start of the tile and the segment; later
    // segments of this CTA start their tile, i.e. way 0
    int way = (int)((tab.q_way[blockIdx.x] >> 9) & 0x7fu);
    int left = per_q;                                 // quads this CTA still owns
    // CTAs sharing the tile this segment lies in (meaningful for cut tiles only): the share's first tile now, the
    // tile it ends in for every later segment
    int n_ways = (int)((tab.q_way[blockIdx.x] >> 16) & 0xffu);
    const int n_ways_end = (int)(tab.q_way[blockIdx.x] >> 24);
    __shared__ int s_is_last;

    while (left > 0 && tile < m_tiles) {
        int q_last = q_first + left;
        if (q_last > n_q) q_last = n_q;
        const int64_t tile_base = (int64_t)tile * (kUBlock * 2 * TP);

        f32x2 qx[TP], qy[TP], qz[TP];
#pragma unroll
        for (int j = 0; j < TP; ++j) {
            float x[2] = {0.f, 0.f}, y[2] = {0.f, 0.f}, z[2] = {0.f, 0.f};
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int64_t m = tile_base + (int64_t)(2 * j + h) * kUBlock + tid;
                if (m < T.M) {
                    const int64_t r = T.tgt_idx ? T.tgt_idx[m] : m;
                    if (!(T.tgt_skip && T.tgt_skip[r])) {
                        const float* q = T.tgt + r * T.ld_t;
                        x[h] = __ldg(q); y[h] = __ldg(q + 1); z[h] = __ldg(q + 2);
                    }
                }
            }
            qx[j] = pack2(x[0], x[1]); qy[j] = pack2(y[0], y[1]); qz[j] = pack2(z[0], z[1]);
        }

        f32x2 tx[TP], ty[TP], tz[TP];
#pragma unroll
        for (int j = 0; j < TP; ++j) { tx[j] = zero2; ty[j] = zero2; tz[j] = zero2; }

        for (int f0 = q_first; f0 < q_last; f0 += kFold / 4) {
            const int f1 = (f0 + kFold / 4 < q_last) ? f0 + kFold / 4 : q_last;
            if (kPotential) {
                f32x2 phi[TP];
#pragma unroll
                for (int j = 0; j < TP; ++j) phi[j] = zero2;
                for (int s4 = 4 * f0; s4 < 4 * f1; s4 += 4) {
                    __builtin_assume((s4 & 3) == 0);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int s = s4 + k;
                        const float px = win[ST * s], py = win[ST * s + 1], pz = win[ST * s + 2];
                        const float nx = win[ST * s + 3], ny = win[ST * s + 4], nz = win[ST * s + 5];
                        const f32x2 px2 = pack2(px, px), py2 = pack2(py, py), pz2 = pack2(pz, pz);
                        const f32x2 nx2 = pack2(nx, nx), ny2 = pack2(ny, ny), nz2 = pack2(nz, nz);
#pragma unroll
                        for (int j = 0; j < TP; ++j) {
                            const f32x2 rx = sub2(px2, qx[j]), ry = sub2(py2, qy[j]), rz = sub2(pz2, qz[j]);
                            const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, fma2(rx, rx, tiny2)));
                            float r2a, r2b;
                            unpack2(r2, r2a, r2b);
                            const f32x2 u = pack2(rsqrt_sfu(r2a), rsqrt_sfu(r2b));      // finite: r2 >= kTinyR2
                            const f32x2 d = fma2(nz2, rz, fma2(ny2, ry, mul2(nx2, rx)));
                            phi[j] = fma2(d, mul2(mul2(u, u), u), phi[j]);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < TP; ++j) tx[j] = add2(tx[j], phi[j]);
            } else {
                // E += w (n + t R),  t = (m.R) u^2,  m = -3 n,  u = rsqrt(r2),  w = 1 / (r2^2 u + eps)
                f32x2 ax[TP], ay[TP], az[TP];
#pragma unroll
                for (int j = 0; j < TP; ++j) { ax[j] = ay[j] = az[j] = zero2; }
                field_fold<TP, false>(win, 4 * f0, 4 * (f1 - f0), qx, qy, qz, eps2, tiny4, ax, ay, az);
                if (__any_sync(0xffffffffu, any_nonfinite<TP>(ax, ay, az))) {    // a duplicate of a target among the sources
#pragma unroll
                    for (int j = 0; j < TP; ++j) { ax[j] = ay[j] = az[j] = zero2; }
                    field_fold<TP, true>(win, 4 * f0, 4 * (f1 - f0), qx, qy, qz, eps2, tiny4, ax, ay, az);
                }
#pragma unroll
                for (int j = 0; j < TP; ++j) {
                    tx[j] = add2(tx[j], ax[j]);
                    ty[j] = add2(ty[j], ay[j]);
                    tz[j] = add2(tz[j], az[j]);
                }
            }
        }

        float* rows = part + (int64_t)way * way_stride;
        if (!fix.enabled) {
            // several launches share the row sets: accumulate (zeroed at the start of the call)
#pragma unroll
            for (int j = 0; j < TP; ++j) {
                float ex[2], ey[2], ez[2];
                unpack2(tx[j], ex[0], ex[1]);
                unpack2(ty[j], ey[0], ey[1]);
                unpack2(tz[j], ez[0], ez[1]);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int64_t m = tile_base + (int64_t)(2 * j + h) * kUBlock + tid;
                    if (m >= T.M) continue;
                    float* o = rows + m * NC;
                    if (kPotential) {
                        o[0] += ex[h];
                    } else {
                        o[0] += ex[h]; o[1] += ey[h]; o[2] += ez[h];
                    }
                }
            }
        } else {
            const bool whole = q_first == 0 && q_last == n_q;
            if (!whole) {
#pragma unroll
                for (int j = 0; j < TP; ++j) {
                    float ex[2], ey[2], ez[2];
                    unpack2(tx[j], ex[0], ex[1]);
                    unpack2(ty[j], ey[0], ey[1]);
                    unpack2(tz[j], ez[0], ez[1]);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int64_t m = tile_base + (int64_t)(2 * j + h) * kUBlock + tid;
                        if (m >= T.M) continue;
                        float* o = rows + m * NC;
                        if (kPotential) {
                            __stcg(o, ex[h]);
                        } else {
                            __stcg(o, ex[h]); __stcg(o + 1, ey[h]); __stcg(o + 2, ez[h]);
                        }
                    }
                }
                __threadfence();
                __syncthreads();
                if (tid == 0) {
                    const int seen = atomicAdd(fix.counters + tile, 1);
                    // the last arriver learns the number of row sets through shared memory (0: not last); routing it
                    // through memory also keeps the table word out of the loop bound below (see the note above)
                    s_is_last = seen == n_ways - 1 ? n_ways : 0;
                    if (s_is_last) fix.counters[tile] = 0;             // everyone has arrived: clean for the next call
                }
                __syncthreads();
            }
            const int n_sum = whole ? 0 : s_is_last;
            if (whole || n_sum) {
                if (!whole) __threadfence();
#pragma unroll
                for (int j = 0; j < TP; ++j) {
                    float ex[2], ey[2], ez[2];
                    unpack2(tx[j], ex[0], ex[1]);
                    unpack2(ty[j], ey[0], ey[1]);
                    unpack2(tz[j], ez[0], ez[1]);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int64_t m = tile_base + (int64_t)(2 * j + h) * kUBlock + tid;
                        if (m >= T.M) continue;
                        const int64_t r = T.tgt_idx ? T.tgt_idx[m] : m;
                        if (T.tgt_skip && T.tgt_skip[r]) continue;
                        float vx = ex[h], vy = ey[h], vz = ez[h];
                        if (!whole) {                                  // fixed order: way 0, 1, ... (L2 reads)
                            vx = vy = vz = 0.f;
                            for (int w = 0; w < n_sum; ++w) {
                                const float* o = part + (int64_t)w * way_stride + m * NC;
                                vx += __ldcg(o);
                                if (!kPotential) { vy += __ldcg(o + 1); vz += __ldcg(o + 2); }
                            }
                        }
                        if (kPotential) apply_potential_epilogue(fix.ep, r, vx);
                        else apply_field_epilogue(fix.ep, r, vx, vy, vz);
                    }
                }
            }
            __syncthreads();                                           // s_is_last is reused by the next segment
        }
        n_ways = n_ways_end;
        left -= q_last - q_first;
        ++tile;
        q_first = 0;
        way = 0;
    }
}


// The same arithmetic with ONE target tile per CTA and the whole slot per CTA (no share table, no cut tiles): the
// launch shape of calls with many chunks and >= 256 target tiles, where 2 or 4 launches in flight (one per lane)
// keep every SM supplied with CTAs that start and end at different times.  `first`: this lane's first launch
// stores its partial rows, later launches accumulate (launches of a lane are serialised by its stream).
// n_src is a multiple of the fold (224 field / 336 potential sources).
// (16 targets per thread for the potential -- 3 LDCU.64 per 8 packed groups instead of 4 -- measured 2.6 % SLOWER at
// 1M x 1M, profiles/README.md r02b: the issue model's 0.25 cycle per group is less than what 160 registers cost.)
// kPotUnroll / kMinB: sources per iteration of the potential loop and the CTAs per SM the registers are capped for
// (the field loop is unrolled by ptxas, its 166 registers allow three CTAs).
// kRagged (field only): n_src is any multiple of 12 and the last fold is shorter -- one launch per PATCH of the
// greedy drivers' slab schedule (run_upairs_patch_slabs below), whose patches are a few hundred rows.
template <bool kPotential, int TP = 4, bool kRagged = false, int kPotUnroll = 2, int kMinB = 3>
__global__ void __launch_bounds__(kUBlock, kMinB)
upairs_tile_kernel(int win_q, int n_src, UTargets T, float eps, float* __restrict__ part, int first) {
    static_assert(!(kPotential && kRagged), "ragged folds exist for the field only");
    constexpr int kFold = kRagged ? 228 : Win<kPotential>::kFold;   // ragged: n_src and the fold are multiples of 12
    const float* __restrict__ win = c_window + 4 * win_q;     // slot offsets in units of 4 floats: 16-byte aligned reads
    const int tid = threadIdx.x;
    const int64_t tile_base = (int64_t)blockIdx.x * (kUBlock * 2 * TP);

    f32x2 qx[TP], qy[TP], qz[TP];
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float x[2] = {0.f, 0.f}, y[2] = {0.f, 0.f}, z[2] = {0.f, 0.f};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t m = tile_base + (int64_t)(2 * j + h) * kUBlock + tid;
            if (m < T.M) {
                const int64_t r = T.tgt_idx ? T.tgt_idx[m] : m;
                if (!(T.tgt_skip && T.tgt_skip[r])) {
                    const float* q = T.tgt + r * T.ld_t;
                    x[h] = __ldg(q); y[h] = __ldg(q + 1); z[h] = __ldg(q + 2);
                }
            }
        }
        qx[j] = pack2(x[0], x[1]); qy[j] = pack2(y[0], y[1]); qz[j] = pack2(z[0], z[1]);
    }

    const f32x2 zero2 = pack2(0.f, 0.f);
    const f32x2 eps2 = pack2(eps, eps);
    const float t4 = c_tiny[0], t2 = c_tiny[1];
    const f32x2 tiny4 = pack2(t4, t4), tiny2 = pack2(t2, t2);
    f32x2 tx[TP], ty[TP], tz[TP];
#pragma unroll
    for (int j = 0; j < TP; ++j) { tx[j] = zero2; ty[j] = zero2; tz[j] = zero2; }

    for (int s0 = 0; s0 < n_src; s0 += kFold) {
        if (kPotential) {
            f32x2 phi[TP];
#pragma unroll
            for (int j = 0; j < TP; ++j) phi[j] = zero2;
#pragma unroll kPotUnroll
            for (int s = s0; s < s0 + kFold; ++s) {
                const float px = win[6 * s], py = win[6 * s + 1], pz = win[6 * s + 2];
                const float nx = win[6 * s + 3], ny = win[6 * s + 4], nz = win[6 * s + 5];
                const f32x2 px2 = pack2(px, px), py2 = pack2(py, py), pz2 = pack2(pz, pz);
                const f32x2 nx2 = pack2(nx, nx), ny2 = pack2(ny, ny), nz2 = pack2(nz, nz);
#pragma unroll
                for (int j = 0; j < TP; ++j) {
                    const f32x2 rx = sub2(px2, qx[j]), ry = sub2(py2, qy[j]), rz = sub2(pz2, qz[j]);
                    const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, fma2(rx, rx, tiny2)));
                    float r2a, r2b;
                    unpack2(r2, r2a, r2b);
                    const f32x2 u = pack2(rsqrt_sfu(r2a), rsqrt_sfu(r2b));              // finite: r2 >= kTinyR2
                    const f32x2 d = fma2(nz2, rz, fma2(ny2, ry, mul2(nx2, rx)));
                    phi[j] = fma2(d, mul2(mul2(u, u), u), phi[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < TP; ++j) tx[j] = add2(tx[j], phi[j]);
        } else {
            f32x2 ax[TP], ay[TP], az[TP];
#pragma unroll
            for (int j = 0; j < TP; ++j) { ax[j] = ay[j] = az[j] = zero2; }
            const int n_fold = kRagged ? min(kFold, n_src - s0) : kFold;
            field_fold<TP, false, kRagged ? 3 : 1>(win, s0, n_fold, qx, qy, qz, eps2, tiny4, ax, ay, az);
            if (__any_sync(0xffffffffu, any_nonfinite<TP>(ax, ay, az))) {        // a duplicate of a target among the sources
#pragma unroll
                for (int j = 0; j < TP; ++j) { ax[j] = ay[j] = az[j] = zero2; }
                field_fold<TP, true>(win, s0, n_fold, qx, qy, qz, eps2, tiny4, ax, ay, az);
            }
#pragma unroll
            for (int j = 0; j < TP; ++j) {
                tx[j] = add2(tx[j], ax[j]);
                ty[j] = add2(ty[j], ay[j]);
                tz[j] = add2(tz[j], az[j]);
            }
        }
    }

#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float ex[2], ey[2], ez[2];
        unpack2(tx[j], ex[0], ex[1]);
        unpack2(ty[j], ey[0], ey[1]);
        unpack2(tz[j], ez[0], ez[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t m = tile_base + (int64_t)(2 * j + h) * kUBlock + tid;
            if (m >= T.M) continue;
            if (kPotential) {
                float* o = part + m;
                *o = first ? ex[h] : *o + ex[h];
            } else {
                float* o = part + m * 3;
                if (first) { o[0] = ex[h]; o[1] = ey[h]; o[2] = ez[h]; }
                else { o[0] += ex[h]; o[1] += ey[h]; o[2] += ez[h]; }
            }
        }
    }
}

// ---------------------------------------------------------------- the per-device window
struct DeviceRing {
    std::mutex mu;
    bool ready = false;
    cudaStream_t lane[kMaxSlots] = {};
    cudaEvent_t lane_done[kMaxSlots] = {};
    cudaEvent_t fork = nullptr;
    cudaEvent_t last_done = nullptr;       // end of the previous user's window accesses
    bool has_last = false;
    cudaStream_t last_stream = nullptr;
    float* window = nullptr;               // global address of c_window on this device
    int* counters = nullptr;               // per-tile arrival counters of the fused fix-up, all zero between calls

    cudaError_t prepare() {
        if (ready) return cudaSuccess;
        cudaError_t e;
        for (int i = 0; i < kMaxSlots; ++i) {
            if ((e = cudaStreamCreateWithFlags(&lane[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&lane_done[i], cudaEventDisableTiming)) != cudaSuccess) return e;
        }
        if ((e = cudaEventCreateWithFlags(&fork, cudaEventDisableTiming)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&last_done, cudaEventDisableTiming)) != cudaSuccess) return e;
        if ((e = cudaGetSymbolAddress(reinterpret_cast<void**>(&window), c_window)) != cudaSuccess) return e;
        if ((e = cudaMalloc(reinterpret_cast<void**>(&counters), (size_t)kFixTiles * sizeof(int))) != cudaSuccess) return e;
        if ((e = cudaMemset(counters, 0, (size_t)kFixTiles * sizeof(int))) != cudaSuccess) return e;
        ready = true;
        return cudaSuccess;
    }
};
DeviceRing g_ring[kMaxDevices];

int env_int_u(const char* name, int dflt) {
    const char* v = getenv(name);
    return v ? atoi(v) : dflt;
}

inline int64_t round_up_u(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

bool capturing(cudaStream_t stream) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) { cudaGetLastError(); return false; }
    return st != cudaStreamCaptureStatusNone;
}

// DNP_RING_FILL: 0 cudaMemcpyToSymbolAsync (default), 1 plain device-to-device copy to the window's global address,
// 2 a copy kernel writing through that address -- all three are seen by the next launch (tests/test_gpu_upairs.py)
cudaError_t fill_slot(DeviceRing& ring, int win_off, const float* from, size_t bytes, cudaStream_t stream) {
    const int mode = env_int_u("DNP_RING_FILL", 0);
    if (mode == 3) return cudaSuccess;          // timing experiment only: stale window, wrong results
    if (mode == 1)
        return cudaMemcpyAsync(ring.window + win_off, from, bytes, cudaMemcpyDeviceToDevice, stream);
    if (mode == 2) {
        const int n4 = (int)(bytes / 16);
        fill_window_kernel<<<(n4 + 255) / 256, 256, 0, stream>>>(
            reinterpret_cast<const float4*>(from), reinterpret_cast<float4*>(ring.window + win_off), n4);
        return cudaGetLastError();
    }
    return cudaMemcpyToSymbolAsync(c_window, from, bytes, (size_t)win_off * sizeof(float), cudaMemcpyDeviceToDevice,
                                   stream);
}

// Cuts the m_tiles x n_src iteration space of one launch into `grid` equal contiguous shares of `per` sources
// (a multiple of 4) and returns the number of partial row sets the cut needs (tab may be null).
int build_shares(int m_tiles, int n_src, int grid, int per, ShareTable* tab) {
    const int n_q = n_src >> 2, per_q = per >> 2;
    const int64_t total_q = (int64_t)m_tiles * n_q;
    // CTAs touching tile t: those whose share [c per_q, (c + 1) per_q) meets [t n_q, (t + 1) n_q)
    auto ctas_of_tile = [&](int64_t t) {
        const int64_t first = (t * n_q) / per_q, last = ((t + 1) * n_q - 1) / per_q;
        return (int)(last - first + 1);
    };
    int max_way = 0;
    for (int c = 0; c < grid; ++c) {
        const int64_t begin = (int64_t)c * per_q;
        int64_t end = begin + per_q;
        if (end > total_q) end = total_q;
        const int64_t tile = begin < total_q ? begin / n_q : m_tiles;
        // a share that starts inside a tile continues it: its way is its distance from the CTA holding the tile's start
        const int way = (end > begin && begin != tile * n_q) ? c - (int)((tile * n_q) / per_q) : 0;
        if (way > max_way) max_way = way;
        if (tab) {
            tab->tile[c] = (int)tile;
            const int wb = end > begin ? ctas_of_tile(tile) : 1, we = end > begin ? ctas_of_tile((end - 1) / n_q) : 1;
            tab->q_way[c] = (unsigned)(begin - tile * n_q) | ((unsigned)way << 9) | ((unsigned)wb << 16) | ((unsigned)we << 24);
        }
    }
    return max_way + 1;
}

template <bool kPotential, int TP>
void launch_shares(const UPairPlan& plan, int win_off, const UTargets& T, float eps, float* part, const FusedFix& fix,
                   cudaStream_t stream) {
    ShareTable tab;
    build_shares(plan.m_tiles, plan.chunk, plan.grid, plan.per, &tab);
    upairs_kernel<kPotential, TP><<<plan.grid, kUBlock, 0, stream>>>(tab, win_off >> 2, plan.chunk >> 2, plan.m_tiles,
                                                                      plan.per >> 2, T, eps, part, plan.way_stride, fix);
}

// one launch over the slot at win_off: whole tiles (plan.whole_tiles) or equal shares
void launch_chunk(bool potential, const UPairPlan& plan, int win_off, const UTargets& T, float eps, float* part,
                  int first, const FusedFix& fix, cudaStream_t stream) {
    if (plan.whole_tiles) {
        // potential: four sources per loop iteration at four CTAs per SM (124 registers) -- 2.898e12 pairs/s at 1M x 1M
        // against 2.832e12 for two sources per iteration; 3 / 4 / 6 / 8 sources at 3 CTAs per SM (134 - 158 registers)
        // and 4 sources at 5 CTAs (94 registers) are all slower (profiles/README.md r02e)
        if (potential) upairs_tile_kernel<true, 4, false, 4, 4><<<plan.m_tiles, kUBlock, 0, stream>>>(win_off >> 2, plan.chunk, T, eps, part, first);
        else upairs_tile_kernel<false><<<plan.m_tiles, kUBlock, 0, stream>>>(win_off >> 2, plan.chunk, T, eps, part, first);
        return;
    }
    if (potential) {
        if (plan.tp == 4) launch_shares<true, 4>(plan, win_off, T, eps, part, fix, stream);
        else launch_shares<true, 1>(plan, win_off, T, eps, part, fix, stream);
    } else {
        if (plan.tp == 4) launch_shares<false, 4>(plan, win_off, T, eps, part, fix, stream);
        else launch_shares<false, 1>(plan, win_off, T, eps, part, fix, stream);
    }
}

}  // namespace

// DNP_RING: 0 = never, 1 = by the size rule (default), 2 = always (tests).
bool use_ring_path(int64_t S, int64_t M, int lattice_n) {
    const int mode = env_int_u("DNP_RING", 1);
    if (mode == 0 || lattice_n > 0) return false;
    if (mode == 2) return S >= 1 && M >= 1;
    // tools/ring_probe.py (profiles/README.md, r02): a call pays ~25 us for the window copy, the zeroing of the
    // partial rows and the reduction pass; below ~2e8 pairs, or with target sets too small to amortise one launch
    // per 896 sources, the shared-memory kernel is faster
    const int64_t min_pairs = (int64_t)env_int_u("DNP_RING_MIN_MPAIRS", 200) * 1000000;
    return S >= 64 && M >= 32768 && S * M >= min_pairs;
}

UPairPlan make_upair_plan(bool potential, int64_t S, int64_t M, int sm_count) {
    UPairPlan p;
    const int stride = potential ? 6 : 9;
    const int fold = potential ? Win<true>::kFold : Win<false>::kFold;
    if (S < 1) S = 1;
    if (M < 1) M = 1;
    const int64_t S4 = round_up_u(S, 4);
    const int64_t tiles4 = (M + (int64_t)kUBlock * 8 - 1) / ((int64_t)kUBlock * 8);
    const int slots3 = sm_count * 3;                         // resident CTAs of the 8-targets-per-thread kernels

    // ---- whole tiles: many chunks x many tiles.  2 lanes when one launch alone is >= ~1.15 waves of CTAs, else 4
    // lanes on quarter slots, so that >= ~2.3 waves are in flight and wave tails overlap the next launch's head.
    int mode = env_int_u("DNP_RING_MODE", 0);                // 1: force shares, 2: force whole tiles (tests)
    const int cap2 = kSlotFloats / stride / fold * fold;     // 896 / 1344
    const int cap4 = kQuarterFloats / stride / fold * fold;  // 448 / 672
    const bool tiles_ok = tiles4 * 4 >= (int64_t)slots3 * 23 / 10 && S4 >= 4 * (int64_t)cap4;
    if (mode == 2 || (mode == 0 && tiles_ok)) {
        p.whole_tiles = true;
        p.tp = 4;
        p.m_tiles = (int)tiles4;
        p.lanes = tiles4 * 2 >= (int64_t)slots3 * 23 / 10 ? 2 : 4;
        { const int forced = env_int_u("DNP_RING_LANES", 0); if (forced == 2 || forced == 4) p.lanes = forced; }
        const int cap = p.lanes == 2 ? cap2 : cap4;
        p.slot_floats = p.lanes == 2 ? kSlotFloats : kQuarterFloats;
        p.n_chunks = (int)((S4 + cap - 1) / cap);
        p.chunk = (int)round_up_u((S4 + p.n_chunks - 1) / p.n_chunks, fold);
        if (p.lanes > p.n_chunks) p.lanes = p.n_chunks;
        p.S_pad = (int64_t)p.n_chunks * p.chunk;
        p.grid = p.m_tiles;
        p.per = p.chunk;
        p.ways = 1;
    } else {
        // ---- equal shares: targets per thread 8 (1024-target tiles) unless that pads a small set by more than ~6 %
        p.tp = 4;
        if ((double)(tiles4 * kUBlock * 8 - M) > 0.06 * (double)M) p.tp = 1;
        { const int forced = env_int_u("DNP_RING_TP", 0); if (forced == 4 || forced == 1) p.tp = forced; }
        const int resident = p.tp == 4 ? 3 : 8;
        const int64_t tile_tgts = (int64_t)kUBlock * 2 * p.tp;
        p.m_tiles = (int)((M + tile_tgts - 1) / tile_tgts);
        const int cap = kSlotFloats / stride;                // 896 / 1344 sources per slot
        p.slot_floats = kSlotFloats;
        p.n_chunks = (int)((S4 + cap - 1) / cap);
        p.chunk = (int)round_up_u((S4 + p.n_chunks - 1) / p.n_chunks, 4);
        p.S_pad = (int64_t)p.n_chunks * p.chunk;
        p.lanes = p.n_chunks > 1 ? 2 : 1;
        // one full wave of shares; shares shorter than ~48 sources are not worth their prologue
        const int64_t total = (int64_t)p.m_tiles * p.chunk;
        int64_t grid = (int64_t)sm_count * resident;
        // a call of ONE launch has no neighbour launch to fill its tail: finer shares (5 per SM instead of the 3
        // resident ones) let the hardware scheduler even out the SMs (tools/ring_probe.py, profiles/README.md r02)
        if (p.n_chunks == 1 && p.tp == 4) grid = (int64_t)sm_count * 5;
        { const int forced = env_int_u("DNP_RING_GRID", 0); if (forced > 0) grid = forced; }
        if (grid > kMaxGrid) grid = kMaxGrid;
        const int64_t min_share = 48;
        if (grid * min_share > total) grid = total / min_share;
        if (grid < 1) grid = 1;
        p.per = (int)round_up_u((total + grid - 1) / grid, 4);
        p.grid = (int)((total + p.per - 1) / p.per);
        p.ways = build_shares(p.m_tiles, p.chunk, p.grid, p.per, nullptr);
    }
    const int ncomp = potential ? 1 : 3;
    p.way_stride = round_up_u(M * ncomp, 64);
    p.packed_bytes = (size_t)round_up_u(p.S_pad * stride * (int64_t)sizeof(float), 256);
    p.partial_bytes = (size_t)round_up_u((int64_t)p.lanes * p.ways * p.way_stride * (int64_t)sizeof(float), 256);
    return p;
}

size_t upairs_workspace_bytes(int64_t S, int64_t M) {
    const UPairPlan f = make_upair_plan(false, S, M, sm_count_cached());
    const UPairPlan p = make_upair_plan(true, S, M, sm_count_cached());
    const size_t a = f.packed_bytes + f.partial_bytes, b = p.packed_bytes + p.partial_bytes;
    return a > b ? a : b;
}

int run_upairs(bool potential, const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx, const float* tgt,
               int64_t M, int64_t ld_t, float eps, const PairEpilogue& ep, void* workspace, size_t workspace_bytes,
               cudaStream_t stream) {
    if (M <= 0) return DNP_OK;
    const UPairPlan plan = make_upair_plan(potential, S, M, sm_count_cached());
    if (!workspace || workspace_bytes < plan.packed_bytes + plan.partial_bytes) {
        set_error("workspace too small: need %zu bytes, got %zu", plan.packed_bytes + plan.partial_bytes,
                  workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    int dev = 0;
    DNP_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("device index %d out of range", dev);
        return DNP_ERR_INVALID;
    }
    float* packed = reinterpret_cast<float*>(workspace);
    float* partial = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + plan.packed_bytes);
    const int stride = potential ? 6 : 9;
    // a call of ONE launch is fused: sources are packed straight into the window and cut tiles are summed by their
    // last-arriving CTA inside the launch -- no staging copy, no zeroing, no reduction pass (DNP_RING_FUSED=0: off)
    const bool fused = plan.lanes == 1 && !plan.whole_tiles && plan.m_tiles <= kFixTiles && env_int_u("DNP_RING_FUSED", 1) != 0;

    const UTargets T{tgt, M, ld_t, ep.tgt_idx, ep.tgt_skip};
    const size_t chunk_bytes = (size_t)plan.chunk * stride * sizeof(float);
    const int64_t lane_stride = (int64_t)plan.ways * plan.way_stride;
    DeviceRing& ring = g_ring[dev];
    int rc = DNP_OK;
    {
        std::lock_guard<std::mutex> lock(ring.mu);
        DNP_CHECK_CUDA(ring.prepare());
        const bool cap = capturing(stream);
        // queue behind the previous user of the window (same stream: already ordered)
        if (!cap && ring.has_last && ring.last_stream != stream) DNP_CHECK_CUDA(cudaStreamWaitEvent(stream, ring.last_done, 0));
        {
            const int threads = 256;
            const unsigned blocks = (unsigned)((plan.S_pad + threads - 1) / threads);
            float* dst = fused ? ring.window : packed;
            if (potential)
                pack_window_kernel<true><<<blocks, threads, 0, stream>>>(src, S < 0 ? 0 : S, ld_s, src_idx, plan.S_pad, dst);
            else
                pack_window_kernel<false><<<blocks, threads, 0, stream>>>(src, S < 0 ? 0 : S, ld_s, src_idx, plan.S_pad, dst);
            count_launch(1);
            DNP_CHECK_CUDA(cudaGetLastError());
            // cut tiles of multi-launch calls accumulate into zeroed row sets; whole-tile launches store on a lane's
            // first launch, fused calls store every segment
            if (!plan.whole_tiles && !fused) DNP_CHECK_CUDA(cudaMemsetAsync(partial, 0, plan.partial_bytes, stream));
        }
        FusedFix fix;
        fix.enabled = fused ? 1 : 0;
        fix.counters = ring.counters;
        fix.ep = ep;
        timer_mark(stream);
        if (plan.lanes == 1) {
            cudaError_t e = fused ? cudaSuccess : fill_slot(ring, 0, packed, chunk_bytes, stream);
            if (e != cudaSuccess) rc = cuda_fail(e, "window fill");
            else launch_chunk(potential, plan, 0, T, eps, partial, 1, fix, stream);
            count_launch(1);
        } else {
            DNP_CHECK_CUDA(cudaEventRecord(ring.fork, stream));
            for (int l = 0; l < plan.lanes; ++l) DNP_CHECK_CUDA(cudaStreamWaitEvent(ring.lane[l], ring.fork, 0));
            for (int k = 0; k < plan.n_chunks && rc == DNP_OK; ++k) {
                const int l = k % plan.lanes;
                const int win_off = l * plan.slot_floats;
                cudaError_t e = fill_slot(ring, win_off, packed + (size_t)k * plan.chunk * stride, chunk_bytes, ring.lane[l]);
                if (e != cudaSuccess) { rc = cuda_fail(e, "window fill"); break; }
                launch_chunk(potential, plan, win_off, T, eps, partial + (int64_t)l * lane_stride, k < plan.lanes, fix,
                             ring.lane[l]);
            }
            count_launch(plan.n_chunks);
            for (int l = 0; l < plan.lanes; ++l) {       // always re-join
                cudaEventRecord(ring.lane_done[l], ring.lane[l]);
                cudaStreamWaitEvent(stream, ring.lane_done[l], 0);
            }
        }
        timer_mark(stream);
        if (rc == DNP_OK) {
            cudaError_t e = cudaGetLastError();
            if (e != cudaSuccess) rc = cuda_fail(e, "upairs launch");
        }
        if (!cap) {
            cudaEventRecord(ring.last_done, stream);
            ring.has_last = true;
            ring.last_stream = stream;
        }
    }
    if (rc != DNP_OK || fused) return rc;
    return run_split_reduce(potential, partial, plan.lanes * plan.ways, M, ep, stream, plan.way_stride);
}


// ---------------------------------------------------------------- one field slab per patch (dnp_slab.cu)
namespace {

// Window rows [x y z | n | -3n] of every patch of a CSR, patch p at row pad_off[p] (a multiple of kSlabQuantum; rows
// between the end of a patch and the next multiple are null dipoles).
__global__ void pack_patches_kernel(const float* __restrict__ src, int64_t ld, const int64_t* __restrict__ perm,
                                    const int64_t* __restrict__ offsets, const int64_t* __restrict__ pad_off, int P,
                                    int64_t total_pad, float* __restrict__ packed) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total_pad) return;
    int lo = 0, hi = P;                                   // last p with pad_off[p] <= i (empty patches repeat a value)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (pad_off[mid] <= i) lo = mid; else hi = mid;
    }
    const int64_t local = i - pad_off[lo], o = offsets[lo], n = offsets[lo + 1] - o;
    float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (local < n) {
        const float* q = src + perm[o + local] * ld;
#pragma unroll
        for (int c = 0; c < 6; ++c) v[c] = __ldg(q + c);
    }
    float* w = packed + 9 * i;
#pragma unroll
    for (int c = 0; c < 6; ++c) w[c] = v[c];
    w[6] = v[3] * -3.f; w[7] = v[4] * -3.f; w[8] = v[5] * -3.f;
}

}  // namespace

int upairs_slab_quantum() { return kSlabQuantum; }

size_t upairs_slab_packed_bytes(int64_t total_pad) {
    return (size_t)round_up_u((total_pad > 0 ? total_pad : 4) * 9 * (int64_t)sizeof(float), 256);
}

// slabs[slot(p)][m][0..3) = field of patch p's rows (current normals) at target m = row tgt_idx[m] of pts, for every
// patch with slot(p) = slot_host[p] >= 0 (slot_host == NULL: slot(p) = p): one whole-tile launch per (patch, chunk of <= one window slot), launches of different
// patches rotating over the lanes -- independent launches, so the tail wave of one overlaps the head of the next and
// the whole batch runs at the rate of one large call.  Empty patches get a zeroed slab.  Asynchronous.
// first_target_host (optional): patch p only acts on the targets m >= first_target_host[p] (rows below stay unwritten)
// -- the symmetric form of the slab schedule, which needs each patch PAIR once; patches are then launched from both
// ends of the list alternately, so that large and small launches share the machine.
int run_upairs_patch_slabs(const float* pts, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                           const int64_t* offsets_host, const int64_t* pad_off_dev, const int64_t* pad_off_host,
                           int64_t P, const int32_t* slot_host, const int64_t* first_target_host,
                           const int64_t* tgt_idx, int64_t M, float eps, float* packed, float* slabs,
                           int64_t slab_stride, cudaStream_t stream) {
    if (P <= 0 || M <= 0) return DNP_OK;
    int dev = 0;
    DNP_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("device index %d out of range", dev);
        return DNP_ERR_INVALID;
    }
    const int64_t total_pad = pad_off_host[P];
    const int64_t tiles = (M + (int64_t)kUBlock * 8 - 1) / ((int64_t)kUBlock * 8);
    const int slots3 = sm_count_cached() * 3;
    // as many launches in flight as the window has room for whole patches: 4 slots of 444 sources, 3 of 588, else 2
    // of 888 (4 quarter slots when a launch is too few CTAs) -- a lane's next launch waits for its previous one, so
    // more lanes leave fewer CTA slots idle at the tail of every launch
    int64_t max_n4 = kSlabQuantum;
    for (int64_t p = 0; p < P; ++p) {
        if (slot_host && slot_host[p] < 0) continue;
        const int64_t n4 = round_up_u(offsets_host[p + 1] - offsets_host[p], kSlabQuantum);
        if (n4 > max_n4) max_n4 = n4;
    }
    int lanes = tiles * 2 >= (int64_t)slots3 * 23 / 10 ? 2 : 4;
    if (max_n4 <= kWindowFloats / 4 / 9 / kSlabQuantum * kSlabQuantum) lanes = 4;
    else if (max_n4 <= kWindowFloats / 3 / 9 / kSlabQuantum * kSlabQuantum) lanes = 3;
    { const int forced = env_int_u("DNP_RING_LANES", 0); if (forced >= 2 && forced <= 4) lanes = forced; }
    const int slot_floats = kWindowFloats / lanes;                            // 8064 / 5376 / 4032: multiples of 4
    const int cap = slot_floats / 9 / kSlabQuantum * kSlabQuantum;            // 888 / 588 / 444 sources per launch
    DeviceRing& ring = g_ring[dev];
    int rc = DNP_OK;
    std::lock_guard<std::mutex> lock(ring.mu);
    DNP_CHECK_CUDA(ring.prepare());
    const bool capt = capturing(stream);
    if (!capt && ring.has_last && ring.last_stream != stream) DNP_CHECK_CUDA(cudaStreamWaitEvent(stream, ring.last_done, 0));
    if (total_pad > 0) {
        pack_patches_kernel<<<(unsigned)((total_pad + 255) / 256), 256, 0, stream>>>(
            pts, ld, perm, offsets_dev, pad_off_dev, (int)P, total_pad, packed);
        count_launch(1);
        DNP_CHECK_CUDA(cudaGetLastError());
    }
    timer_mark(stream);
    DNP_CHECK_CUDA(cudaEventRecord(ring.fork, stream));
    for (int l = 0; l < lanes; ++l) DNP_CHECK_CUDA(cudaStreamWaitEvent(ring.lane[l], ring.fork, 0));
    int64_t launches = 0, turn = 0;
    for (int64_t k = 0; k < P && rc == DNP_OK; ++k) {
        const int64_t p = first_target_host ? ((k & 1) ? P - 1 - (k >> 1) : (k >> 1)) : k;
        if (slot_host && slot_host[p] < 0) continue;
        const int64_t n = offsets_host[p + 1] - offsets_host[p];
        int64_t m0 = first_target_host ? first_target_host[p] : 0;
        if (m0 < 0) m0 = 0;
        if (m0 >= M) continue;
        const UTargets T{tgt_idx ? pts : pts + m0 * ld, M - m0, ld, tgt_idx ? tgt_idx + m0 : nullptr, nullptr};
        const int64_t tiles_p = (M - m0 + (int64_t)kUBlock * 8 - 1) / ((int64_t)kUBlock * 8);
        const int l = (int)(turn++ % lanes);
        float* slab = slabs + (slot_host ? (int64_t)slot_host[p] : p) * slab_stride + 3 * m0;
        if (n <= 0) {
            cudaError_t e = cudaMemsetAsync(slab, 0, (size_t)(M - m0) * 3 * sizeof(float), ring.lane[l]);
            if (e != cudaSuccess) rc = cuda_fail(e, "slab clear");
            continue;
        }
        const int64_t n4 = round_up_u(n, kSlabQuantum);
        const int64_t n_chunks = (n4 + cap - 1) / cap;
        const int64_t even = round_up_u((n4 + n_chunks - 1) / n_chunks, kSlabQuantum);      // equal chunks, not cap + a sliver
        for (int64_t c0 = 0; c0 < n4; c0 += even) {
            const int cn = (int)(n4 - c0 < even ? n4 - c0 : even);
            const int win_off = l * slot_floats;
            cudaError_t e = fill_slot(ring, win_off, packed + (size_t)(pad_off_host[p] + c0) * 9,
                                      (size_t)cn * 9 * sizeof(float), ring.lane[l]);
            if (e != cudaSuccess) { rc = cuda_fail(e, "window fill"); break; }
            upairs_tile_kernel<false, 4, true><<<(unsigned)tiles_p, kUBlock, 0, ring.lane[l]>>>(win_off >> 2, cn, T, eps, slab,
                                                                                                c0 == 0 ? 1 : 0);
            ++launches;
        }
    }
    count_launch((int)launches);
    for (int l = 0; l < lanes; ++l) {            // always re-join
        cudaEventRecord(ring.lane_done[l], ring.lane[l]);
        cudaStreamWaitEvent(stream, ring.lane_done[l], 0);
    }
    timer_mark(stream);
    if (rc == DNP_OK) {
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) rc = cuda_fail(e, "patch slab launch");
    }
    if (!capt) {
        cudaEventRecord(ring.last_done, stream);
        ring.has_last = true;
        ring.last_stream = stream;
    }
    return rc;
}

}  // namespace dnp
