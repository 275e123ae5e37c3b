// dnp_cpairs.cu -- all-pairs field / potential for LARGE source x target sets, sources fed from the
// constant bank (uniform registers).
//
// Why a second kernel (profiles/README.md, round 1): the shared-memory kernel of dnp_pairs.cu is bound
// by register-file read bandwidth, not by the FP32 pipe -- a packed FFMA2 whose three operands are three
// distinct register pairs issues every 3 cycles instead of every 2, and 8 of its 20 instructions per
// pair-of-pairs have that form because the (warp-uniform) source values sit in vector registers.
// A source is the same for every lane, so it belongs in a UNIFORM register: FFMA2/FMUL2/FADD2 take a
// `UR.F32` operand that is broadcast to both halves and costs no register-file read.  The only way to
// fill uniform registers at speed is the constant bank (LDCU), and the only per-launch constant bank a
// re-entrant library can own is the kernel-parameter bank: every launch carries a block of sources BY
// VALUE (896 field sources = 32 256 of the 32 764 parameter bytes), streamed by the kernel with LDCU.64.
//
//   * one launch = one block of sources x all targets; launches rotate over L "lanes" (streams), lane l
//     accumulating chunks l, l+L, ... into its own partial rows, so consecutive launches overlap and
//     the tail of one wave is filled by the head of the next; a fixed-order reduction over the lanes
//     ends the call (deterministic, no float atomics).
//   * the sources have to be on the host to be passed by value: device-resident sources are packed,
//     copied to a pinned staging buffer and the stream is synchronised ONCE per call (24 B/source,
//     0.5 ms per million).  The path is therefore only taken for calls worth milliseconds of GPU time
//     (use_const_path); everything smaller keeps the fully asynchronous shared-memory kernel.
//   * arithmetic per pair: 19 packed FP32 operations + 2 MUFU (field), 12 + 1 (potential), the same
//     coincident-pair handling as dnp_pairs.cu; replaces field_utils.py:96-116 / :46-55.
#include <stdlib.h>
#include <string.h>

#include "dnp_common.cuh"
#include "dnp_pairs.h"

namespace dnp {

namespace {

constexpr int kCBlock = 128;                   // threads per CTA
constexpr int kCTP = 4;                        // target pairs per thread (8 targets)
// Sources per launch and per fold into the second-level totals.  A field source travels as 9 floats
// [x y z | nx ny nz | -3nx -3ny -3nz] (the -3 of E = w (n - 3 (n.R) u^2 R) rides on a second copy of the
// moment so that it costs nothing per pair), a potential source as 6.
template <bool kPotential>
struct Blk {
    static constexpr int kStride = kPotential ? 6 : 9;
    static constexpr int kFold = kPotential ? 256 : 224;
    static constexpr int kChunk = kPotential ? 1280 : 896;      // 30 720 B / 32 256 B of parameters
};
constexpr int kCMaxLanes = 8;
constexpr float kRsqrtClamp = 1.152921504606847e18f;   // 2^60
constexpr float kRsqrtClampPhi = 1.099511627776e12f;   // 2^40

template <bool kPotential>
struct SrcBlock {
    float v[Blk<kPotential>::kChunk * Blk<kPotential>::kStride];
};
static_assert(sizeof(SrcBlock<false>) == 32256 && sizeof(SrcBlock<true>) == 30720, "parameter block size");

struct CTargets {
    const float* tgt;
    int64_t M;
    int64_t ld_t;
    const int64_t* tgt_idx;
    const uint8_t* tgt_skip;
};

// dense, zero-padded 24-byte source rows (the layout SrcBlock carries)
__global__ void pack_sources24_kernel(const float* __restrict__ src, int64_t S, int64_t ld_s,
                                      const int64_t* __restrict__ idx, int64_t S_pad, float* __restrict__ packed) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S_pad) return;
    float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // padding rows: a null dipole
    if (s < S) {
        const float* p = src + (idx ? idx[s] : s) * ld_s;
#pragma unroll
        for (int c = 0; c < 6; ++c) v[c] = __ldg(p + c);
    }
#pragma unroll
    for (int c = 0; c < 6; ++c) packed[6 * s + c] = v[c];
}

__constant__ float c_bank3_probe[8064];   // experiment (DNP_CONST_BANK3=1): same kernel, sources from bank 3

template <bool kPotential, bool kBank3 = false>
__global__ void __launch_bounds__(kCBlock, 3)
cpairs_kernel(const __grid_constant__ SrcBlock<kPotential> blk_param, int n_src, CTargets T, float eps,
              float* __restrict__ part, int first) {
    constexpr int TP = kCTP;
    struct View { const float* v; };
    const View blk{kBank3 ? c_bank3_probe : blk_param.v};
    const int tid = threadIdx.x;
    const int64_t tile_base = (int64_t)blockIdx.x * (kCBlock * 2 * TP);

    f32x2 qx[TP], qy[TP], qz[TP];
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float x[2] = {0.f, 0.f}, y[2] = {0.f, 0.f}, z[2] = {0.f, 0.f};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t m = tile_base + (int64_t)(2 * j + h) * kCBlock + tid;
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
    f32x2 tx[TP], ty[TP], tz[TP];
#pragma unroll
    for (int j = 0; j < TP; ++j) { tx[j] = zero2; ty[j] = zero2; tz[j] = zero2; }

    constexpr int kFold = Blk<kPotential>::kFold;
    for (int s0 = 0; s0 < n_src; s0 += kFold) {
        if (kPotential) {
            f32x2 phi[TP];
#pragma unroll
            for (int j = 0; j < TP; ++j) phi[j] = zero2;
#pragma unroll 2
            for (int s = s0; s < s0 + kFold; ++s) {
                const float px = blk.v[6 * s], py = blk.v[6 * s + 1], pz = blk.v[6 * s + 2];
                const float nx = blk.v[6 * s + 3], ny = blk.v[6 * s + 4], nz = blk.v[6 * s + 5];
                const f32x2 px2 = pack2(px, px), py2 = pack2(py, py), pz2 = pack2(pz, pz);
                const f32x2 nx2 = pack2(nx, nx), ny2 = pack2(ny, ny), nz2 = pack2(nz, nz);
#pragma unroll
                for (int j = 0; j < TP; ++j) {
                    const f32x2 rx = sub2(px2, qx[j]), ry = sub2(py2, qy[j]), rz = sub2(pz2, qz[j]);
                    const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, mul2(rx, rx)));
                    float r2a, r2b;
                    unpack2(r2, r2a, r2b);
                    const f32x2 uc = pack2(fminf(rsqrt_sfu(r2a), kRsqrtClampPhi), fminf(rsqrt_sfu(r2b), kRsqrtClampPhi));
                    const f32x2 d = fma2(nz2, rz, fma2(ny2, ry, mul2(nx2, rx)));
                    phi[j] = fma2(d, mul2(mul2(uc, uc), uc), phi[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < TP; ++j) tx[j] = add2(tx[j], phi[j]);
        } else {
            // 19 packed FP32 operations per (2 targets x 1 source), field_utils.py:96-109 regrouped as
            //   E += w (n + t R),  t = (m.R) u^2,  m = -3 n,  u = rsqrt(r2),  w = 1 / (r2^2 u + eps).
            f32x2 ax[TP], ay[TP], az[TP];
#pragma unroll
            for (int j = 0; j < TP; ++j) { ax[j] = ay[j] = az[j] = zero2; }
#pragma unroll 4
            for (int s = s0; s < s0 + kFold; ++s) {
                const float px = blk.v[9 * s], py = blk.v[9 * s + 1], pz = blk.v[9 * s + 2];
                const float nx = blk.v[9 * s + 3], ny = blk.v[9 * s + 4], nz = blk.v[9 * s + 5];
                const float mx = blk.v[9 * s + 6], my = blk.v[9 * s + 7], mz = blk.v[9 * s + 8];
                const f32x2 px2 = pack2(px, px), py2 = pack2(py, py), pz2 = pack2(pz, pz);
                const f32x2 nx2 = pack2(nx, nx), ny2 = pack2(ny, ny), nz2 = pack2(nz, nz);
                const f32x2 mx2 = pack2(mx, mx), my2 = pack2(my, my), mz2 = pack2(mz, mz);
#pragma unroll
                for (int j = 0; j < TP; ++j) {
                    const f32x2 rx = sub2(px2, qx[j]), ry = sub2(py2, qy[j]), rz = sub2(pz2, qz[j]);
                    const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, mul2(rx, rx)));
                    float r2a, r2b;
                    unpack2(r2, r2a, r2b);
                    const float ua = rsqrt_sfu(r2a), ub = rsqrt_sfu(r2b);           // +inf on a coincident pair
                    const f32x2 u = pack2(ua, ub);
                    const f32x2 uc = pack2(fminf(ua, kRsqrtClamp), fminf(ub, kRsqrtClamp));
                    const f32x2 den = fma2(mul2(r2, r2), u, eps2);                  // |R|^3 + eps; 0*inf = NaN if R = 0
                    float da, db;
                    unpack2(den, da, db);
                    const f32x2 w = pack2(fmaxf(rcp_sfu(da), 0.f), fmaxf(rcp_sfu(db), 0.f));   // NaN -> 0
                    const f32x2 d = fma2(mz2, rz, fma2(my2, ry, mul2(mx2, rx)));    // -3 n.R (exactly 0 when R = 0)
                    const f32x2 t = mul2(d, mul2(uc, uc));                          // -3 t; finite: the clamped u
                    const f32x2 vx = fma2(t, rx, nx2), vy = fma2(t, ry, ny2), vz = fma2(t, rz, nz2);
                    ax[j] = fma2(w, vx, ax[j]); ay[j] = fma2(w, vy, ay[j]); az[j] = fma2(w, vz, az[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < TP; ++j) {
                tx[j] = add2(tx[j], ax[j]);
                ty[j] = add2(ty[j], ay[j]);
                tz[j] = add2(tz[j], az[j]);
            }
        }
    }

    // this lane's partial rows: store on the lane's first chunk, accumulate afterwards (launches of one
    // lane are serialised by its stream, so the read-modify-write is race-free)
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float ex[2], ey[2], ez[2];
        unpack2(tx[j], ex[0], ex[1]);
        unpack2(ty[j], ey[0], ey[1]);
        unpack2(tz[j], ez[0], ez[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t m = tile_base + (int64_t)(2 * j + h) * kCBlock + tid;
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

// ---- per-thread host resources (pinned staging + lane streams), released when the thread exits
struct LaneSet {
    int device = -1;
    cudaStream_t lane[kCMaxLanes] = {};
    cudaEvent_t done[kCMaxLanes] = {};
    cudaEvent_t fork = nullptr;
    float* pinned = nullptr;
    size_t pinned_bytes = 0;
    bool ready = false;

    void release() {
        if (ready) {
            for (int i = 0; i < kCMaxLanes; ++i) {
                if (lane[i]) cudaStreamDestroy(lane[i]);
                if (done[i]) cudaEventDestroy(done[i]);
            }
            if (fork) cudaEventDestroy(fork);
        }
        if (pinned) cudaFreeHost(pinned);
        for (int i = 0; i < kCMaxLanes; ++i) { lane[i] = nullptr; done[i] = nullptr; }
        fork = nullptr;
        pinned = nullptr;
        pinned_bytes = 0;
        ready = false;
        device = -1;
    }
    ~LaneSet() { release(); }

    cudaError_t prepare(int dev) {
        if (ready && device == dev) return cudaSuccess;
        release();
        device = dev;
        cudaError_t e;
        for (int i = 0; i < kCMaxLanes; ++i) {
            if ((e = cudaStreamCreateWithFlags(&lane[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming)) != cudaSuccess) return e;
        }
        if ((e = cudaEventCreateWithFlags(&fork, cudaEventDisableTiming)) != cudaSuccess) return e;
        ready = true;
        return cudaSuccess;
    }
    cudaError_t staging(size_t bytes) {
        if (bytes <= pinned_bytes) return cudaSuccess;
        if (pinned) cudaFreeHost(pinned);
        pinned = nullptr;
        pinned_bytes = 0;
        cudaError_t e = cudaMallocHost(reinterpret_cast<void**>(&pinned), bytes);
        if (e == cudaSuccess) pinned_bytes = bytes;
        return e;
    }
};
thread_local LaneSet t_lanes;

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v ? atoi(v) : dflt;
}

inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

}  // namespace

// Calls worth the one stream synchronisation: enough sources to amortise the per-launch target traffic
// and enough targets to fill the machine with a few lanes.  DNP_CONST_PATH=0 disables, =2 lowers the bar
// (tests).
bool use_const_path(int64_t S, int64_t M, int lattice_n) {
    const int mode = env_int("DNP_CONST_PATH", 1);
    if (mode == 0 || lattice_n > 0) return false;
    if (mode == 2) return S >= 1 && M >= 1;
    // tools/plan_probe.py: 65 536 x 65 536 runs 1.56e12 pairs/s here vs 1.39e12 in the shared-memory kernel, 32 768 x
    // 100 000 1.54e12 vs 1.36e12; at 16 384 x 65 536 they tie and at M = 32 768 this path loses
    return (S >= 16384 && M >= 131072) || (S >= 32768 && M >= 65536);
}

CPairPlan make_cpair_plan(bool potential, int64_t S, int64_t M, int sm_count) {
    CPairPlan p;
    const int fold = potential ? Blk<true>::kFold : Blk<false>::kFold;
    const int chunk = potential ? Blk<true>::kChunk : Blk<false>::kChunk;
    p.S_pad = round_up(S < 1 ? 1 : S, fold);
    p.n_chunks = (int)((p.S_pad + chunk - 1) / chunk);
    p.m_tiles = (int)((M + (int64_t)kCBlock * 2 * kCTP - 1) / ((int64_t)kCBlock * 2 * kCTP));
    if (p.m_tiles < 1) p.m_tiles = 1;
    // enough lanes for >= ~2.2 waves of CTAs in flight, so that wave tails overlap the next launch
    const int64_t want = (int64_t)sm_count * 3 * 22 / 10;
    int lanes = (int)((want + p.m_tiles - 1) / p.m_tiles);
    if (lanes < 2) lanes = 2;
    if (lanes > kCMaxLanes) lanes = kCMaxLanes;
    const int forced = env_int("DNP_CONST_LANES", 0);
    if (forced > 0) lanes = forced > kCMaxLanes ? kCMaxLanes : forced;
    if (lanes > p.n_chunks) lanes = p.n_chunks;
    p.lanes = lanes;
    p.packed_bytes = (size_t)round_up(p.S_pad * 6 * (int64_t)sizeof(float), 256);
    p.partial_bytes = (size_t)round_up((int64_t)p.lanes * (M < 1 ? 1 : M) * (potential ? 1 : 3) * (int64_t)sizeof(float), 256);
    return p;
}

size_t cpairs_workspace_bytes(int64_t S, int64_t M) {
    const CPairPlan f = make_cpair_plan(false, S, M, sm_count_cached());
    const CPairPlan p = make_cpair_plan(true, S, M, sm_count_cached());
    const size_t a = f.packed_bytes + f.partial_bytes, b = p.packed_bytes + p.partial_bytes;
    return a > b ? a : b;
}

namespace {

// one chunk of dense host rows [x y z nx ny nz] -> the by-value block (null dipoles past `n_have`)
template <bool kPotential>
void fill_block(SrcBlock<kPotential>& blk, const float* rows, int64_t n_have, int n_src) {
    constexpr int ST = Blk<kPotential>::kStride;
    if (kPotential) {
        if (n_have > 0) memcpy(blk.v, rows, (size_t)n_have * 6 * sizeof(float));
    } else {
        for (int64_t i = 0; i < n_have; ++i) {
            const float* r = rows + 6 * i;
            float* o = blk.v + ST * i;
            o[0] = r[0]; o[1] = r[1]; o[2] = r[2];
            o[3] = r[3]; o[4] = r[4]; o[5] = r[5];
            o[6] = r[3] * -3.f; o[7] = r[4] * -3.f; o[8] = r[5] * -3.f;
        }
    }
    if (n_have < n_src) memset(blk.v + n_have * ST, 0, (size_t)(n_src - n_have) * ST * sizeof(float));
}

template <bool kPotential>
int launch_chunks(const CPairPlan& plan, const float* host_rows, int64_t host_valid, const CTargets& T, float eps,
                  float* partial) {
    constexpr int kChunk = Blk<kPotential>::kChunk;
    const int ncomp = kPotential ? 1 : 3;
    SrcBlock<kPotential> blk;
    for (int k = 0; k < plan.n_chunks; ++k) {
        const int l = k % plan.lanes;
        const int64_t s_begin = (int64_t)k * kChunk;
        const int n_src = (int)((plan.S_pad - s_begin < kChunk) ? (plan.S_pad - s_begin) : kChunk);
        int64_t n_have = host_valid - s_begin;
        if (n_have < 0) n_have = 0;
        if (n_have > n_src) n_have = n_src;
        fill_block<kPotential>(blk, host_rows + s_begin * 6, n_have, n_src);
        float* part = partial + (size_t)l * T.M * ncomp;
        if (!kPotential && env_int("DNP_CONST_BANK3", 0))
            cpairs_kernel<kPotential, true><<<plan.m_tiles, kCBlock, 0, t_lanes.lane[l]>>>(blk, n_src, T, eps, part,
                                                                                            k < plan.lanes);
        else
        cpairs_kernel<kPotential><<<plan.m_tiles, kCBlock, 0, t_lanes.lane[l]>>>(blk, n_src, T, eps, part,
                                                                                  k < plan.lanes);
    }
    count_launch(plan.n_chunks);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace

// src_host_dense: optional dense [S,6] HOST copy of the sources (the host entry points have one); when
// NULL the device sources are packed and staged through pinned memory (one stream synchronisation).
int run_cpairs(bool potential, const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx,
               const float* src_host_dense, const float* tgt, int64_t M, int64_t ld_t, float eps,
               const PairEpilogue& ep, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    if (M <= 0) return DNP_OK;
    const CPairPlan plan = make_cpair_plan(potential, S, M, sm_count_cached());
    if (!workspace || workspace_bytes < plan.packed_bytes + plan.partial_bytes) {
        set_error("workspace too small: need %zu bytes, got %zu", plan.packed_bytes + plan.partial_bytes,
                  workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    int dev = 0;
    DNP_CHECK_CUDA(cudaGetDevice(&dev));
    DNP_CHECK_CUDA(t_lanes.prepare(dev));
    float* packed = reinterpret_cast<float*>(workspace);
    float* partial = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + plan.packed_bytes);

    const float* host_rows = src_host_dense;
    int64_t host_valid = S < 0 ? 0 : S;          // rows of host_rows that exist (the rest are null dipoles)
    if (!host_rows) {
        const int threads = 256;
        const unsigned blocks = (unsigned)((plan.S_pad + threads - 1) / threads);
        pack_sources24_kernel<<<blocks, threads, 0, stream>>>(src, S < 0 ? 0 : S, ld_s, src_idx, plan.S_pad, packed);
        count_launch(1);
        const size_t bytes = (size_t)plan.S_pad * 6 * sizeof(float);
        DNP_CHECK_CUDA(t_lanes.staging(bytes));
        DNP_CHECK_CUDA(cudaMemcpyAsync(t_lanes.pinned, packed, bytes, cudaMemcpyDeviceToHost, stream));
        DNP_CHECK_CUDA(cudaStreamSynchronize(stream));     // the sources travel by value from here on
        host_rows = t_lanes.pinned;
        host_valid = plan.S_pad;
    }

    const CTargets T{tgt, M, ld_t, ep.tgt_idx, ep.tgt_skip};
    timer_mark(stream);
    DNP_CHECK_CUDA(cudaEventRecord(t_lanes.fork, stream));
    for (int l = 0; l < plan.lanes; ++l) DNP_CHECK_CUDA(cudaStreamWaitEvent(t_lanes.lane[l], t_lanes.fork, 0));
    const int rc = potential ? launch_chunks<true>(plan, host_rows, host_valid, T, eps, partial)
                             : launch_chunks<false>(plan, host_rows, host_valid, T, eps, partial);
    for (int l = 0; l < plan.lanes; ++l) {       // always re-join, also after a failed launch
        cudaEventRecord(t_lanes.done[l], t_lanes.lane[l]);
        cudaStreamWaitEvent(stream, t_lanes.done[l], 0);
    }
    timer_mark(stream);
    if (rc != DNP_OK) return rc;
    return run_split_reduce(potential, partial, plan.lanes, M, ep, stream);
}

}  // namespace dnp
