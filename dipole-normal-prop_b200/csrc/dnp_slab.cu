// dnp_slab.cu -- the greedy patch-propagation loop WITHOUT its chain of dependent field launches ("slab" schedule).
//
// field_utils.strongest_field_propagation{,_reps} (field_utils.py:207-282, :286-348) visit P patches one after the
// other; every visit adds the field of the chosen patch to the targets and the next winner is the arg-max of
// I_p = sum_{t in p} E_t . n_t over the unvisited patches.  The field is linear in the source normals and exactly
// odd in them (a flipped patch's field is the negated field, bit for bit), so the visit-dependent part of the loop
// is only WHICH sign a patch's field enters with and WHEN:
//
//   A  slabs[s][t] = field of patch s (normals as they are on entry) at every target t        (P x T x 12 bytes)
//      -- P independent whole-tile launches of the constant-window kernel rotating over the lanes
//      (run_upairs_patch_slabs, dnp_upairs.cu): the all-pairs work of the whole loop at the rate of ONE large call
//      instead of P dependent one-wave calls;
//   B  W[s][p] = sum_{t in p} slabs[s][t] . n_t                                              (one pass over the slabs)
//   C  the greedy loop itself on the P x P matrix: I += sigma_s W[s][:], first abs-arg-max, sigma = sign -- ONE CTA,
//      P steps of a few microseconds, order / flip decisions as the reference takes them (:244-252, :316-323);
//   D  E_t = sum over the visit order of sigma_s slabs[s][t] (skipping t's own patch; when not diffuse only the
//      visits before t's own patch, :256-265 / :326-335) -- the additions happen in the reference's order -- and the
//      flips are committed to the cloud.
//
// HBM: the slabs are written once (A) and read twice (B, D): 3 x P x T x 12 bytes, i.e. 19 GB for the 1M-point /
// 1000-patch / 524 288-representative case = 3 ms of HBM time against ~150 ms of FP32 work in A.  Sized for the
// 180 GB of a B200: the schedule is taken when the slabs fit DNP_SLAB_MAX_MB (default 64 GB: a third of the device).
// Preconditions as for the persistent kernel: disjoint patches that are exactly the target rows, no "rest" row among
// the patch rows (one 12-byte read per call; not on a capturing stream).
#include <stdlib.h>

#include <vector>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"
#include "dipole.h"

namespace dnp {

namespace {

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

constexpr int kInterBlock = 256;          // threads of slab_interactions_kernel (8 warps)
constexpr int kInterPerBlock = 64;        // source patches per block of that kernel

// target m (CSR position) -> its patch, and its normal as it is on entry (dense copy: B reads it P times)
__global__ void slab_targets_kernel(const float* __restrict__ pts, int64_t ld, const int64_t* __restrict__ perm,
                                    const int64_t* __restrict__ offsets, int P, int64_t T,
                                    int32_t* __restrict__ patch_of, float* __restrict__ nrm) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= T) return;
    int lo = 0, hi = P;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (offsets[mid] <= m) lo = mid; else hi = mid;
    }
    patch_of[m] = lo;
    const float* q = pts + perm[m] * ld + 3;
    nrm[3 * m] = __ldg(q); nrm[3 * m + 1] = __ldg(q + 1); nrm[3 * m + 2] = __ldg(q + 2);
}

// W[s][p] = sum over the rows of patch p of slabs[s][row] . n_row.  Block (p, group of 64 source patches): a warp
// takes one s at a time, its lanes stride over the 3 n_p floats of the patch's slab rows (coalesced), four
// independent partial sums per lane, fixed shuffle tree: deterministic.
__global__ void __launch_bounds__(kInterBlock)
slab_interactions_kernel(const float* __restrict__ slabs, int64_t slab_stride, const float* __restrict__ nrm,
                         const int64_t* __restrict__ offsets, int P, const int32_t* __restrict__ slot, int upper,
                         float* __restrict__ W) {
    const int p = blockIdx.x;
    const int64_t o = offsets[p] * 3, n3 = (offsets[p + 1] - offsets[p]) * 3;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* __restrict__ nn = nrm + o;
    const int s_end = min(P, (int)(blockIdx.y + 1) * kInterPerBlock);
    for (int s = blockIdx.y * kInterPerBlock + warp; s < s_end; s += kInterBlock / 32) {
        const int sl = slot ? slot[s] : s;
        if (sl < 0) continue;                                           // slab of another rank
        if (upper && s >= p) continue;                                  // symmetric form: only s < p was computed
        const float* __restrict__ e = slabs + (int64_t)sl * slab_stride + o;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        int64_t i = lane;
        for (; i + 96 < n3; i += 128) {
            a0 = fmaf(e[i], nn[i], a0);
            a1 = fmaf(e[i + 32], nn[i + 32], a1);
            a2 = fmaf(e[i + 64], nn[i + 64], a2);
            a3 = fmaf(e[i + 96], nn[i + 96], a3);
        }
        for (; i < n3; i += 32) a0 = fmaf(e[i], nn[i], a0);
        const float v = warp_sum((a0 + a1) + (a2 + a3));
        if (lane == 0) W[(int64_t)s * P + p] = v;
    }
}

// symmetric form: W[s][p] = W[p][s] for s > p (the interaction energy of two dipole sets does not depend on which one
// is called the source: n_j . E_i(q_j) = w [n_i . n_j - 3 (n_i . R)(n_j . R) / r^2] is symmetric in i and j)
__global__ void slab_symmetrize_kernel(float* __restrict__ W, int P) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x, s = blockIdx.y;
    if (p >= P) return;
    if (s > p) W[(int64_t)s * P + p] = W[(int64_t)p * P + s];
    else if (s == p) W[(int64_t)s * P + p] = 0.f;
}

// The greedy loop on the matrix (one CTA).  inter[p] = interaction of patch p with the field of the visited patches
// (float32 sums in visit order); per step: first abs-arg-max over the unvisited patches (torch's abs().argmax()),
// sigma = -1 if the winner's interaction is negative (:250,322), then its row enters every unvisited patch's sum.
__global__ void __launch_bounds__(1024)
slab_greedy_kernel(const float* __restrict__ W, int P, int start, float* inter, uint8_t* active, float* __restrict__ sigma, int32_t* __restrict__ rank,
                   int32_t* __restrict__ order, float* __restrict__ value) {
    __shared__ unsigned long long warp_best[32];
    __shared__ int sh_idx;
    __shared__ float sh_sig;
    for (int p = threadIdx.x; p < P; p += blockDim.x) { inter[p] = 0.f; active[p] = 1; sigma[p] = 1.f; rank[p] = P; }
    __syncthreads();
    for (int step = 0; step < P; ++step) {
        unsigned long long best = 0ull;
        if (step > 0) best = block_first_absmax(inter, active, P, warp_best);
        if (threadIdx.x == 0) {
            int idx = start;
            float val = 0.f;
            if (step > 0) {
                idx = best != 0ull ? (int)(0xffffffffu - (uint32_t)best) : -1;
                val = idx >= 0 ? inter[idx] : 0.f;
            }
            sh_idx = idx;
            sh_sig = val < 0.f ? -1.f : 1.f;
            order[step] = idx;
            value[step] = val;
            if (idx >= 0) { active[idx] = 0; sigma[idx] = sh_sig; rank[idx] = step; }
        }
        __syncthreads();
        const int idx = sh_idx;
        if (idx < 0) break;
        const float sig = sh_sig;
        const float* __restrict__ row = W + (int64_t)idx * P;
        for (int p = threadIdx.x; p < P; p += blockDim.x)
            if (active[p]) inter[p] += sig * row[p];
        __syncthreads();
    }
}

// E of target m = the slabs of the visited patches added in visit order with their signs (own patch skipped; when
// not diffuse only the visits before the target's own patch), then the target's flip is committed to the cloud.
// Multi-GPU (slot != null, dense != 0): only the slabs this rank holds (slot[s] >= 0) are added, into a dense [T, 3]
// buffer that is summed over the ranks afterwards.
__global__ void __launch_bounds__(256)
slab_accumulate_kernel(const float* __restrict__ slabs, int64_t slab_stride, int P, int64_t T,
                       const int32_t* __restrict__ patch_of, const int32_t* __restrict__ order,
                       const float* __restrict__ sigma, const int32_t* __restrict__ rank,
                       const int32_t* __restrict__ slot, int diffuse, int no_field, const int64_t* __restrict__ perm,
                       float* __restrict__ E, int64_t ld_e, int dense, float* __restrict__ pts, int64_t ld) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= T) return;
    const int own = patch_of[m];
    const int k_end = no_field ? 0 : (diffuse ? P : min(P, rank[own]));   // symmetric form: half the slabs do not exist
    const float* __restrict__ col = slabs + 3 * m;
    float ex = 0.f, ey = 0.f, ez = 0.f;
#pragma unroll 4
    for (int k = 0; k < k_end; ++k) {
        const int s = order[k];
        if (s < 0) break;
        const int sl = slot ? slot[s] : s;
        if (s == own || sl < 0) continue;
        const float sg = sigma[s];
        const float* __restrict__ v = col + (int64_t)sl * slab_stride;
        ex += sg * v[0]; ey += sg * v[1]; ez += sg * v[2];
    }
    const int64_t r = perm[m];
    float* o = dense ? E + 3 * m : E + r * ld_e;
    o[0] = ex; o[1] = ey; o[2] = ez;
    if (sigma[own] < 0.f) {
        float* q = pts + r * ld + 3;
        q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2];
    }
}

// multi-GPU: the summed field of this rank's own patches goes to the caller's E (rows of other ranks stay untouched)
__global__ void slab_scatter_owned_kernel(const float* __restrict__ dense, int64_t T, const int32_t* __restrict__ patch_of,
                                          const int32_t* __restrict__ slot, const int64_t* __restrict__ perm,
                                          float* __restrict__ E, int64_t ld_e) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= T || slot[patch_of[m]] < 0) return;
    float* o = E + perm[m] * ld_e;
    o[0] = dense[3 * m]; o[1] = dense[3 * m + 1]; o[2] = dense[3 * m + 2];
}

// the rows that follow a patch ("rest", :250-252) take its flip
__global__ void slab_rest_flip_kernel(const int64_t* __restrict__ rest_perm, const int64_t* __restrict__ rest_off,
                                      int P, int64_t R, const float* __restrict__ sigma, float* __restrict__ pts,
                                      int64_t ld) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const int64_t pos = rest_off[0] + i;
    int lo = 0, hi = P;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (rest_off[mid] <= pos) lo = mid; else hi = mid;
    }
    if (sigma[lo] < 0.f) {
        float* q = pts + rest_perm[pos] * ld + 3;
        q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2];
    }
}

struct SlabLayout {
    size_t flags, counts, pad_off, rest_off, patch_of, nrm, W, inter, active, sigma, rank, order, value, slot, dense,
        packed, slabs, total;
    int64_t slab_stride;
};

SlabLayout slab_layout(int64_t N, int64_t P, int64_t T, int64_t n_slabs) {
    SlabLayout L{};
    size_t off = 0;
    L.flags = off; off += up256((size_t)N * 4);
    L.counts = off; off += 256;
    L.pad_off = off; off += up256((size_t)(P + 1) * 8);
    L.rest_off = off; off += up256((size_t)(P + 1) * 8);
    L.patch_of = off; off += up256((size_t)T * 4);
    L.nrm = off; off += up256((size_t)T * 12);
    L.W = off; off += up256((size_t)P * P * 4);
    L.inter = off; off += up256((size_t)P * 4);
    L.active = off; off += up256((size_t)P);
    L.sigma = off; off += up256((size_t)P * 4);
    L.rank = off; off += up256((size_t)P * 4);
    L.order = off; off += up256((size_t)P * 4);
    L.value = off; off += up256((size_t)P * 4);
    L.slot = off; off += up256((size_t)P * 4);
    L.dense = off; off += up256((size_t)T * 12);
    L.packed = off; off += upairs_slab_packed_bytes(T + (upairs_slab_quantum() - 1) * P);
    L.slab_stride = (3 * T + 63) / 64 * 64;
    L.slabs = off; off += up256((size_t)(n_slabs > 0 ? n_slabs : 1) * L.slab_stride * 4);
    L.total = off;
    return L;
}

}  // namespace

// Whether a call of this shape is planned for the slab schedule (the workspace query and the driver must agree).
// DNP_SLAB_MAX_MB: budget for the slabs (default 65536); DNP_SLAB=0 turns the schedule off.
bool slab_schedule_fits(int64_t P, int64_t T, int64_t max_patch, int world) {
    const char* off = getenv("DNP_SLAB");
    if (off && off[0] == '0') return false;
    if (P < 1 || P > 8192 || T < 1 || max_patch < 1) return false;
    const char* cap_env = getenv("DNP_SLAB_MAX_MB");
    const double cap = (cap_env ? atof(cap_env) : 65536.0) * 1048576.0;
    if (world < 1) world = 1;
    return (double)((P + world - 1) / world) * (double)T * 12.0 <= cap;      // per GPU
}

// ... and whether it is WORTH it when nothing is forced (tools/loop_size_probe2.py, profiles/README.md r02d): against
// the device-driven launches it wins from 60 000 rows on with patches of a few hundred rows (5.3 vs 9.3 ms; 400 000
// rows: 95 vs 153-188 ms) -- three or four per-patch launches in flight against a chain of dependent ones.  Below
// ~224 target tiles the machine (444 resident CTAs) is not full, so few LARGE patches on a small target set stay with
// the device-driven launches, whose all-pairs kernel shapes its grid to the target set; patches of >= 128 rows on
// average keep the per-launch overheads small.  (The persistent kernel is tried first by the driver where IT wins.)
bool slab_schedule_pays(int64_t P, int64_t T, int64_t max_patch, int world) {
    if (!slab_schedule_fits(P, T, max_patch, world)) return false;
    const char* min_rows = getenv("DNP_SLAB_MIN_ROWS");                 // tests: any loop of at least that many rows
    if (min_rows) return T >= atoll(min_rows);
    if (T < 128 * P) return false;
    return T >= 224 * 1024 || (T >= 65536 && max_patch <= 2048);
}

// n_slabs: the slabs this call holds -- P on one GPU, the patches the rank owns in the sharded loop
size_t patch_slab_workspace_bytes(int64_t N, int64_t P, int64_t T, int64_t n_slabs) {
    return slab_layout(N, P, T, n_slabs < 0 ? P : n_slabs).total;
}

// Returns DNP_OK and *taken = 1 when the loop ran here; *taken = 0 when the preconditions do not hold.
int run_patch_loop_slab(float* pts, int64_t N, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                        const int64_t* offsets_host, int64_t P, const int64_t* rest_perm,
                        const int64_t* rest_offsets_host, const int64_t* tgt_idx, int64_t n_tgt, int64_t start_patch,
                        float eps, int diffuse, float* E, int64_t ld_e, int32_t* order_host, float* value_host,
                        void* workspace, size_t workspace_bytes, cudaStream_t stream, const ShardHook* shard,
                        int* taken) {
    *taken = 0;
    {   // the precondition check reads 12 bytes back: not on a capturing stream
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) { cudaGetLastError(); return DNP_OK; }
        if (st != cudaStreamCaptureStatusNone) return DNP_OK;
    }
    const int64_t T = offsets_host[P] - offsets_host[0];
    if (P < 1 || offsets_host[0] != 0 || T < 1) return DNP_OK;
    // single GPU: the patches must be the targets.  Sharded: the cloud is replicated, every rank takes all patch
    // rows as targets whatever slice tgt_idx names (the slabs are split by SOURCE patch instead).
    if (!shard && (tgt_idx ? (tgt_idx != perm || n_tgt != T) : T != N)) return DNP_OK;
    int64_t max_patch = 0;
    for (int64_t p = 0; p < P; ++p) {
        const int64_t n = offsets_host[p + 1] - offsets_host[p];
        if (n > max_patch) max_patch = n;
    }
    if (!slab_schedule_fits(P, T, max_patch, shard ? shard->world : 1)) return DNP_OK;
    // multi-GPU: this rank computes and holds the slabs of the patches it owns, slot[p] = its slab or -1
    std::vector<int32_t> slot_host;
    int64_t n_slabs = P;
    if (shard) {
        std::vector<uint8_t> owned((size_t)P);
        DNP_CHECK_CUDA(cudaMemcpyAsync(owned.data(), shard->owned, (size_t)P, cudaMemcpyDeviceToHost, stream));
        DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
        slot_host.resize((size_t)P);
        n_slabs = 0;
        for (int64_t p = 0; p < P; ++p) slot_host[p] = owned[p] ? (int32_t)n_slabs++ : -1;
    }
    const SlabLayout L = slab_layout(N, P, T, n_slabs);
    if (workspace_bytes < L.total) return DNP_OK;

    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    int* flags = reinterpret_cast<int*>(ws + L.flags);
    int* counts = reinterpret_cast<int*>(ws + L.counts);
    int64_t* pad_off_dev = reinterpret_cast<int64_t*>(ws + L.pad_off);
    int64_t* rest_off_dev = reinterpret_cast<int64_t*>(ws + L.rest_off);
    int32_t* patch_of = reinterpret_cast<int32_t*>(ws + L.patch_of);
    float* nrm = reinterpret_cast<float*>(ws + L.nrm);
    float* W = reinterpret_cast<float*>(ws + L.W);
    float* inter = reinterpret_cast<float*>(ws + L.inter);
    uint8_t* active = ws + L.active;
    float* sigma = reinterpret_cast<float*>(ws + L.sigma);
    int32_t* rank = reinterpret_cast<int32_t*>(ws + L.rank);
    int32_t* order_dev = reinterpret_cast<int32_t*>(ws + L.order);
    float* value_dev = reinterpret_cast<float*>(ws + L.value);
    int32_t* slot_dev = reinterpret_cast<int32_t*>(ws + L.slot);
    float* dense = reinterpret_cast<float*>(ws + L.dense);
    float* packed = reinterpret_cast<float*>(ws + L.packed);
    float* slabs = reinterpret_cast<float*>(ws + L.slabs);

    const int64_t R = rest_perm ? rest_offsets_host[P] - rest_offsets_host[0] : 0;
    int ok = 0;
    const int rc_cover = check_patch_cover(perm, T, N, rest_perm ? rest_perm + rest_offsets_host[0] : nullptr, R,
                                           flags, counts, stream, &ok);
    if (rc_cover != DNP_OK) return rc_cover;
    if (!ok) return DNP_OK;

    // patch p's window rows start at a multiple of the kernel's quantum (12 rows: 16-byte aligned, whole iterations)
    const int64_t quantum = upairs_slab_quantum();
    std::vector<int64_t> pad_off((size_t)P + 1);
    pad_off[0] = 0;
    for (int64_t p = 0; p < P; ++p)
        pad_off[p + 1] = pad_off[p] + (offsets_host[p + 1] - offsets_host[p] + quantum - 1) / quantum * quantum;
    DNP_CHECK_CUDA(cudaMemcpyAsync(pad_off_dev, pad_off.data(), (size_t)(P + 1) * 8, cudaMemcpyHostToDevice, stream));
    if (R > 0)
        DNP_CHECK_CUDA(cudaMemcpyAsync(rest_off_dev, rest_offsets_host, (size_t)(P + 1) * 8, cudaMemcpyHostToDevice,
                                       stream));
    if (shard)
        DNP_CHECK_CUDA(cudaMemcpyAsync(slot_dev, slot_host.data(), (size_t)P * 4, cudaMemcpyHostToDevice, stream));
    const int32_t* slot = shard ? slot_dev : nullptr;

    // not diffuse: nobody reads the running field afterwards (field_utils.py:264-276 recomputes what it needs), the
    // decisions only need W, and W is symmetric -- every patch pair once: patch s acts on the targets of the patches
    // AFTER it only (half the pairs), W[p][s] is copied from W[s][p], and E is returned as zeros (sharded: E is not
    // touched; round-robin ownership spreads the long and the short patches over the ranks).
    // DNP_SLAB_SYMMETRIC=0: the full form (E = the running field, as the other schedules leave it).
    const char* sym_env = getenv("DNP_SLAB_SYMMETRIC");
    const bool symmetric = !diffuse && !(sym_env && sym_env[0] == '0');
    std::vector<int64_t> first_target;
    if (symmetric) first_target.assign(offsets_host + 1, offsets_host + P + 1);

    slab_targets_kernel<<<(unsigned)((T + 255) / 256), 256, 0, stream>>>(pts, ld, perm, offsets_dev, (int)P, T, patch_of, nrm);
    count_launch(1);
    // A
    const int rc = run_upairs_patch_slabs(pts, ld, perm, offsets_dev, offsets_host, pad_off_dev, pad_off.data(), P,
                                          shard ? slot_host.data() : nullptr, symmetric ? first_target.data() : nullptr, perm, T,
                                          eps, packed, slabs,
                                          L.slab_stride, stream);
    if (rc != DNP_OK) return rc;
    // (the host vectors above are read by pageable-memory copies, which return once the data is staged)
    // B
    if (shard) DNP_CHECK_CUDA(cudaMemsetAsync(W, 0, (size_t)P * P * 4, stream));     // rows of other ranks: exact zeros
    {
        const dim3 grid((unsigned)P, (unsigned)((P + kInterPerBlock - 1) / kInterPerBlock));
        slab_interactions_kernel<<<grid, kInterBlock, 0, stream>>>(slabs, L.slab_stride, nrm, offsets_dev, (int)P, slot,
                                                                   symmetric ? 1 : 0, W);
        count_launch(1);
    }
    if (shard) {                                                     // x + 0 is exact: every rank gets the owners' rows
        const int rcc = shard->combine(shard->ctx, W, P * P, stream);
        if (rcc != DNP_OK) return rcc;
    }
    if (symmetric) {                                                 // (sharded: once all owners' rows have arrived)
        slab_symmetrize_kernel<<<dim3((unsigned)((P + 255) / 256), (unsigned)P), 256, 0, stream>>>(W, (int)P);
        count_launch(1);
    }
    // C
    slab_greedy_kernel<<<1, 1024, 0, stream>>>(W, (int)P, (int)start_patch, inter, active, sigma, rank, order_dev, value_dev);
    count_launch(1);
    // D
    slab_accumulate_kernel<<<(unsigned)((T + 255) / 256), 256, 0, stream>>>(
        slabs, L.slab_stride, (int)P, T, patch_of, order_dev, sigma, rank, slot, diffuse, symmetric ? 1 : 0, perm,
        shard ? dense : E, ld_e,
        shard ? 1 : 0, pts, ld);
    count_launch(1);
    if (shard && !symmetric) {     // every rank added the slabs it holds: sum over ranks, keep the rows of the own patches
                     // (the sharded call's contract: E of the other ranks' rows is left as it was)
        const int rcc = shard->combine(shard->ctx, dense, T * 3, stream);
        if (rcc != DNP_OK) return rcc;
        slab_scatter_owned_kernel<<<(unsigned)((T + 255) / 256), 256, 0, stream>>>(dense, T, patch_of, slot, perm, E, ld_e);
        count_launch(1);
    }
    if (R > 0) {
        slab_rest_flip_kernel<<<(unsigned)((R + 255) / 256), 256, 0, stream>>>(rest_perm, rest_off_dev, (int)P, R, sigma, pts, ld);
        count_launch(1);
    }
    *taken = 1;
    if (order_host)
        DNP_CHECK_CUDA(cudaMemcpyAsync(order_host, order_dev, (size_t)P * 4, cudaMemcpyDeviceToHost, stream));
    if (value_host)
        DNP_CHECK_CUDA(cudaMemcpyAsync(value_host, value_dev, (size_t)P * 4, cudaMemcpyDeviceToHost, stream));
    if (order_host || value_host) DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
