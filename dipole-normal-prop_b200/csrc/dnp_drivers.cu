// dnp_drivers.cu -- the greedy patch-propagation loop as ONE native call.
//
// field_utils.strongest_field_propagation (field_utils.py:286-348) and strongest_field_propagation_reps
// (:207-282) run P - 1 dependent steps; the reference spends each of them in Python: a list comprehension over
// the remaining patches (:244,316), P host synchronisations, boolean-mask indexing.  The first version of this
// library kept the loop in Python with one call per kernel (about ten launches, several torch index ops and one
// 8-byte read per step, ~100 us of host work).  Here the whole loop lives behind the C ABI: per step
//   commit   flip the chosen patch (and the rows that follow it) if its interaction was negative, mark its rows
//   field    E += field of the chosen patch on the target rows (the all-pairs kernel, sources through src_idx)
//   score    I_p = sum_{t in p} E_t . n_t for the remaining patches + first abs-arg-max (K4)
// Two schedules of the same kernels:
//   * device-driven (patches below 16 384 rows, i.e. nearly always): the winner of a step
//     stays on the device -- the arg-max kernel advances cur = {patch, flip, previous patch}, the commit / pack /
//     all-pairs kernels read their row range from it (launches are shaped for the LARGEST patch; the all-pairs
//     kernel walks only the chosen patch's source tiles).  No host read inside the loop: all P steps are
//     enqueued at once and the GPU runs them back to back.
//   * host-driven (a patch of >= 16 384 rows, served by the constant-window kernel with host-planned launches): one
//     8-byte device->host read per step.
// Decisions are the kernels' (identical to the per-call path); nothing here does arithmetic on the host.
#include <stdlib.h>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"
#include "dipole.h"

namespace dnp {
namespace {

// rows[0..n): negate normals if flip, set settled / scratch; rest[0..n_rest): negate normals if flip
__global__ void patch_commit_kernel(float* __restrict__ normals, int64_t ld, const int64_t* __restrict__ rows,
                                    int64_t n, const int64_t* __restrict__ rest, int64_t n_rest, int flip,
                                    uint8_t* __restrict__ settled, uint8_t* __restrict__ scratch,
                                    uint8_t* __restrict__ active, int64_t patch) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n + n_rest; i += stride) {
        const bool own = i < n;
        const int64_t r = own ? rows[i] : rest[i - n];
        if (flip) {
            float* q = normals + r * ld;
            q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2];
        }
        if (own) {
            settled[r] = 1;
            if (scratch) scratch[r] = 1;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) active[patch] = 0;
}

__global__ void mask_clear_kernel(uint8_t* __restrict__ mask, const int64_t* __restrict__ rows, int64_t n) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) mask[rows[i]] = 0;
}

// ---- device-driven schedule: the same two kernels with the row ranges read from cur = {patch, flip, previous}
__global__ void patch_commit_selected_kernel(float* __restrict__ normals, int64_t ld, const int64_t* __restrict__ perm,
                                             const int64_t* __restrict__ offsets,
                                             const int64_t* __restrict__ rest_perm,
                                             const int64_t* __restrict__ rest_offsets, const int32_t* __restrict__ cur,
                                             uint8_t* __restrict__ settled, uint8_t* __restrict__ scratch,
                                             uint8_t* __restrict__ active) {
    const int patch = cur[0], flip = cur[1];
    if (patch < 0) return;
    const int64_t o = offsets[patch], n = offsets[patch + 1] - o;
    const int64_t ro = rest_perm ? rest_offsets[patch] : 0, rn = rest_perm ? rest_offsets[patch + 1] - ro : 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n + rn; i += stride) {
        const bool own = i < n;
        const int64_t r = own ? perm[o + i] : rest_perm[ro + i - n];
        if (flip) {
            float* q = normals + r * ld;
            q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2];
        }
        if (own) {
            settled[r] = 1;
            if (scratch) scratch[r] = 1;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) active[patch] = 0;
}

__global__ void mask_clear_selected_kernel(uint8_t* __restrict__ mask, const int64_t* __restrict__ perm,
                                           const int64_t* __restrict__ offsets, const int32_t* __restrict__ cur) {
    const int prev = cur[2];
    if (prev < 0) return;
    const int64_t o = offsets[prev], n = offsets[prev + 1] - o;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) mask[perm[o + i]] = 0;
}

// ---- the device-driven step head as ONE single-CTA kernel (patches + rest rows up to kFusedRows): first
// abs-arg-max over the interactions of the previous step -> cur, un-mark the previous patch (diffuse), flip + mark
// the winner's rows, and write them in the all-pairs kernel's packed source layout (null dipoles up to pad_rows).
// Four dependent launches (arg-max, un-mark, commit, pack) become one; at fandisk size a launch costs as much as
// the field kernel itself.
constexpr int64_t kFusedRows = 32768;

__global__ void __launch_bounds__(1024)
patch_step_kernel(const float* __restrict__ inter, int64_t P, int do_argmax, int32_t start, int step,
                  int32_t* __restrict__ cur, int32_t* __restrict__ order, float* __restrict__ value,
                  float* __restrict__ pts, int64_t ld, const int64_t* __restrict__ perm,
                  const int64_t* __restrict__ offsets, const int64_t* __restrict__ rest_perm,
                  const int64_t* __restrict__ rest_offsets, uint8_t* __restrict__ settled,
                  uint8_t* __restrict__ scratch, uint8_t* __restrict__ active, float4* __restrict__ packed,
                  int64_t pad_rows) {
    __shared__ unsigned long long warp_best[32];
    __shared__ int32_t sh[3];
    if (do_argmax) {
        const unsigned long long best = block_first_absmax(inter, active, P, warp_best);
        if (threadIdx.x == 0) {
            int32_t idx = -1;
            float val = 0.f;
            if (best != 0ull) {
                idx = (int32_t)(0xffffffffu - (uint32_t)best);
                val = inter[idx];
            }
            sh[2] = cur[0];
            sh[0] = idx;
            sh[1] = val < 0.f;                                                    // field_utils.py:250,322
            if (value) value[step] = val;
        }
    } else if (threadIdx.x == 0) {
        sh[0] = start; sh[1] = 0; sh[2] = -1;
        if (value) value[step] = 0.f;
    }
    __syncthreads();
    const int patch = sh[0], flip = sh[1], prev = sh[2];
    if (threadIdx.x == 0) {
        cur[0] = patch; cur[1] = flip; cur[2] = prev;
        if (order) order[step] = patch;
    }
    if (scratch && prev >= 0) {                                                   // :256-261: only the source patch is skipped
        const int64_t po = offsets[prev], pn = offsets[prev + 1] - po;
        for (int64_t i = threadIdx.x; i < pn; i += blockDim.x) scratch[perm[po + i]] = 0;
        __syncthreads();
    }
    const int64_t o = patch < 0 ? 0 : offsets[patch], n = patch < 0 ? 0 : offsets[patch + 1] - o;
    const int64_t ro = (rest_perm && patch >= 0) ? rest_offsets[patch] : 0;
    const int64_t rn = (rest_perm && patch >= 0) ? rest_offsets[patch + 1] - ro : 0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const int64_t r = perm[o + i];
        float* q = pts + r * ld;
        float v[6];
#pragma unroll
        for (int c = 0; c < 6; ++c) v[c] = q[c];
        if (flip) {
            v[3] = -v[3]; v[4] = -v[4]; v[5] = -v[5];
            q[3] = v[3]; q[4] = v[4]; q[5] = v[5];
        }
        settled[r] = 1;
        if (scratch) scratch[r] = 1;
        pack_source_row(v, packed + 3 * i);
    }
    if (flip)
        for (int64_t i = threadIdx.x; i < rn; i += blockDim.x) {
            float* q = pts + rest_perm[ro + i] * ld + 3;
            q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2];
        }
    const float zero[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int64_t i = n + threadIdx.x; i < pad_rows; i += blockDim.x) pack_source_row(zero, packed + 3 * i);
    if (threadIdx.x == 0 && patch >= 0) active[patch] = 0;
}

__global__ void cur_init_kernel(int32_t* cur, int32_t start, int32_t* order, float* value) {
    cur[0] = start; cur[1] = 0; cur[2] = -1; cur[3] = 0;
    if (order) order[0] = start;
    if (value) value[0] = 0.f;
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

// DNP_PATCH_LOOP=slab forces the slab schedule wherever it fits; unforced it is taken where it pays
bool slab_wanted(int64_t P, int64_t T, int64_t max_patch, int world = 1) {
    const char* forced = getenv("DNP_PATCH_LOOP");
    if (forced) return forced[0] == 's' && slab_schedule_fits(P, T, max_patch, world);
    return slab_schedule_pays(P, T, max_patch, world);
}

struct DriverLayout {
    size_t settled, scratch, active, inter, best, cur, rest_off, order, value, loop, loop_bytes, pairs, pairs_bytes, total;
};

DriverLayout driver_layout(int64_t N, int64_t P, int64_t max_patch, int64_t M, bool with_slabs = true) {
    DriverLayout L{};
    size_t off = 0;
    L.settled = off; off += up256((size_t)N);
    L.scratch = off; off += up256((size_t)N);
    L.active = off; off += up256((size_t)P);
    L.inter = off; off += up256((size_t)P * sizeof(float));
    L.best = off; off += 256;
    L.cur = off; off += 256;
    L.rest_off = off; off += up256((size_t)(P + 1) * sizeof(int64_t));
    L.order = off; off += up256((size_t)P * sizeof(int32_t));
    L.value = off; off += up256((size_t)P * sizeof(float));
    L.loop_bytes = patch_loop_workspace_bytes(N, P);
    L.loop = off; off += up256(L.loop_bytes);
    L.pairs_bytes = pairs_workspace_bytes_upper(max_patch, M);
    L.pairs = off; off += up256(L.pairs_bytes);
    L.total = off;
    // the slab schedule (dnp_slab.cu) runs INSTEAD of the other two and lays its buffers over theirs
    if (with_slabs && slab_wanted(P, M, max_patch)) {
        const size_t slab_total = L.loop + patch_slab_workspace_bytes(N, P, M);
        if (slab_total > L.total) L.total = slab_total;
    }
    return L;
}

}  // namespace

size_t patch_propagate_workspace_bytes(int64_t N, int64_t P, int64_t max_patch, int64_t n_tgt) {
    if (N < 0 || P < 0 || max_patch < 0) return 0;
    return driver_layout(N, P, max_patch, n_tgt > 0 ? n_tgt : N).total;
}

size_t patch_propagate_sharded_workspace_bytes(int64_t N, int64_t P, int64_t max_patch, int64_t n_tgt,
                                               int64_t patch_rows, int64_t n_owned, int world) {
    if (N < 0 || P < 0 || max_patch < 0 || n_tgt < 0 || patch_rows < 0 || n_owned < 0) return 0;
    // the per-step protocol's buffers for this rank's rows, and -- where the slab schedule applies to the loop over
    // ALL patch rows -- its buffers with one slab per OWNED patch, laid over the same area
    const DriverLayout L = driver_layout(N, P, max_patch, n_tgt > 0 ? n_tgt : 1, false);
    size_t total = L.total;
    if (slab_wanted(P, patch_rows, max_patch, world)) {
        const size_t with_slabs = L.loop + patch_slab_workspace_bytes(N, P, patch_rows, n_owned);
        if (with_slabs > total) total = with_slabs;
    }
    return total;
}

int run_patch_propagate(float* pts, int64_t N, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                        const int64_t* offsets_host, int64_t P, const int64_t* rest_perm,
                        const int64_t* rest_offsets_host, const int64_t* tgt_idx, int64_t n_tgt, int64_t start_patch,
                        float eps, int diffuse, float* E, int64_t ld_e, int32_t* order_host, float* value_host,
                        void* workspace, size_t workspace_bytes, cudaStream_t stream, const ShardHook* shard) {
    if (P <= 0) return DNP_OK;
    if (start_patch < 0 || start_patch >= P) {
        set_error("start patch %lld out of range [0, %lld)", (long long)start_patch, (long long)P);
        return DNP_ERR_INVALID;
    }
    int64_t max_patch = 0;
    for (int64_t p = 0; p < P; ++p) {
        const int64_t n = offsets_host[p + 1] - offsets_host[p];
        if (n < 0) {
            set_error("offsets must be non-decreasing");
            return DNP_ERR_INVALID;
        }
        if (n > max_patch) max_patch = n;
    }
    const int64_t M = tgt_idx ? n_tgt : N;
    // (sharded: M counts this rank's rows only; the slab schedule sizes its own buffers from all patch rows below)
    const DriverLayout L = driver_layout(N, P, max_patch, M, shard == nullptr);
    if (!workspace || workspace_bytes < L.total) {
        set_error("workspace too small: need %zu bytes (dnp_patch_propagate_workspace_bytes), got %zu", L.total,
                  workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    uint8_t* settled = ws + L.settled;
    uint8_t* scratch = ws + L.scratch;
    uint8_t* active = ws + L.active;
    float* inter = reinterpret_cast<float*>(ws + L.inter);
    int32_t* best_idx = reinterpret_cast<int32_t*>(ws + L.best);
    float* best_val = reinterpret_cast<float*>(ws + L.best + 4);
    void* pairs_ws = ws + L.pairs;
    float* normals = pts + 3;

    DNP_CHECK_CUDA(cudaMemsetAsync(settled, 0, (size_t)N, stream));
    if (diffuse) DNP_CHECK_CUDA(cudaMemsetAsync(scratch, 0, (size_t)N, stream));
    DNP_CHECK_CUDA(cudaMemsetAsync(active, 1, (size_t)P, stream));

    PairEpilogue ep;
    ep.out = E;
    ep.ld_out = ld_e;
    ep.tgt_idx = tgt_idx;
    ep.tgt_skip = diffuse ? scratch : settled;       // diffuse: only the source patch itself is skipped (:256-261)

    // DNP_PATCH_LOOP=persistent|device|host overrides the schedule (tests run all of them and compare)
    const char* forced = getenv("DNP_PATCH_LOOP");
    // (sharded over ranks: the targets are this rank's patches only, so the one-kernel loop does not apply)
    // loops of >= 98 304 rows: the slab schedule ahead of the persistent kernel (tools/loop_size_probe2.py, persistent
    // vs slab: diffuse 150 000 rows 22.0 vs 15.8 ms, 300 000 rows 80.6 vs 53.3 ms; not diffuse -- where the slab
    // schedule computes every patch pair once -- 100 000 rows 6.3 vs 5.9 ms, 300 000 rows 47.1 vs 27.0 ms; below
    // that the persistent kernel wins: 60 000 rows 2.6 vs 3.2 ms)
    const int64_t T_rows = offsets_host[P] - offsets_host[0];
    const int world = shard ? shard->world : 1;
    const bool slab_first = !forced && T_rows >= 98304 && slab_wanted(P, T_rows, max_patch, world);
    if (!shard && !slab_first && (!forced || forced[0] == 'p')) {
        // small cloud, disjoint patches covering the targets: the whole loop in one cooperative launch (dnp_loop.cu)
        int taken = 0;
        const int rc = run_patch_loop_persistent(pts, N, ld, perm, offsets_dev, offsets_host, P, rest_perm,
                                                 rest_offsets_host, tgt_idx, n_tgt, start_patch, eps, diffuse, E, ld_e,
                                                 order_host, value_host, ws + L.loop, L.loop_bytes, stream, &taken);
        if (rc != DNP_OK) return rc;
        if (taken) return DNP_OK;
        if (forced) {
            set_error("DNP_PATCH_LOOP=persistent: the preconditions of the persistent loop do not hold");
            return DNP_ERR_INVALID;
        }
    }
    // large loops: all per-patch fields as independent launches, then the loop on the interaction matrix (dnp_slab.cu)
    {
        const int64_t T_all = offsets_host[P] - offsets_host[0];
        if (slab_wanted(P, T_all, max_patch, world) && workspace_bytes > L.loop) {
            int taken = 0;
            const int rc = run_patch_loop_slab(pts, N, ld, perm, offsets_dev, offsets_host, P, rest_perm,
                                               rest_offsets_host, tgt_idx, n_tgt, start_patch, eps, diffuse, E, ld_e,
                                               order_host, value_host, ws + L.loop, workspace_bytes - L.loop, stream,
                                               shard, &taken);
            if (rc != DNP_OK) return rc;
            if (taken) return DNP_OK;
        }
        if (forced && forced[0] == 's') {
            set_error("DNP_PATCH_LOOP=slab: the preconditions of the slab schedule do not hold");
            return DNP_ERR_INVALID;
        }
    }
    // patches of >= 16 384 rows take the constant-window kernel, whose launches are planned on the host for an
    // explicit source range: one 8-byte read of the winner per step
    bool device_driven = max_patch < 16384;
    if (forced && forced[0] == 'h') device_driven = false;
    if (forced && forced[0] == 'd') device_driven = true;
    if (device_driven) {
        int32_t* cur = reinterpret_cast<int32_t*>(ws + L.cur);
        int64_t* rest_off_dev = reinterpret_cast<int64_t*>(ws + L.rest_off);
        int32_t* order_dev = order_host ? reinterpret_cast<int32_t*>(ws + L.order) : nullptr;
        float* value_dev = value_host ? reinterpret_cast<float*>(ws + L.value) : nullptr;
        int64_t max_rest = 0;
        if (rest_perm) {
            for (int64_t p = 0; p < P; ++p) {
                const int64_t n = rest_offsets_host[p + 1] - rest_offsets_host[p];
                if (n < 0) {
                    set_error("rest offsets must be non-decreasing");
                    return DNP_ERR_INVALID;
                }
                if (n > max_rest) max_rest = n;
            }
            DNP_CHECK_CUDA(cudaMemcpyAsync(rest_off_dev, rest_offsets_host, (size_t)(P + 1) * sizeof(int64_t),
                                           cudaMemcpyHostToDevice, stream));
        }
        SourceSelect sel;
        sel.patch = cur;
        sel.offsets = offsets_dev;
        // single-CTA step head when the rows one step touches are few (always, short of 32k-row patches)
        const bool fused = max_patch + max_rest <= kFusedRows && !getenv("DNP_PATCH_LOOP_UNFUSED");
        if (fused) {
            const PairPlan plan = make_pair_plan(max_patch > 0 ? max_patch : 1, M > 0 ? M : 1, sm_count_cached(), 0, 0);
            sel.prepacked = true;
            for (int64_t step = 0; step < P; ++step) {
                patch_step_kernel<<<1, 1024, 0, stream>>>(
                    inter, P, step > 0, (int32_t)start_patch, (int)step, cur, order_dev, value_dev, pts, ld, perm,
                    offsets_dev, rest_perm, rest_perm ? rest_off_dev : nullptr, settled, diffuse ? scratch : nullptr,
                    active, reinterpret_cast<float4*>(pairs_ws), plan.S_pad);
                count_launch(1);
                ep.mode = step == 0 ? DNP_EPI_STORE : DNP_EPI_ACCUM;
                if (max_patch > 0 && M > 0) {
                    const int rc = run_pairs(false, pts, max_patch, ld, perm, pts, M, ld, 0, eps, ep, pairs_ws,
                                             L.pairs_bytes, stream, 0, 0, sel);
                    if (rc != DNP_OK) return rc;
                }
                if (step == P - 1) break;
                const int rc = run_patch_interactions(E, ld_e, normals, ld, perm, offsets_dev, P, active, inter,
                                                      nullptr, nullptr, stream, PatchAdvance(), shard);
                if (rc != DNP_OK) return rc;
            }
        } else {
        cur_init_kernel<<<1, 1, 0, stream>>>(cur, (int32_t)start_patch, order_dev, value_dev);
        count_launch(1);
        const int64_t commit_work = max_patch + max_rest > 0 ? max_patch + max_rest : 1;
        const int64_t commit_want = (commit_work + 255) / 256, clear_want = (max_patch + 255) / 256;
        const unsigned commit_blocks = (unsigned)(commit_want < 1184 ? commit_want : 1184);
        const unsigned clear_blocks = (unsigned)(clear_want < 1184 ? clear_want : 1184);
        for (int64_t step = 0; step < P; ++step) {
            if (diffuse && step > 0 && max_patch > 0) {
                mask_clear_selected_kernel<<<clear_blocks, 256, 0, stream>>>(scratch, perm, offsets_dev, cur);
                count_launch(1);
            }
            patch_commit_selected_kernel<<<commit_blocks, 256, 0, stream>>>(
                normals, ld, perm, offsets_dev, rest_perm, rest_perm ? rest_off_dev : nullptr, cur, settled,
                diffuse ? scratch : nullptr, active);
            count_launch(1);
            ep.mode = step == 0 ? DNP_EPI_STORE : DNP_EPI_ACCUM;
            if (max_patch > 0 && M > 0) {
                const int rc = run_pairs(false, pts, max_patch, ld, perm, pts, M, ld, 0, eps, ep, pairs_ws,
                                         L.pairs_bytes, stream, 0, 0, sel);
                if (rc != DNP_OK) return rc;
            }
            if (step == P - 1) break;
            PatchAdvance adv;
            adv.cur = cur;
            adv.order = order_dev;
            adv.value = value_dev;
            adv.step = (int)(step + 1);
            const int rc = run_patch_interactions(E, ld_e, normals, ld, perm, offsets_dev, P, active, inter, best_idx,
                                                  best_val, stream, adv, shard);
            if (rc != DNP_OK) return rc;
        }
        }
        if (order_host)
            DNP_CHECK_CUDA(cudaMemcpyAsync(order_host, order_dev, (size_t)P * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                           stream));
        if (value_host)
            DNP_CHECK_CUDA(cudaMemcpyAsync(value_host, value_dev, (size_t)P * sizeof(float), cudaMemcpyDeviceToHost,
                                           stream));
        if (order_host || value_host) DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
        DNP_CHECK_CUDA(cudaGetLastError());
        return DNP_OK;
    }

    int64_t j = start_patch, prev = -1;
    int flip = 0;
    for (int64_t step = 0; step < P; ++step) {
        const int64_t o = offsets_host[j], n = offsets_host[j + 1] - o;
        const int64_t ro = rest_perm ? rest_offsets_host[j] : 0;
        const int64_t rn = rest_perm ? rest_offsets_host[j + 1] - ro : 0;
        if (diffuse && prev >= 0) {
            const int64_t po = offsets_host[prev], pn = offsets_host[prev + 1] - po;
            if (pn > 0) {
                mask_clear_kernel<<<(unsigned)((pn + 255) / 256), 256, 0, stream>>>(scratch, perm + po, pn);
                count_launch(1);
            }
        }
        {
            const int64_t work = n + rn > 0 ? n + rn : 1;
            const int64_t blocks = (work + 255) / 256;
            patch_commit_kernel<<<(unsigned)(blocks < 1184 ? blocks : 1184), 256, 0, stream>>>(
                normals, ld, perm + o, n, rest_perm ? rest_perm + ro : nullptr, rn, flip, settled,
                diffuse ? scratch : nullptr, active, j);
            count_launch(1);
        }
        if (order_host) order_host[step] = (int32_t)j;
        ep.mode = step == 0 ? DNP_EPI_STORE : DNP_EPI_ACCUM;                     // :239,311 then :261,265,330,335
        if (n > 0 && M > 0) {
            const int rc = run_pairs(false, pts, n, ld, perm + o, pts, M, ld, 0, eps, ep, pairs_ws, L.pairs_bytes,
                                     stream);
            if (rc != DNP_OK) return rc;
        }
        if (step == P - 1) break;
        const int rc = run_patch_interactions(E, ld_e, normals, ld, perm, offsets_dev, P, active, inter, best_idx,
                                              best_val, stream, PatchAdvance(), shard);   // :244-247,316-319
        if (rc != DNP_OK) return rc;
        struct { int32_t idx; float val; } best;
        DNP_CHECK_CUDA(cudaMemcpyAsync(&best, best_idx, sizeof(best), cudaMemcpyDeviceToHost, stream));
        DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
        if (best.idx < 0 || best.idx >= P) {
            set_error("no remaining patch at step %lld", (long long)step);
            return DNP_ERR_INVALID;
        }
        prev = j;
        j = best.idx;
        flip = best.val < 0.f;                                                    // :250,322
        if (value_host) value_host[step + 1] = best.val;
    }
    if (value_host) value_host[0] = 0.f;
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
