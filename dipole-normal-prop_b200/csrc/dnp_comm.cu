// dnp_comm.cu -- the multi-GPU entry points of the C ABI: one process per GPU, NCCL over
// NVLink / NVSwitch, the collective enqueued on the SAME stream as the shard's kernels so that a non-PyTorch host gets
// the whole sharded call from one function.  The reference is single-process, single-device: this layer is net-new.
//
//   field / potential   targets block-sharded over the ranks (every target is independent), sources replicated;
//                        each rank writes its block of the full result, then the blocks are exchanged in one NCCL
//                        group (a broadcast per rank: blocks may differ by one row)
//   edge interactions    sum_t n_t . E_S(t) per edge (field_utils.py:145-160): every rank sums over ITS slice of each
//                        target patch, ONE ncclAllReduce(sum) over the edge vector combines them
//   lattice mean         sources block-sharded (the potential is linear in the sources), ONE ncclAllReduce of a scalar
//
// NCCL is loaded at run time (dlopen "libnccl.so.2", or $DNP_NCCL_LIB): libdipole.so has no link-time dependency on
// it, single-GPU users never touch it, and inside a PyTorch process the already loaded torch-bundled NCCL is reused.
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"

namespace dnp {

namespace {

// the few NCCL declarations used (stable since NCCL 2.0)
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId { char internal[128]; };
constexpr int kNcclFloat32 = 7, kNcclSum = 0;

struct Nccl {
    void* handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};

std::mutex g_comm_mu;
Nccl g_nccl;
ncclComm_t g_comm = nullptr;
int g_rank = 0, g_world = 1, g_device = -1;

int load_nccl() {
    if (g_nccl.handle) return DNP_OK;
    const char* names[] = {getenv("DNP_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names)
        if (n && (h = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!h) {
        set_error("NCCL not found (dlopen libnccl.so.2 failed: %s); set DNP_NCCL_LIB", dlerror());
        return DNP_ERR_INVALID;
    }
#define DNP_SYM(field, name)                                                 \
    *reinterpret_cast<void**>(&g_nccl.field) = dlsym(h, name);               \
    if (!g_nccl.field) { set_error("NCCL symbol %s missing", name); dlclose(h); return DNP_ERR_INVALID; }
    DNP_SYM(GetUniqueId, "ncclGetUniqueId")
    DNP_SYM(CommInitRank, "ncclCommInitRank")
    DNP_SYM(CommDestroy, "ncclCommDestroy")
    DNP_SYM(AllReduce, "ncclAllReduce")
    DNP_SYM(Broadcast, "ncclBroadcast")
    DNP_SYM(GroupStart, "ncclGroupStart")
    DNP_SYM(GroupEnd, "ncclGroupEnd")
    DNP_SYM(GetErrorString, "ncclGetErrorString")
#undef DNP_SYM
    g_nccl.handle = h;
    return DNP_OK;
}

int nccl_fail(int rc, const char* what) {
    set_error("NCCL error %d (%s) in %s", rc, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?", what);
    return DNP_ERR_CUDA;
}
#define DNP_CHECK_NCCL(expr)                               \
    do {                                                   \
        const int _r = (expr);                             \
        if (_r != 0) return nccl_fail(_r, #expr);          \
    } while (0)

// block partition of range(n): the first n % world ranks get one extra element (dnp_dist.shard_range)
inline void shard_range(int64_t n, int rank, int world, int64_t* a, int64_t* b) {
    const int64_t base = n / world, rem = n % world;
    *a = rank * base + (rank < rem ? rank : rem);
    *b = *a + base + (rank < rem ? 1 : 0);
}

int require_comm() {
    if (!g_comm) {
        set_error("no communicator: call dnp_comm_init first");
        return DNP_ERR_INVALID;
    }
    int dev = -1;
    DNP_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev != g_device) {
        set_error("the communicator was created on device %d, the current device is %d", g_device, dev);
        return DNP_ERR_INVALID;
    }
    return DNP_OK;
}

// this rank's slice of every patch as a second CSR: sliced_perm[new_off[p] ..) = perm[off[p] + a_p .. off[p] + b_p)
__global__ void slice_offsets_kernel(const int64_t* __restrict__ offsets, int64_t P, int rank, int world,
                                     int64_t* __restrict__ counts) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const int64_t n = offsets[p + 1] - offsets[p], base = n / world, rem = n % world;
    counts[p] = base + (rank < rem ? 1 : 0);
}
__global__ void exclusive_scan_kernel(const int64_t* __restrict__ counts, int64_t P, int64_t* __restrict__ new_off) {
    // P is the number of patches (~1e3): one thread, fixed order
    if (blockIdx.x || threadIdx.x) return;
    int64_t acc = 0;
    for (int64_t p = 0; p < P; ++p) { new_off[p] = acc; acc += counts[p]; }
    new_off[P] = acc;
}
__global__ void slice_gather_kernel(const int64_t* __restrict__ perm, const int64_t* __restrict__ offsets,
                                    const int64_t* __restrict__ new_off, int64_t P, int rank, int world,
                                    int64_t* __restrict__ sliced) {
    const int64_t p = blockIdx.x;
    if (p >= P) return;
    const int64_t n = offsets[p + 1] - offsets[p], base = n / world, rem = n % world;
    const int64_t a = rank * base + (rank < rem ? rank : rem);
    const int64_t cnt = new_off[p + 1] - new_off[p];
    for (int64_t i = threadIdx.x; i < cnt; i += blockDim.x) sliced[new_off[p] + i] = perm[offsets[p] + a + i];
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

// ShardHook::combine of the sharded greedy loop: sum the per-patch vectors of all ranks in place, on the loop's stream
int combine_interactions(void*, float* vals, int64_t n, cudaStream_t stream) {
    if (g_world > 1) DNP_CHECK_NCCL(g_nccl.AllReduce(vals, vals, (size_t)n, kNcclFloat32, kNcclSum, g_comm, stream));
    return DNP_OK;
}

}  // namespace

}  // namespace dnp

using namespace dnp;

extern "C" {

int dnp_comm_unique_id(void* id128) {
    if (!id128) { set_error("null id"); return DNP_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(g_comm_mu);
    const int rc = load_nccl();
    if (rc != DNP_OK) return rc;
    ncclUniqueId id;
    DNP_CHECK_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128, id.internal, sizeof(id.internal));
    return DNP_OK;
}

int dnp_comm_init(const void* id128, int rank, int world) {
    if (!id128 || world < 1 || rank < 0 || rank >= world) { set_error("bad communicator arguments"); return DNP_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(g_comm_mu);
    if (g_comm) { set_error("communicator already initialised (dnp_comm_destroy first)"); return DNP_ERR_INVALID; }
    const int rc = load_nccl();
    if (rc != DNP_OK) return rc;
    ncclUniqueId id;
    memcpy(id.internal, id128, sizeof(id.internal));
    DNP_CHECK_CUDA(cudaGetDevice(&g_device));
    DNP_CHECK_NCCL(g_nccl.CommInitRank(&g_comm, world, id, rank));
    g_rank = rank;
    g_world = world;
    return DNP_OK;
}

int dnp_comm_destroy(void) {
    std::lock_guard<std::mutex> lock(g_comm_mu);
    if (g_comm) {
        g_nccl.CommDestroy(g_comm);
        g_comm = nullptr;
    }
    g_rank = 0;
    g_world = 1;
    g_device = -1;
    return DNP_OK;
}

int dnp_comm_rank(int* rank, int* world) {
    if (rank) *rank = g_rank;
    if (world) *world = g_world;
    return g_comm ? DNP_OK : DNP_ERR_INVALID;
}

size_t dnp_sharded_workspace_bytes(int64_t S, int64_t M, int64_t P, int64_t patch_rows) {
    const int64_t Mr = M > 0 ? M : 1;
    return pairs_workspace_bytes(S > 0 ? S : 1, Mr) + up256((size_t)(P + 1) * 16) + up256((size_t)(patch_rows > 0 ? patch_rows : 1) * 8) + 1024;
}

int dnp_field_grad_sharded_f32(const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t,
                               float eps, float* E, int64_t ld_e, int gather, void* workspace, size_t workspace_bytes,
                               void* stream) {
    int rc = require_comm();
    if (rc != DNP_OK) return rc;
    if (S < 0 || M < 0 || !tgt || !E || ld_s < 6 || ld_t < 3 || ld_e < 3 || (gather && ld_e != 3)) {
        set_error("bad argument (a gathered result needs dense rows, ld_e == 3)");
        return DNP_ERR_INVALID;
    }
    int64_t a, b;
    shard_range(M, g_rank, g_world, &a, &b);
    PairEpilogue ep;
    ep.mode = DNP_EPI_STORE;
    ep.out = E + a * ld_e;
    ep.ld_out = ld_e;
    if (b > a) {
        rc = run_pairs(false, src, S, ld_s, nullptr, tgt + a * ld_t, b - a, ld_t, 0, eps, ep, workspace, workspace_bytes,
                       (cudaStream_t)stream);
        if (rc != DNP_OK) return rc;
    }
    if (gather && g_world > 1) {
        DNP_CHECK_NCCL(g_nccl.GroupStart());
        for (int r = 0; r < g_world; ++r) {
            int64_t ra, rb;
            shard_range(M, r, g_world, &ra, &rb);
            if (rb > ra)
                DNP_CHECK_NCCL(g_nccl.Broadcast(E + ra * 3, E + ra * 3, (size_t)(rb - ra) * 3, kNcclFloat32, r, g_comm,
                                                (cudaStream_t)stream));
        }
        DNP_CHECK_NCCL(g_nccl.GroupEnd());
    }
    return DNP_OK;
}

int dnp_edge_interactions_sharded_f32(const float* pts, int64_t ld, const int64_t* perm, const int64_t* offsets, int64_t P,
                                      int64_t patch_rows, const int64_t* edges, int64_t n_edges, float eps, float* out,
                                      void* workspace, size_t workspace_bytes, void* stream) {
    int rc = require_comm();
    if (rc != DNP_OK) return rc;
    if (n_edges <= 0) return DNP_OK;
    if (!pts || !perm || !offsets || !edges || !out || ld < 6 || P < 1 || patch_rows < 0) {
        set_error("bad argument");
        return DNP_ERR_INVALID;
    }
    const size_t need = up256((size_t)(P + 1) * 16) + up256((size_t)(patch_rows > 0 ? patch_rows : 1) * 8);
    if (!workspace || workspace_bytes < need) {
        set_error("workspace too small: need %zu bytes (dnp_sharded_workspace_bytes), got %zu", need, workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int64_t* counts = reinterpret_cast<int64_t*>(workspace);
    int64_t* new_off = counts + (P + 1);
    int64_t* sliced = reinterpret_cast<int64_t*>(reinterpret_cast<unsigned char*>(workspace) + up256((size_t)(P + 1) * 16));
    slice_offsets_kernel<<<(unsigned)((P + 255) / 256), 256, 0, st>>>(offsets, P, g_rank, g_world, counts);
    exclusive_scan_kernel<<<1, 1, 0, st>>>(counts, P, new_off);
    slice_gather_kernel<<<(unsigned)P, 256, 0, st>>>(perm, offsets, new_off, P, g_rank, g_world, sliced);
    count_launch(3);
    DNP_CHECK_CUDA(cudaGetLastError());
    rc = run_edge_interactions_batched(pts, ld, perm, offsets, P, sliced, new_off, edges, n_edges, eps, out, st);
    if (rc != DNP_OK) return rc;
    if (g_world > 1) DNP_CHECK_NCCL(g_nccl.AllReduce(out, out, (size_t)n_edges, kNcclFloat32, kNcclSum, g_comm, st));
    return DNP_OK;
}

size_t dnp_patch_propagate_sharded_workspace_bytes(int64_t N, int64_t P, int64_t max_patch, int64_t n_tgt,
                                                   int64_t patch_rows, int64_t n_owned) {
    return patch_propagate_sharded_workspace_bytes(N, P, max_patch, n_tgt, patch_rows, n_owned, g_world);
}

int dnp_patch_propagate_sharded_f32(float* pts, int64_t N, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                                    const int64_t* offsets_host, int64_t P, const int64_t* rest_perm,
                                    const int64_t* rest_offsets_host, const int64_t* tgt_idx, int64_t n_tgt,
                                    const uint8_t* owned, int64_t start_patch, float eps, int diffuse, float* E,
                                    int64_t ld_e, int32_t* order_host, float* value_host, void* workspace,
                                    size_t workspace_bytes, void* stream) {
    int rc = require_comm();
    if (rc != DNP_OK) return rc;
    if (N < 0 || P < 0 || n_tgt < 0) { set_error("negative size"); return DNP_ERR_INVALID; }
    if (P == 0) return DNP_OK;
    if (!pts || !perm || !offsets_dev || !offsets_host || !E || !owned || ld < 6 || ld_e < 3 || eps < 0.f ||
        (rest_perm == nullptr) != (rest_offsets_host == nullptr) || (n_tgt > 0 && !tgt_idx)) {
        set_error("bad argument (the sharded loop needs the owned-patch mask and this rank's target rows)");
        return DNP_ERR_INVALID;
    }
    ShardHook hook;
    hook.owned = owned;
    hook.combine = combine_interactions;
    hook.world = g_world;
    // a rank that owns no patch still takes part in every all-reduce: its target list is empty, not absent
    static const int64_t no_rows[1] = {0};
    return run_patch_propagate(pts, N, ld, perm, offsets_dev, offsets_host, P, rest_perm, rest_offsets_host,
                               n_tgt > 0 ? tgt_idx : no_rows, n_tgt, start_patch, eps, diffuse, E, ld_e, order_host,
                               value_host, workspace, workspace_bytes, (cudaStream_t)stream, &hook);
}

int dnp_lattice_mean_potential_sharded_f32(const float* src, int64_t S, int64_t ld_s, int n, float* mean_out,
                                           void* workspace, size_t workspace_bytes, void* stream) {
    int rc = require_comm();
    if (rc != DNP_OK) return rc;
    if (S < 0 || !mean_out || ld_s < 6) { set_error("bad argument"); return DNP_ERR_INVALID; }
    int64_t a, b;
    shard_range(S, g_rank, g_world, &a, &b);
    // this rank's share of the sum over sources, already divided by n^3; the means of the shards add up
    rc = run_lattice_mean(src + a * ld_s, b - a, ld_s, n, mean_out, workspace, workspace_bytes, (cudaStream_t)stream);
    if (rc != DNP_OK) return rc;
    if (g_world > 1)
        DNP_CHECK_NCCL(g_nccl.AllReduce(mean_out, mean_out, 1, kNcclFloat32, kNcclSum, g_comm, (cudaStream_t)stream));
    return DNP_OK;
}

}  // extern "C"
