// dnp_abi.cu -- the extern "C" boundary of libdipole.so (see include/dipole.h).
//
// Every entry point validates its arguments, never throws, never allocates an
// output, and reports failures through a negative return code plus a
// thread-local message (dnp_last_error).  The library holds no mutable global
// state besides cached device properties and a launch counter, so it is safe to
// call from many host threads (the reference does: util.py:187-196,308-327).
#include <stdarg.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <errno.h>
#include <unistd.h>

#include <chrono>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"

namespace dnp {

static thread_local char t_error[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_error, sizeof(t_error), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t err, const char* what) {
    set_error("CUDA error %d (%s) in %s", (int)err, cudaGetErrorString(err), what);
    return DNP_ERR_CUDA;
}

static int invalid(const char* msg) {
    set_error("invalid argument: %s", msg);
    return DNP_ERR_INVALID;
}

}  // namespace dnp

using namespace dnp;

#define DNP_STR2(x) #x
#define DNP_STR(x) DNP_STR2(x)
#define DNP_REQUIRE(cond, msg) \
    do {                       \
        if (!(cond)) return invalid(msg); \
    } while (0)

extern "C" {

const char* dnp_last_error(void) { return t_error; }

int dnp_version(void) { return 100; }

#ifndef DNP_SOURCE_HASH
#define DNP_SOURCE_HASH "unknown"
#endif
const char* dnp_build_info(void) {
    return "libdipole sm_100a sources=" DNP_SOURCE_HASH " nvcc=" DNP_STR(__CUDACC_VER_MAJOR__) "." DNP_STR(__CUDACC_VER_MINOR__)
           " built=" __DATE__;
}

int dnp_device_info(int* sm_count, int* sm_clock_khz, int* cc_major, int* cc_minor) {
    int dev = 0;
    DNP_CHECK_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    DNP_CHECK_CUDA(cudaGetDeviceProperties(&prop, dev));
    int clock_khz = 0;
    DNP_CHECK_CUDA(cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, dev));
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (sm_clock_khz) *sm_clock_khz = clock_khz;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    return DNP_OK;
}

size_t dnp_pairs_workspace_bytes(int64_t S, int64_t M) { return pairs_workspace_bytes(S, M); }

int64_t dnp_launch_count(void) { return launch_count(); }

// ------------------------------------------------------------------ potential
int dnp_potential_f32(const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx,
                      const float* tgt, int64_t M, int64_t ld_t,
                      float* phi, int accumulate,
                      void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(S >= 0 && M >= 0, "negative size");
    if (M == 0) return DNP_OK;
    DNP_REQUIRE(tgt && phi, "null target or output pointer");
    DNP_REQUIRE(S == 0 || src, "null source pointer");
    DNP_REQUIRE(ld_s >= 6 && ld_t >= 3, "row stride too small (sources need 6 columns, targets 3)");
    PairEpilogue ep;
    ep.mode = accumulate ? DNP_EPI_ACCUM : DNP_EPI_STORE;
    ep.out = phi;
    ep.ld_out = 1;
    return run_pairs(true, src, S, ld_s, src_idx, tgt, M, ld_t, 0, 0.f, ep, workspace, workspace_bytes,
                     (cudaStream_t)stream);
}

int dnp_lattice_mean_potential_f32(const float* src, int64_t S, int64_t ld_s,
                                   int n, float* phi_opt, float* mean_out,
                                   void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(S >= 0 && n > 0 && n <= 256, "bad size");
    DNP_REQUIRE(mean_out, "null output pointer");
    DNP_REQUIRE(S == 0 || src, "null source pointer");
    DNP_REQUIRE(ld_s >= 6, "row stride too small");
    const int64_t M = (int64_t)n * n * n;
    const size_t need = pairs_workspace_bytes(S, M);
    // callers that only want the mean (every pipeline) take the sources-in-registers kernel; DNP_LATTICE_FAST=0 keeps
    // the generic path (A/B measurements)
    const char* fast = getenv("DNP_LATTICE_FAST");
    if (!phi_opt && n <= 255 && !(fast && fast[0] == '0') && workspace && workspace_bytes >= lattice_mean_workspace_bytes())
        return run_lattice_mean(src, S, ld_s, n, mean_out, workspace, workspace_bytes, (cudaStream_t)stream);
    if (!workspace || workspace_bytes < need) {
        set_error("workspace too small: need %zu bytes, got %zu", need, workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    // per-node potentials live in the caller's buffer or at the tail of the workspace
    const size_t phi_off = (need - 1024 - (size_t)((M * 4 + 255) / 256 * 256));
    float* phi = phi_opt ? phi_opt : reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + phi_off);
    PairEpilogue ep;
    ep.mode = DNP_EPI_STORE;
    ep.out = phi;
    ep.ld_out = 1;
    int rc = run_pairs(true, src, S, ld_s, nullptr, nullptr, M, 3, n, 0.f, ep, workspace, phi_off,
                       (cudaStream_t)stream);
    if (rc != DNP_OK) return rc;
    return run_sum(phi, M, 1, 1.0f / (float)M, mean_out, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ field
int dnp_field_apply_f32(const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx,
                        const float* tgt, int64_t M, int64_t ld_t,
                        const int64_t* tgt_idx, const uint8_t* tgt_skip,
                        float eps, int mode,
                        float* out, int64_t ld_out, float* normals, int64_t ld_n,
                        int32_t* nonfinite,
                        void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(S >= 0 && M >= 0, "negative size");
    if (M == 0) return DNP_OK;
    DNP_REQUIRE(tgt, "null target pointer");
    DNP_REQUIRE(S == 0 || src, "null source pointer");
    DNP_REQUIRE(ld_s >= 6 && ld_t >= 3, "row stride too small (sources need 6 columns, targets 3)");
    DNP_REQUIRE(mode >= DNP_EPI_STORE && mode <= DNP_EPI_DOT, "unknown epilogue mode");
    const bool needs_out = mode == DNP_EPI_STORE || mode == DNP_EPI_ACCUM || mode == DNP_EPI_NORMALIZE ||
                           mode == DNP_EPI_DOT;
    const bool needs_n = mode == DNP_EPI_SIGN_GE || mode == DNP_EPI_SIGN_GT || mode == DNP_EPI_DOT;
    DNP_REQUIRE(!needs_out || (out && ld_out >= (mode == DNP_EPI_DOT ? 1 : 3)), "bad output pointer / stride");
    DNP_REQUIRE(!needs_n || (normals && ld_n >= 3), "bad normals pointer / stride");
    DNP_REQUIRE(eps >= 0.f, "eps must be >= 0");
    PairEpilogue ep;
    ep.mode = mode;
    ep.out = out;
    ep.ld_out = ld_out;
    ep.normals = normals;
    ep.ld_n = ld_n;
    ep.tgt_idx = tgt_idx;
    ep.tgt_skip = tgt_skip;
    ep.nonfinite = nonfinite;
    return run_pairs(false, src, S, ld_s, src_idx, tgt, M, ld_t, 0, eps, ep, workspace, workspace_bytes,
                     (cudaStream_t)stream);
}

int dnp_field_grad_f32(const float* src, int64_t S, int64_t ld_s,
                       const float* tgt, int64_t M, int64_t ld_t,
                       float eps, float* E, int64_t ld_e, int accumulate,
                       void* workspace, size_t workspace_bytes, void* stream) {
    return dnp_field_apply_f32(src, S, ld_s, nullptr, tgt, M, ld_t, nullptr, nullptr, eps,
                               accumulate ? DNP_EPI_ACCUM : DNP_EPI_STORE, E, ld_e, nullptr, 0, nullptr,
                               workspace, workspace_bytes, stream);
}

// ------------------------------------------------------------------ edge score (single pair)
int dnp_edge_interaction_f32(const float* S_pts, int64_t S, int64_t ld_s,
                             const float* T_pts, int64_t T, int64_t ld_t,
                             float eps, float* interaction,
                             void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(S >= 0 && T >= 0, "negative size");
    DNP_REQUIRE(interaction, "null output pointer");
    DNP_REQUIRE(ld_s >= 6 && ld_t >= 6, "both patches need 6 columns");
    DNP_REQUIRE(T == 0 || T_pts, "null target pointer");
    const size_t need = pairs_workspace_bytes(S, T);
    if (!workspace || workspace_bytes < need) {
        set_error("workspace too small: need %zu bytes, got %zu", need, workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    const int64_t Tn = T < 1 ? 1 : T;
    const size_t dot_off = need - 1024 - (size_t)((Tn * 4 + 255) / 256 * 256);
    float* dots = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + dot_off);
    if (T > 0) {
        int rc = dnp_field_apply_f32(S_pts, S, ld_s, nullptr, T_pts, T, ld_t, nullptr, nullptr, eps, DNP_EPI_DOT,
                                     dots, 1, const_cast<float*>(T_pts) + 3, ld_t, nullptr, workspace, dot_off,
                                     stream);
        if (rc != DNP_OK) return rc;
    }
    return run_sum(dots, T, 1, 1.0f, interaction, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ host-buffer entry points
namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int device) {
        if (cudaGetDevice(&prev) == cudaSuccess && cudaSetDevice(device) == cudaSuccess) ok = true;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// Per-host-thread scratch of the host-buffer entry points (released when the thread exits).
struct HostScratch {
    int device = -1;
    cudaStream_t stream = nullptr;
    unsigned char* dev = nullptr;
    size_t dev_cap = 0;
    unsigned char* pin = nullptr;
    size_t pin_cap = 0;

    void release() {
        if (dev) cudaFree(dev);
        if (pin) cudaFreeHost(pin);
        if (stream) cudaStreamDestroy(stream);
        dev = nullptr; pin = nullptr; stream = nullptr;
        dev_cap = pin_cap = 0;
        device = -1;
    }
    ~HostScratch() { release(); }

    cudaError_t prepare(int d, size_t dev_bytes, size_t pin_bytes) {
        cudaError_t e;
        if (device != d) {
            release();
            if ((e = cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking)) != cudaSuccess) return e;
            device = d;
        }
        if (dev_bytes > dev_cap) {
            if (dev) cudaFree(dev);
            dev = nullptr; dev_cap = 0;
            if ((e = cudaMalloc(reinterpret_cast<void**>(&dev), dev_bytes)) != cudaSuccess) return e;
            dev_cap = dev_bytes;
        }
        if (pin_bytes > pin_cap) {
            if (pin) cudaFreeHost(pin);
            pin = nullptr; pin_cap = 0;
            if ((e = cudaMallocHost(reinterpret_cast<void**>(&pin), pin_bytes)) != cudaSuccess) return e;
            pin_cap = pin_bytes;
        }
        return cudaSuccess;
    }
};
thread_local HostScratch t_host;

// copies a strided [rows, cols] host matrix into a dense device matrix
cudaError_t upload_rows(float* dst, const float* src_host, int64_t rows, int64_t cols, int64_t ld,
                        cudaStream_t stream) {
    if (rows == 0) return cudaSuccess;
    if (ld == cols)   // dense rows: one flat copy (a 2-D copy of 12/24-byte rows runs at a fraction of PCIe speed)
        return cudaMemcpyAsync(dst, src_host, (size_t)rows * cols * sizeof(float), cudaMemcpyHostToDevice, stream);
    return cudaMemcpy2DAsync(dst, cols * sizeof(float), src_host, ld * sizeof(float), cols * sizeof(float), rows,
                             cudaMemcpyHostToDevice, stream);
}

int host_pairs(bool potential, const float* src_host, int64_t S, int64_t ld_s, const float* tgt_host, int64_t M,
               int64_t ld_t, float eps, float* out_host, int device) {
    DNP_REQUIRE(S >= 0 && M >= 0, "negative size");
    if (M == 0) return DNP_OK;
    DNP_REQUIRE(tgt_host && out_host, "null target or output pointer");
    DNP_REQUIRE(S == 0 || src_host, "null source pointer");
    DNP_REQUIRE(ld_s >= 6 && ld_t >= 3, "row stride too small");
    DeviceGuard guard(device);
    if (!guard.ok) {
        set_error("cannot select CUDA device %d", device);
        return DNP_ERR_NO_DEVICE;
    }
    const int ncomp = potential ? 1 : 3;
    const size_t ws_bytes = pairs_workspace_bytes(S, M);
    const size_t src_bytes = (size_t)(S > 0 ? S : 1) * 6 * sizeof(float);
    const size_t tgt_bytes = (size_t)M * 3 * sizeof(float);
    const size_t out_bytes = (size_t)M * ncomp * sizeof(float);
    const bool trace = getenv("DNP_HOST_TRACE") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_0 = now();
    // per-thread stream, device arena and pinned result buffer, kept between calls: a call then consists of
    // copies and launches only (no allocation, no stream creation -- both take driver-wide locks and made the
    // end-to-end time sensitive to anything else that talks to the driver, e.g. a monitoring nvidia-smi)
    const size_t a256 = 256;
    auto up = [&](size_t v) { return (v + a256 - 1) / a256 * a256; };
    const size_t total = up(src_bytes) + up(tgt_bytes) + up(out_bytes) + ws_bytes;
    cudaError_t e = t_host.prepare(device, total, out_bytes);
    if (e != cudaSuccess) return cuda_fail(e, "host scratch");
    cudaStream_t stream = t_host.stream;
    unsigned char* base = t_host.dev;
    const double t_alloc = now();
    double t_up = 0, t_launch = 0, t_dl = 0;
    float* d_src = reinterpret_cast<float*>(base);
    float* d_tgt = reinterpret_cast<float*>(base + up(src_bytes));
    float* d_out = reinterpret_cast<float*>(base + up(src_bytes) + up(tgt_bytes));
    void* d_ws = base + up(src_bytes) + up(tgt_bytes) + up(out_bytes);
    int rc = DNP_OK;
    do {
        if ((e = upload_rows(d_src, src_host, S, 6, ld_s, stream)) != cudaSuccess) { rc = cuda_fail(e, "upload sources"); break; }
        if ((e = upload_rows(d_tgt, tgt_host, M, 3, ld_t, stream)) != cudaSuccess) { rc = cuda_fail(e, "upload targets"); break; }
        t_up = now();
        PairEpilogue ep;
        ep.mode = DNP_EPI_STORE;
        ep.out = d_out;
        ep.ld_out = ncomp;
        rc = run_pairs(potential, d_src, S, 6, nullptr, d_tgt, M, 3, 0, eps, ep, d_ws, ws_bytes, stream);
        t_launch = now();
        if (rc != DNP_OK) break;
        // result -> pinned buffer (asynchronous), then one host memcpy into the caller's (pageable) array
        if ((e = cudaMemcpyAsync(t_host.pin, d_out, out_bytes, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) { rc = cuda_fail(e, "download"); break; }
    } while (0);
    t_dl = now();
    e = cudaStreamSynchronize(stream);
    const double t_sync = now();
    if (rc == DNP_OK && e == cudaSuccess) memcpy(out_host, t_host.pin, out_bytes);
    if (trace)
        fprintf(stderr, "[dnp host] scratch %.2f upload %.2f enqueue %.2f download-enqueue %.2f sync %.2f copy-out %.2f ms\n",
                t_alloc - t_0, t_up - t_alloc, t_launch - t_up, t_dl - t_launch, t_sync - t_dl, now() - t_sync);
    if (rc == DNP_OK && e != cudaSuccess) rc = cuda_fail(e, "cudaStreamSynchronize");
    return rc;
}

}  // namespace

int dnp_field_grad_host_f32(const float* src_host, int64_t S, int64_t ld_s,
                            const float* tgt_host, int64_t M, int64_t ld_t,
                            float eps, float* E_host, int device) {
    DNP_REQUIRE(eps >= 0.f, "eps must be >= 0");
    return host_pairs(false, src_host, S, ld_s, tgt_host, M, ld_t, eps, E_host, device);
}

int dnp_potential_host_f32(const float* src_host, int64_t S, int64_t ld_s,
                           const float* tgt_host, int64_t M, int64_t ld_t,
                           float* phi_host, int device) {
    return host_pairs(true, src_host, S, ld_s, tgt_host, M, ld_t, 0.f, phi_host, device);
}

}  // extern "C"

// ------------------------------------------------------------------ scores, drivers, measurement
extern "C" {

int dnp_edge_interactions_batched_f32(const float* pts, int64_t ld,
                                      const int64_t* perm, const int64_t* offsets, int64_t P,
                                      const int64_t* tgt_perm, const int64_t* tgt_offsets,
                                      const int64_t* edges, int64_t n_edges,
                                      float eps, float* out, void* stream) {
    DNP_REQUIRE(n_edges >= 0 && P >= 0, "negative size");
    if (n_edges == 0) return DNP_OK;
    DNP_REQUIRE(pts && perm && offsets && edges && out, "null pointer");
    DNP_REQUIRE((tgt_perm == nullptr) == (tgt_offsets == nullptr), "tgt_perm and tgt_offsets go together");
    DNP_REQUIRE(ld >= 6, "points need 6 columns");
    DNP_REQUIRE(eps >= 0.f, "eps must be >= 0");
    return run_edge_interactions_batched(pts, ld, perm, offsets, P, tgt_perm ? tgt_perm : perm,
                                         tgt_offsets ? tgt_offsets : offsets, edges, n_edges, eps, out,
                                         (cudaStream_t)stream);
}

int dnp_patch_interactions_f32(const float* E, int64_t ld_e, const float* normals, int64_t ld_n,
                               const int64_t* perm, const int64_t* offsets, int64_t P,
                               const uint8_t* active, float* out,
                               int32_t* best_idx, float* best_val, void* stream) {
    DNP_REQUIRE(P >= 0, "negative size");
    if (P == 0) return DNP_OK;
    DNP_REQUIRE(E && normals && perm && offsets && out, "null pointer");
    DNP_REQUIRE(ld_e >= 3 && ld_n >= 3, "row stride too small");
    DNP_REQUIRE((best_idx == nullptr) == (best_val == nullptr), "best_idx and best_val go together");
    return run_patch_interactions(E, ld_e, normals, ld_n, perm, offsets, P, active, out, best_idx, best_val,
                                  (cudaStream_t)stream);
}

int dnp_patch_covariance_f32(const float* pts, int64_t ld, const int64_t* perm, const int64_t* offsets, int64_t P,
                             double* cov, void* stream) {
    DNP_REQUIRE(P >= 0, "negative size");
    if (P == 0) return DNP_OK;
    DNP_REQUIRE(pts && perm && offsets && cov, "null pointer");
    DNP_REQUIRE(ld >= 3, "points need 3 columns");
    return run_patch_covariance(pts, ld, perm, offsets, P, cov, (cudaStream_t)stream);
}

int dnp_field_sign_f32(const float* E, int64_t ld_e, float* normals, int64_t ld_n,
                       const int64_t* idx, int64_t n, int mode, void* stream) {
    DNP_REQUIRE(n >= 0, "negative size");
    if (n == 0) return DNP_OK;
    DNP_REQUIRE(E && normals, "null pointer");
    DNP_REQUIRE(ld_e >= 3 && ld_n >= 3, "row stride too small");
    DNP_REQUIRE(mode == DNP_EPI_SIGN_GE || mode == DNP_EPI_SIGN_GT, "mode must be DNP_EPI_SIGN_GE or DNP_EPI_SIGN_GT");
    return run_field_sign(E, ld_e, normals, ld_n, idx, n, mode == DNP_EPI_SIGN_GE, (cudaStream_t)stream);
}

int64_t dnp_points_max_n(void) { return points_grid_max_n(); }

size_t dnp_points_workspace_bytes(int64_t N) { return points_workspace_bytes(N); }

int dnp_points_propagate_f32(float* pts, int64_t N, int64_t ld, int64_t starting_point,
                             float eps, int diffuse,
                             int32_t* order_opt, uint8_t* flipped_opt,
                             void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(N >= 0, "negative size");
    if (N == 0) return DNP_OK;
    DNP_REQUIRE(pts, "null pointer");
    DNP_REQUIRE(ld >= 6, "points need 6 columns");
    DNP_REQUIRE(starting_point >= 0 && starting_point < N, "starting_point out of range");
    DNP_REQUIRE(N < 0x7ffffffe, "N too large");
    DNP_REQUIRE(eps >= 0.f, "eps must be >= 0");
    return run_points_propagate(pts, N, ld, starting_point, eps, diffuse, order_opt, flipped_opt, workspace,
                                workspace_bytes, (cudaStream_t)stream);
}

size_t dnp_patch_propagate_workspace_bytes(int64_t N, int64_t P, int64_t max_patch, int64_t n_tgt) {
    return patch_propagate_workspace_bytes(N, P, max_patch, n_tgt);
}

int dnp_patch_propagate_f32(float* pts, int64_t N, int64_t ld, const int64_t* perm, const int64_t* offsets_dev,
                            const int64_t* offsets_host, int64_t P, const int64_t* rest_perm,
                            const int64_t* rest_offsets_host, const int64_t* tgt_idx, int64_t n_tgt,
                            int64_t start_patch, float eps, int diffuse, float* E, int64_t ld_e, int32_t* order_host,
                            float* value_host, void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(N >= 0 && P >= 0 && n_tgt >= 0, "negative size");
    if (P == 0) return DNP_OK;
    DNP_REQUIRE(pts && perm && offsets_dev && offsets_host && E, "null pointer");
    DNP_REQUIRE(ld >= 6 && ld_e >= 3, "row stride too small");
    DNP_REQUIRE((rest_perm == nullptr) == (rest_offsets_host == nullptr), "rest_perm and rest_offsets_host go together");
    DNP_REQUIRE(eps >= 0.f, "eps must be >= 0");
    return run_patch_propagate(pts, N, ld, perm, offsets_dev, offsets_host, P, rest_perm, rest_offsets_host, tgt_idx,
                               n_tgt, start_patch, eps, diffuse, E, ld_e, order_host, value_host, workspace,
                               workspace_bytes, (cudaStream_t)stream);
}

int64_t dnp_points_batched_max_n(void) { return points_batched_max_n(); }

int dnp_points_propagate_batched_f32(float* pts, int64_t ld, const int64_t* offsets, int64_t P, int64_t max_patch,
                                     float eps, int diffuse, void* stream) {
    DNP_REQUIRE(P >= 0 && max_patch >= 0, "negative size");
    DNP_REQUIRE(P == 0 || (pts && offsets), "null pointer");
    DNP_REQUIRE(ld >= 6, "points need 6 columns");
    DNP_REQUIRE(eps >= 0.f, "eps must be >= 0");
    return run_points_propagate_batched(pts, ld, offsets, P, max_patch, eps, diffuse, (cudaStream_t)stream);
}

int dnp_patch_lattice_sign_f32(float* pts, int64_t ld, const int64_t* offsets, int64_t P, int n, float* mean_out,
                               int flip, void* stream) {
    DNP_REQUIRE(P >= 0, "negative size");
    DNP_REQUIRE(P == 0 || (pts && offsets && mean_out), "null pointer");
    DNP_REQUIRE(ld >= 6, "points need 6 columns");
    return run_patch_lattice_sign(pts, ld, offsets, P, n, mean_out, flip, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ "xie" interaction (dnp_xie.cu)
static int xie_matrix_abi(bool field, const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M,
                          int64_t ld_t, float C, const double* rad2, float* out, int64_t ld_o, void* stream) {
    DNP_REQUIRE(S >= 0 && M >= 0, "negative size");
    if (S == 0 || M == 0) return DNP_OK;
    DNP_REQUIRE(src && tgt && out, "null pointer");
    DNP_REQUIRE(ld_s >= 6 && ld_t >= (field ? 3 : 6), "row stride too small (sources need 6 columns, targets 6 / 3)");
    DNP_REQUIRE(ld_o >= (field ? 3 * S : S), "output row stride too small");
    return run_xie_matrix(field, src, S, ld_s, tgt, M, ld_t, C, rad2, out, ld_o, (cudaStream_t)stream);
}

int dnp_xie_interaction_f32(const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t,
                            float C, const double* src_radius2_opt, float* inter, int64_t ld_o, void* stream) {
    return xie_matrix_abi(false, src, S, ld_s, tgt, M, ld_t, C, src_radius2_opt, inter, ld_o, stream);
}

int dnp_xie_field_f32(const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t,
                      float C, const double* src_radius2_opt, float* field, int64_t ld_o, void* stream) {
    return xie_matrix_abi(true, src, S, ld_s, tgt, M, ld_t, C, src_radius2_opt, field, ld_o, stream);
}

int64_t dnp_xie_max_n(int knn) { return xie_max_n(knn != 0); }

int dnp_knn_radius2_f64(const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t, int k,
                        double* radius2, void* stream) {
    DNP_REQUIRE(S >= 0 && M >= 0, "negative size");
    if (S == 0) return DNP_OK;
    DNP_REQUIRE(src && tgt && radius2, "null pointer");
    DNP_REQUIRE(ld_s >= 3 && ld_t >= 3, "row stride too small");
    return run_knn_radius2(src, S, ld_s, tgt, M, ld_t, k, radius2, (cudaStream_t)stream);
}

size_t dnp_xie_workspace_bytes(int64_t N, int64_t T) { return xie_workspace_bytes(N, T); }

int dnp_xie_propagate_in_order_f32(const float* pts, int64_t N, int64_t ld, const int64_t* order, int64_t T,
                                   float C, const double* radius2_opt, int diffuse, uint8_t* flips,
                                   float* interactions, void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(N >= 0 && T >= 0, "negative size");
    if (N == 0 || T == 0) return DNP_OK;
    DNP_REQUIRE(pts && order && flips && interactions, "null pointer");
    DNP_REQUIRE(ld >= 6, "points need 6 columns");
    return run_xie_propagate(pts, N, ld, order, T, C, radius2_opt, diffuse, flips, interactions, workspace,
                             workspace_bytes, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ voxel partition (dnp_partition.cu)
size_t dnp_voxel_partition_workspace_bytes(int64_t N) { return voxel_partition_workspace_bytes(N); }

int dnp_voxel_partition_f32(const float* pts, int64_t N, int64_t ld, int n_part, double range_lo, double range_hi,
                            int64_t* perm, int64_t* offsets, int32_t* ijk, int64_t* counts, void* workspace,
                            size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(N >= 0, "negative size");
    DNP_REQUIRE(offsets && counts, "null pointer");
    DNP_REQUIRE(N == 0 || (pts && perm && ijk), "null pointer");
    DNP_REQUIRE(ld >= 3, "points need 3 columns");
    DNP_REQUIRE(range_hi > range_lo, "empty range");
    return run_voxel_partition(pts, N, ld, n_part, range_lo, range_hi, perm, offsets, ijk, counts, workspace,
                               workspace_bytes, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ patch-graph flip solve (dnp_flip.cu)
size_t dnp_flip_solve_workspace_bytes(int64_t n, int exact_max_n, int replicas) {
    return flip_solve_workspace_bytes(n, exact_max_n, replicas);
}

int dnp_flip_solve_f64(const double* A, const double* B, int64_t n, int64_t ld, int exact_max_n, int replicas,
                       int sweeps, uint64_t seed, uint8_t* x_out, double* obj_out, int32_t* status_out,
                       void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(n >= 1, "flip solve needs at least one patch");
    DNP_REQUIRE(A && B && x_out && obj_out && status_out, "null pointer");
    DNP_REQUIRE(ld >= n, "row stride smaller than n");
    return run_flip_solve(A, B, n, ld, exact_max_n, replicas, sweeps, seed, x_out, obj_out, status_out, workspace,
                          workspace_bytes, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ text IO + wire format (dnp_textio.cu)
size_t dnp_xyz_parse_workspace_bytes(int64_t n_bytes, int64_t max_lines) {
    return xyz_parse_workspace_bytes(n_bytes, max_lines);
}

int dnp_xyz_parse_f32(const char* txt, int64_t n_bytes, int64_t max_lines, float* out, int64_t max_rows,
                      int64_t* counts, void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(n_bytes >= 0 && max_lines >= 1 && max_rows >= 0, "bad size");
    DNP_REQUIRE(counts && (n_bytes == 0 || txt) && (max_rows == 0 || out), "null pointer");
    return run_xyz_parse(txt, n_bytes, max_lines, out, max_rows, counts, workspace, workspace_bytes,
                         (cudaStream_t)stream);
}

size_t dnp_xyz_format_workspace_bytes(int64_t n_values) { return xyz_format_workspace_bytes(n_values); }

int dnp_xyz_format_f32(const float* pts, int64_t n_rows, int cols, int64_t ld_row, int64_t ld_col, char* out,
                       int64_t capacity, int64_t* n_bytes_out, void* workspace, size_t workspace_bytes, void* stream) {
    DNP_REQUIRE(n_rows >= 0 && cols >= 1 && capacity >= 0, "bad size");
    DNP_REQUIRE(n_bytes_out && (n_rows == 0 || (pts && out)), "null pointer");
    return run_xyz_format(pts, n_rows, cols, ld_row, ld_col, out, capacity, n_bytes_out, workspace, workspace_bytes,
                          (cudaStream_t)stream);
}

// util.export_pc to an open file: format on the device, then the text goes through two pinned 4 MB buffers straight
// into write(2) -- the copy of chunk k + 1 runs while chunk k is written, and no host buffer of the size of the file
// is ever allocated (its page faults cost more than the whole device side).
namespace {
constexpr size_t kExportChunk = 4u << 20;
inline size_t up256z(size_t v) { return (v + 255) & ~(size_t)255; }
}  // namespace

size_t dnp_xyz_export_workspace_bytes(int64_t n_values) {
    const size_t fmt = xyz_format_workspace_bytes(n_values);
    if (fmt == 0) return 0;
    return up256z(fmt) + up256z((size_t)25 * (size_t)n_values) + 256;
}

int dnp_xyz_export_fd_f32(const float* pts, int64_t n_rows, int cols, int64_t ld_row, int64_t ld_col, int fd,
                          int64_t* n_bytes_host, void* workspace, size_t workspace_bytes, void* stream_) {
    DNP_REQUIRE(n_rows >= 0 && cols >= 1, "bad size");
    DNP_REQUIRE(fd >= 0, "bad file descriptor");
    if (n_bytes_host) *n_bytes_host = 0;
    if (n_rows == 0) return DNP_OK;
    DNP_REQUIRE(pts, "null pointer");
    const int64_t n_values = n_rows * cols;
    const size_t need = dnp_xyz_export_workspace_bytes(n_values);
    DNP_REQUIRE(need != 0, "xyz export: too many values for one call");
    if (!workspace || workspace_bytes < need) {
        set_error("workspace too small: need %zu bytes (dnp_xyz_export_workspace_bytes), got %zu", need, workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t fmt = up256z(xyz_format_workspace_bytes(n_values));
    const size_t cap = (size_t)25 * (size_t)n_values;
    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    char* text = reinterpret_cast<char*>(ws + fmt);
    int64_t* count_dev = reinterpret_cast<int64_t*>(ws + fmt + up256z(cap));
    int rc = run_xyz_format(pts, n_rows, cols, ld_row, ld_col, text, (int64_t)cap, count_dev, ws, fmt, stream);
    if (rc != DNP_OK) return rc;
    int device = 0;
    DNP_CHECK_CUDA(cudaGetDevice(&device));
    cudaError_t e = t_host.prepare(device, 0, 2 * kExportChunk);
    if (e != cudaSuccess) return cuda_fail(e, "pinned staging");
    unsigned char* pin = t_host.pin;
    DNP_CHECK_CUDA(cudaMemcpyAsync(pin, count_dev, sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
    DNP_CHECK_CUDA(cudaStreamSynchronize(stream));
    int64_t total = 0;
    memcpy(&total, pin, sizeof(total));
    if (total < 0 || (size_t)total > cap) {
        set_error("xyz export: formatted size %lld outside the buffer", (long long)total);
        return DNP_ERR_INVALID;
    }
    cudaEvent_t ev[2] = {nullptr, nullptr};
    for (int k = 0; k < 2; ++k)
        if ((e = cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming)) != cudaSuccess) {
            if (ev[0]) cudaEventDestroy(ev[0]);
            return cuda_fail(e, "cudaEventCreate");
        }
    const int64_t n_chunks = (total + (int64_t)kExportChunk - 1) / (int64_t)kExportChunk;
    auto chunk_len = [&](int64_t k) {
        const int64_t left = total - k * (int64_t)kExportChunk;
        return (size_t)(left < (int64_t)kExportChunk ? left : (int64_t)kExportChunk);
    };
    auto issue = [&](int64_t k) {
        cudaError_t err = cudaMemcpyAsync(pin + (k & 1) * kExportChunk, text + k * (int64_t)kExportChunk, chunk_len(k),
                                          cudaMemcpyDeviceToHost, stream);
        if (err == cudaSuccess) err = cudaEventRecord(ev[k & 1], stream);
        return err;
    };
    rc = DNP_OK;
    if (n_chunks > 0 && (e = issue(0)) != cudaSuccess) rc = cuda_fail(e, "download");
    for (int64_t k = 0; k < n_chunks && rc == DNP_OK; ++k) {
        if (k + 1 < n_chunks && (e = issue(k + 1)) != cudaSuccess) { rc = cuda_fail(e, "download"); break; }
        if ((e = cudaEventSynchronize(ev[k & 1])) != cudaSuccess) { rc = cuda_fail(e, "cudaEventSynchronize"); break; }
        const unsigned char* p = pin + (k & 1) * kExportChunk;
        size_t left = chunk_len(k);
        while (left > 0) {
            const ssize_t w = write(fd, p, left);
            if (w < 0) {
                if (errno == EINTR) continue;
                set_error("xyz export: write failed: %s", strerror(errno));
                rc = DNP_ERR_IO;
                break;
            }
            p += w;
            left -= (size_t)w;
        }
    }
    if (rc != DNP_OK) cudaStreamSynchronize(stream);          // no copy may still target the staging buffers
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
    if (rc == DNP_OK && n_bytes_host) *n_bytes_host = total;
    return rc;
}

int dnp_wire_decode_f64(const double* xyz, int64_t n, float* out, int64_t ld, void* stream) {
    DNP_REQUIRE(n >= 0 && ld >= 6, "bad size");
    DNP_REQUIRE(n == 0 || (xyz && out), "null pointer");
    return run_wire_decode(xyz, n, out, ld, (cudaStream_t)stream);
}

int dnp_wire_encode_f64(const float* pts, int64_t n, int64_t ld, double* out, void* stream) {
    DNP_REQUIRE(n >= 0 && ld >= 6, "bad size");
    DNP_REQUIRE(n == 0 || (pts && out), "null pointer");
    return run_wire_encode(pts, n, ld, out, (cudaStream_t)stream);
}

int dnp_profile_begin(void) {
    profile_begin();
    return DNP_OK;
}

int dnp_profile_end(float* kernel_ms_host, int capacity, int* n_kernels) {
    DNP_REQUIRE(capacity >= 0, "negative capacity");
    return profile_end(kernel_ms_host, capacity, n_kernels);
}

int dnp_microbench(int kind, int iters, float* ms_out, double* ops_out, void* stream) {
    return run_microbench(kind, iters, ms_out, ops_out, (cudaStream_t)stream);
}

}  // extern "C"
