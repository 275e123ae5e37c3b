// dnp_common.cuh -- shared device/host helpers for libdipole (sm_100a only).
//
// Dipole conventions follow the reference (field_utils.py:47,97): a point is a
// row [x y z nx ny nz], R = source - target.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libdipole is written for sm_100a (B200) only"
#endif

namespace dnp {

// ---------------------------------------------------------------- error plumbing (dnp_abi.cu)
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t err, const char* what);

#define DNP_CHECK_CUDA(expr)                                   \
    do {                                                       \
        cudaError_t _e = (expr);                               \
        if (_e != cudaSuccess) return ::dnp::cuda_fail(_e, #expr); \
    } while (0)

// ---------------------------------------------------------------- per-device one-time setup
// cudaFuncSetAttribute applies to the CURRENT device only: a process that calls on cuda:0 and then on cuda:1 has
// to opt in on both.  `flags` is a function-local `static std::atomic<int> flags[kMaxDevices]`.
constexpr int kMaxDevices = 64;
#define DNP_ONCE_PER_DEVICE(flags, body)                                             \
    do {                                                                             \
        int _dev = 0;                                                                \
        DNP_CHECK_CUDA(cudaGetDevice(&_dev));                                        \
        const bool _tracked = _dev >= 0 && _dev < ::dnp::kMaxDevices;                \
        if (!_tracked || !(flags)[_dev].load()) {                                    \
            body;                                                                    \
            if (_tracked) (flags)[_dev].store(1);                                    \
        }                                                                            \
    } while (0)

// ---------------------------------------------------------------- packed f32x2 arithmetic
// Blackwell executes two independent FP32 operations per lane with one
// FFMA2/FADD2/FMUL2 instruction on a 64-bit register pair.  The hot loops keep
// the FP32 pipe busy with half the issue slots, leaving the rest for MUFU,
// FMNMX and the shared-memory broadcasts.
typedef unsigned long long f32x2;

__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("sub.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// SFU approximations (MUFU.RSQ / MUFU.RCP): <= 2 ulp, far inside the stated
// 1e-4 relative parity tolerance.
__device__ __forceinline__ float rsqrt_sfu(float x) {
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_sfu(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---------------------------------------------------------------- mbarrier + TMA 1-D bulk copy
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy executed by the TMA unit (SASS: UBLKCP); completion
// is signalled on `bar` as a transaction-byte count.
__device__ __forceinline__ void tma_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                             uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ---------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace dnp
