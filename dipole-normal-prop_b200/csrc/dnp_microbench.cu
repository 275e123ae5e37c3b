// dnp_microbench.cu -- issue-rate microbenchmarks for the roofline denominators.
//
// SURVEY.md section 8(d) asks for the FP32-FMA and SFU peaks to be measured on
// the box rather than taken from the data sheet.  Each kernel keeps 8
// independent dependency chains per thread so the pipes, not latencies, bound it.
//   kind 0: scalar FFMA            ops = FMA lane-operations
//   kind 1: packed FFMA2 (f32x2)   ops = FMA lane-operations (2 per instruction per lane)
//   kind 2: MUFU.RSQ               ops = SFU lane-operations
//   kind 3: the field kernel's mix per 2 pairs: 20 FFMA2 : 4 MUFU : 4 FMNMX   ops = pairs
//   kind 4: FFMA2 with three distinct register operands (register-file bandwidth probe)
//   kind 5: scalar FFMA with three distinct register operands
//   kind 6: FFMA2 with two fresh register operands + one shared by consecutive instructions
#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"

namespace dnp {

namespace {

constexpr int kMbThreads = 256;
constexpr int kChains = 8;

__global__ void __launch_bounds__(kMbThreads) mb_ffma(int iters, float a, float b, float* sink) {
    float x[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = (float)(threadIdx.x + c);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = fmaf(x[c], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) sink[0] = s;
}

__global__ void __launch_bounds__(kMbThreads) mb_ffma2(int iters, float a, float b, float* sink) {
    f32x2 x[kChains];
    const f32x2 a2 = pack2(a, a), b2 = pack2(b, b);
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = pack2((float)(threadIdx.x + c), (float)c);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = fma2(x[c], a2, b2);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) { float lo, hi; unpack2(x[c], lo, hi); s += lo + hi; }
    if (s == 12345.678f) sink[0] = s;
}

__global__ void __launch_bounds__(kMbThreads) mb_mufu(int iters, float* sink) {
    float x[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = 1.5f + (float)(threadIdx.x + c);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = rsqrt_sfu(x[c]);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) sink[0] = s;
}

// per chain and iteration: 10 FFMA2 + 2 MUFU + 2 FMNMX = the field kernel's mix for one pair-of-pairs / 2
__global__ void __launch_bounds__(kMbThreads) mb_mix(int iters, float a, float b, float* sink) {
    f32x2 x[kChains];
    const f32x2 a2 = pack2(a, a), b2 = pack2(b, b);
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = pack2(1.f + (float)(threadIdx.x + c), 2.f + (float)c);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < kChains; ++c) {
            f32x2 v = x[c];
#pragma unroll
            for (int r = 0; r < 10; ++r) v = fma2(v, a2, b2);
            float lo, hi;
            unpack2(v, lo, hi);
            lo = fmaxf(rsqrt_sfu(lo), 0.25f);
            hi = fminf(rcp_sfu(hi), 4.0f);
            x[c] = pack2(lo, hi);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) { float lo, hi; unpack2(x[c], lo, hi); s += lo + hi; }
    if (s == 12345.678f) sink[0] = s;
}

// register-file bandwidth probes: the same FMA count as mb_ffma2 / mb_ffma, but every instruction reads
// THREE distinct register operands (no operand can come from the reuse cache, an immediate or a uniform register)
__global__ void __launch_bounds__(kMbThreads) mb_ffma2_3reg(int iters, float a, float b, float* sink) {
    f32x2 x[kChains], y[kChains], z[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
        x[c] = pack2((float)(threadIdx.x + c), (float)c);
        y[c] = pack2(a + 1e-3f * (float)(threadIdx.x + c), a - 1e-3f * (float)c);
        z[c] = pack2(b + 1e-4f * (float)(threadIdx.x ^ c), b);
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = fma2(x[c], y[(c + r) % kChains], z[(c + 3 * r + 1) % kChains]);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) { float lo, hi; unpack2(x[c], lo, hi); s += lo + hi; }
    if (s == 12345.678f) sink[0] = s;
}

__global__ void __launch_bounds__(kMbThreads) mb_ffma_3reg(int iters, float a, float b, float* sink) {
    float x[kChains], y[kChains], z[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
        x[c] = (float)(threadIdx.x + c);
        y[c] = a + 1e-3f * (float)(threadIdx.x + c);
        z[c] = b + 1e-4f * (float)(threadIdx.x ^ c);
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = fmaf(x[c], y[(c + r) % kChains], z[(c + 3 * r + 1) % kChains]);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) sink[0] = s;
}

// two register operands + one operand shared by consecutive instructions (reuse-cache candidate)
__global__ void __launch_bounds__(kMbThreads) mb_ffma2_shared(int iters, float a, float b, float* sink) {
    f32x2 x[kChains], y[kChains], w[8];
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
        x[c] = pack2((float)(threadIdx.x + c), (float)c);
        y[c] = pack2(b + 1e-4f * (float)(threadIdx.x ^ c), b);
        w[c] = pack2(a + 1e-3f * (float)(threadIdx.x + c), a - 1e-3f * (float)c);
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = fma2(w[r], y[(c + r) % kChains], x[c]);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) { float lo, hi; unpack2(x[c], lo, hi); s += lo + hi; }
    if (s == 12345.678f) sink[0] = s;
}

}  // namespace

int run_microbench(int kind, int iters, float* ms_out, double* ops_out, cudaStream_t stream) {
    if (iters < 1) iters = 1;
    const int blocks = sm_count_cached() * 8;
    float* sink = nullptr;
    DNP_CHECK_CUDA(cudaMalloc(&sink, sizeof(float)));
    cudaEvent_t e0, e1;
    DNP_CHECK_CUDA(cudaEventCreate(&e0));
    DNP_CHECK_CUDA(cudaEventCreate(&e1));
    const double lanes = (double)blocks * kMbThreads;
    double ops = 0.0;
    for (int rep = 0; rep < 2; ++rep) {      // rep 0 warms up
        DNP_CHECK_CUDA(cudaEventRecord(e0, stream));
        switch (kind) {
            case 0: mb_ffma<<<blocks, kMbThreads, 0, stream>>>(iters, 0.999f, 0.001f, sink);
                    ops = lanes * iters * 8.0 * kChains; break;
            case 1: mb_ffma2<<<blocks, kMbThreads, 0, stream>>>(iters, 0.999f, 0.001f, sink);
                    ops = lanes * iters * 8.0 * kChains * 2.0; break;
            case 2: mb_mufu<<<blocks, kMbThreads, 0, stream>>>(iters, sink);
                    ops = lanes * iters * 8.0 * kChains; break;
            case 3: mb_mix<<<blocks, kMbThreads, 0, stream>>>(iters, 0.999f, 0.001f, sink);
                    ops = lanes * iters * (double)kChains; break;   // one (source, target) pair per chain-iteration
            case 4: mb_ffma2_3reg<<<blocks, kMbThreads, 0, stream>>>(iters, 0.999f, 0.001f, sink);
                    ops = lanes * iters * 8.0 * kChains * 2.0; break;
            case 5: mb_ffma_3reg<<<blocks, kMbThreads, 0, stream>>>(iters, 0.999f, 0.001f, sink);
                    ops = lanes * iters * 8.0 * kChains; break;
            case 6: mb_ffma2_shared<<<blocks, kMbThreads, 0, stream>>>(iters, 0.999f, 0.001f, sink);
                    ops = lanes * iters * 8.0 * kChains * 2.0; break;
            default: set_error("unknown microbenchmark kind %d", kind); return DNP_ERR_INVALID;
        }
        count_launch(1);
        DNP_CHECK_CUDA(cudaEventRecord(e1, stream));
        DNP_CHECK_CUDA(cudaEventSynchronize(e1));
    }
    float ms = 0.f;
    DNP_CHECK_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    if (ms_out) *ms_out = ms;
    if (ops_out) *ops_out = ops;
    return DNP_OK;
}

}  // namespace dnp
