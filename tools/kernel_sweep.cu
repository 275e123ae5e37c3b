// kernel_sweep.cu -- development harness (not part of libdipole): times variants of the by-value field
// kernel (targets per thread, unroll, residency, CTA size, arithmetic order) and register-file operand
// probes on exactly full waves.  Build: make -C tools sweep ; run on the GPU box: tools/bin/kernel_sweep
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#include "../dipole-normal-prop_b200/csrc/dnp_common.cuh"

namespace dnp {
void set_error(const char*, ...) {}
int cuda_fail(cudaError_t, const char*) { return -2; }
}  // namespace dnp
using namespace dnp;

constexpr int kChunk = 1280;
struct SrcBlock { float v[kChunk * 6]; };
constexpr float kClamp = 1.152921504606847e18f;

// VAR 0: library order.  VAR 1: b-accumulation right after c (c reuse), a-accumulation grouped.
// VAR 2: tiny-offset r2 (no clamp FMNMX), select w by compare.  VAR 3: scalar (non-packed) arithmetic.
template <int TP, int UNROLL, int BLOCK, int MINB, int VAR>
__global__ void __launch_bounds__(BLOCK, MINB)
field_kernel(const __grid_constant__ SrcBlock blk, int n_src, const float* __restrict__ tgt, int64_t M, float eps,
             float* __restrict__ out) {
    const int tid = threadIdx.x;
    const int64_t tile_base = (int64_t)blockIdx.x * (BLOCK * 2 * TP);
    f32x2 qx[TP], qy[TP], qz[TP];
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float x[2], y[2], z[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t m = (tile_base + (int64_t)(2 * j + h) * BLOCK + tid) % M;
            x[h] = tgt[3 * m]; y[h] = tgt[3 * m + 1]; z[h] = tgt[3 * m + 2];
        }
        qx[j] = pack2(x[0], x[1]); qy[j] = pack2(y[0], y[1]); qz[j] = pack2(z[0], z[1]);
    }
    const f32x2 zero2 = pack2(0.f, 0.f), eps2 = pack2(eps, eps), m3 = pack2(-3.f, -3.f);
    __shared__ float4 sm[(VAR == 7) ? 768 * 3 : 1];
    if (VAR == 7) {
        for (int i = tid; i < 768; i += BLOCK) {
            sm[3 * i] = make_float4(blk.v[9 * i], blk.v[9 * i + 1], blk.v[9 * i + 2], blk.v[9 * i + 3]);
            sm[3 * i + 1] = make_float4(blk.v[9 * i + 4], blk.v[9 * i + 5], blk.v[9 * i + 6], blk.v[9 * i + 7]);
            sm[3 * i + 2] = make_float4(blk.v[9 * i + 8], 0.f, 0.f, 0.f);
        }
        __syncthreads();
    }
    f32x2 tx[TP], ty[TP], tz[TP];
#pragma unroll
    for (int j = 0; j < TP; ++j) { tx[j] = zero2; ty[j] = zero2; tz[j] = zero2; }
    for (int s0 = 0; s0 < n_src; s0 += 256) {
        f32x2 ax[TP], ay[TP], az[TP], bx[TP], by[TP], bz[TP];
#pragma unroll
        for (int j = 0; j < TP; ++j) { ax[j] = ay[j] = az[j] = bx[j] = by[j] = bz[j] = zero2; }
#pragma unroll UNROLL
        for (int s = s0; s < s0 + 256; ++s) {
            constexpr int ST = (VAR == 5) ? 9 : 6;
            float px, py, pz, nx, ny, nz;
            float4 A, B, C;
            if (VAR == 7) {
                A = sm[3 * s]; B = sm[3 * s + 1]; C = sm[3 * s + 2];
                px = A.x; py = A.y; pz = A.z; nx = A.w; ny = B.x; nz = B.y;
            } else {
                px = blk.v[ST * s]; py = blk.v[ST * s + 1]; pz = blk.v[ST * s + 2];
                nx = blk.v[ST * s + 3]; ny = blk.v[ST * s + 4]; nz = blk.v[ST * s + 5];
            }
            const f32x2 px2 = pack2(px, px), py2 = pack2(py, py), pz2 = pack2(pz, pz);
            const f32x2 nx2 = pack2(nx, nx), ny2 = pack2(ny, ny), nz2 = pack2(nz, nz);
            float mx = nx, my = ny, mz = nz;
            if (VAR == 5) { mx = blk.v[ST * s + 6]; my = blk.v[ST * s + 7]; mz = blk.v[ST * s + 8]; }
            if (VAR == 6) { mx = nx * -3.f; my = ny * -3.f; mz = nz * -3.f; }
            if (VAR == 7) { mx = B.z; my = B.w; mz = C.x; }
            const f32x2 mx2 = pack2(mx, mx), my2 = pack2(my, my), mz2 = pack2(mz, mz);
#pragma unroll
            for (int j = 0; j < TP; ++j) {
                const f32x2 rx = sub2(px2, qx[j]), ry = sub2(py2, qy[j]), rz = sub2(pz2, qz[j]);
                if (VAR >= 3) {
                    // 19-op form: E += w * (n3 - t R) with t = (n.R) u^2 (x3 folded at the tile fold)
                    const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, mul2(rx, rx)));
                    float r2a, r2b;
                    unpack2(r2, r2a, r2b);
                    const float ua = rsqrt_sfu(r2a), ub = rsqrt_sfu(r2b);
                    const f32x2 u = pack2(ua, ub);
                    const f32x2 uc = pack2(fminf(ua, kClamp), fminf(ub, kClamp));
                    f32x2 den;
                    if (VAR >= 4) den = fma2(mul2(r2, r2), u, eps2);
                    else den = fma2(r2, mul2(r2, u), eps2);
                    float da, db;
                    unpack2(den, da, db);
                    const f32x2 w = pack2(fmaxf(rcp_sfu(da), 0.f), fmaxf(rcp_sfu(db), 0.f));
                    const f32x2 d = fma2(mz2, rz, fma2(my2, ry, mul2(mx2, rx)));
                    const f32x2 t = mul2(d, mul2(uc, uc));
                    // blk normals hold -n/3 in this variant: v = t R - n/3, E_tile += w v, total = -3 E_tile
                    const f32x2 vx = fma2(t, rx, nx2), vy = fma2(t, ry, ny2), vz = fma2(t, rz, nz2);
                    ax[j] = fma2(w, vx, ax[j]); ay[j] = fma2(w, vy, ay[j]); az[j] = fma2(w, vz, az[j]);
                } else if (VAR == 2) {
                    const f32x2 tiny2 = pack2(1e-30f, 1e-30f);
                    const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, fma2(rx, rx, tiny2)));
                    float r2a, r2b;
                    unpack2(r2, r2a, r2b);
                    const f32x2 u = pack2(rsqrt_sfu(r2a), rsqrt_sfu(r2b));
                    const f32x2 r = mul2(r2, u);
                    const f32x2 den = fma2(r2, r, eps2);
                    float da, db;
                    unpack2(den, da, db);
                    const f32x2 w = pack2(r2a > 2e-30f ? rcp_sfu(da) : 0.f, r2b > 2e-30f ? rcp_sfu(db) : 0.f);
                    const f32x2 d = fma2(nz2, rz, fma2(ny2, ry, mul2(nx2, rx)));
                    const f32x2 c = mul2(mul2(d, w), mul2(u, u));
                    ax[j] = fma2(w, nx2, ax[j]); ay[j] = fma2(w, ny2, ay[j]); az[j] = fma2(w, nz2, az[j]);
                    bx[j] = fma2(c, rx, bx[j]);  by[j] = fma2(c, ry, by[j]);  bz[j] = fma2(c, rz, bz[j]);
                } else {
                    const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, mul2(rx, rx)));
                    float r2a, r2b;
                    unpack2(r2, r2a, r2b);
                    const float ua = rsqrt_sfu(r2a), ub = rsqrt_sfu(r2b);
                    const f32x2 u = pack2(ua, ub);
                    const f32x2 uc = pack2(fminf(ua, kClamp), fminf(ub, kClamp));
                    const f32x2 r = mul2(r2, u);
                    const f32x2 den = fma2(r2, r, eps2);
                    float da, db;
                    unpack2(den, da, db);
                    const f32x2 w = pack2(fmaxf(rcp_sfu(da), 0.f), fmaxf(rcp_sfu(db), 0.f));
                    if (VAR == 1) {
                        // d as a running product with the clamped u^2 first, so c follows d immediately
                        const f32x2 g = mul2(w, mul2(uc, uc));
                        const f32x2 d = fma2(nz2, rz, fma2(ny2, ry, mul2(nx2, rx)));
                        const f32x2 c = mul2(d, g);
                        bx[j] = fma2(c, rx, bx[j]);  by[j] = fma2(c, ry, by[j]);  bz[j] = fma2(c, rz, bz[j]);
                        ax[j] = fma2(w, nx2, ax[j]); ay[j] = fma2(w, ny2, ay[j]); az[j] = fma2(w, nz2, az[j]);
                    } else {
                        const f32x2 d = fma2(nz2, rz, fma2(ny2, ry, mul2(nx2, rx)));
                        const f32x2 c = mul2(mul2(d, w), mul2(uc, uc));
                        ax[j] = fma2(w, nx2, ax[j]); ay[j] = fma2(w, ny2, ay[j]); az[j] = fma2(w, nz2, az[j]);
                        bx[j] = fma2(c, rx, bx[j]);  by[j] = fma2(c, ry, by[j]);  bz[j] = fma2(c, rz, bz[j]);
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < TP; ++j) {
            tx[j] = add2(tx[j], fma2(bx[j], m3, ax[j]));
            ty[j] = add2(ty[j], fma2(by[j], m3, ay[j]));
            tz[j] = add2(tz[j], fma2(bz[j], m3, az[j]));
        }
    }
#pragma unroll
    for (int j = 0; j < TP; ++j) {
        float ex[2], ey[2], ez[2];
        unpack2(tx[j], ex[0], ex[1]); unpack2(ty[j], ey[0], ey[1]); unpack2(tz[j], ez[0], ez[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t m = (tile_base + (int64_t)(2 * j + h) * BLOCK + tid) % M;
            out[3 * m] = ex[h]; out[3 * m + 1] = ey[h]; out[3 * m + 2] = ez[h];
        }
    }
}

// ---------------------------------------------------------------- register-file operand probes (FFMA2, 8 chains)
template <int KIND>
__global__ void __launch_bounds__(256) rf_probe(int iters, float a, float b, float* sink) {
    constexpr int C = 8;
    f32x2 x[C], y[C], z[C];
    const f32x2 a2 = pack2(a, a), b2 = pack2(b, b);
    const f32x2 w = pack2(a + 1e-3f * threadIdx.x, a - 1e-3f * threadIdx.x);
#pragma unroll
    for (int c = 0; c < C; ++c) {
        x[c] = pack2((float)(threadIdx.x + c), (float)c);
        y[c] = pack2(a + 1e-3f * (float)(threadIdx.x + c), a - 1e-3f * (float)c);
        z[c] = pack2(b + 1e-4f * (float)(threadIdx.x ^ c), b);
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < C; ++c) {
                if (KIND == 0) x[c] = fma2(x[c], a2, b2);
                if (KIND == 1) x[c] = fma2(y[c], a2, x[c]);
                if (KIND == 2) x[c] = fma2(y[c], y[c], x[c]);
                if (KIND == 3) x[c] = fma2(y[c], z[c], x[c]);
                if (KIND == 4) x[c] = fma2(w, y[c], x[c]);
                if (KIND == 5) x[c] = fma2(x[c], x[c], b2);
                if (KIND == 6) x[c] = fma2(y[c], z[c], b2) , y[c] = x[c];
            }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < C; ++c) { float lo, hi; unpack2(x[c], lo, hi); s += lo + hi; unpack2(y[c], lo, hi); s += lo; }
    if (s == 12345.678f) sink[0] = s;
}

static int g_sms = 148;
static float* d_tgt; static float* d_out; static float* d_sink;
static SrcBlock h_blk;
static const int64_t kM = 1 << 22;

template <int TP, int UNROLL, int BLOCK, int MINB, int VAR>
void run_field(const char* name) {
    auto k = field_kernel<TP, UNROLL, BLOCK, MINB, VAR>;
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k);
    int resident = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, k, BLOCK, 0);
    const int waves = 4;
    const int grid = g_sms * resident * waves;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int reps = 6;
    const int n_src = (VAR == 5 || VAR == 7) ? 768 : kChunk;
    k<<<grid, BLOCK>>>(h_blk, n_src, d_tgt, kM, 1e-5f, d_out);
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) k<<<grid, BLOCK>>>(h_blk, n_src, d_tgt, kM, 1e-5f, d_out);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double pairs = (double)reps * grid * BLOCK * 2 * TP * n_src;
    const double pps = pairs / (ms * 1e-3);
    printf("{\"kind\":\"field\",\"name\":\"%s\",\"tp\":%d,\"unroll\":%d,\"block\":%d,\"minb\":%d,\"var\":%d,\"regs\":%d,"
           "\"resident\":%d,\"pairs_per_s\":%.4e,\"fma_pipe_frac\":%.4f,\"err\":\"%s\"}\n",
           name, TP, UNROLL, BLOCK, MINB, VAR, fa.numRegs, resident, pps, pps * 20 / (g_sms * 128 * 1.965e9),
           cudaGetErrorString(cudaGetLastError()));
    fflush(stdout);
}

template <int KIND>
void run_probe(const char* name) {
    const int iters = 4000, blocks = g_sms * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    rf_probe<KIND><<<blocks, 256>>>(iters, 0.999f, 0.001f, d_sink);
    cudaEventRecord(e0);
    rf_probe<KIND><<<blocks, 256>>>(iters, 0.999f, 0.001f, d_sink);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double inst = (double)blocks * 256 / 32 * iters * 64.0;          // warp-level FFMA2 instructions
    const double cyc_per_inst = (ms * 1e-3 * 1.965e9) * (g_sms * 4) / inst;  // per SM sub-partition
    printf("{\"kind\":\"rf_probe\",\"name\":\"%s\",\"ms\":%.3f,\"cycles_per_ffma2_per_smsp\":%.3f}\n", name, ms, cyc_per_inst);
    fflush(stdout);
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0); g_sms = prop.multiProcessorCount;
    std::vector<float> t(kM * 3);
    srand(1);
    for (auto& v : t) v = rand() / (float)RAND_MAX - 0.5f;
    for (int i = 0; i < kChunk; ++i) {
        for (int c = 0; c < 3; ++c) h_blk.v[6 * i + c] = rand() / (float)RAND_MAX - 0.5f;
        h_blk.v[6 * i + 3] = 0.6f; h_blk.v[6 * i + 4] = 0.0f; h_blk.v[6 * i + 5] = 0.8f;
    }
    cudaMalloc(&d_tgt, kM * 12); cudaMalloc(&d_out, kM * 12); cudaMalloc(&d_sink, 4);
    cudaMemcpy(d_tgt, t.data(), kM * 12, cudaMemcpyHostToDevice);
    run_probe<0>("x=x*U+U (1 reg pair)");
    run_probe<1>("x=y*U+x (2 pairs)");
    run_probe<2>("x=y*y+x (2 pairs, one twice)");
    run_probe<3>("x=y*z+x (3 pairs)");
    run_probe<4>("x=w*y+x (3 pairs, w shared by consecutive)");
    run_probe<5>("x=x*x+U (1 pair twice)");
    run_probe<6>("x=y*z+U (2 pairs, fresh dst)");
#define F(TP, U, B, MB, V) run_field<TP, U, B, MB, V>(#TP "," #U "," #B "," #MB "," #V)
    F(4, 2, 128, 3, 0);
    F(4, 4, 128, 3, 5);
    F(4, 2, 128, 3, 7);
    F(4, 4, 128, 3, 7);
    F(2, 2, 128, 5, 7);
    F(2, 4, 128, 4, 7);
    F(1, 4, 128, 8, 7);
    F(1, 2, 128, 8, 7);
    F(2, 4, 256, 2, 7);
    return 0;
}
