// dnp_pairmath.cuh -- the per-pair arithmetic of the dipole field / potential on packed f32x2 operands, shared by the
// all-pairs kernel (dnp_pairs.cu) and the persistent patch loop (dnp_loop.cu).  field_utils.py:96-116 / :46-55.
#pragma once

#include "dnp_common.cuh"

namespace dnp {

constexpr float kRsqrtClamp = 1.152921504606847e18f;   // 2^60: clamp^2 is finite (field: t = d u^2)
// Issue-slot economy (profiles/README.md r02): on B200 a packed FFMA2 / FMUL2 / FADD2 costs TWO issue cycles and every
// other instruction one, and the all-pairs loops are issue-bound (field: 2 x 19 + 4 MUFU + 4 FMNMX + LDCU = 47 cycles
// per pair of pairs, measured).  Two of the four FMNMX were only there to turn NaN into 0 on a coincident pair; the
// two constants below get the same exact zero out of the FMA pipe for free.
// field: |R|^4 is formed as r2 * r2 + kTinyR4, so that at R = 0 (u = rsqrt(0) = +inf) den = kTinyR4 * inf + eps = +inf
// and w = rcp(+inf) = +0 exactly, with no 0 * inf NaN to launder.  1e-37 is a normal float; it changes r^4 by less
// than 1 ulp whenever |R| > 2e-8.
constexpr float kTinyR4 = 1.0e-37f;
// potential: r2 is formed as ... + rx * rx + kTinyR2, so that u = rsqrt(r2) stays finite (<= 1e12, u^3 <= 1e36) and a
// coincident pair adds (n.R) u^3 = 0 * finite = 0 exactly.  1e-24 changes r2 by less than 1 ulp whenever |R| > 4e-9
// (two distinct float32 coordinates of a unit-box cloud are at least 6e-8 apart away from the origin).
constexpr float kTinyR2 = 1.0e-24f;

// ------------------------------------------------------------------ per-pair arithmetic, packed
// one source, broadcast to both halves of every packed operation (ptxas emits the `R.F32` operand form)
struct SrcX2 {
    f32x2 px, py, pz, nx, ny, nz, mx, my, mz;
    __device__ __forceinline__ SrcX2(const float4& A, const float4& B, const float4& C)
        : px(pack2(A.x, A.x)), py(pack2(A.y, A.y)), pz(pack2(A.z, A.z)), nx(pack2(A.w, A.w)),
          ny(pack2(B.x, B.x)), nz(pack2(B.y, B.y)), mx(pack2(B.z, B.z)), my(pack2(B.w, B.w)),
          mz(pack2(C.x, C.x)) {}
    // two DIFFERENT sources, one per half (the persistent loop pairs sources, not targets)
    __device__ __forceinline__ SrcX2(const float4& A0, const float4& B0, const float4& C0, const float4& A1,
                                     const float4& B1, const float4& C1)
        : px(pack2(A0.x, A1.x)), py(pack2(A0.y, A1.y)), pz(pack2(A0.z, A1.z)), nx(pack2(A0.w, A1.w)),
          ny(pack2(B0.x, B1.x)), nz(pack2(B0.y, B1.y)), mx(pack2(B0.z, B1.z)), my(pack2(B0.w, B1.w)),
          mz(pack2(C0.x, C1.x)) {}
};

// field: E += w (n + t R) with w = 1/(r^3+eps), t = (m.R) / r^2, m = -3 n            (19 packed FP32 ops)
// kSafe = false leaves out the clamp of u: a coincident pair then makes t = 0 * inf = NaN, which poisons the caller's
// partial sums -- the caller checks them (sums_nonfinite) and re-runs the same sources with kSafe = true, where the
// pair adds exactly 0.  Duplicates of a target among the sources are rare; the two FMNMX are paid on every pair.
template <bool kSafe = true>
__device__ __forceinline__ void pair_field_x2(const SrcX2& s, f32x2 qx, f32x2 qy, f32x2 qz, f32x2 eps2,
                                              f32x2& ax, f32x2& ay, f32x2& az) {
    const f32x2 rx = sub2(s.px, qx), ry = sub2(s.py, qy), rz = sub2(s.pz, qz);
    const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, mul2(rx, rx)));
    float r2a, r2b;
    unpack2(r2, r2a, r2b);
    const float ua = rsqrt_sfu(r2a), ub = rsqrt_sfu(r2b);                 // +inf on a coincident pair
    const f32x2 u = pack2(ua, ub);
    const f32x2 uc = kSafe ? pack2(fminf(ua, kRsqrtClamp), fminf(ub, kRsqrtClamp)) : u;
    const f32x2 tiny4 = pack2(kTinyR4, kTinyR4);
    const f32x2 den = fma2(fma2(r2, r2, tiny4), u, eps2);                 // |R|^3 + eps; tiny * inf = +inf if R = 0
    float da, db;
    unpack2(den, da, db);
    const f32x2 w = pack2(rcp_sfu(da), rcp_sfu(db));                      // rcp(+inf) = +0: a coincident pair adds 0
    const f32x2 d = fma2(s.mz, rz, fma2(s.my, ry, mul2(s.mx, rx)));       // -3 n.R (exactly 0 when R = 0)
    const f32x2 t = mul2(d, mul2(uc, uc));
    const f32x2 vx = fma2(t, rx, s.nx), vy = fma2(t, ry, s.ny), vz = fma2(t, rz, s.nz);
    ax = fma2(w, vx, ax); ay = fma2(w, vy, ay); az = fma2(w, vz, az);
}

// true if any of the three partial sums is NaN or Inf (x * 0 is NaN for both)
__device__ __forceinline__ bool sums_nonfinite(f32x2 ax, f32x2 ay, f32x2 az) {
    float a, b, chk = 0.f;
    unpack2(ax, a, b); chk = fmaf(a, 0.f, chk); chk = fmaf(b, 0.f, chk);
    unpack2(ay, a, b); chk = fmaf(a, 0.f, chk); chk = fmaf(b, 0.f, chk);
    unpack2(az, a, b); chk = fmaf(a, 0.f, chk); chk = fmaf(b, 0.f, chk);
    return chk != 0.f;
}

// potential: phi += (n.R) / r^3                                         (12 packed FP32 ops)
__device__ __forceinline__ void pair_potential_x2(const SrcX2& s, f32x2 qx, f32x2 qy, f32x2 qz, f32x2& phi) {
    const f32x2 rx = sub2(s.px, qx), ry = sub2(s.py, qy), rz = sub2(s.pz, qz);
    const f32x2 tiny2 = pack2(kTinyR2, kTinyR2);
    const f32x2 r2 = fma2(rz, rz, fma2(ry, ry, fma2(rx, rx, tiny2)));
    float r2a, r2b;
    unpack2(r2, r2a, r2b);
    // coincident pair: n.R is exactly 0 and rsqrt(kTinyR2)^3 is finite, so the pair adds 0
    const f32x2 u = pack2(rsqrt_sfu(r2a), rsqrt_sfu(r2b));
    const f32x2 d = fma2(s.nz, rz, fma2(s.ny, ry, mul2(s.nx, rx)));
    phi = fma2(d, mul2(mul2(u, u), u), phi);
}

}  // namespace dnp
