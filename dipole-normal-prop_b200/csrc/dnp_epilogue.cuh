// dnp_epilogue.cuh -- the fused per-target epilogues (K7) shared by the all-pairs kernels (dnp_pairs.cu,
// dnp_upairs.cu).  field_utils.py:110-115 (non-finite -> 0), :190-199, :267-276, :337-342 (normalise / sign).
#pragma once

#include "dnp_common.cuh"
#include "dnp_pairs.h"

namespace dnp {

// ------------------------------------------------------------------ epilogue (K7)
__device__ __forceinline__ float zero_nonfinite(float v, int& bad) {
    // (v - v) is NaN for both Inf and NaN
    if (!(fabsf(v) <= 3.402823466e38f)) { ++bad; return 0.f; }
    return v;
}

__device__ __forceinline__ void apply_field_epilogue(const PairEpilogue& ep, int64_t r,
                                                     float ex, float ey, float ez) {
    int bad = 0;
    ex = zero_nonfinite(ex, bad);          // field_utils.py:114-115
    ey = zero_nonfinite(ey, bad);
    ez = zero_nonfinite(ez, bad);
    if (bad && ep.nonfinite) atomicAdd(ep.nonfinite, bad);
    switch (ep.mode) {
        case DNP_EPI_STORE: {
            float* o = ep.out + r * ep.ld_out;
            o[0] = ex; o[1] = ey; o[2] = ez;
        } break;
        case DNP_EPI_ACCUM: {
            float* o = ep.out + r * ep.ld_out;
            o[0] += ex; o[1] += ey; o[2] += ez;
        } break;
        case DNP_EPI_SIGN_GE:
        case DNP_EPI_SIGN_GT: {
            float* n = ep.normals + r * ep.ld_n;
            const float nx = n[0], ny = n[1], nz = n[2];
            // same evaluation order as (E * n).sum(-1): ((ex*nx + ey*ny) + ez*nz), no contraction
            const float dot = __fadd_rn(__fadd_rn(__fmul_rn(ex, nx), __fmul_rn(ey, ny)), __fmul_rn(ez, nz));
            const bool keep = ep.mode == DNP_EPI_SIGN_GE ? (dot >= 0.f) : (dot > 0.f);
            const float sgn = keep ? 1.f : -1.f;
            n[0] = nx * sgn; n[1] = ny * sgn; n[2] = nz * sgn;
        } break;
        case DNP_EPI_NORMALIZE: {
            float* o = ep.out + r * ep.ld_out;
            const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(ez, ez)));
            if (len != 0.f) { ex = __fdiv_rn(ex, len); ey = __fdiv_rn(ey, len); ez = __fdiv_rn(ez, len); }
            o[0] = ex; o[1] = ey; o[2] = ez;
        } break;
        case DNP_EPI_DOT: {
            const float* n = ep.normals + r * ep.ld_n;
            ep.out[r * ep.ld_out] =
                __fadd_rn(__fadd_rn(__fmul_rn(ex, n[0]), __fmul_rn(ey, n[1])), __fmul_rn(ez, n[2]));
        } break;
        default: break;
    }
}

__device__ __forceinline__ void apply_potential_epilogue(const PairEpilogue& ep, int64_t r, float phi) {
    int bad = 0;
    phi = zero_nonfinite(phi, bad);        // field_utils.py:53-54
    if (bad && ep.nonfinite) atomicAdd(ep.nonfinite, bad);
    float* o = ep.out + r * ep.ld_out;
    if (ep.mode == DNP_EPI_ACCUM) *o += phi; else *o = phi;
}

}  // namespace dnp
