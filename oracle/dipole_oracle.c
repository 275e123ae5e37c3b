/*
 * dipole_oracle.c -- CPU restatement of the dipole-field hot path of
 * crazyMessi/dipole-normal-prop (reference commit 431bb62).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the shipped product path may link,
 * import or call this file; it is the checker for tests/, for
 * __graft_entry__.smoke() and for bench.py's cpu_baseline / --impl reference
 * legs.  Parity status: PINNED -- tests/test_oracle.py checks every function
 * here against the tests/golden fixtures, which were produced by running the
 * unmodified reference (tests/golden/make_golden.py).
 *
 * Conventions (reference field_utils.py:47,97):
 *   a point is a row [x y z nx ny nz]; R = source - target;
 *   phi(q) = sum_s n_s . R / |R|^3 ;  E(q) = -sum_s (3 (n_s.Rh) Rh - n_s) / (|R|^3 + eps).
 * All arrays are float32 with a row stride in elements (ld*), so column views
 * such as pc[:, :3] need no copy.
 *
 * Two arithmetic flavours:
 *   *_f64  : everything in double -- the accuracy referee (the reference run on
 *            .double() inputs).
 *   *_f32  : every per-pair operation rounded to float in the order the
 *            reference's torch expression evaluates it, pair terms accumulated
 *            in double and rounded once.  Compile with -ffp-contract=off.
 * The sequential point driver (orc_points_propagate_f32) is float end to end
 * because its running field is a float tensor in the reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ---------------------------------------------------------------- potential
 * reference: field_utils.py:46-55 (leaf).  mode 0 = reference leaf semantics:
 * a coincident pair gives 0/0 = NaN, and a NaN/Inf total is replaced by 0
 * (:53-54).  mode 1 = product semantics: only the coincident pair is dropped
 * (documented divergence on degenerate input). */
ORC_API void orc_potential_f64(const float *src, int64_t S, int64_t lds,
                               const float *tgt, int64_t M, int64_t ldt,
                               double *phi, int mode)
{
#pragma omp parallel for schedule(static)
    for (int64_t m = 0; m < M; ++m) {
        const double qx = tgt[m * ldt], qy = tgt[m * ldt + 1], qz = tgt[m * ldt + 2];
        double acc = 0.0;
        for (int64_t s = 0; s < S; ++s) {
            const float *p = src + s * lds;
            const double rx = p[0] - qx, ry = p[1] - qy, rz = p[2] - qz;
            const double r2 = rx * rx + ry * ry + rz * rz;
            if (mode == 1 && r2 == 0.0) continue;
            const double r = sqrt(r2);
            acc += (p[3] * rx + p[4] * ry + p[5] * rz) / (r * r * r);
        }
        if (isnan(acc) || isinf(acc)) acc = 0.0;
        phi[m] = acc;
    }
}

ORC_API void orc_potential_f32(const float *src, int64_t S, int64_t lds,
                               const float *tgt, int64_t M, int64_t ldt,
                               float *phi, int mode)
{
#pragma omp parallel for schedule(static)
    for (int64_t m = 0; m < M; ++m) {
        const float qx = tgt[m * ldt], qy = tgt[m * ldt + 1], qz = tgt[m * ldt + 2];
        double acc = 0.0;
        for (int64_t s = 0; s < S; ++s) {
            const float *p = src + s * lds;
            const float rx = p[0] - qx, ry = p[1] - qy, rz = p[2] - qz;
            const float r2 = rx * rx + ry * ry + rz * rz;
            if (mode == 1 && r2 == 0.0f) continue;
            const float r = sqrtf(r2);
            const float num = p[3] * rx + p[4] * ry + p[5] * rz;
            acc += (double)(num / (r * r * r));
        }
        float v = (float)acc;
        if (isnan(v) || isinf(v)) v = 0.0f;
        phi[m] = v;
    }
}

/* ---------------------------------------------------------------- field
 * reference: field_utils.py:96-116 (leaf): coincident pairs contribute exactly
 * zero (:99-106); E = (3 (p.Rh) Rh - p) / (|R|^3 + eps); total * -1 (:109). */
ORC_API void orc_field_grad_f64(const float *src, int64_t S, int64_t lds,
                                const float *tgt, int64_t M, int64_t ldt,
                                double eps, double *E)
{
#pragma omp parallel for schedule(static)
    for (int64_t m = 0; m < M; ++m) {
        const double qx = tgt[m * ldt], qy = tgt[m * ldt + 1], qz = tgt[m * ldt + 2];
        double ex = 0.0, ey = 0.0, ez = 0.0;
        for (int64_t s = 0; s < S; ++s) {
            const float *p = src + s * lds;
            const double rx = p[0] - qx, ry = p[1] - qy, rz = p[2] - qz;
            const double r = sqrt(rx * rx + ry * ry + rz * rz);
            if (r == 0.0) continue;
            const double ux = rx / r, uy = ry / r, uz = rz / r;
            const double d3 = 3.0 * (p[3] * ux + p[4] * uy + p[5] * uz);
            const double den = r * r * r + eps;
            ex += (d3 * ux - p[3]) / den;
            ey += (d3 * uy - p[4]) / den;
            ez += (d3 * uz - p[5]) / den;
        }
        E[3 * m] = -ex; E[3 * m + 1] = -ey; E[3 * m + 2] = -ez;
    }
}

/* one (source,target) term of field_grad in float, torch evaluation order */
static inline void pair_field_f32(const float *p, float qx, float qy, float qz, float eps,
                                  float *tx, float *ty, float *tz)
{
    const float rx = p[0] - qx, ry = p[1] - qy, rz = p[2] - qz;
    const float r = sqrtf(rx * rx + ry * ry + rz * rz);           /* R.norm(dim=-1)      :98  */
    if (r == 0.0f) { *tx = *ty = *tz = 0.0f; return; }            /* zero_mask           :98-106 */
    const float ux = rx / r, uy = ry / r, uz = rz / r;            /* R_unit              :102 */
    const float d3 = 3.0f * (p[3] * ux + p[4] * uy + p[5] * uz);  /* 3 * (p*R_unit).sum  :105 */
    const float den = r * r * r + eps;                            /* norm ** 3 + eps     :108 */
    *tx = (d3 * ux - p[3]) / den;
    *ty = (d3 * uy - p[4]) / den;
    *tz = (d3 * uz - p[5]) / den;
}

ORC_API void orc_field_grad_f32(const float *src, int64_t S, int64_t lds,
                                const float *tgt, int64_t M, int64_t ldt,
                                float eps, float *E)
{
#pragma omp parallel for schedule(static)
    for (int64_t m = 0; m < M; ++m) {
        const float qx = tgt[m * ldt], qy = tgt[m * ldt + 1], qz = tgt[m * ldt + 2];
        double ex = 0.0, ey = 0.0, ez = 0.0;
        for (int64_t s = 0; s < S; ++s) {
            float tx, ty, tz;
            pair_field_f32(src + s * lds, qx, qy, qz, eps, &tx, &ty, &tz);
            ex += tx; ey += ty; ez += tz;
        }
        E[3 * m] = (float)(-ex); E[3 * m + 1] = (float)(-ey); E[3 * m + 2] = (float)(-ez);
    }
}

/* ---------------------------------------------------------------- edge score
 * reference: field_utils.py:146-155: I = sum_t n_t . E_S(t).  The caller forms
 * w = (2 I) / |S| * |T| (left-to-right precedence as written at :155). */
ORC_API double orc_edge_interaction_f64(const float *S_, int64_t S, int64_t lds,
                                        const float *T_, int64_t T, int64_t ldt, double eps)
{
    double *E = (double *)malloc(sizeof(double) * 3 * (size_t)(T > 0 ? T : 1));
    orc_field_grad_f64(S_, S, lds, T_, T, ldt, eps, E);
    double acc = 0.0;
    for (int64_t t = 0; t < T; ++t)
        acc += E[3 * t] * T_[t * ldt + 3] + E[3 * t + 1] * T_[t * ldt + 4] + E[3 * t + 2] * T_[t * ldt + 5];
    free(E);
    return acc;
}

/* ---------------------------------------------------------------- lattice
 * reference: util.py:26-36 (gen_grid) + field_utils.py:7-9: mean potential on
 * the n^3 lattice {-1, -1+2/n, ...}^3, x-major (x = i/n^2, y = i/n % n, z = i % n). */
ORC_API void orc_lattice_points(int n, float *out /* [n^3,3] */)
{
    for (int i = 0; i < n * n * n; ++i) {
        const int z = i % n, xy = i / n, y = xy % n, x = xy / n;
        out[3 * i + 0] = ((float)x / (float)n - 0.5f) * 2.0f;
        out[3 * i + 1] = ((float)y / (float)n - 0.5f) * 2.0f;
        out[3 * i + 2] = ((float)z / (float)n - 0.5f) * 2.0f;
    }
}

ORC_API double orc_mean_potential_f64(const float *src, int64_t S, int64_t lds, int n, int mode)
{
    const int64_t M = (int64_t)n * n * n;
    float *g = (float *)malloc(sizeof(float) * 3 * (size_t)M);
    double *phi = (double *)malloc(sizeof(double) * (size_t)M);
    orc_lattice_points(n, g);
    orc_potential_f64(src, S, lds, g, M, 3, phi, mode);
    double acc = 0.0;
    for (int64_t i = 0; i < M; ++i) acc += phi[i];
    free(g); free(phi);
    return acc / (double)M;
}

/* ---------------------------------------------------------------- per-point greedy propagation
 * reference: field_utils.py:353-388.  pts[N,6] (contiguous) is updated in
 * place; order[k] = index chosen at step k (order[0] = start), flipped[k] = 1
 * if that point's normal was negated.  Arithmetic is float in torch's order:
 *   E[others] += field_grad(pts[i:i+1], pts[others,:3], eps)         :379-380
 *   interaction = (E[~visited] * n[~visited]).sum(-1)                 :369
 *   k = interaction.abs().argmax()  (first maximum in index order)    :372
 *   if interaction[k] < 0: n[idx] *= -1                               :375-376
 * With one source the leaf's sum over sources is the identity, so every step
 * is element-wise and this loop reproduces the reference bit for bit as long
 * as the compiler does not contract a*b+c (build with -ffp-contract=off). */
static void points_propagate_impl(float *pts, int64_t N, int64_t start, float eps, int diffuse,
                                  int64_t max_steps, int32_t *order, uint8_t *flipped);

ORC_API void orc_points_propagate_f32(float *pts, int64_t N, int64_t start, float eps,
                                      int diffuse, int32_t *order, uint8_t *flipped)
{
    points_propagate_impl(pts, N, start, eps, diffuse, N, order, flipped);
}

/* The first `max_steps` greedy steps only (no diffuse epilogue): the prefix of the decision trace of a
 * cloud too large to run to the end on the CPU.  order/flipped entries 0..max_steps are written. */
ORC_API void orc_points_propagate_prefix_f32(float *pts, int64_t N, int64_t start, float eps,
                                             int64_t max_steps, int32_t *order, uint8_t *flipped)
{
    points_propagate_impl(pts, N, start, eps, 0, max_steps, order, flipped);
}

static void points_propagate_impl(float *pts, int64_t N, int64_t start, float eps, int diffuse,
                                  int64_t max_steps, int32_t *order, uint8_t *flipped)
{
    float *E = (float *)calloc((size_t)(3 * N), sizeof(float));
    uint8_t *visited = (uint8_t *)calloc((size_t)N, 1);
    int64_t cur = start;
    visited[cur] = 1;
    if (order) order[0] = (int32_t)cur;
    if (flipped) flipped[0] = 0;
    for (int64_t step = 1; step <= N && step <= max_steps; ++step) {
        /* add the field of `cur` to every other point, then pick the next one */
        const float *p = pts + 6 * cur;
        float best = -1.0f; int64_t best_i = -1; float best_val = 0.0f;
        for (int64_t m = 0; m < N; ++m) {
            if (m != cur) {
                float tx, ty, tz;
                const float *q = pts + 6 * m;
                pair_field_f32(p, q[0], q[1], q[2], eps, &tx, &ty, &tz);
                E[3 * m] += -tx; E[3 * m + 1] += -ty; E[3 * m + 2] += -tz;
            }
            if (!visited[m]) {
                const float *q = pts + 6 * m;
                const float v = E[3 * m] * q[3] + E[3 * m + 1] * q[4] + E[3 * m + 2] * q[5];
                const float a = fabsf(v);
                if (a > best || best_i < 0) {
                    if (best_i < 0 || a > best) { best = a; best_i = m; best_val = v; }
                }
            }
        }
        if (best_i < 0) break;                  /* everything visited */
        if (best_val < 0.0f) {
            pts[6 * best_i + 3] *= -1.0f; pts[6 * best_i + 4] *= -1.0f; pts[6 * best_i + 5] *= -1.0f;
        }
        if (order) order[step] = (int32_t)best_i;
        if (flipped) flipped[step] = best_val < 0.0f;
        visited[best_i] = 1;
        cur = best_i;
    }
    if (diffuse) {                              /* :382-385, strict > 0 */
        for (int64_t m = 0; m < N; ++m) {
            float *q = pts + 6 * m;
            const float v = E[3 * m] * q[3] + E[3 * m + 1] * q[4] + E[3 * m + 2] * q[5];
            const float sgn = v > 0.0f ? 1.0f : -1.0f;
            q[3] *= sgn; q[4] *= sgn; q[5] *= sgn;
        }
    }
    free(E); free(visited);
}

/* ---------------------------------------------------------------------------
 * Patch-graph flip solve (graph.py:431-470).  The reference maximises
 *   cal_loss(x) = sum_ij A_ij (1 - (x_i-x_j)^2) + B_ij (x_i-x_j)^2,  x in {0,1}^n
 * with Gurobi (gurobipy: a third-party exact solver, not pinned by the reference
 * and absent here).  An exact solver returns a global maximiser, so the oracle
 * is plain enumeration: every x in ascending integer order (bit i = x_i),
 * cal_loss evaluated term by term in the reference's loop order, first maximum
 * kept.  n <= 26.
 * ------------------------------------------------------------------------- */
ORC_API double orc_cal_loss(const uint8_t *x, const double *A, const double *B, int64_t n)
{
    double obj = 0.0;
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = 0; j < n; ++j) {
            const double d = (double)x[i] - (double)x[j];
            obj += A[i * n + j] * (1.0 - d * d) + B[i * n + j] * d * d;
        }
    return obj;
}

ORC_API double orc_miqp_bruteforce(const double *A, const double *B, int64_t n, uint8_t *x_best,
                                   int64_t *n_optimal /* states within 1e-12 relative of the maximum */)
{
    const uint64_t states = 1ull << n;
    double best = -INFINITY;
    uint64_t arg = 0;
    double *vals = n_optimal ? (double *)malloc(sizeof(double) * states) : NULL;
    #pragma omp parallel
    {
        uint8_t x[64];
        double lbest = -INFINITY;
        uint64_t larg = 0;
        #pragma omp for schedule(static)
        for (uint64_t s = 0; s < states; ++s) {
            for (int64_t i = 0; i < n; ++i) x[i] = (uint8_t)((s >> i) & 1u);
            const double v = orc_cal_loss(x, A, B, n);
            if (vals) vals[s] = v;
            if (v > lbest) { lbest = v; larg = s; }
        }
        #pragma omp critical
        if (lbest > best || (lbest == best && larg < arg)) { best = lbest; arg = larg; }
    }
    for (int64_t i = 0; i < n; ++i) x_best[i] = (uint8_t)((arg >> i) & 1u);
    if (n_optimal) {
        int64_t c = 0;
        const double tol = 1e-12 * (fabs(best) > 1.0 ? fabs(best) : 1.0);
        for (uint64_t s = 0; s < states; ++s) c += vals[s] >= best - tol;
        *n_optimal = c;
        free(vals);
    }
    return best;
}
