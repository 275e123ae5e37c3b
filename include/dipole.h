/*
 * dipole.h -- C ABI of libdipole.so, the B200 (sm_100a) implementation of the
 * dipole-field hot path of crazyMessi/dipole-normal-prop.
 *
 * The reference has no FFI: its boundary is the Python function set of
 * field_utils.py.  Each entry point below names the
 * reference function (file:line, commit 431bb62) whose device work it replaces.
 * INTEGRATION.md shows the ctypes binding a maintainer of the reference adds.
 *
 * Rules common to every entry point
 *   - plain pointers and sizes only; no torch / C++ types cross the boundary.
 *   - pointers are DEVICE pointers unless the name ends in _host; the caller
 *     owns every buffer, the library never allocates outputs.
 *   - `ld_*` are row strides in float elements, so views such as pc[:, :3]
 *     (stride 6) or E[:, :3] need no copy.  A point row is [x y z nx ny nz].
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *     Work is enqueued asynchronously; nothing synchronises unless stated, and
 *     every device-pointer entry point may be recorded into a CUDA graph.
 *     Potential / field calls above ~2e8 pairs stream their sources through a
 *     64 KB constant window that exists once per device (dnp_upairs.cu): calls
 *     from different host threads / streams are queued on it in submission
 *     order (an event wait on the caller's stream, never a host wait).  A graph
 *     that contains such a call must not be REPLAYED concurrently with other
 *     large potential / field calls on the same device (replays bypass that
 *     queue); DNP_RING=0 in the environment disables the window path.
 *   - scratch memory comes from the caller: query dnp_pairs_workspace_bytes()
 *     and pass a device buffer at least that large (256-byte aligned).
 *   - return value: DNP_OK (0) or a negative DNP_ERR_* code; dnp_last_error()
 *     returns a thread-local message.  Nothing throws across the ABI.
 *   - re-entrant: the only shared state is the per-device constant window and
 *     its lane streams (a mutex held while enqueueing); per-host-thread scratch
 *     of the *_host entry points is released when the thread exits (the
 *     reference calls this path from many Python threads, util.py:187-196,308-327).
 *   - float32 only (the file pipelines of the reference are float32,
 *     util.py:61,66); there is no CPU fallback.
 */
#ifndef DIPOLE_H_
#define DIPOLE_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define DNP_API __attribute__((visibility("default")))
#else
#define DNP_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define DNP_OK 0
#define DNP_ERR_INVALID (-1)   /* bad argument */
#define DNP_ERR_CUDA (-2)      /* a CUDA runtime call failed */
#define DNP_ERR_WORKSPACE (-3) /* workspace missing or too small */
#define DNP_ERR_NO_DEVICE (-4) /* no sm_100 device visible */
#define DNP_ERR_IO (-5)        /* a write to the caller's file descriptor failed */

/* fused epilogues of the all-pairs field kernel (K7 in SURVEY.md section 2.2) */
#define DNP_EPI_STORE 0     /* out[r,0:3]  = E_r                      field_utils.py:96-116        */
#define DNP_EPI_ACCUM 1     /* out[r,0:3] += E_r                      field_utils.py:261,330,335,380 */
#define DNP_EPI_SIGN_GE 2   /* n_r *= (E_r.n_r >= 0 ? 1 : -1)         field_utils.py:196-199       */
#define DNP_EPI_SIGN_GT 3   /* n_r *= (E_r.n_r >  0 ? 1 : -1)         field_utils.py:273-276       */
#define DNP_EPI_NORMALIZE 4 /* out[r,0:3] = E_r/|E_r| (raw E if 0)    field_utils.py:191-194       */
#define DNP_EPI_DOT 5       /* out[r]      = E_r.n_r                  field_utils.py:147-148       */

DNP_API const char* dnp_last_error(void);
DNP_API int dnp_version(void);
/* "libdipole sm_100a sources=<16 hex digits> nvcc=<version> built=<date>": the hash covers every .cu / .cuh / .h of
 * csrc/ plus this header, in sorted order (csrc/Makefile); dnp_native.load() refuses a library whose hash differs from
 * the source tree it sits in. */
DNP_API const char* dnp_build_info(void);

/* SM count and clocks of the current device (for roofline arithmetic). */
DNP_API int dnp_device_info(int* sm_count, int* sm_clock_khz, int* cc_major, int* cc_minor);

/* Scratch bytes needed by any dnp_* pairs call with at most S sources and M targets. */
DNP_API size_t dnp_pairs_workspace_bytes(int64_t S, int64_t M);

/* ------------------------------------------------------------------ a1: potential
 * replaces field_utils.potential (field_utils.py:12-55):
 *   phi[m] (+)= sum_s n_s . (p_s - q_m) / |p_s - q_m|^3.
 * `eps` does not exist here because the reference never uses it.  The recursive
 * halving (`recursive`, `max_pts`) has no equivalent: memory is O(S + M).
 * A coincident (source, target) pair contributes 0 (the reference zeroes the
 * whole leaf total of that target instead -- degenerate input).
 * Non-finite totals are written as 0 (field_utils.py:53-54).
 * src_idx (optional, may be NULL): gather sources src[src_idx[i]], i < S.      */
DNP_API int dnp_potential_f32(const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx,
                      const float* tgt, int64_t M, int64_t ld_t,
                      float* phi, int accumulate,
                      void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ a3: lattice test
 * replaces field_utils.measure_mean_potential + util.gen_grid
 * (field_utils.py:7-9, util.py:26-36): potential on the n^3 lattice
 * ((i/n^2, i/n%n, i%n)/n - 0.5)*2 generated in-kernel, and its mean.
 * phi_opt: optional [n^3] output (may be NULL); mean_out: 1 float (device).     */
DNP_API int dnp_lattice_mean_potential_f32(const float* src, int64_t S, int64_t ld_s,
                                   int n, float* phi_opt, float* mean_out,
                                   void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ a2: field
 * replaces field_utils.field_grad (field_utils.py:61-116):
 *   E[m] = -sum_s (3 (n_s.Rh) Rh - n_s) / (|R|^3 + eps),  R = p_s - q_m,
 * coincident pairs contribute exactly 0 (:98-106); non-finite components are
 * written as 0 (:114-115) and counted in *nonfinite (optional device int).      */
DNP_API int dnp_field_grad_f32(const float* src, int64_t S, int64_t ld_s,
                       const float* tgt, int64_t M, int64_t ld_t,
                       float eps, float* E, int64_t ld_e, int accumulate,
                       void* workspace, size_t workspace_bytes, void* stream);

/* General form used by the drivers (a5-a9): gathers, target mask, fused epilogue.
 *   sources : src[src_idx[i]] (i < S) or src[i] if src_idx == NULL
 *   targets : row r = tgt_idx[k] (k < M) or r = k; skipped if tgt_skip && tgt_skip[r]
 *   mode    : DNP_EPI_*; `out` / `normals` are addressed by the same row r
 *             (out: ld_out, normals: ld_n; normals needed for SIGN_* and DOT).
 *   nonfinite: optional device counter incremented per zeroed component.        */
DNP_API int dnp_field_apply_f32(const float* src, int64_t S, int64_t ld_s, const int64_t* src_idx,
                        const float* tgt, int64_t M, int64_t ld_t,
                        const int64_t* tgt_idx, const uint8_t* tgt_skip,
                        float eps, int mode,
                        float* out, int64_t ld_out, float* normals, int64_t ld_n,
                        int32_t* nonfinite,
                        void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ a4: patch-pair flip score
 * replaces the device work of field_utils.field_edge_calculator
 * (field_utils.py:145-160): interaction[0] = sum_t n_t . E_S(t) (device float).
 * The caller forms w = (2 I) / |S| * |T| and (w, -w) as the reference does.    */
DNP_API int dnp_edge_interaction_f32(const float* S_pts, int64_t S, int64_t ld_s,
                             const float* T_pts, int64_t T, int64_t ld_t,
                             float eps, float* interaction,
                             void* workspace, size_t workspace_bytes, void* stream);

/* Batched form (K3): patches in CSR (perm[offsets[p] .. offsets[p+1]) are rows of
 * pts), edges[e] = (source patch, target patch); out[e] = sum_t n_t . E_S(t).
 * One CTA per edge, deterministic reduction, no atomics.
 * tgt_perm / tgt_offsets (optional, both or neither): a second CSR used for the
 * TARGET side of every edge -- a rank of a multi-GPU job passes its slice of each
 * patch here, and the per-edge partial sums are combined with one allreduce.   */
DNP_API int dnp_edge_interactions_batched_f32(const float* pts, int64_t ld,
                                      const int64_t* perm, const int64_t* offsets, int64_t P,
                                      const int64_t* tgt_perm, const int64_t* tgt_offsets,
                                      const int64_t* edges, int64_t n_edges,
                                      float eps, float* out, void* stream);

/* ------------------------------------------------------------------ a9: per-patch interaction
 * replaces the list comprehension at field_utils.py:244,316 and the
 * abs().argmax() at :247,319:  out[p] = sum_{t in patch p} E_t . n_t for every
 * patch with active[p] != 0 (0 elsewhere);  best[0] = index of the first
 * maximal |out[p]| among active patches (as float bits of an int32 in best_idx),
 * best_val[0] = its signed value.  One launch, no host synchronisation.        */
DNP_API int dnp_patch_interactions_f32(const float* E, int64_t ld_e, const float* normals, int64_t ld_n,
                               const int64_t* perm, const int64_t* offsets, int64_t P,
                               const uint8_t* active, float* out,
                               int32_t* best_idx, float* best_val, void* stream);

/* Streamed per-point sign (K7 without a field evaluation): for k < n, r = idx ? idx[k] : k,
 *   n_r *= (E_r . n_r > 0 ? 1 : -1)    mode DNP_EPI_SIGN_GT  field_utils.py:267-271,337-342,382-385
 *   n_r *= (E_r . n_r >= 0 ? 1 : -1)   mode DNP_EPI_SIGN_GE  field_utils.py:196-199
 * HBM-bound: 24 bytes read + 12 written per point.                              */
DNP_API int dnp_field_sign_f32(const float* E, int64_t ld_e, float* normals, int64_t ld_n,
                       const int64_t* idx, int64_t n, int mode, void* stream);

/* Start-patch selection support: cov[p] (row-major 3x3, double) = covariance of
 * the xyz of patch p about its mean, divided by the patch size -- the matrix
 * util.pca_eigen_values (util.py:495-497) hands to symeig at
 * field_utils.py:230,303.  One CTA per patch, fixed summation order.            */
DNP_API int dnp_patch_covariance_f32(const float* pts, int64_t ld,
                             const int64_t* perm, const int64_t* offsets, int64_t P,
                             double* cov, void* stream);

/* ------------------------------------------------------------------ a6: per-point greedy propagation
 * replaces field_utils.strongest_field_propagation_points (field_utils.py:353-388)
 * entirely on the device: N-1 sequential steps of {add the chosen point's field
 * to every other point, masked E.n, abs-argmax (first maximum), flip if < 0},
 * IEEE arithmetic in the reference's evaluation order so that the decision
 * sequence is the reference's.  pts[N,6] (ld) is updated in place.
 * order_opt[N] / flipped_opt[N] (optional) receive the visit order and flips.
 * All state lives in shared memory: one thread-block cluster (<= 16 CTAs, DSMEM
 * exchange, about 97 000 points) or, beyond that, a cooperative grid of one CTA
 * per SM that exchanges the step winner through L2 (dnp_points_max_n() points,
 * about 900 000 on a B200).  The grid variant needs dnp_points_workspace_bytes(N)
 * bytes of device scratch (0 for the cluster variant; workspace may then be NULL). */
DNP_API int64_t dnp_points_max_n(void);
DNP_API size_t dnp_points_workspace_bytes(int64_t N);
DNP_API int dnp_points_propagate_f32(float* pts, int64_t N, int64_t ld, int64_t starting_point,
                             float eps, int diffuse,
                             int32_t* order_opt, uint8_t* flipped_opt,
                             void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ greedy patch propagation, whole loop
 * replaces the loops of field_utils.strongest_field_propagation (field_utils.py:311-335) and
 * strongest_field_propagation_reps (:239-265): starting from `start_patch`, P - 1 times: per-patch interactions
 * sum E.n over the remaining patches, first abs-arg-max, flip the chosen patch if its interaction is negative,
 * add its field to the running field E.  Patches are a CSR: perm (device int64) with offsets given BOTH on the
 * device (offsets_dev) and on the host (offsets_host, the launch shapes depend on the patch sizes).
 * rest_perm / rest_offsets_host (optional): rows that are flipped together with patch p but are neither sources nor
 * targets (the non-representative points of a patch, field_utils.py:250-252).
 * tgt_idx (optional, device, n_tgt rows): the target rows (all representatives, :239); NULL = all N rows.
 * diffuse != 0: only the source patch is skipped as a target (:256-261,326-331), else every visited patch.
 * E: [N, ld_e >= 3] running field, ZEROED by the caller, holds the final field on return (diffuse loops; a loop that
 * is not diffuse leaves each row at the field accumulated until its patch was visited, or at zero -- see the slab schedule).
 * order_host / value_host (optional, host, P entries): visit order and the signed interaction that chose each
 * patch.  Schedule: the step winner stays on the device.  Small clouds (largest patch <= 1024 rows, disjoint patches
 * that are exactly the target rows -- checked with one 8-byte read per call) run all P steps inside ONE persistent
 * cooperative kernel (loop state in shared memory, one grid barrier per step).  Otherwise every launch reads its row
 * range from the device-resident winner, all P steps are enqueued without a host read and the call returns after the
 * last launch has been enqueued.  Only when a patch is large enough for the constant-window all-pairs kernel (>= 16 384
 * rows; its launches are planned on the host) does the loop read the winner back once per step (8 bytes).  With a trace
 * requested the call synchronises `stream` once at the end.  Large loops (>= 229 376 target rows, or >= 65 536 with patches of <= 2048
 * rows; patches of >= 128 rows on average; same disjointness check; from 98 304 rows on ahead of the persistent
 * kernel) take the SLAB schedule when P x T x 12 bytes fit DNP_SLAB_MAX_MB (default
 * 65536): the field of every patch at every target is computed first -- P independent whole-tile launches at the rate
 * of one large call, one [T, 3] slab per patch inside the workspace -- then the greedy decisions run on the P x P
 * matrix of per-patch sums in one CTA and E is assembled from the slabs in visit order; same decisions, no chain of
 * dependent one-wave field launches (dnp_slab.cu).  When the loop is NOT diffuse the slab schedule computes every patch
 * pair once (the interaction matrix is symmetric) and returns E as ZEROS -- nothing in the reference reads the running
 * field of such a loop; DNP_SLAB_SYMMETRIC=0 keeps the full form.  DNP_PATCH_LOOP=persistent|slab|device|host forces a schedule
 * (the workspace query reads the same variable), DNP_SLAB=0 turns the slab schedule off.                       */
DNP_API size_t dnp_patch_propagate_workspace_bytes(int64_t N, int64_t P, int64_t max_patch, int64_t n_tgt);
DNP_API int dnp_patch_propagate_f32(float* pts, int64_t N, int64_t ld,
                            const int64_t* perm, const int64_t* offsets_dev, const int64_t* offsets_host, int64_t P,
                            const int64_t* rest_perm, const int64_t* rest_offsets_host,
                            const int64_t* tgt_idx, int64_t n_tgt,
                            int64_t start_patch, float eps, int diffuse,
                            float* E, int64_t ld_e, int32_t* order_host, float* value_host,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ per-patch greedy orientation, batched
 * The per-patch stage of the patch graph (util.py:185-196 calling graph_dipole.st_propagation_points_file,
 * graph_dipole.py:27-36, once per patch from Python threads) as ONE launch: patch c = rows offsets[c] ..
 * offsets[c+1] of `pts` (patches stored contiguously, offsets: device int64[P+1]); every patch runs
 * strongest_field_propagation_points (field_utils.py:353-388) from its first row on its own thread-block
 * cluster, in place.  max_patch = the largest patch size (<= dnp_points_batched_max_n()).
 * dnp_patch_lattice_sign_f32: measure_mean_potential (field_utils.py:7-9) of every patch on the n^3 lattice
 * (n^3 <= 2048) -> mean_out[P]; with flip != 0 the normals of patches with a negative mean are negated
 * (graph_dipole.py:30-31).  Both fully asynchronous.                                                     */
DNP_API int64_t dnp_points_batched_max_n(void);
DNP_API int dnp_points_propagate_batched_f32(float* pts, int64_t ld, const int64_t* offsets, int64_t P,
                                     int64_t max_patch, float eps, int diffuse, void* stream);
DNP_API int dnp_patch_lattice_sign_f32(float* pts, int64_t ld, const int64_t* offsets, int64_t P, int n,
                               float* mean_out, int flip, void* stream);

/* ------------------------------------------------------------------ "xie" reflected-normal interaction
 * (the fork's N x N propagation variant.)
 * dnp_xie_field_f32 replaces field_utils.xie_field (field_utils.py:431-469):
 *   field[t, s, 0:3] = (n_s - C (n_s.Rh) Rh) / |R|^3,  R = p_s - q_t,  Rh = R/|R|;
 *   a coincident pair keeps the un-normalised n_s (:456-460).  `eps` is unused by the reference.
 * dnp_xie_interaction_f32 replaces field_utils.xie_intersaction (:509-519):
 *   inter[t, s] = field[t, s] . n_t, non-finite entries written as 0.
 * Outputs are row-major over targets: inter[t * ld_o + s], field[t * ld_o + 3 s + c].
 * src_radius2_opt (device doubles [S], may be NULL) carries the kNN mask of :447-455: source s only
 * acts on targets with |R|^2 <= src_radius2_opt[s] (squared distance to its k-th nearest target,
 * evaluated in double from the float coordinates).                                              */
DNP_API int dnp_xie_interaction_f32(const float* src, int64_t S, int64_t ld_s,
                            const float* tgt, int64_t M, int64_t ld_t,
                            float C, const double* src_radius2_opt,
                            float* inter, int64_t ld_o, void* stream);
DNP_API int dnp_xie_field_f32(const float* src, int64_t S, int64_t ld_s,
                      const float* tgt, int64_t M, int64_t ld_t,
                      float C, const double* src_radius2_opt,
                      float* field, int64_t ld_o, void* stream);

/* replaces field_utils.xie_propagation_points_in_order (field_utils.py:569-605) without
 * materialising the [N, N] matrix: T independent chains, chain t visits order[t * N + i],
 * i = 0..N-1; the visited point's interaction with the points already weighted is summed (in
 * double), its weight becomes -1 if the sum is < 0, else +1.  flips[t * N + p] = (sum < 0) and
 * interactions[t * N + p] = the sum, as of the visit of p; with diffuse != 0 both are
 * recomputed from the final weights (inter = mat w, :598-604).  One launch: one thread-block
 * cluster per chain with the points in shared memory (N <= dnp_xie_max_n()).
 * workspace: dnp_xie_workspace_bytes(N, T) device bytes, only needed with diffuse.            */
DNP_API int64_t dnp_xie_max_n(int knn);

/* the kNN mask of field_utils.xie_field / xie_intersaction (field_utils.py:447-455: a scipy KD-tree query of the
 * knn_mask nearest targets per source) as ONE squared radius per source: radius2[s] = squared distance (double, exact
 * differences of the float coordinates) from source s to its k-th nearest target, 1 <= k <= min(64, M).  The xie
 * kernels keep a pair iff |R|^2 <= radius2[s].  One thread per source, targets staged through shared memory.       */
DNP_API int dnp_knn_radius2_f64(const float* src, int64_t S, int64_t ld_s, const float* tgt, int64_t M, int64_t ld_t,
                        int k, double* radius2, void* stream);
DNP_API size_t dnp_xie_workspace_bytes(int64_t N, int64_t T);
DNP_API int dnp_xie_propagate_in_order_f32(const float* pts, int64_t N, int64_t ld,
                                   const int64_t* order, int64_t T,
                                   float C, const double* radius2_opt, int diffuse,
                                   uint8_t* flips, float* interactions,
                                   void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ voxel partition into patches
 *   replaces util._divide_pc (util.py:110-150): cell (i, j, k),
 * 0 <= i, j, k <= n_part, holds the points with l_t < x <= l_t + edge on every axis,
 * l_t = edge * t + range_lo, edge = (range_hi - range_lo) / n_part (double arithmetic, float compare,
 * as torch does).  Output is the CSR of the non-empty cells in lexicographic (i, j, k) order:
 *   perm[offsets[c] .. offsets[c+1])  rows of cell c, ascending        (perm: [N] int64)
 *   offsets[0 .. n_cells]                                              (offsets: [min(N, (n_part+1)^3) + 1] int64)
 *   ijk[3c .. 3c+2]                   grid coordinates of cell c       (ijk: 3 x as many int32)
 *   counts[0] = n_cells, counts[1] = points inside the box (= offsets[n_cells]); device int64[2].
 * Points outside the box (or NaN) are in no cell, as in the reference.  n_part <= 255.
 * Fully asynchronous; workspace: dnp_voxel_partition_workspace_bytes(N) device bytes.              */
DNP_API size_t dnp_voxel_partition_workspace_bytes(int64_t N);
DNP_API int dnp_voxel_partition_f32(const float* pts, int64_t N, int64_t ld,
                            int n_part, double range_lo, double range_hi,
                            int64_t* perm, int64_t* offsets, int32_t* ijk, int64_t* counts,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ patch-graph flip solve
 *   replaces graph.MIQP(A, B) (graph.py:444-470; objective graph.cal_loss,
 * graph.py:431-442), which the reference hands to Gurobi (gurobipy, unpinned, absent here):
 *     maximise  sum_ij A_ij (1 - (x_i - x_j)^2) + B_ij (x_i - x_j)^2   over x in {0,1}^n.
 * A, B: device float64 [n, n] row-major with row stride ld (the matrices of BidGraph.to_matrix,
 * graph.py:136-152; need not be symmetric).  n <= min(exact_max_n, 40): exhaustive enumeration (optimal up to
 * float32 rounding of the energies: the walk is float32, its 32 best candidates are re-evaluated in double);
 * larger n (<= 4096): `replicas` annealing chains of `sweeps` sweeps + descent, deterministic in `seed`.
 * x_out[n] (uint8, 1 = flip the patch), obj_out[0] = objective of x_out (double, as cal_loss),
 * obj_out[1] = the bound sum of the larger of A, B per pair; status_out[0] = 1 exhaustive, 2 bound reached
 * (both optimal), 0 best found.  Of the two globally flipped twins the one with fewer flips is returned
 * (tie: x_0 = 0).  Fully asynchronous; all outputs are device pointers.                                   */
DNP_API size_t dnp_flip_solve_workspace_bytes(int64_t n, int exact_max_n, int replicas);
DNP_API int dnp_flip_solve_f64(const double* A, const double* B, int64_t n, int64_t ld,
                       int exact_max_n, int replicas, int sweeps, uint64_t seed,
                       uint8_t* x_out, double* obj_out, int32_t* status_out,
                       void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ point-cloud text IO and wire format
 * 
 * dnp_xyz_parse_f32 replaces util.xyz2tensor (util.py:53-69): `txt` = the file's bytes on the device; lines are
 * split on '\n' and stripped, lines containing "nan" are skipped, tokens are split on single spaces, lines with 3 or 6
 * tokens become rows; each token is converted like Python's float() (correctly rounded to double) and then to
 * float32.  out: [max_rows, 6] float32 (3-column lines get zero normals); max_lines >= number of '\n' + 1.
 * counts (device int64[4]): rows written, column mask (bit 0: 3-column lines seen, bit 1: 6-column lines seen),
 * index of the first line holding a literal the parser does not resolve (-1 = none; malformed, underscores, more
 * than 38 significant digits on a rounding boundary, decimal exponent outside [-76, 60]), number of lines.
 * dnp_xyz_format_f32 replaces util.export_pc (util.py:46-51): value (r, c) = pts[r * ld_row + c * ld_col]; every value
 * is written as CPython's str(float(value)) (shortest round-trip repr of the widened double), ' ' between
 * columns, '\n' between rows, no trailing newline.  out: device bytes (capacity >= 25 * n_rows * cols suffices),
 * n_bytes_out: device int64.
 * dnp_wire_decode_f64 / dnp_wire_encode_f64: the server payloads (socket_server_para.py:145-166,191):
 * N x 3 float64 -> [N, 6] float32 rows with zero normals, and [N, 6] float32 rows -> N x 6 float64.
 * All fully asynchronous, except dnp_xyz_export_fd_f32: the same text as dnp_xyz_format_f32, written to the open file
 * descriptor `fd` (util.export_pc's `open(dest, 'w+')` + write, util.py:46-51).  The text is formatted in `workspace`
 * (device, dnp_xyz_export_workspace_bytes) and streamed through two pinned 4 MB buffers kept per host thread, the
 * download of one chunk overlapping the write(2) of the previous one; returns after the last write,
 * *n_bytes_host (optional) = bytes written.                                                                      */
DNP_API size_t dnp_xyz_parse_workspace_bytes(int64_t n_bytes, int64_t max_lines);
DNP_API int dnp_xyz_parse_f32(const char* txt, int64_t n_bytes, int64_t max_lines,
                      float* out, int64_t max_rows, int64_t* counts,
                      void* workspace, size_t workspace_bytes, void* stream);
DNP_API size_t dnp_xyz_format_workspace_bytes(int64_t n_values);
DNP_API int dnp_xyz_format_f32(const float* pts, int64_t n_rows, int cols, int64_t ld_row, int64_t ld_col,
                       char* out, int64_t capacity, int64_t* n_bytes_out,
                       void* workspace, size_t workspace_bytes, void* stream);
DNP_API size_t dnp_xyz_export_workspace_bytes(int64_t n_values);
DNP_API int dnp_xyz_export_fd_f32(const float* pts, int64_t n_rows, int cols, int64_t ld_row, int64_t ld_col,
                          int fd, int64_t* n_bytes_host, void* workspace, size_t workspace_bytes, void* stream);
DNP_API int dnp_wire_decode_f64(const double* xyz, int64_t n, float* out, int64_t ld, void* stream);
DNP_API int dnp_wire_encode_f64(const float* pts, int64_t n, int64_t ld, double* out, void* stream);

/* ------------------------------------------------------------------ host-buffer entry points
 * Same math with HOST pointers: the call stages inputs to the device, runs the
 * kernels and copies the result back before returning (synchronous).  This is
 * what a CPU-tensor caller of the reference API reaches, and what bench.py
 * times as `e2e`.  `device` = CUDA ordinal.                                     */
DNP_API int dnp_field_grad_host_f32(const float* src_host, int64_t S, int64_t ld_s,
                            const float* tgt_host, int64_t M, int64_t ld_t,
                            float eps, float* E_host, int device);
DNP_API int dnp_potential_host_f32(const float* src_host, int64_t S, int64_t ld_s,
                           const float* tgt_host, int64_t M, int64_t ld_t,
                           float* phi_host, int device);

/* ------------------------------------------------------------------ measurement helpers
 * Launch counter: number of libdipole kernels launched by this process
 * (bench.py reports it as gpu_launches).                                        */
DNP_API int64_t dnp_launch_count(void);
/* Per-thread kernel timer for the roofline numbers: between dnp_profile_begin() and
 * dnp_profile_end(), every all-pairs kernel launched by the CALLING host thread is
 * bracketed by CUDA events on its own stream (no synchronisation is added).
 * dnp_profile_end() waits for them and writes one duration (ms) per launch into
 * kernel_ms_host[0 .. min(capacity, *n_kernels)).  State is thread-local.        */
DNP_API int dnp_profile_begin(void);
DNP_API int dnp_profile_end(float* kernel_ms_host, int capacity, int* n_kernels);
/* Issue-rate microbenchmarks for the roofline denominators (kind: 0 = FFMA,
 * 1 = FFMA2 (packed f32x2), 2 = MUFU.RSQ, 3 = the field kernel's instruction mix,
 * 4/5/6 = register-file bandwidth probes: FFMA2 / FFMA with three distinct register
 * operands, FFMA2 with one operand shared by consecutive instructions).
 * Returns elapsed ms and the operation count.                                     */
DNP_API int dnp_microbench(int kind, int iters, float* ms_out, double* ops_out, void* stream);

/* ------------------------------------------------------------------ multi-GPU (dnp_comm.cu)
 * The reference is single-process, single-device: this block is net-new.  One process per GPU;
 * the collective is enqueued on the SAME stream as the shard's kernels, so a non-PyTorch host gets the whole sharded
 * call from one function.  NCCL is loaded at run time (dlopen "libnccl.so.2" or $DNP_NCCL_LIB): single-GPU users never
 * touch it.  dnp_comm_unique_id: rank 0 creates the 128-byte id and hands it to the other ranks by any means;
 * dnp_comm_init: every rank, on the device it will compute on; one communicator per process.
 *   dnp_field_grad_sharded_f32   field_grad (field_utils.py:61-116) with the TARGETS block-sharded (the first M % world
 *       ranks own one row more), sources replicated: this rank fills its block of E[M, 3]; gather != 0: the blocks are
 *       exchanged in one NCCL group so that every rank ends with the full array (needs ld_e == 3).
 *   dnp_edge_interactions_sharded_f32   the per-edge sums of field_edge_calculator (field_utils.py:145-160) with every
 *       target patch sliced over the ranks and ONE ncclAllReduce(sum) over out[n_edges]; patch_rows = offsets[P].
 *   dnp_lattice_mean_potential_sharded_f32   measure_mean_potential (field_utils.py:7-9) with the SOURCES sharded
 *       (1000 nodes are too few to shard) and ONE ncclAllReduce of the scalar.
 *   dnp_patch_propagate_sharded_f32   the greedy patch loop of dnp_patch_propagate_f32 (field_utils.py:239-265,
 *       311-335) with the TARGETS sharded by patch ownership ("inner field updates can be
 *       target-sharded, the arg-max then needs one all-reduce per step"): the cloud is replicated, every patch is owned
 *       by exactly one rank (`owned`: device uint8[P], this rank's patches), tgt_idx / n_tgt = the rows of this rank's
 *       patches.  Per step every rank adds the chosen patch's field to ITS rows, sums E.n over ITS patches, and ONE
 *       ncclAllReduce(sum) of the P-vector (zeros for patches owned elsewhere: the sum is exact, all ranks see identical
 *       values) precedes the arg-max, which every rank takes itself -- so all ranks visit and flip the same patches and
 *       end with identical pts.  E holds the final field on this rank's rows only.  All ranks pass the same
 *       start_patch.  workspace: dnp_patch_propagate_sharded_workspace_bytes(N, P, max_patch, n_tgt, patch_rows =
 *       offsets[P], n_owned = patches this rank owns; call it after dnp_comm_init: DNP_SLAB_MAX_MB is a budget per
 *       GPU, i.e. for ~P / world slabs) -- every rank must size its workspace with this call, because
 *       large loops take the slab schedule instead (all ranks or none), split by SOURCE patch: every rank computes and
 *       holds the slabs of the patches it owns (n_owned x patch_rows x 12 bytes) against all patch rows, ONE
 *       ncclAllReduce of the P x P matrix (rows of other ranks are zeros) replaces the per-step all-reduces, every
 *       rank takes the same decisions on it, and one all-reduce of the [T, 3] partial fields gives E.
 * workspace of the first three: dnp_sharded_workspace_bytes(S, M, P, patch_rows) device bytes.                         */
DNP_API int dnp_comm_unique_id(void* id128);
DNP_API int dnp_comm_init(const void* id128, int rank, int world);
DNP_API int dnp_comm_destroy(void);
DNP_API int dnp_comm_rank(int* rank, int* world);
DNP_API size_t dnp_sharded_workspace_bytes(int64_t S, int64_t M, int64_t P, int64_t patch_rows);
DNP_API int dnp_field_grad_sharded_f32(const float* src, int64_t S, int64_t ld_s,
                               const float* tgt, int64_t M, int64_t ld_t,
                               float eps, float* E, int64_t ld_e, int gather,
                               void* workspace, size_t workspace_bytes, void* stream);
DNP_API int dnp_edge_interactions_sharded_f32(const float* pts, int64_t ld,
                                      const int64_t* perm, const int64_t* offsets, int64_t P, int64_t patch_rows,
                                      const int64_t* edges, int64_t n_edges,
                                      float eps, float* out,
                                      void* workspace, size_t workspace_bytes, void* stream);
DNP_API int dnp_lattice_mean_potential_sharded_f32(const float* src, int64_t S, int64_t ld_s, int n, float* mean_out,
                                           void* workspace, size_t workspace_bytes, void* stream);
DNP_API size_t dnp_patch_propagate_sharded_workspace_bytes(int64_t N, int64_t P, int64_t max_patch, int64_t n_tgt,
                                                           int64_t patch_rows, int64_t n_owned);
DNP_API int dnp_patch_propagate_sharded_f32(float* pts, int64_t N, int64_t ld,
                                    const int64_t* perm, const int64_t* offsets_dev, const int64_t* offsets_host,
                                    int64_t P, const int64_t* rest_perm, const int64_t* rest_offsets_host,
                                    const int64_t* tgt_idx, int64_t n_tgt, const uint8_t* owned,
                                    int64_t start_patch, float eps, int diffuse,
                                    float* E, int64_t ld_e, int32_t* order_host, float* value_host,
                                    void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ allocation-free variant
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * dnp_lattice_mean_potential_noalloc_f32 = dnp_lattice_mean_potential_f32(phi_opt = NULL) -- the global-orientation
 * test's mean dipole potential over the n^3 lattice -- WITHOUT a workspace argument: the per-CTA partial sums live in a
 * file-scope `static __device__` array of the library (1200 floats = 4800 bytes, held by every GPU that loads the
 * library until the process exits).  The call makes no cudaMalloc / cudaFree / allocating library call; the caller
 * still owns src and mean_out (device).  Limits, rejected with DNP_ERR_INVALID and a message: S <=
 * DNP_NOALLOC_MAX_SOURCES, 1 <= n <= DNP_NOALLOC_MAX_LATTICE.  All calls on a device SHARE the array (it keeps the last
 * call's contents; nothing needs clearing, a call overwrites every slot it reads): calls that may overlap on the GPU
 * -- different streams, or host threads on different streams -- RACE on it and must be serialised by the caller.
 * Asynchronous on `stream`, like the workspace variant.                                                            */
#define DNP_NOALLOC_MAX_SOURCES ((int64_t)1 << 24)   /* 16 777 216 sources */
#define DNP_NOALLOC_MAX_LATTICE 255
DNP_API int dnp_lattice_mean_potential_noalloc_f32(const float* src, int64_t S, int64_t ld_s, int n, float* mean_out,
                                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DIPOLE_H_ */
