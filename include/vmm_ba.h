/* vmm_ba.h — C ABI of the B200-native bundle-adjustment engine (libvmm_ba.so).
 *
 * This is the drop-in boundary for ONE path of schmidtniko/visual_marker_mapping: the joint
 * Levenberg-Marquardt refinement of camera and marker poses,
 *     TagReconstructor::doBundleAdjustment            src/TagReconstructor.cpp:646-784
 * with its cost functor
 *     TagReconstructionCostFunction::operator()       include/visual_marker_mapping/TagReconstructionCostFunction.h:101-159
 * and the reprojection statistics that gate it
 *     computeReprojectionErrorPerImg / PerTag         src/TagReconstructor.cpp:340-428.
 *
 * The reference has no FFI of its own (the seam is that one protected member function); the C++
 * facade in visual_marker_mapping_b200/csrc/host/ flattens its std::map state into the dense arrays
 * below and calls these entry points (INTEGRATION.md shows the binding).  Plain pointers and sizes
 * only; every function returns a vmm_ba_status (0 = ok) and never throws.  All pointer arguments
 * are HOST memory unless the name ends in _dev.  There is no CPU fallback: every entry point that
 * computes needs a CUDA device (sm_100a) and fails with VMM_BA_ERR_CUDA otherwise.
 *
 * Conventions
 *   quaternions (w,x,y,z)                               Camera.h:13, TagReconstructor.h:19
 *   camera pose maps world -> camera, marker pose maps marker -> world   TagReconstructionCostFunction.h:107-122
 *   tag observation = 4 corners (LL, LR, UR, UL), u,v interleaved: to_uv[8*k + 2*c + {0,1}]   TagReconstructor.h:44-52
 *   indices are dense 0-based positions; ids stay on the host side
 *   tangent vectors / gradients: camera i at [6i, 6i+6) = (t, delta), marker j at 6*Nc + [6j, 6j+6)
 *   local Jacobian of one corner: 2 x 12 row-major, columns (t_c, delta_c, t_m, delta_m)  — parameter
 *   block order of AddResidualBlock at src/TagReconstructor.cpp:720-722, delta = the 3-vector of
 *   ceres::QuaternionParameterization (src/TagReconstructor.cpp:661).
 */
#ifndef VMM_BA_H_
#define VMM_BA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VMM_BA_ABI_VERSION 2

typedef struct vmm_ba_engine* vmm_ba_handle;

typedef enum {
    VMM_BA_OK = 0,
    VMM_BA_ERR_INVALID_ARGUMENT = 1,
    VMM_BA_ERR_CUDA = 2,          /* no device / CUDA runtime failure */
    VMM_BA_ERR_STATE = 3,         /* call order (no problem set, ...) */
    VMM_BA_ERR_NCCL = 4,
    VMM_BA_ERR_NUMERIC = 5
} vmm_ba_status;

/* replaces: new ceres::HuberLoss(1.0) / nullptr / (commented) CauchyLoss(3)   src/TagReconstructor.cpp:721 */
typedef enum { VMM_BA_LOSS_NONE = 0, VMM_BA_LOSS_HUBER = 1, VMM_BA_LOSS_CAUCHY = 2 } vmm_ba_loss;

/* same numbering as ceres::TerminationType, which the reference prints at src/TagReconstructor.cpp:740 */
typedef enum { VMM_BA_CONVERGENCE = 0, VMM_BA_NO_CONVERGENCE = 1, VMM_BA_FAILURE = 2 } vmm_ba_termination;

/* JACOBI: inverse of B_i + lambda per camera.  SCHUR_JACOBI: inverse of the diagonal block of the reduced system.
 * TWO_LEVEL (default): JACOBI plus a coarse space of rigid motions of camera clusters (additive Schwarz); what
 * the "block-Jacobi-preconditioned CG" of the path needs to stay at tens of iterations on site-sized maps. */
typedef enum { VMM_BA_PRECOND_JACOBI = 0, VMM_BA_PRECOND_SCHUR_JACOBI = 1, VMM_BA_PRECOND_TWO_LEVEL = 2 } vmm_ba_preconditioner;

/* replaces: ceres::Solver::Options as configured at src/TagReconstructor.cpp:725-735
 * (everything the reference leaves at the Ceres default is a field here with that default). */
typedef struct {
    int max_num_iterations;               /* reference: 400 / 1500 */
    int loss_kind;                        /* vmm_ba_loss; reference: HUBER when robustify */
    double loss_scale;                    /* reference: 1.0 */
    double function_tolerance;            /* 1e-6  */
    double gradient_tolerance;            /* 1e-10 */
    double parameter_tolerance;           /* 1e-8  */
    double initial_trust_region_radius;   /* 1e4   */
    double max_trust_region_radius;       /* 1e16  */
    double min_trust_region_radius;       /* 1e-32 */
    double min_relative_decrease;         /* 1e-3  */
    double min_lm_diagonal;               /* 1e-6  */
    double max_lm_diagonal;               /* 1e32  */
    int max_num_consecutive_invalid_steps;/* 5     */
    int jacobi_scaling;                   /* 1     */
    /* linear solver: block-Jacobi PCG on the reduced camera system (markers eliminated) */
    int preconditioner;                   /* vmm_ba_preconditioner, default TWO_LEVEL */
    double pcg_tolerance;                 /* stop when sqrt(r.M^-1 r / r0.M^-1 r0) <= tol; default 1e-2 (5e-2 saved 40 % of the PCG
                                             iterations on C2/C3 but sent c3/8 into another local minimum; Ceres' eta default is 1e-1) */
    int pcg_max_iterations;               /* default 500 */
    int pcg_min_iterations;               /* default 1 */
    int parity_mode;                      /* 1: solve every LM step to pcg_parity_tolerance so that the
                                             iterates track an exact-Cholesky (Ceres) trajectory */
    double pcg_parity_tolerance;          /* default 1e-13 */
    int pcg_parity_max_iterations;        /* default 20000 */
    int verbose;                          /* 1: one line per LM iteration on stderr */
    int coarse_refresh_iterations;        /* TWO_LEVEL: the inverse coarse matrix of an earlier LM iteration is kept
                                             while the last PCG solve took at most this many iterations more than the
                                             first solve after the last refresh (a stale preconditioner changes the
                                             iteration count, not the solution); 0 = refresh every LM iteration.
                                             default 8 (the set-up costs about 17 PCG iterations at C3) */
    double pcg_confirm_tolerance;         /* the reference takes EXACT steps (sparse Cholesky), so its convergence tests
                                             (function / parameter tolerance) judge exact steps.  When one of them fires on
                                             an inexact step (pcg_tolerance above this value), the same damped system is
                                             solved again to this tolerance and the test is repeated on that step: the
                                             solve only stops where a (nearly) exact step would also have stopped it.
                                             default 1e-3; 0 = off (stop on the inexact step).  Unused in parity mode.  On one GPU the
                                             confirming solve CONTINUES the PCG iteration that produced the inexact step. */
    double pcg_polish_tolerance;          /* forcing term once the solve is in its polishing phase, i.e. after an accepted step
                                             lowered the cost by less than pcg_polish_threshold (relative); 0 = the same tolerance
                                             throughout; default 5e-2.  A loose tolerance is only dangerous while the steps are large. */
    double pcg_polish_threshold;          /* default 0.9: all but the first, large steps */
} vmm_ba_options;

/* replaces: ceres::Solver::Summary (the reference only prints termination_type / FullReport) */
typedef struct {
    int termination_type;                 /* vmm_ba_termination */
    double initial_cost;
    double final_cost;
    int num_successful_steps;
    int num_unsuccessful_steps;
    int num_iterations;                   /* LM iterations performed, iteration 0 excluded */
    int64_t num_linearizations;           /* residual+Jacobian(+Schur blocks) passes */
    int64_t num_cost_evaluations;         /* residual-only passes */
    int64_t pcg_iterations;               /* total over the solve (confirmation solves included) */
    int64_t kernel_launches;              /* launches of this library's kernels during the solve */
    double time_total_s;                  /* host wall clock of vmm_ba_solve */
    double time_linearize_s;              /* CUDA-event time on the engine's stream */
    double time_linear_solver_s;
    double time_cost_eval_s;
    double gradient_max_norm;
    double trust_region_radius;
    int num_confirmations;                /* convergence tests repeated on a (nearly) exact step (pcg_confirm_tolerance) */
    int reserved;
    char message[256];
} vmm_ba_summary;

/* Fill with the Ceres defaults in force at src/TagReconstructor.cpp:725-735. */
void vmm_ba_options_init(vmm_ba_options* options);

int vmm_ba_abi_version(void);

/* Number of visible CUDA devices (0 when there is none; never fails). */
int vmm_ba_device_count(void);

/* ---- lifetime ------------------------------------------------------------------------------ */

/* Single-GPU engine on CUDA device `device`. */
int vmm_ba_create(int device, vmm_ba_handle* out);

/* One engine per process per GPU; `nccl_unique_id` is the 128-byte ncclUniqueId that rank 0 got
 * from vmm_ba_nccl_unique_id() and broadcast out of band.  Cameras and their observations are
 * sharded over ranks, markers are replicated; partial marker blocks, gradients and the PCG
 * reductions travel by ncclAllReduce over NVLink. */
int vmm_ba_create_distributed(int device, int rank, int world_size, const void* nccl_unique_id, size_t id_bytes,
                              vmm_ba_handle* out);
int vmm_ba_nccl_unique_id(void* out_id, size_t id_bytes);
int vmm_ba_destroy(vmm_ba_handle h);

/* Error text of the last failing call on this handle (or of the last failing create when h == NULL). */
const char* vmm_ba_last_error(vmm_ba_handle h);

/* Run on the caller's CUDA stream (a cudaStream_t passed as void*) instead of the engine's own. */
int vmm_ba_set_stream(vmm_ba_handle h, void* cuda_stream);

/* ---- problem ------------------------------------------------------------------------------- */

/* replaces: camModel.getK() / distortionCoefficients copied per corner   src/TagReconstructor.cpp:710-711
 * dist = k1, k2, p1, p2, k3 (src/CameraUtilities.cpp:45-58). */
int vmm_ba_set_intrinsics(vmm_ba_handle h, double fx, double fy, double cx, double cy, const double dist[5]);

/* replaces: the problem assembly loop   src/TagReconstructor.cpp:663-724
 * Every rank passes the SAME full problem; a distributed engine keeps only its camera shard.
 * fixed_tag_index: the origin tag (SetParameterBlockConstant, :669-673) or -1.
 * Tag observations may come in any order; they are stably sorted by camera internally. */
int vmm_ba_set_problem(vmm_ba_handle h, int64_t num_cameras, const double* cam_q, const double* cam_t,
                       int64_t num_markers, const double* tag_q, const double* tag_t, const double* tag_wh,
                       int64_t fixed_tag_index, int64_t num_tag_obs, const int32_t* to_cam, const int32_t* to_tag,
                       const double* to_uv);

/* Bytes copied host -> device by this handle's set_problem / set_poses calls so far (bench.py: h2d_bytes_per_step).  A
 * vmm_ba_set_problem call whose sizes and index arrays equal the previous call's keeps every derived plan (device orders,
 * segment lists, camera clusters, shards) and copies only this rank's observations, the marker sizes and the poses. */
int64_t vmm_ba_h2d_bytes(vmm_ba_handle h);

/* Overwrite the current poses (same sizes as in set_problem). */
int vmm_ba_set_poses(vmm_ba_handle h, const double* cam_q, const double* cam_t, const double* tag_q,
                     const double* tag_t);

/* replaces: the in-place write-back through the pointers handed to Ceres   src/TagReconstructor.cpp:665-666,692-693
 * Full-size outputs on every rank (a distributed engine gathers the camera shards). */
int vmm_ba_get_poses(vmm_ba_handle h, double* cam_q, double* cam_t, double* tag_q, double* tag_t);

/* The trust-region step of the LAST Levenberg-Marquardt iteration of the last vmm_ba_solve, in the solver's tangent
 * coordinates (what Plus is applied with: translation increment, then the half-angle rotation increment of
 * ceres::QuaternionParameterization, src/TagReconstructor.cpp:661): delta_cam [6 * num_cameras], delta_tag
 * [6 * num_markers] (zeros for blocks without columns).  Whether that step was accepted does not matter.  For the
 * parity tests: an independent check that (J^T J + D) delta = -g at full problem size.  Full-size outputs on every
 * rank of a distributed engine. */
int vmm_ba_get_step(vmm_ba_handle h, double* delta_cam, double* delta_tag);

/* ---- the hot path -------------------------------------------------------------------------- */

/* replaces: ceres::Solve(options, &markerBAProblem, &summary)   src/TagReconstructor.cpp:737-738 */
int vmm_ba_solve(vmm_ba_handle h, const vmm_ba_options* options, vmm_ba_summary* summary);

/* replaces: one ProgramEvaluator::Evaluate over all residual blocks (AutoDiffCostFunction<...,2,3,4,3,4>
 * of TagReconstructionCostFunction.h:163-170 + QuaternionParameterization + loss Corrector).
 * Outputs (any may be NULL), in the CALLER's tag-observation order:
 *   cost        sum over corners of 0.5*rho(|r|^2)
 *   residuals   [4*num_tag_obs][2]        corrected residuals
 *   jacobians   [4*num_tag_obs][2][12]    corrected local Jacobians
 *   gradient    [6*(num_cameras+num_markers)]   J^T r, zeros for the fixed tag
 * Distributed engines return the cost/gradient of the whole problem and fill residuals/jacobians
 * only for their own shard (other entries untouched). */
int vmm_ba_evaluate(vmm_ba_handle h, int loss_kind, double loss_scale, double* cost, double* residuals,
                    double* jacobians, double* gradient);

/* Device-resident residual+Jacobian pass (what the HBM roofline is quoted on): launches the
 * materialising kernel `iterations` times on the engine's stream into device buffers
 * (component-major: res_dev[2][Nobs_local], jac_dev[24][Nobs_local]) and returns the CUDA-event
 * time per launch in milliseconds.  No host transfer. */
int vmm_ba_evaluate_device(vmm_ba_handle h, int loss_kind, double loss_scale, int iterations, double* ms_per_launch);

/* replaces: computeReprojectionErrorPerImg / PerTag   src/TagReconstructor.cpp:340-428 (projection of
 * src/CameraModel.cpp:6-26, quirk included).  per_tag_obs_sum[k] = sum over the 4 corners of the
 * Euclidean reprojection error of tag observation k at the current poses (caller's order). */
int vmm_ba_reprojection_error_sums(vmm_ba_handle h, double* per_tag_obs_sum);

/* replaces: the ceres::Covariance block of doBundleAdjustment   src/TagReconstructor.cpp:744-783
 * cov[9*j .. 9*j+8] = the 3x3 block (row-major) of (J^T J)^-1 that belongs to the translation of marker j at the
 * current poses, under the given loss (the reference calls it after the final solve, which runs without a loss);
 * zeros for the constant origin tag (as ceres::Covariance returns for constant blocks) and for unobserved markers.
 * Cameras are eliminated first.  Up to 1024 markers the marker-reduced matrix is inverted densely; larger maps go through
 * the breadth-first levels of the co-observation graph, which make that matrix block tridiagonal, and the exact two-sweep
 * recursion for the diagonal blocks of its inverse (csrc/ba_cov_block.cuh) — the reference calls this on the final, full
 * problem.  Single-GPU engines.  VMM_BA_ERR_NUMERIC when the normal equations are rank deficient (Ceres: Compute()
 * returns false); VMM_BA_ERR_INVALID_ARGUMENT when the level blocks do not fit the device memory. */
int vmm_ba_marker_translation_covariance(vmm_ba_handle h, int loss_kind, double loss_scale, double* cov);

typedef enum { VMM_BA_COV_AUTO = 0, VMM_BA_COV_DENSE = 1 /* <= 4096 markers */, VMM_BA_COV_BLOCK_TRIDIAGONAL = 2 } vmm_ba_covariance_method;
/* The same with the route chosen by the caller (tests compare the two). */
int vmm_ba_marker_translation_covariance_ex(vmm_ba_handle h, int loss_kind, double loss_scale, int method, double* cov);

/* Test entry (one GPU): the coarse matrix A_c = P^T S P of the two-level preconditioner at the current poses (damping
 * for `radius`, Jacobi scaling as in the first LM iteration) and the inverse the hand-written kernel
 * (csrc/ba_coarse_inv.cuh) computes from it.  Row-major, *leading_dimension x *leading_dimension doubles each (the
 * order rounded up to the kernel's block size; the padding block is the identity), the lower triangle of `matrix` is
 * what the inversion reads.  *order = 0 when the problem has no coarse space.  matrix == NULL: sizes only. */
int vmm_ba_debug_coarse_matrix(vmm_ba_handle h, int loss_kind, double loss_scale, double radius, int64_t capacity, double* matrix,
                               double* inverse, int* order, int* leading_dimension);

/* Target number of cameras per cluster of the two-level preconditioner's coarse space (default 192; the
 * coarse matrix is 6 x clusters squared and dense, the library caps it at 256 clusters).  Takes effect at
 * the next vmm_ba_set_problem. */
int vmm_ba_set_coarse_cluster_size(vmm_ba_handle h, int cameras_per_cluster);

/* ---- introspection for tests / benchmarks -------------------------------------------------- */

/* One product with the reduced camera system at the current poses (single-GPU handles only):
 *   y = (B + diag(lambda_c)) x - E (C + diag(lambda_m))^-1 E^T x,   x, y: 6 doubles per camera,
 * with B, C, E of the normal equations of src/TagReconstructor.cpp:699-724 under the given loss and the
 * LM damping that vmm_ba_solve would use at trust-region radius `radius` (Jacobi scaling on).
 * path 0: the one-pass kernel the PCG uses on one GPU; path 1: gather / marker solve / scatter (the
 * sharded form).  The parity tests check both against the product formed from the oracle's Jacobian. */
int vmm_ba_schur_multiply(vmm_ba_handle h, int loss_kind, double loss_scale, double radius, int path, const double* x, double* y);

/* Camera range [begin,end) — positions in the engine's internal camera sequence, which is the caller's on one GPU
 * and the spatial order of vmm_ba_shard_plan on a distributed engine — and tag-observation count owned by this engine. */
int vmm_ba_shard_info(vmm_ba_handle h, int64_t* cam_begin, int64_t* cam_end, int64_t* num_local_tag_obs);

/* The cut used by distributed engines on a camera sequence: `world_size` contiguous ranges balanced by
 * tag-observation count.  cam_begin has world_size+1 entries.  Needs no GPU.  (A distributed engine applies it to
 * its spatially renumbered camera sequence, see vmm_ba_shard_plan.) */
int vmm_ba_partition_cameras(int64_t num_cameras, int64_t num_tag_obs, const int32_t* to_cam, int world_size,
                             int64_t* cam_begin);

/* The sharding plan a distributed engine of `world_size` ranks applies to this problem (needs no GPU): cameras are
 * renumbered in a spatially coherent order (k-d leaves of the camera centres, `cameras_per_cluster` per leaf, 0 = the
 * default) and that sequence is cut into `world_size` ranges balanced by tag-observation count, so that most markers
 * are seen from one rank only.  cam_rank_out[i] = owner of the caller's camera i; marker_boundary_out[j] = -1 when all
 * blocks of marker j live on one rank (handled there in one pass), else its position in the compact boundary vector
 * that is all-reduced once per PCG iteration.  Either output may be NULL. */
int vmm_ba_shard_plan(int64_t num_cameras, const double* cam_q, const double* cam_t, int64_t num_markers, int64_t num_tag_obs,
                      const int32_t* to_cam, const int32_t* to_tag, int world_size, int cameras_per_cluster, int32_t* cam_rank_out,
                      int32_t* marker_boundary_out);

/* Number of kernel launches issued by this library on this handle since creation. */
int64_t vmm_ba_kernel_launches(vmm_ba_handle h);

/* Per-kernel device time: when enabled every kernel this library launches on the handle is bracketed
 * by a CUDA-event pair on the launching stream.  enable(…) also clears the counters.  Kernel indices
 * run over [0, vmm_ba_profile_num_kernels()); a kernel appears once it has been launched. */
int vmm_ba_profile_enable(vmm_ba_handle h, int on);
/* Phase clock of the persistent PCG kernel, summed over the launches since vmm_ba_profile_enable(h, 1): nanoseconds CTA 0
 * spent in the streamed product, the update + restriction, the coarse mat-vec, the direction [0..3] and in grid
 * barriers (waiting for the slowest CTA included) [4]; [5..7]: the streamed product per CTA, minimum / mean / maximum
 * over the grid (its load balance); [8], [9]: sharded engines / split markers — the exchange of the boundary vector and
 * the boundary scatter pass ([0] then covers the interior product and the boundary gather pass, [2] the exchange of
 * the restricted residual and the coarse mat-vec).  out must hold 10 doubles. */
int vmm_ba_pcg_phase_ns(vmm_ba_handle h, double* out10);
/* The same clock per CTA: nanoseconds CTA b (b < n, at most the grid of the kernel: one CTA per SM) spent in the streamed
 * product since vmm_ba_profile_enable(h, 1) — what the load-balance model of the segment lists is fitted on (tools/). */
int vmm_ba_pcg_cta_product_ns(vmm_ba_handle h, double* out, int n);
int vmm_ba_profile_num_kernels(void);
const char* vmm_ba_profile_kernel_name(int index);
int vmm_ba_profile_get(vmm_ba_handle h, int index, int64_t* launches, double* total_ms);

/* Measured DFMA throughput of the device (TFLOP/s, 2 flops per FMA) — the FP64 roofline denominator. */
int vmm_ba_measure_fp64_peak(vmm_ba_handle h, double* tflops);

/* Measured device-to-device copy bandwidth (GB/s, read+write bytes) with this library's own copy kernel. */
int vmm_ba_measure_copy_bandwidth(vmm_ba_handle h, size_t bytes, double* gbps);

/* ---- no allocation --------------------------------------------------------------------------------------------- */

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * The robust cost of a small problem — sum over corners of 0.5*rho(|r|^2), the number vmm_ba_evaluate returns in
 * *cost — computed on DEVICE buffers the caller owns, without a handle.  This call performs no cudaMalloc / cudaFree
 * and calls no allocating library routine: the scratch the engine allocates for this computation (prepared pose blocks,
 * one cost partial per block) lives in a static __device__ array of the library (499,712 bytes, held by every GPU
 * that loads the library until the process exits).  Hence the limits, checked before anything is launched
 * (VMM_BA_ERR_INVALID_ARGUMENT with the numbers in vmm_ba_last_error(NULL)):
 *     num_cameras <= VMM_BA_NO_ALLOC_MAX_CAMERAS, num_markers <= VMM_BA_NO_ALLOC_MAX_MARKERS,
 *     num_tag_obs <= VMM_BA_NO_ALLOC_MAX_TAG_OBS.
 * All d_* pointers are device pointers on `device`, with the layouts of vmm_ba_set_problem (cam_q [4*num_cameras] w,x,y,z;
 * cam_t [3*..]; tag_wh [2*num_markers]; to_uv [8*num_tag_obs]); the index arrays are NOT validated (they are on the device).
 * The four kernels are enqueued on cuda_stream (a cudaStream_t, NULL = the default stream) and the call returns without
 * synchronising; *d_cost (device, one double) is valid once the stream has reached that point.  loss_kind / loss_scale
 * as in vmm_ba_options.  The static scratch is shared by all calls in the process and keeps the last call's contents
 * (every call rewrites all it reads): calls that overlap on the GPU — different streams without a dependency — race. */
#define VMM_BA_NO_ALLOC_MAX_CAMERAS 4096
#define VMM_BA_NO_ALLOC_MAX_MARKERS 1024
#define VMM_BA_NO_ALLOC_MAX_TAG_OBS 131072
int vmm_ba_cost_no_alloc(int device, void* cuda_stream, double fx, double fy, double cx, double cy, const double dist[5],
                         int64_t num_cameras, const double* d_cam_q, const double* d_cam_t, int64_t num_markers,
                         const double* d_tag_q, const double* d_tag_t, const double* d_tag_wh, int64_t num_tag_obs,
                         const int32_t* d_to_cam, const int32_t* d_to_tag, const double* d_to_uv, int loss_kind,
                         double loss_scale, double* d_cost);

#ifdef __cplusplus
}
#endif
#endif /* VMM_BA_H_ */
