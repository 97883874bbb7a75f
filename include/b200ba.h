/* b200ba — C-ABI of the B200-native Levenberg–Marquardt bundle-adjustment / batched single-pose
 * hot path.  Plain pointers and sizes only; every array pointer is a DEVICE pointer owned by the
 * caller unless the name ends in `_host`; `stream` is a cudaStream_t passed as void*.
 * Every function returns 0 on success, non-zero otherwise (b200ba_last_error() has the text);
 * nothing throws.  There is no CPU fallback behind any entry point.
 *
 * The reference has no FFI layer: its boundary is two Python classes that hand arrays to
 * jaxopt.  Each entry point names the reference interface it stands in for (paths relative to
 * /root/reference/src/reconstruction/bundle_adjustment/).
 */
#ifndef B200BA_H
#define B200BA_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200ba_plan b200ba_plan;

enum { B200BA_OK = 0, B200BA_EINVAL = 1, B200BA_ECUDA = 2, B200BA_ENCCL = 3 };
/* per-solve status bits (b200ba_lm_result.status / pose status[]) */
enum { B200BA_ST_CONVERGED = 0, B200BA_ST_MAXITER = 1, B200BA_ST_NAN = 2, B200BA_ST_PCG_MAXITER = 4,
       B200BA_ST_XCHG_TIMEOUT = 8 /* multi-rank: a peer's partial sums never arrived; the solve returns B200BA_ENCCL */,
       B200BA_ST_DENSE_NOT_SPD = 16 /* dense reduced-camera Cholesky met a non-positive pivot (NaN step, like LAPACK through jax) */ };
enum { B200BA_LOSS_L2 = 0, B200BA_LOSS_CAUCHY = 1 }; /* loss.py:7-14 */

const char* b200ba_last_error(void);
int b200ba_version(void);

/* ---- problem structure (replaces the dense-padded builder the caller feeds to
 * JaxBundleAdjustment.optimize: benchmark_bundle_adjustment.py:45-106; here one row per VALID
 * observation, any order, duplicates allowed).  Sorts rows by (point, camera), builds the
 * camera-major permutation and CSR offsets, sizes all workspaces.  C is the GLOBAL camera
 * count; P/N are this rank's points / observations (multi-GPU: shard by point). */
int b200ba_plan_create(int C, int P, int64_t N, const int32_t* cam_idx, const int32_t* pt_idx,
                       int jac_f32 /*1: Jacobian rows stored fp32, fp64 accumulation (default of the facade); 0: fp64 rows*/,
                       void* stream, b200ba_plan** out);
int b200ba_plan_destroy(b200ba_plan* plan);
/* obs: N x 2 pixels in the SAME row order as cam_idx/pt_idx given to plan_create
 * (points_2d_all of bundle_adjustment.py:104-111 with mask==1 rows only). */
int b200ba_plan_set_observations(b200ba_plan* plan, const double* obs_xy, void* stream);
/* Optional: sum camera blocks / gradients across ranks (NCCL over NVLink) and exchange the per-PCG-iteration
 * partial camera sums through peer memory (CUDA IPC block per rank; falls back to an NCCL all-reduce when a peer
 * cannot be mapped).  Collective: every rank calls it for every plan.
 * unique_id_host: 128 bytes from b200ba_nccl_unique_id on rank 0, broadcast by the caller; the
 * communicator is created once per process and shared by later plans (pass NULL to re-use it). */
int b200ba_nccl_unique_id(char* unique_id_host /*128 B*/);
int b200ba_plan_set_comm(b200ba_plan* plan, int rank, int nranks, const char* unique_id_host);

/* Residual model of a plan.  0 (default): the reference's ONE scalar residual per observation, (dx^2 + dy^2)/w^2
 * (bundle_adjustment.py:17-21,74).  1: the 2-vector model of the reference's COLMAP comparator, r = (u - x, v - y)/w with the
 * principal point fixed (colmap/src/base/cost_functions.h:44-86; refine_principal_point = 0,
 * src/benchmark/colmap_benchmark/bundle_adjuster.py:25-38) run as the same pipeline over 2N rows: list every observation twice in
 * plan_create and pass (x, NaN) for its x row and (NaN, y) for its y row to plan_set_observations.  Objective
 * 0.5 sum r^2 = 0.5 sum |projection - observation|^2 / w^2 (TrivialLoss).  GPU parity: tests/test_gpu_vector2.py (fp64 rows are the
 * facade's default for this large-residual model). */
int b200ba_plan_set_residual_model(b200ba_plan* plan, int model);

/* The same 2-vector model under the ROBUSTIFIER the reference's COLMAP comparator selects
 * (BundleAdjustmentOptions::CreateLossFunction, colmap/src/optim/bundle_adjustment.cc:53-68): ceres::CauchyLoss(a),
 * rho(s) = b log(1 + s/b), or ceres::SoftLOneLoss(a), rho(s) = 2b (sqrt(1 + s/b) - 1), b = a^2, s = |projection - observation|^2 in
 * pixels^2.  Applied as Ceres applies a loss (Triggs' correction of the residual block and its Jacobian, Corrector in Ceres'
 * corrector.cc; both losses have rho'' < 0, where the correction is a scaling of the block and its rows by sqrt(rho'(s))); the LM
 * loss becomes Ceres' cost 0.5 sum rho(s) / w^2.  Row contract (differs from model 1): list every observation as an ADJACENT pair
 * of rows (x component, then y component) in plan_create (N even) and pass the FULL observation (x, y) on both rows.
 * loss_scale: a, in pixels, > 0.  Ceres itself is an external dependency of the reference (find_package, no sources): the
 * oracle restates its published algorithm (oracle/ba_oracle.py: ceres_rho, ceres_corrector); GPU parity tests/test_gpu_vector2.py. */
#define B200BA_ROBUST_CAUCHY 1
#define B200BA_ROBUST_SOFT_L1 2
int b200ba_plan_set_robust_loss(b200ba_plan* plan, int loss, double loss_scale);

/* ---- K1: residual + analytic Jacobian rows (BundleAdjustment.get_residuals,
 * bundle_adjustment.py:52-74 with reproject_point :17-21, plus the Jacobian jaxopt obtains by
 * jvp/vjp).  Outputs in the caller's ORIGINAL row order.  cam8: C x 8 = [w(3), t(3), fx, fy];
 * cam_const3: C x 3 = [cx, cy, skew] (prepare_params :93-102).  Any output may be NULL. */
int b200ba_residual_jacobian(b200ba_plan* plan, const double* cam8, const double* cam_const3,
                             const double* pts, double inv_w2, double* r /*N*/,
                             double* Jc /*N x 8*/, double* Jp /*N x 3*/,
                             double* cost_host /*0.5*sum r^2*/, void* stream);

/* ---- full LM solve (JaxBundleAdjustment.optimize, bundle_adjustment.py:104-120 =
 * jaxopt.LevenbergMarquardt(tol=1e-6, maxiter=100).run with its inner linear solve replaced by
 * Schur elimination of the points + block-Jacobi PCG on the reduced camera system). */
typedef struct {
  int maxiter;              /* 100  (bundle_adjustment.py:88) */
  double tol;               /* 1e-6 (bundle_adjustment.py:86), stop on ||J^T r||_2 <= tol */
  double damping_parameter; /* 1e-6 (jaxopt default) */
  double pcg_rtol;          /* ||res||_2 <= pcg_rtol * ||rhs||_2 */
  int pcg_maxiter;
  int fixed_pcg_iters;      /* >0: run exactly this many PCG iterations (benchmarking) */
  int verbose;
} b200ba_lm_opts;

typedef struct {
  int iter_num;
  int status;
  double loss, grad_l2, damping_factor, increase_factor;
  int64_t n_jac_evals, n_res_evals, n_pcg_iters_total;
  /* optional host arrays (may be NULL); capacity maxiter+1 / maxiter */
  double* loss_trajectory_host; /* [0]=initial loss, [k]=loss after update k */
  double* grad_trajectory_host;
  double* mu_trajectory_host;
  int* pcg_iters_host;
  int* accepted_host;
} b200ba_lm_result;

void b200ba_lm_default_opts(b200ba_lm_opts* opts);
int b200ba_lm_solve(b200ba_plan* plan, double* cam8_inout, const double* cam_const3,
                    double* pts_inout, double inv_w2, const b200ba_lm_opts* opts,
                    b200ba_lm_result* result, void* stream);

/* ---- building blocks exposed for parity tests and per-kernel benchmarks.  All operate on the
 * plan's internal state as left by b200ba_linearize. */
/* K1+K2: residual, Jacobian rows, U_c (C x 64 row-major, symmetric), g_c (C x 8),
 * V_p (P x 6: xx,xy,xz,yy,yz,zz), g_p (P x 3).  Output pointers may be NULL. */
int b200ba_linearize(b200ba_plan* plan, const double* cam8, const double* cam_const3,
                     const double* pts, double inv_w2, double* U, double* gc, double* V,
                     double* gp, double* scalars_host /*[loss, |g|^2, max diag]*/, void* stream);
/* (V+mu I)^-1, reduced rhs, block-Jacobi blocks; then y = S x for a caller vector (C x 8). */
int b200ba_schur_prepare(b200ba_plan* plan, double mu, double* rhs /*C x 8 or NULL*/, void* stream);
int b200ba_schur_matvec(b200ba_plan* plan, const double* x, double* y, void* stream);
/* Solve S v_c = rhs by PCG and back-substitute; v = [v_c (C x 8) | v_p (P x 3)]. */
int b200ba_schur_solve(b200ba_plan* plan, double rtol, int maxiter, double* vc, double* vp,
                       int* iters_host, void* stream);
/* trial cost 0.5*sum r^2 at given parameters (no Jacobian) */
int b200ba_cost(b200ba_plan* plan, const double* cam8, const double* cam_const3,
                const double* pts, double inv_w2, double* cost_host, void* stream);

/* The same trial cost WITHOUT a plan and WITHOUT any allocation, for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  This entry point has NO counterpart in the reference.
 * cam_idx / pt_idx / obs_xy (N x 2): the observations in the caller's
 * order, DEVICE arrays like cam8 (C x 8), cam_const3 (C x 3) and pts (P x 3); residual_model as in
 * b200ba_plan_set_residual_model; cost_dev: ONE double in device memory, written asynchronously on `stream` (0 for N = 0).
 * Scratch (camera tables + one partial sum per 1,024 rows) is a fixed 4.25 MiB file-scope device array that every GPU
 * loading the library holds until exit.  LIMITS: C <= 16,384 cameras and N <= 33,554,432 observations; beyond either the
 * call returns B200BA_EINVAL with the limit in b200ba_last_error() (use b200ba_cost with a plan).  Calls no cudaMalloc /
 * cudaFree / allocating routine.  All calls share the scratch: calls that OVERLAP on the GPU (different streams without an
 * ordering event) race; each call overwrites everything it reads, so nothing carries over from the previous call.  An index
 * outside [0, C) / [0, P) makes the cost NaN instead of faulting.  The sum order differs from b200ba_cost's (caller-order
 * blocks of 1,024 rows instead of camera runs): the two agree to rounding (~1e-13 relative), not bit for bit. */
int b200ba_cost_noalloc(int C, int P, int64_t N, const int32_t* cam_idx, const int32_t* pt_idx,
                        const double* obs_xy, const double* cam8, const double* cam_const3, const double* pts,
                        double inv_w2, int residual_model, double* cost_dev, void* stream);

/* ---- reprojection-error evaluator: the step after the solve that feeds every metric and plot of the reference
 * (Dataset.compute_reprojection_errors_alt, src/dataset/dataset.py:116-123, which calls
 * Camera.compute_projection_errors_alt, src/dataset/camera.py:78-85, once per camera in a Python loop).
 * Per observation e = loss(dx^2) + loss(dy^2) in pixels^2; loss 0 = TRIVIAL_LOSS (x), 1 = CAUCHY_LOSS (log(1 + x))
 * (src/dataset/loss_functions.py:7-17).  loss 2 / 3: the cost the reference's COLMAP comparator minimises, e = rho(dx^2 + dy^2)
 * of the 2-vector residual (colmap/src/base/cost_functions.h:44-86) with ceres::CauchyLoss(loss_scale) = b log(1 + s/b) or
 * ceres::SoftLOneLoss(loss_scale) = 2b (sqrt(1 + s/b) - 1), b = loss_scale^2 (colmap/src/optim/bundle_adjustment.cc:53-68;
 * Ceres reports half the sum); loss_scale is ignored for loss 0 / 1.  err_rows: N values in the caller's ORIGINAL row order (or NULL);
 * cam_sum: C ordered per-camera sums over this rank's rows (or NULL).  Asynchronous on `stream`. */
int b200ba_reprojection_errors(b200ba_plan* plan, const double* cam8, const double* cam_const3,
                               const double* pts, int loss, double loss_scale, double* err_rows /*N*/,
                               double* cam_sum /*C*/, void* stream);

/* ---- noise model of the benchmarks: Dataset.with_noise (src/dataset/dataset.py:43-63) applied IN PLACE to device arrays.
 * poses: C x 12 world-to-camera [R|t], row-major; intr: C x 5 = [fx, fy, cx, cy, skew]; pts: P x 3; obs: N x 2.
 * Every perturbation = magnitude x random unit direction in the positive orthant (dataset.py:32-41); intrinsics += U[0, intr_noise)
 * per entry (camera.py:30-47); pose: apply_move then apply_transform_3d(from_rotvec) (camera_pose.py:93-106).  move_cameras = 0
 * reproduces the reference as it BEHAVES (its apply_move is a no-op on W2C poses), 1 displaces the camera centre as intended.
 * Counter-based generator: the result depends on `seed` only, not on the launch geometry.  The optional *_vec / intr_add arrays
 * (P x 3, C x 3, C x 3, C x 5, N x 2; NULL = generate) replace the generated perturbations. */
int b200ba_with_noise(int C, int P, int64_t N, double* poses, double* intr, double* pts, double* obs, uint64_t seed,
                      double point3d_noise, double rot_noise, double trans_noise, double intr_noise, double point2d_noise,
                      int move_cameras, const double* pts_vec, const double* move_vec, const double* rot_vec, const double* intr_add,
                      const double* obs_vec, void* stream);

/* ---- K4: batched single-pose LM (JaxPoseOptimizer.optimize, pose_optimization.py:86-89 =
 * jit(vmap(LevenbergMarquardt(tol=1e-7, maxiter=100, solver="cholesky").run)) over
 * PoseOptimization.get_residuals :33-52).  One CTA per pose, early exit per pose.
 * params8: B x 8 in/out; const3: B x 3; points: B x Npts x 3; obs: B x Npts x 2; mask: B x Npts. */
int b200ba_pose_lm_batched(int B, int Npts, double* params8_inout, const double* const3,
                           const double* points, const double* obs, const double* mask,
                           double inv_w2, int loss, int maxiter, double tol,
                           double damping_parameter, int32_t* iter_num /*B*/,
                           double* loss_out /*B*/, double* grad_l2 /*B*/,
                           double* damping_out /*B or NULL*/, int32_t* status /*B*/, void* stream);

/* How the plan will run (diagnostics for bench.py / tests): out[8] = {point side tiled (ba_ps_tile.cuh), tile rows, tiles,
 * fused residual/Jacobian + camera blocks, point blocks of the B order, dense reduced-camera solve selected, point chunks,
 * split camera passes (multi-rank sequence)}. */
int b200ba_plan_info(b200ba_plan* plan, int* out /*[8]*/);

/* ---- measurement: per-kernel-class CUDA-event timing on the launching stream.  Classes:
 * 0 residual+Jacobian (K1), 1 point blocks, 2 camera blocks, 3 point inverses, 4 Schur camera
 * corrections, 5 block-Jacobi inverses, 6 matvec point pass, 7 matvec camera pass, 8 PCG vector
 * kernels, 9 back-substitution, 10 trial cost, 11 other, 12 reprojection-error evaluator, 13 dense S assembly, 14 dense Cholesky,
 * 15 dense triangular solves. */
int b200ba_profile_enable(b200ba_plan* plan, int on);
int b200ba_profile_get(b200ba_plan* plan, double* ms_out /*[16]*/, int64_t* count_out /*[16]*/);

/* number of kernels this library has launched since load (bench.py's gpu_launches) */
int64_t b200ba_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif
