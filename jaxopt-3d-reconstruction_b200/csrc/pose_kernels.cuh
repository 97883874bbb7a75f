// K4: batched single-pose Levenberg-Marquardt, one CTA per pose.
// Restates jit(vmap(LevenbergMarquardt(tol, maxiter, solver="cholesky").run)) over
// PoseOptimization.get_residuals (pose_optimization.py:33-52, 62-71): the correspondences of a
// pose are staged in shared memory once and stay there for every LM iteration; each iteration is
// ONE fused pass (residual + analytic Jacobian rows + J^T J + J^T r + loss, fp64), an ordered CTA
// reduction, and an 8x8 Cholesky solve by one thread.  Poses exit independently (no lock-step).
#pragma once
#include "common.cuh"

#define POSE_THREADS 128

struct PoseShared {
  double x[8], xn[8], k3[3];
  double T[TAB_STRIDE], Tn[TAB_STRIDE];
  double H[36], g[8];
  double loss, mu, nu, denom, err;
  int it, cont;
  double red[45 * 32];
};

template <int LOSS>
__device__ __forceinline__ void pose_eval(const double* __restrict__ T, int npts,
                                          const double* __restrict__ pts, const double* __restrict__ obs,
                                          const double* __restrict__ mask, double inv_w2, double* acc) {
#pragma unroll
  for (int k = 0; k < 45; ++k) acc[k] = 0.0;
  for (int n = threadIdx.x; n < npts; n += POSE_THREADS) {
    const double X = pts[3 * n], Y = pts[3 * n + 1], Z = pts[3 * n + 2];
    const double m = mask[n] * inv_w2;
    const Proj pr = project(T, X, Y, Z);
    const double dx = pr.u - obs[2 * n], dy = pr.v - obs[2 * n + 1];
    double r, ex, ey;
    if (LOSS == 1) {  // cauchy_loss, loss.py:12-14
      const double sx = 1.0 + dx * dx, sy = 1.0 + dy * dy;
      r = (log(sx) + log(sy)) * m;
      ex = 2.0 * dx / sx * m;
      ey = 2.0 * dy / sy * m;
    } else {  // l2_loss, loss.py:7-9
      r = (dx * dx + dy * dy) * m;
      ex = 2.0 * dx * m;
      ey = 2.0 * dy * m;
    }
    double a[8], b[3];
    jac_row(T, pr, X, Y, Z, ex, ey, a, b);
    int t = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
#pragma unroll
      for (int l = k; l < 8; ++l) acc[t++] += a[k] * a[l];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[36 + k] += a[k] * r;
    acc[44] += r * r;
  }
}

template <int LOSS>
__global__ void __launch_bounds__(POSE_THREADS)
k_pose_lm(int B, int npts, int stage_smem, double* __restrict__ params, const double* __restrict__ const3,
          const double* __restrict__ points, const double* __restrict__ obs, const double* __restrict__ mask,
          double inv_w2, int maxiter, double tol, double damping, int32_t* __restrict__ iter_num,
          double* __restrict__ loss_out, double* __restrict__ grad_out, double* __restrict__ mu_out,
          int32_t* __restrict__ status) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  PoseShared& S = *reinterpret_cast<PoseShared*>(smem_raw);
  double* stage = reinterpret_cast<double*>(smem_raw + sizeof(PoseShared));
  const int pose = blockIdx.x;
  if (pose >= B) return;
  const double* P = points + (size_t)pose * npts * 3;
  const double* O = obs + (size_t)pose * npts * 2;
  const double* Mk = mask + (size_t)pose * npts;
  if (stage_smem) {
    double* sp = stage;
    double* so = stage + 3 * (size_t)npts;
    double* sk = stage + 5 * (size_t)npts;
    for (int i = threadIdx.x; i < 3 * npts; i += POSE_THREADS) sp[i] = __ldg(P + i);
    for (int i = threadIdx.x; i < 2 * npts; i += POSE_THREADS) so[i] = __ldg(O + i);
    for (int i = threadIdx.x; i < npts; i += POSE_THREADS) sk[i] = __ldg(Mk + i);
    P = sp; O = so; Mk = sk;
  }
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 8; ++k) S.x[k] = params[8 * (size_t)pose + k];
#pragma unroll
    for (int k = 0; k < 3; ++k) S.k3[k] = const3[3 * (size_t)pose + k];
    build_cam_table(S.x, S.k3, S.T);
  }
  __syncthreads();
  double acc[45];
  pose_eval<LOSS>(S.T, npts, P, O, Mk, inv_w2, acc);
  block_sum<45>(acc, S.red);
  if (threadIdx.x == 0) {  // init_state
    double md = 0.0, g2 = 0.0;
    int t = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
#pragma unroll
      for (int l = k; l < 8; ++l) {
        if (l == k) md = fmax(md, acc[t]);
        S.H[t] = acc[t];
        ++t;
      }
#pragma unroll
    for (int k = 0; k < 8; ++k) { S.g[k] = acc[36 + k]; g2 += acc[36 + k] * acc[36 + k]; }
    S.loss = 0.5 * acc[44];
    S.mu = damping * md;
    S.nu = 2.0;
    S.err = sqrt(g2);
    S.it = 0;
    S.cont = (maxiter > 0) ? 1 : 0;
  }
  __syncthreads();
  while (S.cont) {
    if (threadIdx.x == 0) {
      double A[64], v[8];
      int t = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k)
#pragma unroll
        for (int l = k; l < 8; ++l) {
          A[8 * k + l] = S.H[t];
          A[8 * l + k] = S.H[t];
          ++t;
        }
#pragma unroll
      for (int k = 0; k < 8; ++k) { A[9 * k] += S.mu; v[k] = S.g[k]; }
      chol8_solve(A, v);
      double vv = 0.0, vg = 0.0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        S.xn[k] = S.x[k] - v[k];
        vv += v[k] * v[k];
        vg += v[k] * S.g[k];
      }
      S.denom = 0.5 * (S.mu * vv + vg);
      build_cam_table(S.xn, S.k3, S.Tn);
    }
    __syncthreads();
    pose_eval<LOSS>(S.Tn, npts, P, O, Mk, inv_w2, acc);
    block_sum<45>(acc, S.red);
    if (threadIdx.x == 0) {
      const double loss_next = 0.5 * acc[44];
      const double rho = (S.loss - loss_next) / S.denom;
      if (rho > 0.0) {
        double g2 = 0.0;
#pragma unroll
        for (int k = 0; k < 36; ++k) S.H[k] = acc[k];
#pragma unroll
        for (int k = 0; k < 8; ++k) { S.g[k] = acc[36 + k]; S.x[k] = S.xn[k]; g2 += acc[36 + k] * acc[36 + k]; }
        S.loss = loss_next;
        S.err = sqrt(g2);
        const double t = 2.0 * rho - 1.0;
        S.mu = S.mu * fmax(1.0 / 3.0, 1.0 - t * t * t);
        S.nu = 2.0;
      } else {
        S.mu = S.mu * S.nu;
        S.nu = 2.0 * S.nu;
      }
      S.it += 1;
      S.cont = (S.it < maxiter && S.err > tol) ? 1 : 0;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    bool nan = false;
#pragma unroll
    for (int k = 0; k < 8; ++k) { params[8 * (size_t)pose + k] = S.x[k]; nan |= !(S.x[k] == S.x[k]); }
    iter_num[pose] = S.it;
    loss_out[pose] = S.loss;
    grad_out[pose] = S.err;
    if (mu_out) mu_out[pose] = S.mu;
    int st = 0;
    if (S.err > tol) st |= 1;
    if (nan || !(S.loss == S.loss)) st |= 2;
    status[pose] = st;
  }
}
