// K4: batched single-pose Levenberg-Marquardt, one CTA per pose.
// Restates jit(vmap(LevenbergMarquardt(tol, maxiter, solver="cholesky").run)) over
// PoseOptimization.get_residuals (pose_optimization.py:33-52, 62-71): the correspondences of a
// pose are staged in shared memory once and stay there for every LM iteration; each iteration is
// ONE fused pass (residual + analytic Jacobian rows + J^T J + J^T r + loss, fp64), an ordered CTA
// reduction, and an 8x8 Cholesky solve by one thread.  Poses exit independently (no lock-step).
#pragma once
#include "common.cuh"

#define POSE_THREADS 128
#define POSE_WARPS (POSE_THREADS / 32)
#define POSE_WST 9  // doubles per stashed row: a[8], r (odd stride: conflict-free fragment reads)

// Round 2: the normal equations are accumulated on the FP64 TENSOR CORE.  Round 1 kept 45 fp64 accumulators per thread
// (J^T J upper triangle, J^T r, loss): 205 registers, 8 resident warps per SM, FP64 pipe 21 % busy (profiles/
// ncu_r2_summary.md).  Now every warp stashes the 32 rows it has just computed in a private shared-memory slab and feeds
// them to mma.sync.m8n8k4.f64 four rows at a time: lane (g, t) holds a_{row t}[g], which is the A fragment (8 components x
// 4 rows) and the B fragment (4 rows x 8 components) of J^T J at once; a second DMMA with B = [r, 0, ...] gives J^T r.
// Four accumulator registers per product instead of 90, no 45-value block reduction.
struct PoseShared {
  double x[8], xn[8], k3[3];
  double T[TAB_STRIDE], Tn[TAB_STRIDE];
  double Hf[64], g[8];           // J^T J (full 8x8, row-major), J^T r of the CURRENT point
  double Ht[73];                 // reduced result of the last evaluation: 64 + 8 + loss
  double loss, mu, nu, denom, err;
  int it, cont;
  double red[POSE_WARPS * 73];
  double wst[POSE_WARPS * 32 * POSE_WST];
};

__device__ __forceinline__ void pose_dmma(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// 8x8 SPD solve by the 8 low lanes of a warp (all 32 lanes must call): lane r holds row r of A (lower triangle used) and
// b_r; returns x_r.  Left-looking Cholesky with the pivot row travelling by shuffle, then the two substitutions — the same
// operation order per entry as chol8_solve (common.cuh), 16 registers instead of a 64-double local matrix in ONE thread
// (which had set the register count of the whole kernel).  NaN on a non-positive pivot, like LAPACK posv through jax.
__device__ __forceinline__ double warp_chol8_solve(double (&a)[8], double b, int lane) {
  const unsigned full = 0xffffffffu;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    double s = a[c];
#pragma unroll
    for (int q = 0; q < c; ++q) s -= a[q] * __shfl_sync(full, a[q], c);
    const double d = sqrt(__shfl_sync(full, s, c));
    a[c] = (lane == c) ? d : s / d;
  }
  double y = b;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const double yc = __shfl_sync(full, y, c) / __shfl_sync(full, a[c], c);
    if (lane == c) y = yc;
    else if (lane > c) y -= a[c] * yc;
  }
  double x = y;
#pragma unroll
  for (int c = 7; c >= 0; --c) {
    double t = (lane > c && lane < 8) ? a[c] * x : 0.0;
    t += __shfl_xor_sync(full, t, 1);
    t += __shfl_xor_sync(full, t, 2);
    t += __shfl_xor_sync(full, t, 4);
    const double xc = (__shfl_sync(full, x, c) - __shfl_sync(full, t, 0)) / __shfl_sync(full, a[c], c);
    if (lane == c) x = xc;
  }
  return x;
}

// One evaluation (residuals + analytic Jacobian rows + J^T J + J^T r + sum r^2) by the whole CTA; result in S.Ht
// (valid for every thread after the trailing __syncthreads).
template <int LOSS>
__device__ __forceinline__ void pose_eval(PoseShared& S, const double* __restrict__ T, int npts,
                                          const double* __restrict__ pts, const double* __restrict__ obs,
                                          const double* __restrict__ mask, double inv_w2) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  double* wst = S.wst + wid * 32 * POSE_WST;
  double u0 = 0.0, u1 = 0.0, v0 = 0.0, v1 = 0.0, loss = 0.0;
  for (int base = wid * 32; base < npts; base += POSE_THREADS) {
    const int n = base + lane;
    double a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[3], r = 0.0;
    if (n < npts) {
      const double X = pts[3 * n], Y = pts[3 * n + 1], Z = pts[3 * n + 2];
      const double m = mask[n] * inv_w2;
      const Proj pr = project(T, X, Y, Z);
      const double dx = pr.u - obs[2 * n], dy = pr.v - obs[2 * n + 1];
      double ex, ey;
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
      jac_row(T, pr, X, Y, Z, ex, ey, a, b);
    }
    loss += r * r;
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 8; ++k) wst[lane * POSE_WST + k] = a[k];
    wst[lane * POSE_WST + 8] = r;
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const double av = wst[(4 * q + t) * POSE_WST + g];
      const double rv = g == 0 ? wst[(4 * q + t) * POSE_WST + 8] : 0.0;
      pose_dmma(u0, u1, av, av);
      pose_dmma(v0, v1, av, rv);
    }
  }
  loss = warp_sum(loss);
  double* red = S.red + wid * 73;
  red[8 * g + 2 * t] = u0;
  red[8 * g + 2 * t + 1] = u1;
  if (t == 0) red[64 + g] = v0;
  if (lane == 0) red[72] = loss;
  __syncthreads();
  if (threadIdx.x < 73) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < POSE_WARPS; ++w) s += S.red[w * 73 + threadIdx.x];  // fixed order
    S.Ht[threadIdx.x] = s;
  }
  __syncthreads();
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
  pose_eval<LOSS>(S, S.T, npts, P, O, Mk, inv_w2);
  if (threadIdx.x == 0) {  // init_state
    double md = 0.0, g2 = 0.0;
#pragma unroll
    for (int k = 0; k < 64; ++k) S.Hf[k] = S.Ht[k];
#pragma unroll
    for (int k = 0; k < 8; ++k) { md = fmax(md, S.Ht[9 * k]); S.g[k] = S.Ht[64 + k]; g2 += S.Ht[64 + k] * S.Ht[64 + k]; }
    S.loss = 0.5 * S.Ht[72];
    S.mu = damping * md;
    S.nu = 2.0;
    S.err = sqrt(g2);
    S.it = 0;
    S.cont = (maxiter > 0) ? 1 : 0;
  }
  __syncthreads();
  while (S.cont) {
    if (threadIdx.x < 32) {  // warp 0: damped 8x8 solve, trial parameters, predicted reduction
      const int lane = threadIdx.x, r = lane & 7;
      double a[8];
#pragma unroll
      for (int l = 0; l < 8; ++l) a[l] = S.Hf[8 * r + l] + (l == r ? S.mu : 0.0);
      const double gr = S.g[r];
      const double v = warp_chol8_solve(a, gr, lane);
      double vv = lane < 8 ? v * v : 0.0, vg = lane < 8 ? v * gr : 0.0;
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) {
        vv += __shfl_xor_sync(0xffffffffu, vv, o);
        vg += __shfl_xor_sync(0xffffffffu, vg, o);
      }
      if (lane < 8) S.xn[lane] = S.x[lane] - v;
      if (lane == 0) S.denom = 0.5 * (S.mu * vv + vg);
      __syncwarp();
      if (lane == 0) build_cam_table(S.xn, S.k3, S.Tn);
    }
    __syncthreads();
    pose_eval<LOSS>(S, S.Tn, npts, P, O, Mk, inv_w2);
    if (threadIdx.x == 0) {
      const double loss_next = 0.5 * S.Ht[72];
      const double rho = (S.loss - loss_next) / S.denom;
      if (rho > 0.0) {
        double g2 = 0.0;
#pragma unroll
        for (int k = 0; k < 64; ++k) S.Hf[k] = S.Ht[k];
#pragma unroll
        for (int k = 0; k < 8; ++k) { S.g[k] = S.Ht[64 + k]; S.x[k] = S.xn[k]; g2 += S.Ht[64 + k] * S.Ht[64 + k]; }
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
