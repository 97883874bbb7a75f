// Test scaffolding (NOT a fallback, never loaded by the package): compiles the per-row arithmetic of the CUDA source
// jaxopt-3d-reconstruction_b200/csrc/common.cuh with g++ so that `-m "not gpu"` tests can check the very expressions the
// kernels evaluate (camera table, projection, Jacobian row, evaluator losses, 8x8 Cholesky) against the oracle on a
// machine without a GPU.  Thread / warp / PTX code is compiled out by B200BA_HOST_MATH.
#ifndef _GNU_SOURCE
#define _GNU_SOURCE  // sincos
#endif
#include <math.h>
#define B200BA_HOST_MATH 1
#define __device__
#define __forceinline__ inline
#define __restrict__
#include "../../jaxopt-3d-reconstruction_b200/csrc/common.cuh"

extern "C" {
int hm_tab_stride(void) { return TAB_STRIDE; }
// camera table (TAB_STRIDE doubles) from cam8 = [w, t, fx, fy] and k3 = [cx, cy, skew]
void hm_cam_table(const double* c8, const double* k3, double* T) { build_cam_table(c8, k3, T); }
// BA row: residual r = (dx^2 + dy^2) * inv_w2 and its Jacobian row, exactly as k_resjac forms them
void hm_ba_row(const double* T, const double* X, const double* obs, double inv_w2, double* r, double* dxdy, double* a, double* b) {
  const Proj pr = project(T, X[0], X[1], X[2]);
  const double dx = pr.u - obs[0], dy = pr.v - obs[1];
  *r = (dx * dx + dy * dy) * inv_w2;
  dxdy[0] = dx;
  dxdy[1] = dy;
  jac_row(T, pr, X[0], X[1], X[2], 2.0 * dx * inv_w2, 2.0 * dy * inv_w2, a, b);
}
// pose row as pose_eval (pose_kernels.cuh) forms it: loss 0 = L2, 1 = Cauchy
void hm_pose_row(const double* T, const double* X, const double* obs, double mask, double inv_w2, int loss, double* r, double* a) {
  const double m = mask * inv_w2;
  const Proj pr = project(T, X[0], X[1], X[2]);
  const double dx = pr.u - obs[0], dy = pr.v - obs[1];
  double ex, ey, b[3];
  if (loss == 1) {
    const double sx = 1.0 + dx * dx, sy = 1.0 + dy * dy;
    *r = (log(sx) + log(sy)) * m;
    ex = 2.0 * dx / sx * m;
    ey = 2.0 * dy / sy * m;
  } else {
    *r = (dx * dx + dy * dy) * m;
    ex = 2.0 * dx * m;
    ey = 2.0 * dy * m;
  }
  jac_row(T, pr, X[0], X[1], X[2], ex, ey, a, b);
}
// row of the 2-vector model exactly as k_resjac_v2 forms it
void hm_ba_row_v2(const double* T, const double* X, const double* obs, double inv_w, double* r, double* a, double* b) {
  const Proj pr = project(T, X[0], X[1], X[2]);
  double ex, ey;
  *r = v2_residual(pr, obs[0], obs[1], inv_w, ex, ey);
  jac_row(T, pr, X[0], X[1], X[2], ex, ey, a, b);
}
// row of the robustified 2-vector model exactly as k_resjac<JT, 2 | 3, .> forms it (loss 1 = Cauchy, 2 = soft L1; comp 0 = x row, 1 = y row)
void hm_ba_row_robust(const double* T, const double* X, const double* obs, double inv_w, double rho_b, int loss, int comp,
                      double* r, double* cterm, double* a, double* b) {
  const Proj pr = project(T, X[0], X[1], X[2]);
  double ex, ey;
  *r = loss == 1 ? robust_residual<2>(pr, obs[0], obs[1], inv_w, rho_b, comp, ex, ey, *cterm)
                 : robust_residual<3>(pr, obs[0], obs[1], inv_w, rho_b, comp, ex, ey, *cterm);
  jac_row(T, pr, X[0], X[1], X[2], ex, ey, a, b);
}
double hm_reproj_loss(int loss, double dx, double dy, double b) {
  switch (loss) {
    case 0: return reproj_loss<0>(dx, dy, b);
    case 1: return reproj_loss<1>(dx, dy, b);
    case 2: return reproj_loss<2>(dx, dy, b);
    default: return reproj_loss<3>(dx, dy, b);
  }
}
void hm_chol8_solve(double* A, double* b) { chol8_solve(A, b); }
}
