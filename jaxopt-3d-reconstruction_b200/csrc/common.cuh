// Shared device helpers for the b200ba kernels (sm_100a).
#pragma once
// B200BA_HOST_MATH: tests/host_math/ compiles the per-row arithmetic below (camera table, projection, Jacobian row,
// evaluator losses, 8x8 Cholesky) with g++ so the CPU test suite can check the CUDA SOURCE against known answers on a machine without a GPU.
// It is test scaffolding only: everything that touches threads, warps, barriers or PTX is compiled out of that build,
// and nothing in the library or the Python package ever defines the macro.
#ifndef B200BA_HOST_MATH
#include <cuda_runtime.h>
#endif
#include <stdio.h>
#include <stdint.h>

#define TAB_STRIDE 32  // doubles per camera-table row (256 B, sector aligned)
// table layout: R[0..8] row-major, t[9..11], M[12..20] row-major, fx 21, fy 22, cx 23, cy 24, skew 25

// Device-side solver state, shared by all LM / PCG kernels (one per plan).
struct SolverState {
  double mu, nu, loss, loss_trial, grad2, maxdiag, rho, denom;
  double rz, rz_new, pAp, rr, rhs2, alpha, beta, vg, vv;
  double cg[2][2];  // double-buffered (gamma, alpha) of the Chronopoulos-Gear recurrence
  double rr_best, rr_true;
  int accepted, pcg_done, pcg_iters, status;
  int pcg_first, pcg_stop /*0 converged, 1 maxiter, 2 NaN, 3 stagnated*/, since_best, pad_;
  int lm_it, dn_epoch;  // device-resident LM loop (CUDA graph): iterations done; epoch of the dense solver's tile flags
};

#ifndef B200BA_HOST_MATH
// Programmatic dependent launch (PDL): a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may become
// resident while its predecessor in the stream is still draining, once every CTA of the predecessor has executed
// griddep_launch(); it must not touch anything the predecessor writes before griddep_wait() (= predecessor complete and
// its memory visible).  Both are no-ops in an ordinary launch.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Deterministic CTA-wide sum of NV values per thread. Result valid in thread 0 (all threads of
// warp 0 hold it). `sm` needs NV * 32 doubles. All threads must call.
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* sm) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int k = 0; k < NV; ++k) v[k] = warp_sum(v[k]);
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) sm[k * 32 + wid] = v[k];
  }
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double x = (lane < nw) ? sm[k * 32 + lane] : 0.0;
      v[k] = warp_sum(x);
    }
  }
}

__device__ __forceinline__ double block_max1(double v, double* sm) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) sm[wid] = v;
  __syncthreads();
  double x = -1.0;
  if (wid == 0) {
    x = (lane < nw) ? sm[lane] : -1.0;
    x = warp_max(x);
  }
  return x;
}

#endif  // !B200BA_HOST_MATH

// Projection chain shared by the BA and pose kernels.
struct Proj {
  double q0, q1, q2, iz, u, v;
};
__device__ __forceinline__ Proj project(const double* __restrict__ T, double X, double Y, double Z) {
  Proj p;
  p.q0 = fma(T[0], X, fma(T[1], Y, fma(T[2], Z, T[9])));
  p.q1 = fma(T[3], X, fma(T[4], Y, fma(T[5], Z, T[10])));
  p.q2 = fma(T[6], X, fma(T[7], Y, fma(T[8], Z, T[11])));
  p.iz = 1.0 / p.q2;
  p.u = (T[21] * p.q0 + T[25] * p.q1) * p.iz + T[23];
  p.v = T[22] * p.q1 * p.iz + T[24];
  return p;
}

// Jacobian row from d r/d(u,v) = (ex, ey): a[8] = camera part [w, t, fx, fy], b[3] = point part.
__device__ __forceinline__ void jac_row(const double* __restrict__ T, const Proj& p, double X, double Y,
                                        double Z, double ex, double ey, double* a, double* b) {
  const double fx = T[21], fy = T[22], s = T[25];
  const double e0 = ex * fx * p.iz;
  const double e1 = (ex * s + ey * fy) * p.iz;
  const double e2 = -(ex * (fx * p.q0 + s * p.q1) + ey * fy * p.q1) * p.iz * p.iz;
  b[0] = T[0] * e0 + T[3] * e1 + T[6] * e2;  // R^T eq
  b[1] = T[1] * e0 + T[4] * e1 + T[7] * e2;
  b[2] = T[2] * e0 + T[5] * e1 + T[8] * e2;
  const double c0 = Y * b[2] - Z * b[1];  // X x b
  const double c1 = Z * b[0] - X * b[2];
  const double c2 = X * b[1] - Y * b[0];
  a[0] = c0 * T[12] + c1 * T[15] + c2 * T[18];  // (X x b)^T M
  a[1] = c0 * T[13] + c1 * T[16] + c2 * T[19];
  a[2] = c0 * T[14] + c1 * T[17] + c2 * T[20];
  a[3] = e0;
  a[4] = e1;
  a[5] = e2;
  a[6] = ex * p.q0 * p.iz;
  a[7] = ey * p.q1 * p.iz;
}

// Camera table row from [w,t,fx,fy] + [cx,cy,skew]: Rodrigues as utils.py:21-28 (no small-angle
// guard: NaN at w=0 like the reference) and M = (w w^T + (R^T - I)[w]x)/|w|^2.
__device__ __forceinline__ void build_cam_table(const double* __restrict__ c8, const double* __restrict__ k3,
                                                double* __restrict__ T) {
  const double w0 = c8[0], w1 = c8[1], w2 = c8[2];
  const double th2 = w0 * w0 + w1 * w1 + w2 * w2;
  const double th = sqrt(th2);
  const double k0 = w0 / th, k1 = w1 / th, k2 = w2 / th;
  double sn, cs;
  sincos(th, &sn, &cs);
  const double oc = 1.0 - cs;
  double R[9];
  R[0] = cs + oc * k0 * k0;
  R[1] = oc * k0 * k1 - sn * k2;
  R[2] = oc * k0 * k2 + sn * k1;
  R[3] = oc * k1 * k0 + sn * k2;
  R[4] = cs + oc * k1 * k1;
  R[5] = oc * k1 * k2 - sn * k0;
  R[6] = oc * k2 * k0 - sn * k1;
  R[7] = oc * k2 * k1 + sn * k0;
  R[8] = cs + oc * k2 * k2;
#pragma unroll
  for (int i = 0; i < 9; ++i) T[i] = R[i];
  T[9] = c8[3];
  T[10] = c8[4];
  T[11] = c8[5];
  // A = R^T - I ; [w]x = [[0,-w2,w1],[w2,0,-w0],[-w1,w0,0]]
  const double A[9] = {R[0] - 1.0, R[3], R[6], R[1], R[4] - 1.0, R[7], R[2], R[5], R[8] - 1.0};
  const double w[3] = {w0, w1, w2};
  const double Wx[9] = {0.0, -w2, w1, w2, 0.0, -w0, -w1, w0, 0.0};
  const double inv = 1.0 / th2;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double m = w[i] * w[j];
#pragma unroll
      for (int k = 0; k < 3; ++k) m += A[i * 3 + k] * Wx[k * 3 + j];
      T[12 + i * 3 + j] = m * inv;
    }
  T[21] = c8[6];
  T[22] = c8[7];
  T[23] = k3[0];
  T[24] = k3[1];
  T[25] = k3[2];
}

// In-place Cholesky solve of the 8x8 SPD system A x = b held in registers/local arrays by ONE
// thread.  A is full row-major (only the lower triangle is read). NaN on a non-positive pivot,
// like LAPACK posv through jax (sqrt of a negative).
__device__ __forceinline__ void chol8_solve(double* A, double* b) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    double d = A[j * 8 + j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= A[j * 8 + k] * A[j * 8 + k];
    d = sqrt(d);
    A[j * 8 + j] = d;
    const double id = 1.0 / d;
#pragma unroll
    for (int i = j + 1; i < 8; ++i) {
      double s = A[i * 8 + j];
#pragma unroll
      for (int k = 0; k < j; ++k) s -= A[i * 8 + k] * A[j * 8 + k];
      A[i * 8 + j] = s * id;
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    double s = b[i];
#pragma unroll
    for (int k = 0; k < i; ++k) s -= A[i * 8 + k] * b[k];
    b[i] = s / A[i * 8 + i];
  }
#pragma unroll
  for (int i = 7; i >= 0; --i) {
    double s = b[i];
#pragma unroll
    for (int k = i + 1; k < 8; ++k) s -= A[k * 8 + i] * b[k];
    b[i] = s / A[i * 8 + i];
  }
}

// Row of the 2-vector residual model (k_resjac_v2): the observation (x, NaN) marks the x component, (NaN, y) the y component.
// Returns r = (u - x)/w or (v - y)/w and dr/d(u,v) in (ex, ey).
__device__ __forceinline__ double v2_residual(const Proj& pr, double ox, double oy, double inv_w, double& ex, double& ey) {
  const bool xrow = oy != oy;
  ex = xrow ? inv_w : 0.0;
  ey = xrow ? 0.0 : inv_w;
  return (xrow ? pr.u - ox : pr.v - oy) * inv_w;
}

// MODEL 2 / 3: the 2-vector model under ceres::CauchyLoss(a) / ceres::SoftLOneLoss(a), the robustifiers the reference's
// COLMAP comparator selects (colmap/src/optim/bundle_adjustment.cc:53-68), applied the way Ceres applies them: Triggs'
// correction of the residual block and its Jacobian (Corrector, corrector.cc).  Both losses have rho'' < 0, where the
// corrector reduces to scaling the block r = (dx, dy)/w and its Jacobian rows by sqrt(rho'(s)), s = |r|^2; the LM loss is
// Ceres' cost 0.5 sum rho(s).  Row contract: every observation as an ADJACENT pair of rows (x component, then y), both with
// the full observation; the stable sorts and the 4-aligned run starts of the plan keep the pair on an (even, odd) B position,
// so the component is the parity of the row's B position.  rho_b = (a / w)^2.  cterm: the row's share of 2 x cost.
template <int MODEL>
__device__ __forceinline__ double robust_residual(const Proj& pr, double ox, double oy, double inv_w, double rho_b, int comp,
                                                  double& ex, double& ey, double& cterm) {
  const double dx = (pr.u - ox) * inv_w, dy = (pr.v - oy) * inv_w;
  const double s = dx * dx + dy * dy;
  const double q = s / rho_b;
  double rho, rho1;
  if (MODEL == 2) {
    rho = rho_b * log1p(q);        // b log(1 + s/b)
    rho1 = 1.0 / (1.0 + q);
  } else {
    const double t = sqrt(1.0 + q);
    rho = 2.0 * s / (t + 1.0);     // 2b (sqrt(1 + s/b) - 1)
    rho1 = 1.0 / t;
  }
  const double sq = sqrt(rho1);
  cterm = 0.5 * rho;
  ex = comp ? 0.0 : sq * inv_w;
  ey = comp ? sq * inv_w : 0.0;
  return sq * (comp ? dy : dx);
}
// Per-observation value of the reprojection-error evaluator (k_reproj_err): LOSS 0 TRIVIAL_LOSS, 1 CAUCHY_LOSS
// (src/dataset/loss_functions.py:7-17, applied per coordinate), 2 / 3 ceres::CauchyLoss / SoftLOneLoss of |r|^2 with b = a^2
// (colmap/src/optim/bundle_adjustment.cc:53-68).
template <int LOSS>
__device__ __forceinline__ double reproj_loss(double dx, double dy, double b) {
  const double s2 = dx * dx + dy * dy;
  if (LOSS == 0) return s2;
  if (LOSS == 1) return log(1.0 + dx * dx) + log(1.0 + dy * dy);
  if (LOSS == 2) return b * log(1.0 + s2 / b);
  return 2.0 * b * (sqrt(1.0 + s2 / b) - 1.0);
}

#ifndef B200BA_HOST_MATH
// ---------------------------------------------------------------- mbarrier + bulk async copy (TMA, 1-D)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16 B aligned; completes on `bar` (complete_tx)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// L2 eviction-priority hints (opt-in build: -DB200BA_L2HINT): the sigma / omega exchange arrays (80 MB at Venice, written by
// one kernel of the PCG iteration and read by the other) are stored evict_last and the streamed data (Jacobian rows, indices,
// the exchange values once consumed) is loaded evict_first, so that the exchange stays in the 126 MB L2 between the kernels.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}
template <int NT>
__device__ __forceinline__ void named_bar_sync(int id) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(NT) : "memory");
}
#endif  // !B200BA_HOST_MATH
