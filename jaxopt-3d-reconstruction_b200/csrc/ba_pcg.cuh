// Implicit-Schur matvec, block-Jacobi PCG vector kernels, back-substitution, trial step and the
// LM accept/reject logic (K3b-K3d of SURVEY.md §7.2b).  Device-side control: every kernel reads
// the SolverState flags, so the host never has to synchronise inside an iteration.
#pragma once
#include "ba_kernels.cuh"

// ---------------------------------------------------------------- K3b pass 1 (point side)
// t_p = sum_i b_i (a_i . x_c(i));  z_p = Vinv_p t_p;  w_i = b_i . z_p      (ps order)
// BACKSUB: instead v_p = Vinv_p (g_p - t_p) -> vp; part[2*block] = v_p.g_p, [2*block+1] = |v_p|^2
template <typename JT, bool BACKSUB>
__global__ void __launch_bounds__(256)
k_mv_pt(int P, const SolverState* __restrict__ st, const int32_t* __restrict__ pt_ptr,
        const int32_t* __restrict__ cam_ps, const JT* __restrict__ Jc_ps, const JT* __restrict__ Jp_ps,
        const double* __restrict__ Vinv, const double* __restrict__ x, double* __restrict__ w_ps,
        const double* __restrict__ gp, double* __restrict__ vp, double* __restrict__ part) {
  __shared__ double sm[64];
  if (!BACKSUB && st->pcg_done) return;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  double acc[2] = {0.0, 0.0};
  if (p < P) {
    const int beg = pt_ptr[p], end = pt_ptr[p + 1];
    double t0 = 0, t1 = 0, t2 = 0;
    for (int i = beg; i < end; ++i) {
      double a[8], b[3];
      load8<JT>(Jc_ps + 8 * (size_t)i, a);
      load3<JT>(Jp_ps + 4 * (size_t)i, b);
      const double2* xc = reinterpret_cast<const double2*>(x + 8 * (size_t)__ldg(cam_ps + i));
      const double2 x0 = xc[0], x1 = xc[1], x2 = xc[2], x3 = xc[3];
      const double s = a[0] * x0.x + a[1] * x0.y + a[2] * x1.x + a[3] * x1.y + a[4] * x2.x +
                       a[5] * x2.y + a[6] * x3.x + a[7] * x3.y;
      t0 += b[0] * s; t1 += b[1] * s; t2 += b[2] * s;
    }
    double vi[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) vi[k] = __ldg(Vinv + 6 * (size_t)p + k);
    if (BACKSUB) {
      const double g0 = gp[3 * (size_t)p], g1 = gp[3 * (size_t)p + 1], g2 = gp[3 * (size_t)p + 2];
      t0 = g0 - t0; t1 = g1 - t1; t2 = g2 - t2;
      const double z0 = vi[0] * t0 + vi[1] * t1 + vi[2] * t2;
      const double z1 = vi[1] * t0 + vi[3] * t1 + vi[4] * t2;
      const double z2 = vi[2] * t0 + vi[4] * t1 + vi[5] * t2;
      vp[3 * (size_t)p] = z0; vp[3 * (size_t)p + 1] = z1; vp[3 * (size_t)p + 2] = z2;
      acc[0] = z0 * g0 + z1 * g1 + z2 * g2;
      acc[1] = z0 * z0 + z1 * z1 + z2 * z2;
    } else {
      const double z0 = vi[0] * t0 + vi[1] * t1 + vi[2] * t2;
      const double z1 = vi[1] * t0 + vi[3] * t1 + vi[4] * t2;
      const double z2 = vi[2] * t0 + vi[4] * t1 + vi[5] * t2;
      for (int i = beg; i < end; ++i) {
        double b[3];
        load3<JT>(Jp_ps + 4 * (size_t)i, b);
        w_ps[i] = b[0] * z0 + b[1] * z1 + b[2] * z2;
      }
    }
  }
  if (BACKSUB) {
    block_sum<2>(acc, sm);
    if (threadIdx.x == 0) {
      part[2 * blockIdx.x] = acc[0];
      part[2 * blockIdx.x + 1] = acc[1];
    }
  }
}

// ---------------------------------------------------------------- K3b pass 2 (camera side)
// ypart_c = - sum_{j in cs run of c} a_j w[perm[j]].  FUSE (single rank): y_c = (U_c + mu I) p_c +
// ypart_c and part[c] = p_c . y_c, so no separate vector kernel is needed.
template <typename JT, bool FUSE>
__global__ void __launch_bounds__(256)
k_mv_cam(const SolverState* __restrict__ st, const int32_t* __restrict__ cam_ptr,
         const int32_t* __restrict__ perm_cs, const JT* __restrict__ Jc_cs,
         const double* __restrict__ w_ps, const double* __restrict__ U, const double* __restrict__ pv,
         double* __restrict__ y, double* __restrict__ part) {
  __shared__ double sm[8 * 32];
  if (st->pcg_done) return;
  const int c = blockIdx.x;
  const int beg = cam_ptr[c], end = cam_ptr[c + 1];
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int j = beg + threadIdx.x; j < end; j += blockDim.x) {
    double a[8];
    load8<JT>(Jc_cs + 8 * (size_t)j, a);
    const double w = __ldg(w_ps + __ldg(perm_cs + j));
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] -= a[k] * w;
  }
  block_sum<8>(acc, sm);
  if (threadIdx.x < 32) {
    if (FUSE) {
      // lanes 0..7: row k of (U+mu I) p
      const int k = threadIdx.x & 7;
      const double mu = st->mu;
      double s = 0.0;
#pragma unroll
      for (int l = 0; l < 8; ++l) s += U[64 * (size_t)c + 8 * k + l] * pv[8 * (size_t)c + l];
      const double pk = pv[8 * (size_t)c + k];
      s += mu * pk;
      double yk = 0.0;
#pragma unroll
      for (int l = 0; l < 8; ++l) {
        const double al = __shfl_sync(0xffffffffu, acc[l], 0);
        if (l == k) yk = al;
      }
      yk += s;
      double d = (threadIdx.x < 8) ? yk * pk : 0.0;
      d = warp_sum(d);
      if (threadIdx.x < 8) y[8 * (size_t)c + k] = yk;
      if (threadIdx.x == 0) part[c] = d;
    } else if (threadIdx.x == 0) {
#pragma unroll
      for (int k = 0; k < 8; ++k) y[8 * (size_t)c + k] = acc[k];
    }
  }
}

// multi-rank variant of the fused epilogue, run after the all-reduce of ypart
__global__ void k_pcg_y(int C, const SolverState* __restrict__ st, const double* __restrict__ U,
                        const double* __restrict__ pv, double* __restrict__ y, double* __restrict__ part) {
  if (st->pcg_done) return;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double mu = st->mu;
  double d = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    double s = mu * pv[8 * (size_t)c + k];
#pragma unroll
    for (int l = 0; l < 8; ++l) s += U[64 * (size_t)c + 8 * k + l] * pv[8 * (size_t)c + l];
    const double yk = y[8 * (size_t)c + k] + s;
    y[8 * (size_t)c + k] = yk;
    d += yk * pv[8 * (size_t)c + k];
  }
  part[c] = d;
}

// deterministic sum of n values by one CTA (all threads get the result)
__device__ __forceinline__ double cta_sum_strided(const double* __restrict__ in, int n, int stride,
                                                  double* sm) {
  double v[1] = {0.0};
  for (int i = threadIdx.x; i < n; i += blockDim.x) v[0] += in[(size_t)i * stride];
  block_sum<1>(v, sm);
  __shared__ double bc;
  if (threadIdx.x == 0) bc = v[0];
  __syncthreads();
  const double r = bc;
  __syncthreads();
  return r;
}

// ---------------------------------------------------------------- K3c: PCG vector updates
// Every CTA re-reduces the per-camera partials of p.y (ordered, identical in all CTAs), then
// alpha = rz / pAp; x += alpha p; r -= alpha y; z = Minv r; part2[2c] = r.z, part2[2c+1] = r.r
__global__ void __launch_bounds__(256)
k_pcg_update(int C, SolverState* __restrict__ st, const double* __restrict__ part,
             const double* __restrict__ Minv, const double* __restrict__ pv,
             const double* __restrict__ y, double* __restrict__ x, double* __restrict__ r,
             double* __restrict__ z, double* __restrict__ part2) {
  __shared__ double sm[32];
  if (st->pcg_done) return;
  const double pAp = cta_sum_strided(part, C, 1, sm);
  const double alpha = st->rz / pAp;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->pAp = pAp;
    st->alpha = alpha;
  }
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double rr[8];
  double r2 = 0.0, rz = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const size_t o = 8 * (size_t)c + k;
    x[o] += alpha * pv[o];
    rr[k] = r[o] - alpha * y[o];
    r[o] = rr[k];
    r2 += rr[k] * rr[k];
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    double s = 0.0;
#pragma unroll
    for (int l = 0; l < 8; ++l) s += Minv[64 * (size_t)c + 8 * k + l] * rr[l];
    z[8 * (size_t)c + k] = s;
    rz += s * rr[k];
  }
  part2[2 * c] = rz;
  part2[2 * c + 1] = r2;
}

// beta = rz_new / rz; p = z + beta p; convergence flag.  INIT: only reduce rz / |rhs|^2.
template <bool INIT>
__global__ void __launch_bounds__(256)
k_pcg_dir(int C, SolverState* __restrict__ st, const double* __restrict__ part2,
          const double* __restrict__ z, double* __restrict__ pv, double rtol, int maxiter) {
  __shared__ double sm[32];
  if (!INIT && st->pcg_done) return;
  const double rz_new = cta_sum_strided(part2, C, 2, sm);
  const double rr = cta_sum_strided(part2 + 1, C, 2, sm);
  if (INIT) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      st->rz = rz_new;
      st->rr = rr;
      st->rhs2 = rr;
      st->pcg_iters = 0;
      st->pcg_done = (rr == 0.0 || !(rr == rr)) ? 1 : 0;
    }
    return;
  }
  const double beta = rz_new / st->rz;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const size_t o = 8 * (size_t)c + k;
      pv[o] = z[o] + beta * pv[o];
    }
  }
  // last CTA-independent state write must come after every CTA has read st->rz: deferred to
  // k_pcg_commit (next kernel in stream order).
  (void)rr; (void)rtol; (void)maxiter;
}

__global__ void k_pcg_commit(int C, SolverState* __restrict__ st, const double* __restrict__ part2,
                             double rtol, int maxiter) {
  __shared__ double sm[32];
  if (st->pcg_done) return;
  const double rz_new = cta_sum_strided(part2, C, 2, sm);
  const double rr = cta_sum_strided(part2 + 1, C, 2, sm);
  if (threadIdx.x == 0) {
    st->beta = rz_new / st->rz;
    st->rz = rz_new;
    st->rr = rr;
    st->pcg_iters += 1;
    const bool conv = rr <= rtol * rtol * st->rhs2;
    const bool bad = !(rr == rr);
    if (conv || bad || st->pcg_iters >= maxiter) st->pcg_done = 1;
    if (!conv && !bad && st->pcg_iters >= maxiter) st->status |= 4;
  }
}

// ---------------------------------------------------------------- K3d: trial step + LM logic
// trial = current - v; cam partials part[2c] = v_c.g_c, part[2c+1] = |v_c|^2
__global__ void k_trial_cams(int C, const double* __restrict__ cam8, const double* __restrict__ vc,
                             const double* __restrict__ gc, double* __restrict__ cam8_trial,
                             double* __restrict__ part) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double vg = 0.0, vv = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const size_t o = 8 * (size_t)c + k;
    const double v = vc[o];
    cam8_trial[o] = cam8[o] - v;
    vg += v * gc[o];
    vv += v * v;
  }
  part[2 * c] = vg;
  part[2 * c + 1] = vv;
}
__global__ void k_trial_pts(int64_t n3, const double* __restrict__ pts, const double* __restrict__ vp,
                            double* __restrict__ pts_trial) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n3) pts_trial[i] = pts[i] - vp[i];
}

// generic ordered reduction: out[k] = sum_i in[i*stride + k], k < nout (single CTA)
__global__ void __launch_bounds__(1024)
k_reduce(const double* __restrict__ in, int n, int stride, int nout, double* __restrict__ out, int is_max_mask) {
  __shared__ double sm[32];
  for (int k = 0; k < nout; ++k) {
    if ((is_max_mask >> k) & 1) {
      double m = -1.0;
      for (int i = threadIdx.x; i < n; i += blockDim.x) m = fmax(m, in[(size_t)i * stride + k]);
      m = block_max1(m, sm);
      if (threadIdx.x == 0) out[k] = m;
    } else {
      const double s = cta_sum_strided(in + k, n, stride, sm);
      if (threadIdx.x == 0) out[k] = s;
    }
    __syncthreads();
  }
}

// scal: [0] loss_trial*2 (sum r^2), [1] vg (cams+points), [2] vv.  Gain ratio and damping update
// exactly as jaxopt's update / update_state_using_gain_ratio, with a select
// instead of the arithmetic blend.
__global__ void k_lm_decide(SolverState* __restrict__ st, const double* __restrict__ scal) {
  if (threadIdx.x || blockIdx.x) return;
  const double loss_next = 0.5 * scal[0];
  const double mu = st->mu;
  // delta = -v:  0.5 * delta^T (mu delta - g) = 0.5 * (mu |v|^2 + v.g)
  const double denom = 0.5 * (mu * scal[2] + scal[1]);
  const double rho = (st->loss - loss_next) / denom;
  st->loss_trial = loss_next;
  st->denom = denom;
  st->rho = rho;
  st->vg = scal[1];
  st->vv = scal[2];
  if (rho > 0.0) {
    st->accepted = 1;
    const double t = 2.0 * rho - 1.0;
    st->mu = mu * fmax(1.0 / 3.0, 1.0 - t * t * t);
    st->nu = 2.0;
    st->loss = loss_next;
  } else {
    st->accepted = 0;
    st->mu = mu * st->nu;
    st->nu = 2.0 * st->nu;
  }
}

__global__ void k_accept_copy(const SolverState* __restrict__ st, int64_t n, const double* __restrict__ src,
                              double* __restrict__ dst) {
  if (!st->accepted) return;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i];
}

// row gather for the API that returns rows in the caller's order: out[order[i]*w + k] = in[i*ws + k]
template <typename TI>
__global__ void k_unpermute_rows(int64_t N, const int32_t* __restrict__ order_ps, const TI* __restrict__ in,
                                 int ws, int w, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int64_t o = order_ps[i];
  for (int k = 0; k < w; ++k) out[o * w + k] = (double)in[i * ws + k];
}
