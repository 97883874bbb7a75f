// Implicit-Schur matvec, block-Jacobi PCG vector kernels, back-substitution, trial step and the
// LM accept/reject logic (K3b-K3d of SURVEY.md §7.2b).  Device-side control: every kernel reads
// the SolverState flags, so the host never has to synchronise inside an iteration.
#pragma once
#include "ba_kernels.cuh"

// ---------------------------------------------------------------- K3b pass 1 (point side)
// Rows are processed in CHUNKS of whole points (<= MV_T*RPT rows and <= MV_T points, built once per
// plan), one CTA per chunk, RPT rows per thread (row = r0 + tid + j*MV_T: coalesced SoA loads, all
// issued before the first barrier so a CTA costs ~one DRAM round trip + the x gather):
//   phase 1  s_i = a_i . x_c(i);  b_i s_i -> shared memory
//   phase 2  one thread per point sums its run in order, z_p = Vinv_p t_p
//   phase 3  w_i = b_i . z_p, written at the row's CAMERA-SORTED position (w_cs[inv_cs[i]]) so that
//            pass 2 is a pure stream; the 40 MB/5M-row w array stays L2 resident between the passes.
// BACKSUB: phase 2 instead does v_p = Vinv_p (g_p - t_p) -> vp; part[2*chunk] = v_p.g_p, [+1] = |v_p|^2
#define MV_T 256
template <typename JT> struct MvCfg;
#ifndef MV_RPT_F32
#define MV_RPT_F32 2
#endif
template <> struct MvCfg<float>  { static constexpr int RPT = MV_RPT_F32; static constexpr int MINB = MV_RPT_F32 == 1 ? 5 : 3; };
template <> struct MvCfg<double> { static constexpr int RPT = 2; static constexpr int MINB = 3; };

template <typename JT, bool BACKSUB, bool WSCAT>
__global__ void __launch_bounds__(MV_T, MvCfg<JT>::MINB)
k_mv_rows(size_t N, const SolverState* __restrict__ st, const int32_t* __restrict__ chunk_row,
          const int32_t* __restrict__ chunk_pt, const int32_t* __restrict__ pt_ptr,
          const int32_t* __restrict__ cam_ps, const int32_t* __restrict__ pt_ps,
          const int32_t* __restrict__ inv_cs, const JT* __restrict__ Jc_ps, const JT* __restrict__ Jp_ps,
          const double* __restrict__ Vinv, const double* __restrict__ x, double* __restrict__ w_cs,
          const double* __restrict__ gp, double* __restrict__ vp, double* __restrict__ part) {
  constexpr int RPT = MvCfg<JT>::RPT;
  constexpr int CAP = MV_T * RPT;
  __shared__ double sb[3][CAP];
  __shared__ double sz[3][MV_T];
  __shared__ double sred[3 * 32];
  if (!BACKSUB && st->pcg_done) return;
  const int ch = blockIdx.x, tid = threadIdx.x;
  const int r0 = chunk_row[ch], nrows = chunk_row[ch + 1] - r0;
  const int p0 = chunk_pt[ch], npts = chunk_pt[ch + 1] - p0;
  double acc[2] = {0.0, 0.0};
  if (nrows <= CAP) {
    double vi[6] = {0, 0, 0, 0, 0, 0}, g0 = 0, g1 = 0, g2 = 0;
    int beg = 0, end = 0;
    if (tid < npts) {
      const int p = p0 + tid;
      beg = __ldg(pt_ptr + p) - r0;
      end = __ldg(pt_ptr + p + 1) - r0;
#pragma unroll
      for (int k = 0; k < 6; ++k) vi[k] = __ldg(Vinv + 6 * (size_t)p + k);
      if (BACKSUB) { g0 = gp[3 * (size_t)p]; g1 = gp[3 * (size_t)p + 1]; g2 = gp[3 * (size_t)p + 2]; }
    }
    JT a[RPT][8], b[RPT][3];  // kept in storage precision until used (register budget)
    int cam[RPT], lp[RPT], dst[RPT];
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
      const int i = tid + j * MV_T;
      const size_t row = (size_t)r0 + i;
      cam[j] = 0; lp[j] = 0; dst[j] = 0;
      if (i < nrows) {
        cam[j] = __ldg(cam_ps + row);
        if (!BACKSUB) { lp[j] = __ldg(pt_ps + row) - p0; dst[j] = WSCAT ? __ldg(inv_cs + row) : (int)row; }
#pragma unroll
        for (int k = 0; k < 8; ++k) a[j][k] = __ldg(Jc_ps + (size_t)k * N + row);
#pragma unroll
        for (int k = 0; k < 3; ++k) b[j][k] = __ldg(Jp_ps + (size_t)k * N + row);
      }
    }
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
      const int i = tid + j * MV_T;
      if (i < nrows) {
        double xv[8];
        ld256(x + 8 * (size_t)cam[j], xv);
        ld256(x + 8 * (size_t)cam[j] + 4, xv + 4);
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += (double)a[j][k] * xv[k];
        sb[0][i] = (double)b[j][0] * s; sb[1][i] = (double)b[j][1] * s; sb[2][i] = (double)b[j][2] * s;
      }
    }
    __syncthreads();
    if (tid < npts) {
      const int p = p0 + tid;
      double t0 = 0, t1 = 0, t2 = 0;
      for (int i = beg; i < end; ++i) { t0 += sb[0][i]; t1 += sb[1][i]; t2 += sb[2][i]; }
      if (BACKSUB) {
        t0 = g0 - t0; t1 = g1 - t1; t2 = g2 - t2;
        const double z0 = vi[0] * t0 + vi[1] * t1 + vi[2] * t2;
        const double z1 = vi[1] * t0 + vi[3] * t1 + vi[4] * t2;
        const double z2 = vi[2] * t0 + vi[4] * t1 + vi[5] * t2;
        vp[3 * (size_t)p] = z0; vp[3 * (size_t)p + 1] = z1; vp[3 * (size_t)p + 2] = z2;
        acc[0] = z0 * g0 + z1 * g1 + z2 * g2;
        acc[1] = z0 * z0 + z1 * z1 + z2 * z2;
      } else {
        sz[0][tid] = vi[0] * t0 + vi[1] * t1 + vi[2] * t2;
        sz[1][tid] = vi[1] * t0 + vi[3] * t1 + vi[4] * t2;
        sz[2][tid] = vi[2] * t0 + vi[4] * t1 + vi[5] * t2;
      }
    }
    if (!BACKSUB) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < RPT; ++j) {
        const int i = tid + j * MV_T;
        if (i < nrows)
          w_cs[dst[j]] = (double)b[j][0] * sz[0][lp[j]] + (double)b[j][1] * sz[1][lp[j]] + (double)b[j][2] * sz[2][lp[j]];
      }
    }
  } else {
    // a single point with more than CAP observations: CTA-wide loop + ordered reduction
    const int p = p0;
    double t[3] = {0, 0, 0};
    for (int i = tid; i < nrows; i += MV_T) {
      const size_t row = (size_t)r0 + i;
      const double2* xc = reinterpret_cast<const double2*>(x + 8 * (size_t)__ldg(cam_ps + row));
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const double2 xv = xc[k];
        s += jget<JT>(Jc_ps, N, 2 * k, row) * xv.x + jget<JT>(Jc_ps, N, 2 * k + 1, row) * xv.y;
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) t[k] += jget<JT>(Jp_ps, N, k, row) * s;
    }
    block_sum<3>(t, sred);
    if (tid == 0) { sz[0][0] = t[0]; sz[1][0] = t[1]; sz[2][0] = t[2]; }
    __syncthreads();
    double t0 = sz[0][0], t1 = sz[1][0], t2 = sz[2][0];
    double vi[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) vi[k] = __ldg(Vinv + 6 * (size_t)p + k);
    if (BACKSUB) {
      if (tid == 0) {
        const double g0 = gp[3 * (size_t)p], g1 = gp[3 * (size_t)p + 1], g2 = gp[3 * (size_t)p + 2];
        t0 = g0 - t0; t1 = g1 - t1; t2 = g2 - t2;
        const double z0 = vi[0] * t0 + vi[1] * t1 + vi[2] * t2;
        const double z1 = vi[1] * t0 + vi[3] * t1 + vi[4] * t2;
        const double z2 = vi[2] * t0 + vi[4] * t1 + vi[5] * t2;
        vp[3 * (size_t)p] = z0; vp[3 * (size_t)p + 1] = z1; vp[3 * (size_t)p + 2] = z2;
        acc[0] = z0 * g0 + z1 * g1 + z2 * g2;
        acc[1] = z0 * z0 + z1 * z1 + z2 * z2;
      }
    } else {
      const double z0 = vi[0] * t0 + vi[1] * t1 + vi[2] * t2;
      const double z1 = vi[1] * t0 + vi[3] * t1 + vi[4] * t2;
      const double z2 = vi[2] * t0 + vi[4] * t1 + vi[5] * t2;
      for (int i = tid; i < nrows; i += MV_T) {
        const size_t row = (size_t)r0 + i;
        w_cs[WSCAT ? (size_t)__ldg(inv_cs + row) : row] = jget<JT>(Jp_ps, N, 0, row) * z0 + jget<JT>(Jp_ps, N, 1, row) * z1 +
                                    jget<JT>(Jp_ps, N, 2, row) * z2;
      }
    }
  }
  if (BACKSUB) {
    block_sum<2>(acc, sred);
    if (tid == 0) {
      part[2 * ch] = acc[0];
      part[2 * ch + 1] = acc[1];
    }
  }
}

// ---------------------------------------------------------------- K3b pass 2 (camera side)
// ypart_c = - sum_{j in cs run of c} a_j w_cs[j]  (pure stream).  FUSE (single rank): y_c = (U_c + mu I) x_c +
// ypart_c and part[c] = x_c . y_c, so no separate vector kernel is needed.
template <typename JT, bool FUSE, bool WGATHER>
__global__ void __launch_bounds__(256)
k_mv_cam(const SolverState* __restrict__ st, const int32_t* __restrict__ cam_ptr,
         const int32_t* __restrict__ perm_cs, const JT* __restrict__ Jc_cs,
         const double* __restrict__ w_cs, const double* __restrict__ U, const double* __restrict__ xin,
         double* __restrict__ y, double* __restrict__ part) {
  __shared__ double sm[8 * 32];
  if (st->pcg_done) return;
  const int c = blockIdx.x;
  const int beg = cam_ptr[c], end = cam_ptr[c + 1];
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 4
  for (int j = beg + threadIdx.x; j < end; j += blockDim.x) {
    double a[8];
    load8<JT>(Jc_cs + 8 * (size_t)j, a);
    const double w = WGATHER ? __ldg(w_cs + __ldg(perm_cs + j)) : __ldg(w_cs + j);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] -= a[k] * w;
  }
  block_sum<8>(acc, sm);
  if (threadIdx.x < 32) {
    if (FUSE) {
      const int k = threadIdx.x & 7;  // lanes 0..7: row k of (U + mu I) x
      const double mu = st->mu;
      double s = 0.0;
#pragma unroll
      for (int l = 0; l < 8; ++l) s += U[64 * (size_t)c + 8 * k + l] * xin[8 * (size_t)c + l];
      const double xk = xin[8 * (size_t)c + k];
      s += mu * xk;
      double yk = 0.0;
#pragma unroll
      for (int l = 0; l < 8; ++l)
        if (l == k) yk = acc[l];
      yk += s;
      double d = (threadIdx.x < 8) ? yk * xk : 0.0;
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
                        const double* __restrict__ xin, double* __restrict__ y, double* __restrict__ part) {
  if (st->pcg_done) return;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double mu = st->mu;
  double d = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    double s = mu * xin[8 * (size_t)c + k];
#pragma unroll
    for (int l = 0; l < 8; ++l) s += U[64 * (size_t)c + 8 * k + l] * xin[8 * (size_t)c + l];
    const double yk = y[8 * (size_t)c + k] + s;
    y[8 * (size_t)c + k] = yk;
    d += yk * xin[8 * (size_t)c + k];
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

// ---------------------------------------------------------------- K3c: PCG vector kernel
// Chronopoulos-Gear form of preconditioned CG: ONE fused reduction point per iteration, so an
// iteration is three launches (this kernel, matvec pass 1, matvec pass 2) with no grid sync:
//   gamma = (r,u), delta = (w,u) with u = Minv r, w = S u        [partials from earlier kernels]
//   beta = gamma/gamma_old, alpha = gamma / (delta - beta gamma / alpha_old)   (it 0: beta 0, alpha gamma/delta)
//   p = u + beta p;  s = w + beta s;  x += alpha p;  r -= alpha s;  u = Minv r
// Every CTA re-reduces the per-camera partials in the same order (bit-identical alpha/beta in all
// CTAs).  Scalars carried between iterations are double-buffered in st->cg[it & 1] so no CTA can
// read a value another CTA of the same launch has already overwritten.  8 threads per camera.
__global__ void __launch_bounds__(256)
k_pcg_vec(int C, int it, SolverState* __restrict__ st, const double* __restrict__ part_gr /*2C: gamma_c, rr_c*/,
          const double* __restrict__ part_d /*C: delta_c*/, const double* __restrict__ Minv,
          const double* __restrict__ w, double* __restrict__ x, double* __restrict__ r, double* __restrict__ u,
          double* __restrict__ pv, double* __restrict__ sv, double* __restrict__ part_out /*2C*/, double rtol,
          int maxiter) {
  __shared__ double sm[32];
  if (st->pcg_done) return;
  __shared__ double sm3[3 * 32];
  __shared__ double bc3[3];
  double v3[3] = {0.0, 0.0, 0.0};
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    v3[0] += part_gr[2 * (size_t)i];
    v3[1] += part_gr[2 * (size_t)i + 1];
    v3[2] += part_d[i];
  }
  block_sum<3>(v3, sm3);
  if (threadIdx.x == 0) { bc3[0] = v3[0]; bc3[1] = v3[1]; bc3[2] = v3[2]; }
  __syncthreads();
  const double gamma = bc3[0], rr = bc3[1], delta = bc3[2];
  (void)sm;
  const bool conv = rr <= rtol * rtol * st->rhs2;
  const bool bad = !(rr == rr) || !(delta == delta);
  const bool stop = conv || bad || it >= maxiter;
  double alpha, beta;
  if (it == 0) {
    beta = 0.0;
    alpha = gamma / delta;
  } else {
    const double g_old = st->cg[it & 1][0], a_old = st->cg[it & 1][1];
    beta = gamma / g_old;
    alpha = gamma / (delta - beta * gamma / a_old);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->rr = rr;
    st->pcg_iters = it;
    if (stop) {
      st->pcg_done = 1;
      if (!conv && !bad) st->status |= 4;
    } else {
      st->cg[(it + 1) & 1][0] = gamma;
      st->cg[(it + 1) & 1][1] = alpha;
      st->alpha = alpha;
      st->beta = beta;
    }
  }
  if (stop) return;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;  // element index: camera e>>3, row e&7
  const int c = e >> 3, k = e & 7;
  double rk = 0.0, uk = 0.0;
  const bool act = c < C;
  if (act) {
    const double pk = u[e] + beta * pv[e];
    const double sk = w[e] + beta * sv[e];
    pv[e] = pk;
    sv[e] = sk;
    x[e] += alpha * pk;
    rk = r[e] - alpha * sk;
    r[e] = rk;
  }
  // u_k = sum_l Minv[k][l] r_l : r of the camera lives in the 8-lane group
  const int base = (threadIdx.x & 31) & ~7;
#pragma unroll
  for (int l = 0; l < 8; ++l) {
    const double rl = __shfl_sync(0xffffffffu, rk, base + l);
    if (act) uk += Minv[64 * (size_t)c + 8 * k + l] * rl;
  }
  if (act) u[e] = uk;
  double g = rk * uk, q = rk * rk;
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {
    g += __shfl_xor_sync(0xffffffffu, g, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (act && k == 0) {
    part_out[2 * c] = g;
    part_out[2 * c + 1] = q;
  }
}

// PCG start-up reduction: rhs2 = |rhs|^2
__global__ void __launch_bounds__(256)
k_pcg_init(int C, SolverState* __restrict__ st, const double* __restrict__ part_gr) {
  __shared__ double sm[32];
  const double rr = cta_sum_strided(part_gr + 1, C, 2, sm);
  if (threadIdx.x == 0) {
    st->rr = rr;
    st->rhs2 = rr;
    st->pcg_iters = 0;
    st->pcg_done = (rr == 0.0 || !(rr == rr)) ? 1 : 0;
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

// rows back in the caller's order: out[order[i]*w + k] = in[k*kstride + i]   (SoA ps arrays)
template <typename TI>
__global__ void k_unpermute_rows(int64_t N, const int32_t* __restrict__ order_ps, const TI* __restrict__ in,
                                 size_t kstride, int w, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int64_t o = order_ps[i];
  for (int k = 0; k < w; ++k) out[o * w + k] = (double)in[(size_t)k * kstride + i];
}
