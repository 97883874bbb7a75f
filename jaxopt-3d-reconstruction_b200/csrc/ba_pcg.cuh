// Implicit-Schur matvec + pipelined block-Jacobi PCG, back-substitution, trial step and the LM
// accept/reject logic (K3b-K3d of SURVEY.md §7.2b).  Device-side control: every kernel reads the
// SolverState flags, so the host never has to synchronise inside an iteration.
//
// Row orders (see b200ba.cu, plan construction):
//   A = (point, camera)                 — the point-side kernels stream it
//   B = (point block, camera, point)    — the camera-side kernels stream it: per camera nPB RUNS
// A and B are permutations of each other INSIDE a point block (a few 100 K rows), so the per-row
// scalars exchanged between the two sides (sigma: B->A, omega: A->B) are gathered / scattered
// within an L2-resident window instead of across the whole 5 M-row array.
//
// One PCG iteration is TWO launches (Ghysels & Vanroose 2014, pipelined PCG: the matvec input
// m = Minv w is camera-local, so the camera kernel can finish matvec i, update all vectors and
// start matvec i+1 while its rows are still on chip):
//   k_ps : (block 0: reduce the dot partials -> alpha, beta, convergence flag)
//          omega_i = b_i . Vinv_p sum_{j in p} b_j sigma_j
//   k_cs : n_c = (U_c + mu) m_c - sum_{i in c} a_i omega_i        [rows read from HBM ONCE,
//          z,q,s,p,x,r,u,w recurrences; dot partials; m_c = Minv_c w_c    stashed in shared memory]
//          sigma_i = a_i . m_c
#pragma once
#include "ba_kernels.cuh"

// PCG vectors live camera-major: cv[c][slot][8] (CV_STRIDE doubles per camera), so everything the
// camera-local update needs is one contiguous 576 B record (one bulk copy in the ring kernel).
enum { CV_X = 0, CV_R, CV_U, CV_W, CV_Z, CV_Q, CV_S, CV_P, CV_M, CV_SLOTS };
#define CV_STRIDE (8 * CV_SLOTS)

enum {
  CS_C = 1,        // accumulate -sum a_i omega_i over the camera's rows
  CS_UX = 2,       // add (U_c + mu I) vin_c
  CS_WRITE_Y = 4,  // y_c = n_c
  CS_READ_Y = 8,   // n_c = y_c (multi-rank: after the all-reduce of the partial sums)
  CS_INIT = 16,    // PCG (re)start: w = n, z=q=s=p=0, x kept, dots, m = Minv w
  CS_ITER = 32,    // pipelined PCG recurrences with st->alpha/beta, dots, m = Minv w
  CS_A = 64,       // sigma_i = a_i . v  (v = m after INIT/ITER, else vin)
  CS_IGNORE_DONE = 128
};

#define PCG_STAG 25
#define CS_T 512
#define CS_NW (CS_T / 32)

// ---------------------------------------------------------------- K3b camera side (fused)
// Generic version (any camera size, fp32 or fp64 rows): one CTA per camera, rows through registers, optional
// shared-memory stash.  The hot single-rank path uses the TMA ring kernel in ba_cs_ring.cuh instead, the
// multi-rank iteration and every sigma-only pass the warp-task kernels in ba_cs_warp.cuh.
// Camera-local update shared by both kernels; executed by the 8-lane groups of one warp (lane&7 = component).
// sv: [0,72) the camera's cv record, [72,136) U_c, [136,200) Minv_c.  Returns m_k (the next matvec input).
__device__ __forceinline__ double cs_update(int flags, int k, int lane, double nk, const double* sv, double mu,
                                            double alpha, double beta, double* __restrict__ cvc,
                                            double* __restrict__ part4, int vin_slot) {
  const double vk = sv[8 * vin_slot + k];
  if (flags & CS_UX) {
    double s = mu * vk;
#pragma unroll
    for (int l = 0; l < 8; ++l) s += sv[72 + 8 * k + l] * sv[8 * vin_slot + l];
    nk += s;
  }
  double xk = sv[8 * CV_X + k], rk = sv[8 * CV_R + k], uk = sv[8 * CV_U + k], wk = sv[8 * CV_W + k];
  double zk, qk, sk, pk;
  if (flags & CS_INIT) {
    wk = nk; zk = qk = sk = pk = 0.0;
  } else {
    zk = nk + beta * sv[8 * CV_Z + k];
    qk = vk + beta * sv[8 * CV_Q + k];
    sk = wk + beta * sv[8 * CV_S + k];
    pk = uk + beta * sv[8 * CV_P + k];
    xk += alpha * pk;
    rk -= alpha * sk;
    uk -= alpha * qk;
    wk -= alpha * zk;
  }
  // m = Minv_c w : w of the camera lives in the 8-lane group
  double mk = 0.0;
  const int base = lane & ~7;
#pragma unroll
  for (int l = 0; l < 8; ++l) mk += sv[136 + 8 * k + l] * __shfl_sync(0xffffffffu, wk, base + l);
  double g = rk * uk, d = wk * uk, q2 = rk * rk;
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {
    g += __shfl_xor_sync(0xffffffffu, g, o);
    d += __shfl_xor_sync(0xffffffffu, d, o);
    q2 += __shfl_xor_sync(0xffffffffu, q2, o);
  }
  if (lane < 8) {
    cvc[8 * CV_X + k] = xk; cvc[8 * CV_R + k] = rk; cvc[8 * CV_U + k] = uk; cvc[8 * CV_W + k] = wk;
    cvc[8 * CV_Z + k] = zk; cvc[8 * CV_Q + k] = qk; cvc[8 * CV_S + k] = sk; cvc[8 * CV_P + k] = pk;
    cvc[8 * CV_M + k] = mk;
    if (k == 0) { part4[0] = g; part4[1] = d; part4[2] = q2; }
  }
  return mk;
}

template <typename JT>
__global__ void __launch_bounds__(CS_T, 2)
k_cs(int C, int nPB, int flags, int stash_rows, const SolverState* __restrict__ st,
     const int32_t* __restrict__ camtab, const JT* __restrict__ Jc_B, const double* __restrict__ omg,
     double* __restrict__ sig, const double* __restrict__ U, const double* __restrict__ Minv,
     int vin_slot, double* __restrict__ y, double* __restrict__ cv, double* __restrict__ part) {
  extern __shared__ __align__(128) unsigned char stash_raw[];
  JT* stash = reinterpret_cast<JT*>(stash_raw);
  __shared__ CamRuns runs;
  __shared__ double sred[8 * 32];
  __shared__ __align__(16) double svec[CV_STRIDE + 128];  // cv record | U_c | Minv_c
  __shared__ double sm_m[8];
  if (!(flags & CS_IGNORE_DONE) && st->pcg_done) return;
  const int c = blockIdx.x, tid = threadIdx.x;
  const bool upd = flags & (CS_INIT | CS_ITER);
  // asynchronous prefetch of everything the camera-local vector update needs
  if (tid < 200) {
    const double* src = nullptr;
    if (tid < 72) src = cv + (size_t)CV_STRIDE * c + tid;
    else if (tid < 136) { if (flags & CS_UX) src = U + 64 * (size_t)c + (tid - 72); }
    else if (upd) src = Minv + 64 * (size_t)c + (tid - 136);
    if (src) {
      const unsigned dst = (unsigned)__cvta_generic_to_shared(&svec[tid]);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src));
    } else {
      svec[tid] = 0.0;
    }
  }
  asm volatile("cp.async.commit_group;");
  const int L = runs.load(c, nPB, camtab);  // includes a __syncthreads
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (flags & CS_C) {
    int r0 = 0, r1 = 0;
    for (int l0 = tid; l0 < L; l0 += 2 * CS_T) {
      const int l1 = l0 + CS_T;
      const bool h1 = l1 < L;
      const size_t j0 = runs.locate(l0, r0);
      r1 = r0;
      const size_t j1 = h1 ? runs.locate(l1, r1) : j0;
      JT a0[8], a1[8];
      ldrow(Jc_B + 8 * j0, a0);
      const double w0 = __ldg(omg + j0);
      double w1 = 0.0;
      if (h1) { ldrow(Jc_B + 8 * j1, a1); w1 = __ldg(omg + j1); }
      if (l0 < stash_rows) strow(stash + 8 * (size_t)l0, a0);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] -= (double)a0[k] * w0;
      if (h1) {
        if (l1 < stash_rows) strow(stash + 8 * (size_t)l1, a1);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] -= (double)a1[k] * w1;
      }
      r0 = r1;
    }
    block_sum<8>(acc, sred);
  }
  asm volatile("cp.async.wait_all;");
  __syncthreads();
  if (tid < 32) {
    const int k = tid & 7;
    double nk = 0.0;
    if (flags & CS_C) {
#pragma unroll
      for (int l = 0; l < 8; ++l)
        if (l == k) nk = acc[l];
    }
    if (flags & CS_READ_Y) nk = y[8 * (size_t)c + k];
    if (upd) {
      const double mk = cs_update(flags, k, tid, nk, svec, st->mu, st->alpha, st->beta, cv + (size_t)CV_STRIDE * c,
                                  part + 4 * (size_t)c, vin_slot);
      if (tid < 8) sm_m[k] = mk;
    } else {
      const double vk = vin_slot >= 0 ? svec[8 * vin_slot + k] : 0.0;
      if (flags & CS_UX) {
        double s = st->mu * vk;
#pragma unroll
        for (int l = 0; l < 8; ++l) s += svec[72 + 8 * k + l] * svec[8 * vin_slot + l];
        nk += s;
      }
      if (tid < 8) sm_m[k] = vk;
    }
    if ((flags & CS_WRITE_Y) && tid < 8) y[8 * (size_t)c + k] = nk;
  }
  if (!(flags & CS_A)) return;
  __syncthreads();
  double m[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) m[k] = sm_m[k];
  int r0 = 0;
  const bool stashed = flags & CS_C;
  for (int l = tid; l < L; l += CS_T) {
    const size_t j = runs.locate(l, r0);
    JT a[8];
    if (stashed && l < stash_rows) ldrow_shared(stash + 8 * (size_t)l, a);
    else ldrow(Jc_B + 8 * j, a);
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += (double)a[k] * m[k];
    sig[j] = s;
  }
}

// ---------------------------------------------------------------- PCG scalars (one CTA)
// gamma = (r,u), delta = (w,u), rr = (r,r) from the per-camera partials of the last k_cs;
// beta = gamma/gamma_old, alpha = gamma/(delta - beta gamma/alpha_old)  (first iteration: 0, gamma/delta).
// Ordered reduction: the result does not depend on the launch geometry of anything else.
template <int NBATCH = 8>
__device__ __forceinline__ void pcg_scalars(int C, int it, SolverState* __restrict__ st,
                                            const double* __restrict__ part, double rtol, int maxiter,
                                            double* sm3 /*3*32*/) {
  // per-camera partials {gamma, delta, rr, -}: one 32 B load each, 8 in flight per thread (a serial loop over
  // 13682 cameras made block 0 the tail of the whole point kernel); per-thread summation order unchanged
  double v3[3] = {0.0, 0.0, 0.0};
  const int bd = blockDim.x;
  int i = threadIdx.x;
  for (; i + (NBATCH - 1) * bd < C; i += NBATCH * bd) {
    double q[NBATCH][4];
#pragma unroll
    for (int u = 0; u < NBATCH; ++u) ld256(part + 4 * (size_t)(i + u * bd), q[u]);
#pragma unroll
    for (int u = 0; u < NBATCH; ++u) { v3[0] += q[u][0]; v3[1] += q[u][1]; v3[2] += q[u][2]; }
  }
  for (; i < C; i += bd) {
    double q[4];
    ld256(part + 4 * (size_t)i, q);
    v3[0] += q[0]; v3[1] += q[1]; v3[2] += q[2];
  }
  block_sum<3>(v3, sm3);
  if (threadIdx.x == 0) {
    const double gamma = v3[0], delta = v3[1], rr = v3[2];
    const bool conv = rr <= rtol * rtol * st->rhs2;
    const bool bad = !(rr == rr) || !(delta == delta) || (st->status & 8);  // NaN, or a peer was lost (exchange timeout)
    // the recurrences of pipelined CG lose accuracy before the classical method does: no new minimum of
    // |r|^2 for PCG_STAG iterations = stagnation; the host then verifies / restarts from the true residual
    int since = st->since_best + 1;
    if (rr < st->rr_best) { st->rr_best = rr; since = 0; }
    st->since_best = since;
    const bool stag = since >= PCG_STAG;
    st->rr = rr;
    st->pcg_iters = it;
    if (conv || bad || stag || it >= maxiter) {
      st->pcg_done = 1;
      st->pcg_stop = bad ? 2 : conv ? 0 : stag ? 3 : 1;
      if (!conv && !bad && !stag) st->status |= 4;
    } else {
      double alpha, beta;
      if (st->pcg_first) {
        beta = 0.0;
        alpha = gamma / delta;
        st->pcg_first = 0;
      } else {
        beta = gamma / st->cg[0][0];
        alpha = gamma / (delta - beta * gamma / st->cg[0][1]);
      }
      st->cg[0][0] = gamma;
      st->cg[0][1] = alpha;
      st->alpha = alpha;
      st->beta = beta;
    }
  }
}

// After y = S x (plain matvec): true residual r = rhs - y, u = Minv r; part[4c+2] = |r_c|^2.
// 8 threads per camera.
__global__ void __launch_bounds__(256)
k_pcg_true_residual(int C, const double* __restrict__ rhs, const double* __restrict__ y, const double* __restrict__ Minv,
                    double* __restrict__ cv, double* __restrict__ part) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int c = e >> 3, k = e & 7;
  const bool act = c < C;
  const double rk = act ? rhs[e] - y[e] : 0.0;
  double uk = 0.0;
  const int base = (threadIdx.x & 31) & ~7;
#pragma unroll
  for (int l = 0; l < 8; ++l) {
    const double rl = __shfl_sync(0xffffffffu, rk, base + l);
    if (act) uk += Minv[64 * (size_t)c + 8 * k + l] * rl;
  }
  double q = rk * rk;
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  if (act) {
    cv[(size_t)CV_STRIDE * c + 8 * CV_R + k] = rk;
    cv[(size_t)CV_STRIDE * c + 8 * CV_U + k] = uk;
    if (k == 0) { part[4 * (size_t)c] = 0.0; part[4 * (size_t)c + 1] = 0.0; part[4 * (size_t)c + 2] = q; }
  }
}
// verdict on the true residual: converged -> done stays set (pcg_stop 0); else restart PCG from x
__global__ void __launch_bounds__(256)
k_pcg_verify(int C, SolverState* __restrict__ st, const double* __restrict__ part, double rtol) {
  __shared__ double sm[32];
  double v[1] = {0.0};
  for (int i = threadIdx.x; i < C; i += blockDim.x) v[0] += part[4 * (size_t)i + 2];
  block_sum<1>(v, sm);
  if (threadIdx.x == 0) {
    const double rr = v[0];
    st->rr_true = rr;
    if (rr <= rtol * rtol * st->rhs2 || !(rr == rr)) {
      st->pcg_done = 1;
      st->pcg_stop = (rr == rr) ? 0 : 2;
    } else {
      st->pcg_done = 0;
      st->pcg_first = 1;
      st->rr_best = rr;
      st->since_best = 0;
      st->rr = rr;
    }
  }
}

// ---------------------------------------------------------------- K3b point side
// Rows in A order are processed in CHUNKS of whole points (<= PS_T*PS_RPT rows and <= PS_T points,
// built once per plan), one CTA per chunk, PS_RPT rows per thread (row = r0 + tid + j*PS_T):
//   phase 1  sigma_i gathered from the B-ordered exchange array;  b_i sigma_i -> shared memory
//   phase 2  one thread per point sums its run in order, z_p = Vinv_p t_p
//   phase 3  omega_i = b_i . z_p scattered to the row's B position
// BACKSUB: phase 2 instead does v_p = Vinv_p (g_p - t_p) -> vp; part[2*chunk] = v_p.g_p, [+1] = |v_p|^2
// do_scalars bit 0: block 0 first turns the dot partials of the previous k_cs into alpha/beta/convergence;
// bit 1: run even if the PCG-finished flag is set (plain matvec).
#ifndef PS_T
#define PS_T 128
#endif
#ifndef PS_RPT
#define PS_RPT 4
#endif
#ifndef PS_PRELOAD
#define PS_PRELOAD 0  // 1: row positions / point parts / point blocks are loaded BEFORE the wait for the preceding kernel (PDL) — measured slower (70.5 vs 67.7 us at Venice) and an early-started point kernel costs 30 % of the whole solve, see b200ba.cu pdl_mask()
#endif
#ifndef PS_OWN
#define PS_OWN 1  // row -> local point through a shared-memory byte/short map written by the point lanes (no pt_A load)
#endif
#ifndef PS_GATHER
#define PS_GATHER 1  // sigma gather: 0 = ld.global.nc, 1 = ld.global.nc.L1::no_allocate (measured 72.5 -> 69.9 us at Venice)
#endif
#ifndef PS_SCATTER
#define PS_SCATTER 2  // omega scatter: 0 = st.global, 1 = st.global.cg (L2 only), 2 = st.global.L1::no_allocate
#endif
#ifdef B200BA_L2HINT
__device__ __forceinline__ double ps_gather_h(const double* p, uint64_t pol) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ void ps_scatter_h(double* p, double v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.f64 [%0], %1, %2;" ::"l"(p), "d"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ float ldg_h(const float* p, uint64_t pol) {
  float v;
  asm volatile("ld.global.nc.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ double ldg_h(const double* p, uint64_t pol) {
  double v;
  asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ int ldg_h(const int* p, uint64_t pol) {
  int v;
  asm volatile("ld.global.nc.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
  return v;
}
#endif
__device__ __forceinline__ double ps_gather(const double* p) {
#if PS_GATHER == 1
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
#else
  return __ldg(p);
#endif
}
__device__ __forceinline__ void ps_scatter(double* p, double v) {
#if PS_SCATTER == 1
  asm volatile("st.global.cg.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
#elif PS_SCATTER == 2
  asm volatile("st.global.L1::no_allocate.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
#else
  *p = v;
#endif
}
// z = Vinv t and omega = b . z with an explicit operation order (fma chains are never re-associated): shared by the chunk
// kernel and the tiled kernel (ba_ps_tile.cuh), which therefore agree to the last bit
__device__ __forceinline__ void ps_apply_vinv(const double* vi, double t0, double t1, double t2, double& z0, double& z1, double& z2) {
  z0 = fma(vi[2], t2, fma(vi[1], t1, vi[0] * t0));
  z1 = fma(vi[4], t2, fma(vi[3], t1, vi[1] * t0));
  z2 = fma(vi[5], t2, fma(vi[4], t1, vi[2] * t0));
}
__device__ __forceinline__ double ps_dot3(double b0, double b1, double b2, double z0, double z1, double z2) {
  return fma(b2, z2, fma(b1, z1, b0 * z0));
}
template <typename JT, bool BACKSUB>
#ifdef PS_MINB
__global__ void __launch_bounds__(PS_T, PS_MINB)
#else
__global__ void __launch_bounds__(PS_T)
#endif
k_ps(size_t N, int nchunks, int C, int it, int do_scalars, double rtol, int maxiter, SolverState* __restrict__ st,
     const double* __restrict__ cpart, const int32_t* __restrict__ chunk_row, const int32_t* __restrict__ chunk_pt,
     const int32_t* __restrict__ pt_ptr, const int32_t* __restrict__ pt_A, const int32_t* __restrict__ posB,
     const JT* __restrict__ Jp_A, const double* __restrict__ Vinv /*SoA: [k * Pst + p]*/, size_t Pst, const double* __restrict__ sig,
     double* __restrict__ omg, const double* __restrict__ gp, double* __restrict__ vp, double* __restrict__ part) {
  constexpr int CAP = PS_T * PS_RPT;
  __shared__ double sb[3][CAP];
  __shared__ double sz[3][PS_T];
  __shared__ unsigned short sown[PS_OWN ? CAP : 1];
  __shared__ double sred[3 * 32];
  const int ch = blockIdx.x, tid = threadIdx.x;
  // chunk geometry first: these loads overlap the round trip of the finished-flag below
  int r0 = 0, nrows = 0, p0 = 0, npts = 0;
  if (ch < nchunks) {
    r0 = __ldg(chunk_row + ch); nrows = __ldg(chunk_row + ch + 1) - r0;
    p0 = __ldg(chunk_pt + ch); npts = __ldg(chunk_pt + ch + 1) - p0;
  }
#ifdef B200BA_L2HINT
  const uint64_t pol_first = l2_policy_evict_first(), pol_last = l2_policy_evict_last();
#endif
  // PDL: everything up to griddep_wait() reads only what is constant during a solve (row positions, point parts of the
  // rows, point blocks) — it runs while the preceding camera-side kernel is still draining its last cameras
  griddep_launch();
#if !PS_PRELOAD
  griddep_wait();
  // CTA-uniform early exit: block 0 of this very launch may set the flag while other CTAs are starting
  if (!BACKSUB && !(do_scalars & 2) && __syncthreads_or(tid == 0 ? st->pcg_done : 0)) return;
  if (!BACKSUB && (do_scalars & 1) && ch == 0) {
    pcg_scalars(C, it, st, cpart, rtol, maxiter, sred);
    __syncthreads();
  }
  if (ch >= nchunks) return;
#endif
  const bool small = ch < nchunks && nrows <= CAP;
  double vi[6] = {0, 0, 0, 0, 0, 0}, g0 = 0, g1 = 0, g2 = 0;
  int beg = 0, end = 0;
  JT b[PS_RPT][3];
  int dst[PS_RPT], lp[PS_RPT];
  double s[PS_RPT];
#pragma unroll
  for (int j = 0; j < PS_RPT; ++j) {
    dst[j] = 0; lp[j] = 0; s[j] = 0.0;
    b[j][0] = b[j][1] = b[j][2] = (JT)0;
  }
  if (small) {
#pragma unroll
    for (int j = 0; j < PS_RPT; ++j) {
      const int i = tid + j * PS_T;
      const size_t row = (size_t)r0 + i;
      if (i < nrows) {
#ifdef B200BA_L2HINT
        dst[j] = ldg_h(posB + row, pol_first);
#pragma unroll
        for (int k = 0; k < 3; ++k) b[j][k] = ldg_h(Jp_A + (size_t)k * N + row, pol_first);
#else
        dst[j] = __ldg(posB + row);
#pragma unroll
        for (int k = 0; k < 3; ++k) b[j][k] = __ldg(Jp_A + (size_t)k * N + row);
#endif
        if (!BACKSUB && !PS_OWN) lp[j] = __ldg(pt_A + row) - p0;
      }
    }
    if (tid < npts) {
      const int p = p0 + tid;
      beg = __ldg(pt_ptr + p) - r0;
      end = __ldg(pt_ptr + p + 1) - r0;
#pragma unroll
      for (int k = 0; k < 6; ++k) vi[k] = __ldg(Vinv + (size_t)k * Pst + p);
      if (BACKSUB) { g0 = gp[3 * (size_t)p]; g1 = gp[3 * (size_t)p + 1]; g2 = gp[3 * (size_t)p + 2]; }
    }
  }
#if PS_PRELOAD
  griddep_wait();
  // CTA-uniform early exit: block 0 of this very launch may set the flag while other CTAs are starting
  if (!BACKSUB && !(do_scalars & 2) && __syncthreads_or(tid == 0 ? st->pcg_done : 0)) return;
  if (!BACKSUB && (do_scalars & 1) && ch == 0) {
    pcg_scalars(C, it, st, cpart, rtol, maxiter, sred);
    __syncthreads();
  }
  if (ch >= nchunks) return;
#endif
  double acc[2] = {0.0, 0.0};
  if (nrows <= CAP) {
#pragma unroll
    for (int j = 0; j < PS_RPT; ++j)
#ifdef B200BA_L2HINT
      if (tid + j * PS_T < nrows) s[j] = ps_gather_h(sig + dst[j], pol_first);
#else
      if (tid + j * PS_T < nrows) s[j] = ps_gather(sig + dst[j]);
#endif
#pragma unroll
    for (int j = 0; j < PS_RPT; ++j) {
      const int i = tid + j * PS_T;
      if (i < nrows) { sb[0][i] = (double)b[j][0] * s[j]; sb[1][i] = (double)b[j][1] * s[j]; sb[2][i] = (double)b[j][2] * s[j]; }
    }
    __syncthreads();
    if (tid < npts) {
      const int p = p0 + tid;
      double t0 = 0, t1 = 0, t2 = 0;
      for (int i = beg; i < end; ++i) {
        t0 += sb[0][i]; t1 += sb[1][i]; t2 += sb[2][i];
        if (!BACKSUB && PS_OWN) sown[i] = (unsigned short)tid;
      }
      if (BACKSUB) {
        t0 = g0 - t0; t1 = g1 - t1; t2 = g2 - t2;
        double z0, z1, z2;
        ps_apply_vinv(vi, t0, t1, t2, z0, z1, z2);
        vp[3 * (size_t)p] = z0; vp[3 * (size_t)p + 1] = z1; vp[3 * (size_t)p + 2] = z2;
        acc[0] = ps_dot3(z0, z1, z2, g0, g1, g2);
        acc[1] = ps_dot3(z0, z1, z2, z0, z1, z2);
      } else {
        double z0, z1, z2;
        ps_apply_vinv(vi, t0, t1, t2, z0, z1, z2);
        sz[0][tid] = z0; sz[1][tid] = z1; sz[2][tid] = z2;
      }
    }
    if (!BACKSUB) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < PS_RPT; ++j) {
        const int i = tid + j * PS_T;
        if (i < nrows) {
          int q = PS_OWN ? (int)sown[i] : lp[j];
#ifdef PS_DEBUG
          if (q < 0 || q >= npts || dst[j] < 0) { printf("k_ps: ch %d tid %d i %d q %d npts %d nrows %d dst %d\n", ch, tid, i, q, npts, nrows, dst[j]); q = 0; }
#endif
#ifdef B200BA_L2HINT
          ps_scatter_h(omg + dst[j], ps_dot3((double)b[j][0], (double)b[j][1], (double)b[j][2], sz[0][q], sz[1][q], sz[2][q]), pol_last);
#else
          ps_scatter(omg + dst[j], ps_dot3((double)b[j][0], (double)b[j][1], (double)b[j][2], sz[0][q], sz[1][q], sz[2][q]));
#endif
        }
      }
    }
  } else {
    // a single point with more than CAP observations: CTA-wide loop + ordered reduction
    const int p = p0;
    double t[3] = {0, 0, 0};
    for (int i = tid; i < nrows; i += PS_T) {
      const size_t row = (size_t)r0 + i;
      const double s = __ldg(sig + __ldg(posB + row));
#pragma unroll
      for (int k = 0; k < 3; ++k) t[k] += jget<JT>(Jp_A, N, k, row) * s;
    }
    block_sum<3>(t, sred);
    if (tid == 0) { sz[0][0] = t[0]; sz[1][0] = t[1]; sz[2][0] = t[2]; }
    __syncthreads();
    double t0 = sz[0][0], t1 = sz[1][0], t2 = sz[2][0];
    double vi[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) vi[k] = __ldg(Vinv + (size_t)k * Pst + p);
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
      for (int i = tid; i < nrows; i += PS_T) {
        const size_t row = (size_t)r0 + i;
        omg[__ldg(posB + row)] = jget<JT>(Jp_A, N, 0, row) * z0 + jget<JT>(Jp_A, N, 1, row) * z1 +
                                 jget<JT>(Jp_A, N, 2, row) * z2;
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

// PCG start-up reduction: rhs2 = |rhs|^2 from part[4c+2]
__global__ void __launch_bounds__(256)
k_pcg_init(int C, SolverState* __restrict__ st, const double* __restrict__ part) {
  __shared__ double sm[32];
  double v[1] = {0.0};
  for (int i = threadIdx.x; i < C; i += blockDim.x) v[0] += part[4 * (size_t)i + 2];
  block_sum<1>(v, sm);
  if (threadIdx.x == 0) {
    const double rr = v[0];
    st->rr = rr;
    st->rhs2 = rr;
    st->pcg_iters = 0;
    st->pcg_done = (rr == 0.0 || !(rr == rr)) ? 1 : 0;
    st->pcg_stop = (rr == rr) ? 0 : 2;
    st->pcg_first = 1;
    st->rr_best = rr;
    st->since_best = 0;
    st->rr_true = rr;
  }
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
// ---------------------------------------------------------------- K3d: trial step + LM logic
// trial = current - v; cam partials part[2c] = v_c.g_c, part[2c+1] = |v_c|^2
__global__ void k_trial_cams(int C, const double* __restrict__ cam8, const double* __restrict__ vc, int vc_stride,
                             const double* __restrict__ gc, double* __restrict__ cam8_trial,
                             double* __restrict__ part) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double vg = 0.0, vv = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const size_t o = 8 * (size_t)c + k;
    const double v = vc[(size_t)vc_stride * c + k];
    cam8_trial[o] = cam8[o] - v;
    vg += v * gc[o];
    vv += v * v;
  }
  part[2 * c] = vg;
  part[2 * c + 1] = vv;
}
// points are 32 B records {X, Y, Z, 0}; vp is P x 3
__global__ void k_trial_pts(int64_t n4, const double* __restrict__ pts4, const double* __restrict__ vp,
                            double* __restrict__ pts4_trial) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n4) pts4_trial[i] = (i & 3) == 3 ? 0.0 : pts4[i] - vp[3 * (i >> 2) + (i & 3)];
}

// trial cameras and trial points in ONE launch: the first cam_blocks CTAs do k_trial_cams' work, the rest k_trial_pts'
__global__ void __launch_bounds__(256)
k_trial(int C, int cam_blocks, const double* __restrict__ cam8, const double* __restrict__ vc, int vc_stride,
        const double* __restrict__ gc, double* __restrict__ cam8_trial, double* __restrict__ part, int64_t n4,
        const double* __restrict__ pts4, const double* __restrict__ vp, double* __restrict__ pts4_trial) {
  if ((int)blockIdx.x < cam_blocks) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double vg = 0.0, vv = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const size_t o = 8 * (size_t)c + k;
      const double v = vc[(size_t)vc_stride * c + k];
      cam8_trial[o] = cam8[o] - v;
      vg += v * gc[o];
      vv += v * v;
    }
    part[2 * c] = vg;
    part[2 * c + 1] = vv;
    return;
  }
  const int64_t i = (int64_t)(blockIdx.x - cam_blocks) * blockDim.x + threadIdx.x;
  if (i < n4) pts4_trial[i] = (i & 3) == 3 ? 0.0 : pts4[i] - vp[3 * (i >> 2) + (i & 3)];
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
__device__ __forceinline__ void lm_decide_body(SolverState* __restrict__ st, const double* __restrict__ scal) {
  if (scal[3] > 0.0) {  // a rank's peer-memory exchange timed out (k_status_flag, summed over the ranks): the step was built
    st->status |= 8;    // from incomplete sums -> reject on every rank, nothing is copied, the host returns B200BA_ENCCL
    st->accepted = 0;
    return;
  }
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
__global__ void k_lm_decide(SolverState* __restrict__ st, const double* __restrict__ scal) {
  if (threadIdx.x || blockIdx.x) return;
  lm_decide_body(st, scal);
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

// camera part of the rows back in the caller's order: out[order[i]*8 + k] = Jc_B[posB[i]*8 + k]
template <typename TI>
__global__ void k_unpermute_jc(int64_t N, const int32_t* __restrict__ order_ps, const int32_t* __restrict__ posB,
                               const TI* __restrict__ Jc_B, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int64_t o = order_ps[i], j = posB[i];
  for (int k = 0; k < 8; ++k) out[o * 8 + k] = (double)Jc_B[j * 8 + k];
}

// residuals back in the caller's order from the A-order row records (4 doubles each; r at double index r_off)
__global__ void k_unpermute_rec_r(int64_t N, const int32_t* __restrict__ order_ps, const double* __restrict__ rec, int r_off,
                                  double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) out[order_ps[i]] = rec[4 * i + r_off];
}
