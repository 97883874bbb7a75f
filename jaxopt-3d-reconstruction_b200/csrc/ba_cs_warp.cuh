// Camera side of the Schur matvec when the rows of a camera are FEW per launch: the multi-rank (point-sharded)
// PCG iteration, where the all-reduce of the partial sums separates matvec tail and head, and the plain
// sigma = rows . v passes of prepare / back-substitution / verification.
//
// The ring kernel (ba_cs_ring.cuh) walks one camera at a time per CTA; with 1/8 of the rows per rank its
// per-camera latency chain (record wait, reduce, barrier, update) is all that is left.  Here the unit of work is
// a WARP TASK: <= CW_ROWS consecutive local rows of one camera (cameras in descending size, tasks of a camera
// adjacent), thousands of tasks in flight, no block-level synchronisation at all:
//   k_cs_warp  CS_C  : partial n_c = -sum a_i omega_i  (cameras cut into several tasks: ordered last-arriver sum)
//              CS_A  : sigma_i = a_i . v_c,  v = slot vin_slot of the camera record
//   k_cs_vec         : the camera-local vector work between the two (8 lanes per camera): n_c from y (all-reduced),
//                      + (U_c + mu) v_c, PCG recurrences / start, dot partials, m_c = Minv_c w_c
#pragma once
#include "ba_pcg.cuh"

#define CW_T 256
#ifndef CW_ROWS
#define CW_ROWS 512
#endif

// 8 per-lane accumulators -> lane (l & ~3 == component pattern) holds the warp total of ONE component:
// 18 shuffles instead of 40.  Returns the total of component comp8(lane) (valid in every lane).
__device__ __forceinline__ int comp8(int lane) { return ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1); }
__device__ __forceinline__ double warp_tsum8(const double (&acc)[8], int lane) {
  double v4[4], v2[2], v1;
  {
    const bool up = lane & 16;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const double send = up ? acc[j] : acc[j + 4], keepv = up ? acc[j + 4] : acc[j];
      v4[j] = keepv + __shfl_xor_sync(0xffffffffu, send, 16);
    }
  }
  {
    const bool up = lane & 8;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const double send = up ? v4[j] : v4[j + 2], keepv = up ? v4[j + 2] : v4[j];
      v2[j] = keepv + __shfl_xor_sync(0xffffffffu, send, 8);
    }
  }
  {
    const bool up = lane & 4;
    const double send = up ? v2[0] : v2[1], keepv = up ? v2[1] : v2[0];
    v1 = keepv + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  v1 += __shfl_xor_sync(0xffffffffu, v1, 2);
  v1 += __shfl_xor_sync(0xffffffffu, v1, 1);
  return v1;
}

// ---------------------------------------------------------------- partial-sum exchange over peer memory
// Every rank owns one exchange block that all peers map (CUDA IPC): XF_INTS arrival flags, then two buffers of
// `ycap` doubles.  Iteration `seq`: the partial sums go to buffer seq&1 of the rank's OWN block, the LAST CTA of that
// k_cs_warp launch publishes seq in every peer's flag slot of this rank, and k_cs_vec — once all flags show seq — adds the R partial vectors in
// rank order with loads over NVLink (same order everywhere: all ranks get the same bits).  Two buffers are enough:
// a rank can only reach the write of seq+2 after every peer has signalled seq+1, i.e. finished reading seq.
#define XF_INTS 64
#define X_MAXR 8
struct XPeers { double* base[X_MAXR]; };  // exchange block of each rank as mapped into THIS process
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_relaxed_sys(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__global__ void k_xsignal(XPeers px, int rank, int nr, int seq) {  // stand-alone signal (A/B switch B200BA_XSIGNAL=kernel)
  const int r = threadIdx.x;
  if (r >= nr || r == rank) return;
  __threadfence_system();
  int* f = reinterpret_cast<int*>(px.base[r]) + rank;
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(f), "r"(seq) : "memory");
}
// called by every thread of the CTA at the end of the partial-sum pass; ctr: zero between launches
__device__ __forceinline__ void xchg_signal_when_last(const XPeers& px, int rank, int nr, int seq, int* __restrict__ ctr) {
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int old = atomicAdd(ctr, 1);
    s_last = old == (int)gridDim.x - 1;
    if (s_last) *ctr = 0;
  }
  __syncthreads();
  if (s_last && threadIdx.x < nr && threadIdx.x != rank) {
    __threadfence_system();
    int* f = reinterpret_cast<int*>(px.base[threadIdx.x]) + rank;
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(f), "r"(seq) : "memory");
  }
}

// One 32 B descriptor per task; wpart: 8 doubles per task; wcount: one arrival counter per camera (zero between
// launches: the last arriver resets it).
struct __align__(32) WarpTask {
  int cam, lo, hi;  // local rows [lo, hi) of the camera (B order, padded runs flattened)
  int run0;         // run that holds row lo
  int nparts, t0;   // tasks of this camera, index of its first task
  int pad[2];
};
#ifdef B200BA_L2HINT
// Opt-in build: keep (a fraction of) the camera part of the Jacobian rows in L2 across the two passes of an iteration and
// across iterations — at 8 ranks one rank's share of Final-13682 is 116 MB of rows against 126 MB of L2.
#ifndef B200BA_L2FRAC
#define B200BA_L2FRAC "0.75"
#endif
__device__ __forceinline__ uint64_t cw_policy_rows() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.L2::evict_first.b64 %0, " B200BA_L2FRAC ";" : "=l"(p));
  return p;
}
__device__ __forceinline__ void ldrow_h(const float* __restrict__ p, float* v, uint64_t pol) {
  asm volatile("ld.global.nc.L2::cache_hint.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p), "l"(pol));
}
__device__ __forceinline__ void ldrow_h(const double* __restrict__ p, double* v, uint64_t pol) {
  asm volatile("ld.global.nc.L2::cache_hint.v4.f64 {%0,%1,%2,%3}, [%4], %5;" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p), "l"(pol));
  asm volatile("ld.global.nc.L2::cache_hint.v4.f64 {%0,%1,%2,%3}, [%4], %5;" : "=d"(v[4]), "=d"(v[5]), "=d"(v[6]), "=d"(v[7]) : "l"(p + 4), "l"(pol));
}
#define CW_LDROW(p, v) ldrow_h(p, v, cw_pol)
#else
#define CW_LDROW(p, v) ldrow(p, v)
#endif
template <typename JT, int CW_U>
__device__ __forceinline__ void
k_cs_warp_task(int t, int lane, int nPB, int flags, const int32_t* __restrict__ camtab,
               const WarpTask* __restrict__ wtask, const JT* __restrict__ Jc_B, const double* __restrict__ omg,
               double* __restrict__ sig, const double* __restrict__ cv, int vin_slot, double* __restrict__ y,
               double* __restrict__ wpart, int* __restrict__ wcount) {
  int tk[8];
  {
    const int4* q = reinterpret_cast<const int4*>(wtask + t);
    const int4 a = __ldg(q), b = __ldg(q + 1);
    tk[0] = a.x; tk[1] = a.y; tk[2] = a.z; tk[3] = a.w; tk[4] = b.x; tk[5] = b.y;
  }
  const int c = tk[0], lo = tk[1], hi = tk[2], run0 = tk[3], H = tk[4];
#ifdef B200BA_L2HINT
  const uint64_t cw_pol = cw_policy_rows();
#endif
  const int nb = (nPB + 3) & ~3;
  const int32_t* __restrict__ beg = camtab + (size_t)c * (nb + ((nPB + 4) & ~3));
  const int32_t* __restrict__ pre = beg + nb;
  // current run of this lane: local rows [pr0, pr1) start at B position bg
  int r = run0, pr0 = __ldg(pre + run0), pr1 = __ldg(pre + run0 + 1), bg = __ldg(beg + run0);
  auto locate = [&](int l) -> size_t {
    while (l >= pr1) { ++r; pr0 = pr1; pr1 = __ldg(pre + r + 1); bg = __ldg(beg + r); }
    return (size_t)bg + (size_t)(l - pr0);
  };
  if (flags & CS_C) {
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int l0 = lo + lane; l0 < hi; l0 += 32 * CW_U) {
      size_t j[CW_U];
      bool h[CW_U];
#pragma unroll
      for (int u = 0; u < CW_U; ++u) {
        const int l = l0 + 32 * u;
        h[u] = l < hi;
        j[u] = h[u] ? locate(l) : 0;
      }
      JT a[CW_U][8];
      double w[CW_U];
#pragma unroll
      for (int u = 0; u < CW_U; ++u) {
        if (h[u]) { CW_LDROW(Jc_B + 8 * j[u], a[u]); w[u] = __ldg(omg + j[u]); }
        else {
#pragma unroll
          for (int k = 0; k < 8; ++k) a[u][k] = (JT)0;
          w[u] = 0.0;
        }
      }
#pragma unroll
      for (int u = 0; u < CW_U; ++u)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] -= (double)a[u][k] * w[u];
    }
    const double v = warp_tsum8(acc, lane);
    const int k = comp8(lane);
    if (H <= 1) {
      if ((lane & 3) == 0) y[8 * (size_t)c + k] = v;
    } else {
      if ((lane & 3) == 0) wpart[8 * (size_t)t + k] = v;
      __threadfence();
      __syncwarp();
      int old = 0;
      if (lane == 0) old = atomicAdd(wcount + c, 1);
      old = __shfl_sync(0xffffffffu, old, 0);
      if (old == H - 1) {  // every part of this camera is in: sum them in task order
        __threadfence();
        const size_t t0 = (size_t)tk[5];
        if (lane < 8) {
          double s = 0.0;
          for (int hh = 0; hh < H; ++hh) s += __ldcg(wpart + 8 * (t0 + hh) + lane);
          y[8 * (size_t)c + lane] = s;
        }
        if (lane == 0) wcount[c] = 0;
      }
    }
  }
  if (flags & CS_A) {
    double m[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) m[k] = cv[(size_t)CV_STRIDE * c + 8 * vin_slot + k];
    r = run0; pr0 = __ldg(pre + run0); pr1 = __ldg(pre + run0 + 1); bg = __ldg(beg + run0);
    for (int l0 = lo + lane; l0 < hi; l0 += 32 * CW_U) {
      size_t j[CW_U];
      bool h[CW_U];
#pragma unroll
      for (int u = 0; u < CW_U; ++u) {
        const int l = l0 + 32 * u;
        h[u] = l < hi;
        j[u] = h[u] ? locate(l) : 0;
      }
      JT a[CW_U][8];
#pragma unroll
      for (int u = 0; u < CW_U; ++u)
        if (h[u]) CW_LDROW(Jc_B + 8 * j[u], a[u]);
#pragma unroll
      for (int u = 0; u < CW_U; ++u) {
        if (h[u]) {
          double s = 0.0;
#pragma unroll
          for (int k = 0; k < 8; ++k) s += (double)a[u][k] * m[k];
          sig[j[u]] = s;
        }
      }
    }
  }
}

// CW_U = row loads in flight per lane: 3 (<= 64 registers with fp32 rows, 4 CTAs/SM) for many small tasks, 6 (2 CTAs/SM)
// for few large ones — the bytes in flight per SM are the same, the task granularity is not (tools/cs_warp_probe.py)
template <typename JT, int CW_U>
__global__ void __launch_bounds__(CW_T, sizeof(JT) == 4 ? (CW_U <= 3 ? 4 : CW_U == 4 ? 3 : 2) : (CW_U <= 4 ? 2 : 1))
k_cs_warp(int ntasks, int nPB, int flags, const SolverState* __restrict__ st, const int32_t* __restrict__ camtab,
          const WarpTask* __restrict__ wtask, const JT* __restrict__ Jc_B, const double* __restrict__ omg,
          double* __restrict__ sig, const double* __restrict__ cv, int vin_slot, double* __restrict__ y,
          double* __restrict__ wpart, int* __restrict__ wcount, XPeers px, int rank, int nr, int seq, int* __restrict__ xctr) {
  const int lane = threadIdx.x & 31;
  const int t = blockIdx.x * (CW_T / 32) + (threadIdx.x >> 5);
  griddep_launch();  // PDL (common.cuh): the next kernel of the iteration may become resident while this one drains
  griddep_wait();    // omega / the camera records / the finished flag come from the kernels before
  const bool work = ((flags & CS_IGNORE_DONE) || !st->pcg_done) && t < ntasks;  // warp-uniform
  if (work) k_cs_warp_task<JT, CW_U>(t, lane, nPB, flags, camtab, wtask, Jc_B, omg, sig, cv, vin_slot, y, wpart, wcount);
  // multi-rank: tell the peers that this rank's partial sums of iteration seq are complete (also when the solve is
  // finished and nothing was written: the receiving side waits for every seq)
  if (nr > 1) xchg_signal_when_last(px, rank, nr, seq, xctr);
}

// Camera-local vector step, 8 lanes per camera (lane&7 = component), registers only: every load of the camera's
// 200-double neighbourhood (own components of the 9 PCG vectors, row k of U_c and Minv_c) is issued up front, the
// matvec inputs travel by shuffles inside the 8-lane group.  Same arithmetic and summation order as cs_update().
#define CV_T 128
__global__ void __launch_bounds__(CV_T)
k_cs_vec(int C, int flags, SolverState* __restrict__ st, const double* __restrict__ U, const double* __restrict__ Minv,
         int vin_slot, double* __restrict__ y, double* __restrict__ cv, double* __restrict__ part,
         XPeers px, int rank, int nr, int seq, size_t xoff) {
  const int e = blockIdx.x * CV_T + threadIdx.x;
  const int c = e >> 3, k = e & 7, lane = threadIdx.x & 31, base = lane & ~7;
  const bool act = c < C;
  const size_t cc = act ? c : C - 1;  // idle groups shadow the last camera (convergent shuffles), never write
  const bool upd = flags & (CS_INIT | CS_ITER);
  double* rec = cv + (size_t)CV_STRIDE * cc + k;
  double v[CV_SLOTS], u8[8] = {0, 0, 0, 0, 0, 0, 0, 0}, m8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  // PDL: U / Minv are constant during a solve and may be loaded while the kernel before drains.  The camera records are
  // NOT: with every kernel of the iteration allowed to start early, "the launch that wrote them has completed" only holds
  // after this kernel's own wait
  griddep_launch();
  if (flags & CS_UX) load8<double>(U + 64 * cc + 8 * k, u8);
  if (upd) load8<double>(Minv + 64 * cc + 8 * k, m8);
  griddep_wait();
#pragma unroll
  for (int sl = 0; sl < CV_SLOTS; ++sl) v[sl] = rec[8 * sl];
  const double vk = vin_slot >= 0 ? rec[8 * vin_slot] : 0.0;
  double nk = ((flags & CS_READ_Y) && nr <= 1) ? y[8 * cc + k] : 0.0;
  const int done = (flags & CS_IGNORE_DONE) ? 0 : st->pcg_done;  // set by an earlier launch: uniform over grid and ranks
  const double mu = st->mu, alpha = st->alpha, beta = st->beta;
  const bool xsum = (flags & CS_READ_Y) && nr > 1;
  if (xsum) {
    // wait for every peer's signal of this iteration — also when the solve is already finished, so that the
    // two-buffer invariant above holds for whatever comes next (bounded: a lost peer must not hang the GPU)
    if (threadIdx.x < nr && threadIdx.x != rank) {
      const int* f = reinterpret_cast<const int*>(px.base[rank]) + threadIdx.x;
      long long t0, t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
      while (ld_acquire_sys(f) - seq < 0) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 2000000000LL) { atomicOr(&st->status, 8); break; }  // 2 s of wall time; k_lm_decide then rejects on every rank
        __nanosleep(100);
      }
    }
    __syncthreads();
  }
  if (done) return;
  if (xsum) {
    double pv[X_MAXR];
#pragma unroll
    for (int r = 0; r < X_MAXR; ++r)
      if (r < nr) pv[r] = ld_relaxed_sys(px.base[r] + xoff + 8 * cc + k);
#pragma unroll
    for (int r = 0; r < X_MAXR; ++r)
      if (r < nr) nk += pv[r];
  }
  if (flags & CS_UX) {
    double s = mu * vk;
#pragma unroll
    for (int l = 0; l < 8; ++l) s += u8[l] * __shfl_sync(0xffffffffu, vk, base + l);
    nk += s;
  }
  if (!upd) {
    if ((flags & CS_WRITE_Y) && act) y[8 * cc + k] = nk;
    return;
  }
  double xk = v[CV_X], rk = v[CV_R], uk = v[CV_U], wk = v[CV_W], zk, qk, sk, pk;
  if (flags & CS_INIT) {
    wk = nk; zk = qk = sk = pk = 0.0;
  } else {
    zk = nk + beta * v[CV_Z];
    qk = vk + beta * v[CV_Q];
    sk = wk + beta * v[CV_S];
    pk = uk + beta * v[CV_P];
    xk += alpha * pk;
    rk -= alpha * sk;
    uk -= alpha * qk;
    wk -= alpha * zk;
  }
  double mk = 0.0;
#pragma unroll
  for (int l = 0; l < 8; ++l) mk += m8[l] * __shfl_sync(0xffffffffu, wk, base + l);
  double g = rk * uk, d = wk * uk, q2 = rk * rk;
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {
    g += __shfl_xor_sync(0xffffffffu, g, o);
    d += __shfl_xor_sync(0xffffffffu, d, o);
    q2 += __shfl_xor_sync(0xffffffffu, q2, o);
  }
  if (act) {
    rec[8 * CV_X] = xk; rec[8 * CV_R] = rk; rec[8 * CV_U] = uk; rec[8 * CV_W] = wk;
    rec[8 * CV_Z] = zk; rec[8 * CV_Q] = qk; rec[8 * CV_S] = sk; rec[8 * CV_P] = pk;
    rec[8 * CV_M] = mk;
    if (k == 0) { part[4 * cc] = g; part[4 * cc + 1] = d; part[4 * cc + 2] = q2; }
  }
}
