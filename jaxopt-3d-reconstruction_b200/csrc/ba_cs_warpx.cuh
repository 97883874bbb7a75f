// Camera side of the pipelined PCG on SEVERAL ranks, warp-per-camera version: ONE launch per iteration for what the default
// sequence does in three (k_cs_warp pass C | k_cs_vec | k_cs_warp pass A, ba_cs_warp.cuh).
//
// At 8 ranks every kernel of the three-launch sequence sits on its latency floor (13-16 us each at Venice-1778 for ~3 us of
// memory traffic) and the rows are read from HBM twice.  The fused ring kernel (ba_cs_ringx.cuh) removes both but walks ONE
// camera at a time per CTA: 2.4-3.9 us of serial per-camera latency, which loses at 8 ranks.  Here every WARP owns a camera:
//   pass C   partial n_c = -sum a_i omega_i over this rank's rows of the camera (rows from global memory, CW_U in flight per lane;
//            the same loop and reduction as k_cs_warp on an unsplit camera: the same bits)
//   push     lanes (peer, component) write the partial into every peer's exchange block as self-validating 16 B packets
//            {seq:lo32, seq:hi32} (no fence, no flag; ba_cs_ringx.cuh)
//   wait     lanes (source rank, component) poll this rank's block for the peers' packets of this camera
//   update   the R partial vectors are added IN RANK ORDER (identical bits on every rank), cs_update() does the eight
//            recurrences and m_c = Minv_c w_c from the camera record staged in shared memory (1.6 KB per warp)
//   pass A   sigma_i = a_i . m_c over the same rows again (they were just read: L1 / L2 hits)
// A warp with several cameras runs them as a pipeline of depth 1: pass C + push of the next camera before the wait for the
// current one, so the packets travel while rows stream.
// Persistent warps: warp w takes cameras cam_order[w], cam_order[w + W], ... in the GLOBAL order that is the same on every
// rank, and the launch is COOPERATIVE (all W warps resident on every rank, same W on identical GPUs): the warp that owns camera
// c on a peer is always either running or done, so the wait cannot deadlock; a lost peer ends in the 2 s time-out of the
// other exchange kernels (status bit 8 -> the step is rejected on every rank).
// Two packet buffers (parity of the sequence number) suffice for the same reason as in ba_cs_warp.cuh.
#pragma once
#include "ba_cs_ringx.cuh"

template <typename JT, int CW_U>
__global__ void __launch_bounds__(CW_T, sizeof(JT) == 4 ? (CW_U <= 3 ? 4 : CW_U == 4 ? 3 : 2) : (CW_U <= 4 ? 2 : 1))
k_cs_warpx(int C, int nPB, int flags, SolverState* __restrict__ st, const int32_t* __restrict__ camtab,
           const int32_t* __restrict__ cam_order, const JT* __restrict__ Jc_B, const double* __restrict__ omg,
           double* __restrict__ sig, const double* __restrict__ U, const double* __restrict__ Minv, int vin_slot,
           double* __restrict__ cv, double* __restrict__ part, XPeers px, XCam xc, int rank, int nr, int seq,
           long long timeout_ns, int pipe) {
  __shared__ __align__(16) double s_rec[CW_T / 32][CV_STRIDE + 128];  // per warp: cv record | U_c | Minv_c
  if (st->pcg_done) return;  // identical on every rank: nobody pushes, nobody waits
  bool lost = (st->status & 8) != 0;  // a peer was lost earlier in this solve: do not wait again
  const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
  const int w = blockIdx.x * (CW_T / 32) + wl, W = gridDim.x * (CW_T / 32);
  const double mu = st->mu, alpha = st->alpha, beta = st->beta;
  const int parity = seq & 1;
  const int nb = (nPB + 3) & ~3, tab_stride = nb + ((nPB + 4) & ~3);
  double* sv = s_rec[wl];
  const int k = lane & 7;
  // Software pipeline of depth 1 over the warp's cameras: pass C + push of camera i, THEN the finish (poll, update, pass A)
  // of the camera before — the peers' packets travel while the next camera's rows stream.
  auto row_table = [&](int c, const int32_t*& beg, const int32_t*& pre) {
    beg = camtab + (size_t)c * tab_stride;
    pre = beg + nb;
  };
  auto finish = [&](int c, double own) {
    const int32_t *beg, *pre;
    row_table(c, beg, pre);
    const int L = __ldg(pre + nPB);
    // the camera's record for the update: issued before the poll, consumed after it
    __syncwarp();
    for (int q = lane; q < CV_STRIDE; q += 32) sv[q] = cv[(size_t)CV_STRIDE * c + q];
    for (int q = lane; q < 64; q += 32) {
      sv[CV_STRIDE + q] = __ldg(U + 64 * (size_t)c + q);
      sv[CV_STRIDE + 64 + q] = __ldg(Minv + 64 * (size_t)c + q);
    }
    // all partial sums in rank order (every 8-lane group does the same work; a packet is valid when both words carry this
    // launch's sequence number)
    double nk = 0.0;
    long long t0 = 0;
#pragma unroll
    for (int rr = 0; rr < X_MAXR; ++rr) {
      if (rr >= nr) continue;
      if (rr == rank) { nk += own; continue; }
      const double* src = px.base[rank] + xc.data_off + ((((size_t)parity * nr + rr) * xc.Ccap + c) * 8 + k) * 2;
      unsigned long long q0 = 0, q1 = 0;
      int spins = 0;
      for (;;) {
        asm volatile("ld.relaxed.sys.global.v2.u64 {%0,%1}, [%2];" : "=l"(q0), "=l"(q1) : "l"(src) : "memory");
        if (((unsigned)(q0 >> 32) == (unsigned)seq && (unsigned)(q1 >> 32) == (unsigned)seq) || lost) break;
        if (++spins == 256 && t0 == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        if (spins > 256 && (spins & 255) == 0) {
          long long t1;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
          if (t1 - t0 > timeout_ns) { atomicOr(&st->status, 8); lost = true; }
        }
      }
      nk += __longlong_as_double((long long)((q1 << 32) | (q0 & 0xffffffffull)));
    }
    lost = __any_sync(0xffffffffu, lost);
    __syncwarp();  // the staged record is complete (written by this warp's lanes above)
    const double mk = cs_update(flags, k, lane, nk, sv, mu, alpha, beta, cv + (size_t)CV_STRIDE * c, part + 4 * (size_t)c, vin_slot);
    double m[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) m[q] = __shfl_sync(0xffffffffu, mk, q);
    // pass A: the camera's rows again (L1 / L2)
    int r = 0, pr0 = __ldg(pre), pr1 = __ldg(pre + 1), bg = __ldg(beg);
    auto locate = [&](int l) -> size_t {
      while (l >= pr1) { ++r; pr0 = pr1; pr1 = __ldg(pre + r + 1); bg = __ldg(beg + r); }
      return (size_t)bg + (size_t)(l - pr0);
    };
    for (int l0 = lane; l0 < L; l0 += 32 * CW_U) {
      size_t j[CW_U];
      bool h[CW_U];
#pragma unroll
      for (int u = 0; u < CW_U; ++u) {
        const int l = l0 + 32 * u;
        h[u] = l < L;
        j[u] = h[u] ? locate(l) : 0;
      }
      JT a[CW_U][8];
#pragma unroll
      for (int u = 0; u < CW_U; ++u)
        if (h[u]) ldrow(Jc_B + 8 * j[u], a[u]);
#pragma unroll
      for (int u = 0; u < CW_U; ++u) {
        if (h[u]) {
          double s = 0.0;
#pragma unroll
          for (int q = 0; q < 8; ++q) s += (double)a[u][q] * m[q];
          sig[j[u]] = s;
        }
      }
    }
  };
  int c_prev = -1;
  double own_prev = 0.0;
  for (int i = w; i < C; i += W) {
    const int c = __ldg(cam_order + i);
    const int32_t *beg, *pre;
    row_table(c, beg, pre);
    const int L = __ldg(pre + nPB);
    int r = 0, pr0 = __ldg(pre), pr1 = __ldg(pre + 1), bg = __ldg(beg);
    auto locate = [&](int l) -> size_t {
      while (l >= pr1) { ++r; pr0 = pr1; pr1 = __ldg(pre + r + 1); bg = __ldg(beg + r); }
      return (size_t)bg + (size_t)(l - pr0);
    };
    // ---- pass C
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int l0 = lane; l0 < L; l0 += 32 * CW_U) {
      size_t j[CW_U];
      bool h[CW_U];
#pragma unroll
      for (int u = 0; u < CW_U; ++u) {
        const int l = l0 + 32 * u;
        h[u] = l < L;
        j[u] = h[u] ? locate(l) : 0;
      }
      JT a[CW_U][8];
      double om[CW_U];
#pragma unroll
      for (int u = 0; u < CW_U; ++u) {
        if (h[u]) { ldrow(Jc_B + 8 * j[u], a[u]); om[u] = __ldg(omg + j[u]); }
        else {
#pragma unroll
          for (int q = 0; q < 8; ++q) a[u][q] = (JT)0;
          om[u] = 0.0;
        }
      }
#pragma unroll
      for (int u = 0; u < CW_U; ++u)
#pragma unroll
        for (int q = 0; q < 8; ++q) acc[q] -= (double)a[u][q] * om[u];
    }
    const double vsum = warp_tsum8(acc, lane);                       // component comp8(lane); comp8(4 g) = g
    const double own = __shfl_sync(0xffffffffu, vsum, 4 * k);        // component k = lane & 7 in every lane
    // ---- push this rank's partial to every peer: lane = (peer slot, component), 4 peers per pass
    if (nr > 1) {
      const unsigned long long bits = (unsigned long long)__double_as_longlong(own);
      const unsigned long long p0 = ((unsigned long long)(unsigned)seq << 32) | (bits & 0xffffffffull);
      const unsigned long long p1 = ((unsigned long long)(unsigned)seq << 32) | (bits >> 32);
      for (int q0 = 0; q0 < nr; q0 += 4) {
        const int rr = q0 + (lane >> 3);
        if (rr < nr && rr != rank) {
          double* dst = px.base[rr] + xc.data_off + ((((size_t)parity * nr + rank) * xc.Ccap + c) * 8 + k) * 2;
          asm volatile("st.global.v2.u64 [%0], {%1,%2};" ::"l"(dst), "l"(p0), "l"(p1) : "memory");
        }
      }
    }
    if (!pipe) { finish(c, own); continue; }  // depth 0: finish at once (the rows of pass C are still in L1 / L2 for pass A)
    if (c_prev >= 0) finish(c_prev, own_prev);
    c_prev = c;
    own_prev = own;
  }
  if (c_prev >= 0) finish(c_prev, own_prev);
}
