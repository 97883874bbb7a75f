// Camera side of the pipelined PCG for the POINT-SHARDED (multi-rank) solve: ONE kernel that does the matvec tail, the
// exchange of the partial camera sums over peer memory (NVLink / NVSwitch), the camera-local vector update and the
// next matvec head — the exchange overlaps the streaming of the following cameras, camera by camera.
//
// Round 1 split this step into three launches (k_cs_warp partial sums | k_cs_vec: wait for every peer, sum, update |
// k_cs_warp sigma pass): the Jacobian rows were read from HBM twice per iteration and the wait for the slowest peer sat
// exposed between two launches (Final-13682 on 8 ranks: 2 x 39 us + 22 us against 50 us for an eighth of the single-rank
// ring kernel).  Here the TMA ring of ba_cs_ring.cuh is kept — rows are read once — and every CTA keeps D cameras in
// flight:
//   iteration i   pass C of camera i from its ring slots (slots are NOT released), reduce, and warp 0 PUSHES the partial
//                 8-vector into every peer's exchange block as self-validating 16 B packets (each 32-bit half of a value
//                 travels with the launch sequence number; plain stores over NVLink, no fence)
//                 finish camera i - D: warp 0 polls that camera's packets (they have had D cameras' worth of streaming
//                 time to arrive), adds the R partial vectors IN RANK ORDER (identical bits on every
//                 rank: the PCG scalars stay in lock-step without further communication), runs the camera-local update,
//                 and all warps do pass A from the rows still sitting in shared memory; only then the slots go back to
//                 the producer.
// Two exchange buffers (parity of the sequence number) suffice for the same reason as in ba_cs_warp.cuh.
// With nr == 1 (B200BA_FORCE_SPLIT on one GPU) the same pipeline runs without peers — that is how the single-GPU tests
// cover it.
// ORDER: the wait for camera c is a wait for the same-numbered CTA of every peer to reach camera c, so all ranks deal the
// cameras to the CTAs in ONE order (cam_order_x: descending global row count, built in b200ba_plan_set_comm).  The first
// version used each rank's own order (by local row count): correct on the small test scenes (<= 1 camera per CTA), a
// circular wait — a hang until the 2 s time-outs — on Venice at 2 ranks.
#pragma once
#include "ba_cs_ring.cuh"
#include "ba_cs_warp.cuh"

#define RX_DMAX 8  // cameras in flight per CTA (bounded by the ring: (D + 1) x slots per camera + 1 <= slots)

struct XCam {            // camera-level region of every rank's exchange block (same layout on all ranks)
  size_t flag_off;       // (unused since the packets validate themselves)
  size_t data_off;       // in doubles from the block base: packets[parity][src][Ccap][8] of 16 B {seq:lo32, seq:hi32}
  int Ccap;
};

// dynamic record of one in-flight camera: sv[CV_STRIDE + 128] | yv[8] (unused) | cam, pad | tab[tab_stride]
__host__ __device__ inline size_t rx_rec_bytes(int nPB) {
  const int nb = (nPB + 3) & ~3, ts = nb + ((nPB + 4) & ~3);
  return ((size_t)(CV_STRIDE + 128) * 8 + 16 + (size_t)ts * 4 + 15) & ~(size_t)15;
}
__host__ __device__ inline size_t rx_fixed_bytes(int nPB, int D) {
  return (size_t)(D + 3) * rx_rec_bytes(nPB) + (RING_CW * 8 + 8 + (RX_DMAX + 1) * 8) * 8 + 2048;
}

template <typename JT>
__global__ void __maxnreg__(RING_MAXREG)
k_cs_ringx(int C, int nPB, int flags, int nslots, int D, SolverState* __restrict__ st, const int32_t* __restrict__ camtab,
           const int32_t* __restrict__ cam_order, const JT* __restrict__ Jc_B, const double* __restrict__ omg,
           double* __restrict__ sig, const double* __restrict__ U, const double* __restrict__ Minv, int vin_slot,
           double* __restrict__ cv, double* __restrict__ part, XPeers px, XCam xc, int rank, int nr, int seq,
           long long timeout_ns) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int NCAMX = D + 3;
  const size_t recb = rx_rec_bytes(nPB);
  JT* s_rows = reinterpret_cast<JT*>(smem_raw);
  double* s_omg = reinterpret_cast<double*>(smem_raw + (size_t)nslots * RING_R * 8 * sizeof(JT));
  unsigned char* s_rec = reinterpret_cast<unsigned char*>(s_omg + (size_t)nslots * RING_R);
  double* s_red = reinterpret_cast<double*>(s_rec + (size_t)NCAMX * recb);
  double* s_m = s_red + RING_CW * 8;
  double* s_pend = s_m + 8;  // [(RX_DMAX + 1)][8]: this rank's own partial sums of the cameras in flight
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(s_pend + (RX_DMAX + 1) * 8);
  uint64_t* bar_empty = bar_full + nslots;
  uint64_t* bar_cfull = bar_empty + nslots;
  uint64_t* bar_cempty = bar_cfull + (RX_DMAX + 3);
  int* s_run0 = reinterpret_cast<int*>(bar_cempty + (RX_DMAX + 3));
  if (st->pcg_done) return;  // identical on every rank: nobody pushes, nobody waits
  // a peer was lost earlier in this solve (exchange timeout): do not wait again — the PCG is being wound down
  // (pcg_scalars stops on the status bit) and k_lm_decide rejects the step on every rank
  const bool lost = (st->status & 8) != 0;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  __shared__ double s_scal[4];
  if (tid == 0) {
    s_scal[0] = st->mu; s_scal[1] = st->alpha; s_scal[2] = st->beta;
    for (int i = 0; i < nslots; ++i) { mbar_init(bar_full + i, 1); mbar_init(bar_empty + i, RING_CW); }
    for (int i = 0; i < NCAMX; ++i) { mbar_init(bar_cfull + i, 1); mbar_init(bar_cempty + i, RING_CW); }
    mbar_fence_init();
  }
  __syncthreads();
  const int nb = (nPB + 3) & ~3, tab_stride = nb + ((nPB + 4) & ~3);
  constexpr uint32_t ROWB = 8 * sizeof(JT);
  const int ncam = ((int)blockIdx.x < C) ? (C - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  auto rec_sv = [&](int cb) { return reinterpret_cast<double*>(s_rec + (size_t)cb * recb); };
  auto rec_cam = [&](int cb) { return reinterpret_cast<int*>(s_rec + (size_t)cb * recb + (size_t)(CV_STRIDE + 128) * 8); };
  auto rec_tab = [&](int cb) { return reinterpret_cast<int*>(s_rec + (size_t)cb * recb + (size_t)(CV_STRIDE + 128) * 8 + 16); };

  if (wid == RING_CW) {
    // ------------------------------------------------------------ producer warp (as in k_cs_ring)
    auto issue_record = [&](int i) {
      const int cb = i % NCAMX, ph = (i / NCAMX) & 1;
      mbar_wait(bar_cempty + cb, ph ^ 1);
      if (lane == 0) {
        const int c = __ldg(cam_order + blockIdx.x + (size_t)i * gridDim.x);
        rec_cam(cb)[0] = c;
        mbar_arrive_expect_tx(bar_cfull + cb, (uint32_t)((CV_STRIDE + 128) * sizeof(double) + tab_stride * sizeof(int32_t)));
        bulk_g2s(rec_sv(cb), cv + (size_t)CV_STRIDE * c, CV_STRIDE * sizeof(double), bar_cfull + cb);
        bulk_g2s(rec_sv(cb) + CV_STRIDE, U + 64 * (size_t)c, 64 * sizeof(double), bar_cfull + cb);
        bulk_g2s(rec_sv(cb) + CV_STRIDE + 64, Minv + 64 * (size_t)c, 64 * sizeof(double), bar_cfull + cb);
        bulk_g2s(rec_tab(cb), camtab + (size_t)c * tab_stride, tab_stride * sizeof(int32_t), bar_cfull + cb);
      }
    };
    int ring = 0, ring_ph = 0;
    if (ncam > 0) issue_record(0);
    for (int i = 0; i < ncam; ++i) {
      if (i + 1 < ncam) issue_record(i + 1);
      const int cb = i % NCAMX, ph = (i / NCAMX) & 1;
      mbar_wait(bar_cfull + cb, ph);
      const int* beg = rec_tab(cb);
      const int* pre = beg + nb;
      const int L = pre[nPB];
      const int nchunks = (L + RING_R - 1) / RING_R;
      for (int k = 0; k < nchunks; ++k) {
        mbar_wait(bar_empty + ring, ring_ph ^ 1);
        const int l_lo = k * RING_R, l_hi = min(L, l_lo + RING_R);
        int first_run = nPB;
        for (int r = lane; r < nPB; r += 32) {
          const int lo = max(pre[r], l_lo), hi = min(pre[r + 1], l_hi);
          if (lo < hi) {
            first_run = min(first_run, r);
            const size_t src = (size_t)beg[r] + (size_t)(lo - pre[r]);
            const size_t dst = (size_t)ring * RING_R + (size_t)(lo - l_lo);
            bulk_g2s(s_rows + 8 * dst, Jc_B + 8 * src, (uint32_t)(hi - lo) * ROWB, bar_full + ring);
            bulk_g2s(s_omg + dst, omg + src, (uint32_t)(hi - lo) * 8, bar_full + ring);
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) first_run = min(first_run, __shfl_xor_sync(0xffffffffu, first_run, o));
        if (lane == 0) {
          s_run0[ring] = first_run;
          mbar_arrive_expect_tx(bar_full + ring, (uint32_t)(l_hi - l_lo) * (ROWB + 8));
        }
        if (++ring == nslots) { ring = 0; ring_ph ^= 1; }
      }
    }
    return;
  }

  // -------------------------------------------------------------- consumer warps
  const int parity = seq & 1;
  int ringC = 0, phC = 0, ringA = 0;
  for (int i = 0; i < ncam + D; ++i) {
    // warp 0: the packets of the camera that will be finished in this iteration are requested NOW, so that the ~1-2 us of a
    // load of remotely written memory overlaps pass C instead of sitting on warp 0's critical path once per camera
    // (first packet version: 309 us instead of 2 x 109 + 18 us on Final-13682 at 2 ranks)
    unsigned long long pq0[X_MAXR], pq1[X_MAXR];
    if (wid == 0 && nr > 1 && i >= D) {
      const int c = rec_cam((i - D) % NCAMX)[0];
#pragma unroll
      for (int r = 0; r < X_MAXR; ++r) {
        pq0[r] = pq1[r] = 0;
        if (r < nr && r != rank) {
          const double* src = px.base[rank] + xc.data_off + ((((size_t)parity * nr + r) * xc.Ccap + c) * 8 + (lane & 7)) * 2;
          asm volatile("ld.relaxed.sys.global.v2.u64 {%0,%1}, [%2];" : "=l"(pq0[r]), "=l"(pq1[r]) : "l"(src) : "memory");
        }
      }
    }
    if (i < ncam) {
      // ---- pass C of camera i: n_c(partial) = -sum a_i omega_i over the camera's slots (kept for pass A)
      const int cb = i % NCAMX, ph = (i / NCAMX) & 1;
      mbar_wait(bar_cfull + cb, ph);
      const int* pre = rec_tab(cb) + nb;
      const int L = pre[nPB];
      const int nchunks = (L + RING_R - 1) / RING_R;
      double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int k = 0; k < nchunks; ++k) {
        mbar_wait(bar_full + ringC, phC);
        const int l = k * RING_R + tid;
        if (l < L) {
          JT a[8];
          ldrow_shared(s_rows + 8 * ((size_t)ringC * RING_R + tid), a);
          const double w = s_omg[(size_t)ringC * RING_R + tid];
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] -= (double)a[j] * w;
        }
        if (++ringC == nslots) { ringC = 0; phC ^= 1; }
      }
      const double v1 = warp_tsum8(acc, lane);
      if ((lane & 3) == 0) s_red[wid * 8 + comp8(lane)] = v1;
    }
    named_bar_sync<RING_CW * 32>(1);
    if (wid == 0) {
      const int k = lane & 7;
      if (i < ncam) {
        // own partial, fixed order over the warps; kept for the rank-ordered sum D iterations later and pushed to the peers
        double n0 = 0.0, n1 = 0.0;
#pragma unroll
        for (int w = 0; w < RING_CW; w += 2) { n0 += s_red[w * 8 + k]; n1 += s_red[(w + 1) * 8 + k]; }
        const double nk = n0 + n1;
        if (lane < 8) s_pend[(i % (RX_DMAX + 1)) * 8 + k] = nk;
        if (nr > 1) {
          // self-validating 16 B packets {seq : low word, seq : high word} (the LL idea of NCCL): no fence, no separate flag —
          // a __threadfence_system() here cost ~5 us per camera on the critical path of warp 0 (first version: 133 us
          // instead of 2 x 27 us at Venice on 2 ranks).  lane = (peer slot, component k): 4 peers per pass
          const int c = rec_cam(i % NCAMX)[0];
          const unsigned long long bits = (unsigned long long)__double_as_longlong(nk);
          const unsigned long long p0 = ((unsigned long long)(unsigned)seq << 32) | (bits & 0xffffffffull);
          const unsigned long long p1 = ((unsigned long long)(unsigned)seq << 32) | (bits >> 32);
          for (int q0 = 0; q0 < nr; q0 += 4) {
            const int r = q0 + (lane >> 3);
            if (r < nr && r != rank) {
              double* dst = px.base[r] + xc.data_off + ((((size_t)parity * nr + rank) * xc.Ccap + c) * 8 + k) * 2;
              asm volatile("st.global.v2.u64 [%0], {%1,%2};" ::"l"(dst), "l"(p0), "l"(p1) : "memory");
            }
          }
        }
      }
      if (i >= D) {
        // ---- finish camera j = i - D: all partial sums in rank order, camera-local update, m_c
        const int j = i - D, cbj = j % NCAMX;
        const int c = rec_cam(cbj)[0];
        // every lane (k = lane & 7; the four 8-lane groups do the same work) collects component k of every rank's partial
        // in rank order; a peer's packet is valid when both words carry this launch's sequence number
        double nk = 0.0;
        long long t0 = 0;
#pragma unroll
        for (int r = 0; r < X_MAXR; ++r) {
          if (r >= nr) continue;
          if (r == rank) { nk += s_pend[(j % (RX_DMAX + 1)) * 8 + k]; continue; }
          const double* src = px.base[rank] + xc.data_off + ((((size_t)parity * nr + r) * xc.Ccap + c) * 8 + k) * 2;
          unsigned long long q0 = pq0[r], q1 = pq1[r];  // requested at the top of the iteration
          int spins = 0;
          while (!(((unsigned)(q0 >> 32) == (unsigned)seq && (unsigned)(q1 >> 32) == (unsigned)seq) || lost)) {
            if (++spins == 256 && t0 == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            if (spins > 256 && (spins & 255) == 0) {
              long long t1;
              asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
              if (t1 - t0 > timeout_ns) { atomicOr(&st->status, 8); break; }
            }
            asm volatile("ld.relaxed.sys.global.v2.u64 {%0,%1}, [%2];" : "=l"(q0), "=l"(q1) : "l"(src) : "memory");
          }
          nk += __longlong_as_double((long long)((q1 << 32) | (q0 & 0xffffffffull)));
        }
        const double mu = s_scal[0], alpha = s_scal[1], beta = s_scal[2];
        const double mk = cs_update(flags, k, lane, nk, rec_sv(cbj), mu, alpha, beta, cv + (size_t)CV_STRIDE * c,
                                    part + 4 * (size_t)c, vin_slot);
        if (lane < 8) s_m[k] = mk;
      }
    }
    named_bar_sync<RING_CW * 32>(1);
    if (i >= D) {
      // ---- pass A of camera j from the rows still in shared memory; then the slots and the record go back
      const int j = i - D, cbj = j % NCAMX;
      const int* beg = rec_tab(cbj);
      const int* pre = beg + nb;
      const int L = pre[nPB];
      const int nchunks = (L + RING_R - 1) / RING_R;
      double m[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) m[q] = s_m[q];
      for (int k = 0; k < nchunks; ++k) {
        const int l = k * RING_R + tid;
        if (l < L) {
          JT a[8];
          ldrow_shared(s_rows + 8 * ((size_t)ringA * RING_R + tid), a);
          int r = s_run0[ringA];
          while (l >= pre[r + 1]) ++r;
          double s0 = 0.0;
#pragma unroll
          for (int q = 0; q < 8; ++q) s0 += (double)a[q] * m[q];
          sig[(size_t)beg[r] + (size_t)(l - pre[r])] = s0;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + ringA);
        if (++ringA == nslots) ringA = 0;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_cempty + cbj);
    }
  }
}
