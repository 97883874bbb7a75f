// Camera-side kernel of the pipelined PCG, TMA-ring version (the single-rank hot path).
//
// Persistent CTAs (one per SM), cameras dealt round-robin in descending row count.  One PRODUCER
// warp streams the cameras' Jacobian rows + omega values into a ring of shared-memory slots with 1-D
// bulk async copies (cp.async.bulk, completion on mbarriers) — no registers and no warps are tied up
// by loads in flight; the rows of the NEXT camera are already landing while the consumers finish the
// current one, and the per-camera record (PCG vectors, U, Minv, run table) is fetched two cameras ahead.
// The CONSUMER warps (RING_CW = 12) do, per camera:
//   pass C  n_c = -sum a_i omega_i over the camera's slots            (shared memory only)
//   update  n_c += (U_c + mu) m_c; z,q,s,p,x,r,u,w recurrences; dot partials; m_c = Minv_c w_c
//   pass A  sigma_i = a_i . m_c from the same slots -> global (coalesced per run), slots released
// Every camera is processed by exactly one CTA in a fixed arithmetic order, so results do not
// depend on which CTA picked it up.
#pragma once
#include "ba_pcg.cuh"

#ifndef RING_CW
#define RING_CW 12                 // consumer warps.  12 warps x 128 registers (8 stashed rows per thread, no spills) beat
#endif                             // 16 x 96 (6 rows, spills): 62.5 vs 64.1 us at Venice; 8 x 168 (12 rows): 74 us
#ifndef RING_MAXREG
#define RING_MAXREG 128            // 13 warps -> at most 4 on one SM sub-partition: 4 x 32 x 128 = its 16 K registers
#endif
#ifndef RING_TRIGGER_LATE
#define RING_TRIGGER_LATE 1  // PDL trigger of the following kernel: 0 at the start of the launch, 1 when a CTA starts its last camera
#endif
#ifndef RING_MINB
#define RING_MINB 1                // CTAs per SM (2: two half-size CTAs interleave their per-camera reduce/update bubbles)
#endif
#define RING_R (32 * RING_CW)      // rows per slot: one row per consumer thread
#define RING_T ((RING_CW + 1) * 32)

struct __align__(16) RingCam {  // everything about one in-flight camera
  double sv[CV_STRIDE + 128];   // cv record | U_c | Minv_c   (bulk copied)
  int tab[2 * MAXPB + 8];       // run table: beg[nPB] (padded to 4) | pre[nPB+1] (padded to 4)   (bulk copied)
  double yv[8];                 // multi-rank: the all-reduced partial sums n_c   (bulk copied)
  int cam, pad[3];
};

template <typename JT>
struct RingCfg {
  static constexpr size_t slot_bytes = (size_t)RING_R * (8 * sizeof(JT) + 8);
  static constexpr size_t fixed_bytes = 3 * sizeof(RingCam) + RING_CW * 8 * sizeof(double) + 8 * sizeof(double) + 1024;
  __host__ __device__ static constexpr int slots(size_t smem_budget) { return (int)((smem_budget - fixed_bytes) / slot_bytes); }
};

#ifndef RING_EXP
#define RING_EXP 0  // timing experiments only: 1 no pass-A math, 2 no update, 4 no pass-C math, 8 no cross-lane reduce
#endif
// slots of rows a consumer thread keeps in registers between pass C and pass A (fp32 rows: 8 x 8 regs; fp64 rows: 4 x 16)
#ifndef RING_KEEP32
#define RING_KEEP32 8
#endif
template <typename JT> struct RingKeep { static constexpr int N = sizeof(JT) == 4 ? RING_KEEP32 : RING_KEEP32 / 2; };
#define RING_KEEP (RingKeep<JT>::N)
#define RING_NCAM 3  // camera records in flight: consumers on i, rows streaming for i+1, record/table of i+2 landing

// Compile-time unrolled per-slot steps of the consumer (recursion instead of `#pragma unroll`: the bodies
// contain mbarrier wait loops, and the row stash must be indexed by constants to live in registers).
struct RingCtx {
  int ring, ring_ph, nslots, nchunks, L, tid, lane, do_c;
  const int *beg, *pre, *run0;
  uint64_t *bar_full, *bar_empty;
};
template <typename JT, int K>
struct RingSteps {
  static __device__ __forceinline__ void pass_c(RingCtx& cx, const JT* s_rows, const double* s_omg, JT (&keep)[RING_KEEP][8],
                                                int (&pos)[RING_KEEP], double (&acc)[8]) {
    RingSteps<JT, K - 1>::pass_c(cx, s_rows, s_omg, keep, pos, acc);
    constexpr int k = K - 1;
    pos[k] = -1;
    if (k < cx.nchunks) {
      mbar_wait(cx.bar_full + cx.ring, cx.ring_ph);
      const int l = k * RING_R + cx.tid;
      if (l < cx.L) {
        ldrow_shared(s_rows + 8 * ((size_t)cx.ring * RING_R + cx.tid), keep[k]);
        const double w = cx.do_c ? s_omg[(size_t)cx.ring * RING_R + cx.tid] : 0.0;
        int r = cx.run0[cx.ring];
        while (l >= cx.pre[r + 1]) ++r;
        pos[k] = cx.beg[r] + (l - cx.pre[r]);
        if (!(RING_EXP & 4) && cx.do_c) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] -= (double)keep[k][j] * w;
        }
      }
      __syncwarp();
      if (cx.lane == 0) mbar_arrive(cx.bar_empty + cx.ring);
      if (++cx.ring == cx.nslots) { cx.ring = 0; cx.ring_ph ^= 1; }
    }
  }
  static __device__ __forceinline__ void pass_a(const JT (&keep)[RING_KEEP][8], const int (&pos)[RING_KEEP], const double (&m)[8],
                                                double* __restrict__ sig) {
    RingSteps<JT, K - 1>::pass_a(keep, pos, m, sig);
    constexpr int k = K - 1;
    if (pos[k] >= 0) {
      double s0 = 0.0;
#pragma unroll
      for (int j = 0; j < 8; ++j) s0 += (double)keep[k][j] * m[j];
#ifdef B200BA_L2HINT
      asm volatile("st.global.L2::cache_hint.f64 [%0], %1, %2;" ::"l"(sig + pos[k]), "d"(s0), "l"(l2_policy_evict_last()) : "memory");
#else
      sig[pos[k]] = s0;
#endif
    }
  }
};
template <typename JT>
struct RingSteps<JT, 0> {
  static __device__ __forceinline__ void pass_c(RingCtx&, const JT*, const double*, JT (&)[RING_KEEP][8], int (&)[RING_KEEP], double (&)[8]) {}
  static __device__ __forceinline__ void pass_a(const JT (&)[RING_KEEP][8], const int (&)[RING_KEEP], const double (&)[8], double* __restrict__) {}
};

#if RING_MAXREG > 0
#define RING_BOUNDS __maxnreg__(RING_MAXREG)
#else
#define RING_BOUNDS __launch_bounds__(RING_T, RING_MINB)
#endif
template <typename JT>
__global__ void RING_BOUNDS
k_cs_ring(int C, int nPB, int flags, int nslots, SolverState* __restrict__ st, const int32_t* __restrict__ camtab,
          const int32_t* __restrict__ cam_order, const JT* __restrict__ Jc_B, const double* __restrict__ omg,
          double* __restrict__ sig, const double* __restrict__ U, const double* __restrict__ Minv, int vin_slot,
          double* __restrict__ y, double* __restrict__ cv, double* __restrict__ part) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  // layout: rows[nslots][RING_R][8] | omega[nslots][RING_R] | cams[RING_NCAM] | sred | sm_m | barriers
  JT* s_rows = reinterpret_cast<JT*>(smem_raw);
  double* s_omg = reinterpret_cast<double*>(smem_raw + (size_t)nslots * RING_R * 8 * sizeof(JT));
  RingCam* s_cam = reinterpret_cast<RingCam*>(s_omg + (size_t)nslots * RING_R);
  double* s_red = reinterpret_cast<double*>(s_cam + RING_NCAM);
  double* s_m = s_red + RING_CW * 8;
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(s_m + 8);
  uint64_t* bar_empty = bar_full + nslots;
  uint64_t* bar_cfull = bar_empty + nslots;
  uint64_t* bar_cempty = bar_cfull + RING_NCAM;
  int* s_run0 = reinterpret_cast<int*>(bar_cempty + RING_NCAM);  // first run of the chunk held by each slot
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  // PDL: the CTA may become resident while the point-side kernel before it drains; the barrier set-up needs nothing from it
  if (tid == 0) {
    for (int i = 0; i < nslots; ++i) { mbar_init(bar_full + i, 1); mbar_init(bar_empty + i, RING_CW); }
    for (int i = 0; i < RING_NCAM; ++i) { mbar_init(bar_cfull + i, 1); mbar_init(bar_cempty + i, RING_CW); }
    mbar_fence_init();
  }
#if RING_TRIGGER_LATE
  // the point-side kernel that follows may be scheduled once EVERY CTA has started its last camera (trigger below): its CTAs
  // fill the SMs that fall idle during the last round of cameras instead of being resident through the whole launch
#else
  griddep_launch();
#endif
  griddep_wait();
  if (st->pcg_done) return;
  // PCG scalars once into shared memory: kept as plain `st->x` the compiler re-loads them from GLOBAL memory at
  // their use inside every camera's serial update (register pressure) — two dependent round trips per camera,
  // 68 -> 64 us for the whole kernel at Venice
  __shared__ double s_scal[4];
  if (tid == 0) { s_scal[0] = st->mu; s_scal[1] = st->alpha; s_scal[2] = st->beta; }
  __syncthreads();
  const int nb = (nPB + 3) & ~3, tab_stride = nb + ((nPB + 4) & ~3);
  constexpr uint32_t ROWB = 8 * sizeof(JT);
  // static assignment: this CTA takes cameras cam_order[blockIdx.x + i * gridDim.x] (cam_order = descending
  // row count, so every CTA gets a similar mix); results never depend on the assignment
  const int ncam = ((int)blockIdx.x < C) ? (C - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (wid == RING_CW) {
    // ------------------------------------------------------------ producer warp
    __shared__ int s_ids[64];
    for (int i = lane; i < min(ncam, 64); i += 32) s_ids[i] = __ldg(cam_order + blockIdx.x + (size_t)i * gridDim.x);
    __syncwarp();
    // record(i): camera i's PCG record, U, Minv and run table -> cam buffer i % RING_NCAM (all bulk copies)
    auto issue_record = [&](int i) {
      const int cb = i % RING_NCAM, ph = (i / RING_NCAM) & 1;
      mbar_wait(bar_cempty + cb, ph ^ 1);
      if (lane == 0) {
        RingCam& B = s_cam[cb];
        const int c = i < 64 ? s_ids[i] : __ldg(cam_order + blockIdx.x + (size_t)i * gridDim.x);
        B.cam = c;
        mbar_arrive_expect_tx(bar_cfull + cb, (uint32_t)((CV_STRIDE + 128) * sizeof(double) + tab_stride * sizeof(int32_t) +
                                                         ((flags & CS_READ_Y) ? 64 : 0)));
        if (flags & CS_READ_Y) bulk_g2s(B.yv, y + 8 * (size_t)c, 64, bar_cfull + cb);
        bulk_g2s(B.sv, cv + (size_t)CV_STRIDE * c, CV_STRIDE * sizeof(double), bar_cfull + cb);
        bulk_g2s(B.sv + CV_STRIDE, U + 64 * (size_t)c, 64 * sizeof(double), bar_cfull + cb);
        bulk_g2s(B.sv + CV_STRIDE + 64, Minv + 64 * (size_t)c, 64 * sizeof(double), bar_cfull + cb);
        bulk_g2s(B.tab, camtab + (size_t)c * tab_stride, tab_stride * sizeof(int32_t), bar_cfull + cb);
      }
    };
    int ring = 0, ring_ph = 0;
#ifdef B200BA_L2HINT
    const uint64_t pol_first = l2_policy_evict_first();
#endif
    if (ncam > 0) issue_record(0);
    for (int i = 0; i < ncam; ++i) {
      if (i + 1 < ncam) issue_record(i + 1);
      const int cb = i % RING_NCAM, ph = (i / RING_NCAM) & 1;
      mbar_wait(bar_cfull + cb, ph);  // the producer needs the run table too
      const RingCam& B = s_cam[cb];
      const int* beg = B.tab;
      const int* pre = B.tab + nb;
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
#ifdef B200BA_L2HINT
            bulk_g2s_hint(s_rows + 8 * dst, Jc_B + 8 * src, (uint32_t)(hi - lo) * ROWB, bar_full + ring, pol_first);
            if (flags & CS_C) bulk_g2s_hint(s_omg + dst, omg + src, (uint32_t)(hi - lo) * 8, bar_full + ring, pol_first);
#else
            bulk_g2s(s_rows + 8 * dst, Jc_B + 8 * src, (uint32_t)(hi - lo) * ROWB, bar_full + ring);
            if (flags & CS_C) bulk_g2s(s_omg + dst, omg + src, (uint32_t)(hi - lo) * 8, bar_full + ring);
#endif
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) first_run = min(first_run, __shfl_xor_sync(0xffffffffu, first_run, o));
        if (lane == 0) {
          s_run0[ring] = first_run;  // published by the arrive below
          mbar_arrive_expect_tx(bar_full + ring, (uint32_t)(l_hi - l_lo) * (ROWB + ((flags & CS_C) ? 8 : 0)));
        }
        if (++ring == nslots) { ring = 0; ring_ph ^= 1; }
      }
    }
    return;
  }

  // -------------------------------------------------------------- consumer warps
  // Each thread keeps the rows it met in pass C in REGISTERS (RING_KEEP slots' worth), so a slot goes back
  // to the producer as soon as pass C has read it: the ring is pure prefetch depth (~2 cameras ahead) and
  // HBM streaming never waits for the serial reduce / update / pass-A phases of a camera.
  int ring = 0, ring_ph = 0;
#if RING_TRIGGER_LATE
  if (ncam == 0) griddep_launch();
#endif
  for (int i = 0; i < ncam; ++i) {
#if RING_TRIGGER_LATE
    if (i == ncam - 1) griddep_launch();
#endif
    const int cb = i % RING_NCAM, ph = (i / RING_NCAM) & 1;
    mbar_wait(bar_cfull + cb, ph);
    const RingCam& B = s_cam[cb];
    const int c = B.cam;
    const int* beg = B.tab;
    const int* pre = B.tab + nb;
    const int L = pre[nPB];
    const int nchunks = (L + RING_R - 1) / RING_R;
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    JT keep[RING_KEEP][8];
    int pos[RING_KEEP];
    {
      RingCtx cx{ring, ring_ph, nslots, nchunks, L, tid, lane, flags & CS_C, beg, pre, s_run0, bar_full, bar_empty};
      RingSteps<JT, RING_KEEP>::pass_c(cx, s_rows, s_omg, keep, pos, acc);
      ring = cx.ring; ring_ph = cx.ring_ph;
    }
    // cameras with more rows than the register stash holds: pass C from the slot, pass A re-reads from HBM
    for (int k = RING_KEEP; k < nchunks; ++k) {
      mbar_wait(bar_full + ring, ring_ph);
      const int l = k * RING_R + tid;
      if (l < L && (flags & CS_C)) {
        JT a[8];
        ldrow_shared(s_rows + 8 * ((size_t)ring * RING_R + tid), a);
        const double w = s_omg[(size_t)ring * RING_R + tid];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] -= (double)a[j] * w;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + ring);
      if (++ring == nslots) { ring = 0; ring_ph ^= 1; }
    }
    // cross-lane sum of 8 values: halve the number of live values at every butterfly level
    // (18 shuffles instead of 80); lane l ends up holding component (l >> 2) & 7 summed over the warp
    if (!(flags & CS_C)) {
      // multi-rank second half: n_c arrives all-reduced in the camera record
    } else if (!(RING_EXP & 8)) {
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
      // component index held by this lane: bit4 -> +4, bit3 -> +2, bit2 -> +1
      if ((lane & 3) == 0) s_red[wid * 8 + ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)] = v1;
    } else if (lane == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) s_red[wid * 8 + j] = acc[j];
    }
    if (flags & CS_C) named_bar_sync<RING_CW * 32>(1);
    if (wid == 0) {
      const int k = lane & 7;
      double n0 = 0.0, n1 = 0.0;
      if (flags & CS_C) {
#pragma unroll
        for (int w = 0; w < RING_CW; w += 2) { n0 += s_red[w * 8 + k]; n1 += s_red[(w + 1) * 8 + k]; }  // fixed order
      } else {
        n0 = B.yv[k];
      }
      if ((flags & CS_WRITE_Y) && lane < 8) y[8 * (size_t)c + k] = n0 + n1;  // multi-rank first half: partial sums out
      if (flags & (CS_INIT | CS_ITER)) {
        const double mu = s_scal[0], alpha = s_scal[1], beta = s_scal[2];
        const double mk = (RING_EXP & 2) ? n0 + n1 + B.sv[k]
                                         : cs_update(flags, k, lane, n0 + n1, B.sv, mu, alpha, beta, cv + (size_t)CV_STRIDE * c,
                                                     part + 4 * (size_t)c, vin_slot);
        if (lane < 8) s_m[k] = mk;
      }
    }
    if (!(flags & CS_A)) {  // no matvec head in this launch: the camera is done
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_cempty + cb);
      continue;
    }
    named_bar_sync<RING_CW * 32>(1);
    double m[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = s_m[j];
    if (!(RING_EXP & 1)) {
      RingSteps<JT, RING_KEEP>::pass_a(keep, pos, m, sig);
      int r = 0;
      for (int k = RING_KEEP; k < nchunks; ++k) {
        const int l = k * RING_R + tid;
        if (l < L) {
          while (l >= pre[r + 1]) ++r;
          const size_t p = (size_t)beg[r] + (size_t)(l - pre[r]);
          JT a[8];
          ldrow(Jc_B + 8 * p, a);
          double s0 = 0.0;
#pragma unroll
          for (int j = 0; j < 8; ++j) s0 += (double)a[j] * m[j];
          sig[p] = s0;
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_cempty + cb);
  }
}
