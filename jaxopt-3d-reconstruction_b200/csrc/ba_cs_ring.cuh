// Camera-side kernel of the pipelined PCG, TMA-ring version (the single-rank hot path).
//
// Persistent CTAs (one per SM), cameras dealt round-robin in descending row count.  One PRODUCER
// warp streams the cameras' Jacobian rows + omega values into a ring of shared-memory slots with 1-D
// bulk async copies (cp.async.bulk, completion on mbarriers) — no registers and no warps are tied up
// by loads in flight; the rows of the NEXT camera are already landing while the consumers finish the
// current one, and the per-camera record (PCG vectors, U, Minv, run table) is fetched two cameras ahead.  16 CONSUMER warps do, per camera:
//   pass C  n_c = -sum a_i omega_i over the camera's slots            (shared memory only)
//   update  n_c += (U_c + mu) m_c; z,q,s,p,x,r,u,w recurrences; dot partials; m_c = Minv_c w_c
//   pass A  sigma_i = a_i . m_c from the same slots -> global (coalesced per run), slots released
// Every camera is processed by exactly one CTA in a fixed arithmetic order, so results do not
// depend on which CTA picked it up.
#pragma once
#include "ba_pcg.cuh"

#define RING_R 512                 // rows per slot
#define RING_CW 16                 // consumer warps (RING_R == 32 * RING_CW: one row per thread per slot)
#define RING_T ((RING_CW + 1) * 32)

struct __align__(16) RingCam {  // everything about one in-flight camera
  double sv[CV_STRIDE + 128];   // cv record | U_c | Minv_c   (bulk copied)
  int tab[2 * MAXPB + 8];       // run table: beg[nPB] (padded to 4) | pre[nPB+1] (padded to 4)   (bulk copied)
  int cam, pad[3];
};

template <typename JT>
struct RingCfg {
  static constexpr size_t slot_bytes = (size_t)RING_R * (8 * sizeof(JT) + 8);
  static constexpr size_t fixed_bytes = 3 * sizeof(RingCam) + RING_CW * 8 * sizeof(double) + 8 * sizeof(double) + 1024;
  __host__ __device__ static constexpr int slots(size_t smem_budget) { return (int)((smem_budget - fixed_bytes) / slot_bytes); }
};

#define RING_NCAM 3  // camera records in flight: consumers on i, rows streaming for i+1, record/table of i+2 landing

template <typename JT>
__global__ void __launch_bounds__(RING_T, 1)
k_cs_ring(int C, int nPB, int flags, int nslots, SolverState* __restrict__ st, const int32_t* __restrict__ camtab,
          const int32_t* __restrict__ cam_order, const JT* __restrict__ Jc_B, const double* __restrict__ omg,
          double* __restrict__ sig, const double* __restrict__ U, const double* __restrict__ Minv, int vin_slot,
          double* __restrict__ cv, double* __restrict__ part) {
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
  if (st->pcg_done) return;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) {
    for (int i = 0; i < nslots; ++i) { mbar_init(bar_full + i, 1); mbar_init(bar_empty + i, RING_CW); }
    for (int i = 0; i < RING_NCAM; ++i) { mbar_init(bar_cfull + i, 1); mbar_init(bar_cempty + i, RING_CW); }
    mbar_fence_init();
  }
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
        mbar_arrive_expect_tx(bar_cfull + cb, (uint32_t)((CV_STRIDE + 128) * sizeof(double) + tab_stride * sizeof(int32_t)));
        bulk_g2s(B.sv, cv + (size_t)CV_STRIDE * c, CV_STRIDE * sizeof(double), bar_cfull + cb);
        bulk_g2s(B.sv + CV_STRIDE, U + 64 * (size_t)c, 64 * sizeof(double), bar_cfull + cb);
        bulk_g2s(B.sv + CV_STRIDE + 64, Minv + 64 * (size_t)c, 64 * sizeof(double), bar_cfull + cb);
        bulk_g2s(B.tab, camtab + (size_t)c * tab_stride, tab_stride * sizeof(int32_t), bar_cfull + cb);
      }
    };
    int ring = 0, ring_ph = 0;
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
        if (lane == 0) mbar_arrive_expect_tx(bar_full + ring, (uint32_t)(l_hi - l_lo) * (ROWB + 8));
        for (int r = lane; r < nPB; r += 32) {
          const int lo = max(pre[r], l_lo), hi = min(pre[r + 1], l_hi);
          if (lo < hi) {
            const size_t src = (size_t)beg[r] + (size_t)(lo - pre[r]);
            const size_t dst = (size_t)ring * RING_R + (size_t)(lo - l_lo);
            bulk_g2s(s_rows + 8 * dst, Jc_B + 8 * src, (uint32_t)(hi - lo) * ROWB, bar_full + ring);
            bulk_g2s(s_omg + dst, omg + src, (uint32_t)(hi - lo) * 8, bar_full + ring);
          }
        }
        if (++ring == nslots) { ring = 0; ring_ph ^= 1; }
      }
    }
    return;
  }

  // -------------------------------------------------------------- consumer warps
  int ring = 0, ring_ph = 0;
  const double mu = st->mu, alpha = st->alpha, beta = st->beta;
  for (int i = 0; i < ncam; ++i) {
    const int cb = i % RING_NCAM, ph = (i / RING_NCAM) & 1;
    mbar_wait(bar_cfull + cb, ph);
    const RingCam& B = s_cam[cb];
    const int c = B.cam;
    const int* beg = B.tab;
    const int* pre = B.tab + nb;
    const int L = pre[nPB];
    const int nchunks = (L + RING_R - 1) / RING_R;
    const int ring0 = ring;
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 0; k < nchunks; ++k) {
      mbar_wait(bar_full + ring, ring_ph);
      const int l = k * RING_R + tid;
      if (l < L) {
        JT a[8];
        ldrow_shared(s_rows + 8 * ((size_t)ring * RING_R + tid), a);
        const double w = s_omg[(size_t)ring * RING_R + tid];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] -= (double)a[j] * w;
      }
      if (++ring == nslots) { ring = 0; ring_ph ^= 1; }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = warp_sum(acc[j]);
    if (lane == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) s_red[wid * 8 + j] = acc[j];
    }
    named_bar_sync<RING_CW * 32>(1);
    if (wid == 0) {
      const int k = lane & 7;
      double n0 = 0.0, n1 = 0.0;
#pragma unroll
      for (int w = 0; w < RING_CW; w += 2) { n0 += s_red[w * 8 + k]; n1 += s_red[(w + 1) * 8 + k]; }  // fixed order
      const double mk = cs_update(flags, k, lane, n0 + n1, B.sv, mu, alpha, beta, cv + (size_t)CV_STRIDE * c,
                                  part + 4 * (size_t)c, vin_slot);
      if (lane < 8) s_m[k] = mk;
    }
    named_bar_sync<RING_CW * 32>(1);
    double m[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = s_m[j];
    int r = 0, slot = ring0;
    for (int k = 0; k < nchunks; ++k) {
      const int l = k * RING_R + tid;
      if (l < L) {
        while (l >= pre[r + 1]) ++r;
        const size_t pos = (size_t)beg[r] + (size_t)(l - pre[r]);
        JT a[8];
        ldrow_shared(s_rows + 8 * ((size_t)slot * RING_R + tid), a);
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += (double)a[j] * m[j];
        sig[pos] = s;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + slot);
      if (++slot == nslots) slot = 0;
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_cempty + cb);
  }
}
