// Point side of the Schur matvec, TILED version (opt-in: B200BA_PS_TILE=1; the chunk kernel k_ps stays the default).
//
// Why: the chunk kernel k_ps (ba_pcg.cuh) gathers sigma_i and scatters omega_i with one fully divergent 8 B access per
// row each.  L1TEX serialises a warp request into one wavefront per 128 B line it touches (~1.3-2 cycles each,
// B300_MICROARCH.md "L1tex wavefront queue"): 2 x 32 wavefronts per 32 rows = ~68 K wavefronts per SM per launch at
// Venice-1778 — the measured 68 us of k_ps is that, not HBM (34 % of peak in ncu).
// Here a CTA owns a TILE of consecutive A rows (whole points, <= tile_rows) and exchanges through SHARED memory:
//   load    X[j] = sigma[gpos[j]]      j = 0..nb): the tile's rows in B order — gpos is piecewise CONTIGUOUS (the rows a
//                                      camera has inside the tile are neighbours in B), so a warp request touches
//                                      ~32 / (rows per camera per tile) lines instead of 32
//   points  one lane per point: t_p = sum b_i X[lpos_i]; z_p = Vinv_p t_p; X[lpos_i] = b_i . z_p   (in place; lpos = 16-bit
//                                      slot of the A row inside the tile; shared-memory accesses, no global divergence)
//   store   omega[gpos[j]] = X[j]      same piecewise-contiguous pattern
// BACKSUB: the points phase computes v_p = Vinv_p (g_p - t_p) instead and nothing is stored back.
// Summation order per point = row order of the point (as in k_ps): results are bit-identical to the chunk kernel.
// First version (one tile per CTA, 4 gathers in flight, 85-register cap with spills): 125 us at Venice against 67 us of
// the chunk kernel — every phase of a CTA was a short chain of dependent HBM latencies and 614 tiles met 444 CTA slots
// (profiles/r2_point_side.md).  This version: persistent CTAs + ticket, PT_DEPTH-deep index->value chains, the next point's
// row offsets prefetched, stored values kept as raw 32-bit registers between the two passes: 66.8 us with 4 096-row tiles,
// 76 / 110 us with 5 632 / 8 192 rows — EQUAL to the chunk kernel at best.  ncu (profiles/ncu_r2_summary.md): LSU wavefronts
// fall from 65 % to 32 % of peak as designed, but the kernel is bound by exposed load latency (long-scoreboard 12.9 of 30
// stall cycles per issue, 17 % issue utilisation), and the lane-per-point row loops live off L1, which SHRINKS as the tile
// grows (228 KB - 3 x tile).  Not adopted as default; what it needs is the streaming part (Jp, slots) staged by bulk copies
// ahead of the math — see DESIGN.md section 6.
#pragma once
#include "ba_pcg.cuh"

#ifndef PT_T
#define PT_T 256
#endif
#ifndef PT_MINB
#define PT_MINB 3   // CTAs per SM the register budget allows (shared memory: tile_rows x 8 B each)
#endif
#ifndef PT_KEEP
#define PT_KEEP 6   // rows of a point kept in registers between the two passes (longer points re-read the rest)
#endif
#ifndef PT_DEPTH
#define PT_DEPTH 16  // exchange gathers / scatters in flight per thread (two dependent HBM/L2 latencies per batch)
#endif

__device__ __forceinline__ double ldg_stream(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void stg_stream(double* p, double v) {
  asm volatile("st.global.L1::no_allocate.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
// raw bits of a stored Jacobian value: kept as an opaque 32/64-bit register between the two passes over a point's rows so
// that the compiler converts at each use instead of keeping 24 converted doubles alive (that version spilled)
__device__ __forceinline__ void opaque(float& v) { asm volatile("" : "+f"(v)); }
__device__ __forceinline__ void opaque(double& v) { asm volatile("" : "+d"(v)); }

// Persistent CTAs: tiles are handed out by an atomic ticket (tctr[0]); the last CTA to leave resets the two counters for
// the next launch (every other CTA has drawn its final ticket by then).  No wave quantisation: a CTA that gets a short
// tile simply draws the next one.
template <typename JT, bool BACKSUB>
__global__ void __launch_bounds__(PT_T, PT_MINB)
k_ps_tile(size_t N, size_t Pst, int ntiles, int C, int it, int do_scalars, double rtol, int maxiter,
          SolverState* __restrict__ st, const double* __restrict__ cpart, const int32_t* __restrict__ tile_row,
          const int32_t* __restrict__ tile_pt, const int32_t* __restrict__ pt_ptr, const int32_t* __restrict__ gpos,
          const uint16_t* __restrict__ lpos, const JT* __restrict__ Jp_A, const double* __restrict__ Vinv,
          const double* __restrict__ sig, double* __restrict__ omg, const double* __restrict__ gp,
          double* __restrict__ vp, double* __restrict__ part, int* __restrict__ tctr) {
  extern __shared__ __align__(16) double X[];
  __shared__ double sred[3 * 32];
  __shared__ int s_tile;
  const int tid = threadIdx.x;
  if (!BACKSUB && !(do_scalars & 2) && __syncthreads_or(tid == 0 ? st->pcg_done : 0)) return;
  if (!BACKSUB && (do_scalars & 1) && blockIdx.x == 0) {
#ifndef PT_NOSCAL
    pcg_scalars<4>(C, it, st, cpart, rtol, maxiter, sred);  // same per-thread summation order as the 8-deep version
#endif
    __syncthreads();
  }
  for (;;) {
    if (tid == 0) s_tile = atomicAdd(tctr, 1);
    __syncthreads();
    const int t = s_tile;
    if (t >= ntiles) break;
    const int ra = __ldg(tile_row + t), nb = __ldg(tile_row + t + 1) - ra;
    const int p0 = __ldg(tile_pt + t), np = __ldg(tile_pt + t + 1) - p0;
    // ---- load: sigma of the tile's rows, B order; PT_DEPTH independent (index -> value) chains per thread
    const int32_t* gp_t = gpos + ra;
    for (int j0 = tid; j0 < nb; j0 += PT_DEPTH * PT_T) {
      int g[PT_DEPTH];
#pragma unroll
      for (int u = 0; u < PT_DEPTH; ++u) g[u] = (j0 + u * PT_T < nb) ? __ldg(gp_t + j0 + u * PT_T) : -1;
      double sv[PT_DEPTH];
#pragma unroll
      for (int u = 0; u < PT_DEPTH; ++u) sv[u] = g[u] >= 0 ? ldg_stream(sig + g[u]) : 0.0;
#pragma unroll
      for (int u = 0; u < PT_DEPTH; ++u)
        if (g[u] >= 0) X[j0 + u * PT_T] = sv[u];
    }
    __syncthreads();
    // ---- points: one lane per point
    double acc[2] = {0.0, 0.0};
    const uint16_t* lp_t = lpos + ra;
    const JT* J0 = Jp_A + ra;
    int q = tid;
    int beg = 0, end = 0;
    if (q < np) { beg = __ldg(pt_ptr + p0 + q) - ra; end = __ldg(pt_ptr + p0 + q + 1) - ra; }
    while (q < np) {
      const int p = p0 + q;
      // row offsets of this lane's NEXT point: in flight while the current one is processed
      const int qn = q + PT_T;
      int begn = 0, endn = 0;
      if (qn < np) { begn = __ldg(pt_ptr + p0 + qn) - ra; endn = __ldg(pt_ptr + p0 + qn + 1) - ra; }
      JT b[PT_KEEP][3];
      int lp[PT_KEEP];
#pragma unroll
      for (int u = 0; u < PT_KEEP; ++u) {
        lp[u] = 0; b[u][0] = b[u][1] = b[u][2] = (JT)0;
        if (beg + u < end) {
          lp[u] = __ldg(lp_t + beg + u);
          b[u][0] = __ldg(J0 + beg + u); b[u][1] = __ldg(J0 + N + beg + u); b[u][2] = __ldg(J0 + 2 * N + beg + u);
        }
      }
      double vi[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) vi[k] = __ldg(Vinv + (size_t)k * Pst + p);
      double t0 = 0, t1 = 0, t2 = 0;
#pragma unroll
      for (int u = 0; u < PT_KEEP; ++u) {
        if (beg + u < end) {
          // product rounded, then added (no FMA contraction): the same two roundings as the chunk kernel, which passes
          // the products through shared memory — the two kernels stay bit-identical (tests/test_gpu_ba.py)
          const double s = X[lp[u]];
          t0 = __dadd_rn(t0, __dmul_rn((double)b[u][0], s)); t1 = __dadd_rn(t1, __dmul_rn((double)b[u][1], s));
          t2 = __dadd_rn(t2, __dmul_rn((double)b[u][2], s));
        }
      }
      for (int i = beg + PT_KEEP; i < end; ++i) {
        const double s = X[__ldg(lp_t + i)];
        t0 = __dadd_rn(t0, __dmul_rn((double)__ldg(J0 + i), s)); t1 = __dadd_rn(t1, __dmul_rn((double)__ldg(J0 + N + i), s));
        t2 = __dadd_rn(t2, __dmul_rn((double)__ldg(J0 + 2 * N + i), s));
      }
      if (BACKSUB) {
        const double g0 = gp[3 * (size_t)p], g1 = gp[3 * (size_t)p + 1], g2 = gp[3 * (size_t)p + 2];
        t0 = g0 - t0; t1 = g1 - t1; t2 = g2 - t2;
        double z0, z1, z2;
        ps_apply_vinv(vi, t0, t1, t2, z0, z1, z2);
        vp[3 * (size_t)p] = z0; vp[3 * (size_t)p + 1] = z1; vp[3 * (size_t)p + 2] = z2;
        acc[0] += ps_dot3(z0, z1, z2, g0, g1, g2);
        acc[1] += ps_dot3(z0, z1, z2, z0, z1, z2);
      } else {
        double z0, z1, z2;
        ps_apply_vinv(vi, t0, t1, t2, z0, z1, z2);
#pragma unroll
        for (int u = 0; u < PT_KEEP; ++u) {
          opaque(b[u][0]); opaque(b[u][1]); opaque(b[u][2]);
          if (beg + u < end) X[lp[u]] = ps_dot3((double)b[u][0], (double)b[u][1], (double)b[u][2], z0, z1, z2);
        }
        for (int i = beg + PT_KEEP; i < end; ++i)
          X[__ldg(lp_t + i)] = ps_dot3((double)__ldg(J0 + i), (double)__ldg(J0 + N + i), (double)__ldg(J0 + 2 * N + i), z0, z1, z2);
      }
      q = qn; beg = begn; end = endn;
    }
    if (BACKSUB) {
      // ordered per-tile partials: part[2t], part[2t+1] (summed in tile order by k_reduce: independent of which CTA ran the tile)
      block_sum<2>(acc, sred);
      if (tid == 0) { part[2 * t] = acc[0]; part[2 * t + 1] = acc[1]; }
      __syncthreads();
      continue;
    }
    __syncthreads();
    // ---- store: omega of the tile's rows, B order
    for (int j0 = tid; j0 < nb; j0 += PT_DEPTH * PT_T) {
      int g[PT_DEPTH];
#pragma unroll
      for (int u = 0; u < PT_DEPTH; ++u) g[u] = (j0 + u * PT_T < nb) ? __ldg(gp_t + j0 + u * PT_T) : -1;
#pragma unroll
      for (int u = 0; u < PT_DEPTH; ++u)
        if (g[u] >= 0) stg_stream(omg + g[u], X[j0 + u * PT_T]);
    }
    __syncthreads();  // X and s_tile are reused by the next tile
  }
  // last CTA out resets the ticket counters for the next launch
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(tctr + 1, 1) == (int)gridDim.x - 1) { tctr[0] = 0; tctr[1] = 0; __threadfence(); }
  }
}

// ---------------------------------------------------------------- plan construction of the tile tables
// key of A row i: (tile of the row) << 32 | B position; sorting by it lists every tile's rows in B order
__global__ void k_tile_keys(int64_t N, int ntiles, const int32_t* __restrict__ tile_row, const int32_t* __restrict__ posB,
                            int64_t* __restrict__ keys, int32_t* __restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  int lo = 0, hi = ntiles;  // last tile with tile_row[t] <= i
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(tile_row + mid) <= i) lo = mid; else hi = mid;
  }
  keys[i] = ((int64_t)lo << 32) | (uint32_t)posB[i];
  vals[i] = (int32_t)i;
}
// j-th row in (tile, B position) order: gpos[j] = its B position; lpos[A row] = j - first row of the tile
__global__ void k_tile_fill(int64_t N, const int32_t* __restrict__ tile_row, const int64_t* __restrict__ keys_sorted,
                            const int32_t* __restrict__ rows_sorted, int32_t* __restrict__ gpos, uint16_t* __restrict__ lpos) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= N) return;
  const int64_t k = keys_sorted[j];
  gpos[j] = (int32_t)(k & 0xffffffffLL);
  lpos[rows_sorted[j]] = (uint16_t)(j - __ldg(tile_row + (int)(k >> 32)));
}
