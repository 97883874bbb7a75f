// Point side of the Schur matvec, TILED version (the default whenever a tile still holds a few rows of every camera).
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
#pragma once
#include "ba_pcg.cuh"

#ifndef PT_T
#define PT_T 256
#endif
#ifndef PT_MINB
#define PT_MINB 3   // CTAs per SM the register budget allows (shared memory: tile_rows x 8 B each)
#endif
#ifndef PT_KEEP
#define PT_KEEP 8   // rows of a point kept in registers between the two passes (longer points re-read the rest)
#endif

__device__ __forceinline__ double ldg_stream(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void stg_stream(double* p, double v) {
  asm volatile("st.global.L1::no_allocate.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

template <typename JT, bool BACKSUB>
__global__ void __launch_bounds__(PT_T, PT_MINB)
k_ps_tile(size_t N, size_t Pst, int ntiles, int C, int it, int do_scalars, double rtol, int maxiter,
          SolverState* __restrict__ st, const double* __restrict__ cpart, const int32_t* __restrict__ tile_row,
          const int32_t* __restrict__ tile_pt, const int32_t* __restrict__ pt_ptr, const int32_t* __restrict__ gpos,
          const uint16_t* __restrict__ lpos, const JT* __restrict__ Jp_A, const double* __restrict__ Vinv,
          const double* __restrict__ sig, double* __restrict__ omg, const double* __restrict__ gp,
          double* __restrict__ vp, double* __restrict__ part) {
  extern __shared__ __align__(16) double X[];
  __shared__ double sred[3 * 32];
  const int t = blockIdx.x, tid = threadIdx.x;
  int ra = 0, nb = 0, p0 = 0, np = 0;
  if (t < ntiles) {
    ra = __ldg(tile_row + t); nb = __ldg(tile_row + t + 1) - ra;
    p0 = __ldg(tile_pt + t); np = __ldg(tile_pt + t + 1) - p0;
  }
  if (!BACKSUB && !(do_scalars & 2) && __syncthreads_or(tid == 0 ? st->pcg_done : 0)) return;
  if (!BACKSUB && (do_scalars & 1) && t == 0) {
#ifndef PT_NOSCAL
    pcg_scalars(C, it, st, cpart, rtol, maxiter, sred);
#endif
    __syncthreads();
  }
  if (t >= ntiles) return;
  // ---- load: sigma of the tile's rows, B order
  const int32_t* gp_t = gpos + ra;
  {
    int j = tid;
    for (; j + 3 * PT_T < nb; j += 4 * PT_T) {
      const int g0 = __ldg(gp_t + j), g1 = __ldg(gp_t + j + PT_T), g2 = __ldg(gp_t + j + 2 * PT_T), g3 = __ldg(gp_t + j + 3 * PT_T);
      const double s0 = ldg_stream(sig + g0), s1 = ldg_stream(sig + g1), s2 = ldg_stream(sig + g2), s3 = ldg_stream(sig + g3);
      X[j] = s0; X[j + PT_T] = s1; X[j + 2 * PT_T] = s2; X[j + 3 * PT_T] = s3;
    }
    for (; j < nb; j += PT_T) X[j] = ldg_stream(sig + __ldg(gp_t + j));
  }
  __syncthreads();
  // ---- points: one lane per point
  double acc[2] = {0.0, 0.0};
  const uint16_t* lp_t = lpos + ra;
  const JT* J0 = Jp_A + ra;
  for (int q = tid; q < np; q += PT_T) {
    const int p = p0 + q;
    const int beg = __ldg(pt_ptr + p) - ra, end = __ldg(pt_ptr + p + 1) - ra;
    double vi[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) vi[k] = __ldg(Vinv + (size_t)k * Pst + p);
    double g0 = 0, g1 = 0, g2 = 0;
    if (BACKSUB) { g0 = gp[3 * (size_t)p]; g1 = gp[3 * (size_t)p + 1]; g2 = gp[3 * (size_t)p + 2]; }
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
    double t0 = 0, t1 = 0, t2 = 0;
#pragma unroll
    for (int u = 0; u < PT_KEEP; ++u) {
      if (beg + u < end) {
        const double s = X[lp[u]];
        t0 += (double)b[u][0] * s; t1 += (double)b[u][1] * s; t2 += (double)b[u][2] * s;
      }
    }
    for (int i = beg + PT_KEEP; i < end; ++i) {
      const double s = X[__ldg(lp_t + i)];
      t0 += (double)__ldg(J0 + i) * s; t1 += (double)__ldg(J0 + N + i) * s; t2 += (double)__ldg(J0 + 2 * N + i) * s;
    }
    if (BACKSUB) {
      t0 = g0 - t0; t1 = g1 - t1; t2 = g2 - t2;
      const double z0 = vi[0] * t0 + vi[1] * t1 + vi[2] * t2;
      const double z1 = vi[1] * t0 + vi[3] * t1 + vi[4] * t2;
      const double z2 = vi[2] * t0 + vi[4] * t1 + vi[5] * t2;
      vp[3 * (size_t)p] = z0; vp[3 * (size_t)p + 1] = z1; vp[3 * (size_t)p + 2] = z2;
      acc[0] += z0 * g0 + z1 * g1 + z2 * g2;
      acc[1] += z0 * z0 + z1 * z1 + z2 * z2;
    } else {
      const double z0 = vi[0] * t0 + vi[1] * t1 + vi[2] * t2;
      const double z1 = vi[1] * t0 + vi[3] * t1 + vi[4] * t2;
      const double z2 = vi[2] * t0 + vi[4] * t1 + vi[5] * t2;
#pragma unroll
      for (int u = 0; u < PT_KEEP; ++u)
        if (beg + u < end) X[lp[u]] = (double)b[u][0] * z0 + (double)b[u][1] * z1 + (double)b[u][2] * z2;
      for (int i = beg + PT_KEEP; i < end; ++i)
        X[__ldg(lp_t + i)] = (double)__ldg(J0 + i) * z0 + (double)__ldg(J0 + N + i) * z1 + (double)__ldg(J0 + 2 * N + i) * z2;
    }
  }
  if (BACKSUB) {
    block_sum<2>(acc, sred);
    if (tid == 0) { part[2 * t] = acc[0]; part[2 * t + 1] = acc[1]; }
    return;
  }
  __syncthreads();
  // ---- store: omega of the tile's rows, B order
  {
    int j = tid;
    for (; j + 3 * PT_T < nb; j += 4 * PT_T) {
      const int g0 = __ldg(gp_t + j), g1 = __ldg(gp_t + j + PT_T), g2 = __ldg(gp_t + j + 2 * PT_T), g3 = __ldg(gp_t + j + 3 * PT_T);
      stg_stream(omg + g0, X[j]); stg_stream(omg + g1, X[j + PT_T]); stg_stream(omg + g2, X[j + 2 * PT_T]); stg_stream(omg + g3, X[j + 3 * PT_T]);
    }
    for (; j < nb; j += PT_T) stg_stream(omg + __ldg(gp_t + j), X[j]);
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
