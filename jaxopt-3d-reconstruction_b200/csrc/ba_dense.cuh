// K3e — dense reduced-camera solve for SMALL scenes (north_star subsystem 3).
//
// For 8C <= DENSE_MAX_N the reduced camera system S v_c = rhs is formed DENSELY and factored by a
// Cholesky whose trailing updates run on the FP64 tensor cores (mma.sync.m8n8k4.f64, "DMMA"), instead of ~50-120
// launch-latency-bound PCG iterations.  This is the path of the reference's own benchmark size (15 cameras x <=400 points,
// src/benchmark_implementation/benchmark_impl_bundle_adjustment.py:28-32), of Ladybug-49 and of Trafalgar-257; the
// reference itself solves its pose problems with solver="cholesky" (pose_optimization.py:62-71).
//
//   plan time   pair list: for every point, every ordered pair (i, j) of its rows with cam_i >= cam_j, sorted by the
//               camera pair -> one contiguous segment per non-zero 8x8 block of the lower triangle of S
//   assemble    8 lanes per block (c, d), d <= c:  S_cd = [c == d](U_c + mu I) - sum_pairs a_i (b_i^T Vinv_p b_j) a_j^T
//               (rows and Vinv gathered from the linearisation's arrays; fixed pair order: deterministic, no atomics)
//   factor      tile dataflow, 32 x 32 tiles, ONE cooperative persistent launch: tile (i, j) is owned by one CTA, which
//               accumulates -sum_k L_ik L_jk^T on the tensor core as the L tiles of earlier columns are published
//               (per-tile flags, st.release / ld.acquire at gpu scope), then finishes with a warp-level Cholesky
//               (i == j, rows in registers, shuffles) or a triangular solve against L_jj; no grid-wide barrier anywhere
//   solve       L y = rhs, L^T x = y with the same flag dataflow over block rows; x -> slot X of the camera records, so
//               back-substitution, trial step and gain ratio run unchanged
// All of it is exact up to rounding (no tolerance), like jaxopt's solve_cholesky.
#pragma once
#include "ba_pcg.cuh"

#define DN_NB 32          // tile size
#define DN_T 128          // threads per CTA of the factorisation / solve kernels (4 warps)
#define DN_LD 36          // shared-memory tile stride in doubles (== 4 mod 16: the DMMA fragment loads are conflict-free)
#define DENSE_MAX_N 2560  // hard 8C limit (320 cameras): S is n^2 * 8 B = 52 MB at the limit
#define DENSE_DEFAULT_CAMS 128  // selected automatically up to this many cameras (measured crossover with the PCG path)

__device__ __forceinline__ int dn_ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void dn_st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// 256-bit load through L2 only (.cg): tiles published by OTHER CTAs during the same launch must not come from the
// non-coherent read-only path
__device__ __forceinline__ void ld256cg(const double* p, double* v) {
  asm volatile("ld.global.cg.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p) : "memory");
}
__device__ __forceinline__ void dn_wait(const int* f, int epoch) {
  while (dn_ld_acquire(f) < epoch) __nanosleep(64);
}

// ---------------------------------------------------------------- plan time: the pair list
// pairs of point p: rows ia > ib of its A run (A rows of a point are sorted by camera, so cam[ia] >= cam[ib]).  The self
// pairs ia == ib (one per row: they would make every diagonal block a chain as long as the camera has rows) are not
// listed: their sum  -sum_i beta_i a_i a_i^T  is the block-Jacobi correction Sd_c that k_cam_blocks already reduces.
__global__ void k_dn_pair_count(int P, const int32_t* __restrict__ pt_ptr, int64_t* __restrict__ cnt) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int64_t d = pt_ptr[p + 1] - pt_ptr[p];
  cnt[p] = d * (d - 1) / 2;
}
__global__ void k_dn_pair_fill(int P, int C, const int32_t* __restrict__ pt_ptr, const int32_t* __restrict__ cam_ps,
                               const int64_t* __restrict__ off, int32_t* __restrict__ keys, int2* __restrict__ vals) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int beg = pt_ptr[p], end = pt_ptr[p + 1];
  int64_t o = off[p];
  for (int ia = beg; ia < end; ++ia)
    for (int ib = beg; ib < ia; ++ib) {
      keys[o] = cam_ps[ia] * C + cam_ps[ib];
      vals[o] = make_int2(ia, ib);
      ++o;
    }
}
// identity on the padded part of S (cameras C .. Cpad-1), written once per plan: the factorisation leaves it alone
__global__ void k_dn_init_pad(int n, int npad, double* __restrict__ S) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t tot = (int64_t)npad * npad;
  if (e >= tot) return;
  const int r = (int)(e / npad), c = (int)(e % npad);
  if (r >= n || c >= n) S[e] = (r == c) ? 1.0 : 0.0;
}

// ---------------------------------------------------------------- assemble
// One WARP per block (c, d), d <= c: four groups of 8 lanes take every fourth pair of the block's segment (lane k of a
// group owns row k of the 8x8 block), the four partial blocks are added by two shuffle steps in a fixed order.
// S_cd = [c == d](U_c + mu I + Sd_c) - sum_pairs a_i (b_i^T Vinv_p b_j) a_j^T   (Sd_c: the self pairs, from k_cam_blocks)
#define DN_ASM_T 256
template <typename JT>
__global__ void __launch_bounds__(DN_ASM_T)
k_dn_assemble(int C, int npad, size_t N, size_t Pst, const double* __restrict__ mu_ptr, const int32_t* __restrict__ blk_ptr,
              const int2* __restrict__ pairs, const int32_t* __restrict__ posB, const int32_t* __restrict__ pt_ps,
              const JT* __restrict__ Jc_B, const JT* __restrict__ Jp_ps, const double* __restrict__ Vinv,
              const double* __restrict__ U, const double* __restrict__ Sd, double* __restrict__ S, SolverState* __restrict__ st) {
  const int c = blockIdx.y;
  const int d = blockIdx.x * (DN_ASM_T / 32) + (threadIdx.x >> 5);
  // next epoch of the tile / block-row flags of the factorisation and the sweeps (they read it at their start, after this
  // launch): one thread of the grid; was a launch of its own
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) st->dn_epoch += 1;
  if (d > c) return;  // warp-uniform
  const int lane = threadIdx.x & 31, k = lane & 7, sg = lane >> 3;
  const int64_t gq = (int64_t)c * C + d;
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int beg = __ldg(blk_ptr + gq), end = __ldg(blk_ptr + gq + 1);
  for (int q = beg + sg; q < end; q += 4) {
    const int2 pr = __ldg(pairs + q);
    const int ia = pr.x, ib = pr.y;
    const int p = __ldg(pt_ps + ia);
    const int ja = __ldg(posB + ia), jb = __ldg(posB + ib);
    double vi[6];
#pragma unroll
    for (int m = 0; m < 6; ++m) vi[m] = __ldg(Vinv + (size_t)m * Pst + p);
    const double a0 = (double)__ldg(Jp_ps + ia), a1 = (double)__ldg(Jp_ps + N + ia), a2 = (double)__ldg(Jp_ps + 2 * N + ia);
    const double b0 = (double)__ldg(Jp_ps + ib), b1 = (double)__ldg(Jp_ps + N + ib), b2 = (double)__ldg(Jp_ps + 2 * N + ib);
    double ra[8], rb[8];
    load8<JT>(Jc_B + 8 * (size_t)ja, ra);
    load8<JT>(Jc_B + 8 * (size_t)jb, rb);
    const double t0 = vi[0] * b0 + vi[1] * b1 + vi[2] * b2;
    const double t1 = vi[1] * b0 + vi[3] * b1 + vi[4] * b2;
    const double t2 = vi[2] * b0 + vi[4] * b1 + vi[5] * b2;
    const double beta = a0 * t0 + a1 * t1 + a2 * t2;
    double ak = 0.0, bk = 0.0;
#pragma unroll
    for (int l = 0; l < 8; ++l) { if (l == k) { ak = ra[l]; bk = rb[l]; } }
    const double wa = ak * beta;
#pragma unroll
    for (int l = 0; l < 8; ++l) acc[l] += wa * rb[l];
    if (c == d) {  // the same camera sees the point twice (ia != ib): both orders belong to the diagonal block
      const double wb = bk * beta;
#pragma unroll
      for (int l = 0; l < 8; ++l) acc[l] += wb * ra[l];
    }
  }
#pragma unroll
  for (int l = 0; l < 8; ++l) {
    acc[l] += __shfl_xor_sync(0xffffffffu, acc[l], 8);
    acc[l] += __shfl_xor_sync(0xffffffffu, acc[l], 16);
  }
  if (sg) return;
  const double mu = *mu_ptr;
  double* out = S + (size_t)(8 * c + k) * npad + 8 * d;
#pragma unroll
  for (int l = 0; l < 8; ++l) {
    double v = -acc[l];
    if (c == d) v += U[64 * (size_t)c + 8 * k + l] + Sd[64 * (size_t)c + 8 * k + l] + (l == k ? mu : 0.0);
    out[l] = v;
  }
}

// ---------------------------------------------------------------- factor
// Warp-level Cholesky (Cholesky-Crout, left-looking) of a 32x32 tile in SHARED memory (stride DN_LD); lane = row, kept in
// REGISTERS with static indices (the column loop is unrolled at compile time by template recursion), finished columns are
// published to shared memory as they appear and read back as broadcasts.  Only the lower triangle is read.  A non-positive
// pivot gives NaN (like LAPACK potrf through jax).  The dot products run as four independent chains.
// History (GPU, per tile): right-looking with shared-memory read-modify-write chains 18 us; left-looking with run-time
// loops over shared memory ~7 us; a right-looking register version with 496 unrolled shuffles compiled to 19 K instructions.
template <int C>
struct DnCol {
  static __device__ __forceinline__ double dot(const double (&row)[DN_NB], const double* __restrict__ rc) {
    double s[4] = {row[C], 0.0, 0.0, 0.0};
#pragma unroll
    for (int q = 0; q < C; ++q) s[q & 3] -= row[q] * rc[q];
    return (s[0] + s[1]) + (s[2] + s[3]);
  }
  static __device__ __forceinline__ void potrf(double (&row)[DN_NB], double* __restrict__ A, int lane) {
    DnCol<C - 1>::potrf(row, A, lane);
    const double sv = dot(row, A + C * DN_LD);
    if (lane == C) A[C * DN_LD + C] = sqrt(sv);
    __syncwarp();
    const double v = (lane == C) ? A[C * DN_LD + C] : sv / A[C * DN_LD + C];
    row[C] = v;
    if (lane > C) A[lane * DN_LD + C] = v;
    __syncwarp();
  }
  // X L^T = R for one row per lane: row[C] = (row[C] - sum_{q<C} row[q] L[C][q]) / L[C][C]; L in shared memory (broadcasts)
  static __device__ __forceinline__ void trsm(double (&row)[DN_NB], const double* __restrict__ Ls) {
    DnCol<C - 1>::trsm(row, Ls);
    row[C] = dot(row, Ls + C * DN_LD) / Ls[C * DN_LD + C];
  }
};
template <>
struct DnCol<-1> {
  static __device__ __forceinline__ void potrf(double (&)[DN_NB], double* __restrict__, int) {}
  static __device__ __forceinline__ void trsm(double (&)[DN_NB], const double* __restrict__) {}
};

// One persistent cooperative launch.  Lower tiles in column-major order idx = 0, 1, ...; CTA b owns idx = b, b + G, ...
// and processes them in that order (every dependency of a tile has a smaller idx: no cycles, all CTAs resident).
__global__ void __launch_bounds__(DN_T)
k_dn_factor(int T, int npad, double* __restrict__ S, int* __restrict__ flags, SolverState* __restrict__ st) {
  const int epoch = st->dn_epoch;  // device-side counter (k_dn_next_epoch): the launch sequence can be replayed from a CUDA graph
  __shared__ __align__(16) double As[DN_NB * DN_LD];
  __shared__ __align__(16) double Bs[DN_NB * DN_LD];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int lr = tid >> 2, lc = (tid & 3) * 8;  // this thread's 8 consecutive elements of a tile when it is moved
  const int ntl = T * (T + 1) / 2;
  int j = 0, colstart = 0;                       // column of the current idx, idx of tile (j, j)
  for (int idx = blockIdx.x; idx < ntl; idx += gridDim.x) {
    while (idx >= colstart + (T - j)) { colstart += T - j; ++j; }
    const int i = j + (idx - colstart);
    double* Cg = S + (size_t)(DN_NB * i) * npad + DN_NB * j;
    // accumulators: warp w owns rows 8w..8w+7, four 8x8 n-tiles; init from S
    double acc[4][2];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const double2 v = *reinterpret_cast<const double2*>(Cg + (size_t)(8 * w + g) * npad + 8 * nt + 2 * t);
      acc[nt][0] = v.x; acc[nt][1] = v.y;
    }
    double va[8], vb[8];
    auto wait_k = [&](int k) {
      if (tid == 0) dn_wait(flags + i * T + k, epoch);
      if (tid == 1 && i != j) dn_wait(flags + j * T + k, epoch);
    };
    auto load_k = [&](int k) {
      const double* Ag = S + (size_t)(DN_NB * i + lr) * npad + DN_NB * k + lc;
      ld256cg(Ag, va); ld256cg(Ag + 4, va + 4);
      if (i != j) {
        const double* Bg = S + (size_t)(DN_NB * j + lr) * npad + DN_NB * k + lc;
        ld256cg(Bg, vb); ld256cg(Bg + 4, vb + 4);
      }
    };
    if (j > 0) { wait_k(0); __syncthreads(); load_k(0); }
    for (int k = 0; k < j; ++k) {
#pragma unroll
      for (int e = 0; e < 8; ++e) As[lr * DN_LD + lc + e] = va[e];
      if (i != j) {
#pragma unroll
        for (int e = 0; e < 8; ++e) Bs[lr * DN_LD + lc + e] = vb[e];
      }
      if (k + 1 < j) wait_k(k + 1);
      __syncthreads();
      if (k + 1 < j) load_k(k + 1);  // in flight while the tensor core works on step k
      const double* Bt = (i != j) ? Bs : As;
#pragma unroll
      for (int kk = 0; kk < DN_NB; kk += 4) {
        const double a = -As[(8 * w + g) * DN_LD + kk + t];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) dmma_m8n8k4(acc[nt][0], acc[nt][1], a, Bt[(8 * nt + g) * DN_LD + kk + t]);
      }
      __syncthreads();
    }
    // epilogue: updated tile -> shared memory -> one row per lane of warp 0
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      As[(8 * w + g) * DN_LD + 8 * nt + 2 * t] = acc[nt][0];
      As[(8 * w + g) * DN_LD + 8 * nt + 2 * t + 1] = acc[nt][1];
    }
    if (i != j) {
      if (tid == 0) dn_wait(flags + j * T + j, epoch);
      __syncthreads();
      const double* Lg = S + (size_t)(DN_NB * j + lr) * npad + DN_NB * j + lc;
      double vb[8];
      ld256cg(Lg, vb); ld256cg(Lg + 4, vb + 4);
#pragma unroll
      for (int e = 0; e < 8; ++e) Bs[lr * DN_LD + lc + e] = vb[e];
    }
    __syncthreads();
    if (w == 0) {
      double row[DN_NB];
#pragma unroll
      for (int cc = 0; cc < DN_NB; ++cc) row[cc] = As[lane * DN_LD + cc];
      if (i == j) {
        DnCol<DN_NB - 1>::potrf(row, As, lane);
        bool bad = false;
#pragma unroll
        for (int cc = 0; cc < DN_NB; ++cc) {
          if (cc > lane) row[cc] = 0.0;  // clean upper triangle of the diagonal tile
          else bad |= !(row[cc] == row[cc]);
        }
        if (bad) atomicOr(&st->status, 16);  // not positive definite -> NaN, reported as B200BA_ST_DENSE_NOT_SPD
      } else {
        DnCol<DN_NB - 1>::trsm(row, Bs);
      }
      double* out = Cg + (size_t)lane * npad;
#pragma unroll
      for (int cc = 0; cc < DN_NB; cc += 2) *reinterpret_cast<double2*>(out + cc) = make_double2(row[cc], row[cc + 1]);
      __threadfence();
      __syncwarp();
      if (lane == 0) dn_st_release(flags + i * T + j, epoch);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- triangular solves
// BACKWARD = false: L y = b (block rows ascending);  true: L^T x = y (descending).  CTA b owns block rows b, b + G, ...
// in processing order.  vec: in/out (npad doubles); vflag[T]: block row published (epoch).
template <bool BACKWARD>
__global__ void __launch_bounds__(DN_T)
k_dn_trsv(int T, int npad, const SolverState* __restrict__ st, const double* __restrict__ L, double* __restrict__ vec, int* __restrict__ vflag,
          int n, const double* __restrict__ rhs, double* __restrict__ cv) {
  const int epoch = st->dn_epoch;
  __shared__ double Ls[DN_NB * (DN_NB + 1)];
  __shared__ double xs[DN_NB];
  __shared__ double part[DN_T];
  const int tid = threadIdx.x, lane = tid & 31;
  const int rr = tid >> 2, cs = (tid & 3) * 8;  // row of the tile / first of 8 columns this thread multiplies
  for (int o = blockIdx.x; o < T; o += gridDim.x) {
    const int j = BACKWARD ? T - 1 - o : o;
    double s = 0.0;
    const int nk = BACKWARD ? T - 1 - j : j;
    for (int kk = 0; kk < nk; ++kk) {
      const int k = BACKWARD ? T - 1 - kk : kk;  // block whose solution is consumed
      // forward: tile (j, k), row rr, columns cs..cs+7;  backward: tile (k, j) transposed: element (row cs+e of tile, column rr)
      double lv[8];
      if (!BACKWARD) {
        const double* Lg = L + (size_t)(DN_NB * j + rr) * npad + DN_NB * k + cs;
        ld256(Lg, lv); ld256(Lg + 4, lv + 4);
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) lv[e] = __ldg(L + (size_t)(DN_NB * k + cs + e) * npad + DN_NB * j + rr);
      }
      if (tid == 0) dn_wait(vflag + k, epoch);
      __syncthreads();
      if (tid < DN_NB) xs[tid] = __ldcg(vec + DN_NB * k + tid);
      __syncthreads();
#pragma unroll
      for (int e = 0; e < 8; ++e) s += lv[e] * xs[cs + e];
      __syncthreads();
    }
    // diagonal tile -> shared memory (forward: L_jj; backward: its transpose, so that the solve below is the same loop)
    {
      const double* Lg = L + (size_t)(DN_NB * j + rr) * npad + DN_NB * j + cs;
      double lv[8];
      ld256(Lg, lv); ld256(Lg + 4, lv + 4);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        if (!BACKWARD) Ls[rr * (DN_NB + 1) + cs + e] = lv[e];
        else Ls[(cs + e) * (DN_NB + 1) + rr] = lv[e];
      }
    }
    part[tid] = s;
    __syncthreads();
    if (tid < DN_NB) {
      // lane = row of the block: b_r - (the four partial sums of this row, fixed order), then substitution by shuffles.
      // forward: lower-triangular solve in ascending order; backward: Ls holds L_jj^T (upper): descending order
      // forward sweep: the right-hand side comes straight from the reduced rhs (C x 8, zero beyond n); backward: y from vec
      const int e0 = DN_NB * j + lane;
      const double b0 = BACKWARD ? __ldcg(vec + e0) : (e0 < n ? rhs[e0] : 0.0);
      double v = b0 - (((part[4 * lane] + part[4 * lane + 1]) + part[4 * lane + 2]) + part[4 * lane + 3]);
#pragma unroll
      for (int cc0 = 0; cc0 < DN_NB; ++cc0) {
        const int cc = BACKWARD ? DN_NB - 1 - cc0 : cc0;
        const double xc = __shfl_sync(0xffffffffu, v, cc) / Ls[cc * (DN_NB + 1) + cc];
        if (lane == cc) v = xc;
        else if (BACKWARD ? lane < cc : lane > cc) v -= Ls[lane * (DN_NB + 1) + cc] * xc;
      }
      vec[DN_NB * j + lane] = v;
      if (BACKWARD && e0 < n) cv[(size_t)CV_STRIDE * (e0 >> 3) + 8 * CV_X + (e0 & 7)] = v;  // the solution -> slot X of the camera records
      __threadfence();
      __syncwarp();
      if (lane == 0) dn_st_release(vflag + j, epoch);
    }
    __syncthreads();
  }
}
