// Bundle-adjustment kernels (K0-K3 of SURVEY.md §7.2b), hand-written for sm_100a.
// Observation rows live in two orders: "ps" = sorted by (point, camera), "cs" = sorted by
// (camera, point).  JT is the storage type of Jacobian rows (double: parity default; float).
// All reductions are ordered (no floating-point atomics) so results are bit-reproducible.
#pragma once
#include "common.cuh"

// 256-bit read-only loads (sm_100: LDG.E.256): one request per 32 B sector instead of two 128-bit
// ones — halves the L1 wavefronts of every 32/64 B row gather.  p must be 32 B aligned.
__device__ __forceinline__ void ld256(const double* __restrict__ p, double* v) {
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}
__device__ __forceinline__ void ld256(const float* __restrict__ p, float* v) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p));
}
template <typename JT>
__device__ __forceinline__ void load8(const JT* __restrict__ p, double* a);
template <>
__device__ __forceinline__ void load8<double>(const double* __restrict__ p, double* a) {
  ld256(p, a);
  ld256(p + 4, a + 4);
}
template <>
__device__ __forceinline__ void load8<float>(const float* __restrict__ p, double* a) {
  float v[8];
  ld256(p, v);
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = v[k];
}
template <typename JT>
__device__ __forceinline__ void store8(JT* __restrict__ p, const double* a);
template <>
__device__ __forceinline__ void store8<double>(double* __restrict__ p, const double* a) {
  double2* q = reinterpret_cast<double2*>(p);
  q[0] = make_double2(a[0], a[1]); q[1] = make_double2(a[2], a[3]);
  q[2] = make_double2(a[4], a[5]); q[3] = make_double2(a[6], a[7]);
}
template <>
__device__ __forceinline__ void store8<float>(float* __restrict__ p, const double* a) {
  float4* q = reinterpret_cast<float4*>(p);
  q[0] = make_float4((float)a[0], (float)a[1], (float)a[2], (float)a[3]);
  q[1] = make_float4((float)a[4], (float)a[5], (float)a[6], (float)a[7]);
}
// ps-order rows are stored SoA: value k of row i lives at base[k * N + i], so a warp that maps
// one lane to one row issues perfectly coalesced loads/stores (8 camera + 3 point arrays).
template <typename JT>
__device__ __forceinline__ double jget(const JT* __restrict__ base, size_t N, int k, size_t i) {
  return (double)__ldg(base + (size_t)k * N + i);
}
template <typename JT>
__device__ __forceinline__ void jput(JT* __restrict__ base, size_t N, int k, size_t i, double v) {
  base[(size_t)k * N + i] = (JT)v;
}

// raw (storage-precision) 8-value row moves: global -> registers -> shared -> registers
__device__ __forceinline__ void ldrow(const float* __restrict__ p, float* v) { ld256(p, v); }
__device__ __forceinline__ void ldrow(const double* __restrict__ p, double* v) { ld256(p, v); ld256(p + 4, v + 4); }
__device__ __forceinline__ void strow(float* p, const float* v) {
  float4* q = reinterpret_cast<float4*>(p);
  q[0] = make_float4(v[0], v[1], v[2], v[3]);
  q[1] = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void strow(double* p, const double* v) {
  double2* q = reinterpret_cast<double2*>(p);
  q[0] = make_double2(v[0], v[1]); q[1] = make_double2(v[2], v[3]);
  q[2] = make_double2(v[4], v[5]); q[3] = make_double2(v[6], v[7]);
}
__device__ __forceinline__ void ldrow_shared(const float* p, float* v) {
  const float4* q = reinterpret_cast<const float4*>(p);
  const float4 a = q[0], b = q[1];
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void ldrow_shared(const double* p, double* v) {
  const double2* q = reinterpret_cast<const double2*>(p);
#pragma unroll
  for (int k = 0; k < 4; ++k) { const double2 a = q[k]; v[2 * k] = a.x; v[2 * k + 1] = a.y; }
}

// Rows of one camera in B order = nPB runs (each padded to a multiple of 4 rows with all-zero rows, so every
// run is 16 B-aligned in all per-row arrays and can be moved by bulk copies).  A CTA flattens them:
// logical row l in [0, L) -> B position via the run prefix kept in shared memory (locate() scans
// forward from the caller's previous run, l only grows).
#define MAXPB 256
struct CamRuns {
  int beg[MAXPB];
  int pre[MAXPB + 4];
  // camtab: per camera CAMTAB_STRIDE(nPB) ints = beg[nPB] (padded to 4) | pre[nPB+1] (padded to 4), built on the host
  __device__ __forceinline__ int load(int c, int nPB, const int32_t* __restrict__ camtab) {
    const int nb = (nPB + 3) & ~3;
    const int32_t* t = camtab + (size_t)c * (nb + ((nPB + 4) & ~3));
    for (int r = threadIdx.x; r < nPB; r += blockDim.x) beg[r] = __ldg(t + r);
    for (int r = threadIdx.x; r <= nPB; r += blockDim.x) pre[r] = __ldg(t + nb + r);
    __syncthreads();
    return pre[nPB];
  }
  __device__ __forceinline__ size_t locate(int l, int& r) const {
    while (l >= pre[r + 1]) ++r;
    return (size_t)beg[r] + (size_t)(l - pre[r]);
  }
};

// 32 B row records {b[3], r} of the A order: one 256-bit store / load each (one request per sector).  fp64 in BOTH builds
// (the record is one sector either way): V_p and g_p = sum b r are accumulated from unrounded rows, so the gradient is
// exact also with fp32-STORED Jacobian rows (the 2-vector model does not converge to tol on a gradient with fp32 noise).
struct __align__(32) RowRec { double b[3]; double r; };
static_assert(sizeof(RowRec) == 32, "row record = one 32 B sector");
__device__ __forceinline__ void st_rec(RowRec* p, const double* b, double r) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(r) : "memory");
}
__device__ __forceinline__ void ld_rec(const RowRec* p, double* b, double& r) {
  double v[4];
  ld256(reinterpret_cast<const double*>(p), v);
  b[0] = v[0]; b[1] = v[1]; b[2] = v[2]; r = v[3];
}

// ---------------------------------------------------------------- K0+K1+K2b: residual + Jacobian rows + camera blocks
// One CTA per camera walks the camera's rows in B order (runs of consecutive points inside a point
// block).  The 26-double camera table is built once per CTA in shared memory (every lane reads the same
// entries: broadcasts), the only gather is the 32 B point record (ONE 256-bit request per row), and it stays inside the
// block's point range (L2-resident).  Per row it writes, coalesced in B order, the camera part of the Jacobian row
// (8 x JT), and SCATTERS one 32 B record {b (3 x fp64), r} to the row's A position (a whole sector, inside the L2
// window of the point block) from which k_point_blocks builds V_p, g_p and the SoA point part.
// FUSED camera blocks (round 2): the rows never leave the SM between K1 and K2b.  g_c = sum a r is accumulated from the
// UNROUNDED fp64 row (exact gradient also with fp32-stored rows); the row as STORED (rounded to JT) is stashed in a
// shared-memory tile and U_c = sum a a^T is accumulated from that in a second phase per tile — the camera block is
// therefore exactly the one of the stored operator, (U - W V^-1 W^T) stays the Schur complement of J~^T J~ + mu I.
// Replaces the separate k_cam_blocks launch over the linearisation (40 B/row re-read + 8 B/row r_B write gone).
// MODEL 0: the reference's scalar residual r = (dx^2 + dy^2)/w^2 (bundle_adjustment.py:17-21,74), scale = 1/w^2.
// MODEL 1: the 2-vector residual of the reference's COLMAP comparator over 2N rows (colmap/src/base/cost_functions.h:44-86,
//          principal point fixed): observation (x, NaN) marks the x row, (NaN, y) the y row, scale = 1/w.
// cost_part[c] = sum r^2 over the camera.  WITH_JAC = false: trial cost only (reads 24 B/row, writes nothing).
#define RJ_THREADS 256
#define RJ_TILE_BYTES (64 * 1024)  // shared-memory row stash of the fused kernel: 2048 fp32 rows / 1024 fp64 rows
template <typename JT> __device__ __forceinline__ void strow_cvt(JT* p, const double* a);
template <> __device__ __forceinline__ void strow_cvt<float>(float* p, const double* a) {
  float4* q = reinterpret_cast<float4*>(p);
  q[0] = make_float4((float)a[0], (float)a[1], (float)a[2], (float)a[3]);
  q[1] = make_float4((float)a[4], (float)a[5], (float)a[6], (float)a[7]);
}
template <> __device__ __forceinline__ void strow_cvt<double>(double* p, const double* a) { strow(p, a); }

// (robust_residual<MODEL>: the 2-vector model under Ceres' Cauchy / soft-L1 loss — common.cuh, next to v2_residual)
template <int MODEL>
__device__ __forceinline__ double row_residual(const Proj& pr, double ox, double oy, double scale, double& ex, double& ey) {
  if (MODEL == 1) return v2_residual(pr, ox, oy, scale, ex, ey);
  const double dx = pr.u - ox, dy = pr.v - oy;
  ex = 2.0 * dx * scale;
  ey = 2.0 * dy * scale;
  return (dx * dx + dy * dy) * scale;
}

// D (8x8, fp64) += A (8x4) * B (4x8) on the FP64 tensor core.  Fragment layout (PTX ISA, mma.m8n8k4 .f64), g = lane >> 2,
// t = lane & 3:  a = A[g][t],  b = B[t][g],  d0 = D[g][2t],  d1 = D[g][2t+1].
__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double stash_get(const float* p) { return (double)*p; }
__device__ __forceinline__ double stash_get(const double* p) { return *p; }

template <typename JT, int MODEL, bool WITH_JAC>
__global__ void __launch_bounds__(RJ_THREADS, WITH_JAC ? 3 : 4)
k_resjac(int nPB, const int32_t* __restrict__ camtab, const double* __restrict__ cam8, const double* __restrict__ k3,
         const int32_t* __restrict__ pt_B, const double2* __restrict__ obs_B, const int32_t* __restrict__ posA,
         const double* __restrict__ pts4, double scale, JT* __restrict__ Jc_B, RowRec* __restrict__ rec_A,
         double* __restrict__ cost_part, double* __restrict__ U, double* __restrict__ gc, int exact_grad, double rho_b) {
  extern __shared__ __align__(128) unsigned char rj_stash_raw[];
  JT* stash = reinterpret_cast<JT*>(rj_stash_raw);
  constexpr int TILE = RJ_TILE_BYTES / (8 * (int)sizeof(JT));
  constexpr int NW = RJ_THREADS / 32;
  __shared__ CamRuns runs;
  __shared__ double T[TAB_STRIDE];
  __shared__ double sm[NW * 72];
  const int c = blockIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == RJ_THREADS - 1) build_cam_table(cam8 + 8 * (size_t)c, k3 + 3 * (size_t)c, T);
  const int L = runs.load(c, nPB, camtab);  // includes the __syncthreads that publishes T
  double v[1] = {0.0};
  double g8[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // this thread's share of g_c
  double u0 = 0.0, u1 = 0.0;                // this lane's two entries of the warp's 8x8 partial of U_c (DMMA accumulator)
  int run = 0;
  for (int base = 0; base < L; base += (WITH_JAC ? TILE : L)) {
    const int hi = WITH_JAC ? min(L, base + TILE) : L;
    const int hi4 = WITH_JAC ? min(base + TILE, (hi + 3) & ~3) : hi;  // the tensor-core phase eats rows in groups of 4
    for (int l = base + threadIdx.x; l < hi4; l += RJ_THREADS) {
      int p = -1;
      size_t j = 0;
      if (l < hi) { j = runs.locate(l, run); p = __ldg(pt_B + j); }
      if (p < 0) {  // padding row of a run / of the last group of 4
        if (WITH_JAC) {
          const double z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
          strow_cvt<JT>(stash + 8 * (size_t)(l - base), z);
        }
        continue;
      }
      double X4[4];
      ld256(pts4 + 4 * (size_t)p, X4);  // one 32 B record {X, Y, Z, -}
      const double X = X4[0], Y = X4[1], Z = X4[2];
      const double2 o = __ldg(obs_B + j);
      const Proj pr = project(T, X, Y, Z);
      double ex, ey, r;
      if (MODEL >= 2) {
        double ct;
        r = robust_residual<MODEL>(pr, o.x, o.y, scale, rho_b, (int)(j & 1), ex, ey, ct);
        v[0] += ct;
      } else {
        r = row_residual<MODEL>(pr, o.x, o.y, scale, ex, ey);
        v[0] += r * r;
      }
      if (WITH_JAC) {
        double a[8], b[3];
        jac_row(T, pr, X, Y, Z, ex, ey, a, b);
        store8<JT>(Jc_B + 8 * j, a);
        st_rec(rec_A + __ldg(posA + j), b, r);
        strow_cvt<JT>(stash + 8 * (size_t)(l - base), a);
        // gradient: from the row as STORED (g = J~^T r: the step is the Gauss-Newton step of ONE perturbed least-squares
        // problem, error ~ cond(J) eps — measured 7e-8 trajectory deviation at Venice) or from the unrounded row
        // (exact_grad: (J~^T J~ + mu)^-1 J^T r, error ~ cond(J)^2 eps in the weak directions — measured 4e-6, and 7.5e-5 on
        // the `small` scene); the default is the stored row
#pragma unroll
        for (int k = 0; k < 8; ++k) g8[k] += (exact_grad ? a[k] : (double)(JT)a[k]) * r;
      }
    }
    if (WITH_JAC) {
      __syncthreads();
      // U_c += sum over the tile of a a^T on the FP64 tensor core: for a group of 4 stashed rows, lane (g, t) holds
      // a_{row t}[g] — as the A fragment (8 components x 4 rows) AND as the B fragment (4 rows x 8 components).  One shared
      // load, one conversion and one DMMA per lane per 4 rows; warp w takes groups w, w + NW, ... (fixed order)
      const int ngroups = (hi4 - base) >> 2;
      const int g = lane >> 2, t = lane & 3;
      for (int q = wid; q < ngroups; q += NW) {
        const double av = stash_get(stash + 8 * (size_t)(4 * q + t) + g);
        dmma_m8n8k4(u0, u1, av, av);
      }
      __syncthreads();
    }
  }
  block_sum<1>(v, sm);
  if (threadIdx.x == 0) cost_part[c] = v[0];
  if (!WITH_JAC) return;
  __syncthreads();
  // ordered reduction: per warp 64 entries of U (from the accumulator fragments) and 8 of g (butterfly), then a fixed-order
  // sum over the warps
  {
    const int g = lane >> 2, t = lane & 3;
    sm[wid * 72 + 8 * g + 2 * t] = u0;
    sm[wid * 72 + 8 * g + 2 * t + 1] = u1;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const double s = warp_sum(g8[k]);
      if (lane == k) sm[wid * 72 + 64 + k] = s;
    }
  }
  __syncthreads();
  if (threadIdx.x < 72) {
    double s = 0.0;
    for (int w = 0; w < NW; ++w) s += sm[w * 72 + threadIdx.x];  // fixed order
    if (threadIdx.x < 64) U[64 * (size_t)c + threadIdx.x] = s;
    else gc[8 * (size_t)c + (threadIdx.x - 64)] = s;
  }
}

// ---------------------------------------------------------------- reprojection-error evaluator
// Camera.compute_projection_errors_alt (src/dataset/camera.py:78-85) for every camera at once: per observation
// e = loss(dx^2) + loss(dy^2) in pixels^2, loss = identity (LOSS 0, TRIVIAL_LOSS) or log(1 + x) (LOSS 1, CAUCHY_LOSS,
// src/dataset/loss_functions.py:7-17).  LOSS 2 / 3 = what the reference's COLMAP comparator minimises (cost
// side only): rho(|r|^2) of the 2-vector residual r = projection - observation (colmap/src/base/cost_functions.h:44-86) with
// ceres::CauchyLoss(a) = b log(1 + s/b) or ceres::SoftLOneLoss(a) = 2b (sqrt(1 + s/b) - 1), b = a^2
// (colmap/src/optim/bundle_adjustment.cc:53-68); Ceres' cost is half the sum.  Same traversal as the trial-cost kernel (one CTA per camera over its B runs,
// camera table in shared memory, point gather inside the block window); e goes to the CALLER's row
// (order_B[j], an 8 B scatter) and the ordered per-camera sum to cam_sum[c].  24 B read + 8 B written per row.
template <int LOSS>
__global__ void __launch_bounds__(RJ_THREADS)
k_reproj_err(int nPB, const int32_t* __restrict__ camtab, const double* __restrict__ cam8, const double* __restrict__ k3,
             const int32_t* __restrict__ pt_B, const double2* __restrict__ obs_B,
             const int32_t* __restrict__ order_B, const double* __restrict__ pts4, double b_scale, double* __restrict__ err_rows,
             double* __restrict__ cam_sum) {
  __shared__ CamRuns runs;
  __shared__ double T[TAB_STRIDE];
  __shared__ double sm[32];
  const int c = blockIdx.x;
  if (threadIdx.x == RJ_THREADS - 1) build_cam_table(cam8 + 8 * (size_t)c, k3 + 3 * (size_t)c, T);
  const int L = runs.load(c, nPB, camtab);  // includes the __syncthreads that publishes T
  double v[1] = {0.0};
  int run = 0;
  for (int l = threadIdx.x; l < L; l += RJ_THREADS) {
    const size_t j = runs.locate(l, run);
    const int p = __ldg(pt_B + j);
    if (p < 0) continue;  // padding row of a run
    double X4[4];
    ld256(pts4 + 4 * (size_t)p, X4);  // one 32 B record {X, Y, Z, -}: ONE divergent request per row instead of three
    const double X = X4[0], Y = X4[1], Z = X4[2];
    const double2 o = __ldg(obs_B + j);
    const Proj pr = project(T, X, Y, Z);
    const double dx = pr.u - o.x, dy = pr.v - o.y;
    const double e = reproj_loss<LOSS>(dx, dy, b_scale);
    v[0] += e;
    if (err_rows) err_rows[__ldg(order_B + j)] = e;  // order_B: caller's row of every B position (built once per plan)
  }
  block_sum<1>(v, sm);
  if (threadIdx.x == 0 && cam_sum) cam_sum[c] = v[0];
}

// ---------------------------------------------------------------- K2a: 3x3 point blocks
// One thread per point walks its (short) A run of 32 B row records.  V_p = sum b b^T (6 unique),
// g_p = sum b r; also writes the point part of the rows as SoA arrays (what the PCG point kernel streams).
// part[2*block+0] = sum |g_p|^2, part[2*block+1] = max diag V.
template <typename JT>
__global__ void __launch_bounds__(256)
k_point_blocks(int P, size_t N, const int32_t* __restrict__ pt_ptr, const RowRec* __restrict__ rec_A,
               JT* __restrict__ Jp_ps, double* __restrict__ V, double* __restrict__ gp, double* __restrict__ part, int exact_grad) {
  __shared__ double sm[32];
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  double g2[1] = {0.0};
  double md = 0.0;
  if (p < P) {
    double v[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
    const int e = pt_ptr[p + 1];
    for (int i0 = pt_ptr[p]; i0 < e; i0 += 4) {  // 4 records in flight per thread, accumulated in row order
      double b[4][3], r[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        b[u][0] = b[u][1] = b[u][2] = 0.0; r[u] = 0.0;
        if (i0 + u < e) ld_rec(rec_A + i0 + u, b[u], r[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (i0 + u >= e) break;
#pragma unroll
        for (int k = 0; k < 3; ++k) jput<JT>(Jp_ps, N, k, (size_t)(i0 + u), b[u][k]);
        // V_p from the row as STORED (rounded to JT): the operator is exactly J~^T J~; g_p from the unrounded row
        const double q0 = (double)(JT)b[u][0], q1 = (double)(JT)b[u][1], q2 = (double)(JT)b[u][2];
        v[0] += q0 * q0; v[1] += q0 * q1; v[2] += q0 * q2;
        v[3] += q1 * q1; v[4] += q1 * q2; v[5] += q2 * q2;
        if (exact_grad) { g[0] += b[u][0] * r[u]; g[1] += b[u][1] * r[u]; g[2] += b[u][2] * r[u]; }
        else { g[0] += q0 * r[u]; g[1] += q1 * r[u]; g[2] += q2 * r[u]; }
      }
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) V[6 * (size_t)p + k] = v[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) gp[3 * (size_t)p + k] = g[k];
    g2[0] = g[0] * g[0] + g[1] * g[1] + g[2] * g[2];
    md = fmax(v[0], fmax(v[3], v[5]));
  }
  block_sum<1>(g2, sm);
  const double m = block_max1(md, sm);
  if (threadIdx.x == 0) {
    part[2 * blockIdx.x] = g2[0];
    part[2 * blockIdx.x + 1] = m;
  }
}

// ---------------------------------------------------------------- K2b: 8x8 camera blocks
// One step of a transpose-reduce: N values per lane -> N/2, the lanes with bit `off` set keep the upper half.
template <int N>
__device__ __forceinline__ void tr_fold(double* v, int lane, int off) {
  const bool up = lane & off;
#pragma unroll
  for (int j = 0; j < N / 2; ++j) {
    const double send = up ? v[j] : v[j + N / 2], keepv = up ? v[j + N / 2] : v[j];
    v[j] = keepv + __shfl_xor_sync(0xffffffffu, send, off);
  }
}

// One CTA per camera (64..256 threads, by the average camera size) streams its B-order runs.  If wb_B != nullptr the rows are weighted by its {w_i, beta_i} records:
// out = -sum beta_i a a^T and -sum a w_i (the Schur corrections of the block-Jacobi blocks / reduced rhs);
// otherwise U_c = sum a a^T and g_c = sum a r.
// (An fp64 tensor-core version — one mma.sync.m8n8k4.f64 per 4 rows, the SYRK mapped 1:1 onto the DMMA tile — was
// measured at ~2.4 TFLOP/s on B200 and ran 2.3x SLOWER than this FMA kernel, see profiles/ncu_r1_v5_summary.md; a
// variant with 4 lanes per row and 36 accumulators per lane was slower too: 146 us vs 114 us.)
template <typename JT>
__global__ void __launch_bounds__(256)
k_cam_blocks(int nPB, const int32_t* __restrict__ camtab, const JT* __restrict__ Jc_B,
             const double* __restrict__ s_B, const double2* __restrict__ wb_B,
             double* __restrict__ U, double* __restrict__ gc) {
  __shared__ double sm[44 * 8];
  __shared__ CamRuns runs;
  const int c = blockIdx.x;
  const int L = runs.load(c, nPB, camtab);
  double acc[44];
#pragma unroll
  for (int k = 0; k < 44; ++k) acc[k] = 0.0;
  int rr = 0;
  for (int l = threadIdx.x; l < L; l += blockDim.x) {
    const size_t j = runs.locate(l, rr);
    double a[8];
    load8<JT>(Jc_B + 8 * j, a);
    double s, wgt;
    if (wb_B) {
      const double2 wb = __ldg(wb_B + j);
      s = -wb.x;
      wgt = -wb.y;
    } else {
      s = __ldg(s_B + j);
      wgt = 1.0;
    }
    int t = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const double ak = a[k] * wgt;
#pragma unroll
      for (int l2 = k; l2 < 8; ++l2) acc[t++] += ak * a[l2];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[36 + k] += a[k] * s;
  }
  // transpose-reduce inside the warp (47 shuffles instead of 44 x 5): afterwards lane i holds the warp total of
  // value i (i < 32) and lanes 2j, 2j+1 both hold value 32 + j (j < 12)
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  double hi[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) hi[k] = k < 12 ? acc[32 + k] : 0.0;
  tr_fold<32>(acc, lane, 16); tr_fold<16>(acc, lane, 8); tr_fold<8>(acc, lane, 4); tr_fold<4>(acc, lane, 2); tr_fold<2>(acc, lane, 1);
  tr_fold<16>(hi, lane, 16); tr_fold<8>(hi, lane, 8); tr_fold<4>(hi, lane, 4); tr_fold<2>(hi, lane, 2);
  hi[0] += __shfl_xor_sync(0xffffffffu, hi[0], 1);
  sm[wid * 44 + lane] = acc[0];
  if (!(lane & 1) && lane < 24) sm[wid * 44 + 32 + (lane >> 1)] = hi[0];
  __syncthreads();
  if (threadIdx.x < 44) {
    double v = 0.0;
    for (int w = 0; w < nw; ++w) v += sm[w * 44 + threadIdx.x];  // fixed order
    const int t = threadIdx.x;
    if (t >= 36) {
      gc[8 * (size_t)c + (t - 36)] = v;
    } else {
      int k = 0, rem = t;
      while (rem >= 8 - k) { rem -= 8 - k; ++k; }
      const int l = k + rem;
      U[64 * (size_t)c + 8 * k + l] = v;
      U[64 * (size_t)c + 8 * l + k] = v;
    }
  }
}

// per-camera scalars after the (all-reduced) blocks: part[2c] = |g_c|^2, part[2c+1] = max diag U_c
__global__ void k_cam_scalars(int C, const double* __restrict__ U, const double* __restrict__ gc,
                              double* __restrict__ part) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double g2 = 0.0, md = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const double g = gc[8 * (size_t)c + k];
    g2 += g * g;
    md = fmax(md, U[64 * (size_t)c + 9 * k]);
  }
  part[2 * c] = g2;
  part[2 * c + 1] = md;
}

__device__ __forceinline__ void inv3_sym(const double* v, double mu, double* o) {
  const double a = v[0] + mu, b = v[1], c = v[2], d = v[3] + mu, e = v[4], f = v[5] + mu;
  const double c00 = d * f - e * e, c01 = c * e - b * f, c02 = b * e - c * d;
  const double det = a * c00 + b * c01 + c * c02;
  const double id = 1.0 / det;
  o[0] = c00 * id; o[1] = c01 * id; o[2] = c02 * id;
  o[3] = (a * f - c * c) * id; o[4] = (b * c - a * e) * id; o[5] = (a * d - b * b) * id;
}

// ---------------------------------------------------------------- K3a: point-block inverses
// Vinv_p = (V_p + mu I)^-1; per row w_i = b_i . (Vinv g_p) and beta_i = b_i^T Vinv b_i, scattered to the
// row's B position (the camera-side Schur corrections stream them).
template <typename JT>
__global__ void __launch_bounds__(256)
k_point_prepare(int P, size_t N, const double* __restrict__ mu_ptr, const int32_t* __restrict__ pt_ptr,
                const int32_t* __restrict__ posB, const JT* __restrict__ Jp_ps, const double* __restrict__ V,
                const double* __restrict__ gp, double* __restrict__ Vinv, double2* __restrict__ wb_B) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const double mu = *mu_ptr;
  double v[6], vi[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) v[k] = V[6 * (size_t)p + k];
  inv3_sym(v, mu, vi);
#pragma unroll
  for (int k = 0; k < 6; ++k) Vinv[(size_t)k * P + p] = vi[k];  // SoA: the point-side kernels read it one lane per point
  const double g0 = gp[3 * (size_t)p], g1 = gp[3 * (size_t)p + 1], g2 = gp[3 * (size_t)p + 2];
  const double z0 = vi[0] * g0 + vi[1] * g1 + vi[2] * g2;
  const double z1 = vi[1] * g0 + vi[3] * g1 + vi[4] * g2;
  const double z2 = vi[2] * g0 + vi[4] * g1 + vi[5] * g2;
  const int e = pt_ptr[p + 1];
  for (int i0 = pt_ptr[p]; i0 < e; i0 += 4) {  // 4 rows in flight per thread
    double b[4][3];
    int j[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      j[u] = -1;
      if (i0 + u < e) {
        j[u] = __ldg(posB + i0 + u);
#pragma unroll
        for (int k = 0; k < 3; ++k) b[u][k] = jget<JT>(Jp_ps, N, k, (size_t)(i0 + u));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (j[u] < 0) continue;
      const double wv = b[u][0] * z0 + b[u][1] * z1 + b[u][2] * z2;
      const double t0 = vi[0] * b[u][0] + vi[1] * b[u][1] + vi[2] * b[u][2];
      const double t1 = vi[1] * b[u][0] + vi[3] * b[u][1] + vi[4] * b[u][2];
      const double t2 = vi[2] * b[u][0] + vi[4] * b[u][1] + vi[5] * b[u][2];
      wb_B[j[u]] = make_double2(wv, b[u][0] * t0 + b[u][1] * t1 + b[u][2] * t2);  // ONE 16 B scatter per row
    }
  }
}

// rhs = g_c + rhs_part; Minv_c = (U_c + mu I + Sd_part)^-1 (8x8 Cholesky; 8 threads per camera, thread j factors
// the block and solves for column j of the inverse — the arithmetic per column is what one thread per camera did,
// 8x shorter); PCG start: x = 0, r = rhs, u = Minv r, all other vectors 0; part[4c+2] = r.r
#define CF_T 256
__global__ void __launch_bounds__(CF_T)
k_cam_finish_prepare(int C, const double* __restrict__ mu_ptr, const double* __restrict__ U,
                     const double* __restrict__ gc, const double* __restrict__ Sd_part,
                     const double* __restrict__ rhs_part, double* __restrict__ Minv,
                     double* __restrict__ rhs, double* __restrict__ cv, double* __restrict__ part) {
  __shared__ double sMi[CF_T / 8][64];
  __shared__ double szz[CF_T / 8][8];
  const int e0 = blockIdx.x * CF_T + threadIdx.x;
  const int c = e0 >> 3, col = e0 & 7, g = threadIdx.x >> 3;
  const bool act = c < C;
  const size_t cc = act ? c : C - 1;
  const double mu = *mu_ptr;
  double L[64], e[8];
#pragma unroll
  for (int k = 0; k < 64; ++k) L[k] = U[64 * cc + k] + Sd_part[64 * cc + k];
#pragma unroll
  for (int k = 0; k < 8; ++k) L[9 * k] += mu;
#pragma unroll
  for (int k = 0; k < 8; ++k) e[k] = (k == col) ? 1.0 : 0.0;
  chol8_solve(L, e);
#pragma unroll
  for (int k = 0; k < 8; ++k) sMi[g][8 * k + col] = e[k];
  __syncwarp();
  // thread col -> row col of the symmetrised inverse
  double row[8];
#pragma unroll
  for (int l = 0; l < 8; ++l) row[l] = 0.5 * (sMi[g][8 * col + l] + sMi[g][8 * l + col]);
  double rr[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) rr[k] = gc[8 * cc + k] + rhs_part[8 * cc + k];
  double zk = 0.0;
#pragma unroll
  for (int l = 0; l < 8; ++l) zk += row[l] * rr[l];
  szz[g][col] = zk;
  __syncwarp();
  if (!act) return;
#pragma unroll
  for (int l = 0; l < 8; ++l) Minv[64 * cc + 8 * col + l] = row[l];
  double* v = cv + (size_t)72 * cc;  // camera-major PCG record: x r u w z q s p m
  rhs[8 * cc + col] = gc[8 * cc + col] + rhs_part[8 * cc + col];
  v[col] = 0.0; v[8 + col] = gc[8 * cc + col] + rhs_part[8 * cc + col]; v[16 + col] = zk;
  v[24 + col] = 0.0; v[32 + col] = 0.0; v[40 + col] = 0.0; v[48 + col] = 0.0; v[56 + col] = 0.0; v[64 + col] = 0.0;
  if (col == 0) {
    double rz = 0.0, r2 = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { rz += szz[g][k] * rr[k]; r2 += rr[k] * rr[k]; }
    part[4 * cc] = rz;
    part[4 * cc + 1] = 0.0;
    part[4 * cc + 2] = r2;
  }
}
