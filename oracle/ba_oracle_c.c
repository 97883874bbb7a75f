/* ORACLE (test infrastructure, NOT product code) — plain C (OpenMP) restatement of the reference's
 * bundle-adjustment path in ITS OWN formulation: residual + Jacobian rows, then jaxopt 0.7's
 * LevenbergMarquardt with the default inner solver (unpreconditioned CG on the full (8C+3P) normal
 * equations, rel. tol 1e-5, maxiter 10n, x0 = 0).  PARITY: residual side pinned through ba_oracle.py, LM state
 * machine UNPINNED (see ba_oracle.py header); this
 * file is checked against ba_oracle.py in tests/test_oracle_cpu.py and exists so that the CPU
 * baseline uses every host core.
 *
 * Follows (relative to /root/reference/src/reconstruction/bundle_adjustment/):
 *   rot_mat_from_vec utils.py:21-28; parse_intrinsics utils.py:47-57; reproject_point
 *   bundle_adjustment.py:17-21; get_residuals :52-74; parameter layout prepare_params :93-102;
 *   LM state machine: jaxopt LevenbergMarquardt.init_state/update/run;
 *   CG: jax.scipy.sparse.linalg.cg as called by jaxopt.linear_solve.solve_cg.
 * Build: make -C oracle   (gcc -O3 -fopenmp -shared -fPIC)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#define PAR_MIN 100000 /* below this, OpenMP fork/join costs more than the loop */

typedef struct {
  int C, P;
  int64_t N;
  const int32_t *cam, *pt;
  const double *obs, *k3;
  double w2;
  /* row lists per point / per camera (counting sort) so J^T products need no atomics */
  int64_t *pt_ptr, *cam_ptr;
  int64_t *pt_rows, *cam_rows;
  double *a, *b, *r; /* N x 8, N x 3, N */
} Prob;

int ba_c_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void ba_c_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

static void build_lists(Prob* q) {
  q->pt_ptr = (int64_t*)calloc((size_t)q->P + 1, sizeof(int64_t));
  q->cam_ptr = (int64_t*)calloc((size_t)q->C + 1, sizeof(int64_t));
  q->pt_rows = (int64_t*)malloc((size_t)q->N * sizeof(int64_t));
  q->cam_rows = (int64_t*)malloc((size_t)q->N * sizeof(int64_t));
  for (int64_t i = 0; i < q->N; ++i) { q->pt_ptr[q->pt[i] + 1]++; q->cam_ptr[q->cam[i] + 1]++; }
  for (int p = 0; p < q->P; ++p) q->pt_ptr[p + 1] += q->pt_ptr[p];
  for (int c = 0; c < q->C; ++c) q->cam_ptr[c + 1] += q->cam_ptr[c];
  int64_t* fp = (int64_t*)malloc((size_t)q->P * sizeof(int64_t));
  int64_t* fc = (int64_t*)malloc((size_t)q->C * sizeof(int64_t));
  memcpy(fp, q->pt_ptr, (size_t)q->P * sizeof(int64_t));
  memcpy(fc, q->cam_ptr, (size_t)q->C * sizeof(int64_t));
  for (int64_t i = 0; i < q->N; ++i) { q->pt_rows[fp[q->pt[i]]++] = i; q->cam_rows[fc[q->cam[i]]++] = i; }
  free(fp); free(fc);
}

static void rodrigues(const double* w, double* R) {
  const double th = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  const double k0 = w[0] / th, k1 = w[1] / th, k2 = w[2] / th, c = cos(th), s = sin(th), o = 1.0 - c;
  R[0] = c + o * k0 * k0; R[1] = o * k0 * k1 - s * k2; R[2] = o * k0 * k2 + s * k1;
  R[3] = o * k1 * k0 + s * k2; R[4] = c + o * k1 * k1; R[5] = o * k1 * k2 - s * k0;
  R[6] = o * k2 * k0 - s * k1; R[7] = o * k2 * k1 + s * k0; R[8] = c + o * k2 * k2;
}

/* x in the reference layout [C*6 | C*2 | P*3]; returns 0.5*sum r^2 */
static double eval_rows(Prob* q, const double* x, int want_jac) {
  const int C = q->C;
  double* tab = (double*)malloc((size_t)C * 18 * sizeof(double)); /* R(9) M(9) */
#pragma omp parallel for schedule(static) if (q->N > PAR_MIN)
  for (int c = 0; c < C; ++c) {
    const double* w = x + 6 * (size_t)c;
    double* R = tab + 18 * (size_t)c;
    double* M = R + 9;
    rodrigues(w, R);
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    const double A[9] = {R[0] - 1, R[3], R[6], R[1], R[4] - 1, R[7], R[2], R[5], R[8] - 1};
    const double Wx[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        double m = w[i] * w[j];
        for (int k = 0; k < 3; ++k) m += A[3 * i + k] * Wx[3 * k + j];
        M[3 * i + j] = m / th2;
      }
  }
  double cost = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : cost) if (q->N > PAR_MIN)
  for (int64_t i = 0; i < q->N; ++i) {
    const int c = q->cam[i], p = q->pt[i];
    const double *R = tab + 18 * (size_t)c, *M = R + 9, *t = x + 6 * (size_t)c + 3;
    const double fx = x[6 * (size_t)C + 2 * (size_t)c], fy = x[6 * (size_t)C + 2 * (size_t)c + 1];
    const double cx = q->k3[3 * (size_t)c], cy = q->k3[3 * (size_t)c + 1], sk = q->k3[3 * (size_t)c + 2];
    const double* X = x + 8 * (size_t)C + 3 * (size_t)p;
    const double q0 = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + t[0];
    const double q1 = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + t[1];
    const double q2 = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + t[2];
    const double iz = 1.0 / q2;
    const double dx = (fx * q0 + sk * q1) * iz + cx - q->obs[2 * i];
    const double dy = fy * q1 * iz + cy - q->obs[2 * i + 1];
    const double r = (dx * dx + dy * dy) / q->w2;
    q->r[i] = r;
    cost += r * r;
    if (!want_jac) continue;
    const double ex = 2.0 * dx / q->w2, ey = 2.0 * dy / q->w2;
    const double e0 = ex * fx * iz, e1 = (ex * sk + ey * fy) * iz;
    const double e2 = -(ex * (fx * q0 + sk * q1) + ey * fy * q1) * iz * iz;
    double* a = q->a + 8 * i;
    double* b = q->b + 3 * i;
    b[0] = R[0] * e0 + R[3] * e1 + R[6] * e2;
    b[1] = R[1] * e0 + R[4] * e1 + R[7] * e2;
    b[2] = R[2] * e0 + R[5] * e1 + R[8] * e2;
    const double c0 = X[1] * b[2] - X[2] * b[1], c1 = X[2] * b[0] - X[0] * b[2], c2 = X[0] * b[1] - X[1] * b[0];
    a[0] = c0 * M[0] + c1 * M[3] + c2 * M[6];
    a[1] = c0 * M[1] + c1 * M[4] + c2 * M[7];
    a[2] = c0 * M[2] + c1 * M[5] + c2 * M[8];
    a[3] = e0; a[4] = e1; a[5] = e2;
    a[6] = ex * q0 * iz;
    a[7] = ey * q1 * iz;
  }
  free(tab);
  return 0.5 * cost;
}

/* column index of camera entry k of camera c in the reference layout */
static inline size_t ccol(int C, int c, int k) { return k < 6 ? 6 * (size_t)c + k : 6 * (size_t)C + 2 * (size_t)c + (k - 6); }

/* y = J^T s  (s: N) */
static void jt_apply(const Prob* q, const double* s, double* y) {
  const int C = q->C;
#pragma omp parallel for schedule(dynamic, 8) if (q->N > PAR_MIN)
  for (int c = 0; c < C; ++c) {
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int64_t j = q->cam_ptr[c]; j < q->cam_ptr[c + 1]; ++j) {
      const int64_t i = q->cam_rows[j];
      for (int k = 0; k < 8; ++k) acc[k] += q->a[8 * i + k] * s[i];
    }
    for (int k = 0; k < 8; ++k) y[ccol(C, c, k)] = acc[k];
  }
#pragma omp parallel for schedule(static, 1024) if (q->N > PAR_MIN)
  for (int p = 0; p < q->P; ++p) {
    double acc[3] = {0, 0, 0};
    for (int64_t j = q->pt_ptr[p]; j < q->pt_ptr[p + 1]; ++j) {
      const int64_t i = q->pt_rows[j];
      for (int k = 0; k < 3; ++k) acc[k] += q->b[3 * i + k] * s[i];
    }
    for (int k = 0; k < 3; ++k) y[8 * (size_t)C + 3 * (size_t)p + k] = acc[k];
  }
}

/* s = J v */
static void j_apply(const Prob* q, const double* v, double* s) {
  const int C = q->C;
#pragma omp parallel for schedule(static) if (q->N > PAR_MIN)
  for (int64_t i = 0; i < q->N; ++i) {
    const int c = q->cam[i], p = q->pt[i];
    double t = 0.0;
    for (int k = 0; k < 8; ++k) t += q->a[8 * i + k] * v[ccol(C, c, k)];
    for (int k = 0; k < 3; ++k) t += q->b[3 * i + k] * v[8 * (size_t)C + 3 * (size_t)p + k];
    s[i] = t;
  }
}

static double dot(const double* a, const double* b, size_t n) {
  double s = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : s) if (n > PAR_MIN)
  for (size_t i = 0; i < n; ++i) s += a[i] * b[i];
  return s;
}

/* jax.scipy.sparse.linalg.cg: x0 = 0, M = I, stop r.r <= tol^2 b.b, maxiter. returns iterations */
static int cg_jaxopt(const Prob* q, const double* g, double mu, double tol, int64_t maxiter, double* x, double* ws, double* srow) {
  const size_t n = 8 * (size_t)q->C + 3 * (size_t)q->P;
  double *r = ws, *p = ws + n, *Ap = ws + 2 * n;
  memset(x, 0, n * sizeof(double));
  memcpy(r, g, n * sizeof(double));
  memcpy(p, g, n * sizeof(double));
  double gamma = dot(r, r, n);
  const double atol2 = tol * tol * gamma;
  int64_t k = 0;
  while (gamma > atol2 && k < maxiter) {
    j_apply(q, p, srow);
    jt_apply(q, srow, Ap);
#pragma omp parallel for schedule(static) if (n > PAR_MIN)
    for (size_t i = 0; i < n; ++i) Ap[i] += mu * p[i];
    const double alpha = gamma / dot(p, Ap, n);
#pragma omp parallel for schedule(static) if (n > PAR_MIN)
    for (size_t i = 0; i < n; ++i) { x[i] += alpha * p[i]; r[i] -= alpha * Ap[i]; }
    const double gn = dot(r, r, n);
    const double beta = gn / gamma;
#pragma omp parallel for schedule(static) if (n > PAR_MIN)
    for (size_t i = 0; i < n; ++i) p[i] = r[i] + beta * p[i];
    gamma = gn;
    ++k;
  }
  return (int)k;
}

/* residual + Jacobian rows in the caller's row order; returns 0.5 sum r^2 */
double ba_c_eval(int C, int P, int64_t N, const double* x, const int32_t* cam, const int32_t* pt, const double* obs,
                 const double* k3, double w2, double* r, double* a, double* b, int want_jac) {
  Prob q = {C, P, N, cam, pt, obs, k3, w2, 0, 0, 0, 0, a, b, r};
  return eval_rows(&q, x, want_jac);
}

/* Full LM (jaxopt semantics).  x: in/out.  Trajectories have capacity maxiter+1 / maxiter.
 * cg_cap > 0 truncates the inner CG (bounded benchmarking samples).  Returns iter_num. */
int ba_c_lm(int C, int P, int64_t N, double* x, const int32_t* cam, const int32_t* pt, const double* obs, const double* k3,
            double w2, double tol, int maxiter, double damping, double cg_tol, int64_t cg_cap, double* loss_traj,
            double* grad_traj, int* cg_iters, int* accepted) {
  const size_t n = 8 * (size_t)C + 3 * (size_t)P;
  Prob q = {C, P, N, cam, pt, obs, k3, w2, 0, 0, 0, 0, 0, 0, 0};
  q.a = (double*)malloc((size_t)N * 8 * sizeof(double));
  q.b = (double*)malloc((size_t)N * 3 * sizeof(double));
  q.r = (double*)malloc((size_t)N * sizeof(double));
  build_lists(&q);
  double* g = (double*)malloc(n * sizeof(double));
  double* v = (double*)malloc(n * sizeof(double));
  double* xn = (double*)malloc(n * sizeof(double));
  double* ws = (double*)malloc(3 * n * sizeof(double));
  double* srow = (double*)malloc((size_t)N * sizeof(double));
  double* rsave = (double*)malloc((size_t)N * sizeof(double));
  /* init_state */
  double loss = eval_rows(&q, x, 1);
  jt_apply(&q, q.r, g);
  double maxdiag = 0.0;
  { /* diag(J^T J) = column sums of squares */
    double* d = ws;
    memset(d, 0, n * sizeof(double));
    for (int64_t i = 0; i < N; ++i) {
      for (int k = 0; k < 8; ++k) d[ccol(C, cam[i], k)] += q.a[8 * i + k] * q.a[8 * i + k];
      for (int k = 0; k < 3; ++k) d[8 * (size_t)C + 3 * (size_t)pt[i] + k] += q.b[3 * i + k] * q.b[3 * i + k];
    }
    for (size_t i = 0; i < n; ++i) if (d[i] > maxdiag) maxdiag = d[i];
  }
  double mu = damping * maxdiag, nu = 2.0, err = sqrt(dot(g, g, n));
  int it = 0;
  if (loss_traj) loss_traj[0] = loss;
  if (grad_traj) grad_traj[0] = err;
  const int64_t cgmax = cg_cap > 0 ? cg_cap : 10 * (int64_t)n;
  while (maxiter > 0) {
    const int inner = cg_jaxopt(&q, g, mu, cg_tol, cgmax, v, ws, srow);
#pragma omp parallel for schedule(static) if (n > PAR_MIN)
    for (size_t i = 0; i < n; ++i) xn[i] = x[i] - v[i];
    memcpy(rsave, q.r, (size_t)N * sizeof(double));
    const double loss_next = eval_rows(&q, xn, 0);
    double vv = dot(v, v, n), vg = dot(v, g, n);
    const double rho = (loss - loss_next) / (0.5 * (mu * vv + vg));
    const int ok = rho > 0.0;
    if (ok) {
      memcpy(x, xn, n * sizeof(double));
      loss = eval_rows(&q, x, 1);
      jt_apply(&q, q.r, g);
      const double t = 2.0 * rho - 1.0;
      const double f = 1.0 - t * t * t;
      mu *= f > 1.0 / 3.0 ? f : 1.0 / 3.0;
      nu = 2.0;
    } else {
      memcpy(q.r, rsave, (size_t)N * sizeof(double));
      mu *= nu;
      nu *= 2.0;
    }
    err = sqrt(dot(g, g, n));
    if (cg_iters) cg_iters[it] = inner;
    if (accepted) accepted[it] = ok;
    ++it;
    if (loss_traj) loss_traj[it] = loss;
    if (grad_traj) grad_traj[it] = err;
    if (!(it < maxiter && err > tol)) break;
  }
  free(q.a); free(q.b); free(q.r); free(q.pt_ptr); free(q.cam_ptr); free(q.pt_rows); free(q.cam_rows);
  free(g); free(v); free(xn); free(ws); free(srow); free(rsave);
  return it;
}

/* ---------------------------------------------------------------------------------------------------------------
 * Exact inner solve at Venice scale: the reduced camera system of SURVEY.md App. A.6 formed DENSELY,
 *   S = blockdiag(U_c + mu I) - W (V + mu I)^-1 W^T   (8C x 8C, row-major, full symmetric; S must arrive zeroed)
 *   rhs = g_c - W (V + mu I)^-1 g_p,
 * from the Jacobian rows a (N x 8), b (N x 3) and residuals r of ba_c_eval (caller's row order).  Also returns
 * Vinv (P x 9, full 3x3) and g_p (P x 3) for the back-substitution.  The caller factors S with LAPACK (Cholesky) —
 * arithmetic-equivalent to a direct solve of (J^T J + mu I) v = g, the same mathematics as ba_oracle._solve_schur_exact
 * (which goes through scipy.sparse and cannot reach 5 M rows).  Camera-major 8-blocks: index 8c+k, k = [w, t, fx, fy].
 * Threads own the block rows c_i % T == tid, so no two threads ever write the same entry (no atomics, fixed order). */
int ba_c_schur_dense(int C, int P, int64_t N, const int32_t* cam, const int32_t* pt, const double* a, const double* b,
                     const double* r, double mu, double* S, double* rhs, double* Vinv, double* gp) {
  Prob q = {C, P, N, cam, pt, 0, 0, 0.0, 0, 0, 0, 0, (double*)a, (double*)b, (double*)r};
  build_lists(&q);
  const size_t n = 8 * (size_t)C;
  int bad = 0;
#pragma omp parallel for schedule(static, 1024) reduction(| : bad)
  for (int p = 0; p < P; ++p) {
    double V[9] = {mu, 0, 0, 0, mu, 0, 0, 0, mu}, g[3] = {0, 0, 0};
    for (int64_t j = q.pt_ptr[p]; j < q.pt_ptr[p + 1]; ++j) {
      const int64_t i = q.pt_rows[j];
      for (int k = 0; k < 3; ++k) {
        g[k] += b[3 * i + k] * r[i];
        for (int l = 0; l < 3; ++l) V[3 * k + l] += b[3 * i + k] * b[3 * i + l];
      }
    }
    const double c00 = V[4] * V[8] - V[5] * V[7], c01 = V[5] * V[6] - V[3] * V[8], c02 = V[3] * V[7] - V[4] * V[6];
    const double det = V[0] * c00 + V[1] * c01 + V[2] * c02;
    if (!(det > 0.0)) bad = 1;
    const double id = 1.0 / det;
    double* I = Vinv + 9 * (size_t)p;
    I[0] = c00 * id; I[1] = (V[2] * V[7] - V[1] * V[8]) * id; I[2] = (V[1] * V[5] - V[2] * V[4]) * id;
    I[3] = c01 * id; I[4] = (V[0] * V[8] - V[2] * V[6]) * id; I[5] = (V[2] * V[3] - V[0] * V[5]) * id;
    I[6] = c02 * id; I[7] = (V[1] * V[6] - V[0] * V[7]) * id; I[8] = (V[0] * V[4] - V[1] * V[3]) * id;
    for (int k = 0; k < 3; ++k) gp[3 * (size_t)p + k] = g[k];
  }
  memset(rhs, 0, n * sizeof(double));
#pragma omp parallel
  {
#ifdef _OPENMP
    const int T = omp_get_num_threads(), tid = omp_get_thread_num();
#else
    const int T = 1, tid = 0;
#endif
    for (int c = tid; c < C; c += T)
      for (int k = 0; k < 8; ++k) S[(8 * (size_t)c + k) * n + 8 * (size_t)c + k] += mu;
    for (int p = 0; p < P; ++p) {
      const double* I = Vinv + 9 * (size_t)p;
      const double* g = gp + 3 * (size_t)p;
      for (int64_t ji = q.pt_ptr[p]; ji < q.pt_ptr[p + 1]; ++ji) {
        const int64_t i = q.pt_rows[ji];
        const int ci = cam[i];
        if (ci % T != tid) continue;
        const double *ai = a + 8 * i, *bi = b + 3 * i;
        double* Srow = S + 8 * (size_t)ci * n;
        /* U_c and g_c */
        for (int k = 0; k < 8; ++k) {
          rhs[8 * (size_t)ci + k] += ai[k] * r[i];
          for (int l = 0; l < 8; ++l) Srow[(size_t)k * n + 8 * (size_t)ci + l] += ai[k] * ai[l];
        }
        const double t0 = I[0] * bi[0] + I[3] * bi[1] + I[6] * bi[2]; /* Vinv symmetric: t = Vinv b_i */
        const double t1 = I[1] * bi[0] + I[4] * bi[1] + I[7] * bi[2];
        const double t2 = I[2] * bi[0] + I[5] * bi[1] + I[8] * bi[2];
        const double wg = t0 * g[0] + t1 * g[1] + t2 * g[2];
        for (int k = 0; k < 8; ++k) rhs[8 * (size_t)ci + k] -= ai[k] * wg;
        for (int64_t jj = q.pt_ptr[p]; jj < q.pt_ptr[p + 1]; ++jj) {
          const int64_t j = q.pt_rows[jj];
          const double* aj = a + 8 * j;
          const double beta = t0 * b[3 * j] + t1 * b[3 * j + 1] + t2 * b[3 * j + 2];
          double* Sb = Srow + 8 * (size_t)cam[j];
          for (int k = 0; k < 8; ++k) {
            const double ab = ai[k] * beta;
            for (int l = 0; l < 8; ++l) Sb[(size_t)k * n + l] -= ab * aj[l];
          }
        }
      }
    }
  }
  free(q.pt_ptr); free(q.cam_ptr); free(q.pt_rows); free(q.cam_rows);
  return bad;
}
