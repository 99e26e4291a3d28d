/* CPU oracle (plain C, OpenMP optional) for the LogModel hot path -- TEST INFRASTRUCTURE ONLY.
 *
 * Matrix-free restatement of the reference solver for sizes where the NumPy/SciPy oracle is too
 * slow (n = 2,000 ... 20,000).  Same iteration as graph_learn/smooth_learning.py:125-169 of
 * LTS4/graph-learning, scheduled as ONE sweep over the packed edge vector per iteration
 * (oracle/logmodel_oracle.py::solve_one_pass is the NumPy twin and is pinned to the reference's
 * golden vectors; tests/test_oracle_c.py pins this file to that).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this library.
 * Build: make -C oracle   (gcc -O3 -fopenmp -shared -fPIC)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int orc_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* packed offset of row i: edges (i, i+1..n-1) start here (operators.py:39-41 order) */
static inline int64_t row_base(int64_t i, int64_t n) { return i * (n - 1) - i * (i - 1) / 2; }

/* Squared Euclidean distances between the COLUMNS of x (m rows, n columns, row-major), packed.
 * smooth_learning.py:278 (pdist(x.T)**2). */
void orc_pdist_sq(const double *x, int64_t m, int64_t n, double *out) {
#pragma omp parallel for schedule(dynamic, 4)
  for (int64_t i = 0; i < n; ++i) {
    double *o = out + row_base(i, n) - (i + 1);
    for (int64_t j = i + 1; j < n; ++j) o[j] = 0.0;
    for (int64_t t = 0; t < m; ++t) {
      const double *r = x + t * n;
      const double xi = r[i];
      for (int64_t j = i + 1; j < n; ++j) {
        const double df = xi - r[j];
        o[j] += df * df;
      }
    }
  }
}

/* S w : weighted degrees (smooth_learning.py:13-44 applied to a vector). */
void orc_sum_op(const double *w, int64_t n, double *deg) {
  memset(deg, 0, sizeof(double) * (size_t)n);
  for (int64_t i = 0; i < n; ++i) {
    const double *r = w + row_base(i, n) - (i + 1);
    double acc = 0.0;
    for (int64_t j = i + 1; j < n; ++j) {
      acc += r[j];
      deg[j] += r[j];
    }
    deg[i] += acc;
  }
}

/* S^T d : d_i + d_j per edge. */
void orc_sum_op_adj(const double *d, int64_t n, double *edge) {
  for (int64_t i = 0; i < n; ++i) {
    double *r = edge + row_base(i, n) - (i + 1);
    for (int64_t j = i + 1; j < n; ++j) r[j] = d[i] + d[j];
  }
}

/* One-sweep FBF primal-dual solver, single graph, float64.
 * z: packed (theta-scaled) squared distances, E = n(n-1)/2.  w: in = initial point (or zeros),
 * out = result.  Returns the number of iterations performed. */
int orc_logmodel_solve(const double *z, int64_t n, double alpha, double beta, double *w, int maxit,
                       double tol, double step_size, double edge_tol, int threshold,
                       double *rel_primal_out, double *rel_dual_out) {
  const int64_t n_edges = n * (n - 1) / 2;
  const double g = step_size / (2.0 * beta + sqrt(2.0 * (double)(n - 1))); /* :135-137 */
  const double c = 1.0 - 2.0 * g * beta;
  const int nt = orc_threads();
  double *d = (double *)calloc((size_t)n, sizeof(double));
  double *sw = (double *)calloc((size_t)n, sizeof(double));
  double *sp = (double *)calloc((size_t)n, sizeof(double));
  double *yb = (double *)calloc((size_t)n, sizeof(double));
  double *pb = (double *)calloc((size_t)n, sizeof(double));
  double *col = (double *)calloc((size_t)n * 2 * (size_t)nt, sizeof(double));
  orc_sum_op(w, n, d); /* :144 */
  memcpy(sw, d, sizeof(double) * (size_t)n);
  double rp = INFINITY, rd = INFINITY;
  int it = 0;
  for (it = 1; it <= maxit; ++it) {
    for (int64_t v = 0; v < n; ++v) {
      yb[v] = d[v] + g * sw[v];                                        /* :148 */
      pb[v] = (yb[v] - sqrt(yb[v] * yb[v] + 4.0 * alpha * g)) / 2.0;   /* :151 */
    }
    memset(col, 0, sizeof(double) * (size_t)n * 2 * (size_t)nt);
    memset(sp, 0, sizeof(double) * (size_t)n);
    memset(sw, 0, sizeof(double) * (size_t)n);
    double res = 0.0, wn = 0.0;
#pragma omp parallel reduction(+ : res, wn)
    {
#ifdef _OPENMP
      const int tid = omp_get_thread_num();
#else
      const int tid = 0;
#endif
      double *cp = col + (size_t)tid * 2 * (size_t)n, *cw = cp + n;
#pragma omp for schedule(dynamic, 8)
      for (int64_t i = 0; i < n; ++i) {
        const int64_t off = row_base(i, n) - (i + 1);
        double *wr = w + off;
        const double *zr = z + off;
        double rsp = 0.0, rsw = 0.0;
        for (int64_t j = i + 1; j < n; ++j) {
          const double wo = wr[j];
          const double y = c * wo - g * (d[i] + d[j]);     /* :147 */
          double p = y - 2.0 * g * zr[j];                  /* :150 */
          p = p > 0.0 ? p : 0.0;
          const double q = c * p - g * (pb[i] + pb[j]);    /* :153 */
          const double r = q - y;
          const double wnew = wo - y + q;                  /* :163 */
          res += r * r;
          wn += wo * wo;
          rsp += p;
          rsw += wnew;
          cp[j] += p;
          cw[j] += wnew;
          wr[j] = wnew;
        }
        cp[i] += rsp;
        cw[i] += rsw;
      }
    }
    for (int t = 0; t < nt; ++t)
      for (int64_t v = 0; v < n; ++v) {
        sp[v] += col[(size_t)t * 2 * n + v];
        sw[v] += col[(size_t)t * 2 * n + n + v];
      }
    double resd = 0.0, dn = 0.0;
    for (int64_t v = 0; v < n; ++v) {
      const double qb = pb[v] + g * sp[v];                 /* :154 */
      const double rb = qb - yb[v];
      resd += rb * rb;
      dn += d[v] * d[v];
      d[v] = d[v] - yb[v] + qb;                            /* :164 */
    }
    if (it > 1) {                                          /* :156-161 */
      rp = sqrt(res) / sqrt(wn);
      rd = sqrt(resd) / sqrt(dn);
    }
    if (rp < tol && rd < tol) break;                       /* :166 */
  }
  if (it > maxit) it = maxit;
  if (threshold)
    for (int64_t e = 0; e < n_edges; ++e)
      if (w[e] < edge_tol) w[e] = 0.0;                     /* :169 */
  if (rel_primal_out) *rel_primal_out = rp;
  if (rel_dual_out) *rel_dual_out = rd;
  free(d); free(sw); free(sp); free(yb); free(pb); free(col);
  return it;
}
