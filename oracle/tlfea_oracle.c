/*
 * tlfea_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A plain-C FP64 restatement of the reference's (uwsbel/Total-Lagrangian-FEA)
 * hot-path algorithm, written from the behaviour of the CUDA device functions
 * and following their summation order.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library.
 * The product (libtlfea_b200.so) never links, loads or calls it.
 *
 * Parity pin (tests/test_oracle_golden.py; fixtures committed in tests/golden/
 * with the scripts that made them):
 *   FEAT10 precompute, mass, SVK and Kelvin-Voigt force + tangent: the reference's
 *     numpy prototypes test-scripts/T10-tets/f-form-T10-beam-newton.py and
 *     ...-newton-damped.py, imported in place (oracle/ref_proto.py);
 *   ANCF3243 / ANCF3443 B^-1, mass, SVK force + tangent: the prototypes
 *     test-scripts/3243-beam/f-form-3243-nesterov.py and
 *     test-scripts/3443-shell/f-form-3443-nesterov.py run in place up to their
 *     time loop (tests/golden/make_golden_ancf.py), plus the known-answer
 *     matrices data/utest/mass_matrix_{2,3}_beam.csv;
 *   Mooney-Rivlin P and element tangent: the strain energy stated in
 *     test-scripts/validation/fenics/beam_dynamic_backward_euler.py:332-345.
 * Parity unpinned: Kelvin-Voigt and Mooney-Rivlin on the ANCF elements (same
 * S/Q-generic code as the pinned FEAT10 cases, no reference fixture).
 * The reference's C++/CUDA cannot be built here (needs Eigen, cuDSS, Bazel, a
 * GPU) so there is no oracle/_ref binary.
 *
 * Layout conventions restated from the reference (SURVEY.md section 7,
 * "Exact reference semantics"):
 *   conn      column-major  conn[n*n_elem + e]           FEAT10Data.cuh:36-39
 *   grad_N    [(e*Q+q)*3S + j*S + a]   (a = shape fn, j = x/y/z)  :41-45
 *   F, P      [(e*Q+q)*9 + j*3 + i]    (column-major 3x3)        :115-130
 *   f_int     [3*coef + i]
 *   mass CSR  node/coefficient level, sorted unique columns
 *   H CSR     DOF level: row 3i+d owns 3*deg(i) entries, columns 3j..3j+2
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define T10_S 10
#define T10_Q 5

typedef struct {
  int model;        /* 0 = SVK, 1 = Mooney-Rivlin  (MaterialModel.cuh:14-17) */
  double lambda, mu;
  double mu10, mu01, kappa;
  double eta_damp, lambda_damp;
} orc_material;

/* ------------------------------------------------------------------ */
/* Keast 5-point rule, quadrature_utils.h:135-158                      */
/* ------------------------------------------------------------------ */
static const double T10_BARY[5][4] = {
    {0.25, 0.25, 0.25, 0.25},
    {0.5, 1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0},
    {1.0 / 6.0, 0.5, 1.0 / 6.0, 1.0 / 6.0},
    {1.0 / 6.0, 1.0 / 6.0, 0.5, 1.0 / 6.0},
    {1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0, 0.5}};

void orc_t10_quadrature(double *qx, double *qy, double *qz, double *qw) {
  const double w[5] = {-4.0 / 5.0, 9.0 / 20.0, 9.0 / 20.0, 9.0 / 20.0,
                       9.0 / 20.0};
  for (int q = 0; q < 5; q++) {
    qx[q] = T10_BARY[q][1];
    qy[q] = T10_BARY[q][2];
    qz[q] = T10_BARY[q][3];
    qw[q] = w[q] * (1.0 / 6.0);
  }
}

/* ------------------------------------------------------------------ */
/* 3x3 solve with partial pivoting, FEAT10DataFunc.cuh:30-83           */
/* ------------------------------------------------------------------ */
static void solve3(const double A[3][3], const double b[3], double x[3]) {
  double aug[3][4];
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) aug[i][j] = A[i][j];
    aug[i][3] = b[i];
  }
  for (int k = 0; k < 3; k++) {
    int piv = k;
    double mx = fabs(aug[k][k]);
    for (int i = k + 1; i < 3; i++)
      if (fabs(aug[i][k]) > mx) {
        mx = fabs(aug[i][k]);
        piv = i;
      }
    if (piv != k)
      for (int j = 0; j < 4; j++) {
        double t = aug[k][j];
        aug[k][j] = aug[piv][j];
        aug[piv][j] = t;
      }
    if (fabs(aug[k][k]) < 1e-14) {
      x[0] = x[1] = x[2] = 0.0;
      return;
    }
    for (int i = k + 1; i < 3; i++) {
      double f = aug[i][k] / aug[k][k];
      for (int j = k; j < 4; j++) aug[i][j] -= f * aug[k][j];
    }
  }
  x[2] = aug[2][3] / aug[2][2];
  x[1] = (aug[1][3] - aug[1][2] * x[2]) / aug[1][1];
  x[0] = (aug[0][3] - aug[0][2] * x[2] - aug[0][1] * x[1]) / aug[0][0];
}

/* ------------------------------------------------------------------ */
/* dn_du_pre_kernel, FEAT10Data.cu:97-204                              */
/* ------------------------------------------------------------------ */
void orc_t10_precompute(int n_elem, const int *conn, const double *x,
                        const double *y, const double *z, const double *qx,
                        const double *qy, const double *qz, double *grad_N,
                        double *detJ_out) {
  static const int edges[6][2] = {{0, 1}, {1, 2}, {0, 2},
                                  {0, 3}, {1, 3}, {2, 3}};
  static const double dL[4][3] = {
      {-1.0, -1.0, -1.0}, {1.0, 0.0, 0.0}, {0.0, 1.0, 0.0}, {0.0, 0.0, 1.0}};
#pragma omp parallel for schedule(static)
  for (int t = 0; t < n_elem * T10_Q; t++) {
    int e = t / T10_Q, q = t % T10_Q;
    double xi = qx[q], eta = qy[q], zeta = qz[q];
    double L[4] = {1.0 - xi - eta - zeta, xi, eta, zeta};
    double dN[10][3];
    for (int i = 0; i < 4; i++) {
      double f = 4.0 * L[i] - 1.0;
      for (int j = 0; j < 3; j++) dN[i][j] = f * dL[i][j];
    }
    for (int k = 0; k < 6; k++) {
      int i = edges[k][0], j = edges[k][1];
      for (int d = 0; d < 3; d++)
        dN[k + 4][d] = 4.0 * (L[i] * dL[j][d] + L[j] * dL[i][d]);
    }
    double X[10][3];
    for (int n = 0; n < 10; n++) {
      int g = conn[n * n_elem + e];
      X[n][0] = x[g];
      X[n][1] = y[g];
      X[n][2] = z[g];
    }
    double J[3][3] = {{0}};
    for (int a = 0; a < 10; a++)
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) J[i][j] += X[a][i] * dN[a][j];
    double det = J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) -
                 J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
                 J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
    detJ_out[e * T10_Q + q] = det;
    double JT[3][3];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) JT[i][j] = J[j][i];
    double *g = grad_N + (size_t)(e * T10_Q + q) * 30;
    for (int a = 0; a < 10; a++) {
      double h[3];
      solve3(JT, dN[a], h);
      g[0 * 10 + a] = h[0];
      g[1 * 10 + a] = h[1];
      g[2 * 10 + a] = h[2];
    }
  }
}

/* ------------------------------------------------------------------ */
/* Coefficient-level adjacency pattern (sorted unique (i,j) pairs):    */
/* BuildMassCSRPattern, FEAT10Data.cu:372-440 (keys :32-56).           */
/* `ecoef` is the per-element coefficient list, row-major [e][S].      */
/* Returns nnz; *columns_out is malloc'ed (free with orc_free).        */
/* ------------------------------------------------------------------ */
static int cmp_u64(const void *a, const void *b) {
  uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
  return (x > y) - (x < y);
}

int64_t orc_coef_pattern(int n_elem, int S, const int *ecoef, int n_coef,
                         int *offsets, int **columns_out) {
  size_t total = (size_t)n_elem * S * S;
  uint64_t *keys = (uint64_t *)malloc(total * sizeof(uint64_t));
  size_t k = 0;
  for (int e = 0; e < n_elem; e++)
    for (int i = 0; i < S; i++)
      for (int j = 0; j < S; j++)
        keys[k++] = ((uint64_t)(uint32_t)ecoef[e * S + i] << 32) |
                    (uint64_t)(uint32_t)ecoef[e * S + j];
  qsort(keys, total, sizeof(uint64_t), cmp_u64);
  size_t nnz = 0;
  for (size_t i = 0; i < total; i++)
    if (i == 0 || keys[i] != keys[i - 1]) keys[nnz++] = keys[i];
  int *cols = (int *)malloc((nnz ? nnz : 1) * sizeof(int));
  memset(offsets, 0, (size_t)(n_coef + 1) * sizeof(int));
  for (size_t i = 0; i < nnz; i++) {
    cols[i] = (int)(keys[i] & 0xffffffffULL);
    offsets[(int)(keys[i] >> 32) + 1]++;
  }
  for (int i = 0; i < n_coef; i++) offsets[i + 1] += offsets[i];
  free(keys);
  *columns_out = cols;
  return (int64_t)nnz;
}

void orc_free(void *p) { free(p); }

static int bsearch_col(const int *cols, int n, int target) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    int mid = lo + ((hi - lo) >> 1);
    int v = cols[mid];
    if (v == target) return mid;
    if (v < target)
      lo = mid + 1;
    else
      hi = mid - 1;
  }
  return -1;
}

/* FEAT10 element -> coefficient list (row-major [e][10]) from col-major conn */
void orc_t10_ecoef(int n_elem, const int *conn, int *ecoef) {
  for (int e = 0; e < n_elem; e++)
    for (int n = 0; n < 10; n++) ecoef[e * 10 + n] = conn[n * n_elem + e];
}

/* ------------------------------------------------------------------ */
/* mass_matrix_qp_kernel, FEAT10Data.cu:206-278                        */
/* values must be zeroed by the caller (CalcMassMatrix :356-361)       */
/* ------------------------------------------------------------------ */
void orc_t10_mass(int n_elem, const int *conn, const double *detJ,
                  const double *qx, const double *qy, const double *qz,
                  const double *qw, double rho, const int *offsets,
                  const int *columns, double *values) {
  static const int edges[6][2] = {{0, 1}, {1, 2}, {0, 2},
                                  {0, 3}, {1, 3}, {2, 3}};
  for (int e = 0; e < n_elem; e++)
    for (int il = 0; il < 10; il++)
      for (int jl = 0; jl < 10; jl++) {
        int ig = conn[il * n_elem + e], jg = conn[jl * n_elem + e];
        double m = 0.0;
        for (int q = 0; q < T10_Q; q++) {
          double xi = qx[q], eta = qy[q], zeta = qz[q], wq = qw[q];
          double L[4] = {1.0 - xi - eta - zeta, xi, eta, zeta};
          double N[10];
          for (int k = 0; k < 4; k++) N[k] = L[k] * (2.0 * L[k] - 1.0);
          for (int k = 0; k < 6; k++)
            N[k + 4] = 4.0 * L[edges[k][0]] * L[edges[k][1]];
          m += rho * N[il] * N[jl] * detJ[e * T10_Q + q] * wq;
        }
        int rs = offsets[ig], re = offsets[ig + 1];
        int p = bsearch_col(columns + rs, re - rs, jg);
        if (p >= 0) values[rs + p] += m;
      }
}

/* ------------------------------------------------------------------ */
/* Materials: SVK.cuh:14-55, MooneyRivlin.cuh:17-225                   */
/* ------------------------------------------------------------------ */
static void svk_P(const double F[3][3], double trFtF, const double FFtF[3][3],
                  double lambda, double mu, double P[3][3]) {
  double lf = lambda * (0.5 * trFtF - 1.5);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      P[i][j] = lf * F[i][j] + mu * (FFtF[i][j] - F[i][j]);
}

static void svk_tangent_block(const double Fh_i[3], const double Fh_j[3],
                              double hij, double trE, double Fhj_dot_Fhi,
                              const double FFT[3][3], double lambda, double mu,
                              double dV, double K[3][3]) {
  for (int d = 0; d < 3; d++)
    for (int e = 0; e < 3; e++) {
      double delta = (d == e) ? 1.0 : 0.0;
      double A = lambda * Fh_i[d] * Fh_j[e];
      double B = lambda * trE * hij * delta;
      double C1 = mu * Fhj_dot_Fhi * delta;
      double D = mu * Fh_j[d] * Fh_i[e];
      double Et = mu * hij * FFT[d][e];
      double Ft = -mu * hij * delta;
      K[d][e] = (A + B + C1 + D + Et + Ft) * dV;
    }
}

static double det3(const double A[3][3]) {
  return A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) -
         A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
         A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
}

static void invT3(const double A[3][3], double detA, double o[3][3]) {
  const double eps = 1e-12;
  double sd = detA;
  if (fabs(sd) < eps) sd = (sd >= 0.0) ? eps : -eps;
  double id = 1.0 / sd;
  o[0][0] = (A[1][1] * A[2][2] - A[1][2] * A[2][1]) * id;
  o[0][1] = (A[1][2] * A[2][0] - A[1][0] * A[2][2]) * id;
  o[0][2] = (A[1][0] * A[2][1] - A[1][1] * A[2][0]) * id;
  o[1][0] = (A[0][2] * A[2][1] - A[0][1] * A[2][2]) * id;
  o[1][1] = (A[0][0] * A[2][2] - A[0][2] * A[2][0]) * id;
  o[1][2] = (A[0][1] * A[2][0] - A[0][0] * A[2][1]) * id;
  o[2][0] = (A[0][1] * A[1][2] - A[0][2] * A[1][1]) * id;
  o[2][1] = (A[0][2] * A[1][0] - A[0][0] * A[1][2]) * id;
  o[2][2] = (A[0][0] * A[1][1] - A[0][1] * A[1][0]) * id;
}

typedef struct {
  double C[3][3], FC[3][3], FFT[3][3], FinvT[3][3];
  double I1, I2, J, t1, t2, t3;
} mr_common;

static void mr_setup(const double F[3][3], double mu10, double mu01,
                     double kappa, mr_common *m) {
  memset(m, 0, sizeof(*m));
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      for (int k = 0; k < 3; k++) m->C[i][j] += F[k][i] * F[k][j];
  m->I1 = m->C[0][0] + m->C[1][1] + m->C[2][2];
  double C2[3][3] = {{0}};
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      for (int k = 0; k < 3; k++) C2[i][j] += m->C[i][k] * m->C[k][j];
  double trC2 = C2[0][0] + C2[1][1] + C2[2][2];
  m->I2 = 0.5 * (m->I1 * m->I1 - trC2);
  m->J = det3(F);
  invT3(F, m->J, m->FinvT);
  double J13 = cbrt(m->J);
  double Jm23 = 1.0 / (J13 * J13);
  double Jm43 = Jm23 * Jm23;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      for (int k = 0; k < 3; k++) m->FC[i][j] += F[i][k] * m->C[k][j];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      for (int k = 0; k < 3; k++) m->FFT[i][j] += F[i][k] * F[j][k];
  m->t1 = 2.0 * mu10 * Jm23;
  m->t2 = 2.0 * mu01 * Jm43;
  m->t3 = kappa * (m->J - 1.0) * m->J;
}

static void mr_P(const double F[3][3], double mu10, double mu01, double kappa,
                 double P[3][3]) {
  mr_common m;
  mr_setup(F, mu10, mu01, kappa, &m);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      double term1 = F[i][j] - (m.I1 / 3.0) * m.FinvT[i][j];
      double term2 =
          m.I1 * F[i][j] - m.FC[i][j] - (2.0 * m.I2 / 3.0) * m.FinvT[i][j];
      double term3 = m.FinvT[i][j];
      P[i][j] = m.t1 * term1 + m.t2 * term2 + m.t3 * term3;
    }
}

static void mr_tangent(const double F[3][3], double mu10, double mu01,
                       double kappa, double A[3][3][3][3]) {
  mr_common m;
  mr_setup(F, mu10, mu01, kappa, &m);
  double term1[3][3], term2[3][3];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      term1[i][j] = F[i][j] - (m.I1 / 3.0) * m.FinvT[i][j];
      term2[i][j] =
          m.I1 * F[i][j] - m.FC[i][j] - (2.0 * m.I2 / 3.0) * m.FinvT[i][j];
    }
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      for (int k = 0; k < 3; k++)
        for (int l = 0; l < 3; l++) {
          double dik = (i == k) ? 1.0 : 0.0, djl = (j == l) ? 1.0 : 0.0;
          double dFinvT = -m.FinvT[i][l] * m.FinvT[k][j];
          double dt1 = (-2.0 / 3.0) * m.t1 * m.FinvT[k][l];
          double dt2 = (-4.0 / 3.0) * m.t2 * m.FinvT[k][l];
          double dt3 = (kappa * (2.0 * m.J - 1.0) * m.J) * m.FinvT[k][l];
          double dT1 = dik * djl - (2.0 / 3.0) * F[k][l] * m.FinvT[i][j] +
                       (m.I1 / 3.0) * m.FinvT[i][l] * m.FinvT[k][j];
          double dT2 =
              2.0 * F[k][l] * F[i][j] + m.I1 * dik * djl -
              (dik * m.C[l][j] + F[i][l] * F[k][j] + djl * m.FFT[i][k]) -
              (4.0 / 3.0) * (m.I1 * F[k][l] - m.FC[k][l]) * m.FinvT[i][j] +
              (2.0 * m.I2 / 3.0) * m.FinvT[i][l] * m.FinvT[k][j];
          A[i][j][k][l] = dt1 * term1[i][j] + m.t1 * dT1 + dt2 * term2[i][j] +
                          m.t2 * dT2 + dt3 * m.FinvT[i][j] + m.t3 * dFinvT;
        }
}

/* ------------------------------------------------------------------ */
/* Generic per-Gauss-point kernels.  All three elements share the same */
/* arithmetic once grad_N[e][q] (S x 3, column-major) and the element  */
/* coefficient list are known:                                         */
/*   compute_p            FEAT10DataFunc.cuh:85-293                    */
/*                        ANCF3243DataFunc.cuh:189-412                 */
/*                        ANCF3443DataFunc.cuh:281-495                 */
/*   compute_internal_force  :397-458 / :414-454 / :498-535            */
/*   compute_hessian_assemble_csr :512-791 / :642-940 / :724-1033      */
/* `cx,cy,cz` are per-coefficient current values (FEAT10: nodal        */
/* positions; ANCF: position + gradient coefficients).                 */
/* ------------------------------------------------------------------ */
#define MAXS 16

static void gp_F(int S, const int *ec, const double *cx, const double *cy,
                 const double *cz, const double *g, double F[3][3]) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) F[i][j] = 0.0;
  for (int a = 0; a < S; a++) {
    double xa[3] = {cx[ec[a]], cy[ec[a]], cz[ec[a]]};
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) F[i][j] += xa[i] * g[j * S + a];
  }
}

/* v (may be NULL) is the interleaved velocity [3*coef + i]; writes F, P and,
 * when Fdot/Pvis are non-NULL, the viscous buffers too. */
void orc_compute_p(int n_elem, int S, int Q, const int *ecoef,
                   const double *cx, const double *cy, const double *cz,
                   const double *grad_N, const orc_material *mat,
                   const double *v, double *F_out, double *P_out,
                   double *Fdot_out, double *Pvis_out) {
#pragma omp parallel for schedule(static)
  for (int t = 0; t < n_elem * Q; t++) {
    int e = t / Q;
    const int *ec = ecoef + (size_t)e * S;
    const double *g = grad_N + (size_t)t * 3 * S;
    double F[3][3];
    gp_F(S, ec, cx, cy, cz, g, F);
    double *Fo = F_out + (size_t)t * 9;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) Fo[j * 3 + i] = F[i][j];

    double Pvis[3][3] = {{0}};
    double Fdot[3][3] = {{0}};
    int do_damp =
        (v != NULL) && (mat->eta_damp != 0.0 || mat->lambda_damp != 0.0);
    if (do_damp) {
      for (int a = 0; a < S; a++) {
        const double *va = v + 3 * (size_t)ec[a];
        for (int i = 0; i < 3; i++)
          for (int j = 0; j < 3; j++) Fdot[i][j] += va[i] * g[j * S + a];
      }
      double A1[3][3] = {{0}}, A2[3][3] = {{0}};
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          for (int k = 0; k < 3; k++) {
            A1[i][j] += Fdot[k][i] * F[k][j];
            A2[i][j] += F[k][i] * Fdot[k][j];
          }
      double Edot[3][3];
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) Edot[i][j] = 0.5 * (A1[i][j] + A2[i][j]);
      double trEd = Edot[0][0] + Edot[1][1] + Edot[2][2];
      double Sv[3][3];
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          Sv[i][j] = 2.0 * mat->eta_damp * Edot[i][j] +
                     mat->lambda_damp * trEd * (i == j ? 1.0 : 0.0);
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          for (int k = 0; k < 3; k++) Pvis[i][j] += F[i][k] * Sv[k][j];
    }
    if (Fdot_out && Pvis_out) {
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
          Fdot_out[(size_t)t * 9 + j * 3 + i] = Fdot[i][j];
          Pvis_out[(size_t)t * 9 + j * 3 + i] = Pvis[i][j];
        }
    }

    double FtF[3][3] = {{0}}, FFt[3][3] = {{0}}, FFtF[3][3] = {{0}};
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < 3; k++) FtF[i][j] += F[k][i] * F[k][j];
    double trFtF = FtF[0][0] + FtF[1][1] + FtF[2][2];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < 3; k++) FFt[i][j] += F[i][k] * F[j][k];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < 3; k++) FFtF[i][j] += FFt[i][k] * F[k][j];
    double Pel[3][3];
    if (mat->model == 1)
      mr_P(F, mat->mu10, mat->mu01, mat->kappa, Pel);
    else
      svk_P(F, trFtF, FFtF, mat->lambda, mat->mu, Pel);
    double *Po = P_out + (size_t)t * 9;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) Po[j * 3 + i] = Pel[i][j] + Pvis[i][j];
  }
}

/* Accumulates into f_int (the caller clears it when the reference does:
 * SyncedNewton.cu:1049).  Serial, ascending (e, a) order: one fixed order out
 * of the many the reference's atomicAdd can produce. */
void orc_internal_force(int n_elem, int S, int Q, const int *ecoef,
                        const double *grad_N, const double *detJ,
                        const double *qw, const double *P, double *f_int) {
  for (int e = 0; e < n_elem; e++)
    for (int a = 0; a < S; a++) {
      double f[3] = {0, 0, 0};
      for (int q = 0; q < Q; q++) {
        size_t t = (size_t)e * Q + q;
        const double *Pq = P + t * 9;
        const double *g = grad_N + t * 3 * S;
        double dV = detJ[t] * qw[q];
        for (int i = 0; i < 3; i++) {
          double c = 0.0;
          for (int j = 0; j < 3; j++) c += Pq[j * 3 + i] * g[j * S + a];
          f[i] += c * dV;
        }
      }
      for (int i = 0; i < 3; i++) f_int[3 * (size_t)ecoef[e * S + a] + i] += f[i];
    }
}

/* DOF-level pattern from the coefficient pattern:
 * build_dof_row_nnz_from_coef_csr / fill_dof_cols_from_coef_csr,
 * SyncedNewton.cu:162-205.  Offsets are 64-bit here because nnz(H) of the
 * 10M-tet block exceeds 2^31. */
void orc_dof_pattern(int n_coef, const int *offsets, const int *columns,
                     int64_t *row_offsets, int *col_indices) {
  row_offsets[0] = 0;
  for (int i = 0; i < n_coef; i++) {
    int64_t deg = offsets[i + 1] - offsets[i];
    for (int d = 0; d < 3; d++)
      row_offsets[3 * (size_t)i + d + 1] = row_offsets[3 * (size_t)i + d] + 3 * deg;
  }
  if (!col_indices) return;
  for (int r = 0; r < 3 * n_coef; r++) {
    int i = r / 3;
    int64_t out = row_offsets[r];
    for (int idx = offsets[i]; idx < offsets[i + 1]; idx++) {
      int b = 3 * columns[idx];
      col_indices[out++] = b;
      col_indices[out++] = b + 1;
      col_indices[out++] = b + 2;
    }
  }
}

static int64_t find_dof_slot(const int *offsets, const int *columns, int gi,
                             int d, int gj, int e_) {
  /* row 3*gi+d begins at 9*offsets[gi] + d*3*deg; column 3*gj+e_ sits at
   * 3*pos + e_ where pos is gj's rank in the sorted coefficient row -- the
   * same slot the reference's per-entry binary search finds
   * (FEAT10DataFunc.cuh:487-502,678-683). */
  int rs = offsets[gi], deg = offsets[gi + 1] - rs;
  int pos = bsearch_col(columns + rs, deg, gj);
  if (pos < 0) return -1;
  return 9 * (int64_t)rs + (int64_t)d * 3 * deg + 3 * pos + e_;
}

/* assemble_sparse_hessian_tangent -> compute_hessian_assemble_csr.
 * Per (e,q): K_elem[3S][3S] then search-and-add h*K (and, if damping, the
 * Kelvin-Voigt C_elem unscaled).  With OpenMP the adds are `omp atomic`
 * (the reference's atomicAdd); single-threaded the order is (e,q) ascending. */
void orc_tangent_assemble(int n_elem, int S, int Q, const int *ecoef,
                          const double *cx, const double *cy, const double *cz,
                          const double *grad_N, const double *detJ,
                          const double *qw, const orc_material *mat, double h,
                          const int *offsets, const int *columns,
                          double *H_values) {
  const int D = 3 * S;
#pragma omp parallel
  {
    double *K = (double *)malloc(sizeof(double) * D * D);
    double *Cv = (double *)malloc(sizeof(double) * D * D);
#pragma omp for schedule(static)
    for (int t = 0; t < n_elem * Q; t++) {
      int e = t / Q, q = t % Q;
      const int *ec = ecoef + (size_t)e * S;
      const double *g = grad_N + (size_t)t * 3 * S;
      double F[3][3];
      gp_F(S, ec, cx, cy, cz, g, F);
      double C[3][3] = {{0}}, FFT[3][3] = {{0}};
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          for (int k = 0; k < 3; k++) C[i][j] += F[k][i] * F[k][j];
      double trE = 0.5 * (C[0][0] + C[1][1] + C[2][2] - 3.0);
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          for (int k = 0; k < 3; k++) FFT[i][j] += F[i][k] * F[j][k];
      double Fh[MAXS][3], hN[MAXS][3];
      for (int a = 0; a < S; a++) {
        for (int j = 0; j < 3; j++) hN[a][j] = g[j * S + a];
        for (int r = 0; r < 3; r++) {
          Fh[a][r] = 0.0;
          for (int c = 0; c < 3; c++) Fh[a][r] += F[r][c] * hN[a][c];
        }
      }
      double dV = detJ[t] * qw[q];
      double A[3][3][3][3];
      if (mat->model == 1) mr_tangent(F, mat->mu10, mat->mu01, mat->kappa, A);
      for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) {
          double hij =
              hN[j][0] * hN[i][0] + hN[j][1] * hN[i][1] + hN[j][2] * hN[i][2];
          double FhFh =
              Fh[j][0] * Fh[i][0] + Fh[j][1] * Fh[i][1] + Fh[j][2] * Fh[i][2];
          double Kb[3][3];
          if (mat->model == 1) {
            for (int d = 0; d < 3; d++)
              for (int e_ = 0; e_ < 3; e_++) {
                double s = 0.0;
                for (int Jx = 0; Jx < 3; Jx++)
                  for (int Lx = 0; Lx < 3; Lx++)
                    s += A[d][Jx][e_][Lx] * hN[i][Jx] * hN[j][Lx];
                Kb[d][e_] = s * dV;
              }
          } else {
            svk_tangent_block(Fh[i], Fh[j], hij, trE, FhFh, FFT, mat->lambda,
                              mat->mu, dV, Kb);
          }
          for (int d = 0; d < 3; d++)
            for (int e_ = 0; e_ < 3; e_++)
              K[(3 * i + d) * D + 3 * j + e_] = Kb[d][e_];
        }
      int damp = !(mat->eta_damp == 0.0 && mat->lambda_damp == 0.0);
      if (damp) {
        double et = mat->eta_damp, ld = mat->lambda_damp;
        for (int a = 0; a < S; a++)
          for (int b = 0; b < S; b++) {
            double hd = hN[a][0] * hN[b][0] + hN[a][1] * hN[b][1] +
                        hN[a][2] * hN[b][2];
            for (int d = 0; d < 3; d++)
              for (int e_ = 0; e_ < 3; e_++)
                Cv[(3 * a + d) * D + 3 * b + e_] =
                    (et * (Fh[b][d] * Fh[a][e_]) + et * FFT[d][e_] * hd +
                     ld * (Fh[a][d] * Fh[b][e_])) *
                    dV;
          }
      }
      for (int i = 0; i < S; i++)
        for (int d = 0; d < 3; d++)
          for (int j = 0; j < S; j++)
            for (int e_ = 0; e_ < 3; e_++) {
              int64_t s = find_dof_slot(offsets, columns, ec[i], d, ec[j], e_);
              if (s < 0) continue;
              double val = h * K[(3 * i + d) * D + 3 * j + e_];
#pragma omp atomic
              H_values[s] += val;
              if (damp) {
                double cv = Cv[(3 * i + d) * D + 3 * j + e_];
#pragma omp atomic
                H_values[s] += cv;
              }
            }
    }
    free(K);
    free(Cv);
  }
}

/* Same quantity, element-local: the dense 3S x 3S tangent of one element
 * (sum over its Q points), h = 1, no scatter.  Used by tests to compare with
 * the numpy prototype's tet10_tangent_svk. */
void orc_element_tangent(int S, int Q, const int *ec, const double *cx,
                         const double *cy, const double *cz,
                         const double *grad_N_e, const double *detJ_e,
                         const double *qw, const orc_material *mat,
                         double *K_out) {
  int off[MAXS + 1], col[MAXS * MAXS];
  /* local pattern: dense S x S over local ids */
  int lec[MAXS];
  for (int a = 0; a < S; a++) lec[a] = a;
  for (int a = 0; a <= S; a++) off[a] = a * S;
  for (int a = 0; a < S; a++)
    for (int b = 0; b < S; b++) col[a * S + b] = b;
  double lx[MAXS], ly[MAXS], lz[MAXS];
  for (int a = 0; a < S; a++) {
    lx[a] = cx[ec[a]];
    ly[a] = cy[ec[a]];
    lz[a] = cz[ec[a]];
  }
  memset(K_out, 0, sizeof(double) * 9 * S * S);
  orc_tangent_assemble(1, S, Q, lec, lx, ly, lz, grad_N_e, detJ_e, qw, mat,
                       1.0, off, col, K_out);
}

/* assemble_sparse_hessian_mass, SyncedNewton.cu:213-259 */
void orc_hessian_mass(int n_coef, const int *offsets, const int *columns,
                      const double *mass_values, double h, double *H_values) {
  double inv_h = 1.0 / h;
  for (int i = 0; i < n_coef; i++)
    for (int idx = offsets[i]; idx < offsets[i + 1]; idx++) {
      double c = mass_values[idx] * inv_h;
      for (int d = 0; d < 3; d++) {
        int64_t s = find_dof_slot(offsets, columns, i, d, columns[idx], d);
        if (s >= 0) H_values[s] += c;
      }
    }
}

/* assemble_sparse_hessian_constraints, SyncedNewton.cu:291-341.
 * J is CSR with rows = constraints, columns = DOFs. */
void orc_hessian_constraints(int n_constraints, const int *j_offsets,
                             const int *j_columns, const double *j_values,
                             double h, double rho, const int *offsets,
                             const int *columns, double *H_values) {
  double factor = h * h * rho;
  for (int c = 0; c < n_constraints; c++)
    for (int a = j_offsets[c]; a < j_offsets[c + 1]; a++)
      for (int b = j_offsets[c]; b < j_offsets[c + 1]; b++) {
        int di = j_columns[a], dj = j_columns[b];
        int64_t s =
            find_dof_slot(offsets, columns, di / 3, di % 3, dj / 3, dj % 3);
        if (s >= 0) H_values[s] += factor * j_values[a] * j_values[b];
      }
}

/* solver_grad_L, SyncedNewton.cu:343-407.  JT is CSR with rows = DOFs. */
void orc_grad_L(int n_coef, const int *offsets, const int *columns,
                const double *mass_values, const double *v_guess,
                const double *v_prev, const double *f_int, const double *f_ext,
                double h, int n_constraints, const int *jt_offsets,
                const int *jt_columns, const double *jt_values,
                const double *lam, const double *con, double rho, double *g) {
  double inv_dt = 1.0 / h;
  for (int tid = 0; tid < 3 * n_coef; tid++) {
    int ni = tid / 3, di = tid % 3;
    double res = 0.0;
    for (int idx = offsets[ni]; idx < offsets[ni + 1]; idx++) {
      int tj = columns[idx] * 3 + di;
      res += mass_values[idx] * (v_guess[tj] - v_prev[tj]) * inv_dt;
    }
    res -= (-f_int[tid]);
    res -= f_ext[tid];
    if (n_constraints > 0)
      for (int idx = jt_offsets[tid]; idx < jt_offsets[tid + 1]; idx++) {
        int c = jt_columns[idx];
        res += h * jt_values[idx] * (lam[c] + rho * con[c]);
      }
    g[tid] = res;
  }
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ================================================================== */
/* ANCF3243 / ANCF3443 set-up path                                    */
/*   basis + derivatives  ANCF3243DataFunc.cuh:115-152, ANCF3443DataFunc.cuh:114-238 */
/*   B^-T                 lib_utils/cpu_utils.cc:125-209, 211-441 (Eigen inverse -> Gauss-Jordan) */
/*   precompute           ANCF3243Data.cu:102-198, ANCF3443Data.cu:96-182   */
/*   mass                 ANCF3243Data.cu:200-288, ANCF3443Data.cu:184-254  */
/* ================================================================== */
static void ancf_b(int S, double u, double v, double w, double *b, double *bu,
                   double *bv, double *bw) {
  memset(bu, 0, S * sizeof(double));
  memset(bv, 0, S * sizeof(double));
  memset(bw, 0, S * sizeof(double));
  if (S == 8) {
    b[0] = 1.0; b[1] = u; b[2] = v; b[3] = w; b[4] = u * v; b[5] = u * w;
    b[6] = u * u; b[7] = u * u * u;
    bu[1] = 1.0; bu[4] = v; bu[5] = w; bu[6] = 2.0 * u; bu[7] = 3.0 * u * u;
    bv[2] = 1.0; bv[4] = u;
    bw[3] = 1.0; bw[5] = u;
  } else {
    b[0] = 1.0; b[1] = u; b[2] = v; b[3] = w; b[4] = u * v; b[5] = u * w;
    b[6] = v * w; b[7] = u * v * w; b[8] = u * u; b[9] = v * v;
    b[10] = (u * u) * v; b[11] = u * (v * v); b[12] = u * u * u;
    b[13] = v * v * v; b[14] = (u * u * u) * v; b[15] = u * (v * v * v);
    bu[1] = 1.0; bu[4] = v; bu[5] = w; bu[7] = v * w; bu[8] = 2.0 * u;
    bu[10] = 2.0 * u * v; bu[11] = v * v; bu[12] = 3.0 * u * u;
    bu[14] = 3.0 * u * u * v; bu[15] = v * v * v;
    bv[2] = 1.0; bv[4] = u; bv[6] = w; bv[7] = u * w; bv[9] = 2.0 * v;
    bv[10] = u * u; bv[11] = 2.0 * u * v; bv[13] = 3.0 * v * v;
    bv[14] = u * u * u; bv[15] = 3.0 * u * v * v;
    bw[3] = 1.0; bw[5] = u; bw[6] = v; bw[7] = u * v;
  }
}

/* out: column-major S x S,  out[j*S+i] = (B^T)^-1 (i,j) */
int orc_ancf_binv(int S, double L, double W, double H, double *out) {
  (void)H;
  double pu[4], pv[4];
  int np = (S == 8) ? 2 : 4;
  if (S == 8) {
    pu[0] = -L / 2.0; pu[1] = L / 2.0; pv[0] = pv[1] = 0.0;
  } else {
    pu[0] = -L / 2.0; pv[0] = -W / 2.0; pu[1] = L / 2.0; pv[1] = -W / 2.0;
    pu[2] = L / 2.0;  pv[2] = W / 2.0;  pu[3] = -L / 2.0; pv[3] = W / 2.0;
  }
  double a[16 * 16], inv[16 * 16]; /* a = B^T row-major */
  for (int p = 0; p < np; p++) {
    double b[16], bu[16], bv[16], bw[16];
    ancf_b(S, pu[p], pv[p], 0.0, b, bu, bv, bw);
    for (int c = 0; c < S; c++) {
      a[c * S + 4 * p + 0] = b[c];
      a[c * S + 4 * p + 1] = bu[c];
      a[c * S + 4 * p + 2] = bv[c];
      a[c * S + 4 * p + 3] = bw[c];
    }
  }
  for (int i = 0; i < S * S; i++) inv[i] = 0.0;
  for (int i = 0; i < S; i++) inv[i * S + i] = 1.0;
  for (int c = 0; c < S; c++) {
    int piv = c;
    for (int r = c + 1; r < S; r++)
      if (fabs(a[r * S + c]) > fabs(a[piv * S + c])) piv = r;
    if (a[piv * S + c] == 0.0) return -1;
    if (piv != c)
      for (int k = 0; k < S; k++) {
        double t = a[c * S + k]; a[c * S + k] = a[piv * S + k]; a[piv * S + k] = t;
        t = inv[c * S + k]; inv[c * S + k] = inv[piv * S + k]; inv[piv * S + k] = t;
      }
    double d = 1.0 / a[c * S + c];
    for (int k = 0; k < S; k++) { a[c * S + k] *= d; inv[c * S + k] *= d; }
    for (int r = 0; r < S; r++) {
      if (r == c) continue;
      double f = a[r * S + c];
      if (f == 0.0) continue;
      for (int k = 0; k < S; k++) {
        a[r * S + k] -= f * a[c * S + k];
        inv[r * S + k] -= f * inv[c * S + k];
      }
    }
  }
  for (int i = 0; i < S; i++)
    for (int j = 0; j < S; j++) out[j * S + i] = inv[i * S + j];
  return 0;
}

static void ancf_shape(int S, const double *Binv, double L, double W, double H,
                       double xi, double eta, double zeta, double *s,
                       double *dxi, double *deta, double *dzeta) {
  double b[16], bu[16], bv[16], bw[16];
  ancf_b(S, 0.5 * L * xi, 0.5 * W * eta, 0.5 * H * zeta, b, bu, bv, bw);
  for (int i = 0; i < S; i++) { bu[i] *= 0.5 * L; bv[i] *= 0.5 * W; bw[i] *= 0.5 * H; }
  for (int i = 0; i < S; i++) {
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    for (int j = 0; j < S; j++) {
      double m = Binv[j * S + i];
      a0 += m * b[j]; a1 += m * bu[j]; a2 += m * bv[j]; a3 += m * bw[j];
    }
    s[i] = a0; dxi[i] = a1; deta[i] = a2; dzeta[i] = a3;
  }
}

static void ancf_J(int S, const int *ec, const double *x, const double *y,
                   const double *z, const double *dxi, const double *deta,
                   const double *dzeta, double J[3][3]) {
  for (int i = 0; i < 3; i++) J[i][0] = J[i][1] = J[i][2] = 0.0;
  for (int a = 0; a < S; a++) {
    double X[3] = {x[ec[a]], y[ec[a]], z[ec[a]]};
    for (int i = 0; i < 3; i++) {
      J[i][0] += X[i] * dxi[a];
      J[i][1] += X[i] * deta[a];
      J[i][2] += X[i] * dzeta[a];
    }
  }
}

/* rule: n[3], points p[3][7], weights w[3][7]; q -> (q/(n1*n2), (q/n2)%n1, q%n2) */
void orc_ancf_precompute(int n_elem, int S, const int *ecoef, const double *x,
                         const double *y, const double *z, const double *LWH,
                         const double *Binv, const int *n, const double *pts,
                         double *grad_N, double *detJ) {
  const int Q = n[0] * n[1] * n[2];
  for (int e = 0; e < n_elem; e++)
    for (int q = 0; q < Q; q++) {
      int ix = q / (n[1] * n[2]), ie = (q / n[2]) % n[1], iz = q % n[2];
      double s[16], dxi[16], deta[16], dzeta[16], J[3][3], JT[3][3];
      ancf_shape(S, Binv + (size_t)e * S * S, LWH[e * 3], LWH[e * 3 + 1],
                 LWH[e * 3 + 2], pts[ix], pts[7 + ie], pts[14 + iz], s, dxi,
                 deta, dzeta);
      ancf_J(S, ecoef + (size_t)e * S, x, y, z, dxi, deta, dzeta, J);
      detJ[(size_t)e * Q + q] = det3(J);
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) JT[i][j] = J[j][i];
      double *g = grad_N + ((size_t)e * Q + q) * 3 * S;
      for (int a = 0; a < S; a++) {
        double rhs[3] = {dxi[a], deta[a], dzeta[a]}, gr[3];
        solve3(JT, rhs, gr);
        g[a] = gr[0]; g[S + a] = gr[1]; g[2 * S + a] = gr[2];
      }
    }
}

void orc_ancf_mass(int n_elem, int S, const int *ecoef, const double *x,
                   const double *y, const double *z, const double *LWH,
                   const double *Binv, const int *n, const double *pts,
                   const double *wts, double rho, const int *offsets,
                   const int *columns, double *values) {
  const int Q = n[0] * n[1] * n[2];
  for (int e = 0; e < n_elem; e++) {
    const int *ec = ecoef + (size_t)e * S;
    for (int q = 0; q < Q; q++) {
      int ix = q / (n[1] * n[2]), ie = (q / n[2]) % n[1], iz = q % n[2];
      double weight = wts[ix] * wts[7 + ie] * wts[14 + iz];
      double s[16], dxi[16], deta[16], dzeta[16], J[3][3];
      ancf_shape(S, Binv + (size_t)e * S * S, LWH[e * 3], LWH[e * 3 + 1],
                 LWH[e * 3 + 2], pts[ix], pts[7 + ie], pts[14 + iz], s, dxi,
                 deta, dzeta);
      ancf_J(S, ec, x, y, z, dxi, deta, dzeta, J);
      double dj = det3(J);
      for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) {
          int rs = offsets[ec[i]], re = offsets[ec[i] + 1];
          int p = bsearch_col(columns + rs, re - rs, ec[j]);
          if (p >= 0) values[rs + p] += rho * s[i] * s[j] * weight * dj;
        }
    }
  }
}
