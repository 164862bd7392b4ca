/*
 * bp_oracle.c -- CPU restatement, in plain C, of the reference's (biklooost/PROPhet) Behler-Parrinello
 * hot path.  TEST INFRASTRUCTURE ONLY.
 *
 *   Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 *   library.  Nothing under prophet_b200/ (the product) links, imports or executes it.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks every function below against
 * tests/golden/*.npz, which are outputs of the reference's own code (oracle/ref_driver.cpp linked against
 * the unmodified reference objects built by oracle/build_ref.sh), for nine cases covering the shipped
 * tutorial (doc/tutorial), the shipped ternary fixture (doc/PROPhet.xml) and synthetic Si / TiO2 / ABO4 /
 * all-variants structures.
 *
 * The loops deliberately keep the reference's *shape* (brute-force image scan, ordered (j,k) double visit of
 * every triplet, one descriptor row at a time, per-neuron evaluate) so that this file is an independent
 * check of the factorised GPU kernels rather than a copy of them.  Every function cites the reference
 * file:line it follows.  All arithmetic is IEEE double; compile with -ffp-contract=off so the neighbour
 * distance arithmetic is bit-identical to the reference's SSE2 build (no FMA).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------------------------------------ */
/* cutoff function and derivative: Structure.h:157-178 (fc), Structure.h:421-443 (d_fc) */
static double fc(double frac, int fcswitch)
{
  if (frac >= 1.0) return 0.0;
  if (fcswitch == 0) {
    double x = M_PI * frac;
    if (x < 1.047197551196598) {
      return (0.043539571484819 * x - 0.273567195834312) * x * x + 1.0;
    } else if (x < 2.094395102393195) {
      x -= 1.047197551196598;
      return ((0.087079142969639 * x - 0.136783597917156) * x - 0.429718346348117) * x + 0.75;
    } else {
      x -= 2.094395102393195;
      return ((0.043539571484819 * x + 0.136783597917156) * x - 0.429718346348117) * x + 0.25;
    }
  }
  return pow(tanh(1 - frac), 3);
}

static double d_fc(double Rc, double Rij, int fcswitch)
{
  if (Rij / Rc >= 1.0) return 0.0;
  if (fcswitch == 0) {
    double x = M_PI * Rij / Rc;
    if (x < 1.047197551196598) {
      return x * (0.410350793751469 * x - 1.718873385392471) / Rc;
    } else if (x < 2.094395102393195) {
      x -= 1.047197551196598;
      return (x * (0.820701587502935 * x - 0.859436692696234) - 1.35) / Rc;
    } else {
      x -= 2.094395102393195;
      return (x * (0.410350793751469 * x + 0.859436692696233) - 1.35) / Rc;
    }
  }
  return (-3 * tanh(1 - Rij / Rc) * tanh(1 - Rij / Rc) * (1 / cosh(1 - Rij / Rc)) * (1 / cosh(1 - Rij / Rc))) / (2 * Rij * Rc);
}

/* Structure.h:281-296 */
static double int_power(double base, int exponent)
{
  base = (exponent >= 0 ? base : 1.0 / base);
  unsigned int power = (unsigned int)abs(exponent);
  double result = 1;
  while (power) {
    if (power & 1) result *= base;
    power >>= 1;
    base *= base;
  }
  return result;
}

/* Structure.h:251-279; taylor[i] = 1/(i+1) for i<300, 0 above (Structure.cpp:72-75) */
static double angular_term(double theta, const double* row)
{
  theta *= row[4];
  if (theta <= -1) return 0;
  if (row[3] == (int)(row[3])) return int_power(1 + theta, (int)(row[3]));
  double new_xi = row[3];
  double alpha = new_xi * theta;
  int term = 1;
  double result = 1.0;
  while (fabs(alpha) > 1e-8) {
    result += alpha;
    double tay = term < 300 ? 1.0 / (double)(term + 1) : 0.0;
    alpha *= tay * theta * (new_xi - term);
    ++term;
  }
  if (term >= 250) return pow(1 + theta, row[3]);
  return result + alpha;
}

/* Structure.h:351-360 */
static double G1f(const double* row, double R) { return fc(R / row[0], (int)row[1]); }
static double G2f(const double* row, double R) { return exp(-row[2] * pow(R - row[3], 2)) * fc(R / row[0], (int)row[1]); }
/* Structure.h:361-367 */
static double G3f(const double* row, double R, double Ru, double Rjk, double theta)
{
  double term = 1 + (row[4] * theta);
  if (term <= 0) return 0.0;
  return pow(2, 1 - row[3]) * pow(1 + (row[4] * theta), row[3]) * exp(-row[2] * (R * R + Ru * Ru + Rjk * Rjk)) *
         fc(R / row[0], (int)row[1]) * fc(Ru / row[0], (int)row[1]) * fc(Rjk / row[0], (int)row[1]);
}
/* Structure.h:368-418 (the exp/fc memoisation there is value-neutral and is not restated) */
static double G4f(const double* row, double R, double Ru, double theta)
{
  double pref = pow(2, 1 - row[3]);
  double term = (1.0 + row[4] * theta);
  if (term <= 0) return 0;
  double ang = angular_term(theta, row);
  double e1 = exp(-row[2] * (R * R));
  double e2 = exp(-row[2] * (Ru * Ru));
  double f1 = fc(R / row[0], (int)row[1]);
  double f2 = fc(Ru / row[0], (int)row[1]);
  return pref * ang * e1 * e2 * f1 * f2;
}

/* ------------------------------------------------------------------------------------------------ */
/* Structure::Set_Convert + invert/transpose (Structure.cpp:1159-1186, Structure.h:180-197,239-254) and
 * Structure::Normalize (Structure.cpp:1239-1279): Cartesian -> fractional, single wrap by +-1.
 * cell = rows a,b,c.  FtoA = cell^T, AtoF = (cell^T)^-1 by cofactors. */
static void make_convert(const double* cell, double FtoA[3][3], double AtoF[3][3])
{
  double c[3][3];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { FtoA[j][i] = cell[3 * i + j]; }
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) c[i][j] = FtoA[i][j];
  double det = c[0][0] * (c[1][1] * c[2][2] - c[2][1] * c[1][2]) - c[0][1] * (c[1][0] * c[2][2] - c[1][2] * c[2][0]) +
               c[0][2] * (c[1][0] * c[2][1] - c[1][1] * c[2][0]);
  double invdet = 1 / det;
  AtoF[0][0] = (c[1][1] * c[2][2] - c[2][1] * c[1][2]) * invdet;
  AtoF[0][1] = (c[0][2] * c[2][1] - c[0][1] * c[2][2]) * invdet;
  AtoF[0][2] = (c[0][1] * c[1][2] - c[0][2] * c[1][1]) * invdet;
  AtoF[1][0] = (c[1][2] * c[2][0] - c[1][0] * c[2][2]) * invdet;
  AtoF[1][1] = (c[0][0] * c[2][2] - c[0][2] * c[2][0]) * invdet;
  AtoF[1][2] = (c[1][0] * c[0][2] - c[0][0] * c[1][2]) * invdet;
  AtoF[2][0] = (c[1][0] * c[2][1] - c[2][0] * c[1][1]) * invdet;
  AtoF[2][1] = (c[2][0] * c[0][1] - c[0][0] * c[2][1]) * invdet;
  AtoF[2][2] = (c[0][0] * c[1][1] - c[1][0] * c[0][1]) * invdet;
}

/* Structure.h:320-334 To_Cart / :336-350 To_Frac */
static void mat_apply(double M[3][3], double* p)
{
  double t1 = M[0][0] * p[0] + M[0][1] * p[1] + M[0][2] * p[2];
  double t2 = M[1][0] * p[0] + M[1][1] * p[1] + M[1][2] * p[2];
  double t3 = M[2][0] * p[0] + M[2][1] * p[1] + M[2][2] * p[2];
  p[0] = t1; p[1] = t2; p[2] = t3;
}

void orc_cart_to_frac(const double* cell, const double* cart, int n, double* frac)
{
  double FtoA[3][3], AtoF[3][3];
  make_convert(cell, FtoA, AtoF);
  for (int i = 0; i < n; i++) {
    double p[3] = {cart[3 * i], cart[3 * i + 1], cart[3 * i + 2]};
    mat_apply(AtoF, p);
    for (int j = 0; j < 3; j++) {
      if (p[j] < 0) p[j] += 1;
      else if (p[j] > 1) p[j] -= 1;
      frac[3 * i + j] = p[j];
    }
  }
}

/* ------------------------------------------------------------------------------------------------ */
/* Structure::Set_Neighbor, periodic branch (Structure.cpp:1298-1353).
 * Outputs (caller-allocated, capacities given; returns 0 or -1 on overflow):
 *   translations[nT][3]; neigh_offset[n+1]; neigh_j[] unique j ascending; image_offset[#pairs+1]; image_t[] */
static int translation_table(const double* cell, double Rcut, int* mx, int* my, int* mz)
{
  double na = sqrt(pow(cell[0], 2) + pow(cell[1], 2) + pow(cell[2], 2));
  double nb = sqrt(pow(cell[3], 2) + pow(cell[4], 2) + pow(cell[5], 2));
  double nc = sqrt(pow(cell[6], 2) + pow(cell[7], 2) + pow(cell[8], 2));
  *mx = (int)ceil(1.1 * Rcut / na);
  *my = (int)ceil(1.1 * Rcut / nb);
  *mz = (int)ceil(1.1 * Rcut / nc);
  return (2 * *mx + 1) * (2 * *my + 1) * (2 * *mz + 1);
}

int orc_num_translations(const double* cell, double Rcut)
{
  int mx, my, mz;
  return translation_table(cell, Rcut, &mx, &my, &mz);
}

int orc_neighbors_periodic(const double* cell, const double* frac, int n, double Rcut,
                           int* translations, int* neigh_offset, int* neigh_j, int cap_j,
                           int* image_offset, int* image_t, int cap_t)
{
  double FtoA[3][3], AtoF[3][3];
  make_convert(cell, FtoA, AtoF);
  int mx, my, mz;
  int nT = translation_table(cell, Rcut, &mx, &my, &mz);
  int t = 0;
  for (int ix = -mx; ix <= mx; ix++)
    for (int iy = -my; iy <= my; iy++)
      for (int iz = -mz; iz <= mz; iz++) {
        translations[3 * t] = ix; translations[3 * t + 1] = iy; translations[3 * t + 2] = iz; t++;
      }
  int nj = 0, nt = 0;
  neigh_offset[0] = 0;
  image_offset[0] = 0;
  for (int i = 0; i < n; i++) {
    for (int j = 0; j < n; j++) {
      int found = 0;
      for (t = 0; t < nT; t++) {
        double del[3];
        for (int z = 0; z < 3; z++) {
          double jr = frac[3 * j + z] + translations[3 * t + z];
          del[z] = jr - frac[3 * i + z];
        }
        mat_apply(FtoA, del);
        double distance = sqrt(del[0] * del[0] + del[1] * del[1] + del[2] * del[2]);
        if (distance > 0.01 && distance <= Rcut) {
          if (!found) {
            if (nj >= cap_j) return -1;
            neigh_j[nj] = j;
            found = 1;
          }
          if (nt >= cap_t) return -1;
          image_t[nt++] = t;
        }
      }
      if (found) { nj++; image_offset[nj] = nt; }
    }
    neigh_offset[i + 1] = nj;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------------ */
/* descriptor slot maps: Structure::init_type_G_map (Structure.cpp:1491-1512), G_index (Structure.h:490-500) */
typedef struct {
  int n_elem, nG1, nG2, nG3, nG4;
  const double *G1, *G2, *G3, *G4;
  double Rcut;
} rows_t;

static int slot_pair(const rows_t* r, int row, int e) { return (r->nG1 + r->nG2) * e + row; }
static int slot_trip(const rows_t* r, int row, int e1, int e2)
{
  int lo = e1 < e2 ? e1 : e2, hi = e1 < e2 ? e2 : e1;
  int index = 0;
  for (int i = 0; i < r->n_elem; i++)
    for (int j = i; j < r->n_elem; j++) {
      if (i == lo && j == hi) return (r->nG1 + r->nG2) * r->n_elem + index + row;
      index += r->nG3 + r->nG4;
    }
  return -1;
}

int orc_num_descriptors(int n_elem, int nG1, int nG2, int nG3, int nG4)
{
  return n_elem * (nG1 + nG2) + n_elem * (n_elem + 1) / 2 * (nG3 + nG4);
}

/* Structure::Calc_G periodic branch (Structure.cpp:985-1071) for every atom of one structure.
 * elem[i] = index of atom i's element in the ascending-Z element list.  G_out is [NG][n] like Structure::G. */
int orc_descriptors_periodic(const double* cell, const double* frac, const int* elem, int n,
                             int n_elem, int nG1, const double* G1, int nG2, const double* G2,
                             int nG3, const double* G3, int nG4, const double* G4, double Rcut, double* G_out)
{
  rows_t R_ = {n_elem, nG1, nG2, nG3, nG4, G1, G2, G3, G4, Rcut};
  const rows_t* rw = &R_;
  int NG = orc_num_descriptors(n_elem, nG1, nG2, nG3, nG4);
  int nT = orc_num_translations(cell, Rcut);
  int cap_j = n * n, cap_t = n * n * nT;
  int* translations = (int*)malloc(sizeof(int) * 3 * nT);
  int* neigh_offset = (int*)malloc(sizeof(int) * (n + 1));
  int* neigh_j = (int*)malloc(sizeof(int) * cap_j);
  int* image_offset = (int*)malloc(sizeof(int) * (cap_j + 1));
  int* image_t = (int*)malloc(sizeof(int) * cap_t);
  int rc = orc_neighbors_periodic(cell, frac, n, Rcut, translations, neigh_offset, neigh_j, cap_j, image_offset, image_t, cap_t);
  if (rc) goto done;
  double FtoA[3][3], AtoF[3][3];
  make_convert(cell, FtoA, AtoF);
  double* G = (double*)malloc(sizeof(double) * NG);
  for (int i = 0; i < n; i++) {
    for (int g = 0; g < NG; g++) G[g] = 0.0;
    const double* ir = frac + 3 * i;
    for (int p = neigh_offset[i]; p < neigh_offset[i + 1]; p++) {
      int j = neigh_j[p];
      for (int q = image_offset[p]; q < image_offset[p + 1]; q++) {
        const int* T = translations + 3 * image_t[q];
        double jr[3], del[3];
        for (int z = 0; z < 3; z++) { jr[z] = frac[3 * j + z] + T[z]; del[z] = jr[z] - ir[z]; }
        mat_apply(FtoA, del);
        double R = sqrt(del[0] * del[0] + del[1] * del[1] + del[2] * del[2]);
        if (R < 0.01) continue;
        for (int r1 = 0; r1 < nG1; r1++) G[slot_pair(rw, r1, elem[j])] += G1f(G1 + 2 * r1, R);
        for (int r1 = 0; r1 < nG2; r1++) G[slot_pair(rw, r1 + nG1, elem[j])] += G2f(G2 + 4 * r1, R);
        if (nG3 == 0 && nG4 == 0) continue;
        for (int p2 = neigh_offset[i]; p2 < neigh_offset[i + 1]; p2++) {
          int j2 = neigh_j[p2];
          for (int q2 = image_offset[p2]; q2 < image_offset[p2 + 1]; q2++) {
            const int* T2 = translations + 3 * image_t[q2];
            double del2[3], del3[3];
            for (int z = 0; z < 3; z++) {
              del2[z] = (frac[3 * j2 + z] + T2[z]) - ir[z];
              del3[z] = (frac[3 * j2 + z] + T2[z]) - jr[z];
            }
            mat_apply(FtoA, del2);
            mat_apply(FtoA, del3);
            double Ru = sqrt(del2[0] * del2[0] + del2[1] * del2[1] + del2[2] * del2[2]);
            double Rjk = sqrt(del3[0] * del3[0] + del3[1] * del3[1] + del3[2] * del3[2]);
            if (Ru < 0.01 || Rjk < 0.01) continue;
            double theta = 0;
            for (int z = 0; z < 3; z++) theta += del2[z] * del[z];
            theta = theta / (R * Ru);
            for (int r1 = 0; r1 < nG3; r1++) G[slot_trip(rw, r1, elem[j], elem[j2])] += G3f(G3 + 5 * r1, R, Ru, Rjk, theta);
            for (int r1 = 0; r1 < nG4; r1++) G[slot_trip(rw, r1 + nG3, elem[j], elem[j2])] += G4f(G4 + 5 * r1, R, Ru, theta);
          }
        }
      }
    }
    for (int g = 0; g < NG; g++) G_out[(size_t)g * n + i] = G[g];
  }
  free(G);
done:
  free(translations); free(neigh_offset); free(neigh_j); free(image_offset); free(image_t);
  return rc;
}

/* Structure::Calc_G non-periodic (LAMMPS) branch (Structure.cpp:1073-1145) for local atoms 0..nlocal-1.
 * x[ntotal][3] Cartesian, elem[ntotal], neighbour list as offsets/flat.  G_out is [nlocal][NG]. */
void orc_descriptors_md(const double* x, const int* elem, int nlocal, const long long* offsets, const int* neigh,
                        int n_elem, int nG1, const double* G1, int nG2, const double* G2,
                        int nG3, const double* G3, int nG4, const double* G4, double Rcut, double* G_out)
{
  rows_t R_ = {n_elem, nG1, nG2, nG3, nG4, G1, G2, G3, G4, Rcut};
  const rows_t* rw = &R_;
  int NG = orc_num_descriptors(n_elem, nG1, nG2, nG3, nG4);
  for (int i = 0; i < nlocal; i++) {
    double* G = G_out + (size_t)i * NG;
    for (int g = 0; g < NG; g++) G[g] = 0.0;
    const double* ir = x + 3 * i;
    for (long long p = offsets[i]; p < offsets[i + 1]; p++) {
      int j = neigh[p];
      if (j == i) continue;
      double jr[3], del[3];
      for (int z = 0; z < 3; z++) { jr[z] = x[3 * j + z]; del[z] = jr[z] - ir[z]; }
      double R = sqrt(del[0] * del[0] + del[1] * del[1] + del[2] * del[2]);
      if (R > Rcut) continue;
      for (int r1 = 0; r1 < nG1; r1++) G[slot_pair(rw, r1, elem[j])] += G1f(G1 + 2 * r1, R);
      for (int r1 = 0; r1 < nG2; r1++) G[slot_pair(rw, r1 + nG1, elem[j])] += G2f(G2 + 4 * r1, R);
      if (nG3 == 0 && nG4 == 0) continue;
      for (long long p2 = offsets[i]; p2 < offsets[i + 1]; p2++) {
        int j2 = neigh[p2];
        if (j2 == j || j2 == i) continue;
        double del2[3], del3[3];
        for (int z = 0; z < 3; z++) { double ju = x[3 * j2 + z]; del2[z] = ju - ir[z]; del3[z] = jr[z] - ju; }
        double Ru = sqrt(del2[0] * del2[0] + del2[1] * del2[1] + del2[2] * del2[2]);
        if (Ru > Rcut) continue;
        double Rjk = sqrt(del3[0] * del3[0] + del3[1] * del3[1] + del3[2] * del3[2]);
        double theta = 0;
        for (int z = 0; z < 3; z++) theta += del2[z] * del[z];
        theta = theta / (R * Ru);
        if (nG3) {
          if (Rjk > Rcut) continue;  /* also skips the G4 rows: Structure.cpp:1129-1132 */
          for (int r1 = 0; r1 < nG3; r1++) G[slot_trip(rw, r1, elem[j], elem[j2])] += G3f(G3 + 5 * r1, R, Ru, Rjk, theta);
        }
        for (int r1 = 0; r1 < nG4; r1++) G[slot_trip(rw, r1 + nG3, elem[j], elem[j2])] += G4f(G4 + 5 * r1, R, Ru, theta);
      }
    }
  }
}

/* ------------------------------------------------------------------------------------------------ */
/* Transfer functions: Neural_network_node (NN_node.h:144-254).  ids are this repository's own enumeration. */
enum { TF_LINEAR = 0, TF_TANH = 1, TF_TANH_SPLINE = 2, TF_TANH_CRAZY = 3, TF_CUBIC = 4, TF_PULSE = 5, TF_SIN = 6,
       TF_PERIODIC_TANH = 7, TF_RELU = 8 };

/* Tables("tanh") (Tables.cpp:212-221, Tables.h:74-128): C2 cubic spline of tanh through the 1001 knots
 * X_i = -17 + 0.034 i.  The reference ships the coefficients as data (Table_tanh_accurate.data); they are
 * regenerated here by solving the natural-spline tridiagonal system, which reproduces the shipped values to
 * ~7e-12 in the cubic coefficient (2e-16 in the function value).  Saturation: below X_0 -> a[0][3]; at or above
 * X_N -> a[N-1][0]+a[N-1][1]+a[N-1][2]+a[N-1][3] (Tables.cpp:216-218). */
#define TANH_N 1000
static double tX[TANH_N + 1], tA[TANH_N][4];
static int tanh_table_ready = 0;
static void tanh_table_init(void)
{
  if (tanh_table_ready) return;
  static double y[TANH_N + 1], M[TANH_N + 1], cp[TANH_N + 1], dp[TANH_N + 1];
  for (int i = 0; i <= TANH_N; i++) { tX[i] = (double)(-17000 + 34 * i) / 1000.0; y[i] = tanh(tX[i]); }
  /* natural spline second derivatives M: h/6 M[i-1] + 2h/3 M[i] + h/6 M[i+1] = (y[i+1]-2y[i]+y[i-1])/h */
  M[0] = M[TANH_N] = 0.0;
  cp[0] = 0.0; dp[0] = 0.0;
  for (int i = 1; i < TANH_N; i++) {
    double h0 = tX[i] - tX[i - 1], h1 = tX[i + 1] - tX[i];
    double a = h0 / 6.0, b = (h0 + h1) / 3.0, c = h1 / 6.0;
    double d = (y[i + 1] - y[i]) / h1 - (y[i] - y[i - 1]) / h0;
    double m = b - a * cp[i - 1];
    cp[i] = c / m;
    dp[i] = (d - a * dp[i - 1]) / m;
  }
  for (int i = TANH_N - 1; i >= 1; i--) M[i] = dp[i] - cp[i] * M[i + 1];
  for (int i = 0; i < TANH_N; i++) {
    double h = tX[i + 1] - tX[i];
    tA[i][0] = (M[i + 1] - M[i]) / (6.0 * h);
    tA[i][1] = M[i] / 2.0;
    tA[i][2] = (y[i + 1] - y[i]) / h - h * (2.0 * M[i] + M[i + 1]) / 6.0;
    tA[i][3] = y[i];
  }
  tanh_table_ready = 1;
}

static int tanh_row(double v)
{
  /* Tables::fast_row (Tables.h:74-95) with slow_row's result (the bin with X[bin] <= v < X[bin+1]) as fallback */
  if (v < tX[0]) return -1;
  if (v >= tX[TANH_N]) return TANH_N + 1;
  double m = (double)TANH_N / (tX[TANH_N] - tX[0]);
  double b = -m * tX[0];
  int bin = (int)(m * v + b);
  if (bin < 0) bin = 0;
  if (bin > TANH_N - 1) bin = TANH_N - 1;
  if (v < tX[bin] || v > tX[bin + 1]) {
    while (bin > 0 && v < tX[bin]) bin--;
    while (bin < TANH_N - 1 && v >= tX[bin + 1]) bin++;
  }
  return bin;
}

static double transfer(int id, double in, double* deriv)
{
  switch (id) {
    case TF_LINEAR: *deriv = 1.0; return in;                                      /* NN_node.h:232-241 */
    case TF_TANH: {                                                               /* NN_node.h:193-206 */
      if (fabs(in) > 6) {
        int sign = (0.0 < in) - (in < 0.0);
        double exptl = exp(-2 * sign * in);
        *deriv = 4.0 * exptl * (1 - 2 * exptl);
        return sign * (1 - 2 * exptl * (1 - exptl));
      }
      *deriv = 1.0 / pow(cosh(in), 2);
      return tanh(in);
    }
    case TF_TANH_SPLINE: {                                                        /* NN_node.h:144-153 */
      tanh_table_init();
      int r = tanh_row(in);
      if (r < 0) { *deriv = 0; return tA[0][3]; }
      if (r > TANH_N) { *deriv = 0; return tA[TANH_N - 1][0] + tA[TANH_N - 1][1] + tA[TANH_N - 1][2] + tA[TANH_N - 1][3]; }
      double x = in - tX[r];
      *deriv = (3 * tA[r][0] * x + 2 * tA[r][1]) * x + tA[r][2];
      return ((tA[r][0] * x + tA[r][1]) * x + tA[r][2]) * x + tA[r][3];
    }
    case TF_TANH_CRAZY:                                                           /* NN_node.h:208-215 */
      *deriv = (1 - pow(0.1 * in, 4)) / pow(cosh(in), 2) - tanh(in) * 0.1 * 4 * pow(0.1 * in, 3);
      return (1 - pow(0.1 * in, 4)) * tanh(in);
    case TF_CUBIC: { double x2 = in * in; *deriv = 3 * x2; return x2 * in; }       /* NN_node.h:218-225 */
    case TF_PULSE:                                                                /* NN_node.h:181-190, a=b=c=1 */
      *deriv = exp(-in * in) * (-2 * in * sin(in) + cos(in));
      return sin(in) * exp(-in * in);
    case TF_SIN: *deriv = cos(in); return sin(in);                                /* NN_node.h:173-179 */
    case TF_PERIODIC_TANH: {                                                      /* NN_node.h:155-162 */
      double inner = 4.0 * sin(in);
      *deriv = 4.0 * cos(in) / pow(cosh(inner), 2);
      return tanh(inner);
    }
    case TF_RELU: *deriv = in <= 0.0 ? 0.0 : 1.0; return in <= 0.0 ? 0.0 : in;     /* NN_node.h:243-257 */
  }
  *deriv = 0;
  return 0;
}

double orc_transfer(int id, double in, double* deriv) { return transfer(id, in, deriv); }

/* ------------------------------------------------------------------------------------------------ */
/* One atom through one element's net.
 *   sizes[n_layers+1] = {NG, hidden..., 1}; tf[] one id per node (all layers, last = linear);
 *   params layer-major, node-major, [w_0..w_{n-1}, b] (NN.cpp:83-94, deriv_offsets).
 * Follows Neural_network::train_MD (NN.cpp:626-697) / evaluate_MD(dE_dG) (NN.cpp:516-570) with
 * Neural_network_node::evaluate (NN_node.h:64-81) inlined.
 * Returns out_scale*y.  dE_dG (may be NULL) gets out_scale/scale_k * dy/dx~_k; dY_dP (may be NULL) gets
 * ACCUMULATED += out_scale * dy/dparam. */
#define ORC_MAX_WIDTH 1024
double orc_net_atom(const int* sizes, int n_layers, const int* tf, const double* params,
                    const double* mean, const double* scale, double out_scale,
                    const double* Gin, double* dE_dG, double* dY_dP)
{
  static double vals[2][ORC_MAX_WIDTH];
  int NG = sizes[0];
  int total_nodes = 0;
  for (int l = 1; l <= n_layers; l++) total_nodes += sizes[l];
  double* act = (double*)malloc(sizeof(double) * (NG + total_nodes));   /* inputs to each layer, concatenated */
  double* dT = (double*)malloc(sizeof(double) * total_nodes);           /* T'(X) per node */
  for (int i = 0; i < NG; i++) {
    double v = Gin[i] - mean[i];
    act[i] = (!v) ? 0.0 : v / scale[i];                                 /* NN.cpp:648-654 */
  }
  int in_off = 0, out_off = NG, node = 0, p = 0;
  for (int l = 0; l < n_layers; l++) {
    int n_in = sizes[l], n_out = sizes[l + 1];
    for (int k = 0; k < n_out; k++) {
      double X = 0.0;
      for (int i = 0; i < n_in; i++) X += act[in_off + i] * params[p + i];
      X += params[p + n_in];
      double d;
      act[out_off + k] = transfer(tf[node], X, &d);
      dT[node] = d;
      node++;
      p += n_in + 1;
    }
    in_off = out_off;
    out_off += n_out;
  }
  double y = act[in_off];
  (void)vals;
  /* backprop: delta[node] = dy/dX_node */
  if (dE_dG || dY_dP) {
    double* delta = (double*)malloc(sizeof(double) * total_nodes);
    double* din = (double*)malloc(sizeof(double) * NG);
    /* walk layers from the top */
    int* layer_node0 = (int*)malloc(sizeof(int) * (n_layers + 1));
    int* layer_p0 = (int*)malloc(sizeof(int) * (n_layers + 1));
    int* layer_in0 = (int*)malloc(sizeof(int) * (n_layers + 1));
    int nn = 0, pp = 0, ii = 0;
    for (int l = 0; l < n_layers; l++) {
      layer_node0[l] = nn; layer_p0[l] = pp; layer_in0[l] = ii;
      nn += sizes[l + 1]; pp += sizes[l + 1] * (sizes[l] + 1); ii += sizes[l];
    }
    delta[layer_node0[n_layers - 1]] = dT[layer_node0[n_layers - 1]];   /* output node: dy/dX = T' (linear -> 1) */
    for (int l = n_layers - 1; l >= 0; l--) {
      int n_in = sizes[l], n_out = sizes[l + 1];
      if (dY_dP) {
        for (int k = 0; k < n_out; k++) {
          double dl = delta[layer_node0[l] + k];
          int pk = layer_p0[l] + k * (n_in + 1);
          for (int i = 0; i < n_in; i++) dY_dP[pk + i] += out_scale * dl * act[layer_in0[l] + i];
          dY_dP[pk + n_in] += out_scale * dl;
        }
      }
      /* dy/d(input_i of layer l) = sum_k delta_k w_ki */
      for (int i = 0; i < n_in; i++) {
        double s = 0.0;
        for (int k = 0; k < n_out; k++) s += delta[layer_node0[l] + k] * params[layer_p0[l] + k * (n_in + 1) + i];
        if (l > 0) delta[layer_node0[l - 1] + i] = s * dT[layer_node0[l - 1] + i];
        else din[i] = s;
      }
    }
    if (dE_dG) for (int i = 0; i < NG; i++) dE_dG[i] = din[i] * out_scale / scale[i];   /* NN.cpp:562-566 */
    free(delta); free(din); free(layer_node0); free(layer_p0); free(layer_in0);
  }
  free(act); free(dT);
  return out_scale * y;
}

/* ------------------------------------------------------------------------------------------------ */
/* One pass of Potential::train's loop body (Potential.cpp:626-670) over structures whose descriptors are
 * given atom-major: G[total_atoms][NG].  Per element e: sizes_e, tf_e, params_e, mean_e, scale_e passed as
 * arrays of pointers flattened by the Python side into offsets.
 *   atom_offset[n_struct+1], elem[total_atoms], target[n_struct], tag[n_struct] (0 train, 1 val, 2 other)
 *   net_sizes[n_elem][max_layers+1] with n_layers[n_elem]; tf_off/param_off[n_elem+1] offsets into tf_all/params_all
 *   mean_all/scale_all [n_elem][NG]
 * Outputs: E_struct[n_struct] (0 for 'other'), *SSE, *vSSE, *nval, gradient[P_total], E_atom[total_atoms] (may be NULL). */
void orc_epoch(int n_struct, const int* atom_offset, const int* elem, const double* G, int NG,
               const double* target, const int* tag, int n_elem, const int* n_layers, const int* net_sizes,
               int max_layers, const int* tf_all, const int* tf_off, const double* params_all, const int* param_off,
               const double* mean_all, const double* scale_all, double out_mean, double out_scale,
               double* E_struct, double* SSE, double* vSSE, int* nval, double* gradient, double* E_atom)
{
  int P = param_off[n_elem];
  double* dYdP = (double*)malloc(sizeof(double) * (P > 0 ? P : 1));
  for (int i = 0; i < P; i++) gradient[i] = 0.0;
  *SSE = 0; *vSSE = 0; *nval = 0;
  for (int s = 0; s < n_struct; s++) {
    E_struct[s] = 0.0;
    if (tag[s] != 0 && tag[s] != 1) {
      if (E_atom) for (int a = atom_offset[s]; a < atom_offset[s + 1]; a++) {
        int e = elem[a];
        E_atom[a] = orc_net_atom(net_sizes + e * (max_layers + 1), n_layers[e], tf_all + tf_off[e], params_all + param_off[e],
                                 mean_all + (size_t)e * NG, scale_all + (size_t)e * NG, out_scale, G + (size_t)a * NG, NULL, NULL);
      }
      continue;
    }
    for (int i = 0; i < P; i++) dYdP[i] = 0.0;
    double output = 0;
    for (int a = atom_offset[s]; a < atom_offset[s + 1]; a++) {
      int e = elem[a];
      double Ea = orc_net_atom(net_sizes + e * (max_layers + 1), n_layers[e], tf_all + tf_off[e], params_all + param_off[e],
                               mean_all + (size_t)e * NG, scale_all + (size_t)e * NG, out_scale, G + (size_t)a * NG, NULL,
                               tag[s] == 0 ? dYdP + param_off[e] : NULL);
      if (E_atom) E_atom[a] = Ea;
      output += Ea;
    }
    double Error = (output + out_mean) - target[s];
    E_struct[s] = output + out_mean;
    if (tag[s] == 0) {
      *SSE += Error * Error;
      for (int i = 0; i < P; i++) gradient[i] += 2 * Error * dYdP[i];
    } else {
      *vSSE += Error * Error;
      (*nval)++;
    }
  }
  free(dYdP);
}

/* ------------------------------------------------------------------------------------------------ */
/* Potential::Precondition input part (Potential.cpp:404-455) with Structure::mean/variance (Structure.cpp:1443-1476):
 * per element over atoms of 'train' structures: mean_k, scale_k = max(sqrt(sum (G-mean)^2 / count), 1e-5). */
void orc_precondition(int n_struct, const int* atom_offset, const int* elem, const double* G, int NG, const int* tag,
                      int n_elem, double* mean_all, double* scale_all)
{
  for (int e = 0; e < n_elem; e++) {
    double* mean = mean_all + (size_t)e * NG;
    double* var = scale_all + (size_t)e * NG;
    int count = 0;
    for (int k = 0; k < NG; k++) mean[k] = 0.0;
    for (int s = 0; s < n_struct; s++) {
      if (tag[s] != 0) continue;
      /* per-structure partial sums first, like my_sums (Structure.cpp:884-892), then summed over structures */
      double* part = (double*)calloc(NG, sizeof(double));
      for (int a = atom_offset[s]; a < atom_offset[s + 1]; a++)
        if (elem[a] == e) { count++; for (int k = 0; k < NG; k++) part[k] += G[(size_t)a * NG + k]; }
      for (int k = 0; k < NG; k++) mean[k] += part[k];
      free(part);
    }
    if (count != 0) for (int k = 0; k < NG; k++) mean[k] /= count;
    for (int k = 0; k < NG; k++) var[k] = 0.0;
    for (int s = 0; s < n_struct; s++) {
      if (tag[s] != 0) continue;
      double* part = (double*)calloc(NG, sizeof(double));
      for (int a = atom_offset[s]; a < atom_offset[s + 1]; a++)
        if (elem[a] == e) for (int k = 0; k < NG; k++) part[k] += pow(G[(size_t)a * NG + k] - mean[k], 2);
      for (int k = 0; k < NG; k++) var[k] += part[k];
      free(part);
    }
    if (count != 0) for (int k = 0; k < NG; k++) {
      var[k] = sqrt(var[k] / count);
      if (var[k] < 1e-5) var[k] = 1e-5;
    }
  }
}

/* ------------------------------------------------------------------------------------------------ */
/* Structure::Get_Forces, non-periodic branch (Structure.cpp:429-713).  dE_dG[nlocal][NG]; f[ntotal][3] is
 * accumulated into.  The G4 row-to-row memoisation of :656-698 is restated as the intended math (every row
 * recomputed), which equals the reference for eta-major sorted rows with one Rc. */
void orc_forces_md(const double* x, const int* elem, int nlocal, const long long* offsets, const int* neigh,
                   int n_elem, int nG1, const double* G1, int nG2, const double* G2,
                   int nG3, const double* G3, int nG4, const double* G4, double Rcut,
                   const double* dE_dG, double* f)
{
  rows_t R_ = {n_elem, nG1, nG2, nG3, nG4, G1, G2, G3, G4, Rcut};
  const rows_t* rw = &R_;
  int NG = orc_num_descriptors(n_elem, nG1, nG2, nG3, nG4);
  for (int i = 0; i < nlocal; i++) {
    const double* w = dE_dG + (size_t)i * NG;
    for (long long p = offsets[i]; p < offsets[i + 1]; p++) {
      int j = neigh[p];
      if (j == i) continue;
      double del[3], del2[3], del3[3];
      for (int z = 0; z < 3; z++) del[z] = x[3 * j + z] - x[3 * i + z];
      double R = sqrt(del[0] * del[0] + del[1] * del[1] + del[2] * del[2]);
      if (R > Rcut) continue;
      for (int z = 0; z < 3; z++) del[z] /= R;
      for (int r1 = 0; r1 < nG1; r1++) {
        const double* row = G1 + 2 * r1;
        if (R >= row[0]) continue;
        double force = w[slot_pair(rw, r1, elem[j])] * d_fc(row[0], R, (int)row[1]);
        for (int d = 0; d < 3; d++) { f[3 * i + d] += force * del[d]; f[3 * j + d] -= force * del[d]; }
      }
      for (int r1 = 0; r1 < nG2; r1++) {
        const double* row = G2 + 4 * r1;
        if (R >= row[0]) continue;
        double delta_R = R - row[3];
        double dE_dR = w[slot_pair(rw, r1 + nG1, elem[j])] * exp(-row[2] * delta_R * delta_R) *
                       (d_fc(row[0], R, (int)row[1]) - 2 * row[2] * delta_R * fc(R / row[0], (int)row[1]));
        for (int d = 0; d < 3; d++) { f[3 * i + d] += dE_dR * del[d]; f[3 * j + d] -= dE_dR * del[d]; }
      }
      if (nG3 == 0 && nG4 == 0) continue;
      for (long long p2 = offsets[i]; p2 < offsets[i + 1]; p2++) {
        int j2 = neigh[p2];
        if (j2 == j || j2 == i) continue;
        for (int z = 0; z < 3; z++) { del2[z] = x[3 * j2 + z] - x[3 * i + z]; del3[z] = x[3 * j + z] - x[3 * j2 + z]; }
        double Ru = sqrt(del2[0] * del2[0] + del2[1] * del2[1] + del2[2] * del2[2]);
        double Rjk = sqrt(del3[0] * del3[0] + del3[1] * del3[1] + del3[2] * del3[2]);
        if (Ru > Rcut) continue;
        for (int z = 0; z < 3; z++) { del2[z] /= Ru; del3[z] /= Rjk; }
        double theta = 0;
        for (int z = 0; z < 3; z++) theta += del2[z] * del[z];
        if (nG3) {
          if (Rjk > Rcut) continue;
          for (int r1 = 0; r1 < nG3; r1++) {
            const double* row = G3 + 5 * r1;
            if (R >= row[0] || Ru >= row[0] || Rjk >= row[0]) continue;
            double term = 1 + row[4] * theta;
            if (term <= 0) continue;
            double prefactor = pow(2, 1 - row[3]) * w[slot_trip(rw, r1, elem[j], elem[j2])] * angular_term(theta, row);
            double e1 = exp(-row[2] * (R * R + Ru * Ru + Rjk * Rjk));
            int fs = (int)row[1];
            double fc_R = fc(R / row[0], fs), fc_Ru = fc(Ru / row[0], fs), fc_Rjk = fc(Rjk / row[0], fs);
            double dG_dR = e1 * fc_Ru * fc_Rjk * (d_fc(row[0], R, fs) - 2 * R * row[2] * fc_R);
            double dG_dRu = e1 * fc_R * fc_Rjk * (d_fc(row[0], Ru, fs) - 2 * Ru * row[2] * fc_Ru);
            double dG_dRjk = e1 * fc_R * fc_Ru * (d_fc(row[0], Rjk, fs) - 2 * Rjk * row[2] * fc_Rjk);
            double dG_dcos = e1 * (row[3] * row[4] / term) * fc_R * fc_Ru * fc_Rjk / (R * Ru);
            for (int d = 0; d < 3; d++) {
              double Fjk = prefactor * dG_dRjk * del3[d];
              double Fij = prefactor * (dG_dR * del[d] + Ru * dG_dcos * (del2[d] - theta * del[d]));
              double Fij2 = prefactor * (dG_dRu * del2[d] + R * dG_dcos * (del[d] - theta * del2[d]));
              f[3 * j + d] -= Fij + Fjk;
              f[3 * j2 + d] -= Fij2 - Fjk;
              f[3 * i + d] += Fij + Fij2;
            }
          }
        }
        for (int r1 = 0; r1 < nG4; r1++) {
          const double* row = G4 + 5 * r1;
          if (R >= row[0] || Ru >= row[0]) continue;
          double term = 1 + row[4] * theta;
          if (term <= 0) continue;
          int fs = (int)row[1];
          double prefactor = pow(2, 1 - row[3]) * w[slot_trip(rw, r1 + nG3, elem[j], elem[j2])] * angular_term(theta, row);
          double e1 = exp(-row[2] * (R * R + Ru * Ru));
          double fc_R = fc(R / row[0], fs), fc_Ru = fc(Ru / row[0], fs);
          double dG_dR = e1 * fc_Ru * (d_fc(row[0], R, fs) - 2 * R * row[2] * fc_R);
          double dG_dRu = e1 * fc_R * (d_fc(row[0], Ru, fs) - 2 * Ru * row[2] * fc_Ru);
          double dG_dcos = e1 * (row[3] * row[4] / term) * fc_R * fc_Ru / (R * Ru);
          for (int d = 0; d < 3; d++) {
            double Fij = prefactor * (dG_dR * del[d] + Ru * dG_dcos * (del2[d] - theta * del[d]));
            double Fij2 = prefactor * (dG_dRu * del2[d] + R * dG_dcos * (del[d] - theta * del2[d]));
            f[3 * j + d] -= Fij;
            f[3 * j2 + d] -= Fij2;
            f[3 * i + d] += Fij + Fij2;
          }
        }
      }
    }
  }
}

/* ------------------------------------------------------------------------------------------------ */
/* Optimizer::rprop (Optimizer.cpp:521-593) without the algorithm=default switching branch (:529-547).
 * State carried by the caller between calls (all length P unless noted):
 *   weights (init 1e-4), step (init 0), dir (search_direction; n_dir==0 before the first call),
 *   g_prev1 = gradient of the previous call, g_prev2 = the one before; n_grad = gradients seen incl. current.
 * x is updated in place. */
static int isign(double v) { return (0.0 < v) - (v < 0.0); }

void orc_rprop_step(int P, double* x, const double* g, const double* g_prev1, const double* g_prev2, int n_grad,
                    double* weights, double* step, double* dir, int* dir_ready)
{
  const double wmax = 0.5, wmin = 1000.0 * 2.220446049250313e-16;   /* Optimizer.cpp:285-286 */
  if (!*dir_ready) {
    for (int i = 0; i < P; i++) {
      dir[i] = 1;
      step[i] = -weights[i] * isign(dir[i]);
      x[i] = x[i] + step[i];
    }
    *dir_ready = 1;
    return;
  }
  if (n_grad > 1) {
    for (int i = 0; i < P; i++) {
      int sign_change = (int)(dir[i] * isign(g[i] * g_prev1[i]));
      if (sign_change < 0) {
        dir[i] = 0;
      } else if (sign_change > 0) {
        weights[i] *= 1.2;
        if (weights[i] > wmax && i != P - 1) weights[i] = wmax;
        step[i] = -isign(g[i]) * weights[i];
        x[i] = x[i] + step[i];
      } else if (sign_change == 0 && n_grad >= 3) {
        double D2 = (g[i] - g_prev2[i]) / step[i];
        if (D2 > 0) weights[i] = fabs(g[i]) / D2;
        if (weights[i] < wmin) weights[i] = wmin;
        if (weights[i] > wmax && i != P - 1) weights[i] = wmax;
        step[i] = -isign(g[i]) * weights[i];
        x[i] = x[i] + step[i];
        dir[i] = 1;
      }
    }
  }
}
