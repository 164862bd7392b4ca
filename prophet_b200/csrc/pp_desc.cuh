// Descriptor (G1/G2/G4) and force-chain-rule device code shared by the periodic (training) and the
// LAMMPS-list (MD) kernels.
//
// Replaces Structure::Set_Neighbor (Structure.cpp:1287-1381), Structure::Calc_G (:974-1147) and the
// non-periodic branch of Structure::Get_Forces (:429-713).
//
// Design (B200, FP64-FMA bound -- see DESIGN.md):
//   * one warp per central atom; its neighbour entries (unit vector, R, element, per-radial-key factors
//     A_r(R) = exp(-eta_r R^2) fc_r(R)) are staged in shared memory, sorted by element;
//   * G4 is factorised over a (radial key) x (angular key) grid so that one (j,k) pair costs
//     3 FMA (cos) + a few multiplies (integer powers of 1 + lam c) + NR*NA FMAs, instead of NR*NA exps;
//   * every UNORDERED pair is visited once (the reference visits both orders, Structure.cpp:1027-1066; the
//     factor 2 is applied at the end); lanes own entry a and walk partners b on a circulant schedule
//     b = (a + t) mod m, which balances the triangle and -- in the force kernel -- makes the shared-memory
//     read-modify-write of the partner's force conflict-free without atomics;
//   * accumulators live in registers ([NR_T][NA_T] per lane) and are reduced across the warp once per
//     element-pair block.
#pragma once
#include "pp_common.cuh"
#include "pp_fastmath.cuh"

// ---------------------------------------------------------------------------------------------------------
// per-warp shared-memory views
struct WarpNbr {
  double *ux, *uy, *uz, *R;   // [cap] unit vector i->entry and distance
  double *A;                  // [NR][cap]   radial factors of the G4 grid
  double *dA;                 // [NR][cap]   dA/dR (force kernels only)
  double *fx, *fy, *fz;       // [cap]       force-gradient accumulated on entries by partner lanes (force only)
  double *sx, *sy, *sz, *sR;  // [cap]       staging (pre-sort)
  int *elem, *selem;          // [cap]
  int *idx, *sidx;            // [cap]       neighbour atom index (MD scatter / hook)
  double *buf;                // [max(NG, ...)] G row (desc) or dE/dG row (force)
  double *cell;               // [PP_MAX_RAD*PP_MAX_ANG] reduced grid cells
  int n;                      // entries
  const double* etab;         // 2^(j/64), j = 0..63 (per block, shared memory): table of pp_fast_exp_nonpos
  int* start;                 // [PP_MAX_ELEM + 1] element block boundaries after the sort.  Kept in shared memory:
                              // a run-time indexed array member would push the whole struct (and with it every
                              // pointer above) into local memory and turn all shared loads into generic ones.
};

// per-warp scratch for the cells of one G4 block: NR_T x NA_T (<= 8 x 4) on the general path, NR_T x 2 x NB padded to a
// multiple of 32 (<= 96) on the power-basis path
#define PP_CELL_DOUBLES 128
__host__ __device__ inline size_t pp_warp_smem_bytes(int cap, int NR, int NG, bool force)
{
  size_t d = 0;
  d += 4 * (size_t)cap;                 // ux uy uz R
  d += 4 * (size_t)cap;                 // staging
  d += (size_t)NR * cap;                // A
  if (force) d += (size_t)NR * cap + 3 * (size_t)cap;
  d += (size_t)NG;
  d += PP_CELL_DOUBLES;
  size_t bytes = d * sizeof(double) + 4 * (size_t)cap * sizeof(int) + 8 * sizeof(int);
  return (bytes + 15) & ~(size_t)15;
}

__device__ inline void pp_warp_views(WarpNbr& w, unsigned char* base, int cap, int NR, int NG, bool force)
{
  double* p = reinterpret_cast<double*>(base);
  w.ux = p; p += cap; w.uy = p; p += cap; w.uz = p; p += cap; w.R = p; p += cap;
  w.sx = p; p += cap; w.sy = p; p += cap; w.sz = p; p += cap; w.sR = p; p += cap;
  w.A = p; p += (size_t)NR * cap;
  if (force) { w.dA = p; p += (size_t)NR * cap; w.fx = p; p += cap; w.fy = p; p += cap; w.fz = p; p += cap; }
  else { w.dA = nullptr; w.fx = w.fy = w.fz = nullptr; }
  w.cell = p; p += PP_CELL_DOUBLES;      // first: 16-byte aligned whatever NG is (cap is even), see pp_g4_block_force_basis
  w.buf = p; p += NG;
  int* q = reinterpret_cast<int*>(p);
  w.elem = q; q += cap; w.selem = q; q += cap; w.idx = q; q += cap; w.sidx = q; q += cap;
  w.start = q; q += 8;
  w.n = 0;
  w.etab = nullptr;
}

// the block's table for pp_gauss_exp: call at the top of a kernel, before any thread leaves
__device__ __forceinline__ const double* pp_block_exp_table()
{
  __shared__ double etab[64];
  for (int i = threadIdx.x; i < 64; i += blockDim.x) etab[i] = exp2((double)i * 0.015625);
  __syncthreads();
  return etab;
}

// ---------------------------------------------------------------------------------------------------------
// K1a: periodic image scan for central atom i of one structure, reference arithmetic and order
// (Structure.cpp:1316-1352): for j ascending, for image t ascending:
//   del = (pos[j] + T_t) - pos[i]  (fractional)  ->  To_Cart  ->  distance;  keep 0.01 < d <= Rcut.
// No FMA contraction: the kept set must be bit-identical to the reference's SSE2 build.
// Entries are appended to the staging arrays (sx,sy,sz,sR,selem,sidx) in order; sidx packs j, and the
// image index goes to `timg` when the caller wants it (parity hook).  Returns the count (may exceed cap: then
// nothing beyond cap is stored).
__device__ inline int pp_scan_periodic(WarpNbr& w, int cap, int lane, const double* __restrict__ frac,
                                       const int* __restrict__ elem, int natom, int i, const double* F,
                                       int mx, int my, int mz, const DescSpec* __restrict__ S, int* timg)
{
  const int ny = 2 * my + 1, nz = 2 * mz + 1, nyz = ny * nz;
  const int nT = (2 * mx + 1) * nyz;
  const double ix = frac[3 * i], iy = frac[3 * i + 1], iz = frac[3 * i + 2];
  // d2 thresholds equivalent to the reference's tests on d = sqrt(d2) (correctly rounded sqrt is monotonic):
  //   d <= Rcut  <=>  d2 <= d2_hi ;  d > 0.01  <=>  d2 > d2_lo      (computed on the host, pp_set_descriptors)
  const double d2_hi = S->scan_d2_hi, d2_lo = S->scan_d2_lo;
  const float inv_nyz = 1.0f / (float)nyz, inv_nz = 1.0f / (float)nz;
  // candidate c = j*nT + t, visited 32 at a time in order; (j,t) advanced incrementally (no integer division)
  int j = lane / nT, t = lane - j * nT;
  const int dj = 32 / nT, dt = 32 - dj * nT;
  int count = 0;
  for (;;) {
    bool keep = false;
    double dx = 0, dy = 0, dz = 0, d2 = 0;
    if (j < natom) {
      const int tx = (int)(((float)t + 0.5f) * inv_nyz);
      const int rem = t - tx * nyz;
      const int ty = (int)(((float)rem + 0.5f) * inv_nz);
      const int tz = rem - ty * nz;
      const double fx = __dsub_rn(__dadd_rn(frac[3 * j], (double)(tx - mx)), ix);
      const double fy = __dsub_rn(__dadd_rn(frac[3 * j + 1], (double)(ty - my)), iy);
      const double fz = __dsub_rn(__dadd_rn(frac[3 * j + 2], (double)(tz - mz)), iz);
      dx = __dadd_rn(__dadd_rn(__dmul_rn(F[0], fx), __dmul_rn(F[1], fy)), __dmul_rn(F[2], fz));
      dy = __dadd_rn(__dadd_rn(__dmul_rn(F[3], fx), __dmul_rn(F[4], fy)), __dmul_rn(F[5], fz));
      dz = __dadd_rn(__dadd_rn(__dmul_rn(F[6], fx), __dmul_rn(F[7], fy)), __dmul_rn(F[8], fz));
      d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
      keep = (d2 > d2_lo) && (d2 <= d2_hi);
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      const int pos = count + __popc(m & ((1u << lane) - 1u));
      if (pos < cap) {
        w.sx[pos] = dx; w.sy[pos] = dy; w.sz[pos] = dz; w.sR[pos] = __dsqrt_rn(d2);
        w.selem[pos] = elem[j];
        w.sidx[pos] = j;
        if (timg) timg[pos] = t;
      }
    }
    count += __popc(m);
    // advance by 32 candidates
    j += dj; t += dt;
    if (t >= nT) { t -= nT; j++; }
    if (__shfl_sync(0xffffffffu, j, 0) >= natom) break;   // lane 0 holds the smallest candidate: warp-uniform exit
  }
  return count;
}

// K1b: LAMMPS-list gather (Structure.cpp:1080-1096): skip j == i, skip R > Rcut.
__device__ inline int pp_gather_list(WarpNbr& w, int cap, int lane, const double* __restrict__ x,
                                     const int* __restrict__ elem_of_atom, int i, const int* __restrict__ list,
                                     int nlist, double Rcut)
{
  const double ix = x[3 * (size_t)i], iy = x[3 * (size_t)i + 1], iz = x[3 * (size_t)i + 2];
  int count = 0;
  // four chunks of 32 list entries in flight: all indices first, then all positions, then the tests in list order (the
  // dependent global loads of one chunk at a time left the warp waiting on memory three times per atom)
  constexpr int U = 4;
  for (int p0 = 0; p0 < nlist; p0 += 32 * U) {
    int j[U], el[U];
    double dx[U], dy[U], dz[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int p = p0 + 32 * u + lane;
      j[u] = (p < nlist) ? list[p] : -1;
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const bool ok = j[u] >= 0;
      const size_t jj = ok ? (size_t)j[u] : (size_t)i;
      dx[u] = x[3 * jj]; dy[u] = x[3 * jj + 1]; dz[u] = x[3 * jj + 2];
      el[u] = elem_of_atom[jj];
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (p0 + 32 * u < nlist) {        // warp-uniform
        const double ddx = __dsub_rn(dx[u], ix), ddy = __dsub_rn(dy[u], iy), ddz = __dsub_rn(dz[u], iz);
        const double d = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)), __dmul_rn(ddz, ddz)));
        const bool keep = (j[u] >= 0) && (j[u] != i) && !(d > Rcut);
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) {
          const int pos = count + __popc(m & ((1u << lane) - 1u));
          if (pos < cap) {
            w.sx[pos] = ddx; w.sy[pos] = ddy; w.sz[pos] = ddz; w.sR[pos] = d;
            w.selem[pos] = el[u];
            w.sidx[pos] = j[u];
          }
        }
        count += __popc(m);
      }
    }
  }
  return count;
}

// exp(x) of the Gaussian factors: x = -eta d^2 <= 0 for every sensible row; the table form costs about a third of exp()
__device__ __forceinline__ double pp_gauss_exp(double x, const double* __restrict__ etab)
{
#ifdef PP_DESC_LIBM_EXP
  return exp(x);
#else
  if (x > 0.0) return exp(x);        // negative eta: uniform over the warp, never taken by a physical row set
  return pp_fast_exp_nonpos(x, etab);
#endif
}

// x^e for a small non-negative integer e (square and multiply; e is warp-uniform wherever this is used)
__device__ __forceinline__ double pp_ipow(double x, int e)
{
  double r = (e & 1) ? x : 1.0;
  e >>= 1;
  while (e) {
    x *= x;
    if (e & 1) r *= x;
    e >>= 1;
  }
  return r;
}

// the same for two values (one walk over the exponent's bits)
__device__ __forceinline__ void pp_ipow2(double& x0, double& x1, int e)
{
  double r0 = (e & 1) ? x0 : 1.0, r1 = (e & 1) ? x1 : 1.0;
  e >>= 1;
  while (e) {
    x0 *= x0; x1 *= x1;
    if (e & 1) { r0 *= x0; r1 *= x1; }
    e >>= 1;
  }
  x0 = r0; x1 = r1;
}

// Stable counting sort of the staged entries by element, then unit vectors and radial factors.
template <bool FORCE>
__device__ inline void pp_finalize_entries(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S, int n)
{
  const int n_elem = S->n_elem;
  w.n = n;
  __syncwarp();
  // destination index of every staged entry
  int base = 0;
  for (int e = 0; e < n_elem; e++) {
    if (lane == 0) w.start[e] = base;
    int cnt = 0;
    for (int p0 = 0; p0 < n; p0 += 32) {
      int p = p0 + lane;
      bool is_e = (p < n) && (n_elem == 1 || w.selem[p] == e);
      unsigned m = __ballot_sync(0xffffffffu, is_e);
      if (is_e) {
        int dst = base + cnt + __popc(m & ((1u << lane) - 1u));
        double R = w.sR[p];
        double inv = 1.0 / R;
        w.ux[dst] = w.sx[p] * inv; w.uy[dst] = w.sy[p] * inv; w.uz[dst] = w.sz[p] * inv;
        w.R[dst] = R;
        w.elem[dst] = e;
        w.idx[dst] = w.sidx[p];
      }
      cnt += __popc(m);
    }
    base += cnt;
  }
  if (lane == 0)
    for (int e = n_elem; e <= PP_MAX_ELEM; e++) w.start[e] = base;
  __syncwarp();
  // radial factors of the G4 grid: A_r = exp(-eta_r R^2) fc_r(R)  (0 beyond the row's own Rc: fc returns 0).
  // Two entries per lane at a time (p and p + 32): their exponentials are independent chains of dependent FP64 operations,
  // and a shell of 33..64 neighbours would otherwise walk the keys twice, the second time with half the lanes idle.
  const int NR = S->NR;
  for (int p0 = lane; p0 < n; p0 += 64) {
    const int p1 = p0 + 32;
    const bool two = p1 < n;
    const double R0 = w.R[p0], R1 = two ? w.R[p1] : 1.0;
    double key_rc = -1.0, f0 = 0.0, df0 = 0.0, f1 = 0.0, df1 = 0.0;      // cutoff factor and slope: once per (Rc, fc type), usually one key
    int key_fs = -1;
    double ex0 = 1.0, ex1 = 1.0;
    for (int k = 0; k < NR; k++) {
      const int r = S->rad_order[k], pw = S->rad_chain[k];      // ascending eta; pw: integer power of the previous exponential
      const double rc = S->rad_rc[r], eta = S->rad_eta[r];
      const int fs = S->rad_fc[r];
      if (pw) pp_ipow2(ex0, ex1, pw);
      else { ex0 = pp_gauss_exp(-eta * (R0 * R0), w.etab); ex1 = pp_gauss_exp(-eta * (R1 * R1), w.etab); }
      if (rc != key_rc || fs != key_fs) {
        key_rc = rc; key_fs = fs;
        f0 = pp_fc(R0 / rc, fs); f1 = pp_fc(R1 / rc, fs);
        if (FORCE) { df0 = pp_dfc(rc, R0, fs); df1 = pp_dfc(rc, R1, fs); }
      }
      w.A[(size_t)r * cap + p0] = ex0 * f0;
      if (FORCE) w.dA[(size_t)r * cap + p0] = ex0 * (df0 - 2.0 * R0 * eta * f0);
      if (two) {
        w.A[(size_t)r * cap + p1] = ex1 * f1;
        if (FORCE) w.dA[(size_t)r * cap + p1] = ex1 * (df1 - 2.0 * R1 * eta * f1);
      }
    }
    if (FORCE) {      // staging is free now: keep 1/R there
      w.fx[p0] = 0.0; w.fy[p0] = 0.0; w.fz[p0] = 0.0; w.sR[p0] = 1.0 / R0;
      if (two) { w.fx[p1] = 0.0; w.fy[p1] = 0.0; w.fz[p1] = 0.0; w.sR[p1] = 1.0 / R1; }
    }
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------------------
// (1 + lam c)^xi for the angular keys [a0, a0+NA_T) of the grid (Structure.h:251-296).  Raw power, without
// the 2^(1-xi) prefactor.  Integer xi: square-and-multiply with a warp-uniform exponent (no divergence);
// non-integer xi: the reference's binomial series with its 1e-8 stopping rule and pow() fallback.
__device__ inline double pp_series_pow(double th, double xi)
{
  // th = lam*c > -1
  double alpha = xi * th;
  int term = 1;
  double result = 1.0;
  while (fabs(alpha) > 1e-8) {
    result += alpha;
    double tay = term < 300 ? 1.0 / (double)(term + 1) : 0.0;
    alpha *= tay * th * (xi - term);
    ++term;
  }
  if (term >= 250) return pow(1.0 + th, xi);
  return result + alpha;
}

// Angular keys of one grid chunk, hoisted into registers once per block.
// prog[a] (from DescSpec::ang_prog): bits 0-1 mode (0 padding, 1 power-of-two chain, 2 general integer, 3 series),
// bit 2 "fresh" (start from x = 1 + lam c instead of the previous key's power), bits 3-5 number of squarings.
template <int NA_T>
struct AngKeys {
  double lam[NA_T];
  double xi[NA_T];
  int xe[NA_T];
  int prog[NA_T];
};

template <int NA_T>
__device__ __forceinline__ void pp_load_ang(const DescSpec* __restrict__ S, int a0, AngKeys<NA_T>& K)
{
#pragma unroll
  for (int a = 0; a < NA_T; a++) {
    int ga = a0 + a;
    bool ok = ga < S->NA;
    K.lam[a] = ok ? S->ang_lam[ga] : 0.0;
    K.xi[a] = ok ? S->ang_xi[ga] : 0.0;
    K.xe[a] = ok ? S->ang_xi_int[ga] : 0;
    int pr = ok ? S->ang_prog[ga] : 0;
    if (ok && a == 0) pr |= 4;            // a chunk always starts a fresh chain
    K.prog[a] = pr;
  }
}

// p[a] = (1 + lam_a c)^xi_a (0 where 1 + lam c <= 0); optionally dp[a] = d p[a] / dc = xi lam p / (1 + lam c).
template <int NA_T, bool DERIV>
__device__ __forceinline__ void pp_angular(const AngKeys<NA_T>& K, double c, double (&p)[NA_T], double (&dp)[NA_T])
{
  double run = 0.0, inv = 0.0;
#pragma unroll
  for (int a = 0; a < NA_T; a++) {
    const int pr = K.prog[a];
    const int mode = pr & 3;
    double v = 0.0, d = 0.0;
    if (mode == 1) {
      if (pr & 4) {
        run = fma(K.lam[a], c, 1.0);
        run = run > 0.0 ? run : 0.0;
        if (DERIV) inv = run > 0.0 ? 1.0 / run : 0.0;
        const int xe = K.xe[a];
        // first key of a chain: xi = 2^ns squarings from x
        const int ns = (pr >> 3) & 7;
        if (ns > 0) run *= run;
        if (ns > 1) run *= run;
        if (ns > 2) run *= run;
        if (ns > 3) run *= run;
        if (ns > 4) run *= run;
        (void)xe;
      } else {
        const int ns = (pr >> 3) & 7;
        if (ns > 0) run *= run;
        if (ns > 1) run *= run;
        if (ns > 2) run *= run;
        if (ns > 3) run *= run;
        if (ns > 4) run *= run;
      }
      v = run;
      if (DERIV) d = K.xi[a] * K.lam[a] * v * inv;
    } else if (mode >= 2) {
      const double x = fma(K.lam[a], c, 1.0);
      if (x > 0.0) {
        v = (mode == 2) ? pp_ipow(x, K.xe[a]) : pp_series_pow(K.lam[a] * c, K.xi[a]);
        if (DERIV) d = K.xi[a] * K.lam[a] * v / x;
      }
    }
    p[a] = v;
    if (DERIV) dp[a] = d;
  }
}

// ---------------------------------------------------------------------------------------------------------
// G1/G2 rows of one central atom (Structure.cpp:1005-1014 / 1097-1106), accumulated into w.buf[slot].
// Rows in the order of DescSpec::pair_order, one at a time, two entries per lane together (independent chains).  The
// Gaussian of a row whose eta is an integer multiple of its predecessor's is a power of the predecessor's, kept per
// neighbour in the staging array sy (each lane reads back its own entries).  Tried and dropped: eight rows per sweep with
// eight partial sums reduced together (the unrolled sweep ran 20 % slower than this form).
__device__ inline void pp_pair_rows_desc(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S)
{
  const int npair = S->npair, nG1 = S->nG1;
  for (int e = 0; e < S->n_elem; e++) {
    const int s = w.start[e], m = w.start[e + 1] - s;
    double key_rc = -1.0;
    int key_fs = -1;
    for (int k = 0; k < npair; k++) {
      const int r = S->pair_order[k], pw = S->pair_chain[k];
      const double rc = S->pair_rc[r], eta = S->pair_eta[r], rs = S->pair_rs[r];
      const int fs = S->pair_fc[r];
      const bool g1 = r < nG1;
      if (rc != key_rc || fs != key_fs) {
        key_rc = rc; key_fs = fs;
        for (int a = lane; a < m; a += 32) w.sx[s + a] = pp_fc(w.R[s + a] / rc, fs);
      }
      double acc = 0.0;
      for (int a0 = lane; a0 < m; a0 += 64) {
        const int a1 = a0 + 32;
        const bool two = a1 < m;
        double t0 = w.sx[s + a0], t1 = two ? w.sx[s + a1] : 0.0;
        if (!g1) {
          double ex0, ex1;
          if (pw) { ex0 = w.sy[s + a0]; ex1 = two ? w.sy[s + a1] : 1.0; pp_ipow2(ex0, ex1, pw); }
          else {
            const double d0 = w.R[s + a0] - rs, d1 = (two ? w.R[s + a1] : 0.0) - rs;
            ex0 = pp_gauss_exp(-eta * (d0 * d0), w.etab); ex1 = pp_gauss_exp(-eta * (d1 * d1), w.etab);
          }
          w.sy[s + a0] = ex0;
          if (two) w.sy[s + a1] = ex1;
          t0 *= ex0; t1 *= ex1;
        }
        acc += t0 + t1;
      }
      acc = pp_warp_sum(acc);
      if (lane == 0) w.buf[npair * e + r] = acc;
    }
  }
  __syncwarp();
}
// G4 rows for the element-pair block (e1 <= e2), grid chunk [r0,r0+NR_T) x [a0,a0+NA_T).
// skip_rjk: the non-periodic branch skips the pair when G3 rows exist and Rjk > Rcut (Structure.cpp:1129-1132).
template <int NR_T, int NA_T>
__device__ inline void pp_g4_block_desc(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S,
                                        int e1, int e2, int r0, int a0, bool skip_rjk)
{
  const int s1 = w.start[e1], m1 = w.start[e1 + 1] - s1;
  const int s2 = w.start[e2], m2 = w.start[e2 + 1] - s2;
  const bool same = (e1 == e2);
  const double rcut2 = S->Rcut * S->Rcut;
  double acc[NR_T][NA_T];
#pragma unroll
  for (int r = 0; r < NR_T; r++)
#pragma unroll
    for (int a = 0; a < NA_T; a++) acc[r][a] = 0.0;
  AngKeys<NA_T> K;
  pp_load_ang<NA_T>(S, a0, K);
  const int NR = S->NR;
  const double* Arow[NR_T];     // padded radial keys alias the last valid row; their own-entry factor is 0
#pragma unroll
  for (int r = 0; r < NR_T; r++) Arow[r] = w.A + (size_t)min(r0 + r, NR - 1) * cap;

  for (int a = lane; a < m1; a += 32) {
    const int ja = s1 + a;
    const double ax = w.ux[ja], ay = w.uy[ja], az = w.uz[ja], Ra = w.R[ja];
    double Aa[NR_T];
#pragma unroll
    for (int r = 0; r < NR_T; r++) Aa[r] = (r0 + r < NR) ? Arow[r][ja] : 0.0;
    // partner schedule: same block -> circulant half ring; cross block -> every entry of block 2
    int nt = same ? ((m1 - 1) >> 1) + (((m1 & 1) == 0 && a < (m1 >> 1)) ? 1 : 0) : m2;
    for (int t = 0; t < nt; t++) {
      int jb;
      if (same) { int b = a + t + 1; if (b >= m1) b -= m1; jb = s1 + b; }
      else jb = s2 + t;
      double c = ax * w.ux[jb] + ay * w.uy[jb] + az * w.uz[jb];
      if (skip_rjk) {
        double Rb = w.R[jb];
        double rjk2 = Ra * Ra + Rb * Rb - 2.0 * Ra * Rb * c;
        if (rjk2 > rcut2) continue;
      }
      double p[NA_T];
      pp_angular<NA_T, false>(K, c, p, p);
#pragma unroll
      for (int r = 0; r < NR_T; r++) {
        double tr = Aa[r] * Arow[r][jb];
#pragma unroll
        for (int q = 0; q < NA_T; q++) acc[r][q] = fma(p[q], tr, acc[r][q]);
      }
    }
  }
  // warp reduction of the grid cells; factor 2 = both (j,k) orders of the reference
#pragma unroll
  for (int r = 0; r < NR_T; r++)
#pragma unroll
    for (int q = 0; q < NA_T; q++) {
      double v = pp_warp_sum(acc[r][q]);
      if (lane == 0) w.cell[r * NA_T + q] = 2.0 * v;
    }
  __syncwarp();
  const int base = S->trip_base[e1][e2] + S->nG3;
  for (int row = lane; row < S->nG4; row += 32) {
    int rr = S->g4_rad[row] - r0, qq = S->g4_ang[row] - a0;
    if (rr >= 0 && rr < NR_T && qq >= 0 && qq < NA_T) w.buf[base + row] = S->ang_pref[S->g4_ang[row]] * w.cell[rr * NA_T + qq];
  }
  __syncwarp();
}


// ---------------------------------------------------------------------------------------------------------
// Power-basis fast path for the G4 grid.
//
// When every angular key has an integer power-of-two xi = 2^b (b < NB) and at most two distinct lambdas occur
// (the usual Behler sets: xi in {1,2,4,8,16}, lambda = +-1), the angular factors are members of the small basis
//     pw_g[b] = max(1 + lam_g c, 0)^(2^b),   g in {0,1}, b in [0, NB)
// obtained by NB-1 successive squarings per lambda.  Accumulating on the basis, acc[r][g][b] += pw_g[b] * A_r A_r',
// keeps the pair loop free of branches, selects and run-time exponents: 3 FMA (cos) + 2 FMA + 2(NB-1) squarings +
// NR multiplies + 2 NR NB FMAs.  Rows pick their (r, g, b) slot once per atom at the end.  Basis slots no row
// uses are wasted FMAs (level b = 3 for xi in {1,2,4,16}), which is far cheaper than evaluating powers per key.

// Sum `v[i]` over the warp for all i at once by recursive halving: after the call lane l holds in v[0 .. NPAD/32)
// the warp totals of the slots (NPAD/32)*l + p.  NPAD/2 + NPAD/4 + ... shuffles instead of 5 per value.
template <int NPAD>
__device__ __forceinline__ void pp_reduce_scatter(double (&v)[NPAD], int lane)
{
  static_assert(NPAD == 32 || NPAD == 64 || NPAD == 128, "NPAD must be 32, 64 or 128");
#pragma unroll
  for (int step = 0; step < 5; step++) {
    const int h = NPAD >> (step + 1);
    const int m = 16 >> step;
    const bool up = (lane & m) != 0;
#pragma unroll
    for (int i = 0; i < h; i++) {
      const double send = up ? v[i] : v[i + h];
      const double keep = up ? v[i + h] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
    }
  }
}

// NBX = number of power levels NB, or 6 = five levels of which level 3 (xi = 8) is used by no row: the usual
// xi in {1, 2, 4, 16} set.  That level is still squared through but gets no accumulator and no weight.
template <int NBX> struct PowerLevels {
  static constexpr int NB = NBX == 6 ? 5 : NBX;         // levels in the squaring chain (and in the cell layout)
  static constexpr int NU = NBX == 6 ? 4 : NBX;         // levels that carry rows
  __host__ __device__ static constexpr bool used(int k) { return !(NBX == 6 && k == 3); }
  __host__ __device__ static constexpr int slot(int k) { return (NBX == 6 && k == 4) ? 3 : k; }      // level -> compact index
  __host__ __device__ static constexpr int level(int u) { return (NBX == 6 && u == 3) ? 4 : u; }     // compact index -> level
};

template <int NR_T, int NBX>
__device__ inline void pp_g4_block_desc_basis(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S,
                                              int e1, int e2, int r0, bool skip_rjk)
{
  // One pass over the pairs per lambda group: NR_T * NU accumulators live instead of twice that many (the pair geometry
  // is cheap to redo), which is what lets a fourth block onto the SM.
  using PL = PowerLevels<NBX>;
  constexpr int NB = PL::NB, NU = PL::NU;
  constexpr int NACC = NR_T * NU;
  constexpr int NPAD = NACC <= 32 ? 32 : 64;
  const int s1 = w.start[e1], m1 = w.start[e1 + 1] - s1;
  const int s2 = w.start[e2], m2 = w.start[e2 + 1] - s2;
  const bool same = (e1 == e2);
  const double rcut2 = S->Rcut * S->Rcut;
  const int NR = S->NR;
  const double* Arow[NR_T];
#pragma unroll
  for (int r = 0; r < NR_T; r++) Arow[r] = w.A + (size_t)min(r0 + r, NR - 1) * cap;

#pragma unroll 1
  for (int gq = 0; gq < 2; gq++) {
    const double lam = S->basis_lam[gq];
    double acc[NPAD];
#pragma unroll
    for (int i = 0; i < NPAD; i++) acc[i] = 0.0;
    for (int a0 = 0; a0 < m1; a0 += 32) {
      // Lanes own neighbour a and walk its partners b = a+1, a+2, ... (each unordered pair once).  A pass with at most 16
      // neighbours left (the tail of a 33..48-neighbour shell) would idle half the warp, so its partner range is split
      // over up to four lane groups that share the same a.
      const int na = min(32, m1 - a0);
      int G = 1, grp = 0, sub = lane;
      if (same && na <= 16) { G = min(4, 32 / na); grp = lane / na; sub = lane - grp * na; }
      const bool act = (G == 1) ? (sub < na) : (grp < G);
      const int a = a0 + (act ? sub : 0);
      const int ja = s1 + a;
      const double ax = w.ux[ja], ay = w.uy[ja], az = w.uz[ja], Ra = w.R[ja];
      double Aa[NR_T];
#pragma unroll
      for (int r = 0; r < NR_T; r++) Aa[r] = (r0 + r < NR) ? Arow[r][ja] : 0.0;
      const int nt = same ? ((m1 - 1) >> 1) + (((m1 & 1) == 0 && a < (m1 >> 1)) ? 1 : 0) : m2;
      const int chunk = ((same ? (m1 >> 1) : m2) + G - 1) / G;
      const int t0 = grp * chunk;
      const int t1 = act ? min(nt, t0 + chunk) : t0;
      const int sb = same ? s1 : s2, mb = same ? m1 : m2;
      int b = same ? (a + 1 + t0) % m1 : 0;
      for (int t = t0; t < t1; t++) {
        const int jb = sb + b;
        b++;
        if (b >= mb) b -= mb;
        const double c = ax * w.ux[jb] + ay * w.uy[jb] + az * w.uz[jb];
        if (skip_rjk) {
          const double Rb = w.R[jb];
          const double rjk2 = Ra * Ra + Rb * Rb - 2.0 * Ra * Rb * c;
          if (rjk2 > rcut2) continue;
        }
        double pw[NB];
        pw[0] = fmax(fma(lam, c, 1.0), 0.0);
#pragma unroll
        for (int k = 1; k < NB; k++) pw[k] = pw[k - 1] * pw[k - 1];
#pragma unroll
        for (int r = 0; r < NR_T; r++) {
          const double tr = Aa[r] * Arow[r][jb];
#pragma unroll
          for (int u = 0; u < NU; u++) acc[r * NU + u] = fma(pw[PL::level(u)], tr, acc[r * NU + u]);
        }
      }
    }
    pp_reduce_scatter<NPAD>(acc, lane);
    constexpr int PER = NPAD / 32;
#pragma unroll
    for (int p = 0; p < PER; p++) {
      const int idx = PER * lane + p;                     // = r * NU + u
      if (idx < NACC) w.cell[((idx / NU) * 2 + gq) * NB + PL::level(idx % NU)] = 2.0 * acc[p];     // factor 2 = both (j,k) orders of the reference
    }
  }
  __syncwarp();
  const int base = S->trip_base[e1][e2] + S->nG3;
  for (int row = lane; row < S->nG4; row += 32) {
    const int rr = S->g4_rad[row] - r0;
    if (rr >= 0 && rr < NR_T) w.buf[base + row] = S->ang_pref[S->g4_ang[row]] * w.cell[(rr * 2 + S->g4_bg[row]) * NB + S->g4_bb[row]];
  }
  __syncwarp();
}

// Force chain rule on the power basis.  W[r][g][b] = 2 * 2^(1-xi) * dE/dG(row) at the row's slot (0 elsewhere).
//   d/dc pw_g[b] = 2^b lam_g x^(2^b - 1) = 2^b lam_g * prod_{k<b} pw_g[k]        (no division)
template <int NR_T, int NBX>
__device__ inline void pp_g4_block_force_basis(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S,
                                               int e1, int e2, int r0, bool skip_rjk)
{
  using PL = PowerLevels<NBX>;
  constexpr int NB = PL::NB;
  constexpr int NW = NR_T * 2 * NB;
  const int s1 = w.start[e1], m1 = w.start[e1 + 1] - s1;
  const int s2 = w.start[e2], m2 = w.start[e2 + 1] - s2;
  if (m1 == 0 || m2 == 0) return;
  const bool same = (e1 == e2);
  const double rcut2 = S->Rcut * S->Rcut;
  const double lam0 = S->basis_lam[0], lam1 = S->basis_lam[1];
  for (int i = lane; i < NW; i += 32) w.cell[i] = 0.0;
  __syncwarp();
  const int base = S->trip_base[e1][e2] + S->nG3;
  // weights dE/dG of the basis cells, layout [k][r][g]: the two lambda groups of a (power, radial key) cell sit side by
  // side, and the pair loop fetches each from shared memory when it needs it.  Keeping all
  // NR_T * 2 * NB of them in registers instead costs 100 registers and caps the kernel at 8 warps per SM.
  for (int row = lane; row < S->nG4; row += 32) {
    const int rr = S->g4_rad[row] - r0;
    if (rr >= 0 && rr < NR_T)
      atomicAdd(&w.cell[(S->g4_bb[row] * NR_T + rr) * 2 + S->g4_bg[row]], 2.0 * S->ang_pref[S->g4_ang[row]] * w.buf[base + row]);
  }
  __syncwarp();
  const unsigned cell_s = (unsigned)__cvta_generic_to_shared(w.cell);
  auto Wld = [&](int k, int r, int gq) {      // volatile: must not be hoisted back into registers
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(cell_s + 8u * (unsigned)((k * NR_T + r) * 2 + gq)));
    return v;
  };
  const int NR = S->NR;
  const double *Arow[NR_T], *dArow[NR_T];
#pragma unroll
  for (int r = 0; r < NR_T; r++) {
    Arow[r] = w.A + (size_t)min(r0 + r, NR - 1) * cap;
    dArow[r] = w.dA + (size_t)min(r0 + r, NR - 1) * cap;
  }
  const int width = same ? 32 : (m2 < 32 ? m2 : 32);
  const int npass = (m1 + width - 1) / width;
  for (int pass = 0; pass < npass; pass++) {
    // Lanes own neighbour a and walk its partners; in one step the lanes' partners are distinct, so the shared-memory
    // force accumulators need no atomics.  A pass with at most 16 neighbours left (the tail of a 33..48-neighbour shell)
    // would idle half the warp: its partner range is split over up to four lane groups that share the same a, in chunks
    // at least `na` long so that the groups' partner windows stay disjoint.
    const int na = min(width, m1 - pass * width);
    int G = 1, grp = 0, sub = lane;
    if (same && na <= 16) { G = min(4, 32 / na); grp = lane / na; sub = lane - grp * na; }
    const int a = pass * width + sub;
    const bool act = (G == 1) ? ((lane < width) && (a < m1)) : (grp < G);
    const int ja = s1 + (act ? a : 0);
    const double ax = w.ux[ja], ay = w.uy[ja], az = w.uz[ja], Ra = w.R[ja];
    const double invRa = 1.0 / Ra;
    double Aa[NR_T], dAa[NR_T];
#pragma unroll
    for (int r = 0; r < NR_T; r++) {
      const bool ok = (r0 + r < NR);
      Aa[r] = ok ? Arow[r][ja] : 0.0;
      dAa[r] = ok ? dArow[r][ja] : 0.0;
    }
    double gx = 0.0, gy = 0.0, gz = 0.0;
    const int nt_all = same ? (m1 >> 1) : m2;
    const int chunk = (G == 1) ? nt_all : max((nt_all + G - 1) / G, na);
    const int nt_max = min(chunk, nt_all);
    const int t0 = grp * chunk;
    // this lane's own step count within its chunk
    const int nt_a = same ? ((m1 - 1) >> 1) + (((m1 & 1) == 0 && a < (m1 >> 1)) ? 1 : 0) : m2;
    const int nt = min(nt_a - t0, chunk);
    const int sb = same ? s1 : s2, mb = same ? m1 : m2;
    int b = same ? (a + 1 + t0) % m1 : lane;
    if (b >= mb) b -= mb;
    for (int t = 0; t < nt_max; t++) {
      if (act && t < nt) {
        const int jb = sb + b;
        const double bx = w.ux[jb], by = w.uy[jb], bz = w.uz[jb], Rb = w.R[jb];
        const double c = ax * bx + ay * by + az * bz;
        bool skip = false;
        if (skip_rjk) { const double rjk2 = Ra * Ra + Rb * Rb - 2.0 * Ra * Rb * c; skip = rjk2 > rcut2; }
        if (!skip) {
          // S_r = sum_{g,k} W[k][r][g] p_g^(2^k), T_r the same with d/dc of the powers; the two lambda groups one after
          // the other, the powers produced on the fly (a chain of squarings and its prefix products): 2 NR_T accumulators
          // live instead of 4 NR_T plus 4 NB powers, which is what lets a fourth block onto the SM.  (Both groups in step, so
          // that one 16-byte load of a cell's two weights feeds four FMAs, was measured: 20 instead of 40 shared-memory loads
          // per pair, 70 more bytes of spills, 3 % slower.)
          double Ab[NR_T], dAb[NR_T], Ss[NR_T], Ts[NR_T];
#pragma unroll
          for (int r = 0; r < NR_T; r++) { Ab[r] = Arow[r][jb]; dAb[r] = dArow[r][jb]; Ss[r] = 0.0; Ts[r] = 0.0; }
#pragma unroll
          for (int gq = 0; gq < 2; gq++) {
            const double lam = gq ? lam1 : lam0;
            double pk = fmax(fma(lam, c, 1.0), 0.0);          // p^(2^k)
            double pp = pk > 0.0 ? lam : 0.0;                 // lam * prod_{j<k} p^(2^j)
            double dk = pp;                                   // d/dc p^(2^k) = 2^k * pp
#pragma unroll
            for (int k = 0; k < NB; k++) {
              if (PL::used(k)) {                              // compile-time: an unused level has zero weights
#pragma unroll
                for (int r = 0; r < NR_T; r++) {
                  const double wv = Wld(k, r, gq);
                  Ss[r] = fma(wv, pk, Ss[r]);
                  Ts[r] = fma(wv, dk, Ts[r]);
                }
              }
              if (k + 1 < NB) {
                pp *= pk;
                pk *= pk;
                dk = (double)(2 << k) * pp;
              }
            }
          }
          double dRa = 0.0, dRb = 0.0, dc = 0.0;
#pragma unroll
          for (int r = 0; r < NR_T; r++) {
            const double Sr = Ss[r], Tr = Ts[r];
            dRa = fma(Sr * dAa[r], Ab[r], dRa);
            dRb = fma(Sr * Aa[r], dAb[r], dRb);
            dc = fma(Tr * Aa[r], Ab[r], dc);
          }
          const double ca = dc * invRa, cb = dc * w.sR[jb];          // sR holds 1/R after pp_finalize_entries
          gx += dRa * ax + ca * (bx - c * ax);
          gy += dRa * ay + ca * (by - c * ay);
          gz += dRa * az + ca * (bz - c * az);
          w.fx[jb] += dRb * bx + cb * (ax - c * bx);
          w.fy[jb] += dRb * by + cb * (ay - c * by);
          w.fz[jb] += dRb * bz + cb * (az - c * bz);
        }
      }
      b++;
      if (b >= mb) b -= mb;
      __syncwarp();
    }
    for (int g = 1; g < G; g++) {                     // gather the groups' partial sums for a into group 0
      const int src = (sub + g * na) & 31;
      const double ox = __shfl_sync(0xffffffffu, gx, src), oy = __shfl_sync(0xffffffffu, gy, src),
                   oz = __shfl_sync(0xffffffffu, gz, src);
      if (grp == 0) { gx += ox; gy += oy; gz += oz; }
    }
    if (act && grp == 0) { w.fx[ja] += gx; w.fy[ja] += gy; w.fz[ja] += gz; }
    __syncwarp();
  }
}


// ---------------------------------------------------------------------------------------------------------
// G3 rows (Structure.h:361-367): 2^(1-xi) (1+lam c)^xi exp(-eta (R^2+Ru^2+Rjk^2)) fc(R) fc(Ru) fc(Rjk).
// The Rjk dependence rules out the factorisation used for G4, so each pair evaluates each row directly (exp, pow,
// three cutoffs).  Kept out of line: potentials without G3 rows (every BASELINE config) pay nothing for it.
__device__ __forceinline__ double pp_rjk(double ax, double ay, double az, double Ra, double bx, double by, double bz, double Rb,
                                          double& dx, double& dy, double& dz)
{
  dx = Ra * ax - Rb * bx; dy = Ra * ay - Rb * by; dz = Ra * az - Rb * bz;     // r_j - r_j2
  return sqrt(dx * dx + dy * dy + dz * dz);
}

__device__ __noinline__ void pp_g3_block_desc(WarpNbr w, int cap, int lane, const DescSpec* __restrict__ S, int e1, int e2,
                                              bool nonperiodic)
{
  const int s1 = w.start[e1], m1 = w.start[e1 + 1] - s1;
  const int s2 = w.start[e2], m2 = w.start[e2 + 1] - s2;
  const bool same = (e1 == e2);
  const int nG3 = S->nG3;
  double acc[PP_MAX_G3_ROWS];
#pragma unroll
  for (int r = 0; r < PP_MAX_G3_ROWS; r++) acc[r] = 0.0;
  for (int a = lane; a < m1; a += 32) {
    const int ja = s1 + a;
    const double ax = w.ux[ja], ay = w.uy[ja], az = w.uz[ja], Ra = w.R[ja];
    const int nt = same ? ((m1 - 1) >> 1) + (((m1 & 1) == 0 && a < (m1 >> 1)) ? 1 : 0) : m2;
    for (int t = 0; t < nt; t++) {
      int jb;
      if (same) { int b = a + t + 1; if (b >= m1) b -= m1; jb = s1 + b; }
      else jb = s2 + t;
      const double bx = w.ux[jb], by = w.uy[jb], bz = w.uz[jb], Rb = w.R[jb];
      const double c = ax * bx + ay * by + az * bz;
      double dx, dy, dz;
      const double Rjk = pp_rjk(ax, ay, az, Ra, bx, by, bz, Rb, dx, dy, dz);
      if (nonperiodic ? (Rjk > S->Rcut) : (Rjk < 0.01)) continue;          // Structure.cpp:1129-1132 / :1048-1050
#pragma unroll
      for (int r = 0; r < PP_MAX_G3_ROWS; r++) {
        if (r < nG3) {
          const double term = 1.0 + S->g3_lam[r] * c;
          if (term > 0.0) {
            const double rc = S->g3_rc[r];
            const int fs = S->g3_fc[r];
            acc[r] += pow(term, S->g3_xi[r]) * exp(-S->g3_eta[r] * (Ra * Ra + Rb * Rb + Rjk * Rjk)) * pp_fc(Ra / rc, fs) *
                      pp_fc(Rb / rc, fs) * pp_fc(Rjk / rc, fs);
          }
        }
      }
    }
  }
  const int base = S->trip_base[e1][e2];
#pragma unroll
  for (int r = 0; r < PP_MAX_G3_ROWS; r++) {
    if (r < nG3) {
      const double v = pp_warp_sum(acc[r]);
      if (lane == 0) w.buf[base + r] = 2.0 * pow(2.0, 1.0 - S->g3_xi[r]) * v;
    }
  }
  __syncwarp();
}

// angular_term of the reference (Structure.h:251-279) for one row: integer xi by squaring, else the series
__device__ __forceinline__ double pp_angular_term(double c, double lam, double xi)
{
  const double th = lam * c;
  if (th <= -1.0) return 0.0;
  if (xi == (double)(int)xi) return pp_ipow(xi >= 0 ? 1.0 + th : 1.0 / (1.0 + th), abs((int)xi));
  return pp_series_pow(th, xi);
}

__device__ __noinline__ void pp_g3_block_force(WarpNbr w, int cap, int lane, const DescSpec* __restrict__ S, int e1, int e2)
{
  // Structure.cpp:510-607 (non-periodic branch), unordered pairs x 2 on the circulant schedule of the G4 kernel
  const int s1 = w.start[e1], m1 = w.start[e1 + 1] - s1;
  const int s2 = w.start[e2], m2 = w.start[e2 + 1] - s2;
  if (m1 == 0 || m2 == 0) return;
  const bool same = (e1 == e2);
  const int nG3 = S->nG3;
  const int base = S->trip_base[e1][e2];
  const int width = same ? 32 : (m2 < 32 ? m2 : 32);
  const int npass = (m1 + width - 1) / width;
  const int nt_max = same ? (m1 >> 1) : m2;
  for (int pass = 0; pass < npass; pass++) {
    const int a = pass * width + lane;
    const bool act = (lane < width) && (a < m1);
    const int ja = s1 + (act ? a : 0);
    const double ax = w.ux[ja], ay = w.uy[ja], az = w.uz[ja], Ra = w.R[ja];
    double gx = 0.0, gy = 0.0, gz = 0.0;
    const int nt = same ? ((m1 - 1) >> 1) + (((m1 & 1) == 0 && a < (m1 >> 1)) ? 1 : 0) : m2;
    const int sb = same ? s1 : s2, mb = same ? m1 : m2;
    int b = same ? a + 1 : lane;
    if (b >= mb) b -= mb;
    for (int t = 0; t < nt_max; t++) {
      if (act && t < nt) {
        const int jb = sb + b;
        const double bx = w.ux[jb], by = w.uy[jb], bz = w.uz[jb], Rb = w.R[jb];
        const double c = ax * bx + ay * by + az * bz;
        double d3x, d3y, d3z;
        const double Rjk = pp_rjk(ax, ay, az, Ra, bx, by, bz, Rb, d3x, d3y, d3z);
        if (!(Rjk > S->Rcut)) {
          const double irjk = 1.0 / Rjk;
          d3x *= irjk; d3y *= irjk; d3z *= irjk;                    // unit vector j2 -> j
          double fax = 0, fay = 0, faz = 0, fbx = 0, fby = 0, fbz = 0;
          for (int r = 0; r < nG3; r++) {
            const double rc = S->g3_rc[r];
            if (Ra >= rc || Rb >= rc || Rjk >= rc) continue;
            const double lam = S->g3_lam[r], xi = S->g3_xi[r], eta = S->g3_eta[r];
            const int fs = S->g3_fc[r];
            const double term = 1.0 + lam * c;
            if (term <= 0.0) continue;
            const double pref = 2.0 * pow(2.0, 1.0 - xi) * w.buf[base + r] * pp_angular_term(c, lam, xi);   // x2: both orders
            const double ex = exp(-eta * (Ra * Ra + Rb * Rb + Rjk * Rjk));
            const double fR = pp_fc(Ra / rc, fs), fU = pp_fc(Rb / rc, fs), fJ = pp_fc(Rjk / rc, fs);
            const double dG_dR = ex * fU * fJ * (pp_dfc(rc, Ra, fs) - 2.0 * Ra * eta * fR);
            const double dG_dRu = ex * fR * fJ * (pp_dfc(rc, Rb, fs) - 2.0 * Rb * eta * fU);
            const double dG_dRjk = ex * fR * fU * (pp_dfc(rc, Rjk, fs) - 2.0 * Rjk * eta * fJ);
            const double dG_dcos = ex * (xi * lam / term) * fR * fU * fJ / (Ra * Rb);
            const double Fjkx = pref * dG_dRjk * d3x, Fjky = pref * dG_dRjk * d3y, Fjkz = pref * dG_dRjk * d3z;
            fax += pref * (dG_dR * ax + Rb * dG_dcos * (bx - c * ax)) + Fjkx;
            fay += pref * (dG_dR * ay + Rb * dG_dcos * (by - c * ay)) + Fjky;
            faz += pref * (dG_dR * az + Rb * dG_dcos * (bz - c * az)) + Fjkz;
            fbx += pref * (dG_dRu * bx + Ra * dG_dcos * (ax - c * bx)) - Fjkx;
            fby += pref * (dG_dRu * by + Ra * dG_dcos * (ay - c * by)) - Fjky;
            fbz += pref * (dG_dRu * bz + Ra * dG_dcos * (az - c * bz)) - Fjkz;
          }
          gx += fax; gy += fay; gz += faz;
          w.fx[jb] += fbx; w.fy[jb] += fby; w.fz[jb] += fbz;
        }
      }
      b++;
      if (b >= mb) b -= mb;
      __syncwarp();
    }
    if (act) { w.fx[ja] += gx; w.fy[ja] += gy; w.fz[ja] += gz; }
    __syncwarp();
  }
}

// NX is the angular register tile NA_T on the general path and the power-level code NBX (PowerLevels) on the basis path.
template <int NR_T, int NX, bool BASIS>
__device__ inline void pp_descriptor_row(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S, bool skip_rjk)
{
  for (int g = lane; g < S->NG; g += 32) w.buf[g] = 0.0;
  __syncwarp();
#ifndef PP_SKIP_PAIR_ROWS      // timing experiments only
  if (S->npair) pp_pair_rows_desc(w, cap, lane, S);
#endif
  if (S->nG3) {
    for (int e1 = 0; e1 < S->n_elem; e1++)
      for (int e2 = e1; e2 < S->n_elem; e2++) pp_g3_block_desc(w, cap, lane, S, e1, e2, skip_rjk);
  }
#ifndef PP_SKIP_G4
  if (S->nG4) {
    for (int e1 = 0; e1 < S->n_elem; e1++)
      for (int e2 = e1; e2 < S->n_elem; e2++)
        for (int r0 = 0; r0 < S->NR; r0 += NR_T) {
          if constexpr (BASIS) pp_g4_block_desc_basis<NR_T, NX>(w, cap, lane, S, e1, e2, r0, skip_rjk);
          else
            for (int a0 = 0; a0 < S->NA; a0 += NX) pp_g4_block_desc<NR_T, NX>(w, cap, lane, S, e1, e2, r0, a0, skip_rjk);
        }
  }
#endif
}

// ---------------------------------------------------------------------------------------------------------
// Force chain rule, G1/G2 rows (Structure.cpp:449-481): gradient of sum_r w_r G_r wrt each entry's position
// is (dE/dR) * u; accumulated into the lane's own entries fx/fy/fz.
__device__ inline void pp_pair_rows_force(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S)
{
  // two entries per lane at a time (p and p + 32): independent chains, and one walk over the rows for a shell of 33..64
  const int npair = S->npair, nG1 = S->nG1;
  for (int p0 = lane; p0 < w.n; p0 += 64) {
    const int p1 = p0 + 32;
    const bool two = p1 < w.n;
    const double R0 = w.R[p0], R1 = two ? w.R[p1] : 1.0;
    const double* wt0 = w.buf + npair * w.elem[p0];
    const double* wt1 = w.buf + npair * (two ? w.elem[p1] : 0);
    double dE0 = 0.0, dE1 = 0.0;
    // cutoff factor and slope depend on (Rc, fc type) only, usually shared by all rows: evaluated when that key changes
    double key_rc = -1.0, f0 = 0.0, df0 = 0.0, f1 = 0.0, df1 = 0.0;
    int key_fs = -1;
    double ex0 = 1.0, ex1 = 1.0;
    for (int k = 0; k < npair; k++) {
      const int r = S->pair_order[k], pw = S->pair_chain[k];      // see pp_pair_rows_desc
      const double rc = S->pair_rc[r], eta = S->pair_eta[r], rs = S->pair_rs[r];
      const int fs = S->pair_fc[r];
      if (rc != key_rc || fs != key_fs) {
        key_rc = rc; key_fs = fs;
        f0 = pp_fc(R0 / rc, fs); df0 = pp_dfc(rc, R0, fs);
        f1 = pp_fc(R1 / rc, fs); df1 = pp_dfc(rc, R1, fs);
      }
      const double w0 = wt0[r], w1 = wt1[r];
      if (r < nG1) {
        if (R0 < rc) dE0 += w0 * df0;
        if (R1 < rc) dE1 += w1 * df1;
      } else {
        const double d0 = R0 - rs, d1 = R1 - rs;
        // also for an entry out of the row's range: a successor in the chain may build on the value
        if (pw) pp_ipow2(ex0, ex1, pw);
        else { ex0 = pp_gauss_exp(-eta * d0 * d0, w.etab); ex1 = pp_gauss_exp(-eta * d1 * d1, w.etab); }
        if (R0 < rc) dE0 += w0 * ex0 * (df0 - 2.0 * eta * d0 * f0);
        if (R1 < rc) dE1 += w1 * ex1 * (df1 - 2.0 * eta * d1 * f1);
      }
    }
    w.fx[p0] += dE0 * w.ux[p0]; w.fy[p0] += dE0 * w.uy[p0]; w.fz[p0] += dE0 * w.uz[p0];
    if (two) { w.fx[p1] += dE1 * w.ux[p1]; w.fy[p1] += dE1 * w.uy[p1]; w.fz[p1] += dE1 * w.uz[p1]; }
  }
  __syncwarp();
}

// Force chain rule, G4 rows (Structure.cpp:645-708), factorised:
//   Phi = sum_pairs sum_{r,q} W[r][q] P_q(c) A_r(Ra) A_r(Rb),   W = 2 * 2^(1-xi) * dE/dG(row)
//   dPhi/dRa = sum_r S_r dA_r(Ra) A_r(Rb),  S_r = sum_q W[r][q] P_q
//   dPhi/dc  = sum_r T_r  A_r(Ra) A_r(Rb),  T_r = sum_q W[r][q] P'_q,  P'_q = xi lam P_q / (1 + lam c)
//   grad_a = dPhi/dRa u_a + dPhi/dc (u_b - c u_a)/Ra   (and a<->b)
// The lane accumulates grad_a of its own entry in registers; grad_b goes to shared memory, conflict-free because
// at any step the active lanes hold distinct b.
template <int NR_T, int NA_T>
__device__ inline void pp_g4_block_force(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S,
                                         int e1, int e2, int r0, int a0, bool skip_rjk)
{
  const int s1 = w.start[e1], m1 = w.start[e1 + 1] - s1;
  const int s2 = w.start[e2], m2 = w.start[e2 + 1] - s2;
  if (m1 == 0 || m2 == 0) return;
  const bool same = (e1 == e2);
  const double rcut2 = S->Rcut * S->Rcut;
  // weights of the grid chunk for this element pair
  for (int i = lane; i < NR_T * NA_T; i += 32) w.cell[i] = 0.0;
  __syncwarp();
  const int base = S->trip_base[e1][e2] + S->nG3;
  for (int row = lane; row < S->nG4; row += 32) {
    int rr = S->g4_rad[row] - r0, qq = S->g4_ang[row] - a0;
    // rows sharing a grid cell (duplicated rows) are summed by one lane at a time
    if (rr >= 0 && rr < NR_T && qq >= 0 && qq < NA_T)
      atomicAdd(&w.cell[rr * NA_T + qq], 2.0 * S->ang_pref[S->g4_ang[row]] * w.buf[base + row]);
  }
  __syncwarp();
  double W[NR_T][NA_T];
#pragma unroll
  for (int r = 0; r < NR_T; r++)
#pragma unroll
    for (int q = 0; q < NA_T; q++) W[r][q] = w.cell[r * NA_T + q];
  AngKeys<NA_T> K;
  pp_load_ang<NA_T>(S, a0, K);
  const int NR = S->NR;
  const double *Arow[NR_T], *dArow[NR_T];
#pragma unroll
  for (int r = 0; r < NR_T; r++) {
    Arow[r] = w.A + (size_t)min(r0 + r, NR - 1) * cap;
    dArow[r] = w.dA + (size_t)min(r0 + r, NR - 1) * cap;
  }

  // lanes active per pass: distinct partner indices require #lanes <= m2 in cross blocks
  const int width = same ? 32 : (m2 < 32 ? m2 : 32);
  const int npass = (m1 + width - 1) / width;
  const int nt_max = same ? (m1 >> 1) : m2;
  for (int pass = 0; pass < npass; pass++) {
    const int a = pass * width + lane;
    const bool act = (lane < width) && (a < m1);
    const int ja = s1 + (act ? a : 0);
    const double ax = w.ux[ja], ay = w.uy[ja], az = w.uz[ja], Ra = w.R[ja];
    const double invRa = 1.0 / Ra;
    double Aa[NR_T], dAa[NR_T];
#pragma unroll
    for (int r = 0; r < NR_T; r++) {
      bool ok = (r0 + r < NR);
      Aa[r] = ok ? Arow[r][ja] : 0.0;
      dAa[r] = ok ? dArow[r][ja] : 0.0;
    }
    double gx = 0.0, gy = 0.0, gz = 0.0;
    const int nt = same ? ((m1 - 1) >> 1) + (((m1 & 1) == 0 && a < (m1 >> 1)) ? 1 : 0) : m2;
    for (int t = 0; t < nt_max; t++) {
      if (act && t < nt) {
        int jb;
        if (same) { int b = a + t + 1; if (b >= m1) b -= m1; jb = s1 + b; }
        else { int b = lane + t; if (b >= m2) b -= m2; jb = s2 + b; }
        const double bx = w.ux[jb], by = w.uy[jb], bz = w.uz[jb], Rb = w.R[jb];
        double c = ax * bx + ay * by + az * bz;
        bool skip = false;
        if (skip_rjk) { double rjk2 = Ra * Ra + Rb * Rb - 2.0 * Ra * Rb * c; skip = rjk2 > rcut2; }
        if (!skip) {
          double p[NA_T], dp[NA_T];
          pp_angular<NA_T, true>(K, c, p, dp);
          double dRa = 0.0, dRb = 0.0, dc = 0.0;
#pragma unroll
          for (int r = 0; r < NR_T; r++) {
            double Sr = 0.0, Tr = 0.0;
#pragma unroll
            for (int q = 0; q < NA_T; q++) { Sr = fma(W[r][q], p[q], Sr); Tr = fma(W[r][q], dp[q], Tr); }
            double Ab = Arow[r][jb];
            double dAb = dArow[r][jb];
            dRa = fma(Sr * dAa[r], Ab, dRa);
            dRb = fma(Sr * Aa[r], dAb, dRb);
            dc = fma(Tr * Aa[r], Ab, dc);
          }
          const double ca = dc * invRa, cb = dc / Rb;
          gx += dRa * ax + ca * (bx - c * ax);
          gy += dRa * ay + ca * (by - c * ay);
          gz += dRa * az + ca * (bz - c * az);
          w.fx[jb] += dRb * bx + cb * (ax - c * bx);
          w.fy[jb] += dRb * by + cb * (ay - c * by);
          w.fz[jb] += dRb * bz + cb * (az - c * bz);
        }
      }
      __syncwarp();
    }
    if (act) { w.fx[ja] += gx; w.fy[ja] += gy; w.fz[ja] += gz; }
    __syncwarp();
  }
}

template <int NR_T, int NX, bool BASIS>
__device__ inline void pp_force_entries(WarpNbr& w, int cap, int lane, const DescSpec* __restrict__ S, bool skip_rjk)
{
#ifndef PP_SKIP_PAIR_ROWS
  if (S->npair) pp_pair_rows_force(w, cap, lane, S);
#endif
  if (S->nG3) {
    for (int e1 = 0; e1 < S->n_elem; e1++)
      for (int e2 = e1; e2 < S->n_elem; e2++) pp_g3_block_force(w, cap, lane, S, e1, e2);
  }
#ifndef PP_SKIP_G4
  if (S->nG4) {
    for (int e1 = 0; e1 < S->n_elem; e1++)
      for (int e2 = e1; e2 < S->n_elem; e2++)
        for (int r0 = 0; r0 < S->NR; r0 += NR_T) {
          if constexpr (BASIS) pp_g4_block_force_basis<NR_T, NX>(w, cap, lane, S, e1, e2, r0, skip_rjk);
          else
            for (int a0 = 0; a0 < S->NA; a0 += NX) pp_g4_block_force<NR_T, NX>(w, cap, lane, S, e1, e2, r0, a0, skip_rjk);
        }
  }
#endif
}
