// Shared definitions for the prophet_b200 CUDA library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#define PP_MAX_ELEM 4          // elements per potential
#define PP_MAX_PAIR_ROWS 32    // G1 + G2 rows
#define PP_MAX_G4_ROWS 64
#define PP_MAX_G3_ROWS 16
#define PP_MAX_RAD 64          // distinct (Rc, fc, eta) among G4 rows: up to every row its own (the -optimize search perturbs them all)
#define PP_MAX_ANG 64          // distinct (xi, lambda) among G4 rows
#define PP_MAX_LAYERS 8        // weight layers per net (hidden + output)
#define PP_WARP 32

void pp_set_error(const char* fmt, ...);

#define PP_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      pp_set_error("CUDA error '%s' at %s:%d (%s)", cudaGetErrorString(e_), __FILE__, __LINE__, #call); \
      return 1;                                                                                    \
    }                                                                                              \
  } while (0)

#define PP_REQUIRE(cond, ...)        \
  do {                               \
    if (!(cond)) {                   \
      pp_set_error(__VA_ARGS__);     \
      return 1;                      \
    }                                \
  } while (0)

// ---------------------------------------------------------------------------------------------------------
// Descriptor specification, preprocessed on the host (pp_set_descriptors) and read by every descriptor /
// force kernel.  G4 rows are factorised: G4 = 2^(1-xi) (1+lam c)^xi * [exp(-eta R^2) fc(R)] * [exp(-eta Ru^2) fc(Ru)]
// so each row is a (radial key, angular key) cell of a small grid (Structure.h:368-418).
struct DescSpec {
  int n_elem, nG1, nG2, nG3, nG4, NG, npair, ntrip;
  double Rcut;
  // pair rows: G1 then G2 (Structure.h:351-360)
  double pair_rc[PP_MAX_PAIR_ROWS], pair_eta[PP_MAX_PAIR_ROWS], pair_rs[PP_MAX_PAIR_ROWS];
  int pair_fc[PP_MAX_PAIR_ROWS];
  // G3 rows (not factorisable: depends on Rjk), Structure.h:361-367
  double g3_rc[PP_MAX_G3_ROWS], g3_eta[PP_MAX_G3_ROWS], g3_xi[PP_MAX_G3_ROWS], g3_lam[PP_MAX_G3_ROWS];
  int g3_fc[PP_MAX_G3_ROWS];
  // G4 grid
  int NR, NA;
  double rad_rc[PP_MAX_RAD], rad_eta[PP_MAX_RAD];
  int rad_fc[PP_MAX_RAD];
  double ang_xi[PP_MAX_ANG], ang_lam[PP_MAX_ANG], ang_pref[PP_MAX_ANG];   // pref = 2^(1-xi)
  int ang_xi_int[PP_MAX_ANG];     // xi as integer exponent, or -1 when xi is not an integer (binomial series)
  int ang_prog[PP_MAX_ANG];       // evaluation program word, see pp_desc.cuh (AngKeys)
  double scan_d2_hi, scan_d2_lo;  // squared-distance thresholds equivalent to d <= Rcut and d > 0.01 on d = sqrt(d2)
  int g4_rad[PP_MAX_G4_ROWS], g4_ang[PP_MAX_G4_ROWS];
  // power-basis fast path (pp_desc.cuh): basis_nb = number of power levels (0 = not applicable), the two lambdas,
  // which levels each lambda group uses,
  // and each row's (lambda group, level)
  int basis_nb;
  int basis_used[2];              // per lambda group: bit k set when some row uses power level k
  double basis_lam[2];
  int g4_bg[PP_MAX_G4_ROWS], g4_bb[PP_MAX_G4_ROWS];
  // Evaluation order of the Gaussian factors exp(-eta (R - Rs)^2) of the pair rows and exp(-eta R^2) of the radial keys:
  // entries sharing Rs sorted by eta, so that an eta which is an integer multiple m of its predecessor's costs the m-th
  // power of the predecessor's exponential (two multiplies for the usual ratio-4 ladders) instead of another exp.
  // chain[k] = 0: fresh exp; m >= 1: (exponential of order[k-1])^m.  The accumulated power is capped (pp_set_descriptors),
  // which keeps the relative error below 1e-12.
  int pair_order[PP_MAX_PAIR_ROWS], pair_chain[PP_MAX_PAIR_ROWS];
  int rad_order[PP_MAX_RAD], rad_chain[PP_MAX_RAD];
  // slot of the first triplet row for element pair (e1,e2): Structure::init_type_G_map (Structure.cpp:1500-1510)
  int trip_base[PP_MAX_ELEM][PP_MAX_ELEM];
};

// ---------------------------------------------------------------------------------------------------------
// cutoff function fc and its derivative d fc/dR (Structure.h:157-178, 421-443)
__device__ __forceinline__ double pp_fc(double frac, int fcswitch)
{
  if (frac >= 1.0) return 0.0;
  if (fcswitch == 0) {
    double x = 3.14159265358979323846 * frac;
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
  double t = tanh(1.0 - frac);
  return t * t * t;
}

__device__ __forceinline__ double pp_dfc(double Rc, double R, int fcswitch)
{
  if (R / Rc >= 1.0) return 0.0;
  if (fcswitch == 0) {
    double x = 3.14159265358979323846 * R / Rc;
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
  // the reference's tanh^3 derivative carries an extra 1/(2 R Rc) instead of 1/Rc (Structure.h:439); kept as is
  double t = tanh(1.0 - R / Rc);
  double ch = cosh(1.0 - R / Rc);
  return (-3.0 * t * t * (1.0 / ch) * (1.0 / ch)) / (2.0 * R * Rc);
}

// warp-wide sum, result valid in every lane
__device__ __forceinline__ double pp_warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One row group's share of a column sum over the per-CTA partial rows (reduce_partials_kernel, reduce_push_kernel): rows
// grp, grp + 32, grp + 64, ... of one column, ten at a time with all ten loads in flight, added in a FIXED order (a
// balanced tree over the ten, chunks in sequence), so that the sum is bit-reproducible.
__device__ __forceinline__ double pp_group_column_sum(const double* __restrict__ src, int grp, int n_cta, int stride)
{
  double total = 0.0;
  for (int c0 = grp; c0 < n_cta; c0 += 320) {
    double v[10];
#pragma unroll
    for (int j = 0; j < 10; j++) {
      const int c = c0 + 32 * j;
      v[j] = (c < n_cta) ? src[(size_t)c * stride] : 0.0;
    }
    total += (((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]))) + (v[8] + v[9]);
  }
  return total;
}
