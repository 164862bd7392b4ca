// Short FP64 exp / reciprocal / tanh for the transfer-function phase of mlp_kernel.
//
// The transfer phase is bound by the FP64 pipe (a warp-wide FP64 instruction takes two issue slots of the 16-lane
// unit), and the library exp() + division cost ~70 instructions per tanh.  Here exp(x), x <= 0, uses a 64-entry
// table of 2^(j/64) (shared memory) and a degree-5 polynomial on |r| <= ln2/128, and 1/d for d in [1,2] is the
// hardware seed (MUFU.RCP64H) plus Newton steps without the library's special-case paths: ~25 FP64 instructions per
// tanh together.  Relative error of tanh and of sech^2 < 1e-15 (tools/tanh_accuracy.cu), five orders of magnitude
// inside the 1e-10 parity tolerance.
#pragma once

// exp(x) for x <= 0.  tab[j] = 2^(j/64), j = 0..63.  Returns 0 below -700 (true value < 1e-304).
__device__ __forceinline__ double pp_fast_exp_nonpos(double x, const double* __restrict__ tab)
{
  const double SHIFT = 6755399441055744.0;                  // 1.5 * 2^52: adding it rounds to the nearest integer
  const double t = fma(x, 92.33248261689366 /* 64/ln2 */, SHIFT);
  const int n = __double2loint(t);                          // n = rint(64 x / ln2)  (two's complement in the low word)
  const double nf = t - SHIFT;
  double r = fma(nf, -0.01083042469326756 /* ln2/64, leading 32 bits */, x);
  r = fma(nf, -2.9815858269852933e-12 /* ln2/64, remainder */, r);
  double p = fma(r, 8.3333333333333332e-03, 4.1666666666666664e-02);
  p = fma(p, r, 1.6666666666666666e-01);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  const double v = p * tab[n & 63];
  const int hi = __double2hiint(v) + ((n >> 6) << 20);     // * 2^(n div 64)
  const double res = __hiloint2double(hi, __double2loint(v));
  return x < -700.0 ? 0.0 : res;
}

// 1/d for d in [1, 2]
__device__ __forceinline__ double pp_fast_rcp_1to2(double d)
{
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  // the seed is good to ~2^-23; two Newton steps square that twice (tools/tanh_accuracy.cu: no change from a third)
  double e = fma(-d, y, 1.0);
  y = fma(y, e, y);
  e = fma(-d, y, 1.0);
  y = fma(y, e, y);
  return y;
}

// tanh(x) and its derivative as the reference's "tanh" node computes them (NN_node.h:193-206): the closed form
// up to |x| = 6, the asymptotic expansion in e = exp(-2|x|) beyond.
__device__ __forceinline__ double pp_fast_tanh(double in, double& deriv, const double* __restrict__ tab)
{
  const double ax = fabs(in);
  const double e = pp_fast_exp_nonpos(-2.0 * ax, tab);
  const double r = pp_fast_rcp_1to2(1.0 + e);
  const bool far = ax > 6.0;
  const double er = e * r;
  deriv = far ? 4.0 * e * (1.0 - 2.0 * e) : 4.0 * er * r;
  const double t = far ? 1.0 - 2.0 * e * (1.0 - e) : (1.0 - e) * r;
  return copysign(in == 0.0 ? 0.0 : t, in);
}

// The reference's default transfer function "tanh_spline" (Tables.h:74-128, Tables.cpp:212-221): a C2 cubic spline
// through tanh on the knots X_i = -17 + 0.034 i, i = 0..1000, saturating outside.  table = X[1001], one pad, then
// a[1000][4] (cubic, quadratic, linear, constant in x - X_i) at PP_SPLINE_COEF, 16-byte aligned.
// Branch-free form for the unrolled epilogues: the bin from the uniform spacing with one correction step either way
// (what the reference's search loops come to; the knots are recomputed instead of loaded -- a bin chosen differently
// within an ulp of a knot changes nothing at 1e-10, the spline being C2), two 16-byte loads for the coefficients, the
// saturated ends by selection.
#define PP_SPLINE_COEF 1002
__device__ __forceinline__ double pp_fast_tanh_spline(double in, double& deriv, const double* __restrict__ table)
{
  int bin = (int)((in + 17.0) * (1000.0 / 34.0));
  bin = bin < 0 ? 0 : (bin > 999 ? 999 : bin);
  const double xl = fma(0.034, (double)bin, -17.0);
  bin = (in < xl && bin > 0) ? bin - 1 : ((in >= xl + 0.034 && bin < 999) ? bin + 1 : bin);
  const double2* a = reinterpret_cast<const double2*>(table + PP_SPLINE_COEF + 4 * bin);
  const double2 c01 = __ldg(a), c23 = __ldg(a + 1);
  const double x = in - fma(0.034, (double)bin, -17.0);
  double d = (3.0 * c01.x * x + 2.0 * c01.y) * x + c23.x;
  double v = ((c01.x * x + c01.y) * x + c23.x) * x + c23.y;
  if (in < -17.0) { v = __ldg(table + PP_SPLINE_COEF + 3); d = 0.0; }
  if (in >= 17.0) {
    const double2 e01 = __ldg(reinterpret_cast<const double2*>(table + PP_SPLINE_COEF + 4 * 999));
    const double2 e23 = __ldg(reinterpret_cast<const double2*>(table + PP_SPLINE_COEF + 4 * 999) + 1);
    v = e01.x + e01.y + e23.x + e23.y;
    d = 0.0;
  }
  deriv = d;
  return v;
}
