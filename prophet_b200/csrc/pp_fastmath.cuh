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
  double e = fma(-d, y, 1.0);
  y = fma(y, e, y);
  e = fma(-d, y, 1.0);
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
