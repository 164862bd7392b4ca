// Device helpers shared by the per-element MLP kernels (pp_mlp.cu, pp_mlpw.cu): FP64 DMMA, asynchronous copies
// (cp.async, bulk / tensor copies on mbarriers), transfer functions, weight-gradient blocks.
#pragma once
#include "pp_common.cuh"
#include "pp_kernels.h"
#include "pp_fastmath.cuh"
#include <cuda.h>

// One weight-gradient work item: an nm x nn block (nm, nn in {1, 2}) of 8x8 DMMA tiles
//   dW_l[8 (mi+i) .. +8][8 (ni+j) .. +8] += delta_{l+1}^T [H_l, 1]        (K = atoms of the tile)
// A 2x2 block loads two A and two B fragments per k-step for four independent DMMAs: half the shared-memory
// traffic of single tiles and four accumulator chains in flight instead of one.  Single tiles have nm = nn = 1.
// The output node is the "layer" L-1 with delta = the per-atom error factor c (cvec, pitch 8, column 0).
struct DwTile {
  int a_off;      // delta_{l+1} base + 8*mi          (doubles)
  int b_off;      // H_l base + 8*ni  (layer 0: column offset inside an X buffer)
  int dp, kp;     // pitches
  int layer, mi, ni;
  int nm, nn;     // block shape; nm = 0: empty slot
  int pad[3];
};

__host__ __device__ static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}
__device__ __forceinline__ void pair_barrier(int mt) { asm volatile("bar.sync %0, 64;" ::"r"(1 + mt) : "memory"); }
__device__ __forceinline__ void cp_async16(void* dst, const void* src)
{
  unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src));
}
__device__ __forceinline__ void cp_async8(void* dst, const void* src)
{
  unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

// Bulk asynchronous copies (the TMA engine, 1-D form): one instruction per descriptor row instead of one cp.async per
// 16 bytes per thread; completion is counted in bytes on a shared-memory mbarrier.
__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity)
{
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}

// 2-D tiled tensor copy: box (kp0 columns x TM rows) of the [atoms][NG] descriptor array starting at row `row`;
// columns beyond NG and rows beyond the array are zero-filled by the copy engine.
__device__ __forceinline__ void tma_load_tile(void* dst, const CUtensorMap* map, int row, unsigned long long* bar)
{
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_addr(dst)),
      "l"(map), "r"(0), "r"(row), "r"(smem_addr(bar))
      : "memory");
}

// Transfer functions and derivatives: Neural_network_node (NN_node.h:144-254).
// Every transfer function but tanh: one out-of-line copy (the library tanh / cosh / sin / cos / exp bodies are long, and
// the callers are unrolled over the thread's 2*NT values)
static __device__ __noinline__ double pp_transfer_other(int id, double in, double* deriv_out, const double* __restrict__ table)
{
  double deriv = 0.0;
  double res = 0.0;
  switch (id) {
    case 0: deriv = 1.0; { res = in; break; }
    case 2: {   // cubic-spline table on [-17,17] (Tables.h:74-128, Tables.cpp:212-221)
      const double* X = table;
      const double* A = table + PP_SPLINE_COEF;
      if (in < X[0]) { deriv = 0.0; { res = A[3]; break; } }
      if (in >= X[1000]) { deriv = 0.0; { res = A[4 * 999] + A[4 * 999 + 1] + A[4 * 999 + 2] + A[4 * 999 + 3]; break; } }
      double m = 1000.0 / (X[1000] - X[0]);
      int bin = (int)(m * in - m * X[0]);
      bin = bin < 0 ? 0 : (bin > 999 ? 999 : bin);
      while (bin > 0 && in < X[bin]) bin--;
      while (bin < 999 && in >= X[bin + 1]) bin++;
      double x = in - X[bin];
      const double* a = A + 4 * bin;
      deriv = (3.0 * a[0] * x + 2.0 * a[1]) * x + a[2];
      { res = ((a[0] * x + a[1]) * x + a[2]) * x + a[3]; break; }
    }
    case 3: {   // tanh_crazy (NN_node.h:208-215)
      double t = tanh(in), ch = cosh(in);
      double u = 0.1 * in, u2 = u * u, u4 = u2 * u2;
      deriv = (1.0 - u4) / (ch * ch) - t * 0.1 * 4.0 * (u2 * u);
      { res = (1.0 - u4) * t; break; }
    }
    case 4: { double x2 = in * in; deriv = 3.0 * x2; { res = x2 * in; break; } }
    case 5: {   // pulse, a=b=c=1 (NN_node.h:181-190)
      double ex = exp(-in * in), s = sin(in), c = cos(in);
      deriv = ex * (-2.0 * in * s + c);
      { res = s * ex; break; }
    }
    case 6: deriv = cos(in); { res = sin(in); break; }
    case 7: {   // periodic_tanh (NN_node.h:155-162)
      double inner = 4.0 * sin(in);
      double ch = cosh(inner);
      deriv = 4.0 * cos(in) / (ch * ch);
      { res = tanh(inner); break; }
    }
    case 8: deriv = (in <= 0.0) ? 0.0 : 1.0; { res = (in <= 0.0) ? 0.0 : in; break; }
  }
  *deriv_out = deriv;
  return res;
}

__device__ __forceinline__ double pp_transfer(int id, double in, double& deriv, const double* __restrict__ table,
                                              const double* __restrict__ exptab)
{
  if (id == 1) return pp_fast_tanh(in, deriv, exptab);   // tanh, asymptotic branch beyond |x| > 6 (NN_node.h:193-206)
  if (id == 2) return pp_fast_tanh_spline(in, deriv, table);      // the reference's default
  return pp_transfer_other(id, in, &deriv, table);
}

// One DMMA GEMM for the 8 rows of m-tile `mt` and the n-tiles [nt0, nt0+NT):  C[8][8 NT] = A[8][K] * B, with
//   A(row,k) = As[(8*mt+row)*apitch + k]
//   B(k,n)   = BT ? Bs[k*bpitch + n] : Bs[n*bpitch + k]
// K is padded to a multiple of 4 and N to 8*NT per warp by zero-filled shared memory, so the loops carry no
// bounds tests (GUARD = true restores a run-time tile count ntc <= NT for the one product whose B operand is not
// padded that far).  The next k-step's fragments are loaded before the current step's DMMAs are issued.
template <bool BT, bool GUARD, int NT>
__device__ __forceinline__ void tile_gemm(double (&acc)[NT][2], const double* __restrict__ As, int apitch,
                                          const double* __restrict__ Bs, int bpitch, int mt, int K4, int nt0, int ntc, int lane)
{
  const int g = lane >> 2, q = lane & 3;
#pragma unroll
  for (int nt = 0; nt < NT; nt++) { acc[nt][0] = 0.0; acc[nt][1] = 0.0; }
  const double* ap = As + (size_t)(8 * mt + g) * apitch + q;
  const double* bp = BT ? Bs + (size_t)q * bpitch + 8 * nt0 + g : Bs + (size_t)(8 * nt0 + g) * bpitch + q;
  const int bstep_k = BT ? 4 * bpitch : 4;
  const int bstep_n = BT ? 8 : 8 * bpitch;
  double a_cur = ap[0];
  double b_cur[NT];
#pragma unroll
  for (int nt = 0; nt < NT; nt++) b_cur[nt] = (!GUARD || nt < ntc) ? bp[nt * bstep_n] : 0.0;
  for (int k0 = 4; k0 < K4; k0 += 4) {
    ap += 4; bp += bstep_k;
    double a_nxt = ap[0];
    double b_nxt[NT];
#pragma unroll
    for (int nt = 0; nt < NT; nt++) b_nxt[nt] = (!GUARD || nt < ntc) ? bp[nt * bstep_n] : 0.0;
#pragma unroll
    for (int nt = 0; nt < NT; nt++)
      if (!GUARD || nt < ntc) dmma884(acc[nt][0], acc[nt][1], a_cur, b_cur[nt]);
    a_cur = a_nxt;
#pragma unroll
    for (int nt = 0; nt < NT; nt++) b_cur[nt] = b_nxt[nt];
  }
#pragma unroll
  for (int nt = 0; nt < NT; nt++)
    if (!GUARD || nt < ntc) dmma884(acc[nt][0], acc[nt][1], a_cur, b_cur[nt]);
}

// Weight-gradient block: NM x NN DMMA tiles sharing their A (delta) and B (activation) fragments; accumulators are
// dW[BASE + i*NN + j].  pa / pb point at this lane's element of the first k-step, sa / sb advance four atoms.
template <int OWN, int BASE, int NM, int NN>
__device__ __forceinline__ void dw_block(double (&dW)[OWN][2], const double* __restrict__ pa, const double* __restrict__ pb,
                                         int sa, int sb, int TM)
{
#pragma unroll 2
  for (int k0 = 0; k0 < TM; k0 += 4) {
    double av[NM], bv[NN];
#pragma unroll
    for (int i = 0; i < NM; i++) av[i] = pa[8 * i];
#pragma unroll
    for (int j = 0; j < NN; j++) bv[j] = pb[8 * j];
    pa += sa; pb += sb;
#pragma unroll
    for (int i = 0; i < NM; i++) {
#pragma unroll
      for (int j = 0; j < NN; j++) dmma884(dW[BASE + i * NN + j][0], dW[BASE + i * NN + j][1], av[i], bv[j]);
    }
  }
}
