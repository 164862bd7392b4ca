// Per-element MLP kernels (sm_100a): forward, backprop to parameters, backprop to inputs.
//
// Replaces Neural_network::train_MD / evaluate_MD (NN.cpp:516-697) built on Neural_network_node::evaluate
// (NN_node.h:64-81), batched over all atoms of one element.
//
// Design: the layer products are skinny FP64 GEMMs (K <= ~128, N <= ~64).  On B200 the FP64
// tensor-core path (DMMA, mma.sync m8n8k4) measures 37.2 TFLOP/s against 33.5 for the FMA pipe
// (profiles/r01_peaks_microbench.json) with 4x fewer operand loads per flop, so every product -- forward,
// delta back-propagation and the weight-gradient outer products (K = atoms) -- is issued as DMMA on tiles
// staged in shared memory.  tcgen05 has no FP64 kind, so the legacy mma.sync form is the only tensor path
// for this precision.  A CTA keeps its element's weights resident in shared memory, walks atom tiles
// persistently, and keeps the whole parameter gradient in registers (distributed over its warps as 8x8
// DMMA accumulator tiles); per-CTA partial gradients go to a scratch buffer and are summed by a second
// kernel in a fixed order, so the gradient is bit-reproducible run to run (Rprop is sign based).
//
// The bias is folded in as an extra input column equal to 1, which is exactly the reference's parameter
// layout [w_0..w_{n-1}, b] per node (NN.cpp:83-94): a layer's block of the flat parameter vector IS its
// augmented weight matrix [n_out][n_in+1].
#include "pp_common.cuh"
#include "pp_kernels.h"
#include <cmath>

#define MLP_WARPS 8
#define MLP_THREADS (MLP_WARPS * 32)
#define MLP_MAX_NT 16        // n-tiles (8 wide) per layer: layer width <= 128
#define MLP_MAX_OWN 32       // dW tiles owned by one warp
// the kernel is instantiated for NT in {8,16} and OWN in {12,32} (register budget)

struct MlpPlan {
  int TM, L;
  int n[PP_MAX_LAYERS + 1];
  int kpitch[PP_MAX_LAYERS];          // Hs_l / Ws_l pitch (doubles), l = 0..L-1
  int dpitch[PP_MAX_LAYERS + 1];      // Ds_l pitch, l = 1..L-1
  int ws_off[PP_MAX_LAYERS], hs_off[PP_MAX_LAYERS], ds_off[PP_MAX_LAYERS + 1];
  int cvec_off, mean_off, scale_off, tf_off_bytes;
  int tile_prefix[PP_MAX_LAYERS + 1]; // dW tiles of hidden weight layers l = 0..L-2
  int n_tiles_per_cta_hint;
  size_t smem_bytes;
};

__host__ __device__ static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

static bool make_plan(const NetDev& net, int TM, MlpPlan& p)
{
  p.TM = TM;
  p.L = net.n_layers;
  for (int l = 0; l <= p.L; l++) p.n[l] = net.sizes[l];
  size_t off = 0;
  for (int l = 0; l < p.L; l++) {
    p.kpitch[l] = round_up(p.n[l] + 1, 8) + 4;
    p.ws_off[l] = (int)off;
    int rows = (l == p.L - 1) ? 1 : round_up(p.n[l + 1], 8);
    off += (size_t)rows * p.kpitch[l];
  }
  for (int l = 0; l < p.L; l++) { p.hs_off[l] = (int)off; off += (size_t)TM * p.kpitch[l]; }
  for (int l = 1; l < p.L; l++) {
    p.dpitch[l] = round_up(p.n[l], 8) + 4;
    p.ds_off[l] = (int)off;
    off += (size_t)TM * p.dpitch[l];
  }
  p.cvec_off = (int)off; off += TM;
  p.mean_off = (int)off; off += p.n[0];
  p.scale_off = (int)off; off += p.n[0];
  p.tf_off_bytes = (int)(off * sizeof(double));
  int nodes = 0;
  for (int l = 1; l <= p.L; l++) nodes += p.n[l];
  p.smem_bytes = off * sizeof(double) + (size_t)nodes * sizeof(int);
  p.smem_bytes = (p.smem_bytes + 15) & ~(size_t)15;
  int t = 0;
  for (int l = 0; l < p.L - 1; l++) {
    p.tile_prefix[l] = t;
    t += (round_up(p.n[l + 1], 8) / 8) * (round_up(p.n[l] + 1, 8) / 8);
  }
  for (int l = (p.L - 1 > 0 ? p.L - 1 : 0); l <= PP_MAX_LAYERS; l++) p.tile_prefix[l] = t;
  return true;
}

// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// Transfer functions and derivatives: Neural_network_node (NN_node.h:144-254)
__device__ __forceinline__ double pp_transfer(int id, double in, double& deriv, const double* __restrict__ table)
{
  switch (id) {
    case 0: deriv = 1.0; return in;
    case 1: {   // tanh with the reference's asymptotic branch beyond |x| > 6 (NN_node.h:193-206)
      double ax = fabs(in);
      double sgn = (in > 0.0) ? 1.0 : ((in < 0.0) ? -1.0 : 0.0);
      double e = exp(-2.0 * ax);
      if (ax > 6.0) {
        deriv = 4.0 * e * (1.0 - 2.0 * e);
        return sgn * (1.0 - 2.0 * e * (1.0 - e));
      }
      double r = 1.0 / (1.0 + e);
      deriv = 4.0 * e * r * r;          // 1/cosh^2
      return sgn * (1.0 - e) * r;       // tanh
    }
    case 2: {   // cubic-spline table on [-17,17] (Tables.h:74-128, Tables.cpp:212-221)
      const double* X = table;
      const double* A = table + 1001;
      if (in < X[0]) { deriv = 0.0; return A[3]; }
      if (in >= X[1000]) { deriv = 0.0; return A[4 * 999] + A[4 * 999 + 1] + A[4 * 999 + 2] + A[4 * 999 + 3]; }
      double m = 1000.0 / (X[1000] - X[0]);
      int bin = (int)(m * in - m * X[0]);
      bin = bin < 0 ? 0 : (bin > 999 ? 999 : bin);
      while (bin > 0 && in < X[bin]) bin--;
      while (bin < 999 && in >= X[bin + 1]) bin++;
      double x = in - X[bin];
      const double* a = A + 4 * bin;
      deriv = (3.0 * a[0] * x + 2.0 * a[1]) * x + a[2];
      return ((a[0] * x + a[1]) * x + a[2]) * x + a[3];
    }
    case 3: {   // tanh_crazy (NN_node.h:208-215)
      double t = tanh(in), ch = cosh(in);
      double u = 0.1 * in, u2 = u * u, u4 = u2 * u2;
      deriv = (1.0 - u4) / (ch * ch) - t * 0.1 * 4.0 * (u2 * u);
      return (1.0 - u4) * t;
    }
    case 4: { double x2 = in * in; deriv = 3.0 * x2; return x2 * in; }
    case 5: {   // pulse, a=b=c=1 (NN_node.h:181-190)
      double ex = exp(-in * in), s = sin(in), c = cos(in);
      deriv = ex * (-2.0 * in * s + c);
      return s * ex;
    }
    case 6: deriv = cos(in); return sin(in);
    case 7: {   // periodic_tanh (NN_node.h:155-162)
      double inner = 4.0 * sin(in);
      double ch = cosh(inner);
      deriv = 4.0 * cos(in) / (ch * ch);
      return tanh(inner);
    }
    case 8: deriv = (in <= 0.0) ? 0.0 : 1.0; return (in <= 0.0) ? 0.0 : in;
  }
  deriv = 0.0;
  return 0.0;
}

// One DMMA GEMM for the 8 rows of m-tile `mt`:  C[8][N] = A[8][K] * B, with
//   A(row,k)  = As[(8*mt+row)*apitch + k]
//   B(k,n)    = BT ? Bs[k*bpitch + n] : Bs[n*bpitch + k]
// K is padded to a multiple of 4 and N to a multiple of 8 by zero-filled shared memory.
template <bool BT, int NT>
__device__ __forceinline__ void tile_gemm(double (&acc)[NT][2], const double* __restrict__ As, int apitch,
                                          const double* __restrict__ Bs, int bpitch, int mt, int K4, int ntiles, int lane)
{
  const int g = lane >> 2, q = lane & 3;
#pragma unroll
  for (int nt = 0; nt < NT; nt++) { acc[nt][0] = 0.0; acc[nt][1] = 0.0; }
  const double* arow = As + (size_t)(8 * mt + g) * apitch + q;
  for (int k0 = 0; k0 < K4; k0 += 4) {
    const double a = arow[k0];
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
      if (nt < ntiles) {
        const double b = BT ? Bs[(size_t)(k0 + q) * bpitch + 8 * nt + g] : Bs[(size_t)(8 * nt + g) * bpitch + k0 + q];
        dmma884(acc[nt][0], acc[nt][1], a, b);
      }
    }
  }
}

template <int MODE, int NT, int OWN>
__global__ void __launch_bounds__(MLP_THREADS, 1)
mlp_kernel(MlpArgs a, MlpPlan p)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sm = reinterpret_cast<double*>(smem_raw);
  int* tfs = reinterpret_cast<int*>(smem_raw + p.tf_off_bytes);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  const int L = p.L, TM = p.TM, NG = p.n[0];
  const NetDev& net = a.net;

  // ---- one-time staging: weights (zero padded), constant columns, transfer ids, input scaling ----------------
  for (int l = 0; l < L; l++) {
    const int n_in = p.n[l], n_out = p.n[l + 1], kp = p.kpitch[l];
    const int rows = (l == L - 1) ? 1 : round_up(n_out, 8);
    double* Ws = sm + p.ws_off[l];
    const double* src = net.params + net.w_off[l];
    for (int i = tid; i < rows * kp; i += MLP_THREADS) {
      int r = i / kp, k = i - r * kp;
      Ws[i] = (r < n_out && k <= n_in) ? src[(size_t)r * (n_in + 1) + k] : 0.0;
    }
    double* Hs = sm + p.hs_off[l];
    for (int i = tid; i < TM * kp; i += MLP_THREADS) {
      int k = i % kp;
      Hs[i] = (k == n_in) ? 1.0 : 0.0;
    }
    if (l >= 1) {
      double* Ds = sm + p.ds_off[l];
      for (int i = tid; i < TM * p.dpitch[l]; i += MLP_THREADS) Ds[i] = 0.0;
    }
  }
  {
    int nodes = 0;
    for (int l = 1; l <= L; l++) nodes += p.n[l];
    for (int i = tid; i < nodes; i += MLP_THREADS) tfs[i] = net.tf[i];
    for (int i = tid; i < NG; i += MLP_THREADS) { sm[p.mean_off + i] = net.mean[i]; sm[p.scale_off + i] = net.scale[i]; }
  }
  __syncthreads();

  // ---- parameter-gradient accumulators (TRAIN) ---------------------------------------------------------------
  double dW[OWN][2];
  double dwout = 0.0;     // thread t <= n_{L-1}: gradient of output-node parameter t
  const int n_dw_tiles = p.tile_prefix[L - 1 > 0 ? L - 1 : 0];
#pragma unroll
  for (int t = 0; t < OWN; t++) { dW[t][0] = 0.0; dW[t][1] = 0.0; }

  double* cvec = sm + p.cvec_off;
  const int n_tiles = (a.M + TM - 1) / TM;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int row0 = tile * TM;
    const int nvalid = min(TM, a.M - row0);
    // ---- load + normalise inputs: (G - mean)/scale with the reference's exact-zero guard (NN.cpp:648-654) ----
    {
      double* H0 = sm + p.hs_off[0];
      const int kp = p.kpitch[0];
      for (int i = tid; i < TM * NG; i += MLP_THREADS) {
        int r = i / NG, k = i - r * NG;
        double v = 0.0;
        if (r < nvalid) {
          size_t atom = a.rows ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
          double d = a.G[atom * NG + k] - sm[p.mean_off + k];
          v = (d == 0.0) ? 0.0 : d / sm[p.scale_off + k];
        }
        H0[r * kp + k] = v;
      }
      if (MODE != PP_MLP_FORWARD) {
        for (int r = tid; r < TM; r += MLP_THREADS) {
          double c = 0.0;
          if (r < nvalid) {
            size_t atom = a.rows ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
            c = (MODE == PP_MLP_TRAIN) ? a.cstruct[a.atom_struct[atom]] : a.out_scale;
          }
          cvec[r] = c;
        }
      }
    }
    __syncthreads();

    // ---- forward through the hidden layers; each warp owns whole m-tiles (8 atoms) --------------------------
    for (int mt = warp; mt < TM / 8; mt += MLP_WARPS) {
      double acc[NT][2];
      for (int l = 0; l < L - 1; l++) {
        const int n_in = p.n[l], n_out = p.n[l + 1];
        const int ntiles = (n_out + 7) >> 3;
        tile_gemm<false, NT>(acc, sm + p.hs_off[l], p.kpitch[l], sm + p.ws_off[l], p.kpitch[l], mt, round_up(n_in + 1, 4), ntiles, lane);
        double* Hn = sm + p.hs_off[l + 1] + (size_t)(8 * mt + g) * p.kpitch[l + 1];
        double* Dn = sm + p.ds_off[l + 1] + (size_t)(8 * mt + g) * p.dpitch[l + 1];
        const int* tfl = tfs + net.node_off[l];
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
          if (nt < ntiles) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              int n = 8 * nt + 2 * q + h;
              if (n < n_out) {
                double d;
                double t = pp_transfer(tfl[n], acc[nt][h], d, a.tanh_table);
                Hn[n] = t;
                Dn[n] = d;
              }
            }
          }
        }
        __syncwarp();
      }
      // ---- output node (always linear, NN.cpp:76): y = w_out . [H, 1];  4 lanes per atom ----------------------
      {
        const int n_in = p.n[L - 1];
        const double* Hl = sm + p.hs_off[L - 1] + (size_t)(8 * mt + g) * p.kpitch[L - 1];
        const double* wo = sm + p.ws_off[L - 1];
        double y = 0.0;
        for (int k = q; k <= n_in; k += 4) y = fma(Hl[k], wo[k], y);
        y += __shfl_xor_sync(0xffffffffu, y, 1);
        y += __shfl_xor_sync(0xffffffffu, y, 2);
        const int r = 8 * mt + g;
        if (MODE != PP_MLP_TRAIN && a.E_atom != nullptr && q == 0 && r < nvalid) {
          size_t atom = a.rows ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
          double d;
          a.E_atom[atom] = a.out_scale * pp_transfer(tfs[net.node_off[L - 1]], y, d, a.tanh_table);
        }
      }
      if (MODE == PP_MLP_FORWARD) continue;

      // ---- backward: delta of the last hidden layer, then back through the hidden layers -----------------------
      // (output node linear: dE/dy = c; delta_{L-1}[n] = c * w_out[n] * T'(z_{L-1}[n]))
      if (L >= 2) {
        const int n_h = p.n[L - 1];
        double* Dl = sm + p.ds_off[L - 1] + (size_t)(8 * mt) * p.dpitch[L - 1];
        const double* wo = sm + p.ws_off[L - 1];
        for (int i = lane; i < 8 * n_h; i += 32) {
          int r = i / n_h, n = i - r * n_h;
          Dl[r * p.dpitch[L - 1] + n] *= cvec[8 * mt + r] * wo[n];
        }
        __syncwarp();
        for (int l = L - 2; l >= 1; l--) {
          // delta_l = (delta_{l+1} * W_l[:, :n_l]) .* T'(z_l)
          const int n_l = p.n[l], n_up = p.n[l + 1];
          const int ntiles = (n_l + 7) >> 3;
          tile_gemm<true, NT>(acc, sm + p.ds_off[l + 1], p.dpitch[l + 1], sm + p.ws_off[l], p.kpitch[l], mt, round_up(n_up, 4), ntiles, lane);
          double* Dr = sm + p.ds_off[l] + (size_t)(8 * mt + g) * p.dpitch[l];
#pragma unroll
          for (int nt = 0; nt < NT; nt++) {
            if (nt < ntiles) {
#pragma unroll
              for (int h = 0; h < 2; h++) {
                int n = 8 * nt + 2 * q + h;
                if (n < n_l) Dr[n] *= acc[nt][h];
              }
            }
          }
          __syncwarp();
        }
      }
      if (MODE == PP_MLP_INPUT_GRAD) {
        // dE/dG_k = (delta_1 * W_0[:, k]) / scale_k   (NN.cpp:546-566); no hidden layer: c * w_out[k] / scale_k
        const int r = 8 * mt + g;
        size_t atom = 0;
        if (r < nvalid) atom = a.rows ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
        if (L >= 2) {
          const int ntiles = (NG + 7) >> 3;
          for (int nt0 = 0; nt0 < ntiles; nt0 += NT) {
            const int ntc = min(NT, ntiles - nt0);
            tile_gemm<true, NT>(acc, sm + p.ds_off[1], p.dpitch[1], sm + p.ws_off[0] + 8 * nt0, p.kpitch[0], mt, round_up(p.n[1], 4), ntc, lane);
            if (r < nvalid) {
#pragma unroll
              for (int nt = 0; nt < NT; nt++) {
                if (nt < ntc) {
#pragma unroll
                  for (int h = 0; h < 2; h++) {
                    int k = 8 * (nt0 + nt) + 2 * q + h;
                    if (k < NG) a.dEdG[atom * NG + k] = acc[nt][h] / sm[p.scale_off + k];
                  }
                }
              }
            }
          }
        } else if (r < nvalid) {
          const double* wo = sm + p.ws_off[0];
          for (int k = q; k < NG; k += 4) a.dEdG[atom * NG + k] = cvec[r] * wo[k] / sm[p.scale_off + k];
        }
      }
    }
    __syncthreads();

    // ---- parameter gradients: dW_l += delta_{l+1}^T * [H_l, 1]  (K = atoms of the tile) ------------------------
    if constexpr (MODE == PP_MLP_TRAIN) {
      // output node: thread t <= n_{L-1}
      {
        const int n_in = p.n[L - 1];
        if (tid <= n_in) {
          const double* Hl = sm + p.hs_off[L - 1] + tid;
          const int kp = p.kpitch[L - 1];
          double s = 0.0;
          for (int r = 0; r < TM; r++) s = fma(cvec[r], Hl[(size_t)r * kp], s);
          dwout += s;
        }
      }
#pragma unroll
      for (int t = 0; t < OWN; t++) {
        const int tile_id = warp + t * MLP_WARPS;
        if (tile_id < n_dw_tiles) {
          int l = 0;
          while (tile_id >= p.tile_prefix[l + 1]) l++;
          const int local = tile_id - p.tile_prefix[l];
          const int ncols = round_up(p.n[l] + 1, 8) >> 3;
          const int mi = local / ncols, ni = local - mi * ncols;
          const double* Dl = sm + p.ds_off[l + 1] + 8 * mi + g;      // A(m = neuron, k = atom) = delta[atom][neuron]
          const double* Hl = sm + p.hs_off[l] + 8 * ni + g;          // B(k = atom, n = input) = H[atom][input]
          const int dp = p.dpitch[l + 1], kp = p.kpitch[l];
          double c0 = dW[t][0], c1 = dW[t][1];
          for (int k0 = 0; k0 < TM; k0 += 4) {
            const double av = Dl[(size_t)(k0 + q) * dp];
            const double bv = Hl[(size_t)(k0 + q) * kp];
            dmma884(c0, c1, av, bv);
          }
          dW[t][0] = c0; dW[t][1] = c1;
        }
      }
      __syncthreads();
    }
  }

  // ---- per-CTA partial gradient -> scratch -----------------------------------------------------------------
  if constexpr (MODE == PP_MLP_TRAIN) {
    double* out = a.partial + (size_t)blockIdx.x * a.P_total + a.p_off;
    {
      const int n_in = p.n[L - 1];
      if (tid <= n_in) out[net.w_off[L - 1] + tid] = dwout;
    }
#pragma unroll
    for (int t = 0; t < OWN; t++) {
      const int tile_id = warp + t * MLP_WARPS;
      if (tile_id < n_dw_tiles) {
        int l = 0;
        while (tile_id >= p.tile_prefix[l + 1]) l++;
        const int local = tile_id - p.tile_prefix[l];
        const int ncols = round_up(p.n[l] + 1, 8) >> 3;
        const int mi = local / ncols, ni = local - mi * ncols;
        const int neuron = 8 * mi + g;
        const int n_in = p.n[l];
        if (neuron < p.n[l + 1]) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            int k = 8 * ni + 2 * q + h;
            if (k <= n_in) out[net.w_off[l] + (size_t)neuron * (n_in + 1) + k] = dW[t][h];
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
static int g_sm_count = 0;
static int sm_count()
{
  if (!g_sm_count) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

int ppk_mlp_max_ctas(void) { return sm_count(); }

int ppk_mlp_launch(int mode, const MlpArgs& a, cudaStream_t st, long long* launches)
{
  if (a.M == 0) {
    return 0;
  }
  const NetDev& net = a.net;
  PP_REQUIRE(net.n_layers >= 1 && net.n_layers <= PP_MAX_LAYERS, "network has %d weight layers (max %d)", net.n_layers, PP_MAX_LAYERS);
  int max_hidden = 0;
  for (int l = 1; l < net.n_layers; l++) {
    PP_REQUIRE(net.sizes[l] <= 8 * MLP_MAX_NT, "hidden layer %d has %d nodes (max %d)", l, net.sizes[l], 8 * MLP_MAX_NT);
    if (net.sizes[l] > max_hidden) max_hidden = net.sizes[l];
  }
  MlpPlan plan;
  int TM = 64;
  for (;;) {
    make_plan(net, TM, plan);
    if (plan.smem_bytes <= 227 * 1024) break;
    TM -= 8;
    PP_REQUIRE(TM >= 8, "network too large for shared memory (%zu bytes at 8 atoms per tile)", plan.smem_bytes);
  }
  const int n_dw_tiles = plan.tile_prefix[net.n_layers - 1 > 0 ? net.n_layers - 1 : 0];
  PP_REQUIRE(n_dw_tiles <= MLP_MAX_OWN * MLP_WARPS, "network has %d weight-gradient tiles (max %d)", n_dw_tiles, MLP_MAX_OWN * MLP_WARPS);
  const int n_tiles = (a.M + TM - 1) / TM;
  int grid = sm_count();
  if (mode != PP_MLP_TRAIN && grid > n_tiles) grid = n_tiles;   // TRAIN: every CTA writes its (possibly zero) partial
  const bool wide = max_hidden > 64;
  const bool many = (n_dw_tiles + MLP_WARPS - 1) / MLP_WARPS > 12;
#define PP_LAUNCH(MODE_, NT_, OWN_)                                                                             \
  {                                                                                                             \
    auto k = mlp_kernel<MODE_, NT_, OWN_>;                                                                      \
    PP_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem_bytes));        \
    k<<<grid, MLP_THREADS, plan.smem_bytes, st>>>(a, plan);                                                     \
  }
  if (mode == PP_MLP_FORWARD) {
    if (wide) PP_LAUNCH(PP_MLP_FORWARD, 16, 1) else PP_LAUNCH(PP_MLP_FORWARD, 8, 1)
  } else if (mode == PP_MLP_INPUT_GRAD) {
    if (wide) PP_LAUNCH(PP_MLP_INPUT_GRAD, 16, 1) else PP_LAUNCH(PP_MLP_INPUT_GRAD, 8, 1)
  } else {
    if (wide && many) PP_LAUNCH(PP_MLP_TRAIN, 16, 32)
    else if (wide) PP_LAUNCH(PP_MLP_TRAIN, 16, 12)
    else if (many) PP_LAUNCH(PP_MLP_TRAIN, 8, 32)
    else PP_LAUNCH(PP_MLP_TRAIN, 8, 12)
  }
#undef PP_LAUNCH
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Per-structure energy, error and loss terms: Potential.cpp:641-668.  One CTA, fixed summation order
// (deterministic).  cstruct[s] = 2*Err*out_scale for train structures, else 0.
__global__ void __launch_bounds__(1024) struct_errors_kernel(StructArgs a)
{
  __shared__ double s_sse[32], s_vsse[32], s_nval[32];
  double sse = 0.0, vsse = 0.0, nval = 0.0;
  for (int s = threadIdx.x; s < a.n_struct; s += blockDim.x) {
    double e = 0.0;
    for (long long at = a.struct_atom0[s]; at < a.struct_atom0[s + 1]; at++) e += a.E_atom[at];
    e += a.out_mean;
    a.E_struct[s] = e;
    double err = e - a.target[s];
    int tag = a.tag[s];
    double c = 0.0;
    if (tag == 0) { sse += err * err; c = 2.0 * err * a.out_scale; }
    else if (tag == 1) { vsse += err * err; nval += 1.0; }
    a.cstruct[s] = c;
  }
  sse = pp_warp_sum(sse); vsse = pp_warp_sum(vsse); nval = pp_warp_sum(nval);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_sse[warp] = sse; s_vsse[warp] = vsse; s_nval[warp] = nval; }
  __syncthreads();
  if (warp == 0) {
    int nw = blockDim.x >> 5;
    sse = lane < nw ? s_sse[lane] : 0.0;
    vsse = lane < nw ? s_vsse[lane] : 0.0;
    nval = lane < nw ? s_nval[lane] : 0.0;
    sse = pp_warp_sum(sse); vsse = pp_warp_sum(vsse); nval = pp_warp_sum(nval);
    if (lane == 0) { a.result_tail[0] = sse; a.result_tail[1] = vsse; a.result_tail[2] = nval; }
  }
}

int ppk_struct_errors(const StructArgs& a, cudaStream_t st, long long* launches)
{
  struct_errors_kernel<<<1, 1024, 0, st>>>(a);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

__global__ void reduce_partials_kernel(const double* __restrict__ partial, int n_cta, int P, double* __restrict__ grad)
{
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  double s = 0.0;
  for (int c = 0; c < n_cta; c++) s += partial[(size_t)c * P + p];
  grad[p] = s;
}

int ppk_reduce_partials(const double* partial, int n_cta, int P, double* grad, cudaStream_t st, long long* launches)
{
  if (P == 0) return 0;
  reduce_partials_kernel<<<(P + 127) / 128, 128, 0, st>>>(partial, n_cta, P, grad);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Column statistics for Potential::Precondition.  Each CTA sums a contiguous chunk of rows per column into
// partial[cta][k]; a second kernel adds the CTAs in order.
__global__ void __launch_bounds__(256)
column_stats_kernel(const double* __restrict__ G, int NG, const int* __restrict__ rows, int M, const int* __restrict__ atom_struct,
                    const int* __restrict__ tag, const double* __restrict__ mean, int pass, double* __restrict__ partial,
                    unsigned long long* __restrict__ count)
{
  const int per = (M + gridDim.x - 1) / gridDim.x;
  const int r0 = blockIdx.x * per, r1 = min(M, r0 + per);
  unsigned long long cnt = 0;
  for (int k = threadIdx.x; k < NG; k += blockDim.x) {
    double s = 0.0;
    const double mu = pass ? mean[k] : 0.0;
    unsigned long long c = 0;
    for (int r = r0; r < r1; r++) {
      size_t atom = rows ? (size_t)rows[r] : (size_t)r;
      if (tag[atom_struct[atom]] != 0) continue;
      double v = G[atom * NG + k];
      if (pass) { double d = v - mu; s += d * d; }
      else s += v;
      c++;
    }
    partial[(size_t)blockIdx.x * NG + k] = s;
    if (k == 0) cnt = c;
  }
  if (threadIdx.x == 0 && pass == 0) atomicAdd(count, cnt);
}

int ppk_column_stats(const double* G, int NG, const int* rows, int M, const int* atom_struct, const int* tag,
                     const double* mean, int pass, double* partial, int n_cta, double* sums, long long* count,
                     cudaStream_t st, long long* launches)
{
  if (M == 0) {
    PP_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * NG, st));
    return 0;
  }
  column_stats_kernel<<<n_cta, 256, 0, st>>>(G, NG, rows, M, atom_struct, tag, mean, pass, partial,
                                             reinterpret_cast<unsigned long long*>(count));
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return ppk_reduce_partials(partial, n_cta, NG, sums, st, launches);
}

// ---------------------------------------------------------------------------------------------------------
// tanh spline table of the reference's Tables("tanh") (Tables.cpp:212-221): C2 cubic spline through tanh at the
// knots -17 + 0.034 i, i = 0..1000.  The reference ships the coefficients as data; they are regenerated here
// (natural-spline tridiagonal solve) and agree with the shipped table to ~7e-12 in the cubic coefficient.
// Layout: X[1001] followed by a[1000][4] (cubic, quadratic, linear, constant in (x - X_i)).
void ppk_make_tanh_table(std::vector<double>& t)
{
  const int N = 1000;
  t.assign(1001 + 4 * N, 0.0);
  std::vector<double> y(N + 1), M2(N + 1, 0.0), cp(N + 1, 0.0), dp(N + 1, 0.0);
  for (int i = 0; i <= N; i++) { t[i] = (double)(-17000 + 34 * i) / 1000.0; y[i] = std::tanh(t[i]); }
  for (int i = 1; i < N; i++) {
    double h0 = t[i] - t[i - 1], h1 = t[i + 1] - t[i];
    double sub = h0 / 6.0, diag = (h0 + h1) / 3.0, sup = h1 / 6.0;
    double rhs = (y[i + 1] - y[i]) / h1 - (y[i] - y[i - 1]) / h0;
    double m = diag - sub * cp[i - 1];
    cp[i] = sup / m;
    dp[i] = (rhs - sub * dp[i - 1]) / m;
  }
  for (int i = N - 1; i >= 1; i--) M2[i] = dp[i] - cp[i] * M2[i + 1];
  for (int i = 0; i < N; i++) {
    double h = t[i + 1] - t[i];
    double* a = &t[1001 + 4 * i];
    a[0] = (M2[i + 1] - M2[i]) / (6.0 * h);
    a[1] = M2[i] / 2.0;
    a[2] = (y[i + 1] - y[i]) / h - h * (2.0 * M2[i] + M2[i + 1]) / 6.0;
    a[3] = y[i];
  }
}
