// FP32 mode (pp_create(..., fp_mode = 32)): the single-pass training epoch in single precision (sm_100a).
//
// The reference has one arithmetic type, REAL (Common.h:53-61; its float build does not compile), and north_star
// states the second tolerance against the FP64 path: relative 1e-5.  In this mode the descriptors are kept in HBM as
// FP32 (half the bytes the epoch streams) and every layer product runs on the tensor cores as 3xTF32: each FP32 operand
// is split into hi = tf32(x) and lo = tf32(x - hi), and  A B ~= Ahi Bhi + Alo Bhi + Ahi Blo  is accumulated in FP32
// (the dropped Alo Blo term is 2^-22 relative), so the products keep FP32 accuracy although the multipliers are 10-bit.
// What stays FP64: the per-structure energy sums, errors and loss terms (E_s - target cancels ~3 digits), and the sum
// of the per-CTA partial gradients.
//
// Structure (same as mlp_fused2_kernel, pp_mlp.cu): persistent CTAs of 16 warps, tiles of whole structures up to 256
// atoms, one 2-D tensor copy per tile; a warp owns 16 rows and all columns of every layer, so the forward pass and the
// delta pass need no inter-warp synchronisation; deltas overwrite the activations in place (T' = 1 - H^2 is accurate
// enough at this tolerance); CTA barriers only surround the weight-gradient products, whose K dimension runs over the
// atoms of the tile.  Fragments are those of mma.sync.m16n8k8.tf32 (register operands): tcgen05.mma kind::tf32 was
// probed (tools/umma_probe.cu: K-major operands, TMEM read-back) but its operands live in shared memory in canonical
// layouts, which for the transposed products (K = atoms) means a second, transposed copy of every activation and delta
// tile; with hi / lo pairs that does not fit in shared memory at a useful tile size, so this kernel stays on register
// fragments.
#include "pp_mlp_dev.cuh"
#include <algorithm>
#include <cstring>
#include <vector>

#define M32_WARPS 16
#define M32_THREADS (M32_WARPS * 32)
#define M32_TM (16 * M32_WARPS)      // 256 rows: a warp owns 16

struct Mlp32Plan {
  int L;                               // weight layers
  int n[4];                            // sizes[0..L]
  int NGp;                             // row pitch of the FP32 descriptor array in HBM (multiple of 4 floats)
  int kp[3];                           // pitch of layer l's input buffer (l = 0: X) and of W_l (floats, == 4 mod 8)
  int x_off, h_off[3];                 // floats
  int w_off[3];                        // W_l [rows padded to 8 NT][kp[l]] in FP32 (split into hi / lo when loaded)
  int t_off, tkp;                      // W_1^T [n_1 padded][tkp] for the delta product of a second hidden layer
  int dbl_off_bytes;                   // doubles: erow [TM], cvec [TM], sttgt [TM], red [..]
  int stint_off_bytes, mbar_off_bytes;
  int n_dw_tiles, tab_off_bytes;       // weight-gradient tiles (m16 x n8), dealt round the warps
  size_t smem_bytes;
};

struct Dw32Tile { int layer, mi, ni, pad; };      // rows 16 mi .. of layer l's matrix, columns 8 ni ..

__device__ __forceinline__ unsigned tf32_of(float x)
{
  unsigned r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void split_tf32(float x, unsigned& hi, unsigned& lo)
{
  hi = tf32_of(x);
  lo = tf32_of(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], unsigned b0, unsigned b1)
{
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// tanh and its derivative in FP32 (relative error ~1e-6): e = exp(-2|x|), t = (1 - e) / (1 + e)
__device__ __forceinline__ float tanh32(float x)
{
  const float ax = fabsf(x);
  const float e = __expf(-2.0f * ax);
  const float t = __fdividef(1.0f - e, 1.0f + e);
  return copysignf(t, x);
}

// C[16 rows of this warp][8 NT columns] = A[16][K] * W^T, A = FP32 rows in shared memory, W = FP32 weight rows [n][k];
// both are split into hi / lo as they are loaded.  3 x TF32.
template <int NT>
__device__ __forceinline__ void gemm32(float (&acc)[NT][4], const float* __restrict__ As, int apitch, const float* __restrict__ W, int wpitch,
                                       int K8, int lane)
{
  const int g = lane >> 2, q = lane & 3;
#pragma unroll
  for (int nt = 0; nt < NT; nt++) { acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.0f; }
  const float* a0p = As + (size_t)g * apitch + q;
  const float* a1p = As + (size_t)(g + 8) * apitch + q;
  const float* bp = W + (size_t)g * wpitch + q;
  for (int k0 = 0; k0 < K8; k0 += 8) {
    unsigned ah[4], al[4];
    split_tf32(a0p[k0], ah[0], al[0]);
    split_tf32(a1p[k0], ah[1], al[1]);
    split_tf32(a0p[k0 + 4], ah[2], al[2]);
    split_tf32(a1p[k0 + 4], ah[3], al[3]);
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
      unsigned bh0, bl0, bh1, bl1;
      split_tf32(bp[(size_t)(8 * nt) * wpitch + k0], bh0, bl0);
      split_tf32(bp[(size_t)(8 * nt) * wpitch + k0 + 4], bh1, bl1);
      mma_tf32(acc[nt], al, bh0, bh1);
      mma_tf32(acc[nt], ah, bl0, bl1);
      mma_tf32(acc[nt], ah, bh0, bh1);
    }
  }
}

template <int NT, int NLW>
__global__ void __launch_bounds__(M32_THREADS, 1)
mlp32_fused_kernel(MlpArgs a, Mlp32Plan p, const float* __restrict__ G32, const __grid_constant__ CUtensorMap tmap)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw);
  double* smd = reinterpret_cast<double*>(smem_raw + p.dbl_off_bytes);
  const Dw32Tile* tab = reinterpret_cast<const Dw32Tile*>(smem_raw + p.tab_off_bytes);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  const int L = p.L, NG = p.n[0], kp0 = p.kp[0];
  const NetDev& net = a.net;
  constexpr int TM = M32_TM;
  (void)G32;

  // ---- one-time staging: weights in FP32 (zero padded), bias columns, tables -----------------------------------------------
  for (int l = 0; l < L; l++) {
    const int n_in = p.n[l], n_out = p.n[l + 1], kp = p.kp[l];
    const int rows = (l == L - 1) ? 1 : 8 * NT;
    const double* src = net.params + net.w_off[l];
    for (int r = warp; r < rows; r += M32_WARPS)
      for (int k = lane; k < kp; k += 32) sm[p.w_off[l] + r * kp + k] = (r < n_out && k <= n_in) ? (float)src[(size_t)r * (n_in + 1) + k] : 0.0f;
    if (l >= 1) {
      float* Hs = sm + p.h_off[l];
      for (int r = warp; r < TM; r += M32_WARPS)
        for (int k = lane; k < kp; k += 32) Hs[r * kp + k] = (k == n_in) ? 1.0f : 0.0f;
    }
  }
  if (NLW == 2) {          // W_1^T [n_1][n_2]: the B operand of delta_1 = delta_2 W_1
    const int n1 = p.n[1], n2 = p.n[2];
    const double* src = net.params + net.w_off[1];
    for (int r = warp; r < 8 * NT; r += M32_WARPS)
      for (int k = lane; k < p.tkp; k += 32) sm[p.t_off + r * p.tkp + k] = (r < n1 && k < n2) ? (float)src[(size_t)k * (n1 + 1) + r] : 0.0f;
  }
  for (int i = tid; i < TM * kp0; i += M32_THREADS) sm[p.x_off + i] = 0.0f;
  {
    const int nwords = p.n_dw_tiles * (int)(sizeof(Dw32Tile) / sizeof(int));
    int* dst = reinterpret_cast<int*>(smem_raw + p.tab_off_bytes);
    const int* srcw = reinterpret_cast<const int*>(a.dw_tiles2);
    for (int i = tid; i < nwords; i += M32_THREADS) dst[i] = srcw[i];
  }
  double* erow = smd;                      // [TM]
  double* cvec = smd + TM;                 // [TM]
  double* st_tgt = smd + 2 * TM;           // [TM]
  double* red = reinterpret_cast<double*>(sm + p.x_off);      // [M32_WARPS * 96], epilogue only: aliases the dead descriptor tile
  int* st_lo = reinterpret_cast<int*>(smem_raw + p.stint_off_bytes);
  int* st_tag = st_lo + TM + 1;
  unsigned long long* mbar = reinterpret_cast<unsigned long long*>(smem_raw + p.mbar_off_bytes);
  float* X = sm + p.x_off;

  // weight-gradient accumulators: this warp's tiles tab[warp], tab[warp + 8], ...
  constexpr int MAXT = 4;
  float dW[MAXT][4];
#pragma unroll
  for (int t = 0; t < MAXT; t++) { dW[t][0] = dW[t][1] = dW[t][2] = dW[t][3] = 0.0f; }
  float dwout[3] = {0.0f, 0.0f, 0.0f};
  double sse_w = 0.0, vsse_w = 0.0, nval_w = 0.0;

  const int n_tiles = a.n_tiles;
  struct Geom { int s0, s1, row0, nvalid; };
  auto tile_geom = [&](int tile) {
    Geom gm;
    gm.s0 = gm.s1 = gm.row0 = gm.nvalid = 0;
    if (tile < n_tiles) {
      gm.s0 = a.tile_struct0[tile];
      gm.s1 = a.tile_struct0[tile + 1];
      const long long r0 = a.struct_atom0[gm.s0];
      gm.row0 = (int)r0;
      gm.nvalid = (int)(a.struct_atom0[gm.s1] - r0);
    }
    return gm;
  };
  auto issue_load = [&](const Geom& gm) {
    if (tid == 0) {
      fence_proxy_async();
      mbar_arrive_expect_tx(mbar, (unsigned)TM * (unsigned)kp0 * 4u);
      tma_load_tile(X, &tmap, gm.row0, mbar);
    }
  };
  if (tid == 0) mbar_init(mbar, 1);
  fence_proxy_async();
  __syncthreads();
  Geom g_cur = tile_geom(blockIdx.x), g_nxt = tile_geom(blockIdx.x + gridDim.x);
  if ((int)blockIdx.x < n_tiles) issue_load(g_cur);

  const int rb = 16 * warp;                // this warp's rows
  int n_done = 0;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, n_done++) {
    const int row0 = g_cur.row0, nvalid = g_cur.nvalid;
    const int s_first = g_cur.s0, s_last = g_cur.s1;
    const Geom g_nn = tile_geom(tile + 2 * gridDim.x);
    (void)row0;
    // structure table of this tile (read after barrier A)
    for (int i = tid; i <= s_last - s_first; i += M32_THREADS) {
      const int sidx = s_first + i;
      st_lo[i] = (int)(a.struct_atom0[sidx] - g_cur.row0);
      if (sidx < s_last) { st_tgt[i] = a.target[sidx]; st_tag[i] = a.tag[sidx]; }
    }
    mbar_wait(mbar, (unsigned)n_done & 1u);

    // ---- the rows arrive normalised (NN.cpp:647-651 is applied in FP64 when the FP32 copy is made); column NG is the bias input
    if (lane < 16) X[(size_t)(rb + lane) * kp0 + NG] = 1.0f;
    __syncwarp();

    // ---- forward: this warp's rows through every hidden layer ------------------------------------------------------------
    {
      const float* wo = sm + p.w_off[L - 1];          // output node
      double yp0 = 0.0, yp1 = 0.0;                     // rows rb + g and rb + g + 8
      for (int l = 0; l < L - 1; l++) {
        const int n_in = p.n[l], n_out = p.n[l + 1];
        float acc[NT][4];
        gemm32<NT>(acc, (l == 0 ? X : sm + p.h_off[l]) + (size_t)rb * p.kp[l], p.kp[l], sm + p.w_off[l], p.kp[l], (n_in + 1 + 7) & ~7, lane);
        float* H0 = sm + p.h_off[l + 1] + (size_t)(rb + g) * p.kp[l + 1];
        float* H1 = H0 + (size_t)8 * p.kp[l + 1];
        const bool last = (l == L - 2);
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int n = 8 * nt + 2 * q + h;
            if (n < n_out) {
              const float t0 = tanh32(acc[nt][h]), t1 = tanh32(acc[nt][2 + h]);
              H0[n] = t0;
              H1[n] = t1;
              if (last) { yp0 += (double)t0 * (double)wo[n]; yp1 += (double)t1 * (double)wo[n]; }
            }
          }
        }
        __syncwarp();
      }
      yp0 += __shfl_xor_sync(0xffffffffu, yp0, 1); yp0 += __shfl_xor_sync(0xffffffffu, yp0, 2);
      yp1 += __shfl_xor_sync(0xffffffffu, yp1, 1); yp1 += __shfl_xor_sync(0xffffffffu, yp1, 2);
      if (q == 0) {
        const double b = (double)wo[p.n[L - 1]];
        const int r0 = rb + g, r1 = rb + g + 8;
        erow[r0] = (r0 < nvalid) ? a.out_scale * (yp0 + b) : 0.0;
        erow[r1] = (r1 < nvalid) ? a.out_scale * (yp1 + b) : 0.0;
      }
    }
    __syncthreads();                                                                   // barrier A: erow, structure table

    // ---- per-structure energy, error and loss terms in FP64 (Potential.cpp:641-668): the factor c of this warp's rows ---
    {
      const int re = rb + 16;
      const int ns = s_last - s_first;
      double c_mine = 0.0;                     // lanes 0..15: c of row rb + lane
      for (int i = 0; i < ns; i++) {
        const int r_lo = st_lo[i], r_hi = st_lo[i + 1];
        if (r_hi <= rb) continue;
        if (r_lo >= re) break;
        double e = 0.0;
        for (int r = r_lo + lane; r < r_hi; r += 32) e += erow[r];
        e = pp_warp_sum(e) + a.out_mean;
        const double err = e - st_tgt[i];
        const int tag = st_tag[i];
        const double c = (tag == 0) ? 2.0 * err * a.out_scale : 0.0;
        if (r_lo >= rb && r_lo < re) {
          if (tag == 0) sse_w += err * err;
          else if (tag == 1) { vsse_w += err * err; nval_w += 1.0; }
          if (lane == 0) a.E_struct[s_first + i] = e;
        }
        const int r = rb + lane;
        if (lane < 16 && r >= r_lo && r < r_hi) c_mine = c;
      }
      if (lane < 16) cvec[rb + lane] = c_mine;
      __syncwarp();
    }

    // ---- output-node gradient over this warp's rows, then delta_{L-1} = c w_out (1 - H^2) IN PLACE of H_{L-1} ----------------
    {
      const int n_h = p.n[L - 1], kph = p.kp[L - 1];
      float* Hl = sm + p.h_off[L - 1];
      const float* wo = sm + p.w_off[L - 1];
#pragma unroll
      for (int j = 0; j < 3; j++) {
        const int k = lane + 32 * j;
        if (k <= n_h) {
          float s_ = 0.0f;
          for (int r = 0; r < 16; r++) s_ = fmaf((float)cvec[rb + r], Hl[(size_t)(rb + r) * kph + k], s_);
          dwout[j] += s_;
        }
      }
      __syncwarp();
      for (int r = 0; r < 16; r++) {
        const float c = (float)cvec[rb + r];
        float* hr = Hl + (size_t)(rb + r) * kph;
        for (int n = lane; n < n_h; n += 32) { const float h = hr[n]; hr[n] = c * wo[n] * (1.0f - h * h); }
      }
      __syncwarp();
    }

    // ---- backward: weight gradients dW_l += delta_{l+1}^T [H_l, 1] over the tile's atoms, deltas in place ----------------------
#pragma unroll
    for (int li = 0; li < NLW; li++) {
      const int l = NLW - 1 - li;
      __syncthreads();                                                                 // delta_{l+1} complete in all rows
      float accb[NT][4];
      if (l >= 1)      // (delta_2 W_1) for this warp's rows, kept in registers until the barrier below
        gemm32<NT>(accb, sm + p.h_off[l + 1] + (size_t)rb * p.kp[l + 1], p.kp[l + 1], sm + p.t_off, p.tkp, (p.n[l + 1] + 7) & ~7, lane);
      {
        const float* Dl = sm + p.h_off[l + 1];                          // delta_{l+1} [atom][neuron]
        const float* Bl = (l == 0) ? X : sm + p.h_off[l];               // [H_l, 1]   [atom][input]
        const int dp = p.kp[l + 1], bp = p.kp[l];
#pragma unroll
        for (int t = 0; t < MAXT; t++) {
          const int ti = warp + t * M32_WARPS;
          if (ti < p.n_dw_tiles) {
            const Dw32Tile d = tab[ti];
            if (d.layer == l) {
              const float* pa = Dl + (size_t)q * dp + 16 * d.mi + g;    // A(m = neuron, k = atom) = delta[atom][neuron]
              const float* pb = Bl + (size_t)q * bp + 8 * d.ni + g;     // B(k = atom, n = input)
#pragma unroll 2
              for (int k0 = 0; k0 < TM; k0 += 8) {
                unsigned ah[4], al[4], bh0, bh1, bl0, bl1;
                split_tf32(pa[(size_t)k0 * dp], ah[0], al[0]);
                split_tf32(pa[(size_t)k0 * dp + 8], ah[1], al[1]);
                split_tf32(pa[(size_t)(k0 + 4) * dp], ah[2], al[2]);
                split_tf32(pa[(size_t)(k0 + 4) * dp + 8], ah[3], al[3]);
                split_tf32(pb[(size_t)k0 * bp], bh0, bl0);
                split_tf32(pb[(size_t)(k0 + 4) * bp], bh1, bl1);
                mma_tf32(dW[t], al, bh0, bh1);
                mma_tf32(dW[t], ah, bl0, bl1);
                mma_tf32(dW[t], ah, bh0, bh1);
              }
            }
          }
        }
      }
      if (l >= 1) {
        __syncthreads();                                                               // every product reading H_l is done
        const int n_l = p.n[l];
        float* H0 = sm + p.h_off[l] + (size_t)(rb + g) * p.kp[l];
        float* H1 = H0 + (size_t)8 * p.kp[l];
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int n = 8 * nt + 2 * q + h;
            if (n < n_l) {
              const float h0 = H0[n], h1 = H1[n];
              H0[n] = accb[nt][h] * (1.0f - h0 * h0);
              H1[n] = accb[nt][2 + h] * (1.0f - h1 * h1);
            }
          }
        }
        __syncwarp();
      }
    }
    __syncthreads();                                                                   // X is free: next tile's copy
    if (tile + (int)gridDim.x < n_tiles) issue_load(g_nxt);
    g_cur = g_nxt;
    g_nxt = g_nn;
  }

  // ---- per-CTA partial gradient and loss terms -> scratch (FP64, summed in a fixed order by the reduction kernel) ------------
  double* out = a.partial + (size_t)blockIdx.x * (a.P_total + 3) + a.p_off;
#pragma unroll
  for (int t = 0; t < MAXT; t++) {
    const int ti = warp + t * M32_WARPS;
    if (ti < p.n_dw_tiles) {
      const Dw32Tile d = tab[ti];
      const int n_in = p.n[d.layer], n_out = p.n[d.layer + 1];
#pragma unroll
      for (int h = 0; h < 4; h++) {
        const int neuron = 16 * d.mi + g + ((h & 2) ? 8 : 0);
        const int k = 8 * d.ni + 2 * q + (h & 1);
        if (neuron < n_out && k <= n_in) out[net.w_off[d.layer] + (size_t)neuron * (n_in + 1) + k] = (double)dW[t][h];
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 3; j++) red[warp * 96 + 32 * j + lane] = (double)dwout[j];
  __syncthreads();
  if (tid <= p.n[L - 1]) {
    double s_ = 0.0;
    for (int w = 0; w < M32_WARPS; w++) s_ += red[w * 96 + tid];
    out[net.w_off[L - 1] + tid] = s_;
  }
  __syncthreads();
  if (lane == 0) { red[3 * warp] = sse_w; red[3 * warp + 1] = vsse_w; red[3 * warp + 2] = nval_w; }
  __syncthreads();
  if (tid < 3) {
    double s_ = 0.0;
    for (int w = 0; w < M32_WARPS; w++) s_ += red[3 * w + tid];
    a.partial[(size_t)blockIdx.x * (a.P_total + 3) + a.P_total + tid] = s_;
  }
}

// ---------------------------------------------------------------------------------------------------------
// FP64 descriptors [atoms][NG] -> NORMALISED FP32 [atoms][NGp] (rows padded with zeros to a 16-byte multiple for the
// tensor copy).  The input scaling (x - mean) / scale of NN.cpp:647-651 is applied here in FP64: subtracting the mean
// after rounding to FP32 would amplify the rounding error by |x| / scale.
__global__ void __launch_bounds__(256)
g_to_f32_kernel(const double* __restrict__ G, int NG, int NGp, long long n_atoms, const int* __restrict__ elem, const double* __restrict__ mean,
                const double* __restrict__ scale, float* __restrict__ out)
{
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_atoms * NGp) return;
  const long long atom = i / NGp;
  const int k = (int)(i - atom * NGp);
  const int e = elem ? elem[atom] : 0;
  out[i] = (k < NG) ? (float)((G[atom * NG + k] - mean[(size_t)e * NG + k]) * (1.0 / scale[(size_t)e * NG + k])) : 0.0f;
}

// FP64 -> normalised FP64: exactly the per-tile arithmetic of the training kernels, (x - mean) * (1 / scale), done once
__global__ void __launch_bounds__(256)
g_normalise_kernel(const double* __restrict__ G, int NG, long long n_atoms, const int* __restrict__ elem, const double* __restrict__ mean,
                   const double* __restrict__ inv_scale, double* __restrict__ out)
{
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_atoms * NG) return;
  const long long atom = i / NG;
  const int k = (int)(i - atom * NG);
  const int e = elem ? elem[atom] : 0;
  out[i] = (G[i] - mean[(size_t)e * NG + k]) * inv_scale[(size_t)e * NG + k];
}

int ppk_g_normalise(const double* G, int NG, long long n_atoms, const int* elem, const double* mean, const double* inv_scale, double* out,
                    cudaStream_t st, long long* launches)
{
  if (n_atoms == 0) return 0;
  const long long n = n_atoms * NG;
  g_normalise_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(G, NG, n_atoms, elem, mean, inv_scale, out);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

// mean / scale: [n_elem][NG] on the device
int ppk_g_to_f32(const double* G, int NG, int NGp, long long n_atoms, const int* elem, const double* mean, const double* scale, float* out,
                 cudaStream_t st, long long* launches)
{
  if (n_atoms == 0) return 0;
  const long long n = n_atoms * NGp;
  g_to_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(G, NG, NGp, n_atoms, elem, mean, scale, out);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

static int m32_pitch(int v)
{
  int p = v;
  while ((p & 7) != 4) p++;
  return p;
}

static bool mlp32_make_plan(const NetDev& net, Mlp32Plan& p, std::vector<Dw32Tile>* table)
{
  memset(&p, 0, sizeof(p));
  const int L = net.n_layers;
  if (L < 2 || L > 3) return false;
  int mx = 0;
  for (int l = 1; l < L; l++) {
    if (!((net.tanh_layers >> l) & 1)) return false;         // FP32 mode: tanh hidden layers only
    mx = std::max(mx, net.sizes[l]);
  }
  if (mx > 64) return false;
  p.L = L;
  for (int l = 0; l <= L; l++) p.n[l] = net.sizes[l];
  const int NT = (mx + 7) / 8 <= 5 ? 5 : 8;
  p.NGp = (p.n[0] + 3) & ~3;
  size_t off = 0;
  for (int l = 0; l < L; l++) p.kp[l] = m32_pitch(((p.n[l] + 1 + 7) & ~7));
  if (p.kp[0] > 256) return false;
  p.x_off = 0; off += (size_t)M32_TM * p.kp[0];
  for (int l = 1; l < L; l++) { off = (off + 31) & ~(size_t)31; p.h_off[l] = (int)off; off += (size_t)M32_TM * p.kp[l]; }
  for (int l = 0; l < L; l++) {
    const int rows = (l == L - 1) ? 1 : 8 * NT;
    p.w_off[l] = (int)off; off += (size_t)rows * p.kp[l];
  }
  p.tkp = m32_pitch((mx + 7) & ~7);
  if (L == 3) { p.t_off = (int)off; off += (size_t)8 * NT * p.tkp; }
  size_t bytes = off * sizeof(float);
  bytes = (bytes + 15) & ~(size_t)15;
  p.dbl_off_bytes = (int)bytes;
  bytes += (size_t)(3 * M32_TM) * sizeof(double);
  p.stint_off_bytes = (int)bytes;
  bytes += (size_t)(2 * M32_TM + 2) * sizeof(int);
  bytes = (bytes + 15) & ~(size_t)15;
  p.mbar_off_bytes = (int)bytes;
  bytes += 16;
  p.tab_off_bytes = (int)bytes;
  std::vector<Dw32Tile> tiles;
  for (int l = L - 2; l >= 0; l--) {
    const int rows = (p.n[l + 1] + 15) / 16, cols = (p.n[l] + 1 + 7) / 8;
    for (int mi = 0; mi < rows; mi++)
      for (int ni = 0; ni < cols; ni++) tiles.push_back(Dw32Tile{l, mi, ni, 0});
  }
  if ((int)tiles.size() > 4 * M32_WARPS) return false;
  if ((size_t)M32_WARPS * 96 * sizeof(double) > (size_t)M32_TM * p.kp[0] * sizeof(float)) return false;
  p.n_dw_tiles = (int)tiles.size();
  bytes += tiles.size() * sizeof(Dw32Tile);
  p.smem_bytes = (bytes + 15) & ~(size_t)15;
  if (table) *table = tiles;
  return p.smem_bytes <= 227 * 1024;
}

bool ppk_mlp32_ok(const NetDev& net)
{
  Mlp32Plan p;
  return mlp32_make_plan(net, p, nullptr);
}
int ppk_mlp32_pitch(const NetDev& net) { return (net.sizes[0] + 3) & ~3; }
int ppk_mlp32_tile_atoms(void) { return M32_TM; }
int ppk_mlp32_table(const NetDev& net, void* host_out, int* n_entries)
{
  Mlp32Plan p;
  std::vector<Dw32Tile> t;
  *n_entries = 0;
  if (!mlp32_make_plan(net, p, &t)) return 0;
  *n_entries = (int)t.size();
  if (host_out && !t.empty()) memcpy(host_out, t.data(), t.size() * sizeof(Dw32Tile));
  return (int)sizeof(Dw32Tile);
}

int ppk_mlp32_launch_fused(const MlpArgs& a, const float* G32, long long n_atoms, cudaStream_t st, long long* launches, int* n_cta)
{
  Mlp32Plan p;
  PP_REQUIRE(mlp32_make_plan(a.net, p, nullptr), "internal: network not eligible for the FP32 training kernel");
  PP_REQUIRE(a.dw_tiles2 != nullptr && a.rows == nullptr && a.n_tiles > 0, "internal: FP32 training kernel needs whole-structure tiles");
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  PP_REQUIRE(ppk_tiled_map_f32(G32, p.n[0], p.NGp, n_atoms, p.kp[0], M32_TM, &tmap), "cuTensorMapEncodeTiled failed for the FP32 descriptor array");
  const int grid = ppk_sm_count();
  const int NT = (std::max(p.n[1], p.L == 3 ? p.n[2] : 0) + 7) / 8 <= 5 ? 5 : 8;
  const int NLW = p.L - 1;
#define PP_LAUNCH32(NT_, NLW_)                                                                                  \
  {                                                                                                             \
    auto k = mlp32_fused_kernel<NT_, NLW_>;                                                                     \
    PP_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_bytes));           \
    k<<<grid, M32_THREADS, p.smem_bytes, st>>>(a, p, G32, tmap);                                                \
  }
  if (NT == 5 && NLW == 1) PP_LAUNCH32(5, 1)
  else if (NT == 5) PP_LAUNCH32(5, 2)
  else if (NLW == 1) PP_LAUNCH32(8, 1)
  else PP_LAUNCH32(8, 2)
#undef PP_LAUNCH32
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  *n_cta = grid;
  return 0;
}
