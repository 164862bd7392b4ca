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
#include "pp_mlp_dev.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <type_traits>
#include <vector>

#define MLP_WARPS 16
#define MLP_THREADS (MLP_WARPS * 32)
#define MLP_MAX_WIDTH 128    // widest hidden layer supported (16 n-tiles, 8 per warp of a pair)

// Shared-memory plan of one CTA (all offsets in doubles unless stated).
struct MlpPlan {
  int TM, L;
  int n[PP_MAX_LAYERS + 1];
  int kpitch[PP_MAX_LAYERS];          // pitch of layer l's input buffer (l = 0: raw descriptor tiles X) and of Ws_l
  int dpitch[PP_MAX_LAYERS + 1];      // pitch of Ds_l (T' then delta), l = 1..L-1
  int ws_off[PP_MAX_LAYERS], hs_off[PP_MAX_LAYERS], ds_off[PP_MAX_LAYERS + 1];
  int x_off[2];                       // double-buffered raw descriptor tiles
  int ws_rows[PP_MAX_LAYERS];         // rows of Ws_l held in shared memory (zero padded)
  int cvec_off, erow_off, mu_off, is_off, red_off, sttgt_off, exptab_off, mbar_off;
  int all_tanh[PP_MAX_LAYERS + 1];    // layer l (1..L-1) consists of tanh nodes only
  int stint_off_bytes;                // per-tile structure table: row offsets [TM+1] and tags [TM] (ints)
  int tf_off_bytes, desc_off_bytes;
  int n_dw_tiles;                     // entries of the weight-gradient table (tiles, or 2 * MLP_WARPS block slots)
  int tma;                            // 1: descriptor tiles arrive as ONE 2-D tensor copy per tile (set at launch)
  int blocked;                        // 1: the table holds register-blocked work items, two slots per warp
  size_t smem_bytes;
};


static int plan_nt(const NetDev& net)
{
  int mx = 1;
  for (int l = 1; l < net.n_layers; l++) mx = std::max(mx, net.sizes[l]);
  int nt = ((mx + 7) / 8 + 1) / 2;              // n-tiles per warp of a pair
  return nt <= 2 ? 2 : nt <= 3 ? 3 : nt <= 4 ? 4 : 8;
}

static void make_plan(const NetDev& net, int TM, MlpPlan& p, std::vector<DwTile>* tiles)
{
  const int NTW = 16 * plan_nt(net);             // columns a warp pair covers (zero padded)
  p.TM = TM;
  p.tma = 0;
  p.L = net.n_layers;
  for (int l = 0; l <= p.L; l++) { p.n[l] = net.sizes[l]; p.all_tanh[l] = (net.tanh_layers >> l) & 1; }
  size_t off = 0;
  for (int l = 0; l < p.L; l++) {
    p.kpitch[l] = (l == 0 ? round_up(p.n[l] + 1, 8) : std::max(round_up(p.n[l] + 1, 8), NTW)) + 4;
    p.ws_off[l] = (int)off;
    int rows = (l == p.L - 1) ? 1 : std::max(round_up(p.n[l + 1], 8), NTW);
    p.ws_rows[l] = rows;
    off += (size_t)rows * p.kpitch[l];
  }
  for (int b = 0; b < 2; b++) {          // 128-byte aligned: destination of the tensor copy
    off = (off + 15) & ~(size_t)15;
    p.x_off[b] = (int)off; off += (size_t)TM * p.kpitch[0];
  }
  p.hs_off[0] = p.x_off[0];
  for (int l = 1; l < p.L; l++) { p.hs_off[l] = (int)off; off += (size_t)TM * p.kpitch[l]; }
  for (int l = 1; l < p.L; l++) {
    p.dpitch[l] = std::max(round_up(p.n[l], 8), NTW) + 4;
    p.ds_off[l] = (int)off;
    off += (size_t)TM * p.dpitch[l];
  }
  p.cvec_off = (int)off; off += 8 * (size_t)TM;      // pitch 8: column 0 = c, columns 1..7 stay zero (DMMA A operand)
  p.erow_off = (int)off; off += TM;
  p.mu_off = (int)off; off += p.kpitch[0];
  p.is_off = (int)off; off += p.kpitch[0];
  p.red_off = (int)off; off += 3 * MLP_WARPS;
  p.sttgt_off = (int)off; off += TM;
  p.exptab_off = (int)off; off += 64;          // 2^(j/64) for pp_fast_exp_nonpos
  p.mbar_off = (int)off; off += 2;             // two mbarriers (one per descriptor-tile buffer)
  p.tf_off_bytes = (int)(off * sizeof(double));
  int nodes = 0;
  for (int l = 1; l <= p.L; l++) nodes += p.n[l];
  size_t bytes = off * sizeof(double) + (size_t)nodes * sizeof(int);
  bytes = (bytes + 15) & ~(size_t)15;
  p.stint_off_bytes = (int)bytes;
  bytes += (size_t)(2 * TM + 2) * sizeof(int);
  bytes = (bytes + 15) & ~(size_t)15;
  p.desc_off_bytes = (int)bytes;
  // Work items.  Blocked form when it fits two slots per warp: slot 0 takes any block, slot 1 at most two tiles
  // (so a warp's accumulators stay within 6 tiles = 12 doubles); otherwise single tiles, warp w owning w, w+16, ...
  std::vector<DwTile> blocks;
  for (int l = 0; l < p.L; l++) {
    const int rows = (l == p.L - 1) ? 1 : round_up(p.n[l + 1], 8) / 8, cols = round_up(p.n[l] + 1, 8) / 8;
    for (int mi = 0; mi < rows; mi += 2)
      for (int ni = 0; ni < cols; ni += 2) {
        DwTile d;
        memset(&d, 0, sizeof(d));
        d.a_off = (l == p.L - 1 ? p.cvec_off : p.ds_off[l + 1]) + 8 * mi;
        d.b_off = (l == 0 ? 0 : p.hs_off[l]) + 8 * ni;
        d.dp = (l == p.L - 1) ? 8 : p.dpitch[l + 1];
        d.kp = p.kpitch[l];
        d.layer = l; d.mi = mi; d.ni = ni;
        d.nm = std::min(2, rows - mi); d.nn = std::min(2, cols - ni);
        blocks.push_back(d);
      }
  }
  std::stable_sort(blocks.begin(), blocks.end(), [](const DwTile& x, const DwTile& y) { return x.nm * x.nn > y.nm * y.nn; });
  int n_big = 0;
  for (const DwTile& d : blocks) n_big += (d.nm * d.nn > 2);
  p.blocked = (n_big <= MLP_WARPS && (int)blocks.size() <= 2 * MLP_WARPS) ? 1 : 0;
  int t = 0;
  if (p.blocked) {
    std::vector<DwTile> table(2 * MLP_WARPS);
    for (DwTile& d : table) memset(&d, 0, sizeof(d));
    int load[MLP_WARPS] = {0}, smsp[4] = {0, 0, 0, 0};
    size_t b = 0;
    for (; b < blocks.size() && b < (size_t)MLP_WARPS; b++) {      // largest first into the slot 0 of warp 0, 1, ...
      table[2 * b] = blocks[b];
      load[b] += blocks[b].nm * blocks[b].nn;
      smsp[b & 3] += blocks[b].nm * blocks[b].nn;
    }
    for (; b < blocks.size(); b++) {                                 // the rest: least loaded scheduler, then least loaded warp
      int best = -1;
      for (int w = 0; w < MLP_WARPS; w++) {
        if (table[2 * w + 1].nm) continue;
        if (best < 0 || smsp[w & 3] < smsp[best & 3] || (smsp[w & 3] == smsp[best & 3] && load[w] < load[best])) best = w;
      }
      table[2 * best + 1] = blocks[b];
      load[best] += blocks[b].nm * blocks[b].nn;
      smsp[best & 3] += blocks[b].nm * blocks[b].nn;
    }
    if (tiles) *tiles = table;
    t = 2 * MLP_WARPS;
  } else {
    for (int l = 0; l < p.L - 1; l++) {
      const int nm = round_up(p.n[l + 1], 8) / 8, nn = round_up(p.n[l] + 1, 8) / 8;
      for (int mi = 0; mi < nm; mi++)
        for (int ni = 0; ni < nn; ni++) {
          if (tiles) {
            DwTile d;
            memset(&d, 0, sizeof(d));
            d.a_off = p.ds_off[l + 1] + 8 * mi;
            d.b_off = (l == 0 ? 0 : p.hs_off[l]) + 8 * ni;
            d.dp = p.dpitch[l + 1];
            d.kp = p.kpitch[l];
            d.layer = l; d.mi = mi; d.ni = ni; d.nm = 1; d.nn = 1;
            tiles->push_back(d);
          }
          t++;
        }
    }
  }
  p.n_dw_tiles = t;
  bytes += (size_t)t * sizeof(DwTile);
  p.smem_bytes = (bytes + 15) & ~(size_t)15;
}

// ---------------------------------------------------------------------------------------------------------

// Phase timing for tuning (build with -DPP_MLP_TIMING; not part of the product build): warp 0 and the last warp of
// CTA 0 accumulate clock64() deltas per phase.
#ifdef PP_MLP_TIMING
__device__ unsigned long long pp_mlp_cycles[2][16];
#define PP_T(i)                                                                              \
  do {                                                                                       \
    if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == MLP_WARPS - 1)) {              \
      const long long now_ = clock64();                                                      \
      pp_mlp_cycles[warp ? 1 : 0][i] += (unsigned long long)(now_ - tprev_);                 \
      tprev_ = now_;                                                                         \
    }                                                                                        \
  } while (0)
extern "C" int pp_debug_mlp_cycles(unsigned long long* out, int reset)
{
  if (out && cudaMemcpyFromSymbol(out, pp_mlp_cycles, sizeof(pp_mlp_cycles)) != cudaSuccess) return 1;
  if (reset) {
    unsigned long long z[2][16] = {};
    if (cudaMemcpyToSymbol(pp_mlp_cycles, z, sizeof(z)) != cudaSuccess) return 1;
  }
  return 0;
}
#else
#define PP_T(i)
#endif

// MODE: PP_MLP_FORWARD (energies only), PP_MLP_TRAIN (two-phase backward: c per structure given),
//       PP_MLP_INPUT_GRAD (dE/dG, MD), PP_MLP_FUSED (single pass: forward, per-structure error, backward; needs
//       tiles made of whole structures of a single-element system)
template <int MODE, int NT, int OWN>
__global__ void __launch_bounds__(MLP_THREADS, 1)
mlp_kernel(MlpArgs a, MlpPlan p, const __grid_constant__ CUtensorMap tmap)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sm = reinterpret_cast<double*>(smem_raw);
  int* tfs = reinterpret_cast<int*>(smem_raw + p.tf_off_bytes);
  const DwTile* dwt = reinterpret_cast<const DwTile*>(smem_raw + p.desc_off_bytes);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  const int L = p.L, TM = p.TM, NG = p.n[0];
  const NetDev& net = a.net;
  constexpr bool BACKWARD = (MODE != PP_MLP_FORWARD);
  constexpr bool GRADW = (MODE == PP_MLP_TRAIN || MODE == PP_MLP_FUSED);

  // ---- one-time staging: weights (zero padded), constant columns, transfer ids, input scaling ----------------
  for (int l = 0; l < L; l++) {
    const int n_in = p.n[l], n_out = p.n[l + 1], kp = p.kpitch[l];
    const int rows = p.ws_rows[l];
    double* Ws = sm + p.ws_off[l];
    const double* src = net.params + net.w_off[l];
    for (int r = warp; r < rows; r += MLP_WARPS)
      for (int k = lane; k < kp; k += 32) Ws[r * kp + k] = (r < n_out && k <= n_in) ? src[(size_t)r * (n_in + 1) + k] : 0.0;
    if (l == 0) {
      // raw descriptor tiles: pad column NG holds 1 (the bias input), the rest 0; mu/is make (x - mu)*is the identity there
      for (int b = 0; b < 2; b++) {
        double* X = sm + p.x_off[b];
        for (int r = warp; r < TM; r += MLP_WARPS)
          for (int k = lane; k < kp; k += 32) X[r * kp + k] = (k == n_in) ? 1.0 : 0.0;
      }
      for (int k = tid; k < kp; k += MLP_THREADS) {
        sm[p.mu_off + k] = (k < NG) ? net.mean[k] : 0.0;
        sm[p.is_off + k] = (k < NG) ? 1.0 / net.scale[k] : (k == NG ? 1.0 : 0.0);
      }
    } else {
      double* Hs = sm + p.hs_off[l];
      for (int r = warp; r < TM; r += MLP_WARPS)
        for (int k = lane; k < kp; k += 32) Hs[r * kp + k] = (k == n_in) ? 1.0 : 0.0;
      double* Ds = sm + p.ds_off[l];
      for (int i = tid; i < TM * p.dpitch[l]; i += MLP_THREADS) Ds[i] = 0.0;
    }
  }
  {
    int nodes = 0;
    for (int l = 1; l <= L; l++) nodes += p.n[l];
    for (int i = tid; i < nodes; i += MLP_THREADS) tfs[i] = net.tf[i];
    const int nwords = p.n_dw_tiles * (int)(sizeof(DwTile) / sizeof(int));
    int* dst = reinterpret_cast<int*>(smem_raw + p.desc_off_bytes);
    const int* srcw = reinterpret_cast<const int*>(a.dw_tiles);
    for (int i = tid; i < nwords; i += MLP_THREADS) dst[i] = srcw[i];
  }
  for (int i = tid; i < 64; i += MLP_THREADS) sm[p.exptab_off + i] = exp2((double)i * 0.015625);
  for (int i = tid; i < 8 * TM; i += MLP_THREADS) sm[p.cvec_off + i] = 0.0;      // columns 1..7 stay zero
  const double* mu = sm + p.mu_off;
  const double* is = sm + p.is_off;
  const double* exptab = sm + p.exptab_off;

  // ---- parameter-gradient accumulators ------------------------------------------------------------------------
  double dW[OWN][2];
  double dwout = 0.0;                         // thread t <= n_{L-1}: gradient of output-node parameter t
  double sse_w = 0.0, vsse_w = 0.0, nval_w = 0.0;   // FUSED: per-warp loss terms (lane 0)
#pragma unroll
  for (int t = 0; t < OWN; t++) { dW[t][0] = 0.0; dW[t][1] = 0.0; }

  double* cvec = sm + p.cvec_off;
  double* erow = sm + p.erow_off;
  const int n_tiles = (MODE == PP_MLP_FUSED) ? a.n_tiles : (a.M + TM - 1) / TM;
  const int kp0 = p.kpitch[0];
  const bool vec16 = (NG & 1) == 0;

  // tile geometry (first / one-past-last structure, first row, row count); FUSED tiles come from two dependent global
  // loads, so the geometry of tile t+2 is requested while tile t is being processed
  struct Geom { int s0, s1, row0, nvalid; };
  auto tile_geom = [&](int tile) {
    Geom gm;
    gm.s0 = gm.s1 = gm.row0 = gm.nvalid = 0;
    if (tile < n_tiles) {
      if (MODE == PP_MLP_FUSED) {
        gm.s0 = a.tile_struct0[tile];
        gm.s1 = a.tile_struct0[tile + 1];
        const long long r0 = a.struct_atom0[gm.s0];
        gm.row0 = (int)r0;
        gm.nvalid = (int)(a.struct_atom0[gm.s1] - r0);
      } else {
        gm.row0 = tile * TM;
        gm.nvalid = min(TM, a.M - gm.row0);
      }
    }
    return gm;
  };
  int* st_lo = reinterpret_cast<int*>(smem_raw + p.stint_off_bytes);     // [TM+1] first row of each structure of the tile
  int* st_tag = st_lo + TM + 1;                                            // [TM]
  double* st_tgt = sm + p.sttgt_off;                                       // [TM]
  // Asynchronous copy of one tile's raw descriptor rows into X[buf]: one tensor copy per tile, issued by one thread and
  // counted in bytes on the buffer's mbarrier (p.tma), or cp.async per thread when the rows are gathered.
  unsigned long long* mbar = reinterpret_cast<unsigned long long*>(sm + p.mbar_off);
  const bool bulk = p.tma != 0;
  auto issue_load = [&](const Geom& gm, int buf) {
    const int row0 = gm.row0, nvalid = gm.nvalid;
    double* X = sm + p.x_off[buf];
    if (bulk) {
      // rows are consecutive atoms (no gather): one tensor copy of the whole box; rows past nvalid belong to the next
      // structure (or are zero past the end of the array) and are masked further down like any padding row
      if (tid == 0) {
        fence_proxy_async();            // the buffer was last read through the generic proxy (previous tile's gradients)
        mbar_arrive_expect_tx(mbar + buf, (unsigned)TM * (unsigned)kp0 * 8u);
        tma_load_tile(X, &tmap, row0, mbar + buf);
      }
      return;
    }
    for (int r = warp; r < TM; r += MLP_WARPS) {
      double* dst = X + (size_t)r * kp0;
      if (r < nvalid) {
        const size_t atom = a.rows ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
        const double* src = a.G + atom * NG;
        if (vec16) { for (int c = lane; c < (NG >> 1); c += 32) cp_async16(dst + 2 * c, src + 2 * c); }
        else { for (int c = lane; c < NG; c += 32) cp_async8(dst + c, src + c); }
      } else {
        for (int c = lane; c < NG; c += 32) dst[c] = 0.0;
      }
    }
    cp_async_commit();
  };
  if (bulk && tid == 0) {
    mbar_init(mbar, 1);
    mbar_init(mbar + 1, 1);
    fence_proxy_async();
  }
  __syncwarp();

  int buf = 0;
  Geom g_cur = tile_geom(blockIdx.x), g_nxt = tile_geom(blockIdx.x + gridDim.x);
  if ((int)blockIdx.x < n_tiles) issue_load(g_cur, 0);
  __syncthreads();

  // warps w and w + 8 share m-tile w; with warp % 4 = scheduler, every scheduler hosts two first and two second halves
  const int mt = warp & (MLP_WARPS / 2 - 1), half = warp / (MLP_WARPS / 2);
#ifdef PP_MLP_TIMING
  const long long tstart_ = clock64();
#endif
  int n_done = 0;      // tiles this CTA has finished: buffer = n_done & 1, mbarrier phase = (n_done >> 1) & 1
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, buf ^= 1, n_done++) {
    const int row0 = g_cur.row0, nvalid = g_cur.nvalid;
    const int s_first = g_cur.s0, s_last = g_cur.s1;
    const Geom g_nn = tile_geom(tile + 2 * gridDim.x);      // consumed one iteration from now
#ifdef PP_MLP_TIMING
    long long tprev_ = clock64();
#endif
    if (bulk) mbar_wait(mbar + buf, (unsigned)(n_done >> 1) & 1u);
    else cp_async_wait_all();
    __syncthreads();
    PP_T(0);
    if (tile + (int)gridDim.x < n_tiles) issue_load(g_nxt, buf ^ 1);
    if (MODE == PP_MLP_FUSED) {
      // structure table of this tile; its global-load latency hides behind the forward pass (read after the next barrier)
      for (int i = tid; i <= s_last - s_first; i += MLP_THREADS) {
        const int sidx = s_first + i;
        st_lo[i] = (int)(a.struct_atom0[sidx] - row0);
        if (sidx < s_last) { st_tgt[i] = a.target[sidx]; st_tag[i] = a.tag[sidx]; }
      }
    }
    const double* X = sm + p.x_off[buf];
    PP_T(1);
    if (MODE == PP_MLP_TRAIN || MODE == PP_MLP_INPUT_GRAD) {
      for (int r = tid; r < TM; r += MLP_THREADS) {
        double c = 0.0;
        if (r < nvalid) {
          const size_t atom = a.rows ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
          c = (MODE == PP_MLP_TRAIN) ? a.cstruct[a.atom_struct[atom]] : a.out_scale;
        }
        cvec[8 * r] = c;
      }
    }

    // ---- descriptor normalisation (NN.cpp:647-651), in place and once per tile: the warp pair that owns 8 rows
    // rewrites them, so the layer-0 product and the layer-0 weight gradients read plain operands
    if (mt < TM / 8) {
      double* xr = sm + p.x_off[buf] + (size_t)(8 * mt + g) * kp0;
      if (!a.prenorm) {
#pragma unroll 4
        for (int c = 4 * half + q; c < NG; c += 8) xr[c] = (xr[c] - mu[c]) * is[c];
      }
      if (bulk && half == 0 && q == 0) xr[NG] = 1.0;      // the tensor copy zero-fills the padding; column NG is the bias input
      pair_barrier(mt);
    }
    PP_T(14);
    // ---- forward through the hidden layers: a pair of warps per m-tile (8 atoms), n-tiles split between them ----
    if (mt < TM / 8) {
      // one hidden layer for this warp's n-tiles [nt0, nt0 + NTC)
      auto fwd_layer = [&](int l, auto ntc_tag) {
        constexpr int NTC = decltype(ntc_tag)::value;
        const int n_in = p.n[l], n_out = p.n[l + 1];
        const int nt0 = half * NT;
        double acc[NTC][2];
        if (l == 0) tile_gemm<false, false, NTC>(acc, X, kp0, sm + p.ws_off[0], kp0, mt, round_up(n_in + 1, 4), nt0, NTC, lane);
        else tile_gemm<false, false, NTC>(acc, sm + p.hs_off[l], p.kpitch[l], sm + p.ws_off[l], p.kpitch[l], mt, round_up(n_in + 1, 4), nt0, NTC, lane);
        PP_T(2);
        double* Hn = sm + p.hs_off[l + 1] + (size_t)(8 * mt + g) * p.kpitch[l + 1];
        double* Dn = sm + p.ds_off[l + 1] + (size_t)(8 * mt + g) * p.dpitch[l + 1];
        if (p.all_tanh[l + 1]) {
          // every node of the layer is a tanh: evaluate the thread's 2*NTC values as straight-line code, so that
          // their dependent FP64 chains (exp, reciprocal) overlap instead of running one after the other
          double tv[NTC][2], dv[NTC][2];
#pragma unroll
          for (int nt = 0; nt < NTC; nt++) {
#pragma unroll
            for (int h = 0; h < 2; h++) tv[nt][h] = pp_fast_tanh(acc[nt][h], dv[nt][h], exptab);
          }
#pragma unroll
          for (int nt = 0; nt < NTC; nt++) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const int n = 8 * (nt0 + nt) + 2 * q + h;
              if (n < n_out) {
                Hn[n] = tv[nt][h];
                if (BACKWARD) Dn[n] = dv[nt][h];
              }
            }
          }
        } else {
          const int* tfl = tfs + net.node_off[l];
#pragma unroll
          for (int nt = 0; nt < NTC; nt++) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const int n = 8 * (nt0 + nt) + 2 * q + h;
              if (n < n_out) {
                double d;
                double t = pp_transfer(tfl[n], acc[nt][h], d, a.tanh_table, exptab);
                Hn[n] = t;
                if (BACKWARD) Dn[n] = d;
              }
            }
          }
        }
      };
      for (int l = 0; l < L - 1; l++) {
        // a warp whose last n-tile lies wholly in the padding (the second half of an odd tile count) skips it
        if (NT > 1 && 8 * (half * NT + NT - 1) >= p.n[l + 1]) fwd_layer(l, std::integral_constant<int, (NT > 1 ? NT - 1 : 1)>{});
        else fwd_layer(l, std::integral_constant<int, NT>{});
        PP_T(3);
        pair_barrier(mt);
        PP_T(4);
      }
      // ---- output node (always linear, NN.cpp:76): y = w_out . [H, 1];  4 lanes per atom, first warp of the pair ----
      if (half == 0) {
        const int n_in = p.n[L - 1];
        const double* wo = sm + p.ws_off[L - 1];
        double y = 0.0;
        if (L == 1) {
          const double* xr = X + (size_t)(8 * mt + g) * kp0;
          for (int k = q; k <= n_in; k += 4) y = fma(xr[k], wo[k], y);
        } else {
          const double* Hl = sm + p.hs_off[L - 1] + (size_t)(8 * mt + g) * p.kpitch[L - 1];
          for (int k = q; k <= n_in; k += 4) y = fma(Hl[k], wo[k], y);
        }
        y += __shfl_xor_sync(0xffffffffu, y, 1);
        y += __shfl_xor_sync(0xffffffffu, y, 2);
        const int r = 8 * mt + g;
        if (q == 0) {
          const double e = a.out_scale * y;
          if (MODE == PP_MLP_FUSED) erow[r] = (r < nvalid) ? e : 0.0;
          else if (MODE != PP_MLP_TRAIN && a.E_atom != nullptr && r < nvalid) {
            const size_t atom = a.rows ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
            a.E_atom[atom] = e;
          }
        }
      }
    }
    PP_T(5);
    if (MODE == PP_MLP_FUSED) {
      // ---- per-structure energy, error and loss terms (Potential.cpp:641-668) --------------------------------
      // After this one barrier every warp works out the error factor c of ITS OWN 8 rows by summing the energies
      // of the structures that overlap them (redundantly with the other warps those structures touch: a few dozen
      // shared-memory loads), so no second barrier is needed before the backward pass.  The structure's outputs
      // are written by the warp whose rows contain the structure's first atom.
      __syncthreads();
      PP_T(6);
      if (mt < TM / 8) {
        const int rb = 8 * mt, re = 8 * mt + 8;
        const int ns = s_last - s_first;
        if (lane < 8) cvec[8 * (rb + lane)] = 0.0;
        __syncwarp();
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
          if (half == 0 && r_lo >= rb && r_lo < re) {
            if (tag == 0) sse_w += err * err;
            else if (tag == 1) { vsse_w += err * err; nval_w += 1.0; }
            if (lane == 0) a.E_struct[s_first + i] = e;
          }
          const int r = rb + lane;
          if (lane < 8 && r >= r_lo && r < r_hi) cvec[8 * r] = c;
        }
        __syncwarp();
      }
    } else if (BACKWARD) {
      __syncthreads();   // cvec
    }

    PP_T(7);
    if (BACKWARD && mt < TM / 8) {
      double acc[NT][2];
      // ---- delta of the last hidden layer: c * w_out[n] * T'(z[n])  (output node linear) ------------------------
      if (L >= 2) {
        {
          const int n_h = p.n[L - 1];
          const int nt0 = half * NT, ntc = NT;
          double* Dr = sm + p.ds_off[L - 1] + (size_t)(8 * mt + g) * p.dpitch[L - 1];
          const double* wo = sm + p.ws_off[L - 1];
          const double c = cvec[8 * (8 * mt + g)];
          const int n_end = min(n_h, 8 * (nt0 + ntc));
          for (int n = 8 * nt0 + q; n < n_end; n += 4) Dr[n] *= c * wo[n];
        }
        pair_barrier(mt);
        PP_T(8);
        // delta_l = (delta_{l+1} * W_l[:, :n_l]) .* T'(z_l) for this warp's n-tiles [nt0, nt0 + NTC)
        auto bwd_layer = [&](int l, auto ntc_tag) {
          constexpr int NTC = decltype(ntc_tag)::value;
          const int n_l = p.n[l], n_up = p.n[l + 1];
          const int nt0 = half * NT;
          double accb[NTC][2];
          tile_gemm<true, false, NTC>(accb, sm + p.ds_off[l + 1], p.dpitch[l + 1], sm + p.ws_off[l], p.kpitch[l], mt, round_up(n_up, 4), nt0, NTC, lane);
          double* Dr = sm + p.ds_off[l] + (size_t)(8 * mt + g) * p.dpitch[l];
#pragma unroll
          for (int nt = 0; nt < NTC; nt++) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const int n = 8 * (nt0 + nt) + 2 * q + h;
              if (n < n_l) Dr[n] *= accb[nt][h];
            }
          }
        };
        for (int l = L - 2; l >= 1; l--) {
          if (NT > 1 && 8 * (half * NT + NT - 1) >= p.n[l]) bwd_layer(l, std::integral_constant<int, (NT > 1 ? NT - 1 : 1)>{});
          else bwd_layer(l, std::integral_constant<int, NT>{});
          PP_T(9);
          pair_barrier(mt);
          PP_T(10);
        }
      }
      if (MODE == PP_MLP_INPUT_GRAD) {
        // dE/dG_k = (delta_1 * W_0[:, k]) / scale_k   (NN.cpp:546-566); without hidden layers: c * w_out[k] / scale_k
        const int r = 8 * mt + g;
        size_t atom = 0;
        if (r < nvalid) atom = a.rows ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
        if (L >= 2) {
          const int ntl = (NG + 7) >> 3;
          const int h0 = half ? (ntl + 1) >> 1 : 0;
          const int hc = half ? ntl - h0 : (ntl + 1) >> 1;
          for (int c0 = 0; c0 < hc; c0 += NT) {
            const int ntc = min(NT, hc - c0);
            tile_gemm<true, true, NT>(acc, sm + p.ds_off[1], p.dpitch[1], sm + p.ws_off[0], kp0, mt, round_up(p.n[1], 4), h0 + c0, ntc, lane);
            if (r < nvalid) {
#pragma unroll
              for (int nt = 0; nt < NT; nt++) {
                if (nt < ntc) {
#pragma unroll
                  for (int h = 0; h < 2; h++) {
                    const int k = 8 * (h0 + c0 + nt) + 2 * q + h;
                    if (k < NG) a.dEdG[atom * NG + k] = acc[nt][h] * is[k];
                  }
                }
              }
            }
          }
        } else if (r < nvalid && half == 0) {
          const double* wo = sm + p.ws_off[0];
          for (int k = q; k < NG; k += 4) a.dEdG[atom * NG + k] = cvec[8 * r] * wo[k] * is[k];
        }
      }
    }

    // ---- parameter gradients: dW_l += delta_{l+1}^T * [H_l, 1]  (K = atoms of the tile) ------------------------
    if constexpr (GRADW) {
      __syncthreads();
      PP_T(11);
      if constexpr (OWN == 6) {
        // register-blocked work items, two slots per warp (make_plan); the output node is one of the "layers"
                {
          const DwTile d = dwt[2 * warp];
          if (d.nm) {
            const double* pa = sm + d.a_off + g + (size_t)q * d.dp;        // A(m = neuron, k = atom) = delta[atom][neuron]
            const double* pb = (d.layer == 0 ? X : sm) + d.b_off + g + (size_t)q * d.kp;   // B(k = atom, n = input)
            const int sa = 4 * d.dp, sb = 4 * d.kp;
            if (d.nm == 2 && d.nn == 2) dw_block<OWN, 0, 2, 2>(dW, pa, pb, sa, sb, TM);
            else if (d.nn == 2) dw_block<OWN, 0, 1, 2>(dW, pa, pb, sa, sb, TM);
            else if (d.nm == 2) dw_block<OWN, 0, 2, 1>(dW, pa, pb, sa, sb, TM);
            else dw_block<OWN, 0, 1, 1>(dW, pa, pb, sa, sb, TM);
          }
        }
        PP_T(12);
        {
          const DwTile d = dwt[2 * warp + 1];
          if (d.nm) {
            const double* pa = sm + d.a_off + g + (size_t)q * d.dp;
            const double* pb = (d.layer == 0 ? X : sm) + d.b_off + g + (size_t)q * d.kp;
            const int sa = 4 * d.dp, sb = 4 * d.kp;
            if (d.nn == 2) dw_block<OWN, 4, 1, 2>(dW, pa, pb, sa, sb, TM);
            else if (d.nm == 2) dw_block<OWN, 4, 2, 1>(dW, pa, pb, sa, sb, TM);
            else dw_block<OWN, 4, 1, 1>(dW, pa, pb, sa, sb, TM);
          }
        }
      } else {
        {
          // output-node gradient on the threads of the LAST warps (they own the fewest weight-gradient tiles)
          const int n_in = p.n[L - 1];
          const int j = MLP_THREADS - 1 - tid;
          if (j <= n_in) {
            double s = 0.0;
            const double* Hl = (L == 1 ? X : sm + p.hs_off[L - 1]) + j;
            const int kp = (L == 1) ? kp0 : p.kpitch[L - 1];
            for (int r = 0; r < TM; r++) s = fma(cvec[8 * r], Hl[(size_t)r * kp], s);
            dwout += s;
          }
        }
        PP_T(12);
#pragma unroll
        for (int t = 0; t < OWN; t++) {
          const int tile_id = warp + t * MLP_WARPS;
          if (tile_id < p.n_dw_tiles) {
            const DwTile d = dwt[tile_id];
            const double* pa = sm + d.a_off + g + (size_t)q * d.dp;
            const double* pb = (d.layer == 0 ? X : sm) + d.b_off + g + (size_t)q * d.kp;
            const int sa = 4 * d.dp, sb = 4 * d.kp;
            double c0 = dW[t][0], c1 = dW[t][1];
#pragma unroll 4
            for (int k0 = 0; k0 < TM; k0 += 4) {
              const double av = *pa;
              const double bv = *pb;
              pa += sa; pb += sb;
              dmma884(c0, c1, av, bv);
            }
            dW[t][0] = c0; dW[t][1] = c1;
          }
        }
      }
      PP_T(13);
    }
    g_cur = g_nxt;
    g_nxt = g_nn;
  }
  cp_async_wait_all();
#ifdef PP_MLP_TIMING
  if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == MLP_WARPS - 1)) {
    pp_mlp_cycles[warp ? 1 : 0][15] += (unsigned long long)(clock64() - tstart_);
  }
#endif

  // ---- per-CTA partial gradient (+ loss terms) -> scratch ------------------------------------------------------
  if constexpr (GRADW) {
    double* out = a.partial + (size_t)blockIdx.x * (a.P_total + 3) + a.p_off;
    if constexpr (OWN == 6) {
      // slot 0 holds tiles 0..3 (t = i*nn + j), slot 1 tiles 4..5
#pragma unroll
      for (int t = 0; t < 6; t++) {
        const DwTile d = dwt[2 * warp + (t >> 2)];
        const int tt = t & 3;
        if (tt < d.nm * d.nn) {
          const int bi = (d.nn == 2) ? (tt >> 1) : tt, bj = (d.nn == 2) ? (tt & 1) : 0;
          const int neuron = 8 * (d.mi + bi) + g;
          const int n_in = p.n[d.layer];
          if (neuron < p.n[d.layer + 1]) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const int k = 8 * (d.ni + bj) + 2 * q + h;
              if (k <= n_in) out[net.w_off[d.layer] + (size_t)neuron * (n_in + 1) + k] = dW[t][h];
            }
          }
        }
      }
    } else {
      {
        const int n_in = p.n[L - 1];
        const int j = MLP_THREADS - 1 - tid;
        if (j <= n_in) out[net.w_off[L - 1] + j] = dwout;
      }
#pragma unroll
      for (int t = 0; t < OWN; t++) {
        const int tile_id = warp + t * MLP_WARPS;
        if (tile_id < p.n_dw_tiles) {
          const DwTile d = dwt[tile_id];
          const int neuron = 8 * d.mi + g;
          const int n_in = p.n[d.layer];
          if (neuron < p.n[d.layer + 1]) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const int k = 8 * d.ni + 2 * q + h;
              if (k <= n_in) out[net.w_off[d.layer] + (size_t)neuron * (n_in + 1) + k] = dW[t][h];
            }
          }
        }
      }
    }
    if (MODE == PP_MLP_FUSED) {
      double* red = sm + p.red_off;
      __syncthreads();
      if (lane == 0) { red[3 * warp] = sse_w; red[3 * warp + 1] = vsse_w; red[3 * warp + 2] = nval_w; }
      __syncthreads();
      if (tid < 3) {
        double s = 0.0;
        for (int w = 0; w < MLP_WARPS; w++) s += red[3 * w + tid];
        a.partial[(size_t)blockIdx.x * (a.P_total + 3) + a.P_total + tid] = s;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// cuTensorMapEncodeTiled through the runtime (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled()
{
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)ptr;
  }
  return fn;
}

// Tensor map of an [atoms][NG] FP64 array with box (box_cols x box_rows); the last few are cached, since an epoch loop
// asks for the same one every time (encoding costs a few microseconds of host time in front of the launch).
bool ppk_tiled_map(const double* G, int NG, long long atoms, int box_cols, int box_rows, void* out_map)
{
  CUtensorMap* out = reinterpret_cast<CUtensorMap*>(out_map);
  struct Entry { const double* G; int NG, bc, br; long long atoms; CUtensorMap map; bool used; };
  static Entry cache[64];
  static int next = 0;
  for (int i = 0; i < 64; i++) {
    const Entry& e = cache[i];
    if (e.used && e.G == G && e.NG == NG && e.atoms == atoms && e.bc == box_cols && e.br == box_rows) { *out = e.map; return true; }
  }
  if (encode_tiled() == nullptr) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)NG, (cuuint64_t)atoms};
  const cuuint64_t gstride[1] = {(cuuint64_t)NG * sizeof(double)};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  if (encode_tiled()(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(G), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  Entry& e = cache[next];
  next = (next + 1) & 63;
  e.G = G; e.NG = NG; e.atoms = atoms; e.bc = box_cols; e.br = box_rows; e.map = *out; e.used = true;
  return true;
}

// the same for an FP32 array whose rows are `pitch_floats` apart (FP32 mode); not cached: one per epoch launch
bool ppk_tiled_map_f32(const float* base, int cols, int pitch_floats, long long rows, int box_cols, int box_rows, void* out_map)
{
  if (encode_tiled() == nullptr) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)pitch_floats * sizeof(float)};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return encode_tiled()(reinterpret_cast<CUtensorMap*>(out_map), CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box,
                        estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static int g_sm_count = 0;
int ppk_sm_count(void)
{
  if (!g_sm_count) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

int ppk_mlp_single_ctas(void) { return ppk_sm_count(); }       // CTAs (= partial rows) of the one-CTA-per-SM kernels
int ppk_mlp_max_ctas(void) { return 2 * ppk_sm_count(); }      // rows of the partial-gradient scratch (two-CTA form)

// Largest tile (multiple of 8 atoms, <= 64) whose shared-memory plan fits; 0 if none.
static int choose_tm(const NetDev& net, MlpPlan& plan, std::vector<DwTile>* tiles)
{
  for (int TM = 64; TM >= 8; TM -= 8) {
    if (tiles) tiles->clear();
    make_plan(net, TM, plan, tiles);
    if (plan.smem_bytes <= 227 * 1024) return TM;
  }
  return 0;
}

int ppk_mlp_tile_atoms(const NetDev& net)
{
  MlpPlan plan;
  return choose_tm(net, plan, nullptr);
}

int ppk_mlp_num_dw_tiles(const NetDev& net, void* host_tiles_out)
{
  MlpPlan plan;
  std::vector<DwTile> tiles;
  if (!choose_tm(net, plan, &tiles)) return -1;
  if (host_tiles_out && !tiles.empty()) memcpy(host_tiles_out, tiles.data(), tiles.size() * sizeof(DwTile));
  return (int)tiles.size();
}

int ppk_mlp_dw_tile_bytes(void) { return (int)sizeof(DwTile); }

// =========================================================================================================
// Fused training kernel, second form: TWO resident CTAs per SM.
//
// The first form keeps activations H_l, derivative/delta buffers D_l and a double-buffered descriptor tile in
// shared memory (~200 KB), so one CTA owns an SM and its 16 warps move through the phases of a tile in lockstep:
// while they evaluate transfer functions or the per-structure error the DMMA pipe idles.  Here the derivatives T' wait
// in a per-CTA scratch in global memory (L2 resident, see `dsc` in the kernel) and the deltas overwrite the
// activations IN PLACE once the products that need the activations are done.  The working set shrinks to
// X + H_1 .. H_{L-1} + the weights (~100 KB for 50-40-40-1), two CTAs fit on an SM, and one CTA's scalar phases
// overlap the other's tensor phases.  A warp owns 8 atoms (rows) and ALL columns of every layer, so the forward and
// delta passes need no inter-warp synchronisation at all; CTA barriers only surround the weight-gradient products,
// whose K dimension runs over all atoms of the tile.
//
// Eligibility (checked on the host, anything else runs the first form): an even descriptor count (16-byte rows for
// the copy engine), 1 or 2 hidden layers, each all tanh or all tanh_spline, no wider than 8*NTW, a shared-memory plan of
// at most 113 KB, 64-atom tiles.  FUSED additionally needs a single-element system (tiles of whole structures);
// multi-element systems run FORWARD and TRAIN through it with their rows gathered one bulk copy per atom.
#define MLP2_WARPS 8
#define MLP2_THREADS (MLP2_WARPS * 32)
#define MLP2_TM 64

struct Mlp2Plan {
  int L;
  int n[PP_MAX_LAYERS + 1];
  int kp[PP_MAX_LAYERS];              // pitch of layer l's input buffer (l = 0: X) and of W_l
  int w_off[PP_MAX_LAYERS], w_rows[PP_MAX_LAYERS];
  int x_off, h_off[PP_MAX_LAYERS];    // h_off[l]: H_l, l = 1 .. L-1 (later delta_l)
  int mu_off, is_off, exptab_off, cvec_off, erow_off, sttgt_off, red_off, mbar_off;
  int spline[PP_MAX_LAYERS + 1];      // hidden layer l uses tanh_spline (else tanh)
  int stint_off_bytes, tab_off_bytes;
  size_t smem_bytes;
};

// smallest pitch >= v that keeps 8-row fragment loads free of bank conflicts (pitch mod 16 in {4, 12})
static int mlp2_pitch(int v)
{
  int p = v;
  while (!((p & 15) == 4 || (p & 15) == 12)) p++;
  return p;
}

// Work table of the weight-gradient products: for every hidden-layer matrix l < L-1, two slots per warp.
static bool mlp2_make_plan(const NetDev& net, int NTW, Mlp2Plan& p, std::vector<DwTile>* table)
{
  memset(&p, 0, sizeof(p));
  p.L = net.n_layers;
  for (int l = 0; l <= p.L; l++) { p.n[l] = net.sizes[l]; p.spline[l] = ((net.tanh_layers >> l) & 1) ? 0 : (int)((net.spline_layers >> l) & 1); }
  size_t off = 0;
  for (int l = 0; l < p.L; l++) {
    p.kp[l] = mlp2_pitch(p.n[l] + 1);
    p.w_off[l] = (int)off;
    p.w_rows[l] = (l == p.L - 1) ? 1 : 8 * NTW;
    off += (size_t)p.w_rows[l] * p.kp[l];
  }
  off = (off + 15) & ~(size_t)15;
  p.x_off = (int)off; off += (size_t)MLP2_TM * p.kp[0];
  for (int l = 1; l < p.L; l++) { p.h_off[l] = (int)off; off += (size_t)MLP2_TM * p.kp[l]; }
  p.mu_off = (int)off; off += p.kp[0];
  p.is_off = (int)off; off += p.kp[0];
  p.exptab_off = (int)off; off += 64;
  p.cvec_off = (int)off; off += MLP2_TM;
  p.erow_off = (int)off; off += MLP2_TM;
  p.sttgt_off = (int)off; off += MLP2_TM;
  p.red_off = p.x_off;                                     // epilogue only (cross-warp sums): aliases the dead descriptor tile
  p.mbar_off = (int)off; off += 2;
  size_t bytes = off * sizeof(double);
  p.stint_off_bytes = (int)bytes;
  bytes += (size_t)(2 * MLP2_TM + 2) * sizeof(int);
  bytes = (bytes + 15) & ~(size_t)15;
  p.tab_off_bytes = (int)bytes;
  std::vector<DwTile> all;
  for (int l = 0; l < p.L - 1; l++) {
    std::vector<DwTile> blocks;
    const int rows = round_up(p.n[l + 1], 8) / 8, cols = round_up(p.n[l] + 1, 8) / 8;
    for (int mi = 0; mi < rows; mi += 2)
      for (int ni = 0; ni < cols; ni += 2) {
        DwTile d;
        memset(&d, 0, sizeof(d));
        d.a_off = p.h_off[l + 1] + 8 * mi;
        d.b_off = (l == 0 ? p.x_off : p.h_off[l]) + 8 * ni;
        d.dp = p.kp[l + 1];
        d.kp = p.kp[l];
        d.layer = l; d.mi = mi; d.ni = ni;
        d.nm = std::min(2, rows - mi); d.nn = std::min(2, cols - ni);
        blocks.push_back(d);
      }
    std::stable_sort(blocks.begin(), blocks.end(), [](const DwTile& x, const DwTile& y) { return x.nm * x.nn > y.nm * y.nn; });
    int n_big = 0;
    for (const DwTile& d : blocks) n_big += (d.nm * d.nn > 2);
    if (n_big > MLP2_WARPS || (int)blocks.size() > 2 * MLP2_WARPS) return false;
    std::vector<DwTile> slots(2 * MLP2_WARPS);
    for (DwTile& d : slots) memset(&d, 0, sizeof(d));
    int load[MLP2_WARPS] = {0}, smsp[4] = {0, 0, 0, 0};
    size_t b = 0;
    for (; b < blocks.size() && b < (size_t)MLP2_WARPS; b++) {
      slots[2 * b] = blocks[b];
      load[b] += blocks[b].nm * blocks[b].nn;
      smsp[b & 3] += blocks[b].nm * blocks[b].nn;
    }
    for (; b < blocks.size(); b++) {
      int best = -1;
      for (int w = 0; w < MLP2_WARPS; w++) {
        if (slots[2 * w + 1].nm) continue;
        if (best < 0 || smsp[w & 3] < smsp[best & 3] || (smsp[w & 3] == smsp[best & 3] && load[w] < load[best])) best = w;
      }
      slots[2 * best + 1] = blocks[b];
      load[best] += blocks[b].nm * blocks[b].nn;
      smsp[best & 3] += blocks[b].nm * blocks[b].nn;
    }
    all.insert(all.end(), slots.begin(), slots.end());
  }
  bytes += all.size() * sizeof(DwTile);
  p.smem_bytes = (bytes + 15) & ~(size_t)15;
  if (table) *table = all;
  return true;
}

// Stored TRAIN launch of mlp_fused2_kernel: this thread's entries of every H_l tile (row 8 warp + g, columns 8 nt + 2 q + h,
// the ones it owns in the forward pass) by asynchronous 16-byte copies from the activation store, and the matching T'
// lines pulled into L2 for the loads that follow later in the tile.  A function of its own, reached through `if constexpr`:
// as a lambda inside the kernel its mere definition cost the single-pass instantiation 48 bytes of spills.
template <int NTW>
__device__ __forceinline__ void mlp2_issue_hidden(const MlpArgs& a, const Mlp2Plan& p, double* sm, int row0, int nvalid, int warp, int g, int q,
                                                  int SW, bool store_vec)
{
  const int r = 8 * warp + g;
  const bool rv = r < nvalid;
  const size_t rowoff = (size_t)(row0 + r) * SW;
  int off = 0;
  for (int l = 1; l < p.L; l++) {
    const int n_l = p.n[l];
    double* Hr = sm + p.h_off[l] + (size_t)r * p.kp[l];
    const bool vec = store_vec && ((off | n_l) & 1) == 0;
#pragma unroll
    for (int nt = 0; nt < NTW; nt++) {
      const int n = 8 * nt + 2 * q;
      if (n < n_l) {
        if (!rv) { Hr[n] = 0.0; if (n + 1 < n_l) Hr[n + 1] = 0.0; }
        else if (vec) {
          cp_async16(Hr + n, a.hstore + rowoff + off + n);
          asm volatile("prefetch.global.L2 [%0];" ::"l"(a.tstore + rowoff + off + n));
        } else {
          cp_async8(Hr + n, a.hstore + rowoff + off + n);
          if (n + 1 < n_l) cp_async8(Hr + n + 1, a.hstore + rowoff + off + n + 1);
        }
      }
    }
    off += n_l;
  }
  cp_async_commit();
}

// MODE: PP_MLP_FUSED (single pass over tiles of whole structures), PP_MLP_TRAIN (backward of the two-phase epoch: the
// error factor of every atom's structure is given, rows may be gathered through a.rows), PP_MLP_FORWARD (energies).
// NTW: n-tiles (8 columns) of the widest hidden layer; NLW: hidden-layer weight matrices (L - 1)
// SPLINE: some hidden layer uses tanh_spline (compiled out otherwise, so the all-tanh kernel keeps its registers)
template <int MODE, int NTW, int NLW, bool SPLINE>
__global__ void __launch_bounds__(MLP2_THREADS, 2)
mlp_fused2_kernel(MlpArgs a, Mlp2Plan p, const __grid_constant__ CUtensorMap tmap)
{
  constexpr bool FUSED = (MODE == PP_MLP_FUSED);
  // PP_MLP_TRAIN_STORED: the TRAIN launch of the stored form as an instantiation of its own -- the forward products and
  // transfer functions are not compiled in, which frees the registers the early loads of T' and of the row indices need
  constexpr bool TRAIN_ = (MODE == PP_MLP_TRAIN || MODE == PP_MLP_TRAIN_STORED);
  constexpr bool ST_ = (MODE == PP_MLP_TRAIN_STORED);
  constexpr bool GRADW = (MODE != PP_MLP_FORWARD);
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sm = reinterpret_cast<double*>(smem_raw);
  const DwTile* tab = reinterpret_cast<const DwTile*>(smem_raw + p.tab_off_bytes);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  const int L = p.L, NG = p.n[0], kp0 = p.kp[0];
  const NetDev& net = a.net;
  constexpr int TM = MLP2_TM;

  // ---- one-time staging ------------------------------------------------------------------------------------------
  for (int l = 0; l < L; l++) {
    const int n_in = p.n[l], n_out = p.n[l + 1], kp = p.kp[l];
    double* Ws = sm + p.w_off[l];
    const double* src = net.params + net.w_off[l];
    for (int r = warp; r < p.w_rows[l]; r += MLP2_WARPS)
      for (int k = lane; k < kp; k += 32) Ws[r * kp + k] = (r < n_out && k <= n_in) ? src[(size_t)r * (n_in + 1) + k] : 0.0;
    if (l >= 1) {
      double* Hs = sm + p.h_off[l];       // column n_in is the bias input; the in-place deltas never touch it
      for (int r = warp; r < TM; r += MLP2_WARPS)
        for (int k = lane; k < kp; k += 32) Hs[r * kp + k] = (k == n_in) ? 1.0 : 0.0;
    }
  }
  for (int k = tid; k < kp0; k += MLP2_THREADS) {
    sm[p.mu_off + k] = (k < NG) ? net.mean[k] : 0.0;
    sm[p.is_off + k] = (k < NG) ? 1.0 / net.scale[k] : 0.0;
  }
  for (int i = tid; i < 64; i += MLP2_THREADS) sm[p.exptab_off + i] = exp2((double)i * 0.015625);
  for (int i = tid; i < TM * kp0; i += MLP2_THREADS) sm[p.x_off + i] = 0.0;      // padding columns (row copies leave them alone)
  {
    const int nwords = NLW * 2 * MLP2_WARPS * (int)(sizeof(DwTile) / sizeof(int));
    int* dst = reinterpret_cast<int*>(smem_raw + p.tab_off_bytes);
    const int* srcw = reinterpret_cast<const int*>(a.dw_tiles2);
    for (int i = tid; i < nwords; i += MLP2_THREADS) dst[i] = srcw[i];
  }
  const double* mu = sm + p.mu_off;
  const double* is = sm + p.is_off;
  const double* exptab = sm + p.exptab_off;
  double* cvec = sm + p.cvec_off;
  double* erow = sm + p.erow_off;
  double* red = sm + p.red_off;
  int* st_lo = reinterpret_cast<int*>(smem_raw + p.stint_off_bytes);
  int* st_tag = st_lo + TM + 1;
  double* st_tgt = sm + p.sttgt_off;
  unsigned long long* mbar = reinterpret_cast<unsigned long long*>(sm + p.mbar_off);
  double* X = sm + p.x_off;
  // The derivative T' of every activation is parked in a per-CTA scratch in global memory (2 * NTW * NLW doubles per
  // thread, element-major so that a warp's accesses are contiguous; 20 KB per layer and CTA, which never leaves L2) and
  // read back by the SAME thread when it forms that activation's delta -- no synchronisation involved.  T' could be
  // recomputed as 1 - H^2, but that is 0 or 2e-16 for a saturated node: harmless for the 1e-10 tolerance, not for
  // Rprop, which follows the SIGN of a dead neuron's gradient, a sum of nothing but such terms.
  double* dsc = a.scratch2 + (size_t)blockIdx.x * (NLW * NTW * 2 * MLP2_THREADS);
  // Two-phase epoch with STORED activations (several elements): FORWARD also writes H_l and T'_l to HBM, row-major
  // [M][store_w]; TRAIN copies the H_l tiles back (cp.async, each thread the entries it owns in the forward pass) and takes
  // T' from the store where it forms the deltas, so that the forward products and the transfer functions run once.
  const bool prenorm_ = FUSED && a.prenorm != 0;
  const bool stored = ST_ || (MODE == PP_MLP_FORWARD && a.hstore != nullptr);      // compile-time in the gradient launches
  const int SW = a.store_w;
  const bool store_vec = (SW & 1) == 0;      // 16-byte accesses need even offsets; every layer width is checked below
  double dW[NLW][6][2];                       // per hidden-layer matrix: slot 0 = tiles 0..3, slot 1 = tiles 4..5
  double dwout[2] = {0.0, 0.0};               // output node: this warp's rows, parameters lane and lane + 32
  double sse_w = 0.0, vsse_w = 0.0, nval_w = 0.0;
#pragma unroll
  for (int l = 0; l < NLW; l++)
#pragma unroll
    for (int t = 0; t < 6; t++) { dW[l][t][0] = 0.0; dW[l][t][1] = 0.0; }

  const int n_tiles = FUSED ? a.n_tiles : (a.M + TM - 1) / TM;
  struct Geom { int s0, s1, row0, nvalid; };
  auto tile_geom = [&](int tile) {
    Geom gm;
    gm.s0 = gm.s1 = gm.row0 = gm.nvalid = 0;
    if (!FUSED) {
      if (tile < n_tiles) { gm.row0 = tile * TM; gm.nvalid = min(TM, a.M - gm.row0); }
      return gm;
    }
    if (tile < n_tiles) {
      gm.s0 = a.tile_struct0[tile];
      gm.s1 = a.tile_struct0[tile + 1];
      const long long r0 = a.struct_atom0[gm.s0];
      gm.row0 = (int)r0;
      gm.nvalid = (int)(a.struct_atom0[gm.s1] - r0);
    }
    return gm;
  };
  // entry `tid` of a tile's structure table, fetched one tile ahead into registers
  int pf_lo = 0, pf_tag = 0;
  double pf_tgt = 0.0;
  auto st_fetch = [&](const Geom& gm) {
    if (FUSED && tid <= gm.s1 - gm.s0) {
      const int sidx = gm.s0 + tid;
      pf_lo = (int)(a.struct_atom0[sidx] - gm.row0);
      if (sidx < gm.s1) { pf_tgt = a.target[sidx]; pf_tag = a.tag[sidx]; }
    }
  };
  // One tensor copy per tile when the rows are consecutive atoms; gathered rows (an element of a multi-element system)
  // come as one bulk copy per row, issued by the warp that owns the row (lanes 0..7), all counted on the same mbarrier.
  const bool gather = !FUSED && (a.rows != nullptr);      // fused tiles are whole structures of consecutive atoms
  int pf_row = -1;                            // gathered rows: atom index of this lane's row of the NEXT tile
  auto issue_load = [&](const Geom& gm) {
    if (!gather) {
      if (tid == 0) {
        fence_proxy_async();            // X was last written through the generic proxy (normalisation, bias column)
        mbar_arrive_expect_tx(mbar, (unsigned)TM * (unsigned)kp0 * 8u);
        tma_load_tile(X, &tmap, gm.row0, mbar);
      }
      return;
    }
    if (tid == 0) mbar_arrive_expect_tx(mbar, (unsigned)gm.nvalid * (unsigned)NG * 8u);
    if (lane < 8) {
      fence_proxy_async();              // this lane's row was rewritten in place through the generic proxy
      const int r = 8 * warp + lane;
      double* dst = X + (size_t)r * kp0;
      // the row's atom index was fetched a tile ago (pf_row): the copy does not wait on a dependent global load
      if (r < gm.nvalid) bulk_copy_g2s(dst, a.G + (size_t)(pf_row >= 0 ? pf_row : a.rows[gm.row0 + r]) * NG, (unsigned)NG * 8u, mbar);
      else for (int c = 0; c < NG; c++) dst[c] = 0.0;
    }
  };
  if (tid == 0) mbar_init(mbar, 1);
  fence_proxy_async();                         // staging stores (and the barrier's init) before the copy engine's writes
  __syncthreads();
  Geom g_cur = tile_geom(blockIdx.x), g_nxt = tile_geom(blockIdx.x + gridDim.x);
  if ((int)blockIdx.x < n_tiles) {
    issue_load(g_cur);
    if constexpr (TRAIN_) { if (stored) mlp2_issue_hidden<NTW>(a, p, sm, g_cur.row0, g_cur.nvalid, warp, g, q, SW, store_vec); }
  }
  st_fetch(g_cur);

  const int rb = 8 * warp;                     // this warp's rows
  int n_done = 0;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, n_done++) {
    const int row0 = g_cur.row0, nvalid = g_cur.nvalid;
    const int s_first = g_cur.s0, s_last = g_cur.s1;
    const Geom g_nn = tile_geom(tile + 2 * gridDim.x);
    if constexpr (MODE == PP_MLP_FORWARD || ST_) {      // (in the recomputing TRAIN launch one more live value costs 200 bytes of spills per tile)
      if (gather && lane < 8) {
        const int r = 8 * warp + lane;
        pf_row = (r < g_nxt.nvalid) ? a.rows[g_nxt.row0 + r] : -1;
      }
    }
    if (FUSED) {
      // structure table of this tile (read after barrier A); then the next tile's entries start their flight
      if (tid <= s_last - s_first) {
        st_lo[tid] = pf_lo;
        if (tid < s_last - s_first) { st_tgt[tid] = pf_tgt; st_tag[tid] = pf_tag; }
      }
      st_fetch(g_nxt);
    } else if (TRAIN_) {
      // error factor of each row's structure (struct_errors_kernel); read after barrier A
      if (tid < TM) {
        double c = 0.0;
        if (tid < nvalid) {
          const size_t atom = gather ? (size_t)a.rows[row0 + tid] : (size_t)(row0 + tid);
          c = a.cstruct[a.atom_struct[atom]];
        }
        cvec[tid] = c;
      }
    }
    // stored TRAIN: T' of the last hidden layer is needed right after barrier A; its loads fly while the tile's copies land
    double tpl[(TRAIN_) ? NTW : 1][2];
    if constexpr (TRAIN_) {
      if (stored) {
        int off_last0 = 0;
        for (int l = 1; l < L - 1; l++) off_last0 += p.n[l];
        const int n_h = p.n[L - 1];
        const bool ok = (rb + g) < nvalid, vec = store_vec && ((off_last0 | n_h) & 1) == 0;
        const double* src = a.tstore + (size_t)(row0 + rb + g) * SW + off_last0;
#pragma unroll
        for (int nt = 0; nt < NTW; nt++) {
          const int n = 8 * nt + 2 * q;
          tpl[nt][0] = tpl[nt][1] = 0.0;
          if (ok && n < n_h) {
            if (vec) { const double2 v = *reinterpret_cast<const double2*>(src + n); tpl[nt][0] = v.x; tpl[nt][1] = v.y; }
            else { tpl[nt][0] = src[n]; if (n + 1 < n_h) tpl[nt][1] = src[n + 1]; }
          }
        }
      }
    }
    mbar_wait(mbar, (unsigned)n_done & 1u);

    // ---- normalise this warp's 8 rows in place (NN.cpp:647-651); column NG is the bias input ---------------------
    {
      double* xr = X + (size_t)(rb + g) * kp0;
      // (a normalised copy of the descriptors, when the caller keeps one, saves this pass.  The flag is read once at the top
      // of the single-pass instantiation and at this point in the others: each placement is the one ptxas does not answer
      // with spills -- 112 bytes per tile and 4 % of the epoch in the single-pass kernel otherwise)
      if (FUSED ? !prenorm_ : !a.prenorm) {
#pragma unroll 4
        for (int c = q; c < NG; c += 4) xr[c] = (xr[c] - mu[c]) * is[c];
      }
      if (q == 0) xr[NG] = 1.0;
      __syncwarp();
    }

    // ---- forward: this warp's rows through every hidden layer, no inter-warp synchronisation -------------------------
    const size_t srow = (size_t)(row0 + rb + g) * SW;      // this thread's row of the activation store
    const bool srow_ok = (rb + g) < nvalid;
    auto store_pair = [&](double* base, int off, int n_l, int nt, double v0, double v1) {
      const int n = 8 * nt + 2 * q;
      if (!srow_ok || n >= n_l) return;
      double* dst = base + srow + off + n;
      if (store_vec && ((off | n_l) & 1) == 0) *reinterpret_cast<double2*>(dst) = make_double2(v0, v1);
      else { dst[0] = v0; if (n + 1 < n_l) dst[1] = v1; }
    };
    // T' of this thread's activations (nt, h = 0 / 1) of hidden layer lh (0-based): from the store, else from the scratch
    auto tprime = [&](int lh, int off, int n_l, int nt, double& t0, double& t1) {
      if (TRAIN_ && stored) {
        const int n = 8 * nt + 2 * q;
        t0 = t1 = 0.0;
        if (srow_ok && n < n_l) {
          const double* src = a.tstore + srow + off + n;
          if (store_vec && ((off | n_l) & 1) == 0) { const double2 v = *reinterpret_cast<const double2*>(src); t0 = v.x; t1 = v.y; }
          else { t0 = src[0]; if (n + 1 < n_l) t1 = src[1]; }
        }
      } else {
        t0 = dsc[(size_t)((lh * NTW + nt) * 2 + 0) * MLP2_THREADS + tid];
        t1 = dsc[(size_t)((lh * NTW + nt) * 2 + 1) * MLP2_THREADS + tid];
      }
    };
    int off_last = 0;                            // column offset of the last hidden layer in the store
    if (TRAIN_)
      for (int l = 1; l < L - 1; l++) off_last += p.n[l];
    if (TRAIN_ && stored) {
      cp_async_wait_all();                      // this thread's entries of the H_l tiles have landed
      __syncwarp();
    } else {
      const double* wo = sm + p.w_off[L - 1];
      double yp = 0.0;
      int soff = 0;
      for (int l = 0; l < L - 1; l++) {
        const int n_in = p.n[l], n_out = p.n[l + 1];
        double acc[NTW][2];
        tile_gemm<false, false, NTW>(acc, l == 0 ? X : sm + p.h_off[l], p.kp[l], sm + p.w_off[l], p.kp[l], warp,
                                     round_up(n_in + 1, 4), 0, NTW, lane);
        double* Hn = sm + p.h_off[l + 1] + (size_t)(rb + g) * p.kp[l + 1];
        const bool spline_layer = SPLINE && p.spline[l + 1] != 0;      // uniform per layer: tanh_spline (the reference's default) or tanh
        double tv[NTW][2];
        if (MODE == PP_MLP_FORWARD) {           // energies only; the stored form keeps T' (and H, below) for the TRAIN launch
#pragma unroll
          for (int nt = 0; nt < NTW; nt++) {
            double dv0, dv1;
            if (!spline_layer) { tv[nt][0] = pp_fast_tanh(acc[nt][0], dv0, exptab); tv[nt][1] = pp_fast_tanh(acc[nt][1], dv1, exptab); }
            else { tv[nt][0] = pp_fast_tanh_spline(acc[nt][0], dv0, a.tanh_table); tv[nt][1] = pp_fast_tanh_spline(acc[nt][1], dv1, a.tanh_table); }
            if (stored) store_pair(a.tstore, soff, n_out, nt, dv0, dv1);
          }
        } else if (!spline_layer) {
#pragma unroll
          for (int nt = 0; nt < NTW; nt++) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              double dv;
              tv[nt][h] = pp_fast_tanh(acc[nt][h], dv, exptab);
              dsc[(size_t)((l * NTW + nt) * 2 + h) * MLP2_THREADS + tid] = dv;      // T' of this activation, see `dsc`
            }
          }
        } else {
#pragma unroll
          for (int nt = 0; nt < NTW; nt++) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              double dv;
              tv[nt][h] = pp_fast_tanh_spline(acc[nt][h], dv, a.tanh_table);
              dsc[(size_t)((l * NTW + nt) * 2 + h) * MLP2_THREADS + tid] = dv;
            }
          }
        }
        if (MODE == PP_MLP_FORWARD && stored) {
#pragma unroll
          for (int nt = 0; nt < NTW; nt++) store_pair(a.hstore, soff, n_out, nt, tv[nt][0], tv[nt][1]);
        }
        soff += n_out;
        const bool last = (l == L - 2);
#pragma unroll
        for (int nt = 0; nt < NTW; nt++) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int n = 8 * nt + 2 * q + h;
            if (n < n_out) {
              Hn[n] = tv[nt][h];
              if (last) yp = fma(tv[nt][h], wo[n], yp);      // output node (linear, NN.cpp:76) folded in
            }
          }
        }
        __syncwarp();
      }
      if (MODE != PP_MLP_TRAIN) {
        yp += __shfl_xor_sync(0xffffffffu, yp, 1);
        yp += __shfl_xor_sync(0xffffffffu, yp, 2);
        if (q == 0) {
          const int r = rb + g;
          const double e = a.out_scale * (yp + wo[p.n[L - 1]]);
          if (FUSED) erow[r] = (r < nvalid) ? e : 0.0;
          else if (r < nvalid && a.E_atom != nullptr) a.E_atom[gather ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r)] = e;
        }
      }
    }
    __syncthreads();                                                                   // barrier A: erow / cvec, structure table
    if (MODE == PP_MLP_FORWARD) {              // energies only: X is free again
      if (tile + (int)gridDim.x < n_tiles) issue_load(g_nxt);
      g_cur = g_nxt;
      g_nxt = g_nn;
      continue;
    }

    // ---- per-structure energy, error and loss terms (Potential.cpp:641-668): the factor c of this warp's own rows --
    if (FUSED) {
      const int re = rb + 8;
      const int ns = s_last - s_first;
      double c_mine = 0.0;                     // lanes 0..7: c of row rb + lane
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
        if (lane < 8 && r >= r_lo && r < r_hi) c_mine = c;
      }
      if (lane < 8) cvec[rb + lane] = c_mine;
      __syncwarp();
    }

    // ---- output-node gradient over this warp's rows, then delta_{L-1} = c w_out T' IN PLACE of H_{L-1} ---------------
    {
      const int n_h = p.n[L - 1], kph = p.kp[L - 1];
      double* Hl = sm + p.h_off[L - 1];
      const double* wo = sm + p.w_off[L - 1];
#pragma unroll
      for (int j = 0; j < 2; j++) {
        const int k = lane + 32 * j;
        if (k <= n_h) {
          double s_ = 0.0;
#pragma unroll
          for (int r = 0; r < 8; r++) s_ = fma(cvec[rb + r], Hl[(size_t)(rb + r) * kph + k], s_);
          dwout[j] += s_;
        }
      }
      __syncwarp();
      double* Hr = Hl + (size_t)(rb + g) * kph;
      const double c = cvec[rb + g];
#pragma unroll
      for (int nt = 0; nt < NTW; nt++) {
        if (TRAIN_) {
          double tp[2];
          if (stored) { tp[0] = tpl[nt][0]; tp[1] = tpl[nt][1]; }
          else tprime(L - 2, off_last, n_h, nt, tp[0], tp[1]);
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int n = 8 * nt + 2 * q + h;
            if (n < n_h) Hr[n] = c * wo[n] * tp[h];
          }
        } else {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int n = 8 * nt + 2 * q + h;
            const double tp = dsc[(size_t)(((L - 2) * NTW + nt) * 2 + h) * MLP2_THREADS + tid];
            if (n < n_h) Hr[n] = c * wo[n] * tp;
          }
        }
      }
      __syncwarp();
    }

    // ---- backward: weight gradients dW_l += delta_{l+1}^T [H_l, 1] (all atoms of the tile), deltas in place --------------
#pragma unroll
    for (int li = 0; li < NLW; li++) {
      const int l = NLW - 1 - li;
      __syncthreads();                                                                 // delta_{l+1} complete in all rows
      double accb[NTW][2];
      // stored TRAIN: T' of layer l is fetched from the store while the product below runs and multiplied in right after it:
      // its latency hides behind the product and only the deltas stay live across the weight-gradient blocks
      double tpm[ST_ ? NTW : 1][2];
      if constexpr (ST_) {
        if (l >= 1) {
          int off_l0 = 0;
          for (int l2 = 1; l2 < l; l2++) off_l0 += p.n[l2];
          const int n_l = p.n[l];
#pragma unroll
          for (int nt = 0; nt < NTW; nt++) tprime(l - 1, off_l0, n_l, nt, tpm[nt][0], tpm[nt][1]);
        }
      }
      if (l >= 1)      // (delta_{l+1} W_l) for this warp's rows, kept in registers until the barrier below
        tile_gemm<true, false, NTW>(accb, sm + p.h_off[l + 1], p.kp[l + 1], sm + p.w_off[l], p.kp[l], warp, round_up(p.n[l + 1], 4), 0,
                                    NTW, lane);
      if constexpr (ST_) {
        if (l >= 1) {
#pragma unroll
          for (int nt = 0; nt < NTW; nt++) { accb[nt][0] *= tpm[nt][0]; accb[nt][1] *= tpm[nt][1]; }
        }
      }
      {
        const DwTile d = tab[(l * MLP2_WARPS + warp) * 2];
        if (d.nm) {
          const double* pa = sm + d.a_off + g + (size_t)q * d.dp;
          const double* pb = sm + d.b_off + g + (size_t)q * d.kp;
          const int sa = 4 * d.dp, sb = 4 * d.kp;
          if (d.nm == 2 && d.nn == 2) dw_block<6, 0, 2, 2>(dW[l], pa, pb, sa, sb, TM);
          else if (d.nn == 2) dw_block<6, 0, 1, 2>(dW[l], pa, pb, sa, sb, TM);
          else if (d.nm == 2) dw_block<6, 0, 2, 1>(dW[l], pa, pb, sa, sb, TM);
          else dw_block<6, 0, 1, 1>(dW[l], pa, pb, sa, sb, TM);
        }
      }
      {
        const DwTile d = tab[(l * MLP2_WARPS + warp) * 2 + 1];
        if (d.nm) {
          const double* pa = sm + d.a_off + g + (size_t)q * d.dp;
          const double* pb = sm + d.b_off + g + (size_t)q * d.kp;
          const int sa = 4 * d.dp, sb = 4 * d.kp;
          if (d.nn == 2) dw_block<6, 4, 1, 2>(dW[l], pa, pb, sa, sb, TM);
          else if (d.nm == 2) dw_block<6, 4, 2, 1>(dW[l], pa, pb, sa, sb, TM);
          else dw_block<6, 4, 1, 1>(dW[l], pa, pb, sa, sb, TM);
        }
      }
      if (l >= 1) {
        __syncthreads();                                                               // every product reading H_l is done
        const int n_l = p.n[l];
        double* Hr = sm + p.h_off[l] + (size_t)(rb + g) * p.kp[l];
        int off_l = 0;
        if (TRAIN_)
          for (int l2 = 1; l2 < l; l2++) off_l += p.n[l2];
#pragma unroll
        for (int nt = 0; nt < NTW; nt++) {
          if (TRAIN_) {
            double tp[2] = {1.0, 1.0};                     // stored form: T' is in the deltas already (multiplied in above)
            if (!ST_) tprime(l - 1, off_l, n_l, nt, tp[0], tp[1]);
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const int n = 8 * nt + 2 * q + h;
              if (n < n_l) Hr[n] = accb[nt][h] * tp[h];
            }
          } else {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const int n = 8 * nt + 2 * q + h;
              const double tp = dsc[(size_t)(((l - 1) * NTW + nt) * 2 + h) * MLP2_THREADS + tid];
              if (n < n_l) Hr[n] = accb[nt][h] * tp;
            }
          }
        }
        __syncwarp();
      }
    }
    __syncthreads();                                                                   // X is free: next tile's copy
    if (tile + (int)gridDim.x < n_tiles) {
      issue_load(g_nxt);
      if constexpr (TRAIN_) { if (stored) mlp2_issue_hidden<NTW>(a, p, sm, g_nxt.row0, g_nxt.nvalid, warp, g, q, SW, store_vec); }
    }
    g_cur = g_nxt;
    g_nxt = g_nn;
  }

  // ---- per-CTA partial gradient and loss terms -> scratch (summed in a fixed order by reduce_partials_kernel) -----------
  if (!GRADW) return;
  double* out = a.partial + (size_t)blockIdx.x * (a.P_total + 3) + a.p_off;
#pragma unroll
  for (int l = 0; l < NLW; l++) {
#pragma unroll
    for (int t = 0; t < 6; t++) {
      const DwTile d = tab[(l * MLP2_WARPS + warp) * 2 + (t >> 2)];
      const int tt = t & 3;
      if (tt < d.nm * d.nn) {
        const int bi = (d.nn == 2) ? (tt >> 1) : tt, bj = (d.nn == 2) ? (tt & 1) : 0;
        const int neuron = 8 * (d.mi + bi) + g;
        const int n_in = p.n[l];
        if (neuron < p.n[l + 1]) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int k = 8 * (d.ni + bj) + 2 * q + h;
            if (k <= n_in) out[net.w_off[l] + (size_t)neuron * (n_in + 1) + k] = dW[l][t][h];
          }
        }
      }
    }
  }
  __syncthreads();
  red[warp * 64 + lane] = dwout[0];
  red[warp * 64 + 32 + lane] = dwout[1];
  __syncthreads();
  if (tid <= p.n[L - 1]) {
    double s_ = 0.0;
    for (int w = 0; w < MLP2_WARPS; w++) s_ += red[w * 64 + tid];
    out[net.w_off[L - 1] + tid] = s_;
  }
  if (!FUSED) return;                          // two-phase epoch: the loss terms come from struct_errors_kernel
  __syncthreads();
  if (lane == 0) { red[3 * warp] = sse_w; red[3 * warp + 1] = vsse_w; red[3 * warp + 2] = nval_w; }
  __syncthreads();
  if (tid < 3) {
    double s_ = 0.0;
    for (int w = 0; w < MLP2_WARPS; w++) s_ += red[3 * w + tid];
    a.partial[(size_t)blockIdx.x * (a.P_total + 3) + a.P_total + tid] = s_;
  }
}

// Eligibility of the two-CTA form and its launch; *n_cta = rows of the partial buffer the launched kernel writes.
static int mlp2_ntw(const NetDev& net)
{
  int mx = 1;
  for (int l = 1; l < net.n_layers; l++) mx = std::max(mx, net.sizes[l]);
  const int nt = (mx + 7) / 8;
  return nt <= 5 ? 5 : nt <= 8 ? 8 : 0;
}
static bool mlp2_eligible(const NetDev& net, Mlp2Plan& plan, std::vector<DwTile>* table)
{
  const int L = net.n_layers;
  if (L < 2 || L > 3) return false;
  for (int l = 1; l < L; l++)
    if (!(((net.tanh_layers | net.spline_layers) >> l) & 1)) return false;      // per layer: all tanh or all tanh_spline
  if (net.sizes[0] & 1) return false;
  if (net.sizes[L - 1] + 1 > 64) return false;
  const int NTW = mlp2_ntw(net);
  if (!NTW) return false;
  if (!mlp2_make_plan(net, NTW, plan, table)) return false;
  if (plan.kp[0] > 256) return false;
  return plan.smem_bytes <= 113 * 1024;
}
// Shared-memory plans of a network shape, computed once (the launchers run every epoch / MD step)
struct CachedPlans { int n_layers, sizes[PP_MAX_LAYERS + 1]; unsigned tanh_layers, spline_layers; int TM; MlpPlan p1; bool ok2; Mlp2Plan p2; bool used; };
static const CachedPlans& cached_plans(const NetDev& net)
{
  static CachedPlans cache[8];
  static int next = 0;
  for (int i = 0; i < 8; i++) {
    const CachedPlans& c = cache[i];
    if (c.used && c.n_layers == net.n_layers && c.tanh_layers == net.tanh_layers && c.spline_layers == net.spline_layers &&
        memcmp(c.sizes, net.sizes, sizeof(int) * (net.n_layers + 1)) == 0)
      return c;
  }
  CachedPlans& c = cache[next];
  next = (next + 1) & 7;
  c.used = true;
  c.n_layers = net.n_layers;
  c.tanh_layers = net.tanh_layers;
  c.spline_layers = net.spline_layers;
  memset(c.sizes, 0, sizeof(c.sizes));
  memcpy(c.sizes, net.sizes, sizeof(int) * (net.n_layers + 1));
  c.TM = choose_tm(net, c.p1, nullptr);
  c.ok2 = mlp2_eligible(net, c.p2, nullptr);
  return c;
}

// doubles of derivative scratch the two-CTA form needs (0: not eligible)
size_t ppk_mlp2_scratch_doubles(const NetDev& net)
{
  const CachedPlans& cp = cached_plans(net);
  if (!cp.ok2) return 0;
  return (size_t)2 * ppk_sm_count() * (net.n_layers - 1) * mlp2_ntw(net) * 2 * MLP2_THREADS;
}

int ppk_mlp2_table(const NetDev& net, void* host_out, int* n_entries)
{
  Mlp2Plan plan;
  std::vector<DwTile> table;
  *n_entries = 0;
  if (!mlp2_eligible(net, plan, &table)) return 0;
  *n_entries = (int)table.size();
  if (host_out && !table.empty()) memcpy(host_out, table.data(), table.size() * sizeof(DwTile));
  return 1;
}
// The two-CTA form of `mode` (FUSED, TRAIN or FORWARD) if the network and data layout allow it, else the one-CTA form;
// *n_cta = rows of the partial-gradient buffer the launched kernel writes (gradient modes).
bool ppk_mlp_two_cta_ok(const NetDev& net)
{
  const CachedPlans& cp = cached_plans(net);
  return cp.TM == MLP2_TM && cp.ok2 && getenv("PP_MLP_ONE_CTA") == nullptr;
}

bool ppk_mlp2_available(void) { return encode_tiled() != nullptr; }

int ppk_mlp_launch2(int mode, bool allow_two_cta, const MlpArgs& a, cudaStream_t st, long long* launches, int* n_cta)
{
  *n_cta = ppk_sm_count();
  const NetDev& net = a.net;
  const CachedPlans& cp = cached_plans(net);
  const bool try2 = allow_two_cta && (mode == PP_MLP_FUSED || mode == PP_MLP_TRAIN || mode == PP_MLP_FORWARD) &&
                    a.dw_tiles2 != nullptr && a.scratch2 != nullptr && a.M > 0 && encode_tiled() != nullptr &&
                    (mode != PP_MLP_FUSED || a.rows == nullptr) && ppk_mlp_two_cta_ok(net);
  if (!try2) return ppk_mlp_launch(mode, a, st, launches);
  Mlp2Plan plan = cp.p2;
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  if (a.rows == nullptr && !ppk_tiled_map(a.G, net.sizes[0], a.M, plan.kp[0], MLP2_TM, &tmap)) return ppk_mlp_launch(mode, a, st, launches);
  const int n_tiles = (mode == PP_MLP_FUSED) ? a.n_tiles : (a.M + MLP2_TM - 1) / MLP2_TM;
  int grid = 2 * ppk_sm_count();
  if (mode == PP_MLP_FORWARD) grid = std::max(1, std::min(grid, n_tiles));     // gradient modes: every CTA writes its partial
  const int NTW = mlp2_ntw(net), NLW = net.n_layers - 1;
  bool any_spline = false;
  for (int l = 1; l < net.n_layers; l++) any_spline = any_spline || plan.spline[l] != 0;
#define PP_LAUNCH2(MODE_, NTW_, NLW_)                                                                           \
  {                                                                                                             \
    auto k = any_spline ? mlp_fused2_kernel<MODE_, NTW_, NLW_, true> : mlp_fused2_kernel<MODE_, NTW_, NLW_, false>;  \
    PP_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem_bytes));        \
    k<<<grid, MLP2_THREADS, plan.smem_bytes, st>>>(a, plan, tmap);                                              \
  }
#define PP_LAUNCH2_SHAPE(MODE_)                                  \
  {                                                              \
    if (NTW == 5 && NLW == 1) PP_LAUNCH2(MODE_, 5, 1)            \
    else if (NTW == 5 && NLW == 2) PP_LAUNCH2(MODE_, 5, 2)       \
    else if (NTW == 8 && NLW == 1) PP_LAUNCH2(MODE_, 8, 1)       \
    else PP_LAUNCH2(MODE_, 8, 2)                                 \
  }
  if (mode == PP_MLP_FUSED) PP_LAUNCH2_SHAPE(PP_MLP_FUSED)
  else if (mode == PP_MLP_TRAIN) { if (a.hstore != nullptr) PP_LAUNCH2_SHAPE(PP_MLP_TRAIN_STORED) else PP_LAUNCH2_SHAPE(PP_MLP_TRAIN) }
  else PP_LAUNCH2_SHAPE(PP_MLP_FORWARD)
#undef PP_LAUNCH2_SHAPE
#undef PP_LAUNCH2
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  *n_cta = grid;
  return 0;
}

int ppk_mlp_launch_fused(const MlpArgs& a, cudaStream_t st, long long* launches, int* n_cta)
{
  return ppk_mlp_launch2(PP_MLP_FUSED, true, a, st, launches, n_cta);
}

int ppk_mlp_launch(int mode, const MlpArgs& a, cudaStream_t st, long long* launches)
{
  if (a.M == 0 && mode != PP_MLP_FUSED && mode != PP_MLP_TRAIN) return 0;
  const NetDev& net = a.net;
  PP_REQUIRE(net.n_layers >= 1 && net.n_layers <= PP_MAX_LAYERS, "network has %d weight layers (max %d)", net.n_layers, PP_MAX_LAYERS);
  int max_hidden = 0;
  for (int l = 1; l < net.n_layers; l++) {
    PP_REQUIRE(net.sizes[l] <= MLP_MAX_WIDTH, "hidden layer %d has %d nodes (max %d)", l, net.sizes[l], MLP_MAX_WIDTH);
    if (net.sizes[l] > max_hidden) max_hidden = net.sizes[l];
  }
  MlpPlan plan = cached_plans(net).p1;
  const int TM = cached_plans(net).TM;
  PP_REQUIRE(TM > 0, "network too large for shared memory (%zu bytes at 8 atoms per tile)", plan.smem_bytes);
  PP_REQUIRE(plan.n_dw_tiles <= 16 * MLP_WARPS, "network has %d weight-gradient tiles (max %d)", plan.n_dw_tiles, 16 * MLP_WARPS);
  PP_REQUIRE(a.dw_tiles != nullptr || plan.n_dw_tiles == 0, "internal: weight-gradient tile table missing");
  const int n_tiles = (mode == PP_MLP_FUSED) ? a.n_tiles : (a.M + TM - 1) / TM;
  int grid = ppk_sm_count();
  const bool gradw = (mode == PP_MLP_TRAIN || mode == PP_MLP_FUSED);
  if (!gradw && grid > n_tiles) grid = n_tiles;   // gradient modes: every CTA writes its (possibly zero) partial
  if (grid == 0) return 0;
  // Tensor map of the [atoms][NG] descriptor array for the per-tile copy (box: kpitch[0] columns x TM rows).  Only
  // when rows are consecutive atoms and the row size is a 16-byte multiple; anything else keeps the cp.async path.
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  if (a.rows == nullptr && (net.sizes[0] & 1) == 0 && a.M > 0 && plan.kpitch[0] <= 256)
    plan.tma = ppk_tiled_map(a.G, net.sizes[0], a.M, plan.kpitch[0], TM, &tmap) ? 1 : 0;
  const int NTsel = plan_nt(net);
  const bool many = !plan.blocked;      // OWN = 6: register-blocked table; OWN = 16: single tiles, up to 16 per warp
#define PP_LAUNCH(MODE_, NT_, OWN_)                                                                             \
  {                                                                                                             \
    auto k = mlp_kernel<MODE_, NT_, OWN_>;                                                                      \
    PP_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem_bytes));        \
    k<<<grid, MLP_THREADS, plan.smem_bytes, st>>>(a, plan, tmap);                                               \
  }
#define PP_LAUNCH_NT(MODE_, OWN_)                                \
  {                                                              \
    if (NTsel == 2) PP_LAUNCH(MODE_, 2, OWN_)                    \
    else if (NTsel == 3) PP_LAUNCH(MODE_, 3, OWN_)               \
    else if (NTsel == 4) PP_LAUNCH(MODE_, 4, OWN_)               \
    else PP_LAUNCH(MODE_, 8, OWN_)                               \
  }
  if (mode == PP_MLP_FORWARD) PP_LAUNCH_NT(PP_MLP_FORWARD, 1)
  else if (mode == PP_MLP_INPUT_GRAD) PP_LAUNCH_NT(PP_MLP_INPUT_GRAD, 1)
  else if (mode == PP_MLP_TRAIN) { if (many) PP_LAUNCH_NT(PP_MLP_TRAIN, 16) else PP_LAUNCH_NT(PP_MLP_TRAIN, 6) }
  else { if (many) PP_LAUNCH_NT(PP_MLP_FUSED, 16) else PP_LAUNCH_NT(PP_MLP_FUSED, 6) }
#undef PP_LAUNCH
#undef PP_LAUNCH_NT
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Per-structure energy, error and loss terms: Potential.cpp:641-668.  cstruct[s] = 2*Err*out_scale for train
// structures, else 0.  One warp per structure (lanes stride over its atoms), a block owns a contiguous range of
// structures, and every sum has a FIXED order -- lanes, then the block's warps, then the blocks by index (added up by
// whichever block finishes last) -- so SSE is bit-reproducible run to run.
#define PP_ERR_WARPS 8
__global__ void __launch_bounds__(PP_ERR_WARPS * 32) struct_errors_kernel(StructArgs a)
{
  __shared__ double s_part[PP_ERR_WARPS][3];
  __shared__ bool s_last;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per = (a.n_struct + gridDim.x - 1) / gridDim.x;
  const int s0 = blockIdx.x * per, s1 = min(a.n_struct, s0 + per);
  double sse = 0.0, vsse = 0.0, nval = 0.0;
  for (int s = s0 + warp; s < s1; s += PP_ERR_WARPS) {
    const long long b = a.struct_atom0[s], e_ = a.struct_atom0[s + 1];
    double e = 0.0;
    for (long long at = b + lane; at < e_; at += 32) e += a.E_atom[at];
    e = pp_warp_sum(e) + a.out_mean;
    const double err = e - a.target[s];
    const int tag = a.tag[s];
    double c = 0.0;
    if (tag == 0) { sse += err * err; c = 2.0 * err * a.out_scale; }
    else if (tag == 1) { vsse += err * err; nval += 1.0; }
    if (lane == 0) { a.E_struct[s] = e; a.cstruct[s] = c; }
  }
  if (lane == 0) { s_part[warp][0] = sse; s_part[warp][1] = vsse; s_part[warp][2] = nval; }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w = 0; w < PP_ERR_WARPS; w++) t += s_part[w][threadIdx.x];
    a.scratch[3 * blockIdx.x + threadIdx.x] = t;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (s_last && threadIdx.x < 3) {
    __threadfence();
    double t = 0.0;
    for (unsigned b = 0; b < gridDim.x; b++) t += ((volatile double*)a.scratch)[3 * b + threadIdx.x];
    a.result_tail[threadIdx.x] = t;
    if (threadIdx.x == 0) *a.ticket = 0;
  }
}

int ppk_struct_errors_blocks(int n_struct)
{
  return std::max(1, std::min(2 * ppk_sm_count(), (n_struct + 4 * PP_ERR_WARPS - 1) / (4 * PP_ERR_WARPS)));
}

int ppk_struct_errors(const StructArgs& a, cudaStream_t st, long long* launches)
{
  struct_errors_kernel<<<ppk_struct_errors_blocks(a.n_struct), PP_ERR_WARPS * 32, 0, st>>>(a);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

// Sum of the per-CTA partial rows, in a FIXED order (the result is bit-reproducible run to run): 32 columns x 32 row
// groups per block; group g adds rows g, g+32, g+64, ... (two independent chains to cover the load latency), then the
// 32 group sums are added in order.  reduce_push_kernel (pp_comm.cu) uses the same order.
#define PP_RED_GROUPS 32
__global__ void __launch_bounds__(32 * PP_RED_GROUPS)
reduce_partials_kernel(const double* __restrict__ partial, int n_cta, int stride, int ncols, double* __restrict__ out)
{
  __shared__ double part[PP_RED_GROUPS][33];
  const int col = threadIdx.x & 31, grp = threadIdx.x >> 5;
  const int p = blockIdx.x * 32 + col;
  part[grp][col] = (p < ncols) ? pp_group_column_sum(partial + p, grp, n_cta, stride) : 0.0;
  __syncthreads();
  if (grp == 0 && p < ncols) {
    double s = part[0][col];
#pragma unroll
    for (int j = 1; j < PP_RED_GROUPS; j++) s += part[j][col];
    out[p] = s;
  }
}

int ppk_reduce_partials(const double* partial, int n_cta, int stride, int ncols, double* out, cudaStream_t st, long long* launches)
{
  if (ncols == 0) return 0;
  reduce_partials_kernel<<<(ncols + 31) / 32, 32 * PP_RED_GROUPS, 0, st>>>(partial, n_cta, stride, ncols, out);
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
  // a block owns a contiguous range of rows; warp w takes rows r0+w, r0+w+8, ...; lanes take columns (coalesced);
  // the 8 warps' sums are combined in shared memory in a fixed order
  __shared__ double acc[8][33];
  __shared__ unsigned long long cnt_s[8];
  const int per = (M + gridDim.x - 1) / gridDim.x;
  const int r0 = blockIdx.x * per, r1 = min(M, r0 + per);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned long long cnt = 0;
  for (int k0 = 0; k0 < NG; k0 += 32) {
    const int k = k0 + lane;
    const double mu = (pass && k < NG) ? mean[k] : 0.0;
    double s = 0.0;
    for (int r = r0 + warp; r < r1; r += 8) {
      const size_t atom = rows ? (size_t)rows[r] : (size_t)r;
      if (tag[atom_struct[atom]] != 0) continue;
      if (k0 == 0 && lane == 0) cnt++;
      if (k < NG) {
        const double v = G[atom * NG + k];
        if (pass) { const double d = v - mu; s += d * d; }
        else s += v;
      }
    }
    acc[warp][lane] = s;
    __syncthreads();
    if (warp == 0 && k < NG) {
      double t = 0.0;
      for (int w = 0; w < 8; w++) t += acc[w][lane];
      partial[(size_t)blockIdx.x * NG + k] = t;
    }
    __syncthreads();
  }
  if (lane == 0) cnt_s[warp] = cnt;
  __syncthreads();
  if (threadIdx.x == 0 && pass == 0) {
    unsigned long long t = 0;
    for (int w = 0; w < 8; w++) t += cnt_s[w];
    atomicAdd(count, t);
  }
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
  return ppk_reduce_partials(partial, n_cta, NG, NG, sums, st, launches);
}

// ---------------------------------------------------------------------------------------------------------
// tanh spline table of the reference's Tables("tanh") (Tables.cpp:212-221): C2 cubic spline through tanh at the
// knots -17 + 0.034 i, i = 0..1000.  The reference ships the coefficients as data; they are regenerated here
// (natural-spline tridiagonal solve) and agree with the shipped table to ~7e-12 in the cubic coefficient.
// Layout: X[1001], one pad, then a[1000][4] at PP_SPLINE_COEF (cubic, quadratic, linear, constant in (x - X_i)).
void ppk_make_tanh_table(std::vector<double>& t)
{
  const int N = 1000;
  t.assign(PP_SPLINE_COEF + 4 * N, 0.0);
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
    double* a = &t[PP_SPLINE_COEF + 4 * i];
    a[0] = (M2[i + 1] - M2[i]) / (6.0 * h);
    a[1] = M2[i] / 2.0;
    a[2] = (y[i + 1] - y[i]) / h - h * (2.0 * M2[i] + M2[i + 1]) / 6.0;
    a[3] = y[i];
  }
}
