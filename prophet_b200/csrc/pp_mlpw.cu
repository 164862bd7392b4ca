// Two-phase training epoch with STORED activations, for systems of several elements and for wide / deep networks
// (sm_100a).  Replaces Neural_network::train_MD (NN.cpp:626-697) batched over all atoms of one element, like pp_mlp.cu.
//
// Why two phases: the error of a structure couples the networks of all its elements (Potential.cpp:641-657), so a
// system with several elements needs every element's forward pass before any backward pass.  The first round's
// kernels ran the forward pass AGAIN inside the backward launch (4 F flops for 3 F of work) and kept every weight
// matrix in shared memory, which left room for 16-24 atoms per tile at BASELINE config 4 (three nets 120-60-60-60-1).
// Here
//   * phase 1 (mlpw_forward_kernel) writes each hidden layer's activations H_l and exact derivatives T'_l to HBM
//     ([atoms of the element][n_l], read back once: 16 B per node and atom against ~400 flops per node and atom);
//   * phase 2 (mlpw_backward_kernel) loads H_l tiles with one tensor copy each, reads T'_l straight into registers
//     where the deltas are formed, and runs only the delta and weight-gradient products;
//   * no weight matrix lives in shared memory: a warp owns ONE 8-column n-tile of a layer and streams that slice of
//     the weights from global memory (L1/L2 resident, 4 k-steps ahead) into registers as the B operand of its DMMAs,
//     re-using each fragment for all m-tiles (8-atom row groups) it serves.  Shared memory holds activations only, so
//     tiles are 64 atoms whatever the widths (120-60-60-60-1: 168 KB), and the weight-gradient products get K = 64.
// Phase 1 splits the CTA into MG independent teams of NTW warps (NTW = n-tiles of the widest hidden layer); a team
// walks its own 8*MT-row tiles with its own buffers, copy barrier and named barrier, so one team's transfer-function
// epilogue overlaps another team's products.  Phase 2 shares the weight-gradient accumulators (registers of all 16
// warps, K = all 64 rows), so its stages are separated by CTA barriers.
#include "pp_mlp_dev.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#define MLPW_WARPS 16
#define MLPW_THREADS (MLPW_WARPS * 32)
#define MLPW_TM 64            // rows of a phase-2 tile
#define MLPW_MAX_HIDDEN 4     // hidden layers

struct MlpwPlan {
  int L;                              // weight layers (hidden + output)
  int n[PP_MAX_LAYERS + 1];
  int kp0, kph;                       // phase 2: pitch of a descriptor tile / of a hidden-layer tile (doubles)
  int f_kp0, f_kph;                   // phase 1: the same, chosen for conflict-free 16-byte fragment loads
  int NTW, MG, MT;                    // n-tiles of the widest hidden layer, m-groups, m-tiles per group
  int tfmode;                         // 1: every hidden node tanh, 2: every hidden node tanh_spline, 0: anything else
  int skew;                           // phase 1: team mg starts mg * skew cycles late, so that the teams' product and
                                      // transfer phases interleave instead of running in lockstep
  // phase 1: per team X [8 MT][kp0] and two hidden buffers [8 MT][kph]; then shared tables
  int f_team_stride, f_x_off, f_h_off[2], f_y_off;
  int f_mu_off, f_is_off, f_wout_off, f_exptab_off, f_mbar_off;
  size_t f_smem;
  // phase 2: X [64][kp0], H_l [64][kph] (l = 1 .. L-1)
  int b_x_off, b_h_off[MLPW_MAX_HIDDEN + 1];
  int b_t_off;                        // T'_{L-1} tile [64][kph] brought in by its own tensor copy, or -1: read from global memory
  int b_mu_off, b_is_off, b_wout_off, b_cvec_off, b_red_off, b_mbar_off, b_tab_off_bytes;
  int nb;                             // weight-gradient block slots per warp
  size_t b_smem;
};

struct MlpwMaps { CUtensorMap m[MLPW_MAX_HIDDEN + 2]; };      // 0: descriptors (consecutive rows only), l: H_l, last: T'_{L-1}
#define MLPW_TMAP (MLPW_MAX_HIDDEN + 1)

static int mlpw_pitch16(int v)    // smallest pitch >= v whose 8-row 16-byte fragment loads are free of bank conflicts
{
  int p = v;
  while ((p & 15) != 8) p++;
  return p;
}
static int mlpw_pitch(int v)      // smallest pitch >= v whose 8-row fragment loads are free of bank conflicts
{
  int p = v;
  while (!((p & 15) == 4 || (p & 15) == 12)) p++;
  return p;
}

// ---------------------------------------------------------------------------------------------------------
// C[8 rows of each m-tile][8 columns of n-tile j] = A * B with A in shared memory and B streamed from global memory.
//   A(row, k)  = As[(8 mt + row) * apitch + k]                                  mt = 0 .. nmt-1
//   B(k, n)    = BT ? W[k * ld + n] : W[n * ld + k]     for k < kmax, n < nmax, else 0
// A DMMA sums over any four k; here two consecutive k-steps take k = 8t + 2q and k = 8t + 2q + 1 from lane q (instead of
// 4s + q and 4s + 4 + q), so ONE 16-byte shared-memory load feeds two DMMAs per m-tile -- the loop is bound by issue
// slots (DMMA + its NOP + the loads of four warps per scheduler), not by the pipe.  The weights of KP such pairs are
// fetched one chunk ahead; loads never leave the matrix (indices are clamped) and carry no branch, what lies outside
// k < kmax, n < nmax is replaced by zero.  K is walked in whole pairs without bounds tests: past kmax the weights are
// zeros and the A operand reads finite shared memory (padding columns, the next row, or the buffer that follows).
template <int KP, bool BT>
__device__ __forceinline__ void stream_weights(double (&w)[2 * KP], const double* __restrict__ wb, int kstride, int t0, int q, int kmax,
                                               bool nok)
{
#pragma unroll
  for (int i = 0; i < KP; i++) {
    const int k0 = 8 * (t0 + i) + 2 * q;
    const double v0 = __ldg(wb + (size_t)min(k0, kmax - 1) * kstride);
    const double v1 = __ldg(wb + (size_t)min(k0 + 1, kmax - 1) * kstride);
    w[2 * i] = (nok && k0 < kmax) ? v0 : 0.0;
    w[2 * i + 1] = (nok && k0 + 1 < kmax) ? v1 : 0.0;
  }
}
template <int MT, int KP>
__device__ __forceinline__ void stream_products(double (&acc)[MT][2], const double* const (&ap)[MT], const double (&w)[2 * KP], int t0)
{
#pragma unroll
  for (int i = 0; i < KP; i++) {
    double2 av[MT];
#pragma unroll
    for (int mt = 0; mt < MT; mt++) av[mt] = *reinterpret_cast<const double2*>(ap[mt] + 8 * (t0 + i));
#pragma unroll
    for (int mt = 0; mt < MT; mt++) dmma884(acc[mt][0], acc[mt][1], av[mt].x, w[2 * i]);
#pragma unroll
    for (int mt = 0; mt < MT; mt++) dmma884(acc[mt][0], acc[mt][1], av[mt].y, w[2 * i + 1]);
  }
}
// The same product with the usual k order (k-step s: k = 4 s + q, one 8-byte load per DMMA): for buffers whose pitch is
// chosen for the weight-gradient products (pitch mod 16 in {4, 12}), where 16-byte fragment loads would conflict.
template <int MT, bool BT, int KC, bool FULL>
__device__ __forceinline__ void stream_gemm_k4(double (&acc)[MT][2], const double* __restrict__ As, int apitch, int nmt,
                                               const double* __restrict__ W, int ld, int kmax, int nmax, int j, int lane)
{
  const int g = lane >> 2, q = lane & 3;
  const int ncol = 8 * j + g;
  const int nsteps = (ncol < nmax) ? ((kmax - q + 3) >> 2) : 0;             // k-steps with k < kmax for this lane
  const int kstride = BT ? ld : 1;
  const double* wb = W + (BT ? (size_t)min(ncol, nmax - 1) : (size_t)min(ncol, nmax - 1) * ld);
  const int klast = kmax - 1;
  const int nks = (kmax + 3) >> 2;
#pragma unroll
  for (int mt = 0; mt < MT; mt++) { acc[mt][0] = 0.0; acc[mt][1] = 0.0; }
  const double* ap[MT];
#pragma unroll
  for (int mt = 0; mt < MT; mt++) ap[mt] = As + (size_t)(8 * ((FULL || mt < nmt) ? mt : 0) + g) * apitch + q;
  double wcur[KC], wnxt[KC];
#pragma unroll
  for (int i = 0; i < KC; i++) {
    const double v = __ldg(wb + (size_t)min(4 * i + q, klast) * kstride);
    wcur[i] = (i < nsteps) ? v : 0.0;
  }
  for (int ks0 = 0; ks0 < nks; ks0 += KC) {
#pragma unroll
    for (int i = 0; i < KC; i++) {
      const int s_ = ks0 + KC + i;
      const double v = __ldg(wb + (size_t)min(4 * s_ + q, klast) * kstride);
      wnxt[i] = (s_ < nsteps) ? v : 0.0;
    }
#pragma unroll
    for (int i = 0; i < KC; i++) {
      double av[MT];
#pragma unroll
      for (int mt = 0; mt < MT; mt++) av[mt] = ap[mt][4 * (ks0 + i)];
#pragma unroll
      for (int mt = 0; mt < MT; mt++) dmma884(acc[mt][0], acc[mt][1], av[mt], wcur[i]);
    }
#pragma unroll
    for (int i = 0; i < KC; i++) wcur[i] = wnxt[i];
  }
}

// Where this lane's weights of a product start, and their first chunk: split from the product itself so that the chunk
// can be requested BEFORE the barrier that precedes the product (its L2 latency then hides behind the wait).
struct StreamPos { const double* wb; int kstride; bool nok; };
template <bool BT>
__device__ __forceinline__ StreamPos stream_pos(const double* __restrict__ W, int ld, int nmax, int j, int lane)
{
  const int ncol = 8 * j + (lane >> 2);
  StreamPos sp;
  sp.nok = ncol < nmax;
  sp.kstride = BT ? ld : 1;
  sp.wb = W + (BT ? (size_t)min(ncol, nmax - 1) : (size_t)min(ncol, nmax - 1) * ld);
  return sp;
}
template <int MT, bool BT, int KP, bool FULL>
__device__ __forceinline__ void stream_gemm(double (&acc)[MT][2], const double* __restrict__ As, int apitch, int nmt,
                                            const StreamPos& sp, double (&wa)[2 * KP], int kmax, int lane)
{
  const int g = lane >> 2, q = lane & 3;
  const int npairs = (kmax + 7) >> 3;
#pragma unroll
  for (int mt = 0; mt < MT; mt++) { acc[mt][0] = 0.0; acc[mt][1] = 0.0; }
  const double* ap[MT];
#pragma unroll
  for (int mt = 0; mt < MT; mt++) ap[mt] = As + (size_t)(8 * ((FULL || mt < nmt) ? mt : 0) + g) * apitch + 2 * q;
  double wb2[2 * KP];
  for (int t0 = 0; t0 < npairs; t0 += 2 * KP) {
    stream_weights<KP, BT>(wb2, sp.wb, sp.kstride, t0 + KP, q, kmax, sp.nok);
    stream_products<MT, KP>(acc, ap, wa, t0);
    if (t0 + KP < npairs) {
      stream_weights<KP, BT>(wa, sp.wb, sp.kstride, t0 + 2 * KP, q, kmax, sp.nok);
      stream_products<MT, KP>(acc, ap, wb2, t0 + KP);
    }
  }
}

// Phase timing for tuning (build with -DPP_MLPW_TIMING; not part of the product build): warps 0 and 8 of CTA 0
// accumulate clock64() deltas per phase; [0] forward kernel, [1] backward kernel.
#ifdef PP_MLPW_TIMING
__device__ unsigned long long pp_mlpw_cycles[2][2][16];
#define PW_T(k, i)                                                                           \
  do {                                                                                       \
    if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 8)) {                          \
      const long long now_ = clock64();                                                      \
      pp_mlpw_cycles[k][warp ? 1 : 0][i] += (unsigned long long)(now_ - tprev_);             \
      tprev_ = now_;                                                                         \
    }                                                                                        \
  } while (0)
#define PW_T0 long long tprev_ = clock64()
extern "C" int pp_debug_mlpw_cycles(unsigned long long* out, int reset)
{
  if (out && cudaMemcpyFromSymbol(out, pp_mlpw_cycles, sizeof(pp_mlpw_cycles)) != cudaSuccess) return 1;
  if (reset) {
    unsigned long long z[2][2][16] = {};
    if (cudaMemcpyToSymbol(pp_mlpw_cycles, z, sizeof(z)) != cudaSuccess) return 1;
  }
  return 0;
}
#else
#define PW_T(k, i)
#define PW_T0
#endif

__device__ __forceinline__ void team_barrier(int id, int threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory"); }

__device__ __forceinline__ void tma_load_tile_at(void* dst, const CUtensorMap* map, int col, int row, unsigned long long* bar)
{
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_addr(dst)),
      "l"(map), "r"(col), "r"(row), "r"(smem_addr(bar))
      : "memory");
}

// ---------------------------------------------------------------------------------------------------------
// Phase 1: forward pass of one element's atoms; energies to E_atom, and (st.H != nullptr) H_l / T'_l to HBM.
template <int NTW, int MG, int MT, int TFMODE>
__global__ void __launch_bounds__(MLPW_THREADS, 1)
mlpw_forward_kernel(MlpArgs a, MlpwPlan p, MlpwStore st, const __grid_constant__ CUtensorMap tmapX)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sm = reinterpret_cast<double*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  const int L = p.L, NG = p.n[0], kp0 = p.f_kp0, kph = p.f_kph;
  const NetDev& net = a.net;
  constexpr int TMG = 8 * MT;             // rows of a team tile
  constexpr int TT = NTW * 32;            // threads of a team

  // ---- one-time staging ------------------------------------------------------------------------------------------
  for (int k = tid; k < kp0; k += MLPW_THREADS) {
    sm[p.f_mu_off + k] = (k < NG) ? net.mean[k] : 0.0;
    sm[p.f_is_off + k] = (k < NG) ? 1.0 / net.scale[k] : 0.0;
  }
  for (int k = tid; k < kph; k += MLPW_THREADS) sm[p.f_wout_off + k] = (k <= p.n[L - 1]) ? net.params[net.w_off[L - 1] + k] : 0.0;
  for (int i = tid; i < 64; i += MLPW_THREADS) sm[p.f_exptab_off + i] = exp2((double)i * 0.015625);
  for (int i = tid; i < MG * p.f_team_stride; i += MLPW_THREADS) sm[p.f_x_off + i] = 0.0;      // padding columns stay zero
  unsigned long long* mbars = reinterpret_cast<unsigned long long*>(sm + p.f_mbar_off);
  if (tid < MG) mbar_init(mbars + tid, 1);
  fence_proxy_async();
  __syncthreads();
  if (warp >= NTW * MG) return;             // spare warps (NTW * MG < 16): no CTA-wide barrier below this line

  const int j = warp % NTW, mg = warp / NTW;
  const int tt = j * 32 + lane;             // thread index inside the team
  double* X = sm + p.f_x_off + mg * p.f_team_stride;
  double* Hb0 = sm + p.f_h_off[0] + mg * p.f_team_stride;
  double* Hb1 = sm + p.f_h_off[1] + mg * p.f_team_stride;
  double* ypart = sm + p.f_y_off + mg * p.f_team_stride;      // [TMG][8]
  unsigned long long* mbar = mbars + mg;
  const double* mu = sm + p.f_mu_off;
  const double* is = sm + p.f_is_off;
  const double* wout = sm + p.f_wout_off;
  const double* exptab = sm + p.f_exptab_off;
  const bool gather = (a.rows != nullptr);
  const bool store = (st.H != nullptr);
  const int n_tiles = (a.M + TMG - 1) / TMG;
  const int tstride = gridDim.x * MG;

  auto issue_load = [&](int tile) {         // every thread of the team calls this
    const int row0 = tile * TMG;
    const int nvalid = min(TMG, a.M - row0);
    if (!gather) {
      if (tt == 0) {
        fence_proxy_async();
        mbar_arrive_expect_tx(mbar, (unsigned)TMG * (unsigned)kp0 * 8u);
        tma_load_tile(X, &tmapX, row0, mbar);
      }
      return;
    }
    if (tt == 0) mbar_arrive_expect_tx(mbar, (unsigned)nvalid * (unsigned)NG * 8u);
    if (tt < TMG) {
      double* dst = X + (size_t)tt * kp0;
      if (tt < nvalid) {
        fence_proxy_async();
        bulk_copy_g2s(dst, a.G + (size_t)a.rows[row0 + tt] * NG, (unsigned)NG * 8u, mbar);
      } else {
        for (int c = 0; c < NG; c++) dst[c] = 0.0;
      }
    }
  };

  // first weight chunk of the product that comes next, requested one barrier early
  constexpr int KP = 4;
  double wfirst[2 * KP];
  StreamPos spn = stream_pos<false>(net.params + net.w_off[0], p.n[0] + 1, p.n[1], j, lane);
  stream_weights<KP, false>(wfirst, spn.wb, spn.kstride, 0, q, p.n[0] + 1, spn.nok && 8 * j < p.n[1]);
  int tile = blockIdx.x * MG + mg;
  if (tile < n_tiles) issue_load(tile);
  if (mg > 0 && p.skew > 0) {
    const long long t0 = clock64();
    while (clock64() - t0 < (long long)mg * p.skew) __nanosleep(200);
  }
  team_barrier(1 + mg, TT);
  unsigned phase = 0;
  for (; tile < n_tiles; tile += tstride, phase ^= 1u) {
    const int row0 = tile * TMG;
    const int nvalid = min(TMG, a.M - row0);
    PW_T0;
    mbar_wait(mbar, phase);
    PW_T(0, 0);
    // ---- normalise the rows in place (NN.cpp:647-651); column NG is the bias input ---------------------------------
    for (int r = j; r < TMG; r += NTW) {
      double* xr = X + (size_t)r * kp0;
      if (!a.prenorm) {                                       // (skipped when the descriptors come normalised, pp_api.cu)
        for (int c = 2 * lane; c < NG; c += 64) {             // NG is even, rows and tables are 16-byte aligned
          const double2 x = *reinterpret_cast<const double2*>(xr + c);
          const double2 m = *reinterpret_cast<const double2*>(mu + c);
          const double2 v = *reinterpret_cast<const double2*>(is + c);
          *reinterpret_cast<double2*>(xr + c) = make_double2((x.x - m.x) * v.x, (x.y - m.y) * v.y);
        }
      }
      if (lane == 0) xr[NG] = 1.0;
    }
    PW_T(0, 7);
    team_barrier(1 + mg, TT);
    PW_T(0, 1);
    // ---- hidden layers -------------------------------------------------------------------------------------------------
    size_t soff = 0;                        // offset of layer l + 1's block in the stores
    for (int l = 0; l < L - 1; l++) {
      const int n_in = p.n[l], n_out = p.n[l + 1];
      const double* A = (l == 0) ? X : ((l & 1) ? Hb0 : Hb1);
      const int apitch = (l == 0) ? kp0 : kph;
      double* Hn = (l & 1) ? Hb1 : Hb0;
      if (8 * j < n_out) {
        double acc[MT][2];
        stream_gemm<MT, false, KP, true>(acc, A, apitch, MT, spn, wfirst, n_in + 1, lane);
        PW_T(0, 2);
        double tv[MT][2], dv[MT][2];
        if (TFMODE == 1) {
#pragma unroll
          for (int mt = 0; mt < MT; mt++)
#pragma unroll
            for (int h = 0; h < 2; h++) tv[mt][h] = pp_fast_tanh(acc[mt][h], dv[mt][h], exptab);
        } else if (TFMODE == 2) {
#pragma unroll
          for (int mt = 0; mt < MT; mt++)
#pragma unroll
            for (int h = 0; h < 2; h++) tv[mt][h] = pp_fast_tanh_spline(acc[mt][h], dv[mt][h], a.tanh_table);
        } else {
          const int* tfl = net.tf + net.node_off[l];
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int n = 8 * j + 2 * q + h;
            const int id = (n < n_out) ? __ldg(tfl + n) : 0;
#pragma unroll
            for (int mt = 0; mt < MT; mt++) tv[mt][h] = pp_transfer(id, acc[mt][h], dv[mt][h], a.tanh_table, exptab);
          }
        }
        const int n0 = 8 * j + 2 * q;
        const bool last = (l == L - 2);
        if (last) {
          // output node (always linear, NN.cpp:76) folded in: this warp's share of w_out . H, summed over the four
          // lanes that hold a row's columns; the team's n-tiles meet in `ypart` after the barrier
          const double w0 = (n0 < n_out) ? wout[n0] : 0.0, w1 = (n0 < n_out) ? wout[n0 + 1] : 0.0;
#pragma unroll
          for (int mt = 0; mt < MT; mt++) {
            double y = fma(tv[mt][0], w0, tv[mt][1] * w1);
            y += __shfl_xor_sync(0xffffffffu, y, 1);
            y += __shfl_xor_sync(0xffffffffu, y, 2);
            if (q == 0) ypart[(8 * mt + g) * 8 + j] = y;
          }
        }
        if (n0 < n_out) {                    // widths are even: columns n0 and n0 + 1 exist together
#pragma unroll
          for (int mt = 0; mt < MT; mt++) {
            const int r = 8 * mt + g;
            if (!last) *reinterpret_cast<double2*>(Hn + (size_t)r * kph + n0) = make_double2(tv[mt][0], tv[mt][1]);
            if (store && r < nvalid) {
              const size_t o = soff + (size_t)(row0 + r) * n_out + n0;
              *reinterpret_cast<double2*>(st.H + o) = make_double2(tv[mt][0], tv[mt][1]);
              *reinterpret_cast<double2*>(st.T + o) = make_double2(dv[mt][0], dv[mt][1]);
            }
          }
        }
      }
      if (j == 0 && q == 0 && l < L - 2) {
#pragma unroll
        for (int mt = 0; mt < MT; mt++) Hn[(size_t)(8 * mt + g) * kph + n_out] = 1.0;      // bias input of the next layer
      }
      soff += (size_t)a.M * n_out;
      {
        // the next product: layer l + 1 of this tile, or layer 0 of the next one
        const int ln = (l + 1 < L - 1) ? l + 1 : 0;
        spn = stream_pos<false>(net.params + net.w_off[ln], p.n[ln] + 1, p.n[ln + 1], j, lane);
        stream_weights<KP, false>(wfirst, spn.wb, spn.kstride, 0, q, p.n[ln] + 1, spn.nok && 8 * j < p.n[ln + 1]);
      }
      PW_T(0, 3);
      team_barrier(1 + mg, TT);
      PW_T(0, 4);
      if (l == 0 && tile + tstride < n_tiles) issue_load(tile + tstride);      // X is free: the next tile's rows
    }
    // ---- energies: the n-tiles' partial sums of the output node, in a fixed order, plus its bias ---------------------
    if (tt < TMG && tt < nvalid && a.E_atom != nullptr) {
      const int n_h = p.n[L - 1];
      const int ntl = (n_h + 7) >> 3;
      double y = wout[n_h];
      for (int t = 0; t < ntl; t++) y += ypart[tt * 8 + t];
      a.E_atom[gather ? (size_t)a.rows[row0 + tt] : (size_t)(row0 + tt)] = a.out_scale * y;
    }
    PW_T(0, 5);
    team_barrier(1 + mg, TT);               // the hidden buffers are rewritten by the next tile
    PW_T(0, 6);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Weight-gradient block with the bias input supplied in registers: the H_l tiles come straight from HBM and carry no
// "1" column, so the lane that would load column n_l of [H_l, 1] uses 1.0 instead (bj = block column holding it, or -1).
template <int OWN, int BASE, int NM, int NN>
__device__ __forceinline__ void dw_block_bias(double (&dW)[OWN][2], const double* __restrict__ pa, const double* __restrict__ pb,
                                              int sa, int sb, int bias_j, bool bias_lane)
{
  bool one[NN];
#pragma unroll
  for (int jn = 0; jn < NN; jn++) one[jn] = bias_lane && (jn == bias_j);
#pragma unroll 4
  for (int k0 = 0; k0 < MLPW_TM; k0 += 4) {
    double av[NM], bv[NN];
#pragma unroll
    for (int i = 0; i < NM; i++) av[i] = pa[8 * i];
#pragma unroll
    for (int jn = 0; jn < NN; jn++) bv[jn] = one[jn] ? 1.0 : pb[8 * jn];
    pa += sa; pb += sb;
#pragma unroll
    for (int i = 0; i < NM; i++) {
#pragma unroll
      for (int jn = 0; jn < NN; jn++) dmma884(dW[BASE + i * NN + jn][0], dW[BASE + i * NN + jn][1], av[i], bv[jn]);
    }
  }
}

// Phase 2: parameter gradients of one element from the stored activations; cstruct[s] = 2 Err_s out_scale is given.
template <int NTW, int MG, int MT, int NB>
__global__ void __launch_bounds__(MLPW_THREADS, 1)
mlpw_backward_kernel(MlpArgs a, MlpwPlan p, MlpwStore st, const __grid_constant__ MlpwMaps maps)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sm = reinterpret_cast<double*>(smem_raw);
  const DwTile* tab = reinterpret_cast<const DwTile*>(smem_raw + p.b_tab_off_bytes);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  const int L = p.L, NG = p.n[0], kp0 = p.kp0, kph = p.kph;
  const NetDev& net = a.net;
  constexpr int TM = MLPW_TM;

  // ---- one-time staging ------------------------------------------------------------------------------------------
  for (int k = tid; k < kp0; k += MLPW_THREADS) {
    sm[p.b_mu_off + k] = (k < NG) ? net.mean[k] : 0.0;
    sm[p.b_is_off + k] = (k < NG) ? 1.0 / net.scale[k] : 0.0;
  }
  for (int k = tid; k < kph; k += MLPW_THREADS) sm[p.b_wout_off + k] = (k <= p.n[L - 1]) ? net.params[net.w_off[L - 1] + k] : 0.0;
  for (int i = tid; i < TM * kp0; i += MLPW_THREADS) sm[p.b_x_off + i] = 0.0;
  for (int i = tid; i < 2 * TM; i += MLPW_THREADS) sm[p.b_cvec_off + i] = 0.0;
  {
    const int nwords = MLPW_WARPS * NB * (int)(sizeof(DwTile) / sizeof(int));
    int* dst = reinterpret_cast<int*>(smem_raw + p.b_tab_off_bytes);
    const int* srcw = reinterpret_cast<const int*>(a.dw_tiles2);
    for (int i = tid; i < nwords; i += MLPW_THREADS) dst[i] = srcw[i];
  }
  unsigned long long* mbar = reinterpret_cast<unsigned long long*>(sm + p.b_mbar_off);      // [0]: X, [l]: H_l, [L]: T'_{L-1}
  if (tid <= L) mbar_init(mbar + tid, 1);
  fence_proxy_async();
  __syncthreads();

  const double* mu = sm + p.b_mu_off;
  const double* is = sm + p.b_is_off;
  const double* wout = sm + p.b_wout_off;
  double* cvec = sm + p.b_cvec_off;         // [2][TM]
  double* X = sm + p.b_x_off;
  const bool gather = (a.rows != nullptr);
  const int n_tiles = (a.M + TM - 1) / TM;
  const int j = warp % NTW, mg = warp / NTW;
  const bool gemm_warp = warp < NTW * MG;
  const int mt0 = MT * mg;                  // first m-tile of this warp's group
  const int nmt = gemm_warp ? max(0, min(MT, TM / 8 - mt0)) : 0;

  size_t off_last = 0;                      // offset of layer L - 1's block in the stores
  for (int l = 1; l < L - 1; l++) off_last += (size_t)a.M * p.n[l];
  double dW[4 * NB][2];
  double dwout[3] = {0.0, 0.0, 0.0};        // output node: parameters lane, lane + 32, lane + 64 over this warp's rows
#pragma unroll
  for (int t = 0; t < 4 * NB; t++) { dW[t][0] = 0.0; dW[t][1] = 0.0; }

  auto issue_h = [&](int l, int tile) {     // thread 0: H_l of `tile` (one tensor copy; rows past the end arrive as zeros)
    fence_proxy_async();
    mbar_arrive_expect_tx(mbar + l, (unsigned)TM * (unsigned)kph * 8u);
    tma_load_tile(sm + p.b_h_off[l], &maps.m[l], tile * TM, mbar + l);
  };
  const bool t_smem = p.b_t_off >= 0;
  auto issue_t = [&](int tile) {            // thread 0: T'_{L-1} of `tile`
    fence_proxy_async();
    mbar_arrive_expect_tx(mbar + L, (unsigned)TM * (unsigned)kph * 8u);
    tma_load_tile(sm + p.b_t_off, &maps.m[MLPW_TMAP], tile * TM, mbar + L);
  };
  auto issue_x = [&](int tile) {            // every thread calls this
    const int row0 = tile * TM;
    const int nvalid = min(TM, a.M - row0);
    if (!gather) {
      if (tid == 0) {
        fence_proxy_async();
        mbar_arrive_expect_tx(mbar, (unsigned)TM * (unsigned)kp0 * 8u);
        tma_load_tile(X, &maps.m[0], row0, mbar);
      }
      return;
    }
    if (tid == 0) mbar_arrive_expect_tx(mbar, (unsigned)nvalid * (unsigned)NG * 8u);
    if (tid < TM) {
      double* dst = X + (size_t)tid * kp0;
      if (tid < nvalid) {
        fence_proxy_async();
        bulk_copy_g2s(dst, a.G + (size_t)a.rows[row0 + tid] * NG, (unsigned)NG * 8u, mbar);
      } else {
        for (int c = 0; c < NG; c++) dst[c] = 0.0;
      }
    }
  };
  auto row_factor = [&](int tile, int r) {  // 2 Err_s out_scale of the structure that owns row r of `tile`
    const int row0 = tile * TM;
    if (tile >= n_tiles || row0 + r >= a.M) return 0.0;
    const size_t atom = gather ? (size_t)a.rows[row0 + r] : (size_t)(row0 + r);
    return a.cstruct[a.atom_struct[atom]];
  };
  auto normalise_x = [&](unsigned ph) {     // NN.cpp:647-651 on this tile's rows, in place
    mbar_wait(mbar, ph);
    if (a.prenorm) return;                  // the descriptors come normalised (pp_api.cu keeps such a copy)
    for (int r = warp; r < TM; r += MLPW_WARPS) {
      double* xr = X + (size_t)r * kp0;
      for (int c = 2 * lane; c < NG; c += 64) {
        const double2 x = *reinterpret_cast<const double2*>(xr + c);
        const double2 m = *reinterpret_cast<const double2*>(mu + c);
        const double2 v = *reinterpret_cast<const double2*>(is + c);
        *reinterpret_cast<double2*>(xr + c) = make_double2((x.x - m.x) * v.x, (x.y - m.y) * v.y);
      }
    }
  };

  int tile = blockIdx.x;
  if (tile < n_tiles) {
    issue_x(tile);
    if (tid == 0) {
      for (int l = 1; l < L; l++) issue_h(l, tile);
      if (t_smem) issue_t(tile);
    }
    if (tid < TM) cvec[tid] = row_factor(tile, tid);
  }
  __syncthreads();

  unsigned ph = 0;
  int cbuf = 0;
  for (; tile < n_tiles; tile += gridDim.x, ph ^= 1u, cbuf ^= 1) {
    const int row0 = tile * TM;
    const int nvalid = min(TM, a.M - row0);
    const int nxt = tile + gridDim.x;
    const bool has_next = nxt < n_tiles;
    double c_next = 0.0;                    // next tile's error factors start their two dependent loads now
    if (tid < TM) c_next = row_factor(nxt, tid);
    const double* cv = cvec + cbuf * TM;
    PW_T0;

    // ---- output node: gradient over this warp's four rows, then delta_{L-1} = c w_out T' in place of H_{L-1} ---------
    {
      const int n_h = p.n[L - 1];
      double* Hl = sm + p.b_h_off[L - 1];
      const double* Tl = st.T + off_last;
      const double* Tt = sm + (t_smem ? p.b_t_off : 0);
      mbar_wait(mbar + (L - 1), ph);
      if (t_smem) mbar_wait(mbar + L, ph);
      PW_T(1, 0);
#pragma unroll
      for (int rr = 0; rr < TM / MLPW_WARPS; rr++) {
        const int r = (TM / MLPW_WARPS) * warp + rr;
        const double c = cv[r];
        const bool rv = r < nvalid;
#pragma unroll
        for (int kk = 0; kk < 3; kk++) {
          const int k = lane + 32 * kk;
          if (k < n_h) {
            const double tp = t_smem ? Tt[(size_t)r * kph + k] : (rv ? Tl[(size_t)(row0 + r) * n_h + k] : 0.0);
            const double h = Hl[(size_t)r * kph + k];
            dwout[kk] = fma(c, h, dwout[kk]);
            Hl[(size_t)r * kph + k] = c * wout[k] * tp;
          } else if (k == n_h) {
            dwout[kk] += c;
          }
        }
      }
      if (L == 2) normalise_x(ph);
    }
    PW_T(1, 1);
    __syncthreads();                                                                   // delta_{L-1} complete
    PW_T(1, 2);
    if (t_smem && tid == 0 && has_next) issue_t(nxt);

    size_t off_l = off_last;                // offset of layer l's block, walked downwards
    for (int l = L - 2; l >= 0; l--) {
      const int n_l = p.n[l], n_up = p.n[l + 1];
      if (l >= 1) off_l -= (size_t)a.M * n_l;
      // ---- dW_l += delta_{l+1}^T [H_l, 1] over the 64 rows of the tile ---------------------------------------------
      if (l >= 1) mbar_wait(mbar + l, ph);
      PW_T(1, 5);
#pragma unroll
      for (int s = 0; s < NB; s++) {
        const DwTile d = tab[warp * NB + s];
        if (d.nm && d.layer == l) {
          const double* pa = sm + d.a_off + g + (size_t)q * d.dp;
          const double* pb = sm + d.b_off + g + (size_t)q * d.kp;
          const int sa = 4 * d.dp, sb = 4 * d.kp;
          const bool bl = (g == d.pad[1]);
          if (s == 0) {
            if (d.nm == 2 && d.nn == 2) dw_block_bias<4 * NB, 0, 2, 2>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else if (d.nn == 2) dw_block_bias<4 * NB, 0, 1, 2>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else if (d.nm == 2) dw_block_bias<4 * NB, 0, 2, 1>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else dw_block_bias<4 * NB, 0, 1, 1>(dW, pa, pb, sa, sb, d.pad[0], bl);
          } else if (s == 1) {
            if (d.nm == 2 && d.nn == 2) dw_block_bias<4 * NB, 4, 2, 2>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else if (d.nn == 2) dw_block_bias<4 * NB, 4, 1, 2>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else if (d.nm == 2) dw_block_bias<4 * NB, 4, 2, 1>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else dw_block_bias<4 * NB, 4, 1, 1>(dW, pa, pb, sa, sb, d.pad[0], bl);
          } else if (NB > 2 && s == 2) {
            if (d.nm == 2 && d.nn == 2) dw_block_bias<4 * NB, (NB > 2 ? 8 : 0), 2, 2>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else if (d.nn == 2) dw_block_bias<4 * NB, (NB > 2 ? 8 : 0), 1, 2>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else if (d.nm == 2) dw_block_bias<4 * NB, (NB > 2 ? 8 : 0), 2, 1>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else dw_block_bias<4 * NB, (NB > 2 ? 8 : 0), 1, 1>(dW, pa, pb, sa, sb, d.pad[0], bl);
          } else if (NB > 2) {
            if (d.nm == 2 && d.nn == 2) dw_block_bias<4 * NB, (NB > 2 ? 12 : 0), 2, 2>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else if (d.nn == 2) dw_block_bias<4 * NB, (NB > 2 ? 12 : 0), 1, 2>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else if (d.nm == 2) dw_block_bias<4 * NB, (NB > 2 ? 12 : 0), 2, 1>(dW, pa, pb, sa, sb, d.pad[0], bl);
            else dw_block_bias<4 * NB, (NB > 2 ? 12 : 0), 1, 1>(dW, pa, pb, sa, sb, d.pad[0], bl);
          }
        }
      }
      PW_T(1, 6);
      if (l == 0 && tid < TM) cvec[(cbuf ^ 1) * TM + tid] = c_next;                   // the next tile's factors, published by the barrier below
      __syncthreads();                                                                 // every reader of H_l is done
      PW_T(1, 7);
      if (l == 0) break;
      // ---- delta_l = (delta_{l+1} W_l) .* T'_l, written over H_l: this warp's n-tile and m-group --------------------
      if (gemm_warp && (8 * j < n_l) && nmt > 0) {
        const int n0 = 8 * j + 2 * q;
        double2 tp[MT];
#pragma unroll
        for (int mt = 0; mt < MT; mt++) {
          const int r = 8 * (mt0 + mt) + g;
          tp[mt] = (mt < nmt && r < nvalid && n0 < n_l)
                       ? *reinterpret_cast<const double2*>(st.T + off_l + (size_t)(row0 + r) * n_l + n0)
                       : make_double2(0.0, 0.0);
        }
        double acc[MT][2];
        stream_gemm_k4<MT, true, 4, (MT * MG == MLPW_TM / 8)>(acc, sm + p.b_h_off[l + 1] + (size_t)(8 * mt0) * kph, kph, nmt,
                                                           net.params + net.w_off[l], n_l + 1, n_up, n_l, j, lane);
        if (n0 < n_l) {
#pragma unroll
          for (int mt = 0; mt < MT; mt++) {
            if (mt < nmt) {
              double* hr = sm + p.b_h_off[l] + (size_t)(8 * (mt0 + mt) + g) * kph + n0;
              *reinterpret_cast<double2*>(hr) = make_double2(acc[mt][0] * tp[mt].x, acc[mt][1] * tp[mt].y);
            }
          }
        }
      }
      PW_T(1, 3);
      if (l == 1) normalise_x(ph);          // visible to the layer-0 products after the barrier below
      PW_T(1, 4);
      __syncthreads();                                                                 // delta_l complete, delta_{l+1} no longer read
      PW_T(1, 9);
      if (tid == 0 && has_next) issue_h(l + 1, nxt);                                   // the buffer of delta_{l+1} is free
    }
    if (tid == 0 && has_next) issue_h(1, nxt);                                         // so is the buffer of delta_1
    if (has_next) issue_x(nxt);             // X is free as well
  }

  // ---- per-CTA partial gradient -> scratch (summed in a fixed order by reduce_partials_kernel) -------------------------
  double* out = a.partial + (size_t)blockIdx.x * (a.P_total + 3) + a.p_off;
#pragma unroll
  for (int t = 0; t < 4 * NB; t++) {
    const DwTile d = tab[warp * NB + (t >> 2)];
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
  double* red = sm + p.b_red_off;           // [16][96]
  __syncthreads();
#pragma unroll
  for (int kk = 0; kk < 3; kk++) red[warp * 96 + 32 * kk + lane] = dwout[kk];
  __syncthreads();
  if (tid <= p.n[L - 1]) {
    double s_ = 0.0;
    for (int w = 0; w < MLPW_WARPS; w++) s_ += red[w * 96 + tid];
    out[net.w_off[L - 1] + tid] = s_;
  }
}

// =========================================================================================================
// host side
static bool mlpw_make_plan(const NetDev& net, MlpwPlan& p, std::vector<DwTile>* table)
{
  memset(&p, 0, sizeof(p));
  const int L = net.n_layers;
  if (L < 2 || L - 1 > MLPW_MAX_HIDDEN) return false;
  if (net.sizes[0] & 1) return false;                       // 16-byte rows for the copy engine
  int mx = 0;
  for (int l = 1; l < L; l++) {
    if (net.sizes[l] & 1) return false;                     // double2 stores of column pairs, 16-byte rows of H_l
    mx = std::max(mx, net.sizes[l]);
  }
  if (mx > 64) return false;
  p.L = L;
  for (int l = 0; l <= L; l++) p.n[l] = net.sizes[l];
  const int nt = (mx + 7) / 8;
  if (nt <= 5) { p.NTW = 5; p.MG = 3; p.MT = 3; }
  else { p.NTW = 8; p.MG = 2; p.MT = 4; }
  p.kp0 = mlpw_pitch(p.n[0] + 1);
  p.kph = mlpw_pitch(mx + 1);
  p.f_kp0 = mlpw_pitch16(p.n[0] + 1);
  p.f_kph = mlpw_pitch16(mx + 1);
  if (p.kp0 > 256 || p.kph > 256 || p.f_kp0 > 256) return false;      // box limit of the tensor copy
  bool all_tanh = true, all_spline = true;
  for (int l = 1; l < L; l++) {
    all_tanh = all_tanh && ((net.tanh_layers >> l) & 1);
    all_spline = all_spline && ((net.spline_layers >> l) & 1);
  }
  p.tfmode = all_tanh ? 1 : (all_spline ? 2 : 0);
  // phase 1
  {
    const int TMG = 8 * p.MT;
    size_t off = 0;
    p.f_x_off = 0;
    size_t team = (size_t)TMG * p.f_kp0;
    team = (team + 15) & ~(size_t)15;
    p.f_h_off[0] = (int)team; team += (size_t)TMG * p.f_kph;
    team = (team + 15) & ~(size_t)15;
    p.f_h_off[1] = (int)team; team += (size_t)TMG * p.f_kph;
    team = (team + 15) & ~(size_t)15;
    p.f_y_off = (int)team; team += (size_t)TMG * 8;          // output-node partial sums [row][n-tile]
    team = (team + 15) & ~(size_t)15;
    p.f_team_stride = (int)team;
    off = team * p.MG;
    p.f_mu_off = (int)off; off += p.f_kp0;
    p.f_is_off = (int)off; off += p.f_kp0;
    p.f_wout_off = (int)off; off += p.f_kph;
    p.f_exptab_off = (int)off; off += 64;
    p.f_mbar_off = (int)off; off += 4;
    p.f_smem = off * sizeof(double);
  }
  // phase 2
  {
    size_t off = 0;
    p.b_x_off = 0; off += (size_t)MLPW_TM * p.kp0;
    for (int l = 1; l < L; l++) {
      off = (off + 15) & ~(size_t)15;
      p.b_h_off[l] = (int)off; off += (size_t)MLPW_TM * p.kph;
    }
    off = (off + 15) & ~(size_t)15;
    const size_t small = (size_t)2 * p.kp0 + p.kph + 2 * MLPW_TM + MLPW_MAX_HIDDEN + 4 + (size_t)MLPW_WARPS * 4 * sizeof(DwTile) / 8 + 64;
    p.b_t_off = -1;
    if ((off + (size_t)MLPW_TM * p.kph + small) * sizeof(double) <= 227 * 1024) {
      p.b_t_off = (int)off; off += (size_t)MLPW_TM * p.kph;
      off = (off + 15) & ~(size_t)15;
    }
    p.b_mu_off = (int)off; off += p.kp0;
    p.b_is_off = (int)off; off += p.kp0;
    p.b_wout_off = (int)off; off += p.kph;
    p.b_cvec_off = (int)off; off += 2 * MLPW_TM;
    p.b_red_off = p.b_x_off;                                  // epilogue only: aliases the dead descriptor tile
    p.b_mbar_off = (int)off; off += MLPW_MAX_HIDDEN + 4;
    size_t bytes = off * sizeof(double);
    bytes = (bytes + 15) & ~(size_t)15;
    p.b_tab_off_bytes = (int)bytes;
    // weight-gradient blocks: the blocks of each layer are dealt round the warps (a warp's blocks of one layer run in
    // the same stage, so every stage is spread over all 16 warps)
    std::vector<std::vector<DwTile>> per_warp(MLPW_WARPS);
    int next = 0;
    for (int l = L - 2; l >= 0; l--) {
      const int rows = round_up(p.n[l + 1], 8) / 8, cols = round_up(p.n[l] + 1, 8) / 8;
      std::vector<DwTile> blocks;
      for (int mi = 0; mi < rows; mi += 2)
        for (int ni = 0; ni < cols; ni += 2) {
          DwTile d;
          memset(&d, 0, sizeof(d));
          d.a_off = p.b_h_off[l + 1] + 8 * mi;
          d.b_off = (l == 0 ? p.b_x_off : p.b_h_off[l]) + 8 * ni;
          d.dp = p.kph;
          d.kp = (l == 0) ? p.kp0 : p.kph;
          d.layer = l; d.mi = mi; d.ni = ni;
          d.nm = std::min(2, rows - mi); d.nn = std::min(2, cols - ni);
          const int bt = p.n[l] / 8 - ni;                     // block column holding the bias input (column n_l)
          d.pad[0] = (bt >= 0 && bt < d.nn) ? bt : -1;
          d.pad[1] = p.n[l] % 8;
          blocks.push_back(d);
        }
      std::stable_sort(blocks.begin(), blocks.end(), [](const DwTile& x, const DwTile& y) { return x.nm * x.nn > y.nm * y.nn; });
      for (const DwTile& d : blocks) { per_warp[next].push_back(d); next = (next + 1) % MLPW_WARPS; }
    }
    size_t mxs = 0;
    for (const auto& v : per_warp) mxs = std::max(mxs, v.size());
    if (mxs > 4) return false;
    p.nb = mxs <= 2 ? 2 : 4;
    if (table) {
      table->assign((size_t)MLPW_WARPS * p.nb, DwTile());
      for (DwTile& d : *table) memset(&d, 0, sizeof(d));
      for (int w = 0; w < MLPW_WARPS; w++)
        for (size_t s = 0; s < per_warp[w].size(); s++) (*table)[(size_t)w * p.nb + s] = per_warp[w][s];
    }
    bytes += (size_t)MLPW_WARPS * p.nb * sizeof(DwTile);
    p.b_smem = (bytes + 15) & ~(size_t)15;
    if ((size_t)16 * 96 > (size_t)MLPW_TM * p.kp0) return false;      // the epilogue's cross-warp sums alias the descriptor tile
  }
  if (p.f_smem > 227 * 1024 || p.b_smem > 227 * 1024) return false;
  return true;
}

struct CachedW { int n_layers, sizes[PP_MAX_LAYERS + 1]; unsigned tanh_layers, spline_layers; bool ok; MlpwPlan p; bool used; };
static const CachedW& cached_w(const NetDev& net)
{
  static CachedW cache[8];
  static int next = 0;
  for (int i = 0; i < 8; i++) {
    const CachedW& c = cache[i];
    if (c.used && c.n_layers == net.n_layers && c.tanh_layers == net.tanh_layers && c.spline_layers == net.spline_layers &&
        memcmp(c.sizes, net.sizes, sizeof(int) * (net.n_layers + 1)) == 0)
      return c;
  }
  CachedW& c = cache[next];
  next = (next + 1) & 7;
  c.used = true;
  c.n_layers = net.n_layers;
  c.tanh_layers = net.tanh_layers;
  c.spline_layers = net.spline_layers;
  memset(c.sizes, 0, sizeof(c.sizes));
  memcpy(c.sizes, net.sizes, sizeof(int) * (net.n_layers + 1));
  c.ok = mlpw_make_plan(net, c.p, nullptr);
  return c;
}

bool ppk_mlpw_ok(const NetDev& net) { return cached_w(net).ok && getenv("PP_MLP_NO_STORE") == nullptr; }

int ppk_mlpw_table(const NetDev& net, void* host_out, int* n_entries)
{
  MlpwPlan p;
  std::vector<DwTile> table;
  *n_entries = 0;
  if (!mlpw_make_plan(net, p, &table)) return 0;
  *n_entries = (int)table.size();
  if (host_out && !table.empty()) memcpy(host_out, table.data(), table.size() * sizeof(DwTile));
  return 1;
}

// doubles of H (and as many of T') one atom of this network stores
int ppk_mlpw_store_width(const NetDev& net)
{
  int zw = 0;
  for (int l = 1; l < net.n_layers; l++) zw += net.sizes[l];
  return zw;
}

int ppk_mlpw_forward(const MlpArgs& a, const MlpwStore& st, cudaStream_t stream, long long* launches)
{
  if (a.M == 0) return 0;
  const CachedW& cw = cached_w(a.net);
  PP_REQUIRE(cw.ok, "internal: network not eligible for the stored two-phase kernels");
  MlpwPlan p = cw.p;
  {
    const char* sk = getenv("PP_MLPW_SKEW");
    p.skew = sk ? atoi(sk) : 0;
  }
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  const int TMG = 8 * p.MT;
  if (a.rows == nullptr)
    PP_REQUIRE(ppk_tiled_map(a.G, p.n[0], a.M, p.f_kp0, TMG, &tmap), "cuTensorMapEncodeTiled failed for the descriptor array");
  const int n_tiles = (a.M + TMG - 1) / TMG;
  const int grid = std::max(1, std::min(ppk_sm_count(), (n_tiles + p.MG - 1) / p.MG));
#define PP_LAUNCHF(NTW_, MG_, MT_, TF_)                                                                         \
  {                                                                                                             \
    auto k = mlpw_forward_kernel<NTW_, MG_, MT_, TF_>;                                                          \
    PP_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.f_smem));               \
    k<<<grid, MLPW_THREADS, p.f_smem, stream>>>(a, p, st, tmap);                                                \
  }
#define PP_LAUNCHF_TF(NTW_, MG_, MT_)                            \
  {                                                              \
    if (p.tfmode == 1) PP_LAUNCHF(NTW_, MG_, MT_, 1)             \
    else if (p.tfmode == 2) PP_LAUNCHF(NTW_, MG_, MT_, 2)        \
    else PP_LAUNCHF(NTW_, MG_, MT_, 0)                           \
  }
  if (p.NTW == 5) PP_LAUNCHF_TF(5, 3, 3)
  else PP_LAUNCHF_TF(8, 2, 4)
#undef PP_LAUNCHF_TF
#undef PP_LAUNCHF
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

int ppk_mlpw_backward(const MlpArgs& a, const MlpwStore& st, cudaStream_t stream, long long* launches, int* n_cta)
{
  const CachedW& cw = cached_w(a.net);
  PP_REQUIRE(cw.ok, "internal: network not eligible for the stored two-phase kernels");
  PP_REQUIRE(a.M > 0 && a.dw_tiles2 != nullptr, "internal: stored two-phase backward without rows or work table");
  MlpwPlan p = cw.p;
  MlpwMaps maps;
  memset(&maps, 0, sizeof(maps));
  if (a.rows == nullptr)
    PP_REQUIRE(ppk_tiled_map(a.G, p.n[0], a.M, p.kp0, MLPW_TM, &maps.m[0]), "cuTensorMapEncodeTiled failed for the descriptor array");
  size_t off = 0;
  for (int l = 1; l < p.L; l++) {
    PP_REQUIRE(ppk_tiled_map(st.H + off, p.n[l], a.M, p.kph, MLPW_TM, &maps.m[l]), "cuTensorMapEncodeTiled failed for the stored activations");
    if (l == p.L - 1 && p.b_t_off >= 0)
      PP_REQUIRE(ppk_tiled_map(st.T + off, p.n[l], a.M, p.kph, MLPW_TM, &maps.m[MLPW_TMAP]),
                 "cuTensorMapEncodeTiled failed for the stored derivatives");
    off += (size_t)a.M * p.n[l];
  }
  const int grid = ppk_sm_count();          // every CTA writes its (possibly zero) partial row
#define PP_LAUNCHB(NTW_, MG_, MT_, NB_)                                                                         \
  {                                                                                                             \
    auto k = mlpw_backward_kernel<NTW_, MG_, MT_, NB_>;                                                         \
    PP_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.b_smem));               \
    k<<<grid, MLPW_THREADS, p.b_smem, stream>>>(a, p, st, maps);                                                \
  }
  if (p.NTW == 5) { if (p.nb == 2) PP_LAUNCHB(5, 3, 3, 2) else PP_LAUNCHB(5, 3, 3, 4) }
  else { if (p.nb == 2) PP_LAUNCHB(8, 2, 4, 2) else PP_LAUNCHB(8, 2, 4, 4) }
#undef PP_LAUNCHB
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  *n_cta = grid;
  return 0;
}
