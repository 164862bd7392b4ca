// FP32 mode on the fifth-generation tensor cores (sm_100a): the single-pass training epoch of pp_mlp32.cu with the forward
// products and the delta product issued as tcgen05.mma (kind::tf32) from shared memory into TMEM.
//
// Same arithmetic contract as pp_mlp32.cu (north_star's second tolerance, relative 1e-5 against the FP64 path; the
// reference has one arithmetic type, Common.h:53-61): every FP32 operand is split into hi = tf32(x) and lo = tf32(x - hi),
// and A B ~= Alo Bhi + Ahi Blo + Ahi Bhi is accumulated in FP32 -- here in TMEM; energies, errors and the sums of the
// per-CTA partial gradients stay FP64.
//
// Layout of one tile (128 atoms, one CTA per SM, 16 warps):
//   * the descriptor tile arrives by one 2-D tensor copy as FP32 rows (already normalised, pp_mlp32.cu: g_to_f32_kernel);
//   * A operands (X, then H_1, then delta_2) are written by all threads as hi / lo pairs in the canonical K-major layout
//     without swizzle (8-row x 16-byte core matrices; validated on the B200 by tools/umma_probe.cu), B operands (W_l and
//     W_1^T, hi / lo) once per CTA;
//   * ONE thread issues the MMAs of a layer (M = 128, N = padded width, K = 8 per instruction, three per k-step) and commits
//     them to an mbarrier; the accumulator (128 lanes x N columns of TMEM) is read back with tcgen05.ld.32x32b, thread =
//     (row, quarter of the columns), which runs tanh and writes the activations both as FP32 rows (for the weight
//     gradients) and as the next layer's hi / lo operand;
//   * the weight gradients dW_l += delta_{l+1}^T [H_l, 1] have K = atoms: on tcgen05 that needs every activation and delta
//     tile a second time, transposed, in canonical layout (MN-major descriptors without swizzle returned zeros in the
//     probe), so they stay on mma.sync.m16n8k8.tf32 register fragments as in pp_mlp32.cu, and run WHILE the delta product
//     is in flight on the tensor pipe.
#include "pp_mlp_dev.cuh"
#include <algorithm>
#include <cstring>
#include <vector>

#define U32_TM 128
#define U32_WARPS 16
#define U32_THREADS (U32_WARPS * 32)
#define U32_GEO 64                    // tiles of a CTA whose geometry is cached in shared memory

struct Mlp32uPlan {
  int L;                               // weight layers (2 or 3)
  int n[4];                            // sizes[0..L]
  int NGp;                             // row pitch of the FP32 descriptor array in HBM
  int kp[3];                           // FP32 row pitch of X / H_l in shared memory (== 4 mod 8)
  int K[3];                            // MMA depth of layer l: n_l + 1 (bias) rounded up to 8
  int Kd;                              // MMA depth of the delta product: n_2 rounded up to 8
  int NP;                              // padded output width: a multiple of 16 holding max(n_l) + 1
  int x_off, h_off[3];                 // floats: FP32 rows
  int a_hi_off, a_lo_off;              // floats: A operand [128][K] canonical
  int w_hi_off[2], w_lo_off[2];        // floats: W_l [NP][K_l] canonical
  int t_hi_off, t_lo_off;              // floats: W_1^T [NP][Kd] canonical
  int wout_off;                        // floats: output node [n_{L-1} + 1]
  int ypart_off_bytes;                 // doubles [4][128]
  int dbl_off_bytes;                   // doubles: erow, cvec, sttgt [128] each
  int stint_off_bytes, mbar_off_bytes, tmem_off_bytes, tgeo_off_bytes, tab_off_bytes;
  int n_dw_tiles;
  size_t smem_bytes;
};

struct Dw32uTile { int layer, mi, ni, pad; };      // same records as pp_mlp32.cu's table (ppk_mlp32_table)

__host__ __device__ inline int u32_kmaj(int r, int k, int K) { return ((r >> 3) * (K >> 2) + (k >> 2)) * 32 + (r & 7) * 4 + (k & 3); }

__device__ __forceinline__ unsigned u32_tf32(float x)
{
  unsigned r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void u32_split(float x, float& hi, float& lo)
{
  hi = __uint_as_float(u32_tf32(x));
  lo = __uint_as_float(u32_tf32(x - hi));
}
__device__ __forceinline__ void u32_split_u(float x, unsigned& hi, unsigned& lo)
{
  hi = u32_tf32(x);
  lo = u32_tf32(x - __uint_as_float(hi));
}
__device__ __forceinline__ void u32_mma_sync(float (&c)[4], const unsigned (&a)[4], unsigned b0, unsigned b1)
{
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float u32_tanh(float x)
{
  const float ax = fabsf(x);
  const float e = __expf(-2.0f * ax);
  const float t = __fdividef(1.0f - e, 1.0f + e);
  return copysignf(t, x);
}

// shared-memory matrix descriptor, K-major, no swizzle: start address, leading (between the two 16-byte K chunks of one
// instruction) and stride (between 8-row groups) byte offsets, all >> 4; descriptor version 1
__device__ __forceinline__ unsigned long long u32_desc(unsigned addr, unsigned lbo_bytes, unsigned sbo_bytes)
{
  unsigned long long d = 0;
  d |= (unsigned long long)((addr >> 4) & 0x3FFF);
  d |= (unsigned long long)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (unsigned long long)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= 1ull << 46;
  return d;
}
// instruction descriptor: FP32 accumulator, TF32 x TF32, both operands K-major, N >> 3, M >> 4
__host__ __device__ inline unsigned u32_idesc(int M, int N) { return (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24); }

__device__ __forceinline__ void u32_umma(unsigned tmem_d, unsigned long long da, unsigned long long db, unsigned idesc, unsigned accumulate)
{
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void u32_commit(unsigned long long* bar)
{
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(bar)) : "memory");
}
// wait for a phase of the MMA barrier; bounded, so that a descriptor mistake ends in wrong numbers and an error flag
// instead of a hung GPU
__device__ __forceinline__ bool u32_wait(unsigned long long* bar, unsigned parity)
{
  const unsigned addr = smem_addr(bar);
  for (int it = 0; it < (1 << 22); it++) {
    unsigned done;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (done) return true;
  }
  return false;
}

// the three products of one layer: D[128 x NP] (+)= A_lo B_hi + A_hi B_lo + A_hi B_hi over K in steps of 8.  The four
// descriptors are built once; a k-step moves each start address by 256 bytes (16 in the descriptor's 16-byte units --
// shared-memory addresses stay below 2^18, so the 14-bit field never carries).
__device__ __forceinline__ void u32_layer_mma(unsigned tmem_d, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, int K, int NP)
{
  const unsigned idesc = u32_idesc(U32_TM, NP);
  const unsigned sbo = (unsigned)(K >> 2) * 128u;
  unsigned long long dah = u32_desc(smem_addr(a_hi), 128u, sbo), dal = u32_desc(smem_addr(a_lo), 128u, sbo);
  unsigned long long dbh = u32_desc(smem_addr(b_hi), 128u, sbo), dbl = u32_desc(smem_addr(b_lo), 128u, sbo);
  const int steps = K >> 3;
  u32_umma(tmem_d, dal, dbh, idesc, 0u);
  u32_umma(tmem_d, dah, dbl, idesc, 1u);
  u32_umma(tmem_d, dah, dbh, idesc, 1u);
  for (int s = 1; s < steps; s++) {
    dah += 16ull; dal += 16ull; dbh += 16ull; dbl += 16ull;
    u32_umma(tmem_d, dal, dbh, idesc, 1u);
    u32_umma(tmem_d, dah, dbl, idesc, 1u);
    u32_umma(tmem_d, dah, dbh, idesc, 1u);
  }
}

// this thread's accumulator columns of its row (12 or 16), loads and the wait in ONE asm statement so that no use of the
// results can be scheduled ahead of tcgen05.wait::ld
__device__ __forceinline__ void u32_ld_acc(float (&z)[12], unsigned taddr)
{
  unsigned v[12];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%12];\n"
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%4,%5,%6,%7}, [%13];\n"
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%8,%9,%10,%11}, [%14];\n"
      "tcgen05.wait::ld.sync.aligned;\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11])
      : "r"(taddr), "r"(taddr + 4u), "r"(taddr + 8u)
      : "memory");
#pragma unroll
  for (int i = 0; i < 12; i++) z[i] = __uint_as_float(v[i]);
}
__device__ __forceinline__ void u32_ld_acc(float (&z)[16], unsigned taddr)
{
  unsigned v[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%16];\n"
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%4,%5,%6,%7}, [%17];\n"
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%8,%9,%10,%11}, [%18];\n"
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%12,%13,%14,%15}, [%19];\n"
      "tcgen05.wait::ld.sync.aligned;\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]),
        "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr), "r"(taddr + 4u), "r"(taddr + 8u), "r"(taddr + 12u)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; i++) z[i] = __uint_as_float(v[i]);
}

// NC4: a thread's accumulator columns / 4 (NP = 16 NC4 = 48 or 64); NH: hidden layers (L - 1)
template <int NC4, int NH>
__global__ void __launch_bounds__(U32_THREADS, 1)
mlp32_umma_kernel(MlpArgs a, Mlp32uPlan p, const __grid_constant__ CUtensorMap tmap, int* __restrict__ err_flag)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* sm = reinterpret_cast<float*>(smem_raw);
  double* smd = reinterpret_cast<double*>(smem_raw + p.dbl_off_bytes);
  double* ypart = reinterpret_cast<double*>(smem_raw + p.ypart_off_bytes);
  const Dw32uTile* tab = reinterpret_cast<const Dw32uTile*>(smem_raw + p.tab_off_bytes);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  constexpr int TM = U32_TM, NC = 4 * NC4, NP = 16 * NC4;
  const int L = p.L, NG = p.n[0], kp0 = p.kp[0];
  const NetDev& net = a.net;
  // epilogue ownership: TMEM lane = row, a quarter of the accumulator's columns
  const int erow_ = 32 * (warp & 3) + lane, ecol0 = NC * (warp >> 2);

  // ---- one-time staging ---------------------------------------------------------------------------------------------
  for (int l = 0; l < NH; l++) {          // W_l [NP][K_l], hi / lo, canonical K-major
    const int n_in = p.n[l], n_out = p.n[l + 1], K = p.K[l];
    const double* src = net.params + net.w_off[l];
    for (int i = tid; i < NP * K; i += U32_THREADS) {
      const int r = i / K, k = i - r * K;
      const float w = (r < n_out && k <= n_in) ? (float)src[(size_t)r * (n_in + 1) + k] : 0.0f;
      float hi, lo;
      u32_split(w, hi, lo);
      sm[p.w_hi_off[l] + u32_kmaj(r, k, K)] = hi;
      sm[p.w_lo_off[l] + u32_kmaj(r, k, K)] = lo;
    }
    if (l >= 1) {                         // FP32 rows of H_l: bias input in column n_l
      float* Hs = sm + p.h_off[l];
      for (int i = tid; i < TM * p.kp[l]; i += U32_THREADS) Hs[i] = ((i % p.kp[l]) == n_in) ? 1.0f : 0.0f;
    }
  }
  {                                       // FP32 rows of the last hidden layer
    float* Hs = sm + p.h_off[NH];
    for (int i = tid; i < TM * p.kp[NH]; i += U32_THREADS) Hs[i] = ((i % p.kp[NH]) == p.n[NH]) ? 1.0f : 0.0f;
  }
  if (NH == 2) {                          // W_1^T [NP (n_1)][Kd (n_2)]: B operand of delta_1 = delta_2 W_1
    const int n1 = p.n[1], n2 = p.n[2], K = p.Kd;
    const double* src = net.params + net.w_off[1];
    for (int i = tid; i < NP * K; i += U32_THREADS) {
      const int r = i / K, k = i - r * K;
      const float w = (r < n1 && k < n2) ? (float)src[(size_t)k * (n1 + 1) + r] : 0.0f;
      float hi, lo;
      u32_split(w, hi, lo);
      sm[p.t_hi_off + u32_kmaj(r, k, K)] = hi;
      sm[p.t_lo_off + u32_kmaj(r, k, K)] = lo;
    }
  }
  {
    const int n_h = p.n[NH];
    const double* src = net.params + net.w_off[NH];
    for (int i = tid; i <= n_h; i += U32_THREADS) sm[p.wout_off + i] = (float)src[i];
  }
  for (int i = tid; i < TM * kp0; i += U32_THREADS) sm[p.x_off + i] = 0.0f;
  {
    const int nwords = p.n_dw_tiles * (int)(sizeof(Dw32uTile) / sizeof(int));
    int* dst = reinterpret_cast<int*>(smem_raw + p.tab_off_bytes);
    const int* srcw = reinterpret_cast<const int*>(a.dw_tiles2);
    for (int i = tid; i < nwords; i += U32_THREADS) dst[i] = srcw[i];
  }
  double* erow = smd;                      // [TM]
  double* cvec = smd + TM;                 // [TM]
  double* st_tgt = smd + 2 * TM;           // [TM]
  double* red = reinterpret_cast<double*>(sm + p.a_hi_off);      // epilogue only: aliases the dead operand buffer
  int* st_lo = reinterpret_cast<int*>(smem_raw + p.stint_off_bytes);
  int* st_tag = st_lo + TM + 1;
  unsigned long long* mbar = reinterpret_cast<unsigned long long*>(smem_raw + p.mbar_off_bytes);      // [0] tile copy, [1] MMAs
  unsigned* tmem_slot = reinterpret_cast<unsigned*>(smem_raw + p.tmem_off_bytes);
  float* X = sm + p.x_off;
  float* A_hi = sm + p.a_hi_off;
  float* A_lo = sm + p.a_lo_off;
  const float* wout = sm + p.wout_off;

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(tmem_slot)), "r"(64) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) { mbar_init(&mbar[0], 1); mbar_init(&mbar[1], 1); }
  fence_proxy_async();
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned tmem = *tmem_slot;
  const unsigned taddr = tmem + ((unsigned)(32 * (warp & 3)) << 16) + (unsigned)ecol0;

  constexpr int MAXT = 4;
  float dW[MAXT][4];
#pragma unroll
  for (int t = 0; t < MAXT; t++) { dW[t][0] = dW[t][1] = dW[t][2] = dW[t][3] = 0.0f; }
  float dwout[3] = {0.0f, 0.0f, 0.0f};
  double sse_w = 0.0, vsse_w = 0.0, nval_w = 0.0;
  unsigned mma_phase = 0;
  bool bad = false;
#ifdef PP_U32_TIMING
  long long tph[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, tlast = clock64();
#define U32_T(i) { const long long t_ = clock64(); tph[i] += t_ - tlast; tlast = t_; }
#else
#define U32_T(i)
#endif

  const int n_tiles = a.n_tiles;
  struct Geom { int s0, s1, row0, nvalid; };
  // geometry of this CTA's first U32_GEO tiles, looked up once: walking tile -> first structure -> first atom per tile is a
  // chain of dependent global loads that a single resident CTA cannot hide
  int4* tgeo = reinterpret_cast<int4*>(smem_raw + p.tgeo_off_bytes);
  if (tid < U32_GEO) {
    const int tile = blockIdx.x + tid * gridDim.x;
    int4 v = make_int4(0, 0, 0, 0);
    if (tile < n_tiles) {
      v.x = a.tile_struct0[tile];
      v.y = a.tile_struct0[tile + 1];
      const long long r0 = a.struct_atom0[v.x];
      v.z = (int)r0;
      v.w = (int)(a.struct_atom0[v.y] - r0);
    }
    tgeo[tid] = v;
  }
  __syncthreads();
  auto tile_geom = [&](int tile) {
    Geom gm;
    gm.s0 = gm.s1 = gm.row0 = gm.nvalid = 0;
    const int slot = (tile - (int)blockIdx.x) / (int)gridDim.x;
    if (slot < U32_GEO) {
      const int4 v = tgeo[slot];
      gm.s0 = v.x; gm.s1 = v.y; gm.row0 = v.z; gm.nvalid = v.w;
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
  auto issue_load = [&](const Geom& gm) {
    if (tid == 0) {
      fence_proxy_async();
      mbar_arrive_expect_tx(&mbar[0], (unsigned)TM * (unsigned)kp0 * 4u);
      tma_load_tile(X, &tmap, gm.row0, &mbar[0]);
    }
  };
  // operands written through the generic proxy -> visible to the tensor pipe; then one thread issues the layer's MMAs
  auto publish_and_issue = [&](const float* b_hi, const float* b_lo, int K) {
    fence_proxy_async();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      u32_layer_mma(tmem, A_hi, A_lo, b_hi, b_lo, K, NP);
      u32_commit(&mbar[1]);
    }
  };
  auto wait_mma = [&]() {
    if (!u32_wait(&mbar[1], mma_phase & 1u)) bad = true;
    mma_phase++;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  };
  // this thread's NC values of its row as the next product's A operand: hi / lo, canonical layout of depth K
  auto stage_row = [&](const float (&v)[NC], int K) {
#pragma unroll
    for (int c = 0; c < NC4; c++) {
      const int k0 = ecol0 + 4 * c;
      if (k0 < K) {
        float4 hi, lo;
        u32_split(v[4 * c], hi.x, lo.x); u32_split(v[4 * c + 1], hi.y, lo.y);
        u32_split(v[4 * c + 2], hi.z, lo.z); u32_split(v[4 * c + 3], hi.w, lo.w);
        const int o = u32_kmaj(erow_, k0, K);
        *reinterpret_cast<float4*>(A_hi + o) = hi;
        *reinterpret_cast<float4*>(A_lo + o) = lo;
      }
    }
  };

  long long pf_a0 = 0;                     // raw: the subtraction waits until the value is used, a tile later
  int pf_tag = 0;
  double pf_tgt = 0.0;
  auto st_fetch = [&](const Geom& gm) {
    if (tid <= gm.s1 - gm.s0) {
      const int sidx = gm.s0 + tid;
      pf_a0 = a.struct_atom0[sidx];
      if (sidx < gm.s1) { pf_tgt = a.target[sidx]; pf_tag = a.tag[sidx]; }
    }
  };
  Geom g_cur = tile_geom(blockIdx.x), g_nxt = tile_geom(blockIdx.x + gridDim.x);
  if ((int)blockIdx.x < n_tiles) issue_load(g_cur);
  st_fetch(g_cur);

  const int rb = 8 * warp;                 // this warp's rows in the row-owned phases (structure errors, output node)
  int n_done = 0;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, n_done++) {
    const int nvalid = g_cur.nvalid;
    const int s_first = g_cur.s0, s_last = g_cur.s1;
    const Geom g_nn = tile_geom(tile + 2 * gridDim.x);
    // structure table of this tile (read after barrier A) from the registers filled one tile ago; then the next tile's entries
    // start their flight (a tile holds at most 128 structures: one entry per thread)
    if (tid <= s_last - s_first) {
      st_lo[tid] = (int)(pf_a0 - g_cur.row0);
      if (tid < s_last - s_first) { st_tgt[tid] = pf_tgt; st_tag[tid] = pf_tag; }
    }
    st_fetch(g_nxt);
    U32_T(0)
    mbar_wait(&mbar[0], (unsigned)n_done & 1u);
    U32_T(1)

    // ---- X (normalised FP32 rows; column NG = bias input) -> hi / lo operand of depth K_0 ------------------------------------
    {
      const int K = p.K[0], nchunk = K >> 2;
      for (int u = tid; u < TM * nchunk; u += U32_THREADS) {
        const int r = u & (TM - 1), c = u >> 7;
        float4 v = *reinterpret_cast<const float4*>(X + (size_t)r * kp0 + 4 * c);
        const int k0 = 4 * c;
        if (NG >= k0 && NG < k0 + 4) {       // the bias column lies in this chunk
          if (NG == k0) v.x = 1.0f; else if (NG == k0 + 1) v.y = 1.0f; else if (NG == k0 + 2) v.z = 1.0f; else v.w = 1.0f;
          X[(size_t)r * kp0 + NG] = 1.0f;     // the FP32 rows feed the weight gradients
        }
        float4 hi, lo;
        u32_split(v.x, hi.x, lo.x); u32_split(v.y, hi.y, lo.y); u32_split(v.z, hi.z, lo.z); u32_split(v.w, hi.w, lo.w);
        const int o = u32_kmaj(r, k0, K);
        *reinterpret_cast<float4*>(A_hi + o) = hi;
        *reinterpret_cast<float4*>(A_lo + o) = lo;
      }
    }

    U32_T(2)
    // ---- forward: one MMA batch per hidden layer; the epilogue thread owns (row, a quarter of the columns) ---------------------
    float hlast[NC];                         // this thread's activations of the last hidden layer
    double yp = 0.0;
#pragma unroll
    for (int l = 0; l < NH; l++) {
      publish_and_issue(sm + p.w_hi_off[l], sm + p.w_lo_off[l], p.K[l]);
      U32_T(3)
      wait_mma();
      U32_T(4)
      float z[NC];
      u32_ld_acc(z, taddr);
      const int n_out = p.n[l + 1];
      float* Hr = sm + p.h_off[l + 1] + (size_t)erow_ * p.kp[l + 1];
#pragma unroll
      for (int c = 0; c < NC; c++) {
        const int n = ecol0 + c;
        z[c] = (n < n_out) ? u32_tanh(z[c]) : ((n == n_out) ? 1.0f : 0.0f);      // column n_out: the next layer's bias input
        if (n < n_out) Hr[n] = z[c];
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");          // the accumulator may be overwritten after the next barrier
      if (l + 1 < NH) stage_row(z, p.K[l + 1]);
      else {
#pragma unroll
        for (int c = 0; c < NC; c++) {
          const int n = ecol0 + c;
          hlast[c] = z[c];
          if (n < n_out) yp += (double)z[c] * (double)wout[n];
        }
      }
    }
    U32_T(5)
    ypart[(warp >> 2) * TM + erow_] = yp;
    __syncthreads();
    if (tid < TM) {
      const double y = ypart[tid] + ypart[TM + tid] + ypart[2 * TM + tid] + ypart[3 * TM + tid] + (double)wout[p.n[NH]];
      erow[tid] = (tid < nvalid) ? a.out_scale * y : 0.0;
    }
    __syncthreads();                                                                   // barrier A: erow, structure table

    // ---- per-structure energy, error and loss terms in FP64 (Potential.cpp:641-668): the factor c of this warp's rows ---
    {
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
    }
    __syncthreads();                                                                   // cvec of every row

    // ---- output-node gradient over this warp's rows (reads the FP32 rows of the last hidden layer) ---------------------------
    {
      const int n_h = p.n[NH], kph = p.kp[NH];
      const float* Hl = sm + p.h_off[NH];
#pragma unroll
      for (int j = 0; j < 3; j++) {
        const int k = lane + 32 * j;
        if (k <= n_h) {
          float s_ = 0.0f;
#pragma unroll
          for (int r = 0; r < 8; r++) s_ = fmaf((float)cvec[rb + r], Hl[(size_t)(rb + r) * kph + k], s_);
          dwout[j] += s_;
        }
      }
    }
    __syncthreads();                                                                   // those rows are about to be overwritten

    U32_T(6)
    // ---- delta of the last hidden layer, c w_out (1 - H^2), in place of its FP32 rows and as the delta product's operand ------
    float dlast[NC];
    {
      const int n_h = p.n[NH];
      const float c = (float)cvec[erow_];
      float* Hr = sm + p.h_off[NH] + (size_t)erow_ * p.kp[NH];
#pragma unroll
      for (int cc = 0; cc < NC; cc++) {
        const int n = ecol0 + cc;
        const float h = hlast[cc];
        dlast[cc] = (n < n_h) ? c * wout[n] * (1.0f - h * h) : 0.0f;
        if (n < n_h) Hr[n] = dlast[cc];
      }
    }
    if (NH == 2) {
      stage_row(dlast, p.Kd);
      publish_and_issue(sm + p.t_hi_off, sm + p.t_lo_off, p.Kd);      // (its barrier also publishes the FP32 rows of delta_2)
    } else {
      __syncthreads();
    }

    // ---- weight gradients dW_l += delta_{l+1}^T [H_l, 1] over the tile's atoms (mma.sync, 3 x TF32); with two hidden layers the
    //      delta product runs on the tensor pipe meanwhile -------------------------------------------------------------------------
    auto dw_layer = [&](int l) {
      const float* Dl = sm + p.h_off[l + 1];                          // delta_{l+1} [atom][neuron]
      const float* Bl = (l == 0) ? X : sm + p.h_off[l];               // [H_l, 1]   [atom][input]
      const int dp = p.kp[l + 1], bp = p.kp[l];
#pragma unroll
      for (int t = 0; t < MAXT; t++) {
        const int ti = warp + t * U32_WARPS;
        if (ti < p.n_dw_tiles) {
          const Dw32uTile d = tab[ti];
          if (d.layer == l) {
            const float* pa = Dl + (size_t)q * dp + 16 * d.mi + g;    // A(m = neuron, k = atom) = delta[atom][neuron]
            const float* pb = Bl + (size_t)q * bp + 8 * d.ni + g;     // B(k = atom, n = input)
#pragma unroll 4
            for (int k0 = 0; k0 < TM; k0 += 8) {
              unsigned ah[4], al[4], bh0, bh1, bl0, bl1;
              u32_split_u(pa[(size_t)k0 * dp], ah[0], al[0]);
              u32_split_u(pa[(size_t)k0 * dp + 8], ah[1], al[1]);
              u32_split_u(pa[(size_t)(k0 + 4) * dp], ah[2], al[2]);
              u32_split_u(pa[(size_t)(k0 + 4) * dp + 8], ah[3], al[3]);
              u32_split_u(pb[(size_t)k0 * bp], bh0, bl0);
              u32_split_u(pb[(size_t)(k0 + 4) * bp], bh1, bl1);
              u32_mma_sync(dW[t], al, bh0, bh1);
              u32_mma_sync(dW[t], ah, bl0, bl1);
              u32_mma_sync(dW[t], ah, bh0, bh1);
            }
          }
        }
      }
    };
    U32_T(7)
    dw_layer(NH - 1);
    U32_T(8)
    if (NH == 2) {
      wait_mma();
      U32_T(9)
      float z[NC];
      u32_ld_acc(z, taddr);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      const int n_1 = p.n[1];
      float* Hr = sm + p.h_off[1] + (size_t)erow_ * p.kp[1];
#pragma unroll
      for (int c = 0; c < NC; c++) {
        const int n = ecol0 + c;
        const float h = (n < n_1) ? Hr[n] : 0.0f;
        z[c] *= (1.0f - h * h);
      }
      __syncthreads();                                                                 // every product reading H_1 is done
#pragma unroll
      for (int c = 0; c < NC; c++) {
        const int n = ecol0 + c;
        if (n < n_1) Hr[n] = z[c];
      }
      __syncthreads();                                                                 // delta_1 complete in all rows
      U32_T(10)
      dw_layer(0);
    }
    __syncthreads();                                                                   // X and the operand buffers are free
    U32_T(11)
    if (tile + (int)gridDim.x < n_tiles) issue_load(g_nxt);
    g_cur = g_nxt;
    g_nxt = g_nn;
  }

  // ---- per-CTA partial gradient and loss terms -> scratch (FP64, summed in a fixed order by the reduction kernel) ------------
  if (bad) atomicExch(err_flag, 1);
#ifdef PP_U32_TIMING
  if (blockIdx.x == 0 && (tid == 0 || tid == 255))
    printf("u32 timing (thread %d, %d tiles) cycles/tile: table %lld, copy wait %lld, stage X %lld, publish+issue %lld, mma wait %lld, epilogue %lld, "
           "errors+dwout %lld, delta+issue %lld, dW_last %lld, mma wait %lld, delta1 %lld, dW0+sync %lld\n", tid, n_done,
           tph[0] / n_done, tph[1] / n_done, tph[2] / n_done, tph[3] / n_done, tph[4] / n_done, tph[5] / n_done, tph[6] / n_done, tph[7] / n_done,
           tph[8] / n_done, tph[9] / n_done, tph[10] / n_done, tph[11] / n_done);
#endif
  double* out = a.partial + (size_t)blockIdx.x * (a.P_total + 3) + a.p_off;
#pragma unroll
  for (int t = 0; t < MAXT; t++) {
    const int ti = warp + t * U32_WARPS;
    if (ti < p.n_dw_tiles) {
      const Dw32uTile d = tab[ti];
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
  if (tid <= p.n[NH]) {
    double s_ = 0.0;
    for (int w = 0; w < U32_WARPS; w++) s_ += red[w * 96 + tid];
    out[net.w_off[NH] + tid] = s_;
  }
  __syncthreads();
  if (lane == 0) { red[3 * warp] = sse_w; red[3 * warp + 1] = vsse_w; red[3 * warp + 2] = nval_w; }
  __syncthreads();
  if (tid < 3) {
    double s_ = 0.0;
    for (int w = 0; w < U32_WARPS; w++) s_ += red[3 * w + tid];
    a.partial[(size_t)blockIdx.x * (a.P_total + 3) + a.P_total + tid] = s_;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64) : "memory");
}

// ---------------------------------------------------------------------------------------------------------
static int u32_pitch(int v)
{
  int p = v;
  while ((p & 7) != 4) p++;
  return p;
}

static bool mlp32u_make_plan(const NetDev& net, Mlp32uPlan& p)
{
  memset(&p, 0, sizeof(p));
  const int L = net.n_layers;
  if (L < 2 || L > 3) return false;
  int mx = 0;
  for (int l = 1; l < L; l++) {
    if (!((net.tanh_layers >> l) & 1)) return false;         // FP32 mode: tanh hidden layers only
    mx = std::max(mx, net.sizes[l]);
  }
  if (mx + 1 > 64 || net.sizes[0] + 1 > 248) return false;
  p.L = L;
  for (int l = 0; l <= L; l++) p.n[l] = net.sizes[l];
  p.NGp = (p.n[0] + 3) & ~3;
  p.NP = ((mx + 1 + 15) / 16) * 16;
  if (p.NP < 48) p.NP = 48;
  for (int l = 0; l < L - 1; l++) {
    p.K[l] = (p.n[l] + 1 + 7) & ~7;
    p.kp[l] = u32_pitch(p.K[l]);
  }
  p.kp[L - 1] = u32_pitch((p.n[L - 1] + 1 + 7) & ~7);
  p.Kd = (L == 3) ? ((p.n[2] + 7) & ~7) : 8;
  const int Kmax = std::max(std::max(p.K[0], L == 3 ? p.K[1] : 0), p.Kd);
  auto align32 = [](size_t v) { return (v + 31) & ~(size_t)31; };      // floats: 128 bytes
  size_t off = 0;
  p.x_off = 0; off += (size_t)U32_TM * p.kp[0];
  for (int l = 1; l < L; l++) { off = align32(off); p.h_off[l] = (int)off; off += (size_t)U32_TM * p.kp[l]; }
  off = align32(off); p.a_hi_off = (int)off; off += (size_t)U32_TM * Kmax;
  off = align32(off); p.a_lo_off = (int)off; off += (size_t)U32_TM * Kmax;
  for (int l = 0; l < L - 1; l++) {
    off = align32(off); p.w_hi_off[l] = (int)off; off += (size_t)p.NP * p.K[l];
    off = align32(off); p.w_lo_off[l] = (int)off; off += (size_t)p.NP * p.K[l];
  }
  if (L == 3) {
    off = align32(off); p.t_hi_off = (int)off; off += (size_t)p.NP * p.Kd;
    off = align32(off); p.t_lo_off = (int)off; off += (size_t)p.NP * p.Kd;
  }
  off = align32(off); p.wout_off = (int)off; off += 64;
  size_t bytes = off * sizeof(float);
  bytes = (bytes + 15) & ~(size_t)15;
  p.ypart_off_bytes = (int)bytes; bytes += (size_t)4 * U32_TM * sizeof(double);
  p.dbl_off_bytes = (int)bytes; bytes += (size_t)3 * U32_TM * sizeof(double);
  p.stint_off_bytes = (int)bytes; bytes += (size_t)(2 * U32_TM + 2) * sizeof(int);
  bytes = (bytes + 15) & ~(size_t)15;
  p.mbar_off_bytes = (int)bytes; bytes += 16;
  p.tmem_off_bytes = (int)bytes; bytes += 16;
  p.tgeo_off_bytes = (int)bytes; bytes += (size_t)U32_GEO * 4 * sizeof(int);
  p.tab_off_bytes = (int)bytes;
  int n_tiles = 0;
  for (int l = L - 2; l >= 0; l--) n_tiles += ((p.n[l + 1] + 15) / 16) * ((p.n[l] + 1 + 7) / 8);
  if (n_tiles > 4 * U32_WARPS) return false;
  if ((size_t)U32_WARPS * 96 * sizeof(double) > (size_t)U32_TM * Kmax * sizeof(float)) return false;
  p.n_dw_tiles = n_tiles;
  bytes += (size_t)n_tiles * sizeof(Dw32uTile);
  p.smem_bytes = (bytes + 15) & ~(size_t)15;
  return p.smem_bytes <= 227 * 1024;
}

// the network can run the tcgen05 form (PP_FP32_NO_UMMA=1 keeps the mma.sync kernel of pp_mlp32.cu)
bool ppk_mlp32u_ok(const NetDev& net)
{
  Mlp32uPlan p;
  return getenv("PP_FP32_NO_UMMA") == nullptr && mlp32u_make_plan(net, p);
}
int ppk_mlp32u_tile_atoms(void) { return U32_TM; }

static int* g_u32_err[16] = {nullptr};
static int* u32_err_flag()
{
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 16) dev = 0;
  if (!g_u32_err[dev]) {
    if (cudaMalloc((void**)&g_u32_err[dev], sizeof(int)) != cudaSuccess) return nullptr;
    cudaMemset(g_u32_err[dev], 0, sizeof(int));
  }
  return g_u32_err[dev];
}
// 1 when some launch since the last call saw a tensor-pipe batch that never completed (synchronises the stream)
int ppk_mlp32u_failed(cudaStream_t st)
{
  int* f = u32_err_flag();
  if (!f) return 0;
  int h = 0;
  if (cudaMemcpyAsync(&h, f, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess) return 0;
  cudaStreamSynchronize(st);
  if (h) cudaMemsetAsync(f, 0, sizeof(int), st);
  return h;
}

int ppk_mlp32u_launch_fused(const MlpArgs& a, const float* G32, long long n_atoms, cudaStream_t st, long long* launches, int* n_cta)
{
  Mlp32uPlan p;
  PP_REQUIRE(mlp32u_make_plan(a.net, p), "internal: network not eligible for the tcgen05 FP32 training kernel");
  PP_REQUIRE(a.dw_tiles2 != nullptr && a.rows == nullptr && a.n_tiles > 0, "internal: FP32 training kernel needs whole-structure tiles");
  int* flag = u32_err_flag();
  PP_REQUIRE(flag != nullptr, "cudaMalloc failed for the FP32 kernel's error flag");
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  PP_REQUIRE(ppk_tiled_map_f32(G32, p.n[0], p.NGp, n_atoms, p.kp[0], U32_TM, &tmap), "cuTensorMapEncodeTiled failed for the FP32 descriptor array");
  const int grid = ppk_sm_count();
  const int NH = p.L - 1;
#define PP_LAUNCH32U(NC4_, NH_)                                                                                 \
  {                                                                                                             \
    auto k = mlp32_umma_kernel<NC4_, NH_>;                                                                      \
    PP_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_bytes));           \
    k<<<grid, U32_THREADS, p.smem_bytes, st>>>(a, p, tmap, flag);                                               \
  }
  if (p.NP == 48 && NH == 1) PP_LAUNCH32U(3, 1)
  else if (p.NP == 48) PP_LAUNCH32U(3, 2)
  else if (NH == 1) PP_LAUNCH32U(4, 1)
  else PP_LAUNCH32U(4, 2)
#undef PP_LAUNCH32U
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  *n_cta = grid;
  return 0;
}
