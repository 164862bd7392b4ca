// Descriptor / force kernels and their launchers (sm_100a).  See pp_desc.cuh for the algorithm.
#include "pp_desc.cuh"
#include "pp_kernels.h"
#include "pp_mlp_dev.cuh"

#define DESC_WARPS 4
#ifndef DESC_BLOCKS_PER_SM
#ifndef DESC_BLOCKS_PER_SM
#define DESC_BLOCKS_PER_SM 4      // 16 warps per SM (<= 128 registers)
#endif
#endif

// ---------------------------------------------------------------------------------------------------------
// K1+K2+K3 fused, periodic structures (training set-up): one warp per atom.
template <int NR_T, int NX, bool BASIS>
__global__ void __launch_bounds__(DESC_WARPS * 32, DESC_BLOCKS_PER_SM)
desc_periodic_kernel(PeriodicArgs a)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const double* etab = pp_block_exp_table();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long rel = (long long)blockIdx.x * DESC_WARPS + warp;
  if (rel >= a.n_atoms) return;
  const long long atom = a.atom_begin + rel;
  const DescSpec* S = a.spec;
  const size_t per_warp = pp_warp_smem_bytes(a.cap, S->NR, S->NG, false);
  WarpNbr w;
  pp_warp_views(w, smem + per_warp * warp, a.cap, S->NR, S->NG, false);
  w.etab = etab;
  const int s = a.atom_struct[atom];
  const long long a0 = a.struct_atom0[s];
  const int natom = (int)(a.struct_atom0[s + 1] - a0);
  const int i = (int)(atom - a0);
  const double* F = a.FtoA + 9 * (size_t)s;
  double Fm[9];
#pragma unroll
  for (int k = 0; k < 9; k++) Fm[k] = F[k];
  const int* m = a.mxyz + 3 * (size_t)s;
  int n = pp_scan_periodic(w, a.cap, lane, a.frac + 3 * a0, a.elem + a0, natom, i, Fm, m[0], m[1], m[2], S, nullptr);
  if (n > a.cap) {
    if (lane == 0) atomicMax(a.overflow, n);
    return;
  }
  pp_finalize_entries<false>(w, a.cap, lane, S, n);
  pp_descriptor_row<NR_T, NX, BASIS>(w, a.cap, lane, S, false);
  double* out = a.G + (size_t)atom * S->NG;
  for (int g = lane; g < S->NG; g += 32) out[g] = w.buf[g];
}

// Parity hook: the (j, image) list of every atom of one structure, in the reference's order.
// pass 0: counts[i]; pass 1: pairs written at offsets[i].
__global__ void __launch_bounds__(DESC_WARPS * 32)
neighbors_periodic_kernel(PeriodicArgs a, int s, int pass, int* counts, const int* offsets, int* out_j, int* out_t)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const DescSpec* S = a.spec;
  const long long a0 = a.struct_atom0[s];
  const int natom = (int)(a.struct_atom0[s + 1] - a0);
  const int i = blockIdx.x * DESC_WARPS + warp;
  if (i >= natom) return;
  const size_t per_warp = pp_warp_smem_bytes(a.cap, S->NR, S->NG, false);
  WarpNbr w;
  pp_warp_views(w, smem + per_warp * warp, a.cap, S->NR, S->NG, false);
  double Fm[9];
#pragma unroll
  for (int k = 0; k < 9; k++) Fm[k] = a.FtoA[9 * (size_t)s + k];
  const int* m = a.mxyz + 3 * (size_t)s;
  int* timg = w.idx;   // unused by the scan itself: receives the image index
  int n = pp_scan_periodic(w, a.cap, lane, a.frac + 3 * a0, a.elem + a0, natom, i, Fm, m[0], m[1], m[2], S, timg);
  if (pass == 0) {
    if (lane == 0) counts[i] = n;
    if (n > a.cap && lane == 0) atomicMax(a.overflow, n);
    return;
  }
  __syncwarp();
  if (n > a.cap) return;
  for (int p = lane; p < n; p += 32) {
    out_j[offsets[i] + p] = w.sidx[p];
    out_t[offsets[i] + p] = timg[p];
  }
}

// ---------------------------------------------------------------------------------------------------------
// MD (LAMMPS list) descriptors: Structure::Calc_G non-periodic branch (Structure.cpp:1073-1145).
template <int NR_T, int NX, bool BASIS>
__global__ void __launch_bounds__(DESC_WARPS * 32, DESC_BLOCKS_PER_SM)
md_desc_kernel(MdArgs a)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const double* etab = pp_block_exp_table();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * DESC_WARPS + warp;
  if (i >= a.nlocal) return;
  const DescSpec* S = a.spec;
  const size_t per_warp = pp_warp_smem_bytes(a.cap, S->NR, S->NG, false);
  WarpNbr w;
  pp_warp_views(w, smem + per_warp * warp, a.cap, S->NR, S->NG, false);
  w.etab = etab;
  const long long p0 = a.neigh_offsets[i];
  const int nlist = (int)(a.neigh_offsets[i + 1] - p0);
  int n = pp_gather_list(w, a.cap, lane, a.x, a.elem, i, a.neigh + p0, nlist, S->Rcut);
  if (n > a.cap) {
    if (lane == 0) atomicMax(a.overflow, n);
    return;
  }
  pp_finalize_entries<false>(w, a.cap, lane, S, n);
#ifndef PP_MD_SETUP_ONLY      // timing experiments only
  pp_descriptor_row<NR_T, NX, BASIS>(w, a.cap, lane, S, S->nG3 > 0);
#endif
  double* out = a.G + (size_t)i * S->NG;
  for (int g = lane; g < S->NG; g += 32) out[g] = w.buf[g];
}

// MD forces: Structure::Get_Forces non-periodic branch (Structure.cpp:429-713) with dE/dG given.
template <int NR_T, int NX, bool BASIS>
__global__ void __launch_bounds__(DESC_WARPS * 32, DESC_BLOCKS_PER_SM)
md_force_kernel(MdArgs a)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const double* etab = pp_block_exp_table();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * DESC_WARPS + warp;
  if (i >= a.nlocal) return;
  const DescSpec* S = a.spec;
  const size_t per_warp = pp_warp_smem_bytes(a.cap, S->NR, S->NG, true);
  WarpNbr w;
  pp_warp_views(w, smem + per_warp * warp, a.cap, S->NR, S->NG, true);
  w.etab = etab;
  const long long p0 = a.neigh_offsets[i];
  const int nlist = (int)(a.neigh_offsets[i + 1] - p0);
  int n = pp_gather_list(w, a.cap, lane, a.x, a.elem, i, a.neigh + p0, nlist, S->Rcut);
  if (n > a.cap) {
    if (lane == 0) atomicMax(a.overflow, n);
    return;
  }
  pp_finalize_entries<true>(w, a.cap, lane, S, n);
  const double* wrow = a.dEdG + (size_t)i * S->NG;
  for (int g = lane; g < S->NG; g += 32) w.buf[g] = wrow[g];
  __syncwarp();
#ifndef PP_MD_SETUP_ONLY
  pp_force_entries<NR_T, NX, BASIS>(w, a.cap, lane, S, S->nG3 > 0);
#endif
  // f[j] -= grad_j ; f[i] += sum_j grad_j   (Structure.cpp:700-706)
  double sx = 0.0, sy = 0.0, sz = 0.0;
  for (int p = lane; p < n; p += 32) {
    double gx = w.fx[p], gy = w.fy[p], gz = w.fz[p];
    size_t j = (size_t)w.idx[p];
    atomicAdd(a.f + 3 * j, -gx);
    atomicAdd(a.f + 3 * j + 1, -gy);
    atomicAdd(a.f + 3 * j + 2, -gz);
    sx += gx; sy += gy; sz += gz;
  }
  sx = pp_warp_sum(sx); sy = pp_warp_sum(sy); sz = pp_warp_sum(sz);
  if (lane == 0) {
    atomicAdd(a.f + 3 * (size_t)i, sx);
    atomicAdd(a.f + 3 * (size_t)i + 1, sy);
    atomicAdd(a.f + 3 * (size_t)i + 2, sz);
  }
}

// ---------------------------------------------------------------------------------------------------------
// One atom's network inside its warp: Neural_network::evaluate_MD(dE_dG) (NN.cpp:516-570).  Descriptors come from and
// dE/dG goes back to w.buf; activations live in the staging arrays (free after the descriptor pass), derivatives T' in
// registers (a lane owns neurons lane and lane + 32 of every layer).  Forward products read the TRANSPOSED weights
// (lanes = consecutive neurons: coalesced), backward products the original ones (lanes = consecutive inputs).
// ~2 P FMAs per atom against ~10^5 in the pair loops: the point is not speed but that G and dE/dG never leave the SM
// and the neighbour gather, unit vectors and radial factors are built once per atom instead of once per kernel.
__device__ inline double pp_warp_network(WarpNbr& w, int lane, int NG, const NetDev& net, const double* __restrict__ WT, double out_scale,
                                         const double* __restrict__ table, const double* __restrict__ exptab)
{
  const int L = net.n_layers;
  double* act[3] = {w.sx, w.sy, w.sz};
  double tp[3][2];
  // input scaling (NN.cpp:547-553, the exact-zero guard included), in place
  for (int k = lane; k < NG; k += 32) {
    const double d = w.buf[k] - net.mean[k];
    w.buf[k] = (d == 0.0) ? 0.0 : d / net.scale[k];
  }
  __syncwarp();
  const double* in = w.buf;
  for (int l = 0; l < L - 1; l++) {
    const int n_in = net.sizes[l], n_out = net.sizes[l + 1];
    const double* wt = WT + net.w_off[l];
    const int* tf = net.tf + net.node_off[l];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int n = lane + 32 * h;
      tp[l][h] = 0.0;
      if (n < n_out) {
        // four independent partial sums: the dependent FMA chain is what this loop would otherwise wait on
        double z0 = 0.0, z1 = 0.0, z2 = 0.0, z3 = 0.0;
        int k = 0;
        for (; k + 3 < n_in; k += 4) {
          z0 = fma(__ldg(wt + (size_t)k * n_out + n), in[k], z0);
          z1 = fma(__ldg(wt + (size_t)(k + 1) * n_out + n), in[k + 1], z1);
          z2 = fma(__ldg(wt + (size_t)(k + 2) * n_out + n), in[k + 2], z2);
          z3 = fma(__ldg(wt + (size_t)(k + 3) * n_out + n), in[k + 3], z3);
        }
        for (; k < n_in; k++) z0 = fma(__ldg(wt + (size_t)k * n_out + n), in[k], z0);
        double z = (z0 + z1) + (z2 + z3);
        z += __ldg(wt + (size_t)n_in * n_out + n);
        double d;
        act[l][n] = pp_transfer(__ldg(tf + n), z, d, table, exptab);
        tp[l][h] = d;
      }
    }
    __syncwarp();
    in = act[l];
  }
  // output node (linear, NN.cpp:76)
  const int n_h = net.sizes[L - 1];
  const double* wo = net.params + net.w_off[L - 1];
  double y = 0.0;
  for (int n = lane; n < n_h; n += 32) y = fma(__ldg(wo + n), in[n], y);
  y = pp_warp_sum(y) + __ldg(wo + n_h);
  // back to the inputs: delta_{L-1} = w_out T', delta_l = (delta_{l+1} W_l) T'_l, dE/dG = (delta_1 W_0) out_scale / scale
  if (L == 1) {
    for (int k = lane; k < NG; k += 32) w.buf[k] = __ldg(wo + k) * out_scale / net.scale[k];
    __syncwarp();
    return out_scale * y;
  }
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int n = lane + 32 * h;
    if (n < n_h) act[L - 2][n] = __ldg(wo + n) * tp[L - 2][h];
  }
  __syncwarp();
  for (int l = L - 2; l >= 1; l--) {
    const int n_lo = net.sizes[l], n_up = net.sizes[l + 1];
    const double* wl = net.params + net.w_off[l];
    const double* dup = act[l];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int k = lane + 32 * h;
      if (k < n_lo) {
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int n = 0;
        for (; n + 3 < n_up; n += 4) {
          s0 = fma(dup[n], __ldg(wl + (size_t)n * (n_lo + 1) + k), s0);
          s1 = fma(dup[n + 1], __ldg(wl + (size_t)(n + 1) * (n_lo + 1) + k), s1);
          s2 = fma(dup[n + 2], __ldg(wl + (size_t)(n + 2) * (n_lo + 1) + k), s2);
          s3 = fma(dup[n + 3], __ldg(wl + (size_t)(n + 3) * (n_lo + 1) + k), s3);
        }
        for (; n < n_up; n++) s0 = fma(dup[n], __ldg(wl + (size_t)n * (n_lo + 1) + k), s0);
        act[l - 1][k] = ((s0 + s1) + (s2 + s3)) * tp[l - 1][h];
      }
    }
    __syncwarp();
  }
  {
    const int n1 = net.sizes[1];
    const double* w0 = net.params + net.w_off[0];
    const double* d1 = act[0];
    for (int k = lane; k < NG; k += 32) {
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
      int n = 0;
      for (; n + 3 < n1; n += 4) {
        s0 = fma(d1[n], __ldg(w0 + (size_t)n * (NG + 1) + k), s0);
        s1 = fma(d1[n + 1], __ldg(w0 + (size_t)(n + 1) * (NG + 1) + k), s1);
        s2 = fma(d1[n + 2], __ldg(w0 + (size_t)(n + 2) * (NG + 1) + k), s2);
        s3 = fma(d1[n + 3], __ldg(w0 + (size_t)(n + 3) * (NG + 1) + k), s3);
      }
      for (; n < n1; n++) s0 = fma(d1[n], __ldg(w0 + (size_t)n * (NG + 1) + k), s0);
      w.buf[k] = ((s0 + s1) + (s2 + s3)) * out_scale / net.scale[k];
    }
  }
  __syncwarp();
  return out_scale * y;
}

// Fused MD step for one atom per warp: gather, radial factors and their slopes ONCE, descriptors (Calc_G non-periodic
// branch, Structure.cpp:1073-1145), the atom's network (NN.cpp:516-570), force chain rule (Get_Forces, :429-713).
template <int NR_T, int NX, bool BASIS>
__global__ void __launch_bounds__(DESC_WARPS * 32, DESC_BLOCKS_PER_SM)
md_fused_kernel(MdArgs a, MdNets nets)
{
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ double exptab[64];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 64; i += DESC_WARPS * 32) exptab[i] = exp2((double)i * 0.015625);
  __syncthreads();
  const int i = blockIdx.x * DESC_WARPS + warp;
  if (i >= a.nlocal) return;
  const DescSpec* S = a.spec;
  const size_t per_warp = pp_warp_smem_bytes(a.cap, S->NR, S->NG, true);
  WarpNbr w;
  pp_warp_views(w, smem + per_warp * warp, a.cap, S->NR, S->NG, true);
  w.etab = exptab;
  const long long p0 = a.neigh_offsets[i];
  const int nlist = (int)(a.neigh_offsets[i + 1] - p0);
  int n = pp_gather_list(w, a.cap, lane, a.x, a.elem, i, a.neigh + p0, nlist, S->Rcut);
  if (n > a.cap) {
    if (lane == 0) atomicMax(a.overflow, n);
    return;
  }
  if (n <= nets.cap_lo) return;                       // done by an earlier launch with a smaller capacity
  pp_finalize_entries<true>(w, a.cap, lane, S, n);
  pp_descriptor_row<NR_T, NX, BASIS>(w, a.cap, lane, S, S->nG3 > 0);
  __syncwarp();
  if (nets.G_store) {
    double* out = nets.G_store + (size_t)i * S->NG;
    for (int g = lane; g < S->NG; g += 32) out[g] = w.buf[g];
    __syncwarp();
  }
  const int e = a.elem[i];
#ifdef PP_FUSED_SKIP_NET      // timing experiments only: forces are then meaningless
  const double en = 0.0;
#else
  const double en = pp_warp_network(w, lane, S->NG, nets.net[e], nets.WT[e], nets.out_scale, nets.tanh_table, exptab);
#endif
  if (lane == 0) nets.eatom[i] = en;
  if (nets.dEdG_store) {
    double* out = nets.dEdG_store + (size_t)i * S->NG;
    for (int g = lane; g < S->NG; g += 32) out[g] = w.buf[g];
  }
  __syncwarp();
  pp_force_entries<NR_T, NX, BASIS>(w, a.cap, lane, S, S->nG3 > 0);
  // f[j] -= grad_j ; f[i] += sum_j grad_j   (Structure.cpp:700-706)
  double sx = 0.0, sy = 0.0, sz = 0.0;
  for (int p = lane; p < n; p += 32) {
    double gx = w.fx[p], gy = w.fy[p], gz = w.fz[p];
    size_t j = (size_t)w.idx[p];
    atomicAdd(a.f + 3 * j, -gx);
    atomicAdd(a.f + 3 * j + 1, -gy);
    atomicAdd(a.f + 3 * j + 2, -gz);
    sx += gx; sy += gy; sz += gz;
  }
  sx = pp_warp_sum(sx); sy = pp_warp_sum(sy); sz = pp_warp_sum(sz);
  if (lane == 0) {
    atomicAdd(a.f + 3 * (size_t)i, sx);
    atomicAdd(a.f + 3 * (size_t)i + 1, sy);
    atomicAdd(a.f + 3 * (size_t)i + 2, sz);
  }
}

// ---------------------------------------------------------------------------------------------------------
// template dispatch on the register tile of the G4 grid.
//   basis path (power-of-two xi, <= 2 lambdas): NR_T in {1,3,5,8} x NB in {3,5} (+ 5 levels without xi = 8)
//   general path (any xi / lambda):             NR_T in {3,8}     x NA_T = 4, chunked over the grid
template <typename F>
static int dispatch_tile(const DescSpec& hs, bool skip_variant, F&& f)
{
  const int NR = hs.NR;
#define PP_CASE(R_, X_, B_) return f(std::integral_constant<int, R_>(), std::integral_constant<int, X_>(), std::integral_constant<bool, B_>());
  if (hs.basis_nb > 0) {
    const int nr = NR <= 1 ? 1 : NR <= 3 ? 3 : NR <= 5 ? 5 : 8;
    if (hs.basis_nb <= 3) {
      if (nr == 1) PP_CASE(1, 3, true) if (nr == 3) PP_CASE(3, 3, true) if (nr == 5) PP_CASE(5, 3, true) PP_CASE(8, 3, true)
    } else if (skip_variant && ((hs.basis_used[0] | hs.basis_used[1]) & 8) == 0) {      // xi = 8 unused, e.g. xi in {1, 2, 4, 16}
      if (nr == 1) PP_CASE(1, 6, true) if (nr == 3) PP_CASE(3, 6, true) if (nr == 5) PP_CASE(5, 6, true) PP_CASE(8, 6, true)
    } else {
      if (nr == 1) PP_CASE(1, 5, true) if (nr == 3) PP_CASE(3, 5, true) if (nr == 5) PP_CASE(5, 5, true) PP_CASE(8, 5, true)
    }
  }
  if (NR <= 3) PP_CASE(3, 4, false)
  PP_CASE(8, 4, false)
#undef PP_CASE
}

static int max_cap_for(int NR, int NG, bool force)
{
  int cap = 32;
  while (pp_warp_smem_bytes(cap * 2, NR, NG, force) * DESC_WARPS <= 110 * 1024) cap *= 2;      // two blocks per SM at least
  // many distinct radial keys (every row its own after the -optimize search perturbed them): one block per SM rather
  // than refusing shells of more than 32 neighbours
  while (cap < 128 && pp_warp_smem_bytes(cap * 2, NR, NG, force) * DESC_WARPS <= 220 * 1024) cap *= 2;
  return cap;
}

template <typename K>
static int set_smem(K kernel, size_t bytes)
{
  PP_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

int ppk_desc_periodic(const DescSpec& hs, PeriodicArgs a, int* cap_io, cudaStream_t st, long long* launches)
{
  if (a.n_atoms == 0) return 0;
  int cap = *cap_io;
  const int cap_max = max_cap_for(hs.NR, hs.NG, false);
  for (;;) {
    if (cap > cap_max) {
      pp_set_error("an atom has more than %d neighbour images within Rcut: does not fit shared memory", cap_max);
      return 1;
    }
    a.cap = cap;
    PP_CUDA(cudaMemsetAsync(a.overflow, 0, sizeof(int), st));
    size_t bytes = pp_warp_smem_bytes(cap, hs.NR, hs.NG, false) * DESC_WARPS;
    unsigned grid = (unsigned)((a.n_atoms + DESC_WARPS - 1) / DESC_WARPS);
    // no xi = 8 variant here: measured 15 % slower in this kernel (29 -> 34 ms for C2's 10 000 structures)
    int rc = dispatch_tile(hs, false, [&](auto R_, auto X_, auto B_) {
      auto k = desc_periodic_kernel<decltype(R_)::value, decltype(X_)::value, decltype(B_)::value>;
      if (set_smem(k, bytes)) return 1;
      k<<<grid, DESC_WARPS * 32, bytes, st>>>(a);
      return 0;
    });
    if (rc) return rc;
    (*launches)++;
    PP_CUDA(cudaGetLastError());
    int over = 0;
    PP_CUDA(cudaMemcpyAsync(&over, a.overflow, sizeof(int), cudaMemcpyDeviceToHost, st));
    PP_CUDA(cudaStreamSynchronize(st));
    if (over == 0) break;
    while (cap < over) cap *= 2;
  }
  *cap_io = cap;
  return 0;
}

int ppk_neighbors_periodic(const DescSpec& hs, PeriodicArgs a, int s, int natom, int pass, int* counts, const int* offsets,
                           int* out_j, int* out_t, int cap, cudaStream_t st, long long* launches)
{
  a.cap = cap;
  size_t bytes = pp_warp_smem_bytes(cap, hs.NR, hs.NG, false) * DESC_WARPS;
  if (set_smem(neighbors_periodic_kernel, bytes)) return 1;
  unsigned grid = (unsigned)((natom + DESC_WARPS - 1) / DESC_WARPS);
  neighbors_periodic_kernel<<<grid, DESC_WARPS * 32, bytes, st>>>(a, s, pass, counts, offsets, out_j, out_t);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

int ppk_max_cap(const DescSpec& hs, bool force) { return max_cap_for(hs.NR, hs.NG, force); }

int ppk_md_desc(const DescSpec& hs, MdArgs a, int* cap_io, cudaStream_t st, long long* launches)
{
  if (a.nlocal == 0) return 0;
  int cap = *cap_io;
  const int cap_max = max_cap_for(hs.NR, hs.NG, false);
  for (;;) {
    if (cap > cap_max) {
      pp_set_error("an atom has more than %d neighbours within Rcut: does not fit shared memory", cap_max);
      return 1;
    }
    a.cap = cap;
    PP_CUDA(cudaMemsetAsync(a.overflow, 0, sizeof(int), st));
    size_t bytes = pp_warp_smem_bytes(cap, hs.NR, hs.NG, false) * DESC_WARPS;
    unsigned grid = (unsigned)((a.nlocal + DESC_WARPS - 1) / DESC_WARPS);
    int rc = dispatch_tile(hs, true, [&](auto R_, auto X_, auto B_) {
      auto k = md_desc_kernel<decltype(R_)::value, decltype(X_)::value, decltype(B_)::value>;
      if (set_smem(k, bytes)) return 1;
      k<<<grid, DESC_WARPS * 32, bytes, st>>>(a);
      return 0;
    });
    if (rc) return rc;
    (*launches)++;
    PP_CUDA(cudaGetLastError());
    int over = 0;
    PP_CUDA(cudaMemcpyAsync(&over, a.overflow, sizeof(int), cudaMemcpyDeviceToHost, st));
    PP_CUDA(cudaStreamSynchronize(st));
    if (over == 0) break;
    while (cap < over) cap *= 2;
  }
  *cap_io = cap;
  return 0;
}

// The force kernel runs after md_desc on the same list, so `cap` is already known to fit.
int ppk_md_force(const DescSpec& hs, MdArgs a, int cap, cudaStream_t st, long long* launches)
{
  if (a.nlocal == 0) return 0;
  const int cap_max = max_cap_for(hs.NR, hs.NG, true);
  if (cap > cap_max) {
    pp_set_error("an atom has more than %d neighbours within Rcut: force kernel does not fit shared memory", cap_max);
    return 1;
  }
  a.cap = cap;
  size_t bytes = pp_warp_smem_bytes(cap, hs.NR, hs.NG, true) * DESC_WARPS;
  unsigned grid = (unsigned)((a.nlocal + DESC_WARPS - 1) / DESC_WARPS);
  int rc = dispatch_tile(hs, true, [&](auto R_, auto X_, auto B_) {
    auto k = md_force_kernel<decltype(R_)::value, decltype(X_)::value, decltype(B_)::value>;
    if (set_smem(k, bytes)) return 1;
    k<<<grid, DESC_WARPS * 32, bytes, st>>>(a);
    return 0;
  });
  if (rc) return rc;
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// The fused MD step.  Networks of up to three hidden layers, each up to 64 wide (a lane owns two neurons per layer).
bool ppk_md_fused_ok(const NetDev& net)
{
  if (net.n_layers > 4) return false;
  for (int l = 1; l < net.n_layers; l++)
    if (net.sizes[l] > 64) return false;
  return true;
}

int ppk_md_fused(const DescSpec& hs, MdArgs a, const MdNets& nets_in, int* cap_io, cudaStream_t st, long long* launches)
{
  if (a.nlocal == 0) return 0;
  MdNets nets = nets_in;
  int cap = std::max(*cap_io, 64);                     // the network's activations use the 64-entry staging arrays
  const int cap_max = max_cap_for(hs.NR, hs.NG, true);
  nets.cap_lo = -1;
  for (;;) {
    if (cap > cap_max) {
      pp_set_error("an atom has more than %d neighbours within Rcut: does not fit shared memory", cap_max);
      return 1;
    }
    a.cap = cap;
    PP_CUDA(cudaMemsetAsync(a.overflow, 0, sizeof(int), st));
    size_t bytes = pp_warp_smem_bytes(cap, hs.NR, hs.NG, true) * DESC_WARPS;
    unsigned grid = (unsigned)((a.nlocal + DESC_WARPS - 1) / DESC_WARPS);
    int rc = dispatch_tile(hs, true, [&](auto R_, auto X_, auto B_) {
      auto k = md_fused_kernel<decltype(R_)::value, decltype(X_)::value, decltype(B_)::value>;
      if (set_smem(k, bytes)) return 1;
      k<<<grid, DESC_WARPS * 32, bytes, st>>>(a, nets);
      return 0;
    });
    if (rc) return rc;
    (*launches)++;
    PP_CUDA(cudaGetLastError());
    int over = 0;
    PP_CUDA(cudaMemcpyAsync(&over, a.overflow, sizeof(int), cudaMemcpyDeviceToHost, st));
    PP_CUDA(cudaStreamSynchronize(st));
    if (over == 0) break;
    nets.cap_lo = cap;                                 // atoms that fitted are done: the next launch takes the others
    while (cap < over) cap *= 2;
  }
  *cap_io = cap;
  return 0;
}
