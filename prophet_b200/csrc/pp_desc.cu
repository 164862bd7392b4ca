// Descriptor / force kernels and their launchers (sm_100a).  See pp_desc.cuh for the algorithm.
#include "pp_desc.cuh"
#include "pp_kernels.h"

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
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long rel = (long long)blockIdx.x * DESC_WARPS + warp;
  if (rel >= a.n_atoms) return;
  const long long atom = a.atom_begin + rel;
  const DescSpec* S = a.spec;
  const size_t per_warp = pp_warp_smem_bytes(a.cap, S->NR, S->NG, false);
  WarpNbr w;
  pp_warp_views(w, smem + per_warp * warp, a.cap, S->NR, S->NG, false);
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
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * DESC_WARPS + warp;
  if (i >= a.nlocal) return;
  const DescSpec* S = a.spec;
  const size_t per_warp = pp_warp_smem_bytes(a.cap, S->NR, S->NG, false);
  WarpNbr w;
  pp_warp_views(w, smem + per_warp * warp, a.cap, S->NR, S->NG, false);
  const long long p0 = a.neigh_offsets[i];
  const int nlist = (int)(a.neigh_offsets[i + 1] - p0);
  int n = pp_gather_list(w, a.cap, lane, a.x, a.elem, i, a.neigh + p0, nlist, S->Rcut);
  if (n > a.cap) {
    if (lane == 0) atomicMax(a.overflow, n);
    return;
  }
  pp_finalize_entries<false>(w, a.cap, lane, S, n);
  pp_descriptor_row<NR_T, NX, BASIS>(w, a.cap, lane, S, S->nG3 > 0);
  double* out = a.G + (size_t)i * S->NG;
  for (int g = lane; g < S->NG; g += 32) out[g] = w.buf[g];
}

// MD forces: Structure::Get_Forces non-periodic branch (Structure.cpp:429-713) with dE/dG given.
template <int NR_T, int NX, bool BASIS>
__global__ void __launch_bounds__(DESC_WARPS * 32, DESC_BLOCKS_PER_SM)
md_force_kernel(MdArgs a)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * DESC_WARPS + warp;
  if (i >= a.nlocal) return;
  const DescSpec* S = a.spec;
  const size_t per_warp = pp_warp_smem_bytes(a.cap, S->NR, S->NG, true);
  WarpNbr w;
  pp_warp_views(w, smem + per_warp * warp, a.cap, S->NR, S->NG, true);
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
