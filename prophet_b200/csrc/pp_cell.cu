// GPU cell list + full neighbour list for the MD harness ("K1 in LAMMPS mode").
//
// In LAMMPS the neighbour list is an input to pair_style nn (built by LAMMPS with cutoff Rcut + skin and the test
// rsq < cutneigh^2); PROPhet only filters it with R > Rcut.  A harness that drives the pair style without LAMMPS
// (the 1M-atom C5 case) has to build that list itself.  This file does it on the GPU:
//   1. bin all atoms (locals + ghosts) into cells of edge >= cutneigh          (cell_index_kernel, counting sort)
//   2. sort the atoms inside each cell by index so that the result is deterministic (cell_sort_kernel)
//   3. one warp per local atom walks its 27 cells, counts and then writes the neighbours (neigh_kernel, two passes)
// The distance test uses the same non-contracted FP64 arithmetic as a CPU brute-force check, so the neighbour SETS are
// bit-identical to an O(N^2) reference with the same comparison (tests/test_gpu_md.py).
#include "pp_common.cuh"
#include "pp_kernels.h"

__global__ void cell_index_kernel(const double* __restrict__ x, int n, CellGrid g, int* __restrict__ cell_of, int* __restrict__ count)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int cx = (int)((x[3 * (size_t)i] - g.lo[0]) * g.inv[0]);
  int cy = (int)((x[3 * (size_t)i + 1] - g.lo[1]) * g.inv[1]);
  int cz = (int)((x[3 * (size_t)i + 2] - g.lo[2]) * g.inv[2]);
  cx = min(max(cx, 0), g.n[0] - 1);
  cy = min(max(cy, 0), g.n[1] - 1);
  cz = min(max(cz, 0), g.n[2] - 1);
  int c = (cx * g.n[1] + cy) * g.n[2] + cz;
  cell_of[i] = c;
  atomicAdd(count + c, 1);
}

// exclusive scan of int counts -> int offsets, three small kernels (block scan, scan of block sums, add)
#define SCAN_THREADS 1024
__global__ void __launch_bounds__(SCAN_THREADS) scan_block_kernel(const int* __restrict__ in, long long n, long long* __restrict__ out,
                                                                  long long* __restrict__ block_sums)
{
  __shared__ long long warp_tot[32];
  const long long i = (long long)blockIdx.x * SCAN_THREADS + threadIdx.x;
  long long v = i < n ? in[i] : 0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    long long w = warp_tot[lane];
    long long wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      long long t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    warp_tot[lane] = wi - w;     // exclusive
    if (lane == 31) block_sums[blockIdx.x] = wi;
  }
  __syncthreads();
  if (i < n) out[i] = warp_tot[warp] + incl - v;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_sums_kernel(long long* __restrict__ block_sums, int nblocks, long long* __restrict__ total)
{
  // single block; nblocks <= a few thousand: serial chunks of SCAN_THREADS with a running carry
  __shared__ long long warp_tot[32];
  __shared__ long long carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < nblocks; base += SCAN_THREADS) {
    const int i = base + threadIdx.x;
    long long v = i < nblocks ? block_sums[i] : 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    long long incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      long long t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      long long w = warp_tot[lane], wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        long long t = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += t;
      }
      warp_tot[lane] = wi - w;
    }
    __syncthreads();
    const long long carry = carry_s;
    if (i < nblocks) block_sums[i] = carry + warp_tot[warp] + incl - v;
    __syncthreads();
    if (threadIdx.x == SCAN_THREADS - 1) carry_s = carry + warp_tot[warp] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry_s;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_add_kernel(long long* __restrict__ out, long long n, const long long* __restrict__ block_sums,
                                                                const long long* __restrict__ total)
{
  const long long i = (long long)blockIdx.x * SCAN_THREADS + threadIdx.x;
  if (i < n) out[i] += block_sums[blockIdx.x];
  if (i == n) out[n] = *total;      // out has n+1 entries
}

static int exclusive_scan(const int* in, long long n, long long* out /*[n+1]*/, long long* block_sums, long long* total_dev,
                          cudaStream_t st, long long* launches)
{
  const int nb = (int)((n + 1 + SCAN_THREADS - 1) / SCAN_THREADS);
  scan_block_kernel<<<nb, SCAN_THREADS, 0, st>>>(in, n, out, block_sums);
  scan_sums_kernel<<<1, SCAN_THREADS, 0, st>>>(block_sums, nb, total_dev);
  scan_add_kernel<<<nb, SCAN_THREADS, 0, st>>>(out, n, block_sums, total_dev);
  *launches += 3;
  PP_CUDA(cudaGetLastError());
  return 0;
}

__global__ void cell_fill_kernel(int n, const int* __restrict__ cell_of, const long long* __restrict__ cell_start, int* __restrict__ cursor,
                                 int* __restrict__ sorted)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int c = cell_of[i];
  int p = atomicAdd(cursor + c, 1);
  sorted[cell_start[c] + p] = i;
}

// ascending atom index inside every cell (insertion sort, cells hold a few tens of atoms)
__global__ void cell_sort_kernel(int ncells, const long long* __restrict__ cell_start, int* __restrict__ sorted)
{
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncells) return;
  const long long b = cell_start[c], e = cell_start[c + 1];
  for (long long i = b + 1; i < e; i++) {
    int v = sorted[i];
    long long j = i - 1;
    while (j >= b && sorted[j] > v) { sorted[j + 1] = sorted[j]; j--; }
    sorted[j + 1] = v;
  }
}

// One warp per local atom.  pass 0: counts[i]; pass 1: write neighbours at offsets[i] (cell order, ascending index in a cell).
__global__ void __launch_bounds__(256) neigh_kernel(const double* __restrict__ x, int nlocal, CellGrid g, const int* __restrict__ cell_of,
                                                    const long long* __restrict__ cell_start, const int* __restrict__ sorted, double cutsq,
                                                    int pass, int* __restrict__ counts, const long long* __restrict__ offsets, int* __restrict__ out)
{
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * 8 + warp;
  if (i >= nlocal) return;
  const double xi = x[3 * (size_t)i], yi = x[3 * (size_t)i + 1], zi = x[3 * (size_t)i + 2];
  const int c = cell_of[i];
  const int cz = c % g.n[2], cy = (c / g.n[2]) % g.n[1], cx = c / (g.n[2] * g.n[1]);
  int count = 0;
  const long long base = pass ? offsets[i] : 0;
  for (int dx = -1; dx <= 1; dx++) {
    const int ax = cx + dx;
    if (ax < 0 || ax >= g.n[0]) continue;
    for (int dy = -1; dy <= 1; dy++) {
      const int ay = cy + dy;
      if (ay < 0 || ay >= g.n[1]) continue;
      for (int dz = -1; dz <= 1; dz++) {
        const int az = cz + dz;
        if (az < 0 || az >= g.n[2]) continue;
        const int cc = (ax * g.n[1] + ay) * g.n[2] + az;
        const long long b = cell_start[cc], e = cell_start[cc + 1];
        for (long long p0 = b; p0 < e; p0 += 32) {
          const long long p = p0 + lane;
          bool keep = false;
          int j = 0;
          if (p < e) {
            j = sorted[p];
            const double ddx = __dsub_rn(xi, x[3 * (size_t)j]);
            const double ddy = __dsub_rn(yi, x[3 * (size_t)j + 1]);
            const double ddz = __dsub_rn(zi, x[3 * (size_t)j + 2]);
            const double rsq = __dadd_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)), __dmul_rn(ddz, ddz));
            keep = (j != i) && (rsq < cutsq);
          }
          const unsigned m = __ballot_sync(0xffffffffu, keep);
          if (pass && keep) out[base + count + __popc(m & ((1u << lane) - 1u))] = j;
          count += __popc(m);
        }
      }
    }
  }
  if (pass == 0 && lane == 0) counts[i] = count;
}

int ppk_build_cell_neighbors(const double* x_dev, int nlocal, int ntotal, const CellGrid& g, double cutneigh, CellWork& w,
                             cudaStream_t st, long long* launches)
{
  const int ncells = g.n[0] * g.n[1] * g.n[2];
  PP_CUDA(cudaMemsetAsync(w.cell_count, 0, sizeof(int) * (size_t)ncells, st));
  cell_index_kernel<<<(ntotal + 255) / 256, 256, 0, st>>>(x_dev, ntotal, g, w.cell_of, w.cell_count);
  (*launches)++;
  if (exclusive_scan(w.cell_count, ncells, w.cell_start, w.block_sums, w.total, st, launches)) return 1;
  PP_CUDA(cudaMemsetAsync(w.cell_count, 0, sizeof(int) * (size_t)ncells, st));
  cell_fill_kernel<<<(ntotal + 255) / 256, 256, 0, st>>>(ntotal, w.cell_of, w.cell_start, w.cell_count, w.sorted);
  cell_sort_kernel<<<(ncells + 127) / 128, 128, 0, st>>>(ncells, w.cell_start, w.sorted);
  *launches += 2;
  const double cutsq = cutneigh * cutneigh;
  neigh_kernel<<<(nlocal + 7) / 8, 256, 0, st>>>(x_dev, nlocal, g, w.cell_of, w.cell_start, w.sorted, cutsq, 0, w.counts, nullptr, nullptr);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

int ppk_scan_counts(const int* counts, long long n, long long* offsets, CellWork& w, cudaStream_t st, long long* launches)
{
  return exclusive_scan(counts, n, offsets, w.block_sums, w.total, st, launches);
}

int ppk_fill_cell_neighbors(const double* x_dev, int nlocal, const CellGrid& g, double cutneigh, CellWork& w, const long long* offsets,
                            int* neigh_out, cudaStream_t st, long long* launches)
{
  const double cutsq = cutneigh * cutneigh;
  neigh_kernel<<<(nlocal + 7) / 8, 256, 0, st>>>(x_dev, nlocal, g, w.cell_of, w.cell_start, w.sorted, cutsq, 1, w.counts, offsets, neigh_out);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

// sum of n doubles -> out[0] (+ add), fixed order: per-block partials then one thread
__global__ void __launch_bounds__(256) block_sum_kernel(const double* __restrict__ v, long long n, double* __restrict__ partial)
{
  __shared__ double ws[8];
  const long long per = (n + gridDim.x - 1) / gridDim.x;
  const long long b = (long long)blockIdx.x * per, e = b + per < n ? b + per : n;
  double s = 0.0;
  for (long long i = b + threadIdx.x; i < e; i += blockDim.x) s += v[i];
  s = pp_warp_sum(s);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < 8; k++) t += ws[k];
    partial[blockIdx.x] = t;
  }
}
__global__ void final_sum_kernel(const double* __restrict__ partial, int nb, double add, double* __restrict__ out)
{
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < nb; k++) t += partial[k];
    out[0] = t + add;
  }
}

int ppk_sum(const double* v, long long n, double add, double* partial /*[128]*/, double* out, cudaStream_t st, long long* launches)
{
  block_sum_kernel<<<128, 256, 0, st>>>(v, n, partial);
  final_sum_kernel<<<1, 32, 0, st>>>(partial, 128, add, out);
  *launches += 2;
  PP_CUDA(cudaGetLastError());
  return 0;
}
