// Reverse communication of ghost forces for the spatially partitioned MD harness (sm_100a).
//
// PairNN::compute leaves forces on ghost atoms (Structure::Get_Forces writes f[j] for every neighbour j, pair_nn.cpp:182)
// and relies on the host engine's reverse communication to return them to the atoms that own them.  When this library
// drives a partitioned box itself (the C5 benchmark: one box cut into slabs, one per GPU), that step is these kernels:
//   pack     sendbuf[k] = f[send_idx[k]]                      (ghosts owned by another rank, grouped by owner)
//   add      f[dst] += sum of a SEGMENT of values, in a fixed order (one thread per destination atom: an atom can own
//            several ghosts of the same sender -- its y/z periodic images -- and the fixed order keeps the forces
//            bit-reproducible; no atomics)
// Ghosts owned by the rank itself (periodic images across the faces the slab does not share) use the same add kernel
// with the values gathered from f.  The transfers are NCCL send/recv on the library's stream (pp_comm.cu).
#include "pp_common.cuh"
#include "pp_kernels.h"

__global__ void __launch_bounds__(256) pack_forces_kernel(const double* __restrict__ f, const int* __restrict__ idx, int n, double* __restrict__ out)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const size_t a = (size_t)idx[k];
  out[3 * (size_t)k + 0] = f[3 * a + 0];
  out[3 * (size_t)k + 1] = f[3 * a + 1];
  out[3 * (size_t)k + 2] = f[3 * a + 2];
}

// segment s: destination atom seg_dst[s] receives the values base[3 idx[k] ..], k in [seg_off[s], seg_off[s+1]), added in
// that order.  base is f itself (ghosts this rank owns: idx = ghost index) or a receive buffer (idx = position in it).
__global__ void __launch_bounds__(256)
segment_add_kernel(double* __restrict__ f, const int* __restrict__ seg_off, const int* __restrict__ seg_dst, int n_seg,
                   const double* __restrict__ base, const int* __restrict__ idx)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_seg) return;
  double ax = 0.0, ay = 0.0, az = 0.0;
  for (int k = seg_off[s]; k < seg_off[s + 1]; k++) {
    const double* v = base + 3 * (size_t)idx[k];
    ax += v[0]; ay += v[1]; az += v[2];
  }
  const size_t d = (size_t)seg_dst[s];
  f[3 * d + 0] += ax; f[3 * d + 1] += ay; f[3 * d + 2] += az;
}

int ppk_pack_forces(const double* f, const int* idx, int n, double* out, cudaStream_t st, long long* launches)
{
  if (n == 0) return 0;
  pack_forces_kernel<<<(n + 255) / 256, 256, 0, st>>>(f, idx, n, out);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

int ppk_segment_add(double* f, const int* seg_off, const int* seg_dst, int n_seg, const double* base, const int* idx, cudaStream_t st,
                    long long* launches)
{
  if (n_seg == 0) return 0;
  segment_add_kernel<<<(n_seg + 255) / 256, 256, 0, st>>>(f, seg_off, seg_dst, n_seg, base, idx);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}

__global__ void add_scalar_kernel(double* v, double a) { *v += a; }
int ppk_add_scalar(double* v, double a, cudaStream_t st, long long* launches)
{
  add_scalar_kernel<<<1, 1, 0, st>>>(v, a);
  (*launches)++;
  PP_CUDA(cudaGetLastError());
  return 0;
}
