// Parallel (Parallel.h:93-214) on NCCL: the only collective on the hot path is the per-epoch sum of
// [gradient, SSE, vSSE, nval] (Optimizer.cpp:887-888, 969-970) plus the two preconditioning sums
// (Potential.cpp:419-420, 442).  NCCL is loaded at run time (dlopen) so the library has no link-time
// dependency on a particular NCCL build: the copy already loaded into the process (e.g. torch's) is reused,
// else $PP_NCCL_LIB, else libnccl.so.2 from the loader path.
#pragma once
#include <cuda_runtime.h>

struct PPComm {
  int nranks = 1, rank = 0;
  void* comm = nullptr;
  double* d_stage = nullptr;
  size_t stage_cap = 0;

  static int unique_id(void* out128);
  int init(int nranks_, int rank_, const void* id128);
  int allreduce_device(double* buf_dev, int n, cudaStream_t st);   // in place, sum; no-op for one rank
  int allreduce_host(double* buf_host, int n, cudaStream_t st);    // in place, sum; no-op for one rank
  void destroy();
};
