// Parallel (Parallel.h:93-214) on NCCL: the only collective on the hot path is the per-epoch sum of
// [gradient, SSE, vSSE, nval] (Optimizer.cpp:887-888, 969-970) plus the two preconditioning sums
// (Potential.cpp:419-420, 442).  NCCL is loaded at run time (dlopen) so the library has no link-time
// dependency on a particular NCCL build: the copy already loaded into the process (e.g. torch's) is reused,
// else $PP_NCCL_LIB, else libnccl.so.2 from the loader path.
#pragma once
#include <cuda_runtime.h>

#define PP_PEER_MAX 8
// Receive slots of the one-shot sum over peer memory (pp_comm.cu): every rank's region holds, per epoch parity, one
// slot of `cap` doubles per source rank, and one arrival flag per source rank.
struct PPPeerTable {
  double* slots[PP_PEER_MAX];                 // base of rank r's slot area, as mapped into THIS process
  unsigned long long* flags[PP_PEER_MAX];     // base of rank r's flag array
};

struct PPComm {
  int nranks = 1, rank = 0;
  void* comm = nullptr;
  double* d_stage = nullptr;
  size_t stage_cap = 0;
  // one-shot sum over peer memory (NVLink): state
  int peer_state = 0;                         // 0 not tried, 1 ready, -1 unavailable (NCCL is used)
  size_t peer_cap = 0;                        // doubles per slot
  void* peer_region = nullptr;                // this rank's region (flags, then slots)
  void* peer_mapped[PP_PEER_MAX] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  PPPeerTable peers;
  unsigned long long peer_epoch = 0;
  unsigned* d_ticket = nullptr;
  int* d_peer_err = nullptr;

  static int unique_id(void* out128);
  int init(int nranks_, int rank_, const void* id128);
  int allreduce_device(double* buf_dev, int n, cudaStream_t st);   // in place, sum; no-op for one rank
  int allreduce_host(double* buf_host, int n, cudaStream_t st);    // in place, sum; no-op for one rank
  // out[0..ncols) = sum over ranks of (column sums of this rank's partial rows, then tail[ncols_partial..ncols) taken
  // from out itself), identical on every rank.  One rank: the local reduction only.
  int reduce_sum(const double* partial, int n_cta, int stride, int ncols_partial, int ncols, double* out, cudaStream_t st,
                 long long* launches);
  int peer_setup(size_t need, cudaStream_t st);
  // grouped point-to-point exchange of doubles with n_peers ranks (ncclGroupStart / ncclSend / ncclRecv / ncclGroupEnd)
  int sendrecv(int n_peers, const int* peer, double* const* sendbuf, const size_t* send_n, double* const* recvbuf, const size_t* recv_n,
               cudaStream_t st);
  const char* path() const { return nranks == 1 ? "single rank" : (peer_state == 1 ? "peer memory" : "nccl"); }
  void destroy();
};
