#include "pp_comm.h"
#include "pp_common.cuh"
#include "pp_kernels.h"
#include <dlfcn.h>
#include <algorithm>
#include <vector>

namespace {
struct NcclId { char internal[128]; };
typedef int (*fn_get_id)(NcclId*);
typedef int (*fn_init_rank)(void**, int, NcclId, int);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_allgather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*fn_send)(const void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_recv)(void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_group)(void);
typedef int (*fn_destroy)(void*);
typedef const char* (*fn_errstr)(int);

struct NcclApi {
  void* lib = nullptr;
  fn_get_id get_id = nullptr;
  fn_init_rank init_rank = nullptr;
  fn_allreduce allreduce = nullptr;
  fn_allgather allgather = nullptr;
  fn_send send = nullptr;
  fn_recv recv = nullptr;
  fn_group group_start = nullptr, group_end = nullptr;
  fn_destroy destroy = nullptr;
  fn_errstr errstr = nullptr;
} g_nccl;

int load_nccl()
{
  if (g_nccl.lib) return 0;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h) {
    const char* env = getenv("PP_NCCL_LIB");
    if (env && *env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
  }
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  PP_REQUIRE(h, "NCCL not found: set PP_NCCL_LIB to libnccl.so.2 (%s)", dlerror());
  g_nccl.get_id = (fn_get_id)dlsym(h, "ncclGetUniqueId");
  g_nccl.init_rank = (fn_init_rank)dlsym(h, "ncclCommInitRank");
  g_nccl.allreduce = (fn_allreduce)dlsym(h, "ncclAllReduce");
  g_nccl.allgather = (fn_allgather)dlsym(h, "ncclAllGather");
  g_nccl.send = (fn_send)dlsym(h, "ncclSend");
  g_nccl.recv = (fn_recv)dlsym(h, "ncclRecv");
  g_nccl.group_start = (fn_group)dlsym(h, "ncclGroupStart");
  g_nccl.group_end = (fn_group)dlsym(h, "ncclGroupEnd");
  g_nccl.destroy = (fn_destroy)dlsym(h, "ncclCommDestroy");
  g_nccl.errstr = (fn_errstr)dlsym(h, "ncclGetErrorString");
  PP_REQUIRE(g_nccl.get_id && g_nccl.init_rank && g_nccl.allreduce && g_nccl.destroy, "NCCL library lacks required symbols");
  g_nccl.lib = h;
  return 0;
}
const char* nccl_err(int rc) { return g_nccl.errstr ? g_nccl.errstr(rc) : "unknown"; }
}  // namespace

int PPComm::unique_id(void* out128)
{
  PP_REQUIRE(out128, "pp_comm_unique_id: NULL buffer");
  if (load_nccl()) return 1;
  NcclId id;
  int rc = g_nccl.get_id(&id);
  PP_REQUIRE(rc == 0, "ncclGetUniqueId failed: %s", nccl_err(rc));
  memcpy(out128, &id, sizeof(id));
  return 0;
}

int PPComm::init(int nranks_, int rank_, const void* id128)
{
  PP_REQUIRE(nranks_ >= 1 && rank_ >= 0 && rank_ < nranks_, "pp_comm_init: bad rank %d of %d", rank_, nranks_);
  destroy();
  nranks = nranks_;
  rank = rank_;
  if (nranks == 1) return 0;
  PP_REQUIRE(id128, "pp_comm_init: NULL unique id");
  if (load_nccl()) return 1;
  NcclId id;
  memcpy(&id, id128, sizeof(id));
  int rc = g_nccl.init_rank(&comm, nranks, id, rank);
  PP_REQUIRE(rc == 0, "ncclCommInitRank failed: %s", nccl_err(rc));
  return 0;
}

int PPComm::allreduce_device(double* buf_dev, int n, cudaStream_t st)
{
  if (nranks == 1 || n == 0) return 0;
  PP_REQUIRE(comm, "communicator not initialised");
  int rc = g_nccl.allreduce(buf_dev, buf_dev, (size_t)n, /*ncclDouble*/ 8, /*ncclSum*/ 0, comm, st);
  PP_REQUIRE(rc == 0, "ncclAllReduce failed: %s", nccl_err(rc));
  return 0;
}

int PPComm::allreduce_host(double* buf_host, int n, cudaStream_t st)
{
  if (nranks == 1 || n == 0) return 0;
  if ((size_t)n > stage_cap) {
    if (d_stage) cudaFree(d_stage);
    PP_CUDA(cudaMalloc(&d_stage, (size_t)n * sizeof(double)));
    stage_cap = n;
  }
  PP_CUDA(cudaMemcpyAsync(d_stage, buf_host, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, st));
  if (allreduce_device(d_stage, n, st)) return 1;
  PP_CUDA(cudaMemcpyAsync(buf_host, d_stage, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
  PP_CUDA(cudaStreamSynchronize(st));
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// One-shot sum over peer memory.  The per-epoch payload is (P + 3) doubles (30 - 350 KB): far below the bandwidth
// regime, so what an epoch pays for the reduction is latency -- a library all-reduce costs 17 us on two GPUs and 41 us
// on eight, next to a 90 - 700 us epoch.  Here every rank's reduction kernel stores its column sums straight into a
// receive slot in EACH peer's memory (plain stores through the NVLink mapping of the peer's allocation), the last block
// raises this rank's arrival flag on every peer, and a second kernel on each rank waits for the N flags and adds the N
// slots in rank order -- so every rank holds the bit-identical vector (the host optimizer runs redundantly on it,
// replacing Optimizer.cpp:887-888 and the broadcast of :957).  Slots alternate with the epoch's parity: a peer can run
// at most one epoch ahead of a rank's summing kernel (its next push needs that rank's next flag).
#define PP_RED_GROUPS 32      // row groups of a reduction block (x 32 columns = 1024 threads)
__global__ void __launch_bounds__(32 * PP_RED_GROUPS)
reduce_push_kernel(const double* __restrict__ partial, int n_cta, int stride, int ncols_partial, int ncols, const double* __restrict__ tail,
                   PPPeerTable peers, int nranks, int rank, size_t cap, unsigned long long epoch, unsigned* __restrict__ ticket)
{
  __shared__ double part[PP_RED_GROUPS][33];
  __shared__ bool last;
  const int col = threadIdx.x & 31, grp = threadIdx.x >> 5;
  const int p = blockIdx.x * 32 + col;
  // fixed order: group g adds rows g, g + 32, g + 64, ... (two chains), then the 32 group sums are added in order
  part[grp][col] = (p < ncols_partial) ? pp_group_column_sum(partial + p, grp, n_cta, stride) : 0.0;
  __syncthreads();
  if (grp == 0 && p < ncols) {
    double s = part[0][col];
#pragma unroll
    for (int j = 1; j < PP_RED_GROUPS; j++) s += part[j][col];
    if (p >= ncols_partial) s = tail[p];
    const size_t off = ((size_t)(epoch & 1) * nranks + rank) * cap + p;
    for (int r = 0; r < nranks; r++) peers.slots[r][off] = s;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (last && threadIdx.x < nranks) {
    __threadfence_system();
    *reinterpret_cast<volatile unsigned long long*>(peers.flags[threadIdx.x] + rank) = epoch;
    if (threadIdx.x == 0) *ticket = 0;
  }
}

__global__ void __launch_bounds__(256)
wait_sum_kernel(const double* __restrict__ slots, const unsigned long long* __restrict__ flags, int nranks, size_t cap, int ncols,
                unsigned long long epoch, double* __restrict__ out, int* __restrict__ err)
{
  __shared__ int bad;
  if (threadIdx.x == 0) bad = 0;
  __syncthreads();
  if (threadIdx.x < nranks) {
    const volatile unsigned long long* f = flags + threadIdx.x;
    const long long t0 = clock64();
    while (*f < epoch) {
      if (clock64() - t0 > 8000000000LL) { bad = 1; break; }      // ~4 s: a peer died; do not hang the GPU
      __nanosleep(100);
    }
  }
  __syncthreads();
  __threadfence_system();
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < ncols) {
    const double* base = slots + (size_t)(epoch & 1) * nranks * cap + p;
    double s = 0.0;
    for (int r = 0; r < nranks; r++) s += __ldcg(base + (size_t)r * cap);
    out[p] = s;
  }
  if (bad && threadIdx.x == 0) *err = 1;
}

int PPComm::peer_setup(size_t need, cudaStream_t st)
{
  if (peer_state != 0) return 0;
  peer_state = -1;
  if (nranks < 2 || nranks > PP_PEER_MAX || getenv("PP_COMM_NCCL_ONLY") != nullptr || !g_nccl.allgather) return 0;
  // every step below is taken by all ranks in the same order; a rank that fails still joins the collectives and the
  // final vote, so that all ranks agree on the path
  int ok = 1;
  peer_cap = (std::max<size_t>(need, 65536) + 31) & ~(size_t)31;
  const size_t flag_bytes = 256;
  const size_t bytes = flag_bytes + (size_t)2 * nranks * peer_cap * sizeof(double);
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (cudaMalloc(&peer_region, bytes) != cudaSuccess) { ok = 0; peer_region = nullptr; }
  if (ok && cudaMemset(peer_region, 0, bytes) != cudaSuccess) ok = 0;
  if (ok && cudaIpcGetMemHandle(&mine, peer_region) != cudaSuccess) ok = 0;
  if (cudaMalloc(&d_ticket, sizeof(unsigned)) != cudaSuccess || cudaMemset(d_ticket, 0, sizeof(unsigned)) != cudaSuccess) ok = 0;
  if (cudaMalloc(&d_peer_err, sizeof(int)) != cudaSuccess || cudaMemset(d_peer_err, 0, sizeof(int)) != cudaSuccess) ok = 0;
  cudaGetLastError();
  // all-gather of the 64-byte handles through the communicator
  unsigned char* d_h = nullptr;
  std::vector<cudaIpcMemHandle_t> all(nranks);
  if (cudaMalloc(&d_h, sizeof(mine) * (nranks + 1)) != cudaSuccess) { pp_set_error("peer_setup: cudaMalloc failed"); return 1; }
  PP_CUDA(cudaMemcpyAsync(d_h, &mine, sizeof(mine), cudaMemcpyHostToDevice, st));
  int rc = g_nccl.allgather(d_h, d_h + sizeof(mine), sizeof(mine), /*ncclUint8*/ 1, comm, st);
  PP_REQUIRE(rc == 0, "ncclAllGather failed: %s", nccl_err(rc));
  PP_CUDA(cudaMemcpyAsync(all.data(), d_h + sizeof(mine), sizeof(mine) * nranks, cudaMemcpyDeviceToHost, st));
  PP_CUDA(cudaStreamSynchronize(st));
  cudaFree(d_h);
  for (int r = 0; r < nranks && ok; r++) {
    void* base = peer_region;
    if (r != rank) {
      if (cudaIpcOpenMemHandle(&base, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cudaGetLastError(); break; }
      peer_mapped[r] = base;
    }
    peers.flags[r] = reinterpret_cast<unsigned long long*>(base);
    peers.slots[r] = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(base) + flag_bytes);
  }
  // vote: the peer path is used only if every rank mapped every peer
  double vote = ok ? 0.0 : 1.0;
  if (allreduce_host(&vote, 1, st)) return 1;
  if (vote == 0.0) peer_state = 1;
  peer_epoch = 0;
  return 0;
}

int PPComm::reduce_sum(const double* partial, int n_cta, int stride, int ncols_partial, int ncols, double* out, cudaStream_t st,
                       long long* launches)
{
  if (nranks > 1 && peer_state == 0 && peer_setup((size_t)ncols, st)) return 1;
  if (nranks == 1 || peer_state != 1 || (size_t)ncols > peer_cap) {
    if (ppk_reduce_partials(partial, n_cta, stride, ncols_partial, out, st, launches)) return 1;
    return allreduce_device(out, ncols, st);
  }
  peer_epoch++;
  reduce_push_kernel<<<(ncols + 31) / 32, 32 * PP_RED_GROUPS, 0, st>>>(partial, n_cta, stride, ncols_partial, ncols, out, peers, nranks, rank, peer_cap,
                                                        peer_epoch, d_ticket);
  wait_sum_kernel<<<(ncols + 255) / 256, 256, 0, st>>>(peers.slots[rank], peers.flags[rank], nranks, peer_cap, ncols, peer_epoch, out,
                                                       d_peer_err);
  *launches += 2;
  PP_CUDA(cudaGetLastError());
  return 0;
}

int PPComm::sendrecv(int n_peers, const int* peer, double* const* sendbuf, const size_t* send_n, double* const* recvbuf, const size_t* recv_n,
                     cudaStream_t st)
{
  if (n_peers == 0) return 0;
  PP_REQUIRE(comm && g_nccl.send && g_nccl.recv && g_nccl.group_start && g_nccl.group_end, "point-to-point exchange needs an initialised communicator");
  int rc = g_nccl.group_start();
  PP_REQUIRE(rc == 0, "ncclGroupStart failed: %s", nccl_err(rc));
  for (int i = 0; i < n_peers; i++) {
    if (send_n[i]) { rc = g_nccl.send(sendbuf[i], send_n[i], /*ncclDouble*/ 8, peer[i], comm, st); PP_REQUIRE(rc == 0, "ncclSend failed: %s", nccl_err(rc)); }
    if (recv_n[i]) { rc = g_nccl.recv(recvbuf[i], recv_n[i], /*ncclDouble*/ 8, peer[i], comm, st); PP_REQUIRE(rc == 0, "ncclRecv failed: %s", nccl_err(rc)); }
  }
  rc = g_nccl.group_end();
  PP_REQUIRE(rc == 0, "ncclGroupEnd failed: %s", nccl_err(rc));
  return 0;
}

void PPComm::destroy()
{
  for (int r = 0; r < PP_PEER_MAX; r++) {
    if (peer_mapped[r]) cudaIpcCloseMemHandle(peer_mapped[r]);
    peer_mapped[r] = nullptr;
  }
  if (peer_region) cudaFree(peer_region);
  if (d_ticket) cudaFree(d_ticket);
  if (d_peer_err) cudaFree(d_peer_err);
  peer_region = nullptr; d_ticket = nullptr; d_peer_err = nullptr;
  peer_state = 0;
  peer_cap = 0;
  if (comm && g_nccl.destroy) g_nccl.destroy(comm);
  comm = nullptr;
  if (d_stage) cudaFree(d_stage);
  d_stage = nullptr;
  stage_cap = 0;
  nranks = 1;
  rank = 0;
}
