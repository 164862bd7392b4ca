#include "pp_comm.h"
#include "pp_common.cuh"
#include <dlfcn.h>

namespace {
struct NcclId { char internal[128]; };
typedef int (*fn_get_id)(NcclId*);
typedef int (*fn_init_rank)(void**, int, NcclId, int);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_destroy)(void*);
typedef const char* (*fn_errstr)(int);

struct NcclApi {
  void* lib = nullptr;
  fn_get_id get_id = nullptr;
  fn_init_rank init_rank = nullptr;
  fn_allreduce allreduce = nullptr;
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

void PPComm::destroy()
{
  if (comm && g_nccl.destroy) g_nccl.destroy(comm);
  comm = nullptr;
  if (d_stage) cudaFree(d_stage);
  d_stage = nullptr;
  stage_cap = 0;
  nranks = 1;
  rank = 0;
}
