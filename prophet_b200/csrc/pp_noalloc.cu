// pp_column_sums_static: column sums of a device-resident matrix with NO memory allocation.
//
// Some users of this library are not allowed to call cudaMalloc when they need descriptor statistics: a host
// application (an MD engine, a training job) that has already handed the whole GPU to its own memory pool leaves
// nothing for the driver allocator, so a late cudaMalloc fails or forces the pool to shrink; and code that runs in
// a time-critical loop cannot afford cudaMalloc / cudaFree because both may synchronise the whole device.  Those
// callers still need the two reductions of Potential::Precondition (sum of a descriptor column, and sum of squared
// deviations from a per-column shift).  This entry point does exactly that work with the per-block partial sums
// kept in a fixed array that lives in the library image instead of in allocated memory.
//
// Consequences of that design, all of which matter to the caller:
//   * the array costs PP_STATIC_SCRATCH_BYTES on every GPU that loads the library, from load until exit;
//   * its capacity is fixed: matrices wider than PP_STATIC_MAX_COLS columns are refused with an error;
//   * there is one array per GPU, shared by every call on that GPU.  It keeps whatever the previous call left in
//     it; each call overwrites all the slots it later reads, so nothing has to be cleared, but two calls whose
//     kernels overlap on the same GPU (different streams, different host threads) would write the same slots and
//     corrupt each other.  Callers must serialise calls per GPU;
//   * the first call on a GPU looks up the array's device address and caches it.  That lookup is a set-up step:
//     make this first call early, before starting any stream capture on that GPU.
#include "../../include/prophet_b200.h"
#include "pp_common.cuh"

#define PP_STATIC_MAX_COLS 512
#define PP_STATIC_BLOCKS 64
#define PP_STATIC_SCRATCH_BYTES (PP_STATIC_BLOCKS * PP_STATIC_MAX_COLS * 8)
#define PP_STATIC_MAX_DEVICES 64

// per-block partial sums: [PP_STATIC_BLOCKS][PP_STATIC_MAX_COLS]
static __device__ __align__(256) double g_static_partial[PP_STATIC_BLOCKS * PP_STATIC_MAX_COLS];

// Device address of g_static_partial on each GPU.  In host code the symbol name is only a handle for the runtime,
// not an address a kernel could use, so the address is asked for once per device and passed to the kernels.
static double* g_static_partial_addr[PP_STATIC_MAX_DEVICES] = {nullptr};

// Each block sums a contiguous range of rows for every column into partial[block][col].
__global__ void __launch_bounds__(256)
column_partial_static_kernel(const double* __restrict__ A, long long n_rows, int n_cols, const double* __restrict__ shift,
                             int squared, double* __restrict__ partial)
{
  const long long per = (n_rows + gridDim.x - 1) / gridDim.x;
  const long long r0 = (long long)blockIdx.x * per;
  const long long r1 = r0 + per < n_rows ? r0 + per : n_rows;
  for (int k = threadIdx.x; k < n_cols; k += blockDim.x) {
    const double sh = (squared && shift) ? shift[k] : 0.0;
    double s = 0.0;
    for (long long r = r0; r < r1; r++) {
      const double v = A[(size_t)r * n_cols + k];
      if (squared) { const double d = v - sh; s += d * d; }
      else s += v;
    }
    partial[(size_t)blockIdx.x * PP_STATIC_MAX_COLS + k] = s;    // every slot read below is written here first
  }
}

// Adds the blocks in a fixed order, so the result does not change from run to run.
__global__ void column_final_static_kernel(const double* __restrict__ partial, int n_blocks, int n_cols, double* __restrict__ out)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_cols) return;
  double s = 0.0;
  for (int b = 0; b < n_blocks; b++) s += partial[(size_t)b * PP_STATIC_MAX_COLS + k];
  out[k] = s;
}

extern "C" int pp_column_sums_static(int device, const double* matrix_dev, long long n_rows, int n_cols,
                                     const double* shift_dev, int squared, double* sums_out_dev, void* stream)
{
  PP_REQUIRE(device >= 0 && device < PP_STATIC_MAX_DEVICES, "pp_column_sums_static: device %d out of range", device);
  PP_REQUIRE(matrix_dev != nullptr && sums_out_dev != nullptr, "pp_column_sums_static: NULL matrix or output pointer");
  PP_REQUIRE(n_rows >= 0, "pp_column_sums_static: negative row count");
  PP_REQUIRE(n_cols >= 1, "pp_column_sums_static: need at least one column");
  PP_REQUIRE(n_cols <= PP_STATIC_MAX_COLS,
             "pp_column_sums_static: %d columns exceed the fixed scratch capacity of %d columns; use pp_precondition "
             "(which allocates) for wider matrices", n_cols, PP_STATIC_MAX_COLS);
  int caller_device = device;
  PP_CUDA(cudaGetDevice(&caller_device));      // the caller's current device is put back before returning
  PP_CUDA(cudaSetDevice(device));
  if (g_static_partial_addr[device] == nullptr) {
    // one-time set-up on this GPU (with it current): resolve the array's device address
    double* addr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&addr, g_static_partial);
    PP_REQUIRE(e == cudaSuccess && addr != nullptr, "pp_column_sums_static: cudaGetSymbolAddress failed on device %d: %s", device,
               cudaGetErrorString(e));
    g_static_partial_addr[device] = addr;
  }
  double* partial = g_static_partial_addr[device];
  cudaStream_t st = (cudaStream_t)stream;
  column_partial_static_kernel<<<PP_STATIC_BLOCKS, 256, 0, st>>>(matrix_dev, n_rows, n_cols, shift_dev, squared ? 1 : 0, partial);
  PP_CUDA(cudaGetLastError());
  column_final_static_kernel<<<(n_cols + 127) / 128, 128, 0, st>>>(partial, PP_STATIC_BLOCKS, n_cols, sums_out_dev);
  PP_CUDA(cudaGetLastError());
  if (caller_device != device) PP_CUDA(cudaSetDevice(caller_device));
  return 0;
}
