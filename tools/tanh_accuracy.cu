// Accuracy check of the fast tanh / sech^2 used by mlp_kernel against the CUDA math library (tuning aid).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I prophet_b200/csrc tools/tanh_accuracy.cu -o /tmp/tanh_acc && /tmp/tanh_acc
#include <cstdio>
#include <cmath>
#include "pp_fastmath.cuh"

__global__ void k(int n, double lo, double hi, double* err)
{
  __shared__ double tab[64];
  for (int i = threadIdx.x; i < 64; i += blockDim.x) tab[i] = exp2((double)i / 64.0);
  __syncthreads();
  double e_t = 0, e_d = 0, e_e = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double x = lo + (hi - lo) * (double)i / (double)(n - 1);
    double d;
    double t = pp_fast_tanh(x, d, tab);
    double tr = tanh(x), c = cosh(x), dr = 1.0 / (c * c);
    if (fabs(x) > 6.0) {   // the reference's asymptotic branch (NN_node.h:193-206)
      double e = exp(-2.0 * fabs(x));
      tr = (x > 0 ? 1.0 : -1.0) * (1.0 - 2.0 * e * (1.0 - e));
      dr = 4.0 * e * (1.0 - 2.0 * e);
    }
    if (tr != 0) e_t = fmax(e_t, fabs(t - tr) / fabs(tr));
    if (dr > 1e-300) e_d = fmax(e_d, fabs(d - dr) / dr);
    double ex = pp_fast_exp_nonpos(-fabs(x) * 2.0, tab), er = exp(-fabs(x) * 2.0);
    if (er > 1e-300) e_e = fmax(e_e, fabs(ex - er) / er);
  }
  atomicMax((unsigned long long*)&err[0], (unsigned long long)__double_as_longlong(e_t));
  atomicMax((unsigned long long*)&err[1], (unsigned long long)__double_as_longlong(e_d));
  atomicMax((unsigned long long*)&err[2], (unsigned long long)__double_as_longlong(e_e));
}

int main()
{
  double* err;
  cudaMallocManaged(&err, 3 * sizeof(double));
  const double ranges[4][2] = {{-1e-3, 1e-3}, {-6.5, 6.5}, {-40, 40}, {-800, 800}};
  for (int r = 0; r < 4; r++) {
    err[0] = err[1] = err[2] = 0;
    k<<<296, 256>>>(20000001, ranges[r][0], ranges[r][1], err);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("cuda error\n"); return 1; }
    printf("x in [%g, %g]: max rel err tanh %.3e  sech2 %.3e  exp(-2|x|) %.3e\n", ranges[r][0], ranges[r][1], err[0], err[1], err[2]);
  }
  return 0;
}
