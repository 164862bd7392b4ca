// FMA-chain microbenchmark: measures the FP64 / FP32 vector-FMA and FP64 DMMA (mma.sync) peaks of the
// GPU it runs on.  SURVEY.md section 8(d): MEASURED_PEAKS.json holds only HBM and bf16 GEMM figures, so
// the FP64/FP32 roofline denominators for the descriptor / force / MLP kernels are measured here.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o peaks_microbench peaks_microbench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

template <typename T, int CHAINS>
__global__ void fma_chain(T* out, int iters, T a, T b) {
  T acc[CHAINS];
#pragma unroll
  for (int c = 0; c < CHAINS; c++) acc[c] = (T)(threadIdx.x + c);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CHAINS; c++) acc[c] = acc[c] * a + b;
  }
  T s = 0;
#pragma unroll
  for (int c = 0; c < CHAINS; c++) s += acc[c];
  if (s == (T)123456.789) out[0] = s;  // never true; defeats DCE
}

template <int CHAINS>
__global__ void dmma_chain(double* out, int iters, double a, double b) {
  double c0[CHAINS], c1[CHAINS];
#pragma unroll
  for (int c = 0; c < CHAINS; c++) { c0[c] = threadIdx.x + c; c1[c] = c; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CHAINS; c++) {
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                   : "+d"(c0[c]), "+d"(c1[c]) : "d"(a), "d"(b));
    }
  }
  double s = 0;
#pragma unroll
  for (int c = 0; c < CHAINS; c++) s += c0[c] + c1[c];
  if (s == 123456.789) out[0] = s;
}

template <int CHAINS>
__global__ void dmma16816_chain(double* out, int iters, double a, double b) {
  double c0[CHAINS], c1[CHAINS], c2[CHAINS], c3[CHAINS];
#pragma unroll
  for (int c = 0; c < CHAINS; c++) { c0[c] = threadIdx.x + c; c1[c] = c; c2[c] = 1; c3[c] = 2; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int c = 0; c < CHAINS; c++) {
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
                   : "+d"(c0[c]), "+d"(c1[c]), "+d"(c2[c]), "+d"(c3[c])
                   : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b), "d"(b), "d"(a), "d"(b), "d"(a));
    }
  }
  double s = 0;
#pragma unroll
  for (int c = 0; c < CHAINS; c++) s += c0[c] + c1[c] + c2[c] + c3[c];
  if (s == 123456.789) out[0] = s;
}

__global__ void exp_chain(double* out, int iters, double a) {
  double x0 = threadIdx.x * 1e-3, x1 = x0 + 0.1, x2 = x0 + 0.2, x3 = x0 + 0.3;
  for (int it = 0; it < iters; it++) {
    x0 = exp(-x0 * a); x1 = exp(-x1 * a); x2 = exp(-x2 * a); x3 = exp(-x3 * a);
  }
  if (x0 + x1 + x2 + x3 == 123456.789) out[0] = x0;
}

__global__ void sqrtdiv_chain(double* out, int iters, double a) {
  double x0 = threadIdx.x + 1.0, x1 = x0 + 0.1, x2 = x0 + 0.2, x3 = x0 + 0.3;
  for (int it = 0; it < iters; it++) {
    x0 = sqrt(x0 + a) / x1; x1 = sqrt(x1 + a) / x2; x2 = sqrt(x2 + a) / x3; x3 = sqrt(x3 + a) / x0;
  }
  if (x0 + x1 + x2 + x3 == 123456.789) out[0] = x0;
}

template <typename F>
static float time_ms(F launch, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int w = 0; w < 3; w++) launch();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    CK(cudaEventRecord(e0));
    launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  double* out; CK(cudaMalloc(&out, 8));
  const int threads = 256, iters = 4096;
  const int blocks = sms * 8;
  printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, sms);
  {
    constexpr int CH = 16;
    float ms = time_ms([&] { fma_chain<double, CH><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); }, 10);
    double fl = 2.0 * CH * (double)iters * threads * blocks;
    printf(", \"fp64_fma_tflops\": %.2f", fl / ms * 1e-9);
  }
  {
    constexpr int CH = 16;
    float ms = time_ms([&] { fma_chain<float, CH><<<blocks, threads>>>((float*)out, iters, 1.0000001f, 1e-9f); }, 10);
    double fl = 2.0 * CH * (double)iters * threads * blocks;
    printf(", \"fp32_fma_tflops\": %.2f", fl / ms * 1e-9);
  }
  {
    constexpr int CH = 8;
    float ms = time_ms([&] { dmma_chain<CH><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); }, 10);
    double fl = 2.0 * 8 * 8 * 4 * CH * (double)iters * (threads / 32) * blocks;
    printf(", \"fp64_dmma_m8n8k4_tflops\": %.2f", fl / ms * 1e-9);
  }
  {
    constexpr int CH = 4;
    float ms = time_ms([&] { dmma16816_chain<CH><<<blocks, threads>>>(out, iters / 4, 1.0000001, 1e-9); }, 10);
    double fl = 2.0 * 16 * 8 * 16 * CH * (double)(iters / 4) * (threads / 32) * blocks;
    printf(", \"fp64_dmma_m16n8k16_tflops\": %.2f", fl / ms * 1e-9);
  }
  {
    float ms = time_ms([&] { exp_chain<<<blocks, threads>>>(out, iters / 8, 0.5); }, 5);
    double n = 4.0 * (iters / 8) * threads * (double)blocks;
    printf(", \"fp64_exp_gops\": %.2f", n / ms * 1e-6);
  }
  {
    float ms = time_ms([&] { sqrtdiv_chain<<<blocks, threads>>>(out, iters / 8, 0.5); }, 5);
    double n = 4.0 * (iters / 8) * threads * (double)blocks;
    printf(", \"fp64_sqrt_plus_div_gops\": %.2f", n / ms * 1e-6);
  }
  printf("}\n");
  return 0;
}
