// Do FP64 DMMA (mma.sync.m8n8k4.f64) and scalar FP64 FMA share one datapath on this GPU?
// Three launches of the same grid: every warp on DMMA chains; every warp on DFMA chains; even warps on DMMA and odd
// warps on DFMA.  If the two instruction kinds had pipes of their own, the mixed launch would sustain both solo rates
// at once (sum of the fractions ~ 2 x 0.5 of each = each kind at its full solo rate for half the warps); on a shared
// datapath the fractions of the solo rates add up to ~1.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_share_probe pipe_share_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

// mode 0: all DMMA, 1: all DFMA, 2: even warps DMMA / odd warps DFMA; it_mma / it_fma: iterations of each kind
template <int CH>
__global__ void __launch_bounds__(256) mixed(double* out, int mode, int it_mma, int it_fma, double a, double b)
{
  const int warp = threadIdx.x >> 5;
  const bool do_mma = (mode == 0) || (mode == 2 && (warp & 1) == 0);
  double s = 0;
  if (do_mma) {
    double c0[CH], c1[CH];
#pragma unroll
    for (int c = 0; c < CH; c++) { c0[c] = threadIdx.x + c; c1[c] = c; }
    for (int it = 0; it < it_mma; it++) {
#pragma unroll
      for (int c = 0; c < CH; c++)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0[c]), "+d"(c1[c]) : "d"(a), "d"(b));
    }
#pragma unroll
    for (int c = 0; c < CH; c++) s += c0[c] + c1[c];
  } else {
    double acc[2 * CH];
#pragma unroll
    for (int c = 0; c < 2 * CH; c++) acc[c] = threadIdx.x + c;
    for (int it = 0; it < it_fma; it++) {
#pragma unroll
      for (int c = 0; c < 2 * CH; c++) acc[c] = acc[c] * a + b;
    }
#pragma unroll
    for (int c = 0; c < 2 * CH; c++) s += acc[c];
  }
  if (s == 123456.789) out[0] = s;
}

template <typename F>
static float time_ms(F launch, int reps)
{
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

int main()
{
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int sms = p.multiProcessorCount, threads = 256, blocks = sms * 8, warps = threads / 32;
  constexpr int CH = 8;
  double* out; CK(cudaMalloc(&out, 8));
  // a DMMA is 256 FMAs per warp, a DFMA warp instruction 32: 8 DFMA iterations of 2 CH chains ~ the pipe time of 2 DMMA iterations of CH chains
  const int it_mma = 4096, it_fma = 4 * it_mma;
  const float t_mma = time_ms([&] { mixed<CH><<<blocks, threads>>>(out, 0, it_mma, it_fma, 1.0000001, 1e-9); }, 10);
  const float t_fma = time_ms([&] { mixed<CH><<<blocks, threads>>>(out, 1, it_mma, it_fma, 1.0000001, 1e-9); }, 10);
  const float t_mix = time_ms([&] { mixed<CH><<<blocks, threads>>>(out, 2, it_mma, it_fma, 1.0000001, 1e-9); }, 10);
  const double fl_mma = 2.0 * 256 * CH * (double)it_mma * warps * blocks;           // all warps
  const double fl_fma = 2.0 * 32 * 2 * CH * (double)it_fma * warps * blocks;
  const double r_mma = fl_mma / t_mma * 1e-9, r_fma = fl_fma / t_fma * 1e-9;
  const double m_mma = 0.5 * fl_mma / t_mix * 1e-9, m_fma = 0.5 * fl_fma / t_mix * 1e-9;   // half the warps each
  printf("{\"gpu\": \"%s\", \"dmma_solo_tflops\": %.2f, \"dfma_solo_tflops\": %.2f, \"mixed_ms\": %.3f, \"dmma_solo_ms\": %.3f, \"dfma_solo_ms\": %.3f, "
         "\"mixed_dmma_tflops\": %.2f, \"mixed_dfma_tflops\": %.2f, \"sum_of_solo_fractions\": %.3f}\n",
         p.name, r_mma, r_fma, t_mix, t_mma, t_fma, m_mma, m_fma, m_mma / r_mma + m_fma / r_fma);
  return 0;
}
