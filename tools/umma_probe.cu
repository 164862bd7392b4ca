// Probe of the tcgen05 (UMMA) building blocks the FP32-mode training kernel uses, checked against a CPU product:
//   test 1: D[128 x 64] = A[128 x K] * B[64 x K]^T, kind::tf32, A and B K-major, no swizzle (canonical "interleave"
//           layout: 8-row x 16-byte core matrices), accumulator in TMEM, read back with tcgen05.ld.32x32b
//   test 2: D[64 x 64] = A^T * B with A stored [K x 64] and B stored [K x 64] (both MN-major descriptors over the SAME
//           kind of storage), M = 64: prints which TMEM lanes hold which rows
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_probe umma_probe.cu ; run on a B200.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor, no swizzle: start address, leading / stride byte offsets (all >> 4), version 1
__device__ __forceinline__ unsigned long long umma_desc(unsigned addr, unsigned lbo_bytes, unsigned sbo_bytes)
{
  unsigned long long d = 0;
  d |= (unsigned long long)((addr >> 4) & 0x3FFF);
  d |= (unsigned long long)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (unsigned long long)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= 1ull << 46;                       // version = 1 (Blackwell)
  return d;
}
// instruction descriptor: F32 accumulate, TF32 x TF32, majors, N >> 3, M >> 4
__host__ __device__ inline unsigned umma_idesc(int M, int N, int a_mn, int b_mn)
{
  return (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)a_mn << 15) | ((unsigned)b_mn << 16) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(unsigned tmem_d, unsigned long long da, unsigned long long db, unsigned idesc, unsigned accumulate)
{
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar)
{
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity)
{
  asm volatile(
      "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// element (r, k) of a K-major operand with `rows` rows and K columns (K multiple of 8): core matrices of 8 rows x 4 floats
__host__ __device__ inline int kmajor_off(int r, int k, int K) { return ((r >> 3) * (K >> 2) + (k >> 2)) * 32 + (r & 7) * 4 + (k & 3); }

// A: a_mn == 0: [M][K] (K-major operand); a_mn == 1: [K][M] (MN-major operand over the row-major-by-atom storage)
// B: b_mn == 0: [64][K];                     b_mn == 1: [K][64]
__global__ void __launch_bounds__(128) probe_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int K, int M,
                                                    int a_mn, int b_mn, int variant)
{
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ unsigned tmem_base_s;
  __shared__ __align__(8) unsigned long long bar;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = 64;
  float* As = reinterpret_cast<float*>(smem);
  float* Bs = As + 128 * 128;
  if (!a_mn) { for (int i = tid; i < M * K; i += 128) As[kmajor_off(i / K, i % K, K)] = A[i]; }
  else { for (int i = tid; i < K * M; i += 128) As[kmajor_off(i / M, i % M, M)] = A[i]; }
  if (!b_mn) { for (int i = tid; i < N * K; i += 128) Bs[kmajor_off(i / K, i % K, K)] = B[i]; }
  else { for (int i = tid; i < K * N; i += 128) Bs[kmajor_off(i / N, i % N, N)] = B[i]; }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(64) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) mbar_init(&bar, 1);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned tmem = tmem_base_s;
  if (tid == 0) {
    const unsigned idesc = umma_idesc(M, N, a_mn, b_mn);
    for (int s = 0; s < K / 8; s++) {
      // K-major: LBO = distance between the two 16-byte K chunks of one instruction (128 B), SBO = between 8-row groups.
      // MN-major over [K][cols] storage: MN chunks (4 columns) 128 B apart (SBO); a group of 8 atoms is (cols / 4) * 128 B (LBO)
      // variant 1: leading / stride offsets of the MN-major descriptors swapped
      const unsigned a_l = variant ? 128u : (unsigned)(M >> 2) * 128u, a_s = variant ? (unsigned)(M >> 2) * 128u : 128u;
      const unsigned b_l = variant ? 128u : (unsigned)(N >> 2) * 128u, b_s = variant ? (unsigned)(N >> 2) * 128u : 128u;
      const unsigned long long da = a_mn ? umma_desc(smem_u32(As) + s * (M >> 2) * 128, a_l, a_s)
                                         : umma_desc(smem_u32(As) + s * 256, 128, (K >> 2) * 128);
      const unsigned long long db = b_mn ? umma_desc(smem_u32(Bs) + s * (N >> 2) * 128, b_l, b_s)
                                         : umma_desc(smem_u32(Bs) + s * 256, 128, (K >> 2) * 128);
      umma_tf32(tmem, da, db, idesc, s > 0);
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  unsigned v[64];
  const unsigned taddr = tmem + ((unsigned)(32 * warp) << 16);
#pragma unroll
  for (int c = 0; c < 64; c += 8) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[c]), "=r"(v[c + 1]), "=r"(v[c + 2]), "=r"(v[c + 3]), "=r"(v[c + 4]), "=r"(v[c + 5]), "=r"(v[c + 6]), "=r"(v[c + 7])
                 : "r"(taddr + c));
  }
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int c = 0; c < 64; c++) D[(size_t)tid * 64 + c] = __uint_as_float(v[c]);      // D is indexed by TMEM LANE here
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64) : "memory");
}

static float tf32r(float x)
{
  unsigned u;
  memcpy(&u, &x, 4);
  u = (u + 0x1000u) & 0xFFFFE000u;      // round to nearest (ties away), keep 10 mantissa bits
  float y;
  memcpy(&y, &u, 4);
  return y;
}

static void run_case(int M, int K, int a_mn, int b_mn, float* dA, float* dB, float* dD, size_t smem, int variant = 0)
{
  std::vector<float> A(M * K), B(64 * K), D(128 * 64);
  for (auto& v : A) v = tf32r((float)rand() / RAND_MAX - 0.5f);       // logical A[m][k], B[n][k]
  for (auto& v : B) v = tf32r((float)rand() / RAND_MAX - 0.5f);
  std::vector<float> As(M * K), Bs(64 * K);
  for (int m = 0; m < M; m++) for (int k = 0; k < K; k++) As[a_mn ? k * M + m : m * K + k] = A[m * K + k];
  for (int n = 0; n < 64; n++) for (int k = 0; k < K; k++) Bs[b_mn ? k * 64 + n : n * K + k] = B[n * K + k];
  CK(cudaMemcpy(dA, As.data(), As.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, Bs.data(), Bs.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0, 128 * 64 * 4));
  probe_kernel<<<1, 128, smem>>>(dA, dB, dD, K, M, a_mn, b_mn, variant);
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  std::vector<double> R(M * 64);
  for (int m = 0; m < M; m++)
    for (int n = 0; n < 64; n++) {
      double r = 0;
      for (int k = 0; k < K; k++) r += (double)A[m * K + k] * B[n * K + k];
      R[m * 64 + n] = r;
    }
  int found = 0;
  std::vector<int> lane_of_row(M, -1);
  for (int m = 0; m < M; m++)
    for (int l = 0; l < 128; l++) {
      double e = 0;
      for (int n = 0; n < 64; n++) e = fmax(e, fabs(R[m * 64 + n] - D[l * 64 + n]));
      if (e < 1e-4) { lane_of_row[m] = l; found++; break; }
    }
  printf("variant %d M=%d K=%d A %s B %s: %d of %d rows found; lanes of rows 0,1,15,16,17,31,32,48,63: %d %d %d %d %d %d %d %d %d\n", variant, M, K, a_mn ? "MN" : "K",
         b_mn ? "MN" : "K", found, M, lane_of_row[0], lane_of_row[1], lane_of_row[15], lane_of_row[16], lane_of_row[17], lane_of_row[31],
         lane_of_row[32], lane_of_row[48], lane_of_row[63]);
  if (found == 0) printf("   D[lane 0][0..3] = %g %g %g %g ; R[0][0..3] = %g %g %g %g\n", D[0], D[1], D[2], D[3], R[0], R[1], R[2], R[3]);
}

int main()
{
  float *dA, *dB, *dD;
  CK(cudaMalloc(&dA, 128 * 128 * 4)); CK(cudaMalloc(&dB, 128 * 128 * 4)); CK(cudaMalloc(&dD, 128 * 64 * 4));
  const size_t smem = 2 * 128 * 128 * 4;
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  srand(1);
  run_case(128, 56, 0, 0, dA, dB, dD, smem);
  run_case(128, 64, 1, 0, dA, dB, dD, smem);
  run_case(128, 64, 0, 1, dA, dB, dD, smem);
  run_case(128, 128, 1, 1, dA, dB, dD, smem);
  run_case(64, 64, 0, 0, dA, dB, dD, smem);
  run_case(64, 128, 1, 1, dA, dB, dD, smem);
  run_case(128, 64, 1, 0, dA, dB, dD, smem, 1);
  run_case(128, 64, 0, 1, dA, dB, dD, smem, 1);
  run_case(128, 128, 1, 1, dA, dB, dD, smem, 1);
  run_case(64, 128, 1, 1, dA, dB, dD, smem, 1);
  run_case(128, 8, 1, 0, dA, dB, dD, smem, 0);
  run_case(128, 8, 1, 0, dA, dB, dD, smem, 1);
  run_case(128, 8, 0, 1, dA, dB, dD, smem, 0);
  run_case(128, 8, 0, 1, dA, dB, dD, smem, 1);
  return 0;
}
