// Probe of the tcgen05 facts the fused model kernel (csrc/model_fused.cu) relies on, run once on a B200:
//   * kind::tf32 MMA with M = 64 (cta_group::1): accumulator row m lands in TMEM lane 32*(m/16) + m%16,
//     and a second M = 64 tile may be placed at lane offset 16 (CUTLASS "interleaved" allocation);
//   * K-major no-swizzle operands with a leading-dimension byte offset other than 128 (144: spreads
//     the 16 K-chunks of a row over all 32 banks for channel-per-thread stores);
//   * N = 24 / N = 16 tiles, D at a non-zero column.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I keyword-spotting_b200/csrc -I include -o tools/umma_probe tools/umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "mfcc_dev.cuh"

using namespace kws;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int M = 64, K = 64;

struct Cfg { int N, lbo, lane_off, col_off; };

__global__ void probe(const float* A, const float* Bm, float* out, Cfg c) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t sbo = (uint32_t)c.lbo * (K / 4);
  unsigned char* sA = smem;
  unsigned char* sB = smem + 8 * sbo;
  for (int i = tid; i < M * K; i += blockDim.x) {
    const int m = i / K, k = i % K;
    *reinterpret_cast<float*>(sA + (m >> 3) * sbo + (k >> 2) * c.lbo + (m & 7) * 16 + (k & 3) * 4) = A[i];
  }
  for (int i = tid; i < c.N * K; i += blockDim.x) {
    const int n = i / K, k = i % K;
    *reinterpret_cast<float*>(sB + (n >> 3) * sbo + (k >> 2) * c.lbo + (n & 7) * 16 + (k & 3) * 4) = Bm[i];
  }
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(128));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  // poison the accumulator region so untouched lanes are visible
  {
    uint32_t z = __float_as_uint(-777.f);
    for (int col = 0; col < 64; ++col)
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(tmem + ((uint32_t)(warp * 32) << 16) + col), "r"(z) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(c.N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const uint32_t d = tmem + ((uint32_t)c.lane_off << 16) + (uint32_t)c.col_off;
    for (int ks = 0; ks < K / 8; ++ks)
      tc_mma_tf32(d, umma_desc(smem_u32(sA) + ks * 2 * c.lbo, c.lbo, sbo), umma_desc(smem_u32(sB) + ks * 2 * c.lbo, c.lbo, sbo),
                  idesc, ks > 0);
    tc_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  for (int col = 0; col < 64; ++col) {
    uint32_t r;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(tmem + ((uint32_t)(warp * 32) << 16) + col) : "memory");
    tc_wait_ld();
    out[(warp * 32 + lane) * 64 + col] = __uint_as_float(r);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(128));
}

int main() {
  std::vector<float> A(M * K), B(32 * K);
  for (int i = 0; i < M * K; ++i) A[i] = (float)((i * 7 + 3) % 17 - 8) * 0.25f;     // exact in tf32
  for (int i = 0; i < 32 * K; ++i) B[i] = (float)((i * 5 + 1) % 13 - 6) * 0.5f;
  float *dA, *dB, *dO;
  CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dO, 128 * 64 * 4));
  CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
  const Cfg cfgs[] = {{24, 128, 0, 0}, {24, 128, 16, 0}, {24, 144, 0, 8}, {24, 144, 16, 24}, {16, 144, 16, 40}, {24, 272, 16, 5}};
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  int bad = 0;
  for (const Cfg& c : cfgs) {
    probe<<<1, 128, 100 * 1024>>>(dA, dB, dO, c);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("N=%d lbo=%d lane_off=%d col_off=%d: CUDA error %s\n", c.N, c.lbo, c.lane_off, c.col_off, cudaGetErrorString(e)); return 1; }
    std::vector<float> O(128 * 64);
    CK(cudaMemcpy(O.data(), dO, O.size() * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0; int touched_wrong = 0;
    for (int lane = 0; lane < 128; ++lane) {
      const int in_tile = ((lane & 31) >> 4) == (c.lane_off >> 4);
      const int m = (lane >> 5) * 16 + (lane & 15);
      for (int col = 0; col < 64; ++col) {
        const float got = O[lane * 64 + col];
        const int n = col - c.col_off;
        if (in_tile && n >= 0 && n < c.N) {
          double ref = 0;
          for (int k = 0; k < K; ++k) ref += (double)A[m * K + k] * B[n * K + k];
          const double err = fabs(ref - got);
          if (err > maxerr) maxerr = err;
        } else if (got != -777.f) {
          ++touched_wrong;
        }
      }
    }
    printf("N=%d lbo=%d lane_off=%d col_off=%d: max err %.3g, cells touched outside the tile %d  %s\n", c.N, c.lbo, c.lane_off,
           c.col_off, maxerr, touched_wrong, (maxerr < 1e-3 && touched_wrong == 0) ? "OK" : "MISMATCH");
    if (!(maxerr < 1e-3 && touched_wrong == 0)) ++bad;
  }
  printf(bad ? "PROBE FAILED\n" : "PROBE PASSED\n");
  return 0;
}
