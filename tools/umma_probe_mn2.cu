// Probe 2: tcgen05.mma kind::tf32 with MN-major operands in the 128-byte-swizzle layout.
// Tile [pixel][64 channels] stored as two 32-channel halves; half h, pixel p, 16-byte chunk q (4 channels):
//   byte = h * (P/8 * 1024) + (p >> 3) * 1024 + (p & 7) * 128 + ((q ^ (p & 7)) * 16)
// read (a) MN-major as A (M = 64 ch) and B (N = 64 ch), K = pixels; (b) K-major as A (M = 128 px), K = channels.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "mfcc_dev.cuh"

using namespace kws;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int C = 64, P = 128;
constexpr uint32_t kHalf = P / 8 * 1024, kTile = 2 * kHalf;  // 32 KB

__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3ffff) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}

__global__ void probe(const float* G, const float* X, const float* Wt, float* out1, float* out2, uint32_t lbo, uint32_t sbo) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* sG = smem;
  unsigned char* sX = smem + kTile;
  unsigned char* sW = smem + 2 * kTile;  // K-major no-swizzle [ci][co]
  for (int i = tid; i < P * C; i += blockDim.x) {
    const int px = i / C, c = i % C, h = c >> 5, q = (c >> 2) & 7;
    const uint32_t off = h * kHalf + (px >> 3) * 1024 + (px & 7) * 128 + ((q ^ (px & 7)) * 16) + (c & 3) * 4;
    *reinterpret_cast<float*>(sG + off) = G[i];
    *reinterpret_cast<float*>(sX + off) = X[i];
  }
  for (int i = tid; i < C * C; i += blockDim.x) {
    const int n = i / C, k = i % C;
    *reinterpret_cast<float*>(sW + (n >> 3) * 2048 + (k >> 2) * 128 + (n & 7) * 16 + (k & 3) * 4) = Wt[i];
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
  {  // clear
    for (int col = 0; col < 128; ++col)
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(tmem + ((uint32_t)(warp * 32) << 16) + col), "r"(__float_as_uint(-777.f)) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) {
    const uint32_t idesc1 = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(C >> 3) << 17) | ((64u >> 4) << 24);
    for (int ks = 0; ks < P / 8; ++ks)  // 8 pixels per MMA = one 1024-byte swizzle atom
      tc_mma_tf32_elect(tmem, desc_sw128(smem_u32(sG) + ks * 1024, lbo, sbo), desc_sw128(smem_u32(sX) + ks * 1024, lbo, sbo), idesc1,
                        ks > 0);
    // (b) K-major SW128: rows = pixels, 8-row groups 1024 B apart; 8 channels (32 B) per MMA inside the 128-byte row,
    //     second half of the channels in the second half tile; B = no-swizzle K-major weights
    const uint32_t idesc2 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(C >> 3) << 17) | ((128u >> 4) << 24);
    for (int ks = 0; ks < C / 8; ++ks)
      tc_mma_tf32_elect(tmem + 64, desc_sw128(smem_u32(sG) + (ks >> 2) * kHalf + (ks & 3) * 32, 16u, 1024u),
                        umma_desc(smem_u32(sW) + ks * 256, 128u, 2048u), idesc2, ks > 0);
    tc_commit_elect(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  for (int col = 0; col < 128; ++col) {
    uint32_t r;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(tmem + ((uint32_t)(warp * 32) << 16) + col) : "memory");
    tc_wait_ld();
    if (col < 64) out1[(warp * 32 + lane) * 64 + col] = __uint_as_float(r);
    else out2[(warp * 32 + lane) * 64 + col - 64] = __uint_as_float(r);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(128));
}

int main() {
  std::vector<float> G(P * C), X(P * C), Wt(C * C);
  for (int i = 0; i < P * C; ++i) { G[i] = (float)((i * 7 + 3) % 17 - 8) * 0.25f; X[i] = (float)((i * 5 + 1) % 13 - 6) * 0.5f; }
  for (int i = 0; i < C * C; ++i) Wt[i] = (float)((i * 11 + 5) % 19 - 9) * 0.125f;
  float *dG, *dX, *dW, *dO1, *dO2;
  CK(cudaMalloc(&dG, G.size() * 4)); CK(cudaMalloc(&dX, X.size() * 4)); CK(cudaMalloc(&dW, Wt.size() * 4));
  CK(cudaMalloc(&dO1, 128 * 64 * 4)); CK(cudaMalloc(&dO2, 128 * 64 * 4));
  CK(cudaMemcpy(dG, G.data(), G.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dW, Wt.data(), Wt.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  int ok = 0;
  const uint32_t cand[][2] = {{kHalf, 1024}, {1024, kHalf}, {16, 1024}, {kHalf, kHalf}};
  for (auto& cd : cand) {
    probe<<<1, 128, 100 * 1024>>>(dG, dX, dW, dO1, dO2, cd[0], cd[1]);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
    std::vector<float> O1(128 * 64), O2(128 * 64);
    CK(cudaMemcpy(O1.data(), dO1, O1.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(O2.data(), dO2, O2.size() * 4, cudaMemcpyDeviceToHost));
    double e1 = 0, e2 = 0;
    int untouched = 0;
    for (int co = 0; co < C; ++co) {
      const int lane = 32 * (co / 16) + co % 16;
      for (int ci = 0; ci < C; ++ci) {
        double ref = 0;
        for (int px = 0; px < P; ++px) ref += (double)G[px * C + co] * X[px * C + ci];
        if (O1[lane * 64 + ci] == -777.f) ++untouched;
        e1 = fmax(e1, fabs(ref - O1[lane * 64 + ci]));
      }
    }
    for (int px = 0; px < P; ++px)
      for (int ci = 0; ci < C; ++ci) {
        double ref = 0;
        for (int co = 0; co < C; ++co) ref += (double)G[px * C + co] * Wt[ci * C + co];
        e2 = fmax(e2, fabs(ref - O2[px * 64 + ci]));
      }
    printf("LBO=%u SBO=%u: MN-major SW128 A,B: max err %.3g (untouched %d) %s | K-major SW128 A M=128: max err %.3g %s\n", cd[0], cd[1], e1,
           untouched, e1 < 1e-3 ? "OK" : "MISMATCH", e2, e2 < 1e-3 ? "OK" : "MISMATCH");
    if (e1 < 1e-3 && e2 < 1e-3) ++ok;
  }
  printf(ok ? "PROBE PASSED\n" : "PROBE FAILED\n");
  return 0;
}
