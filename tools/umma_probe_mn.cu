// Probe: does tcgen05.mma kind::tf32 accept MN-major (transposed) operands?
// RESULT on B200 (round 1): no -- with the a_major/b_major bits set the MMA writes zeros for every LBO/SBO
// combination, in this interleaved layout and in the 128-byte-swizzle layout of umma_probe_mn2.cu, while the
// K-major reads of the same tiles (form (b)) are exact.  The weight-gradient GEMM of the training step would
// therefore need transposed staging; see DESIGN.md 4.6.  One smem tile stored as
// [channel/4][pixel][4 channels] (16-byte units) is read
//   (a) MN-major as A (M = 64 channels, K = 128 pixels) and MN-major as B (N = 64 channels, K = 128 pixels):
//       D1[co, ci] = sum_px G[px, co] X[px, ci]                       (the weight-gradient GEMM)
//   (b) K-major as A (M = 128 pixels, K = 64 channels) against a K-major B (N = 64):
//       D2[px, ci] = sum_co G[px, co] Wt[ci, co]                      (the input-gradient GEMM)
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I keyword-spotting_b200/csrc -I include -o tools/umma_probe_mn tools/umma_probe_mn.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "mfcc_dev.cuh"

using namespace kws;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int C = 64, P = 128;
constexpr uint32_t kTile = C / 4 * P * 16;  // 32 KB

__global__ void probe(const float* G, const float* X, const float* Wt, float* out1, float* out2, uint32_t lbo, uint32_t sbo) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* sG = smem;
  unsigned char* sX = smem + kTile;
  unsigned char* sW = smem + 2 * kTile;  // K-major [ci][co]: 8-row groups of 16 chunks x 128 B
  for (int i = tid; i < P * C; i += blockDim.x) {
    const int px = i / C, c = i % C;
    *reinterpret_cast<float*>(sG + (c >> 2) * (P * 16) + px * 16 + (c & 3) * 4) = G[i];
    *reinterpret_cast<float*>(sX + (c >> 2) * (P * 16) + px * 16 + (c & 3) * 4) = X[i];
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
  if (warp == 0) {
    // (a) both operands MN-major: bits 15 (A) and 16 (B) of the instruction descriptor
    const uint32_t idesc1 = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(C >> 3) << 17) | ((64u >> 4) << 24);
    for (int ks = 0; ks < P / 8; ++ks)
      tc_mma_tf32_elect(tmem, umma_desc(smem_u32(sG) + ks * 128, lbo, sbo), umma_desc(smem_u32(sX) + ks * 128, lbo, sbo),
                        idesc1, ks > 0);
    // (b) the same G tile K-major as a 128-row A: chunk pitch (LBO) = P*16, 8-row group pitch (SBO) = 128
    const uint32_t idesc2 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(C >> 3) << 17) | ((128u >> 4) << 24);
    for (int ks = 0; ks < C / 8; ++ks)
      tc_mma_tf32_elect(tmem + 64, umma_desc(smem_u32(sG) + ks * 2 * (P * 16), P * 16, 128u),
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
  const uint32_t cand[][2] = {{128, P * 16}, {P * 16, 128}, {128, 128}, {P * 16, P * 16}};
  for (auto& cd : cand) {
    probe<<<1, 128, 100 * 1024>>>(dG, dX, dW, dO1, dO2, cd[0], cd[1]);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
    std::vector<float> O1(128 * 64), O2(128 * 64);
    CK(cudaMemcpy(O1.data(), dO1, O1.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(O2.data(), dO2, O2.size() * 4, cudaMemcpyDeviceToHost));
    double e1 = 0, e2 = 0;
    for (int co = 0; co < C; ++co) {
      const int lane = 32 * (co / 16) + co % 16;  // M = 64 accumulator row -> TMEM lane
      for (int ci = 0; ci < C; ++ci) {
        double ref = 0;
        for (int px = 0; px < P; ++px) ref += (double)G[px * C + co] * X[px * C + ci];
        e1 = fmax(e1, fabs(ref - O1[lane * 64 + ci]));
      }
    }
    for (int px = 0; px < P; ++px)
      for (int ci = 0; ci < C; ++ci) {
        double ref = 0;
        for (int co = 0; co < C; ++co) ref += (double)G[px * C + co] * Wt[ci * C + co];
        e2 = fmax(e2, fabs(ref - O2[px * 64 + ci]));
      }
    if (cd[0] == 128 && cd[1] == (uint32_t)P * 16) {
      for (int co = 0; co < 3; ++co) {
        printf("co=%d got:", co);
        for (int ci = 0; ci < 8; ++ci) printf(" %8.3f", O1[(32 * (co / 16) + co % 16) * 64 + ci]);
        printf("\n      ref:");
        for (int ci = 0; ci < 8; ++ci) {
          double ref = 0;
          for (int px = 0; px < P; ++px) ref += (double)G[px * C + co] * X[px * C + ci];
          printf(" %8.3f", ref);
        }
        printf("\n");
      }
    }
    printf("LBO=%u SBO=%u: MN-major A and B (wgrad form): max err %.3g %s | K-major M=128 (dgrad form): max err %.3g %s\n", cd[0], cd[1], e1,
           e1 < 1e-3 ? "OK" : "MISMATCH", e2, e2 < 1e-3 ? "OK" : "MISMATCH");
    if (e1 < 1e-3 && e2 < 1e-3) ++ok;
  }
  printf(ok ? "PROBE PASSED\n" : "PROBE FAILED\n");
  return 0;
}
