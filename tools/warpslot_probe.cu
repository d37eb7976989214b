// Which hardware warp slot (%warpid) and SM each warp of a 6-warp, 2-CTAs-per-SM persistent launch lands on.
// The warp scheduler (SM sub-partition) of a warp is its slot modulo 4; the MFCC kernel's role layout
// (mfcc_ws.cu) is chosen from this.  build: nvcc -arch=sm_100a -o tools/warpslot_probe tools/warpslot_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(192, 2) probe(int* out, long long spin) {
  extern __shared__ unsigned char smem[];
  unsigned smid, wid;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  asm volatile("mov.u32 %0, %%warpid;" : "=r"(wid));
  const long long t0 = clock64();
  while (clock64() - t0 < spin) smem[threadIdx.x] = (unsigned char)wid;  // keep every CTA resident together
  if ((threadIdx.x & 31) == 0) {
    int* o = out + (blockIdx.x * 6 + (threadIdx.x >> 5)) * 2;
    o[0] = (int)smid;
    o[1] = (int)wid;
  }
}
int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int grid = 2 * sms, smem = 110 * 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(probe, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  int* d;
  cudaMalloc(&d, grid * 12 * sizeof(int));
  probe<<<grid, 192, smem>>>(d, 2000000);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("launch failed\n"); return 1; }
  int* h = new int[grid * 12];
  cudaMemcpy(h, d, grid * 12 * sizeof(int), cudaMemcpyDeviceToHost);
  int hist[8][4] = {};
  for (int b = 0; b < grid; ++b) {
    if (b < 4 || (b >= sms && b < sms + 4)) {
      printf("block %3d sm %3d slots:", b, h[b * 12]);
      for (int w = 0; w < 6; ++w) printf(" %2d", h[(b * 6 + w) * 2 + 1]);
      printf("\n");
    }
    for (int w = 0; w < 6; ++w) hist[w][h[(b * 6 + w) * 2 + 1] & 3]++;
  }
  // per SM: which blocks share it
  int shown = 0;
  for (int b = 0; b < grid && shown < 4; ++b)
    for (int c = b + 1; c < grid; ++c)
      if (h[b * 12] == h[c * 12]) { printf("sm %d holds blocks %d and %d\n", h[b * 12], b, c); ++shown; break; }
  for (int w = 0; w < 6; ++w) printf("cta warp %d: slot%%4 histogram %d %d %d %d\n", w, hist[w][0], hist[w][1], hist[w][2], hist[w][3]);
  return 0;
}
