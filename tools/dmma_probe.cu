// fp64 tensor-core (DMMA) latency / throughput on B200, and a layout check of the larger f64 shapes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_probe dmma_probe.cu && ./dmma_probe
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double (&c)[4], const double (&a)[4], const double (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
template <int CH>
__global__ void k_dmma884(double* out, long long* cyc, int iters) {
  double c0[CH], c1[CH];
  for (int i = 0; i < CH; ++i) { c0[i] = threadIdx.x + i; c1[i] = i; }
  const double a = 1.0 + 1e-9 * threadIdx.x, b = 1.0 - 1e-9 * threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < CH; ++i) dmma884(c0[i], c1[i], a, b);
  long long t1 = clock64();
  double s = 0; for (int i = 0; i < CH; ++i) s += c0[i] + c1[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int CH>
__global__ void k_dmma1688(double* out, long long* cyc, int iters) {
  double c[CH][4];
  for (int i = 0; i < CH; ++i) for (int j = 0; j < 4; ++j) c[i][j] = threadIdx.x + i + j;
  const double a[4] = {1.0, 1.0 + 1e-9, 1.0 - 1e-9, 1.0}, b[2] = {1.0 - 1e-9 * threadIdx.x, 1.0};
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < CH; ++i) dmma1688(c[i], a, b);
  long long t1 = clock64();
  double s = 0; for (int i = 0; i < CH; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_dfma(double* out, long long* cyc, int iters) {
  double r[8];
  for (int i = 0; i < 8; ++i) r[i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = fma(r[i], 1.0000001, 0.5);
  long long t1 = clock64();
  double s = 0; for (int i = 0; i < 8; ++i) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// layout check of m16n8k8: D = A[16x8] . B[8x8] with the layout assumed in the comment of mfcc_ws_body.cuh
__global__ void k_layout(const double* A, const double* B, double* D) {
  const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
  double a[4] = {A[g * 8 + t], A[(g + 8) * 8 + t], A[g * 8 + t + 4], A[(g + 8) * 8 + t + 4]};
  double b[2] = {B[t * 8 + g], B[(t + 4) * 8 + g]};
  double c[4] = {0, 0, 0, 0};
  dmma1688(c, a, b);
  D[g * 8 + 2 * t] = c[0]; D[g * 8 + 2 * t + 1] = c[1]; D[(g + 8) * 8 + 2 * t] = c[2]; D[(g + 8) * 8 + 2 * t + 1] = c[3];
}
int main() {
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  int clk = 0; CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
  double* out; long long* cyc; CK(cudaMalloc(&out, 1 << 24)); CK(cudaMallocManaged(&cyc, 8));
  const int iters = 2048;
  printf("{\"sms\": %d, \"clock_mhz\": %d", sms, clk / 1000);
#define RUN(name, kern, warps, flops_per_instr, ch)                                                        \
  { kern<<<1, 32 * warps>>>(out, cyc, 64); CK(cudaDeviceSynchronize());                                    \
    kern<<<1, 32 * warps>>>(out, cyc, iters); CK(cudaDeviceSynchronize());                                 \
    printf(", \"%s_w%d_cyc_per_instr\": %.2f", name, warps, (double)*cyc / ((double)iters * ch)); }
  RUN("dmma884_chain1", k_dmma884<1>, 1, 512, 1)
  RUN("dmma884_chain4", k_dmma884<4>, 1, 512, 4)
  RUN("dmma884_chain8", k_dmma884<8>, 1, 512, 8)
  RUN("dmma884_chain8", k_dmma884<8>, 4, 512, 8)
  RUN("dmma884_chain8", k_dmma884<8>, 16, 512, 8)
  RUN("dmma1688_chain1", k_dmma1688<1>, 1, 2048, 1)
  RUN("dmma1688_chain4", k_dmma1688<4>, 1, 2048, 4)
  RUN("dmma1688_chain4", k_dmma1688<4>, 4, 2048, 4)
  RUN("dmma1688_chain4", k_dmma1688<4>, 16, 2048, 4)
  RUN("dfma_chain8", k_dfma, 1, 64, 8)
  RUN("dfma_chain8", k_dfma, 4, 64, 8)
  RUN("dfma_chain8", k_dfma, 16, 64, 8)
  // full-chip DMMA throughput
  { cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_dmma884<8><<<sms * 2, 512>>>(out, cyc, iters); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); k_dmma884<8><<<sms * 2, 512>>>(out, cyc, iters); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    printf(", \"dmma884_tflops\": %.2f", (double)sms * 2 * 16 * iters * 8 * 512.0 / (ms * 1e-3) * 1e-12);
    k_dmma1688<4><<<sms * 2, 512>>>(out, cyc, iters); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); k_dmma1688<4><<<sms * 2, 512>>>(out, cyc, iters); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf(", \"dmma1688_tflops\": %.2f", (double)sms * 2 * 16 * iters * 4 * 2048.0 / (ms * 1e-3) * 1e-12);
    k_dfma<<<sms * 2, 512>>>(out, cyc, iters); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); k_dfma<<<sms * 2, 512>>>(out, cyc, iters); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf(", \"dfma_tflops\": %.2f", (double)sms * 2 * 512 * iters * 8 * 2.0 / (ms * 1e-3) * 1e-12); }
  // layout check
  { double *A, *B, *D; CK(cudaMallocManaged(&A, 128 * 8)); CK(cudaMallocManaged(&B, 64 * 8)); CK(cudaMallocManaged(&D, 128 * 8));
    for (int i = 0; i < 128; ++i) A[i] = sin(0.37 * i) ; for (int i = 0; i < 64; ++i) B[i] = cos(0.11 * i + 1);
    k_layout<<<1, 32>>>(A, B, D); CK(cudaDeviceSynchronize());
    double err = 0; for (int m = 0; m < 16; ++m) for (int n = 0; n < 8; ++n) { double r = 0; for (int k = 0; k < 8; ++k) r += A[m * 8 + k] * B[k * 8 + n]; err = fmax(err, fabs(r - D[m * 8 + n])); }
    printf(", \"m16n8k8_layout_max_err\": %.3e", err); }
  printf("}\n");
  return 0;
}
