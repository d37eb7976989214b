// Pipe-throughput microbenchmarks for B200 (sm_100a): the denominators behind the MFCC kernel's
// design (fp32 FMA scalar vs packed f32x2, shared-memory LDS.64/128, warp shuffle).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu && ./microbench
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
constexpr int ITERS = 4096;

__global__ void k_ffma(float* out, float a, float b) {
  float r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = fmaf(r[i], a, b);
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_fadd(float* out, float a) {
  float r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = r[i] + a;
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma2(float* out, float a, float b) {
  float2 r[8];
  const float2 A = make_float2(a, a * 1.01f), Bv = make_float2(b, b * 0.99f);
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = __ffma2_rn(r[i], A, Bv);
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += r[i].x + r[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_fadd2(float* out, float a) {
  float2 r[8];
  const float2 A = make_float2(a, a * 1.01f);
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = __fadd2_rn(r[i], A);
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += r[i].x + r[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// mixed: 1 FFMA2 + 1 FADD2 per slot pair, like a butterfly network
__global__ void k_mix2(float* out, float a, float b) {
  float2 r[8];
  const float2 A = make_float2(a, a * 1.01f), Bv = make_float2(b, b * 0.99f);
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; i += 2) { r[i] = __ffma2_rn(r[i], A, Bv); r[i + 1] = __fadd2_rn(r[i + 1], A); }
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += r[i].x + r[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename T>
__global__ void k_lds(float* out) {
  extern __shared__ __align__(16) unsigned char sm[];
  T* s = reinterpret_cast<T*>(sm);
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = i;
  __syncthreads();
  float acc = 0;
  int idx = threadIdx.x;
  const int n = 16384 / sizeof(T);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      T v = s[(idx + i * 32) & (n - 1)];
      acc += *reinterpret_cast<float*>(&v);
    }
    idx += 1;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
__global__ void k_shfl(float* out) {
  float r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = __shfl_xor_sync(0xffffffffu, r[i], 1 + (i & 3));
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float timeit(F f) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  f(); f(); CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a)); for (int i = 0; i < 5; ++i) f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
  float ms; CK(cudaEventElapsedTime(&ms, a, b)); return ms / 5;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int clk_khz = 0; CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  const int sms = p.multiProcessorCount, threads = 512, blocks = sms * 4;
  float* out; CK(cudaMalloc(&out, (size_t)blocks * threads * 4));
  const double n_thread_ops = (double)blocks * threads * ITERS * 8;
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_mhz_attr\": %d", p.name, sms, clk_khz / 1000);
  float ms;
  ms = timeit([&] { k_ffma<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
  printf(", \"ffma_tflops\": %.2f, \"ffma_lane_ops_per_clk_sm_at_max\": %.1f", 2 * n_thread_ops / ms * 1e-9, n_thread_ops / (ms * 1e-3) / sms / (clk_khz * 1e3));
  ms = timeit([&] { k_fadd<<<blocks, threads>>>(out, 1.0001f); });
  printf(", \"fadd_gops\": %.1f", n_thread_ops / ms * 1e-6);
  ms = timeit([&] { k_ffma2<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
  printf(", \"ffma2_tflops\": %.2f", 4 * n_thread_ops / ms * 1e-9);
  ms = timeit([&] { k_fadd2<<<blocks, threads>>>(out, 1.0001f); });
  printf(", \"fadd2_gflops\": %.1f", 2 * n_thread_ops / ms * 1e-6);
  ms = timeit([&] { k_mix2<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
  printf(", \"mix_ffma2_fadd2_ginstr\": %.1f", n_thread_ops / ms * 1e-6);
  ms = timeit([&] { k_lds<float><<<blocks, threads, 16384>>>(out); });
  printf(", \"lds32_TBps\": %.2f", n_thread_ops * 4 / ms * 1e-9);
  ms = timeit([&] { k_lds<float2><<<blocks, threads, 16384>>>(out); });
  printf(", \"lds64_TBps\": %.2f", n_thread_ops * 8 / ms * 1e-9);
  ms = timeit([&] { k_lds<float4><<<blocks, threads, 16384>>>(out); });
  printf(", \"lds128_TBps\": %.2f", n_thread_ops * 16 / ms * 1e-9);
  ms = timeit([&] { k_shfl<<<blocks, threads>>>(out); });
  printf(", \"shfl_ginstr_lane\": %.1f}\n", n_thread_ops / ms * 1e-6);
  return 0;
}
