// Device probes used by bench.py for its roofline denominators: the fp32 FMA pipe peak is not in
// MEASURED_PEAKS.json (driver-written: HBM copy bandwidth and cuBLAS bf16 only), so it is measured
// live on the device the bench runs on.  Not part of the hot path.
#include "common.cuh"

using namespace kws;

namespace {
constexpr int kProbeIters = 4096;
// 8 independent packed f32x2 FMA chains per thread: FFMA2 issues 64 fp32 FMAs per warp instruction
__global__ void __launch_bounds__(512) probe_ffma2_kernel(float* out, float a, float b) {
  float2 r[8];
  const float2 A = make_float2(a, a * 1.01f), Bv = make_float2(b, b * 0.99f);
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = make_float2((float)(threadIdx.x + i), (float)((int)threadIdx.x - i));
  for (int it = 0; it < kProbeIters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = __ffma2_rn(r[i], A, Bv);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += r[i].x + r[i].y;
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(512) probe_ffma_kernel(float* out, float a, float b) {
  float r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = (float)(threadIdx.x + i);
  for (int it = 0; it < kProbeIters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = fmaf(r[i], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += r[i];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace

extern "C" int kws_probe_fp32_tflops(kws_handle_t h, double* ffma_tflops, double* ffma2_tflops) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  KWS_CUDA(cudaSetDevice(h->device));
  const int threads = 512, blocks = h->sm_count * 4;
  float* out = nullptr;
  KWS_CUDA(cudaMalloc((void**)&out, (size_t)blocks * threads * 4));
  cudaEvent_t e0, e1;
  KWS_CUDA(cudaEventCreate(&e0));
  KWS_CUDA(cudaEventCreate(&e1));
  double res[2] = {0, 0};
  for (int v = 0; v < 2; ++v) {
    float best = 1e30f;
    for (int rep = 0; rep < 6; ++rep) {
      KWS_CUDA(cudaEventRecord(e0, 0));
      if (v == 0) probe_ffma_kernel<<<blocks, threads>>>(out, 1.0001f, 0.5f);
      else probe_ffma2_kernel<<<blocks, threads>>>(out, 1.0001f, 0.5f);
      KWS_CUDA(cudaEventRecord(e1, 0));
      KWS_CUDA(cudaEventSynchronize(e1));
      float ms = 0.f;
      KWS_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      if (rep >= 2 && ms < best) best = ms;  // two warm-up launches
    }
    const double fmas = (double)blocks * threads * kProbeIters * 8 * (v == 0 ? 1 : 2);
    res[v] = 2.0 * fmas / (best * 1e-3) * 1e-12;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  if (ffma_tflops) *ffma_tflops = res[0];
  if (ffma2_tflops) *ffma2_tflops = res[1];
  return KWS_OK;
}
