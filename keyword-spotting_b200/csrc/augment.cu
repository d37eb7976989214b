// GPU augmentation: the stages AudioProcessor.transform applies between loading a
// clip and extract_features (kws/libs/audio_processor.py:150-159):
//   silence label -> zeros (:151-152); _apply_time_shift (:172-188): out[i] = x[i - s], zero fill;
//   _add_background_noise (:190-233): out += noise[(offset + i) mod len] * volume   (np.tile semantics)
// The reference draws shift / noise clip / offset / volume from the global numpy & random generators per
// clip; here the caller supplies the draws as per-clip arrays, so the kernel is a pure function and is
// bit-exact against the numpy restatement: float32 product rounded, then float32 sum (numpy's float32
// array times a Python float stays float32).  Element-wise, HBM-bound: 8 B read + 4 B written per sample.
#include <cuda_runtime.h>

#include "common.cuh"

namespace kws {

__global__ void augment_kernel(const float* __restrict__ pcm, long long B, int N, long long stride,
                               const int* __restrict__ shift, const unsigned char* __restrict__ silence,
                               const float* __restrict__ bank, const long long* __restrict__ nstart,
                               const long long* __restrict__ nlen, const long long* __restrict__ noff,
                               const float* __restrict__ nvol, float* __restrict__ out) {
  const long long total = B * (long long)N;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long b = idx / N;
    const int i = (int)(idx - b * N);
    float v = 0.f;
    if (!(silence && silence[b])) {
      const int src = i - (shift ? shift[b] : 0);
      if (src >= 0 && src < N) v = pcm[b * stride + src];
    }
    if (bank) {
      const long long len = nlen[b];
      if (len > 0) {
        const float vol = nvol[b];
        const float nz = bank[nstart[b] + (noff[b] + i) % len];
        v = __fadd_rn(v, __fmul_rn(nz, vol));
      }
    }
    out[idx] = v;
  }
}

}  // namespace kws

using namespace kws;

extern "C" int kws_augment(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                           const int32_t* shift_dev, const uint8_t* silence_dev, const float* noise_bank_dev,
                           const int64_t* noise_start_dev, const int64_t* noise_len_dev, const int64_t* noise_offset_dev,
                           const float* noise_volume_dev, float* out_dev, void* stream) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (B < 0 || N < 1 || N > (1ll << 30) || pcm_stride < N) return fail(KWS_ERR_INVALID, "bad B/N/stride");
  if (B == 0) return KWS_OK;
  if (!pcm_dev || !out_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  if (noise_bank_dev && (!noise_start_dev || !noise_len_dev || !noise_offset_dev || !noise_volume_dev))
    return fail(KWS_ERR_INVALID, "noise bank given without per-clip start/len/offset/volume");
  KWS_CUDA(cudaSetDevice(h->device));
  long long g = (B * N + 255) / 256;
  if (g > h->sm_count * 32LL) g = h->sm_count * 32LL;
  augment_kernel<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>(
      pcm_dev, B, (int)N, pcm_stride, shift_dev, silence_dev, noise_bank_dev, (const long long*)noise_start_dev,
      (const long long*)noise_len_dev, (const long long*)noise_offset_dev, noise_volume_dev, out_dev);
  KWS_CHECK_LAUNCH(h);
  return KWS_OK;
}
