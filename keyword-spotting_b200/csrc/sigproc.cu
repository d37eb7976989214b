// Stand-alone signal-processing kernels behind the reference's torch helpers
// (kws/libs/speech_features/sigproc.py): preemphasis (:93-103), framesig (:14-52),
// magspec (:55-77), powspec (:80-90).  They follow the input dtype (float32 / float64)
// exactly like the torch code does.  The fused MFCC kernel (mfcc.cu) does not use these;
// they exist so the helper API is a drop-in on CUDA tensors.
#include <cuda_runtime.h>

#include "common.cuh"

namespace kws {

template <typename T> __device__ __forceinline__ T mul_rn(T a, T b);
template <> __device__ __forceinline__ float mul_rn<float>(float a, float b) { return __fmul_rn(a, b); }
template <> __device__ __forceinline__ double mul_rn<double>(double a, double b) { return __dmul_rn(a, b); }
template <typename T> __device__ __forceinline__ T sub_rn(T a, T b);
template <> __device__ __forceinline__ float sub_rn<float>(float a, float b) { return __fsub_rn(a, b); }
template <> __device__ __forceinline__ double sub_rn<double>(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ void sincospi_t(float x, float* s, float* c) { sincospif(x, s, c); }
__device__ __forceinline__ void sincospi_t(double x, double* s, double* c) { sincospi(x, s, c); }

// out[b, 0] = in[b, 0]; out[b, n] = in[b, n] - coeff * in[b, n-1]   (product rounded first)
template <typename T>
__global__ void preemphasis_kernel(const T* __restrict__ in, T* __restrict__ out, long long B,
                                   long long N, T coeff) {
  const long long total = B * N;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long n = i % N;
    const T x = in[i];
    out[i] = (n == 0) ? x : sub_rn<T>(x, mul_rn<T>(coeff, in[i - 1]));
  }
}

// out[b, f, i] = (f*step + i < N ? in[b, f*step + i] : 0) * win[i]
template <typename T>
__global__ void framesig_kernel(const T* __restrict__ in, T* __restrict__ out, long long B,
                                long long N, int nframes, int frame_len, int frame_step,
                                const T* __restrict__ win) {
  const long long per_clip = (long long)nframes * frame_len;
  const long long total = B * per_clip;
  for (long long o = blockIdx.x * (long long)blockDim.x + threadIdx.x; o < total;
       o += (long long)gridDim.x * blockDim.x) {
    const long long b = o / per_clip;
    const int r = (int)(o - b * per_clip);
    const int f = r / frame_len, i = r - f * frame_len;
    const long long s = (long long)f * frame_step + i;
    T v = (s < N) ? in[b * N + s] : (T)0;
    if (win) v = v * win[i];
    out[o] = v;
  }
}

// One CTA per frame row.  Power-of-two nfft: shared-memory radix-2 FFT (bit-reversed load,
// twiddles from sincospi in the working precision).  Output |X| or |X|^2 / nfft.
template <typename T>
__global__ void spec_pow2_kernel(const T* __restrict__ frames, T* __restrict__ out, long long R,
                                 int L, int nfft, int log2n, int power) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* re = reinterpret_cast<T*>(smem_raw);
  T* im = re + nfft;
  const int nb = nfft / 2 + 1;
  for (long long row = blockIdx.x; row < R; row += gridDim.x) {
    const T* fr = frames + row * (long long)L;
    __syncthreads();
    for (int i = threadIdx.x; i < nfft; i += blockDim.x) {
      const int j = (int)(__brev((unsigned)i) >> (32 - log2n));
      re[j] = (i < L) ? fr[i] : (T)0;  // crop to nfft or right-zero-pad (sigproc.py:65-74)
      im[j] = (T)0;
    }
    __syncthreads();
    for (int s = 1; s <= log2n; ++s) {
      const int hm = 1 << (s - 1);
      for (int bfly = threadIdx.x; bfly < nfft / 2; bfly += blockDim.x) {
        const int pos = bfly & (hm - 1);
        const int i0 = ((bfly >> (s - 1)) << s) + pos, i1 = i0 + hm;
        T sn, cs;
        sincospi_t((T)pos / (T)hm, &sn, &cs);  // W = exp(-i pi pos / hm)
        const T xr = re[i1], xi = im[i1];
        const T tr = xr * cs + xi * sn, ti = xi * cs - xr * sn;
        const T ur = re[i0], ui = im[i0];
        re[i0] = ur + tr; im[i0] = ui + ti;
        re[i1] = ur - tr; im[i1] = ui - ti;
      }
      __syncthreads();
    }
    for (int k = threadIdx.x; k < nb; k += blockDim.x) {
      const T a = re[k], b = im[k];
      T v;
      if (power) v = (a * a + b * b) / (T)nfft;   // (1/NFFT) * |X|^2
      else v = sizeof(T) == 4 ? (T)hypotf((float)a, (float)b) : (T)hypot((double)a, (double)b);
      out[row * (long long)nb + k] = v;
    }
  }
}

// Any nfft: direct DFT, one CTA per row, thread per bin; exact twiddle phase by integer k*n mod nfft.
template <typename T>
__global__ void spec_dft_kernel(const T* __restrict__ frames, T* __restrict__ out, long long R, int L,
                                int nfft, int power) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* x = reinterpret_cast<T*>(smem_raw);
  const int nb = nfft / 2 + 1, len = L < nfft ? L : nfft;
  for (long long row = blockIdx.x; row < R; row += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < len; i += blockDim.x) x[i] = frames[row * (long long)L + i];
    __syncthreads();
    for (int k = threadIdx.x; k < nb; k += blockDim.x) {
      T a = 0, b = 0;
      int ph = 0;
      for (int n = 0; n < len; ++n) {
        T sn, cs;
        sincospi_t((T)2 * (T)ph / (T)nfft, &sn, &cs);
        a += x[n] * cs;
        b -= x[n] * sn;
        ph += k;
        if (ph >= nfft) ph -= nfft;
      }
      T v;
      if (power) v = (a * a + b * b) / (T)nfft;
      else v = sizeof(T) == 4 ? (T)hypotf((float)a, (float)b) : (T)hypot((double)a, (double)b);
      out[row * (long long)nb + k] = v;
    }
  }
}

static unsigned grid_for(long long total, int threads) {
  long long g = (total + threads - 1) / threads;
  const long long cap = 148LL * 32;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace kws

using namespace kws;

extern "C" {

// one launch counter for the handle-free sigproc entry points
static int64_t g_sigproc_launches = 0;
int64_t kws_sigproc_launch_count(void) { return g_sigproc_launches; }

#define SIG_CHECK_LAUNCH()                                                                     \
  do {                                                                                         \
    cudaError_t _e = cudaGetLastError();                                                       \
    if (_e != cudaSuccess) return fail(KWS_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(_e)); \
    g_sigproc_launches++;                                                                      \
  } while (0)

int kws_sigproc_preemphasis(int dtype, const void* in_dev, void* out_dev, int64_t B, int64_t N,
                            double coeff, void* stream) {
  if (dtype != 0 && dtype != 1) return fail(KWS_ERR_INVALID, "dtype must be 0 (f32) or 1 (f64)");
  if (B < 0 || N < 0) return fail(KWS_ERR_INVALID, "bad B/N");
  if (B * N == 0) return KWS_OK;
  if (!in_dev || !out_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = grid_for(B * N, 256);
  if (dtype == 0)
    preemphasis_kernel<float><<<grid, 256, 0, st>>>((const float*)in_dev, (float*)out_dev, B, N, (float)coeff);
  else
    preemphasis_kernel<double><<<grid, 256, 0, st>>>((const double*)in_dev, (double*)out_dev, B, N, coeff);
  SIG_CHECK_LAUNCH();
  return KWS_OK;
}

int kws_sigproc_framesig(int dtype, const void* in_dev, int64_t B, int64_t N, int64_t frame_len,
                         int64_t frame_step, const void* window_dev, void* out_dev, void* stream) {
  if (dtype != 0 && dtype != 1) return fail(KWS_ERR_INVALID, "dtype must be 0 (f32) or 1 (f64)");
  if (B < 0 || N < 0 || frame_len < 1 || frame_step < 1) return fail(KWS_ERR_INVALID, "bad framesig arguments");
  if (B == 0) return KWS_OK;
  if ((N > 0 && !in_dev) || !out_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  const int64_t nf = kws_num_frames(N, frame_len, frame_step);
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = grid_for(B * nf * frame_len, 256);
  if (dtype == 0)
    framesig_kernel<float><<<grid, 256, 0, st>>>((const float*)in_dev, (float*)out_dev, B, N, (int)nf,
                                                 (int)frame_len, (int)frame_step, (const float*)window_dev);
  else
    framesig_kernel<double><<<grid, 256, 0, st>>>((const double*)in_dev, (double*)out_dev, B, N, (int)nf,
                                                  (int)frame_len, (int)frame_step, (const double*)window_dev);
  SIG_CHECK_LAUNCH();
  return KWS_OK;
}

int kws_sigproc_spec(int dtype, const void* frames_dev, int64_t R, int64_t L, int64_t nfft, int power,
                     void* out_dev, void* stream) {
  if (dtype != 0 && dtype != 1) return fail(KWS_ERR_INVALID, "dtype must be 0 (f32) or 1 (f64)");
  if (R < 0 || L < 1 || nfft < 1) return fail(KWS_ERR_INVALID, "bad spec arguments");
  if (R == 0) return KWS_OK;
  if (!frames_dev || !out_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t es = dtype == 0 ? 4 : 8;
  const unsigned grid = (unsigned)(R < 148 * 16 ? R : 148 * 16);
  const bool pow2 = (nfft & (nfft - 1)) == 0 && nfft >= 2;
  if (pow2 && (size_t)nfft * 2 * es <= 200 * 1024) {
    int lg = 0;
    while ((1LL << lg) < nfft) ++lg;
    const size_t sm = (size_t)nfft * 2 * es;
    const int threads = nfft / 2 < 64 ? 64 : (nfft / 2 > 512 ? 512 : (int)nfft / 2);
    if (dtype == 0) {
      if (int rc_ = ensure_dyn_smem(spec_pow2_kernel<float>, (size_t)sm)) return rc_;
      spec_pow2_kernel<float><<<grid, threads, sm, st>>>((const float*)frames_dev, (float*)out_dev, R, (int)L,
                                                         (int)nfft, lg, power);
    } else {
      if (int rc_ = ensure_dyn_smem(spec_pow2_kernel<double>, (size_t)sm)) return rc_;
      spec_pow2_kernel<double><<<grid, threads, sm, st>>>((const double*)frames_dev, (double*)out_dev, R, (int)L,
                                                          (int)nfft, lg, power);
    }
  } else {
    const size_t sm = (size_t)(L < nfft ? L : nfft) * es;
    if (sm > 200 * 1024) return fail(KWS_ERR_INVALID, "frame too long for the DFT kernel");
    if (dtype == 0) {
      if (int rc_ = ensure_dyn_smem(spec_dft_kernel<float>, (size_t)sm)) return rc_;
      spec_dft_kernel<float><<<grid, 256, sm, st>>>((const float*)frames_dev, (float*)out_dev, R, (int)L, (int)nfft, power);
    } else {
      if (int rc_ = ensure_dyn_smem(spec_dft_kernel<double>, (size_t)sm)) return rc_;
      spec_dft_kernel<double><<<grid, 256, sm, st>>>((const double*)frames_dev, (double*)out_dev, R, (int)L, (int)nfft, power);
    }
  }
  SIG_CHECK_LAUNCH();
  return KWS_OK;
}

}  // extern "C"
