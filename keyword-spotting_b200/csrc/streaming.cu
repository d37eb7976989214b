// Incremental features for sliding-window inference (the batch form of the reference's
// VAD-gated 1 s window loop, kws/inference/inference_local.py:114-188).
//
// Windows of N samples taken every `hop` samples (hop a multiple of the frame step) share all but two of their
// frames with the recording's own frame grid: frame f of window w is recording frame w*k + f (k = hop / frame_step)
// for 1 <= f <= T-2.  Only the two EDGE frames of a window differ from the recording's:
//   frame 0    starts the window: psf's pre-emphasis has no predecessor there (y[0] = x[0]);
//   frame T-1  runs past the window end: samples >= N are zero AFTER pre-emphasis (psf pads y, not x).
// So the recording is transformed ONCE (kws_mfcc on the whole recording as one long clip), the 2*nwin edge frames
// are pre-emphasised here into a dense [2*nwin, frame_len] buffer and transformed by a second feature plan with
// frame_step = frame_len and preemph = 0 (every row is exactly one frame), and the per-window feature maps the
// network reads are assembled by a gather.  MFCC work per window drops from T frames to k + 2.
#include <cuda_runtime.h>

#include "common.cuh"

namespace kws {

// out[2w]   = pre-emphasised first frame of window w, out[2w+1] = pre-emphasised, zero-tailed last frame.
// fp32 with the product rounded before the subtraction: psf's float32 pre-emphasis step, bit for bit.
__global__ void edge_frames_kernel(const float* __restrict__ rec, long long nwin, long long hop, int N, int frame_len,
                                   int last_start, float pre, float* __restrict__ out) {
  const long long total = 2 * nwin * frame_len;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long e = idx / frame_len;
    const int i = (int)(idx - e * frame_len);
    const long long w = e >> 1;
    const float* x = rec + w * hop;
    float v;
    if ((e & 1) == 0) {
      v = (i == 0) ? x[0] : __fsub_rn(x[i], __fmul_rn(pre, x[i - 1]));
    } else {
      const int g = last_start + i;   // window-local sample index
      v = (g < N) ? (g == 0 ? x[0] : __fsub_rn(x[g], __fmul_rn(pre, x[g - 1]))) : 0.f;
    }
    out[idx] = v;
  }
}

// win[w][f][:] = f == 0 ? edge[2w] : f == T-1 ? edge[2w+1] : rec_feats[w*k + f]      (rows of F floats)
__global__ void window_gather_kernel(const float* __restrict__ rec_feats, const float* __restrict__ edge, long long nwin,
                                     int T, int F, long long k, float* __restrict__ win) {
  const long long total = nwin * (long long)T * F;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long row = idx / F;
    const int c = (int)(idx - row * F);
    const long long w = row / T;
    const int f = (int)(row - w * T);
    float v;
    if (f == 0) v = edge[(2 * w) * F + c];
    else if (f == T - 1) v = edge[(2 * w + 1) * F + c];
    else v = rec_feats[(w * k + f) * F + c];
    win[idx] = v;
  }
}

}  // namespace kws

using namespace kws;

extern "C" {

int kws_stream_edge_frames(kws_handle_t h, const float* rec_dev, int64_t L, int64_t nwin, int64_t hop, int64_t N,
                           int64_t frame_len, int64_t frame_step, float preemph, float* out_dev, void* stream) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (nwin < 0 || hop < 1 || N < frame_len || frame_len < 1 || frame_step < 1 || N > (1ll << 30))
    return fail(KWS_ERR_INVALID, "bad window geometry");
  if (nwin == 0) return KWS_OK;
  if (!rec_dev || !out_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  if ((nwin - 1) * hop + N > L) return fail(KWS_ERR_INVALID, "windows run past the end of the recording");
  KWS_CUDA(cudaSetDevice(h->device));
  const int64_t T = kws_num_frames(N, frame_len, frame_step);
  const int last_start = (int)((T - 1) * frame_step);
  long long g = (2 * nwin * frame_len + 255) / 256;
  if (g > h->sm_count * 32LL) g = h->sm_count * 32LL;
  edge_frames_kernel<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>(rec_dev, nwin, hop, (int)N, (int)frame_len, last_start,
                                                                   preemph, out_dev);
  KWS_CHECK_LAUNCH(h);
  return KWS_OK;
}

int kws_stream_gather_windows(kws_handle_t h, const float* rec_feats_dev, int64_t rec_frames, const float* edge_feats_dev,
                              int64_t nwin, int64_t T, int64_t F, int64_t hop_frames, float* win_dev, void* stream) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (nwin < 0 || T < 2 || F < 1 || hop_frames < 1) return fail(KWS_ERR_INVALID, "bad window geometry");
  if (nwin == 0) return KWS_OK;
  if (!rec_feats_dev || !edge_feats_dev || !win_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  if (T > 2 && (nwin - 1) * hop_frames + (T - 2) >= rec_frames)
    return fail(KWS_ERR_INVALID, "windows need recording frame %lld, the recording has %lld",
                (long long)((nwin - 1) * hop_frames + T - 2), (long long)rec_frames);
  KWS_CUDA(cudaSetDevice(h->device));
  long long g = (nwin * T * F + 255) / 256;
  if (g > h->sm_count * 32LL) g = h->sm_count * 32LL;
  window_gather_kernel<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>(rec_feats_dev, edge_feats_dev, nwin, (int)T, (int)F,
                                                                     hop_frames, win_dev);
  KWS_CHECK_LAUNCH(h);
  return KWS_OK;
}

}  // extern "C"
