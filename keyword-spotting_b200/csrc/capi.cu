// extern "C" surface of libkws_b200.so: handle lifetime, feature entry points and the
// host-buffer (end-to-end) variants.  See include/kws_b200.h for the reference interfaces
// each function replaces.
#include <cstdlib>
#include <algorithm>
#include <cstring>

#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"

using namespace kws;

namespace kws {
int ensure_dyn_smem_impl(const void* func, size_t bytes, bool max_carveout) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, size_t> granted;
  int dev = 0;
  KWS_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  size_t& g = granted[{func, dev}];
  if (bytes <= g) return KWS_OK;
  KWS_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  if (max_carveout)
    KWS_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  g = bytes;
  return KWS_OK;
}
}  // namespace kws

namespace {

// clips per H2D/compute/D2H pipeline stage of the *_host entry points: small enough that the un-overlapped first
// copy and last compute are a small part of a call, large enough to fill the persistent kernels (1536 clips =
// 21 tiles per SM in the MFCC kernel).  Measured on one B200, 32768-clip calls: 4096 -> 811 K clips/s,
// 2048 -> 828 K, 1536 -> 840 K, 1024 -> 831 K (float32 PCM -> MFCC).  KWS_HOST_CHUNK overrides.
static int64_t host_chunk_clips() {
  static const int64_t v = [] {
    const char* e = std::getenv("KWS_HOST_CHUNK");
    const long long x = e ? std::atoll(e) : 0;
    return (int64_t)(x >= 64 && x <= (1 << 20) ? x : 1536);
  }();
  return v;
}

int ensure_streams(kws_handle_t h) {
  for (int i = 0; i < 2; ++i)
    if (!h->s_copy[i]) KWS_CUDA(cudaStreamCreateWithFlags(&h->s_copy[i], cudaStreamNonBlocking));
  return KWS_OK;
}

int ensure_stage(kws_handle_t h, size_t in_bytes, size_t out_bytes) {
  if (in_bytes > h->stage_in_bytes) {
    for (int i = 0; i < 2; ++i) {
      if (h->d_stage_in[i]) cudaFree(h->d_stage_in[i]);
      h->d_stage_in[i] = nullptr;
      KWS_CUDA(cudaMalloc(&h->d_stage_in[i], in_bytes));
    }
    h->stage_in_bytes = in_bytes;
  }
  if (out_bytes > h->stage_out_bytes) {
    for (int i = 0; i < 2; ++i) {
      if (h->d_stage_out[i]) cudaFree(h->d_stage_out[i]);
      h->d_stage_out[i] = nullptr;
      KWS_CUDA(cudaMalloc(&h->d_stage_out[i], out_bytes));
    }
    h->stage_out_bytes = out_bytes;
  }
  return KWS_OK;
}

}  // namespace

extern "C" {

int kws_create(int device, kws_handle_t* out) {
  if (!out) return fail(KWS_ERR_INVALID, "null handle pointer");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(KWS_ERR_CUDA, "no CUDA device (%s); kws_b200 has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail(KWS_ERR_INVALID, "device %d out of range [0,%d)", device, ndev);
  KWS_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  KWS_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(KWS_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device,
                prop.major, prop.minor);
  kws_handle_t h = new (std::nothrow) kws_handle_s();
  if (!h) return fail(KWS_ERR_ALLOC, "out of host memory");
  h->device = device;
  h->sm_count = prop.multiProcessorCount;
  *out = h;
  return KWS_OK;
}

void kws_model_free_internal(kws_handle_t h);  // model.cu
void kws_crnn_free_internal(kws_handle_t h);   // crnn.cu

int kws_destroy(kws_handle_t h) {
  if (!h) return KWS_OK;
  cudaSetDevice(h->device);
  free_feature_tables(h->feat);
  kws_model_free_internal(h);
  kws_crnn_free_internal(h);
  for (int i = 0; i < 2; ++i) {
    if (h->s_copy[i]) cudaStreamDestroy(h->s_copy[i]);
    if (h->d_stage_in[i]) cudaFree(h->d_stage_in[i]);
    if (h->d_stage_out[i]) cudaFree(h->d_stage_out[i]);
  }
  for (int i = 0; i < 4; ++i)
    if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  if (h->d_feat_tmp) cudaFree(h->d_feat_tmp);
  if (h->d_ring) cudaFree(h->d_ring);
  if (h->train.base) cudaFree(h->train.base);
  delete h;
  return KWS_OK;
}

int64_t kws_launch_count(kws_handle_t h) { return h ? h->launches : -1; }

int kws_features_configure(kws_handle_t h, const kws_feature_config* cfg) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  KWS_CUDA(cudaSetDevice(h->device));
  return build_feature_tables(h, cfg);
}

int kws_features_ncep(kws_handle_t h) {
  if (!h || !h->feat.ready) return fail(KWS_ERR_STATE, "features not configured");
  return h->feat.ncep;
}

int kws_mfcc(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
             float* out_dev, void* stream) {
  return launch_mfcc(h, pcm_dev, false, B, N, pcm_stride, out_dev, kModeMfcc, (cudaStream_t)stream);
}

int kws_logfbank(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                 float* out_dev, void* stream) {
  return launch_mfcc(h, pcm_dev, false, B, N, pcm_stride, out_dev, kModeLogMel, (cudaStream_t)stream);
}

int kws_logfbank_i16(kws_handle_t h, const int16_t* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                     float* out_dev, void* stream) {
  return launch_mfcc(h, pcm_dev, true, B, N, pcm_stride, out_dev, kModeLogMel, (cudaStream_t)stream);
}

int kws_mfcc_i16(kws_handle_t h, const int16_t* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                 float* out_dev, void* stream) {
  return launch_mfcc(h, pcm_dev, true, B, N, pcm_stride, out_dev, kModeMfcc, (cudaStream_t)stream);
}

static int mfcc_host_impl(kws_handle_t h, const void* pcm_host, bool is_i16, int64_t B, int64_t N, float* out_host) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (!h->feat.ready) return fail(KWS_ERR_STATE, "features not configured");
  if (B < 0 || N < 1) return fail(KWS_ERR_INVALID, "bad B/N");
  if (B == 0) return KWS_OK;
  if (!pcm_host || !out_host) return fail(KWS_ERR_INVALID, "null host pointer");
  KWS_CUDA(cudaSetDevice(h->device));
  int rc = ensure_streams(h);
  if (rc) return rc;
  const size_t esz = is_i16 ? 2 : 4;
  const int64_t nframes = kws_num_frames(N, h->feat.cfg.frame_len, h->feat.cfg.frame_step);
  const int64_t ocols = h->feat.ncep;
  const int64_t chunk = std::min<int64_t>(B, host_chunk_clips());
  rc = ensure_stage(h, (size_t)chunk * N * 4, (size_t)chunk * nframes * ocols * 4);
  if (rc) return rc;
  int i = 0;
  for (int64_t b0 = 0; b0 < B; b0 += chunk, ++i) {
    const int64_t nb = std::min<int64_t>(chunk, B - b0);
    const int s = i & 1;
    cudaStream_t st = h->s_copy[s];
    KWS_CUDA(cudaMemcpyAsync(h->d_stage_in[s], (const char*)pcm_host + (size_t)b0 * N * esz, (size_t)nb * N * esz,
                             cudaMemcpyHostToDevice, st));
    rc = launch_mfcc(h, h->d_stage_in[s], is_i16, nb, N, N, (float*)h->d_stage_out[s], kModeMfcc, st);
    if (rc) return rc;
    KWS_CUDA(cudaMemcpyAsync(out_host + b0 * nframes * ocols, h->d_stage_out[s],
                             (size_t)nb * nframes * ocols * 4, cudaMemcpyDeviceToHost, st));
  }
  KWS_CUDA(cudaStreamSynchronize(h->s_copy[0]));
  KWS_CUDA(cudaStreamSynchronize(h->s_copy[1]));
  return KWS_OK;
}

int kws_mfcc_host(kws_handle_t h, const float* pcm_host, int64_t B, int64_t N, float* out_host) {
  return mfcc_host_impl(h, pcm_host, false, B, N, out_host);
}

int kws_mfcc_host_i16(kws_handle_t h, const int16_t* pcm_host, int64_t B, int64_t N, float* out_host) {
  return mfcc_host_impl(h, pcm_host, true, B, N, out_host);
}


static int pcm_to_logits_impl(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N,
                             int64_t pcm_stride, float* logits_dev, void* stream) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (!h->feat.ready) return fail(KWS_ERR_STATE, "features not configured");
  if (!h->model.finalized) return fail(KWS_ERR_STATE, "model weights not finalized");
  if (B < 0 || N < 1) return fail(KWS_ERR_INVALID, "bad B/N");
  if (B == 0) return KWS_OK;
  KWS_CUDA(cudaSetDevice(h->device));
  if (!pcm_dev || !logits_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  if (pcm_stride < 1) return fail(KWS_ERR_INVALID, "bad stride");
  // one launch: every CTA alternates between the MFCC of its next chunk of clips and the network on that chunk, the
  // features stay in the CTA's L2-resident scratch (pcm_logits_fused.cu)
  if (B >= h->pipe_min_clips) {
    const int rc = launch_pcm_logits_fused(h, pcm_dev, is_i16, B, N, pcm_stride, logits_dev, (cudaStream_t)stream);
    if (rc != 0) return rc < 0 ? rc : KWS_OK;
  }
  const int64_t nframes = kws_num_frames(N, h->feat.cfg.frame_len, h->feat.cfg.frame_step);
  const int64_t F = h->feat.ncep;
  const size_t need = (size_t)B * nframes * F * 4;
  if (need > h->feat_tmp_bytes) {  // grows geometrically: a sweep of growing batches reallocates O(log) times
    const size_t want = h->feat_tmp_bytes ? std::max(need, std::min<size_t>(h->feat_tmp_bytes * 2, need + (need >> 1))) : need;
    if (h->d_feat_tmp) KWS_CUDA(cudaFree(h->d_feat_tmp));
    h->d_feat_tmp = nullptr;
    h->feat_tmp_bytes = 0;
    KWS_CUDA(cudaMalloc((void**)&h->d_feat_tmp, want));
    h->feat_tmp_bytes = want;
  }
  int rc = launch_mfcc(h, pcm_dev, is_i16, B, N, pcm_stride, h->d_feat_tmp, kModeMfcc, (cudaStream_t)stream);
  if (rc) return rc;
  return kws_model_forward(h, h->d_feat_tmp, B, nframes, F, logits_dev, stream);
}

int kws_pcm_to_logits_configure(kws_handle_t h, int64_t min_clips, int chunk_clips) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (chunk_clips < 0 || chunk_clips > 4096) return fail(KWS_ERR_INVALID, "chunk_clips must be in [0, 4096]");
  h->pipe_min_clips = min_clips < 1 ? (int64_t)1 << 62 : min_clips;   // < 1: never use the single-launch kernel
  if (chunk_clips > 0) h->pipe_chunk_clips = chunk_clips;
  return KWS_OK;
}

int kws_pcm_to_logits(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                      float* logits_dev, void* stream) {
  return pcm_to_logits_impl(h, pcm_dev, false, B, N, pcm_stride, logits_dev, stream);
}

int kws_pcm_to_logits_i16(kws_handle_t h, const int16_t* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                          float* logits_dev, void* stream) {
  return pcm_to_logits_impl(h, pcm_dev, true, B, N, pcm_stride, logits_dev, stream);
}

static int pcm_to_logits_host_impl(kws_handle_t h, const void* pcm_host, bool is_i16, int64_t B, int64_t N,
                                   float* logits_host) {
  const size_t esz = is_i16 ? 2 : 4;
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (!h->feat.ready) return fail(KWS_ERR_STATE, "features not configured");
  if (!h->model.finalized) return fail(KWS_ERR_STATE, "model weights not finalized");
  if (B < 0 || N < 1) return fail(KWS_ERR_INVALID, "bad B/N");
  if (B == 0) return KWS_OK;
  if (!pcm_host || !logits_host) return fail(KWS_ERR_INVALID, "null host pointer");
  KWS_CUDA(cudaSetDevice(h->device));
  int rc = ensure_streams(h);
  if (rc) return rc;
  const int64_t K = h->model.num_classes;
  const int64_t chunk = std::min<int64_t>(B, host_chunk_clips());
  rc = ensure_stage(h, (size_t)chunk * N * 4, (size_t)chunk * K * 4);
  if (rc) return rc;
  // the feature scratch and the model workspace are shared, so the two streams only overlap
  // copies with compute: compute of chunk i+1 waits for compute of chunk i.
  if (!h->ev[0]) KWS_CUDA(cudaEventCreateWithFlags(&h->ev[0], cudaEventDisableTiming));  // cached on the handle
  cudaEvent_t done = h->ev[0];
  int i = 0;
  for (int64_t b0 = 0; b0 < B; b0 += chunk, ++i) {
    const int64_t nb = std::min<int64_t>(chunk, B - b0);
    const int s = i & 1;
    cudaStream_t st = h->s_copy[s];
    KWS_CUDA(cudaMemcpyAsync(h->d_stage_in[s], (const char*)pcm_host + (size_t)b0 * N * esz, (size_t)nb * N * esz,
                             cudaMemcpyHostToDevice, st));
    if (i > 0) KWS_CUDA(cudaStreamWaitEvent(st, done, 0));
    rc = pcm_to_logits_impl(h, h->d_stage_in[s], is_i16, nb, N, N, (float*)h->d_stage_out[s], st);
    if (rc) return rc;
    KWS_CUDA(cudaEventRecord(done, st));
    KWS_CUDA(cudaMemcpyAsync(logits_host + b0 * K, h->d_stage_out[s], (size_t)nb * K * 4, cudaMemcpyDeviceToHost, st));
  }
  KWS_CUDA(cudaStreamSynchronize(h->s_copy[0]));
  KWS_CUDA(cudaStreamSynchronize(h->s_copy[1]));
  return KWS_OK;
}

int kws_pcm_to_logits_host(kws_handle_t h, const float* pcm_host, int64_t B, int64_t N, float* logits_host) {
  return pcm_to_logits_host_impl(h, pcm_host, false, B, N, logits_host);
}

int kws_pcm_to_logits_host_i16(kws_handle_t h, const int16_t* pcm_host, int64_t B, int64_t N, float* logits_host) {
  return pcm_to_logits_host_impl(h, pcm_host, true, B, N, logits_host);
}

}  // extern "C"
