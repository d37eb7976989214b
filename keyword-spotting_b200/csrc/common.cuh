// Internal declarations shared by the translation units of libkws_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "kws_b200.h"

namespace kws {

// (clip, row, column) of a linear NHWC pixel index.  32-bit unsigned division when the index fits (it does for
// every realistic batch: 64-bit integer division costs ~100 instructions and used to dominate the per-tile
// staging of the training kernels), 64-bit otherwise.
__host__ __device__ __forceinline__ void split_pixel(long long gp, int Wd, int Hd, long long* b, int* row, int* col) {
  if (gp < 0x7fffffffLL) {
    const unsigned g = (unsigned)gp, r = g / (unsigned)Wd;
    *col = (int)(g - r * (unsigned)Wd);
    const unsigned bb = r / (unsigned)Hd;
    *row = (int)(r - bb * (unsigned)Hd);
    *b = (long long)bb;
  } else {
    const long long r = gp / Wd;
    *col = (int)(gp - r * Wd);
    *b = r / Hd;
    *row = (int)(r - *b * Hd);
  }
}


void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);

#define KWS_CUDA(expr)                                                                      \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return ::kws::fail(KWS_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                         __FILE__, __LINE__);                                               \
  } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is sticky per (function, device): issue it only the first
// time, or when a launch needs more than was granted before (capi.cu keeps the table)
int ensure_dyn_smem_impl(const void* func, size_t bytes, bool max_carveout);
template <typename K>
inline int ensure_dyn_smem(K kern, size_t bytes, bool max_carveout = false) {
  return ensure_dyn_smem_impl(reinterpret_cast<const void*>(kern), bytes, max_carveout);
}

#define KWS_CHECK_LAUNCH(h)                                                       \
  do {                                                                            \
    cudaError_t _e = cudaGetLastError();                                          \
    if (_e != cudaSuccess)                                                        \
      return ::kws::fail(KWS_ERR_CUDA, "kernel launch failed: %s (%s:%d)",        \
                         cudaGetErrorString(_e), __FILE__, __LINE__);             \
    (h)->launches++;                                                              \
  } while (0)

// ---- feature tables -------------------------------------------------------------------
constexpr int kNfft = 512;          // fast path: 512-point real FFT as a 256-point complex FFT
constexpr int kBins = kNfft / 2 + 1;  // 257
constexpr int kMaxFilt = 80;        // scratch budget of the fused kernel (see mfcc.cu)
constexpr int kFramesPerTile = 20;  // frames one CTA pass consumes (5 warps x 4 frames)

struct FeatureTables {
  bool ready = false;
  kws_feature_config cfg{};
  int ncep = 0;              // min(numcep, nfilt)
  int ncep_pad = 0;          // DCT matrix row pitch
  int load_len = 0;          // min(frame_len, nfft): samples of a frame that enter the FFT
  // device tables
  float* d_window = nullptr;    // [frame_len] or null
  float2* d_tw256 = nullptr;    // W256^t, t = 0..255 (cos, -sin)
  float2* d_tw512 = nullptr;    // W512^t, t = 0..255
  float2* d_melw = nullptr;     // [257] (rising weight, falling weight) of the bin's segment
  int* d_segedge = nullptr;     // [nfilt+2] bin edges
  float* d_dct = nullptr;       // [nfilt][ncep_pad], lifter folded in
  float* d_dctB = nullptr;      // tensor-core DCT B operand image (hi, lo), null when nfilt > 48
  int dct_kpad = 0;             // K of that GEMM: nfilt rounded up to 8
  std::vector<int> h_edges;
};

// ---- model ----------------------------------------------------------------------------
struct ModelWeights {
  bool created = false, finalized = false;
  int num_classes = 0, channels = 0, num_blocks = 0;
  std::vector<std::vector<float>> host;   // staged tensors by slot
  std::vector<std::string> names;
  std::vector<bool> loaded;
  // packed device weights
  float* d_conv1_w = nullptr;   // [100][C]   (tap-major, channel-minor)
  float* d_conv1_b = nullptr;   // [C]
  float* d_dw_w = nullptr;      // [nb][9][C]
  float* d_dw_b = nullptr;      // [nb][C]
  float* d_pw_w = nullptr;      // [nb][Cin][Cout]  (k-major)
  float* d_pw_b = nullptr;      // [nb][C]
  float* d_pw_beff = nullptr;   // [nb][C]   pointwise bias with the depthwise bias folded through
  float* d_fc_w = nullptr;      // [C][num_classes]
  float* d_fc_b = nullptr;      // [num_classes]
  // fused kernel images (model_fused.cu), built when channels == 64 and num_blocks == 4
  float* d_fused_a = nullptr;       // tf32 A-operand images: conv1, then the four pointwise layers
  float* d_fused_dw = nullptr;      // [4][9][64]
  float* d_fused_actb = nullptr;    // [5][64]
  float* d_fused_border = nullptr;  // [5][64]
  int path = 0;                     // 0 = auto (fused / tensor-core GEMM when applicable), 1 = layered fp32 kernels
  void* gemm = nullptr;             // kws::GemmModel (model_gemm.cu): chunked tf32 weight images for any channel count
  // workspace
  void* d_ws = nullptr;
  size_t ws_bytes = 0;
};

// ---- training step workspace (train.cu) -------------------------------------------------
struct TrainWorkspace {
  void* base = nullptr;
  size_t bytes = 0;
  float* L[5] = {};      // saved activations: conv1 output and the four block outputs (post-ReLU), NHWC
  float* D[4] = {};      // saved depthwise outputs (pointwise inputs); overwritten by their gradients
  float* dX[2] = {};     // ping-pong gradient w.r.t. a block input
  float* pooled = nullptr;   // [B, C]
  float* dpooled = nullptr;  // [B, C]
  float* logits = nullptr;   // [B, K]
};

}  // namespace kws

struct kws_handle_s {
  int device = 0;
  int sm_count = 0;
  int64_t launches = 0;
  kws::FeatureTables feat;
  kws::ModelWeights model;
  // staging for the *_host entry points
  cudaStream_t s_copy[2] = {nullptr, nullptr};
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  void* d_stage_in[2] = {nullptr, nullptr};
  void* d_stage_out[2] = {nullptr, nullptr};
  size_t stage_in_bytes = 0, stage_out_bytes = 0;
  float* d_feat_tmp = nullptr;   // features scratch of the two-kernel pcm->logits path (small batches)
  size_t feat_tmp_bytes = 0;
  // single-launch pcm->logits (pcm_logits_fused.cu): per-CTA, L2-resident feature scratch (one chunk of clips each)
  float* d_ring = nullptr;
  size_t ring_bytes = 0;
  int pipe_chunk_clips = 32;     // clips a CTA transforms before it runs the network on them
  int64_t pipe_min_clips = 1;    // smaller batches run the two kernels back to back (see capi.cu)
  kws::TrainWorkspace train;
  void* crnn = nullptr;          // kws::CrnnWeights, owned by crnn.cu
};

namespace kws {
// mfcc.cu
enum FeatureMode { kModeMfcc = 0, kModeLogMel = 1 };
int launch_mfcc(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N,
                int64_t pcm_stride, float* out_dev, int mode, cudaStream_t stream);
// mfcc_fast.cu: 1 = handled, 0 = configuration needs the generic kernel, <0 = error
int launch_mfcc_fast(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N, int64_t pcm_stride,
                     float* out_dev, int mode, cudaStream_t stream);
// model_fused.cu: 1 = handled, 0 = configuration outside the fused kernel, <0 = error
int launch_dscnn_fused(kws_handle_t h, const float* feats_dev, int64_t B, int64_t T, int64_t F, float* logits_dev,
                       cudaStream_t stream);
int build_fused_model(ModelWeights& m, const std::vector<float>& conv1_w, const std::vector<float>& conv1_b,
                      const std::vector<const float*>& dw_w, const std::vector<const float*>& dw_b,
                      const std::vector<const float*>& pw_w, const std::vector<const float*>& pw_b);
// pcm_logits_fused.cu: PCM -> logits in one launch (1 = handled, 0 = n/a, <0 = error)
int launch_pcm_logits_fused(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N, int64_t pcm_stride,
                            float* logits_dev, cudaStream_t stream);
// model.cu: layered forward that keeps every activation (training), logits/pooled into the workspace
int train_forward_layers(kws_handle_t h, const float* feats_dev, int64_t B, int T, int F, TrainWorkspace& w,
                         cudaStream_t st);
// train_tc.cu: pointwise + ReLU backward of one 64-channel block on the tensor core (1 = launched, <0 = error)
int launch_pw_bwd_tc(kws_handle_t h, const float* out, const float* dout, const float* dpooled, float inv_area, float* d_io,
                     const float* w_flat, float* g_w, float* g_b, long long B, int H, int W, cudaStream_t st);
// train_tc.cu: pointwise + bias + ReLU forward of one 64-channel block on the tensor core, from the saved depthwise output
int launch_pw_fwd_tc(kws_handle_t h, const float* d, float* out, const float* w_kmaj, const float* bias, long long B, int H,
                     int W, cudaStream_t st);
// model.cu: conv 10x10 s2 p2 (one input channel) + ReLU, NHWC output [B, H1, W1, C]
int launch_conv1_relu(kws_handle_t h, const float* in, float* out, const float* w_packed, const float* bias, int64_t B,
                      int T, int F, int H1, int W1, int C, cudaStream_t st);
// model_gemm.cu: layer-by-layer tcgen05 forward for any channel count (1 = handled, 0 = n/a, <0 = error)
int build_gemm_model(ModelWeights& m, const std::vector<float>& conv1_w, const std::vector<float>& conv1_b,
                     const std::vector<const float*>& dw_b, const std::vector<const float*>& pw_w,
                     const std::vector<const float*>& pw_b);
void free_gemm_model(void* gm);
int launch_dscnn_gemm(kws_handle_t h, const float* feats_dev, int64_t B, int T, int F, float* buf0, float* buf1,
                      float** last_out, int* last_rows, cudaStream_t st);
// tables.cu
int build_feature_tables(kws_handle_t h, const kws_feature_config* cfg);
void free_feature_tables(FeatureTables& t);
}  // namespace kws
