/* kws_b200.h -- C ABI of libkws_b200.so: the B200 (sm_100a) hot path of z430/keyword-spotting.
 *
 * PCM -> MFCC (pre-emphasis, framing, window, 512-point real FFT, power spectrum, mel bank,
 * log, DCT-II, lifter) -> depthwise-separable KWS network -> logits.
 *
 * The reference has no FFI for this path: its boundary is three in-process Python call
 * sites.  Each entry point below names the reference interface it stands in for
 * (paths relative to the reference repository):
 *
 *   kws_features_*, kws_mfcc*      kws/libs/audio_processor.py:235-278  AudioProcessor.extract_features
 *                                  kws/libs/signal_handler.py:112-131   AudioProcessor.to_mfcc (legacy twin)
 *                                  -> python_speech_features==0.6 mfcc/fbank/logfbank (third party)
 *   kws_sigproc_*                  kws/libs/speech_features/sigproc.py:14-103  framesig/magspec/powspec/preemphasis
 *   kws_model_*                    kws/libs/models.py:75-183  DepthwiseSeparableConvBlock / DepthwiseSeparableConv
 *   kws_pcm_to_logits              audio_processor.py:235-278 followed by models.py:160-183 (fused, no HBM round trip)
 *
 * Conventions
 *   - plain C types only; every pointer named *_dev is a CUDA device pointer owned by the
 *     caller, *_host is host memory owned by the caller.  The library allocates nothing
 *     persistent except what hangs off an opaque handle.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - every function returns 0 on success and a negative kws_status otherwise;
 *     kws_last_error() returns a thread-local message.  CUDA errors never abort the process.
 *   - one host thread per handle; handles are independent (one per device / per stream).
 *   - there is NO CPU fallback: compute entry points fail with KWS_ERR_CUDA without a GPU.
 */
#ifndef KWS_B200_H_
#define KWS_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KWS_B200_VERSION 100 /* 0.1.0 */

typedef enum {
  KWS_OK = 0,
  KWS_ERR_INVALID = -1,     /* bad argument / unsupported configuration */
  KWS_ERR_CUDA = -2,        /* CUDA runtime error (message holds cudaGetErrorString) */
  KWS_ERR_STATE = -3,       /* call order: features not configured, weights missing ... */
  KWS_ERR_ALLOC = -4
} kws_status;

typedef struct kws_handle_s* kws_handle_t;

/* Feature parameters = the arguments python_speech_features.mfcc receives from
 * extract_features (audio_processor.py:270-278) plus the psf defaults it leaves alone. */
typedef struct {
  int32_t sample_rate;   /* 16000 */
  int32_t frame_len;     /* samples; psf round_half_up(winlen*sr) = 400 */
  int32_t frame_step;    /* samples; psf round_half_up(winstep*sr) = 160 */
  int32_t nfft;          /* 512 (fast path).  extract_features forces max(512,int(winlen*sr)) */
  int32_t nfilt;         /* mel filters M (26 default, 40 for "40-MFCC") */
  int32_t numcep;        /* C; output has min(C, M) columns like psf */
  float preemph;         /* 0.97; 0 disables */
  int32_t ceplifter;     /* 22; 0 disables */
  int32_t append_energy; /* 1: c0 := log(frame energy) */
  double lowfreq;        /* 0 */
  double highfreq;       /* <=0 means sample_rate/2 */
  const float* window_host; /* frame_len window values or NULL = rectangular (psf default) */
} kws_feature_config;

void kws_feature_config_default(kws_feature_config* cfg);

/* ---- library / handle ------------------------------------------------------------- */
int kws_version(void);
const char* kws_last_error(void);
int kws_create(int device, kws_handle_t* out);
int kws_destroy(kws_handle_t h);
/* number of this library's kernels launched through the handle since creation */
int64_t kws_launch_count(kws_handle_t h);

/* measurement aid (bench.py roofline): fp32 FMA throughput of the handle's device in TFLOP/s, scalar FFMA
 * and packed FFMA2 chains, best of four timed launches.  No reference counterpart. */
int kws_probe_fp32_tflops(kws_handle_t h, double* ffma_tflops, double* ffma2_tflops);

/* ---- host-side helpers (pure CPU arithmetic, usable without a GPU) ------------------ */
/* psf sigproc.round_half_up (decimal ROUND_HALF_UP) */
int64_t kws_round_half_up(double v);
/* frame count rule of psf framesig == sigproc.py:31-35 */
int64_t kws_num_frames(int64_t n, int64_t frame_len, int64_t frame_step);
/* nfilt+2 FFT-bin edges of psf get_filterbanks: floor((nfft+1)*mel2hz(linspace)/sr) */
int kws_mel_bin_edges(int nfilt, int nfft, int sample_rate, double lowfreq, double highfreq,
                      int32_t* edges_out /* [nfilt+2] */);
/* dense [nfilt, nfft/2+1] float64 filterbank of psf get_filterbanks */
int kws_mel_filterbank(int nfilt, int nfft, int sample_rate, double lowfreq, double highfreq,
                       double* fb_out);
/* orthonormal DCT-II matrix [nfilt, ncep] with the lifter folded in (float64) */
int kws_dct_lifter_matrix(int nfilt, int ncep, int ceplifter, double* out);

/* ---- features: replaces AudioProcessor.extract_features / to_mfcc ------------------- */
int kws_features_configure(kws_handle_t h, const kws_feature_config* cfg);
/* output columns = min(numcep, nfilt) */
int kws_features_ncep(kws_handle_t h);
/* pcm_dev: [B, N] float32 rows `pcm_stride` floats apart; out_dev: [B, nframes, ncep] float32.
 * Alignment: none required.  16-byte aligned pcm_dev / out_dev with pcm_stride, N and the hop multiples of
 * 4 samples (8 for int16) run on the TMA-fed tensor-core kernel; anything else takes the generic kernel.
 * Rows may overlap (1 <= pcm_stride < N): B sliding windows of one long recording, hop = pcm_stride, are
 * read in place -- the batch form of the reference's 1 s sliding-window inference concept
 * (kws/inference/inference_local.py:114-188).  The same holds for kws_pcm_to_logits. */
int kws_mfcc(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
             float* out_dev, void* stream);
/* psf logfbank ("log-mel"): out_dev [B, nframes, nfilt] */
int kws_logfbank(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                 float* out_dev, void* stream);
/* int16 PCM ingest (librosa.load scaling x/32768, audio_processor.py:145) fused in front */
int kws_logfbank_i16(kws_handle_t h, const int16_t* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                     float* out_dev, void* stream);
int kws_mfcc_i16(kws_handle_t h, const int16_t* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                 float* out_dev, void* stream);
/* end-to-end with HOST buffers: chunked H2D -> kernel -> D2H, double-buffered on two streams.
 * pcm_host / out_host should be pinned for full PCIe rate but need not be. */
int kws_mfcc_host(kws_handle_t h, const float* pcm_host, int64_t B, int64_t N, float* out_host);

/* same with int16 HOST buffers (the on-disk sample format of Speech Commands; halves the PCIe bytes) */
int kws_mfcc_host_i16(kws_handle_t h, const int16_t* pcm_host, int64_t B, int64_t N, float* out_host);

/* ---- augmentation: replaces the silence / time-shift / background-noise stages of
 * AudioProcessor.transform (kws/libs/audio_processor.py:150-159, 172-233).  The random draws are the
 * caller's (per-clip arrays), the kernel is a pure function: out[b,i] = (silence[b] ? 0 : x[b, i - shift[b]])
 * + bank[start[b] + (offset[b] + i) % len[b]] * volume[b]; NULL shift / silence / bank skip that stage. */
int kws_augment(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                const int32_t* shift_dev, const uint8_t* silence_dev, const float* noise_bank_dev,
                const int64_t* noise_start_dev, const int64_t* noise_len_dev, const int64_t* noise_offset_dev,
                const float* noise_volume_dev, float* out_dev, void* stream);

/* ---- incremental features for sliding windows (kws/inference/inference_local.py:114-188).
 * Windows of N samples every `hop` samples (hop a multiple of frame_step) share frames 1..T-2 with the recording's
 * own frame grid; only a window's first frame (no pre-emphasis predecessor) and last frame (zero tail after the
 * window end) are its own.  kws_stream_edge_frames writes those 2*nwin frames, pre-emphasised exactly like psf's
 * float32 step, as rows of frame_len floats (row 2w = first, 2w+1 = last frame of window w) for a feature plan with
 * frame_step = frame_len and preemph = 0; kws_stream_gather_windows assembles [nwin, T, F] feature maps from the
 * recording's features [rec_frames, F] (kws_mfcc on the whole recording) and the edge features [2*nwin, F]. */
int kws_stream_edge_frames(kws_handle_t h, const float* rec_dev, int64_t L, int64_t nwin, int64_t hop, int64_t N,
                           int64_t frame_len, int64_t frame_step, float preemph, float* out_dev, void* stream);
int kws_stream_gather_windows(kws_handle_t h, const float* rec_feats_dev, int64_t rec_frames, const float* edge_feats_dev,
                              int64_t nwin, int64_t T, int64_t F, int64_t hop_frames, float* win_dev, void* stream);

/* ---- sigproc: replaces kws/libs/speech_features/sigproc.py -------------------------- */
/* dtype: 0 = float32, 1 = float64 (the torch helpers follow the input dtype) */
int kws_sigproc_preemphasis(int dtype, const void* in_dev, void* out_dev, int64_t B, int64_t N,
                            double coeff, void* stream);
/* in [B,N] -> out [B, nframes, frame_len]; window_dev: frame_len values of `dtype` or NULL */
int kws_sigproc_framesig(int dtype, const void* in_dev, int64_t B, int64_t N, int64_t frame_len,
                         int64_t frame_step, const void* window_dev, void* out_dev, void* stream);
/* frames [R, L] -> [R, nfft/2+1]; power=0: |rfft|, power=1: |rfft|^2/nfft */
int kws_sigproc_spec(int dtype, const void* frames_dev, int64_t R, int64_t L, int64_t nfft,
                     int power, void* out_dev, void* stream);

/* kernels launched by the handle-free kws_sigproc_* entry points since load */
int64_t kws_sigproc_launch_count(void);

/* ---- model: replaces DepthwiseSeparableConv (models.py:122-183) --------------------- */
/* channels=64, num_blocks=4 is the reference network; other values are the "DS-CNN-L" style
 * extension that keeps every quirk (pointwise padding=1, no activation after depthwise). */
int kws_model_create(kws_handle_t h, int num_classes, int channels, int num_blocks);
/* name = reference state_dict key (models.py:135-144), data_host float32 in torch layout */
int kws_model_load_tensor(kws_handle_t h, const char* name, const float* data_host,
                          const int64_t* shape, int ndim);
int kws_model_finalize(kws_handle_t h);
/* kernel selection: 0 = auto (the fused tcgen05 kernel for the reference shape: 64 channels, 4 blocks,
 * F in {40, 10}; layered kernels otherwise), 1 = always the layered kernels (cross-check path) */
int kws_model_set_path(kws_handle_t h, int path);
/* feats_dev [B,1,T,F] float32 contiguous -> logits_dev [B,num_classes] */
int kws_model_forward(kws_handle_t h, const float* feats_dev, int64_t B, int64_t T, int64_t F,
                      float* logits_dev, void* stream);
/* optional epilogue: row-wise softmax of [B,K] logits (the reference's torch path never applies one:
 * logits go to CrossEntropyLoss / torch.max, kws/libs/training.py:293,301) */
int kws_softmax(kws_handle_t h, const float* logits_dev, int64_t B, int64_t K, float* out_dev, void* stream);
/* bytes of scratch the forward needs for batch B (allocated lazily on the handle) */
int64_t kws_model_workspace_bytes(kws_handle_t h, int64_t B, int64_t T, int64_t F);

/* ---- CRNN / GRU model: BASELINE config 4 ("log-mel + CRNN/GRU").  The reference has no recurrent model;
 * this is an extension in its idiom (a KeywordSpottingModel subclass, kws/libs/models.py:13-72): the
 * reference's conv1 (10x10 s2 p2 + ReLU, models.py:135,170) -> torch.nn.GRU(conv_channels * W1, hidden,
 * num_layers, batch_first) over the time axis -> fc on the last hidden state.  state_dict keys:
 * conv1.{weight,bias}, gru.{weight_ih,weight_hh,bias_ih,bias_hh}_l{k}, fc.{weight,bias}. -------------- */
int kws_crnn_create(kws_handle_t h, int num_classes, int conv_channels, int hidden, int num_layers, int n_features);
int kws_crnn_load_tensor(kws_handle_t h, const char* name, const float* data_host, const int64_t* shape, int ndim);
int kws_crnn_finalize(kws_handle_t h);
/* feats_dev [B,1,T,F] float32 (e.g. kws_logfbank output) -> logits_dev [B,num_classes] */
int kws_crnn_forward(kws_handle_t h, const float* feats_dev, int64_t B, int64_t T, int64_t F, float* logits_dev,
                     void* stream);

/* ---- training step: replaces the body of KWSTrainer.train_epoch (kws/libs/training.py:285-297;
 * train.py:39-49): forward, CrossEntropyLoss (mean), backward, Adam.  Reference network only
 * (64 channels, 4 blocks).  Parameters / gradients / Adam moments are caller-owned flat float32
 * device buffers in state_dict order (models.py:135-144, torch layouts), kws_train_num_params()
 * elements each, so the data-parallel exchange is one allreduce of the gradient buffer issued by
 * the caller between the two calls (grad_scale = 1/world_size folds the mean into Adam). ------- */
int64_t kws_train_num_params(kws_handle_t h);
/* feats_dev [B,1,T,F] float32, labels_dev [B] int64 -> grads_dev (overwritten), loss_dev[1] (mean CE),
 * logits_out_dev [B,num_classes] or NULL.  A label outside [0, num_classes) (torch raises there) contributes no
 * gradient and makes loss_dev NaN. */
int kws_train_forward_backward(kws_handle_t h, const float* params_dev, const float* feats_dev,
                               const int64_t* labels_dev, int64_t B, int64_t T, int64_t F, float* grads_dev,
                               float* loss_dev, float* logits_out_dev, void* stream);
/* torch.optim.Adam (weight_decay 0, amsgrad off) update number `step` (1-based) on the flat buffers */
int kws_train_adam(kws_handle_t h, float* params_dev, const float* grads_dev, float* m_dev, float* v_dev, int64_t n,
                   int64_t step, float lr, float beta1, float beta2, float eps, float grad_scale, void* stream);

/* ---- fused path: PCM -> logits ------------------------------------------------------
 * Replaces extract_features (audio_processor.py:235-278) + DepthwiseSeparableConv.forward (models.py:160-183)
 * back to back, as ONE kernel launch: every CTA alternates between the MFCC of its next chunk of clips (features into
 * the CTA's own L2-resident scratch) and the network on that chunk; HBM sees PCM in and logits out.  Configurations
 * the single-launch kernel does not cover (other channel counts, unaligned PCM, window functions ...) run the MFCC
 * kernel and the network kernel back to back.  Both give the same logits bit for bit. */
int kws_pcm_to_logits(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N,
                      int64_t pcm_stride, float* logits_dev, void* stream);
/* min_clips: smallest batch that takes the single launch (default 1; < 1 disables it: always two kernels);
 * chunk_clips: clips per CTA and phase (default 32; 0 keeps the current value) */
int kws_pcm_to_logits_configure(kws_handle_t h, int64_t min_clips, int chunk_clips);
/* The same single-launch PCM -> logits call for callers that cannot allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * It calls no cudaMalloc / cudaFree: the kernel's feature scratch is a 10 MiB static __device__ array that
 * every GPU loading the library holds until the process exits, cut into one slice per SM.  Limits: the single-launch
 * configurations only (reference network, 10 or 40 coefficients, aligned float32 PCM), and a clip's feature map must
 * fit a slice at least twice (four times at 10 coefficients): nframes * ncep * 4 bytes * 2 (or 4) * SM count <= 10 MiB,
 * i.e. up to 221 frames of 40 coefficients (2.2 s of audio) or 442 frames of 10 on a 148-SM B200; larger clips return
 * KWS_ERR_INVALID with a message, and the batch size B is not limited.  Handle set-up (kws_features_configure, kws_model_finalize) still allocates and must
 * happen before.  All calls share the array: calls that overlap on the GPU (streams, handles, threads) race --
 * serialise them.  Logits are bit-identical to kws_pcm_to_logits. */
int kws_pcm_to_logits_static(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                             float* logits_dev, void* stream);
int kws_pcm_to_logits_host(kws_handle_t h, const float* pcm_host, int64_t B, int64_t N,
                           float* logits_host);
/* the same two calls fed 16-bit PCM (x/32768 = librosa.load, audio_processor.py:145, happens inside the
 * MFCC kernel; logits are bit-equal to the float32 calls on x/32768) */
int kws_pcm_to_logits_i16(kws_handle_t h, const int16_t* pcm_dev, int64_t B, int64_t N,
                          int64_t pcm_stride, float* logits_dev, void* stream);
int kws_pcm_to_logits_host_i16(kws_handle_t h, const int16_t* pcm_host, int64_t B, int64_t N,
                               float* logits_host);

#ifdef __cplusplus
}
#endif
#endif /* KWS_B200_H_ */
