// PCM -> logits in ONE launch (north_star: "the whole clip stays on-chip from PCM to logits").
//
// Reference chain replaced: AudioProcessor.extract_features (kws/libs/audio_processor.py:235-278 -> psf.mfcc)
// followed by DepthwiseSeparableConv.forward (kws/libs/models.py:160-183).
//
// One persistent CTA per SM, two ROLES in the same grid:
//   * CTAs [0, n_mfcc): the warp-specialised MFCC pipeline of mfcc_ws_body.cuh (TMA ring -> register FFT ->
//     mel/log -> tcgen05 DCT), writing each tile of 52 frames into slot `clip % R` of a feature RING;
//   * CTAs [n_mfcc, grid): the row-streamed tcgen05 network of model_fused_body.cuh, reading a clip's features
//     from that slot as soon as all of its tiles are published, and handing the slot back after the last read.
// The ring (R = 1024 clips x 15.8 KB = 16 MB by default) lives in the 126 MB L2: a slot is overwritten ~R clips
// later, long before its lines would be evicted, so the features never travel to HBM and back -- HBM sees the
// PCM (64 KB/clip) and the logits (48 B/clip).  Handshake per slot, two monotonic counters in global memory:
//     prod_cnt[slot] += 1   by each MFCC epilogue warp after its rows of a tile are stored (red.release.gpu)
//     cons_cnt[slot] += 1   by the model CTA after its last load of the clip                (red.release.gpu)
//   clip c (generation g = c / R) may be READ  when prod_cnt[slot] >= prod_per_clip * (g + 1)   (ld.acquire.gpu)
//                                  may be WRITTEN when cons_cnt[slot] >= g
// Both roles walk their work in increasing clip order with a static round-robin assignment, so the smallest
// unconsumed clip can always be produced and then consumed: no deadlock for any R >= 1 as long as all CTAs are
// co-resident (grid <= SM count, one CTA per SM by shared memory).  The split n_mfcc : n_model follows the
// measured per-clip cost of the two roles (handle setting; default from the rates in DESIGN.md section 4).
#include "model_fused_body.cuh"

namespace kws {

constexpr int kPipeThreads = kWsThreads > kFThreads ? kWsThreads : kFThreads;
static_assert(kPipeThreads == kWsThreads, "the block is sized for the MFCC role; the model role idles the rest");

template <typename T, int F, int NPIX, int PACK>
__global__ void __launch_bounds__(kPipeThreads, 1)
pcm_logits_fused_kernel(const FastParams mp, const FusedParams fp, const unsigned n_mfcc) {
  extern __shared__ __align__(128) unsigned char smem[];
  if (blockIdx.x < n_mfcc) {
    mfcc512_ws_body<400, kWsRounds, T, true>(mp, smem, blockIdx.x, n_mfcc);
  } else {
    dscnn_fused_body<F, NPIX, PACK, true>(fp, smem, blockIdx.x - n_mfcc, gridDim.x - n_mfcc);
  }
}

template <typename T, int F, int NPIX, int PACK>
static int launch_t(kws_handle_t h, FastParams& mp, FusedParams& fp, size_t mfcc_smem, cudaStream_t stream) {
  using L = FusedLayout<F, NPIX, PACK>;
  const size_t model_smem = L::smem_bytes(fp.T);
  if (model_smem > 227 * 1024) return 0;
  const size_t smem_bytes = mfcc_smem > model_smem ? mfcc_smem : model_smem;
  const long long ngroups = (fp.B + L::CL - 1) / L::CL;
  if (ngroups > 0x7fffffffLL / 64) return 0;
  // role split: share of the SMs the MFCC role needs so that both roles run at the same clips/s
  const double share = h->pipe_mfcc_share > 0.0 ? h->pipe_mfcc_share : (F == 40 ? 0.218 : 0.371);
  long long grid = h->sm_count;
  long long n_mfcc = (long long)(grid * share + 0.5);
  if (n_mfcc < 1) n_mfcc = 1;
  if (n_mfcc > grid - 1) n_mfcc = grid - 1;
  if (n_mfcc > mp.total_tiles) n_mfcc = mp.total_tiles;
  long long n_model = grid - n_mfcc;
  if (n_model > ngroups) n_model = ngroups;
  grid = n_mfcc + n_model;
  auto kern = pcm_logits_fused_kernel<T, F, NPIX, PACK>;
  if (int rc = ensure_dyn_smem(kern, smem_bytes, true)) return rc;
  KWS_CUDA(cudaMemsetAsync(h->d_ring_cnt, 0, (size_t)2 * h->ring_slots * sizeof(unsigned), stream));
  kern<<<(unsigned)grid, kPipeThreads, smem_bytes, stream>>>(mp, fp, (unsigned)n_mfcc);
  KWS_CHECK_LAUNCH(h);
  return 1;
}

// 1 = handled, 0 = configuration outside the single-launch kernel (caller runs the two kernels back to back),
// < 0 = error
int launch_pcm_logits_fused(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N, int64_t pcm_stride,
                            float* logits_dev, cudaStream_t stream) {
  const FeatureTables& t = h->feat;
  if (h->model.path != 0) return 0;
  const int64_t nframes = kws_num_frames(N, t.cfg.frame_len, t.cfg.frame_step);
  const int F = t.ncep;
  // ring: power-of-two slot count, allocated once per handle (re-allocated when the feature shape grows)
  unsigned shift = 10;
  if (const char* e = std::getenv("KWS_RING_SLOTS")) {
    const long v = std::atol(e);
    if (v >= 2 && v <= (1 << 20)) {
      shift = 0;
      while ((1L << (shift + 1)) <= v) ++shift;
    }
  }
  const size_t slot_bytes = (size_t)nframes * F * 4;
  const size_t need = slot_bytes << shift;
  if (need > h->ring_bytes || h->ring_slots != (1u << shift)) {
    if (h->d_ring) KWS_CUDA(cudaFree(h->d_ring));
    if (h->d_ring_cnt) KWS_CUDA(cudaFree(h->d_ring_cnt));
    h->d_ring = nullptr;
    h->d_ring_cnt = nullptr;
    h->ring_bytes = 0;
    KWS_CUDA(cudaMalloc((void**)&h->d_ring, need));
    KWS_CUDA(cudaMalloc((void**)&h->d_ring_cnt, (size_t)2 * sizeof(unsigned) << shift));
    h->ring_bytes = need;
    h->ring_slots = 1u << shift;
  }
  FastParams mp;
  size_t mfcc_smem = 0;
  if (!fill_fast_params(h, pcm_dev, is_i16, B, N, pcm_stride, h->d_ring, kModeMfcc, &mp, &mfcc_smem)) return 0;
  FusedParams fp;
  if (!fill_fused_params(h, h->d_ring, B, nframes, F, logits_dev, &fp)) return 0;
  mp.ring_shift = fp.ring_shift = shift;
  mp.prod_cnt = h->d_ring_cnt;
  mp.cons_cnt = h->d_ring_cnt + h->ring_slots;
  fp.prod_cnt = h->d_ring_cnt;
  fp.cons_cnt = h->d_ring_cnt + h->ring_slots;
  fp.prod_per_clip = 2u * (unsigned)mp.tiles_per_clip;   // two epilogue warps publish each tile
  if (F == 40) {
    return is_i16 ? launch_t<short, 40, 24, 1>(h, mp, fp, mfcc_smem, stream)
                  : launch_t<float, 40, 24, 1>(h, mp, fp, mfcc_smem, stream);
  }
  if (F == 10) {
    const int rc = is_i16 ? launch_t<short, 10, 24, 2>(h, mp, fp, mfcc_smem, stream)
                          : launch_t<float, 10, 24, 2>(h, mp, fp, mfcc_smem, stream);
    if (rc != 0) return rc;
    return is_i16 ? launch_t<short, 10, 16, 1>(h, mp, fp, mfcc_smem, stream)
                  : launch_t<float, 10, 16, 1>(h, mp, fp, mfcc_smem, stream);
  }
  return 0;
}

}  // namespace kws
