// PCM -> logits in ONE launch (north_star: "the whole clip stays on-chip from PCM to logits").
//
// Reference chain replaced: AudioProcessor.extract_features (kws/libs/audio_processor.py:235-278 -> psf.mfcc)
// followed by DepthwiseSeparableConv.forward (kws/libs/models.py:160-183).
//
// One persistent CTA per SM.  Every CTA owns a contiguous range of clips and walks it in CHUNKS (32 clips by
// default), alternating between the two phases it shares its shared memory and tensor memory between:
//   1. MFCC phase   : the warp-specialised pipeline of mfcc_ws_body.cuh (TMA ring -> register FFT -> float64 low
//                     bins -> mel/log -> tcgen05 DCT) turns the chunk's PCM into features in the CTA's OWN scratch
//                     (chunk x 15.8 KB; 148 CTAs x 32 clips = 75 MB, inside the 126 MB L2, rewritten every chunk);
//   2. network phase: the row-streamed tcgen05 network of model_fused_body.cuh reads those features back with
//                     L2-coherent loads and writes the chunk's logits.
// A CTA only ever consumes what it produced itself, so the phases need no handshake beyond the CTA barrier between
// them: no flags, no inter-CTA ordering, nothing to balance (each CTA does the same work per clip), and a small
// batch is simply one short chunk per CTA.  HBM sees the PCM (64 KB/clip) and the logits (48 B/clip); the feature
// scratch lives in L2.  Per chunk the phase switch costs one re-initialisation of each role's shared-memory tables
// (a few microseconds against ~2 ms of work per 32-clip chunk).
//
// Round-2 history, measured on B200 at 65 536 clips: a producer/consumer split of the grid (MFCC CTAs feeding
// network CTAs through an L2 ring with per-slot counters) reached 2.17 M clips/s against 2.34 M for the two kernels
// back to back -- the network CTAs ran 5-10 % slower next to polling producers and the two roles had to be balanced
// to within one CTA; this design replaced it.
#include <cstdlib>

#include "model_fused_body.cuh"

namespace kws {

constexpr int kPipeThreads = kWsThreads > kFThreads ? kWsThreads : kFThreads;
static_assert(kPipeThreads == kWsThreads, "the block is sized for the MFCC phase; the network phase idles one warp");

// The network phase is a separate (non-inlined) device function: with both phases inlined into one loop body ptxas
// allocated registers for an 11 K-instruction function and spilled the stencil state and FFT registers inside the
// hot loops (-45 %).  The MFCC phase stays inline in the kernel so that its parameter record keeps its place in the
// constant bank (as a function argument it lives in local memory: +26 % instructions, measured); the network reads
// its few parameters by value (-2 %).  The dynamic shared memory is re-declared inside the function (same array),
// so that the address space stays known.
template <int F, int NPIX, int PACK>
__device__ __noinline__ void network_phase(const FusedParams fp, long long nc, const float* feats, float* logits,
                                           uint32_t tmem) {
  extern __shared__ __align__(128) unsigned char smem[];
  dscnn_fused_body<F, NPIX, PACK, true>(fp, smem, 0u, 1u, nc, feats, logits, tmem);
}

template <typename T, int F, int NPIX, int PACK>
__global__ void __launch_bounds__(kPipeThreads, 1)
pcm_logits_fused_kernel(const FastParams mp, const FusedParams fp, float* scratch, const int chunk_clips, const int skip) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint32_t s_tmem_all;
  using L = FusedLayout<F, NPIX, PACK>;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  if (warp == 0) {  // all of tensor memory, once: the MFCC phase uses 64 columns of it, the network all 512
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem_all)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem_all;
  // this CTA's clips: groups of CL clips split evenly over the grid (a group is the network's unit of work)
  const long long ngroups = (fp.B + L::CL - 1) / L::CL;
  const long long lo = L::CL * (ngroups * blockIdx.x / gridDim.x);
  long long hi = L::CL * (ngroups * (blockIdx.x + 1) / gridDim.x);
  if (hi > fp.B) hi = fp.B;
  float* my = scratch + (size_t)blockIdx.x * chunk_clips * fp.T * F;
  for (long long c0 = lo; c0 < hi; c0 += chunk_clips) {
    const long long nc = hi - c0 < chunk_clips ? hi - c0 : chunk_clips;
    if (!(skip & 1)) mfcc512_ws_body<400, kWsRounds, T, true>(mp, smem, 0u, 1u, c0, nc, my, tmem);
    __threadfence();  // the features were written through L2 by this CTA's epilogue warps
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (!(skip & 2)) network_phase<F, NPIX, PACK>(fp, nc, my, fp.logits + c0 * fp.ncls, tmem);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512));
}

template <typename T, int F, int NPIX, int PACK>
static int launch_t(kws_handle_t h, FastParams& mp, FusedParams& fp, size_t mfcc_smem, int chunk, cudaStream_t stream) {
  using L = FusedLayout<F, NPIX, PACK>;
  const size_t model_smem = L::smem_bytes(fp.T);
  if (model_smem > 227 * 1024) return 0;
  const size_t smem_bytes = mfcc_smem > model_smem ? mfcc_smem : model_smem;
  const long long ngroups = (fp.B + L::CL - 1) / L::CL;
  if (ngroups > 0x7fffffffLL / 64) return 0;
  long long grid = h->sm_count;
  if (grid > ngroups) grid = ngroups;
  chunk = (chunk + L::CL - 1) / L::CL * L::CL;
  // scratch: one chunk of features per CTA, allocated once per handle (grows with the chunk size / feature shape)
  const size_t need = (size_t)h->sm_count * chunk * fp.T * F * sizeof(float);
  if (need > h->ring_bytes) {
    if (h->d_ring) KWS_CUDA(cudaFree(h->d_ring));
    h->d_ring = nullptr;
    h->ring_bytes = 0;
    KWS_CUDA(cudaMalloc((void**)&h->d_ring, need));
    h->ring_bytes = need;
  }
  auto kern = pcm_logits_fused_kernel<T, F, NPIX, PACK>;
  if (int rc = ensure_dyn_smem(kern, smem_bytes, true)) return rc;
  // KWS_PHASE_SKIP (timing aid only; the logits are then meaningless): 1 = skip the MFCC phase, 2 = skip the network phase
  static const int skip = [] { const char* e = std::getenv("KWS_PHASE_SKIP"); return e ? std::atoi(e) : 0; }();
  kern<<<(unsigned)grid, kPipeThreads, smem_bytes, stream>>>(mp, fp, h->d_ring, chunk, skip);
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
  int chunk = h->pipe_chunk_clips;
  if (const char* e = std::getenv("KWS_CHUNK_CLIPS")) {
    const long v = std::atol(e);
    if (v >= 1 && v <= 4096) chunk = (int)v;
  }
  FastParams mp;
  size_t mfcc_smem = 0;
  // out_dev is replaced by the per-CTA scratch inside the kernel; any aligned non-null pointer passes the checks
  if (!fill_fast_params(h, pcm_dev, is_i16, B, N, pcm_stride, logits_dev, kModeMfcc, &mp, &mfcc_smem)) return 0;
  FusedParams fp;
  if (!fill_fused_params(h, nullptr, B, nframes, F, logits_dev, &fp)) return 0;
  if (F == 40) {
    return is_i16 ? launch_t<short, 40, 24, 1>(h, mp, fp, mfcc_smem, chunk, stream)
                  : launch_t<float, 40, 24, 1>(h, mp, fp, mfcc_smem, chunk, stream);
  }
  if (F == 10) {
    const int rc = is_i16 ? launch_t<short, 10, 24, 2>(h, mp, fp, mfcc_smem, chunk, stream)
                          : launch_t<float, 10, 24, 2>(h, mp, fp, mfcc_smem, chunk, stream);
    if (rc != 0) return rc;
    return is_i16 ? launch_t<short, 10, 16, 1>(h, mp, fp, mfcc_smem, chunk, stream)
                  : launch_t<float, 10, 16, 1>(h, mp, fp, mfcc_smem, chunk, stream);
  }
  return 0;
}

}  // namespace kws

// ---------------------------------------------------------------------------------------------------------------
// kws_pcm_to_logits_static: the same single-launch kernel with its feature scratch in a file-scope device array
// instead of a cudaMalloc'ed buffer -- for callers that cannot allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// The array is kStaticScratchBytes (10 MiB) on every GPU that loads the library, held until the process exits.  It is
// cut into one slice per CTA; a CTA's chunk is as many clips as its slice holds (at 99 x 40 features on 148 SMs: 4
// clips per CTA and phase instead of 32, ~3 % slower).  Clips whose feature map does not fit even one group of clips
// per CTA are rejected with KWS_ERR_INVALID.  Nothing in the scratch has to start at zero: the MFCC phase writes every
// feature value of a chunk before the network phase reads it.  Every call shares the array, so calls that overlap on
// the GPU (two streams, two handles, two host threads on one device) race; serialise them.
namespace kws {

constexpr size_t kStaticScratchBytes = 10u << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
constexpr int kStaticMaxDevices = 64;

// In host code `g_static_scratch` is only a host-side stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int static_scratch_ptr(int device, float** out) {
  static float* ptrs[kStaticMaxDevices] = {};
  if (device < 0 || device >= kStaticMaxDevices) return fail(KWS_ERR_INVALID, "device %d out of range", device);
  if (!ptrs[device]) {
    void* p = nullptr;
    KWS_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) return fail(KWS_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
    ptrs[device] = static_cast<float*>(p);
  }
  *out = ptrs[device];
  return KWS_OK;
}

template <typename T, int F, int NPIX, int PACK>
static int launch_static_t(kws_handle_t h, FastParams& mp, FusedParams& fp, size_t mfcc_smem, float* scratch,
                           cudaStream_t stream) {
  using L = FusedLayout<F, NPIX, PACK>;
  // clips per CTA slice: a multiple of the network's group size, at most the default chunk
  const size_t clip_bytes = (size_t)fp.T * F * sizeof(float);
  long long chunk = (long long)(kStaticScratchBytes / ((size_t)h->sm_count * clip_bytes));
  chunk = chunk / L::CL * L::CL;
  if (chunk > 32) chunk = 32;
  if (chunk < L::CL)
    return fail(KWS_ERR_INVALID,
                "kws_pcm_to_logits_static: %d frames x %d coefficients need %zu bytes of scratch for %d clips on each of "
                "%d SMs, the static scratch holds %zu bytes; use kws_pcm_to_logits",
                fp.T, F, (size_t)L::CL * clip_bytes, L::CL, h->sm_count, kStaticScratchBytes);
  const size_t model_smem = L::smem_bytes(fp.T);
  if (model_smem > 227 * 1024) return 0;
  const size_t smem_bytes = mfcc_smem > model_smem ? mfcc_smem : model_smem;
  const long long ngroups = (fp.B + L::CL - 1) / L::CL;
  if (ngroups > 0x7fffffffLL / 64) return 0;
  long long grid = h->sm_count;
  if (grid > ngroups) grid = ngroups;
  auto kern = pcm_logits_fused_kernel<T, F, NPIX, PACK>;
  if (int rc = ensure_dyn_smem(kern, smem_bytes, true)) return rc;
  kern<<<(unsigned)grid, kPipeThreads, smem_bytes, stream>>>(mp, fp, scratch, (int)chunk, 0);
  KWS_CHECK_LAUNCH(h);
  return 1;
}

}  // namespace kws

extern "C" int kws_pcm_to_logits_static(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                                        float* logits_dev, void* stream) {
  using namespace kws;
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (!h->feat.ready) return fail(KWS_ERR_STATE, "features not configured");
  if (!h->model.finalized) return fail(KWS_ERR_STATE, "model weights not finalized");
  if (B < 0 || N < 1 || pcm_stride < 1) return fail(KWS_ERR_INVALID, "bad B/N/stride");
  if (B == 0) return KWS_OK;
  if (!pcm_dev || !logits_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  KWS_CUDA(cudaSetDevice(h->device));
  float* scratch = nullptr;
  if (int rc = static_scratch_ptr(h->device, &scratch)) return rc;
  const FeatureTables& t = h->feat;
  const int64_t nframes = kws_num_frames(N, t.cfg.frame_len, t.cfg.frame_step);
  const int F = t.ncep;
  FastParams mp;
  size_t mfcc_smem = 0;
  FusedParams fp;
  int rc = 0;
  if (h->model.path == 0 && (F == 40 || F == 10) &&
      fill_fast_params(h, pcm_dev, false, B, N, pcm_stride, logits_dev, kModeMfcc, &mp, &mfcc_smem) &&
      fill_fused_params(h, nullptr, B, nframes, F, logits_dev, &fp)) {
    cudaStream_t st = (cudaStream_t)stream;
    if (F == 40) {
      rc = launch_static_t<float, 40, 24, 1>(h, mp, fp, mfcc_smem, scratch, st);
    } else {
      rc = launch_static_t<float, 10, 24, 2>(h, mp, fp, mfcc_smem, scratch, st);
      if (rc == 0) rc = launch_static_t<float, 10, 16, 1>(h, mp, fp, mfcc_smem, scratch, st);
    }
  }
  if (rc < 0) return rc;
  if (rc == 0)
    return fail(KWS_ERR_INVALID,
                "kws_pcm_to_logits_static covers the single-launch configurations only (reference network 64 ch / 4 blocks, "
                "10 or 40 coefficients, 400-sample frames, rectangular window, 16-byte aligned float32 PCM); use "
                "kws_pcm_to_logits");
  return KWS_OK;
}
