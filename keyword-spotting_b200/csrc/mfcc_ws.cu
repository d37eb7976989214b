// Fast, warp-specialised PCM -> MFCC kernel: stand-alone launch (body in mfcc_ws_body.cuh).
#include <cstdlib>

#include "mfcc_ws_body.cuh"

namespace kws {

template <int LOAD_LEN, int ROUNDS, typename T>
__global__ void __launch_bounds__(kWsThreads, 1) mfcc512_ws_kernel(const FastParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  mfcc512_ws_body<LOAD_LEN, ROUNDS, T, false>(p, smem, blockIdx.x, gridDim.x, 0, p.B, p.out, 0u);
}

// Fills the kernel parameters and the dynamic shared-memory size for a configuration the fast kernel covers:
// returns 1, or 0 when the configuration needs the generic kernel (declared in mfcc_ws_body.cuh).
int fill_fast_params(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N, int64_t pcm_stride,
                     float* out_dev, int mode, FastParams* pp, size_t* smem_out) {
  const FeatureTables& t = h->feat;
  const int M = t.cfg.nfilt, hop = t.cfg.frame_step;
  if (t.d_window != nullptr || t.load_len != 400 || M > 48 || t.d_dctB == nullptr) return 0;
  // TMA bulk copies need 16-byte aligned sources and sizes (4 float or 8 int16 samples); the interleaved
  // tile needs an even hop
  const int64_t amask = is_i16 ? 7 : 3;
  const size_t esz = is_i16 ? 2 : 4;
  if ((hop & amask) != 0 || (N & amask) != 0 || (pcm_stride & amask) != 0 ||
      (reinterpret_cast<size_t>(pcm_dev) & 15) != 0)
    return 0;
  // the epilogue writes rows with float4 stores when ncep % 4 == 0: the output must then be 16-byte aligned
  // (an offset into a larger feature buffer need not be) - otherwise the generic kernel's scalar stores
  if ((t.ncep & 3) == 0 && mode == kModeMfcc && (reinterpret_cast<size_t>(out_dev) & 15) != 0) return 0;
  if (N >= (1ll << 30)) return 0;
  constexpr int ROUNDS = kWsRounds;
  FastParams& p = *pp;
  p = FastParams{};
  p.pcm = pcm_dev;
  p.out = out_dev;
  p.B = B;
  p.N = N;
  p.stride = pcm_stride;
  p.nframes = (int)kws_num_frames(N, t.cfg.frame_len, t.cfg.frame_step);
  p.tiles_per_clip = (p.nframes + kTileFrames - 1) / kTileFrames;
  p.total_tiles = (long long)B * p.tiles_per_clip;
  if (p.total_tiles >= (1ll << 31) - 4096) return 0;
  p.hop = hop;
  p.M = M;
  p.ncep = t.ncep;
  p.mode = mode;
  p.append_energy = t.cfg.append_energy;
  p.preemph = t.cfg.preemph;
  p.tw256 = t.d_tw256;
  p.twp = t.d_tw512;
  p.melw = t.d_melw;
  p.edges = t.d_segedge;
  p.dctB = t.d_dctB;
  p.kpad = t.dct_kpad;
  p.tile_len = (kTileFrames - 1) * hop + 400;  // samples a tile's frames span
  const uint32_t sbo = (uint32_t)(p.kpad / 4) * 128u;
  p.raw_stride = (int)(((size_t)p.tile_len * esz + 16 + 127) & ~(size_t)127);
  size_t off = (size_t)2 * p.raw_stride;           // two raw stages at offset 0
  p.off_a = (int)off;                              // A hi row groups, then A lo row groups
  off += (size_t)2 * kAGroups * sbo;
  p.off_scratch = (int)off;
  off += (size_t)kWsCWarps * kWarpScratch;
  p.off_b = (int)off;                              // B hi [48 x kpad], then B lo
  off += (size_t)2 * (kDctN / 8) * sbo;
  p.off_melw = (int)off;
  off += (size_t)kBins * 8 + 8;
  p.off_edges = (int)off;
  off += ((size_t)(M + 2) * 4 + 15) & ~(size_t)15;
  p.off_twp = (int)off;
  off += 128 * 8;
  p.off_c0 = (int)off;
  off += 2 * kTileFrames * 4;
  off = (off + 15) & ~(size_t)15;
  p.off_lows = (int)off;                           // low bins: float64 block sums, B fragments, frame phases
  off += kLowSBytes;
  p.off_lowbw = (int)off;
  off += kLowBwBytes;
  p.off_lowfw = (int)off;
  off += kLowFwBytes;
  const size_t smem_bytes = off;
  // the low-bin warp's block decomposition is specific to hop 160 / frame 400 (blocks of gcd = 80 samples);
  // KWS_LOW_BINS=0 switches it off (A/B timing, and the tail measurement of tools/mfcc_tail.py)
  static const bool low_env = [] { const char* e = std::getenv("KWS_LOW_BINS"); return !(e && e[0] == '0'); }();
  p.low_bins = (hop == 2 * kLowBlock && low_env) ? 1 : 0;
  static_assert(2064 + 2 * (16 * ROUNDS + 1) * 8 <= kWarpScratch / 2, "scratch layout overflow");
  // the tensor core reads 128 rows of A: 16 row groups past each base must stay inside the CTA's
  // shared memory (rows >= 24 are don't-care values that only reach accumulator rows we never read)
  if ((size_t)p.off_a + (size_t)(kAGroups + 16) * sbo > smem_bytes || smem_bytes > 227 * 1024 - 256) return 0;

  *smem_out = smem_bytes;
  return 1;
}

// returns 1 when the fast kernel handled the launch, 0 when the configuration needs the generic
// kernel, negative on error
int launch_mfcc_fast(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N, int64_t pcm_stride,
                     float* out_dev, int mode, cudaStream_t stream) {
  FastParams p;
  size_t smem_bytes = 0;
  if (!fill_fast_params(h, pcm_dev, is_i16, B, N, pcm_stride, out_dev, mode, &p, &smem_bytes)) return 0;
  auto kern = is_i16 ? mfcc512_ws_kernel<400, kWsRounds, short> : mfcc512_ws_kernel<400, kWsRounds, float>;
  if (int rc = ensure_dyn_smem(kern, smem_bytes, true)) return rc;
  long long grid = (long long)h->sm_count;  // one persistent CTA per SM
  if (grid > p.total_tiles) grid = p.total_tiles;
  kern<<<(unsigned)grid, kWsThreads, smem_bytes, stream>>>(p);
  KWS_CHECK_LAUNCH(h);
  return 1;
}

}  // namespace kws
