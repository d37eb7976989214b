// Fast, warp-specialised variant of the fused PCM -> MFCC kernel (see mfcc.cu for the algorithm
// and the reference call chain it replaces: kws/libs/audio_processor.py:235-278 -> psf.mfcc).
//
// Same arithmetic as mfcc.cu (warp = 4 frames, half-warp = 2 frames packed in f32x2 registers,
// 256-point complex FFT = 16 lanes x radix-16 twice).  The kernel is bound by exposed latency,
// instruction issue and shared-memory bandwidth, not HBM, so this variant removes
// instructions and every CTA-wide barrier and keeps as many compute warps resident as fit:
//   * one persistent CTA per SM, 15 warps (128 registers): 13 COMPUTE warps that never wait for each
//     other - they only meet at mbarriers (tile landed / A rows written / previous MMA retired) - and
//     2 SERVICE warps.  Service warp 0 streams raw PCM tiles (52 frames of one clip; float32 or int16
//     samples) into a 2-stage shared-memory ring with TMA bulk copies (cp.async.bulk + mbarrier
//     full/empty pairs) and issues the tensor-core DCT of each finished tile; both service warps drain
//     the accumulator from TMEM to global memory (rows 0..31 and 32..51: a warp reads the TMEM lane
//     quarter of its index);
//   * pre-emphasis (fp32, two roundings = psf's float32 step) on the fly when a warp reads its samples
//     from the raw stage;
//   * FADD2/FFMA2 operand modifiers (negate, scalar broadcast) keep twiddles and weights scalar;
//   * the real-FFT split computes X[k] and X[256-k] together from one shuffled mirror value;
//   * DCT-II (+lifter) of the tile, [52 x M] x [M x C], on the tensor core: tcgen05.mma kind::tf32
//     with a 3-term hi/lo split (fp32-grade), K-major no-swizzle operands, accumulator in TMEM,
//     tcgen05.commit -> mbarrier, tcgen05.ld epilogue in the service warps;
//   * log is a range-reduced MUFU.LG2.
// Configurations it does not cover (window function, frame_len != 400, nfilt > 48, unaligned
// input, a hop too long for the shared-memory tile) fall back to the generic kernel in mfcc.cu.
#pragma once
#include "mfcc_dev.cuh"

namespace kws {

struct FastParams {
  const void* pcm;      // float32, or int16 (scaled by 1/32768 = librosa.load) when the kernel is the int16 variant
  float* out;
  long long B, N, stride, total_tiles;
  int nframes, tiles_per_clip, hop, M, ncep, mode, append_energy;
  float preemph;
  const float2* tw256;
  const float2* twp;    // [128] W512^k
  const float2* melw;   // [257] (rising, falling) weight of the bin's segment
  const int* edges;     // [M+2]
  const float* dctB;    // B operand image: hi then lo, canonical K-major layout [48 x Kpad] each
  int kpad, tile_len, raw_stride;  // raw_stride: bytes between the two raw stages
  int off_a, off_scratch, off_b, off_melw, off_edges, off_twp, off_c0;
  int off_lows, off_lowbw, off_lowfw;   // low bins: block sums, DMMA B fragments, frame-assembly phases
  int low_bins;             // 1: bins 1..4 in float64 on the fp64 tensor core (hop 160 / frame 400 only), 0: fp32 FFT only
};

// One CTA per SM: the kernel is bound by exposed latency, not by issue slots or HBM, so what counts is the
// number of resident compute warps.  ptxas needs 128 registers for the compute path, i.e. 16 warps per SM:
// 13 compute warps (tile = 52 frames, two tiles per 99-frame clip, 25 of 26 warp slots busy) + 2 service
// warps.  (Two CTAs of 5 + 1 warps, the previous layout, kept 10 compute warps resident: 10.8 M clips/s.)
// Warp slots are handed out consecutively (tools/warpslot_probe.cu) and slot % 4 is the scheduler: 12 compute
// warps are three per scheduler, the 13th is a fourth one on one scheduler and makes its tile 19 % slower
// for 8 % more work - at this occupancy the kernel is at a throughput plateau: tiles that
// ignore clip ends, with 12 or 13 always-busy warps, were measured and are no faster.
constexpr int kWsCWarps = 13;                        // compute warps
constexpr int kSvcWarps = 2;                       // service warps: TMEM lanes 0..31 and 32..63
constexpr int kTileFrames = kWsCWarps * 4;           // 52 frames per tile = rows of the DCT GEMM
constexpr int kAGroups = (kTileFrames + 7) / 8;    // 8-row groups of the A operand
constexpr int kWsThreads = (kWsCWarps + kSvcWarps) * 32;
// ---- bins 1..4 in float64 (see low_block_sums below) ----
constexpr int kLowBins = 4;                        // bins 1..4
constexpr int kLowBlock = 80;                      // gcd(hop 160, frame 400): block sums are shared between frames
constexpr int kLowMTiles = kWsCWarps + 1;            // 8-block row tiles of a 52-frame tile (107 blocks): one per compute warp + 1
constexpr int kLowSBytes = kLowMTiles * 8 * kLowBins * 16;   // block sums (re, im) float64
constexpr int kLowBwBytes = 5 * 2 * 32 * 16;       // DMMA B fragments [q][half][lane][2] float64
constexpr int kLowFwBytes = kLowBins * 5 * 16;     // frame-assembly phases [bin][j](re, im) float64
constexpr int kWarpScratch = 2 * 16 * 17 * 16;     // bytes: two half-warp transposes (also holds S2)
constexpr float kFEps = 2.220446049250313e-16f;
constexpr int kWsRounds = 3;                       // mel filters per lane: nfilt <= 16 * kWsRounds


// raw-sample access: the raw stage holds the clip's own sample type.  int16 samples are used as integers-in-
// float: librosa.load's x / 32768 (audio_processor.py:145) is a power of two, so it commutes with every fp32
// rounding of the pre-emphasis and the FFT (nothing here comes near the under/overflow range: |X| < 2^26) and
// is applied once, as 2^-30, to the power spectrum - the same bits as converting first, 6 multiplies per
// sample pair cheaper.
__device__ __forceinline__ float2 ld_pair(const float* q) { return *reinterpret_cast<const float2*>(q); }
__device__ __forceinline__ float ld_one(const float* q) { return *q; }
__device__ __forceinline__ float2 ld_pair(const short* q) {
  const short2 s = *reinterpret_cast<const short2*>(q);
  return make_float2((float)s.x, (float)s.y);
}
__device__ __forceinline__ float ld_one(const short* q) { return (float)*q; }
__device__ __forceinline__ void ld_quad(const float* q, float (&x)[4]) {
  const float4 v = *reinterpret_cast<const float4*>(q);
  x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
}
__device__ __forceinline__ void ld_quad(const short* q, float (&x)[4]) {
  const short4 v = *reinterpret_cast<const short4*>(q);
  x[0] = (float)v.x; x[1] = (float)v.y; x[2] = (float)v.z; x[3] = (float)v.w;
}
// D[8x8] += A[8x4] . B[4x8] in float64 on the tensor core (DMMA): lane holds A[lane/4][lane%4], B[lane%4][lane/4],
// C/D[lane/4][2 (lane%4) + {0, 1}]
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}
template <typename T> __device__ __forceinline__ constexpr float power_scale() { return 1.0f; }
template <> __device__ __forceinline__ constexpr float power_scale<short>() { return 1.0f / 1073741824.0f; }

// T = float or short.  kLead = samples kept in front of a tile (16 bytes, so that the TMA destination of a
// tile that starts inside the clip stays 16-byte aligned while sample s0 - 1 is available)
// The body works on the sub-batch of `nclips` clips starting at clip `clip0` of p.pcm and writes the features of its
// clip c to out_base[c] (c = 0 .. nclips-1); tiles bid, bid + nblk, ... of that sub-batch are this CTA's.
//   CHUNK = false: the stand-alone kernel (clip0 = 0, nclips = p.B, out_base = p.out, grid-strided tiles).
//   CHUNK = true : one phase of the single-launch PCM -> logits kernel (pcm_logits_fused.cu): the CTA transforms
//                  its next chunk of clips into its own L2-resident scratch (bid = 0, nblk = 1) before the network
//                  phase consumes it; tensor memory is allocated once by the caller (tmem_ext) and the mbarriers
//                  are invalidated at the end so that the next phase can re-initialise them.
template <int LOAD_LEN, int ROUNDS, typename T, bool CHUNK>
__device__ __forceinline__ void mfcc512_ws_body(const FastParams& p, unsigned char* smem, const unsigned bid,
                                                const unsigned nblk, const long long clip0, const long long nclips,
                                                float* out_base, const uint32_t tmem_ext) {
  const long long total_tiles = nclips * p.tiles_per_clip;
  constexpr int kLead = 16 / (int)sizeof(T);
  // [0,1] raw stage full, [2,3] raw stage empty, [4] A rows of the tile written, [5] tile MMAs retired,
  // [6] second service warp has drained its accumulator rows, [7] low-bin block sums of the tile written by every
  // compute warp, [8] ... and read by every compute warp
  __shared__ __align__(8) unsigned long long s_bar[9];
  __shared__ uint32_t s_tmem;
  const float* s_inv = reinterpret_cast<const float*>(smem + p.off_melw);  // [M+1] 1 / segment length
  const int* s_edges = reinterpret_cast<const int*>(smem + p.off_edges);
  const float2* s_twp = reinterpret_cast<const float2*>(smem + p.off_twp);
  float* s_c0 = reinterpret_cast<float*>(smem + p.off_c0);  // [2][kTileFrames]

  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const bool do_dct = (p.mode == kModeMfcc);
  const uint32_t bar_full = smem_u32(&s_bar[0]), bar_empty = smem_u32(&s_bar[2]);
  const uint32_t bar_afull = smem_u32(&s_bar[4]), bar_mma = smem_u32(&s_bar[5]), bar_epi = smem_u32(&s_bar[6]);
  const uint32_t bar_sfull = smem_u32(&s_bar[7]), bar_sfree = smem_u32(&s_bar[8]);
  const int hop = p.hop;
  const uint32_t sbo = (uint32_t)(p.kpad / 4) * 128u;  // bytes between 8-row groups of A and B

  if (tid == 0) {
    mbar_init(bar_full, 1);
    mbar_init(bar_full + 8, 1);
    mbar_init(bar_empty, kWsCWarps);
    mbar_init(bar_empty + 8, kWsCWarps);
    mbar_init(bar_sfull, kWsCWarps);
    mbar_init(bar_sfree, kWsCWarps);
    mbar_init(bar_afull, kWsCWarps);
    mbar_init(bar_mma, 1);
    mbar_init(bar_epi, 1);
    fence_mbar_init();
  }
  if (!CHUNK && warp == 0 && do_dct) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)),
                 "n"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  {
    float* d = reinterpret_cast<float*>(smem + p.off_melw);
    for (int i = tid; i <= p.M; i += kWsThreads) d[i] = 1.0f / (float)(p.edges[i + 1] - p.edges[i]);
    int* e = reinterpret_cast<int*>(smem + p.off_edges);
    for (int i = tid; i < p.M + 2; i += kWsThreads) e[i] = p.edges[i];
    float2* t = reinterpret_cast<float2*>(smem + p.off_twp);
    for (int i = tid; i < 128; i += kWsThreads) t[i] = p.twp[i];
    float* b = reinterpret_cast<float*>(smem + p.off_b);
    for (int i = tid; i < 2 * kDctN * p.kpad; i += kWsThreads) b[i] = p.dctB[i];
    float* a = reinterpret_cast<float*>(smem + p.off_a);
    for (int i = tid; i < 2 * kAGroups * 8 * p.kpad; i += kWsThreads) a[i] = 0.f;
    if (p.low_bins) {
      // DMMA B fragments of the block-sum GEMM, [q][half][lane][2]: lane (r = lane/4, c = lane%4) holds, for k-step
      // 4 q + j, the twiddle of sample i = 16 q + 4 c + j and column n = r: bin r/2 + 1, cos for even n, -sin for odd
      double* bwt = reinterpret_cast<double*>(smem + p.off_lowbw);
      for (int i = tid; i < 5 * 2 * 32 * 2; i += kWsThreads) {
        const int e = i & 1, ln = (i >> 1) & 31, hh = (i >> 6) & 1, q = i >> 7;
        const int j = 2 * hh + e, r = ln >> 2, c = ln & 3;
        double sn, cs;
        sincospi((double)((16 * q + 4 * c + j) * ((r >> 1) + 1)) / 256.0, &sn, &cs);
        bwt[i] = (r & 1) ? -sn : cs;
      }
      // phase of block j of a frame for bin b + 1: e^{-2 pi i 80 j (b + 1) / 512}
      double* fwt = reinterpret_cast<double*>(smem + p.off_lowfw);
      for (int i = tid; i < kLowBins * 5; i += kWsThreads) {
        double sn, cs;
        sincospi((double)(kLowBlock * (i % 5) * (i / 5 + 1)) / 256.0, &sn, &cs);
        fwt[2 * i] = cs;
        fwt[2 * i + 1] = -sn;
      }
    }
  }
  fence_proxy_async();  // B (and the zeroed A) were written through the generic proxy
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = CHUNK ? tmem_ext : (do_dct ? s_tmem : 0u);
  // tiles of this CTA: bid + it * nblk, it = 0 .. n_it-1
  const int n_it = (total_tiles > (long long)bid)
                       ? (int)((total_tiles - bid + nblk - 1) / nblk)
                       : 0;

  // Tile `it` of this CTA.  A clip's first tile keeps all 13 compute warps busy, its second one 12, so a CTA
  // that always got the same half (even grid, two tiles per clip) would be the slow or the fast one of its
  // pair: neighbouring CTAs swap tiles every other iteration.
  const bool swap_ok = (nblk & 1u) == 0u;
  auto tile_of = [&](int it) -> unsigned {
    unsigned id = bid + (unsigned)it * nblk;  // launcher: total_tiles < 2^31
    if (swap_ok && (it & 1) && (long long)(id | 1u) < total_tiles) id ^= 1u;
    return id;
  };

  if (warp < kSvcWarps) {
    // =========================== service warps ===========================
    // warp 0: TMA ring producer, MMA issue, epilogue of accumulator rows 0..31; warp 1: epilogue of rows 32..63
    // (a warp reads the TMEM lane quarter of its index)
    auto issue_tma = [&](int it) {  // one thread: raw[stage][j] = x[s0 - kLead + j]
      const unsigned tile_id = tile_of(it);
      const unsigned clip = tile_id / (unsigned)p.tiles_per_clip;
      const int t = (int)(tile_id - clip * (unsigned)p.tiles_per_clip);
      const long long s0 = (long long)t * kTileFrames * hop;
      const long long first = s0 >= kLead ? s0 - kLead : 0;
      long long end = s0 + p.tile_len;
      if (end > p.N) end = p.N;
      const uint32_t bytes = (uint32_t)(end - first) * (uint32_t)sizeof(T);
      const uint32_t bar = bar_full + 8u * (it & 1);
      T* raw = reinterpret_cast<T*>(smem + (size_t)(it & 1) * p.raw_stride);
      mbar_expect_tx(bar, bytes);
      const T* src = reinterpret_cast<const T*>(p.pcm) + (size_t)(clip0 + clip) * p.stride + first;
      if (CHUNK)  // the PCM streams through L2 once: it must not push the CTAs' feature scratch out (evict_first)
        tma_load_1d_hint(smem_u32(raw + (s0 >= kLead ? 0 : kLead)), src, bytes, bar, l2_policy_evict_first());
      else
        tma_load_1d(smem_u32(raw + (s0 >= kLead ? 0 : kLead)), src, bytes, bar);
    };
    auto issue_mma = [&]() {  // converged warp, elected lane: D[128 x 48] = Ahi*Bhi + Alo*Bhi + Ahi*Blo over Kpad
      tc_fence_after();
      const uint32_t a_hi = smem_u32(smem + p.off_a), a_lo = a_hi + (uint32_t)kAGroups * sbo;
      const uint32_t b_hi = smem_u32(smem + p.off_b), b_lo = b_hi + (uint32_t)(kDctN / 8) * sbo;
      const int ksteps = p.kpad / 8;
      uint32_t acc = 0;
#pragma unroll 1
      for (int term = 0; term < 3; ++term) {
        const uint32_t ab = term == 1 ? a_lo : a_hi, bb = term == 2 ? b_lo : b_hi;
        for (int ks = 0; ks < ksteps; ++ks) {
          tc_mma_tf32_elect(tmem, umma_desc(ab + ks * 256u, 128u, sbo), umma_desc(bb + ks * 256u, 128u, sbo), kIdesc, acc);
          acc = 1;
        }
      }
      tc_commit_elect(bar_mma);
    };
    if (warp == 0 && lane == 0) {
      if (n_it > 0) issue_tma(0);
      if (n_it > 1) issue_tma(1);
    }
    for (int it = 0; it < n_it; ++it) {
      const unsigned tile_id = tile_of(it);
      const unsigned clip = tile_id / (unsigned)p.tiles_per_clip;
      const int t = (int)(tile_id - clip * (unsigned)p.tiles_per_clip);
      if (warp == 0 && it + 2 < n_it) {  // refill this stage as soon as every compute warp has converted tile `it`
        mbar_wait_idle(bar_empty + 8u * (it & 1), (it >> 1) & 1);
        if (lane == 0) {
          fence_proxy_async();
          issue_tma(it + 2);
        }
        __syncwarp();
      }
      if (do_dct) {
        if (warp == 0) {
          mbar_wait_idle(bar_afull, it & 1);
          if (it > 0) mbar_wait(bar_epi, (it - 1) & 1);  // warp 1 has read the previous tile's rows
          issue_mma();
          mbar_wait(bar_mma, it & 1);
        } else {
          mbar_wait_idle(bar_mma, it & 1);
        }
        // ---- epilogue (lane = frame row of the tile): TMEM -> registers -> global ----
        float v[kDctN];
        tc_fence_after();
        {
          float q[16];
#pragma unroll
          for (int c = 0; c < kDctN / 16; ++c) {
            tc_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c * 16, q);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[c * 16 + i] = q[i];
          }
        }
        tc_wait_ld();
        tc_fence_before();
        const int row = warp * 32 + lane;
        const int f = t * kTileFrames + row;
        const uint64_t pol_keep = CHUNK ? l2_policy_evict_last() : 0ull;

        // c0 of this tile is read BEFORE warp 1 releases the accumulator: its release lets the next tile's MMAs
        // go, and their completion lets the compute warps overwrite this half of s_c0 two tiles on
        const float c0 = (row < kTileFrames) ? s_c0[(it & 1) * kTileFrames + row] : 0.f;
        if (warp != 0) {
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_epi);
        }
        if (row < kTileFrames && f < p.nframes) {
          v[0] = p.append_energy ? c0 : v[0] + c0;
          float* o = out_base + ((size_t)clip * p.nframes + f) * p.ncep;
          if ((p.ncep & 3) == 0) {
#pragma unroll
            for (int c = 0; c < kDctN; c += 4)
              if (c < p.ncep) {
                const float4 q4 = make_float4(v[c], v[c + 1], v[c + 2], v[c + 3]);
                if (CHUNK) st_global_v4_hint(o + c, q4, pol_keep);  // scratch: stays in L2 until the network phase reads it
                else *reinterpret_cast<float4*>(o + c) = q4;
              }
          } else {
#pragma unroll
            for (int c = 0; c < kDctN; ++c)
              if (c < p.ncep) o[c] = v[c];
          }
        }
      }
    }
  } else {
    // =========================== compute warps ===========================
    const int cw = warp - kSvcWarps, half = lane >> 4, l = lane & 15;
    double* s_S = reinterpret_cast<double*>(smem + p.off_lows);  // [blocks][4 bins][re, im]
    unsigned char* wscratch = smem + p.off_scratch + cw * kWarpScratch;
    unsigned char* scratch = wscratch + half * (kWarpScratch / 2);  // this half-warp's transpose area
    float4* s_tr = reinterpret_cast<float4*>(scratch);
    V* s_P = reinterpret_cast<V*>(scratch);
    V* s_R = reinterpret_cast<V*>(scratch + 2064);
    V* s_F = s_R + (16 * ROUNDS + 1);
    float twr[16], twi[16];  // W256^(l k1); FFMA2 takes a scalar register as a broadcast operand
#pragma unroll
    for (int k = 1; k < 16; ++k) {
      const float2 w = p.tw256[(l * k) & 255];
      twr[k] = w.x;
      twi[k] = w.y;
    }
    const float pre = p.preemph;
    const int n = (int)p.N;

    for (int it = 0; it < n_it; ++it) {
      const unsigned tile_id = tile_of(it);
      const unsigned clip = tile_id / (unsigned)p.tiles_per_clip;
      const int t = (int)(tile_id - clip * (unsigned)p.tiles_per_clip);
      const int s0 = t * kTileFrames * hop;
      T* raw_w = reinterpret_cast<T*>(smem + (size_t)(it & 1) * p.raw_stride);
      const T* raw = raw_w;
      __syncwarp();  // previous tile: every lane is done with the scratch this conversion overwrites
      mbar_wait(bar_full + 8u * (it & 1), (it >> 1) & 1);
      if (s0 == 0 && cw == 0) {  // y[0] = x[0] - pre * 0: give the clip's first sample a zero predecessor
        if (lane == 0) raw_w[kLead - 1] = (T)0;
        __syncwarp();
      }

      const int fa_local = cw * 4 + half * 2;
      const int fA = t * kTileFrames + fa_local, fB = fA + 1;
      const bool active = (t * kTileFrames + cw * 4) < p.nframes;  // warp-uniform

      // ---- bins 1..4 in float64, part 1: block sums on the fp64 tensor core ----
      // Why: pre-emphasised broadband audio leaves bins 1..2 about 30 dB under the rest of the spectrum, so the
      // ~3e-7 * ||frame|| rounding floor of ANY fp32 FFT is a 1e-5 relative error of the lowest mel filters, which
      // the log, the DCT and the x11.5 lifter turn into MFCC errors of up to 7x the north_star bound on white noise
      // (22 of 3.2e7 values outside it).  Frames overlap (hop 160, length 400), so the DFT of the four bins is taken
      // once per 80-sample BLOCK,  S_b[k] = sum_{i<80} y[80 b + i] e^{-2 pi i k / 512}  -- a GEMM [blocks x 80] . [80 x 8]
      // in DMMA m8n8k4 steps, one 8-block row tile per compute warp (the warp's own 640 samples; warp 0 also
      // takes the 14th tile) -- and a frame is five phase-shifted blocks (part 2, below).  The GEMM's K index is
      // permuted (sample 16 q + 4 c + j <-> k-step 4 q + j, column c) so that a lane's four A values of a k-step group
      // are one 16-byte (8-byte for int16) shared-memory read.  On B200 a DMMA issues every 16 cycles per SM
      // sub-core and holds that sub-core's FMA datapath meanwhile (tools/dmma_probe.cu), which is why the work is
      // spread over all compute warps instead of one dedicated warp (measured: 9.1 M -> see DESIGN 4.1).
      if (p.low_bins) {
        int nfr = p.nframes - t * kTileFrames;
        if (nfr > kTileFrames) nfr = kTileFrames;
        const int nmt = (2 * (nfr - 1) + 5 + 7) >> 3;  // row tiles that hold the blocks this tile's frames use
        if (it > 0) mbar_wait(bar_sfree, (it - 1) & 1);  // every warp has assembled the previous tile's frames
        const int r = lane >> 2, c = lane & 3;
        const double2* bwt = reinterpret_cast<const double2*>(smem + p.off_lowbw) + lane;
#pragma unroll 1
        for (int mt = cw; mt < nmt; mt += kWsCWarps) {
          double a00 = 0.0, a01 = 0.0, a10 = 0.0, a11 = 0.0;  // two accumulators: even / odd k-steps
          const int loc = kLowBlock * (8 * mt + r) + 4 * c;  // tile-local index of this lane's first sample
          const T* q0 = raw + kLead + loc;
          const bool inside = s0 + kLowBlock * 8 * (mt + 1) <= n;  // warp-uniform: the row tile ends inside the clip
#pragma unroll
          for (int q = 0; q < 5; ++q) {
            float x[4];
            ld_quad(q0 + 16 * q, x);
            const float xm = ld_one(q0 + 16 * q - 1);
            float y[4];
            y[0] = __fsub_rn(x[0], __fmul_rn(pre, xm));
            y[1] = __fsub_rn(x[1], __fmul_rn(pre, x[0]));
            y[2] = __fsub_rn(x[2], __fmul_rn(pre, x[1]));
            y[3] = __fsub_rn(x[3], __fmul_rn(pre, x[2]));
            if (!inside) {  // samples >= N are zero AFTER pre-emphasis (psf pads y)
              const int g = s0 + loc + 16 * q;
#pragma unroll
              for (int j = 0; j < 4; ++j) y[j] = (g + j < n) ? y[j] : 0.f;
            }
            const double2 b01 = bwt[(2 * q) * 32], b23 = bwt[(2 * q + 1) * 32];
            dmma884(a00, a01, (double)y[0], b01.x);
            dmma884(a10, a11, (double)y[1], b01.y);
            dmma884(a00, a01, (double)y[2], b23.x);
            dmma884(a10, a11, (double)y[3], b23.y);
          }
          *reinterpret_cast<double2*>(s_S + ((8 * mt + r) * kLowBins + c) * 2) = make_double2(a00 + a10, a01 + a11);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_sfull);
      }
      if (!active) {
        if (lane == 0) {
          mbar_arrive(bar_empty + 8u * (it & 1));
          if (p.low_bins) mbar_arrive(bar_sfree);
        }
      }
      V lm[ROUNDS];
      V mean = fsplat(0.f), esum = fsplat(0.f);
      if (active) {
        // ---- pass-1 loads straight from the raw stage with pre-emphasis on the fly:
        //      z[16 m + l] = (y[32m + 2l], y[32m + 2l + 1]),  y[i] = x[i] - pre * x[i-1]  (fp32, the
        //      product rounded before the subtraction = psf's float32 step); frame A in .x, B in .y
        const T* qa = raw + kLead + fa_local * hop + 2 * l;
        const int gA = s0 + fa_local * hop + 2 * l;  // clip sample index of this lane's first sample
        CF x16[16];
        if (gA - 2 * l + hop + LOAD_LEN <= n) {      // both frames entirely inside the clip (half-warp uniform)
#pragma unroll
          for (int m = 0; m < 16; ++m) {
            x16[m].re = fsplat(0.f);
            x16[m].im = fsplat(0.f);
            if (32 * m + 31 < LOAD_LEN || (32 * m < LOAD_LEN && 32 * m + 2 * l + 1 < LOAD_LEN)) {
              const float2 a = ld_pair(qa + 32 * m);
              const float2 b = ld_pair(qa + 32 * m + hop);
              const float am = ld_one(qa + 32 * m - 1), bm = ld_one(qa + 32 * m + hop - 1);
              x16[m].re.x = __fsub_rn(a.x, __fmul_rn(pre, am));
              x16[m].im.x = __fsub_rn(a.y, __fmul_rn(pre, a.x));
              x16[m].re.y = __fsub_rn(b.x, __fmul_rn(pre, bm));
              x16[m].im.y = __fsub_rn(b.y, __fmul_rn(pre, b.x));
            }
          }
        } else {  // the clip ends inside these frames: samples >= N are zero AFTER pre-emphasis (psf pads y)
#pragma unroll
          for (int m = 0; m < 16; ++m) {
            x16[m].re = fsplat(0.f);
            x16[m].im = fsplat(0.f);
            if (32 * m + 31 < LOAD_LEN || (32 * m < LOAD_LEN && 32 * m + 2 * l + 1 < LOAD_LEN)) {
              const int g = gA + 32 * m;
              const float2 a = ld_pair(qa + 32 * m);
              const float2 b = ld_pair(qa + 32 * m + hop);
              const float am = ld_one(qa + 32 * m - 1), bm = ld_one(qa + 32 * m + hop - 1);
              x16[m].re.x = (g < n) ? __fsub_rn(a.x, __fmul_rn(pre, am)) : 0.f;
              x16[m].im.x = (g + 1 < n) ? __fsub_rn(a.y, __fmul_rn(pre, a.x)) : 0.f;
              x16[m].re.y = (g + hop < n) ? __fsub_rn(b.x, __fmul_rn(pre, bm)) : 0.f;
              x16[m].im.y = (g + hop + 1 < n) ? __fsub_rn(b.y, __fmul_rn(pre, b.x)) : 0.f;
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + 8u * (it & 1));  // this warp no longer reads the raw stage
        dft16(x16);
#pragma unroll
        for (int k1 = 0; k1 < 16; ++k1) {
          CF y = x16[Q16(k1)];
          if (k1 > 0) {
            CF z;
            z.re = ffma(fneg(y.im), fsplat(twi[k1]), fmul(y.re, fsplat(twr[k1])));
            z.im = ffma(y.im, fsplat(twr[k1]), fmul(y.re, fsplat(twi[k1])));
            y = z;
          }
          s_tr[k1 * 17 + l] = make_float4(y.re.x, y.re.y, y.im.x, y.im.y);
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float4 v = s_tr[l * 17 + j];
          x16[j].re = make_float2(v.x, v.y);
          x16[j].im = make_float2(v.z, v.w);
        }
        __syncwarp();
        dft16(x16);  // Z[l + 16 k2] at x16[Q16(k2)]

        // ---- real-FFT split, two bins per step: k = l + 16 k2 and 256 - k ----
        const int partner = (lane & 16) | ((16 - l) & 15);
        const bool l0 = (l == 0);
        const V sc = fsplat(power_scale<T>() / 2048.0f);
        // bins 1..4 come from the low-bin warp (float64 DFT): lane l = 1..4 of the first step owns them
        // ---- bins 1..4 in float64, part 2: X_f[k] = sum_{j<5} e^{-2 pi i 80 j k / 512} S_{2f+j}[k]; lanes 0..15 take
        //      (frame lane/4 of this warp, bin lane%4 + 1), the powers are shuffled to the lanes that own those bins ----
        V lowp = fsplat(0.f);
        const bool use_low = p.low_bins && l >= 1 && l <= kLowBins;
        if (p.low_bins) {
          mbar_wait(bar_sfull, it & 1);
          float pw = 0.f;
          if (lane < 16) {
            const int f = cw * 4 + (lane >> 2), b = lane & 3;
            const double2* fw = reinterpret_cast<const double2*>(smem + p.off_lowfw) + b * 5;
            double xr = 0.0, xi = 0.0;
#pragma unroll
            for (int j = 0; j < 5; ++j) {
              const double2 sv = *reinterpret_cast<const double2*>(s_S + ((2 * f + j) * kLowBins + b) * 2);
              const double2 w = fw[j];
              xr = fma(sv.x, w.x, fma(-sv.y, w.y, xr));
              xi = fma(sv.x, w.y, fma(sv.y, w.x, xi));
            }
            pw = (float)((xr * xr + xi * xi) * ((double)power_scale<T>() / 512.0));
          }
          const int src = ((2 * half) << 2) | ((l - 1) & 3);
          lowp.x = __shfl_sync(0xffffffffu, pw, src);
          lowp.y = __shfl_sync(0xffffffffu, pw, src + 4);
          if (lane == 0) mbar_arrive(bar_sfree);
        }
#pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) {
          const CF zk = x16[Q16(k2)];
          const CF own = x16[Q16((16 - k2) & 15)];
          const CF oth = x16[Q16(15 - k2)];
          CF zp;
          zp.re.x = __shfl_sync(0xffffffffu, l0 ? own.re.x : oth.re.x, partner);
          zp.re.y = __shfl_sync(0xffffffffu, l0 ? own.re.y : oth.re.y, partner);
          zp.im.x = __shfl_sync(0xffffffffu, l0 ? own.im.x : oth.im.x, partner);
          zp.im.y = __shfl_sync(0xffffffffu, l0 ? own.im.y : oth.im.y, partner);
          const float2 w = s_twp[l + 16 * k2];
          const V wr = fsplat(w.x), wi = fsplat(w.y);
          const V sr = fadd(zk.re, zp.re), si = fsub(zk.im, zp.im);
          const V dr = fsub(zk.re, zp.re), di = fadd(zk.im, zp.im);
          const V tr = ffma(di, wr, fmul(dr, wi));          // T = -i W D
          const V ti = ffma(di, wi, fneg(fmul(dr, wr)));
          const V ar = fadd(sr, tr), ai = fadd(si, ti);     // 2 X[k]
          const V br = fsub(sr, tr), bi = fsub(si, ti);     // conj(2 X[256-k])
          V pk = fmul(ffma(ai, ai, fmul(ar, ar)), sc);
          const V pm = fmul(ffma(bi, bi, fmul(br, br)), sc);
          if (k2 == 0 && use_low) pk = lowp;
          s_P[l + 16 * k2] = pk;
          s_P[256 - l - 16 * k2] = pm;
          esum = fadd(esum, fadd(pk, pm));
        }
        if (l0) {  // bin 128 is its own mirror: |X[128]|^2 = |Z[128]|^2
          const CF z8 = x16[Q16(8)];
          const V p8 = fmul(ffma(z8.im, z8.im, fmul(z8.re, z8.re)), fsplat(power_scale<T>() / 512.0f));
          s_P[128] = p8;
          esum = fadd(esum, p8);
        }
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) {
          esum.x += __shfl_xor_sync(0xffffffffu, esum.x, o);
          esum.y += __shfl_xor_sync(0xffffffffu, esum.y, o);
        }
        __syncwarp();

        // ---- mel bank by segment ----
        // bin k of segment [e_s, e_{s+1}) feeds filter s with r = (k - e_s)/len (psf's rising edge)
        // and filter s-1 with 1 - r (psf's falling edge): R = sum r P, F = sum P - R.  Four bins per
        // step keep several shared-memory loads in flight.
        for (int s = l; s <= p.M; s += 16) {
          const int k0 = s_edges[s], k1 = s_edges[s + 1];
          const float inv = s_inv[s];
          V a0 = fsplat(0.f), a1 = fsplat(0.f), r0 = fsplat(0.f), r1 = fsplat(0.f);
          float w = 0.f;
          for (int k = k0; k < k1; k += 4) {
            const V z = fsplat(0.f);
            const V p0 = s_P[k];
            const V p1 = (k + 1 < k1) ? s_P[k + 1] : z;
            const V p2 = (k + 2 < k1) ? s_P[k + 2] : z;
            const V p3 = (k + 3 < k1) ? s_P[k + 3] : z;
            a0 = fadd(a0, fadd(p0, p1));
            a1 = fadd(a1, fadd(p2, p3));
            r0 = ffma(p0, fsplat(w * inv), r0);
            r1 = ffma(p1, fsplat((w + 1.f) * inv), r1);
            r0 = ffma(p2, fsplat((w + 2.f) * inv), r0);
            r1 = ffma(p3, fsplat((w + 3.f) * inv), r1);
            w += 4.f;
          }
          const V a = fadd(a0, a1), r = fadd(r0, r1);
          s_R[s] = r;
          s_F[s] = fsub(a, r);
        }
        __syncwarp();
        V msum = fsplat(0.f);
#pragma unroll
        for (int r = 0; r < ROUNDS; ++r) {
          const int j = l + 16 * r;
          lm[r] = fsplat(0.f);
          if (j < p.M) {
            V m = fadd(s_R[j], s_F[j + 1]);
            m.x = (m.x == 0.f) ? kFEps : m.x;
            m.y = (m.y == 0.f) ? kFEps : m.y;
            lm[r] = make_float2(fast_log(m.x), fast_log(m.y));
            msum = fadd(msum, lm[r]);
          }
        }
        if (!do_dct) {
#pragma unroll
          for (int r = 0; r < ROUNDS; ++r) {
            const int j = l + 16 * r;
            if (j < p.M) {
              if (fA < p.nframes) out_base[((size_t)clip * p.nframes + fA) * p.M + j] = lm[r].x;
              if (fB < p.nframes) out_base[((size_t)clip * p.nframes + fB) * p.M + j] = lm[r].y;
            }
          }
        } else {
#pragma unroll
          for (int o = 8; o >= 1; o >>= 1) {
            msum.x += __shfl_xor_sync(0xffffffffu, msum.x, o);
            msum.y += __shfl_xor_sync(0xffffffffu, msum.y, o);
          }
          mean = fmul(msum, fsplat(1.0f / (float)p.M));
        }
      }

      if (do_dct) {
        // the previous tile's MMAs must have retired before its A rows are overwritten
        if (it > 0) mbar_wait(bar_mma, (it - 1) & 1);
        if (active) {
          // A rows fa_local (x) and fa_local+1 (y): centred log-mel split into tf32 hi + lo
          unsigned char* abase = smem + p.off_a;
#pragma unroll
          for (int r = 0; r < ROUNDS; ++r) {
            const int nn = l + 16 * r;
            if (nn < p.M) {
              const V xc = fsub(lm[r], mean);
              const float hx = to_tf32(xc.x), hy = to_tf32(xc.y);
              const float lx = to_tf32(xc.x - hx), ly = to_tf32(xc.y - hy);
              const uint32_t col = (uint32_t)(nn >> 2) * 128u + (uint32_t)(nn & 3) * 4u;
              const uint32_t oa = (uint32_t)(fa_local >> 3) * sbo + (uint32_t)(fa_local & 7) * 16u + col;
              const uint32_t ob = (uint32_t)((fa_local + 1) >> 3) * sbo + (uint32_t)((fa_local + 1) & 7) * 16u + col;
              *reinterpret_cast<float*>(abase + oa) = hx;
              *reinterpret_cast<float*>(abase + ob) = hy;
              *reinterpret_cast<float*>(abase + (uint32_t)kAGroups * sbo + oa) = lx;
              *reinterpret_cast<float*>(abase + (uint32_t)kAGroups * sbo + ob) = ly;
            }
          }
          if (l == 0) {
            const float eA = (esum.x == 0.f) ? kFEps : esum.x, eB = (esum.y == 0.f) ? kFEps : esum.y;
            const float sq = sqrtf((float)p.M);
            float* c0 = s_c0 + (it & 1) * kTileFrames + fa_local;
            c0[0] = p.append_energy ? fast_log(eA) : sq * mean.x;  // else: the removed mean only moves c0
            c0[1] = p.append_energy ? fast_log(eB) : sq * mean.y;
          }
        }
        fence_proxy_async();  // generic-proxy writes of A must be visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_afull);
      }
    }
  }
  __syncthreads();
  if (CHUNK) {  // the next phase of the caller's loop re-initialises these barriers
    if (tid < 9) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&s_bar[tid])) : "memory");
  } else if (warp == 0 && do_dct) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols));
  }
}

// mfcc_ws.cu
int fill_fast_params(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N, int64_t pcm_stride,
                     float* out_dev, int mode, FastParams* pp, size_t* smem_out);

}  // namespace kws
