// Tensor-core forward of the depthwise-separable network for ANY channel count (multiple of 4, <= 512) and
// block count -- the "DS-CNN-L"-style generalisation of kws/libs/models.py:122-183 (SURVEY.md a-X1, BASELINE
// config 2: 276 channels, 5 blocks) and the reference shape at feature widths the fused kernel does not
// cover.  Layer by layer, activations NHWC in HBM (at 276 channels one layer of one clip is 1.6 MB: it cannot
// stay on chip), every GEMM on tcgen05:
//     conv1     D[pixel, c_out] = im2col[pixel, 100 taps]      . W1[c_out, tap]^T
//     block l   D[pixel, c_out] = depthwise3x3(x)[pixel, c_in] . Wpw[c_out, c_in]^T      (+bias, ReLU in the epilogue)
// M = 128 output pixels per tile (flattened over the batch), N = channels padded to 32 (two N halves when
// > 256), K streamed in chunks of 32: for every chunk the CTA copies the chunk of the pre-packed tf32 weight
// image to shared memory, computes the A chunk on the CUDA cores (the 3x3 depthwise stencil straight from
// the NHWC input through L1/L2, or the im2col gather) into the K-major UMMA layout, and one thread issues
// the chunk's tcgen05.mma kind::tf32 into the TMEM accumulator.  A and B chunks are double-buffered and
// tcgen05.commit -> mbarrier releases a buffer, so chunk k+1 is produced while chunk k multiplies.
// Epilogue: tcgen05.ld (lane = pixel) + bias + ReLU -> NHWC stores at the padded position (y+1, x+1) of the
// (H+2) x (W+2) output; the border pixels, ReLU(pointwise bias) (models.py:104-106), are written by
// border_fill_kernel.  The depthwise bias is folded through the pointwise weights on the host.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "mfcc_dev.cuh"

namespace kws {

constexpr int kGT = 512;                   // threads per CTA
constexpr int kGKC = 32;                   // K chunk (4 MMA k-steps)
constexpr int kGSbo = (kGKC / 4) * 128;    // 8-row group pitch of the B (weight) chunks
constexpr int kGLboA = 144;                // K-chunk pitch of the A chunks (bank spreading for the stencil stores)
constexpr int kGSboA = (kGKC / 4) * kGLboA;
constexpr int kGABytes = 16 * kGSboA;      // A chunk: 128 pixels x 32 k

struct GemmParams {
  const float* in;     // MODE 1: [B,H,W,C] NHWC block input; MODE 0: features [B,T,F]
  float* out;          // MODE 1: [B,H+2,W+2,C]; MODE 0: [B,H,W,C]
  const float* wimg;   // weight chunks, tf32, K-major: [nchunks][npad/8][kGSbo bytes]
  const float* dw_w;   // MODE 1: [9][C] depthwise taps
  const float* bias;   // [C] epilogue bias
  long long B;
  int H, W, C, T, F;   // H x W: output (interior) grid of the layer
  int R, nbands;       // tile = R whole rows of one clip (R * W <= 128), nbands = ceil(H / R)
  int kvalid, npad, nh, nhalf, nchunks, tmem_cols, patch_bytes;
  int nslots;  // dscnn_block_ws_kernel: accumulator slots of nhalf TMEM columns
  int fill_border;  // dscnn_block_ws_kernel writes the border pixels itself (warps 18, 19)
  int ldc;     // floats between consecutive pixels of the activations (C, or C padded to a multiple of 32: see launch_dscnn_gemm)
  // last block only: instead of the activation, write per-tile channel sums (the global average pool's partial sums,
  // scaled so that pool_fc_kernel's mean over nbands + 1 rows is the mean over the padded output) -- the largest
  // activation of the network is never written or re-read
  float* pool;           // [B][nbands + 1][C] or null
  const float* border;   // [C] ReLU(pointwise bias): value of every border pixel of the padded output
  const float* zero_row; // [W][C] zeros (warp-specialised kernel: input rows outside the clip)
  float* dump;           // [npad] write-only scratch row: where lanes of a tile without a pixel store
  float pool_scale;      // (nbands + 1) / ((H + 2) * (W + 2))
};

template <int MODE>
__global__ void __launch_bounds__(kGT, 1) dscnn_gemm_kernel(const GemmParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar[4];  // [0,1] MMAs of buffer retired, [2,3] weight chunk landed
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int bbytes = (p.npad / 8) * kGSbo;
  unsigned char* sA = smem;                      // [2][kGABytes]
  unsigned char* sB = smem + 2 * kGABytes;       // [2][bbytes]
  float* sW = reinterpret_cast<float*>(sB + 2 * (size_t)bbytes);  // [9][nchunks * 32] depthwise taps of the layer
  const uint32_t bar0 = smem_u32(&s_bar[0]), barB = smem_u32(&s_bar[2]);
  if (tid == 0) {
    mbar_init(bar0, 1);
    mbar_init(bar0 + 8, 1);
    mbar_init(barB, 1);
    mbar_init(barB + 8, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)),
                 "r"((uint32_t)p.tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  const int kpad = p.nchunks * kGKC;
  if (MODE == 1)
    for (int i = tid; i < 9 * kpad; i += kGT) {  // [chunk][tap][32 channels]: a chunk's taps at immediate offsets
      const int kcw = i / (9 * kGKC), r = i - kcw * 9 * kGKC, t = r >> 5, k = kcw * kGKC + (r & 31);
      sW[i] = k < p.C ? __ldg(p.dw_w + t * p.C + k) : 0.f;
    }
  for (int i = tid; i < p.npad; i += kGT) (sW + 9 * kpad)[i] = i < p.C ? __ldg(p.bias + i) : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.nhalf >> 3) << 17) | ((128u >> 4) << 24);
  const int H = p.H, W = p.W, C = p.C, R = p.R;
  const long long ntiles = p.B * p.nbands;
  // MODE 1 ownership: channel quad q of the chunk, column x, row group rg (rows rg*RG .. of the band)
  // warps 0..14 produce (stage, stencil, im2col); warp 15 only issues TMA / MMA, so that issue latency
  // (descriptor setup, the weight-chunk wait) never delays a producer's arrival at the chunk barrier
  constexpr int kProd = kGT - 32, kIssuer = kGT - 32;
  const bool producer = tid < kProd;
  const int q = tid & 7, cx = tid >> 3, NG = (kProd / 8) / W > 0 ? (kProd / 8) / W : 1, RG = (R + NG - 1) / NG;
  const int rg = cx / W, x = cx - rg * W;
  uint32_t cnt = 0;                              // chunks issued so far: buffer = cnt & 1, its phase = cnt >> 1

  // MODE 1 input staging: the (R+2) x W x 32-channel input patch of a chunk is copied to shared memory with
  // cp.async two chunks ahead of its use (3 patch buffers), so HBM/L2 latency is off the critical path and
  // every input element is fetched once per chunk (the 3x3 reuse happens in shared memory).
  float* sBias = sW + 9 * kpad;                  // [npad] epilogue bias (filled before the first __syncthreads)
  unsigned char* sP = reinterpret_cast<unsigned char*>(sBias + p.npad);
  const int patch_bytes = p.patch_bytes;
  const long long my_tiles = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long total_chunks = my_tiles * p.nchunks;
  // this thread's (up to 4) 16-byte pieces of a patch are the same for every chunk: patch row, global
  // offset relative to the first patch row, channel quad
  // The patch keeps one zero column on either side (W + 2 columns per row, written once here: the copies never
  // touch them), so the stencil reads its three columns without edge predicates.
  constexpr int kPieces = 4;
  int pf_row[kPieces], pf_off[kPieces], pf_dst[kPieces];
  const int np = (R + 2) * W * 8, prow_bytes = (W + 2) * 128;
#pragma unroll
  for (int i = 0; i < kPieces; ++i) {
    const int e = tid + i * kProd, pxl = e >> 3, ii = pxl / W, xx = pxl - ii * W;
    pf_row[i] = (producer && e < np) ? ii : -1;
    pf_off[i] = (ii * W + xx) * C + 4 * (e & 7);
    pf_dst[i] = ii * prow_bytes + (xx + 1) * 128 + (e & 7) * 16;
  }
  if (MODE == 1)
    for (int i = tid; i < 3 * (R + 2) * 16; i += kGT) {
      const int slot = i / ((R + 2) * 16), r = i - slot * (R + 2) * 16, row = r >> 4, side = (r >> 3) & 1;
      *reinterpret_cast<float4*>(sP + (size_t)slot * patch_bytes + row * prow_bytes + side * (W + 1) * 128 + (r & 7) * 16) =
          make_float4(0.f, 0.f, 0.f, 0.f);
    }
  // stencil ownership of this thread: up to 3 output rows rl0.. of column x, channel quad q
  const int rl0 = rg * RG, st_rows = (producer && rg < NG) ? min(RG, R - rl0) : 0;
  const int st_patch = rl0 * prow_bytes + x * 128 + q * 16;  // padded column x = input column x - 1
  int st_a[3];
#pragma unroll
  for (int ro = 0; ro < 3; ++ro) {
    const int pp = (rl0 + ro) * W + x;
    st_a[ro] = q * kGLboA + (pp >> 3) * kGSboA + (pp & 7) * 16;
  }
  // The prefetch stream runs two chunks ahead of the consumer; its position (chunk in tile, clip, band, patch slot)
  // is stepped incrementally -- 64-bit divisions here cost more instructions than the stencil itself.
  const int g_q = (int)(gridDim.x / (unsigned)p.nbands), g_r = (int)(gridDim.x % (unsigned)p.nbands);
  long long pf_left = total_chunks, pf_b = (long long)(blockIdx.x / (unsigned)p.nbands);
  int pf_kc = 0, pf_band = (int)(blockIdx.x % (unsigned)p.nbands), pf_slot = 0;
  auto prefetch_patch = [&]() {
    if (MODE == 1 && producer && pf_left > 0) {
      const int rp = pf_band * R;
      const uint32_t dst0 = smem_u32(sP + (size_t)pf_slot * patch_bytes);
      const float* base = p.in + ((pf_b * H + rp - 1) * W) * (long long)C + pf_kc * kGKC;  // row rp - 1 (may lie outside)
      const int kq = pf_kc * kGKC + 4 * (tid & 7);
#pragma unroll
      for (int i = 0; i < kPieces; ++i) {
        if (pf_row[i] >= 0) {
          const int yy = rp - 1 + pf_row[i];
          const bool ok = yy >= 0 && yy < H && kq < C;
          const float* src = ok ? base + pf_off[i] : p.in;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst0 + (uint32_t)pf_dst[i]), "l"(src),
                       "r"(ok ? 16 : 0)
                       : "memory");
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");  // always commit: keeps the group count in step
    --pf_left;
    pf_slot = pf_slot == 2 ? 0 : pf_slot + 1;
    if (++pf_kc == p.nchunks) {  // next tile of this CTA: tile += gridDim.x
      pf_kc = 0;
      pf_b += g_q;
      pf_band += g_r;
      if (pf_band >= p.nbands) {
        pf_band -= p.nbands;
        ++pf_b;
      }
    }
  };
  prefetch_patch();
  prefetch_patch();
  asm volatile("cp.async.wait_group 1;" ::: "memory");
  __syncthreads();
  long long fchunk = 0;
  int cslot = 0;                                 // patch slot of the chunk being consumed (= fchunk % 3)
  long long b = (long long)(blockIdx.x / (unsigned)p.nbands);
  int band = (int)(blockIdx.x % (unsigned)p.nbands);

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int r0 = band * R;
    for (int kc = 0; kc < p.nchunks; ++kc, ++cnt, ++fchunk, cslot = cslot == 2 ? 0 : cslot + 1) {
      const int buf = cnt & 1;
      prefetch_patch();
      if (cnt >= 2) mbar_wait(bar0 + 8u * buf, ((cnt - 2) >> 1) & 1);  // the MMAs that read this buffer retired
      if (tid == kIssuer && cnt == 0) {  // first weight chunk (later ones are requested one chunk ahead, see below)
        mbar_expect_tx(barB, (uint32_t)bbytes);
        tma_load_1d(smem_u32(sB), p.wimg, (uint32_t)bbytes, barB);
      }
      unsigned char* abuf = sA + buf * kGABytes;
      if (MODE == 1) {
        // depthwise 3x3 (zero padding 1) of channels k..k+3 down this thread's column: the 5 x 3 input window of
        // the row group comes from the padded patch (no edge predicates; rows past a short group are read but
        // never stored), 3 output rows are accumulated with packed f32x2 FMAs over channel pairs.  Channels past
        // the channel count were zero-filled by the copy, so the K padding needs no special case; rows of A that
        // no pixel owns keep stale data, which only reaches accumulator rows the epilogue never reads.
        if (st_rows > 0) {
          const unsigned char* pb = sP + (size_t)cslot * patch_bytes + st_patch;
          const float4* w4 = reinterpret_cast<const float4*>(sW + kc * 9 * kGKC + 4 * q);
          V wlo[9], whi[9];
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            const float4 w = w4[t * 8];
            wlo[t] = make_float2(w.x, w.y);
            whi[t] = make_float2(w.z, w.w);
          }
          float4 vin[5][3];
#pragma unroll
          for (int ii = 0; ii < 5; ++ii) {
            const float4* prow = reinterpret_cast<const float4*>(pb + ii * prow_bytes);
            vin[ii][0] = prow[0];
            vin[ii][1] = prow[8];
            vin[ii][2] = prow[16];
          }
#pragma unroll
          for (int ro = 0; ro < 3; ++ro) {  // output row rl0 + ro from input rows ro, ro+1, ro+2 of the group
            V olo = fsplat(0.f), ohi = fsplat(0.f);
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
              for (int j = 0; j < 3; ++j) {
                const float4 a = vin[ro + i][j];
                olo = ffma(make_float2(a.x, a.y), wlo[i * 3 + j], olo);
                ohi = ffma(make_float2(a.z, a.w), whi[i * 3 + j], ohi);
              }
            uint4 r;  // tf32 round-to-nearest: add half an ulp of the 10-bit mantissa, the MMA drops the low 13 bits
            r.x = __float_as_uint(olo.x) + 0x1000u;
            r.y = __float_as_uint(olo.y) + 0x1000u;
            r.z = __float_as_uint(ohi.x) + 0x1000u;
            r.w = __float_as_uint(ohi.y) + 0x1000u;
            if (ro < st_rows) *reinterpret_cast<uint4*>(abuf + st_a[ro]) = r;
          }
        }
      } else if (producer) {  // im2col of the 10x10 stride-2 padding-2 window: item = (pixel row, 8-wide K slice)
        for (int item = tid; item < 512; item += kProd) {
          const int px = item & 127, kg = item >> 7;
          const int rl = px / W, xq = px - rl * W, y = r0 + rl;
          const bool valid = rl < R && y < H;
          unsigned char* arow = abuf + (px >> 3) * kGSboA + (px & 7) * 16;
#pragma unroll
          for (int g4 = 0; g4 < 2; ++g4) {
            const int kl = kg * 8 + g4 * 4, k = kc * kGKC + kl;
            float t4[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int kk = k + e, ki = kk / 10, kj = kk - 10 * ki;
              const int fr = 2 * y - 2 + ki, fc = 2 * xq - 2 + kj;
              t4[e] = (valid && kk < p.kvalid && fr >= 0 && fr < p.T && fc >= 0 && fc < p.F) ? __ldg(p.in + (b * p.T + fr) * p.F + fc)
                                                                                              : 0.f;
            }
            uint4 r;
            r.x = __float_as_uint(t4[0]) + 0x1000u;
            r.y = __float_as_uint(t4[1]) + 0x1000u;
            r.z = __float_as_uint(t4[2]) + 0x1000u;
            r.w = __float_as_uint(t4[3]) + 0x1000u;
            *reinterpret_cast<uint4*>(arow + (kl >> 2) * kGLboA) = r;
          }
        }
      }
      fence_proxy_async();
      asm volatile("cp.async.wait_group 1;" ::: "memory");  // this thread's share of the NEXT chunk's patch has landed
      __syncthreads();
      if (warp == kIssuer / 32) {  // the issue warp runs converged; one elected lane issues (tc_mma_tf32_elect)
        mbar_wait(barB + 8u * buf, (cnt >> 1) & 1);
        tc_fence_after();
        const uint32_t ab = smem_u32(abuf), bb = smem_u32(sB + (size_t)buf * bbytes);
#pragma unroll 1
        for (int ks = 0; ks < kGKC / 8; ++ks)
          for (int hf = 0; hf < p.nh; ++hf)
            tc_mma_tf32_elect(tmem + (uint32_t)(hf * p.nhalf), umma_desc(ab + ks * 2u * kGLboA, kGLboA, kGSboA),
                              umma_desc(bb + (uint32_t)(hf * (p.nhalf / 8)) * kGSbo + ks * 256u, 128u, kGSbo), idesc,
                              (kc > 0 || ks > 0) ? 1u : 0u);
        tc_commit_elect(bar0 + 8u * buf);
        // request the NEXT chunk's weights now (TMA bulk copy), a full chunk ahead of their use: their buffer
        // is free once the MMAs of chunk cnt - 1 have retired, which happened while this chunk was produced
        if (fchunk + 1 < total_chunks) {
          const int nb2 = buf ^ 1, kn = kc + 1 < p.nchunks ? kc + 1 : 0;
          if (cnt >= 1) mbar_wait(bar0 + 8u * nb2, ((cnt - 1) >> 1) & 1);
          if (lane == 0) {
            mbar_expect_tx(barB + 8u * nb2, (uint32_t)bbytes);
            tma_load_1d(smem_u32(sB + (size_t)nb2 * bbytes),
                        reinterpret_cast<const unsigned char*>(p.wimg) + (size_t)kn * bbytes, (uint32_t)bbytes,
                        barB + 8u * nb2);
          }
          __syncwarp();
        }
      }
    }
    {  // all MMAs of the tile have retired when the last chunk's commit arrives
      const uint32_t last = cnt - 1;
      mbar_wait(bar0 + 8u * (last & 1), (last >> 1) & 1);
      tc_fence_after();
    }
    {  // epilogue on all 16 warps: lane = pixel row of the tile (TMEM lane quarter warp % 4), 16-column groups
       // warp / 4, warp / 4 + 4, ...
      const int row = (warp & 3) * 32 + lane;
      const int rl = row / W, xq = row - rl * W, y = r0 + rl;
      float* orow = nullptr;
      if (rl < R && y < H)
        orow = MODE == 1 ? p.out + ((b * (H + 2) + y + 1) * (W + 2) + xq + 1) * (long long)C
                         : p.out + ((b * H + y) * W + xq) * (long long)C;
      const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16);
      float* s_part = reinterpret_cast<float*>(sA);  // [4 lane quarters][npad] (the A buffers are idle: all MMAs retired)
      for (int c = (warp >> 2) * 16; c < p.npad; c += 64) {
        float v[16];
        tc_ld16(taddr + (uint32_t)c, v);
        tc_wait_ld();
        if (MODE == 1 && p.pool) {  // channel sums over the tile's pixels: lanes without a pixel contribute zero
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            float r = orow ? fmaxf(v[e] + sBias[c + e], 0.f) : 0.f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
            v[e] = r;
          }
          if (lane == 0) {
#pragma unroll
            for (int e = 0; e < 16; e += 4)
              *reinterpret_cast<float4*>(s_part + (warp & 3) * p.npad + c + e) = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
          }
        } else if (orow) {
#pragma unroll
          for (int e = 0; e < 16; e += 4) {
            if (c + e < C) {
              const float4 bz = *reinterpret_cast<const float4*>(sBias + c + e);
              *reinterpret_cast<float4*>(orow + c + e) =
                  make_float4(fmaxf(v[e] + bz.x, 0.f), fmaxf(v[e + 1] + bz.y, 0.f), fmaxf(v[e + 2] + bz.z, 0.f),
                              fmaxf(v[e + 3] + bz.w, 0.f));
            }
          }
        }
      }
      if (MODE == 1 && p.pool) {
        __syncthreads();
        float* prow = p.pool + (b * (p.nbands + 1) + band) * (long long)C;
        for (int c = tid; c < C; c += kGT) {
          prow[c] = ((s_part[c] + s_part[p.npad + c]) + (s_part[2 * p.npad + c] + s_part[3 * p.npad + c])) * p.pool_scale;
          if (band == 0)  // one extra row per clip: the 2 (W + 2) + 2 H border pixels, all ReLU(bias)
            p.pool[(b * (p.nbands + 1) + p.nbands) * (long long)C + c] =
                (float)(2 * (W + 2) + 2 * H) * __ldg(p.border + c) * p.pool_scale;
        }
      }
    }
    tc_fence_before();
    __syncthreads();  // the next tile's first MMA overwrites the accumulator
    b += g_q;
    band += g_r;
    if (band >= p.nbands) {
      band -= p.nbands;
      ++b;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)p.tmem_cols));
}

// ---- warp-specialised block kernel ------------------------------------------------------------------------
// One depthwise-separable block (depthwise 3x3 -> pointwise 1x1 -> bias -> ReLU) as an asynchronous pipeline; the
// chunk-synchronous kernel above stays for conv1 (MODE 0) and for shapes outside this one's thread map.
//
// What bounds the block layers is the data path around the tensor core, not the tensor core: a tf32 MMA of
// M = 128, N = 144, K = 8 reads 8.6 KB of operands for 72 cycles of math, the weight chunks are written once and
// read once per 128 pixels, and the chunk-synchronous kernel on top of that staged every input patch in shared
// memory and read each element five times from there, with all 16 warps meeting at a CTA barrier per K chunk
// (ncu: 4.9 K cycles per chunk, tensor pipe 16 % active, barrier + LSU-queue stalls on top).  Here:
//   * no input patch in shared memory: a producer thread owns (channel pair, NC adjacent columns) of the tile and
//     walks DOWN the rows with the 3x3 window in registers -- one 8-byte global load per input pixel and channel
//     pair (NC + 2 per row, L1 catches the neighbour thread's overlap), each input row updates the three output
//     rows it touches (out = P + w2.v, P = R + w1.v, R = w0.v), so nothing is read twice and nothing is exchanged;
//     out-of-range columns load a clamped (valid) address and are zeroed by a 0/1 multiply when used, rows outside
//     the clip take a warp-uniform branch: no per-load predicates or address selects in the row loop;
//   * four producer groups of 3 warps work on four DIFFERENT K chunks at the same time; item n goes into A-ring stage
//     n % 6 (full / empty mbarriers, tcgen05.commit releases a stage), so there is no CTA-wide barrier anywhere;
//   * the issue warp (16) fires a chunk's MMAs -- the first N half's four, then the second half's -- and pulls the next
//     tile's whole input patch (one contiguous range) into L2 with bulk prefetches half a tile ahead; the loader warp
//     (17) streams the weight chunks through a 3-stage TMA ring; the two spare warps of that warpgroup (18, 19) write
//     the output's border pixels;
//   * the accumulator is drained by four warps of their own (12..15, one TMEM lane quarter each) out of a ring of
//     accumulator slots, so the issue warp (16) never drains and a tile's MMAs overlap the previous tile's drain:
//     with two N halves of 144 columns three half-slots fit the 512 TMEM columns (tile j's halves in slots 2j % 3
//     and (2j + 1) % 3), one-half layers get two whole slots, halves wider than 170 columns two half-slots (each
//     half then waits for the previous tile's same half only).  Five warps per SM sub-partition: 96 registers each at
//     launch, rebalanced with setmaxnreg to 120 (producers) / 96 (drain) / 24 (the last warpgroup).
constexpr int kWsG = 4;     // producer groups
#ifndef KWS_WS_S
#define KWS_WS_S 6
#define KWS_WS_BS 3
#endif
constexpr int kWsS = KWS_WS_S;     // A-ring stages (item n uses stage n % kWsS: one more than groups, so that production runs on
                            // while the accumulator drains and the MMAs of the four waiting chunks catch up)
constexpr int kWsGT = 96;   // threads per group: 16 channel pairs x 6 column groups
constexpr int kWsBS = KWS_WS_BS;    // weight-chunk ring stages
// A-ring stage layout (K-major, no swizzle): K quad q of pixel p at q * kWsLbo + 16 p -- the 8-row core matrices of a
// quad are contiguous along M (SBO = 128), so a producer's NC adjacent pixels are NC adjacent 16-byte slots and its
// store address is one register + an immediate (the chunk-synchronous kernel's layout, quads 144 bytes apart inside an
// 8-row group, cost 8 address instructions per store); the quads are 16 bytes more than 2048 apart so that the
// 16 channel pairs of a warp's store (8 quads x 2 halves) cover all 32 banks.
constexpr int kWsLbo = 16 * 128 + 16;
constexpr int kWsSbo = 128;
constexpr int kWsABytes = (kGKC / 4) * kWsLbo;  // 16512
constexpr int kWsEpi = 4;   // draining warps 12..15 (TMEM lane quarter = warp % 4)
constexpr int kWsT = 640;   // five warpgroups: 12 producer warps | 4 drain warps | the issue warp + 3 idle warps.  Every
                            // SM sub-partition holds 3 producer warps, a drain warp and a warp of the last group: launched
                            // at 96 registers, the last group shrinks to 40 (setmaxnreg) and the producers grow to 112, the
                            // drain warps to 104 (3 x 112 + 104 + 40 = 480 = 5 x 96 per sub-partition lane)

#ifdef KWS_WS_PROFILE
#include <cstdio>
#define WSP_DECL(n) long long wsp_##n = 0
#define WSP_T0() const long long wsp_t0 = clock64()
#define WSP_ADD(n) wsp_##n += clock64() - wsp_t0
#else
#define WSP_DECL(n)
#define WSP_T0()
#define WSP_ADD(n)
#endif
// waits of the issue warp / the drain warps (A/B hook: try_wait with a 2 us suspendTimeHint instead of polling took the
// polling instructions off the schedulers but woke up late: 207 K -> 201 K clips/s)
#ifdef KWS_WS_ISSUEIDLE
#define WS_WAIT(bar, par) mbar_wait_idle(bar, par, KWS_WS_ISSUEIDLE)
#else
#define WS_WAIT(bar, par) mbar_wait(bar, par)
#endif
#ifdef KWS_WS_ACCSPIN
#define WS_WAIT_ACC(bar, par) mbar_wait(bar, par)
#else
#define WS_WAIT_ACC(bar, par) mbar_wait_idle(bar, par, 32)
#endif
// NC = adjacent columns per producer thread, PF = rows the loads run ahead, POOL = last block (channel sums instead of
// the activation; its own instantiation keeps the other's drain code small: the kernel is 60 KB of SASS run by five roles)
template <int NC, int PF, bool POOL>
__global__ void __launch_bounds__(kWsT, 1) dscnn_block_ws_kernel(const GemmParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar[2 * kWsS + 2 * kWsBS + 2 + 3];
  __shared__ uint32_t s_tmem;
  __shared__ volatile int s_tile;  // tile the issue warp has reached: paces the border warps
#ifdef KWS_WS_PROFILE
  __shared__ long long s_prof[8];
#endif
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int bbytes = (p.npad / 8) * kGSbo;
  unsigned char* sA = smem;                                  // [kWsS][kWsABytes]
  unsigned char* sB = smem + kWsS * kWsABytes;                // [kWsBS][bbytes]
  const int kpad = p.nchunks * kGKC;
  float* sW = reinterpret_cast<float*>(sB + kWsBS * (size_t)bbytes);  // [chunk][tap][32 channels]
  float* sBias = sW + 9 * kpad;                              // [npad]
  const uint32_t bFull = smem_u32(&s_bar[0]), bEmpty = bFull + 8 * kWsS, bBfull = bEmpty + 8 * kWsS,
                 bBempty = bBfull + 8 * kWsBS, bAccFull = bBempty + 8 * kWsBS, bFreeS = bAccFull + 16;
  if (tid == 0) {
    for (int g = 0; g < kWsS; ++g) {
      mbar_init(bFull + 8 * g, 3);   // one arrival per producer warp of the group
      mbar_init(bEmpty + 8 * g, 1);  // tcgen05.commit
    }
    for (int s = 0; s < kWsBS; ++s) {
      mbar_init(bBfull + 8 * s, 1);
      mbar_init(bBempty + 8 * s, 1);
    }
    s_tile = -1;
    mbar_init(bAccFull, 1);  // MMAs of an even / odd tile retired
    mbar_init(bAccFull + 8, 1);
    for (int s = 0; s < 3; ++s) mbar_init(bFreeS + 8 * s, kWsEpi);  // accumulator slot read out by its four warps
    fence_mbar_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)),
                 "r"((uint32_t)p.tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (int i = tid; i < 9 * kpad; i += kWsT) {
    const int kcw = i / (9 * kGKC), r = i - kcw * 9 * kGKC, t = r >> 5, k = kcw * kGKC + (r & 31);
    sW[i] = k < p.C ? __ldg(p.dw_w + t * p.C + k) : 0.f;
  }
  for (int i = tid; i < p.npad; i += kWsT) sBias[i] = i < p.C ? __ldg(p.bias + i) : 0.f;
  // A ring starts as zeros: rows no pixel owns and channel quads past the channel count are never written, and
  // what the MMA reads there must be finite (it meets zero weight rows, or lands in accumulator rows nobody reads)
  for (int i = tid; i < kWsS * kWsABytes / 16; i += kWsT) reinterpret_cast<float4*>(sA)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const int H = p.H, W = p.W, C = p.C, R = p.R, nch = p.nchunks;
  const unsigned ntiles = (unsigned)(p.B * p.nbands);  // the host keeps B * nbands * nchunks below 2^31
  const int my_tiles = blockIdx.x < ntiles ? (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  const int total = my_tiles * nch;  // work items (tile, K chunk) of this CTA, in issue order

  // drain of tile j's accumulator by this warp: all columns of TMEM lane quarter warp % 4 (lane = pixel of the
  // tile); bias + ReLU -> NHWC store at the padded position, or (last block) per-(tile, lane quarter) channel sums
  // into the pool rows.  The loads use the 16-lane x 256-bit shape: a thread gets rows lane/4 and lane/4 + 8 of
  // the 16-lane half and columns 2 (lane % 4) + {0, 1} of every 8-column block, so the four threads of a quad hold
  // 8 consecutive channels of one pixel; after one exchange between neighbouring lanes each of them stores 16
  // bytes of a 64-byte run per pixel (the 32x32b shape -- one pixel per thread -- made every store instruction
  // touch 32 different 1104-byte-strided lines, 16 bytes each: 30 % of the kernel's L1 cycles).
  // Accumulator slots: N half h (counted over the CTA's tiles: h = nh j, nh j + 1) lives in slot h % ns at columns
  // (h % ns) nhalf; its MMAs wait until the slot's previous tenant has been read out (bFreeS[slot], one phase per use).
  const int ns = p.nslots;
  const int dr_hp = p.nhalf >> 4;                                       // 16-column pairs per N half
  const int dr_nr = p.npad >> 5;                                        // rounds of two pairs
  const int dr_ra = p.nh == 2 ? (dr_hp - 1) >> 1 : -1;                  // round whose loads complete the first of two halves
  WSP_DECL(aw);
  auto epilogue = [&](int j) {
    { WSP_T0(); WS_WAIT_ACC(bAccFull + 8 * (uint32_t)(j & 1), (uint32_t)(j >> 1) & 1u); WSP_ADD(aw); }
    tc_fence_after();
    const int s0 = (p.nh * j) % ns, s1 = (p.nh * j + 1) % ns;
    const uint32_t col0 = (uint32_t)(s0 * p.nhalf), col1 = (uint32_t)(s1 * p.nhalf);
    const uint32_t bar_a = bFreeS + 8 * (uint32_t)s0, bar_b = bFreeS + 8 * (uint32_t)(p.nh == 2 ? s1 : s0);
    const int rs = 0, re = dr_nr;
    const unsigned tile = blockIdx.x + (unsigned)j * gridDim.x, b = tile / (unsigned)p.nbands;
    const int band = (int)(tile - b * (unsigned)p.nbands), r0 = band * R;
    const int lq = warp & 3, tq = lane & 3, tr = lane >> 2;
    constexpr bool pool = POOL;
    const bool odd = lane & 1;
    // this thread's four pixels (16-lane half hh, row block ss); a lane without a pixel stores into a dump row, so
    // that the stores of all rounds but the last need no predicate
#ifdef KWS_WS_STG64
    const int coff = 2 * tq;
#else
    const int coff = odd ? 8 + 2 * (tq - 1) : 2 * tq;  // first of this lane's four channels within a 16-channel pair
#endif
    float* orow[2][2];
    bool valid[2][2];
#pragma unroll
    for (int hh = 0; hh < 2; ++hh)
#pragma unroll
      for (int ss = 0; ss < 2; ++ss) {
        const int pp = lq * 32 + hh * 16 + ss * 8 + tr, rl = pp / W, xq = pp - rl * W, y = r0 + rl;
        valid[hh][ss] = rl < R && y < H;
        orow[hh][ss] = (valid[hh][ss] ? p.out + (((long long)b * (H + 2) + y + 1) * (W + 2) + xq + 1) * (long long)p.ldc : p.dump) + coff + 32 * rs;
      }
    float* prow = pool ? p.pool + ((long long)b * (4 * p.nbands + 1) + band * 4 + lq) * (long long)C : nullptr;
    const uint32_t ta = tmem + ((uint32_t)(lq * 32) << 16);
    // 16-lane halves without a pixel (the tile's last rows of TMEM lanes, short last bands) are neither loaded nor stored
    const int npix = min(R, H - r0) * W;
    const bool hv0 = lq * 32 < npix, hv1 = lq * 32 + 16 < npix;
    auto load_round = [&](float (&v)[2][2][8], int r) {  // round r = pairs 2r, 2r + 1 of 8-column blocks x two 16-lane halves
#pragma unroll
      for (int pr = 0; pr < 2; ++pr) {
        const int i = 2 * r + pr;  // pair i holds channels 16 i ..: first or second N half
        const uint32_t tc = i < dr_hp ? col0 + 16u * (uint32_t)i : col1 + 16u * (uint32_t)(i - dr_hp);
        if (hv0) tc_ld_16x256b_x2(ta + tc, v[pr][0]);
        if (hv1) tc_ld_16x256b_x2(ta + (16u << 16) + tc, v[pr][1]);
      }
    };
    auto release = [&](int r) {  // after the loads of round r have landed in registers
      if (r == dr_ra || r == dr_nr - 1) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (r == dr_ra) mbar_arrive(bar_a);
          if (r == dr_nr - 1) mbar_arrive(bar_b);
        }
      }
    };
    // bias of a round for this thread: block 0 / block 1 of the two pairs.  Read before the wait for the round's TMEM
    // loads, so that the shared-memory latency is not paid in front of the first add
    auto load_bias = [&](float2 (&bz)[4], int r) {
      const float* b = sBias + 32 * r + 2 * tq;
#pragma unroll
      for (int i = 0; i < 4; ++i) bz[i] = *reinterpret_cast<const float2*>(b + 8 * i);
    };
    auto process = [&](float (&v)[2][2][8], int r, const float2 (&bz)[4]) {
      const int cb = 32 * r;
      const bool guard = cb + 32 > C;  // only the round that holds the channel count's end needs a store predicate
#ifdef KWS_WS_LDONLY  // timing experiment only: TMEM loads without the arithmetic and the stores
      if (v[0][0][0] != 1.2345e30f) return;
#endif
#pragma unroll
      for (int pr = 0; pr < 2; ++pr) {
        const int col = cb + 16 * pr + 2 * tq;  // channel of register e = 0 of block 0 (block 1: + 8)
        const float2 bz0 = bz[2 * pr], bz1 = bz[2 * pr + 1];
#pragma unroll
        for (int hh = 0; hh < 2; ++hh)
#pragma unroll
          for (int ss = 0; ss < 2; ++ss) {
            float* x = v[pr][hh] + 2 * ss;  // registers [4 blk + 2 s + e]
            x[0] = fmaxf(x[0] + bz0.x, 0.f);
            x[1] = fmaxf(x[1] + bz0.y, 0.f);
            x[4] = fmaxf(x[4] + bz1.x, 0.f);
            x[5] = fmaxf(x[5] + bz1.y, 0.f);
          }
        if (!pool) {
#ifdef KWS_WS_STG64  // A/B: 8-byte stores straight from the load layout (32-byte runs per pixel, no exchange)
          const bool cok0 = !guard || cb + 16 * pr + 2 * tq < C, cok1 = !guard || cb + 16 * pr + 8 + 2 * tq < C;
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            if (hh == 0 ? hv0 : hv1) {
#pragma unroll
              for (int ss = 0; ss < 2; ++ss) {
                const float* x = v[pr][hh] + 2 * ss;
                float* o = orow[hh][ss] + 16 * pr;
                if (cok0) *reinterpret_cast<float2*>(o) = make_float2(x[0], x[1]);
                if (cok1) *reinterpret_cast<float2*>(o + 8) = make_float2(x[4], x[5]);
              }
            }
          }
#else
          const bool cok = !guard || cb + 16 * pr + coff < C;
          // even lanes keep block 0 and take the neighbour's, odd lanes block 1: the eight shuffles of the pair are
          // issued back to back, then the four 16-byte stores (one shuffle latency per pair instead of one per store)
          float g[2][2][2];
#pragma unroll
          for (int hh = 0; hh < 2; ++hh)
#pragma unroll
            for (int ss = 0; ss < 2; ++ss) {
              const float* x = v[pr][hh] + 2 * ss;
              g[hh][ss][0] = __shfl_xor_sync(0xffffffffu, odd ? x[0] : x[4], 1);
              g[hh][ss][1] = __shfl_xor_sync(0xffffffffu, odd ? x[1] : x[5], 1);
            }
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            if (hh == 0 ? hv0 : hv1) {
#pragma unroll
              for (int ss = 0; ss < 2; ++ss) {
                const float* x = v[pr][hh] + 2 * ss;
                const float k0 = odd ? x[4] : x[0], k1 = odd ? x[5] : x[1], g0 = g[hh][ss][0], g1 = g[hh][ss][1];
                const float4 o = odd ? make_float4(g0, g1, k0, k1) : make_float4(k0, k1, g0, g1);
#ifdef KWS_WS_NOSTORE  // timing experiment only (results are wrong): the drain without its global stores
                if (o.x == 1.2345e30f) *reinterpret_cast<float4*>(orow[hh][ss] + 16 * pr) = o;
#else
                if (cok) *reinterpret_cast<float4*>(orow[hh][ss] + 16 * pr) = o;
#endif
              }
            }
          }
#endif
        } else {  // channel sums over the warp's 32 pixels in a fixed order: 4 in the thread, then lanes 4, 8, 16 apart
#pragma unroll
          for (int be = 0; be < 4; ++be) {  // be = 2 blk + e
            const int r = 4 * (be >> 1) + (be & 1);
            float sum = ((valid[0][0] ? v[pr][0][r] : 0.f) + (valid[0][1] ? v[pr][0][r + 2] : 0.f)) +
                        ((valid[1][0] ? v[pr][1][r] : 0.f) + (valid[1][1] ? v[pr][1][r + 2] : 0.f));
            sum += __shfl_xor_sync(0xffffffffu, sum, 4);
            sum += __shfl_xor_sync(0xffffffffu, sum, 8);
            sum += __shfl_xor_sync(0xffffffffu, sum, 16);
            const int c1 = col + 8 * (be >> 1) + (be & 1);
            if (tr == 0 && c1 < C) prow[c1] = sum * p.pool_scale;
          }
        }
      }
#pragma unroll
      for (int hh = 0; hh < 2; ++hh)
#pragma unroll
        for (int ss = 0; ss < 2; ++ss) orow[hh][ss] += 32;
    };
    if (!hv0 && !pool) {  // nothing of this tile lives in this lane quarter
      for (int r = rs; r < re; ++r) release(r);
    } else if (rs < re) {  // two register sets in rotation: the next round's TMEM loads are in flight while a round is stored
      float va[2][2][8], vb[2][2][8];
      float2 bz[4];
      load_round(va, rs);
      for (int r = rs;; r += 2) {
        load_bias(bz, r);
        tc_wait_ld();
        if (r + 1 < re) load_round(vb, r + 1);
        release(r);
        process(va, r, bz);
        if (r + 1 >= re) break;
        load_bias(bz, r + 1);
        tc_wait_ld();
        if (r + 2 < re) load_round(va, r + 2);
        release(r + 1);
        process(vb, r + 1, bz);
        if (r + 2 >= re) break;
      }
    }
    if (pool && band == 0 && lq == 0)  // one extra row per clip: the 2 (W + 2) + 2 H border pixels, all ReLU(bias)
      for (int c = lane; c < C; c += 32)
        p.pool[((long long)b * (4 * p.nbands + 1) + 4 * p.nbands) * (long long)C + c] =
            (float)(2 * (W + 2) + 2 * H) * __ldg(p.border + c) * p.pool_scale;
  };

  // The whole input patch of a tile (rows r0 - 1 .. r0 + R of one clip, all channels) is one contiguous range:
  // pulled into L2 one tile ahead by bulk prefetches, one per row, so that the producers' register loads (two
  // rows ahead of their arithmetic, which hides L2 latency but not DRAM latency under load) hit L2.
  auto prefetch_tile = [&](int jj) {
    if (jj < my_tiles) {
      const unsigned tile = blockIdx.x + (unsigned)jj * gridDim.x, b = tile / (unsigned)p.nbands;
      const int r0 = (int)(tile - b * (unsigned)p.nbands) * R;
      const int y0 = r0 > 0 ? r0 - 1 : 0, y1 = r0 + R < H ? r0 + R : H - 1;
      const uint32_t bytes = (uint32_t)(W * p.ldc * 4);
#pragma unroll 1
      for (int y = y0 + lane; y <= y1; y += 32)
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.in + ((long long)b * H + y) * W * (long long)p.ldc), "r"(bytes)
                     : "memory");
    }
    __syncwarp();
  };
  if (warp < kWsG * 3) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 120;");
    // ---- producers ----
    // thread = (channel PAIR d of the chunk, NC adjacent columns): 16 pairs x 6 column groups.  With two channels per
    // thread the nine depthwise taps are 18 registers and stay resident for the item (four channels per thread had to
    // re-read them from shared memory every row: 36 of its 44 wavefronts per row, which is what the MMAs' operand reads
    // were competing with), and a row of the window is NC + 2 eight-byte loads.
    const int g = warp / 3, gt = tid - g * kWsGT;
    const int d = gt & 15, cg = gt >> 4, x0 = cg * NC;
    uint32_t abase0 = smem_u32(sA) + (uint32_t)((d >> 1) * kWsLbo + (d & 1) * 8);
    asm volatile("" : "+r"(abase0));  // opaque, like coff[] below
    // column c of the window = input column x0 - 1 + c: byte offset within a row (clamped into the row) and 0/1 mask
    // the loads are (warp-uniform row pointer) + (this thread's 32-bit offset): no 64-bit address arithmetic per load
    uint32_t coff[NC + 2];
    float cm[NC + 2];
#pragma unroll
    for (int c = 0; c < NC + 2; ++c) {
      const int xx = x0 - 1 + c, xc = xx < 0 ? 0 : (xx >= W ? W - 1 : xx);
      coff[c] = (uint32_t)(xc * p.ldc * 4 + d * 8);
      cm[c] = xx == xc ? 1.f : 0.f;
      asm volatile("" : "+r"(coff[c]));  // opaque: keeps ptxas from recomputing the offset at every load (-5 % time)
    }
    const long long rowbytes = (long long)W * p.ldc * 4;
    uint32_t k = 0;
    WSP_DECL(pw);
#ifdef KWS_WS_PROFILE
    const long long wsp_start = clock64();
#endif
    for (int n = g; n < total; n += kWsG, ++k) {
      const int j = n / nch, kc = n - j * nch;
      const int stage = n % kWsS, use = n / kWsS;  // ring stage of the item and how often it has been used before
      const uint32_t abase = abase0 + (uint32_t)(stage * kWsABytes);
      const unsigned tile = blockIdx.x + (unsigned)j * gridDim.x, b = tile / (unsigned)p.nbands;
      const int band = (int)(tile - b * (unsigned)p.nbands), r0 = band * R;
      const int nrows = min(R, H - r0) + 2;  // input rows r0 - 1 .. of the band (rows outside the clip are zeros)
      const int kq = kc * kGKC + 2 * d;
      const bool act = x0 < W && kq < C;
      // next row to load: row r0 - 1 of the clip, first channel of the chunk (the pointer may run outside the clip:
      // never read there); all of this is warp-uniform
      const char* rp = reinterpret_cast<const char*>(p.in + ((long long)b * H + r0 - 1) * W * (long long)p.ldc + kc * kGKC);
      int ynext = r0 - 1;
      const char* zr = reinterpret_cast<const char*>(p.zero_row + kc * kGKC);  // a row of zeros stands in for rows outside the clip
      auto load_row = [&](V (&v)[NC + 2]) {
        const char* r = (ynext >= 0 && ynext < H) ? rp : zr;
#pragma unroll
        for (int c = 0; c < NC + 2; ++c) v[c] = __ldg(reinterpret_cast<const float2*>(r + coff[c]));
        rp += rowbytes;
        ++ynext;
      };
      V vbuf[PF + 1][NC + 2];
      V w[9];
      if (act) {
#pragma unroll
        for (int r = 0; r < PF; ++r) load_row(vbuf[r]);
        const float2* w2 = reinterpret_cast<const float2*>(sW + kc * 9 * kGKC + 2 * d);
#pragma unroll
        for (int t = 0; t < 9; ++t) w[t] = w2[t * 16];
      }
      {
        WSP_T0();
        if (use >= 1) mbar_wait(bEmpty + 8 * stage, (uint32_t)(use - 1) & 1u);  // the MMAs that read this stage last have retired
        WSP_ADD(pw);
      }
      if (act) {
        V P[NC], Rr[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) P[c] = Rr[c] = fsplat(0.f);
        uint32_t arow = abase + 16u * (uint32_t)(-2 * W + x0);  // A-ring slot of (output row i - 2, column x0)
        // input row i of the band: finishes output row i - 2 (taps 6..8), continues i - 1 (3..5), starts i (0..2)
        auto step = [&](int i, const V (&vin)[NC + 2]) {
          V v[NC + 2];
#pragma unroll
          for (int c = 0; c < NC + 2; ++c) v[c] = c == 1 ? vin[c] : fmul(vin[c], fsplat(cm[c]));  // column x0 always exists
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            V o = P[c], q1 = Rr[c], q0;
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) {
              o = ffma(v[c + dx], w[6 + dx], o);
              q1 = ffma(v[c + dx], w[3 + dx], q1);
              q0 = dx == 0 ? fmul(v[c], w[0]) : ffma(v[c + dx], w[dx], q0);
            }
            P[c] = q1;
            Rr[c] = q0;
            if (i >= 2 && x0 + c < W) {  // tf32 round-to-nearest: add half an ulp of the 10-bit mantissa
              asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(arow + 16u * (uint32_t)c), "r"(__float_as_uint(o.x) + 0x1000u),
                           "r"(__float_as_uint(o.y) + 0x1000u)
                           : "memory");
            }
          }
          arow += 16u * (uint32_t)W;
        };
        [&] {
          int i = 0;
          while (true) {  // PF + 1 row buffers in rotation: the loads run PF rows ahead of the arithmetic
#pragma unroll
            for (int r = 0; r <= PF; ++r) {
              if (i + PF < nrows) load_row(vbuf[(r + PF) % (PF + 1)]);
              step(i, vbuf[r]);
              if (++i >= nrows) return;
            }
          }
        }();
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(bFull + 8 * stage);
    }
#ifdef KWS_WS_PROFILE
    if (blockIdx.x == 3 && (tid % kWsGT) == 0) printf("producer group %d: total %lld wait_empty %lld items %u\n", g, clock64() - wsp_start, wsp_pw, k);
#endif
  } else if (warp < kWsG * 3 + kWsEpi) {
    // ---- warps 12..15: drain only ----
#ifdef KWS_WS_PROFILE
    const long long wsp_start = clock64();
#endif
    for (int j = 0; j < my_tiles; ++j) epilogue(j);
#ifdef KWS_WS_PROFILE
    if (blockIdx.x == 3 && lane == 0) printf("drain warp %d: total %lld wait_accfull %lld\n", warp, clock64() - wsp_start, wsp_aw);
#endif
  } else if (warp > kWsG * 3 + kWsEpi + 1) {
    // ---- warps 18, 19 (they complete the issue warp's warpgroup): the output's border pixels, ReLU(bias) each (the
    // padding = 1 quirk), a tile's border while the issue warp is on that tile (written all at once at the start of the
    // kernel they cost the pipeline as much as the separate border_fill launches they replace) ----
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    if (!POOL && p.fill_border) {
      const int cq = C >> 2;
      for (int j = warp - (kWsG * 3 + kWsEpi + 2); j < my_tiles; j += 2) {
        while (s_tile < j) __nanosleep(500);
        const unsigned tile = blockIdx.x + (unsigned)j * gridDim.x, b = tile / (unsigned)p.nbands;
        const int band = (int)(tile - b * (unsigned)p.nbands), r0 = band * R, nr = min(R, H - r0);
        float* obase = p.out + (long long)b * (H + 2) * (W + 2) * (long long)p.ldc;
        // rows r0 + 1 .. r0 + nr: columns 0 and W + 1; the whole top row with the first band, the bottom row with the last
        const int nside = 2 * nr, ntop = band == 0 ? W + 2 : 0, nbot = r0 + nr == H ? W + 2 : 0;
#pragma unroll 1
        for (int e = 0; e < nside + ntop + nbot; ++e) {
          int y, x;
          if (e < nside) {
            y = r0 + 1 + (e >> 1);
            x = (e & 1) ? W + 1 : 0;
          } else if (e < nside + ntop) {
            y = 0;
            x = e - nside;
          } else {
            y = H + 1;
            x = e - nside - ntop;
          }
          float4* dst = reinterpret_cast<float4*>(obase + ((long long)y * (W + 2) + x) * p.ldc);
#pragma unroll 1
          for (int q = lane; q < cq; q += 32) dst[q] = __ldg(reinterpret_cast<const float4*>(p.border) + q);
        }
      }
    }
  } else if (warp == kWsG * 3 + kWsEpi + 1) {
    // ---- loader warp: weight chunks through the TMA ring, as far ahead as the ring allows (~45 instructions and the
    // wait for the ring stage off every item of the issue loop: 1.1 K -> 0.9 K cycles per item)
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    int kc = 0, bs = 0;
    uint32_t bph = 1;  // parity of the phase before the stage's current use
    for (int n = 0; n < total; ++n) {
      if (n >= kWsBS) mbar_wait_idle(bBempty + 8 * bs, bph, 64);  // the MMAs that read this stage last have retired
      if (lane == 0) {
        mbar_expect_tx(bBfull + 8 * bs, (uint32_t)bbytes);
        tma_load_1d(smem_u32(sB + (size_t)bs * bbytes), reinterpret_cast<const unsigned char*>(p.wimg) + (size_t)kc * bbytes,
                    (uint32_t)bbytes, bBfull + 8 * bs);
      }
      __syncwarp();
      if (++bs == kWsBS) {
        bs = 0;
        bph ^= 1u;
      }
      if (++kc == nch) kc = 0;
    }
  } else {
    // ---- issue warp: MMAs in item order, L2 prefetch of the next tile ----
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.nhalf >> 3) << 17) | ((128u >> 4) << 24);
#pragma unroll 1
    for (int jj = 0; jj < 2; ++jj) prefetch_tile(jj);
    int j = 0, kc = 0;
    const int pf_kc = nch / 2;
    WSP_DECL(free); WSP_DECL(full); WSP_DECL(bfull);
#ifdef KWS_WS_PROFILE
    const long long wsp_start = clock64();
#endif
    // descriptors of ring stage 0; a stage / K step / N half only moves the 14-bit start-address field (>> 4)
    const uint64_t adesc0 = umma_desc(smem_u32(sA), kWsLbo, kWsSbo), bdesc0 = umma_desc(smem_u32(sB), 128u, kGSbo);
    const uint32_t a_stage = kWsABytes >> 4, b_stage = (uint32_t)bbytes >> 4, b_half = (uint32_t)((p.nhalf / 8) * kGSbo) >> 4;
    const bool two_halves = p.nh == 2;
    uint32_t tmem_h0 = tmem, tmem_h1 = tmem + (uint32_t)p.nhalf;  // this tile's accumulator slots
    int bs = 0;        // weight ring stage of item n
    uint32_t bph = 0;  // and its phase
    bool wait_s1 = false;
    uint32_t bar_s1 = 0, par_s1 = 0;
    for (int n = 0; n < total; ++n) {
      const int g = n % kWsS;  // A ring stage of the item
      if (kc == 0) {  // tile boundary: this tile's accumulator slots must have been read out by their four warps
        if (lane == 0) s_tile = j;
        const int h = p.nh * j, s0 = h % ns, s1 = (h + 1) % ns, u0 = h / ns, u1 = (h + 1) / ns;
        WSP_T0();
        if (u0 > 0) WS_WAIT(bFreeS + 8 * (uint32_t)s0, (uint32_t)(u0 - 1) & 1u);
        WSP_ADD(free);
        // the second half's slot (the previous tile's first half, still being read out) is waited for after the first
        // half's MMAs of the tile's first chunk have been issued
        wait_s1 = two_halves && u1 > 0;
        bar_s1 = bFreeS + 8 * (uint32_t)s1;
        par_s1 = (uint32_t)(u1 - 1) & 1u;
        tmem_h0 = tmem + (uint32_t)(s0 * p.nhalf);
        tmem_h1 = tmem + (uint32_t)(s1 * p.nhalf);
      }
      // the next tile's input into L2 half a tile ahead: the producers start on it about four chunks before this warp
      // does.  Timing matters: at the tile's first chunk 245 K clips/s, third 253 K, fourth / fifth 255 K, seventh 245 K, never
      // 230 K; from the loader warp (three chunks earlier still) or two tiles ahead it is slower than at the first chunk
      if (kc == pf_kc && j > 0) prefetch_tile(j + 1);
      { WSP_T0(); WS_WAIT(bFull + 8 * g, (uint32_t)(n / kWsS) & 1u); WSP_ADD(full); }
      { WSP_T0(); WS_WAIT(bBfull + 8 * bs, bph); WSP_ADD(bfull); }
      tc_fence_after();
      const uint64_t ad = adesc0 + (uint64_t)((uint32_t)g * a_stage), bd = bdesc0 + (uint64_t)((uint32_t)bs * b_stage);
#pragma unroll
      for (int ks = 0; ks < kGKC / 8; ++ks) {
#ifdef KWS_WS_NOMMA  // timing experiment only (results are wrong): one K step of MMAs per item instead of four
        if (ks > 0) break;
#endif
        tc_mma_tf32_elect(tmem_h0, ad + (uint64_t)(ks * ((2 * kWsLbo) >> 4)), bd + (uint64_t)(ks * 16), idesc, (kc > 0 || ks > 0) ? 1u : 0u);
      }
      if (two_halves) {
        if (wait_s1) {
          WSP_T0();
          WS_WAIT(bar_s1, par_s1);
          WSP_ADD(free);
          wait_s1 = false;
        }
#pragma unroll
        for (int ks = 0; ks < kGKC / 8; ++ks) {
#ifdef KWS_WS_NOMMA
          if (ks > 0) break;
#endif
          tc_mma_tf32_elect(tmem_h1, ad + (uint64_t)(ks * ((2 * kWsLbo) >> 4)), bd + (uint64_t)(b_half + ks * 16), idesc, (kc > 0 || ks > 0) ? 1u : 0u);
        }
      }
      tc_commit_elect(bEmpty + 8 * g);
      tc_commit_elect(bBempty + 8 * bs);
      if (kc == nch - 1) tc_commit_elect(bAccFull + 8 * (uint32_t)(j & 1));
      if (++bs == kWsBS) {
        bs = 0;
        bph ^= 1u;
      }
      if (++kc == nch) {
        kc = 0;
        ++j;
      }
    }
#ifdef KWS_WS_PROFILE
    if (lane == 0) {  // printed by a drain warp after the last barrier (printf does not fit this warp's 24 registers)
      s_prof[0] = clock64() - wsp_start; s_prof[1] = wsp_free; s_prof[2] = wsp_full; s_prof[3] = wsp_bfull; s_prof[4] = 0;
    }
#endif
  }
  tc_fence_before();
  __syncthreads();
#ifdef KWS_WS_PROFILE
  if (blockIdx.x == 3 && tid == kWsG * kWsGT) printf("issuer: total %lld wait_accfree %lld wait_full %lld wait_bfull %lld wait_bempty %lld tiles %d items %d\n", s_prof[0], s_prof[1], s_prof[2], s_prof[3], s_prof[4], my_tiles, total);
#endif
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)p.tmem_cols));
}

// ---- conv1 (10x10, stride 2, padding 2, one input channel) as an im2col GEMM with resident weights -------------------
// K = 100 taps (+ 4 zero) = 13 MMA k-steps.  The whole weight image (4 chunks) is loaded into shared memory once per
// CTA; a tile's A operand (128 pixels x 104) is gathered by all 16 warps in one pass: an item is (pixel, pair of kernel
// rows 2m, 2m+1) = 20 consecutive k = exactly five 16-byte K quads, read as two runs of 10 features (the chunk-
// synchronous kernel did this tap by tap with a division per element and streamed the weights for every tile: 1.28 ms
// of the 11.5 ms forward at 2048 clips).  Production is ~60 instructions per thread and tile, so the tile loop is simply
// gather -> barrier -> 26 MMAs -> drain by all warps (lane quarter x column quarter, 16x256b loads as in the block
// kernel) -> barrier.
__global__ void __launch_bounds__(kGT, 1) dscnn_conv1_kernel(const GemmParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar[2];  // [0] weights landed, [1] MMAs of the tile retired
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int bbytes = (p.npad / 8) * kGSbo;
  unsigned char* sA = smem;                       // [4][kGABytes]
  unsigned char* sB = smem + 4 * kGABytes;        // [4][bbytes]
  float* sBias = reinterpret_cast<float*>(sB + 4 * (size_t)bbytes);
  const uint32_t barB = smem_u32(&s_bar[0]), barM = barB + 8;
  if (tid == 0) {
    mbar_init(barB, 1);
    mbar_init(barM, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)),
                 "r"((uint32_t)p.tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (int i = tid; i < p.npad; i += kGT) sBias[i] = i < p.C ? __ldg(p.bias + i) : 0.f;
  // K quads 25..31 of the last chunk are never written: zero once (quad 25 = k 100..103 is read by the 13th k-step)
  for (int i = tid; i < 4 * kGABytes / 16; i += kGT) reinterpret_cast<float4*>(sA)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid == 0) {
    mbar_expect_tx(barB, 4u * (uint32_t)bbytes);
    for (int c = 0; c < 4; ++c)
      tma_load_1d(smem_u32(sB + (size_t)c * bbytes), reinterpret_cast<const unsigned char*>(p.wimg) + (size_t)c * bbytes,
                  (uint32_t)bbytes, barB);
  }
  const uint32_t tmem = s_tmem;
  const int H = p.H, W = p.W, C = p.C, R = p.R, T = p.T, F = p.F;
  const unsigned ntiles = (unsigned)(p.B * p.nbands);
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.nhalf >> 3) << 17) | ((128u >> 4) << 24);
  const uint64_t adesc0 = umma_desc(smem_u32(sA), kGLboA, kGSboA), bdesc0 = umma_desc(smem_u32(sB), 128u, kGSbo);
  const uint32_t a_stage = kGABytes >> 4, b_stage = (uint32_t)bbytes >> 4, b_half = (uint32_t)((p.nhalf / 8) * kGSbo) >> 4;
  // operand descriptors of the 13 k-steps (chunk ks / 4, step ks % 4 inside it), built once: in the issue loop the
  // uniform datapath that assembles a descriptor costs more than the MMA it feeds
  uint64_t adv[13], bdv[13];
#pragma unroll
  for (int ks = 0; ks < 13; ++ks) {
    adv[ks] = adesc0 + (uint64_t)((uint32_t)(ks >> 2) * a_stage + (uint32_t)(ks & 3) * ((2 * kGLboA) >> 4));
    bdv[ks] = bdesc0 + (uint64_t)((uint32_t)(ks >> 2) * b_stage + (uint32_t)(ks & 3) * 16u);
  }
  const bool two_halves = p.nh == 2;
  const uint32_t tmem_h1 = tmem + (uint32_t)p.nhalf;
  const bool f_even = (F & 1) == 0;
  const bool pf_ok = (F & 3) == 0 && (reinterpret_cast<size_t>(p.in) & 15) == 0;  // bulk prefetch: 16-byte granules
  // drain share of this warp: TMEM lane quarter warp % 4, column quarter warp / 4 (in pairs of 8-column blocks)
  const int dr_pairs = p.npad >> 4, dr_cs = warp >> 2;
  const int dr_n = dr_pairs / 4 + (dr_cs < (dr_pairs & 3) ? 1 : 0);
  const int dr_c0 = 16 * (dr_cs * (dr_pairs / 4) + min(dr_cs, dr_pairs & 3)), dr_c1 = dr_c0 + 16 * dr_n;
  // gather: item = (pixel pp, kernel row pair m), 128 x 5 items = one per thread + a second one for threads 0..127.
  // The loads of tile i + 1 are issued right after tile i's MMAs and land in registers while the MMAs execute and the
  // accumulator is drained (the gather used to load and store in one go at the top of the tile: ncu had 60 % of the
  // warps' time on that scoreboard); the stores into the A operand follow after the drain.
  auto gather_load = [&](unsigned tl, int item, float (&v)[20]) {
    const unsigned b = tl / (unsigned)p.nbands;
    const int r0 = (int)(tl - b * (unsigned)p.nbands) * R;
    const float* feat = p.in + (long long)b * T * F;
    const int m = item >> 7, pp = item & 127;
    const int rl = pp / W, xq = pp - rl * W, y = r0 + rl;
    const bool valid = rl < R && y < H;
#pragma unroll
    for (int kr = 0; kr < 2; ++kr) {
      const int fr = 2 * y - 2 + 2 * m + kr, fc0 = 2 * xq - 2;
      const bool rok = valid && fr >= 0 && fr < T;
      const float* row = feat + (long long)fr * F + fc0;
      if (rok && f_even && fc0 >= 0 && fc0 + 10 <= F) {  // interior: five aligned 8-byte loads
#pragma unroll
        for (int e = 0; e < 5; ++e) {
          const float2 t2 = __ldg(reinterpret_cast<const float2*>(row) + e);
          v[kr * 10 + 2 * e] = t2.x;
          v[kr * 10 + 2 * e + 1] = t2.y;
        }
      } else {
#pragma unroll
        for (int e = 0; e < 10; ++e) v[kr * 10 + e] = (rok && fc0 + e >= 0 && fc0 + e < F) ? __ldg(row + e) : 0.f;
      }
    }
  };
  auto gather_store = [&](int item, const float (&v)[20]) {
    const int m = item >> 7, pp = item & 127;
    const uint32_t a0 = smem_u32(sA) + (uint32_t)((pp >> 3) * kGSboA + (pp & 7) * 16);
#pragma unroll
    for (int e = 0; e < 5; ++e) {  // K quad 5 m + e: chunk (quad >> 3), quad in chunk (quad & 7)
      const int quad = 5 * m + e;
      const uint32_t a = a0 + (uint32_t)((quad >> 3) * kGABytes + (quad & 7) * kGLboA);
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(__float_as_uint(v[4 * e]) + 0x1000u),
                   "r"(__float_as_uint(v[4 * e + 1]) + 0x1000u), "r"(__float_as_uint(v[4 * e + 2]) + 0x1000u),
                   "r"(__float_as_uint(v[4 * e + 3]) + 0x1000u)
                   : "memory");
    }
  };
  static_assert(128 * 5 == kGT + 128, "gather item map");
  float ga[20], gb[20];
  if (blockIdx.x < ntiles) {
    gather_load(blockIdx.x, tid, ga);
    if (tid < 128) gather_load(blockIdx.x, kGT + tid, gb);
  }
  uint32_t it = 0;
  for (unsigned tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    const unsigned b = tile / (unsigned)p.nbands;
    const int band = (int)(tile - b * (unsigned)p.nbands), r0 = band * R;
    if (tid == 0 && pf_ok && tile + 2 * gridDim.x < ntiles) {  // the feature rows of the tile after next -> L2
      const unsigned tn = tile + 2 * gridDim.x, bn = tn / (unsigned)p.nbands;
      const int rn = (int)(tn - bn * (unsigned)p.nbands) * R;
      const int f0 = max(2 * rn - 2, 0), f1 = min(2 * (rn + R) + 6, T - 1);
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.in + ((long long)bn * T + f0) * F), "r"((uint32_t)((f1 - f0 + 1) * F * 4))
                   : "memory");
    }
    gather_store(tid, ga);
    if (tid < 128) gather_store(kGT + tid, gb);
    fence_proxy_async();
    __syncthreads();
    if (warp == kGT / 32 - 1) {
      if (it == 0) mbar_wait(barB, 0);
      tc_fence_after();
      if (two_halves) {
#pragma unroll
        for (int ks = 0; ks < 13; ++ks) {
          tc_mma_tf32_elect(tmem, adv[ks], bdv[ks], idesc, ks > 0 ? 1u : 0u);
          tc_mma_tf32_elect(tmem_h1, adv[ks], bdv[ks] + (uint64_t)b_half, idesc, ks > 0 ? 1u : 0u);
        }
      } else {
#pragma unroll
        for (int ks = 0; ks < 13; ++ks) tc_mma_tf32_elect(tmem, adv[ks], bdv[ks], idesc, ks > 0 ? 1u : 0u);
      }
      tc_commit_elect(barM);
    }
    if (tile + gridDim.x < ntiles) {  // the next tile's features: in flight during the MMAs and the drain
      gather_load(tile + gridDim.x, tid, ga);
      if (tid < 128) gather_load(tile + gridDim.x, kGT + tid, gb);
    }
    mbar_wait(barM, it & 1u);
    tc_fence_after();
    {  // ---- drain: bias + ReLU -> NHWC [B, H, W, C] ----
      const int lq = warp & 3, tq = lane & 3, tr = lane >> 2;
      const bool odd = lane & 1;
      const int coff = odd ? 8 + 2 * (tq - 1) : 2 * tq;
      float* orow[2][2];
      bool valid[2][2];
#pragma unroll
      for (int hh = 0; hh < 2; ++hh)
#pragma unroll
        for (int ss = 0; ss < 2; ++ss) {
          const int pp = lq * 32 + hh * 16 + ss * 8 + tr, rl = pp / W, xq = pp - rl * W, y = r0 + rl;
          valid[hh][ss] = rl < R && y < H;
          orow[hh][ss] = p.out + (((long long)b * H + y) * W + xq) * (long long)p.ldc + coff + dr_c0;
        }
      const int npix = min(R, H - r0) * W;
      const bool hv0 = lq * 32 < npix, hv1 = lq * 32 + 16 < npix;
      const uint32_t ta = tmem + ((uint32_t)(lq * 32) << 16);
      for (int cb = dr_c0; cb < dr_c1 && hv0; cb += 16) {
        float v[2][8];
        tc_ld_16x256b_x2(ta + (uint32_t)cb, v[0]);
        if (hv1) tc_ld_16x256b_x2(ta + (16u << 16) + (uint32_t)cb, v[1]);
        tc_wait_ld();
        const int col = cb + 2 * tq;
        const float2 bz0 = *reinterpret_cast<const float2*>(sBias + col), bz1 = *reinterpret_cast<const float2*>(sBias + col + 8);
        const bool cok = cb + coff < C;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          if (hh == 0 || hv1) {
#pragma unroll
            for (int ss = 0; ss < 2; ++ss) {
              float* x = v[hh] + 2 * ss;  // registers [4 blk + 2 s + e]
              x[0] = fmaxf(x[0] + bz0.x, 0.f);
              x[1] = fmaxf(x[1] + bz0.y, 0.f);
              x[4] = fmaxf(x[4] + bz1.x, 0.f);
              x[5] = fmaxf(x[5] + bz1.y, 0.f);
              const float k0 = odd ? x[4] : x[0], k1 = odd ? x[5] : x[1], s0 = odd ? x[0] : x[4], s1 = odd ? x[1] : x[5];
              const float g0 = __shfl_xor_sync(0xffffffffu, s0, 1), g1 = __shfl_xor_sync(0xffffffffu, s1, 1);
              if (cok && valid[hh][ss])
                *reinterpret_cast<float4*>(orow[hh][ss]) = odd ? make_float4(g0, g1, k0, k1) : make_float4(k0, k1, g0, g1);
              orow[hh][ss] += 16;
            }
          }
        }
      }
    }
    tc_fence_before();
    __syncthreads();  // the next tile's gather overwrites A, its first MMA the accumulator
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)p.tmem_cols));
}

// border of a block output [B, Ho, Wo, C]: ReLU(pointwise bias)
__global__ void border_fill_kernel(float* __restrict__ out, const float* __restrict__ relu_bias, long long B, int Ho,
                                   int Wo, int C, int ld) {
  const int cq = C / 4, nb = 2 * Wo + 2 * (Ho - 2);
  const long long total = B * nb * (long long)cq;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i % cq);
    const int e = (int)((i / cq) % nb);
    const long long b = i / ((long long)cq * nb);
    int y, x;
    if (e < Wo) { y = 0; x = e; }
    else if (e < 2 * Wo) { y = Ho - 1; x = e - Wo; }
    else { const int r = e - 2 * Wo; y = 1 + (r >> 1); x = (r & 1) ? Wo - 1 : 0; }
    reinterpret_cast<float4*>(out + ((b * Ho + y) * Wo + x) * ld)[q] = reinterpret_cast<const float4*>(relu_bias)[q];
  }
}


// ---- host side ------------------------------------------------------------------------------
struct GemmModel {
  int C = 0, nb = 0, npad = 0, nh = 0, nhalf = 0, tmem_cols = 0;
  std::vector<float*> wimg;     // conv1, then one per block
  std::vector<int> kvalid, nchunks;
  float* d_bias = nullptr;      // [1 + nb][C]: conv1 bias, then pointwise bias with the depthwise bias folded through
  float* d_border = nullptr;    // [nb][C] ReLU(pointwise bias)
  float* d_zero = nullptr;      // [36][C] zeros: one input row of the widest layer the warp-specialised kernel takes
  float* d_dump = nullptr;      // [npad] write-only scratch (GemmParams::dump)
};

static float gemm_host_tf32(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  if ((u & 0x7f800000u) != 0x7f800000u) u = (u + 0x1000u) & ~0x1fffu;
  memcpy(&x, &u, 4);
  return x;
}

void free_gemm_model(void* gm) {
  GemmModel* g = reinterpret_cast<GemmModel*>(gm);
  if (!g) return;
  for (float* p : g->wimg)
    if (p) cudaFree(p);
  if (g->d_bias) cudaFree(g->d_bias);
  if (g->d_border) cudaFree(g->d_border);
  if (g->d_zero) cudaFree(g->d_zero);
  if (g->d_dump) cudaFree(g->d_dump);
  delete g;
}

// weight [N rows][K] (torch [c_out][k]) -> chunked K-major tf32 image
static int upload_wimg(GemmModel& g, const float* w, int N, int K) {
  const int nchunks = (K + kGKC - 1) / kGKC;
  const size_t chunk_floats = (size_t)(g.npad / 8) * kGSbo / 4;
  std::vector<float> img(chunk_floats * nchunks, 0.f);
  for (int n = 0; n < N; ++n)
    for (int k = 0; k < K; ++k) {
      const int kc = k / kGKC, kl = k % kGKC;
      img[kc * chunk_floats + ((size_t)(n >> 3) * kGSbo + (kl >> 2) * 128 + (n & 7) * 16 + (kl & 3) * 4) / 4] =
          gemm_host_tf32(w[(size_t)n * K + k]);
    }
  float* d = nullptr;
  KWS_CUDA(cudaMalloc((void**)&d, img.size() * 4));
  KWS_CUDA(cudaMemcpy(d, img.data(), img.size() * 4, cudaMemcpyHostToDevice));
  g.wimg.push_back(d);
  g.kvalid.push_back(K);
  g.nchunks.push_back(nchunks);
  return KWS_OK;
}

int build_gemm_model(ModelWeights& m, const std::vector<float>& conv1_w, const std::vector<float>& conv1_b,
                     const std::vector<const float*>& dw_b, const std::vector<const float*>& pw_w,
                     const std::vector<const float*>& pw_b) {
  free_gemm_model(m.gemm);
  m.gemm = nullptr;
  const int C = m.channels, nb = m.num_blocks;
  if (C % 4 != 0 || C > 512) return KWS_OK;  // outside the tensor-core path: layered kernels only
  GemmModel* g = new (std::nothrow) GemmModel();
  if (!g) return fail(KWS_ERR_ALLOC, "out of host memory");
  g->C = C;
  g->nb = nb;
  g->npad = (C + 31) / 32 * 32;
  g->nh = g->npad > 256 ? 2 : 1;
  g->nhalf = g->npad / g->nh;
  g->tmem_cols = 32;
  while (g->tmem_cols < g->npad) g->tmem_cols *= 2;
  int rc = upload_wimg(*g, conv1_w.data(), C, 100);
  std::vector<float> bias((size_t)(1 + nb) * C), border((size_t)(nb > 0 ? nb : 1) * C, 0.f);
  for (int c = 0; c < C; ++c) bias[c] = conv1_b[c];
  for (int l = 0; l < nb && rc == KWS_OK; ++l) {
    rc = upload_wimg(*g, pw_w[l], C, C);
    for (int co = 0; co < C; ++co) {
      double be = pw_b[l][co];
      for (int ci = 0; ci < C; ++ci) be += (double)pw_w[l][(size_t)co * C + ci] * (double)dw_b[l][ci];
      bias[(size_t)(l + 1) * C + co] = (float)be;
      border[(size_t)l * C + co] = pw_b[l][co] > 0.f ? pw_b[l][co] : 0.f;
    }
  }
  if (rc == KWS_OK && cudaMalloc((void**)&g->d_bias, bias.size() * 4) != cudaSuccess) rc = fail(KWS_ERR_CUDA, "cudaMalloc failed");
  if (rc == KWS_OK && cudaMalloc((void**)&g->d_border, border.size() * 4) != cudaSuccess) rc = fail(KWS_ERR_CUDA, "cudaMalloc failed");
  if (rc == KWS_OK && (cudaMalloc((void**)&g->d_zero, (size_t)36 * g->npad * 4) != cudaSuccess || cudaMemset(g->d_zero, 0, (size_t)36 * g->npad * 4) != cudaSuccess))
    rc = fail(KWS_ERR_CUDA, "cudaMalloc failed");
  if (rc == KWS_OK && cudaMalloc((void**)&g->d_dump, (size_t)(g->npad + 32) * 4) != cudaSuccess) rc = fail(KWS_ERR_CUDA, "cudaMalloc failed");
  if (rc == KWS_OK) {
    cudaMemcpy(g->d_bias, bias.data(), bias.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(g->d_border, border.data(), border.size() * 4, cudaMemcpyHostToDevice);
  }
  if (rc != KWS_OK) {
    free_gemm_model(g);
    return rc;
  }
  m.gemm = g;
  return KWS_OK;
}

// 1 = handled, 0 = not applicable, < 0 = error.  buf0/buf1: ping-pong NHWC workspaces of the caller.
int launch_dscnn_gemm(kws_handle_t h, const float* feats_dev, int64_t B, int T, int F, float* buf0, float* buf1,
                      float** last_out, int* last_rows, cudaStream_t st) {
  ModelWeights& m = h->model;
  GemmModel* g = reinterpret_cast<GemmModel*>(m.gemm);
  if (!g) return 0;
  const int C = g->C, nb = g->nb;
  const int H1 = (T + 4 - 10) / 2 + 1, W1 = (F + 4 - 10) / 2 + 1;
  if (W1 + 2 * nb > 60) return 0;  // one thread column per pixel column (60 producer columns per CTA)
  // tile = R whole rows of one clip: at most 128 pixels, and at most 3 rows per row group of the stencil's thread map
  auto rows_per_tile = [](int Hh, int Ww) {
    const int ng = 60 / Ww > 0 ? 60 / Ww : 1;
    int r = 128 / Ww > 0 ? 128 / Ww : 1;
    if (r > 3 * ng) r = 3 * ng;
    return r > Hh ? Hh : r;
  };
  int patch_bytes = 0, slack = 0;  // largest (R + 2) x (W + 2) x 32-channel input patch over the layers
  for (int l = 0; l < nb; ++l) {
    const int Ww = W1 + 2 * l, Hh = H1 + 2 * l;
    const int pb = (rows_per_tile(Hh, Ww) + 2) * (Ww + 2) * 128;
    if (pb > patch_bytes) patch_bytes = pb;
    if (2 * (Ww + 2) * 128 > slack) slack = 2 * (Ww + 2) * 128;  // short row groups read (and drop) two rows past their band
  }
  const size_t smem_bytes = 2 * kGABytes + 2 * (size_t)(g->npad / 8) * kGSbo + (size_t)9 * ((C + kGKC - 1) / kGKC * kGKC) * 4 +
                            (size_t)g->npad * 4 + (size_t)3 * patch_bytes + slack;
  if (smem_bytes > 227 * 1024) return 0;
  if (int rc = ensure_dyn_smem(dscnn_gemm_kernel<0>, smem_bytes)) return rc;
  if (int rc = ensure_dyn_smem(dscnn_gemm_kernel<1>, smem_bytes)) return rc;
  const int ctas_per_sm = (g->tmem_cols <= 256 && smem_bytes <= 100 * 1024) ? 2 : 1;
  GemmParams p{};
  p.C = C;
  p.T = T;
  p.F = F;
  p.npad = g->npad;
  p.nh = g->nh;
  p.nhalf = g->nhalf;
  p.tmem_cols = g->tmem_cols;
  p.patch_bytes = patch_bytes;
  p.zero_row = g->d_zero;
  p.dump = g->d_dump;
  p.B = B;
  auto set_grid = [&](int Hh, int Ww) {  // tile = R whole rows; returns the grid size
    p.H = Hh;
    p.W = Ww;
    p.R = rows_per_tile(Hh, Ww);
    p.nbands = (Hh + p.R - 1) / p.R;
    const long long t = B * p.nbands, gmax = (long long)h->sm_count * ctas_per_sm;
    return (unsigned)(t < gmax ? t : gmax);
  };
  // conv1
  p.in = feats_dev;
  p.out = buf0;
  p.wimg = g->wimg[0];
  p.dw_w = nullptr;
  p.bias = g->d_bias;
  p.kvalid = g->kvalid[0];
  p.nchunks = g->nchunks[0];
  // conv1: resident-weight im2col kernel when its shared memory (4 A chunks + the whole weight image) fits
  const size_t smem_c1 = (size_t)4 * kGABytes + (size_t)4 * (g->npad / 8) * kGSbo + (size_t)g->npad * 4;
  static const bool c1_env = [] { const char* e = std::getenv("KWS_GEMM_CONV1"); return !(e && e[0] == '0'); }();
  const bool c1_fast = c1_env && smem_c1 <= 227 * 1024 && W1 <= 128 && g->nchunks[0] == 4 && B * ((H1 + 127) / 1) < 0x7fffffffLL;
  // block layers: the warp-specialised kernel when its thread map (6 column groups of up to 6 columns) and its
  // shared-memory rings fit, else the chunk-synchronous one (KWS_GEMM_WS=0 forces the latter, for A/B runs)
  static const bool ws_env = [] { const char* e = std::getenv("KWS_GEMM_WS"); return !(e && e[0] == '0'); }();
  const int nch_blk = (C + kGKC - 1) / kGKC;
  const size_t smem_ws = (size_t)kWsS * kWsABytes + (size_t)kWsBS * (g->npad / 8) * kGSbo + (size_t)9 * nch_blk * kGKC * 4 +
                         (size_t)g->npad * 4;
  const bool ws_fits = ws_env && smem_ws <= 227 * 1024;
  // Pixel pitch of the activations.  With 276 channels a pixel is 1104 bytes: every 128-byte K-chunk segment a producer
  // warp loads straddles two cache lines and the drain's 64-byte runs start at any multiple of 16 bytes -- ncu showed the
  // L1 -> crossbar request path 60 % busy.  When every layer runs the two fast kernels (only they know the pitch) and the
  // padding is at most an eighth of the traffic, pixels are npad floats apart (a multiple of 128 bytes): one request per
  // segment, and the padding channels are never written or read.  KWS_GEMM_PAD=0: A/B.
  static const bool pad_env = [] { const char* e = std::getenv("KWS_GEMM_PAD"); return !(e && e[0] == '0'); }();
  bool all_fast = c1_fast && ws_fits && nb >= 1;
  for (int l = 0; l < nb && all_fast; ++l) {
    const int Hh = H1 + 2 * l, Ww = W1 + 2 * l, Rr = 128 / Ww < Hh ? 128 / Ww : Hh;
    if (Ww > 36 || B * ((Hh + Rr - 1) / Rr) * (long long)nch_blk >= 0x7fffffffLL) all_fast = false;
  }
  p.ldc = (all_fast && pad_env && g->npad != C && (g->npad - C) * 8 <= C) ? g->npad : C;
  {
    if (c1_fast) {
      if (int rc = ensure_dyn_smem(dscnn_conv1_kernel, smem_c1)) return rc;
      p.H = H1;
      p.W = W1;
      p.R = 128 / W1 < H1 ? 128 / W1 : H1;
      p.nbands = (H1 + p.R - 1) / p.R;
      const long long t = B * p.nbands;
      dscnn_conv1_kernel<<<(unsigned)(t < h->sm_count ? t : h->sm_count), kGT, smem_c1, st>>>(p);
    } else {
      dscnn_gemm_kernel<0><<<set_grid(H1, W1), kGT, smem_bytes, st>>>(p);
    }
  }
  KWS_CHECK_LAUNCH(h);
  float *cur = buf0, *nxt = buf1;
  int H = H1, W = W1;
  if (ws_fits) {
    if (int rc = ensure_dyn_smem(dscnn_block_ws_kernel<4, 2, false>, smem_ws)) return rc;
    if (int rc = ensure_dyn_smem(dscnn_block_ws_kernel<5, 2, false>, smem_ws)) return rc;
    if (int rc = ensure_dyn_smem(dscnn_block_ws_kernel<6, 1, false>, smem_ws)) return rc;
    if (int rc = ensure_dyn_smem(dscnn_block_ws_kernel<4, 2, true>, smem_ws)) return rc;
    if (int rc = ensure_dyn_smem(dscnn_block_ws_kernel<5, 2, true>, smem_ws)) return rc;
    if (int rc = ensure_dyn_smem(dscnn_block_ws_kernel<6, 1, true>, smem_ws)) return rc;
  }
  for (int l = 0; l < nb; ++l) {
    p.in = cur;
    p.out = nxt;
    p.wimg = g->wimg[l + 1];
    p.dw_w = m.d_dw_w + (size_t)l * 9 * C;
    p.bias = g->d_bias + (size_t)(l + 1) * C;
    p.kvalid = g->kvalid[l + 1];
    p.nchunks = g->nchunks[l + 1];
    unsigned grid = set_grid(H, W);
    bool ws = ws_fits && W <= 36;
    p.tmem_cols = g->tmem_cols;
    if (ws) {  // tile = as many whole rows as fit 128 pixels; one persistent CTA per SM
      p.nslots = g->nh == 2 && 3 * g->nhalf <= 512 ? 3 : 2;
      p.tmem_cols = 32;
      while (p.tmem_cols < p.nslots * g->nhalf) p.tmem_cols *= 2;
      p.R = 128 / W < H ? 128 / W : H;
      p.nbands = (H + p.R - 1) / p.R;
      const long long t = B * p.nbands;
      if (t * p.nchunks >= 0x7fffffffLL) {
        ws = false;
        grid = set_grid(H, W);
      } else {
        grid = (unsigned)(t < h->sm_count ? t : h->sm_count);
      }
    }
    const int pool_rows = ws ? 4 * p.nbands + 1 : p.nbands + 1;
    auto launch_block = [&]() {
      if (!ws) dscnn_gemm_kernel<1><<<grid, kGT, smem_bytes, st>>>(p);
      else if (p.pool) {
        if (W <= 24) dscnn_block_ws_kernel<4, 2, true><<<grid, kWsT, smem_ws, st>>>(p);
        else if (W <= 30) dscnn_block_ws_kernel<5, 2, true><<<grid, kWsT, smem_ws, st>>>(p);
        else dscnn_block_ws_kernel<6, 1, true><<<grid, kWsT, smem_ws, st>>>(p);
      } else {
        if (W <= 24) dscnn_block_ws_kernel<4, 2, false><<<grid, kWsT, smem_ws, st>>>(p);
        else if (W <= 30) dscnn_block_ws_kernel<5, 2, false><<<grid, kWsT, smem_ws, st>>>(p);
        else dscnn_block_ws_kernel<6, 1, false><<<grid, kWsT, smem_ws, st>>>(p);
      }
    };
    static const bool bf_env = [] { const char* e = std::getenv("KWS_GEMM_BORDER"); return !(e && e[0] == '0'); }();
    p.border = g->d_border + (size_t)l * C;
    p.fill_border = ws && bf_env ? 1 : 0;
    if (l == nb - 1) {  // the pooled rows replace the last activation (see GemmParams::pool)
      p.pool = nxt;
      p.pool_scale = (float)pool_rows / (float)((H + 2) * (W + 2));
      launch_block();
      KWS_CHECK_LAUNCH(h);
      *last_out = nxt;
      *last_rows = pool_rows;
      return 1;
    }
    launch_block();
    KWS_CHECK_LAUNCH(h);
    if (!p.fill_border) {
      const long long nborder = B * (long long)(2 * (W + 2) + 2 * H) * (C / 4);
      long long gb = (nborder + 255) / 256;
      if (gb > h->sm_count * 8LL) gb = h->sm_count * 8LL;
      border_fill_kernel<<<(unsigned)gb, 256, 0, st>>>(nxt, g->d_border + (size_t)l * C, B, H + 2, W + 2, C, p.ldc);
      KWS_CHECK_LAUNCH(h);
    }
    float* t = cur;
    cur = nxt;
    nxt = t;
    H += 2;
    W += 2;
  }
  *last_out = cur;  // no blocks: conv1 output [B, H1, W1, C]
  *last_rows = H * W;
  return 1;
}

}  // namespace kws
