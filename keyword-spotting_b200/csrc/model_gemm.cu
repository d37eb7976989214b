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
    for (int i = tid; i < 9 * kpad; i += kGT) {
      const int t = i / kpad, k = i - t * kpad;
      sW[i] = k < p.C ? __ldg(p.dw_w + t * p.C + k) : 0.f;
    }
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
  unsigned char* sP = reinterpret_cast<unsigned char*>(sW + 9 * kpad);
  const int patch_bytes = p.patch_bytes;
  const long long my_tiles = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long total_chunks = my_tiles * p.nchunks;
  // this thread's (up to 4) 16-byte pieces of a patch are the same for every chunk: patch row, global
  // offset relative to the first patch row, channel quad
  constexpr int kPieces = 4;
  int pf_row[kPieces], pf_off[kPieces];
  const int np = (R + 2) * W * 8;
#pragma unroll
  for (int i = 0; i < kPieces; ++i) {
    const int e = tid + i * kProd, pxl = e >> 3, ii = pxl / W, xx = pxl - ii * W;
    pf_row[i] = (producer && e < np) ? ii : -1;
    pf_off[i] = (ii * W + xx) * C + 4 * (e & 7);
  }
  auto prefetch_patch = [&](long long fc) {  // flat chunk index of this CTA
    if (MODE == 1 && producer && fc < total_chunks) {
      const long long lt = fc / p.nchunks;
      const int kcp = (int)(fc - lt * p.nchunks);
      const long long tl = blockIdx.x + lt * gridDim.x, bp = tl / p.nbands;
      const int rp = (int)(tl - bp * p.nbands) * R;
      const uint32_t dst0 = smem_u32(sP + (size_t)(fc % 3) * patch_bytes) + (uint32_t)tid * 16u;
      const float* base = p.in + ((bp * H + rp - 1) * W) * (long long)C + kcp * kGKC;  // row rp - 1 (may lie outside)
      const int kq = kcp * kGKC + 4 * (tid & 7);
#pragma unroll
      for (int i = 0; i < kPieces; ++i) {
        if (pf_row[i] >= 0) {
          const int yy = rp - 1 + pf_row[i];
          const bool ok = yy >= 0 && yy < H && kq < C;
          const float* src = ok ? base + pf_off[i] : p.in;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst0 + (uint32_t)(i * kProd) * 16u), "l"(src),
                       "r"(ok ? 16 : 0)
                       : "memory");
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");  // always commit: keeps the group count in step
  };
  prefetch_patch(0);
  prefetch_patch(1);
  asm volatile("cp.async.wait_group 1;" ::: "memory");
  __syncthreads();
  long long fchunk = 0;

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long b = tile / p.nbands;
    const int r0 = (int)(tile - b * p.nbands) * R;
    for (int kc = 0; kc < p.nchunks; ++kc, ++cnt, ++fchunk) {
      const int buf = cnt & 1;
      prefetch_patch(fchunk + 2);
      if (cnt >= 2) mbar_wait(bar0 + 8u * buf, ((cnt - 2) >> 1) & 1);  // the MMAs that read this buffer retired
      if (tid == kIssuer && cnt == 0) {  // first weight chunk (later ones are requested one chunk ahead, see below)
        mbar_expect_tx(barB, (uint32_t)bbytes);
        tma_load_1d(smem_u32(sB), p.wimg, (uint32_t)bbytes, barB);
      }
      unsigned char* abuf = sA + buf * kGABytes;
      if (MODE == 1) {
        // depthwise 3x3 (zero padding 1) of channels k..k+3 down this thread's column: every input row of the
        // row group (<= 3 output rows + 2 halo rows) is fetched up front -- 3 coalesced float4 per row, all in
        // flight together -- and accumulated into the output rows it feeds
        const int k = kc * kGKC + 4 * q;
        if (producer && rg < NG && k < p.kvalid) {
          const int rl0 = rg * RG, rl1 = min(rl0 + RG, R);
          const float4* patch = reinterpret_cast<const float4*>(sP + (size_t)(fchunk % 3) * patch_bytes);
          float4 vin[5][3];
#pragma unroll
          for (int ii = 0; ii < 5; ++ii) {
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            vin[ii][0] = vin[ii][1] = vin[ii][2] = z;
            if (ii <= rl1 - rl0 + 1) {  // patch row rl0 + ii is input row r0 + rl0 - 1 + ii (zero-filled outside the image)
              const float4* prow = patch + ((rl0 + ii) * W + x) * 8 + q;
              vin[ii][1] = prow[0];
              if (x > 0) vin[ii][0] = prow[-8];
              if (x + 1 < W) vin[ii][2] = prow[8];
            }
          }
          unsigned char* acol = abuf + q * kGLboA;
          const float* wq = sW + kc * kGKC + 4 * q;
          V wlo[9], whi[9];  // taps of channels (k, k+1) and (k+2, k+3): packed f32x2 FMAs over channel pairs
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            const float4 w = *reinterpret_cast<const float4*>(wq + t * kpad);
            wlo[t] = make_float2(w.x, w.y);
            whi[t] = make_float2(w.z, w.w);
          }
#pragma unroll
          for (int ro = 0; ro < 3; ++ro) {  // output row rl0 + ro from input rows ro, ro+1, ro+2 of the group
            if (rl0 + ro < rl1) {
              V olo = fsplat(0.f), ohi = fsplat(0.f);
#pragma unroll
              for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                  const float4 a = vin[ro + i][j];
                  olo = ffma(make_float2(a.x, a.y), wlo[i * 3 + j], olo);
                  ohi = ffma(make_float2(a.z, a.w), whi[i * 3 + j], ohi);
                }
              const float4 o = make_float4(olo.x, olo.y, ohi.x, ohi.y);
              const int pp = (rl0 + ro) * W + x;
              uint4 r;  // tf32 round-to-nearest: add half an ulp of the 10-bit mantissa, the MMA drops the low 13 bits
              const bool live = r0 + rl0 + ro < H;
              r.x = live ? __float_as_uint(o.x) + 0x1000u : 0u;
              r.y = live ? __float_as_uint(o.y) + 0x1000u : 0u;
              r.z = live ? __float_as_uint(o.z) + 0x1000u : 0u;
              r.w = live ? __float_as_uint(o.w) + 0x1000u : 0u;
              *reinterpret_cast<uint4*>(acol + (pp >> 3) * kGSboA + (pp & 7) * 16) = r;
            }
          }
        } else if (producer && rg < NG) {  // K padding beyond the channel count
          const int rl0 = rg * RG, rl1 = min(rl0 + RG, R);
          for (int ro = rl0; ro < rl1; ++ro) {
            const int pp = ro * W + x;
            *reinterpret_cast<uint4*>(abuf + q * kGLboA + (pp >> 3) * kGSboA + (pp & 7) * 16) = make_uint4(0, 0, 0, 0);
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
    if (warp < 8) {  // epilogue: lane = pixel row of the tile, warp / 4 = column half
      const int row = (warp & 3) * 32 + lane, chalf = warp >> 2;
      const int rl = row / W, xq = row - rl * W, y = r0 + rl;
      float* orow = nullptr;
      if (rl < R && y < H)
        orow = MODE == 1 ? p.out + ((b * (H + 2) + y + 1) * (W + 2) + xq + 1) * (long long)C
                         : p.out + ((b * H + y) * W + xq) * (long long)C;
      const int c0 = chalf * (p.npad / 2), c1 = c0 + p.npad / 2;
      const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16);
      for (int c = c0; c < c1; c += 16) {
        float v[16];
        tc_ld16(taddr + (uint32_t)c, v);
        tc_wait_ld();
        if (orow) {
#pragma unroll
          for (int e = 0; e < 16; e += 4) {
            if (c + e < C) {
              const float4 bz = __ldg(reinterpret_cast<const float4*>(p.bias + c + e));
              *reinterpret_cast<float4*>(orow + c + e) =
                  make_float4(fmaxf(v[e] + bz.x, 0.f), fmaxf(v[e + 1] + bz.y, 0.f), fmaxf(v[e + 2] + bz.z, 0.f),
                              fmaxf(v[e + 3] + bz.w, 0.f));
            }
          }
        }
      }
    }
    tc_fence_before();
    __syncthreads();  // the next tile's first MMA overwrites the accumulator
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)p.tmem_cols));
}

// border of a block output [B, Ho, Wo, C]: ReLU(pointwise bias)
__global__ void border_fill_kernel(float* __restrict__ out, const float* __restrict__ relu_bias, long long B, int Ho,
                                   int Wo, int C) {
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
    reinterpret_cast<float4*>(out + ((b * Ho + y) * Wo + x) * C)[q] = reinterpret_cast<const float4*>(relu_bias)[q];
  }
}


// ---- host side ------------------------------------------------------------------------------
struct GemmModel {
  int C = 0, nb = 0, npad = 0, nh = 0, nhalf = 0, tmem_cols = 0;
  std::vector<float*> wimg;     // conv1, then one per block
  std::vector<int> kvalid, nchunks;
  float* d_bias = nullptr;      // [1 + nb][C]: conv1 bias, then pointwise bias with the depthwise bias folded through
  float* d_border = nullptr;    // [nb][C] ReLU(pointwise bias)
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
                      float** last_out, cudaStream_t st) {
  ModelWeights& m = h->model;
  GemmModel* g = reinterpret_cast<GemmModel*>(m.gemm);
  if (!g) return 0;
  const int C = g->C, nb = g->nb;
  const int H1 = (T + 4 - 10) / 2 + 1, W1 = (F + 4 - 10) / 2 + 1;
  if (W1 + 2 * nb > 60) return 0;  // one thread column per pixel column (60 producer columns per CTA)
  int patch_bytes = 0;  // largest (R + 2) x W x 32-channel input patch over the layers
  for (int l = 0; l < nb; ++l) {
    const int Ww = W1 + 2 * l, Hh = H1 + 2 * l;
    int Rr = 128 / Ww > 0 ? 128 / Ww : 1;
    if (Rr > Hh) Rr = Hh;
    const int pb = (Rr + 2) * Ww * 128;
    if (pb > patch_bytes) patch_bytes = pb;
  }
  const size_t smem_bytes = 2 * kGABytes + 2 * (size_t)(g->npad / 8) * kGSbo + (size_t)9 * ((C + kGKC - 1) / kGKC * kGKC) * 4 +
                            (size_t)3 * patch_bytes;
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
  p.B = B;
  auto set_grid = [&](int Hh, int Ww) {  // tile = R whole rows; returns the grid size
    p.H = Hh;
    p.W = Ww;
    p.R = 128 / Ww > 0 ? 128 / Ww : 1;
    if (p.R > Hh) p.R = Hh;
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
  dscnn_gemm_kernel<0><<<set_grid(H1, W1), kGT, smem_bytes, st>>>(p);
  KWS_CHECK_LAUNCH(h);
  float *cur = buf0, *nxt = buf1;
  int H = H1, W = W1;
  for (int l = 0; l < nb; ++l) {
    p.in = cur;
    p.out = nxt;
    p.wimg = g->wimg[l + 1];
    p.dw_w = m.d_dw_w + (size_t)l * 9 * C;
    p.bias = g->d_bias + (size_t)(l + 1) * C;
    p.kvalid = g->kvalid[l + 1];
    p.nchunks = g->nchunks[l + 1];
    dscnn_gemm_kernel<1><<<set_grid(H, W), kGT, smem_bytes, st>>>(p);
    KWS_CHECK_LAUNCH(h);
    const long long nborder = B * (long long)(2 * (W + 2) + 2 * H) * (C / 4);
    long long gb = (nborder + 255) / 256;
    if (gb > h->sm_count * 8LL) gb = h->sm_count * 8LL;
    border_fill_kernel<<<(unsigned)gb, 256, 0, st>>>(nxt, g->d_border + (size_t)l * C, B, H + 2, W + 2, C);
    KWS_CHECK_LAUNCH(h);
    float* t = cur;
    cur = nxt;
    nxt = t;
    H += 2;
    W += 2;
  }
  *last_out = cur;
  return 1;
}

}  // namespace kws
