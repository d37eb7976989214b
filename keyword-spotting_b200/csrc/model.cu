// Forward pass of the reference KWS network, DepthwiseSeparableConv
// (kws/libs/models.py:122-183): conv1 10x10 s2 p2 + ReLU -> num_blocks x [depthwise 3x3 p1
// (+bias, no activation) -> pointwise 1x1 with padding=1 (+bias) -> ReLU] -> global average
// pool -> fc.  Each block grows H and W by 2 and its border pixels equal ReLU(pointwise bias)
// (models.py:104-106; SURVEY.md a-M2).  Logits are returned, no softmax (models.py:181).
//
// Layer-by-layer fp32 path: activations NHWC [B, H, W, C] in a ping-pong workspace.
// Weights are repacked at kws_model_finalize() so the channel index is innermost.
#include <cstdlib>
#include <cuda_runtime.h>

#include <cstring>
#include <string>

#include "mfcc_dev.cuh"

namespace kws {

// conv1: in [B, T, F] (single input channel) -> out [B, H1, W1, C], ReLU.
// thread = (pixel, 4 consecutive output channels)
__global__ void conv1_relu_kernel(const float* __restrict__ in, float* __restrict__ out,
                                  const float* __restrict__ w /*[100][C]*/,
                                  const float* __restrict__ bias, long long B, int T, int F, int H1,
                                  int W1, int C) {
  const int cq = C / 4;
  const long long total = B * H1 * W1 * (long long)cq;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i % cq);
    const long long pix = i / cq;
    const int wo = (int)(pix % W1);
    const int ho = (int)((pix / W1) % H1);
    const long long b = pix / ((long long)W1 * H1);
    float4 acc = __ldg(reinterpret_cast<const float4*>(bias) + q);
    const float* img = in + b * (long long)T * F;
#pragma unroll 2
    for (int ki = 0; ki < 10; ++ki) {
      const int hi = 2 * ho - 2 + ki;
      if (hi < 0 || hi >= T) continue;
#pragma unroll
      for (int kj = 0; kj < 10; ++kj) {
        const int wi = 2 * wo - 2 + kj;
        if (wi < 0 || wi >= F) continue;
        const float x = __ldg(img + hi * F + wi);
        const float4 wv = __ldg(reinterpret_cast<const float4*>(w + (ki * 10 + kj) * C) + q);
        acc.x = fmaf(x, wv.x, acc.x);
        acc.y = fmaf(x, wv.y, acc.y);
        acc.z = fmaf(x, wv.z, acc.z);
        acc.w = fmaf(x, wv.w, acc.w);
      }
    }
    acc.x = fmaxf(acc.x, 0.f); acc.y = fmaxf(acc.y, 0.f);
    acc.z = fmaxf(acc.z, 0.f); acc.w = fmaxf(acc.w, 0.f);
    reinterpret_cast<float4*>(out)[i] = acc;
  }
}

// One DS block: in [B,H,W,C] -> out [B,H+2,W+2,C].
// CTA = PIX output pixels x C/4 channel quads; the depthwise result of each pixel is shared
// through shared memory so it is computed once and consumed by all output channels.
constexpr int kBlkPix = 8;
__global__ void dsblock_kernel(const float* __restrict__ in, float* __restrict__ out,
                               const float* __restrict__ dw_w /*[9][C]*/,
                               const float* __restrict__ dw_b, const float* __restrict__ pw_w /*[C][C]*/,
                               const float* __restrict__ pw_b, long long B, int H, int W, int C,
                               float* __restrict__ dsave /* [B,H,W,C] depthwise output for training, or null */) {
  extern __shared__ __align__(16) float s_d[];  // [kBlkPix][C]
  const int cq = C / 4;
  const int Ho = H + 2, Wo = W + 2;
  const long long npix = B * Ho * (long long)Wo;
  const long long ngroups = (npix + kBlkPix - 1) / kBlkPix;
  for (long long grp = blockIdx.x; grp < ngroups; grp += gridDim.x) {
    __syncthreads();
    // phase 1: depthwise 3x3 (zero padding 1) + bias for interior pixels
    for (int t = threadIdx.x; t < kBlkPix * cq; t += blockDim.x) {
      const int pl = t / cq, q = t - pl * cq;
      const long long pix = grp * kBlkPix + pl;
      float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
      if (pix < npix) {
        int wo, ho;
        long long b;
        split_pixel(pix, Wo, Ho, &b, &ho, &wo);
        if (ho >= 1 && ho <= H && wo >= 1 && wo <= W) {
          const int h = ho - 1, wc = wo - 1;
          d = __ldg(reinterpret_cast<const float4*>(dw_b) + q);
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const int hi = h + i - 1;
            if (hi < 0 || hi >= H) continue;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
              const int wi = wc + j - 1;
              if (wi < 0 || wi >= W) continue;
              const float4 x = __ldg(reinterpret_cast<const float4*>(in + ((b * H + hi) * W + wi) * C) + q);
              const float4 k = __ldg(reinterpret_cast<const float4*>(dw_w + (i * 3 + j) * C) + q);
              d.x = fmaf(x.x, k.x, d.x); d.y = fmaf(x.y, k.y, d.y);
              d.z = fmaf(x.z, k.z, d.z); d.w = fmaf(x.w, k.w, d.w);
            }
          }
          if (dsave) reinterpret_cast<float4*>(dsave + ((b * H + h) * W + wc) * C)[q] = d;
        }
      }
      reinterpret_cast<float4*>(s_d + pl * C)[q] = d;
    }
    __syncthreads();
    // phase 2: pointwise 1x1 (+bias) + ReLU; border pixels see zero input -> ReLU(bias)
    for (int t = threadIdx.x; t < kBlkPix * cq; t += blockDim.x) {
      const int pl = t / cq, q = t - pl * cq;
      const long long pix = grp * kBlkPix + pl;
      if (pix >= npix) continue;
      int wo, ho;
      long long b_unused;
      split_pixel(pix, Wo, Ho, &b_unused, &ho, &wo);
      float4 acc = __ldg(reinterpret_cast<const float4*>(pw_b) + q);
      if (ho >= 1 && ho <= H && wo >= 1 && wo <= W) {
        const float* d = s_d + pl * C;
#pragma unroll 4
        for (int c = 0; c < C; ++c) {
          const float dv = d[c];
          const float4 wv = __ldg(reinterpret_cast<const float4*>(pw_w + (size_t)c * C) + q);
          acc.x = fmaf(dv, wv.x, acc.x); acc.y = fmaf(dv, wv.y, acc.y);
          acc.z = fmaf(dv, wv.z, acc.z); acc.w = fmaf(dv, wv.w, acc.w);
        }
      }
      acc.x = fmaxf(acc.x, 0.f); acc.y = fmaxf(acc.y, 0.f);
      acc.z = fmaxf(acc.z, 0.f); acc.w = fmaxf(acc.w, 0.f);
      reinterpret_cast<float4*>(out + pix * C)[q] = acc;
    }
  }
}

// adaptive_avg_pool2d(1,1) + fc: one CTA per clip; the pixels are split over blockDim / C thread groups
__global__ void pool_fc_kernel(const float* __restrict__ in, float* __restrict__ logits,
                               const float* __restrict__ fc_w /*[C][ncls]*/,
                               const float* __restrict__ fc_b, long long B, int HW, int C, int ncls,
                               float* __restrict__ pooled_out /* [B,C] for training, or null */) {
  extern __shared__ __align__(16) float s_pool[];  // [C] means, then [npg][C] partial sums
  float* s_part = s_pool + C;
  const int npg = C <= (int)blockDim.x ? (int)blockDim.x / C : 1;
  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    __syncthreads();
    const float* x = in + b * (long long)HW * C;
    for (int t = threadIdx.x; t < npg * C; t += blockDim.x) {
      const int pg = t / C, c = t - pg * C;
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
      int p = pg;
      for (; p + 3 * npg < HW; p += 4 * npg) {
        s0 += x[(size_t)p * C + c];
        s1 += x[(size_t)(p + npg) * C + c];
        s2 += x[(size_t)(p + 2 * npg) * C + c];
        s3 += x[(size_t)(p + 3 * npg) * C + c];
      }
      for (; p < HW; p += npg) s0 += x[(size_t)p * C + c];
      s_part[pg * C + c] = (s0 + s1) + (s2 + s3);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float s = 0.f;
      for (int pg = 0; pg < npg; ++pg) s += s_part[pg * C + c];
      s_pool[c] = s / (float)HW;
      if (pooled_out) pooled_out[b * C + c] = s_pool[c];
    }
    __syncthreads();
    for (int k = threadIdx.x; k < ncls; k += blockDim.x) {
      float acc = fc_b[k];
      for (int c = 0; c < C; ++c) acc = fmaf(s_pool[c], fc_w[c * ncls + k], acc);
      logits[b * ncls + k] = acc;
    }
  }
}
constexpr int kPoolThreads = 256;
static size_t pool_smem(int C) { return (size_t)((C <= kPoolThreads ? kPoolThreads / C : 1) + 1) * C * 4; }

// Depthwise 3x3 (padding 1) + bias of one 64-channel block, row-walking form (training forward): thread = (clip, column,
// row segment, channel PAIR) walks down its rows with the 3 x 3 window of the input in registers: three 8-byte loads and
// nine packed FMAs per pixel, every access a 256-byte run across the warp.  Writes the depthwise output the backward
// pass needs anyway; dsblock64_kernel<true> then reads it instead of gathering the window nine times per pixel.
constexpr int kDwfSeg = 2;
__global__ void __launch_bounds__(256) dw_fwd_rows_kernel(const float* __restrict__ x, float* __restrict__ d,
                                                          const float* __restrict__ w_packed /*[9][64]*/,
                                                          const float* __restrict__ bias, long long B, int H, int W) {
  constexpr int C = 64;
  const int tid = threadIdx.x, cp = tid & 31, slot = tid >> 5;
  const long long ncol = B * W * (long long)kDwfSeg;
  float2 wv[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) wv[t] = __ldg(reinterpret_cast<const float2*>(w_packed + t * C) + cp);
  const float2 bv = __ldg(reinterpret_cast<const float2*>(bias) + cp), z2 = make_float2(0.f, 0.f);
  const int seg_rows = (H + kDwfSeg - 1) / kDwfSeg, rs = W * (C / 2), cs = C / 2;
  for (long long col = blockIdx.x * 8LL + slot; col < ncol; col += gridDim.x * 8LL) {
    const unsigned c32 = (unsigned)col;
    const unsigned seg = c32 % (unsigned)kDwfSeg, bx = c32 / (unsigned)kDwfSeg, xx = bx % (unsigned)W, b = bx / (unsigned)W;
    const int y0 = (int)seg * seg_rows, y1 = min(H, y0 + seg_rows);
    const bool hasl = xx > 0, hasr = (int)xx + 1 < W;
    const float2* xc = reinterpret_cast<const float2*>(x + ((long long)b * H * W + xx) * C) + cp;
    float2* dc = reinterpret_cast<float2*>(d + ((long long)b * H * W + xx) * C) + cp;
    auto load3 = [&](int y, float2 (&r)[3]) {  // x[y][xx-1 .. xx+1], zeros outside the layer
      const bool ok = y >= 0 && y < H;
      const float2* q = xc + (long long)y * rs;
      r[0] = ok && hasl ? __ldg(q - cs) : z2;
      r[1] = ok ? __ldg(q) : z2;
      r[2] = ok && hasr ? __ldg(q + cs) : z2;
    };
    auto pixel = [&](int y, const float2 (&m)[3], const float2 (&c)[3], const float2 (&n)[3]) {
      float2 acc = bv;
#pragma unroll
      for (int j = 0; j < 3; ++j) {  // tap (i, j) reads x[y + i - 1][xx + j - 1]
        acc = __ffma2_rn(m[j], wv[j], acc);
        acc = __ffma2_rn(c[j], wv[3 + j], acc);
        acc = __ffma2_rn(n[j], wv[6 + j], acc);
      }
      dc[(long long)y * rs] = acc;
    };
    float2 ra[3], rb[3], rc[3];  // rows y - 1, y, y + 1 in rotation
    load3(y0 - 1, ra);
    load3(y0, rb);
    for (int y = y0; y < y1; y += 3) {
      load3(y + 1, rc);
      pixel(y, ra, rb, rc);
      if (y + 1 >= y1) break;
      load3(y + 2, ra);
      pixel(y + 1, rb, rc, ra);
      if (y + 2 >= y1) break;
      load3(y + 3, rb);
      pixel(y + 2, rc, ra, rb);
    }
  }
}

// Training forward of one block for C = 64 (the reference network): depthwise 3x3 into a transposed shared
// tile, then the pointwise 1x1 as a register-tiled fp32 GEMM (4 px x 8 co per thread, packed f32x2 FMAs,
// weights in shared memory).  Same results as dsblock_kernel up to fp32 summation order; saves the depthwise
// output for the backward pass.
constexpr int kB64Pix = 128;
constexpr size_t kB64Smem = (size_t)(64 * 64 + 64 * kB64Pix) * 4;
// DLOAD = true: the depthwise output was computed by dw_fwd_rows_kernel into `dsave`; this kernel only reads it (one
// coalesced 16-byte load per item instead of nine gathers) and runs the pointwise GEMM.
template <bool DLOAD>
__global__ void __launch_bounds__(256, 3) dsblock64_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                        const float* __restrict__ dw_w /*[9][64]*/,
                                                        const float* __restrict__ dw_b,
                                                        const float* __restrict__ pw_w /*[ci][co]*/,
                                                        const float* __restrict__ pw_b, long long B, int H, int W,
                                                        float* __restrict__ dsave) {
  constexpr int C = 64, TP = kB64Pix;
  extern __shared__ __align__(16) float s_b64[];
  float* s_w = s_b64;            // [64 ci][64 co]
  float* s_dT = s_w + C * C;     // [64 ci][TP px], px index XOR-swizzled per channel quad
  const int tid = threadIdx.x;
  const int Ho = H + 2, Wo = W + 2;
  const long long npix = B * Ho * (long long)Wo, ntiles = (npix + TP - 1) / TP;
  for (int i = tid; i < C * C / 4; i += 256) reinterpret_cast<float4*>(s_w)[i] = __ldg(reinterpret_cast<const float4*>(pw_w) + i);
  const int q = tid & 15, px0 = tid >> 4;
  const int rg = tid >> 3, cg = (tid & 7) * 4;  // GEMM ownership: pixels 4rg..4rg+3, co = cg..cg+3 and 32+cg..32+cg+3
  float4 kw[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) kw[t] = __ldg(reinterpret_cast<const float4*>(dw_w + t * C) + q);
  const float4 kb = __ldg(reinterpret_cast<const float4*>(dw_b) + q);
  const float4 b0 = __ldg(reinterpret_cast<const float4*>(pw_b + cg)), b1 = __ldg(reinterpret_cast<const float4*>(pw_b + 32 + cg));
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    __syncthreads();
#pragma unroll 2
    for (int it = 0; it < TP / 16; ++it) {
      const int pl = px0 + 16 * it;
      const long long pix = tile * TP + pl;
      float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
      if (pix < npix) {
        int wo, ho;
        long long b;
        split_pixel(pix, Wo, Ho, &b, &ho, &wo);
        if (ho >= 1 && ho <= H && wo >= 1 && wo <= W) {
          const int h = ho - 1, wc = wo - 1;
          if (DLOAD) {
            d = __ldg(reinterpret_cast<const float4*>(dsave + ((b * H + h) * W + wc) * C) + q);
          } else {
          d = kb;
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const int hi = h + i - 1;
            if (hi < 0 || hi >= H) continue;
#pragma unroll
            for (int jx = 0; jx < 3; ++jx) {
              const int wi = wc + jx - 1;
              if (wi < 0 || wi >= W) continue;
              const float4 x = __ldg(reinterpret_cast<const float4*>(in + ((b * H + hi) * W + wi) * C) + q);
              const float4 k = kw[i * 3 + jx];
              d.x = fmaf(x.x, k.x, d.x); d.y = fmaf(x.y, k.y, d.y);
              d.z = fmaf(x.z, k.z, d.z); d.w = fmaf(x.w, k.w, d.w);
            }
          }
          if (dsave) reinterpret_cast<float4*>(dsave + ((b * H + h) * W + wc) * C)[q] = d;
          }
        }
      }
      const int pls = pl ^ (4 * (q & 7));
      s_dT[(4 * q + 0) * TP + pls] = d.x;
      s_dT[(4 * q + 1) * TP + pls] = d.y;
      s_dT[(4 * q + 2) * TP + pls] = d.z;
      s_dT[(4 * q + 3) * TP + pls] = d.w;
    }
    __syncthreads();
    V acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      acc[i][0] = make_float2(b0.x, b0.y); acc[i][1] = make_float2(b0.z, b0.w);
      acc[i][2] = make_float2(b1.x, b1.y); acc[i][3] = make_float2(b1.z, b1.w);
    }
#pragma unroll 4
    for (int ci = 0; ci < C; ++ci) {
      const float4 a = *reinterpret_cast<const float4*>(s_dT + ci * TP + ((4 * rg) ^ (4 * ((ci >> 2) & 7))));
      const float4 w0 = *reinterpret_cast<const float4*>(s_w + ci * C + cg);
      const float4 w1 = *reinterpret_cast<const float4*>(s_w + ci * C + 32 + cg);
      const V wv[4] = {make_float2(w0.x, w0.y), make_float2(w0.z, w0.w), make_float2(w1.x, w1.y), make_float2(w1.z, w1.w)};
      const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int jx = 0; jx < 4; ++jx) acc[i][jx] = ffma(fsplat(av[i]), wv[jx], acc[i][jx]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {  // border pixels had a zero depthwise row: ReLU(bias), the padding=1 quirk
      const long long pix = tile * TP + 4 * rg + i;
      if (pix < npix) {
        float* dst = out + pix * C;
        *reinterpret_cast<float4*>(dst + cg) = make_float4(fmaxf(acc[i][0].x, 0.f), fmaxf(acc[i][0].y, 0.f),
                                                           fmaxf(acc[i][1].x, 0.f), fmaxf(acc[i][1].y, 0.f));
        *reinterpret_cast<float4*>(dst + 32 + cg) = make_float4(fmaxf(acc[i][2].x, 0.f), fmaxf(acc[i][2].y, 0.f),
                                                                fmaxf(acc[i][3].x, 0.f), fmaxf(acc[i][3].y, 0.f));
      }
    }
  }
}

// optional epilogue (SURVEY a-X4): row-wise softmax of the logits.  The reference never applies it on
// the torch path (logits go to CrossEntropyLoss / torch.max, training.py:293,301); not part of parity.
__global__ void softmax_kernel(const float* __restrict__ logits, float* __restrict__ out, long long B, int K) {
  for (long long b = blockIdx.x * (long long)blockDim.x + threadIdx.x; b < B; b += (long long)gridDim.x * blockDim.x) {
    const float* x = logits + b * K;
    float mx = x[0];
    for (int k = 1; k < K; ++k) mx = fmaxf(mx, x[k]);
    float se = 0.f;
    for (int k = 0; k < K; ++k) se += expf(x[k] - mx);
    const float inv = 1.0f / se;
    for (int k = 0; k < K; ++k) out[b * K + k] = expf(x[k] - mx) * inv;
  }
}

static void conv_out_hw(int T, int F, int nb, int* H1, int* W1, int* Hl, int* Wl) {
  *H1 = (T + 4 - 10) / 2 + 1;
  *W1 = (F + 4 - 10) / 2 + 1;
  *Hl = *H1 + 2 * nb;
  *Wl = *W1 + 2 * nb;
}

static int slot_of(const ModelWeights& m, const std::string& name) {
  for (size_t i = 0; i < m.names.size(); ++i)
    if (m.names[i] == name) return (int)i;
  return -1;
}

static void model_free(ModelWeights& m) {
  void* ptrs[] = {m.d_conv1_w, m.d_conv1_b, m.d_dw_w, m.d_dw_b, m.d_pw_w, m.d_pw_b, m.d_pw_beff,
                  m.d_fc_w,    m.d_fc_b,    m.d_ws,     m.d_fused_a, m.d_fused_dw, m.d_fused_actb, m.d_fused_border};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  free_gemm_model(m.gemm);
  m = ModelWeights();
}

constexpr int64_t kModelChunk = 2048;  // clips per pass through the ping-pong workspace

// conv1 (10x10, stride 2, padding 2, one input channel) + ReLU, tiled: a CTA takes RB output rows of one clip,
// keeps the [100][C] weights in shared memory and double-buffers the 2*RB+8 zero-padded input rows (the next
// tile's rows are fetched into registers while the current tile is computed).  A thread produces 16 channels of
// two horizontally adjacent output pixels: per tap 4 broadcast weight reads + 2 input reads feed 16 packed
// f32x2 FMAs.  Lanes of a warp are consecutive pixel pairs of one channel group, so the weight reads are uniform.
// Same sums in the same order as conv1_relu_kernel (bias first, taps row-major; a padded tap adds 0).  C % 16 == 0.
constexpr int kC1Threads = 128, kC1Slack = 16;
__global__ void __launch_bounds__(kC1Threads) conv1_tile_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                               const float* __restrict__ w /*[100][C]*/,
                                                               const float* __restrict__ bias, long long B, int T, int F,
                                                               int H1, int W1, int C, int RB) {
  extern __shared__ __align__(16) float s_c1[];
  constexpr int kMaxPre = 16;                       // staged floats per thread and tile (launcher guarantees the bound)
  const int FP = F + 4, NR = 2 * RB + 8, npatch = NR * FP;
  float* s_w = s_c1;                                // [100][C]
  float* s_in = s_w + 100 * C;                      // [2][NR * FP + slack]: input rows 2*r0-2 .., columns -2 .. F+1
  const int pstride = npatch + kC1Slack;
  const int tid = threadIdx.x, ng = C / 16, wg = (W1 + 1) / 2;
  for (int i = tid; i < 100 * C / 4; i += kC1Threads) reinterpret_cast<float4*>(s_w)[i] = __ldg(reinterpret_cast<const float4*>(w) + i);
  for (int i = tid; i < 2 * pstride; i += kC1Threads) s_in[i] = 0.f;
  const int nrb = (H1 + RB - 1) / RB;
  const long long ntiles = B * nrb;
  float pre[kMaxPre];
  auto fetch = [&](long long tile) {  // the tile's input rows -> registers (zero outside the image)
    const long long b = tile / nrb;
    const int r0 = (int)(tile - b * nrb) * RB;
    const float* img = in + b * (long long)T * F;
#pragma unroll
    for (int u = 0; u < kMaxPre; ++u) {
      const int i = tid + u * kC1Threads;
      float v = 0.f;
      if (i < npatch) {
        const int rr = i / FP, cc = i - rr * FP;
        const int hi = 2 * r0 - 2 + rr, wi = cc - 2;
        if (hi >= 0 && hi < T && wi >= 0 && wi < F) v = __ldg(img + hi * F + wi);
      }
      pre[u] = v;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int u = 0; u < kMaxPre; ++u) {
      const int i = tid + u * kC1Threads;
      if (i < npatch) s_in[buf * pstride + i] = pre[u];
    }
  };
  __syncthreads();
  if ((long long)blockIdx.x < ntiles) {
    fetch(blockIdx.x);
    stash(0);
  }
  __syncthreads();
  int buf = 0;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, buf ^= 1) {
    const long long b = tile / nrb;
    const int r0 = (int)(tile - b * nrb) * RB;
    const bool more = tile + gridDim.x < ntiles;
    if (more) fetch(tile + gridDim.x);
    const float* patch = s_in + buf * pstride;
    const int rows = min(RB, H1 - r0), nper = rows * wg;
    for (int item = tid; item < nper * ng; item += kC1Threads) {
      const int g = item / nper, pp = item - g * nper;
      const int rl = pp / wg, wo = 2 * (pp - rl * wg);
      V acc[2][8];
      {
        const float4* bq = reinterpret_cast<const float4*>(bias + 16 * g);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 bv = __ldg(bq + i);
          acc[0][2 * i] = acc[1][2 * i] = make_float2(bv.x, bv.y);
          acc[0][2 * i + 1] = acc[1][2 * i + 1] = make_float2(bv.z, bv.w);
        }
      }
      const float* xp = patch + (2 * rl) * FP + 2 * wo;
      const float* wp = s_w + 16 * g;
#pragma unroll 1
      for (int ki = 0; ki < 10; ++ki) {
#pragma unroll
        for (int kj = 0; kj < 10; ++kj) {
          const V xa = fsplat(xp[ki * FP + kj]), xb = fsplat(xp[ki * FP + kj + 2]);
          const float4* wq = reinterpret_cast<const float4*>(wp + (ki * 10 + kj) * C);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 wv = wq[i];
            const V w0 = make_float2(wv.x, wv.y), w1 = make_float2(wv.z, wv.w);
            acc[0][2 * i] = ffma(xa, w0, acc[0][2 * i]);
            acc[0][2 * i + 1] = ffma(xa, w1, acc[0][2 * i + 1]);
            acc[1][2 * i] = ffma(xb, w0, acc[1][2 * i]);
            acc[1][2 * i + 1] = ffma(xb, w1, acc[1][2 * i + 1]);
          }
        }
      }
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        if (wo + e < W1) {
          float4* dst = reinterpret_cast<float4*>(out + ((b * H1 + r0 + rl) * W1 + wo + e) * (long long)C + 16 * g);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dst[i] = make_float4(fmaxf(acc[e][2 * i].x, 0.f), fmaxf(acc[e][2 * i].y, 0.f), fmaxf(acc[e][2 * i + 1].x, 0.f),
                                 fmaxf(acc[e][2 * i + 1].y, 0.f));
        }
      }
    }
    if (more) stash(buf ^ 1);   // the other buffer was last read one tile ago, behind the barrier below
    __syncthreads();
  }
}

int launch_conv1_relu(kws_handle_t h, const float* in, float* out, const float* w_packed, const float* bias, int64_t B,
                      int T, int F, int H1, int W1, int C, cudaStream_t st) {
  if (C % 16 == 0) {
    // rows per tile: about one work item (pixel pair x 16 channels) per thread, tiles of a clip equally tall
    const int per_row = ((W1 + 1) / 2) * (C / 16);
    int nrb = (H1 * per_row + kC1Threads - 1) / kC1Threads;
    if (nrb < 1) nrb = 1;
    if (nrb > H1) nrb = H1;
    const int RB = (H1 + nrb - 1) / nrb;
    const size_t npatch = (size_t)(2 * RB + 8) * (F + 4);
    const size_t smem = ((size_t)100 * C + 2 * (npatch + kC1Slack)) * 4;
    if (smem <= 100 * 1024 && npatch <= (size_t)16 * kC1Threads) {
      const long long ntiles = B * ((H1 + RB - 1) / RB);
      long long g = ntiles < h->sm_count * 12LL ? ntiles : h->sm_count * 12LL;
      if (int rc_ = ensure_dyn_smem(conv1_tile_kernel, (size_t)smem)) return rc_;
      conv1_tile_kernel<<<(unsigned)g, kC1Threads, smem, st>>>(in, out, w_packed, bias, B, T, F, H1, W1, C, RB);
      KWS_CHECK_LAUNCH(h);
      return KWS_OK;
    }
  }
  const long long total = B * H1 * W1 * (long long)(C / 4);
  long long g = (total + 255) / 256;
  if (g > h->sm_count * 16LL) g = h->sm_count * 16LL;
  conv1_relu_kernel<<<(unsigned)g, 256, 0, st>>>(in, out, w_packed, bias, B, T, F, H1, W1, C);
  KWS_CHECK_LAUNCH(h);
  return KWS_OK;
}

int train_forward_layers(kws_handle_t h, const float* feats_dev, int64_t B, int T, int F, TrainWorkspace& w,
                         cudaStream_t st) {
  ModelWeights& m = h->model;
  const int C = m.channels, nb = m.num_blocks, K = m.num_classes, cq = C / 4;
  int H1, W1, Hl, Wl;
  conv_out_hw(T, F, nb, &H1, &W1, &Hl, &Wl);
  {
    const int rc = launch_conv1_relu(h, feats_dev, w.L[0], m.d_conv1_w, m.d_conv1_b, B, T, F, H1, W1, C, st);
    if (rc) return rc;
  }
  int H = H1, W = W1;
  for (int i = 0; i < nb; ++i) {
    const long long npix = B * (long long)(H + 2) * (W + 2);
    long long g = (npix + kBlkPix - 1) / kBlkPix;
    if (g > h->sm_count * 16LL) g = h->sm_count * 16LL;
    if (C == 64) {
      const long long nt = (npix + kB64Pix - 1) / kB64Pix;
      const long long g64 = nt < h->sm_count * 3LL ? nt : h->sm_count * 3LL;
      static const bool dw_rows = [] { const char* e = std::getenv("KWS_TRAIN_DWROWS"); return !(e && e[0] == '0'); }();
      const long long ncol = B * (long long)W * kDwfSeg;
      if (dw_rows && w.D[i] && ncol < 0x7fffffffLL) {  // depthwise by the row-walking kernel, then the GEMM kernel on its output
        long long g2 = (ncol + 7) / 8;
        if (g2 > h->sm_count * 16LL) g2 = h->sm_count * 16LL;
        dw_fwd_rows_kernel<<<(unsigned)g2, 256, 0, st>>>(w.L[i], w.D[i], m.d_dw_w + (size_t)i * 9 * C, m.d_dw_b + (size_t)i * C, B, H, W);
        KWS_CHECK_LAUNCH(h);
        static const bool fwd_tc = [] { const char* e = std::getenv("KWS_TRAIN_TC"); return !(e && e[0] == '0'); }();
        if (fwd_tc) {  // pointwise + bias + ReLU on the tensor core (train_tc.cu)
          const int rc_tc = launch_pw_fwd_tc(h, w.D[i], w.L[i + 1], m.d_pw_w + (size_t)i * C * C, m.d_pw_b + (size_t)i * C, B, H, W, st);
          if (rc_tc < 0) return rc_tc;
          H += 2;
          W += 2;
          continue;
        }
        if (int rc_ = ensure_dyn_smem(dsblock64_kernel<true>, (size_t)kB64Smem)) return rc_;
        dsblock64_kernel<true><<<(unsigned)g64, 256, kB64Smem, st>>>(w.L[i], w.L[i + 1], m.d_dw_w + (size_t)i * 9 * C,
                                                                    m.d_dw_b + (size_t)i * C, m.d_pw_w + (size_t)i * C * C,
                                                                    m.d_pw_b + (size_t)i * C, B, H, W, w.D[i]);
      } else {
        if (int rc_ = ensure_dyn_smem(dsblock64_kernel<false>, (size_t)kB64Smem)) return rc_;
        dsblock64_kernel<false><<<(unsigned)g64, 256, kB64Smem, st>>>(w.L[i], w.L[i + 1], m.d_dw_w + (size_t)i * 9 * C,
                                                                     m.d_dw_b + (size_t)i * C, m.d_pw_w + (size_t)i * C * C,
                                                                     m.d_pw_b + (size_t)i * C, B, H, W, w.D[i]);
      }
    } else {
      int threads = kBlkPix * cq;
      if (threads > 512) threads = 512;
      dsblock_kernel<<<(unsigned)g, threads, kBlkPix * C * 4, st>>>(
          w.L[i], w.L[i + 1], m.d_dw_w + (size_t)i * 9 * C, m.d_dw_b + (size_t)i * C, m.d_pw_w + (size_t)i * C * C,
          m.d_pw_b + (size_t)i * C, B, H, W, C, w.D[i]);
    }
    KWS_CHECK_LAUNCH(h);
    H += 2;
    W += 2;
  }
  long long g = B < h->sm_count * 8LL ? B : h->sm_count * 8LL;
  pool_fc_kernel<<<(unsigned)g, kPoolThreads, pool_smem(C), st>>>(w.L[nb], w.logits, m.d_fc_w, m.d_fc_b, B, H * W, C, K, w.pooled);
  KWS_CHECK_LAUNCH(h);
  return KWS_OK;
}

}  // namespace kws

using namespace kws;

extern "C" {

void kws_model_free_internal(kws_handle_t h) {
  if (h) model_free(h->model);
}

int kws_model_create(kws_handle_t h, int num_classes, int channels, int num_blocks) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (num_classes < 1 || channels < 4 || channels % 4 != 0 || num_blocks < 0 || num_blocks > 16)
    return fail(KWS_ERR_INVALID, "model: need num_classes>=1, channels a multiple of 4, 0<=num_blocks<=16");
  KWS_CUDA(cudaSetDevice(h->device));
  model_free(h->model);
  ModelWeights& m = h->model;
  m.num_classes = num_classes;
  m.channels = channels;
  m.num_blocks = num_blocks;
  m.names = {"conv1.weight", "conv1.bias"};
  for (int i = 1; i <= num_blocks; ++i) {
    const std::string p = "dsconv" + std::to_string(i);
    m.names.push_back(p + ".depthwise.weight");
    m.names.push_back(p + ".depthwise.bias");
    m.names.push_back(p + ".pointwise.weight");
    m.names.push_back(p + ".pointwise.bias");
  }
  m.names.push_back("fc.weight");
  m.names.push_back("fc.bias");
  m.host.assign(m.names.size(), std::vector<float>());
  m.loaded.assign(m.names.size(), false);
  m.created = true;
  return KWS_OK;
}

int kws_model_load_tensor(kws_handle_t h, const char* name, const float* data, const int64_t* shape,
                          int ndim) {
  if (!h || !h->model.created) return fail(KWS_ERR_STATE, "model not created");
  if (!name || !data || !shape) return fail(KWS_ERR_INVALID, "null argument");
  ModelWeights& m = h->model;
  const int s = slot_of(m, name);
  if (s < 0) return fail(KWS_ERR_INVALID, "unexpected key '%s' in state_dict", name);
  const int64_t C = m.channels, K = m.num_classes;
  std::vector<int64_t> want;
  const std::string n(name);
  if (n == "conv1.weight") want = {C, 1, 10, 10};
  else if (n == "fc.weight") want = {K, C};
  else if (n == "fc.bias") want = {K};
  else if (n.find("depthwise.weight") != std::string::npos) want = {C, 1, 3, 3};
  else if (n.find("pointwise.weight") != std::string::npos) want = {C, C, 1, 1};
  else want = {C};
  bool ok = (int)want.size() == ndim;
  int64_t numel = 1;
  for (int i = 0; ok && i < ndim; ++i) {
    ok = shape[i] == want[i];
    numel *= want[i];
  }
  if (!ok) return fail(KWS_ERR_INVALID, "size mismatch for %s", name);
  m.host[s].assign(data, data + numel);
  m.loaded[s] = true;
  m.finalized = false;
  return KWS_OK;
}

int kws_model_finalize(kws_handle_t h) {
  if (!h || !h->model.created) return fail(KWS_ERR_STATE, "model not created");
  ModelWeights& m = h->model;
  for (size_t i = 0; i < m.names.size(); ++i)
    if (!m.loaded[i]) return fail(KWS_ERR_STATE, "missing key '%s' in state_dict", m.names[i].c_str());
  KWS_CUDA(cudaSetDevice(h->device));
  const int C = m.channels, K = m.num_classes, nb = m.num_blocks;
  auto up = [&](float** dst, const std::vector<float>& v) -> int {
    if (*dst) cudaFree(*dst);
    *dst = nullptr;
    if (v.empty()) return KWS_OK;
    KWS_CUDA(cudaMalloc((void**)dst, v.size() * 4));
    KWS_CUDA(cudaMemcpy(*dst, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
    return KWS_OK;
  };
  int rc;
  {  // conv1.weight [C,1,10,10] -> [100][C]
    const std::vector<float>& w = m.host[slot_of(m, "conv1.weight")];
    std::vector<float> p((size_t)100 * C);
    for (int c = 0; c < C; ++c)
      for (int t = 0; t < 100; ++t) p[(size_t)t * C + c] = w[(size_t)c * 100 + t];
    if ((rc = up(&m.d_conv1_w, p))) return rc;
    if ((rc = up(&m.d_conv1_b, m.host[slot_of(m, "conv1.bias")]))) return rc;
  }
  std::vector<float> dww((size_t)nb * 9 * C), dwb((size_t)nb * C), pww((size_t)nb * C * C), pwb((size_t)nb * C);
  for (int i = 0; i < nb; ++i) {
    const std::string pre = "dsconv" + std::to_string(i + 1);
    const std::vector<float>& a = m.host[slot_of(m, pre + ".depthwise.weight")];  // [C,1,3,3]
    const std::vector<float>& ab = m.host[slot_of(m, pre + ".depthwise.bias")];
    const std::vector<float>& b = m.host[slot_of(m, pre + ".pointwise.weight")];  // [Cout,Cin,1,1]
    const std::vector<float>& bb = m.host[slot_of(m, pre + ".pointwise.bias")];
    for (int c = 0; c < C; ++c) {
      for (int t = 0; t < 9; ++t) dww[((size_t)i * 9 + t) * C + c] = a[(size_t)c * 9 + t];
      dwb[(size_t)i * C + c] = ab[c];
      pwb[(size_t)i * C + c] = bb[c];
      for (int ci = 0; ci < C; ++ci) pww[((size_t)i * C + ci) * C + c] = b[(size_t)c * C + ci];
    }
  }
  if ((rc = up(&m.d_dw_w, dww))) return rc;
  if ((rc = up(&m.d_dw_b, dwb))) return rc;
  if ((rc = up(&m.d_pw_w, pww))) return rc;
  if ((rc = up(&m.d_pw_b, pwb))) return rc;
  {  // fc.weight [K,C] -> [C][K]
    const std::vector<float>& w = m.host[slot_of(m, "fc.weight")];
    std::vector<float> p((size_t)C * K);
    for (int k = 0; k < K; ++k)
      for (int c = 0; c < C; ++c) p[(size_t)c * K + k] = w[(size_t)k * C + c];
    if ((rc = up(&m.d_fc_w, p))) return rc;
    if ((rc = up(&m.d_fc_b, m.host[slot_of(m, "fc.bias")]))) return rc;
  }
  {
    std::vector<const float*> a, ab, b, bb;
    for (int i = 0; i < nb; ++i) {
      const std::string pre = "dsconv" + std::to_string(i + 1);
      a.push_back(m.host[slot_of(m, pre + ".depthwise.weight")].data());
      ab.push_back(m.host[slot_of(m, pre + ".depthwise.bias")].data());
      b.push_back(m.host[slot_of(m, pre + ".pointwise.weight")].data());
      bb.push_back(m.host[slot_of(m, pre + ".pointwise.bias")].data());
    }
    if (C == 64 && nb == 4) {  // parameter images of the fused tensor-core kernel (model_fused.cu)
      if ((rc = build_fused_model(m, m.host[slot_of(m, "conv1.weight")], m.host[slot_of(m, "conv1.bias")], a, ab, b, bb)))
        return rc;
    }
    // chunked weight images of the layer-by-layer tensor-core path (model_gemm.cu), any channel count
    if ((rc = build_gemm_model(m, m.host[slot_of(m, "conv1.weight")], m.host[slot_of(m, "conv1.bias")], ab, b, bb)))
      return rc;
  }
  m.finalized = true;
  return KWS_OK;
}

int kws_softmax(kws_handle_t h, const float* logits_dev, int64_t B, int64_t K, float* out_dev, void* stream) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (B < 0 || K < 1) return fail(KWS_ERR_INVALID, "bad B/K");
  if (B == 0) return KWS_OK;
  if (!logits_dev || !out_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  KWS_CUDA(cudaSetDevice(h->device));
  long long g = (B + 127) / 128;
  if (g > h->sm_count * 16LL) g = h->sm_count * 16LL;
  softmax_kernel<<<(unsigned)g, 128, 0, (cudaStream_t)stream>>>(logits_dev, out_dev, B, (int)K);
  KWS_CHECK_LAUNCH(h);
  return KWS_OK;
}

int kws_model_set_path(kws_handle_t h, int path) {
  if (!h || !h->model.created) return fail(KWS_ERR_STATE, "model not created");
  if (path != 0 && path != 1) return fail(KWS_ERR_INVALID, "path must be 0 (auto) or 1 (layered)");
  h->model.path = path;
  return KWS_OK;
}

int64_t kws_model_workspace_bytes(kws_handle_t h, int64_t B, int64_t T, int64_t F) {
  if (!h || !h->model.created) return fail(KWS_ERR_STATE, "model not created");
  int H1, W1, Hl, Wl;
  conv_out_hw((int)T, (int)F, h->model.num_blocks, &H1, &W1, &Hl, &Wl);
  if (H1 < 1 || W1 < 1) return fail(KWS_ERR_INVALID, "input %lldx%lld too small for conv1", (long long)T, (long long)F);
  const int64_t chunk = B < kModelChunk ? B : kModelChunk;
  return 2 * chunk * Hl * Wl * ((h->model.channels + 31) / 32 * 32) * 4;  // the tensor-core path may pad a pixel to 32 channels
}

int kws_model_forward(kws_handle_t h, const float* feats_dev, int64_t B, int64_t T, int64_t F,
                      float* logits_dev, void* stream) {
  if (!h || !h->model.created) return fail(KWS_ERR_STATE, "model not created");
  ModelWeights& m = h->model;
  if (!m.finalized) return fail(KWS_ERR_STATE, "model weights not finalized");
  if (B < 0 || T < 1 || F < 1) return fail(KWS_ERR_INVALID, "bad B/T/F");
  if (B == 0) return KWS_OK;
  if (!feats_dev || !logits_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  KWS_CUDA(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int C = m.channels, nb = m.num_blocks, K = m.num_classes;
  int H1, W1, Hl, Wl;
  conv_out_hw((int)T, (int)F, nb, &H1, &W1, &Hl, &Wl);
  if (H1 < 1 || W1 < 1) return fail(KWS_ERR_INVALID, "input %lldx%lld too small for conv1 (10x10, stride 2, pad 2)", (long long)T, (long long)F);
  if (m.path == 0) {  // fused tensor-core kernel when the configuration is the reference network
    const int rc = launch_dscnn_fused(h, feats_dev, B, T, F, logits_dev, st);
    if (rc != 0) return rc < 0 ? rc : KWS_OK;
  }
  const int64_t need = kws_model_workspace_bytes(h, B, T, F);
  if ((size_t)need > m.ws_bytes) {
    if (m.d_ws) KWS_CUDA(cudaFree(m.d_ws));
    m.d_ws = nullptr;
    m.ws_bytes = 0;
    KWS_CUDA(cudaMalloc(&m.d_ws, (size_t)need));
    m.ws_bytes = (size_t)need;
  }
  const int64_t chunk = B < kModelChunk ? B : kModelChunk;
  float* buf0 = (float*)m.d_ws;
  float* buf1 = buf0 + chunk * Hl * Wl * ((C + 31) / 32 * 32);
  const int cq = C / 4;
  for (int64_t b0 = 0; b0 < B; b0 += chunk) {
    const int64_t nbat = (B - b0 < chunk) ? B - b0 : chunk;
    if (m.path == 0 && m.gemm) {  // layer-by-layer tensor-core path (model_gemm.cu)
      float* last = nullptr;
      int last_rows = 0;  // rows of [C] per clip that pool_fc averages: pixels, or the last block's pooled partial sums
      const int rc = launch_dscnn_gemm(h, feats_dev + b0 * T * F, nbat, (int)T, (int)F, buf0, buf1, &last, &last_rows, st);
      if (rc < 0) return rc;
      if (rc == 1) {
        long long g = nbat < h->sm_count * 8LL ? nbat : h->sm_count * 8LL;
        pool_fc_kernel<<<(unsigned)g, kPoolThreads, pool_smem(C), st>>>(last, logits_dev + b0 * K, m.d_fc_w, m.d_fc_b, nbat, last_rows, C, K,
                                                        nullptr);
        KWS_CHECK_LAUNCH(h);
        continue;
      }
    }
    {
      const long long total = nbat * H1 * W1 * (long long)cq;
      long long g = (total + 255) / 256;
      if (g > h->sm_count * 16LL) g = h->sm_count * 16LL;
      conv1_relu_kernel<<<(unsigned)g, 256, 0, st>>>(feats_dev + b0 * T * F, buf0, m.d_conv1_w, m.d_conv1_b,
                                                     nbat, (int)T, (int)F, H1, W1, C);
      KWS_CHECK_LAUNCH(h);
    }
    float *cur = buf0, *nxt = buf1;
    int H = H1, W = W1;
    for (int i = 0; i < nb; ++i) {
      const long long npix = nbat * (long long)(H + 2) * (W + 2);
      long long g = (npix + kBlkPix - 1) / kBlkPix;
      if (g > h->sm_count * 16LL) g = h->sm_count * 16LL;
      int threads = kBlkPix * cq;
      if (threads > 512) threads = 512;
      dsblock_kernel<<<(unsigned)g, threads, kBlkPix * C * 4, st>>>(
          cur, nxt, m.d_dw_w + (size_t)i * 9 * C, m.d_dw_b + (size_t)i * C, m.d_pw_w + (size_t)i * C * C,
          m.d_pw_b + (size_t)i * C, nbat, H, W, C, nullptr);
      KWS_CHECK_LAUNCH(h);
      float* t = cur; cur = nxt; nxt = t;
      H += 2; W += 2;
    }
    {
      long long g = nbat < h->sm_count * 8LL ? nbat : h->sm_count * 8LL;
      pool_fc_kernel<<<(unsigned)g, kPoolThreads, pool_smem(C), st>>>(cur, logits_dev + b0 * K, m.d_fc_w, m.d_fc_b, nbat,
                                                      H * W, C, K, nullptr);
      KWS_CHECK_LAUNCH(h);
    }
  }
  return KWS_OK;
}

}  // extern "C"
