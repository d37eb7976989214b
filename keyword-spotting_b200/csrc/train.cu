// One training step of the reference KWS network: the body of KWSTrainer.train_epoch
// (kws/libs/training.py:285-297; duplicate in train.py:39-49) --
//     outputs = model(inputs); loss = CrossEntropyLoss()(outputs, labels); loss.backward(); Adam.step()
// -- for the reference shape (64 channels, 4 blocks).  Parameters, gradients and the Adam moments are
// FLAT fp32 buffers in state_dict order (models.py:135-144: conv1.weight, conv1.bias,
// dsconv{l}.depthwise.{weight,bias}, dsconv{l}.pointwise.{weight,bias}, fc.weight, fc.bias; torch
// layouts), owned by the caller, so that the data-parallel gradient exchange is ONE allreduce of
// 26 444 floats issued by the host between kws_train_forward_backward and kws_train_adam,
// and the 1/world scale is folded into the Adam kernel that consumes the reduced buffer.
//
// Forward = the layered fp32 kernels of model.cu keeping every activation (NHWC).  Backward, per block
// (models.py:108-119; out = ReLU(pointwise_pad1(depthwise(x)))):
//   pw_bwd_kernel : dY = dOut * (out > 0) over the (H+2)x(W+2) output -- border pixels feed only the
//                   pointwise bias gradient (their input is the zero padding); interior pixels give
//                   dWpw += dY^T D and dD = dY Wpw (written over the saved depthwise output D)
//   dw_bwd_kernel : dX = correlate(dD, w), dWdw += X (*) dD, dbdw += sum dD
// then conv1_bwd_kernel (im2col outer products) and ce_fc_bwd_kernel (softmax, mean CE, fc grads,
// gradient of the average pool).  Reductions use per-CTA register/shared partials and one atomicAdd
// per partial, so results are deterministic only up to fp32 summation order.
#include <cstdlib>
#include <cstring>

#include "mfcc_dev.cuh"

namespace kws {

constexpr int kTC = 64;  // channels of the reference network (the only shape the training step covers)

struct FlatLayout {
  int C, nb, K;
  __host__ __device__ int c1w() const { return 0; }
  __host__ __device__ int c1b() const { return C * 100; }
  __host__ __device__ int blk(int l) const { return C * 101 + l * (11 * C + C * C); }
  __host__ __device__ int dww(int l) const { return blk(l); }
  __host__ __device__ int dwb(int l) const { return blk(l) + 9 * C; }
  __host__ __device__ int pww(int l) const { return blk(l) + 10 * C; }
  __host__ __device__ int pwb(int l) const { return blk(l) + 10 * C + C * C; }
  __host__ __device__ int fcw() const { return blk(nb); }
  __host__ __device__ int fcb() const { return blk(nb) + K * C; }
  __host__ __device__ int total() const { return blk(nb) + K * C + K; }
};

// flat (torch layout) parameters -> the packed layouts the forward kernels read (model.cu)
__global__ void repack_kernel(const float* __restrict__ flat, FlatLayout fl, float* c1w, float* c1b, float* dww,
                              float* dwb, float* pww, float* pwb, float* fcw, float* fcb) {
  const int C = fl.C, K = fl.K;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < fl.total(); i += gridDim.x * blockDim.x) {
    const float v = flat[i];
    if (i < fl.c1b()) {
      const int c = i / 100, t = i - c * 100;
      c1w[t * C + c] = v;
    } else if (i < fl.blk(0)) {
      c1b[i - fl.c1b()] = v;
    } else if (i < fl.fcw()) {
      const int per = 11 * C + C * C, l = (i - fl.blk(0)) / per, r = (i - fl.blk(0)) - l * per;
      if (r < 9 * C) {
        const int c = r / 9, t = r - c * 9;
        dww[(l * 9 + t) * C + c] = v;
      } else if (r < 10 * C) {
        dwb[l * C + (r - 9 * C)] = v;
      } else if (r < 10 * C + C * C) {
        const int rr = r - 10 * C, co = rr / C, ci = rr - co * C;
        pww[(l * C + ci) * C + co] = v;
      } else {
        pwb[l * C + (r - 10 * C - C * C)] = v;
      }
    } else if (i < fl.fcb()) {
      const int rr = i - fl.fcw(), k = rr / C, c = rr - k * C;
      fcw[c * K + k] = v;
    } else {
      fcb[i - fl.fcb()] = v;
    }
  }
}

// softmax + mean cross-entropy + fc backward + gradient of the global average pool input.
// block = C threads (thread = channel); samples strided over blocks.  K <= 32.
__global__ void ce_fc_bwd_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                 const float* __restrict__ pooled, const float* __restrict__ fc_w_flat /*[K][C]*/,
                                 float* __restrict__ dpooled, float* __restrict__ g_fcw, float* __restrict__ g_fcb,
                                 float* __restrict__ loss, long long B, int C, int K, float invB) {
  __shared__ float s_dl[32];
  const int c = threadIdx.x;
  float accw[32];
#pragma unroll
  for (int k = 0; k < 32; ++k) accw[k] = 0.f;
  float accb = 0.f, lsum = 0.f;
  for (long long b = blockIdx.x; b < B; b += gridDim.x) {
    __syncthreads();
    if (c == 0) {
      const float* lg = logits + b * K;
      float mx = lg[0];
      for (int k = 1; k < K; ++k) mx = fmaxf(mx, lg[k]);
      float se = 0.f;
      for (int k = 0; k < K; ++k) se += expf(lg[k] - mx);
      const long long yl = labels[b];
      const float lse = mx + logf(se);
      if (yl < 0 || yl >= K) {
        // torch's CrossEntropyLoss raises on a label outside [0, K) (ignore_index is not part of the reference
        // step, training.py:115); a kernel cannot: the sample contributes no gradient and the loss turns NaN
        for (int k = 0; k < K; ++k) s_dl[k] = 0.f;
        lsum = __int_as_float(0x7fc00000);
      } else {
        const int y = (int)yl;
        for (int k = 0; k < K; ++k) s_dl[k] = (expf(lg[k] - lse) - (k == y ? 1.f : 0.f)) * invB;
        lsum += (lse - lg[y]) * invB;
      }
    }
    __syncthreads();
    const float pv = pooled[b * C + c];
    float dp = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      if (k < K) {
        const float dl = s_dl[k];
        dp = fmaf(dl, fc_w_flat[k * C + c], dp);
        accw[k] = fmaf(dl, pv, accw[k]);
      }
    }
    dpooled[b * C + c] = dp;
    if (c < K) accb += s_dl[c];
  }
#pragma unroll
  for (int k = 0; k < 32; ++k)
    if (k < K) atomicAdd(&g_fcw[k * C + c], accw[k]);
  if (c < K) atomicAdd(&g_fcb[c], accb);
  if (c == 0) atomicAdd(loss, lsum);
}


// Pointwise (1x1, padding=1) + ReLU backward of one block, kPwTP output pixels per tile.
//   out  [B, H+2, W+2, C]  saved block output (post-ReLU)        d_io [B, H, W, C] saved depthwise output,
//   dout same shape (or, LAST, dpooled[b][c] * inv_area)           overwritten with its gradient dD
// Two register-tiled fp32 GEMMs per tile run side by side on the two halves of the CTA (packed f32x2 FMAs,
// one 16-byte + two 16-byte shared loads per 32 FMAs):
//   threads   0..127  dD[px][ci]  = sum_co dY[px][co] W[co][ci]   4 px x 8 ci per thread, K = co, written per tile
//   threads 128..255  dW[co][ci] += sum_px dY[px][co] D[px][ci]   4 co x 8 ci per thread, K = px, kept in registers
// dY is staged both as [px][co] and transposed [co][px] so that each GEMM reads its K rows as float4.
constexpr int kPwTP = 64;                      // pixels per tile
constexpr int kPwLd = kPwTP;                   // row pitch of the transposed dY tile (px index XOR-swizzled per channel quad)
constexpr size_t kPwSmem = (size_t)(kTC * kTC + 2 * kPwTP * kTC + kTC * kPwLd) * 4;

__global__ void __launch_bounds__(256, 3) pw_bwd_kernel(const float* __restrict__ out, const float* __restrict__ dout,
                                                     const float* __restrict__ dpooled, float inv_area,
                                                     float* __restrict__ d_io, const float* __restrict__ w_flat /*[co][ci]*/,
                                                     float* __restrict__ g_w, float* __restrict__ g_b, long long B, int H,
                                                     int W) {
  constexpr int C = kTC, TP = kPwTP;
  extern __shared__ __align__(16) float sm_pw[];
  float* s_w = sm_pw;                 // [64 co][64 ci]
  float* s_dy = s_w + C * C;          // [TP px][64 co]
  float* s_d = s_dy + TP * C;         // [TP px][64 ci]
  float* s_dyT = s_d + TP * C;        // [64 co][TP (+4) px]
  const int tid = threadIdx.x;
  const int Ho = H + 2, Wo = W + 2;
  const long long npix = B * Ho * (long long)Wo, ntiles = (npix + TP - 1) / TP;
  for (int i = tid; i < C * C / 4; i += 256) reinterpret_cast<float4*>(s_w)[i] = reinterpret_cast<const float4*>(w_flat)[i];
  const bool wgrad = tid >= 128;
  const int t7 = tid & 127, rg = t7 >> 3, cig = (t7 & 7) * 4;  // row group (4 px or 4 co); ci = cig..cig+3 and 32+cig..32+cig+3
  V accw[4][4];                        // wgrad threads: dW[4 co][8 ci] as ci pairs
  float accb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int jx = 0; jx < 4; ++jx) accw[i][jx] = fsplat(0.f);
  const int q = tid & 15, px0 = tid >> 4;  // staging: channel quad, first pixel (then +16, +32, +48)
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    __syncthreads();
#pragma unroll
    for (int it = 0; it < TP / 16; ++it) {
      const int px = px0 + 16 * it;
      const long long gp = tile * TP + px;
      float4 dy = make_float4(0.f, 0.f, 0.f, 0.f), dv = dy;
      if (gp < npix) {
        int wo, ho;
        long long b;
        split_pixel(gp, Wo, Ho, &b, &ho, &wo);
        const float4 o = reinterpret_cast<const float4*>(out + gp * C)[q];
        float4 g;
        if (dpooled) {
          g = reinterpret_cast<const float4*>(dpooled + b * C)[q];
          g.x *= inv_area; g.y *= inv_area; g.z *= inv_area; g.w *= inv_area;
        } else {
          g = reinterpret_cast<const float4*>(dout + gp * C)[q];
        }
        dy = make_float4(o.x > 0.f ? g.x : 0.f, o.y > 0.f ? g.y : 0.f, o.z > 0.f ? g.z : 0.f, o.w > 0.f ? g.w : 0.f);
        if (ho >= 1 && ho <= H && wo >= 1 && wo <= W)
          dv = reinterpret_cast<const float4*>(d_io + ((b * H + ho - 1) * W + wo - 1) * C)[q];
      }
      reinterpret_cast<float4*>(s_dy + px * C)[q] = dy;
      reinterpret_cast<float4*>(s_d + px * C)[q] = dv;
      const int pxs = px ^ (4 * (q & 7));  // swizzle: the 16 channel quads of a warp hit 8 different bank groups
      s_dyT[(4 * q + 0) * kPwLd + pxs] = dy.x;
      s_dyT[(4 * q + 1) * kPwLd + pxs] = dy.y;
      s_dyT[(4 * q + 2) * kPwLd + pxs] = dy.z;
      s_dyT[(4 * q + 3) * kPwLd + pxs] = dy.w;
    }
    __syncthreads();
    if (!wgrad) {  // dD[4 px][8 ci] = sum_co dY[px][co] W[co][ci]
      V acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int jx = 0; jx < 4; ++jx) acc[i][jx] = fsplat(0.f);
#pragma unroll 4
      for (int co = 0; co < C; ++co) {
        const float4 a = *reinterpret_cast<const float4*>(s_dyT + co * kPwLd + ((4 * rg) ^ (4 * ((co >> 2) & 7))));
        const float4 w0 = *reinterpret_cast<const float4*>(s_w + co * C + cig);
        const float4 w1 = *reinterpret_cast<const float4*>(s_w + co * C + 32 + cig);
        const V wv[4] = {make_float2(w0.x, w0.y), make_float2(w0.z, w0.w), make_float2(w1.x, w1.y), make_float2(w1.z, w1.w)};
        const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int jx = 0; jx < 4; ++jx) acc[i][jx] = ffma(fsplat(av[i]), wv[jx], acc[i][jx]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {  // interior pixels only: the border's input is the zero padding
        const long long gp = tile * TP + 4 * rg + i;
        if (gp < npix) {
          int wo, ho;
          long long b;
          split_pixel(gp, Wo, Ho, &b, &ho, &wo);
          if (ho >= 1 && ho <= H && wo >= 1 && wo <= W) {
            float* dst = d_io + ((b * H + ho - 1) * W + wo - 1) * C + cig;
            *reinterpret_cast<float4*>(dst) = make_float4(acc[i][0].x, acc[i][0].y, acc[i][1].x, acc[i][1].y);
            *reinterpret_cast<float4*>(dst + 32) = make_float4(acc[i][2].x, acc[i][2].y, acc[i][3].x, acc[i][3].y);
          }
        }
      }
    } else {  // dW[4 co][8 ci] += sum_px dY[px][co] D[px][ci]  (D is zero for border pixels); bias gradient alongside
#pragma unroll 4
      for (int px = 0; px < TP; ++px) {
        const float4 a = *reinterpret_cast<const float4*>(s_dy + px * C + 4 * rg);
        const float4 d0 = *reinterpret_cast<const float4*>(s_d + px * C + cig);
        const float4 d1 = *reinterpret_cast<const float4*>(s_d + px * C + 32 + cig);
        const V dv[4] = {make_float2(d0.x, d0.y), make_float2(d0.z, d0.w), make_float2(d1.x, d1.y), make_float2(d1.z, d1.w)};
        const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
          for (int jx = 0; jx < 4; ++jx) accw[i][jx] = ffma(fsplat(av[i]), dv[jx], accw[i][jx]);
          accb[i] += av[i];
        }
      }
    }
  }
  if (wgrad) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
      for (int jx = 0; jx < 4; ++jx) {
        const int ci = cig + (jx >> 1) * 32 + (jx & 1) * 2;
        atomicAdd(&g_w[(4 * rg + i) * C + ci], accw[i][jx].x);
        atomicAdd(&g_w[(4 * rg + i) * C + ci + 1], accw[i][jx].y);
      }
      if (cig == 0) atomicAdd(&g_b[4 * rg + i], accb[i]);
    }
  }
}

// Depthwise 3x3 (padding 1) backward: thread = (pixel slot, channel quad).
//   x [B,H,W,C] block input, dd [B,H,W,C] gradient of the depthwise output, dx [B,H,W,C] (pre-ReLU-mask)
__global__ void __launch_bounds__(256) dw_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dd,
                                                     float* __restrict__ dx, const float* __restrict__ w_packed /*[9][C]*/,
                                                     float* __restrict__ g_w /*[C][9]*/, float* __restrict__ g_b,
                                                     long long B, int H, int W) {
  constexpr int C = kTC;
  __shared__ float s_red[8][16 * 40];
  const int tid = threadIdx.x, q = tid & 15, ps = tid >> 4;
  float4 wv[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) wv[t] = reinterpret_cast<const float4*>(w_packed + t * C)[q];
  float4 wacc[9], bacc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int t = 0; t < 9; ++t) wacc[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  const long long npix = B * H * (long long)W;
  for (long long gp = blockIdx.x * 16LL + ps; gp < npix; gp += gridDim.x * 16LL) {
    int xx, yy;
    long long b;
    split_pixel(gp, W, H, &b, &yy, &xx);
    const float4 xv = reinterpret_cast<const float4*>(x + gp * C)[q];
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int y2 = yy + 1 - i;
      if (y2 < 0 || y2 >= H) continue;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int x2 = xx + 1 - j;
        if (x2 < 0 || x2 >= W) continue;
        const float4 g = reinterpret_cast<const float4*>(dd + ((b * H + y2) * W + x2) * C)[q];
        const float4 k = wv[i * 3 + j];
        acc.x = fmaf(g.x, k.x, acc.x); acc.y = fmaf(g.y, k.y, acc.y); acc.z = fmaf(g.z, k.z, acc.z); acc.w = fmaf(g.w, k.w, acc.w);
        float4& wa = wacc[i * 3 + j];
        wa.x = fmaf(g.x, xv.x, wa.x); wa.y = fmaf(g.y, xv.y, wa.y); wa.z = fmaf(g.z, xv.z, wa.z); wa.w = fmaf(g.w, xv.w, wa.w);
        if (i == 1 && j == 1) { bacc.x += g.x; bacc.y += g.y; bacc.z += g.z; bacc.w += g.w; }
      }
    }
    reinterpret_cast<float4*>(dx + gp * C)[q] = acc;
  }
  // reduce the 16 pixel slots of the CTA: lanes l and l+16 share a quad, then 8 warps through shared memory
  float vals[40];
#pragma unroll
  for (int t = 0; t < 9; ++t) { vals[4 * t] = wacc[t].x; vals[4 * t + 1] = wacc[t].y; vals[4 * t + 2] = wacc[t].z; vals[4 * t + 3] = wacc[t].w; }
  vals[36] = bacc.x; vals[37] = bacc.y; vals[38] = bacc.z; vals[39] = bacc.w;
  const int warp = tid >> 5, lane = tid & 31;
#pragma unroll
  for (int i = 0; i < 40; ++i) {
    vals[i] += __shfl_xor_sync(0xffffffffu, vals[i], 16);
    if (lane < 16) s_red[warp][lane * 40 + i] = vals[i];
  }
  __syncthreads();
  for (int i = tid; i < 16 * 40; i += 256) {
    float s = 0.f;
#pragma unroll
    for (int wp = 0; wp < 8; ++wp) s += s_red[wp][i];
    const int qq = i / 40, r = i - qq * 40;
    if (r < 36) atomicAdd(&g_w[(qq * 4 + (r & 3)) * 9 + (r >> 2)], s);
    else atomicAdd(&g_b[qq * 4 + (r - 36)], s);
  }
}


// Depthwise 3x3 backward, row-walking form: thread = (clip, column, row segment, channel PAIR); it walks down its rows with
// the 3 x 3 window of dd in registers (three 8-byte loads of dd and one of x per pixel instead of nine 16-byte gathers with
// bounds tests and a pixel-index division: the gather kernel above spends ~100 instructions per pixel and channel pair,
// this one ~40, all arithmetic packed f32x2).  Lanes = the 32 channel pairs of a pixel: every access is one 256-byte run.
constexpr int kDwSeg = 2;  // row segments per column (more CTAs for the narrow layers; one halo row each side)
__global__ void __launch_bounds__(256) dw_bwd_rows_kernel(const float* __restrict__ x, const float* __restrict__ dd,
                                                          float* __restrict__ dx, const float* __restrict__ w_packed /*[9][C]*/,
                                                          float* __restrict__ g_w /*[C][9]*/, float* __restrict__ g_b,
                                                          long long B, int H, int W) {
  constexpr int C = kTC;
  __shared__ float s_red[8][32 * 20];
  const int tid = threadIdx.x, cp = tid & 31, slot = tid >> 5;
  const long long ncol = B * W * (long long)kDwSeg;
  float2 wv[9], wacc[9], bacc = make_float2(0.f, 0.f);
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    wv[t] = reinterpret_cast<const float2*>(w_packed + t * C)[cp];
    wacc[t] = make_float2(0.f, 0.f);
  }
  const float2 z2 = make_float2(0.f, 0.f);
  const int seg_rows = (H + kDwSeg - 1) / kDwSeg;
  for (long long col = blockIdx.x * 8LL + slot; col < ncol; col += gridDim.x * 8LL) {
    const unsigned c32 = (unsigned)col;  // B * W * kDwSeg < 2^31 (the host checks)
    const unsigned seg = c32 % (unsigned)kDwSeg, bx = c32 / (unsigned)kDwSeg, xx = bx % (unsigned)W, b = bx / (unsigned)W;
    const int y0 = (int)seg * seg_rows, y1 = min(H, y0 + seg_rows);
    const bool hasl = xx > 0, hasr = (int)xx + 1 < W;
    const float2* ddc = reinterpret_cast<const float2*>(dd + ((long long)b * H * W + xx) * C) + cp;  // row 0 of this column
    const float2* xc = reinterpret_cast<const float2*>(x + ((long long)b * H * W + xx) * C) + cp;
    float2* dxc = reinterpret_cast<float2*>(dx + ((long long)b * H * W + xx) * C) + cp;
    const int rs = W * (C / 2), cs = C / 2;  // row / column stride in float2
    auto load3 = [&](int y, float2 (&r)[3]) {  // dd[y][xx-1 .. xx+1], zeros outside the layer
      const bool ok = y >= 0 && y < H;
      const float2* q = ddc + (long long)y * rs;
      r[0] = ok && hasl ? __ldg(q - cs) : z2;
      r[1] = ok ? __ldg(q) : z2;
      r[2] = ok && hasr ? __ldg(q + cs) : z2;
    };
    float2 ra[3], rb[3], rc[3];  // rows y - 1, y, y + 1 in rotation
    load3(y0 - 1, ra);
    load3(y0, rb);
    float2 xv = y0 < y1 ? __ldg(xc + (long long)y0 * rs) : z2;
    // one pixel: m = row y - 1, c = row y, n = row y + 1
    auto pixel = [&](int y, const float2 (&m)[3], const float2 (&c)[3], const float2 (&n)[3], float2 xcur) {
      float2 acc = z2;
#pragma unroll
      for (int j = 0; j < 3; ++j) {  // tap (i, j) reads dd[y + 1 - i][xx + 1 - j]
        acc = __ffma2_rn(n[2 - j], wv[j], acc);
        acc = __ffma2_rn(c[2 - j], wv[3 + j], acc);
        acc = __ffma2_rn(m[2 - j], wv[6 + j], acc);
        wacc[j] = __ffma2_rn(n[2 - j], xcur, wacc[j]);
        wacc[3 + j] = __ffma2_rn(c[2 - j], xcur, wacc[3 + j]);
        wacc[6 + j] = __ffma2_rn(m[2 - j], xcur, wacc[6 + j]);
      }
      bacc = __fadd2_rn(bacc, c[1]);
      dxc[(long long)y * rs] = acc;
    };
    for (int y = y0; y < y1; y += 3) {
      float2 xn;
      load3(y + 1, rc);
      xn = y + 1 < y1 ? __ldg(xc + (long long)(y + 1) * rs) : z2;
      pixel(y, ra, rb, rc, xv);
      if (y + 1 >= y1) break;
      load3(y + 2, ra);
      xv = y + 2 < y1 ? __ldg(xc + (long long)(y + 2) * rs) : z2;
      pixel(y + 1, rb, rc, ra, xn);
      if (y + 2 >= y1) break;
      load3(y + 3, rb);
      xn = y + 3 < y1 ? __ldg(xc + (long long)(y + 3) * rs) : z2;
      pixel(y + 2, rc, ra, rb, xv);
      xv = xn;
    }
  }
  // reduce the 8 columns of the CTA that share a channel pair, then one atomic per value
  float* mine = &s_red[slot][cp * 20];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    mine[2 * t] = wacc[t].x;
    mine[2 * t + 1] = wacc[t].y;
  }
  mine[18] = bacc.x;
  mine[19] = bacc.y;
  __syncthreads();
  for (int i = tid; i < 32 * 20; i += 256) {
    float sum = 0.f;
#pragma unroll
    for (int sl = 0; sl < 8; ++sl) sum += s_red[sl][i];
    const int pr = i / 20, r = i - pr * 20, ch = 2 * pr + (r & 1);
    if (r < 18) atomicAdd(&g_w[ch * 9 + (r >> 1)], sum);
    else atomicAdd(&g_b[ch], sum);
  }
}

// conv1 (10x10, stride 2, padding 2, one input channel) + ReLU backward: dW[c][tap] += dY[px][c] * patch[px][tap]
//   l0 [B,H1,W1,C] saved conv1 output, dl0 gradient w.r.t. it (pre-mask), feats [B,T,F]
__global__ void __launch_bounds__(320) conv1_bwd_kernel(const float* __restrict__ l0, const float* __restrict__ dl0,
                                                        const float* __restrict__ feats, float* __restrict__ g_w /*[C][100]*/,
                                                        float* __restrict__ g_b, long long B, int T, int F, int H1, int W1) {
  constexpr int C = kTC, TP = 32;
  __shared__ __align__(16) float s_dy[TP * C];
  __shared__ __align__(16) float s_col[TP * 100];
  const int tid = threadIdx.x, c = tid & 63, tg = tid >> 6;
  float acc[20];  // 64 channels x 5 groups of 20 taps: a group's taps of a pixel are five 16-byte shared-memory reads
#pragma unroll
  for (int i = 0; i < 20; ++i) acc[i] = 0.f;
  float accb = 0.f;
  const long long npix = B * H1 * (long long)W1, ntiles = (npix + TP - 1) / TP;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    __syncthreads();
    for (int idx = tid; idx < TP * (C / 4); idx += 320) {
      const int px = idx >> 4, q = idx & 15;
      const long long gp = tile * TP + px;
      float4 dy = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gp < npix) {
        const float4 o = reinterpret_cast<const float4*>(l0 + gp * C)[q];
        const float4 g = reinterpret_cast<const float4*>(dl0 + gp * C)[q];
        dy = make_float4(o.x > 0.f ? g.x : 0.f, o.y > 0.f ? g.y : 0.f, o.z > 0.f ? g.z : 0.f, o.w > 0.f ? g.w : 0.f);
      }
      reinterpret_cast<float4*>(s_dy + px * C)[q] = dy;
    }
    for (int idx = tid; idx < TP * 100; idx += 320) {
      const int px = idx / 100, tap = idx - px * 100, ki = tap / 10, kj = tap - ki * 10;
      const long long gp = tile * TP + px;
      float v = 0.f;
      if (gp < npix) {
        int wo, ho;  // 32-bit division when the index fits: the 64-bit % and / here cost more than the gradient's FMAs
        long long b;
        split_pixel(gp, W1, H1, &b, &ho, &wo);
        const int fr = 2 * ho - 2 + ki, fc = 2 * wo - 2 + kj;
        if (fr >= 0 && fr < T && fc >= 0 && fc < F) v = feats[(b * T + fr) * F + fc];
      }
      s_col[idx] = v;
    }
    __syncthreads();
#pragma unroll 2
    for (int px = 0; px < TP; ++px) {
      const float a = s_dy[px * C + c];
      if (tg == 0) accb += a;
      const float4* col = reinterpret_cast<const float4*>(s_col + px * 100 + tg * 20);
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const float4 v = col[i];
        acc[4 * i] = fmaf(a, v.x, acc[4 * i]);
        acc[4 * i + 1] = fmaf(a, v.y, acc[4 * i + 1]);
        acc[4 * i + 2] = fmaf(a, v.z, acc[4 * i + 2]);
        acc[4 * i + 3] = fmaf(a, v.w, acc[4 * i + 3]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 20; ++i) atomicAdd(&g_w[c * 100 + tg * 20 + i], acc[i]);
  if (tg == 0) atomicAdd(&g_b[c], accb);
}

// Adam (torch.optim.Adam defaults as used by training.py:119-123; weight_decay 0) on the flat buffers;
// gscale = 1/world_size folds the data-parallel mean into the consumer of the allreduced gradient.
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int n, float lr, float b1, float b2, float eps, float bc1, float bc2,
                            float gscale) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float gi = g[i] * gscale;
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= (lr / bc1) * mi / (sqrtf(vi) / sqrtf(bc2) + eps);
  }
}

static int ensure_train_ws(kws_handle_t h, int64_t B, int T, int F) {
  const ModelWeights& m = h->model;
  const int C = m.channels, nb = m.num_blocks, K = m.num_classes;
  const int H1 = (T + 4 - 10) / 2 + 1, W1 = (F + 4 - 10) / 2 + 1;
  size_t need = 0;
  auto take = [&](size_t floats) { size_t o = need; need += (floats * 4 + 255) & ~(size_t)255; return o; };
  size_t oL[5], oD[4], oX[2];
  for (int l = 0; l <= nb; ++l) oL[l] = take((size_t)B * (H1 + 2 * l) * (W1 + 2 * l) * C);
  for (int l = 0; l < nb; ++l) oD[l] = take((size_t)B * (H1 + 2 * l) * (W1 + 2 * l) * C);
  for (int i = 0; i < 2; ++i) oX[i] = take((size_t)B * (H1 + 2 * (nb - 1)) * (W1 + 2 * (nb - 1)) * C);
  const size_t oP = take((size_t)B * C), oDP = take((size_t)B * C), oLg = take((size_t)B * K);
  TrainWorkspace& w = h->train;
  if (need > w.bytes) {
    if (w.base) KWS_CUDA(cudaFree(w.base));
    w.base = nullptr;
    w.bytes = 0;
    KWS_CUDA(cudaMalloc(&w.base, need));
    w.bytes = need;
  }
  char* b = (char*)w.base;
  for (int l = 0; l <= nb; ++l) w.L[l] = (float*)(b + oL[l]);
  for (int l = 0; l < nb; ++l) w.D[l] = (float*)(b + oD[l]);
  for (int i = 0; i < 2; ++i) w.dX[i] = (float*)(b + oX[i]);
  w.pooled = (float*)(b + oP);
  w.dpooled = (float*)(b + oDP);
  w.logits = (float*)(b + oLg);
  return KWS_OK;
}

}  // namespace kws

using namespace kws;

extern "C" {

int64_t kws_train_num_params(kws_handle_t h) {
  if (!h || !h->model.created) return fail(KWS_ERR_STATE, "model not created");
  FlatLayout fl{h->model.channels, h->model.num_blocks, h->model.num_classes};
  return fl.total();
}

int kws_train_forward_backward(kws_handle_t h, const float* params_dev, const float* feats_dev,
                               const int64_t* labels_dev, int64_t B, int64_t T, int64_t F, float* grads_dev,
                               float* loss_dev, float* logits_out_dev, void* stream) {
  if (!h || !h->model.created || !h->model.finalized) return fail(KWS_ERR_STATE, "model weights not finalized");
  const ModelWeights& m = h->model;
  if (m.channels != kTC || m.num_blocks != 4 || m.num_classes > 32)
    return fail(KWS_ERR_INVALID, "training step covers the reference network only (64 channels, 4 blocks, <= 32 classes)");
  if (B < 1 || T < 6 || F < 6) return fail(KWS_ERR_INVALID, "bad B/T/F");
  if (!params_dev || !feats_dev || !labels_dev || !grads_dev || !loss_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  KWS_CUDA(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  int rc = ensure_train_ws(h, B, (int)T, (int)F);
  if (rc) return rc;
  TrainWorkspace& w = h->train;
  const int C = kTC, nb = 4, K = m.num_classes;
  FlatLayout fl{C, nb, K};
  const int n = fl.total();
  repack_kernel<<<(n + 255) / 256, 256, 0, st>>>(params_dev, fl, m.d_conv1_w, m.d_conv1_b, m.d_dw_w, m.d_dw_b, m.d_pw_w,
                                                 m.d_pw_b, m.d_fc_w, m.d_fc_b);
  KWS_CHECK_LAUNCH(h);
  rc = train_forward_layers(h, feats_dev, B, (int)T, (int)F, w, st);
  if (rc) return rc;
  KWS_CUDA(cudaMemsetAsync(grads_dev, 0, (size_t)n * 4, st));
  KWS_CUDA(cudaMemsetAsync(loss_dev, 0, 4, st));
  if (logits_out_dev) KWS_CUDA(cudaMemcpyAsync(logits_out_dev, w.logits, (size_t)B * K * 4, cudaMemcpyDeviceToDevice, st));
  const int H1 = (int)((T + 4 - 10) / 2 + 1), W1 = (int)((F + 4 - 10) / 2 + 1);
  {
    long long g = B < h->sm_count * 4LL ? B : h->sm_count * 4LL;
    ce_fc_bwd_kernel<<<(unsigned)g, C, 0, st>>>(w.logits, (const long long*)labels_dev, w.pooled, params_dev + fl.fcw(),
                                                w.dpooled, grads_dev + fl.fcw(), grads_dev + fl.fcb(), loss_dev, B, C, K,
                                                1.0f / (float)B);
    KWS_CHECK_LAUNCH(h);
  }
  const size_t pw_smem = kPwSmem;
  if (int rc_ = ensure_dyn_smem(pw_bwd_kernel, (size_t)pw_smem)) return rc_;
  // KWS_TRAIN_TC=0 selects the CUDA-core pointwise backward (A/B and fallback for debugging); default: tensor core
  static const bool use_tc = [] {
    const char* e = std::getenv("KWS_TRAIN_TC");
    return !(e && e[0] == '0');
  }();
  const float* dout = nullptr;
  for (int l = nb - 1; l >= 0; --l) {
    const int H = H1 + 2 * l, W = W1 + 2 * l;  // block input size; output is (H+2) x (W+2)
    const long long ntiles = (B * (long long)(H + 2) * (W + 2) + kPwTP - 1) / kPwTP;
    long long g = ntiles < h->sm_count * 3LL ? ntiles : h->sm_count * 3LL;
    const bool last = (l == nb - 1);
    if (use_tc) {  // tensor-core twin (train_tc.cu)
      const int rc_tc = launch_pw_bwd_tc(h, w.L[l + 1], dout, last ? w.dpooled : nullptr, 1.0f / (float)((H + 2) * (W + 2)),
                                         w.D[l], params_dev + fl.pww(l), grads_dev + fl.pww(l), grads_dev + fl.pwb(l), B, H, W,
                                         st);
      if (rc_tc < 0) return rc_tc;
    } else {
      pw_bwd_kernel<<<(unsigned)g, 256, pw_smem, st>>>(w.L[l + 1], dout, last ? w.dpooled : nullptr,
                                                       1.0f / (float)((H + 2) * (W + 2)), w.D[l], params_dev + fl.pww(l),
                                                       grads_dev + fl.pww(l), grads_dev + fl.pwb(l), B, H, W);
      KWS_CHECK_LAUNCH(h);
    }
    float* dx = w.dX[l & 1];
    const long long npix = B * (long long)H * W;
    static const bool dw_rows = [] { const char* e = std::getenv("KWS_TRAIN_DWROWS"); return !(e && e[0] == '0'); }();
    const long long ncol = B * (long long)W * kDwSeg;
    if (dw_rows && ncol < 0x7fffffffLL) {
      long long g2 = (ncol + 7) / 8;
      if (g2 > h->sm_count * 16LL) g2 = h->sm_count * 16LL;
      dw_bwd_rows_kernel<<<(unsigned)g2, 256, 0, st>>>(w.L[l], w.D[l], dx, m.d_dw_w + (size_t)l * 9 * C, grads_dev + fl.dww(l),
                                                       grads_dev + fl.dwb(l), B, H, W);
    } else {
      long long g2 = (npix + 15) / 16;
      if (g2 > h->sm_count * 8LL) g2 = h->sm_count * 8LL;
      dw_bwd_kernel<<<(unsigned)g2, 256, 0, st>>>(w.L[l], w.D[l], dx, m.d_dw_w + (size_t)l * 9 * C, grads_dev + fl.dww(l),
                                                  grads_dev + fl.dwb(l), B, H, W);
    }
    KWS_CHECK_LAUNCH(h);
    dout = dx;
  }
  {
    const long long ntiles = (B * (long long)H1 * W1 + 31) / 32;
    long long g = ntiles < h->sm_count * 4LL ? ntiles : h->sm_count * 4LL;
    conv1_bwd_kernel<<<(unsigned)g, 320, 0, st>>>(w.L[0], dout, feats_dev, grads_dev + fl.c1w(), grads_dev + fl.c1b(), B,
                                                  (int)T, (int)F, H1, W1);
    KWS_CHECK_LAUNCH(h);
  }
  return KWS_OK;
}

int kws_train_adam(kws_handle_t h, float* params_dev, const float* grads_dev, float* m_dev, float* v_dev, int64_t n,
                   int64_t step, float lr, float beta1, float beta2, float eps, float grad_scale, void* stream) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (!params_dev || !grads_dev || !m_dev || !v_dev || n < 1 || step < 1) return fail(KWS_ERR_INVALID, "bad argument");
  KWS_CUDA(cudaSetDevice(h->device));
  const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
  adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(params_dev, grads_dev, m_dev, v_dev, (int)n, lr,
                                                                           beta1, beta2, eps, bc1, bc2, grad_scale);
  KWS_CHECK_LAUNCH(h);
  return KWS_OK;
}

}  // extern "C"
