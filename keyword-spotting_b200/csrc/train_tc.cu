// Pointwise (1x1, padding=1) + ReLU backward of one block on the tensor core (tcgen05, kind::tf32 with the 3-term
// hi/lo split = fp32-grade): the tensor-core twin of pw_bwd_kernel in train.cu (reference step body
// kws/libs/training.py:285-297; layer kws/libs/models.py:104-119).
//
//   dY[px][co] = dOut[px][co] * (out[px][co] > 0)                  over the padded (H+2) x (W+2) output grid
//   dD[px][ci] = sum_co dY[px][co] W[co][ci]                       dgrad, written for interior pixels only
//   dW[co][ci] = sum_px dY[px][co] D[px][ci],  db[co] = sum_px dY  wgrad + bias gradient (D = 0 on the border)
//
// Tile = 64 consecutive pixels of the padded grid.  Both GEMMs are M = 64:
//   dgrad  D1[px, ci]  = A1[px, K = co] . B1[ci, K = co]^T     A1 = dY (K-major as it lies in NHWC), B1 = W^T, built once
//   wgrad  D2[co, ci+] = A2[co, K = px] . B2[ci+, K = px]^T    A2 = dY^T, B2 = D^T with an extra all-ones row (n = 64)
//                                                             whose column of D2 is the bias gradient
// kind::tf32 returns zeros for MN-major operands (tools/umma_probe_mn*.cu), so the K = px operands are staged
// TRANSPOSED by the CUDA cores: thread = (16 pixels x 2 channel quads per warp), which makes both the K-major
// 16-byte stores of A1 and the transposed 4-byte stores of A2 / B2 bank-conflict free with the 144-byte K-chunk pitch.
// Every operand is split hi = rna_tf32(x), lo = rna_tf32(x - hi) and each GEMM is Ahi.Bhi + Alo.Bhi + Ahi.Blo.
// The wgrad accumulator stays in tensor memory across all tiles of the CTA and is added to the global gradient once.
#include "mfcc_dev.cuh"

namespace kws {

constexpr int kPC = 64;                        // channels
constexpr int kPTile = 64;                     // pixels per tile
constexpr int kPThreads = 512;
constexpr int kPLbo = 144;                     // K-chunk pitch (16-byte chunks of 8 rows = 128 B, padded)
constexpr int kPSbo = (64 / 4) * kPLbo;        // 8-row group pitch for K = 64
constexpr int kPOp = 8 * kPSbo;                // one 64-row operand image (hi or lo): 18 432 B
constexpr int kPOpB2 = 9 * kPSbo;              // B2: 72 rows (64 ci + ones row + 7 zero rows)
constexpr int kPOffA1 = 0, kPOffB1 = 2 * kPOp, kPOffA2 = 4 * kPOp, kPOffB2 = 6 * kPOp;
constexpr size_t kPSmem = 6 * (size_t)kPOp + 2 * (size_t)kPOpB2;
constexpr int kPN2 = 72;
constexpr uint32_t kPIdesc1 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((64u >> 4) << 24);
constexpr uint32_t kPIdesc2 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kPN2 >> 3) << 17) | ((64u >> 4) << 24);

__device__ __forceinline__ uint32_t kmaj_off(int row, int k) {  // byte offset of element (row, k) in a K-major image
  return (uint32_t)((row >> 3) * kPSbo + (k >> 2) * kPLbo + (row & 7) * 16 + (k & 3) * 4);
}
// pull [ptr, ptr + bytes) into L2 (bytes a multiple of 16, ptr 16-byte aligned): the register loads that follow one tile
// later then see L2 latency, which the MMAs and the drain of a tile do hide, instead of DRAM latency, which they do not
__device__ __forceinline__ void l2_prefetch(const void* ptr, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
  hi = to_tf32(x);
  lo = to_tf32(x - hi);
}

__global__ void __launch_bounds__(kPThreads, 1)
pw_bwd_tc_kernel(const float* __restrict__ out, const float* __restrict__ dout, const float* __restrict__ dpooled,
                 float inv_area, float* __restrict__ d_io, const float* __restrict__ w_flat /*[co][ci]*/,
                 float* __restrict__ g_w, float* __restrict__ g_b, long long B, int H, int W) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const uint32_t bar = smem_u32(&s_bar);
  const int Ho = H + 2, Wo = W + 2;
  const long long npix = B * Ho * (long long)Wo, ntiles = (npix + kPTile - 1) / kPTile;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  // B1 = W^T (rows ci, K = co), hi and lo; B2's constant rows: ones (n = 64) and zeros (n = 65..71)
  for (int i = tid; i < kPC * kPC; i += kPThreads) {
    const int co = i >> 6, ci = i & 63;
    float hi, lo;
    split_tf32(w_flat[i], hi, lo);
    *reinterpret_cast<float*>(smem + kPOffB1 + kmaj_off(ci, co)) = hi;
    *reinterpret_cast<float*>(smem + kPOffB1 + kPOp + kmaj_off(ci, co)) = lo;
  }
  for (int i = tid; i < 8 * kPTile; i += kPThreads) {
    const int n = 64 + (i >> 6), k = i & 63;
    *reinterpret_cast<float*>(smem + kPOffB2 + kmaj_off(n, k)) = (n == 64) ? 1.0f : 0.0f;
    *reinterpret_cast<float*>(smem + kPOffB2 + kPOpB2 + kmaj_off(n, k)) = 0.0f;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const uint32_t sbase = smem_u32(smem);
  const uint64_t dA1h = umma_desc(sbase + kPOffA1, kPLbo, kPSbo), dA1l = umma_desc(sbase + kPOffA1 + kPOp, kPLbo, kPSbo);
  const uint64_t dB1h = umma_desc(sbase + kPOffB1, kPLbo, kPSbo), dB1l = umma_desc(sbase + kPOffB1 + kPOp, kPLbo, kPSbo);
  const uint64_t dA2h = umma_desc(sbase + kPOffA2, kPLbo, kPSbo), dA2l = umma_desc(sbase + kPOffA2 + kPOp, kPLbo, kPSbo);
  const uint64_t dB2h = umma_desc(sbase + kPOffB2, kPLbo, kPSbo), dB2l = umma_desc(sbase + kPOffB2 + kPOpB2, kPLbo, kPSbo);

  // staging item of this thread: 16 pixels x 2 channel quads per warp (see the header), two items per tile
  const int pxl = lane & 15, q = 2 * (warp & 7) + (lane >> 4), pxb = (warp >> 3) * 16 + pxl;  // + 32 for the second item
  uint32_t parity = 0;
  bool first = true;
  // The global loads of tile i + 1 are issued right after the MMAs of tile i and land in registers while those MMAs
  // and the epilogue run (one CTA per SM: nothing else would hide the L2 / HBM latency): dy = masked upstream
  // gradient, dv = saved depthwise output, per item.
  // The loaded values are kept RAW (layer output ro, upstream gradient rg): masking them here would make the thread wait
  // for its own loads right after issuing them (ncu: 40 % of the kernel's warp samples sat on that select).
  float4 ro[2], rg[2], dv[2];
  auto load_tile = [&](long long tile) {
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const long long gp = tile * kPTile + pxb + 32 * it;
      ro[it] = make_float4(0.f, 0.f, 0.f, 0.f);
      rg[it] = ro[it];
      dv[it] = ro[it];
      if (gp < npix) {
        int wo, ho;
        long long b;
        split_pixel(gp, Wo, Ho, &b, &ho, &wo);
        ro[it] = reinterpret_cast<const float4*>(out + gp * kPC)[q];
        rg[it] = dpooled ? reinterpret_cast<const float4*>(dpooled + b * kPC)[q] : reinterpret_cast<const float4*>(dout + gp * kPC)[q];
        if (ho >= 1 && ho <= H && wo >= 1 && wo <= W)
          dv[it] = reinterpret_cast<const float4*>(d_io + ((b * H + ho - 1) * W + wo - 1) * kPC)[q];
      }
    }
  };
  const float gscale = dpooled ? inv_area : 1.0f;
  // tile t's inputs are contiguous ranges: 64 pixels of out / dout, and the interior pixels among them of d_io
  auto prefetch_tile = [&](long long tile) {
    if (tid == 0 && tile < ntiles) {
      const long long gp0 = tile * kPTile;
      const uint32_t bytes = (uint32_t)(min((long long)kPTile, npix - gp0) * kPC * 4);
      l2_prefetch(out + gp0 * kPC, bytes);
      if (!dpooled) l2_prefetch(dout + gp0 * kPC, bytes);
      int wo, ho;
      long long b;
      split_pixel(gp0, Wo, Ho, &b, &ho, &wo);
      const int hi = min(max(ho - 1, 0), H - 1), wi = min(max(wo - 1, 0), W - 1);
      const long long d0 = (b * H + hi) * W + wi, dn = min((long long)kPTile, B * H * (long long)W - d0);
      l2_prefetch(d_io + d0 * kPC, (uint32_t)(dn * kPC * 4));
    }
  };
  if ((long long)blockIdx.x < ntiles) load_tile(blockIdx.x);
  prefetch_tile((long long)blockIdx.x + gridDim.x);
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    prefetch_tile(tile + 2LL * gridDim.x);
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int px = pxb + 32 * it;
      float yh[4], yl[4], dh[4], dl[4];
      const float4 o = ro[it], g = rg[it];  // dY = upstream gradient where the layer's output is positive
      const float4 dyv = make_float4(o.x > 0.f ? g.x * gscale : 0.f, o.y > 0.f ? g.y * gscale : 0.f, o.z > 0.f ? g.z * gscale : 0.f,
                                     o.w > 0.f ? g.w * gscale : 0.f);
      split_tf32(dyv.x, yh[0], yl[0]); split_tf32(dyv.y, yh[1], yl[1]); split_tf32(dyv.z, yh[2], yl[2]); split_tf32(dyv.w, yh[3], yl[3]);
      split_tf32(dv[it].x, dh[0], dl[0]); split_tf32(dv[it].y, dh[1], dl[1]); split_tf32(dv[it].z, dh[2], dl[2]); split_tf32(dv[it].w, dh[3], dl[3]);
      // A1 (rows px, K = co): one 16-byte chunk
      const uint32_t o1 = kmaj_off(px, 4 * q);
      *reinterpret_cast<float4*>(smem + kPOffA1 + o1) = make_float4(yh[0], yh[1], yh[2], yh[3]);
      *reinterpret_cast<float4*>(smem + kPOffA1 + kPOp + o1) = make_float4(yl[0], yl[1], yl[2], yl[3]);
      // A2 (rows co, K = px) and B2 (rows ci, K = px): transposed, 4-byte stores
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t o2 = kmaj_off(4 * q + i, px);
        *reinterpret_cast<float*>(smem + kPOffA2 + o2) = yh[i];
        *reinterpret_cast<float*>(smem + kPOffA2 + kPOp + o2) = yl[i];
        *reinterpret_cast<float*>(smem + kPOffB2 + o2) = dh[i];
        *reinterpret_cast<float*>(smem + kPOffB2 + kPOpB2 + o2) = dl[i];
      }
    }
    fence_proxy_async();  // generic-proxy writes of the operand images must be visible to the tensor core
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {  // converged warp, elected lane issues
      tc_fence_after();
      // descriptors = tile-invariant bases (built once, above) + a K step's 18 sixteen-byte units: assembling a descriptor
      // per MMA on the uniform datapath cost more than the MMA (48 per tile)
#pragma unroll
      for (int term = 0; term < 3; ++term) {
        const uint64_t a1 = term == 1 ? dA1l : dA1h, b1 = term == 2 ? dB1l : dB1h;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc_mma_tf32_elect(tmem, a1 + (uint64_t)(ks * ((2 * kPLbo) >> 4)), b1 + (uint64_t)(ks * ((2 * kPLbo) >> 4)), kPIdesc1,
                            (term | ks) != 0);
      }
#pragma unroll
      for (int term = 0; term < 3; ++term) {
        const uint64_t a2 = term == 1 ? dA2l : dA2h, b2 = term == 2 ? dB2l : dB2h;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc_mma_tf32_elect(tmem + 64u, a2 + (uint64_t)(ks * ((2 * kPLbo) >> 4)), b2 + (uint64_t)(ks * ((2 * kPLbo) >> 4)), kPIdesc2,
                            !(first && term == 0 && ks == 0));
      }
      tc_commit_elect(bar);
    }
    first = false;
    if (tile + gridDim.x < ntiles) load_tile(tile + gridDim.x);  // in flight during the MMAs and the epilogue
    mbar_wait(bar, parity);
    parity ^= 1u;
    tc_fence_after();
    // dgrad epilogue: with M = 64 accumulator row r lives in TMEM lane 32 (r / 16) + r % 16: warp w reads the lane
    // quarter w % 4 (rows 16 (w % 4) + lane, lanes 0..15) and the 16 columns (w / 4) * 16 ..
    {
      float v[16];
      tc_ld16(tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 16), v);
      tc_wait_ld();
      const long long gp = tile * kPTile + 16 * (warp & 3) + lane;
      if (lane < 16 && gp < npix) {
        int wo, ho;
        long long b;
        split_pixel(gp, Wo, Ho, &b, &ho, &wo);
        if (ho >= 1 && ho <= H && wo >= 1 && wo <= W) {  // interior pixels only: the border's input is the zero padding
          float4* dst = reinterpret_cast<float4*>(d_io + ((b * H + ho - 1) * W + wo - 1) * kPC + (warp >> 2) * 16);
#pragma unroll
          for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
      }
    }
    tc_fence_before();
    // the next tile's staging overwrites the operand images (their MMAs have retired) and, before its MMAs are
    // issued, passes the barrier above, which also orders this epilogue's TMEM reads before those MMAs
  }
  // wgrad + bias gradient: rows co (lane layout as above), columns ci 0..63 and the ones column 64
  if (!first && warp < 4) {
    float v[16];
    const int co = 16 * warp + lane;
#pragma unroll 1
    for (int cb = 0; cb < 4; ++cb) {
      tc_ld16(tmem + ((uint32_t)(warp * 32) << 16) + 64u + (uint32_t)(cb * 16), v);
      tc_wait_ld();
      if (lane < 16)
#pragma unroll
        for (int i = 0; i < 16; ++i) atomicAdd(&g_w[co * kPC + cb * 16 + i], v[i]);
    }
    float v8[8];
    tc_ld8(tmem + ((uint32_t)(warp * 32) << 16) + 64u + 64u, v8);
    tc_wait_ld();
    if (lane < 16) atomicAdd(&g_b[co], v8[0]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256));
}

// ---- forward twin: pointwise 1x1 (padding = 1) + bias + ReLU of one block on the tensor core -------------------------------
//   out[px][co] = ReLU(b[co] + sum_ci D[px][ci] W[ci][co])   over the padded (H+2) x (W+2) grid; D = 0 on the border, so a
//   border pixel comes out as ReLU(b) (the padding = 1 quirk, models.py:104-106) with no special case.
// D is the depthwise output dw_fwd_rows_kernel has just written (it is saved for the backward pass anyway).  Tile = 128
// consecutive pixels of the padded grid = one M = 128, N = 64, K = 64 GEMM; both operands are K-major as they lie in memory
// (NHWC rows, [ci][co] weights read transposed once), fp32-grade through the 3-term tf32 split (24 MMAs per tile).  110 KB
// of shared memory and 64 TMEM columns per CTA: two CTAs per SM overlap one's loads and drain with the other's MMAs.
constexpr int kFTile = 128;
constexpr int kFOpA = (kFTile / 8) * kPSbo;  // one 128-row A image (hi or lo): 36 864 B
constexpr size_t kFSmem = 2 * (size_t)kFOpA + 2 * (size_t)kPOp + 64 * 4;
constexpr uint32_t kFIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((128u >> 4) << 24);

__global__ void __launch_bounds__(kPThreads, 2)
pw_fwd_tc_kernel(const float* __restrict__ d, float* __restrict__ out, const float* __restrict__ w_kmaj /*[ci][co]*/,
                 const float* __restrict__ bias, long long B, int H, int W) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const uint32_t bar = smem_u32(&s_bar);
  const int Ho = H + 2, Wo = W + 2;
  const long long npix = B * Ho * (long long)Wo, ntiles = (npix + kFTile - 1) / kFTile;
  unsigned char* sAh = smem;
  unsigned char* sAl = smem + kFOpA;
  unsigned char* sBh = smem + 2 * kFOpA;
  unsigned char* sBl = sBh + kPOp;
  float* sBias = reinterpret_cast<float*>(sBl + kPOp);
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(64));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (int i = tid; i < kPC * kPC; i += kPThreads) {  // B rows = co, K = ci
    const int ci = i >> 6, co = i & 63;
    float hi, lo;
    split_tf32(__ldg(w_kmaj + i), hi, lo);
    *reinterpret_cast<float*>(sBh + kmaj_off(co, ci)) = hi;
    *reinterpret_cast<float*>(sBl + kmaj_off(co, ci)) = lo;
  }
  if (tid < kPC) sBias[tid] = __ldg(bias + tid);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const uint64_t dAh = umma_desc(smem_u32(sAh), kPLbo, kPSbo), dAl = umma_desc(smem_u32(sAl), kPLbo, kPSbo);
  const uint64_t dBh = umma_desc(smem_u32(sBh), kPLbo, kPSbo), dBl = umma_desc(smem_u32(sBl), kPLbo, kPSbo);
  // staging items of this thread: channel quad q of pixels px0 + 32 it; the next tile's loads are in flight while this
  // tile multiplies and drains
  const int q = tid & 15, px0 = tid >> 4;
  float4 dv[4];
  auto load_tile = [&](long long tile) {
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const long long gp = tile * kFTile + px0 + 32 * it;
      dv[it] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gp < npix) {
        int wo, ho;
        long long b;
        split_pixel(gp, Wo, Ho, &b, &ho, &wo);
        if (ho >= 1 && ho <= H && wo >= 1 && wo <= W)
          dv[it] = __ldg(reinterpret_cast<const float4*>(d + ((b * H + ho - 1) * W + wo - 1) * kPC) + q);
      }
    }
  };
  auto prefetch_tile = [&](long long tile) {  // the interior pixels of a tile are (nearly) one contiguous range of d
    if (tid == 0 && tile < ntiles) {
      int wo, ho;
      long long b;
      split_pixel(tile * kFTile, Wo, Ho, &b, &ho, &wo);
      const int hi = min(max(ho - 1, 0), H - 1), wi = min(max(wo - 1, 0), W - 1);
      const long long d0 = (b * H + hi) * W + wi, dn = min((long long)kFTile, B * H * (long long)W - d0);
      l2_prefetch(d + d0 * kPC, (uint32_t)(dn * kPC * 4));
    }
  };
  if ((long long)blockIdx.x < ntiles) load_tile(blockIdx.x);
  prefetch_tile((long long)blockIdx.x + gridDim.x);
  uint32_t parity = 0;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    prefetch_tile(tile + 2LL * gridDim.x);
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      float h4[4], l4[4];
      split_tf32(dv[it].x, h4[0], l4[0]); split_tf32(dv[it].y, h4[1], l4[1]);
      split_tf32(dv[it].z, h4[2], l4[2]); split_tf32(dv[it].w, h4[3], l4[3]);
      const uint32_t o = kmaj_off(px0 + 32 * it, 4 * q);
      *reinterpret_cast<float4*>(sAh + o) = make_float4(h4[0], h4[1], h4[2], h4[3]);
      *reinterpret_cast<float4*>(sAl + o) = make_float4(l4[0], l4[1], l4[2], l4[3]);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {  // converged warp, elected lane issues: Ahi.Bhi + Alo.Bhi + Ahi.Blo
      tc_fence_after();
#pragma unroll
      for (int term = 0; term < 3; ++term) {
        const uint64_t a = term == 1 ? dAl : dAh, b = term == 2 ? dBl : dBh;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc_mma_tf32_elect(tmem, a + (uint64_t)(ks * ((2 * kPLbo) >> 4)), b + (uint64_t)(ks * ((2 * kPLbo) >> 4)), kFIdesc,
                            (term | ks) != 0);
      }
      tc_commit_elect(bar);
    }
    if (tile + gridDim.x < ntiles) load_tile(tile + gridDim.x);
    mbar_wait(bar, parity);
    parity ^= 1u;
    tc_fence_after();
    {  // drain: TMEM lane quarter warp % 4, channels 16 (warp / 4) ..; 16x256b loads + one lane exchange (see model_gemm.cu)
      const int lq = warp & 3, tq = lane & 3, tr = lane >> 2, c0 = (warp >> 2) * 16;
      const bool odd = lane & 1;
      const int coff = odd ? 8 + 2 * (tq - 1) : 2 * tq;
      float v[2][8];
      const uint32_t ta = tmem + ((uint32_t)(lq * 32) << 16) + (uint32_t)c0;
      tc_ld_16x256b_x2(ta, v[0]);
      tc_ld_16x256b_x2(ta + (16u << 16), v[1]);
      tc_wait_ld();
      const float2 bz0 = *reinterpret_cast<const float2*>(sBias + c0 + 2 * tq), bz1 = *reinterpret_cast<const float2*>(sBias + c0 + 8 + 2 * tq);
#pragma unroll
      for (int hh = 0; hh < 2; ++hh)
#pragma unroll
        for (int ss = 0; ss < 2; ++ss) {
          float* x = v[hh] + 2 * ss;  // registers [4 blk + 2 s + e]
          x[0] = fmaxf(x[0] + bz0.x, 0.f);
          x[1] = fmaxf(x[1] + bz0.y, 0.f);
          x[4] = fmaxf(x[4] + bz1.x, 0.f);
          x[5] = fmaxf(x[5] + bz1.y, 0.f);
          const float k0 = odd ? x[4] : x[0], k1 = odd ? x[5] : x[1], s0 = odd ? x[0] : x[4], s1 = odd ? x[1] : x[5];
          const float g0 = __shfl_xor_sync(0xffffffffu, s0, 1), g1 = __shfl_xor_sync(0xffffffffu, s1, 1);
          const long long gp = tile * kFTile + lq * 32 + hh * 16 + ss * 8 + tr;
          if (gp < npix)
            *reinterpret_cast<float4*>(out + gp * kPC + c0 + coff) = odd ? make_float4(g0, g1, k0, k1) : make_float4(k0, k1, g0, g1);
        }
    }
    tc_fence_before();
    // the next tile's staging overwrites the A images (their MMAs have retired); the barrier above orders this drain's
    // TMEM reads before the next tile's MMAs
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(64));
}

// 1 = launched, < 0 = error
int launch_pw_fwd_tc(kws_handle_t h, const float* d, float* out, const float* w_kmaj, const float* bias, long long B, int H,
                     int W, cudaStream_t st) {
  if (int rc = ensure_dyn_smem(pw_fwd_tc_kernel, kFSmem)) return rc;
  const long long ntiles = (B * (long long)(H + 2) * (W + 2) + kFTile - 1) / kFTile;
  const long long g = ntiles < 2LL * h->sm_count ? ntiles : 2LL * h->sm_count;
  pw_fwd_tc_kernel<<<(unsigned)g, kPThreads, kFSmem, st>>>(d, out, w_kmaj, bias, B, H, W);
  KWS_CHECK_LAUNCH(h);
  return 1;
}

// 1 = launched, < 0 = error
int launch_pw_bwd_tc(kws_handle_t h, const float* out, const float* dout, const float* dpooled, float inv_area, float* d_io,
                     const float* w_flat, float* g_w, float* g_b, long long B, int H, int W, cudaStream_t st) {
  if (int rc = ensure_dyn_smem(pw_bwd_tc_kernel, kPSmem)) return rc;
  const long long ntiles = (B * (long long)(H + 2) * (W + 2) + kPTile - 1) / kPTile;
  const long long g = ntiles < h->sm_count ? ntiles : h->sm_count;
  pw_bwd_tc_kernel<<<(unsigned)g, kPThreads, kPSmem, st>>>(out, dout, dpooled, inv_area, d_io, w_flat, g_w, g_b, B, H, W);
  KWS_CHECK_LAUNCH(h);
  return 1;
}

}  // namespace kws
