// Fast variant of the fused PCM -> MFCC kernel (see mfcc.cu for the algorithm and the
// reference call chain it replaces: kws/libs/audio_processor.py:235-278 -> psf.mfcc).
//
// Same decomposition as mfcc.cu (tile = 20 frames, warp = 4 frames, half-warp = 2 frames packed
// in f32x2 registers, 256-point complex FFT = 16 lanes x radix-16 twice).  The kernel is bound
// by instruction issue and shared-memory bandwidth, not HBM (DESIGN.md "MFCC kernel"), so this
// variant removes instructions and exposed latency:
//   * the raw PCM of the NEXT tile is fetched by one TMA bulk copy (cp.async.bulk + mbarrier)
//     while the current tile is being transformed: no thread waits on a global load;
//   * a conversion pass pre-emphasises the raw tile (fp32, two roundings = psf's float32 step)
//     and stores it INTERLEAVED, S2[i] = (y[i], y[i+hop]), so one LDS.128 yields the packed
//     (re, im) pair of both frames with no register shuffling;
//   * FADD2/FFMA2 operand modifiers (negate, scalar broadcast) replace explicit negations and
//     splats, so twiddles and weights stay scalar;
//   * the real-FFT split computes X[k] and X[256-k] together (they share Zk +- conj Zp): each
//     lane handles k2 = 0..7 and fetches 8 mirror values by shuffle instead of 16;
//   * the DCT-II (+lifter) of the whole tile, [20 x M] x [M x C], runs on the tensor core:
//     tcgen05.mma kind::tf32 with a 3-term hi/lo split (error ~2^-21, i.e. fp32-grade),
//     operands in the canonical K-major no-swizzle layout, accumulator in TMEM, read back with
//     tcgen05.ld by one warp one tile later (the epilogue overlaps the next tile's FFTs);
//   * log is a range-reduced MUFU.LG2.
// Configurations it does not cover (window function, frame_len != 400, nfilt > 48, unaligned
// input) fall back to the generic kernel in mfcc.cu.
#include <cuda_runtime.h>

#include "common.cuh"

namespace kws {

typedef float2 V;
__device__ __forceinline__ V fsplat(float s) { return make_float2(s, s); }
__device__ __forceinline__ V fneg(V a) { return make_float2(-a.x, -a.y); }  // folds into an operand modifier
__device__ __forceinline__ V fadd(V a, V b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ V fsub(V a, V b) { return __fadd2_rn(a, fneg(b)); }
__device__ __forceinline__ V fmul(V a, V b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ V ffma(V a, V b, V c) { return __ffma2_rn(a, b, c); }

struct CF {
  V re, im;
};
__device__ __forceinline__ CF cfadd(CF a, CF b) { return CF{fadd(a.re, b.re), fadd(a.im, b.im)}; }
__device__ __forceinline__ CF cfsub(CF a, CF b) { return CF{fsub(a.re, b.re), fsub(a.im, b.im)}; }

__device__ __forceinline__ void r4(CF& a0, CF& a1, CF& a2, CF& a3) {
  CF t0 = cfadd(a0, a2), t1 = cfsub(a0, a2), t2 = cfadd(a1, a3), t3 = cfsub(a1, a3);
  a0 = cfadd(t0, t2);
  a2 = cfsub(t0, t2);
  a1 = CF{fadd(t1.re, t3.im), fsub(t1.im, t3.re)};
  a3 = CF{fsub(t1.re, t3.im), fadd(t1.im, t3.re)};
}
__device__ __forceinline__ CF cmc(CF x, float wr, float wi) {
  CF r;
  r.re = ffma(x.im, fsplat(-wi), fmul(x.re, fsplat(wr)));
  r.im = ffma(x.im, fsplat(wr), fmul(x.re, fsplat(wi)));
  return r;
}
#define Q16(k) (4 * ((k)&3) + ((k) >> 2))
// 16-point forward DFT; X[k] is left at x[Q16(k)]
__device__ __forceinline__ void dft16(CF (&x)[16]) {
  const float c1 = 0.92387953251128674f, s1 = 0.38268343236508977f, hh = 0.70710678118654752f;
#pragma unroll
  for (int i = 0; i < 4; ++i) r4(x[i], x[i + 4], x[i + 8], x[i + 12]);
  x[5] = cmc(x[5], c1, -s1);
  { CF t = x[9];  x[9]  = CF{fmul(fadd(t.re, t.im), fsplat(hh)), fmul(fsub(t.im, t.re), fsplat(hh))}; }
  x[13] = cmc(x[13], s1, -c1);
  { CF t = x[6];  x[6]  = CF{fmul(fadd(t.re, t.im), fsplat(hh)), fmul(fsub(t.im, t.re), fsplat(hh))}; }
  { CF t = x[10]; x[10] = CF{t.im, fneg(t.re)}; }
  { CF t = x[14]; x[14] = CF{fmul(fsub(t.im, t.re), fsplat(hh)), fmul(fadd(t.re, t.im), fsplat(-hh))}; }
  x[7] = cmc(x[7], s1, -c1);
  { CF t = x[11]; x[11] = CF{fmul(fsub(t.im, t.re), fsplat(hh)), fmul(fadd(t.re, t.im), fsplat(-hh))}; }
  x[15] = cmc(x[15], -c1, s1);
#pragma unroll
  for (int q = 0; q < 4; ++q) r4(x[4 * q], x[4 * q + 1], x[4 * q + 2], x[4 * q + 3]);
}

// natural log of a positive finite float: exponent/mantissa split + MUFU.LG2 on [2/3, 4/3)
__device__ __forceinline__ float fast_log(float x) {
  float bias = 0.f;
  if (x < 1.17549435e-38f) { x *= 16777216.f; bias = -24.f; }
  const int bits = __float_as_int(x);
  const int e = (bits - 0x3f2aaaab) >> 23;
  const float m = __int_as_float(bits - (e << 23));
  const float t = __log2f(m);
  const float ef = (float)e + bias;
  // ln2 split: hi has 12 significant bits so ef * hi is exact
  return fmaf(ef, 0.693115234375f, fmaf(ef, 3.194618e-5f, t * 0.69314718055994531f));
}

// ---- PTX: mbarrier, TMA bulk copy, tcgen05 ------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// bounded wait: a protocol bug traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (int spin = 0; spin < (1 << 24); ++spin) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) return;
  }
  __trap();
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
// K-major, no-swizzle UMMA shared-memory descriptor (cute::UMMA::SmemDescriptor): start>>4 in
// bits [0,14), leading (K-chunk) byte offset>>4 in [16,30), stride (8-row group) byte offset>>4
// in [32,46), version 1 in [46,48), layout_type 0 (interleave) in [61,64)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3ffff) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         (1ull << 46);
}

constexpr int kDctN = 48;        // accumulator columns (coefficients padded to a multiple of 16)
constexpr int kTmemCols = 64;    // allocation: power of two >= 32
// instruction descriptor: D=f32 (1<<4), A=B=tf32 (2<<7, 2<<10), K-major both, N>>3 at 17, M>>4 at 24
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kDctN >> 3) << 17) | ((128u >> 4) << 24);

struct FastParams {
  const float* pcm;
  float* out;
  long long B, N, stride, total_tiles;
  int nframes, tiles_per_clip, hop, M, ncep, mode, append_energy;
  float preemph;
  const float2* tw256;
  const float2* twp;    // [128] W512^k
  const float2* melw;   // [257] (rising, falling) weight of the bin's segment
  const int* edges;     // [M+2]
  const float* dctB;    // B operand image: hi then lo, canonical K-major layout [48 x Kpad] each
  int kpad, s2_len, tile_len;
  int off_s2, off_a, off_scratch, off_b, off_melw, off_edges, off_twp, off_c0;
};

constexpr int kFWarps = kFramesPerTile / 4;
constexpr int kFThreads = kFWarps * 32;
constexpr int kFScratchHalf = 16 * 17 * 16;
constexpr float kFEps = 2.220446049250313e-16f;

template <int LOAD_LEN, int ROUNDS>
__global__ void __launch_bounds__(kFThreads, 2) mfcc512_fast_kernel(const FastParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar[2];  // [0] TMA tile landed, [1] tile MMAs retired
  __shared__ uint32_t s_tmem;
  float* raw = reinterpret_cast<float*>(smem);          // raw[j] = x[s0 - 4 + j]
  float2* s2 = reinterpret_cast<float2*>(smem + p.off_s2);
  const float2* s_melw = reinterpret_cast<const float2*>(smem + p.off_melw);
  const int* s_edges = reinterpret_cast<const int*>(smem + p.off_edges);
  const float2* s_twp = reinterpret_cast<const float2*>(smem + p.off_twp);
  float* s_c0 = reinterpret_cast<float*>(smem + p.off_c0);  // [2][20]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, half = lane >> 4, l = lane & 15;
  unsigned char* scratch = smem + p.off_scratch + (warp * 2 + half) * kFScratchHalf;
  float4* s_tr = reinterpret_cast<float4*>(scratch);
  V* s_P = reinterpret_cast<V*>(scratch);
  V* s_R = reinterpret_cast<V*>(scratch + 2064);
  V* s_F = s_R + (16 * ROUNDS + 1);

  const bool do_dct = (p.mode == kModeMfcc);
  const uint32_t bar_tma = smem_u32(&s_bar[0]), bar_mma = smem_u32(&s_bar[1]);
  const int hop = p.hop;
  const uint32_t sbo = (uint32_t)(p.kpad / 4) * 128u;   // bytes between 8-row groups of A and B

  if (tid == 0) {
    mbar_init(bar_tma, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0 && do_dct) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)),
                 "n"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  {
    float2* d = reinterpret_cast<float2*>(smem + p.off_melw);
    for (int i = tid; i < kBins; i += kFThreads) d[i] = p.melw[i];
    int* e = reinterpret_cast<int*>(smem + p.off_edges);
    for (int i = tid; i < p.M + 2; i += kFThreads) e[i] = p.edges[i];
    float2* t = reinterpret_cast<float2*>(smem + p.off_twp);
    for (int i = tid; i < 128; i += kFThreads) t[i] = p.twp[i];
    float* b = reinterpret_cast<float*>(smem + p.off_b);
    for (int i = tid; i < 2 * kDctN * p.kpad; i += kFThreads) b[i] = p.dctB[i];
    float* a = reinterpret_cast<float*>(smem + p.off_a);
    for (int i = tid; i < 2 * 24 * p.kpad; i += kFThreads) a[i] = 0.f;
  }
  float twr[16], twi[16];  // W256^(l k1); FFMA2 takes a scalar register as a broadcast operand
#pragma unroll
  for (int k = 1; k < 16; ++k) {
    const float2 w = p.tw256[(l * k) & 255];
    twr[k] = w.x;
    twi[k] = w.y;
  }
  fence_proxy_async();  // B (and the zeroed A) were written through the generic proxy
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = do_dct ? s_tmem : 0u;
  const float pre = p.preemph;

  auto issue_tma = [&](long long tile_id) {  // one thread
    const long long clip = tile_id / p.tiles_per_clip;
    const int t = (int)(tile_id - clip * p.tiles_per_clip);
    const long long s0 = (long long)t * kFramesPerTile * hop;
    const long long first = s0 >= 4 ? s0 - 4 : 0;
    long long end = s0 + p.tile_len;
    if (end > p.N) end = p.N;
    const uint32_t bytes = (uint32_t)(end - first) * 4u;
    mbar_expect_tx(bar_tma, bytes);
    tma_load_1d(smem_u32(raw + (s0 >= 4 ? 0 : 4)), p.pcm + clip * p.stride + first, bytes, bar_tma);
  };
  auto issue_mma = [&]() {  // one thread: D[128 x 48] = Ahi*Bhi + Alo*Bhi + Ahi*Blo over Kpad
    tc_fence_after();
    const uint32_t a_hi = smem_u32(smem + p.off_a), a_lo = a_hi + 3u * sbo;
    const uint32_t b_hi = smem_u32(smem + p.off_b), b_lo = b_hi + (uint32_t)(kDctN / 8) * sbo;
    const int ksteps = p.kpad / 8;
    uint32_t acc = 0;
#pragma unroll 1
    for (int term = 0; term < 3; ++term) {
      const uint32_t ab = term == 1 ? a_lo : a_hi, bb = term == 2 ? b_lo : b_hi;
      for (int ks = 0; ks < ksteps; ++ks) {
        tc_mma_tf32(tmem, umma_desc(ab + ks * 256u, 128u, sbo), umma_desc(bb + ks * 256u, 128u, sbo), kIdesc, acc);
        acc = 1;
      }
    }
    tc_commit(bar_mma);
  };
  // epilogue of a finished tile (warp 0, lane = frame row): TMEM -> registers -> global
  auto epilogue = [&](long long clip, int t, int buf) {
    float v[kDctN];
    tc_fence_after();
    {
      float q[16];
#pragma unroll
      for (int c = 0; c < kDctN / 16; ++c) {
        tc_ld16(tmem + c * 16, q);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[c * 16 + i] = q[i];
      }
    }
    tc_wait_ld();
    tc_fence_before();
    const int f = t * kFramesPerTile + lane;
    if (lane < kFramesPerTile && f < p.nframes) {
      const float c0 = s_c0[buf * kFramesPerTile + lane];
      v[0] = p.append_energy ? c0 : v[0] + c0;
      float* o = p.out + ((size_t)clip * p.nframes + f) * p.ncep;
      if ((p.ncep & 3) == 0) {
#pragma unroll
        for (int c = 0; c < kDctN; c += 4)
          if (c < p.ncep) *reinterpret_cast<float4*>(o + c) = make_float4(v[c], v[c + 1], v[c + 2], v[c + 3]);
      } else {
#pragma unroll
        for (int c = 0; c < kDctN; ++c)
          if (c < p.ncep) o[c] = v[c];
      }
    }
  };

  if (tid == 0 && (long long)blockIdx.x < p.total_tiles) issue_tma(blockIdx.x);

  int it = 0;
  long long prev_clip = 0;
  int prev_t = 0;
  for (long long tile_id = blockIdx.x; tile_id < p.total_tiles; tile_id += gridDim.x, ++it) {
    const long long clip = tile_id / p.tiles_per_clip;
    const int t = (int)(tile_id - clip * p.tiles_per_clip);
    const long long s0 = (long long)t * kFramesPerTile * hop;
    __syncthreads();  // previous tile: S2 no longer read, its A rows written
    if (tid == 0 && it > 0 && do_dct) issue_mma();
    mbar_wait(bar_tma, it & 1);
    // ---- conversion: S2[i] = (y[i], y[i+hop]), y = pre-emphasised signal ----
    for (int i = tid * 2; i < p.s2_len; i += kFThreads * 2) {
      const long long g = s0 + i;
      const float* r = raw + i + 4;
      const float2 va = *reinterpret_cast<const float2*>(r);
      const float2 vb = *reinterpret_cast<const float2*>(r + hop);
      const float am = r[-1], bm = r[hop - 1];
      float4 o;
      o.x = (g > 0) ? __fsub_rn(va.x, __fmul_rn(pre, am)) : va.x;  // y[g]
      o.y = __fsub_rn(vb.x, __fmul_rn(pre, bm));                    // y[g+hop]
      o.z = __fsub_rn(va.y, __fmul_rn(pre, va.x));                  // y[g+1]
      o.w = __fsub_rn(vb.y, __fmul_rn(pre, vb.x));                  // y[g+hop+1]
      if (g >= p.N) o.x = 0.f;        // beyond the clip the raw buffer holds stale data
      if (g + 1 >= p.N) o.z = 0.f;
      if (g + hop >= p.N) o.y = 0.f;
      if (g + hop + 1 >= p.N) o.w = 0.f;
      *reinterpret_cast<float4*>(s2 + i) = o;
    }
    __syncthreads();  // S2 complete, raw consumed
    if (tid == 0 && tile_id + gridDim.x < p.total_tiles) {
      fence_proxy_async();
      issue_tma(tile_id + gridDim.x);
    }

    const int fa_local = warp * 4 + half * 2;
    const int fA = t * kFramesPerTile + fa_local, fB = fA + 1;
    const bool active = (t * kFramesPerTile + warp * 4) < p.nframes;  // warp-uniform
    V lm[ROUNDS];
    V mean = fsplat(0.f), esum = fsplat(0.f);
    if (active) {
      const float2* pa = s2 + fa_local * hop;
      CF x16[16];
#pragma unroll
      for (int m = 0; m < 16; ++m) {
        const int idx = 32 * m + 2 * l;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (32 * m + 31 < LOAD_LEN) {
          v = *reinterpret_cast<const float4*>(pa + idx);
        } else if (32 * m < LOAD_LEN) {
          if (idx + 1 < LOAD_LEN) v = *reinterpret_cast<const float4*>(pa + idx);
          else if (idx < LOAD_LEN) { const float2 u = pa[idx]; v.x = u.x; v.y = u.y; }
        }
        x16[m].re = make_float2(v.x, v.y);
        x16[m].im = make_float2(v.z, v.w);
      }
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
      const V sc = fsplat(1.0f / 2048.0f);
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
        const V pk = fmul(ffma(ai, ai, fmul(ar, ar)), sc);
        const V pm = fmul(ffma(bi, bi, fmul(br, br)), sc);
        s_P[l + 16 * k2] = pk;
        s_P[256 - l - 16 * k2] = pm;
        esum = fadd(esum, fadd(pk, pm));
      }
      if (l0) {  // bin 128 is its own mirror: |X[128]|^2 = |Z[128]|^2
        const CF z8 = x16[Q16(8)];
        const V p8 = fmul(ffma(z8.im, z8.im, fmul(z8.re, z8.re)), fsplat(1.0f / 512.0f));
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
      for (int s = l; s <= p.M; s += 16) {
        V r = fsplat(0.f), f = fsplat(0.f);
        const int k1 = s_edges[s + 1];
        for (int k = s_edges[s]; k < k1; ++k) {
          const V pw = s_P[k];
          const float2 w = s_melw[k];
          r = ffma(pw, fsplat(w.x), r);
          f = ffma(pw, fsplat(w.y), f);
        }
        s_R[s] = r;
        s_F[s] = f;
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
            if (fA < p.nframes) p.out[((size_t)clip * p.nframes + fA) * p.M + j] = lm[r].x;
            if (fB < p.nframes) p.out[((size_t)clip * p.nframes + fB) * p.M + j] = lm[r].y;
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
      // the previous tile's MMAs must have retired before its A rows are overwritten and before
      // its accumulator is read
      if (it > 0) mbar_wait(bar_mma, (it - 1) & 1);
      if (warp == 0 && it > 0) epilogue(prev_clip, prev_t, (it - 1) & 1);
      if (active) {
        // A rows fa_local (x) and fa_local+1 (y): centred log-mel split into tf32 hi + lo
        unsigned char* abase = smem + p.off_a;
#pragma unroll
        for (int r = 0; r < ROUNDS; ++r) {
          const int n = l + 16 * r;
          if (n < p.M) {
            const V xc = fsub(lm[r], mean);
            const float hx = to_tf32(xc.x), hy = to_tf32(xc.y);
            const float lx = to_tf32(xc.x - hx), ly = to_tf32(xc.y - hy);
            const uint32_t col = (uint32_t)(n >> 2) * 128u + (uint32_t)(n & 3) * 4u;
            const uint32_t oa = (uint32_t)(fa_local >> 3) * sbo + (uint32_t)(fa_local & 7) * 16u + col;
            const uint32_t ob = (uint32_t)((fa_local + 1) >> 3) * sbo + (uint32_t)((fa_local + 1) & 7) * 16u + col;
            *reinterpret_cast<float*>(abase + oa) = hx;
            *reinterpret_cast<float*>(abase + ob) = hy;
            *reinterpret_cast<float*>(abase + 3u * sbo + oa) = lx;
            *reinterpret_cast<float*>(abase + 3u * sbo + ob) = ly;
          }
        }
        if (l == 0) {
          const float eA = (esum.x == 0.f) ? kFEps : esum.x, eB = (esum.y == 0.f) ? kFEps : esum.y;
          const float sq = sqrtf((float)p.M);
          float* c0 = s_c0 + (it & 1) * kFramesPerTile + fa_local;
          c0[0] = p.append_energy ? fast_log(eA) : sq * mean.x;  // else: the removed mean only moves c0
          c0[1] = p.append_energy ? fast_log(eB) : sq * mean.y;
        }
      }
      fence_proxy_async();  // generic-proxy writes of A must be visible to the tensor core
    }
    prev_clip = clip;
    prev_t = t;
  }
  __syncthreads();
  if (do_dct) {
    if (it > 0) {
      if (tid == 0) issue_mma();
      if (warp == 0) {
        mbar_wait(bar_mma, (it - 1) & 1);
        epilogue(prev_clip, prev_t, (it - 1) & 1);
      }
    }
    __syncthreads();
    if (warp == 0)
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols));
  }
}

// returns 1 when the fast kernel handled the launch, 0 when the configuration needs the generic
// kernel, negative on error
int launch_mfcc_fast(kws_handle_t h, const float* pcm_dev, int64_t B, int64_t N, int64_t pcm_stride,
                     float* out_dev, int mode, cudaStream_t stream) {
  const FeatureTables& t = h->feat;
  const int M = t.cfg.nfilt, hop = t.cfg.frame_step;
  if (t.d_window != nullptr || t.load_len != 400 || M > 48 || t.d_dctB == nullptr) return 0;
  // TMA bulk copies need 16-byte aligned sources and sizes; the interleaved tile needs an even hop
  if ((hop & 3) != 0 || (N & 3) != 0 || (pcm_stride & 3) != 0 || (reinterpret_cast<size_t>(pcm_dev) & 15) != 0)
    return 0;
  constexpr int ROUNDS = 3;
  FastParams p{};
  p.pcm = pcm_dev;
  p.out = out_dev;
  p.B = B;
  p.N = N;
  p.stride = pcm_stride;
  p.nframes = (int)kws_num_frames(N, t.cfg.frame_len, t.cfg.frame_step);
  p.tiles_per_clip = (p.nframes + kFramesPerTile - 1) / kFramesPerTile;
  p.total_tiles = (long long)B * p.tiles_per_clip;
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
  p.tile_len = (kFramesPerTile - 1) * hop + 400;           // samples a tile's 20 frames span
  p.s2_len = (kFramesPerTile - 2) * hop + 400;             // S2[i] read for i < 18*hop + 400
  const uint32_t sbo = (uint32_t)(p.kpad / 4) * 128u;
  size_t off = ((size_t)(p.tile_len + 4) * 4 + 127) & ~(size_t)127;
  p.off_s2 = (int)off;
  off += ((size_t)p.s2_len * 8 + 127) & ~(size_t)127;
  p.off_a = (int)off;                                      // A hi rows 0..23, then A lo rows 0..23
  off += (size_t)6 * sbo;
  p.off_scratch = (int)off;
  off += (size_t)kFWarps * 2 * kFScratchHalf;
  p.off_b = (int)off;                                      // B hi [48 x kpad], then B lo
  off += (size_t)2 * (kDctN / 8) * sbo;
  p.off_melw = (int)off;
  off += (size_t)kBins * 8 + 8;
  p.off_edges = (int)off;
  off += ((size_t)(M + 2) * 4 + 15) & ~(size_t)15;
  p.off_twp = (int)off;
  off += 128 * 8;
  p.off_c0 = (int)off;
  off += 2 * kFramesPerTile * 4;
  const size_t smem_bytes = off;
  static_assert(2064 + 2 * (16 * ROUNDS + 1) * 8 <= kFScratchHalf, "scratch layout overflow");
  // the tensor core reads 128 rows of A: 16 row groups past each base must stay inside the CTA's
  // shared memory (rows >= 24 are don't-care values that only reach accumulator rows we never read)
  if ((size_t)p.off_a + (size_t)(3 + 16) * sbo > smem_bytes || smem_bytes > 113 * 1024) return 0;

  auto kern = mfcc512_fast_kernel<400, ROUNDS>;
  KWS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  // two CTAs per SM by construction (<= 113 KB shared memory and <= 204 registers each); ask for
  // the full shared-memory carve-out so the second CTA is not refused
  KWS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  long long grid = (long long)h->sm_count * 2;
  if (grid > p.total_tiles) grid = p.total_tiles;
  kern<<<(unsigned)grid, kFThreads, smem_bytes, stream>>>(p);
  KWS_CHECK_LAUNCH(h);
  return 1;
}

}  // namespace kws
