// Fused PCM -> MFCC / log-mel kernel for sm_100a.
//
// Replaces the per-clip call chain of the reference,
//   AudioProcessor.extract_features (kws/libs/audio_processor.py:235-278)
//     -> python_speech_features.mfcc -> fbank -> sigproc.{preemphasis,framesig,powspec},
//        get_filterbanks, log, scipy dct(type=2, norm='ortho'), lifter, c0 := log(energy)
//  with one launch over the whole batch.  Nothing but the PCM read and
// the feature write touches HBM.
//
// Work decomposition
//   tile  = kFramesPerTile (20) consecutive frames of one clip; a persistent CTA of 5 warps
//           walks tiles with stride gridDim.x.  The tile's samples are pre-emphasised while they
//           are staged into shared memory (fp32, two roundings, exactly psf's float32 step).
//   warp  = 4 frames: each half-warp owns TWO frames, carried in the two halves of packed
//           f32x2 registers so that every butterfly is one FADD2/FFMA2 (Blackwell packed fp32).
//   frame = 512-point real FFT done as a 256-point complex FFT of (even, odd) samples:
//           16 lanes x 16 points, radix-16 in registers, one transpose through shared memory,
//           radix-16 again, then the real-FFT split with the mirror bin fetched by shuffle.
//   The power spectrum goes to shared memory once; the mel bank is applied per SEGMENT
//   between two bin edges (each bin read once, weights (rising, falling)), then log, mean
//   removal, DCT-II with the lifter folded into the matrix, and c0 := log(energy).
#include <cuda_runtime.h>

#include "common.cuh"

namespace kws {

typedef float2 V;  // .x = frame A, .y = frame B of the half-warp

__device__ __forceinline__ V vsplat(float s) { return make_float2(s, s); }
__device__ __forceinline__ V vadd(V a, V b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ V vsub(V a, V b) { return __ffma2_rn(b, vsplat(-1.f), a); }
__device__ __forceinline__ V vmul(V a, V b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ V vfma(V a, V b, V c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ V vneg(V a) { return make_float2(-a.x, -a.y); }

struct C {
  V re, im;
};
__device__ __forceinline__ C cadd(C a, C b) { return C{vadd(a.re, b.re), vadd(a.im, b.im)}; }
__device__ __forceinline__ C csub(C a, C b) { return C{vsub(a.re, b.re), vsub(a.im, b.im)}; }

// forward 4-point DFT in place (W4 = -i)
__device__ __forceinline__ void radix4(C& a0, C& a1, C& a2, C& a3) {
  C t0 = cadd(a0, a2), t1 = csub(a0, a2), t2 = cadd(a1, a3), t3 = csub(a1, a3);
  a0 = cadd(t0, t2);
  a2 = csub(t0, t2);
  a1 = C{vadd(t1.re, t3.im), vsub(t1.im, t3.re)};  // t1 - i t3
  a3 = C{vsub(t1.re, t3.im), vadd(t1.im, t3.re)};  // t1 + i t3
}

// x * (wr + i wi) with compile-time constants
__device__ __forceinline__ C cmulc(C x, float wr, float wi) {
  C r;
  r.re = vfma(x.im, vsplat(-wi), vmul(x.re, vsplat(wr)));
  r.im = vfma(x.im, vsplat(wr), vmul(x.re, vsplat(wi)));
  return r;
}

// 16-point forward DFT in registers.  Input x[n] natural order; output X[k] is left at
// x[4*(k&3) + (k>>2)]  (use P16(k)).
#define P16(k) (4 * ((k)&3) + ((k) >> 2))
__device__ __forceinline__ void fft16(C (&x)[16]) {
  const float c1 = 0.92387953251128674f, s1 = 0.38268343236508977f, hh = 0.70710678118654752f;
#pragma unroll
  for (int i = 0; i < 4; ++i) radix4(x[i], x[i + 4], x[i + 8], x[i + 12]);
  // now x[i + 4q] = A_i[q]; twiddle by W16^(i q)
  x[1 + 4] = cmulc(x[1 + 4], c1, -s1);                                   // W^1
  {                                                                      // W^2 = h(1 - i)
    C t = x[1 + 8];
    x[1 + 8] = C{vmul(vadd(t.re, t.im), vsplat(hh)), vmul(vsub(t.im, t.re), vsplat(hh))};
  }
  x[1 + 12] = cmulc(x[1 + 12], s1, -c1);                                 // W^3
  {                                                                      // W^2
    C t = x[2 + 4];
    x[2 + 4] = C{vmul(vadd(t.re, t.im), vsplat(hh)), vmul(vsub(t.im, t.re), vsplat(hh))};
  }
  {                                                                      // W^4 = -i
    C t = x[2 + 8];
    x[2 + 8] = C{t.im, vneg(t.re)};
  }
  {                                                                      // W^6 = -h(1 + i)
    C t = x[2 + 12];
    x[2 + 12] = C{vmul(vsub(t.im, t.re), vsplat(hh)), vmul(vadd(t.re, t.im), vsplat(-hh))};
  }
  x[3 + 4] = cmulc(x[3 + 4], s1, -c1);                                   // W^3
  {                                                                      // W^6
    C t = x[3 + 8];
    x[3 + 8] = C{vmul(vsub(t.im, t.re), vsplat(hh)), vmul(vadd(t.re, t.im), vsplat(-hh))};
  }
  x[3 + 12] = cmulc(x[3 + 12], -c1, s1);                                 // W^9
#pragma unroll
  for (int q = 0; q < 4; ++q) radix4(x[4 * q], x[4 * q + 1], x[4 * q + 2], x[4 * q + 3]);
}

struct MfccParams {
  const void* pcm;
  float* out;
  long long B, N, stride, total_tiles;
  int nframes, tiles_per_clip, load_len, hop, tile_len;
  int M, ncep, ncep_pad, mode, append_energy;
  float preemph;
  const float* window;
  const float2* tw256;
  const float2* tw512;
  const float2* melw;
  const int* edges;
  const float* dct;
  int off_scratch, off_melw, off_edges, off_dct, off_tw512;  // byte offsets in dynamic smem
};

constexpr int kWarps = kFramesPerTile / 4;   // 5
constexpr int kThreads = kWarps * 32;        // 160
constexpr int kScratchHalf = 16 * 17 * 16;   // bytes of transpose scratch per half-warp
constexpr float kEps = 2.220446049250313e-16f;  // numpy.finfo(float).eps, exact in fp32 (2^-52)

template <bool I16>
__device__ __forceinline__ float load_sample(const void* base, long long i) {
  if (I16) return (float)((const short*)base)[i] * (1.0f / 32768.0f);
  return ((const float*)base)[i];
}

template <bool I16>
__global__ void __launch_bounds__(kThreads, 3) mfcc512_kernel(const MfccParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  float* tile = reinterpret_cast<float*>(smem);
  float2* s_melw = reinterpret_cast<float2*>(smem + p.off_melw);
  int* s_edges = reinterpret_cast<int*>(smem + p.off_edges);
  float* s_dct = reinterpret_cast<float*>(smem + p.off_dct);
  float2* s_tw512 = reinterpret_cast<float2*>(smem + p.off_tw512);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, half = lane >> 4, l = lane & 15;
  unsigned char* scratch = smem + p.off_scratch + (warp * 2 + half) * kScratchHalf;
  float4* s_tr = reinterpret_cast<float4*>(scratch);                 // [16][17] complex pairs
  V* s_P = reinterpret_cast<V*>(scratch);                            // [257] (aliases s_tr)
  V* s_R = reinterpret_cast<V*>(scratch + 2064);                     // [M+1] rising sums
  V* s_F = s_R + (kMaxFilt + 1);                                     // [M+1] falling sums
  V* s_x = s_F + (kMaxFilt + 1);                                     // [M] centred log-mel

  for (int i = tid; i < kBins; i += kThreads) s_melw[i] = p.melw[i];
  for (int i = tid; i < p.M + 2; i += kThreads) s_edges[i] = p.edges[i];
  for (int i = tid; i < p.M * p.ncep_pad; i += kThreads) s_dct[i] = p.dct[i];
  for (int i = tid; i < 256; i += kThreads) s_tw512[i] = p.tw512[i];
  float2 tw[16];  // W256^(l*k): twiddles between the two radix-16 passes, fixed per lane
#pragma unroll
  for (int k = 0; k < 16; ++k) tw[k] = p.tw256[(l * k) & 255];

  const bool hop_even = (p.hop & 1) == 0;
  const bool vec_ok = !I16 && (p.N % 4 == 0) && (p.stride % 4 == 0) &&
                      ((reinterpret_cast<size_t>(p.pcm) & 15) == 0);
  const float pre = p.preemph;

  for (long long tile_id = blockIdx.x; tile_id < p.total_tiles; tile_id += gridDim.x) {
    const long long clip = tile_id / p.tiles_per_clip;
    const int t = (int)(tile_id - clip * p.tiles_per_clip);
    const long long s0 = (long long)t * kFramesPerTile * p.hop;
    __syncthreads();  // everyone is done with the previous tile buffer (and tables are in)

    // ---- stage: global -> pre-emphasis -> shared (psf sigproc.preemphasis in float32) ----
    {
      const char* cbase = (const char*)p.pcm + (size_t)clip * p.stride * (I16 ? 2 : 4);
      for (int i4 = tid * 4; i4 < p.tile_len; i4 += kThreads * 4) {
        const long long g = s0 + i4;
        float4 y;
        if (vec_ok && g + 3 < p.N) {
          const float* x = (const float*)cbase;
          float4 v = __ldg(reinterpret_cast<const float4*>(x + g));
          float xm = (g > 0) ? __ldg(x + g - 1) : 0.f;
          y.x = (g > 0) ? __fsub_rn(v.x, __fmul_rn(pre, xm)) : v.x;
          y.y = __fsub_rn(v.y, __fmul_rn(pre, v.x));
          y.z = __fsub_rn(v.z, __fmul_rn(pre, v.y));
          y.w = __fsub_rn(v.w, __fmul_rn(pre, v.z));
        } else {
          float r[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const long long ge = g + e;
            float val = 0.f;
            if (ge < p.N) {
              float xc = load_sample<I16>(cbase, ge);
              val = (ge > 0) ? __fsub_rn(xc, __fmul_rn(pre, load_sample<I16>(cbase, ge - 1))) : xc;
            }
            r[e] = val;
          }
          y = make_float4(r[0], r[1], r[2], r[3]);
        }
        *reinterpret_cast<float4*>(tile + i4) = y;
      }
    }
    __syncthreads();

    const int fw = t * kFramesPerTile + warp * 4;  // first frame of this warp's item
    if (fw >= p.nframes) continue;                 // warp-uniform
    const int fa_local = warp * 4 + half * 2;
    const int fA = t * kFramesPerTile + fa_local, fB = fA + 1;
    const float* pa = tile + fa_local * p.hop;
    const float* pb = pa + p.hop;

    // ---- pass 1: lane l takes z[16 m + l] = (s[32m + 2l], s[32m + 2l + 1]), m = 0..15 ----
    C x[16];
#pragma unroll
    for (int m = 0; m < 16; ++m) {
      const int idx = 32 * m + 2 * l;
      float2 a = make_float2(0.f, 0.f), b = make_float2(0.f, 0.f);
      if (idx + 1 < p.load_len) {
        if (hop_even) {
          a = *reinterpret_cast<const float2*>(pa + idx);
          b = *reinterpret_cast<const float2*>(pb + idx);
        } else {
          a = make_float2(pa[idx], pa[idx + 1]);
          b = make_float2(pb[idx], pb[idx + 1]);
        }
      } else if (idx < p.load_len) {
        a.x = pa[idx];
        b.x = pb[idx];
      }
      if (p.window != nullptr && idx < p.load_len) {
        float w0 = __ldg(p.window + idx), w1 = (idx + 1 < p.load_len) ? __ldg(p.window + idx + 1) : 0.f;
        a.x *= w0; b.x *= w0; a.y *= w1; b.y *= w1;
      }
      x[m].re = make_float2(a.x, b.x);
      x[m].im = make_float2(a.y, b.y);
    }
    fft16(x);
    // twiddle by W256^(l k1) and transpose through shared memory
#pragma unroll
    for (int k1 = 0; k1 < 16; ++k1) {
      C y = x[P16(k1)];
      if (k1 > 0) {
        const float wr = tw[k1].x, wi = tw[k1].y;
        C z;
        z.re = make_float2(fmaf(-y.im.x, wi, y.re.x * wr), fmaf(-y.im.y, wi, y.re.y * wr));
        z.im = make_float2(fmaf(y.im.x, wr, y.re.x * wi), fmaf(y.im.y, wr, y.re.y * wi));
        y = z;
      }
      s_tr[k1 * 17 + l] = make_float4(y.re.x, y.re.y, y.im.x, y.im.y);
    }
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float4 v = s_tr[l * 17 + j];
      x[j].re = make_float2(v.x, v.y);
      x[j].im = make_float2(v.z, v.w);
    }
    __syncwarp();
    // ---- pass 2: lane l = k1 produces Z[l + 16 k2] at x[P16(k2)] ----
    fft16(x);

    // ---- real-FFT split + power spectrum:  2X[k] = (Zk + conj Zp) - i W512^k (Zk - conj Zp) ----
    const int partner = (lane & 16) | ((16 - l) & 15);
    V esum = vsplat(0.f);
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) {
      const C zk = x[P16(k2)];
      const C own = x[P16((16 - k2) & 15)];
      const C oth = x[P16(15 - k2)];
      C snd = (l == 0) ? own : oth;
      C zp;
      zp.re.x = __shfl_sync(0xffffffffu, snd.re.x, partner);
      zp.re.y = __shfl_sync(0xffffffffu, snd.re.y, partner);
      zp.im.x = __shfl_sync(0xffffffffu, snd.im.x, partner);
      zp.im.y = __shfl_sync(0xffffffffu, snd.im.y, partner);
      const float2 w = s_tw512[l + 16 * k2];
      const V sr = vadd(zk.re, zp.re), si = vsub(zk.im, zp.im);
      const V dr = vsub(zk.re, zp.re), di = vadd(zk.im, zp.im);
      // T = -i W D : Tr = di wr + dr wi ; Ti = di wi - dr wr
      const V xr = vfma(di, vsplat(w.x), vfma(dr, vsplat(w.y), sr));
      const V xi = vfma(di, vsplat(w.y), vfma(dr, vsplat(-w.x), si));
      const V pw = vmul(vfma(xi, xi, vmul(xr, xr)), vsplat(1.0f / 2048.0f));
      s_P[l + 16 * k2] = pw;
      esum = vadd(esum, pw);
      if (k2 == 0 && l == 0) {  // Nyquist bin: X[256] = Re Z0 - Im Z0
        const V ny = vsub(zk.re, zk.im);
        const V pn = vmul(vmul(ny, ny), vsplat(1.0f / 512.0f));
        s_P[256] = pn;
        esum = vadd(esum, pn);
      }
    }
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) {
      esum.x += __shfl_xor_sync(0xffffffffu, esum.x, o);
      esum.y += __shfl_xor_sync(0xffffffffu, esum.y, o);
    }
    __syncwarp();

    // ---- mel bank by segment: bin k in [e_s, e_{s+1}) feeds filter s (rising) and s-1 (falling) ----
    for (int s = l; s <= p.M; s += 16) {
      V r = vsplat(0.f), f = vsplat(0.f);
      const int k1 = s_edges[s + 1];
      for (int k = s_edges[s]; k < k1; ++k) {
        const V pw = s_P[k];
        const float2 w = s_melw[k];
        r = vfma(pw, vsplat(w.x), r);
        f = vfma(pw, vsplat(w.y), f);
      }
      s_R[s] = r;
      s_F[s] = f;
    }
    __syncwarp();
    V lm[(kMaxFilt + 15) / 16];
    V msum = vsplat(0.f);
#pragma unroll
    for (int r = 0; r < (kMaxFilt + 15) / 16; ++r) {
      const int j = l + 16 * r;
      lm[r] = vsplat(0.f);
      if (j < p.M) {
        V m = vadd(s_R[j], s_F[j + 1]);
        m.x = (m.x == 0.f) ? kEps : m.x;
        m.y = (m.y == 0.f) ? kEps : m.y;
        lm[r] = make_float2(logf(m.x), logf(m.y));
        msum = vadd(msum, lm[r]);
      }
    }
    if (p.mode == kModeLogMel) {
#pragma unroll
      for (int r = 0; r < (kMaxFilt + 15) / 16; ++r) {
        const int j = l + 16 * r;
        if (j < p.M) {
          if (fA < p.nframes) p.out[((size_t)clip * p.nframes + fA) * p.M + j] = lm[r].x;
          if (fB < p.nframes) p.out[((size_t)clip * p.nframes + fB) * p.M + j] = lm[r].y;
        }
      }
      continue;
    }
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) {
      msum.x += __shfl_xor_sync(0xffffffffu, msum.x, o);
      msum.y += __shfl_xor_sync(0xffffffffu, msum.y, o);
    }
    const V mean = vmul(msum, vsplat(1.0f / (float)p.M));
#pragma unroll
    for (int r = 0; r < (kMaxFilt + 15) / 16; ++r) {
      const int j = l + 16 * r;
      if (j < p.M) s_x[j] = vsub(lm[r], mean);
    }
    __syncwarp();

    // ---- DCT-II (ortho, lifter folded in) on the centred log-mel vector ----
    V acc[(kMaxFilt + 15) / 16];
#pragma unroll
    for (int r = 0; r < (kMaxFilt + 15) / 16; ++r) acc[r] = vsplat(0.f);
    for (int n = 0; n < p.M; ++n) {
      const V xv = s_x[n];
      const float* drow = s_dct + n * p.ncep_pad;
#pragma unroll
      for (int r = 0; r < (kMaxFilt + 15) / 16; ++r) {
        const int c = l + 16 * r;
        if (c < p.ncep) acc[r] = vfma(xv, vsplat(drow[c]), acc[r]);
      }
    }
    float eA = (esum.x == 0.f) ? kEps : esum.x, eB = (esum.y == 0.f) ? kEps : esum.y;
#pragma unroll
    for (int r = 0; r < (kMaxFilt + 15) / 16; ++r) {
      const int c = l + 16 * r;
      if (c < p.ncep) {
        V v = acc[r];
        if (c == 0) {
          if (p.append_energy) {
            v = make_float2(logf(eA), logf(eB));
          } else {  // the removed mean only affects c0: + sqrt(M) * mean (lifter[0] = 1)
            v = vfma(mean, vsplat(sqrtf((float)p.M)), v);
          }
        }
        if (fA < p.nframes) p.out[((size_t)clip * p.nframes + fA) * p.ncep + c] = v.x;
        if (fB < p.nframes) p.out[((size_t)clip * p.nframes + fB) * p.ncep + c] = v.y;
      }
    }
  }
}

int launch_mfcc(kws_handle_t h, const void* pcm_dev, bool is_i16, int64_t B, int64_t N,
                int64_t pcm_stride, float* out_dev, int mode, cudaStream_t stream) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  const FeatureTables& t = h->feat;
  if (!t.ready) return fail(KWS_ERR_STATE, "features not configured: call kws_features_configure first");
  // rows may overlap (1 <= pcm_stride < N): the sliding windows of one long recording are read in place
  if (B < 0 || N < 1 || pcm_stride < 1) return fail(KWS_ERR_INVALID, "bad B/N/stride");
  if (B == 0) return KWS_OK;
  if (!pcm_dev || !out_dev) return fail(KWS_ERR_INVALID, "null device pointer");

  {
    const int rc = launch_mfcc_fast(h, pcm_dev, is_i16, B, N, pcm_stride, out_dev, mode, stream);
    if (rc != 0) return rc < 0 ? rc : KWS_OK;
  }
  MfccParams p{};
  p.pcm = pcm_dev;
  p.out = out_dev;
  p.B = B;
  p.N = N;
  p.stride = pcm_stride;
  p.nframes = (int)kws_num_frames(N, t.cfg.frame_len, t.cfg.frame_step);
  p.tiles_per_clip = (p.nframes + kFramesPerTile - 1) / kFramesPerTile;
  p.total_tiles = (long long)B * p.tiles_per_clip;
  p.load_len = t.load_len;
  p.hop = t.cfg.frame_step;
  p.tile_len = ((kFramesPerTile - 1) * p.hop + p.load_len + 3) & ~3;
  p.M = t.cfg.nfilt;
  p.ncep = t.ncep;
  p.ncep_pad = t.ncep_pad;
  p.mode = mode;
  p.append_energy = t.cfg.append_energy;
  p.preemph = t.cfg.preemph;
  p.window = t.d_window;
  p.tw256 = t.d_tw256;
  p.tw512 = t.d_tw512;
  p.melw = t.d_melw;
  p.edges = t.d_segedge;
  p.dct = t.d_dct;

  size_t off = (size_t)p.tile_len * 4;
  off = (off + 15) & ~(size_t)15;
  p.off_scratch = (int)off;
  off += (size_t)kWarps * 2 * kScratchHalf;
  p.off_melw = (int)off;
  off += kBins * 8 + 8;
  p.off_edges = (int)off;
  off += ((size_t)(p.M + 2) * 4 + 15) & ~(size_t)15;
  p.off_dct = (int)off;
  off += ((size_t)p.M * p.ncep_pad * 4 + 15) & ~(size_t)15;
  p.off_tw512 = (int)off;
  off += 256 * 8;
  const size_t smem_bytes = off;
  if (smem_bytes > 227 * 1024)
    return fail(KWS_ERR_INVALID, "frame_step=%d needs %zu bytes of shared memory per CTA (> 227 KB)",
                p.hop, smem_bytes);
  static_assert(2064 + (2 * (kMaxFilt + 1) + kMaxFilt) * 8 <= kScratchHalf, "scratch layout overflow");

  auto kern = is_i16 ? mfcc512_kernel<true> : mfcc512_kernel<false>;
  if (int rc = ensure_dyn_smem(kern, smem_bytes)) return rc;
  int occ = 0;
  KWS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, smem_bytes));
  if (occ < 1) return fail(KWS_ERR_CUDA, "mfcc kernel does not fit on an SM");
  long long grid = (long long)h->sm_count * occ;
  if (grid > p.total_tiles) grid = p.total_tiles;
  kern<<<(unsigned)grid, kThreads, smem_bytes, stream>>>(p);
  KWS_CHECK_LAUNCH(h);
  return KWS_OK;
}

}  // namespace kws
