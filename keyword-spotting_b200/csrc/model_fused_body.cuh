// Fused forward of the reference KWS network (kws/libs/models.py:122-183) for channels = 64,
// four blocks: features [B, T, F] -> logits [B, num_classes] in ONE kernel, every activation
// on-chip (tensor memory + registers + shared memory), HBM sees only the features and the logits.
//
// Formulation.  Every GEMM-able layer (conv1 10x10 s2 p2 as an im2col GEMM with K = 100, the
// four pointwise 1x1 convs with K = 64) is computed TRANSPOSED on the tensor core:
//     D[c_out, pixel] = W[c_out, k] . X[pixel, k]^T      tcgen05.mma kind::tf32, M = 64, N = one
// output row of pixels (padded to 24, or 16 for F = 10), accumulator in TMEM.  With M = 64 the
// accumulator row c_out lives in TMEM lane 32*(c/16) + c%16 and a second, independent M = 64 tile
// sits at lane offset 16 (tools/umma_probe.cu pins both facts on the device), so a CTA works on a
// PAIR of clips ("slots") and thread (warp quarter q, lane) of every warp owns
//     slot = lane / 16,  channel = 16 q + lane % 16
// -- the depthwise 3x3 is per channel, so the thread that reads a channel's accumulator row from
// TMEM (tcgen05.ld 32x32b: one lane per thread) applies bias + ReLU and runs the 3x3 stencil on
// rows it holds in registers, with no inter-thread exchange at all, then stores the depthwise
// result as the K-major B operand of the next pointwise GEMM (one 4-byte store per pixel, the
// 144-byte K-chunk pitch puts the 16 channels of a quarter on 16 different banks).
//
// Schedule.  The network is streamed row by row through all layers (a skewed wavefront, see the
// stage table above fused_mma_warp): a step has two phases, each = depthwise stencils of two
// blocks by the twelve compute warps, one CTA barrier, and the GEMMs of the rows just finished,
// issued by the two MMA warps (one per clip slot) and retired while the other phase computes.
// TMEM holds, per layer, a ring of the last two raw accumulator rows; the last layer's row is
// pooled as soon as it retires.
//
// Bias, ReLU and the padding=1 quirk.  The pointwise padding=1 quirk (models.py:104-106: each block
// grows H and W by 2 and the new border pixels are ReLU(bias)) and every bias are folded INTO the
// GEMMs, so the depthwise stage reads a finished padded layer row from TMEM and its activation is
// a bare max(raw, 0) -- no bias add, no border patching, no column masks in the CUDA-core code:
//   * each pointwise GEMM has K = 64 + 8: B-tile K-slots 64..67 hold the indicators
//     (interior, border, interior, border) of the output pixel and the A image holds
//     (folded bias hi, border bias hi, folded bias lo, border bias lo) there (hi + lo = the fp32
//     bias as two tf32 terms; "folded" = pointwise bias + W_pw . depthwise bias, on the host);
//   * the depthwise result of pixel x is stored as B-tile pixel row x + 1; rows 0 and W + 1 carry
//     only the border indicator, rows past the layer carry nothing, so the GEMM row IS the padded
//     layer row: [ReLU-able border bias | W interior pixels | border bias | zeros];
//   * conv1 uses its four spare K slots (100..103) the same way; the margin column left of a ring
//     row is never written and stays zero = the depthwise conv's own zero padding;
//   * the constant top/bottom border rows and the all-zero row that flushes the stencil are
//     constant TMEM rows written once at kernel start; a stage just picks the row address.
//
// Precision: tf32 operands (cvt.rna), fp32 accumulate; logits agree with the fp32 reference to
// ~3e-6 absolute / 2e-4 relative (tests/test_gpu_model.py; north_star tolerance 1e-3).
#pragma once
#include <cstdlib>
#include <cstring>
#include <vector>

#include "mfcc_ws_body.cuh"

namespace kws {

constexpr int kFC = 64;                      // channels of the reference network
constexpr int kLboB = 144;                   // K-chunk pitch of the B tiles (bank spreading)
constexpr int kKC1 = 104;                    // conv1 K: 100 taps + (bias hi, 0, bias lo, 0)
constexpr int kKCp = 72;                     // pointwise K: 64 channels + 4 bias/indicator slots + 4 zero slots
constexpr int kSboA1 = (kKC1 / 4) * 128;     // 8-row group pitch: conv1 weights
// pointwise weights: 17 stored K-chunks; the 18th (K slots 68..71) aliases the next group's first
// chunk -- finite numbers that meet the B tile's real zeros there
constexpr int kSboAp = (kKCp / 4 - 1) * 128;
constexpr int kSboB1 = (kKC1 / 4) * kLboB;   //                    conv1 im2col tile
constexpr int kSboBp = (kKCp / 4) * kLboB;   //                    depthwise-output tile
constexpr int kAImgBytes = 8 * kSboA1 + 4 * 8 * kSboAp;

struct FusedParams {
  const float* feats;     // [B, T, F]
  float* logits;          // [B, ncls]
  long long B;
  int T, H0, ncls;
  const float* a_img;     // tf32 A-operand images: conv1 then 4 pointwise, K-major no-swizzle, biases in the K tail
  const float* dw_w;      // [4][9][64]
  const float* border;    // [5][64] row 0 unused; ReLU(pointwise bias) of block l
  const float* fc_w;      // [64][ncls]
  const float* fc_b;      // [ncls]
};

// CTA-wide barrier of the model role.  Stand-alone kernel: all kFThreads threads of the block.  CHUNK: the block
// has one more (idle, exited) warp than the role uses, so the role's 14 warps meet at named barrier 1.
template <bool CHUNK>
__device__ __forceinline__ void fused_sync();

// One step of the depthwise 3x3 (zero padding 1) of a block for this thread's (slot, channel) on
// its warp's 8-pixel column segment [8j, 8j+8): consume one INPUT row of the block's input layer
// and finish the OUTPUT row above it.  The stencil is accumulated per input row: with v = the
// activated row,
//     out[x]  = P[x] + sum_dx w[2][dx] v[x+dx]     (output row i-1 complete -> B tile of the pointwise GEMM)
//     P'[x]   = R[x] + sum_dx w[1][dx] v[x+dx]     (output row i, two of three rows in)
//     R'[x]   =        sum_dx w[0][dx] v[x+dx]     (output row i+1, one row in)
// so every input row is read from TMEM and activated exactly once and nothing is ever moved.
// `taddr` is the TMEM address of layer column 8j - 1 of the input row (a ring row, a constant border
// row or the zero row -- the caller picks); layer column = TMEM column for every layer, the bias is
// already in the accumulator, so v = max(raw, 0).  The stencil runs on packed f32x2 pairs of
// adjacent pixels: E[p] = (v[2p], v[2p+1]) and the odd-aligned O[p] = (v[2p+1], v[2p+2]) are both
// produced straight by the max, so no register pair is ever re-assembled.
// W = width of the layer (= number of valid output pixels), SHIFT = pixel offset of the result in the
// B tile (1 where the following pointwise output carries a border column).  `par` points at this
// channel's [w0..w8] column.
// Clips per slot: with PACK = 2 (narrow feature maps, F = 10) two clips sit side by side in one GEMM row, the
// second at column kPackOff; the indicator scheme leaves the columns between them zero, which is each clip's
// own zero padding, so the stencil needs no change -- only "which pixels are real" does.
constexpr int kPackOff = 12;
template <int W, int PACK>
__host__ __device__ constexpr bool px_valid(int px) {
  return PACK == 1 ? px < W : (px < W || (px >= kPackOff && px < kPackOff + W));
}

// Shared memory through explicit 32-bit shared-window addresses: with generic pointers ptxas rebuilt the shared
// base (S2R CgaCtaId + LEA) and the per-thread offsets (S2R TID + IMADs) in front of every stencil call -- ~28 of the
// ~110 instructions of a call (ncu r2a, SASS view).  keep_u32() makes a value opaque to the optimiser so that it is
// kept in its register instead of being rematerialised from the thread index.
__device__ __forceinline__ uint32_t keep_u32(uint32_t v) {
  asm volatile("" : "+r"(v));
  return v;
}
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t a) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ float lds32(uint32_t a) {
  float v;
  asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) {  // a = base + constant: ptxas folds it into the STS offset
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}

__device__ __forceinline__ V relu2x(V a) { return __fadd2_rn(a, make_float2(fabsf(a.x), fabsf(a.y))); }

template <int W, int SHIFT, int PACK>
__device__ __forceinline__ void dw_seg(uint32_t taddr, int j, bool st, uint32_t par, uint32_t dst, V (&P)[4], V (&R)[4]) {
  const float4 w0 = lds128(par), w1 = lds128(par + 16u);
  const float w8 = lds32(par + 32u);
  // E pairs from a load at column 8j-1, O pairs from a second load one column to the right; the max is
  // done in place on the load's registers so that each pair stays the aligned register pair it arrived as
  float ra[16], rb[8];
  tc_ld16(taddr, ra);
  tc_ld8(taddr + 1u, rb);
  tc_wait_ld();
  // ReLU of a PAIR in one packed instruction: x + |x| = 2 max(x, 0) exactly (FADD2 takes |.| as an operand modifier);
  // the factor 2 is taken out of the depthwise taps, which are stored halved (a power of two: same bits as
  // max(x, 0) * w).  Nine FADD2 on the FMA pipe instead of eighteen FMNMX on the ALU pipe per call.
  V E[5], O[4];
#pragma unroll
  for (int pp = 0; pp < 5; ++pp) E[pp] = relu2x(make_float2(ra[2 * pp], ra[2 * pp + 1]));
#pragma unroll
  for (int pp = 0; pp < 4; ++pp) O[pp] = relu2x(make_float2(rb[2 * pp], rb[2 * pp + 1]));
  // warp-uniform store predicates.  PACK = 1: pixels x < KL of the last valid segment JL, all pixels of the
  // segments before it; PACK = 2: per segment, from the compile-time validity of its eight columns
  constexpr int JL = (W - 1) / 8, KL = W - 8 * JL;
  const bool st_lo = st && j <= JL, st_hi = st && j < JL;
  const bool s0 = st && j == 0, s1 = st && j == 1, s2 = st && j == 2;
#pragma unroll
  for (int pp = 0; pp < 4; ++pp) {  // output pixels 2pp, 2pp+1 of the segment, packed f32x2
    const V o = ffma(fsplat(w8), E[pp + 1], ffma(fsplat(w1.w), O[pp], ffma(fsplat(w1.z), E[pp], P[pp])));
    P[pp] = ffma(fsplat(w1.y), E[pp + 1], ffma(fsplat(w1.x), O[pp], ffma(fsplat(w0.w), E[pp], R[pp])));
    R[pp] = ffma(fsplat(w0.z), E[pp + 1], ffma(fsplat(w0.y), O[pp], fmul(fsplat(w0.x), E[pp])));
    // tf32 round-to-nearest (ties away) = add half an ulp of the 10-bit mantissa; the MMA drops the low 13 bits
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int x = 2 * pp + h;
      // pixels past the layer are never stored: their B-tile rows stay zero (or hold the border indicator)
      const bool on = PACK == 1 ? (x < KL ? st_lo : st_hi)
                                : ((px_valid<W, PACK>(x) && s0) || (px_valid<W, PACK>(8 + x) && s1) || (px_valid<W, PACK>(16 + x) && s2));
      if (on) sts32(dst + (uint32_t)(((x + SHIFT) >> 3) * kSboBp + ((x + SHIFT) & 7) * 16), __float_as_uint(h ? o.y : o.x) + 0x1000u);
    }
  }
}

constexpr int kFCWarps = 12;                       // compute warps: 3 column segments x 4 lane quarters
constexpr int kCThreads = kFCWarps * 32;
constexpr int kFThreads = (kFCWarps + 2) * 32;     // + one MMA-issue warp per clip slot
template <>
__device__ __forceinline__ void fused_sync<false>() { __syncthreads(); }
// Phase barriers of the row loop (ids 5 = end of phase A, 6 = end of phase B).  Only the MMA-issue warps WAIT for a
// phase to be complete (they may not issue a GEMM before every stencil has stored its B-tile rows); the compute warps
// just ARRIVE and walk on into the next phase, whose first action is the mbarrier wait for the GEMMs that phase
// depends on.  That wait bounds the skew between compute warps to less than one step (the GEMMs of step t + 1 are
// issued only after every warp has arrived at the same barrier of step t), which is what the one-step delays of the
// feature refill and of pool_fc below are sized for.
template <int ID>
__device__ __forceinline__ void phase_arrive() { asm volatile("bar.arrive %0, %1;" ::"n"(ID), "n"(kFThreads) : "memory"); }
template <int ID>
__device__ __forceinline__ void phase_sync() { asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(kFThreads) : "memory"); }
template <>
__device__ __forceinline__ void fused_sync<true>() { asm volatile("bar.sync 1, %0;" ::"n"(kFThreads) : "memory"); }
constexpr int kRingCols = 64;                     // per layer: 2 row slots x (8 margin + 24) columns
// last pointwise layer's row (pooled at once): 8-column margin + 24, like a ring slot.  The margin is load-bearing:
// the right halo of ring 3's second slot (layer column 24 of the widest layer) falls on it and must read zero.
constexpr int kAcc4Col = 4 * kRingCols;
constexpr int kConstCol = kAcc4Col + 32;          // constant border rows of the three padded layers, 32 columns each
constexpr int kZeroCol = kConstCol + 3 * 32;      // the all-zero row that flushes a stencil

template <int F, int NPIX, int PACK>
struct FusedLayout {
  static constexpr int FP = F + 4;  // feature row pitch: 2 zero columns left, >= 2 right (conv1 padding 2)
  static constexpr int CL = 2 * PACK;  // clips per CTA step: 2 slots x PACK side by side
  static constexpr int b1_tile = (NPIX / 8) * kSboB1, b1_stride = b1_tile + 64;
  // slot stride of the depthwise-output tiles.  (A stride of 16 banks mod 32 makes the two clip slots of a warp's store
  // conflict-free -- measured in round 2: shared-memory conflicts halve, run time unchanged -- but costs 256 bytes
  // that the stage-schedule table needs.)
  static constexpr int bp_tile = (NPIX / 8) * kSboBp, bp_stride = bp_tile + 64;
  static constexpr int off_a1 = 0;
  static constexpr int off_ap = 8 * kSboA1;
  static constexpr int off_b1 = kAImgBytes;
  static constexpr int off_bp = off_b1 + 2 * b1_stride;
  static constexpr int off_par = off_bp + 8 * bp_stride;          // [4][64][12] depthwise taps (9 used)
  static constexpr int off_pool = off_par + 4 * kFC * 12 * 4;     // [3][2][PACK][64] pooling partials
  static constexpr int off_feat = off_pool + 3 * CL * kFC * 4;
  // zero-padded feature map of one clip: rows -2 .. T+1
  static constexpr size_t feat_floats(int T) { return (size_t)(T + 4) * FP; }
  // bytes of dynamic shared memory for T frames: ... + feature maps + the per-step stage schedule (uint2 per local step)
  static constexpr size_t smem_bytes(int T) {
    return (((size_t)off_feat + CL * feat_floats(T) * 4 + 7) & ~(size_t)7) + (size_t)((T + 4 - 10) / 2 + 1 + 7) * 8;
  }
  static constexpr size_t off_sched(int T) { return ((size_t)off_feat + CL * feat_floats(T) * 4 + 7) & ~(size_t)7; }
};

// The MMA warps run their issue code CONVERGED and every tcgen05 instruction is predicated on an elected lane
// (see tc_mma_tf32_elect in mfcc_dev.cuh): issued from inside `if (lane == 0)`, MMA issue was the critical
// path of a step.  ACC is an immediate here.
template <bool ACC>
__device__ __forceinline__ void tc_mma_tf32_c(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred p, pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "n"(ACC ? 1 : 0)
      : "memory");
}
// one GEMM: NK k-steps of 8; the descriptors' address fields advance by 256 B (A) / 2 * 144 B (B)
template <int NK>
__device__ __forceinline__ void issue_gemm(uint32_t d, uint64_t adesc, uint64_t bdesc, uint32_t idesc) {
  tc_mma_tf32_c<false>(d, adesc, bdesc, idesc);
#pragma unroll
  for (int ks = 1; ks < NK; ++ks)
    tc_mma_tf32_c<true>(d, adesc + (uint64_t)(ks * (256 >> 4)), bdesc + (uint64_t)(ks * ((2 * kLboB) >> 4)), idesc);
}

// Stage schedule (local step t of a clip pair; phase A then phase B inside a step):
//   conv1 im2col + GEMM row t                    (B)      | dw1: input row t-2 (A), pw1 GEMM row t-3 (end of A)
//   dw2: input row t-3 (B), pw2 GEMM row t-4 (end of B)    | dw3: input row t-5 (A), pw3 GEMM row t-6 (end of A)
//   dw4: input row t-6 (B), pw4 GEMM row t-7 (end of B)    | pooling of pw4 row t-8 (B)
// GEMMs issued at the end of a phase run on the tensor core while the compute warps execute the
// next phase, whose stages never touch what those GEMMs read or write.
//
// The clip pairs of a CTA are streamed back to back with period PER = H0 + 7 steps (the longest stage,
// dw4, has exactly PER rows), so the pipeline never drains between pairs: with tl = global step mod PER
// and n = global step / PER, a stage with lag g works on row i = tl - g of pair n, or, while that is
// negative, on row i + PER of pair n - 1.  Every stage resets itself (the zero flush row clears the
// stencil state), so nothing else marks the boundary between two pairs.
struct StageRow {
  int i;      // row index inside the stage's pair (valid when on)
  bool on;
};
__device__ __forceinline__ StageRow stage_row(int tl, int n, int npair, int per, int g, int len) {
  int i = tl - g, cn = n;
  if (i < 0) {
    i += per;
    cn = n - 1;
  }
  return StageRow{i, cn >= 0 && cn < npair && i < len};
}

template <int F, int NPIX, int PACK, int SLOT, bool CHUNK>
__device__ __forceinline__ void fused_mma_warp(unsigned char* smem, uint32_t tmem, uint32_t barO, uint32_t barE, int H0,
                                               int my_pairs) {
  using L = FusedLayout<F, NPIX, PACK>;
  constexpr uint32_t kIdescF = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(NPIX >> 3) << 17) | ((64u >> 4) << 24);
  const uint32_t sbase = smem_u32(smem);
  const uint64_t a1 = umma_desc(sbase + L::off_a1, 128u, kSboA1);
  const uint64_t b1 = umma_desc(sbase + L::off_b1 + SLOT * L::b1_stride, kLboB, kSboB1);
  uint64_t ap[4], bp[4];
#pragma unroll
  for (int l = 0; l < 4; ++l) {
    ap[l] = umma_desc(sbase + L::off_ap + l * 8 * kSboAp, 128u, kSboAp);
    bp[l] = umma_desc(sbase + L::off_bp + (l * 2 + SLOT) * L::bp_stride, kLboB, kSboBp);
  }
  const uint32_t d0 = tmem + ((uint32_t)(SLOT * 16) << 16) + 8u;  // ring rows start after an 8-column margin
  const int per = H0 + 7, total = my_pairs * per + 7;
  fused_sync<CHUNK>();  // features of the first pair loaded
  int tl = 0, n = 0;
  for (int tg = 0; tg < total; ++tg) {
    phase_sync<5>();  // phase A written
    tc_fence_after();
    {
      StageRow r = stage_row(tl, n, my_pairs, per, 3, H0);
      if (r.on) issue_gemm<kKCp / 8>(d0 + 1 * kRingCols + (r.i & 1) * 32, ap[0], bp[0], kIdescF);
      r = stage_row(tl, n, my_pairs, per, 6, H0 + 4);
      if (r.on) issue_gemm<kKCp / 8>(d0 + 3 * kRingCols + (r.i & 1) * 32, ap[2], bp[2], kIdescF);
      tc_commit_elect(barO);
    }
    phase_sync<6>();  // phase B written
    tc_fence_after();
    {
      StageRow r = stage_row(tl, n, my_pairs, per, 0, H0);
      if (r.on) issue_gemm<kKC1 / 8>(d0 + (r.i & 1) * 32, a1, b1, kIdescF);
      r = stage_row(tl, n, my_pairs, per, 4, H0 + 2);
      if (r.on) issue_gemm<kKCp / 8>(d0 + 2 * kRingCols + (r.i & 1) * 32, ap[1], bp[1], kIdescF);
      r = stage_row(tl, n, my_pairs, per, 7, H0 + 6);
      if (r.on) issue_gemm<kKCp / 8>(d0 + kAcc4Col, ap[3], bp[3], kIdescF);
      tc_commit_elect(barE);
    }
    if (++tl == per) {
      tl = 0;
      ++n;
    }
  }
}

// TMEM column (relative to the lane quarter's base) of input row i of a depthwise stage: `ring` is the
// input layer's ring; BORDER layers have HIN + 2 rows whose first and last are the constant row `cst`;
// row HL (one past the end) is the zero row that flushes the last output row.
template <bool BORDER>
__device__ __forceinline__ uint32_t dw_row_col(int i, int HIN, uint32_t ring, uint32_t cst) {
  const int HL = BORDER ? HIN + 2 : HIN;
  if (i >= HL) return (uint32_t)kZeroCol;
  if (BORDER) {
    if (i == 0 || i == HL - 1) return cst;
    return ring + (uint32_t)(((i - 1) & 1) * 32);
  }
  return ring + (uint32_t)((i & 1) * 32);
}

// Stage schedule of the compute warps, one uint2 per local step tl, built once per kernel: which row of which pair
// each stencil stage works on is a function of tl alone (except "does that pair exist", which depends on the pair
// counter n).  12 bits per stencil stage: [8:0] TMEM column of the input row relative to the lane quarter's base
// (ring slot, constant border row or the zero row), [9] the finished output row is stored, [10] the row belongs to the
// PREVIOUS pair (tl < lag), [11] the stage has a row at this step.  Computing this per step cost ~20 instructions per
// stage and step (stage_row + dw_row_col: 16 % of the executed instructions, ncu r2a).
//   .x: dw1 [0:12), dw3 [12:24), pooling [24:28) = (valid, prev, last row, -)      .y: dw2 [0:12), dw4 [12:24)
constexpr uint32_t kSchSt = 1u << 9, kSchPrev = 1u << 10, kSchOn = 1u << 11;
template <bool BORDER>
__device__ __forceinline__ uint32_t sched_entry(int tl, int per, int g, int len, int HIN, uint32_t ring, uint32_t cst) {
  int i = tl - g;
  uint32_t e = 0;
  if (i < 0) {
    i += per;
    e |= kSchPrev;
  }
  if (i < len) e |= kSchOn | dw_row_col<BORDER>(i, HIN, ring, cst) | (i >= 1 ? kSchSt : 0u);
  return e;
}

// The body runs the network on `nclips` clips whose features are feats_base[c] (c = 0 .. nclips-1) and writes
// logits_base[c]; groups bid, bid + nblk, ... of those clips are this CTA's.
//   CHUNK = false: the stand-alone kernel (nclips = p.B, feats_base = p.feats, logits_base = p.logits).
//   CHUNK = true : one phase of the single-launch PCM -> logits kernel (pcm_logits_fused.cu): the features are the
//                  chunk this CTA's MFCC phase has just written to its L2-resident scratch -- read with L2-coherent
//                  loads, the same addresses are rewritten every chunk --, the block has one more warp than the
//                  role uses, tensor memory belongs to the caller (tmem_ext) and the mbarriers are invalidated at
//                  the end.
template <int F, int NPIX, int PACK, bool CHUNK>
__device__ __forceinline__ void dscnn_fused_body(const FusedParams& p, unsigned char* smem, const unsigned bid,
                                                 const unsigned nblk, const long long nclips, const float* feats_base,
                                                 float* logits_base, const uint32_t tmem_ext) {
  constexpr int W0 = (F - 6) / 2 + 1, W3 = W0 + 6, NSEG = NPIX / 8;
  static_assert(W3 <= NPIX && NPIX % 8 == 0 && NSEG <= 3, "row of the widest layer must fit the MMA N");
  static_assert(PACK == 1 || (PACK == 2 && W3 + 2 < kPackOff && kPackOff + W3 <= NPIX), "two clips per slot need a zero gap");
  using L = FusedLayout<F, NPIX, PACK>;
  constexpr int FP = L::FP, CL = L::CL;
  __shared__ __align__(8) unsigned long long s_bar[2];
  __shared__ uint32_t s_tmem;
  if (CHUNK && threadIdx.x >= kFThreads) return;  // the block is sized for the MFCC role: one warp more
  float* s_feat = reinterpret_cast<float*>(smem + L::off_feat);  // [CL][(T + 4) * FP], zero padded, tf32-rounded
  float* s_pool = reinterpret_cast<float*>(smem + L::off_pool);  // [3][2][PACK][64]
  float* s_par = reinterpret_cast<float*>(smem + L::off_par);    // [4][64][12]

  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const uint32_t barO = smem_u32(&s_bar[0]), barE = smem_u32(&s_bar[1]);
  const int T = p.T, H0 = p.H0, TF = T * F;
  const int feat_stride = (int)L::feat_floats(T);
  const long long npairs = (nclips + CL - 1) / CL;  // groups of CL clips ("pairs" when PACK = 1)

  if (tid == 0) {
    mbar_init(barO, 2);
    mbar_init(barE, 2);
    fence_mbar_init();
  }
  if (!CHUNK && warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(p.a_img);
    uint4* dst = reinterpret_cast<uint4*>(smem);
    for (int i = tid; i < kAImgBytes / 16; i += kFThreads) dst[i] = src[i];
    uint4* z = reinterpret_cast<uint4*>(smem + L::off_b1);
    const int nz = (L::off_feat - L::off_b1) / 16 + (CL * feat_stride * 4 + 15) / 16;  // B tiles, pool, feature pads
    for (int i = tid; i < nz; i += kFThreads) z[i] = make_uint4(0, 0, 0, 0);
  }
  fused_sync<CHUNK>();
  for (int i = tid; i < 4 * kFC * 12; i += kFThreads) {
    const int l = i / (12 * kFC), ch = (i / 12) % kFC, k = i % 12;
    s_par[i] = k < 9 ? 0.5f * p.dw_w[(l * 9 + k) * kFC + ch] : 0.f;   // halved: the stencil's ReLU yields 2 max(x, 0)
  }
  uint2* s_sched = reinterpret_cast<uint2*>(smem + L::off_sched(T));
  for (int tl = tid; tl < H0 + 7; tl += kFThreads) {
    const int per = H0 + 7;
    uint2 e;
    e.x = sched_entry<false>(tl, per, 2, H0 + 1, H0, 0 * kRingCols, 0) |
          (sched_entry<true>(tl, per, 5, H0 + 5, H0 + 2, 2 * kRingCols, kConstCol + 32) << 12);
    int ip = tl - 8;
    uint32_t pe = 0;
    if (ip < 0) {
      ip += per;
      pe |= 2u;
    }
    if (ip < H0 + 6) pe |= 1u | (ip == H0 + 5 ? 4u : 0u);
    e.x |= pe << 24;
    e.y = sched_entry<true>(tl, per, 3, H0 + 3, H0, 1 * kRingCols, kConstCol) |
          (sched_entry<true>(tl, per, 6, H0 + 7, H0 + 4, 3 * kRingCols, kConstCol + 64) << 12);
    s_sched[tl] = e;
  }
  // constant K tail of the B tiles: (interior, border, interior, border) indicators of every pixel row
  for (int i = tid; i < 2 * 5 * NPIX; i += kFThreads) {
    const int s = i / (5 * NPIX), l = (i / NPIX) % 5, px = i % NPIX;  // l = 0: conv1 tile, 1..4: tile of pointwise l
    const int pl = (PACK == 2 && px >= kPackOff) ? px - kPackOff : px;  // column inside its clip
    float in = 0.f, bo = 0.f;
    unsigned char* tile;
    int chunk, sbo;
    if (l == 0) {
      tile = smem + L::off_b1 + s * L::b1_stride, chunk = 25, sbo = kSboB1;
      in = pl < W0 ? 1.f : 0.f;
    } else {
      tile = smem + L::off_bp + ((l - 1) * 2 + s) * L::bp_stride, chunk = 16, sbo = kSboBp;
      const int Wl = W0 + 2 * (l - 1);  // width of the depthwise output feeding pointwise l
      if (l < 4) {
        in = (pl >= 1 && pl <= Wl) ? 1.f : 0.f;
        bo = (pl == 0 || pl == Wl + 1) ? 1.f : 0.f;
      } else {
        in = pl < Wl ? 1.f : 0.f;
      }
    }
    *reinterpret_cast<float4*>(tile + (px >> 3) * sbo + chunk * kLboB + (px & 7) * 16) = make_float4(in, bo, in, bo);
  }
  fence_proxy_async();
  tc_fence_before();
  fused_sync<CHUNK>();
  tc_fence_after();
  const uint32_t tmem = CHUNK ? tmem_ext : s_tmem;
  if (warp < 4) {  // zero all of tensor memory once, then the constant border rows of the three padded layers
    for (int col = 0; col < 512; col += 16) {
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(
              tmem + ((uint32_t)(warp * 32) << 16) + col),
          "r"(0)
          : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    const int ch = warp * 16 + (lane & 15);
    for (int l = 1; l < 4; ++l) {
      const uint32_t bv = __float_as_uint(p.border[l * kFC + ch]);
      const int WL = W0 + 2 * l;
      for (int x0 = 0; x0 < NPIX; x0 += 8) {
        uint32_t v[8];
#pragma unroll
        for (int x = 0; x < 8; ++x) v[x] = (x0 + x < WL || (PACK == 2 && x0 + x >= kPackOff && x0 + x < kPackOff + WL)) ? bv : 0u;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(
                         tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(kConstCol + (l - 1) * 32 + 8 + x0)),
                     "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                     : "memory");
      }
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  fused_sync<CHUNK>();
  tc_fence_after();

  // pairs of this CTA: bid, + nblk, ... (the launcher keeps nblk <= npairs)
  const int my_pairs = (int)((npairs - bid + nblk - 1) / nblk);
  if (warp == kFCWarps) {
    fused_mma_warp<F, NPIX, PACK, 0, CHUNK>(smem, tmem, barO, barE, H0, my_pairs);
  } else if (warp == kFCWarps + 1) {
    fused_mma_warp<F, NPIX, PACK, 1, CHUNK>(smem, tmem, barO, barE, H0, my_pairs);
  } else {
    // =========================== compute warps ===========================
    const int j = warp >> 2, q = warp & 3, sl = lane >> 4, c = q * 16 + (lane & 15);
    const bool seg_on = j < NSEG;
    // TMEM: this warp's lane quarter; ring row slot s of layer l starts at column l*64 + s*32, its data
    // after an 8-column margin; a 16-column load from (segment start - 1) covers the segment and its halo
    const uint32_t my_tmem = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t my_ld = keep_u32(my_tmem + 7u + (uint32_t)(8 * j));
    const uint32_t my_dst = keep_u32(smem_u32(smem) + (uint32_t)(L::off_bp + sl * L::bp_stride + j * kSboBp + (c >> 2) * kLboB + (c & 3) * 4));
    const uint32_t sched_s = keep_u32(smem_u32(s_sched));
    const uint32_t my_par = keep_u32(smem_u32(s_par) + (uint32_t)(c * 48));
    // conv1 im2col item of this thread: clip, 16-byte K-chunk (4 taps), pixel column(s).
    //   PACK = 1: 2 clips x 25 chunks x 7 column groups (columns xg, xg+7, xg+14)
    //   PACK = 2: 4 clips x 25 chunks x W0 columns, one item per thread; clip u of a slot sits at row u * kPackOff
    constexpr int kImItems = PACK == 1 ? 350 : CL * 25 * W0;
    static_assert(kImItems <= kCThreads, "im2col items must fit the compute threads");
    const bool im_on = tid < kImItems;
    const int im_kc = tid % 25;
    const int im_r = tid / 25;                                   // PACK 1: slot * 7 + xg;  PACK 2: clip * W0 + x
    const int im_sc = PACK == 1 ? im_r / 7 : im_r / W0;          // clip of the CTA step (slot * PACK + u)
    const int im_x0 = PACK == 1 ? im_r % 7 : im_r % W0;
    const int im_row0 = PACK == 1 ? 0 : (im_sc % PACK) * kPackOff;
    // taps 4kc .. 4kc+3 are two pairs of horizontally adjacent taps (a window row has 10 taps and 4kc is even),
    // each pair one aligned 8-byte read of the padded feature row
    int im_off[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int k = 4 * im_kc + 2 * e, ki = k / 10, kj = k - 10 * ki;
      im_off[e] = ki * FP + kj;
    }
    const float* im_src = s_feat + im_sc * feat_stride;
    unsigned char* im_dst = smem + L::off_b1 + (im_sc / PACK) * L::b1_stride + im_kc * kLboB;
    // rolling feature load: while pair n runs, the two feature-map rows conv1 has just left behind are
    // refilled with rows 2tl-4, 2tl-3 of pair n+1 (thread = clip, row parity, column)
    const bool fl_on = tid < CL * 2 * F;
    const int fl_s = tid / (2 * F), fl_r = (tid / F) & 1, fl_c = tid % F;

    const int per = H0 + 7, total = my_pairs * per + 7;
    const int H3 = H0 + 6;
    const float nborder = (float)((H3 + 2) * (W3 + 2) - H3 * W3);
    const float inv_area = 1.0f / (float)((H3 + 2) * (W3 + 2));
    uint32_t nO = 0, nE = 0;  // completed waits on the two MMA-group barriers

    // global average pool (interior sum + border constant) and fc of the group whose pooling has finished:
    // warp w takes outputs w, w + 12, ...; a lane sums two channels, the warp reduces
    auto pool_fc = [&](long long pair) {
      for (int o = warp; o < CL * p.ncls; o += kFCWarps) {
        const int sc = o / p.ncls, k = o - sc * p.ncls;   // sc = slot * PACK + u
        float acc = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int ch = lane + 32 * h;
          const float pooled = (s_pool[sc * kFC + ch] + s_pool[(CL + sc) * kFC + ch] + s_pool[(2 * CL + sc) * kFC + ch] +
                                nborder * p.border[4 * kFC + ch]) * inv_area;
          acc = fmaf(pooled, p.fc_w[ch * p.ncls + k], acc);
        }
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        const long long b = CL * pair + sc;
        if (lane == 0 && b < nclips) logits_base[b * p.ncls + k] = acc + p.fc_b[k];
      }
    };

    {  // features of the first group -> zero-padded, tf32-rounded maps in shared memory
      const long long first = (long long)CL * bid;
      const float* src = feats_base + first * TF;
      for (int i = tid; i < CL * TF; i += kCThreads) {
        const int sc = i / TF, rem = i - sc * TF, fr = rem / F, fc = rem - fr * F;
        s_feat[sc * feat_stride + (fr + 2) * FP + fc + 2] =
            (first + sc < nclips) ? to_tf32(CHUNK ? __ldcg(src + i) : __ldg(src + i)) : 0.f;
      }
    }
    float psum[PACK];
#pragma unroll
    for (int u = 0; u < PACK; ++u) psum[u] = 0.f;
    V P[4][4], R[4][4];
#pragma unroll
    for (int l = 0; l < 4; ++l)
#pragma unroll
      for (int x = 0; x < 4; ++x) P[l][x] = R[l][x] = fsplat(0.f);
    fused_sync<CHUNK>();

    int tl = 0, n = 0, fc_n = -1, fc_run = -1;  // fc_n / fc_run: group (CTA-local index) whose pooled sums wait in s_pool
    for (int tg = 0; tg < total; ++tg) {
      // ================= phase A: depthwise of blocks 1 and 3 =================
      float fl_v = 0.f;
      // rows conv1 left behind TWO steps ago: a compute warp may be up to one step ahead of the slowest one, whose
      // im2col of the previous step still reads the two rows a one-step distance would overwrite
      const int fl_fr = 2 * tl - 6 + fl_r;
      const bool fl_do = fl_on && n + 1 < my_pairs && fl_fr >= 0 && fl_fr < T;
      if (fl_do) {
        const long long clip = CL * ((long long)bid + (long long)(n + 1) * nblk) + fl_s;
        if (clip < nclips) {
          const float* src = feats_base + clip * TF + fl_fr * F + fl_c;
          fl_v = CHUNK ? __ldcg(src) : __ldg(src);
        }
      }
      if (tg > 0) {  // the GEMMs issued after the previous phase A (pw1, pw3) have retired
        mbar_wait(barO, nO & 1);
        ++nO;
        tc_fence_after();
      }
      const uint2 sch = lds64(sched_s + 8u * (uint32_t)tl);
      // a stage's row exists if its pair does: the current pair n, or pair n - 1 for rows that lag past the period end
      const uint32_t pair_ok = (n < my_pairs ? kSchOn : 0u) | (n >= 1 ? (kSchOn << 16) : 0u);
      auto stage_on = [&](uint32_t e) { return (e & ((e & kSchPrev) ? (pair_ok >> 16) : pair_ok)) != 0u; };
      if (seg_on) {
        const uint32_t e1 = sch.x & 0xfffu, e3 = (sch.x >> 12) & 0xfffu;
        if (stage_on(e1))
          dw_seg<W0, 1, PACK>(my_ld + (e1 & 0x1ffu), j, (e1 & kSchSt) != 0u, my_par + 0u * 48u * kFC,
                              my_dst + (uint32_t)(0 * 2 * L::bp_stride), P[0], R[0]);
        if (stage_on(e3))
          dw_seg<W0 + 4, 1, PACK>(my_ld + (e3 & 0x1ffu), j, (e3 & kSchSt) != 0u, my_par + 2u * 48u * kFC,
                                  my_dst + (uint32_t)(2 * 2 * L::bp_stride), P[2], R[2]);
      }
      if (fc_run >= 0) {  // pooled sums handed over two phase-A's ago: every warp's partial is in s_pool by now
        pool_fc((long long)bid + (long long)fc_run * nblk);
        fc_run = -1;
      }
      if (fc_n >= 0) {
        fc_run = fc_n;
        fc_n = -1;
      }
      if (fl_do) s_feat[fl_s * feat_stride + (fl_fr + 2) * FP + fl_c + 2] = to_tf32(fl_v);
      fence_proxy_async();
      tc_fence_before();
      phase_arrive<5>();
      // ================= phase B: conv1 im2col, depthwise of blocks 2 and 4, pooling =================
      if (tg > 0) {  // the GEMMs issued after the previous phase B (conv1, pw2, pw4) have retired
        mbar_wait(barE, nE & 1);
        ++nE;
        tc_fence_after();
      }
      if (tl < H0 && n < my_pairs && im_on) {
        const float* rowp = im_src + 2 * tl * FP;
        if (PACK == 1) {
          for (int x = im_x0; x < W0; x += 7) {
            const float* q4 = rowp + 2 * x;
            const float2 lo = *reinterpret_cast<const float2*>(q4 + im_off[0]);
            const float2 hi = *reinterpret_cast<const float2*>(q4 + im_off[1]);
            *reinterpret_cast<float4*>(im_dst + (x >> 3) * kSboB1 + (x & 7) * 16) = make_float4(lo.x, lo.y, hi.x, hi.y);
          }
        } else {
          const float* q4 = rowp + 2 * im_x0;
          const float2 lo = *reinterpret_cast<const float2*>(q4 + im_off[0]);
          const float2 hi = *reinterpret_cast<const float2*>(q4 + im_off[1]);
          const int row = im_row0 + im_x0;
          *reinterpret_cast<float4*>(im_dst + (row >> 3) * kSboB1 + (row & 7) * 16) = make_float4(lo.x, lo.y, hi.x, hi.y);
        }
      }
      {
        const uint32_t e2 = sch.y & 0xfffu, e4 = (sch.y >> 12) & 0xfffu;
        if (seg_on && stage_on(e2))
          dw_seg<W0 + 2, 1, PACK>(my_ld + (e2 & 0x1ffu), j, (e2 & kSchSt) != 0u, my_par + 1u * 48u * kFC,
                                  my_dst + (uint32_t)(1 * 2 * L::bp_stride), P[1], R[1]);
        if (seg_on && stage_on(e4))
          dw_seg<W0 + 6, 0, PACK>(my_ld + (e4 & 0x1ffu), j, (e4 & kSchSt) != 0u, my_par + 3u * 48u * kFC,
                                  my_dst + (uint32_t)(3 * 2 * L::bp_stride), P[3], R[3]);
        const uint32_t ep = sch.x >> 24;   // pooling: bit 0 has a row, bit 1 previous pair, bit 2 last row of the group
        if ((ep & 1u) && ((ep & 2u) ? n >= 1 : n < my_pairs)) {  // pooling of the last block's row (interior pixels; the border is added in pool_fc)
          if (seg_on) {
            float raw[8];
            tc_ld8(my_tmem + (uint32_t)kAcc4Col + 8u + (uint32_t)(8 * j), raw);
            tc_wait_ld();
            // columns past a clip's layer are 0; with two clips per slot columns < kPackOff belong to the first
            float lo4 = 0.f, hi4 = 0.f;
#pragma unroll
            for (int x = 0; x < 4; ++x) {
              lo4 += fmaxf(raw[x], 0.f);
              hi4 += fmaxf(raw[4 + x], 0.f);
            }
            if (PACK == 1) {
              psum[0] += lo4 + hi4;
            } else {
              psum[0] += (j < 2 ? lo4 : 0.f) + (j < 1 ? hi4 : 0.f);
              psum[PACK - 1] += (j >= 2 ? lo4 : 0.f) + (j >= 1 ? hi4 : 0.f);
            }
          }
          if (ep & 4u) {  // last row of the group: hand the sums to pool_fc (next phase A)
#pragma unroll
            for (int u = 0; u < PACK; ++u) {
              s_pool[((j * 2 + sl) * PACK + u) * kFC + c] = psum[u];
              psum[u] = 0.f;
            }
            fc_n = (ep & 2u) ? n - 1 : n;
          }
        }
      }
      fence_proxy_async();
      tc_fence_before();
      phase_arrive<6>();
      if (++tl == per) {
        tl = 0;
        ++n;
      }
    }
    // the last step's GEMM groups are empty; wait for them so that nothing is in flight at dealloc
    mbar_wait(barO, nO & 1);
    mbar_wait(barE, nE & 1);
    tc_fence_after();
    asm volatile("bar.sync 2, %0;" ::"n"(kCThreads) : "memory");  // compute warps: every pooling partial is written
    if (fc_run >= 0) pool_fc((long long)bid + (long long)fc_run * nblk);
    if (fc_n >= 0) pool_fc((long long)bid + (long long)fc_n * nblk);
  }
  tc_fence_before();
  fused_sync<CHUNK>();
  if (CHUNK) {  // the next phase of the caller's loop re-initialises these barriers
    if (tid < 2) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&s_bar[tid])) : "memory");
  } else if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512));
  }
}

// model_fused.cu
int fill_fused_params(kws_handle_t h, const float* feats_dev, int64_t B, int64_t T, int64_t F, float* logits_dev,
                      FusedParams* pp);

}  // namespace kws
