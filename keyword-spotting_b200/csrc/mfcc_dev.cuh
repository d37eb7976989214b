// Device helpers shared by the fast MFCC kernel: packed f32x2 arithmetic, the register radix-16
// DFT, the range-reduced log, and the PTX wrappers for mbarrier / TMA bulk copy / tcgen05.
#pragma once
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
// L2 eviction-priority policies (createpolicy) for data that streams through once (evict_first) or must stay
// resident until it is read back (evict_last)
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void tma_load_1d_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void st_global_v4_hint(float* p, float4 v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w),
               "l"(pol)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// bounded wait: a protocol bug traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
#pragma unroll 1
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
// same, for a warp that expects to wait long (service warp): back off between probes so the
// polling does not take issue slots from the compute warps
__device__ __forceinline__ void mbar_wait_idle(uint32_t bar, uint32_t parity, unsigned ns = 256) {
  uint32_t done = 0;
#pragma unroll 1
  for (int spin = 0; spin < (1 << 22); ++spin) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) return;
    __nanosleep(ns);
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
// Converged-warp variants: the WHOLE warp executes them with warp-uniform operands and one elected lane
// issues.  Issuing from inside an `if (lane == 0)` region instead makes ptxas wrap every MMA in a per-thread
// waterfall loop (4x R2UR + ELECT + branch per instruction); converged, the descriptors stay in uniform registers.
__device__ __forceinline__ void tc_mma_tf32_elect(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_commit_elect(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar)
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
__device__ __forceinline__ void tc_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
// 16 lanes x 256 bits: thread t gets rows t/4 and t/4 + 8 (of the 16 lanes at the address) and columns 2 (t%4) + {0, 1} of
// every 8-column block; register [4 blk + 2 s + e] = D[row t/4 + 8 s][column 8 blk + 2 (t%4) + e]
__device__ __forceinline__ void tc_ld_16x256b_x1(uint32_t taddr, float* v) {
  uint32_t r[4];
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(taddr)
               : "memory");
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tc_ld_16x256b_x2(uint32_t taddr, float* v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
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

}  // namespace kws
