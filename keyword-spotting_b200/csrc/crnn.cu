// CRNN / GRU keyword-spotting model (SURVEY.md a-X3; BASELINE config 4 "log-mel + CRNN/GRU").
// The reference has NO recurrent model (grep GRU|LSTM|RNN over its tree returns nothing), so this is
// an extension authored in the reference's idiom -- a KeywordSpottingModel subclass
// (kws/libs/models.py:13-72) whose front end is the reference's own first layer:
//     conv1 10x10 s2 p2 (1 -> Cc) + ReLU          (models.py:135,170)
//     -> per time step t the feature vector x_t[c * W1 + w] = conv[c, t, w]   (I = Cc * W1)
//     -> num_layers x torch.nn.GRU(I, H, batch_first=True) semantics, h_0 = 0
//     -> fc(h_T)                                   (models.py:144,181)
// Parity is against the builder's PyTorch definition (oracle/crnn_ref.py); the reference pins nothing.
//
// Kernels: conv front end = conv1_relu_kernel of model.cu (NHWC, so x_t is [w][c]-ordered and W_ih's
// columns are permuted to match at finalize); gru_inproj_kernel = the time-parallel half of the GRU, one
// [B*T', I] x [I, 3H] GEMM (+ b_ih) -- shared-memory tiled fp32; gru_recurrent_kernel = the sequential
// half: W_hh resident in shared memory, a tile of clips per CTA, hidden state in shared memory, gates
// in registers (r, z sigmoid; n = tanh(xn + r * (W_hn h + b_hn)); h' = (1 - z) n + z h).
#include <cuda_runtime.h>

#include <cstring>
#include <string>
#include <vector>

#include "mfcc_dev.cuh"

namespace kws {

struct CrnnWeights {
  bool created = false, finalized = false;
  int num_classes = 0, cc = 0, hidden = 0, layers = 0, w1 = 0;  // w1: conv output width the GRU was sized for
  std::vector<std::string> names;
  std::vector<std::vector<float>> host;
  std::vector<bool> loaded;
  float* d_conv_w = nullptr;   // [100][Cc]
  float* d_conv_b = nullptr;   // [Cc]
  float* d_c1img = nullptr;    // tensor-core conv1 (Cc == 16, 18 output columns): [dy][hi|lo][32 x 16] UMMA K-major tiles
  float* d_wih[2] = {};        // [I][3H]  (k-major; layer 0 columns in [w][c] order)
  float* d_wimg[2] = {};       // tensor-core projection: per 32-wide K chunk [tf32 hi tile | tf32 lo tile] of W_ih, UMMA K-major
  float* d_whh[2] = {};        // [H][3H]  (k-major)
  float* d_whimg[2] = {};      // tensor-core recurrence (H == 64): W_hh in the same tile format
  float* d_bih[2] = {};        // [3H]
  float* d_bhh[2] = {};        // [3H]
  float* d_fc_w = nullptr;     // [H][K]
  float* d_fc_b = nullptr;     // [K]
  void* d_ws = nullptr;
  size_t ws_bytes = 0;
};

// C[M, N] = A[M, K] * Bm[K, N] + bias[N]; 64 x 64 tile, 256 threads, 4 x 4 outputs per thread
__global__ void __launch_bounds__(256) gru_inproj_kernel(const float* __restrict__ A, const float* __restrict__ Bm,
                                                         const float* __restrict__ bias, float* __restrict__ Cout,
                                                         long long M, int N, int K) {
  __shared__ float sA[16][64 + 4];
  __shared__ __align__(16) float sB[16][64];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const long long m0 = (long long)blockIdx.x * 64;
  const int n0 = blockIdx.y * 64;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += 16) {
    for (int i = tid; i < 64 * 16; i += 256) {
      const int r = i >> 4, kk = i & 15;
      const long long m = m0 + r;
      sA[kk][r] = (m < M && k0 + kk < K) ? A[m * K + k0 + kk] : 0.f;
    }
    for (int i = tid; i < 16 * 64; i += 256) {
      const int kk = i >> 6, c = i & 63;
      sB[kk][c] = (k0 + kk < K && n0 + c < N) ? Bm[(size_t)(k0 + kk) * N + n0 + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[kk][ty * 4 + i];
      const float4 b = *reinterpret_cast<const float4*>(&sB[kk][tx * 4]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[i][0] = fmaf(a[i], b.x, acc[i][0]);
        acc[i][1] = fmaf(a[i], b.y, acc[i][1]);
        acc[i][2] = fmaf(a[i], b.z, acc[i][2]);
        acc[i][3] = fmaf(a[i], b.w, acc[i][3]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < N) Cout[m * N + n] = acc[i][j] + bias[n];
    }
  }
}

// The same GEMM, register-tiled: 128 x 64 tile, 8 x 4 outputs per thread as packed f32x2 pairs, K in slices
// of 16 through two shared-memory buffers (the next slice's global loads are in flight while the current one
// is multiplied).  Requires K % 16 == 0 and N % 64 == 0 (288 -> 192 and 64 -> 192 of the CRNN do).
__global__ void __launch_bounds__(256) gru_inproj_tiled_kernel(const float* __restrict__ A, const float* __restrict__ Bm,
                                                               const float* __restrict__ bias, float* __restrict__ Cout,
                                                               long long M, int N, int K) {
  constexpr int TM = 128, TN = 64, TK = 16, LDA = TM + 4;
  __shared__ __align__(16) float sA[2][TK][LDA];   // [k][m]
  __shared__ __align__(16) float sB[2][TK][TN];    // [k][n]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const long long m0 = (long long)blockIdx.x * TM;
  const int n0 = blockIdx.y * TN;
  // global -> register staging: A rows (2 float4 along k per thread), B rows (1 float4 along n)
  const int ar = tid >> 2, ak = (tid & 3) * 4;         // rows ar and ar + 64, k offset ak
  const int bk = tid >> 4, bn = (tid & 15) * 4;
  float4 ra[2], rb;
  auto gload = [&](int k0) {
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const long long m = m0 + ar + 64 * hh;
      ra[hh] = m < M ? *reinterpret_cast<const float4*>(A + m * K + k0 + ak) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    rb = *reinterpret_cast<const float4*>(Bm + (size_t)(k0 + bk) * N + n0 + bn);
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int r = ar + 64 * hh;
      sA[buf][ak + 0][r] = ra[hh].x;
      sA[buf][ak + 1][r] = ra[hh].y;
      sA[buf][ak + 2][r] = ra[hh].z;
      sA[buf][ak + 3][r] = ra[hh].w;
    }
    *reinterpret_cast<float4*>(&sB[buf][bk][bn]) = rb;
  };
  V acc[8][2];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i][0] = acc[i][1] = fsplat(0.f);
  gload(0);
  sstore(0);
  __syncthreads();
  const int nk = K / TK;
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload((kt + 1) * TK);
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&sA[buf][kk][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&sA[buf][kk][ty * 8 + 4]);
      const float4 b = *reinterpret_cast<const float4*>(&sB[buf][kk][tx * 4]);
      const V b0 = make_float2(b.x, b.y), b1 = make_float2(b.z, b.w);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        acc[i][0] = ffma(fsplat(av[i]), b0, acc[i][0]);
        acc[i][1] = ffma(fsplat(av[i]), b1, acc[i][1]);
      }
    }
    if (kt + 1 < nk) sstore(buf ^ 1);   // the other buffer was last read one iteration ago (barrier below)
    __syncthreads();
  }
  const float4 bv = *reinterpret_cast<const float4*>(bias + n0 + tx * 4);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long m = m0 + ty * 8 + i;
    if (m < M)
      *reinterpret_cast<float4*>(Cout + m * N + n0 + tx * 4) =
          make_float4(acc[i][0].x + bv.x, acc[i][0].y + bv.y, acc[i][1].x + bv.z, acc[i][1].y + bv.w);
  }
}

// The time-parallel half on the tensor core: C[M, N] = A[M, K] . W[N, K]^T + bias, fp32-grade through a 3-term
// tf32 split (A_hi W_hi + A_lo W_hi + A_hi W_lo, operands rounded with cvt.rna; error ~2^-21 like the MFCC's DCT).
// Both operands are K-major as they lie in memory (activations [row][feature], weights [gate][feature]), so no
// transposition is needed.  Persistent CTAs, tile = 128 rows x N (<= 256) columns, K streamed in chunks of 32
// through two stages, no CTA-wide barrier in the main loop:
//   warps 0..14 producers: read the A chunk (coalesced float4), split hi/lo, store both K-major tiles
//               (144-byte K-chunk pitch: the 8 lanes of a quarter warp hit 32 different banks), arrive on fullA
//   warp  15    issue warp (converged, elected lane): requests the chunk's pre-split weight tiles by TMA bulk
//               copy as soon as the stage is free, waits fullA / fullB, issues the 12 tcgen05.mma of the chunk,
//               tcgen05.commit -> empty[stage] (and -> acc_done after the last chunk of a tile)
//   warps 0..3  epilogue: TMEM -> registers -> + bias -> global, then arrive on acc_free so that the next tile's
//               first MMA may overwrite the accumulator; warps 4..6 already produce the next tile meanwhile.

// Loads the compiler may not sink below a later asm volatile (the stage waits): the next stage's data is
// requested BEFORE the current stage is converted.
__device__ __forceinline__ float4 ldg_early4(const float4* ptr) {
  float4 v;
  asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(ptr));
  return v;
}
__device__ __forceinline__ float ldg_early(const float* ptr) {
  float v;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(ptr));
  return v;
}
constexpr int kDK = 32, kDLboA = 144, kDSboA = (kDK / 4) * kDLboA, kDATile = 16 * kDSboA;   // A: 128 rows x 32 K
constexpr int kDProd = 15;                                                                   // producer warps
constexpr int kDBStages = 3;                                                                 // weight-tile ring of the projection
struct DenseParams {
  const float* A;
  const float* wimg;   // [K/32][2][N/8 * 1024 B]
  const float* bias;
  float* C;
  long long M;
  int N, K;
  int atile;   // 1 (needs T > 0): A is tile-major [tile][K][128 rows] as conv1_tc_kernel writes it
  int T;   // 0: rows as they lie, C row-major.  > 0: "step-major" tiles for the tensor-core recurrence: A rows are
           // (clip, step) pairs, tile = (group of 128 clips, step), C = [group][step][column][128 clips]
};
__global__ void __launch_bounds__((kDProd + 1) * 32, 1) dense_gemm_tf32x3_kernel(const DenseParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar[12];   // fullA[2], empty[2], fullB[3], emptyB[3], acc_done, acc_free
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int N = p.N, K = p.K, nch = K / kDK;
  const uint32_t btile = (uint32_t)(N / 8) * 1024u;            // one weight tile (hi or lo) of a chunk
  // shared memory: two A stages (hi|lo tiles written by the producers), then kDBStages weight stages (TMA)
  const uint32_t boff = 2u * 2u * kDATile;
  const uint32_t bar0 = smem_u32(&s_bar[0]);
  auto fullA = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 16u + 8u * s; };
  auto fullB = [&](int s) { return bar0 + 32u + 8u * s; };
  auto emptyB = [&](int s) { return bar0 + 56u + 8u * s; };
  const uint32_t acc_done = bar0 + 80u, acc_free = bar0 + 88u;
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(fullA(s), kDProd);
      mbar_init(empty(s), 1);
    }
    for (int s = 0; s < kDBStages; ++s) {
      mbar_init(fullB(s), 1);
      mbar_init(emptyB(s), 1);
    }
    mbar_init(acc_done, 1);
    mbar_init(acc_free, 4);
    fence_mbar_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const int Tm = p.T;
  const long long nrows = Tm > 0 ? p.M / Tm : p.M;            // step-major: clips
  const long long ntiles = Tm > 0 ? ((nrows + 127) / 128) * Tm : (p.M + 127) / 128;
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);

  if (warp == kDProd) {
    // ============================ issue warp ============================
    // The weight tiles run kDBStages - 1 chunks ahead of the MMAs in their own ring: a stage is refilled as soon as
    // the chunk before the current one has retired, so the copy has two chunks of MMA time to land.
    const long long mine = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const uint32_t total = (uint32_t)(mine * nch);
    auto fetchB = [&](uint32_t c) {   // chunk c of this CTA's flat sequence
      const int sb = c % kDBStages;
      if (c >= (uint32_t)kDBStages) mbar_wait(emptyB(sb), (c / kDBStages - 1) & 1);
      if (lane == 0) {
        mbar_expect_tx(fullB(sb), 2u * btile);
        tma_load_1d(smem_u32(smem) + boff + (uint32_t)sb * 2u * btile,
                    reinterpret_cast<const unsigned char*>(p.wimg) + (size_t)(c % nch) * 2u * btile, 2u * btile, fullB(sb));
      }
      __syncwarp();
    };
    for (uint32_t c = 0; c < (uint32_t)(kDBStages - 1) && c < total; ++c) fetchB(c);
    uint32_t cnt = 0, it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      for (int kc = 0; kc < nch; ++kc, ++cnt) {
        const int s = cnt & 1, sb = cnt % kDBStages;
        if (cnt + kDBStages - 1 < total) fetchB(cnt + kDBStages - 1);
        const uint32_t sa = smem_u32(smem) + (uint32_t)s * 2u * kDATile;
        mbar_wait(fullA(s), (cnt >> 1) & 1);
        mbar_wait(fullB(sb), (cnt / kDBStages) & 1);
        if (kc == 0 && it > 0) mbar_wait(acc_free, (it - 1) & 1);    // the previous tile's epilogue has read TMEM
        tc_fence_after();
        const uint32_t a_hi = sa, a_lo = sa + kDATile, b_hi = smem_u32(smem) + boff + (uint32_t)sb * 2u * btile, b_lo = b_hi + btile;
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t ab = term == 1 ? a_lo : a_hi, bb = term == 2 ? b_lo : b_hi;
#pragma unroll
          for (int ks = 0; ks < kDK / 8; ++ks)
            tc_mma_tf32_elect(tmem, umma_desc(ab + ks * 2u * kDLboA, kDLboA, kDSboA), umma_desc(bb + ks * 256u, 128u, 1024u),
                              idesc, (kc > 0 || term > 0 || ks > 0) ? 1u : 0u);
        }
        tc_commit_elect(empty(s));
        tc_commit_elect(emptyB(sb));
        if (kc == nch - 1) tc_commit_elect(acc_done);
      }
    }
  } else {
    // ============================ producers (+ epilogue on warps 0..3) ============================
    // The loads of stage c+1 are issued before stage c is converted: two chunks per thread in flight, the global
    // latency is hidden behind the conversion, the stage wait and (warps 0..3) the epilogue.
    constexpr int kIt = (128 * (kDK / 4) + kDProd * 32 - 1) / (kDProd * 32);
    auto load = [&](long long tile, int kc, float4 (&v)[kIt]) {
      const long long m0 = Tm > 0 ? (tile / Tm) * 128 : tile * 128;      // row r of the tile is A row (m0 + r) * mstride + moff
      const long long mstride = Tm > 0 ? Tm : 1, moff = Tm > 0 ? tile % Tm : 0;
      if (p.atile) {   // lanes = rows: four coalesced scalar loads per item
        const float* at = p.A + ((size_t)tile * K + (size_t)kc * kDK) * 128;
#pragma unroll
        for (int u = 0; u < kIt; ++u) {
          const int i = tid + u * kDProd * 32, r = i & 127, c = i >> 7;
          v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (i < 128 * (kDK / 4)) {
            const float* a = at + (size_t)(4 * c) * 128 + r;
            v[u] = make_float4(ldg_early(a), ldg_early(a + 128), ldg_early(a + 256), ldg_early(a + 384));
          }
        }
        return;
      }
#pragma unroll
      for (int u = 0; u < kIt; ++u) {
        const int i = tid + u * kDProd * 32, r = i >> 3, c = i & 7;   // row of the tile, 16-byte K chunk
        const long long m = m0 + r;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < 128 * (kDK / 4) && m < nrows)
          v[u] = ldg_early4(reinterpret_cast<const float4*>(p.A + (m * mstride + moff) * K + kc * kDK) + c);
      }
    };
    uint32_t cnt = 0, it = 0;
    long long tile = blockIdx.x;
    int kc = 0;
    float4 v[kIt], vn[kIt];
    if (tile < ntiles) load(tile, 0, v);
    while (tile < ntiles) {
      long long ntile = tile;
      int nkc = kc + 1;
      if (nkc == nch) nkc = 0, ntile += gridDim.x;
      if (ntile < ntiles) load(ntile, nkc, vn);
      const int s = cnt & 1;
      unsigned char* sa = smem + (size_t)s * 2 * kDATile;
      if (cnt >= 2) mbar_wait(empty(s), ((cnt - 2) >> 1) & 1);
#pragma unroll
      for (int u = 0; u < kIt; ++u) {
        const int i = tid + u * kDProd * 32, r = p.atile ? (i & 127) : (i >> 3), c = p.atile ? (i >> 7) : (i & 7);
        if (i < 128 * (kDK / 4)) {
          const float4 hi = make_float4(to_tf32(v[u].x), to_tf32(v[u].y), to_tf32(v[u].z), to_tf32(v[u].w));
          const float4 lo = make_float4(to_tf32(v[u].x - hi.x), to_tf32(v[u].y - hi.y), to_tf32(v[u].z - hi.z),
                                        to_tf32(v[u].w - hi.w));
          const uint32_t off = (uint32_t)(r >> 3) * kDSboA + (uint32_t)c * kDLboA + (uint32_t)(r & 7) * 16u;
          *reinterpret_cast<float4*>(sa + off) = hi;
          *reinterpret_cast<float4*>(sa + kDATile + off) = lo;
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(fullA(s));
      if (kc == nch - 1) {
        if (warp < 4) {  // epilogue: lane = accumulator row
          mbar_wait(acc_done, it & 1);
          tc_fence_after();
          const long long m = (Tm > 0 ? (tile / Tm) * 128 : tile * 128) + warp * 32 + lane;
          float* crow = p.C + m * N;
          float* ccol = p.C + tile * N * 128 + warp * 32 + lane;   // step-major: column pitch 128 floats, lanes contiguous
          for (int c0 = 0; c0 < N; c0 += 16) {
            float q[16];
            tc_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, q);
            tc_wait_ld();
            if (Tm > 0) {
#pragma unroll
              for (int j = 0; j < 16; ++j) ccol[(size_t)(c0 + j) * 128] = q[j] + __ldg(p.bias + c0 + j);
            } else if (m < nrows) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float4 bv = __ldg(reinterpret_cast<const float4*>(p.bias + c0) + j);
                *reinterpret_cast<float4*>(crow + c0 + 4 * j) =
                    make_float4(q[4 * j] + bv.x, q[4 * j + 1] + bv.y, q[4 * j + 2] + bv.z, q[4 * j + 3] + bv.w);
              }
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acc_free);
        }
        ++it;
      }
#pragma unroll
      for (int u = 0; u < kIt; ++u) v[u] = vn[u];
      tile = ntile, kc = nkc, ++cnt;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256));
}

// conv1 (10 x 10, stride 2, padding 2, + bias, ReLU) of the reference-idiom CRNN (16 channels, 40 features) on the
// tensor core, fp32-grade through the same 3-term tf32 split.  Tile = 128 clips at one output step t; per kernel row
// dy the A operand is the feature row 2t-2+dy of those clips, stored K-major with two zero columns in front (the
// conv's own left padding), so that the 10-wide window of output column w starts at stored column 2w.  Output
// columns are taken in pairs (w0 = 2g, w0+1): their windows cover stored columns 4g .. 4g+11, i.e. they start on a
// 16-byte K chunk - the SAME small weight tile [16 window columns x (2 x 16 channels)] serves all nine pairs, only
// the A descriptor's start moves (a Toeplitz matrix never materialised).  Rows outside the clip (top/bottom
// padding) are simply not issued: a tile is one step, so they are the same for all its rows.
//   warps 0..14 producers: feature row -> tf32 hi/lo K-major tiles (2 stages), arrive on fullA
//   warp  15    issue warp: 30 tcgen05.mma (N = 64) per kernel row, commit -> empty[stage] / acc_done
//   warps 0..3  epilogue: TMEM -> + bias -> ReLU -> conv output [clip][t][w*16 + c]
constexpr int kC1Lbo = 144, kC1Sbo = 12 * kC1Lbo, kC1ATile = 16 * kC1Sbo;   // A: 128 rows x 48 stored columns
constexpr int kC1Stages = 2, kC1WBytes = 10 * 2 * 4096;                     // weights: [dy][hi|lo][64 n x 16 k]
constexpr int kC1Smem = kC1WBytes + kC1Stages * 2 * kC1ATile;
constexpr int kC1Prod = 15, kC1Threads = (kC1Prod + 1) * 32;              // producer warps + the issue warp
struct Conv1TcParams {
  const float* feats;   // [B][T][40]
  const float* wimg;    // kC1WBytes
  const float* bias;    // [16]
  float* out;           // [group * H1 + t][288][128 clips]  (tile-major: what dense_gemm_tf32x3_kernel reads with atile = 1)
  long long B;
  int T, H1;
};
__global__ void __launch_bounds__(kC1Threads, 1) conv1_tc_kernel(const Conv1TcParams p) {
  constexpr int F = 40, NC = 288;
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar[2 * kC1Stages + 3];   // fullA[3], empty[3], wfull, acc_done, acc_free
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31, T = p.T, H1 = p.H1;
  const uint32_t bar0 = smem_u32(&s_bar[0]);
  auto fullA = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (kC1Stages + s); };
  const uint32_t wfull = bar0 + 8u * (2 * kC1Stages), acc_done = wfull + 8u, acc_free = wfull + 16u;
  unsigned char* s_a = smem + kC1WBytes;
  if (tid == 0) {
    for (int s = 0; s < kC1Stages; ++s) {
      mbar_init(fullA(s), kC1Prod);
      mbar_init(empty(s), 1);
    }
    mbar_init(wfull, 1);
    mbar_init(acc_done, 1);
    mbar_init(acc_free, 4);
    fence_mbar_init();
  }
  // stored columns 0, 1 and 42..47 stay zero for the whole kernel (padding / K round-up)
  for (int i = tid; i < kC1Stages * 2 * kC1ATile / 16; i += kC1Threads) reinterpret_cast<float4*>(s_a)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  fence_proxy_async();
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const long long ngroups = (p.B + 127) / 128, ntiles = ngroups * H1;

  if (warp == kC1Prod) {
    // ============================ issue warp ============================
    if (lane == 0) {
      mbar_expect_tx(wfull, kC1WBytes);
      tma_load_1d(smem_u32(smem), p.wimg, kC1WBytes, wfull);
    }
    __syncwarp();
    mbar_wait(wfull, 0);
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    uint32_t cnt = 0, it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int t = (int)(tile % H1);
      const int dy_lo = 2 - 2 * t > 0 ? 2 - 2 * t : 0, dy_hi = T + 1 - 2 * t < 9 ? T + 1 - 2 * t : 9;
      for (int dy = dy_lo; dy <= dy_hi; ++dy, ++cnt) {
        const int s = cnt % kC1Stages;
        mbar_wait(fullA(s), (cnt / kC1Stages) & 1);
        if (dy == dy_lo && it > 0) mbar_wait(acc_free, (it - 1) & 1);
        tc_fence_after();
        const uint32_t a_hi = smem_u32(s_a) + (uint32_t)s * 2u * kC1ATile, a_lo = a_hi + kC1ATile;
        const uint32_t b_hi = smem_u32(smem) + (uint32_t)dy * 8192u, b_lo = b_hi + 4096u;
#pragma unroll
        for (int g = 0; g < 5; ++g) {
#pragma unroll
          for (int term = 0; term < 3; ++term) {
            const uint32_t ab = (term == 1 ? a_lo : a_hi) + (uint32_t)g * 2u * kC1Lbo, bb = term == 2 ? b_lo : b_hi;
#pragma unroll
            for (int kk = 0; kk < 2; ++kk)
              tc_mma_tf32_elect(tmem + 64u * g, umma_desc(ab + kk * 2u * kC1Lbo, kC1Lbo, kC1Sbo), umma_desc(bb + kk * 256u, 128u, 512u),
                                idesc, (dy > dy_lo || term > 0 || kk > 0) ? 1u : 0u);
          }
        }
        tc_commit_elect(empty(s));
        if (dy == dy_hi) tc_commit_elect(acc_done);
      }
    }
  } else {
    // ============================ producers (+ epilogue on warps 0..3) ============================
    // Stages are (tile, kernel row) pairs walked in one flat sequence; a stage's loads are issued before the wait
    // for its shared-memory slot.  15 producer warps: the conversion is issue-latency bound with fewer.
    constexpr int kItems = 128 * (F / 4), kIt = (kItems + kC1Prod * 32 - 1) / (kC1Prod * 32);
    auto dy_first = [&](long long tile) { const int t = (int)(tile % H1); return 2 - 2 * t > 0 ? 2 - 2 * t : 0; };
    auto dy_last = [&](long long tile) { const int t = (int)(tile % H1); return T + 1 - 2 * t < 9 ? T + 1 - 2 * t : 9; };
    auto load = [&](long long tile, int dy, float4 (&v)[kIt]) {
      const long long c0 = (tile / H1) * 128;
      const int y = 2 * (int)(tile % H1) - 2 + dy;
#pragma unroll
      for (int u = 0; u < kIt; ++u) {
        const int i = tid + u * kC1Prod * 32, r = i / (F / 4), c = i - r * (F / 4);
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < kItems && c0 + r < p.B) v[u] = ldg_early4(reinterpret_cast<const float4*>(p.feats + ((c0 + r) * T + y) * F) + c);
      }
    };
    uint32_t cnt = 0, it = 0;
    long long tile = blockIdx.x;
    int dy = tile < ntiles ? dy_first(tile) : 0;
    float4 v[kIt];
    while (tile < ntiles) {
      load(tile, dy, v);
      const bool last = dy == dy_last(tile);
      long long ntile = tile;
      int ndy = dy + 1;
      if (last) {
        ntile += gridDim.x;
        ndy = ntile < ntiles ? dy_first(ntile) : 0;
      }
      const int s = cnt % kC1Stages;
      unsigned char* sa = s_a + (size_t)s * 2 * kC1ATile;
      if (cnt >= kC1Stages) mbar_wait(empty(s), (cnt / kC1Stages - 1) & 1);
#pragma unroll
      for (int u = 0; u < kIt; ++u) {
        const int i = tid + u * kC1Prod * 32, r = i / (F / 4), c = i - r * (F / 4);
        if (i < kItems) {
          const float2 h0 = make_float2(to_tf32(v[u].x), to_tf32(v[u].y)), h1 = make_float2(to_tf32(v[u].z), to_tf32(v[u].w));
          const float2 l0 = make_float2(to_tf32(v[u].x - h0.x), to_tf32(v[u].y - h0.y));
          const float2 l1 = make_float2(to_tf32(v[u].z - h1.x), to_tf32(v[u].w - h1.y));
          // feature columns 4c .. 4c+3 are stored columns 4c+2 .. 4c+5: the upper half of K chunk c, the lower of c+1
          unsigned char* d = sa + (r >> 3) * kC1Sbo + c * kC1Lbo + (r & 7) * 16;
          *reinterpret_cast<float2*>(d + 8) = h0;
          *reinterpret_cast<float2*>(d + kC1Lbo) = h1;
          *reinterpret_cast<float2*>(d + kC1ATile + 8) = l0;
          *reinterpret_cast<float2*>(d + kC1ATile + kC1Lbo) = l1;
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(fullA(s));
      if (last) {
        if (warp < 4) {  // epilogue: lane = accumulator row = clip of the group
          mbar_wait(acc_done, it & 1);
          tc_fence_after();
          // tile-major output [tile][w*16 + c][128 clips]: lanes = clips, every store a full 128-byte line
          float* ocol = p.out + (size_t)tile * NC * 128 + warp * 32 + lane;
          float bv[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) bv[j] = __ldg(p.bias + j);
#pragma unroll 2
          for (int w = 0; w < NC / 16; ++w) {
            float q[16];
            tc_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(16 * w), q);
            tc_wait_ld();
#pragma unroll
            for (int j = 0; j < 16; ++j) ocol[(size_t)(16 * w + j) * 128] = fmaxf(q[j] + bv[j], 0.f);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acc_free);
        }
        ++it;
      }
      tile = ntile, dy = ndy, ++cnt;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512));
}

// Sequential half of one GRU layer.  xp [B, T, 3H] = W_ih x_t + b_ih (gate order r, z, n like torch).
// CTA = kGruNB clips; W_hh [H][3H] and the hidden state (transposed, [H][NB], double-buffered) live in shared
// memory.  Thread = (hidden unit j, group of 8 clips): per k it reads its three gate weights once and eight
// states as two float4, 24 FMAs per 5 shared loads; all lanes of a warp share the clip group, so the state
// reads are broadcasts.  hseq (nullable) [B, T, H] receives every h_t; hlast [B, H] the final state.
// H == 64 (the reference-idiom CRNN); other sizes use gru_recurrent_generic_kernel.
constexpr int kGruNB = 32;
__global__ void __launch_bounds__(256) gru_recurrent_kernel(const float* __restrict__ xp, const float* __restrict__ whh,
                                                            const float* __restrict__ bhh, float* __restrict__ hseq,
                                                            float* __restrict__ hlast, long long B, int T) {
  constexpr int H = 64, NB = kGruNB;
  extern __shared__ __align__(16) float sm_gru[];
  float* s_w = sm_gru;                     // [H][3H]
  float* s_h = s_w + (size_t)H * 3 * H;    // [2][H][NB]
  const int tid = threadIdx.x, j = tid & 63, cg = tid >> 6;   // clips cg*8 .. cg*8+7 of the CTA
  for (int i = tid; i < H * 3 * H; i += 256) s_w[i] = whh[i];
  for (int i = tid; i < 2 * NB * H; i += 256) s_h[i] = 0.f;
  __syncthreads();
  const long long b0 = (long long)blockIdx.x * NB + cg * 8;
  const float br = bhh[j], bz = bhh[H + j], bn = bhh[2 * H + j];
  for (int t = 0; t < T; ++t) {
    const float* hc = s_h + (t & 1) * NB * H;
    float* hn = s_h + ((t + 1) & 1) * NB * H;
    float ar[8], az[8], an[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) ar[c] = br, az[c] = bz, an[c] = bn;
#pragma unroll 4
    for (int k = 0; k < H; ++k) {
      const float wr = s_w[k * 3 * H + j], wz = s_w[k * 3 * H + H + j], wn = s_w[k * 3 * H + 2 * H + j];
      const float4 h0 = *reinterpret_cast<const float4*>(hc + k * NB + cg * 8);
      const float4 h1 = *reinterpret_cast<const float4*>(hc + k * NB + cg * 8 + 4);
      const float hv[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        ar[c] = fmaf(hv[c], wr, ar[c]);
        az[c] = fmaf(hv[c], wz, az[c]);
        an[c] = fmaf(hv[c], wn, an[c]);
      }
    }
    float hnew[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const long long b = b0 + c;
      hnew[c] = 0.f;
      if (b < B) {
        const float* x = xp + (b * T + t) * 3 * H;
        const float r = 1.f / (1.f + expf(-(x[j] + ar[c])));
        const float z = 1.f / (1.f + expf(-(x[H + j] + az[c])));
        const float n = tanhf(x[2 * H + j] + r * an[c]);
        hnew[c] = (1.f - z) * n + z * hc[j * NB + cg * 8 + c];
        if (hseq) hseq[(b * T + t) * H + j] = hnew[c];
        if (t == T - 1) hlast[b * H + j] = hnew[c];
      }
    }
    *reinterpret_cast<float4*>(hn + j * NB + cg * 8) = make_float4(hnew[0], hnew[1], hnew[2], hnew[3]);
    *reinterpret_cast<float4*>(hn + j * NB + cg * 8 + 4) = make_float4(hnew[4], hnew[5], hnew[6], hnew[7]);
    __syncthreads();
  }
}

// The sequential half on the tensor core (H == 64).  CTA = 128 clips = the 128 rows of one accumulator.  Per step
// D[128 x 192] = h_{t-1} . W_hh^T (3-term tf32 split like the projection, K = 64: 24 tcgen05.mma from the issue
// warp); the 8 gate warps (lane = clip row, warp / 4 = which 32 units) then read the r/z/n pre-activations from
// TMEM, add the step-major input projection (lanes = clips: coalesced), apply the gates, keep h in fp32 registers
// and write its tf32 hi/lo split into the K-major A tiles of the next step.  W_hh hi/lo tiles arrive once by TMA.
// sigmoid/tanh through ex2 + fast division (abs error ~1e-7, far inside the 1e-4 logit tolerance).
constexpr int kGtBTile = (192 / 8) * 1024;                     // one W_hh tile (hi or lo) of a 32-wide K chunk
constexpr int kGtSmem = 4 * kGtBTile + 4 * kDATile;            // W_hh: 2 chunks x (hi|lo); h: 2 chunks x (hi|lo)
struct GruTcParams {
  const float* xg;     // [group][T][192][128]  (dense_gemm_tf32x3_kernel, step-major)
  const float* wimg;   // [2][hi tile | lo tile]
  const float* bhh;
  float* hseq;         // nullable, [B][T][64]
  float* hlast;        // [B][64]
  long long B;
  int T;
};
__device__ __forceinline__ float fast_sigmoid(float a) { return __fdividef(1.f, 1.f + __expf(-a)); }
__device__ __forceinline__ float fast_tanh(float a) { return 1.f - __fdividef(2.f, 1.f + __expf(2.f * a)); }

__global__ void __launch_bounds__(288, 1) gru_recurrent_tc_kernel(const GruTcParams p) {
  constexpr int H = 64, N = 3 * H;
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar[3];   // wfull, h_ready, acc_done
  __shared__ uint32_t s_tmem;
  __shared__ float s_b[N];
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31, T = p.T;
  const uint32_t bar0 = smem_u32(&s_bar[0]), wfull = bar0, h_ready = bar0 + 8u, acc_done = bar0 + 16u;
  unsigned char* s_a = smem + 4 * kGtBTile;
  if (tid == 0) {
    mbar_init(wfull, 1);
    mbar_init(h_ready, 8);
    mbar_init(acc_done, 1);
    fence_mbar_init();
  }
  for (int i = tid; i < 4 * kDATile / 16; i += 288) reinterpret_cast<float4*>(s_a)[i] = make_float4(0.f, 0.f, 0.f, 0.f);  // h_0 = 0
  for (int i = tid; i < N; i += 288) s_b[i] = p.bhh[i];
  fence_proxy_async();
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(256));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  if (warp == 8) {
    // ============================ issue warp ============================
    if (lane == 0) {
      mbar_expect_tx(wfull, 4u * kGtBTile);
      tma_load_1d(smem_u32(smem), p.wimg, 2u * kGtBTile, wfull);
      tma_load_1d(smem_u32(smem) + 2u * kGtBTile, reinterpret_cast<const unsigned char*>(p.wimg) + 2u * kGtBTile, 2u * kGtBTile,
                  wfull);
    }
    __syncwarp();
    mbar_wait(wfull, 0);
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    for (int t = 0; t < T; ++t) {
      if (t > 0) mbar_wait(h_ready, (t - 1) & 1);      // h_{t-1} is in the A tiles and the gates have read D
      tc_fence_after();
#pragma unroll
      for (int kc = 0; kc < 2; ++kc) {
        const uint32_t a_hi = smem_u32(s_a) + kc * 2u * kDATile, a_lo = a_hi + kDATile;
        const uint32_t b_hi = smem_u32(smem) + kc * 2u * kGtBTile, b_lo = b_hi + kGtBTile;
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t ab = term == 1 ? a_lo : a_hi, bb = term == 2 ? b_lo : b_hi;
#pragma unroll
          for (int ks = 0; ks < kDK / 8; ++ks)
            tc_mma_tf32_elect(tmem, umma_desc(ab + ks * 2u * kDLboA, kDLboA, kDSboA), umma_desc(bb + ks * 256u, 128u, 1024u),
                              idesc, (kc > 0 || term > 0 || ks > 0) ? 1u : 0u);
        }
      }
      tc_commit_elect(acc_done);
    }
  } else {
    // ============================ gate warps ============================
    const int rg = warp & 3, uh = warp >> 2, row = rg * 32 + lane;
    const long long clip = (long long)blockIdx.x * 128 + row;
    const bool valid = clip < p.B;
    const float* xg = p.xg + (size_t)blockIdx.x * T * N * 128 + row;
    const uint32_t trow = tmem + ((uint32_t)(rg * 32) << 16);
    unsigned char* arow = s_a + uh * 2 * kDATile + (row >> 3) * kDSboA + (row & 7) * 16;
    float h[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) h[j] = 0.f;
    for (int t = 0; t < T; ++t) {
      const float* xt = xg + (size_t)t * N * 128;
      float x[2][24];
      auto fetch = [&](int g, float (&d)[24]) {
        const int u0 = uh * 32 + g * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          d[j] = __ldg(xt + (size_t)(u0 + j) * 128);
          d[8 + j] = __ldg(xt + (size_t)(H + u0 + j) * 128);
          d[16 + j] = __ldg(xt + (size_t)(2 * H + u0 + j) * 128);
        }
      };
      fetch(0, x[0]);                        // in flight while the step's MMAs run
      mbar_wait(acc_done, t & 1);
      tc_fence_after();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        if (g < 3) fetch(g + 1, x[(g + 1) & 1]);
        const int u0 = uh * 32 + g * 8;
        float dr[8], dz[8], dn[8];
        tc_ld8(trow + (uint32_t)u0, dr);
        tc_ld8(trow + (uint32_t)(H + u0), dz);
        tc_ld8(trow + (uint32_t)(2 * H + u0), dn);
        tc_wait_ld();
        float hi[8], lo[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float r = fast_sigmoid(x[g & 1][j] + dr[j] + s_b[u0 + j]);
          const float z = fast_sigmoid(x[g & 1][8 + j] + dz[j] + s_b[H + u0 + j]);
          const float n = fast_tanh(x[g & 1][16 + j] + r * (dn[j] + s_b[2 * H + u0 + j]));
          const float hn = (1.f - z) * n + z * h[g * 8 + j];
          h[g * 8 + j] = hn;
          hi[j] = to_tf32(hn);
          lo[j] = to_tf32(hn - hi[j]);
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          unsigned char* dst = arow + (2 * g + q) * kDLboA;
          *reinterpret_cast<float4*>(dst) = make_float4(hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
          *reinterpret_cast<float4*>(dst + kDATile) = make_float4(lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
        }
        if (valid) {
          if (p.hseq) {
            float4* o = reinterpret_cast<float4*>(p.hseq + (clip * T + t) * H + u0);
            o[0] = make_float4(h[g * 8], h[g * 8 + 1], h[g * 8 + 2], h[g * 8 + 3]);
            o[1] = make_float4(h[g * 8 + 4], h[g * 8 + 5], h[g * 8 + 6], h[g * 8 + 7]);
          }
          if (t == T - 1) {
            float4* o = reinterpret_cast<float4*>(p.hlast + clip * H + u0);
            o[0] = make_float4(h[g * 8], h[g * 8 + 1], h[g * 8 + 2], h[g * 8 + 3]);
            o[1] = make_float4(h[g * 8 + 4], h[g * 8 + 5], h[g * 8 + 6], h[g * 8 + 7]);
          }
        }
      }
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(h_ready);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256));
}

// Any hidden size: thread = (clip, unit) items strided over the CTA, 16 clips per CTA.
constexpr int kGruNBg = 16;
__global__ void __launch_bounds__(256) gru_recurrent_generic_kernel(const float* __restrict__ xp, const float* __restrict__ whh,
                                                                    const float* __restrict__ bhh, float* __restrict__ hseq,
                                                                    float* __restrict__ hlast, long long B, int T, int H) {
  extern __shared__ __align__(16) float sm_gru[];
  float* s_w = sm_gru;                     // [H][3H]
  float* s_h = s_w + (size_t)H * 3 * H;    // [2][NB][H]
  const int tid = threadIdx.x;
  for (int i = tid; i < H * 3 * H; i += 256) s_w[i] = whh[i];
  for (int i = tid; i < 2 * kGruNBg * H; i += 256) s_h[i] = 0.f;
  __syncthreads();
  const long long b0 = (long long)blockIdx.x * kGruNBg;
  const int nitems = kGruNBg * H;           // (clip, unit) pairs of this CTA
  for (int t = 0; t < T; ++t) {
    const float* hc = s_h + (t & 1) * kGruNBg * H;
    float* hn = s_h + ((t + 1) & 1) * kGruNBg * H;
    for (int it = tid; it < nitems; it += 256) {
      const int cb = it / H, j = it - cb * H;
      const long long b = b0 + cb;
      if (b >= B) continue;
      float ar = bhh[j], az = bhh[H + j], an = bhh[2 * H + j];
      const float* hrow = hc + cb * H;
      for (int k = 0; k < H; ++k) {
        const float hv = hrow[k];
        const float* wr = s_w + (size_t)k * 3 * H;
        ar = fmaf(hv, wr[j], ar);
        az = fmaf(hv, wr[H + j], az);
        an = fmaf(hv, wr[2 * H + j], an);
      }
      const float* x = xp + (b * T + t) * 3 * H;
      const float r = 1.f / (1.f + expf(-(x[j] + ar)));
      const float z = 1.f / (1.f + expf(-(x[H + j] + az)));
      const float n = tanhf(x[2 * H + j] + r * an);
      const float hnew = (1.f - z) * n + z * hrow[j];
      hn[cb * H + j] = hnew;
      if (hseq) hseq[(b * T + t) * H + j] = hnew;
      if (t == T - 1) hlast[b * H + j] = hnew;
    }
    __syncthreads();
  }
}

__global__ void crnn_fc_kernel(const float* __restrict__ h, const float* __restrict__ w /*[H][K]*/,
                               const float* __restrict__ bias, float* __restrict__ logits, long long B, int H, int K) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < B * K; i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / K;
    const int k = (int)(i - b * K);
    float acc = bias[k];
    for (int c = 0; c < H; ++c) acc = fmaf(h[b * H + c], w[c * K + k], acc);
    logits[i] = acc;
  }
}


static CrnnWeights* crnn_of(kws_handle_t h) { return reinterpret_cast<CrnnWeights*>(h->crnn); }

static void crnn_free(CrnnWeights* m) {
  if (!m) return;
  void* ptrs[] = {m->d_c1img, m->d_wimg[0], m->d_wimg[1], m->d_whimg[0], m->d_whimg[1], m->d_conv_w, m->d_conv_b, m->d_wih[0], m->d_wih[1], m->d_whh[0], m->d_whh[1], m->d_bih[0],
                  m->d_bih[1],  m->d_bhh[0], m->d_bhh[1], m->d_fc_w,   m->d_fc_b,   m->d_ws};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  delete m;
}

static int crnn_slot(const CrnnWeights& m, const std::string& n) {
  for (size_t i = 0; i < m.names.size(); ++i)
    if (m.names[i] == n) return (int)i;
  return -1;
}

}  // namespace kws

using namespace kws;

extern "C" {

void kws_crnn_free_internal(kws_handle_t h) {
  if (h && h->crnn) {
    crnn_free(crnn_of(h));
    h->crnn = nullptr;
  }
}

int kws_crnn_create(kws_handle_t h, int num_classes, int conv_channels, int hidden, int num_layers, int n_features) {
  if (!h) return fail(KWS_ERR_INVALID, "null handle");
  if (num_classes < 1 || conv_channels < 4 || conv_channels % 4 != 0 || hidden < 1 || hidden > 128 || num_layers < 1 ||
      num_layers > 2 || n_features < 6)
    return fail(KWS_ERR_INVALID, "crnn: need conv_channels a multiple of 4, 1<=hidden<=128, 1<=num_layers<=2, n_features>=6");
  KWS_CUDA(cudaSetDevice(h->device));
  kws_crnn_free_internal(h);
  CrnnWeights* m = new (std::nothrow) CrnnWeights();
  if (!m) return fail(KWS_ERR_ALLOC, "out of host memory");
  m->num_classes = num_classes;
  m->cc = conv_channels;
  m->hidden = hidden;
  m->layers = num_layers;
  m->w1 = (n_features + 4 - 10) / 2 + 1;
  m->names = {"conv1.weight", "conv1.bias"};
  for (int l = 0; l < num_layers; ++l) {
    const std::string s = "_l" + std::to_string(l);
    m->names.push_back("gru.weight_ih" + s);
    m->names.push_back("gru.weight_hh" + s);
    m->names.push_back("gru.bias_ih" + s);
    m->names.push_back("gru.bias_hh" + s);
  }
  m->names.push_back("fc.weight");
  m->names.push_back("fc.bias");
  m->host.assign(m->names.size(), std::vector<float>());
  m->loaded.assign(m->names.size(), false);
  m->created = true;
  h->crnn = m;
  return KWS_OK;
}

int kws_crnn_load_tensor(kws_handle_t h, const char* name, const float* data, const int64_t* shape, int ndim) {
  if (!h || !h->crnn) return fail(KWS_ERR_STATE, "crnn not created");
  if (!name || !data || !shape) return fail(KWS_ERR_INVALID, "null argument");
  CrnnWeights& m = *crnn_of(h);
  const int s = crnn_slot(m, name);
  if (s < 0) return fail(KWS_ERR_INVALID, "unexpected key '%s' in state_dict", name);
  const int64_t H = m.hidden, I0 = (int64_t)m.cc * m.w1;
  const std::string n(name);
  std::vector<int64_t> want;
  if (n == "conv1.weight") want = {m.cc, 1, 10, 10};
  else if (n == "conv1.bias") want = {m.cc};
  else if (n == "fc.weight") want = {m.num_classes, H};
  else if (n == "fc.bias") want = {m.num_classes};
  else if (n.rfind("gru.weight_ih", 0) == 0) want = {3 * H, n.back() == '0' ? I0 : H};
  else if (n.rfind("gru.weight_hh", 0) == 0) want = {3 * H, H};
  else want = {3 * H};
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

int kws_crnn_finalize(kws_handle_t h) {
  if (!h || !h->crnn) return fail(KWS_ERR_STATE, "crnn not created");
  CrnnWeights& m = *crnn_of(h);
  for (size_t i = 0; i < m.names.size(); ++i)
    if (!m.loaded[i]) return fail(KWS_ERR_STATE, "missing key '%s' in state_dict", m.names[i].c_str());
  KWS_CUDA(cudaSetDevice(h->device));
  auto up = [&](float** dst, const std::vector<float>& v) -> int {
    if (*dst) cudaFree(*dst);
    *dst = nullptr;
    KWS_CUDA(cudaMalloc((void**)dst, v.size() * 4));
    KWS_CUDA(cudaMemcpy(*dst, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
    return KWS_OK;
  };
  int rc;
  const int Cc = m.cc, H = m.hidden, K = m.num_classes, W1 = m.w1;
  {
    const std::vector<float>& w = m.host[crnn_slot(m, "conv1.weight")];
    std::vector<float> p((size_t)100 * Cc);
    for (int c = 0; c < Cc; ++c)
      for (int t = 0; t < 100; ++t) p[(size_t)t * Cc + c] = w[(size_t)c * 100 + t];
    if ((rc = up(&m.d_conv_w, p))) return rc;
    if ((rc = up(&m.d_conv_b, m.host[crnn_slot(m, "conv1.bias")]))) return rc;
    if (Cc == 16 && W1 == 18) {
      auto rn = [](float x) {  // cvt.rna.tf32.f32
        uint32_t u;
        memcpy(&u, &x, 4);
        if ((u & 0x7f800000u) != 0x7f800000u) u = (u + 0x1000u) & ~0x1fffu;
        memcpy(&x, &u, 4);
        return x;
      };
      std::vector<float> img(kC1WBytes / 4, 0.f);
      for (int dy = 0; dy < 10; ++dy)
        for (int n = 0; n < 64; ++n)        // n = wl * 16 + c: column wl of the group sees the window 2 * wl to the right
          for (int j = 0; j < 16; ++j) {
            const int wl = n >> 4, c = n & 15, dx = j - 2 * wl;
            const float v = (dx >= 0 && dx < 10) ? w[(size_t)c * 100 + dy * 10 + dx] : 0.f;
            const float hi = rn(v), lo = rn(v - hi);
            const size_t off = ((size_t)(n >> 3) * 512 + (size_t)(j >> 2) * 128 + (size_t)(n & 7) * 16 + (size_t)(j & 3) * 4) / 4;
            img[(size_t)dy * 2048 + off] = hi;
            img[(size_t)dy * 2048 + 1024 + off] = lo;
          }
      if ((rc = up(&m.d_c1img, img))) return rc;
    }
  }
  for (int l = 0; l < m.layers; ++l) {
    const std::string s = "_l" + std::to_string(l);
    const int I = l == 0 ? Cc * W1 : H;
    const std::vector<float>& wih = m.host[crnn_slot(m, "gru.weight_ih" + s)];  // [3H][I], I index = c*W1 + w (layer 0)
    std::vector<float> p((size_t)I * 3 * H);
    for (int g = 0; g < 3 * H; ++g)
      for (int i = 0; i < I; ++i) {
        int dst = i;  // our conv output is NHWC: feature index w*Cc + c
        if (l == 0) {
          const int c = i / W1, w = i - c * W1;
          dst = w * Cc + c;
        }
        p[(size_t)dst * 3 * H + g] = wih[(size_t)g * I + i];
      }
    if ((rc = up(&m.d_wih[l], p))) return rc;
    // tf32 hi/lo weight tiles for the tensor-core kernels, from a k-major [K][N] matrix
    auto wimg_of = [](const std::vector<float>& km, int Kd, int N) {
      const size_t tile = (size_t)(N / 8) * 1024 / 4;   // floats per tile
      std::vector<float> img((size_t)(Kd / kDK) * 2 * tile, 0.f);
      auto rn = [](float x) {  // cvt.rna.tf32.f32
        uint32_t u;
        memcpy(&u, &x, 4);
        if ((u & 0x7f800000u) != 0x7f800000u) u = (u + 0x1000u) & ~0x1fffu;
        memcpy(&x, &u, 4);
        return x;
      };
      for (int n = 0; n < N; ++n)
        for (int k = 0; k < Kd; ++k) {
          const float v = km[(size_t)k * N + n], hi = rn(v), lo = rn(v - hi);
          const int kc = k / kDK, kk = k % kDK;
          const size_t off = ((size_t)(n >> 3) * 1024 + (size_t)(kk >> 2) * 128 + (size_t)(n & 7) * 16 + (size_t)(kk & 3) * 4) / 4;
          img[(size_t)kc * 2 * tile + off] = hi;
          img[(size_t)kc * 2 * tile + tile + off] = lo;
        }
      return img;
    };
    if (I % kDK == 0 && (3 * H) % 16 == 0 && 3 * H <= 256)
      if ((rc = up(&m.d_wimg[l], wimg_of(p, I, 3 * H)))) return rc;
    const std::vector<float>& whh = m.host[crnn_slot(m, "gru.weight_hh" + s)];  // [3H][H]
    std::vector<float> q((size_t)H * 3 * H);
    for (int g = 0; g < 3 * H; ++g)
      for (int k = 0; k < H; ++k) q[(size_t)k * 3 * H + g] = whh[(size_t)g * H + k];
    if ((rc = up(&m.d_whh[l], q))) return rc;
    if (H == 64)
      if ((rc = up(&m.d_whimg[l], wimg_of(q, H, 3 * H)))) return rc;
    if ((rc = up(&m.d_bih[l], m.host[crnn_slot(m, "gru.bias_ih" + s)]))) return rc;
    if ((rc = up(&m.d_bhh[l], m.host[crnn_slot(m, "gru.bias_hh" + s)]))) return rc;
  }
  {
    const std::vector<float>& w = m.host[crnn_slot(m, "fc.weight")];  // [K][H]
    std::vector<float> p((size_t)H * K);
    for (int k = 0; k < K; ++k)
      for (int c = 0; c < H; ++c) p[(size_t)c * K + k] = w[(size_t)k * H + c];
    if ((rc = up(&m.d_fc_w, p))) return rc;
    if ((rc = up(&m.d_fc_b, m.host[crnn_slot(m, "fc.bias")]))) return rc;
  }
  m.finalized = true;
  return KWS_OK;
}

int kws_crnn_forward(kws_handle_t h, const float* feats_dev, int64_t B, int64_t T, int64_t F, float* logits_dev,
                     void* stream) {
  if (!h || !h->crnn) return fail(KWS_ERR_STATE, "crnn not created");
  CrnnWeights& m = *crnn_of(h);
  if (!m.finalized) return fail(KWS_ERR_STATE, "crnn weights not finalized");
  if (B < 0 || T < 6 || F < 6) return fail(KWS_ERR_INVALID, "bad B/T/F");
  const int H1 = (int)((T + 4 - 10) / 2 + 1), W1 = (int)((F + 4 - 10) / 2 + 1);
  if (W1 != m.w1) return fail(KWS_ERR_INVALID, "crnn was built for %d conv columns, input gives %d", m.w1, W1);
  if (B == 0) return KWS_OK;
  if (!feats_dev || !logits_dev) return fail(KWS_ERR_INVALID, "null device pointer");
  KWS_CUDA(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int Cc = m.cc, H = m.hidden, K = m.num_classes, I0 = Cc * W1;
  const int64_t chunk = B < 16384 ? B : 16384;
  // workspace per chunk: conv output [chunk, H1, I0], projections [chunk, H1, 3H], hidden sequence, last state
  size_t need = 0;
  auto take = [&](size_t floats) { size_t o = need; need += (floats * 4 + 255) & ~(size_t)255; return o; };
  const size_t o_conv = take((size_t)((chunk + 127) / 128 * 128) * H1 * I0), o_xp = take((size_t)((chunk + 127) / 128 * 128) * H1 * 3 * H),
               o_seq = take((size_t)chunk * H1 * H), o_last = take((size_t)chunk * H);
  if (need > m.ws_bytes) {
    if (m.d_ws) KWS_CUDA(cudaFree(m.d_ws));
    m.d_ws = nullptr;
    m.ws_bytes = 0;
    KWS_CUDA(cudaMalloc(&m.d_ws, need));
    m.ws_bytes = need;
  }
  char* ws = (char*)m.d_ws;
  float *d_conv = (float*)(ws + o_conv), *d_xp = (float*)(ws + o_xp), *d_seq = (float*)(ws + o_seq),
        *d_last = (float*)(ws + o_last);
  const int gnb = H == 64 ? kGruNB : kGruNBg;
  const size_t gru_smem = ((size_t)H * 3 * H + 2 * (size_t)gnb * H) * 4;
  if (H == 64) {
    if (int rc_ = ensure_dyn_smem(gru_recurrent_kernel, (size_t)gru_smem)) return rc_;
  } else {
    if (int rc_ = ensure_dyn_smem(gru_recurrent_generic_kernel, (size_t)gru_smem)) return rc_;
  }
  for (int64_t b0 = 0; b0 < B; b0 += chunk) {
    const int64_t nb = (B - b0 < chunk) ? B - b0 : chunk;
    // tensor-core conv1 writes tile-major, which only the step-major tensor-core projection reads
    const bool conv_tc = m.d_c1img && F == 40 && Cc == 16 && H == 64 && m.d_wimg[0] && m.d_whimg[0] &&
                         (reinterpret_cast<uintptr_t>(feats_dev) & 15) == 0;   // 16-byte feature loads
    if (conv_tc) {
      Conv1TcParams cp{feats_dev + b0 * T * F, m.d_c1img, m.d_conv_b, d_conv, nb, (int)T, H1};
      if (int rc_ = ensure_dyn_smem(conv1_tc_kernel, (size_t)kC1Smem)) return rc_;
      const long long nt = (nb + 127) / 128 * H1;
      conv1_tc_kernel<<<(unsigned)(nt < h->sm_count ? nt : h->sm_count), kC1Threads, kC1Smem, st>>>(cp);
      KWS_CHECK_LAUNCH(h);
    } else {
      int rc = launch_conv1_relu(h, feats_dev + b0 * T * F, d_conv, m.d_conv_w, m.d_conv_b, nb, (int)T, (int)F, H1, W1, Cc, st);
      if (rc) return rc;
    }
    const float* x = d_conv;
    int I = I0;
    for (int l = 0; l < m.layers; ++l) {
      const long long M = nb * H1;
      const size_t dsmem = 2 * 2 * (size_t)kDATile + kDBStages * 2 * (size_t)(3 * H / 8) * 1024;
      const bool tc_rec = H == 64 && m.d_wimg[l] && m.d_whimg[l];   // both halves of the layer on the tensor core
      if (m.d_wimg[l] && dsmem <= 227 * 1024) {
        DenseParams dp{x, m.d_wimg[l], m.d_bih[l], d_xp, M, 3 * H, I, (l == 0 && conv_tc) ? 1 : 0, tc_rec ? H1 : 0};
        if (int rc_ = ensure_dyn_smem(dense_gemm_tf32x3_kernel, (size_t)dsmem)) return rc_;
        const long long nt = tc_rec ? (nb + 127) / 128 * H1 : (M + 127) / 128;
        dense_gemm_tf32x3_kernel<<<(unsigned)(nt < h->sm_count ? nt : h->sm_count), (kDProd + 1) * 32, dsmem, st>>>(dp);
      } else if (I % 16 == 0 && (3 * H) % 64 == 0) {
        dim3 grid((unsigned)((M + 127) / 128), (unsigned)(3 * H / 64));
        gru_inproj_tiled_kernel<<<grid, 256, 0, st>>>(x, m.d_wih[l], m.d_bih[l], d_xp, M, 3 * H, I);
      } else {
        dim3 grid((unsigned)((M + 63) / 64), (unsigned)((3 * H + 63) / 64));
        gru_inproj_kernel<<<grid, 256, 0, st>>>(x, m.d_wih[l], m.d_bih[l], d_xp, M, 3 * H, I);
      }
      KWS_CHECK_LAUNCH(h);
      const bool more = l + 1 < m.layers;
      if (tc_rec) {
        GruTcParams gp{d_xp, m.d_whimg[l], m.d_bhh[l], more ? d_seq : nullptr, d_last, nb, H1};
        if (int rc_ = ensure_dyn_smem(gru_recurrent_tc_kernel, (size_t)kGtSmem)) return rc_;
        gru_recurrent_tc_kernel<<<(unsigned)((nb + 127) / 128), 288, kGtSmem, st>>>(gp);
      } else if (H == 64)
        gru_recurrent_kernel<<<(unsigned)((nb + kGruNB - 1) / kGruNB), 256, gru_smem, st>>>(
            d_xp, m.d_whh[l], m.d_bhh[l], more ? d_seq : nullptr, d_last, nb, H1);
      else
        gru_recurrent_generic_kernel<<<(unsigned)((nb + kGruNBg - 1) / kGruNBg), 256, gru_smem, st>>>(
            d_xp, m.d_whh[l], m.d_bhh[l], more ? d_seq : nullptr, d_last, nb, H1, H);
      KWS_CHECK_LAUNCH(h);
      x = d_seq;
      I = H;
    }
    long long g = (nb * K + 127) / 128;
    crnn_fc_kernel<<<(unsigned)g, 128, 0, st>>>(d_last, m.d_fc_w, m.d_fc_b, logits_dev + b0 * K, nb, H, K);
    KWS_CHECK_LAUNCH(h);
  }
  return KWS_OK;
}

}  // extern "C"
