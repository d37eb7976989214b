// Fused forward of the reference KWS network: stand-alone launch + host-side packing (body in model_fused_body.cuh).
#include "model_fused_body.cuh"

namespace kws {

template <int F, int NPIX, int PACK>
__global__ void __launch_bounds__(kFThreads, 1) dscnn_fused_kernel(const FusedParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  dscnn_fused_body<F, NPIX, PACK, false>(p, smem, blockIdx.x, gridDim.x, p.B, p.feats, p.logits, 0u);
}

// ---- host side ------------------------------------------------------------------------------
static float host_tf32(float x) {  // cvt.rna.tf32.f32: round to nearest, ties away, 10 mantissa bits
  uint32_t u;
  memcpy(&u, &x, 4);
  if ((u & 0x7f800000u) != 0x7f800000u) u = (u + 0x1000u) & ~0x1fffu;
  memcpy(&x, &u, 4);
  return x;
}

// Pack the fused kernel's parameter images from the staged state_dict tensors (torch layout).
int build_fused_model(ModelWeights& m, const std::vector<float>& conv1_w, const std::vector<float>& conv1_b,
                      const std::vector<const float*>& dw_w, const std::vector<const float*>& dw_b,
                      const std::vector<const float*>& pw_w, const std::vector<const float*>& pw_b) {
  std::vector<float> img(kAImgBytes / 4, 0.f);
  auto put = [&](size_t base, int sbo, int mrow, int k, float v) {
    img[(base + (size_t)(mrow >> 3) * sbo + (size_t)(k >> 2) * 128 + (mrow & 7) * 16 + (k & 3) * 4) / 4] = host_tf32(v);
  };
  // a bias enters its GEMM as two tf32 terms (hi + lo = the fp32 value to ~2^-22) against constant-1 K slots
  auto put_bias = [&](size_t base, int sbo, int mrow, int k, float b) {
    const float hi = host_tf32(b);
    put(base, sbo, mrow, k, hi);
    put(base, sbo, mrow, k + 2, b - hi);
  };
  for (int co = 0; co < kFC; ++co) {
    for (int k = 0; k < 100; ++k) put(0, kSboA1, co, k, conv1_w[(size_t)co * 100 + k]);
    put_bias(0, kSboA1, co, 100, conv1_b[co]);
  }
  std::vector<float> dww(4 * 9 * kFC), actb(5 * kFC), border(5 * kFC, 0.f);
  for (int co = 0; co < kFC; ++co) actb[co] = conv1_b[co];
  for (int l = 0; l < 4; ++l) {
    const size_t base = (size_t)8 * kSboA1 + (size_t)l * 8 * kSboAp;
    for (int co = 0; co < kFC; ++co) {
      double be = pw_b[l][co];
      for (int ci = 0; ci < kFC; ++ci) {
        put(base, kSboAp, co, ci, pw_w[l][(size_t)co * kFC + ci]);
        be += (double)pw_w[l][(size_t)co * kFC + ci] * (double)dw_b[l][ci];
      }
      actb[(l + 1) * kFC + co] = (float)be;
      put_bias(base, kSboAp, co, kFC, (float)be);          // interior pixels: pointwise bias + W_pw . depthwise bias
      put_bias(base, kSboAp, co, kFC + 1, pw_b[l][co]);    // padding=1 border pixels: the pointwise bias alone
      border[(l + 1) * kFC + co] = pw_b[l][co] > 0.f ? pw_b[l][co] : 0.f;
      for (int tap = 0; tap < 9; ++tap) dww[((size_t)l * 9 + tap) * kFC + co] = dw_w[l][(size_t)co * 9 + tap];
    }
  }
  auto up = [&](float** dst, const std::vector<float>& v) -> int {
    if (*dst) cudaFree(*dst);
    *dst = nullptr;
    KWS_CUDA(cudaMalloc((void**)dst, v.size() * 4));
    KWS_CUDA(cudaMemcpy(*dst, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
    return KWS_OK;
  };
  int rc;
  if ((rc = up(&m.d_fused_a, img))) return rc;
  if ((rc = up(&m.d_fused_dw, dww))) return rc;
  if ((rc = up(&m.d_fused_actb, actb))) return rc;
  if ((rc = up(&m.d_fused_border, border))) return rc;
  return KWS_OK;
}

template <int F, int NPIX, int PACK>
static int launch_fused_t(kws_handle_t h, const FusedParams& p, cudaStream_t stream) {
  using L = FusedLayout<F, NPIX, PACK>;
  const size_t smem_bytes = L::smem_bytes(p.T);
  if (smem_bytes > 227 * 1024) return 0;
  auto kern = dscnn_fused_kernel<F, NPIX, PACK>;
  if (int rc = ensure_dyn_smem(kern, smem_bytes)) return rc;
  const long long ngroups = (p.B + L::CL - 1) / L::CL;
  if (ngroups > 0x7fffffffLL / 64) return 0;
  long long grid = h->sm_count;
  if (grid > ngroups) grid = ngroups;
  kern<<<(unsigned)grid, kFThreads, smem_bytes, stream>>>(p);
  KWS_CHECK_LAUNCH(h);
  return 1;
}

// 1 = the fused kernel covers this shape (parameters filled), 0 = not
int fill_fused_params(kws_handle_t h, const float* feats_dev, int64_t B, int64_t T, int64_t F, float* logits_dev,
                      FusedParams* pp) {
  const ModelWeights& m = h->model;
  if (m.channels != kFC || m.num_blocks != 4 || m.d_fused_a == nullptr) return 0;
  if (T < 6 || T > 4096 || (F != 40 && F != 10)) return 0;
  FusedParams& p = *pp;
  p = FusedParams{};
  p.feats = feats_dev;
  p.logits = logits_dev;
  p.B = B;
  p.T = (int)T;
  p.H0 = (int)((T + 4 - 10) / 2 + 1);
  p.ncls = m.num_classes;
  p.a_img = m.d_fused_a;
  p.dw_w = m.d_fused_dw;
  p.border = m.d_fused_border;
  p.fc_w = m.d_fc_w;
  p.fc_b = m.d_fc_b;
  return 1;
}

// 1 = handled, 0 = shape/configuration outside the fused kernel (caller uses the layered path), < 0 = error
int launch_dscnn_fused(kws_handle_t h, const float* feats_dev, int64_t B, int64_t T, int64_t F, float* logits_dev,
                       cudaStream_t stream) {
  FusedParams p;
  if (!fill_fused_params(h, feats_dev, B, T, F, logits_dev, &p)) return 0;
  if (F == 40) return launch_fused_t<40, 24, 1>(h, p, stream);
  // 99x10 features (the reference's default): the widest layer is 9 columns, two clips share a slot's 24-column row
  const int rc = launch_fused_t<10, 24, 2>(h, p, stream);
  return rc != 0 ? rc : launch_fused_t<10, 16, 1>(h, p, stream);
}

}  // namespace kws
