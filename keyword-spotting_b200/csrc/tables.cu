// Host-side table construction for the feature path (float64 on the host, rounded once to
// float32 for the device).  Follows the published algorithm of python_speech_features 0.6
// (base.get_filterbanks, base.lifter, scipy.fftpack.dct(type=2, norm='ortho')) that the
// reference calls at kws/libs/audio_processor.py:270-278; SURVEY.md Appendix A steps 6, 8, 9.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "common.cuh"

namespace kws {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

const char* last_error() { return g_err; }

static double hz2mel(double hz) { return 2595.0 * std::log10(1.0 + hz / 700.0); }
static double mel2hz(double mel) { return 700.0 * (std::pow(10.0, mel / 2595.0) - 1.0); }

// numpy.linspace(start, stop, num): arange(num) * step + start, last point forced to stop.
static void linspace(double start, double stop, int num, std::vector<double>& y) {
  y.resize(num);
  double step = (num > 1) ? (stop - start) / (num - 1) : 0.0;
  for (int i = 0; i < num; ++i) y[i] = i * step + start;
  if (num > 1) y[num - 1] = stop;
}

static int mel_edges(int nfilt, int nfft, int sr, double lowfreq, double highfreq,
                     std::vector<double>& edges) {
  if (nfilt < 1 || nfft < 2 || sr < 1) return fail(KWS_ERR_INVALID, "mel: bad nfilt/nfft/sample_rate");
  if (highfreq <= 0) highfreq = sr / 2.0;
  if (highfreq > sr / 2.0) return fail(KWS_ERR_INVALID, "highfreq is greater than samplerate/2");
  std::vector<double> mp;
  linspace(hz2mel(lowfreq), hz2mel(highfreq), nfilt + 2, mp);
  edges.resize(nfilt + 2);
  for (int i = 0; i < nfilt + 2; ++i) edges[i] = std::floor((nfft + 1) * mel2hz(mp[i]) / sr);
  return KWS_OK;
}

static void free_dev(void* p) {
  if (p) cudaFree(p);
}

void free_feature_tables(FeatureTables& t) {
  free_dev(t.d_window);
  free_dev(t.d_tw256);
  free_dev(t.d_tw512);
  free_dev(t.d_melw);
  free_dev(t.d_segedge);
  free_dev(t.d_dct);
  free_dev(t.d_dctB);
  t = FeatureTables();
}

template <typename T>
static int upload(T** dst, const std::vector<T>& src) {
  KWS_CUDA(cudaMalloc((void**)dst, src.size() * sizeof(T)));
  KWS_CUDA(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  return KWS_OK;
}

int build_feature_tables(kws_handle_t h, const kws_feature_config* cfg) {
  if (!cfg) return fail(KWS_ERR_INVALID, "null feature config");
  if (cfg->nfft != kNfft)
    return fail(KWS_ERR_INVALID, "unsupported nfft=%d: the sm_100a feature kernel is a 512-point FFT",
                cfg->nfft);
  if (cfg->frame_len < 1 || cfg->frame_step < 1)
    return fail(KWS_ERR_INVALID, "frame_len and frame_step must be >= 1");
  if (cfg->nfilt < 1 || cfg->nfilt > kMaxFilt)
    return fail(KWS_ERR_INVALID, "nfilt=%d outside [1,%d]", cfg->nfilt, kMaxFilt);
  if (cfg->numcep < 1) return fail(KWS_ERR_INVALID, "numcep must be >= 1");
  std::vector<double> edges;
  int rc = mel_edges(cfg->nfilt, cfg->nfft, cfg->sample_rate, cfg->lowfreq, cfg->highfreq, edges);
  if (rc) return rc;

  free_feature_tables(h->feat);
  FeatureTables& t = h->feat;
  t.cfg = *cfg;
  t.cfg.window_host = nullptr;
  t.ncep = cfg->numcep < cfg->nfilt ? cfg->numcep : cfg->nfilt;
  t.ncep_pad = (t.ncep + 3) & ~3;
  t.load_len = cfg->frame_len < kNfft ? cfg->frame_len : kNfft;
  const int M = cfg->nfilt;

  std::vector<float2> tw256(256), tw512(256);
  for (int i = 0; i < 256; ++i) {
    double a = 2.0 * M_PI * i / 256.0, b = 2.0 * M_PI * i / 512.0;
    tw256[i] = make_float2((float)std::cos(a), (float)-std::sin(a));
    tw512[i] = make_float2((float)std::cos(b), (float)-std::sin(b));
  }
  // per-bin (rising, falling) weights of the segment [e_s, e_{s+1}) the bin lies in:
  // rising belongs to filter s, falling to filter s-1 (psf get_filterbanks' two loops).
  std::vector<float2> melw(kBins, make_float2(0.f, 0.f));
  std::vector<int> iedges(M + 2);
  for (int s = 0; s < M + 2; ++s) {
    double e = edges[s];
    if (e < 0) e = 0;
    if (e > kBins) e = kBins;
    iedges[s] = (int)e;
  }
  for (int s = 0; s <= M; ++s) {
    for (int k = iedges[s]; k < iedges[s + 1] && k < kBins; ++k) {
      double w = edges[s + 1] - edges[s];
      double r = (s < M) ? (k - edges[s]) / w : 0.0;
      double f = (s >= 1) ? (edges[s + 1] - k) / w : 0.0;
      melw[k] = make_float2((float)r, (float)f);
    }
  }
  t.h_edges = iedges;

  std::vector<double> dct((size_t)M * t.ncep);
  rc = kws_dct_lifter_matrix(M, t.ncep, cfg->ceplifter, dct.data());
  if (rc) return rc;
  std::vector<float> dctf((size_t)M * t.ncep_pad, 0.f);
  for (int n = 0; n < M; ++n)
    for (int k = 0; k < t.ncep; ++k) dctf[(size_t)n * t.ncep_pad + k] = (float)dct[(size_t)n * t.ncep + k];

  if ((rc = upload(&t.d_tw256, tw256))) return rc;
  if ((rc = upload(&t.d_tw512, tw512))) return rc;
  if ((rc = upload(&t.d_melw, melw))) return rc;
  if ((rc = upload(&t.d_segedge, iedges))) return rc;
  if ((rc = upload(&t.d_dct, dctf))) return rc;
  if (M <= 48) {
    // B operand of the tensor-core DCT (mfcc_fast.cu): B[n][k] = lifter*dct[k][n], n = coefficient
    // (48 rows), k = mel index (kpad columns), split into tf32 hi + lo and laid out in the
    // canonical K-major no-swizzle UMMA layout: 8-row groups of kpad/4 core matrices (8 x 16 B).
    auto tf32_rn = [](float v) {
      uint32_t b;
      std::memcpy(&b, &v, 4);
      b = (b + 0x1000u) & 0xffffe000u;  // round to nearest (ties away), like cvt.rna.tf32.f32
      float r;
      std::memcpy(&r, &b, 4);
      return r;
    };
    const int kpad = (M + 7) & ~7;
    const size_t group = (size_t)(kpad / 4) * 32;  // floats per 8-row group
    std::vector<float> img(2 * 6 * group, 0.f);
    for (int n = 0; n < t.ncep; ++n)
      for (int k = 0; k < M; ++k) {
        const float v = (float)dct[(size_t)k * t.ncep + n];
        const float hi = tf32_rn(v), lo = tf32_rn(v - hi);
        const size_t idx = (size_t)(n / 8) * group + (size_t)(k / 4) * 32 + (size_t)(n % 8) * 4 + (k % 4);
        img[idx] = hi;
        img[6 * group + idx] = lo;
      }
    if ((rc = upload(&t.d_dctB, img))) return rc;
    t.dct_kpad = kpad;
  }
  if (cfg->window_host) {
    std::vector<float> w(cfg->window_host, cfg->window_host + cfg->frame_len);
    if ((rc = upload(&t.d_window, w))) return rc;
  }
  t.ready = true;
  return KWS_OK;
}

}  // namespace kws

// ---- exported host helpers --------------------------------------------------------------
extern "C" {

const char* kws_last_error(void) { return kws::last_error(); }
int kws_version(void) { return KWS_B200_VERSION; }

void kws_feature_config_default(kws_feature_config* c) {
  if (!c) return;
  std::memset(c, 0, sizeof(*c));
  c->sample_rate = 16000;
  c->frame_len = 400;
  c->frame_step = 160;
  c->nfft = 512;
  c->nfilt = 26;
  c->numcep = 10;
  c->preemph = 0.97f;
  c->ceplifter = 22;
  c->append_energy = 1;
  c->lowfreq = 0.0;
  c->highfreq = 0.0;
  c->window_host = nullptr;
}

int64_t kws_round_half_up(double v) {
  // decimal ROUND_HALF_UP on the exact binary value: ties go away from zero
  double a = std::fabs(v), r = std::floor(a);
  if (a - r >= 0.5) r += 1.0;
  return (int64_t)(v < 0 ? -r : r);
}

int64_t kws_num_frames(int64_t n, int64_t frame_len, int64_t frame_step) {
  if (frame_len < 1 || frame_step < 1 || n < 0) return -1;
  if (n <= frame_len) return 1;
  return 1 + (n - frame_len + frame_step - 1) / frame_step;
}

int kws_mel_bin_edges(int nfilt, int nfft, int sample_rate, double lowfreq, double highfreq,
                      int32_t* edges_out) {
  if (!edges_out) return kws::fail(KWS_ERR_INVALID, "null output");
  std::vector<double> e;
  int rc = kws::mel_edges(nfilt, nfft, sample_rate, lowfreq, highfreq, e);
  if (rc) return rc;
  for (int i = 0; i < nfilt + 2; ++i) edges_out[i] = (int32_t)e[i];
  return KWS_OK;
}

int kws_mel_filterbank(int nfilt, int nfft, int sample_rate, double lowfreq, double highfreq,
                       double* fb) {
  if (!fb) return kws::fail(KWS_ERR_INVALID, "null output");
  std::vector<double> b;
  int rc = kws::mel_edges(nfilt, nfft, sample_rate, lowfreq, highfreq, b);
  if (rc) return rc;
  const int nb = nfft / 2 + 1;
  std::memset(fb, 0, sizeof(double) * (size_t)nfilt * nb);
  for (int j = 0; j < nfilt; ++j) {
    for (int i = (int)b[j]; i < (int)b[j + 1] && i < nb; ++i)
      fb[(size_t)j * nb + i] = (i - b[j]) / (b[j + 1] - b[j]);
    for (int i = (int)b[j + 1]; i < (int)b[j + 2] && i < nb; ++i)
      fb[(size_t)j * nb + i] = (b[j + 2] - i) / (b[j + 2] - b[j + 1]);
  }
  return KWS_OK;
}

int kws_dct_lifter_matrix(int nfilt, int ncep, int ceplifter, double* out) {
  if (!out || nfilt < 1 || ncep < 1 || ncep > nfilt)
    return kws::fail(KWS_ERR_INVALID, "dct: need 1 <= ncep <= nfilt");
  for (int k = 0; k < ncep; ++k) {
    double scale = (k == 0) ? std::sqrt(1.0 / nfilt) : std::sqrt(2.0 / nfilt);
    double lift = (ceplifter > 0) ? 1.0 + (ceplifter / 2.0) * std::sin(M_PI * k / ceplifter) : 1.0;
    for (int n = 0; n < nfilt; ++n)
      out[(size_t)n * ncep + k] = lift * scale * std::cos(M_PI * k * (2 * n + 1) / (2.0 * nfilt));
  }
  return KWS_OK;
}

}  // extern "C"
