// rvcrefine_* - C ABI (include/rvcrefine.h) of the sibling decoder RefineGANGenerator
// (rvc/lib/algorithm/generators/refinegan.py:268-451).  Host side: weight-norm folding, repacking of every convolution
// into the fp32 back-end's image, workspace carving and the launch sequence of forward (:415-462).
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/rvcrefine.h"
#include "common.cuh"

using namespace rvcdec;

namespace rvcdec {
// kernels of refine_kernels.cu
cudaError_t launch_refine_source(const float* f0, double* before, const float* noise, float* har8, uint64_t seed, uint64_t item0,
                                 int B, int T, int upp, int Lp, float sr, float merge_w, cudaStream_t s);
cudaError_t launch_resample(const float* x, const float* taps, float* y, int B, int C, int L_in, int Lp_in, int L_out, int Lp_out,
                            int rate, int width, int ntaps, cudaStream_t s);
cudaError_t launch_upsample_cat(const float* x, const float* skip, float* out, int B, int C, int Cs, int L, int Lp, int rate,
                                int Lp_out, float slope, cudaStream_t s);
cudaError_t launch_adain(const float* x, const float* noise, const float* w, float* res, float* act, float* acc, int mode, int first,
                         float scale, float slope, uint64_t seed, uint64_t ctr0, int B, int C, int L, int Lp, cudaStream_t s);
}  // namespace rvcdec

namespace {

thread_local std::string g_rerr;

int rfail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_rerr = buf;
  return code;
}

#define RCK(expr)                                                                                  \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return rfail(RVCDEC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

inline size_t r_align(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct RGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit RGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) err = cudaSetDevice(dev); else prev = -1;
  }
  ~RGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct RConv {  // fp32 back-end image: wf[tap][ci][n]
  int Cin = 0, N = 0, ntaps = 0, halo = 0;
  int shift[MAX_TAPS] = {0};
  std::vector<float> wf, bias;
  float* d_wf = nullptr;
  float* d_bias = nullptr;
};

struct RMap {
  std::map<std::string, const rvcdec_tensor*> m;
  const rvcdec_tensor* find(const std::string& n) const {
    auto it = m.find(n);
    return it == m.end() ? nullptr : it->second;
  }
};

int r_get_w(const RMap& tm, const std::string& base, int64_t d0, int64_t rest, std::vector<float>& out) {
  out.resize((size_t)d0 * rest);
  if (const rvcdec_tensor* w = tm.find(base + ".weight")) {
    if (w->numel != d0 * rest) return rfail(RVCDEC_ERR_WEIGHTS, "%s.weight: expected %lld elements, got %lld", base.c_str(), (long long)(d0 * rest), (long long)w->numel);
    memcpy(out.data(), w->data, sizeof(float) * out.size());
    return 0;
  }
  const rvcdec_tensor* g = tm.find(base + ".parametrizations.weight.original0");
  const rvcdec_tensor* v = tm.find(base + ".parametrizations.weight.original1");
  if (!g || !v) { g = tm.find(base + ".weight_g"); v = tm.find(base + ".weight_v"); }
  if (!g || !v) return rfail(RVCDEC_ERR_WEIGHTS, "missing weight for %s", base.c_str());
  if (g->numel != d0 || v->numel != d0 * rest) return rfail(RVCDEC_ERR_WEIGHTS, "%s: weight-norm tensors have wrong size", base.c_str());
  for (int64_t i = 0; i < d0; ++i) {  // w = v * g / ||v|| over all dims but 0 (torch weight_norm, refinegan.py:37,305)
    const float* vi = v->data + i * rest;
    double ss = 0.0;
    for (int64_t k = 0; k < rest; ++k) ss += (double)vi[k] * (double)vi[k];
    const float sc = (float)((double)g->data[i] / std::sqrt(ss));
    for (int64_t k = 0; k < rest; ++k) out[(size_t)(i * rest + k)] = vi[k] * sc;
  }
  return 0;
}

int r_get_v(const RMap& tm, const std::string& name, int64_t n, std::vector<float>& out) {
  const rvcdec_tensor* t = tm.find(name);
  if (!t) return rfail(RVCDEC_ERR_WEIGHTS, "missing tensor %s", name.c_str());
  if (t->numel != n) return rfail(RVCDEC_ERR_WEIGHTS, "%s: expected %lld elements, got %lld", name.c_str(), (long long)n, (long long)t->numel);
  out.assign(t->data, t->data + n);
  return 0;
}

// Conv1d weight (Cout, Cin, k), dilation d, "same" padding -> image; Cin_pad >= Cin pads the input channels with zero
// weights (pre_conv reads the 1-channel source out of an 8-channel blocked tensor).
void r_pack(RConv& L, const std::vector<float>& W, const std::vector<float>& b, int Cout, int Cin, int k, int d, int Cin_pad = 0) {
  const int Cp = Cin_pad ? Cin_pad : Cin;
  L.Cin = Cp; L.N = Cout; L.ntaps = k; L.halo = (k - 1) / 2 * d;
  L.wf.assign((size_t)k * Cp * Cout, 0.f);
  for (int j = 0; j < k; ++j) {
    L.shift[j] = (j - (k - 1) / 2) * d;
    for (int ci = 0; ci < Cin; ++ci)
      for (int co = 0; co < Cout; ++co) L.wf[((size_t)j * Cp + ci) * Cout + co] = W[((size_t)co * Cin + ci) * k + j];
  }
  L.bias = b;
  if (L.bias.empty()) L.bias.assign(Cout, 0.f);
}

struct RBlock {  // one branch of a ParallelResBlock: AdaIN, ResBlock(k, dilations 1/3/5), AdaIN
  std::vector<float> w1, w2;
  float *d_w1 = nullptr, *d_w2 = nullptr;
  RConv c1[3], c2[3];
};

struct RStage {
  RConv input_conv;
  RBlock blk[3];
};

}  // namespace

struct rvcrefine_handle {
  rvcrefine_config cfg{};
  int device = 0;
  int nups = 0, upp = 1;
  bool loaded = false;
  float merge_w = 1.f;
  RConv pre_conv, mel_conv;
  std::vector<RConv> down;
  std::vector<std::vector<float>> taps;
  std::vector<float*> d_taps;
  std::vector<int> widths;
  std::vector<float> cond_w, cond_b, post_w;
  float *d_cond_w = nullptr, *d_cond_b = nullptr, *d_post_w = nullptr;
  std::vector<RStage> stage;
  std::vector<void*> allocs;
  int64_t launches = 0;
};

namespace {

template <typename T>
int r_up(rvcrefine_handle* h, const std::vector<T>& v, T** d) {
  *d = nullptr;
  if (v.empty()) return 0;
  void* p = nullptr;
  RCK(cudaMalloc(&p, v.size() * sizeof(T)));
  h->allocs.push_back(p);
  RCK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  *d = (T*)p;
  return 0;
}
int r_up_conv(rvcrefine_handle* h, RConv& L) {
  int rc;
  if ((rc = r_up(h, L.wf, &L.d_wf))) return rc;
  return r_up(h, L.bias, &L.d_bias);
}

// geometry helpers: channels / lengths of the down path (index 0 = full rate) and of the up path
int down_ch(const rvcrefine_handle* h, int i) { return h->cfg.start_channels << i; }
long long down_len(const rvcrefine_handle* h, int i, int T) {
  long long L = (long long)T * h->upp;
  for (int q = 0; q < i; ++q) L /= h->cfg.upsample_rates[h->nups - 1 - q];
  return L;
}
int up_ch(const rvcrefine_handle* h, int i) { return h->cfg.upsample_initial_channel >> (i + 1); }
long long up_len(const rvcrefine_handle* h, int i, int T) {
  long long L = T;
  for (int q = 0; q <= i; ++q) L *= h->cfg.upsample_rates[q];
  return L;
}

}  // namespace

namespace {

struct RWs {
  double* before;
  float *har8, *rs, *melb, *x512, *condB, *cat, *x0, *ra, *a0, *a1, *xr, *acc;
  std::vector<float*> downs;
  size_t bytes;
};

RWs r_carve(const rvcrefine_handle* h, void* base, int B, int T) {
  RWs w{};
  size_t off = 0;
  auto take = [&](size_t n) { size_t o = off; off += r_align(n, 256); return base ? (void*)((uint8_t*)base + o) : (void*)nullptr; };
  const rvcrefine_config& c = h->cfg;
  const long long n = (long long)T * h->upp;
  w.before = (double*)take((size_t)B * (T + 1) * sizeof(double));
  w.har8 = (float*)take((size_t)B * 8 * padded_rows((int)n) * 4);
  size_t rs_max = 0;
  for (int i = 0; i < h->nups; ++i) {
    w.downs.push_back((float*)take((size_t)B * down_ch(h, i) * padded_rows((int)down_len(h, i, T)) * 4));
    rs_max = std::max(rs_max, (size_t)down_ch(h, i) * padded_rows((int)down_len(h, i + 1, T)));
  }
  w.rs = (float*)take((size_t)B * rs_max * 4);
  w.melb = (float*)take((size_t)B * c.num_mels * padded_rows(T) * 4);
  w.x512 = (float*)take((size_t)B * c.upsample_initial_channel * padded_rows(T) * 4);
  w.condB = (float*)take((size_t)B * (c.upsample_initial_channel / 2) * 4);
  size_t cat_max = 0, st_max = 0;
  for (int i = 0; i < h->nups; ++i) {
    const size_t Lp = padded_rows((int)up_len(h, i, T));
    const int cin = 2 * up_ch(h, i) + down_ch(h, h->nups - 1 - i);
    cat_max = std::max(cat_max, (size_t)cin * Lp);
    st_max = std::max(st_max, (size_t)up_ch(h, i) * Lp);
  }
  w.cat = (float*)take((size_t)B * cat_max * 4);
  w.x0 = (float*)take((size_t)B * st_max * 4);
  w.ra = (float*)take((size_t)B * st_max * 4);
  w.a0 = (float*)take((size_t)B * st_max * 4);
  w.a1 = (float*)take((size_t)B * st_max * 4);
  w.xr = (float*)take((size_t)B * st_max * 4);
  w.acc = (float*)take((size_t)B * st_max * 4);
  w.bytes = off;
  return w;
}

int r_conv(rvcrefine_handle* h, const RConv& L, const float* in, int B, int Lrows, const Epilogue& e, cudaStream_t s) {
  ConvParams p{};
  p.act_in = in; p.act_fmt_in = ACT_F32; p.w = L.d_wf;
  p.B = B; p.Cin = L.Cin; p.L_in = Lrows; p.Lp_in = padded_rows(Lrows);
  p.N_total = L.N; p.ntaps = L.ntaps; p.halo = L.halo;
  for (int t = 0; t < L.ntaps; ++t) p.shift[t] = L.shift[t];
  const int ntiles = (L.N + 63) / 64;
  if (ntiles > MAX_NTILES) return rfail(RVCDEC_ERR_ARG, "convolution with %d columns exceeds the back-end's tap masks", L.N);
  for (int i = 0; i < ntiles; ++i) p.tapmask[i] = (1u << L.ntaps) - 1u;
  p.epi = e;
  RCK(launch_conv_f32(p, s));
  ++h->launches;
  return 0;
}

Epilogue r_epi(const float* bias, int cout_phys, int Lrows) {
  Epilogue e{};
  e.bias = bias; e.scale = 1.f; e.slope = 0.2f; e.up = 1; e.cout = cout_phys; e.L_out = Lrows; e.Lp_out = padded_rows(Lrows);
  e.act_fmt = ACT_F32;
  return e;
}

}  // namespace

extern "C" {

const char* rvcrefine_last_error(void) { return g_rerr.c_str(); }

int rvcrefine_create(const rvcrefine_config* cfg, int device, rvcrefine_handle** out) {
  if (!cfg || !out) return rfail(RVCDEC_ERR_ARG, "null argument");
  *out = nullptr;
  const rvcrefine_config& c = *cfg;
  if (c.num_upsamples < 1 || c.num_upsamples > RVCDEC_MAX_UPS) return rfail(RVCDEC_ERR_ARG, "bad number of upsamples");
  if (c.sample_rate <= 0 || c.num_mels <= 0 || c.num_mels % 8 || c.start_channels % 8 || c.start_channels <= 0 || c.gin_channels < 0)
    return rfail(RVCDEC_ERR_ARG, "num_mels and start_channels must be positive multiples of 8");
  if ((c.start_channels << c.num_upsamples) != c.upsample_initial_channel / 2)
    return rfail(RVCDEC_ERR_ARG, "start_channels * 2^n must equal upsample_initial_channel / 2 (the two halves concatenated at refinegan.py:442)");
  if ((c.upsample_initial_channel >> c.num_upsamples) < 8 || (c.upsample_initial_channel >> c.num_upsamples) % 8)
    return rfail(RVCDEC_ERR_ARG, "the last stage must keep a multiple of 8 channels");
  auto h = new rvcrefine_handle();
  h->cfg = c; h->device = device; h->nups = c.num_upsamples; h->upp = 1;
  for (int i = 0; i < h->nups; ++i) {
    const int u = c.upsample_rates[i];
    if (u < 2 || u % 2) { delete h; return rfail(RVCDEC_ERR_ARG, "upsample rate %d unsupported (even rates: the source's half-frame pieces)", u); }
    h->upp *= u;
  }
  *out = h;
  return 0;
}

void rvcrefine_destroy(rvcrefine_handle* h) {
  if (!h) return;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) == cudaSuccess && h->device < ndev) {
    RGuard g(h->device);
    for (void* p : h->allocs) cudaFree(p);
  }
  delete h;
}

int rvcrefine_load_weights(rvcrefine_handle* h, const rvcdec_tensor* tensors, int32_t n) {
  if (!h || !tensors) return rfail(RVCDEC_ERR_ARG, "null argument");
  const rvcrefine_config& c = h->cfg;
  RMap tm;
  for (int i = 0; i < n; ++i) tm.m[tensors[i].name ? tensors[i].name : ""] = &tensors[i];
  int rc;
  std::vector<float> W, b;
  if ((rc = r_get_v(tm, "m_source.merge.0.weight", 1, W))) return rc;  // Linear(1, 1, bias=False), refinegan.py:213-216
  h->merge_w = W[0];
  if ((rc = r_get_w(tm, "pre_conv", c.start_channels, 7, W)) || (rc = r_get_v(tm, "pre_conv.bias", c.start_channels, b))) return rc;
  r_pack(h->pre_conv, W, b, c.start_channels, 1, 7, 1, 8);
  h->down.assign(h->nups, RConv());
  h->taps.assign(h->nups, std::vector<float>());
  h->widths.assign(h->nups, 0);
  for (int i = 0; i < h->nups; ++i) {
    const int ci = down_ch(h, i), co = 2 * ci;
    const std::string nm = "downsample_blocks." + std::to_string(i);
    if ((rc = r_get_w(tm, nm, co, (int64_t)ci * 7, W)) || (rc = r_get_v(tm, nm + ".bias", co, b))) return rc;
    r_pack(h->down[i], W, b, co, ci, 7, 1);
    const rvcdec_tensor* t = tm.find("resample_taps." + std::to_string(i));
    const int rate = c.upsample_rates[h->nups - 1 - i];
    if (!t || t->numel <= rate || (t->numel - rate) % 2) return rfail(RVCDEC_ERR_WEIGHTS, "resample_taps.%d missing or not 2*width + %d taps", i, rate);
    h->taps[i].assign(t->data, t->data + t->numel);
    h->widths[i] = (int)((t->numel - rate) / 2);
  }
  const int C0 = c.upsample_initial_channel;
  if ((rc = r_get_w(tm, "mel_conv", C0 / 2, (int64_t)c.num_mels * 7, W)) || (rc = r_get_v(tm, "mel_conv.bias", C0 / 2, b))) return rc;
  r_pack(h->mel_conv, W, b, C0 / 2, c.num_mels, 7, 1);
  h->cond_w.clear(); h->cond_b.clear();
  if (c.gin_channels) {
    if ((rc = r_get_w(tm, "cond", C0 / 2, c.gin_channels, h->cond_w)) || (rc = r_get_v(tm, "cond.bias", C0 / 2, h->cond_b))) return rc;
  }
  h->stage.assign(h->nups, RStage());
  const int ks[3] = {3, 7, 11}, ds[3] = {1, 3, 5};
  for (int i = 0; i < h->nups; ++i) {
    RStage& S = h->stage[i];
    const int co = up_ch(h, i), cin = 2 * co + down_ch(h, h->nups - 1 - i);
    const std::string sp = "upsample_conv_blocks." + std::to_string(i);
    if ((rc = r_get_w(tm, sp + ".input_conv", co, (int64_t)cin * 7, W)) || (rc = r_get_v(tm, sp + ".input_conv.bias", co, b))) return rc;
    r_pack(S.input_conv, W, b, co, cin, 7, 1);
    for (int j = 0; j < 3; ++j) {
      RBlock& Bk = S.blk[j];
      const std::string bp = sp + ".blocks." + std::to_string(j);
      if ((rc = r_get_v(tm, bp + ".0.weight", co, Bk.w1)) || (rc = r_get_v(tm, bp + ".2.weight", co, Bk.w2))) return rc;
      for (int m = 0; m < 3; ++m) {
        const std::string c1 = bp + ".1.convs1." + std::to_string(m), c2 = bp + ".1.convs2." + std::to_string(m);
        if ((rc = r_get_w(tm, c1, co, (int64_t)co * ks[j], W)) || (rc = r_get_v(tm, c1 + ".bias", co, b))) return rc;
        r_pack(Bk.c1[m], W, b, co, co, ks[j], ds[m]);
        if (Bk.c1[m].halo > PADL) return rfail(RVCDEC_ERR_ARG, "ResBlock halo exceeds %d rows", PADL);
        if ((rc = r_get_w(tm, c2, co, (int64_t)co * ks[j], W)) || (rc = r_get_v(tm, c2 + ".bias", co, b))) return rc;
        r_pack(Bk.c2[m], W, b, co, co, ks[j], 1);
      }
    }
  }
  const int Cl = up_ch(h, h->nups - 1);
  if ((rc = r_get_w(tm, "conv_post", 1, (int64_t)Cl * 7, W))) return rc;  // bias=False, refinegan.py:376-379
  h->post_w.resize((size_t)7 * Cl);
  for (int cc = 0; cc < Cl; ++cc)
    for (int j = 0; j < 7; ++j) h->post_w[(size_t)j * Cl + cc] = W[(size_t)cc * 7 + j];
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev == 0) { h->loaded = false; return rfail(RVCDEC_ERR_CUDA, "no CUDA device: %s", cudaGetErrorString(ce)); }
  RGuard g(h->device);
  RCK(g.err);
  for (void* p : h->allocs) cudaFree(p);
  h->allocs.clear();
  if ((rc = r_up_conv(h, h->pre_conv)) || (rc = r_up_conv(h, h->mel_conv)) || (rc = r_up(h, h->cond_w, &h->d_cond_w)) ||
      (rc = r_up(h, h->cond_b, &h->d_cond_b)) || (rc = r_up(h, h->post_w, &h->d_post_w)))
    return rc;
  h->d_taps.assign(h->nups, nullptr);
  for (int i = 0; i < h->nups; ++i)
    if ((rc = r_up_conv(h, h->down[i])) || (rc = r_up(h, h->taps[i], &h->d_taps[i]))) return rc;
  for (auto& S : h->stage) {
    if ((rc = r_up_conv(h, S.input_conv))) return rc;
    for (auto& Bk : S.blk) {
      if ((rc = r_up(h, Bk.w1, &Bk.d_w1)) || (rc = r_up(h, Bk.w2, &Bk.d_w2))) return rc;
      for (int m = 0; m < 3; ++m)
        if ((rc = r_up_conv(h, Bk.c1[m])) || (rc = r_up_conv(h, Bk.c2[m]))) return rc;
    }
  }
  h->loaded = true;
  return 0;
}

size_t rvcrefine_workspace_bytes(const rvcrefine_handle* h, int32_t B, int32_t T) {
  if (!h || B < 1 || T < 1) return 0;
  return r_carve(h, nullptr, B, T).bytes;
}

int64_t rvcrefine_noise_floats(const rvcrefine_handle* h, int32_t B, int32_t T) {
  if (!h || B < 1 || T < 1) return 0;
  int64_t n = (int64_t)B * T * h->upp;
  for (int i = 0; i < h->nups; ++i) n += 6 * (int64_t)B * up_ch(h, i) * up_len(h, i, T);
  return n;
}

int64_t rvcrefine_last_launch_count(const rvcrefine_handle* h) { return h ? h->launches : 0; }

}  // extern "C"

extern "C" int rvcrefine_forward(rvcrefine_handle* h, const float* mel, const float* f0, const float* g, const float* noise,
                                 uint64_t seed, float* out, void* workspace, size_t workspace_bytes, int32_t B, int32_t T, void* stream) {
  if (!h || !mel || !f0 || !out || !workspace) return rfail(RVCDEC_ERR_ARG, "null argument");
  if (!h->loaded) return rfail(RVCDEC_ERR_STATE, "rvcrefine_load_weights has not succeeded on this handle");
  if (B < 1 || T < 1) return rfail(RVCDEC_ERR_ARG, "B and T must be >= 1");
  if ((long long)T * h->upp > 0x7fffffffLL / 16) return rfail(RVCDEC_ERR_ARG, "T too large");
  const size_t need = rvcrefine_workspace_bytes(h, B, T);
  if (workspace_bytes < need) return rfail(RVCDEC_ERR_WORKSPACE, "workspace of %zu bytes is smaller than the %zu needed for B=%d, T=%d", workspace_bytes, need, B, T);
  RGuard guard(h->device);
  RCK(guard.err);
  cudaStream_t s = (cudaStream_t)stream;
  const rvcrefine_config& c = h->cfg;
  RWs w = r_carve(h, workspace, B, T);
  const int n = T * h->upp, Lpn = padded_rows(n), LpT = padded_rows(T);
  const float slope = 0.2f;
  h->launches = 0;
  int rc;
  // ---- harmonic source on the interpolated f0 (refinegan.py:417-420), then pre_conv (:421) ----
  RCK(launch_zero_pads(w.har8, nullptr, nullptr, 4, B, 8, n, Lpn, s)); ++h->launches;
  RCK(launch_refine_source(f0, w.before, noise, w.har8, seed, 0, B, T, h->upp, Lpn, (float)c.sample_rate, h->merge_w, s)); h->launches += 2;
  {
    Epilogue e = r_epi(h->pre_conv.d_bias, c.start_channels, n);
    e.out_act = w.downs[0];  // downs[0] = leaky_relu(pre_conv(har), 0.2)  (:425-426)
    if ((rc = r_conv(h, h->pre_conv, w.har8, B, n, e, s))) return rc;
  }
  // ---- downsample path (:424-436): resample the activated tensor, weight-normed k = 7 conv ----
  const int C0 = c.upsample_initial_channel;
  for (int i = 0; i < h->nups; ++i) {
    const int ci = down_ch(h, i), rate = c.upsample_rates[h->nups - 1 - i];
    const int Li = (int)down_len(h, i, T), Lo = (int)down_len(h, i + 1, T);
    RCK(launch_zero_pads(w.rs, nullptr, nullptr, 4, B, ci, Lo, padded_rows(Lo), s)); ++h->launches;
    RCK(launch_resample(w.downs[i], h->d_taps[i], w.rs, B, ci, Li, padded_rows(Li), Lo, padded_rows(Lo), rate, h->widths[i],
                        (int)h->taps[i].size(), s)); ++h->launches;
    if (i < h->nups - 1) {
      Epilogue e = r_epi(h->down[i].d_bias, 2 * ci, Lo);
      e.out_act = w.downs[i + 1];
      if ((rc = r_conv(h, h->down[i], w.rs, B, Lo, e, s))) return rc;
    } else {  // the last block's output is the upper half of the concatenation with the mel branch (:442), not activated yet
      Epilogue e = r_epi(h->down[i].d_bias, C0, Lo);
      e.out_f32 = w.x512 + (size_t)(C0 / 2 / 8) * LpT * 8;
      if ((rc = r_conv(h, h->down[i], w.rs, B, Lo, e, s))) return rc;
    }
  }
  // ---- mel branch: mel_conv + cond(g) (:438-441) into the lower half ----
  RCK(launch_zero_pads(w.melb, nullptr, nullptr, 4, B, c.num_mels, T, LpT, s)); ++h->launches;
  RCK(launch_pack_input(mel, w.melb, ACT_F32, B, c.num_mels, T, LpT, s)); ++h->launches;
  RCK(launch_cond((c.gin_channels && g) ? g : nullptr, h->d_cond_w, h->d_cond_b, h->mel_conv.d_bias, w.condB, B, c.gin_channels, C0 / 2, s)); ++h->launches;
  {
    Epilogue e = r_epi(w.condB, C0, T);
    e.bias_bstride = C0 / 2;
    e.out_f32 = w.x512;
    if ((rc = r_conv(h, h->mel_conv, w.melb, B, T, e, s))) return rc;
  }
  // ---- upsample path (:444-455) ----
  const float* x_prev = w.x512;
  int C_prev = C0, L_prev = T;
  const float* nz = noise ? noise + (size_t)B * n : nullptr;
  uint64_t region = 1;
  for (int i = 0; i < h->nups; ++i) {
    const RStage& S = h->stage[i];
    const int rate = c.upsample_rates[i], co = up_ch(h, i), cs = down_ch(h, h->nups - 1 - i);
    const int L = L_prev * rate, Lp = padded_rows(L);
    const size_t nblk = (size_t)B * co * L;
    // x = cat([Upsample(leaky_relu(x)), skip])
    RCK(launch_zero_pads(w.cat, nullptr, nullptr, 4, B, C_prev + cs, L, Lp, s)); ++h->launches;
    RCK(launch_upsample_cat(x_prev, w.downs[h->nups - 1 - i], w.cat, B, C_prev, cs, L_prev, padded_rows(L_prev), rate, Lp, slope, s)); ++h->launches;
    {  // ParallelResBlock.input_conv (:137-144)
      Epilogue e = r_epi(S.input_conv.d_bias, co, L);
      e.out_f32 = w.x0;
      if ((rc = r_conv(h, S.input_conv, w.cat, B, L, e, s))) return rc;
    }
    RCK(launch_zero_pads(w.a0, w.a1, nullptr, 4, B, co, L, Lp, s)); ++h->launches;
    for (int j = 0; j < 3; ++j) {
      const RBlock& Bk = S.blk[j];
      // AdaIN in front: ra = lrelu(x0 + noise * w), a0 = lrelu(ra)  (:98-100, :79)
      RCK(launch_adain(w.x0, nz, Bk.d_w1, w.ra, w.a0, nullptr, 0, 0, 1.f, slope, seed, region << 40, B, co, L, Lp, s)); ++h->launches;
      if (nz) nz += nblk;
      ++region;
      for (int m = 0; m < 3; ++m) {  // ResBlock (:77-86)
        {
          Epilogue e = r_epi(Bk.c1[m].d_bias, co, L);
          e.out_act = w.a1;
          if ((rc = r_conv(h, Bk.c1[m], w.a0, B, L, e, s))) return rc;
        }
        {
          Epilogue e = r_epi(Bk.c2[m].d_bias, co, L);
          e.res = (m == 0) ? w.ra : w.xr;
          e.out_f32 = w.xr;
          if (m < 2) e.out_act = w.a0;
          if ((rc = r_conv(h, Bk.c2[m], w.a1, B, L, e, s))) return rc;
        }
      }
      // AdaIN behind + mean over the three branches (:158)
      RCK(launch_adain(w.xr, nz, Bk.d_w2, nullptr, nullptr, w.acc, 1, j == 0 ? 1 : 0, 1.f / 3.f, slope, seed, region << 40, B, co, L, Lp, s)); ++h->launches;
      if (nz) nz += nblk;
      ++region;
    }
    x_prev = w.acc; C_prev = co; L_prev = L;
  }
  // ---- tail (:457-460): leaky_relu(0.2), conv_post (no bias), tanh ----
  RCK(launch_post(w.acc, h->d_post_w, out, B, C_prev, n, Lpn, 7, slope, s)); ++h->launches;
  return 0;
}
