// librvcdec.so - C ABI (include/rvcdec.h) over the sm_100a decoder kernels.
// Host side: weight-norm folding and repacking at load time, workspace carving, and the
// launch sequence of HiFiGANNSFGenerator.forward (reference hifigan_nsf.py:173-212).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/rvcdec.h"
#include "common.cuh"

using namespace rvcdec;

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CK(expr)                                                                                  \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess)                                                                        \
      return fail(RVCDEC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

struct ConvLayer {
  std::string name;
  int Cin = 0, N_total = 0, cout = 0, up = 1, ntaps = 0, halo = 0;
  int shift[MAX_TAPS] = {0};
  double macs_per_row = 0;  // algorithmic MACs per INPUT row (zero-padded polyphase taps not counted)
  std::vector<float> wf;    // [tap][ci][N_total]
  std::vector<float> bias;  // [N_total]
  std::vector<uint32_t> mask_f32;
  // tensor-core image
  int n_tile = 0, kc = 0;
  std::vector<uint32_t> mask_mma;
  std::vector<uint16_t> wp[2];  // [fmt] f16 / bf16
  // 128 -> 128 channel layers only: the image split by output-channel half for CTA pairs (mrf_pair2.cu):
  // [tap][half 2][slice 16][n 64][8], one contiguous 16 KB block per (tap, CTA)
  std::vector<uint16_t> wp2[2];
  uint16_t* d_wp2[2] = {nullptr, nullptr};
  // device copies
  float* d_wf = nullptr;
  float* d_bias = nullptr;
  uint16_t* d_wp[2] = {nullptr, nullptr};
  // conv2 of a ResBlock only: b2[0] + .. + b2[m] (the fused kernel keeps x un-biased in TMEM)
  std::vector<float> cbias;
  float* d_cbias = nullptr;
};

struct NoiseLayer {
  int C = 0, k = 0, stride = 1, pad = 0;
  std::vector<float> w, bias;  // w: [k][C]
  float* d_w = nullptr;
  float* d_bias = nullptr;
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Every entry point runs on the handle's device and leaves the caller's current device as it found it
// (one process may drive several GPUs, one handle each).
struct DeviceGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) err = cudaSetDevice(dev); else prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

struct rvcdec_handle {
  rvcdec_config cfg{};
  int device = 0;
  int num_sms = 148;
  int upp = 1;
  int nups = 0, nk = 0, nd = 0;
  std::vector<int> ch;  // channels after each upsample
  std::vector<int> ups_pad;  // ConvTranspose1d padding of each upsample (hifigan_nsf.py:113-117)
  bool loaded = false;
  bool fast_ok = false;
  std::string fast_why;
  ConvLayer conv_pre;
  std::vector<ConvLayer> ups;
  std::vector<ConvLayer> ups_zs;  // stride-2 upsamples as ordinary convs over a zero-stuffed input (n_tile == 0: unused)
  std::vector<ConvLayer> ups_p2;  // stride-2 upsamples in true polyphase form, GEMM columns ordered (slice, phase, channel): a thread's
                                  // 16 columns are output rows 2t, 2t+1 of one slice = 64 contiguous bytes (n_tile == 0: unused)
  std::vector<ConvLayer> ups_pm;  // rate-10 / 12 upsamples with phase-major GEMM columns (coalesced epilogue; n_tile == 0: unused)
  std::vector<ConvLayer> rbconv;  // index ((i*nk + j)*nd + m)*2 + which
  std::vector<NoiseLayer> noise;
  std::vector<float> post_w;  // [k][C]
  std::vector<float> cond_w, cond_b;
  float lin_w = 1.f, lin_b = 0.f;
  int H = 1;                 // sines of the harmonic source (cfg.harmonic_num + 1)
  float lin_wv[16] = {0};    // m_source.l_linear weights (H of them)
  float post_bias = 0.f;     // conv_post bias (optional tensor; the MRF sibling has one)
  const float* rand_ini = nullptr;  // rvcdec_set_initial_phases
  float* d_post_w = nullptr;
  float* d_cond_w = nullptr;
  float* d_cond_b = nullptr;
  float* d_zero_bias = nullptr;
  std::vector<void*> dev_allocs;
  int64_t launches = 0;
  int debug_stop = 0;
  // instrumented fused kernel (tools/fused_timing.py): [launch][sm][16] clock64 spans
  long long* d_timing = nullptr;
  int timing_launches = 0;
  static constexpr int kTimingMaxLaunches = 32;
  int no_source = 0;   // 1 = no harmonic source / noise convs (the no-f0 sibling HiFiGANGenerator, hifigan.py:72-97): the source
                       // kernel is skipped and no stage adds noise_convs[i](har)
  int conv_epi_warps = 16;  // (profiles/r1bq: 16 warps 214 ms vs 8 warps 221 ms on the 128-channel stage) epilogue warps of the 128-wide conv_mma tiling (8 over 32-column items, 16 over 16-column items)
  int conv_epi_warps_small = 16;  // same for the 64- / 32-wide tilings (profiles/r1br: ups2 21.8 -> 17.6 ms, ups3 13.9 -> 12.3 ms)
  int conv_lean = 1;   // 1 = ordinary convolutions (up == 1, no fused noise conv, dense output) of the 128-wide tiling run the
                       // instantiation compiled without the noise-conv / zero-stuffing / phase-scatter epilogue code
  int pm_a_stages = 2; // activation ring depth of the rate-10 phase-major upsample tile (2 or 3)
  int ups_pm_on = 1;   // 1 = upsamples that have a phase-major image use it (and carry the stage's 8-tap noise conv + 16-bit copy)
  int poly2 = 0;       // 1 = stride-2 upsamples of fused-MRF stages run their polyphase image (no zeros read or multiplied; takes
                       // precedence over zero_stuff).  Measured on par with the zero-stuffed form (profiles/r2e_*: ups2 19 vs 17.4 ms,
                       // the two MRF stages in front 3-8 ms faster for their dense output), so it stays an option
  int zero_stuff = 1;  // 1 = stride-2 upsamples of fused-MRF stages read a zero-stuffed input (coalesced epilogue)
  int max_ntile = 128;  // (r1ab: N = 128 tiles keep the accumulators double-buffered in TMEM; 256-wide tiles serialise MMA and epilogue) largest GEMM N tile the weight images are built for (takes effect at the next rvcdec_load_weights)
  int a_stages = 2;  // activation ring depth of the 128-wide conv_mma tiling (r1af: a third stage costs 7 % on the 128-channel stage)
  int mrf_items = 0;  // > 0: the per-conv MRF path runs its launches over groups of items whose 16-bit hand-off tensor
                      // stays under this many MiB (L2 reuse between consecutive launches); 0 = all items at once
  int fused_split_k = 0;  // fused ResBlock chains as two launches (less tile-edge recompute): > 0 = those with kernel size >= this,
                          // 0 = automatic (chains whose recompute exceeds a quarter of the tile), -1 = never
  int pair_noise = 0;  // 1 = the pair stage's short noise conv rides in its upsample's epilogue (no noise_add launch)
  int pair_max_k = 0;  // > 0: on a per-conv 128-channel stage, ResBlocks with kernel size <= this run one fused launch per
                       // (conv1, conv2) pair (mrf_pair.cu) instead of two conv launches; the others stay per-conv
  int pair_kernel = 1;  // 1 = mrf_pair.cu (one CTA per 256-row tile), 2 = mrf_pair2.cu (CTA pairs, cta_group::2, two tiles in flight)
  int pair_mrf = 0;  // 1 = fused conv1+conv2 pair kernel (mrf_pair.cu) on the 128-channel stage (r1z: on par with the per-conv path, so off)
  int fused_epi_warps = 8;   // epilogue warps of the fused ResBlock kernel (8 or 16)
  int fused_post = 1; // 1 = conv_post + tanh fused into the last stage's final fused-MRF launch (C = 32, 7 taps)
  int fused_mrf = 1;  // 0 = one kernel per conv everywhere, 1 = fused ResBlock kernel (one launch per ResBlock) on
                      // the 64/32-channel stages, 2 = one launch per stage (all ResBlocks, common margin)
  // profiling (bench.py): one event before every launch, groups aggregated on read
  bool prof_on = false;
  struct ProfMark { int group; double flops; double bytes; cudaEvent_t ev; };
  std::vector<ProfMark> prof_marks;
  std::vector<cudaEvent_t> prof_pool;
  std::vector<std::string> prof_names;
  std::vector<double> prof_ms, prof_flops, prof_bytes;
  std::vector<int64_t> prof_launches;
  // CUDA graphs: the launch sequence of one pass is captured once per (buffers, shape, mode) and replayed; the Philox
  // key is the only value that changes between replays and lives in device memory (d_seed)
  int use_graphs = 1;
  uint64_t epoch = 0;  // bumped by everything that changes the launch sequence (options, weights): drops the cache
  struct GraphEntry {
    const void *x, *f0, *g, *noise, *out, *ws, *ini;
    int B, T, mode;
    uint64_t item0, epoch;
    cudaGraphExec_t exec;
    int64_t launches;
    uint64_t last_use;
  };
  std::vector<GraphEntry> graphs;
  uint64_t graph_clock = 0, graph_hits = 0, graph_captures = 0;
  uint64_t* d_seed = nullptr;
  cudaStream_t cap_stream = nullptr;
  // host-side time of the last rvcdec_forward_host call, seconds: [0] set-up (memory query, buffers), [1] enqueueing
  // the passes, [2] waiting for the streams
  double host_t[3] = {0, 0, 0};
  // forward_host state
  cudaStream_t stream = nullptr;
  // host-buffer entry point: copies of pass i+1 / i-1 overlap the kernels of pass i
  cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
  cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_comp[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
  void* hbuf = nullptr;
  size_t hbuf_bytes = 0;
};

namespace {

struct TensorMap {
  std::map<std::string, const rvcdec_tensor*> m;
  const rvcdec_tensor* find(const std::string& n) const {
    auto it = m.find(n);
    return it == m.end() ? nullptr : it->second;
  }
};

// w = v * (g / ||v||_2 over all dims but 0)  (torch parametrizations.weight_norm, dim=0)
int get_weight(const TensorMap& tm, const std::string& base, int64_t d0, int64_t rest, std::vector<float>& out) {
  out.resize((size_t)d0 * rest);
  if (const rvcdec_tensor* w = tm.find(base + ".weight")) {
    if (w->numel != d0 * rest) return fail(RVCDEC_ERR_WEIGHTS, "%s.weight: expected %lld elements, got %lld", base.c_str(), (long long)(d0 * rest), (long long)w->numel);
    memcpy(out.data(), w->data, sizeof(float) * out.size());
    return 0;
  }
  const rvcdec_tensor* g = tm.find(base + ".parametrizations.weight.original0");
  const rvcdec_tensor* v = tm.find(base + ".parametrizations.weight.original1");
  if (!g || !v) {
    g = tm.find(base + ".weight_g");
    v = tm.find(base + ".weight_v");
  }
  if (!g || !v) return fail(RVCDEC_ERR_WEIGHTS, "missing weight for %s (looked for .weight, .parametrizations.weight.original0/1, .weight_g/.weight_v)", base.c_str());
  if (g->numel != d0 || v->numel != d0 * rest) return fail(RVCDEC_ERR_WEIGHTS, "%s: weight-norm tensors have wrong size", base.c_str());
  for (int64_t i = 0; i < d0; ++i) {
    const float* vi = v->data + i * rest;
    double ss = 0.0;
    for (int64_t k = 0; k < rest; ++k) ss += (double)vi[k] * (double)vi[k];
    const float scale = (float)((double)g->data[i] / std::sqrt(ss));
    for (int64_t k = 0; k < rest; ++k) out[(size_t)(i * rest + k)] = vi[k] * scale;
  }
  return 0;
}

int get_vec(const TensorMap& tm, const std::string& name, int64_t n, std::vector<float>& out) {
  const rvcdec_tensor* t = tm.find(name);
  if (!t) return fail(RVCDEC_ERR_WEIGHTS, "missing tensor %s", name.c_str());
  if (t->numel != n) return fail(RVCDEC_ERR_WEIGHTS, "%s: expected %lld elements, got %lld", name.c_str(), (long long)n, (long long)t->numel);
  out.assign(t->data, t->data + n);
  return 0;
}

// Conv1d weight (Cout, Cin, k), dilation d, "same" zero padding -> taps
void pack_conv(ConvLayer& L, const std::vector<float>& W, int Cout, int Cin, int k, int d) {
  L.Cin = Cin; L.N_total = Cout; L.cout = Cout; L.up = 1; L.ntaps = k; L.halo = (k - 1) / 2 * d;
  L.macs_per_row = (double)Cin * Cout * k;
  L.wf.assign((size_t)k * Cin * Cout, 0.f);
  for (int j = 0; j < k; ++j) {
    L.shift[j] = (j - (k - 1) / 2) * d;
    for (int ci = 0; ci < Cin; ++ci)
      for (int co = 0; co < Cout; ++co) L.wf[((size_t)j * Cin + ci) * Cout + co] = W[((size_t)co * Cin + ci) * k + j];
  }
}

// ConvTranspose1d weight (Cin, Cout, k), stride u, padding p, output length u*L_in.
// Polyphase form: y[t*u + r] = sum_m x[t + q_r - m] * w[:, :, r' + m*u],  r' = (r+p) % u, q_r = (r+p) / u.
// Written as an ordinary convolution over input shifts s = q_r - m with N = u*Cout columns (n = r*Cout + co).
int pack_ups(ConvLayer& L, const std::vector<float>& W, int Cin, int Cout, int k, int u, int p) {
  int smin = 1 << 30, smax = -(1 << 30);
  for (int r = 0; r < u; ++r) {
    const int rp = (r + p) % u, q = (r + p) / u;
    for (int m = 0; rp + m * u < k; ++m) { smin = std::min(smin, q - m); smax = std::max(smax, q - m); }
  }
  const int nt = smax - smin + 1;
  if (nt > MAX_TAPS) return fail(RVCDEC_ERR_ARG, "upsample kernel %d / rate %d needs %d input shifts (max %d)", k, u, nt, MAX_TAPS);
  L.Cin = Cin; L.N_total = u * Cout; L.cout = Cout; L.up = u; L.ntaps = nt;
  L.macs_per_row = (double)Cin * Cout * k;
  L.halo = std::max(std::abs(smin), std::abs(smax));
  L.wf.assign((size_t)nt * Cin * L.N_total, 0.f);
  for (int t = 0; t < nt; ++t) L.shift[t] = smin + t;
  for (int r = 0; r < u; ++r) {
    const int rp = (r + p) % u, q = (r + p) / u;
    for (int m = 0; rp + m * u < k; ++m) {
      const int t = (q - m) - smin;
      for (int ci = 0; ci < Cin; ++ci)
        for (int co = 0; co < Cout; ++co)
          L.wf[((size_t)t * Cin + ci) * L.N_total + r * Cout + co] = W[((size_t)ci * Cout + co) * k + rp + m * u];
    }
  }
  return 0;
}

uint32_t tap_mask(const ConvLayer& L, int col0, int ncols) {
  uint32_t mask = 0;
  for (int t = 0; t < L.ntaps; ++t) {
    bool nz = false;
    for (int ci = 0; ci < L.Cin && !nz; ++ci) {
      const float* row = &L.wf[((size_t)t * L.Cin + ci) * L.N_total];
      for (int n = col0; n < std::min(col0 + ncols, L.N_total); ++n)
        if (row[n] != 0.f) { nz = true; break; }
    }
    if (nz) mask |= 1u << t;
  }
  return mask ? mask : 1u;
}

uint16_t to16(float v, int fmt) {
  if (fmt == ACT_F16) { __half h = __float2half_rn(v); uint16_t u; memcpy(&u, &h, 2); return u; }
  __nv_bfloat16 h = __float2bfloat16_rn(v); uint16_t u; memcpy(&u, &h, 2); return u;
}

void finish_layer(ConvLayer& L, bool want_mma, int max_ntile = 256, int force_ntile = 0) {
  L.mask_f32.clear();
  for (int c0 = 0; c0 < L.N_total; c0 += 64) L.mask_f32.push_back(tap_mask(L, c0, 64));
  L.n_tile = 0; L.kc = 0;
  if (!want_mma) return;
  if (force_ntile) {
    if (L.N_total % force_ntile == 0 && L.Cin % 64 == 0 && L.N_total / force_ntile <= MAX_NTILES && conv_mma_supported(force_ntile, 64)) {
      L.n_tile = force_ntile; L.kc = 64;
    }
  } else
  for (int nt : {256, 128, 64, 32}) {
    if (nt > max_ntile || L.N_total % nt) continue;
    const int kc = (nt == 32) ? 32 : 64;
    if (L.Cin % kc) continue;
    if (L.N_total / nt > MAX_NTILES) continue;
    if (!conv_mma_supported(nt, kc)) continue;
    L.n_tile = nt; L.kc = kc;
    break;
  }
  if (!L.n_tile) return;
  const int nt = L.n_tile, kc = L.kc, nkc = L.Cin / kc, ntiles = L.N_total / nt;
  L.mask_mma.clear();
  for (int i = 0; i < ntiles; ++i) L.mask_mma.push_back(tap_mask(L, i * nt, nt));
  for (int fmt = 0; fmt < 2; ++fmt) {
    std::vector<uint16_t>& wp = L.wp[fmt];
    wp.assign((size_t)L.ntaps * L.Cin * L.N_total, 0);
    size_t o = 0;
    for (int tile = 0; tile < ntiles; ++tile)
      for (int t = 0; t < L.ntaps; ++t)
        for (int c = 0; c < nkc; ++c)
          for (int sl = 0; sl < kc / 8; ++sl)
            for (int n = 0; n < nt; ++n)
              for (int e = 0; e < 8; ++e)
                wp[o++] = to16(L.wf[((size_t)t * L.Cin + c * kc + sl * 8 + e) * L.N_total + tile * nt + n], fmt);
    if (L.Cin == 128 && L.N_total == 128 && L.up == 1) {
      std::vector<uint16_t>& w2 = L.wp2[fmt];
      w2.assign((size_t)L.ntaps * 128 * 128, 0);
      size_t o2 = 0;
      for (int t = 0; t < L.ntaps; ++t)
        for (int half = 0; half < 2; ++half)
          for (int sl = 0; sl < 16; ++sl)
            for (int n = 0; n < 64; ++n)
              for (int e = 0; e < 8; ++e) w2[o2++] = to16(L.wf[((size_t)t * L.Cin + sl * 8 + e) * L.N_total + half * 64 + n], fmt);
    }
  }
}

template <typename T>
int upload(rvcdec_handle* h, const std::vector<T>& v, T** dptr) {
  if (v.empty()) { *dptr = nullptr; return 0; }
  void* d = nullptr;
  CK(cudaMalloc(&d, v.size() * sizeof(T)));
  h->dev_allocs.push_back(d);
  CK(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  *dptr = (T*)d;
  return 0;
}

int upload_layer(rvcdec_handle* h, ConvLayer& L) {
  int rc;
  if ((rc = upload(h, L.wf, &L.d_wf))) return rc;
  if ((rc = upload(h, L.bias, &L.d_bias))) return rc;
  for (int f = 0; f < 2; ++f)
    if ((rc = upload(h, L.wp[f], &L.d_wp[f]))) return rc;
  for (int f = 0; f < 2; ++f)
    if ((rc = upload(h, L.wp2[f], &L.d_wp2[f]))) return rc;
  if ((rc = upload(h, L.cbias, &L.d_cbias))) return rc;
  return 0;
}

void free_device(rvcdec_handle* h) {
  for (void* p : h->dev_allocs) cudaFree(p);
  h->dev_allocs.clear();
}

// ---- per-mode formats ------------------------------------------------------
// mode 0: fp32 everywhere.  mode 1: bf16 operands up to and including stage 1, fp16 after
// (SURVEY.md 8d numerics probe).  mode 2: fp16 everywhere.  mode 3: bf16 everywhere.
int fmt_pre(int mode) { return mode == 0 ? ACT_F32 : (mode == 2 ? ACT_F16 : ACT_BF16); }
int fmt_stage(int mode, int i) {
  if (mode == 0) return ACT_F32;
  if (mode == 2) return ACT_F16;
  if (mode == 3) return ACT_BF16;
  return i <= 1 ? ACT_BF16 : ACT_F16;
}

bool use_fused_mrf(const rvcdec_handle* h, int mode, int i);
bool use_pair_mrf(const rvcdec_handle* h, int mode, int i);
// Stage i's stride-2 upsample runs in polyphase form with the (slice, phase, channel) column order ("poly2").
bool use_p2(const rvcdec_handle* h, int mode, int i) {
  if (mode == 0 || !h->poly2 || i < 1 || i >= h->nups || h->cfg.upsample_rates[i] != 2 || !h->ups_p2[i].n_tile) return false;
  if (!use_fused_mrf(h, mode, i)) return false;  // (the epilogue writes the fp32 stage input only)
  const NoiseLayer& nl = h->noise[i];
  return h->no_source || (nl.k <= 8 && nl.stride <= 4);
}
// Stage i writes its 16-bit output zero-stuffed when stage i+1 is a stride-2 upsample feeding a fused MRF stage.
bool zs_out_stage(const rvcdec_handle* h, int mode, int i) {
  if (mode == 0 || !h->zero_stuff || i + 1 >= h->nups || use_p2(h, mode, i + 1)) return false;
  return h->cfg.upsample_rates[i + 1] == 2 && h->ups_zs[i + 1].n_tile != 0 && use_fused_mrf(h, mode, i + 1);
}

// Stage i's upsample runs the phase-major tensor-core image (fast modes, dense input).
bool use_ups_pm(const rvcdec_handle* h, int mode, int i) {
  if (mode == 0 || !h->ups_pm_on || i < 0 || i >= h->nups || !h->ups_pm[i].n_tile) return false;
  if (h->ch[i] > 256) return false;  // the kernel keeps the per-channel bias of at most 256 channels in shared memory
  if (i > 0 && zs_out_stage(h, mode, i - 1)) return false;
  return true;
}
// ... and also carries the stage's noise conv (8 taps, stride 4) in its epilogue.
bool pm_noise(const rvcdec_handle* h, int mode, int i) {
  if (!use_ups_pm(h, mode, i)) return false;
  if (h->no_source) return true;  // nothing to add: any phase-major upsample qualifies
  const NoiseLayer& nl = h->noise[i];
  return nl.k == 8 && nl.stride == 4 && nl.pad == 2 && h->ch[i] <= 128;  // (8 taps x 128 channels of shared memory)
}
// ... and, when the stage runs the per-conv MRF path, writes the activated 16-bit operand copy (a0) as well, which
// replaces the noise_add pass.  The upsample then must not read its input out of a0: the previous stage (per-conv
// path, all items in one group, dense output) leaves its output in a2 instead.
bool pm_act(const rvcdec_handle* h, int mode, int i) {
  if (!pm_noise(h, mode, i) || use_fused_mrf(h, mode, i) || use_pair_mrf(h, mode, i)) return false;
  if (i == 0) return true;
  return !use_fused_mrf(h, mode, i - 1) && !use_pair_mrf(h, mode, i - 1) && h->mrf_items == 0 && !zs_out_stage(h, mode, i - 1);
}

struct Workspace {
  float* har; float* biasB; void* xin; void* pre;
  float* xs0; float* xr; float* sum; void* a0; void* a1; void* a2;
  size_t bytes;
};

Workspace carve(const rvcdec_handle* h, void* base, int B, int T, int mode) {
  Workspace w{};
  size_t off = 0;
  auto take = [&](size_t n) { size_t o = off; off += align_up(n, 256); return base ? (void*)((uint8_t*)base + o) : (void*)nullptr; };
  const int C0 = h->cfg.upsample_initial_channel;
  const int LpT = padded_rows(T);
  w.har = (float*)take((size_t)B * T * h->upp * 4);
  w.biasB = (float*)take((size_t)B * C0 * 4);
  w.xin = take((size_t)B * h->cfg.initial_channel * LpT * act_bytes(fmt_pre(mode)));
  w.pre = take((size_t)B * C0 * LpT * act_bytes(fmt_pre(mode)));
  size_t emax = 0, emax_a0 = 0;
  int L = T;
  for (int i = 0; i < h->nups; ++i) {
    L *= h->cfg.upsample_rates[i];
    emax = std::max(emax, (size_t)h->ch[i] * padded_rows(L));
    if (h->loaded && zs_out_stage(h, mode, i)) emax_a0 = std::max(emax_a0, (size_t)h->ch[i] * padded_rows(2 * L));
  }
  emax_a0 = std::max(emax_a0, emax);
  const size_t asz = mode == 0 ? 4 : 2;
  w.xs0 = (float*)take((size_t)B * emax * 4);
  w.xr = (float*)take((size_t)B * emax * 4);
  w.sum = (float*)take((size_t)B * emax * 4);
  w.a0 = take((size_t)B * emax_a0 * asz);
  // a1 and a2 are adjacent: together they also serve as one fp32 tensor (the second residual buffer of the
  // fused-pair path, run_pair_mrf)
  w.a1 = take(2 * (size_t)B * emax * asz);
  w.a2 = w.a1 ? (void*)((uint8_t*)w.a1 + (size_t)B * emax * asz) : nullptr;
  w.bytes = off;
  return w;
}

int prof_group(rvcdec_handle* h, const std::string& name) {
  for (size_t i = 0; i < h->prof_names.size(); ++i)
    if (h->prof_names[i] == name) return (int)i;
  h->prof_names.push_back(name);
  h->prof_ms.push_back(0); h->prof_flops.push_back(0); h->prof_bytes.push_back(0); h->prof_launches.push_back(0);
  return (int)h->prof_names.size() - 1;
}

// Record an event in front of the next launch (group < 0 closes the pass).
int prof_mark(rvcdec_handle* h, const char* group, double flops, double bytes, cudaStream_t s) {
  if (!h->prof_on) return 0;
  cudaEvent_t ev;
  if (!h->prof_pool.empty()) { ev = h->prof_pool.back(); h->prof_pool.pop_back(); }
  else CK(cudaEventCreate(&ev));
  CK(cudaEventRecord(ev, s));
  h->prof_marks.push_back({group ? prof_group(h, group) : -1, flops, bytes, ev});
  return 0;
}
#define MARK(group, flops, bytes) do { int _rc = prof_mark(h, group, flops, bytes, s); if (_rc) return _rc; } while (0)

int run_conv(rvcdec_handle* h, const char* group, const ConvLayer& L, int mode, int fmt_in, const void* act_in, int B,
             int L_in, const Epilogue& epi, cudaStream_t s) {
  if (h->prof_on) {
    const double rows_in = (double)B * L_in, el_out = (double)B * L_in * L.N_total;
    double bytes = rows_in * L.Cin * act_bytes(fmt_in) + (epi.res ? el_out * 4 : 0) + (epi.sum_in ? el_out * 4 : 0) +
                   (epi.out_f32 ? el_out * 4 : 0) + (epi.out_act ? el_out * act_bytes(epi.act_fmt) : 0);
    MARK(group, 2.0 * rows_in * L.macs_per_row, bytes);
  }
  ConvParams p{};
  p.act_in = act_in;
  p.act_fmt_in = fmt_in;
  p.B = B; p.Cin = L.Cin; p.L_in = L_in; p.Lp_in = padded_rows(L_in);
  p.N_total = L.N_total; p.ntaps = L.ntaps; p.halo = L.halo;
  for (int t = 0; t < L.ntaps; ++t) p.shift[t] = L.shift[t];
  p.epi = epi;
  p.a_stages = epi.pm ? h->pm_a_stages : h->a_stages;
  p.epi_warps = h->conv_epi_warps;
  p.epi_warps_small = h->conv_epi_warps_small;
  p.lean = h->conv_lean;
  if (mode == 0) {
    p.w = L.d_wf;
    if (L.mask_f32.size() > (size_t)MAX_NTILES)
      return fail(RVCDEC_ERR_ARG, "layer %s: %d GEMM columns exceed the fp32 back-end's %d tap masks of 64 columns", L.name.c_str(), L.N_total, MAX_NTILES);
    for (size_t i = 0; i < L.mask_f32.size(); ++i) p.tapmask[i] = L.mask_f32[i];
    CK(launch_conv_f32(p, s));
  } else {
    if (!L.n_tile) return fail(RVCDEC_ERR_ARG, "layer %s has no tensor-core tiling (N=%d, Cin=%d)", L.name.c_str(), L.N_total, L.Cin);
    p.w = L.d_wp[fmt_in == ACT_BF16 ? 1 : 0];
    for (size_t i = 0; i < L.mask_mma.size(); ++i) p.tapmask[i] = L.mask_mma[i];
    CK(launch_conv_mma(p, L.n_tile, L.kc, h->num_sms, s));
  }
  ++h->launches;
  return 0;
}


// The fused ResBlock-chain kernel (mrf_fused.cu) serves the stages whose channel count makes the
// per-conv kernels HBM-bound (C = 64, 32).
// largest receptive-field margin (rows per side) of stage i's ResBlock chains
int fused_margin(const rvcdec_handle* h, int i) {
  int margin = 0;
  for (int j = 0; j < h->nk; ++j) {
    int mj = 0;
    for (int m = 0; m < h->nd; ++m)
      for (int which = 0; which < 2; ++which) mj += h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + which].halo;
    margin = std::max(margin, mj);
  }
  return margin;
}
bool use_fused_mrf(const rvcdec_handle* h, int mode, int i) {
  if (mode == 0 || h->fused_mrf == 0) return false;
  if (h->nd > FUSED_MAX_D || h->nk > FUSED_MAX_RB) return false;
  int max_taps = 0, max_halo = 0;
  for (int j = 0; j < h->nk; ++j)
    for (int m = 0; m < h->nd; ++m)
      for (int which = 0; which < 2; ++which) {
        const ConvLayer& L = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + which];
        max_taps = std::max(max_taps, L.ntaps);
        max_halo = std::max(max_halo, L.halo);
        if (L.n_tile != h->ch[i]) return false;  // needs the single-N-tile weight image ([tap][C/8][C][8])
      }
  if (!mrf_fused_supported(h->ch[i], max_taps, max_halo)) return false;
  // the chain's receptive field must leave a useful tile interior
  return mrf_fused_tile_rows(h->ch[i]) - 2 * fused_margin(h, i) >= 128;
}

// The fused-pair kernel (mrf_pair.cu) serves the 128-channel stage, where the per-conv path is HBM-bound.
bool use_pair_mrf(const rvcdec_handle* h, int mode, int i) {
  if (mode == 0 || !h->pair_mrf || h->ch[i] != 128) return false;
  for (int j = 0; j < h->nk; ++j)
    for (int m = 0; m < h->nd; ++m) {
      const ConvLayer& c1 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 0];
      const ConvLayer& c2 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 1];
      if (c1.n_tile != 128 || c1.kc != 64 || c2.n_tile != 128 || c2.kc != 64) return false;  // weight image it streams
      if (!(h->pair_kernel == 2 ? mrf_pair2_supported(h->ch[i], c1.halo, c2.halo) : mrf_pair_supported(h->ch[i], c1.halo, c2.halo))) return false;
    }
  return true;
}

// Mixed mode ("pair_max_k"): ResBlock j of per-conv stage i runs the fused-pair kernel.
bool pair_rb(const rvcdec_handle* h, int mode, int i, int j) {
  if (mode == 0 || h->pair_max_k <= 0 || h->ch[i] != 128 || h->mrf_items != 0) return false;
  if (h->cfg.resblock_kernel_sizes[j] > h->pair_max_k) return false;
  for (int m = 0; m < h->nd; ++m) {
    const ConvLayer& c1 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 0];
    const ConvLayer& c2 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 1];
    if (c1.n_tile != 128 || c1.kc != 64 || c2.n_tile != 128 || c2.kc != 64) return false;
    if (!(h->pair_kernel == 2 ? mrf_pair2_supported(h->ch[i], c1.halo, c2.halo) : mrf_pair_supported(h->ch[i], c1.halo, c2.halo))) return false;
  }
  return true;
}

// One ResBlock j of stage i as nd fused-pair launches.  The fp32 residual stream ping-pongs between xr and the
// a1|a2 region (whose 16-bit pad rows are clobbered: the caller re-zeroes them before the next per-conv ResBlock).
int run_pair_rb(rvcdec_handle* h, const char* group, int i, int j, const Workspace& w, int B, int Lo, int Lpo, int fs,
                bool last, bool zs_out, void* stage_out_act, cudaStream_t s) {
  const int C = h->ch[i];
  float* buf[2] = {w.xr, (float*)w.a1};
  const double el = (double)B * Lo * C;
  for (int m = 0; m < h->nd; ++m) {
    const ConvLayer& c1 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 0];
    const ConvLayer& c2 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 1];
    PairParams p{};
    p.x_in = (m == 0) ? w.xs0 : buf[(m - 1) & 1];
    p.w1 = c1.d_wp[fs == ACT_BF16 ? 1 : 0]; p.w2 = c2.d_wp[fs == ACT_BF16 ? 1 : 0];
    p.w1h = c1.d_wp2[fs == ACT_BF16 ? 1 : 0]; p.w2h = c2.d_wp2[fs == ACT_BF16 ? 1 : 0];
    p.b1 = c1.d_bias; p.b2 = c2.d_bias;
    p.k1 = c1.ntaps; p.d1 = c1.ntaps > 1 ? c1.shift[1] - c1.shift[0] : 1;
    p.k2 = c2.ntaps; p.d2 = c2.ntaps > 1 ? c2.shift[1] - c2.shift[0] : 1;
    p.fmt = fs; p.slope = 0.1f; p.B = B; p.L = Lo; p.Lp = Lpo; p.margin = c1.halo + c2.halo;
    p.scale = 1.f; p.out_slope = 0.1f; p.out_fmt = fs; p.out_zs = 1; p.Lp_out_act = Lpo;
    double bytes = el * 8;
    if (m < h->nd - 1) {
      p.out_f32 = buf[m & 1];
    } else {
      p.scale = 1.f / (float)h->nk;
      p.sum_in = (j == 0) ? nullptr : w.sum;
      if (j > 0) bytes += el * 4;
      if (j < h->nk - 1 || last) p.out_f32 = w.sum; else bytes -= el * 4;
      if (j == h->nk - 1 && !last) {
        p.out_act = stage_out_act;
        if (zs_out) { p.out_zs = 2; p.Lp_out_act = padded_rows(2 * Lo); }
        bytes += el * act_bytes(fs);
      }
    }
    MARK(group, 2.0 * B * Lo * (c1.macs_per_row + c2.macs_per_row), bytes);
    if (h->d_timing && h->timing_launches < rvcdec_handle::kTimingMaxLaunches)
      p.timing = h->d_timing + (size_t)(h->timing_launches++) * h->num_sms * 16;
    CK(h->pair_kernel == 2 ? launch_mrf_pair2(p, h->num_sms, s) : launch_mrf_pair(p, h->num_sms, s)); ++h->launches;
  }
  return 0;
}

// xs = sum_j ResBlock_j(x) / nk with one launch per (conv1, conv2) pair.  The fp32 residual stream ping-pongs
// between xr and the a1|a2 region (a launch must not write the tensor its neighbouring tiles still read).
int run_pair_mrf(rvcdec_handle* h, const char* group, int i, const Workspace& w, int B, int Lo, int Lpo, int fs,
                 bool last, bool zs_out, cudaStream_t s) {
  const int C = h->ch[i];
  float* buf[2] = {w.xr, (float*)w.a1};
  const double el = (double)B * Lo * C;
  for (int j = 0; j < h->nk; ++j)
    for (int m = 0; m < h->nd; ++m) {
      const ConvLayer& c1 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 0];
      const ConvLayer& c2 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 1];
      PairParams p{};
      p.x_in = (m == 0) ? w.xs0 : buf[(m - 1) & 1];
      p.w1 = c1.d_wp[fs == ACT_BF16 ? 1 : 0]; p.w2 = c2.d_wp[fs == ACT_BF16 ? 1 : 0];
    p.w1h = c1.d_wp2[fs == ACT_BF16 ? 1 : 0]; p.w2h = c2.d_wp2[fs == ACT_BF16 ? 1 : 0];
      p.b1 = c1.d_bias; p.b2 = c2.d_bias;
      p.k1 = c1.ntaps; p.d1 = c1.ntaps > 1 ? c1.shift[1] - c1.shift[0] : 1;
      p.k2 = c2.ntaps; p.d2 = c2.ntaps > 1 ? c2.shift[1] - c2.shift[0] : 1;
      p.fmt = fs; p.slope = 0.1f; p.B = B; p.L = Lo; p.Lp = Lpo; p.margin = c1.halo + c2.halo;
      p.scale = 1.f; p.out_slope = 0.1f; p.out_fmt = fs; p.out_zs = 1; p.Lp_out_act = Lpo;
      double bytes = el * 8;
      if (m < h->nd - 1) {
        p.out_f32 = buf[m & 1];
      } else {
        p.scale = 1.f / (float)h->nk;
        p.sum_in = (j == 0) ? nullptr : w.sum;
        if (j > 0) bytes += el * 4;
        if (j < h->nk - 1 || last) p.out_f32 = w.sum; else bytes -= el * 4;
        if (j == h->nk - 1 && !last) {
          p.out_act = w.a0;  // the next stage consumes leaky_relu(x, 0.1) in this stage's operand format
          if (zs_out) { p.out_zs = 2; p.Lp_out_act = padded_rows(2 * Lo); }
          bytes += el * act_bytes(fs);
        }
      }
      MARK(group, 2.0 * B * Lo * (c1.macs_per_row + c2.macs_per_row), bytes);
      if (h->d_timing && h->timing_launches < rvcdec_handle::kTimingMaxLaunches)
        p.timing = h->d_timing + (size_t)(h->timing_launches++) * h->num_sms * 16;
      CK(h->pair_kernel == 2 ? launch_mrf_pair2(p, h->num_sms, s) : launch_mrf_pair(p, h->num_sms, s)); ++h->launches;
    }
  return 0;
}

// conv_post + tanh can ride in the final fused-MRF launch of the last stage (not while a debug tap wants the
// stage output in HBM).
bool fuse_post(const rvcdec_handle* h, int mode, int i) {
  return h->fused_post && h->fused_mrf == 1 && h->debug_stop == 0 && i == h->nups - 1 && use_fused_mrf(h, mode, i) &&
         mrf_fused_post_supported(h->ch[i], 7) &&
         mrf_fused_tile_rows(h->ch[i]) - 2 * (fused_margin(h, i) + 3) >= 128;
}

// `post_out` non-null: conv_post + tanh ride in the last ResBlock's epilogue (the caller checked fuse_post()).
int run_fused_mrf(rvcdec_handle* h, const char* group, int i, const Workspace& w, int B, int Lo, int Lpo, int fs,
                  bool last, bool zs_out, float* post_out, cudaStream_t s) {
  const int C = h->ch[i];
  auto layer = [&](int j, int m, int which) -> const ConvLayer& {
    return h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + which];
  };
  auto fill_rb = [&](FusedRB& R, int j) {
    int margin = 0;
    for (int m = 0; m < h->nd; ++m) {
      const ConvLayer& c1 = layer(j, m, 0);
      const ConvLayer& c2 = layer(j, m, 1);
      const int d1 = c1.ntaps > 1 ? c1.shift[1] - c1.shift[0] : 1, d2 = c2.ntaps > 1 ? c2.shift[1] - c2.shift[0] : 1;
      R.c1[m] = FusedConv{c1.d_wp[fs == ACT_BF16 ? 1 : 0], c1.d_bias, c1.ntaps, d1};
      R.c2[m] = FusedConv{c2.d_wp[fs == ACT_BF16 ? 1 : 0], c2.d_cbias, c2.ntaps, d2};
      margin += c1.halo + c2.halo;
    }
    R.sum_in = (j > 0);
    R.is_last = (j == h->nk - 1);
    return margin;
  };
  FusedParams p{};
  auto next_timing = [&]() -> long long* {
    if (!h->d_timing || h->timing_launches >= rvcdec_handle::kTimingMaxLaunches) return nullptr;
    return h->d_timing + (size_t)(h->timing_launches++) * h->num_sms * 16;
  };
  p.x_in = w.xs0; p.sum = w.sum; p.epi_warps = h->fused_epi_warps;
  p.out_f32 = last ? w.sum : nullptr;  // the last stage hands its fp32 average to conv_post in `sum`
  p.out_act = last ? nullptr : w.a0;   // the next stage consumes leaky_relu(x, 0.1) in this stage's format
  p.fmt = fs; p.out_fmt = fs;
  p.out_zs = zs_out ? 2 : 1; p.Lp_out_act = padded_rows(2 * Lo);
  p.scale = 1.f / (float)h->nk; p.slope = 0.1f; p.out_slope = 0.1f;
  p.B = B; p.L = Lo; p.Lp = Lpo; p.C = C; p.nd = h->nd;
  const double el = (double)B * Lo * C;
  auto flops_rb = [&](int j) {
    double f = 0;
    for (int m = 0; m < h->nd; ++m) f += 2.0 * B * Lo * (layer(j, m, 0).macs_per_row + layer(j, m, 1).macs_per_row);
    return f;
  };
  const double post_fl = post_out ? 2.0 * B * Lo * C * 7 : 0.0;
  auto arm_post = [&]() {  // the launch that completes the MRF average also produces the waveform
    p.post_w = h->d_post_w; p.post_out = post_out; p.post_slope = 0.01f; p.post_bias = h->post_bias; p.out_f32 = nullptr; p.margin += 3;
    for (size_t q = 0; q < h->post_w.size() && q < sizeof(p.post_wc) / sizeof(float); ++q) p.post_wc[q] = h->post_w[q];
  };
  if (h->fused_mrf == 2) {
    double fl = 0;
    p.n_rb = h->nk; p.margin = 0;
    for (int j = 0; j < h->nk; ++j) { p.margin = std::max(p.margin, fill_rb(p.rb[j], j)); fl += flops_rb(j); }
    if (post_out) arm_post();
    MARK(group, fl + post_fl, el * (4 + (last ? (post_out ? 0 : 4) : act_bytes(fs))) + (post_out ? 4.0 * B * Lo : 0.0));
    p.timing = next_timing();
    CK(launch_mrf_fused(p, h->num_sms, s)); ++h->launches;
  } else {
    for (int j = 0; j < h->nk; ++j) {
      p.n_rb = 1;
      const bool fin = (j == h->nk - 1);
      // split when asked for by kernel size ("fused_split_k"), or by default when the chain's tile-edge recompute
      // exceeds a quarter of the tile (64-channel stage, k = 11: 120 of 384 rows; A/B profiles/r1ap_*)
      int chain_margin = 0;
      for (int m = 0; m < h->nd; ++m) chain_margin += layer(j, m, 0).halo + layer(j, m, 1).halo;
      const bool split = h->nd >= 2 && (h->fused_split_k > 0 ? h->cfg.resblock_kernel_sizes[j] >= h->fused_split_k
                                                             : (h->fused_split_k == 0 && 8 * chain_margin > mrf_fused_tile_rows(C)));
      if (split) {
        // Long chains in two launches: dilations [0, nd-1) -> xr (fp32), then the last dilation.  Each launch
        // recomputes only its own receptive field at the tile edges (k = 11: 30 + 30 rows instead of 60 per side,
        // i.e. 84 % instead of 69 % useful rows at 384-row tiles) for one extra fp32 round trip through HBM.
        const int ndA = h->nd - 1;
        double flA = 0, flB = 0;
        int marginA = 0;
        for (int m = 0; m < ndA; ++m) {
          const ConvLayer& c1 = layer(j, m, 0);
          const ConvLayer& c2 = layer(j, m, 1);
          const int d1 = c1.ntaps > 1 ? c1.shift[1] - c1.shift[0] : 1, d2 = c2.ntaps > 1 ? c2.shift[1] - c2.shift[0] : 1;
          p.rb[0].c1[m] = FusedConv{c1.d_wp[fs == ACT_BF16 ? 1 : 0], c1.d_bias, c1.ntaps, d1};
          p.rb[0].c2[m] = FusedConv{c2.d_wp[fs == ACT_BF16 ? 1 : 0], c2.d_cbias, c2.ntaps, d2};  // cumulative over m
          marginA += c1.halo + c2.halo;
          flA += 2.0 * B * Lo * (c1.macs_per_row + c2.macs_per_row);
        }
        FusedParams q = p;
        q.nd = ndA; q.margin = marginA; q.rb[0].sum_in = 0; q.rb[0].is_last = 0;
        q.x_in = w.xs0; q.sum = w.xr; q.scale = 1.f; q.post_w = nullptr;
        MARK(group, flA, el * 8);
        q.timing = next_timing();
        CK(launch_mrf_fused(q, h->num_sms, s)); ++h->launches;
        {
          const ConvLayer& c1 = layer(j, ndA, 0);
          const ConvLayer& c2 = layer(j, ndA, 1);
          const int d1 = c1.ntaps > 1 ? c1.shift[1] - c1.shift[0] : 1, d2 = c2.ntaps > 1 ? c2.shift[1] - c2.shift[0] : 1;
          p.rb[0].c1[0] = FusedConv{c1.d_wp[fs == ACT_BF16 ? 1 : 0], c1.d_bias, c1.ntaps, d1};
          p.rb[0].c2[0] = FusedConv{c2.d_wp[fs == ACT_BF16 ? 1 : 0], c2.d_bias, c2.ntaps, d2};  // xr already carries the earlier biases
          p.margin = c1.halo + c2.halo;
          flB = 2.0 * B * Lo * (c1.macs_per_row + c2.macs_per_row);
        }
        p.nd = 1; p.rb[0].sum_in = (j > 0); p.rb[0].is_last = fin;
        p.x_in = w.xr;
        if (fin && post_out) arm_post();
        MARK(group, flB + (fin ? post_fl : 0.0),
             el * (4 + (j > 0 ? 4 : 0) + (fin ? (last ? (post_out ? 0 : 4) : act_bytes(fs)) : 4)) + (fin && post_out ? 4.0 * B * Lo : 0.0));
        p.timing = next_timing();
        CK(launch_mrf_fused(p, h->num_sms, s)); ++h->launches;
        p.nd = h->nd; p.x_in = w.xs0;
        continue;
      }
      p.margin = fill_rb(p.rb[0], j);
      if (fin && post_out) arm_post();
      MARK(group, flops_rb(j) + (fin ? post_fl : 0.0),
           el * (4 + (j > 0 ? 4 : 0) + (fin ? (last ? (post_out ? 0 : 4) : act_bytes(fs)) : 4)) + (fin && post_out ? 4.0 * B * Lo : 0.0));
      p.timing = next_timing();
      CK(launch_mrf_fused(p, h->num_sms, s)); ++h->launches;
    }
  }
  return 0;
}

// One pass over B items (B small enough for the workspace).
int forward_pass(rvcdec_handle* h, const float* x, const float* f0, const float* g, const float* noise, uint64_t seed,
                 uint64_t item0, float* out, void* wsbase, int B, int T, int mode, cudaStream_t s,
                 const uint64_t* seed_dev = nullptr) {
  const rvcdec_config& c = h->cfg;
  Workspace w = carve(h, wsbase, B, T, mode);
  const int stop = h->debug_stop;
  bool post_fused = false;
  const int C0 = c.upsample_initial_channel;
  const int LpT = padded_rows(T);
  const int fpre = fmt_pre(mode);

  SourceParams sp{};
  sp.f0 = f0; sp.noise = noise; sp.har = w.har; sp.seed = seed; sp.item0 = item0; sp.seed_dev = seed_dev;
  sp.B = B; sp.T = T; sp.upp = h->upp; sp.sr = (float)c.sr; sp.lin_w = h->lin_w; sp.lin_b = h->lin_b;
  sp.H = h->H; sp.rand_ini = h->rand_ini ? h->rand_ini + (size_t)item0 * h->H : nullptr;
  for (int k = 0; k < 16; ++k) sp.lin_wv[k] = h->lin_wv[k];
  if (h->no_source) {
    // stages that still run the noise_add pass (for its activated operand copy) multiply har by zero weights
    CK(cudaMemsetAsync(w.har, 0, (size_t)B * T * h->upp * sizeof(float), s));
  } else {
    MARK("source", 0, (double)B * T * h->upp * (noise ? 4 + 4 * h->H : 4));
    CK(launch_source(sp, s)); ++h->launches;
  }
  if (stop == 1) { MARK(nullptr, 0, 0); return 0; }

  MARK("prep", 0, (double)B * c.initial_channel * T * (4 + act_bytes(fpre)));
  CK(launch_cond(c.gin_channels ? g : nullptr, h->d_cond_w, h->d_cond_b, h->conv_pre.d_bias, w.biasB, B,
                 c.gin_channels, C0, s)); ++h->launches;
  CK(launch_zero_pads(w.xin, nullptr, nullptr, act_bytes(fpre), B, c.initial_channel, T, LpT, s)); ++h->launches;
  CK(launch_zero_pads(w.pre, nullptr, nullptr, act_bytes(fpre), B, C0, T, LpT, s)); ++h->launches;
  CK(launch_pack_input(x, w.xin, fpre, B, c.initial_channel, T, LpT, s)); ++h->launches;
  {
    Epilogue e{};
    e.bias = w.biasB; e.bias_bstride = C0; e.scale = 1.f; e.slope = 0.1f;
    e.out_act = w.pre; e.act_fmt = fpre; e.up = 1; e.cout = C0; e.L_out = T; e.Lp_out = LpT;
    int rc = run_conv(h, "conv_pre", h->conv_pre, mode, fpre, w.xin, B, T, e, s);
    if (rc) return rc;
  }
  if (stop == 2) { MARK(nullptr, 0, 0); return 0; }

  const void* stage_in = w.pre;
  int fmt_in = fpre;
  int L = T;
  bool zs_in = false;  // stage_in is zero-stuffed (rows 2t hold the data)
  for (int i = 0; i < h->nups; ++i) {
    const std::string g_ups = "ups" + std::to_string(i), g_noise = "noise_add" + std::to_string(i),
                      g_mrf = "mrf" + std::to_string(i);
    const int u = c.upsample_rates[i];
    const int C = h->ch[i];
    const int Lo = L * u;
    const int Lpo = padded_rows(Lo);
    const int fs = fmt_stage(mode, i);
    const bool last = (i == h->nups - 1);
    const NoiseLayer& nl = h->noise[i];
    // Short noise convs ride in the upsample epilogue when the stage runs the fused MRF kernel (which
    // reads the fp32 stage input only, so no activated 16-bit copy is needed).
    const bool nz_epi = nl.k <= 8 || h->no_source;  // the noise conv can ride in (or is absent from) an upsample epilogue
    const bool fuse_other = nz_epi && (use_fused_mrf(h, mode, i) || (h->pair_noise && use_pair_mrf(h, mode, i)));
    const bool pm = use_ups_pm(h, mode, i) && (!fuse_other || pm_noise(h, mode, i)), pma = pm_act(h, mode, i);
    const bool pmn = pm_noise(h, mode, i) && (pma || use_fused_mrf(h, mode, i) || (h->pair_noise && use_pair_mrf(h, mode, i)));
    const bool fuse_noise = pmn || (nl.k <= 8 && (use_fused_mrf(h, mode, i) || (h->pair_noise && use_pair_mrf(h, mode, i))));
    const bool zs_out = zs_out_stage(h, mode, i);
    const bool next_in_a2 = (i + 1 < h->nups) && i + 1 >= 1 && pm_act(h, mode, i + 1);  // see pm_act
    if (zs_in) {
      // x = ups[i](...) as an ordinary convolution over the zero-stuffed tensor the previous stage wrote
      CK(launch_zero_pads(w.a0, nullptr, nullptr, act_bytes(fmt_in), B, h->ch[i - 1], Lo, Lpo, s)); ++h->launches;
      Epilogue e{};
      e.bias = h->ups_zs[i].d_bias; e.scale = 1.f; e.slope = 0.1f;
      e.out_f32 = w.xs0; e.up = 1; e.cout = C; e.L_out = Lo; e.Lp_out = Lpo;
      if (nl.k <= 8 && !h->no_source) {  // the coalesced epilogue carries noise convs of up to 8 taps
        e.nz_har = w.har; e.nz_w = nl.d_w; e.nz_bias = nl.d_bias;
        e.nz_k = nl.k; e.nz_stride = nl.stride; e.nz_pad = nl.pad; e.nz_src = T * h->upp;
      }
      int rc = run_conv(h, g_ups.c_str(), h->ups_zs[i], mode, fmt_in, stage_in, B, Lo, e, s);
      if (rc) return rc;
    } else
    {  // x = ups[i](leaky_relu(x, 0.1))  (hifigan_nsf.py:185,198); the activation was applied by the producer
      const bool p2 = use_p2(h, mode, i);
      const ConvLayer& UL = p2 ? h->ups_p2[i] : (pm ? h->ups_pm[i] : h->ups[i]);
      Epilogue e{};
      e.bias = UL.d_bias; e.scale = 1.f; e.slope = 0.1f; e.pm = p2 ? 2 : (pm ? 1 : 0);
      e.out_f32 = w.xs0; e.up = u; e.cout = C; e.L_out = Lo; e.Lp_out = Lpo;
      if (fuse_noise && !h->no_source) {  // ... + noise_convs[i](har_source)  (:199)
        e.nz_har = w.har; e.nz_w = nl.d_w; e.nz_bias = nl.d_bias;
        e.nz_k = nl.k; e.nz_stride = nl.stride; e.nz_pad = nl.pad; e.nz_src = T * h->upp;
      }
      if (pma) {  // leaky_relu(x, 0.1) as the ResBlocks' first operand (what noise_add wrote)
        if (stage_in == w.a0) return fail(RVCDEC_ERR_STATE, "phase-major upsample %d would overwrite its own input", i);
        e.out_act = w.a0; e.act_fmt = fs;
      }
      int rc = run_conv(h, g_ups.c_str(), UL, mode, fmt_in, stage_in, B, L, e, s);
      if (rc) return rc;
    }
    // (after the upsample has read the previous stage's tensor out of a0: new geometry, new pad rows)
    if (pma) { CK(launch_zero_pads(w.a0, w.a1, w.a2, act_bytes(fs), B, C, Lo, Lpo, s)); ++h->launches; }
    else if ((fuse_noise || (zs_in && nz_epi)) && !zs_out) { CK(launch_zero_pads(w.a0, nullptr, nullptr, act_bytes(fs), B, C, Lo, Lpo, s)); ++h->launches; }
    if (!fuse_noise && !(zs_in && nz_epi)) {
      MARK(g_noise.c_str(), 2.0 * B * Lo * C * nl.k, (double)B * Lo * C * (8 + act_bytes(fs)));
      CK(launch_zero_pads(w.a0, w.a1, w.a2, act_bytes(fs), B, C, Lo, Lpo, s)); ++h->launches;
      // x = x + noise_convs[i](har_source)  (:199)
      NoiseAddParams np{};
      np.har = w.har; np.w = nl.d_w; np.bias = nl.d_bias; np.x = w.xs0; np.act = w.a0; np.act_fmt = fs;
      np.B = B; np.C = C; np.L = Lo; np.Lp = Lpo; np.n_src = T * h->upp; np.k = nl.k; np.stride = nl.stride;
      np.pad = nl.pad; np.slope = 0.1f;
      CK(launch_noise_add(np, s)); ++h->launches;
    }
    if (stop == 10 + i) { MARK(nullptr, 0, 0); return 0; }
    // xs = sum_j resblocks[i*nk + j](x) / nk  (:200-207), ResBlock.forward residuals.py:71-80
    if (use_fused_mrf(h, mode, i)) {
      post_fused = last && fuse_post(h, mode, i);
      int rc = run_fused_mrf(h, g_mrf.c_str(), i, w, B, Lo, Lpo, fs, last, zs_out, post_fused ? out : nullptr, s);
      if (rc) return rc;
    } else if (use_pair_mrf(h, mode, i)) {
      int rc = run_pair_mrf(h, g_mrf.c_str(), i, w, B, Lo, Lpo, fs, last, zs_out, s);
      if (rc) return rc;
    } else {
      // Per-conv path.  With "mrf_items" = G > 0 (fast mode) the 18 launches run over G items at a time, so the
      // 16-bit tensors handed from one launch to the next (conv1 -> conv2 -> next conv1) are small enough to be
      // re-read out of L2 instead of HBM.  Groups go in DESCENDING item order: the last launch of a group writes
      // the stage output into a0, possibly in the wider zero-stuffed geometry, i.e. over the a0 inputs of
      // higher-numbered items only - which are consumed by then.
      const size_t ab = act_bytes(mode == 0 ? ACT_F32 : fs);
      const size_t item_el = (size_t)C * Lpo;
      int G = B;  // "mrf_items" = budget in MiB for one 16-bit hand-off tensor of a group
      if (mode != 0 && h->mrf_items > 0)
        G = (int)std::min<size_t>((size_t)B, std::max<size_t>(1, ((size_t)h->mrf_items << 20) / (item_el * ab)));
      const int Lp_zs = padded_rows(2 * Lo);
      for (int b0 = ((B - 1) / G) * G; b0 >= 0; b0 -= G) {
        const int nb = std::min(G, B - b0);
        auto A = [&](void* p_) { return (void*)((uint8_t*)p_ + (size_t)b0 * item_el * ab); };
        auto F = [&](float* p_) { return p_ + (size_t)b0 * item_el; };
        bool pads_dirty = false;  // a fused-pair ResBlock used the a1|a2 region as an fp32 buffer
        for (int j = 0; j < h->nk; ++j) {
          if (G == B && pair_rb(h, mode, i, j) && !(j == h->nk - 1 && next_in_a2)) {
            void* oa = zs_out ? w.a0 : (next_in_a2 ? w.a2 : w.a0);
            int rc = run_pair_rb(h, g_mrf.c_str(), i, j, w, B, Lo, Lpo, fs, last, zs_out, oa, s);
            if (rc) return rc;
            pads_dirty = true;
            continue;
          }
          if (pads_dirty) {
            CK(launch_zero_pads(w.a1, w.a2, nullptr, act_bytes(fs), B, C, Lo, Lpo, s)); ++h->launches;
            pads_dirty = false;
          }
          for (int m = 0; m < h->nd; ++m) {
            const ConvLayer& c1 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 0];
            const ConvLayer& c2 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 1];
            {
              Epilogue e{};
              e.bias = c1.d_bias; e.scale = 1.f; e.slope = 0.1f;
              e.out_act = A(w.a1); e.act_fmt = fs; e.up = 1; e.cout = C; e.L_out = Lo; e.Lp_out = Lpo;
              int rc = run_conv(h, g_mrf.c_str(), c1, mode, fs, m == 0 ? A(w.a0) : A(w.a2), nb, Lo, e, s);
              if (rc) return rc;
            }
            {
              Epilogue e{};
              e.bias = c2.d_bias; e.slope = 0.1f; e.up = 1; e.cout = C; e.L_out = Lo; e.Lp_out = Lpo;
              e.res = (m == 0) ? F(w.xs0) : F(w.xr);
              if (m < h->nd - 1) {
                e.scale = 1.f; e.out_f32 = F(w.xr); e.out_act = A(w.a2); e.act_fmt = fs;
              } else {
                e.scale = 1.f / (float)h->nk;
                e.sum_in = (j == 0) ? nullptr : F(w.sum);
                if (j < h->nk - 1 || last) e.out_f32 = F(w.sum);
                if (j == h->nk - 1 && !last) {
                  // next stage consumes leaky_relu(x, 0.1) in ITS operand format; a0 is dead by now
                  e.act_fmt = fmt_stage(mode, i);
                  if (zs_out) {
                    e.act_zs = 2; e.Lp_act = Lp_zs;
                    e.out_act = (uint8_t*)w.a0 + (size_t)b0 * C * Lp_zs * ab;
                  } else {
                    e.out_act = next_in_a2 ? A(w.a2) : A(w.a0);  // a2's last reader (this pair's conv1) is done
                  }
                }
              }
              int rc = run_conv(h, g_mrf.c_str(), c2, mode, fs, A(w.a1), nb, Lo, e, s);
              if (rc) return rc;
            }
          }
        }
      }
    }
    if (stop == 20 + i) { MARK(nullptr, 0, 0); return 0; }
    stage_in = next_in_a2 ? w.a2 : w.a0;
    fmt_in = fs;
    L = Lo;
    zs_in = zs_out;
  }
  // x = tanh(conv_post(leaky_relu(x)))  (:209-210), default slope 0.01
  if (!post_fused) {
    MARK("conv_post", 2.0 * B * L * h->ch[h->nups - 1] * 7, (double)B * L * (h->ch[h->nups - 1] * 4 + 4));
    CK(launch_post(w.sum, h->d_post_w, out, B, h->ch[h->nups - 1], L, padded_rows(L), 7, 0.01f, s, h->post_bias)); ++h->launches;
  }
  MARK(nullptr, 0, 0);
  return 0;
}

void drop_graphs(rvcdec_handle* h) {
  for (auto& e : h->graphs) cudaGraphExecDestroy(e.exec);
  h->graphs.clear();
}

// One pass, replayed from a CUDA graph when the same buffers / shape / mode were seen before (the reference's call
// pattern is one B = 1 chunk after another through the same staging buffers: ~60 dependent launches each).  Anything
// that needs host-visible side effects per launch (profiling events, debug stops, the timing twin) runs directly.
int run_pass(rvcdec_handle* h, const float* x, const float* f0, const float* g, const float* noise, uint64_t seed,
             uint64_t item0, float* out, void* ws, int B, int T, int mode, cudaStream_t s) {
  const bool can = h->use_graphs && !h->prof_on && h->debug_stop == 0 && !h->d_timing;
  if (!can) return forward_pass(h, x, f0, g, noise, seed, item0, out, ws, B, T, mode, s);
  if (!h->d_seed) {
    CK(cudaMalloc((void**)&h->d_seed, sizeof(uint64_t)));
    CK(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
  }
  CK(launch_set_u64(h->d_seed, seed, s));
  rvcdec_handle::GraphEntry* hit = nullptr;
  for (auto& e : h->graphs)
    if (e.x == x && e.f0 == f0 && e.g == g && e.noise == noise && e.out == out && e.ws == ws && e.ini == h->rand_ini && e.B == B && e.T == T &&
        e.mode == mode && e.item0 == item0 && e.epoch == h->epoch) { hit = &e; break; }
  if (!hit) {
    const int64_t before = h->launches;
    cudaGraph_t graph = nullptr;
    cudaError_t ce = cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeRelaxed);
    int rc = 0;
    if (ce == cudaSuccess) {
      rc = forward_pass(h, x, f0, g, noise, seed, item0, out, ws, B, T, mode, h->cap_stream, h->d_seed);
      ce = cudaStreamEndCapture(h->cap_stream, &graph);
    }
    cudaGraphExec_t exec = nullptr;
    if (rc == 0 && ce == cudaSuccess && graph) ce = cudaGraphInstantiate(&exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (rc != 0 || ce != cudaSuccess || !exec) {
      // capture is an optimisation: fall back to direct launches for the rest of this handle's life
      (void)cudaGetLastError();
      h->use_graphs = 0;
      h->launches = before;
      if (rc) return rc;
      return forward_pass(h, x, f0, g, noise, seed, item0, out, ws, B, T, mode, s);
    }
    if (h->graphs.size() >= 32) {  // LRU
      size_t victim = 0;
      for (size_t i = 1; i < h->graphs.size(); ++i) if (h->graphs[i].last_use < h->graphs[victim].last_use) victim = i;
      cudaGraphExecDestroy(h->graphs[victim].exec);
      h->graphs.erase(h->graphs.begin() + victim);
    }
    h->graphs.push_back({x, f0, g, noise, out, ws, h->rand_ini, B, T, mode, item0, h->epoch, exec, h->launches - before, 0});
    h->launches = before;
    hit = &h->graphs.back();
    ++h->graph_captures;
  } else {
    ++h->graph_hits;
  }
  hit->last_use = ++h->graph_clock;
  CK(cudaGraphLaunch(hit->exec, s));
  h->launches += hit->launches + 1;
  return 0;
}

}  // namespace

extern "C" {

const char* rvcdec_last_error(void) { return g_err.c_str(); }
const char* rvcdec_version(void) { return "rvcdec 0.1 (sm_100a)"; }

int rvcdec_create(const rvcdec_config* cfg, int device, rvcdec_handle** out) {
  if (!cfg || !out) return fail(RVCDEC_ERR_ARG, "null argument");
  *out = nullptr;
  const rvcdec_config& c = *cfg;
  if (c.num_upsamples < 1 || c.num_upsamples > RVCDEC_MAX_UPS || c.num_kernels < 1 || c.num_kernels > RVCDEC_MAX_KERNELS ||
      c.num_dilations < 1 || c.num_dilations > RVCDEC_MAX_DILATIONS)
    return fail(RVCDEC_ERR_ARG, "bad layer counts");
  if (c.initial_channel % 8 || c.initial_channel <= 0) return fail(RVCDEC_ERR_ARG, "initial_channel must be a positive multiple of 8");
  if (c.sr <= 0 || c.gin_channels < 0) return fail(RVCDEC_ERR_ARG, "bad sr / gin_channels");
  if (c.harmonic_num < 0 || c.harmonic_num > 15) return fail(RVCDEC_ERR_ARG, "harmonic_num must be in [0, 15]");
  auto h = new rvcdec_handle();
  h->cfg = c; h->device = device; h->nups = c.num_upsamples; h->nk = c.num_kernels; h->nd = c.num_dilations;
  h->H = c.harmonic_num + 1;
  h->upp = 1;
  for (int i = 0; i < h->nups; ++i) {
    const int u = c.upsample_rates[i], k = c.upsample_kernel_sizes[i];
    // hifigan_nsf.py:113-127: even rates pad (k-u)/2, odd rates pad u/2 + 1 with output_padding 1; the output must be
    // exactly u times the input (x + noise_convs[i](har), :199, needs equal lengths in the reference as well)
    const int pad = (u % 2 == 0) ? (k - u) / 2 : u / 2 + 1, opad = u % 2;
    if (u < 1 || k < u || k + opad - 2 * pad != u) {
      delete h;
      return fail(RVCDEC_ERR_ARG, "upsample %d: rate %d / kernel %d gives an output that is not %d x the input (even rates need k - u even, odd rates k = 2u)", i, u, k, u);
    }
    h->ups_pad.push_back(pad);
    const int C = c.upsample_initial_channel >> (i + 1);
    if (C < 8 || C % 8 || (C << (i + 1)) != c.upsample_initial_channel) { delete h; return fail(RVCDEC_ERR_ARG, "channel count %d at stage %d must be a multiple of 8", C, i); }
    h->ch.push_back(C);
    h->upp *= u;
  }
  for (int j = 0; j < h->nk; ++j) {
    const int k = c.resblock_kernel_sizes[j];
    if (k < 1 || k % 2 == 0 || k > MAX_TAPS) { delete h; return fail(RVCDEC_ERR_ARG, "resblock kernel size %d unsupported (odd, <= %d)", k, MAX_TAPS); }
    for (int m = 0; m < h->nd; ++m) {
      const int d = c.resblock_dilation_sizes[j][m];
      if (d < 1 || (k - 1) / 2 * d > PADL) { delete h; return fail(RVCDEC_ERR_ARG, "resblock halo (k=%d, d=%d) exceeds %d rows", k, d, PADL); }
    }
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) == cudaSuccess && device >= 0 && device < ndev) {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) h->num_sms = prop.multiProcessorCount;
  }
  *out = h;
  return 0;
}

void rvcdec_destroy(rvcdec_handle* h) {
  if (!h) return;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) == cudaSuccess && h->device < ndev) {
    DeviceGuard dguard(h->device);
    drop_graphs(h);
    free_device(h);
    if (h->d_seed) cudaFree(h->d_seed);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    if (h->d_timing) cudaFree(h->d_timing);
    if (h->hbuf) cudaFree(h->hbuf);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->h2d_stream) cudaStreamDestroy(h->h2d_stream);
    if (h->d2h_stream) cudaStreamDestroy(h->d2h_stream);
    for (int i = 0; i < 2; ++i)
      for (cudaEvent_t e : {h->ev_in[i], h->ev_comp[i], h->ev_out[i]})
        if (e) cudaEventDestroy(e);
  }
  delete h;
}

// Host-side fold + repack only (no CUDA): shared by load_weights and the CPU tests.
static int pack_all(rvcdec_handle* h, const rvcdec_tensor* tensors, int32_t n) {
  const rvcdec_config& c = h->cfg;
  TensorMap tm;
  for (int i = 0; i < n; ++i) {
    std::string name = tensors[i].name ? tensors[i].name : "";
    if (name.rfind("dec.", 0) == 0) name = name.substr(4);
    tm.m[name] = &tensors[i];
  }
  int rc;
  std::vector<float> W, v;
  const int C0 = c.upsample_initial_channel;
  // m_source.l_linear (hifigan_nsf.py:45)
  if ((rc = get_vec(tm, "m_source.l_linear.weight", h->H, v))) return rc;
  h->lin_w = v[0];
  for (int k = 0; k < 16; ++k) h->lin_wv[k] = k < h->H ? v[k] : 0.f;
  if ((rc = get_vec(tm, "m_source.l_linear.bias", 1, v))) return rc;
  h->lin_b = v[0];
  // conv_pre (:95-97)
  if ((rc = get_weight(tm, "conv_pre", C0, (int64_t)c.initial_channel * 7, W))) return rc;
  h->conv_pre = ConvLayer();
  h->conv_pre.name = "conv_pre";
  pack_conv(h->conv_pre, W, C0, c.initial_channel, 7, 1);
  if ((rc = get_vec(tm, "conv_pre.bias", C0, h->conv_pre.bias))) return rc;
  // cond (:167-168)
  h->cond_w.clear(); h->cond_b.clear();
  if (c.gin_channels) {
    if ((rc = get_weight(tm, "cond", C0, c.gin_channels, h->cond_w))) return rc;
    if ((rc = get_vec(tm, "cond.bias", C0, h->cond_b))) return rc;
  }
  h->ups.assign(h->nups, ConvLayer());
  h->ups_zs.assign(h->nups, ConvLayer());
  h->ups_pm.assign(h->nups, ConvLayer());
  h->ups_p2.assign(h->nups, ConvLayer());
  h->noise.assign(h->nups, NoiseLayer());
  h->rbconv.assign((size_t)h->nups * h->nk * h->nd * 2, ConvLayer());
  for (int i = 0; i < h->nups; ++i) {
    const int Cin = C0 >> i, C = h->ch[i], u = c.upsample_rates[i], k = c.upsample_kernel_sizes[i];
    const std::string un = "ups." + std::to_string(i);
    if ((rc = get_weight(tm, un, Cin, (int64_t)C * k, W))) return rc;  // weight-norm dim 0 = C_in
    h->ups[i].name = un;
    if ((rc = pack_ups(h->ups[i], W, Cin, C, k, u, h->ups_pad[i]))) return rc;
    if ((rc = get_vec(tm, un + ".bias", C, v))) return rc;
    h->ups_zs[i] = ConvLayer();
    if (u == 2 && k <= MAX_TAPS) {
      // ConvTranspose1d(k, stride u, padding p) == Conv1d over the zero-stuffed input z[i*u] = x[i] with the
      // flipped kernel: y[o] = b + sum_j z[o - (k-1-p) + j] * w[:, :, k-1-j]
      ConvLayer& Z = h->ups_zs[i];
      const int p0 = h->ups_pad[i];
      Z.name = un + "(zero-stuffed)";
      Z.Cin = Cin; Z.N_total = C; Z.cout = C; Z.up = 1; Z.ntaps = k; Z.halo = 0;
      Z.macs_per_row = (double)Cin * C * k / u;  // algorithmic MACs per OUTPUT row
      Z.wf.assign((size_t)k * Cin * C, 0.f);
      for (int j = 0; j < k; ++j) {
        Z.shift[j] = j - (k - 1 - p0);
        Z.halo = std::max(Z.halo, std::abs(Z.shift[j]));
        for (int ci = 0; ci < Cin; ++ci)
          for (int co = 0; co < C; ++co) Z.wf[((size_t)j * Cin + ci) * C + co] = W[((size_t)ci * C + co) * k + (k - 1 - j)];
      }
      Z.bias = v;
    }
    h->ups[i].bias.resize((size_t)u * C);
    for (int r = 0; r < u; ++r) memcpy(&h->ups[i].bias[(size_t)r * C], v.data(), sizeof(float) * C);
    if (u == 2 && C % 8 == 0) {
      // Polyphase image of a stride-2 upsample with the columns of a slice's two phases side by side:
      // n_new = ((co/8)*2 + r)*8 + co%8, so input row t's 16 columns of slice co/8 are output rows 2t and 2t+1.
      const ConvLayer& S = h->ups[i];
      ConvLayer& P = h->ups_p2[i];
      P.name = un + "(polyphase-2)";
      P.Cin = S.Cin; P.N_total = S.N_total; P.cout = C; P.up = u; P.ntaps = S.ntaps; P.halo = S.halo;
      P.macs_per_row = S.macs_per_row;
      for (int t = 0; t < S.ntaps; ++t) P.shift[t] = S.shift[t];
      P.wf.assign(S.wf.size(), 0.f);
      P.bias.resize((size_t)u * C);
      for (int r = 0; r < u; ++r)
        for (int co = 0; co < C; ++co) {
          const int n_new = ((co / 8) * 2 + r) * 8 + co % 8;
          P.bias[n_new] = v[co];
          for (int t = 0; t < S.ntaps; ++t)
            for (int ci = 0; ci < Cin; ++ci)
              P.wf[((size_t)t * Cin + ci) * P.N_total + n_new] = S.wf[((size_t)t * Cin + ci) * S.N_total + r * C + co];
        }
    }
    if (u % 2 == 0 && C % 16 == 0 && conv_mma_supported(u / 2 * 16, 64) && u / 2 * 16 != 64 && Cin % 64 == 0) {
      // Phase-major column order: tile (co/16, half) holds the phases r = half*PH + rr of 16 channels as
      // column ((co/8 % 2)*PH + rr)*8 + co%8.  Each half of the phases reads two of the three input shifts, so the
      // per-tile tap masks keep skipping the all-zero weight blocks.
      const ConvLayer& S = h->ups[i];
      ConvLayer& P = h->ups_pm[i];
      const int PH = u / 2, NT = PH * 16;
      P.name = un + "(phase-major)";
      P.Cin = S.Cin; P.N_total = S.N_total; P.cout = C; P.up = u; P.ntaps = S.ntaps; P.halo = S.halo;
      P.macs_per_row = S.macs_per_row;
      for (int t = 0; t < S.ntaps; ++t) P.shift[t] = S.shift[t];
      P.wf.assign(S.wf.size(), 0.f);
      P.bias.resize((size_t)u * C);
      for (int r = 0; r < u; ++r)
        for (int co = 0; co < C; ++co) {
          const int n_new = ((co / 16) * 2 + r / PH) * NT + (((co / 8) % 2) * PH + r % PH) * 8 + co % 8;
          P.bias[n_new] = v[co];
          for (int t = 0; t < S.ntaps; ++t)
            for (int ci = 0; ci < Cin; ++ci)
              P.wf[((size_t)t * Cin + ci) * P.N_total + n_new] = S.wf[((size_t)t * Cin + ci) * S.N_total + r * C + co];
        }
    }
    // noise_convs (:142-154): kernel 2s, stride s, pad s/2 (1/1/0 when s == 1)
    int stride = 1;
    for (int q = i + 1; q < h->nups; ++q) stride *= c.upsample_rates[q];
    NoiseLayer& nl = h->noise[i];
    nl.C = C; nl.stride = stride;
    nl.k = stride == 1 ? 1 : stride * 2 - stride % 2;
    nl.pad = stride == 1 ? 0 : (nl.k - stride) / 2;
    const std::string nn = "noise_convs." + std::to_string(i);
    if ((rc = get_weight(tm, nn, C, nl.k, W))) return rc;
    nl.w.resize((size_t)nl.k * C);
    for (int co = 0; co < C; ++co)
      for (int j = 0; j < nl.k; ++j) nl.w[(size_t)j * C + co] = W[(size_t)co * nl.k + j];
    if ((rc = get_vec(tm, nn + ".bias", C, nl.bias))) return rc;
    for (int j = 0; j < h->nk; ++j)
      for (int m = 0; m < h->nd; ++m)
        for (int which = 0; which < 2; ++which) {
          const int k2 = c.resblock_kernel_sizes[j];
          const int d = which == 0 ? c.resblock_dilation_sizes[j][m] : 1;
          const std::string base = "resblocks." + std::to_string(i * h->nk + j) + ".convs" + std::to_string(which + 1) + "." + std::to_string(m);
          ConvLayer& L = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + which];
          L.name = base;
          if ((rc = get_weight(tm, base, C, (int64_t)C * k2, W))) return rc;
          pack_conv(L, W, C, C, k2, d);
          if ((rc = get_vec(tm, base + ".bias", C, L.bias))) return rc;
        }
  }
  for (int i = 0; i < h->nups; ++i)
    for (int j = 0; j < h->nk; ++j) {
      std::vector<float> run(h->ch[i], 0.f);
      for (int m = 0; m < h->nd; ++m) {
        ConvLayer& L2 = h->rbconv[(((size_t)i * h->nk + j) * h->nd + m) * 2 + 1];
        for (int cc = 0; cc < h->ch[i]; ++cc) run[cc] += L2.bias[cc];
        L2.cbias = run;
      }
    }
  // conv_post (:164): (1, C, 7) no bias -> [j][c]
  const int Cl = h->ch[h->nups - 1];
  if ((rc = get_weight(tm, "conv_post", 1, (int64_t)Cl * 7, W))) return rc;
  h->post_bias = 0.f;
  if (const rvcdec_tensor* pb = tm.find("conv_post.bias")) {  // hifigan_mrf.py:327-329 (the NSF decoder's conv_post has none)
    if (pb->numel != 1) return fail(RVCDEC_ERR_WEIGHTS, "conv_post.bias: expected 1 element");
    h->post_bias = pb->data[0];
  }
  h->post_w.resize((size_t)7 * Cl);
  for (int cc = 0; cc < Cl; ++cc)
    for (int j = 0; j < 7; ++j) h->post_w[(size_t)j * Cl + cc] = W[(size_t)cc * 7 + j];

  // tilings / 16-bit images
  h->fast_ok = true; h->fast_why.clear();
  auto fin = [&](ConvLayer& L) {
    finish_layer(L, true, h->max_ntile);
    if (!L.n_tile && h->fast_ok) { h->fast_ok = false; h->fast_why = L.name + " (N=" + std::to_string(L.N_total) + ", Cin=" + std::to_string(L.Cin) + ")"; }
  };
  fin(h->conv_pre);
  for (auto& L : h->ups) fin(L);
  for (auto& L : h->ups_zs) if (L.ntaps) finish_layer(L, true);
  for (auto& L : h->ups_pm) if (L.ntaps) { finish_layer(L, true, 256, L.up / 2 * 16); L.wf.clear(); L.wf.shrink_to_fit(); }
  for (auto& L : h->ups_p2) if (L.ntaps) finish_layer(L, true, 128);
  for (auto& L : h->rbconv) fin(L);
  return 0;
}

int rvcdec_load_weights(rvcdec_handle* h, const rvcdec_tensor* tensors, int32_t n) {
  if (!h || !tensors) return fail(RVCDEC_ERR_ARG, "null argument");
  int rc = pack_all(h, tensors, n);
  if (rc) return rc;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    // weights are packed on the host (CPU tests can inspect them) but nothing can run
    h->loaded = false;
    return fail(RVCDEC_ERR_CUDA, "no CUDA device: %s", cudaGetErrorString(e));
  }
  DeviceGuard dguard(h->device);
  CK(dguard.err);
  CK(cudaDeviceSynchronize());  // replayed graphs may still read the old images
  drop_graphs(h);
  ++h->epoch;
  free_device(h);
  if ((rc = upload_layer(h, h->conv_pre))) return rc;
  for (auto& L : h->ups) if ((rc = upload_layer(h, L))) return rc;
  for (auto& L : h->ups_zs) if (L.ntaps && (rc = upload_layer(h, L))) return rc;
  for (auto& L : h->ups_pm) if (L.n_tile && (rc = upload_layer(h, L))) return rc;
  for (auto& L : h->ups_p2) if (L.n_tile && (rc = upload_layer(h, L))) return rc;
  for (auto& L : h->rbconv) if ((rc = upload_layer(h, L))) return rc;
  for (auto& nl : h->noise) {
    if ((rc = upload(h, nl.w, &nl.d_w))) return rc;
    if ((rc = upload(h, nl.bias, &nl.d_bias))) return rc;
  }
  if ((rc = upload(h, h->post_w, &h->d_post_w))) return rc;
  if ((rc = upload(h, h->cond_w, &h->d_cond_w))) return rc;
  if ((rc = upload(h, h->cond_b, &h->d_cond_b))) return rc;
  h->loaded = true;
  return 0;
}

size_t rvcdec_workspace_bytes(const rvcdec_handle* h, int32_t B, int32_t T, int32_t mode) {
  if (!h || B < 1 || T < 1) return 0;
  return carve(h, nullptr, B, T, mode).bytes;
}

static int check_forward_args(rvcdec_handle* h, const void* x, const void* f0, const void* g, const void* out, int B,
                              int T, int mode) {
  if (!h || !x || !f0 || !out) return fail(RVCDEC_ERR_ARG, "null argument");
  if (!h->loaded) return fail(RVCDEC_ERR_STATE, "rvcdec_load_weights has not succeeded on this handle");
  if (B < 1 || T < 1) return fail(RVCDEC_ERR_ARG, "B and T must be >= 1");
  if (mode < 0 || mode > 3) return fail(RVCDEC_ERR_ARG, "unknown mode %d", mode);
  if (h->cfg.gin_channels && !g) { /* g=None is legal in the reference: cond is skipped */ }
  if (mode != 0 && !h->fast_ok) return fail(RVCDEC_ERR_ARG, "fast mode unsupported for this configuration: no tensor-core tiling for %s", h->fast_why.c_str());
  if ((long long)T * h->upp > 0x7fffffffLL / 8) return fail(RVCDEC_ERR_ARG, "T too large");
  return 0;
}

int rvcdec_forward(rvcdec_handle* h, const float* x, const float* f0, const float* g, const float* noise,
                   uint64_t seed, float* out, void* workspace, size_t workspace_bytes, int32_t B, int32_t T,
                   int32_t mode, void* stream) {
  int rc = check_forward_args(h, x, f0, g, out, B, T, mode);
  if (rc) return rc;
  if (!workspace) return fail(RVCDEC_ERR_ARG, "null workspace");
  DeviceGuard dguard(h->device);
  CK(dguard.err);
  const size_t per1 = carve(h, nullptr, 1, T, mode).bytes;
  if (workspace_bytes < per1) return fail(RVCDEC_ERR_WORKSPACE, "workspace of %zu bytes cannot hold one item (%zu needed)", workspace_bytes, per1);
  int Bm = B;
  while (Bm > 1 && carve(h, nullptr, Bm, T, mode).bytes > workspace_bytes) Bm = (Bm + 1) / 2;
  while (Bm < B && carve(h, nullptr, Bm + 1, T, mode).bytes <= workspace_bytes) ++Bm;
  h->launches = 0;
  const size_t n = (size_t)T * h->upp;
  for (int b0 = 0; b0 < B; b0 += Bm) {
    const int nb = std::min(Bm, B - b0);
    rc = run_pass(h, x + (size_t)b0 * h->cfg.initial_channel * T, f0 + (size_t)b0 * T,
                  g ? g + (size_t)b0 * h->cfg.gin_channels : nullptr, noise ? noise + b0 * n * h->H : nullptr, seed,
                  (uint64_t)b0, out + b0 * n, workspace, nb, T, mode, (cudaStream_t)stream);
    if (rc) return rc;
  }
  return 0;
}

int rvcdec_forward_host(rvcdec_handle* h, const float* x, const float* f0, const float* g, const float* noise,
                        uint64_t seed, float* out, int32_t B, int32_t T, int32_t mode, size_t max_device_bytes) {
  int rc = check_forward_args(h, x, f0, g, out, B, T, mode);
  if (rc) return rc;
  const auto t_begin = std::chrono::steady_clock::now();
  DeviceGuard dguard(h->device);
  CK(dguard.err);
  if (!h->stream) CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  if (!h->h2d_stream) {
    CK(cudaStreamCreateWithFlags(&h->h2d_stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      CK(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&h->ev_comp[i], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&h->ev_out[i], cudaEventDisableTiming));
    }
  }
  const bool auto_cap = (max_device_bytes == 0);
  const size_t n = (size_t)T * h->upp;
  const int Cx = h->cfg.initial_channel, G = h->cfg.gin_channels;
  // two sets of input / output staging buffers (double buffering across passes) + one workspace
  auto io_bytes = [&](int b) {
    return align_up((size_t)b * Cx * T * 4, 256) + align_up((size_t)b * T * 4, 256) + align_up((size_t)b * std::max(G, 1) * 4, 256) +
           (noise ? align_up((size_t)b * n * h->H * 4, 256) : 0) + align_up((size_t)b * n * 4, 256);
  };
  int Bm = B;
  // at least three passes on a large batch, so that most of the host<->device traffic hides behind kernels
  if (auto_cap && B >= 6) Bm = (B + 2) / 3;
  if (auto_cap) {
    // 60 % of what is free now (plus what this handle already holds) - but the memory query takes a device-wide lock
    // and was seen to cost milliseconds (tools/latency_bench.py), so it is skipped when the scratch this handle
    // already owns covers the preferred split (the steady state of a per-chunk caller)
    if (2 * io_bytes(Bm) + carve(h, nullptr, Bm, T, mode).bytes <= h->hbuf_bytes) {
      max_device_bytes = h->hbuf_bytes;
    } else {
      size_t free_b = 0, total_b = 0;
      CK(cudaMemGetInfo(&free_b, &total_b));
      max_device_bytes = (size_t)(0.6 * (double)(free_b + h->hbuf_bytes));
    }
  }
  while (Bm > 1 && 2 * io_bytes(Bm) + carve(h, nullptr, Bm, T, mode).bytes > max_device_bytes) Bm = (Bm + 1) / 2;
  while (Bm < B && !(auto_cap && B >= 6 && Bm >= (B + 2) / 3) &&
         2 * io_bytes(Bm + 1) + carve(h, nullptr, Bm + 1, T, mode).bytes <= max_device_bytes) ++Bm;
  const size_t need = 2 * io_bytes(Bm) + carve(h, nullptr, Bm, T, mode).bytes;
  if (need > h->hbuf_bytes) {
    if (h->hbuf) { CK(cudaDeviceSynchronize()); drop_graphs(h); CK(cudaFree(h->hbuf)); h->hbuf = nullptr; h->hbuf_bytes = 0; }
    CK(cudaMalloc(&h->hbuf, need));
    h->hbuf_bytes = need;
  }
  uint8_t* p = (uint8_t*)h->hbuf;
  float *dx[2], *df0[2], *dg[2], *dnz[2] = {nullptr, nullptr}, *dout[2];
  for (int s2 = 0; s2 < 2; ++s2) {
    dx[s2] = (float*)p; p += align_up((size_t)Bm * Cx * T * 4, 256);
    df0[s2] = (float*)p; p += align_up((size_t)Bm * T * 4, 256);
    dg[s2] = (float*)p; p += align_up((size_t)Bm * std::max(G, 1) * 4, 256);
    if (noise) { dnz[s2] = (float*)p; p += align_up((size_t)Bm * n * h->H * 4, 256); }
    dout[s2] = (float*)p; p += align_up((size_t)Bm * n * 4, 256);
  }
  void* ws = p;
  h->launches = 0;
  // The pass loop queues copies on three streams; whatever way it ends, nothing may still be in flight into the
  // caller's buffers when this function returns (the caller is free to release `out` / `x` right away).
  auto run_passes = [&]() -> int {
  int pass = 0;
  for (int b0 = 0; b0 < B; b0 += Bm, ++pass) {
    const int nb = std::min(Bm, B - b0);
    const int s2 = pass & 1;
    // inputs of this pass: the staging set is free once the pass that used it two passes ago has run
    if (pass >= 2) CK(cudaStreamWaitEvent(h->h2d_stream, h->ev_comp[s2], 0));
    CK(cudaMemcpyAsync(dx[s2], x + (size_t)b0 * Cx * T, (size_t)nb * Cx * T * 4, cudaMemcpyHostToDevice, h->h2d_stream));
    CK(cudaMemcpyAsync(df0[s2], f0 + (size_t)b0 * T, (size_t)nb * T * 4, cudaMemcpyHostToDevice, h->h2d_stream));
    if (g && G) CK(cudaMemcpyAsync(dg[s2], g + (size_t)b0 * G, (size_t)nb * G * 4, cudaMemcpyHostToDevice, h->h2d_stream));
    if (noise) CK(cudaMemcpyAsync(dnz[s2], noise + b0 * n * h->H, nb * n * h->H * 4, cudaMemcpyHostToDevice, h->h2d_stream));
    CK(cudaEventRecord(h->ev_in[s2], h->h2d_stream));
    // kernels: need the inputs, and the output staging buffer drained by the copy of two passes ago
    CK(cudaStreamWaitEvent(h->stream, h->ev_in[s2], 0));
    if (pass >= 2) CK(cudaStreamWaitEvent(h->stream, h->ev_out[s2], 0));
    rc = run_pass(h, dx[s2], df0[s2], (g && G) ? dg[s2] : nullptr, dnz[s2], seed, (uint64_t)b0, dout[s2], ws, nb, T, mode, h->stream);
    if (rc) return rc;
    CK(cudaEventRecord(h->ev_comp[s2], h->stream));
    // waveform of this pass back to the host while the next pass computes
    CK(cudaStreamWaitEvent(h->d2h_stream, h->ev_comp[s2], 0));
    CK(cudaMemcpyAsync(out + b0 * n, dout[s2], nb * n * 4, cudaMemcpyDeviceToHost, h->d2h_stream));
    CK(cudaEventRecord(h->ev_out[s2], h->d2h_stream));
  }
  return 0;
  };
  const auto t_setup = std::chrono::steady_clock::now();
  rc = run_passes();
  const auto t_enq = std::chrono::steady_clock::now();
  const cudaError_t e1 = cudaStreamSynchronize(h->h2d_stream), e2 = cudaStreamSynchronize(h->stream),
                    e3 = cudaStreamSynchronize(h->d2h_stream);
  const auto t_end = std::chrono::steady_clock::now();
  h->host_t[0] = std::chrono::duration<double>(t_setup - t_begin).count();
  h->host_t[1] = std::chrono::duration<double>(t_enq - t_setup).count();
  h->host_t[2] = std::chrono::duration<double>(t_end - t_enq).count();
  if (rc) return rc;  // (message of the first failure is kept)
  CK(e1); CK(e2); CK(e3);
  return 0;
}

int rvcdec_set_initial_phases(rvcdec_handle* h, const float* phases) {
  if (!h) return fail(RVCDEC_ERR_ARG, "null handle");
  h->rand_ini = phases;
  return 0;
}

int64_t rvcdec_last_launch_count(const rvcdec_handle* h) { return h ? h->launches : 0; }

int rvcdec_profile_enable(rvcdec_handle* h, int32_t on) {
  if (!h) return fail(RVCDEC_ERR_ARG, "null handle");
  h->prof_on = on != 0;
  for (auto& m : h->prof_marks) h->prof_pool.push_back(m.ev);
  h->prof_marks.clear();
  h->prof_names.clear(); h->prof_ms.clear(); h->prof_flops.clear(); h->prof_bytes.clear(); h->prof_launches.clear();
  return 0;
}

int32_t rvcdec_profile_read(rvcdec_handle* h, rvcdec_profile_entry* out, int32_t max_entries) {
  if (!h || !out) return fail(RVCDEC_ERR_ARG, "null argument");
  // fold the recorded marks into the per-group totals: a mark's span ends at the next mark
  if (!h->prof_marks.empty()) {
    CK(cudaEventSynchronize(h->prof_marks.back().ev));
    for (size_t i = 0; i + 1 < h->prof_marks.size(); ++i) {
      const auto& m = h->prof_marks[i];
      if (m.group < 0) continue;
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, m.ev, h->prof_marks[i + 1].ev));
      h->prof_ms[m.group] += ms; h->prof_flops[m.group] += m.flops; h->prof_bytes[m.group] += m.bytes;
      h->prof_launches[m.group] += 1;
    }
    for (auto& m : h->prof_marks) h->prof_pool.push_back(m.ev);
    h->prof_marks.clear();
  }
  const int32_t n = std::min<int32_t>(max_entries, (int32_t)h->prof_names.size());
  for (int32_t i = 0; i < n; ++i) {
    memset(&out[i], 0, sizeof out[i]);
    strncpy(out[i].name, h->prof_names[i].c_str(), sizeof(out[i].name) - 1);
    out[i].ms = h->prof_ms[i]; out[i].flops = h->prof_flops[i]; out[i].bytes = h->prof_bytes[i];
    out[i].spans = h->prof_launches[i];
  }
  return n;
}

int rvcdec_set_option(rvcdec_handle* h, const char* key, int64_t value) {
  if (!h || !key) return fail(RVCDEC_ERR_ARG, "null argument");
  const std::string k = key;
  ++h->epoch;  // any knob may change the launch sequence: captured graphs of older epochs are never replayed
  if (!h->graphs.empty()) {
    DeviceGuard dguard(h->device);
    CK(dguard.err);
    CK(cudaDeviceSynchronize());
    drop_graphs(h);
  }
  if (k == "graphs") {
    if (value < 0 || value > 1) return fail(RVCDEC_ERR_ARG, "graphs must be 0 or 1");
    h->use_graphs = (int)value;
    return 0;
  }
  if (k == "fused_timing") {  // allocate / reset the timing buffer of the instrumented fused kernel
    DeviceGuard dguard(h->device);
    CK(dguard.err);
    if (value && !h->d_timing) {
      CK(cudaMalloc(&h->d_timing, sizeof(long long) * rvcdec_handle::kTimingMaxLaunches * h->num_sms * 16));
    }
    if (!value && h->d_timing) { cudaFree(h->d_timing); h->d_timing = nullptr; }
    if (h->d_timing) CK(cudaMemset(h->d_timing, 0, sizeof(long long) * rvcdec_handle::kTimingMaxLaunches * h->num_sms * 16));
    h->timing_launches = 0;
    return 0;
  }
  if (k == "no_source") {
    if (value < 0 || value > 1) return fail(RVCDEC_ERR_ARG, "no_source must be 0 or 1");
    h->no_source = (int)value;
    return 0;
  }
  if (k == "conv_epi_warps") {
    if (value != 8 && value != 16) return fail(RVCDEC_ERR_ARG, "conv_epi_warps must be 8 or 16");
    h->conv_epi_warps = (int)value;
    return 0;
  }
  if (k == "conv_epi_warps_small") {
    if (value != 8 && value != 16) return fail(RVCDEC_ERR_ARG, "conv_epi_warps_small must be 8 or 16");
    h->conv_epi_warps_small = (int)value;
    return 0;
  }
  if (k == "conv_lean") { h->conv_lean = value ? 1 : 0; return RVCDEC_OK; }
  if (k == "pm_a_stages") {
    if (value != 2 && value != 3) return fail(RVCDEC_ERR_ARG, "pm_a_stages must be 2 or 3");
    h->pm_a_stages = (int)value;
    return 0;
  }
  if (k == "ups_pm") {
    if (value < 0 || value > 1) return fail(RVCDEC_ERR_ARG, "ups_pm must be 0 or 1");
    h->ups_pm_on = (int)value;
    return 0;
  }
  if (k == "poly2") {
    if (value < 0 || value > 1) return fail(RVCDEC_ERR_ARG, "poly2 must be 0 or 1");
    h->poly2 = (int)value;
    return 0;
  }
  if (k == "zero_stuff") {
    if (value < 0 || value > 1) return fail(RVCDEC_ERR_ARG, "zero_stuff must be 0 or 1");
    h->zero_stuff = (int)value;
    return 0;
  }
  if (k == "max_ntile") {
    if (value != 256 && value != 128) return fail(RVCDEC_ERR_ARG, "max_ntile must be 128 or 256");
    h->max_ntile = (int)value;
    return RVCDEC_OK;
  }
  if (k == "a_stages") {
    if (value != 2 && value != 3) return fail(RVCDEC_ERR_ARG, "a_stages must be 2 or 3");
    h->a_stages = (int)value;
    return RVCDEC_OK;
  }
  if (k == "mrf_items") {
    if (value < 0 || value > (1 << 20)) return fail(RVCDEC_ERR_ARG, "mrf_items out of range");
    h->mrf_items = (int)value;
    return RVCDEC_OK;
  }
  if (k == "fused_split_k") {
    if (value < -1 || value > 64) return fail(RVCDEC_ERR_ARG, "fused_split_k out of range");
    h->fused_split_k = (int)value;
    return RVCDEC_OK;
  }
  if (k == "pair_noise") { h->pair_noise = value ? 1 : 0; return RVCDEC_OK; }
  if (k == "pair_mrf") { h->pair_mrf = value ? 1 : 0; drop_graphs(h); return RVCDEC_OK; }
  if (k == "pair_kernel") {
    if (value != 1 && value != 2) return fail(RVCDEC_ERR_ARG, "pair_kernel must be 1 or 2");
    h->pair_kernel = (int)value; drop_graphs(h); return RVCDEC_OK;
  }
  if (k == "pair_max_k") {
    if (value < 0 || value > 64) return fail(RVCDEC_ERR_ARG, "pair_max_k out of range");
    h->pair_max_k = (int)value;
    return RVCDEC_OK;
  }
  if (k == "fused_epi_warps") {
    if (value != 8 && value != 16) return fail(RVCDEC_ERR_ARG, "fused_epi_warps must be 8 or 16");
    h->fused_epi_warps = (int)value;
    return RVCDEC_OK;
  }
  if (k == "fused_post") { h->fused_post = value ? 1 : 0; return RVCDEC_OK; }
  if (k == "fused_mrf") {
    if (value < 0 || value > 2) return fail(RVCDEC_ERR_ARG, "fused_mrf must be 0, 1 or 2");
    h->fused_mrf = (int)value;
    return 0;
  }
  return fail(RVCDEC_ERR_ARG, "unknown option '%s'", key);
}

double rvcdec_debug_stat(const rvcdec_handle* h, const char* name) {
  if (!h || !name) return -1.0;
  const std::string k = name;
  if (k == "graph_hits") return (double)h->graph_hits;
  if (k == "graph_captures") return (double)h->graph_captures;
  if (k == "graphs_cached") return (double)h->graphs.size();
  if (k == "host_setup_s") return h->host_t[0];
  if (k == "host_enqueue_s") return h->host_t[1];
  if (k == "host_wait_s") return h->host_t[2];
  return -1.0;
}

int rvcdec_debug_set_stop(rvcdec_handle* h, int32_t stop_after) {
  if (!h) return fail(RVCDEC_ERR_ARG, "null handle");
  h->debug_stop = stop_after;
  return 0;
}

int64_t rvcdec_debug_read(rvcdec_handle* h, const char* name, float* dst, int64_t dst_numel, void* workspace,
                          int32_t B, int32_t T, int32_t mode, void* stream) {
  if (!h || !name || !dst || !workspace) return fail(RVCDEC_ERR_ARG, "null argument");
  DeviceGuard dguard(h->device);
  CK(dguard.err);
  Workspace w = carve(h, workspace, B, T, mode);
  const std::string nm = name;
  cudaStream_t s = (cudaStream_t)stream;
  auto need = [&](int64_t n) { return n <= dst_numel ? 0 : fail(RVCDEC_ERR_ARG, "destination too small (%lld < %lld)", (long long)dst_numel, (long long)n); };
  if (nm == "har") {
    const int64_t n = (int64_t)B * T * h->upp;
    if (need(n)) return RVCDEC_ERR_ARG;
    CK(cudaMemcpyAsync(dst, w.har, n * 4, cudaMemcpyDeviceToDevice, s));
    return n;
  }
  if (nm == "pre") {  // leaky_relu(conv_pre(x) + cond(g), 0.1) in the operand format
    const int C0 = h->cfg.upsample_initial_channel;
    const int64_t n = (int64_t)B * C0 * T;
    if (need(n)) return RVCDEC_ERR_ARG;
    CK(launch_unpack(w.pre, fmt_pre(mode), dst, B, C0, T, padded_rows(T), s));
    return n;
  }
  int L = T;
  for (int i = 0; i < h->nups; ++i) {
    L *= h->cfg.upsample_rates[i];
    const int64_t n = (int64_t)B * h->ch[i] * L;
    if (nm == "stage" + std::to_string(i) + ".in") {  // ups + noise_conv, fp32
      if (need(n)) return RVCDEC_ERR_ARG;
      CK(launch_unpack(w.xs0, ACT_F32, dst, B, h->ch[i], L, padded_rows(L), s));
      return n;
    }
    if (nm == "stage" + std::to_string(i) + ".out") {
      if (need(n)) return RVCDEC_ERR_ARG;
      if (i == h->nups - 1) CK(launch_unpack(w.sum, ACT_F32, dst, B, h->ch[i], L, padded_rows(L), s));  // pre-activation
      else if (zs_out_stage(h, mode, i)) CK(launch_unpack(w.a0, fmt_stage(mode, i), dst, B, h->ch[i], L, padded_rows(2 * L), s, 2));  // zero-stuffed rows
      else CK(launch_unpack((i + 1 < h->nups && pm_act(h, mode, i + 1)) ? w.a2 : w.a0, fmt_stage(mode, i), dst, B, h->ch[i], L, padded_rows(L), s));  // leaky_relu(.,0.1)
      return n;
    }
  }
  return fail(RVCDEC_ERR_ARG, "unknown intermediate '%s'", name);
}

// Copies the instrumented fused kernel's spans (see "fused_timing") to the host: returns the number of
// launches recorded; dst gets launches * num_sms * 16 int64 values.
int64_t rvcdec_debug_fused_timing(rvcdec_handle* h, int64_t* dst, int64_t dst_numel) {
  if (!h || !dst) return fail(RVCDEC_ERR_ARG, "null argument");
  if (!h->d_timing) return 0;
  const int64_t n = (int64_t)h->timing_launches * h->num_sms * 16;
  if (dst_numel < n) return fail(RVCDEC_ERR_ARG, "destination too small");
  DeviceGuard dguard(h->device);
  CK(dguard.err);
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(dst, h->d_timing, (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost));
  return h->timing_launches;
}

// Host-only packing for CPU tests: runs fold + repack without touching CUDA.
int rvcdec_debug_pack_host(rvcdec_handle* h, const rvcdec_tensor* tensors, int32_t n) {
  if (!h || !tensors) return fail(RVCDEC_ERR_ARG, "null argument");
  return pack_all(h, tensors, n);
}

int64_t rvcdec_debug_packed(const rvcdec_handle* h, const char* layer, int32_t kind, void* dst, int64_t dst_bytes) {
  if (!h || !layer) return fail(RVCDEC_ERR_ARG, "null argument");
  const ConvLayer* L = nullptr;
  const std::string nm = layer;
  if (nm == h->conv_pre.name) L = &h->conv_pre;
  for (auto& u : h->ups) if (u.name == nm) L = &u;
  for (auto& u : h->ups_pm) if (u.ntaps && u.name == nm) L = &u;
  for (auto& u : h->ups_p2) if (u.ntaps && u.name == nm) L = &u;
  for (auto& r : h->rbconv) if (r.name == nm) L = &r;
  if (!L) return fail(RVCDEC_ERR_ARG, "unknown layer '%s'", layer);
  const void* src = nullptr;
  int64_t bytes = 0;
  int32_t meta[8];
  switch (kind) {
    case 0: src = L->wf.data(); bytes = (int64_t)L->wf.size() * 4; break;
    case 1: src = L->wp[0].data(); bytes = (int64_t)L->wp[0].size() * 2; break;
    case 2: src = L->shift; bytes = (int64_t)L->ntaps * 4; break;
    case 3: src = L->mask_mma.data(); bytes = (int64_t)L->mask_mma.size() * 4; break;
    case 4: src = L->wp[1].data(); bytes = (int64_t)L->wp[1].size() * 2; break;
    case 5:
      meta[0] = L->Cin; meta[1] = L->N_total; meta[2] = L->cout; meta[3] = L->up; meta[4] = L->ntaps; meta[5] = L->halo;
      meta[6] = L->n_tile; meta[7] = L->kc;
      src = meta; bytes = sizeof meta; break;
    case 6: src = L->bias.data(); bytes = (int64_t)L->bias.size() * 4; break;
    default: return fail(RVCDEC_ERR_ARG, "unknown kind %d", kind);
  }
  if (!dst) return bytes;
  if (dst_bytes < bytes) return fail(RVCDEC_ERR_ARG, "destination too small");
  memcpy(dst, src, (size_t)bytes);
  return bytes;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free harmonic source (no counterpart in the reference: added for
// callers that cannot call cudaMalloc - code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// Same arithmetic as source_kernel (aux_kernels.cu; hifigan.py:197-228 + hifigan_nsf.py:48-52, one sine), but with the
// frame-level fp64 phase prefix as a real device-wide scan kept in SCRATCH - here a file-scope device array, so the entry
// point allocates nothing: no cudaMalloc / cudaFree, no workspace argument, no graph, no handle.  Every GPU that loads
// the library holds the array (512 KB) until exit.  All calls share it and it keeps the last call's contents: every
// element a call reads is written by that call first (entry 0 of every item is stored as 0 each time), and calls that
// overlap on the GPU (different streams, several host threads) RACE on it - serialise them.
// ---------------------------------------------------------------------------------------------------------------------
#include "philox.cuh"

constexpr int kNoAllocMaxPrefix = 1 << 16;  // B * (T + 1) <= 65536 phase-prefix entries
static __device__ __align__(256) double g_noalloc_prefix[kNoAllocMaxPrefix];

namespace {

constexpr int NA_THREADS = 256;

// prefix[b][l] = sum over frames j < l of the wrapped per-frame phase advance (cycles, modulo 1); one CTA per item
__global__ void __launch_bounds__(NA_THREADS) noalloc_prefix_kernel(const float* __restrict__ f0, double* __restrict__ prefix, int T,
                                                                    int upp, double inv_sr) {
  __shared__ double chunk_sum[NA_THREADS];
  const int b = blockIdx.x;
  const float* f = f0 + (size_t)b * T;
  double* o = prefix + (size_t)b * (T + 1);
  const int per = (T + NA_THREADS - 1) / NA_THREADS;
  const int j0 = min(threadIdx.x * per, T), j1 = min(j0 + per, T);
  const double dupp = (double)upp;
  double s = 0.0;
  for (int j = j0; j < j1; ++j) {
    const double adv = (double)f[j] * inv_sr * dupp + 0.5;
    s += (adv - floor(adv)) - 0.5;
  }
  chunk_sum[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {  // exclusive scan over the 256 chunk sums
    double acc = 0.0;
    for (int i = 0; i < NA_THREADS; ++i) { const double v = chunk_sum[i]; chunk_sum[i] = acc; acc += v; }
    o[0] = 0.0;  // must start at zero whatever the last call left here
  }
  __syncthreads();
  double c = chunk_sum[threadIdx.x];
  for (int j = j0; j < j1; ++j) {
    const double adv = (double)f[j] * inv_sr * dupp + 0.5;
    c += (adv - floor(adv)) - 0.5;
    o[j + 1] = c - floor(c);
  }
}

__global__ void __launch_bounds__(NA_THREADS) noalloc_source_kernel(const float* __restrict__ f0, const double* __restrict__ prefix,
                                                                    const float* __restrict__ noise, float* __restrict__ har,
                                                                    uint64_t seed, int B, int T, int upp, double inv_sr, float lin_w,
                                                                    float lin_b) {
  const size_t n = (size_t)T * upp, total = (size_t)B * n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / n);
    const size_t s = i - (size_t)b * n;
    const int fr = (int)(s / upp), k = (int)(s - (size_t)fr * upp);
    const float fv = f0[(size_t)b * T + fr];
    const double pc = prefix[(size_t)b * (T + 1) + fr];
    const double ph = (double)fv * inv_sr * (double)(k + 1) + (pc - floor(pc));
    const float frac = (float)(ph - floor(ph));
    const float sine = 0.1f * sinpif(2.0f * frac);
    const float z = noise ? __ldg(noise + i) : philox_normal(seed, (uint64_t)i);
    const float v = fv > 0.f ? (sine + 0.003f * z) : ((0.1f / 3.0f) * z);
    har[i] = tanhf(lin_w * v + lin_b);
  }
}

}  // namespace

extern "C" int rvcdec_source_noalloc(const float* f0, const float* noise, uint64_t seed, float* har, int32_t B, int32_t T, int32_t upp,
                                     float sr, float lin_w, float lin_b, void* stream) {
  if (!f0 || !har) return fail(RVCDEC_ERR_ARG, "null argument");
  if (B < 1 || T < 1 || upp < 1 || !(sr > 0.f)) return fail(RVCDEC_ERR_ARG, "B, T, upp must be >= 1 and sr > 0");
  if ((int64_t)B * ((int64_t)T + 1) > (int64_t)kNoAllocMaxPrefix)
    return fail(RVCDEC_ERR_ARG, "rvcdec_source_noalloc: B * (T + 1) = %lld exceeds the static scratch of %d phase-prefix entries; split the batch "
                "or use rvcdec_forward with a caller workspace", (long long)B * ((long long)T + 1), kNoAllocMaxPrefix);
  // The array's name is a host-side stand-in in host code: its device address is looked up once per device, with that
  // device current (the caller's current device is the one f0 / har live on).
  static double* dev_ptr[64] = {nullptr};
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(RVCDEC_ERR_ARG, "device index %d out of range", dev);
  double* prefix = dev_ptr[dev];
  if (!prefix) {
    CK(cudaGetSymbolAddress((void**)&prefix, g_noalloc_prefix));
    dev_ptr[dev] = prefix;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const double inv_sr = 1.0 / (double)sr;
  noalloc_prefix_kernel<<<B, NA_THREADS, 0, s>>>(f0, prefix, T, upp, inv_sr);
  CK(cudaGetLastError());
  const size_t total = (size_t)B * T * upp;
  const int blocks = (int)std::min<size_t>((total + NA_THREADS - 1) / NA_THREADS, (size_t)(148 * 16));
  noalloc_source_kernel<<<blocks, NA_THREADS, 0, s>>>(f0, prefix, noise, har, seed, B, T, upp, inv_sr, lin_w, lin_b);
  CK(cudaGetLastError());
  return RVCDEC_OK;
}
