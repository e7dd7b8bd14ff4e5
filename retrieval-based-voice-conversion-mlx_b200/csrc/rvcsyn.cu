// rvcsyn_* - C ABI (include/rvcsyn.h) of the whole Synthesizer.infer (rvc/lib/algorithm/synthesizers.py:205-243):
// speaker embedding lookup, TextEncoder, z_p sampling, reverse flow, then the decoder through rvcdec_forward.
// Host side: weight-norm folding / repacking of the front end's layers into the fp32 convolution back-end's images
// (the Flip layers of the flow are folded into the coupling layers' weights) and the launch sequence.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/rvcsyn.h"
#include "common.cuh"

using namespace rvcdec;

namespace rvcdec {
// kernels of synth_front.cu
cudaError_t launch_pack_btc(const float* x, float* out, int B, int T, int C, int Lp, cudaStream_t s);
cudaError_t launch_embed_post(float* x, const long long* pitch, const float* emb_pitch, const int* lengths, int B, int T, int H,
                              int Lp, float scale, void* x16, int fmt, cudaStream_t s);
cudaError_t launch_gather_rows(const float* table, const long long* idx, float* out, int B, int D, int rows, cudaStream_t s);
cudaError_t launch_layernorm(const float* x, float* y, const float* gamma, const float* beta, const int* lengths, int B, int T, int C,
                             int Lp, float eps, void* y16, int fmt, cudaStream_t s);
cudaError_t launch_gate(const float* a, float* acts, int B, int T, int H, int Lp, void* acts16, int fmt, cudaStream_t s);
cudaError_t launch_zp(const float* stats, const float* noise, float* zp, const int* lengths, uint64_t seed, uint64_t item0, int B,
                      int T, int C, int Lp, void* zp16, int fmt, cudaStream_t s);
cudaError_t launch_att_pack_rel(const float* ek, const float* ev, void* ek_img, void* ev_img, int fmt, cudaStream_t s);
int att_mma_rel_bytes();
cudaError_t launch_attention_mma(const void* qkv16, float* out, void* out16, const void* ek_img, const void* ev_img, const int* lengths,
                                 int B, int T, int heads, int kc, int window, int Lp, int fmt, cudaStream_t s);
cudaError_t launch_cast16(const float* x, void* y16, int fmt, int B, int T, int C, int Lp, cudaStream_t s);
cudaError_t launch_mask_rows(float* x32, int phys32, void* x16, int phys16, const int* lengths, int B, int T, int C, int Lp, cudaStream_t s);
cudaError_t launch_attention(const float* qkv, float* out, const float* emb_rel_k, const float* emb_rel_v, const int* lengths, int B,
                             int T, int heads, int kc, int window, int Lp, cudaStream_t s);
}  // namespace rvcdec

namespace {

thread_local std::string g_serr;

int sfail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_serr = buf;
  return code;
}

#define SCK(expr)                                                                                  \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return sfail(RVCDEC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct DevGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DevGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) err = cudaSetDevice(dev); else prev = -1;
  }
  ~DevGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// One convolution of the front end in the fp32 back-end's image: wf[tap][ci][n], shifts, bias.
struct FConv {
  int Cin = 0, N = 0, ntaps = 0, halo = 0;
  int shift[MAX_TAPS] = {0};
  std::vector<float> wf, bias;
  float* d_wf = nullptr;
  float* d_bias = nullptr;
  // tensor-core image (conv_mma.cu; mode != 0): [N tile][tap][K chunk][slice kc/8][n n_tile][8], fp16 and bf16
  int n_tile = 0, kc = 0;
  std::vector<uint16_t> wp[2];
  uint16_t* d_wp[2] = {nullptr, nullptr};
};

uint16_t f_to16(float v, int fmt) {
  if (fmt == 1) { __nv_bfloat16 b = __float2bfloat16_rn(v); return *reinterpret_cast<uint16_t*>(&b); }
  __half hv = __float2half_rn(v);
  return *reinterpret_cast<uint16_t*>(&hv);
}

// Tiling of the tensor-core back-end for this layer (widest N tile that divides N; 32-wide tiles with 32-channel chunks
// for the coupling layers' 96-channel halves), and the packed images.
void pack_mma(FConv& L) {
  L.n_tile = 0;
  if (L.Cin % 64 == 0 && L.N % 128 == 0) { L.n_tile = 128; L.kc = 64; }
  else if (L.Cin % 64 == 0 && L.N % 64 == 0) { L.n_tile = 64; L.kc = 64; }
  else if (L.Cin % 32 == 0 && L.N % 32 == 0) { L.n_tile = 32; L.kc = 32; }
  if (!L.n_tile || L.N / L.n_tile > MAX_NTILES || !conv_mma_supported(L.n_tile, L.kc)) { L.n_tile = 0; return; }
  const int nt = L.n_tile, kc = L.kc, nkc = L.Cin / kc, ntiles = L.N / nt;
  for (int fmt = 0; fmt < 2; ++fmt) {
    std::vector<uint16_t>& wp = L.wp[fmt];
    wp.assign((size_t)L.ntaps * L.Cin * L.N, 0);
    size_t o = 0;
    for (int tile = 0; tile < ntiles; ++tile)
      for (int t = 0; t < L.ntaps; ++t)
        for (int c = 0; c < nkc; ++c)
          for (int sl = 0; sl < kc / 8; ++sl)
            for (int n = 0; n < nt; ++n)
              for (int e = 0; e < 8; ++e) wp[o++] = f_to16(L.wf[((size_t)t * L.Cin + c * kc + sl * 8 + e) * L.N + tile * nt + n], fmt);
  }
}

struct TMap {
  std::map<std::string, const rvcdec_tensor*> m;
  const rvcdec_tensor* find(const std::string& n) const {
    auto it = m.find(n);
    return it == m.end() ? nullptr : it->second;
  }
};

int get_w(const TMap& tm, const std::string& base, int64_t d0, int64_t rest, std::vector<float>& out) {
  out.resize((size_t)d0 * rest);
  if (const rvcdec_tensor* w = tm.find(base + ".weight")) {
    if (w->numel != d0 * rest) return sfail(RVCDEC_ERR_WEIGHTS, "%s.weight: expected %lld elements, got %lld", base.c_str(), (long long)(d0 * rest), (long long)w->numel);
    memcpy(out.data(), w->data, sizeof(float) * out.size());
    return 0;
  }
  const rvcdec_tensor* g = tm.find(base + ".parametrizations.weight.original0");
  const rvcdec_tensor* v = tm.find(base + ".parametrizations.weight.original1");
  if (!g || !v) { g = tm.find(base + ".weight_g"); v = tm.find(base + ".weight_v"); }
  if (!g || !v) return sfail(RVCDEC_ERR_WEIGHTS, "missing weight for %s", base.c_str());
  if (g->numel != d0 || v->numel != d0 * rest) return sfail(RVCDEC_ERR_WEIGHTS, "%s: weight-norm tensors have wrong size", base.c_str());
  for (int64_t i = 0; i < d0; ++i) {  // w = v * g / ||v||  (torch weight_norm dim 0; modules.py:44-47, 61-80)
    const float* vi = v->data + i * rest;
    double ss = 0.0;
    for (int64_t k = 0; k < rest; ++k) ss += (double)vi[k] * (double)vi[k];
    const float sc = (float)((double)g->data[i] / std::sqrt(ss));
    for (int64_t k = 0; k < rest; ++k) out[(size_t)(i * rest + k)] = vi[k] * sc;
  }
  return 0;
}

int get_v(const TMap& tm, const std::string& name, int64_t n, std::vector<float>& out) {
  const rvcdec_tensor* t = tm.find(name);
  if (!t) return sfail(RVCDEC_ERR_WEIGHTS, "missing tensor %s", name.c_str());
  if (t->numel != n) return sfail(RVCDEC_ERR_WEIGHTS, "%s: expected %lld elements, got %lld", name.c_str(), (long long)n, (long long)t->numel);
  out.assign(t->data, t->data + n);
  return 0;
}

// Conv1d / Linear weight W (Cout, Cin, k) -> FConv image; rows [co0, co0 + Cout_sub) of W only; optional reversal of the
// input / output channel order (the flow's Flip layers folded into the neighbouring 1x1 convolutions).
void pack_fconv(FConv& L, const std::vector<float>& W, const std::vector<float>& b, int Cout_total, int Cin, int k, int d,
                int co0, int Cout_sub, bool rev_in, bool rev_out) {
  L.Cin = Cin; L.N = Cout_sub; L.ntaps = k; L.halo = (k - 1) / 2 * d;
  L.wf.assign((size_t)k * Cin * Cout_sub, 0.f);
  L.bias.assign(Cout_sub, 0.f);
  (void)Cout_total;
  for (int j = 0; j < k; ++j) {
    L.shift[j] = (j - (k - 1) / 2) * d;
    for (int ci = 0; ci < Cin; ++ci)
      for (int co = 0; co < Cout_sub; ++co) {
        const int sci = rev_in ? Cin - 1 - ci : ci;
        const int sco = co0 + (rev_out ? Cout_sub - 1 - co : co);
        L.wf[((size_t)j * Cin + ci) * Cout_sub + co] = W[((size_t)sco * Cin + sci) * k + j];
      }
  }
  for (int co = 0; co < Cout_sub; ++co) L.bias[co] = b.empty() ? 0.f : b[co0 + (rev_out ? Cout_sub - 1 - co : co)];
}

struct AttnLayer {
  FConv qkv, o, ffn1, ffn2;
  std::vector<float> ek, ev, g1, b1, g2, b2;
  uint8_t* d_rel16[2] = {nullptr, nullptr};  // [fmt] Ek | Ev as 16-bit B operand images of att_mma.cu
  float *d_ek = nullptr, *d_ev = nullptr, *d_g1 = nullptr, *d_b1 = nullptr, *d_g2 = nullptr, *d_b2 = nullptr;
};

struct FlowLayer {
  bool flipped = false;  // operates in the flipped frame: reads the upper channel half, updates the lower one
  FConv pre, post;
  std::vector<FConv> in, res, skip;  // per WaveNet layer (res is empty for the last one)
  std::vector<float> cond_w, cond_b, in_bias_all;  // cond_layer (2*H*n, gin), its bias, concatenated in_layer biases
  float *d_cond_w = nullptr, *d_cond_b = nullptr, *d_in_bias_all = nullptr;
};

}  // namespace

struct rvcsyn_handle {
  rvcsyn_config cfg{};
  int device = 0;
  rvcdec_handle* dec = nullptr;
  bool loaded = false;
  FConv emb_phone, proj;
  std::vector<float> emb_pitch, emb_g;
  float *d_emb_pitch = nullptr, *d_emb_g = nullptr;
  std::vector<AttnLayer> attn;
  std::vector<FlowLayer> flow;
  std::vector<void*> allocs;
  int64_t launches = 0;
  int num_sms = 148;
  // CUDA-graph replay of the front end's launch sequence (small batches)
  struct GraphEntry {
    const void* ws;
    int B, T, mode;
    bool has_pitch;
    uint64_t epoch;
    cudaGraphExec_t exec;
    int64_t launches;
    uint64_t last_use;
  };
  std::vector<GraphEntry> graphs;
  uint64_t epoch = 0, graph_clock = 0, graph_hits = 0, graph_captures = 0;
  cudaStream_t cap_stream = nullptr;
  int use_graphs = 1;                // option "graphs"
  int64_t graph_max_frames = 16384;  // option "graph_max_frames": B * T up to which the sequence is graphed
  int att_mma = 1;     // option "att_mma": 1 = fast modes run the attention on the tensor cores (att_mma.cu), 0 = fp32 kernel
  int front_fp32 = 0;  // option "front_fp32": 1 = the front end stays on the fp32 back-end in every mode
  int debug_stop = 0;  // tests: 1 embed, 2 qkv of layer 0, 3 attention of layer 0, 4 first LayerNorm, 5 layer 0 done,
                       // 6 encoder done (stats), 7 z_p, 8 first coupling layer done, 9 flow done
};

namespace {

template <typename T>
int up(rvcsyn_handle* h, const std::vector<T>& v, T** d) {
  *d = nullptr;
  if (v.empty()) return 0;
  void* p = nullptr;
  SCK(cudaMalloc(&p, v.size() * sizeof(T)));
  h->allocs.push_back(p);
  SCK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  *d = (T*)p;
  return 0;
}
int up_conv(rvcsyn_handle* h, FConv& L) {
  int rc;
  if ((rc = up(h, L.wf, &L.d_wf))) return rc;
  pack_mma(L);
  for (int f = 0; f < 2; ++f)
    if ((rc = up(h, L.wp[f], &L.d_wp[f]))) return rc;
  return up(h, L.bias, &L.d_bias);
}

struct SynWs {
  float *phone, *x, *qkv, *att, *tmp, *h1, *stats, *xf, *hh, *a, *acts, *outacc, *g, *condB, *z_nct;
  // 16-bit operand copies for the tensor-core front end (mode != 0), same blocked-8 layout
  uint16_t *x16, *att16, *h116, *xf16, *hh16, *acts16, *out16, *qkv16;
  // fixed places for the inputs of the graphed front end
  float *in_phone, *in_nzp;
  int32_t* in_len;
  int64_t *in_pitch, *in_sid;
  size_t bytes;
};

SynWs carve_syn(const rvcsyn_handle* h, void* base, int B, int T) {
  const rvcsyn_config& c = h->cfg;
  SynWs w{};
  size_t off = 0;
  const size_t Lp = (size_t)padded_rows(T);
  auto take = [&](size_t n) { size_t o = off; off += align_up(n, 256); return base ? (float*)((uint8_t*)base + o) : (float*)nullptr; };
  const int H = c.hidden_channels, I = c.inter_channels;
  w.phone = take((size_t)B * c.text_dim * Lp * 4);
  w.x = take((size_t)B * H * Lp * 4);
  w.qkv = take((size_t)B * 3 * H * Lp * 4);
  w.att = take((size_t)B * H * Lp * 4);
  w.tmp = take((size_t)B * H * Lp * 4);
  w.h1 = take((size_t)B * c.filter_channels * Lp * 4);
  w.stats = take((size_t)B * 2 * I * Lp * 4);
  w.xf = take((size_t)B * I * Lp * 4);
  w.hh = take((size_t)B * H * Lp * 4);
  w.a = take((size_t)B * 2 * H * Lp * 4);
  w.acts = take((size_t)B * H * Lp * 4);
  w.outacc = take((size_t)B * H * Lp * 4);
  w.g = take((size_t)B * std::max(c.gin_channels, 1) * 4);
  w.condB = take((size_t)B * 2 * H * c.flow_layers * 4);
  w.z_nct = take((size_t)B * I * T * 4);
  w.x16 = (uint16_t*)take((size_t)B * H * Lp * 2);
  w.att16 = (uint16_t*)take((size_t)B * H * Lp * 2);
  w.h116 = (uint16_t*)take((size_t)B * c.filter_channels * Lp * 2);
  w.xf16 = (uint16_t*)take((size_t)B * I * Lp * 2);
  w.hh16 = (uint16_t*)take((size_t)B * H * Lp * 2);
  w.acts16 = (uint16_t*)take((size_t)B * H * Lp * 2);
  w.out16 = (uint16_t*)take((size_t)B * H * Lp * 2);
  w.qkv16 = (uint16_t*)take((size_t)B * 3 * H * Lp * 2);
  w.in_phone = take((size_t)B * T * c.text_dim * 4);
  w.in_nzp = take((size_t)B * I * T * 4);
  w.in_len = (int32_t*)take((size_t)B * 4);
  w.in_pitch = (int64_t*)take((size_t)B * T * 8);
  w.in_sid = (int64_t*)take((size_t)B * 8);
  w.bytes = off;
  return w;
}

// y = conv(x) through the fp32 back-end.  in_phys / out_phys: slices per item of the tensors the pointers point into.
int fconv(rvcsyn_handle* h, const FConv& L, const float* in, int in_phys_slices, int B, int T, const Epilogue& e, cudaStream_t s) {
  ConvParams p{};
  p.act_in = in; p.act_fmt_in = ACT_F32; p.w = L.d_wf;
  p.B = B; p.Cin = L.Cin; p.L_in = T; p.Lp_in = padded_rows(T);
  p.N_total = L.N; p.ntaps = L.ntaps; p.halo = L.halo;
  for (int t = 0; t < L.ntaps; ++t) p.shift[t] = L.shift[t];
  const int ntiles = (L.N + 63) / 64;
  if (ntiles > MAX_NTILES) return sfail(RVCDEC_ERR_ARG, "front-end convolution with %d columns exceeds the back-end's tap masks", L.N);
  for (int i = 0; i < ntiles; ++i) p.tapmask[i] = (1u << L.ntaps) - 1u;
  p.in_slices_phys = in_phys_slices;
  p.epi = e;
  SCK(launch_conv_f32(p, s));
  ++h->launches;
  return 0;
}

// the same through the tensor-core back-end: 16-bit blocked-8 input in format fmt (the epilogue knows nothing of lengths:
// the caller masks the rows beyond an item's length with launch_mask_rows where the reference multiplies by x_mask)
int mconv(rvcsyn_handle* h, const FConv& L, const void* in16, int fmt, int in_phys_slices, int B, int T, const Epilogue& e, cudaStream_t s) {
  if (!L.n_tile) return sfail(RVCDEC_ERR_ARG, "front-end convolution %d -> %d has no tensor-core tiling", L.Cin, L.N);
  ConvParams p{};
  p.act_in = in16; p.act_fmt_in = fmt; p.w = L.d_wp[fmt == ACT_BF16 ? 1 : 0];
  p.B = B; p.Cin = L.Cin; p.L_in = T; p.Lp_in = padded_rows(T);
  p.N_total = L.N; p.ntaps = L.ntaps; p.halo = L.halo;
  for (int t = 0; t < L.ntaps; ++t) p.shift[t] = L.shift[t];
  for (int i = 0; i < L.N / L.n_tile; ++i) p.tapmask[i] = (1u << L.ntaps) - 1u;
  p.in_slices_phys = in_phys_slices;
  p.a_stages = 2; p.epi_warps = 16; p.epi_warps_small = 8; p.lean = 1;
  p.epi = e;
  p.epi.lengths = nullptr;
  SCK(launch_conv_mma(p, L.n_tile, L.kc, h->num_sms, s));
  ++h->launches;
  return 0;
}

Epilogue epi_basic(const float* bias, int cout_phys, int T, const int* lengths) {
  Epilogue e{};
  e.bias = bias; e.scale = 1.f; e.slope = 0.f; e.up = 1; e.cout = cout_phys; e.L_out = T; e.Lp_out = padded_rows(T);
  e.act_fmt = ACT_F32; e.lengths = lengths;
  return e;
}

}  // namespace

extern "C" {

const char* rvcsyn_last_error(void) { return g_serr.c_str(); }

int rvcsyn_create(const rvcsyn_config* cfg, int device, rvcsyn_handle** out) {
  if (!cfg || !out) return sfail(RVCDEC_ERR_ARG, "null argument");
  *out = nullptr;
  const rvcsyn_config& c = *cfg;
  if (c.hidden_channels % 8 || c.inter_channels % 16 || c.filter_channels % 8 || c.text_dim % 8 || c.hidden_channels <= 0)
    return sfail(RVCDEC_ERR_ARG, "channel counts must be multiples of 8 (inter_channels of 16)");
  if (c.n_heads < 1 || c.hidden_channels % c.n_heads || c.hidden_channels / c.n_heads != 96 || c.window_size != 10)
    return sfail(RVCDEC_ERR_ARG, "the attention kernel is built for 96 channels per head and window 10 (got %d heads over %d channels, window %d)",
                 c.n_heads, c.hidden_channels, c.window_size);
  if (c.kernel_size < 1 || c.kernel_size % 2 == 0 || c.kernel_size > MAX_TAPS || c.flow_kernel % 2 == 0 || c.flow_kernel > MAX_TAPS ||
      c.n_layers < 1 || c.flow_layers < 1 || c.n_flows < 1 || c.n_flows % 2)
    return sfail(RVCDEC_ERR_ARG, "unsupported kernel sizes / layer counts (odd kernels <= %d, an even number of flows)", MAX_TAPS);
  // dec.num_upsamples == 0: front end only (the caller runs another decoder on z * x_mask, e.g. the RefineGAN sibling)
  rvcdec_handle* dec = nullptr;
  if (c.dec.num_upsamples != 0) {
    if (c.dec.initial_channel != c.inter_channels || c.dec.gin_channels != c.gin_channels)
      return sfail(RVCDEC_ERR_ARG, "decoder configuration does not match inter_channels / gin_channels");
    int rc = rvcdec_create(&c.dec, device, &dec);
    if (rc) return sfail(rc, "decoder: %s", rvcdec_last_error());
  }
  auto h = new rvcsyn_handle();
  h->cfg = c; h->device = device; h->dec = dec;
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sms > 0) h->num_sms = sms;
  *out = h;
  return 0;
}

void rvcsyn_destroy(rvcsyn_handle* h) {
  if (!h) return;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) == cudaSuccess && h->device < ndev) {
    DevGuard dg(h->device);
    for (auto& e : h->graphs) cudaGraphExecDestroy(e.exec);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    for (void* p : h->allocs) cudaFree(p);
  }
  if (h->dec) rvcdec_destroy(h->dec);
  delete h;
}

rvcdec_handle* rvcsyn_decoder(rvcsyn_handle* h) { return h ? h->dec : nullptr; }

int rvcsyn_load_weights(rvcsyn_handle* h, const rvcdec_tensor* tensors, int32_t n) {
  if (!h || !tensors) return sfail(RVCDEC_ERR_ARG, "null argument");
  const rvcsyn_config& c = h->cfg;
  TMap tm;
  std::vector<rvcdec_tensor> dec_t;
  for (int i = 0; i < n; ++i) {
    const std::string name = tensors[i].name ? tensors[i].name : "";
    tm.m[name] = &tensors[i];
    if (name.rfind("dec.", 0) == 0) dec_t.push_back(tensors[i]);
  }
  int rc;
  const int H = c.hidden_channels, I = c.inter_channels, F = c.filter_channels, K = c.kernel_size;
  std::vector<float> W, b;
  // enc_p.emb_phone: Linear(text_dim, H)  (encoders.py:119)
  if ((rc = get_w(tm, "enc_p.emb_phone", H, c.text_dim, W)) || (rc = get_v(tm, "enc_p.emb_phone.bias", H, b))) return rc;
  pack_fconv(h->emb_phone, W, b, H, c.text_dim, 1, 1, 0, H, false, false);
  h->emb_pitch.clear();
  if (c.use_f0 && (rc = get_v(tm, "enc_p.emb_pitch.weight", (int64_t)256 * H, h->emb_pitch))) return rc;
  if ((rc = get_v(tm, "emb_g.weight", (int64_t)c.spk_embed_dim * c.gin_channels, h->emb_g))) return rc;
  h->attn.assign(c.n_layers, AttnLayer());
  for (int l = 0; l < c.n_layers; ++l) {
    AttnLayer& A = h->attn[l];
    const std::string ap = "enc_p.encoder.attn_layers." + std::to_string(l), fp = "enc_p.encoder.ffn_layers." + std::to_string(l);
    // q | k | v as one convolution with 3H output channels (attentions.py:81)
    std::vector<float> Wq, Wk, Wv, bq, bk, bv;
    if ((rc = get_w(tm, ap + ".conv_q", H, H, Wq)) || (rc = get_w(tm, ap + ".conv_k", H, H, Wk)) || (rc = get_w(tm, ap + ".conv_v", H, H, Wv)) ||
        (rc = get_v(tm, ap + ".conv_q.bias", H, bq)) || (rc = get_v(tm, ap + ".conv_k.bias", H, bk)) || (rc = get_v(tm, ap + ".conv_v.bias", H, bv)))
      return rc;
    W = Wq; W.insert(W.end(), Wk.begin(), Wk.end()); W.insert(W.end(), Wv.begin(), Wv.end());
    b = bq; b.insert(b.end(), bk.begin(), bk.end()); b.insert(b.end(), bv.begin(), bv.end());
    pack_fconv(A.qkv, W, b, 3 * H, H, 1, 1, 0, 3 * H, false, false);
    if ((rc = get_w(tm, ap + ".conv_o", H, H, W)) || (rc = get_v(tm, ap + ".conv_o.bias", H, b))) return rc;
    pack_fconv(A.o, W, b, H, H, 1, 1, 0, H, false, false);
    const int64_t nrel = (int64_t)(2 * c.window_size + 1) * (H / c.n_heads);
    if ((rc = get_v(tm, ap + ".emb_rel_k", nrel, A.ek)) || (rc = get_v(tm, ap + ".emb_rel_v", nrel, A.ev))) return rc;
    if ((rc = get_w(tm, fp + ".conv_1", F, (int64_t)H * K, W)) || (rc = get_v(tm, fp + ".conv_1.bias", F, b))) return rc;
    pack_fconv(A.ffn1, W, b, F, H, K, 1, 0, F, false, false);
    if ((rc = get_w(tm, fp + ".conv_2", H, (int64_t)F * K, W)) || (rc = get_v(tm, fp + ".conv_2.bias", H, b))) return rc;
    pack_fconv(A.ffn2, W, b, H, F, K, 1, 0, H, false, false);
    const std::string n1 = "enc_p.encoder.norm_layers_1." + std::to_string(l), n2 = "enc_p.encoder.norm_layers_2." + std::to_string(l);
    if ((rc = get_v(tm, n1 + ".gamma", H, A.g1)) || (rc = get_v(tm, n1 + ".beta", H, A.b1)) || (rc = get_v(tm, n2 + ".gamma", H, A.g2)) ||
        (rc = get_v(tm, n2 + ".beta", H, A.b2)))
      return rc;
  }
  if ((rc = get_w(tm, "enc_p.proj", 2 * I, H, W)) || (rc = get_v(tm, "enc_p.proj.bias", 2 * I, b))) return rc;
  pack_fconv(h->proj, W, b, 2 * I, H, 1, 1, 0, 2 * I, false, false);
  // flow: ResidualCouplingLayer i sits at flows[2 i]; reverse order, every one preceded by a Flip (residuals.py:162-165).
  // Flips cancel in pairs: layers n_flows-1, n_flows-3, ... run in the flipped frame with channel-reversed 1x1 convs.
  h->flow.assign(c.n_flows, FlowLayer());
  const int half = I / 2, NL = c.flow_layers, FK = c.flow_kernel;
  for (int i = 0; i < c.n_flows; ++i) {
    FlowLayer& Fl = h->flow[i];
    Fl.flipped = ((c.n_flows - 1 - i) % 2 == 0);
    const std::string fp = "flow.flows." + std::to_string(2 * i);
    if ((rc = get_w(tm, fp + ".pre", H, half, W)) || (rc = get_v(tm, fp + ".pre.bias", H, b))) return rc;
    pack_fconv(Fl.pre, W, b, H, half, 1, 1, 0, H, Fl.flipped, false);
    if ((rc = get_w(tm, fp + ".post", half, H, W)) || (rc = get_v(tm, fp + ".post.bias", half, b))) return rc;
    pack_fconv(Fl.post, W, b, half, H, 1, 1, 0, half, false, Fl.flipped);
    if (c.gin_channels) {
      if ((rc = get_w(tm, fp + ".enc.cond_layer", (int64_t)2 * H * NL, c.gin_channels, Fl.cond_w)) ||
          (rc = get_v(tm, fp + ".enc.cond_layer.bias", (int64_t)2 * H * NL, Fl.cond_b)))
        return rc;
    }
    Fl.in.assign(NL, FConv()); Fl.res.assign(NL, FConv()); Fl.skip.assign(NL, FConv());
    Fl.in_bias_all.clear();
    for (int l = 0; l < NL; ++l) {
      const std::string ip = fp + ".enc.in_layers." + std::to_string(l), rp = fp + ".enc.res_skip_layers." + std::to_string(l);
      if ((rc = get_w(tm, ip, 2 * H, (int64_t)H * FK, W)) || (rc = get_v(tm, ip + ".bias", 2 * H, b))) return rc;
      pack_fconv(Fl.in[l], W, std::vector<float>(), 2 * H, H, FK, 1, 0, 2 * H, false, false);  // bias enters per item (with cond)
      Fl.in_bias_all.insert(Fl.in_bias_all.end(), b.begin(), b.end());
      const int rs = (l == NL - 1) ? H : 2 * H;
      if ((rc = get_w(tm, rp, rs, H, W)) || (rc = get_v(tm, rp + ".bias", rs, b))) return rc;
      if (l < NL - 1) {
        pack_fconv(Fl.res[l], W, b, rs, H, 1, 1, 0, H, false, false);
        pack_fconv(Fl.skip[l], W, b, rs, H, 1, 1, H, H, false, false);
      } else {
        pack_fconv(Fl.skip[l], W, b, rs, H, 1, 1, 0, H, false, false);
      }
    }
  }
  // decoder
  if (h->dec) {
    rc = rvcdec_load_weights(h->dec, dec_t.data(), (int32_t)dec_t.size());
    if (rc) return sfail(rc, "decoder: %s", rvcdec_last_error());
  }
  DevGuard dg(h->device);
  SCK(dg.err);
  for (auto& e : h->graphs) cudaGraphExecDestroy(e.exec);  // they hold the old weight pointers
  h->graphs.clear();
  ++h->epoch;
  for (void* p : h->allocs) cudaFree(p);
  h->allocs.clear();
  if ((rc = up_conv(h, h->emb_phone)) || (rc = up_conv(h, h->proj)) || (rc = up(h, h->emb_pitch, &h->d_emb_pitch)) ||
      (rc = up(h, h->emb_g, &h->d_emb_g)))
    return rc;
  for (auto& A : h->attn) {
    if ((rc = up_conv(h, A.qkv)) || (rc = up_conv(h, A.o)) || (rc = up_conv(h, A.ffn1)) || (rc = up_conv(h, A.ffn2)) ||
        (rc = up(h, A.ek, &A.d_ek)) || (rc = up(h, A.ev, &A.d_ev)) || (rc = up(h, A.g1, &A.d_g1)) || (rc = up(h, A.b1, &A.d_b1)) ||
        (rc = up(h, A.g2, &A.d_g2)) || (rc = up(h, A.b2, &A.d_b2)))
      return rc;
  }
  for (auto& A : h->attn)
    for (int f = 0; f < 2; ++f) {  // Ek | Ev as 16-bit B operand images (att_mma.cu)
      void* img = nullptr;
      SCK(cudaMalloc(&img, 2 * (size_t)att_mma_rel_bytes()));
      h->allocs.push_back(img);
      A.d_rel16[f] = (uint8_t*)img;
      SCK(launch_att_pack_rel(A.d_ek, A.d_ev, img, (uint8_t*)img + att_mma_rel_bytes(), f == 1 ? ACT_BF16 : ACT_F16, 0));
    }
  SCK(cudaDeviceSynchronize());
  for (auto& Fl : h->flow) {
    if ((rc = up_conv(h, Fl.pre)) || (rc = up_conv(h, Fl.post)) || (rc = up(h, Fl.cond_w, &Fl.d_cond_w)) ||
        (rc = up(h, Fl.cond_b, &Fl.d_cond_b)) || (rc = up(h, Fl.in_bias_all, &Fl.d_in_bias_all)))
      return rc;
    for (auto& L : Fl.in) if ((rc = up_conv(h, L))) return rc;
    for (auto& L : Fl.res) if (L.N && (rc = up_conv(h, L))) return rc;
    for (auto& L : Fl.skip) if ((rc = up_conv(h, L))) return rc;
  }
  h->loaded = true;
  return 0;
}

size_t rvcsyn_workspace_bytes(const rvcsyn_handle* h, int32_t B, int32_t T, int32_t mode) {
  if (!h || B < 1 || T < 1) return 0;
  return align_up(carve_syn(h, nullptr, B, T).bytes, 256) + (h->dec ? rvcdec_workspace_bytes(h->dec, B, T, mode) : 0);
}

int rvcsyn_set_option(rvcsyn_handle* h, const char* key, int64_t value) {
  if (!h || !key) return sfail(RVCDEC_ERR_ARG, "null argument");
  const std::string k = key;
  ++h->epoch;  // a cached graph was captured under the old options
  if (k == "front_fp32") { h->front_fp32 = value ? 1 : 0; return RVCDEC_OK; }
  if (k == "att_mma") { h->att_mma = value ? 1 : 0; return RVCDEC_OK; }
  if (k == "graphs") { h->use_graphs = value ? 1 : 0; return RVCDEC_OK; }
  if (k == "graph_max_frames") { h->graph_max_frames = value; return RVCDEC_OK; }
  return sfail(RVCDEC_ERR_ARG, "unknown option '%s'", key);
}

int rvcsyn_debug_set_stop(rvcsyn_handle* h, int32_t stop) {
  if (!h) return sfail(RVCDEC_ERR_ARG, "null handle");
  h->debug_stop = stop;
  return 0;
}

// Copies a named intermediate of the last call out of the workspace as plain (B, C, T) fp32 (device): "x" (encoder
// stream, H), "qkv" (3H), "att" (H), "stats" (2*inter = [m_p | logs_p]), "xf" (flow stream, inter), "hh" (H).
int64_t rvcsyn_debug_read(rvcsyn_handle* h, const char* name, float* dst, int64_t dst_numel, void* workspace, int32_t B, int32_t T,
                          void* stream) {
  if (!h || !name || !dst || !workspace) return sfail(RVCDEC_ERR_ARG, "null argument");
  DevGuard dg(h->device);
  SCK(dg.err);
  const rvcsyn_config& c = h->cfg;
  SynWs w = carve_syn(h, workspace, B, T);
  const std::string nm = name;
  const float* src = nullptr;
  int C = 0;
  if (nm == "x") { src = w.x; C = c.hidden_channels; }
  else if (nm == "qkv") { src = w.qkv; C = 3 * c.hidden_channels; }
  else if (nm == "att") { src = w.att; C = c.hidden_channels; }
  else if (nm == "stats") { src = w.stats; C = 2 * c.inter_channels; }
  else if (nm == "xf") { src = w.xf; C = c.inter_channels; }
  else if (nm == "hh") { src = w.hh; C = c.hidden_channels; }
  else return sfail(RVCDEC_ERR_ARG, "unknown intermediate '%s'", name);
  const int64_t n = (int64_t)B * C * T;
  if (dst_numel < n) return sfail(RVCDEC_ERR_ARG, "destination too small");
  SCK(launch_unpack(src, ACT_F32, dst, B, C, T, padded_rows(T), (cudaStream_t)stream));
  return n;
}

int64_t rvcsyn_last_launch_count(const rvcsyn_handle* h) { return h ? h->launches + (h->dec ? rvcdec_last_launch_count(h->dec) : 0) : 0; }

int rvcsyn_infer(rvcsyn_handle* h, const float* phone, const int32_t* lengths, const int64_t* pitch, const float* nsff0,
                 const int64_t* sid, const float* noise_zp, const float* noise_src, uint64_t seed, float* out, float* z_out,
                 void* workspace, size_t workspace_bytes, int32_t B, int32_t T, int32_t mode, void* stream) {
  if (!h || !phone || !lengths || !sid || !workspace) return sfail(RVCDEC_ERR_ARG, "null argument");
  if (h->dec ? !out : !z_out) return sfail(RVCDEC_ERR_ARG, h->dec ? "null output" : "a front-end-only handle needs z_out");
  if (!h->loaded) return sfail(RVCDEC_ERR_STATE, "rvcsyn_load_weights has not succeeded on this handle");
  if (B < 1 || T < 1) return sfail(RVCDEC_ERR_ARG, "B and T must be >= 1");
  const rvcsyn_config& c = h->cfg;
  if (c.use_f0 && (!pitch || (h->dec && !nsff0))) return sfail(RVCDEC_ERR_ARG, "an f0 model needs pitch and nsff0");
  const size_t need = rvcsyn_workspace_bytes(h, B, T, mode);
  if (workspace_bytes < need) return sfail(RVCDEC_ERR_WORKSPACE, "workspace of %zu bytes is smaller than the %zu needed for B=%d, T=%d", workspace_bytes, need, B, T);
  DevGuard dg(h->device);
  SCK(dg.err);
  cudaStream_t s = (cudaStream_t)stream;
  SynWs w = carve_syn(h, workspace, B, T);
  void* dec_ws = (uint8_t*)workspace + align_up(w.bytes, 256);
  const int H = c.hidden_channels, I = c.inter_channels, F = c.filter_channels, Lp = padded_rows(T), half = I / 2;
  const int NL = c.flow_layers;
  h->launches = 0;
  int rc;
  const int stop0 = h->debug_stop;
  // Everything in front of the decoder as one launch sequence on stream `s`, reading the inputs through the pointers given
  // (the caller's, or the workspace copies the CUDA-graph path replays on).
  auto front = [&](const float* phone, const int32_t* lengths, const int64_t* pitch, const int64_t* sid, const float* noise_zp,
                   cudaStream_t s) -> int {
    // mode != 0: every convolution but the phone embedding runs on the tensor-core back-end from 16-bit operand copies
    // (fp16; bf16 in mode 3) with fp32 accumulation, bias, residual and LayerNorm; attention, gate and z_p stay fp32
    const bool fast = mode != 0 && !h->front_fp32;
    const int ff = (mode == 3) ? ACT_BF16 : ACT_F16;
    // y = conv(x): fp32 back-end from `in32`, or tensor cores from `in16`
    auto conv = [&](const FConv& L, const float* in32, const void* in16, int phys, const Epilogue& e) -> int {
      return fast ? mconv(h, L, in16, ff, phys, B, T, e, s) : fconv(h, L, in32, phys, B, T, e, s);
    };
    // x * x_mask behind a tensor-core convolution
    auto mask = [&](float* x32, int phys32, void* x16, int phys16, int C) -> int {
      if (!fast) return 0;
      SCK(launch_mask_rows(x32, phys32, x16, phys16, lengths, B, T, C, Lp, s)); ++h->launches;
      return 0;
    };
    // zero padding rows of every tensor a convolution with taps reads (FFN k = 3, WaveNet k = 5)
    SCK(launch_zero_pads(w.x, w.hh, nullptr, 4, B, H, T, Lp, s)); ++h->launches;
    SCK(launch_zero_pads(w.h1, nullptr, nullptr, 4, B, F, T, Lp, s)); ++h->launches;
    const bool att_tc = fast && h->att_mma;
    if (att_tc) { SCK(launch_zero_pads(w.qkv16, nullptr, nullptr, 2, B, 3 * H, T, Lp, s)); ++h->launches; }  // key / value tiles run past T
    if (fast) {
      SCK(launch_zero_pads(w.x16, w.hh16, nullptr, 2, B, H, T, Lp, s)); ++h->launches;
      SCK(launch_zero_pads(w.h116, nullptr, nullptr, 2, B, F, T, Lp, s)); ++h->launches;
    }
    // g = emb_g(sid)  (synthesizers.py:226)
    SCK(launch_gather_rows(h->d_emb_g, (const long long*)sid, w.g, B, c.gin_channels, c.spk_embed_dim, s)); ++h->launches;
    // ---- TextEncoder (encoders.py:127-144) ----
    SCK(launch_pack_btc(phone, w.phone, B, T, c.text_dim, Lp, s)); ++h->launches;
    {
      Epilogue e = epi_basic(h->emb_phone.d_bias, H, T, nullptr);
      e.out_f32 = w.x;
      if ((rc = fconv(h, h->emb_phone, w.phone, 0, B, T, e, s))) return rc;
    }
    SCK(launch_embed_post(w.x, c.use_f0 ? (const long long*)pitch : nullptr, h->d_emb_pitch, lengths, B, T, H, Lp, sqrtf((float)H),
                          fast ? w.x16 : nullptr, ff, s)); ++h->launches;
    const int stop = h->debug_stop;
    if (stop == 1) return 0;
    for (int l = 0; l < c.n_layers; ++l) {
      const AttnLayer& A = h->attn[l];
      {  // q | k | v
        Epilogue e = epi_basic(A.qkv.d_bias, 3 * H, T, nullptr);
        e.out_f32 = w.qkv;
        if (att_tc) {  // the tensor-core attention reads q | k | v as 16-bit operand tiles
          e.out_act = w.qkv16; e.act_fmt = ff; e.slope = 1.f;
          if (stop != 2) e.out_f32 = nullptr;
        }
        if ((rc = conv(A.qkv, w.x, w.x16, 0, e))) return rc;
      }
      if (stop == 2 && l == 0) return 0;
      if (att_tc) {
        const uint8_t* rel = A.d_rel16[ff == ACT_BF16 ? 1 : 0];
        SCK(launch_attention_mma(w.qkv16, w.att, w.att16, rel, rel + att_mma_rel_bytes(), lengths, B, T, c.n_heads, H / c.n_heads, c.window_size,
                                 Lp, ff, s)); ++h->launches;
      } else {
        SCK(launch_attention(w.qkv, w.att, A.d_ek, A.d_ev, lengths, B, T, c.n_heads, H / c.n_heads, c.window_size, Lp, s)); ++h->launches;
        if (fast) { SCK(launch_cast16(w.att, w.att16, ff, B, T, H, Lp, s)); ++h->launches; }
      }
      if (stop == 3 && l == 0) return 0;
      {  // x + conv_o(att)  -> LayerNorm  (encoders.py:77-79)
        Epilogue e = epi_basic(A.o.d_bias, H, T, nullptr);
        e.res = w.x; e.out_f32 = w.tmp;
        if ((rc = conv(A.o, w.att, w.att16, 0, e))) return rc;
      }
      SCK(launch_layernorm(w.tmp, w.x, A.d_g1, A.d_b1, lengths, B, T, H, Lp, 1e-5f, fast ? w.x16 : nullptr, ff, s)); ++h->launches;
      if (stop == 4 && l == 0) return 0;
      {  // FFN: conv_1 -> relu -> mask  (attentions.py:221-224); x is masked already
        Epilogue e = epi_basic(A.ffn1.d_bias, F, T, lengths);
        e.slope = 0.f;
        if (fast) { e.out_act = w.h116; e.act_fmt = ff; } else e.out_act = w.h1;
        if ((rc = conv(A.ffn1, w.x, w.x16, 0, e)) || (rc = mask(nullptr, 0, w.h116, 0, F))) return rc;
      }
      {  // x + conv_2(...) * mask -> LayerNorm  (:225-226, encoders.py:81-83)
        Epilogue e = epi_basic(A.ffn2.d_bias, H, T, nullptr);
        e.res = w.x; e.out_f32 = w.tmp;
        // (rows beyond the length: conv_2 * mask = 0 and x = 0 there, and the LayerNorm below zeroes them anyway)
        if ((rc = conv(A.ffn2, w.h1, w.h116, 0, e))) return rc;
      }
      SCK(launch_layernorm(w.tmp, w.x, A.d_g2, A.d_b2, lengths, B, T, H, Lp, 1e-5f, fast ? w.x16 : nullptr, ff, s)); ++h->launches;
      if (stop == 5 && l == 0) return 0;
    }
    {  // stats = proj(x) * mask  (encoders.py:141)
      Epilogue e = epi_basic(h->proj.d_bias, 2 * I, T, lengths);
      e.out_f32 = w.stats;
      if ((rc = conv(h->proj, w.x, w.x16, 0, e)) || (rc = mask(w.stats, 0, nullptr, 0, 2 * I))) return rc;
    }
    if (stop == 6) return 0;
    // z_p = (m_p + exp(logs_p) * randn * 0.66666) * mask  (synthesizers.py:228)
    SCK(launch_zp(w.stats, noise_zp, w.xf, lengths, seed, 0, B, T, I, Lp, fast ? w.xf16 : nullptr, ff, s)); ++h->launches;
    if (stop == 7) return 0;
    // ---- flow(reverse=True) (residuals.py:162-165) ----
    const size_t half_off = (size_t)(half / 8) * Lp * 8;  // element offset of the upper channel half inside an item
    for (int i = c.n_flows - 1; i >= 0; --i) {
      const FlowLayer& Fl = h->flow[i];
      const size_t o0 = Fl.flipped ? half_off : 0, o1 = Fl.flipped ? 0 : half_off;
      const float* x0 = w.xf + o0;   // the half the layer conditions on
      float* x1 = w.xf + o1;         // the half it updates
      {  // h = pre(x0) * mask
        Epilogue e = epi_basic(Fl.pre.d_bias, H, T, lengths);
        e.out_f32 = w.hh;
        if (fast) { e.out_act = w.hh16; e.act_fmt = ff; e.slope = 1.f; }
        if ((rc = conv(Fl.pre, x0, w.xf16 + o0, I / 8, e)) || (rc = mask(w.hh, 0, w.hh16, 0, H))) return rc;
      }
      // WaveNet (modules.py:84-117): per-item bias = in_layer bias + cond_layer(g) slice
      SCK(launch_cond(c.gin_channels ? w.g : nullptr, Fl.d_cond_w, Fl.d_cond_b, Fl.d_in_bias_all, w.condB, B, c.gin_channels, 2 * H * NL, s)); ++h->launches;
      for (int l = 0; l < NL; ++l) {
        {
          Epilogue e = epi_basic(w.condB + (size_t)l * 2 * H, 2 * H, T, nullptr);
          e.bias_bstride = 2 * H * NL;
          e.out_f32 = w.a;
          if ((rc = conv(Fl.in[l], w.hh, w.hh16, 0, e))) return rc;
        }
        SCK(launch_gate(w.a, w.acts, B, T, H, Lp, fast ? w.acts16 : nullptr, ff, s)); ++h->launches;
        {  // output += skip part
          Epilogue e = epi_basic(Fl.skip[l].d_bias, H, T, nullptr);
          e.res = (l == 0) ? nullptr : w.outacc; e.out_f32 = w.outacc;
          if (fast && l == NL - 1) { e.out_act = w.out16; e.act_fmt = ff; e.slope = 1.f; }
          if ((rc = conv(Fl.skip[l], w.acts, w.acts16, 0, e))) return rc;
        }
        if (l < NL - 1) {  // x = (x + res part) * mask
          Epilogue e = epi_basic(Fl.res[l].d_bias, H, T, lengths);
          e.res = w.hh; e.out_f32 = w.hh;
          if (fast) { e.out_act = w.hh16; e.act_fmt = ff; e.slope = 1.f; }
          if ((rc = conv(Fl.res[l], w.acts, w.acts16, 0, e)) || (rc = mask(w.hh, 0, w.hh16, 0, H))) return rc;
        }
      }
      {  // x1 = (x1 - post(output * mask)) * mask   (mean-only coupling, reverse; residuals.py:239-255)
        Epilogue e = epi_basic(Fl.post.d_bias, I, T, lengths);
        e.scale = -1.f; e.sum_in = x1; e.out_f32 = x1;
        if (fast) { e.out_act = w.xf16 + o1; e.act_fmt = ff; e.slope = 1.f; }
        if ((rc = conv(Fl.post, w.outacc, w.out16, 0, e)) || (rc = mask(x1, I / 8, w.xf16 + o1, I / 8, half))) return rc;
      }
      if (stop == 8 && i == c.n_flows - 1) return 0;
    }
    if (stop == 9) return 0;
    // z * mask -> (B, inter, T) fp32 for the decoder (and the caller, if asked for)
    SCK(launch_unpack(w.xf, ACT_F32, w.z_nct, B, I, T, Lp, s)); ++h->launches;
    return 0;
  };
  // Small batches are launch-bound here (~130 launches of a few microseconds of work each: 2.2 ms at B = 1 whatever T), so
  // their sequence is captured once per (workspace, B, T, mode) and replayed as a CUDA graph; the inputs are first copied
  // to fixed places in the workspace so that one graph serves every call.  Large batches, debug stops and in-kernel z_p
  // draws (the seed is a kernel argument) launch directly.
  const bool graphable = h->use_graphs && stop0 == 0 && noise_zp != nullptr && (int64_t)B * T <= h->graph_max_frames;
  if (!graphable) {
    if ((rc = front(phone, lengths, pitch, sid, noise_zp, s))) return rc;
    if (stop0) return 0;
  } else {
    SCK(cudaMemcpyAsync(w.in_phone, phone, (size_t)B * T * c.text_dim * 4, cudaMemcpyDeviceToDevice, s));
    SCK(cudaMemcpyAsync(w.in_len, lengths, (size_t)B * 4, cudaMemcpyDeviceToDevice, s));
    if (pitch) SCK(cudaMemcpyAsync(w.in_pitch, pitch, (size_t)B * T * 8, cudaMemcpyDeviceToDevice, s));
    SCK(cudaMemcpyAsync(w.in_sid, sid, (size_t)B * 8, cudaMemcpyDeviceToDevice, s));
    SCK(cudaMemcpyAsync(w.in_nzp, noise_zp, (size_t)B * I * T * 4, cudaMemcpyDeviceToDevice, s));
    rvcsyn_handle::GraphEntry* hit = nullptr;
    for (auto& e : h->graphs)
      if (e.ws == workspace && e.B == B && e.T == T && e.mode == mode && e.has_pitch == (pitch != nullptr) && e.epoch == h->epoch) { hit = &e; break; }
    if (!hit) {
      if (!h->cap_stream) SCK(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
      cudaGraph_t graph = nullptr;
      cudaError_t ce = cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeRelaxed);
      int frc = 0;
      if (ce == cudaSuccess) {
        frc = front(w.in_phone, w.in_len, pitch ? w.in_pitch : nullptr, w.in_sid, w.in_nzp, h->cap_stream);
        ce = cudaStreamEndCapture(h->cap_stream, &graph);
      }
      cudaGraphExec_t exec = nullptr;
      if (frc == 0 && ce == cudaSuccess && graph) ce = cudaGraphInstantiate(&exec, graph, 0);
      if (graph) cudaGraphDestroy(graph);
      if (frc != 0 || ce != cudaSuccess || !exec) {  // capture is an optimisation: direct launches from now on
        (void)cudaGetLastError();
        h->use_graphs = 0;
        h->launches = 0;
        if (frc) return frc;
        if ((rc = front(phone, lengths, pitch, sid, noise_zp, s))) return rc;
      } else {
        if (h->graphs.size() >= 8) {  // LRU
          size_t victim = 0;
          for (size_t i = 1; i < h->graphs.size(); ++i) if (h->graphs[i].last_use < h->graphs[victim].last_use) victim = i;
          cudaGraphExecDestroy(h->graphs[victim].exec);
          h->graphs.erase(h->graphs.begin() + victim);
        }
        h->graphs.push_back({workspace, B, T, mode, pitch != nullptr, h->epoch, exec, h->launches, 0});
        hit = &h->graphs.back();
        ++h->graph_captures;
      }
    } else {
      ++h->graph_hits;
    }
    if (hit) {
      hit->last_use = ++h->graph_clock;
      SCK(cudaGraphLaunch(hit->exec, s));
      h->launches = hit->launches + 1;
    }
  }
  if (z_out) SCK(cudaMemcpyAsync(z_out, w.z_nct, (size_t)B * I * T * 4, cudaMemcpyDeviceToDevice, s));
  if (!h->dec) return 0;
  rc = rvcdec_forward(h->dec, w.z_nct, nsff0, c.gin_channels ? w.g : nullptr, noise_src, seed, out, dec_ws,
                      workspace_bytes - align_up(w.bytes, 256), B, T, mode, stream);
  if (rc) return sfail(rc, "decoder: %s", rvcdec_last_error());
  return 0;
}

}  // extern "C"
