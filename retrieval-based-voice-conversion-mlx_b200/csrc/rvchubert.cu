// C ABI of the HuBERT / ContentVec content embedder (include/rvchubert.h): weight packing (conv taps into GEMM rows, the
// positional conv's weight norm, q | k | v fused) and the layer schedule of transformers' HubertModel for the base
// configuration (modeling_hubert.py:178-477, :889-966; the reference subclasses it in rvc/lib/utils.py:31-34) over the
// kernels of hubert_kernels.cu.
#include <cuda_fp16.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/rvchubert.h"
#include "common.cuh"

using namespace rvcdec;

namespace rvcdec {
cudaError_t launch_gemm128(const float* A, size_t lda, size_t strideA, const float* W, const float* bias, const float* R, float* C, size_t ldc,
                           size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s);
cudaError_t launch_hubert_ln(const float* a, const float* b, const float* gamma, const float* beta, float* out, __half* out16, size_t rows, int C,
                             float eps, cudaStream_t s);
cudaError_t launch_hubert_feat0(const float* audio, size_t n, const float* w, const float* gamma, const float* beta, double* stats, float* scale,
                                float* shift, float* out, __half* out16, int B, int L0, int Lstride, int Cout, float eps, cudaStream_t s);
cudaError_t launch_hubert_posconv(const float* x, const float* wp, const float* bias, float* out, int B, int T, int C, cudaStream_t s);
cudaError_t launch_hubert_attention(const float* qkv, float* ctx, __half* ctx16, int B, int T, int heads, cudaStream_t s);
cudaError_t launch_hubert_att_tc(const __half* qkv16, __half* ctx16, int B, int T, int heads, cudaStream_t s);
cudaError_t launch_hubert_att_tma(const __half* qkv16, __half* ctx16, int B, int T, int heads, cudaStream_t s);
cudaError_t launch_posconv_tma(const __half* x16, const __half* wimg, const float* bias, float* out, int B, int T, int C, cudaStream_t s);
cudaError_t launch_gemm_tma(const __half* A, size_t lda, size_t strideA, const __half* W, const float* bias, const float* R, float* C32, __half* C16,
                            size_t ldc, size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s);
cudaError_t launch_gemm_tc(const __half* A, size_t lda, size_t strideA, const __half* W, const float* bias, const float* R, float* C32, __half* C16,
                           size_t ldc, size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s);
}  // namespace rvcdec

namespace {

thread_local std::string g_herr;
int hfail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_herr = buf;
  return code;
}
#define HCK(expr)                                                                                                     \
  do {                                                                                                                \
    cudaError_t _e = (expr);                                                                                          \
    if (_e != cudaSuccess) return hfail(RVCDEC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

struct HGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit HGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) err = cudaSetDevice(dev); else prev = -1;
  }
  ~HGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

constexpr int NCONV = 7, CONV_DIM = 512, HID = 768, NLAYER = 12, HEADS = 12, FFN = 3072, PROJ = 256, POS_K = 128, POS_G = 16;
constexpr int CONV_K[NCONV] = {10, 3, 3, 3, 3, 2, 2}, CONV_S[NCONV] = {5, 2, 2, 2, 2, 2, 2};
constexpr float LN_EPS = 1e-5f;
constexpr int N_STATS = 65;  // hubert_kernels.cu: S[10] + the 55 entries of R

inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }

struct LayerW { size_t wqkv, bqkv, wo, bo, ln1g, ln1b, w1, b1, w2, b2, ln2g, ln2b; };

void conv_lengths(int n, int L[NCONV]) {
  long long l = n;
  for (int i = 0; i < NCONV; ++i) {
    l = l >= CONV_K[i] ? (l - CONV_K[i]) / CONV_S[i] + 1 : 0;
    L[i] = (int)l;
  }
}

}  // namespace

struct rvchubert_handle {
  int device = 0;
  bool loaded = false, has_proj = false, keep_layer0 = false;
  bool tma = true;           // with tc: operands through tensor maps (gemm_tma.cu) instead of cp.async gathers (gemm_tc.cu)
  bool att_tc = true;        // with tc: attention on the tensor cores too (hubert_att_tc.cu)
  bool tc = true;            // GEMMs on the tensor cores with fp16 operands (gemm_tc.cu); 0: fp32 CUDA-core GEMMs
  float* dev = nullptr;
  __half* dev16 = nullptr;   // the GEMM weights once more in fp16, same offsets as in `dev`
  __half* pc16 = nullptr;    // positional conv weights for the tensor-core kernel: [16 groups][128 taps][48 co][64 ci] (ci >= 48 zero)
  size_t conv_w[NCONV] = {}, gn_g = 0, gn_b = 0, fp_lng = 0, fp_lnb = 0, fp_w = 0, fp_b = 0, pos_w = 0, pos_b = 0, enc_lng = 0, enc_lnb = 0,
         proj_w = 0, proj_b = 0;
  LayerW layer[NLAYER] = {};
  int64_t launches = 0;
};

namespace {

struct TMh {
  std::map<std::string, const rvcdec_tensor*> m;
  const rvcdec_tensor* get(const std::string& n, int64_t numel, int* rc) const {
    auto it = m.find(n);
    if (it == m.end()) { *rc = hfail(RVCDEC_ERR_WEIGHTS, "missing tensor %s", n.c_str()); return nullptr; }
    if (it->second->numel != numel) {
      *rc = hfail(RVCDEC_ERR_WEIGHTS, "%s: expected %lld elements, got %lld", n.c_str(), (long long)numel, (long long)it->second->numel);
      return nullptr;
    }
    return it->second;
  }
  bool has(const std::string& n) const { return m.count(n) != 0; }
};

size_t hpush(std::vector<float>& v, const float* src, size_t n) {
  const size_t off = (v.size() + 63) / 64 * 64;  // 256-byte aligned
  v.resize(off + n);
  if (src) memcpy(v.data() + off, src, n * sizeof(float));
  return off;
}

int push_named(std::vector<float>& host, const TMh& tm, const std::string& name, int64_t numel, size_t& off) {
  int rc = 0;
  const rvcdec_tensor* t = tm.get(name, numel, &rc);
  if (!t) return rc;
  off = hpush(host, t->data, (size_t)numel);
  return 0;
}

struct HWs {
  double* stats;
  float *scale, *shift, *bufA, *bufB, *featln, *xp, *pos, *emb, *x, *x1, *qkv, *ffn, *keep0;
  __half *a16, *b16, *featln16, *in16, *ctx16, *x1_16, *ffn16, *qkv16, *xp16;  // fp16 operands of the tensor-core mode
  size_t bytes;
};

HWs carve_h(void* base, int B, int n) {
  HWs w{};
  int L[NCONV];
  conv_lengths(n, L);
  const size_t T = (size_t)L[NCONV - 1], rows = (size_t)B * T;
  size_t off = 0;
  auto take = [&](size_t bytes) { void* p = base ? (void*)((uint8_t*)base + off) : nullptr; off += up256(bytes); return p; };
  w.stats = (double*)take((size_t)B * N_STATS * sizeof(double));
  w.scale = (float*)take((size_t)B * CONV_DIM * 4);
  w.shift = (float*)take((size_t)B * CONV_DIM * 4);
  // items are L + (L & 1) frames apart: a stride-2 layer then reads the batch as one matrix of rows 2 * 512 apart (gemm_tma.cu)
  w.bufA = (float*)take((size_t)B * (L[0] + 1) * CONV_DIM * 4);
  w.bufB = (float*)take((size_t)B * (L[1] + 1) * CONV_DIM * 4);
  w.featln = (float*)take(rows * CONV_DIM * 4);
  w.xp = (float*)take(rows * HID * 4);
  w.pos = (float*)take(rows * HID * 4);   // positional conv output, then the attention context
  w.emb = (float*)take(rows * HID * 4);
  w.x = (float*)take(rows * HID * 4);
  w.x1 = (float*)take(rows * HID * 4);
  w.qkv = (float*)take(rows * 3 * HID * 4);
  w.ffn = (float*)take(rows * FFN * 4);
  w.keep0 = (float*)take(rows * HID * 4);
  w.a16 = (__half*)take((size_t)B * (L[0] + 1) * CONV_DIM * 2);
  w.b16 = (__half*)take((size_t)B * (L[1] + 1) * CONV_DIM * 2);
  w.featln16 = (__half*)take(rows * CONV_DIM * 2);
  w.in16 = (__half*)take(rows * HID * 2);
  w.ctx16 = (__half*)take(rows * HID * 2);
  w.x1_16 = (__half*)take(rows * HID * 2);
  w.ffn16 = (__half*)take(rows * FFN * 2);
  w.qkv16 = (__half*)take(rows * 3 * HID * 2);
  w.xp16 = (__half*)take(rows * HID * 2);
  w.bytes = off;
  return w;
}

}  // namespace

extern "C" {

const char* rvchubert_last_error(void) { return g_herr.c_str(); }

int rvchubert_create(int device, rvchubert_handle** out) {
  if (!out) return hfail(RVCDEC_ERR_ARG, "null argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return hfail(RVCDEC_ERR_CUDA, "no CUDA device %d (the HuBERT path has no CPU fallback)", device);
  auto h = new rvchubert_handle();
  h->device = device;
  *out = h;
  return 0;
}

void rvchubert_destroy(rvchubert_handle* h) {
  if (!h) return;
  if (h->dev || h->dev16 || h->pc16) {
    HGuard g(h->device);
    cudaFree(h->dev);
    cudaFree(h->dev16);
    cudaFree(h->pc16);
  }
  delete h;
}

int64_t rvchubert_last_launch_count(const rvchubert_handle* h) { return h ? h->launches : 0; }

int32_t rvchubert_frames(int32_t n_samples) {
  int L[NCONV];
  conv_lengths(n_samples, L);
  return L[NCONV - 1];
}

int rvchubert_set_option(rvchubert_handle* h, const char* key, int64_t value) {
  if (!h || !key) return hfail(RVCDEC_ERR_ARG, "null argument");
  if (!strcmp(key, "keep_layer0")) { h->keep_layer0 = value != 0; return 0; }
  if (!strcmp(key, "tc")) { h->tc = value != 0; return 0; }
  if (!strcmp(key, "att_tc")) { h->att_tc = value != 0; return 0; }
  if (!strcmp(key, "tma")) { h->tma = value != 0; return 0; }
  return hfail(RVCDEC_ERR_ARG, "unknown option %s", key);
}

int rvchubert_load_weights(rvchubert_handle* h, const rvcdec_tensor* tensors, int32_t n) {
  if (!h || !tensors) return hfail(RVCDEC_ERR_ARG, "null argument");
  TMh tm;
  for (int i = 0; i < n; ++i) tm.m[tensors[i].name ? tensors[i].name : ""] = &tensors[i];
  std::vector<float> host;
  std::vector<__half> pc_img;
  h->loaded = false;
  int rc = 0;
  // feature extractor (HubertFeatureEncoder, modeling_hubert.py:178-213; conv_bias = False)
  for (int i = 0; i < NCONV; ++i) {
    const std::string p = "feature_extractor.conv_layers." + std::to_string(i);
    if (tm.has(p + ".conv.bias")) return hfail(RVCDEC_ERR_WEIGHTS, "%s.conv.bias: the base configuration has no conv bias", p.c_str());
    const int cin = i == 0 ? 1 : CONV_DIM, k = CONV_K[i];
    const rvcdec_tensor* w = tm.get(p + ".conv.weight", (int64_t)CONV_DIM * cin * k, &rc);
    if (!w) return rc;
    if (i == 0) {
      h->conv_w[0] = hpush(host, w->data, (size_t)w->numel);
    } else {  // (co, ci, j) -> rows of k * 512: [co][j * 512 + ci], the order of k consecutive time-major frames
      h->conv_w[i] = hpush(host, nullptr, (size_t)w->numel);
      float* dst = host.data() + h->conv_w[i];
      for (int co = 0; co < CONV_DIM; ++co)
        for (int ci = 0; ci < CONV_DIM; ++ci)
          for (int j = 0; j < k; ++j) dst[((size_t)co * k + j) * CONV_DIM + ci] = w->data[((size_t)co * CONV_DIM + ci) * k + j];
    }
  }
  if ((rc = push_named(host, tm, "feature_extractor.conv_layers.0.layer_norm.weight", CONV_DIM, h->gn_g))) return rc;
  if ((rc = push_named(host, tm, "feature_extractor.conv_layers.0.layer_norm.bias", CONV_DIM, h->gn_b))) return rc;
  if ((rc = push_named(host, tm, "feature_projection.layer_norm.weight", CONV_DIM, h->fp_lng))) return rc;
  if ((rc = push_named(host, tm, "feature_projection.layer_norm.bias", CONV_DIM, h->fp_lnb))) return rc;
  if ((rc = push_named(host, tm, "feature_projection.projection.weight", (int64_t)HID * CONV_DIM, h->fp_w))) return rc;
  if ((rc = push_named(host, tm, "feature_projection.projection.bias", HID, h->fp_b))) return rc;
  {  // positional conv: weight_norm(dim = 2): w = g[j] * v / ||v[:, :, j]|| (modeling_hubert.py:45-80), packed (g, ci, j, co)
    const std::string p = "encoder.pos_conv_embed.conv.";
    const int cg = HID / POS_G;
    const rvcdec_tensor *g = nullptr, *v = nullptr;
    if (tm.has(p + "parametrizations.weight.original0")) {
      g = tm.get(p + "parametrizations.weight.original0", POS_K, &rc);
      v = tm.get(p + "parametrizations.weight.original1", (int64_t)HID * cg * POS_K, &rc);
    } else if (tm.has(p + "weight_g")) {
      g = tm.get(p + "weight_g", POS_K, &rc);
      v = tm.get(p + "weight_v", (int64_t)HID * cg * POS_K, &rc);
    } else {
      v = tm.get(p + "weight", (int64_t)HID * cg * POS_K, &rc);
      if (!v) return rc;
    }
    if (rc || !v) return rc ? rc : hfail(RVCDEC_ERR_WEIGHTS, "positional conv weight missing");
    std::vector<double> norm(POS_K, 1.0);
    if (g) {
      for (int j = 0; j < POS_K; ++j) {
        double s = 0.0;
        for (size_t r = 0; r < (size_t)HID * cg; ++r) { const double x = v->data[r * POS_K + j]; s += x * x; }
        norm[j] = (double)g->data[j] / std::sqrt(s);
      }
    }
    h->pos_w = hpush(host, nullptr, (size_t)v->numel);
    float* dst = host.data() + h->pos_w;
    for (int grp = 0; grp < POS_G; ++grp)
      for (int ci = 0; ci < cg; ++ci)
        for (int j = 0; j < POS_K; ++j)
          for (int co = 0; co < cg; ++co)
            dst[(((size_t)grp * cg + ci) * POS_K + j) * cg + co] = (float)(v->data[((size_t)(grp * cg + co) * cg + ci) * POS_K + j] * norm[j]);
    if ((rc = push_named(host, tm, p + "bias", HID, h->pos_b))) return rc;
    pc_img.assign((size_t)POS_G * POS_K * cg * 64, __float2half_rn(0.f));
    for (int grp = 0; grp < POS_G; ++grp)
      for (int j = 0; j < POS_K; ++j)
        for (int co = 0; co < cg; ++co)
          for (int ci = 0; ci < cg; ++ci)
            pc_img[(((size_t)grp * POS_K + j) * cg + co) * 64 + ci] =
                __float2half_rn((float)(v->data[((size_t)(grp * cg + co) * cg + ci) * POS_K + j] * norm[j]));
  }
  if ((rc = push_named(host, tm, "encoder.layer_norm.weight", HID, h->enc_lng))) return rc;
  if ((rc = push_named(host, tm, "encoder.layer_norm.bias", HID, h->enc_lnb))) return rc;
  for (int i = 0; i < NLAYER; ++i) {
    const std::string p = "encoder.layers." + std::to_string(i);
    LayerW& L = h->layer[i];
    const rvcdec_tensor* t[6];
    const char* names[6] = {".attention.q_proj.weight", ".attention.k_proj.weight", ".attention.v_proj.weight",
                            ".attention.q_proj.bias", ".attention.k_proj.bias", ".attention.v_proj.bias"};
    for (int k = 0; k < 6; ++k) {
      t[k] = tm.get(p + names[k], k < 3 ? (int64_t)HID * HID : HID, &rc);
      if (!t[k]) return rc;
    }
    L.wqkv = hpush(host, nullptr, (size_t)3 * HID * HID);
    L.bqkv = hpush(host, nullptr, (size_t)3 * HID);
    for (int k = 0; k < 3; ++k) {
      memcpy(host.data() + L.wqkv + (size_t)k * HID * HID, t[k]->data, sizeof(float) * HID * HID);
      memcpy(host.data() + L.bqkv + (size_t)k * HID, t[3 + k]->data, sizeof(float) * HID);
    }
    if ((rc = push_named(host, tm, p + ".attention.out_proj.weight", (int64_t)HID * HID, L.wo))) return rc;
    if ((rc = push_named(host, tm, p + ".attention.out_proj.bias", HID, L.bo))) return rc;
    if ((rc = push_named(host, tm, p + ".layer_norm.weight", HID, L.ln1g))) return rc;
    if ((rc = push_named(host, tm, p + ".layer_norm.bias", HID, L.ln1b))) return rc;
    if ((rc = push_named(host, tm, p + ".feed_forward.intermediate_dense.weight", (int64_t)FFN * HID, L.w1))) return rc;
    if ((rc = push_named(host, tm, p + ".feed_forward.intermediate_dense.bias", FFN, L.b1))) return rc;
    if ((rc = push_named(host, tm, p + ".feed_forward.output_dense.weight", (int64_t)HID * FFN, L.w2))) return rc;
    if ((rc = push_named(host, tm, p + ".feed_forward.output_dense.bias", HID, L.b2))) return rc;
    if ((rc = push_named(host, tm, p + ".final_layer_norm.weight", HID, L.ln2g))) return rc;
    if ((rc = push_named(host, tm, p + ".final_layer_norm.bias", HID, L.ln2b))) return rc;
  }
  h->has_proj = tm.has("final_proj.weight");
  if (h->has_proj) {
    if ((rc = push_named(host, tm, "final_proj.weight", (int64_t)PROJ * HID, h->proj_w))) return rc;
    if ((rc = push_named(host, tm, "final_proj.bias", PROJ, h->proj_b))) return rc;
  }
  HGuard g(h->device);
  HCK(g.err);
  if (h->dev) { HCK(cudaFree(h->dev)); h->dev = nullptr; }
  HCK(cudaMalloc((void**)&h->dev, host.size() * sizeof(float)));
  HCK(cudaMemcpy(h->dev, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice));
  {  // fp16 image of the same array (only the GEMM weights are read from it)
    std::vector<__half> h16(host.size());
    for (size_t i = 0; i < host.size(); ++i) h16[i] = __float2half_rn(host[i]);
    if (h->dev16) { HCK(cudaFree(h->dev16)); h->dev16 = nullptr; }
    HCK(cudaMalloc((void**)&h->dev16, h16.size() * sizeof(__half)));
    HCK(cudaMemcpy(h->dev16, h16.data(), h16.size() * sizeof(__half), cudaMemcpyHostToDevice));
    if (h->pc16) { HCK(cudaFree(h->pc16)); h->pc16 = nullptr; }
    HCK(cudaMalloc((void**)&h->pc16, pc_img.size() * sizeof(__half)));
    HCK(cudaMemcpy(h->pc16, pc_img.data(), pc_img.size() * sizeof(__half), cudaMemcpyHostToDevice));
  }
  h->loaded = true;
  return 0;
}

size_t rvchubert_workspace_bytes(const rvchubert_handle* h, int32_t B, int32_t n_samples) {
  (void)h;
  if (B <= 0 || n_samples <= 0) return 0;
  return carve_h(nullptr, B, n_samples).bytes;
}

int rvchubert_forward(rvchubert_handle* h, const float* audio, float* hidden, void* workspace, size_t workspace_bytes, int32_t B,
                      int32_t n_samples, void* stream) {
  if (!h || !audio || !hidden || !workspace) return hfail(RVCDEC_ERR_ARG, "null argument");
  if (!h->loaded) return hfail(RVCDEC_ERR_WEIGHTS, "rvchubert_load_weights has not succeeded on this handle");
  if (B <= 0) return hfail(RVCDEC_ERR_ARG, "B = %d", B);
  int L[NCONV];
  conv_lengths(n_samples, L);
  const int T = L[NCONV - 1];
  if (T <= 0) return hfail(RVCDEC_ERR_ARG, "%d samples are shorter than one frame (400 samples)", n_samples);
  const HWs w = carve_h(workspace, B, n_samples);
  if (workspace_bytes < w.bytes) return hfail(RVCDEC_ERR_WORKSPACE, "workspace of %zu bytes, %zu needed", workspace_bytes, w.bytes);
  if ((uintptr_t)workspace % 256) return hfail(RVCDEC_ERR_ARG, "workspace must be 256-byte aligned");
  HGuard g(h->device);
  HCK(g.err);
  cudaStream_t s = (cudaStream_t)stream;
  const float* P = h->dev;
  h->launches = 0;
  const size_t rows = (size_t)B * T;
  if (rows > 0x7fffffff) return hfail(RVCDEC_ERR_ARG, "B * T too large");
  const int M = (int)rows;
  const bool tc = h->tc;
  const __half* P16 = h->dev16;
  // one linear layer over rows: fp32 CUDA-core GEMM, or fp16 operands on the tensor cores (fp32 accumulation either way)
  auto linear = [&](const float* a32, const __half* a16, size_t lda, size_t strideA, size_t w, const float* bias, const float* res, float* c32,
                    __half* c16, size_t ldc, size_t strideC, int batch, int m, int n, int k, int act) -> cudaError_t {
    ++h->launches;
    if (tc && h->tma) return launch_gemm_tma(a16, lda, strideA, P16 + w, bias, res, c32, c16, ldc, strideC, batch, m, n, k, act, s);
    if (tc) return launch_gemm_tc(a16, lda, strideA, P16 + w, bias, res, c32, c16, ldc, strideC, batch, m, n, k, act, s);
    return launch_gemm128(a32, lda, strideA, P + w, bias, res, c32, ldc, strideC, batch, m, n, k, act, s);
  };
  // conv feature extractor: layer 0 with its GroupNorm, then six strided convs as GEMMs over k * 512-element rows
  HCK(launch_hubert_feat0(audio, (size_t)n_samples, P + h->conv_w[0], P + h->gn_g, P + h->gn_b, w.stats, w.scale, w.shift, tc ? nullptr : w.bufA,
                          tc ? w.a16 : nullptr, B, L[0], L[0] + (L[0] & 1), CONV_DIM, LN_EPS, s));
  h->launches += 3;
  float *src = w.bufA, *dst = w.bufB;
  __half *src16 = w.a16, *dst16 = w.b16;
  for (int i = 1; i < NCONV; ++i) {
    const bool last = i == NCONV - 1;  // the last layer's output is kept in fp32 (layer norm input, "features" hook)
    const size_t in_stride = (size_t)(L[i - 1] + (L[i - 1] & 1)) * CONV_DIM;
    const size_t out_stride = last ? (size_t)L[i] * CONV_DIM : (size_t)(L[i] + (L[i] & 1)) * CONV_DIM;  // the features stay (B, T, 512)
    HCK(linear(src, src16, (size_t)CONV_S[i] * CONV_DIM, in_stride, h->conv_w[i], nullptr, nullptr, (tc && !last) ? nullptr : dst,
               (tc && !last) ? dst16 : nullptr, CONV_DIM, out_stride, B, L[i], CONV_DIM, CONV_K[i] * CONV_DIM, 1));
    float* t = src; src = dst; dst = t;
    __half* t16 = src16; src16 = dst16; dst16 = t16;
  }
  const float* feat = src;  // (B, T, 512): six swaps leave the result in bufA
  // feature projection (modeling_hubert.py:216-232)
  HCK(launch_hubert_ln(feat, nullptr, P + h->fp_lng, P + h->fp_lnb, tc ? nullptr : w.featln, tc ? w.featln16 : nullptr, rows, CONV_DIM, LN_EPS, s));
  const bool pc_tc = tc && h->tma;
  HCK(linear(w.featln, w.featln16, CONV_DIM, 0, h->fp_w, P + h->fp_b, nullptr, w.xp, pc_tc ? w.xp16 : nullptr, HID, 0, 1, M, HID, CONV_DIM, 0));
  // encoder front (modeling_hubert.py:437-440): x + gelu(pos_conv(x)), LayerNorm
  cudaError_t pe = pc_tc ? launch_posconv_tma(w.xp16, h->pc16, P + h->pos_b, w.pos, B, T, HID, s) : cudaErrorNotSupported;
  if (pe == cudaErrorNotSupported) pe = launch_hubert_posconv(w.xp, P + h->pos_w, P + h->pos_b, w.pos, B, T, HID, s);
  HCK(pe);
  HCK(launch_hubert_ln(w.xp, w.pos, P + h->enc_lng, P + h->enc_lnb, w.emb, tc ? w.in16 : nullptr, rows, HID, LN_EPS, s));
  h->launches += 3;
  const float* in = w.emb;
  for (int i = 0; i < NLAYER; ++i) {  // HubertEncoderLayer.forward (modeling_hubert.py:388-405)
    const LayerW& Lw = h->layer[i];
    const bool atc = tc && h->att_tc;
    HCK(linear(in, w.in16, HID, 0, Lw.wqkv, P + Lw.bqkv, nullptr, atc ? nullptr : w.qkv, atc ? w.qkv16 : nullptr, 3 * HID, 0, 1, M, 3 * HID, HID, 0));
    if (atc) {
      cudaError_t ae = h->tma ? launch_hubert_att_tma(w.qkv16, w.ctx16, B, T, HEADS, s) : cudaErrorNotSupported;
      if (ae == cudaErrorNotSupported) ae = launch_hubert_att_tc(w.qkv16, w.ctx16, B, T, HEADS, s);
      HCK(ae);
    }
    else HCK(launch_hubert_attention(w.qkv, tc ? nullptr : w.pos, tc ? w.ctx16 : nullptr, B, T, HEADS, s));
    HCK(linear(w.pos, w.ctx16, HID, 0, Lw.wo, P + Lw.bo, in, w.x1, nullptr, HID, 0, 1, M, HID, HID, 0));
    HCK(launch_hubert_ln(w.x1, nullptr, P + Lw.ln1g, P + Lw.ln1b, w.x1, tc ? w.x1_16 : nullptr, rows, HID, LN_EPS, s));
    HCK(linear(w.x1, w.x1_16, HID, 0, Lw.w1, P + Lw.b1, nullptr, tc ? nullptr : w.ffn, tc ? w.ffn16 : nullptr, FFN, 0, 1, M, FFN, HID, 1));
    HCK(linear(w.ffn, w.ffn16, FFN, 0, Lw.w2, P + Lw.b2, w.x1, w.x, nullptr, HID, 0, 1, M, HID, FFN, 0));
    float* out = i == NLAYER - 1 ? hidden : w.x;
    HCK(launch_hubert_ln(w.x, nullptr, P + Lw.ln2g, P + Lw.ln2b, out, (tc && i != NLAYER - 1) ? w.in16 : nullptr, rows, HID, LN_EPS, s));
    h->launches += 3;
    if (i == 0 && h->keep_layer0) HCK(cudaMemcpyAsync(w.keep0, out, rows * HID * sizeof(float), cudaMemcpyDeviceToDevice, s));
    in = w.x;
  }
  return 0;
}

int rvchubert_final_proj(rvchubert_handle* h, const float* x, float* out, int64_t rows, void* stream) {
  if (!h || !x || !out) return hfail(RVCDEC_ERR_ARG, "null argument");
  if (!h->loaded || !h->has_proj) return hfail(RVCDEC_ERR_WEIGHTS, "final_proj.weight / final_proj.bias were not loaded");
  if (rows <= 0 || rows > 0x7fffffff) return hfail(RVCDEC_ERR_ARG, "rows = %lld", (long long)rows);
  HGuard g(h->device);
  HCK(g.err);
  HCK(launch_gemm128(x, HID, 0, h->dev + h->proj_w, h->dev + h->proj_b, nullptr, out, PROJ, 0, 1, (int)rows, PROJ, HID, 0, (cudaStream_t)stream));
  h->launches = 1;
  return 0;
}

int64_t rvchubert_debug_read(rvchubert_handle* h, const char* name, float* dst, int64_t dst_numel, void* workspace, int32_t B,
                             int32_t n_samples, void* stream) {
  if (!h || !name || !dst || !workspace) return hfail(RVCDEC_ERR_ARG, "null argument");
  const HWs w = carve_h(workspace, B, n_samples);
  const int64_t T = rvchubert_frames(n_samples);
  const float* src = nullptr;
  int64_t numel = 0;
  if (!strcmp(name, "features")) { src = w.bufA; numel = (int64_t)B * T * CONV_DIM; }
  else if (!strcmp(name, "embedded")) { src = w.emb; numel = (int64_t)B * T * HID; }
  else if (!strcmp(name, "layer0")) {
    if (!h->keep_layer0) return hfail(RVCDEC_ERR_ARG, "layer0 is kept only with the keep_layer0 option");
    src = w.keep0; numel = (int64_t)B * T * HID;
  } else return hfail(RVCDEC_ERR_ARG, "unknown intermediate %s", name);
  if (dst_numel < numel) return hfail(RVCDEC_ERR_ARG, "%s has %lld floats, the buffer %lld", name, (long long)numel, (long long)dst_numel);
  HGuard g(h->device);
  HCK(g.err);
  HCK(cudaMemcpyAsync(dst, src, (size_t)numel * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return numel;
}

}  // extern "C"
