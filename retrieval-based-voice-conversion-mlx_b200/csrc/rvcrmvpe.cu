// C ABI of the RMVPE f0 predictor (include/rvcrmvpe.h): weight folding (BatchNorm -> per-channel scale / shift), the layer
// schedule of E2E(4, 1, (2, 2)) (rvc/lib/predictors/RMVPE.py:246-339) over the kernels of rmvpe_kernels.cu, mel front end and decode.
#include <cuda_fp16.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/rvcrmvpe.h"
#include "common.cuh"

using namespace rvcdec;

namespace rvcdec {
struct Conv2dParams {
  const float* in0; const float* in1; int C0, Cin;
  const float* w; const float* scale; const float* shift; const float* res;
  float* out;
  int B, Cout, H, W, relu, out_btcw, tw;
};
cudaError_t launch_conv2d(const Conv2dParams& p, int ks, cudaStream_t s);
struct ConvTcParams {
  const float* in0; const float* in1; int C0, Cin;
  const __half* wimg;
  const float* scale; const float* shift; const float* res;
  float* out;
  int B, Cout, H, W, relu;
  float* part;
  int ksplit;
};
int conv3x3_tc_splits(int Cin, int Cout, int HW, int B);
bool conv3x3_tc_supported(int Cin, int C0, int Cout);
size_t conv3x3_tc_image_elems(int Cin, int Cout);
void conv3x3_tc_pack_host(const float* w, int Cin, int Cout, __half* img);
cudaError_t launch_conv3x3_tc(const ConvTcParams& p, cudaStream_t s);
bool conv3x3_tma_supported(int Cin, int Cout);
void conv3x3_tma_pack_host(const float* w, int Cin, int Cout, int taps, __half* img);
void convt2x_tma_pack_host(const float* w, int Cin, int Cout, __half* img);
cudaError_t launch_rmvpe_to_nhwc16(const float* in0, const float* in1, int C0, int C, int B, int H, int W, __half* out, cudaStream_t s);
cudaError_t launch_conv3x3_tma(const __half* in16, const __half* wimg, const float* scale, const float* shift, const float* res, float* out32,
                               __half* out16, int B, int Cin, int Cout, int H, int W, int relu, int taps, int up_cout, float* part,
                               size_t part_floats, int* launches, cudaStream_t s);
cudaError_t launch_convt2x(const float* in, const float* w, const float* scale, const float* shift, float* out, int B, int Cin, int Cout,
                           int H, int W, cudaStream_t s);
cudaError_t launch_avgpool2(const float* in, float* out, size_t planes, int H, int W, cudaStream_t s);
cudaError_t launch_mel_in(const float* mel, float* x, int B, int T, float a, float c, cudaStream_t s);
cudaError_t launch_sgemm_nt(const float* A, size_t lda, size_t strideA, const float* Bm, const float* bias, float* C, size_t ldc,
                            size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s);
cudaError_t launch_gru(const float* gi, const float* whhT, const float* bhh, float* out, int B, int T, cudaStream_t s);
cudaError_t launch_reflect_pad(const float* audio, float* padded, int B, int n, int pad, cudaStream_t s);
cudaError_t launch_mel_log(const float* spec, const float* basis, float* mel, int B, int frames, int NB, float clampv, cudaStream_t s);
cudaError_t launch_rmvpe_decode(const float* sal, float thred, float* f0, int T, cudaStream_t s);
}  // namespace rvcdec

namespace {

thread_local std::string g_merr;
int mfail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_merr = buf;
  return code;
}
#define MCK(expr)                                                                                                     \
  do {                                                                                                                \
    cudaError_t _e = (expr);                                                                                          \
    if (_e != cudaSuccess) return mfail(RVCDEC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

struct MGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit MGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) err = cudaSetDevice(dev); else prev = -1;
  }
  ~MGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

constexpr int N_MELS = 128, N_CLASS = 360, N_FFT = 1024, HOP = 160, NBINS = N_FFT / 2 + 1;
constexpr int LEVELS = 5, INTERS = 4, BLOCKS = 4, C_BASE = 16, GRU_HID = 256, GRU_IN = 3 * N_MELS;
constexpr float BN_EPS = 1e-5f;

// one convolution with its folded BatchNorm (or bias): offsets into the flat parameter array
struct ConvW { size_t w = 0, scale = 0, shift = 0, img = 0, img2 = 0; int cin = 0, cout = 0; bool has_scale = false, has_shift = false, has_img = false, has_img2 = false; };
// ConvBlockRes (RMVPE.py:13-57)
struct Block { ConvW c1, c2, sc; bool shortcut = false; };

inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }

}  // namespace

struct rvcrmvpe_handle {
  int device = 0;
  bool loaded = false;
  std::vector<float> host;   // every parameter, folded, in one array
  std::vector<__half> host16;  // fp16 tile images of the 3 x 3 convolutions the tensor-core kernel takes (rmvpe_conv_tc.cu)
  float* dev = nullptr;
  __half* dev16 = nullptr;
  bool keep_inter = false;
  bool tma = true;           // with tc: 3 x 3 convolutions with Cin % 64 == 0 as TMA-fed implicit GEMMs (rmvpe_conv_tma.cu)
  __half *t16 = nullptr, *y16[2] = {nullptr, nullptr};  // channels-last fp16 copies of the current forward (workspace)
  int y16_cur = 0;
  int T_ws = 0;              // frames of the current forward (sizes the fp16 buffers)
  const float* y16_of = nullptr;  // the fp32 tensor whose copy is in y16[y16_cur] (valid between consecutive blocks of a level)
  float* part = nullptr;     // split-K scratch of the current forward (in the caller's workspace)
  size_t part_floats = 0;
  bool tc = true;            // those convolutions on the tensor cores (fp16 operands, fp32 accumulation); 0: all fp32
  float in_a = 1.f, in_c = 0.f;                 // encoder input BatchNorm (one channel)
  std::vector<Block> enc[LEVELS], inter[INTERS], dec[LEVELS];
  ConvW up[LEVELS];                              // decoder ConvTranspose2d + BN
  ConvW cnn;
  size_t wih = 0, bih = 0, whhT = 0, bhh = 0, fcw = 0, fcb = 0, basis = 0, dft = 0;
  int64_t launches = 0;
};

namespace {

struct TM {
  std::map<std::string, const rvcdec_tensor*> m;
  const rvcdec_tensor* get(const std::string& n, int64_t numel, int* rc) const {
    auto it = m.find(n);
    if (it == m.end()) { *rc = mfail(RVCDEC_ERR_WEIGHTS, "missing tensor %s", n.c_str()); return nullptr; }
    if (it->second->numel != numel) {
      *rc = mfail(RVCDEC_ERR_WEIGHTS, "%s: expected %lld elements, got %lld", n.c_str(), (long long)numel, (long long)it->second->numel);
      return nullptr;
    }
    return it->second;
  }
};

size_t push(std::vector<float>& v, const float* src, size_t n) {
  const size_t off = (v.size() + 63) / 64 * 64;  // 256-byte aligned
  v.resize(off + n);
  if (src) memcpy(v.data() + off, src, n * sizeof(float));
  return off;
}

// conv weight + BatchNorm `bn` folded into scale / shift (eval mode: y = (x - mean) / sqrt(var + eps) * w + b)
int load_conv_bn(rvcrmvpe_handle* h, const TM& tm, const std::string& conv, const std::string& bn, int cin, int cout, int ks, ConvW& c,
                 bool pack_tc = false) {
  int rc = 0;
  const rvcdec_tensor* w = tm.get(conv + ".weight", (int64_t)cin * cout * ks * ks, &rc);
  if (!w) return rc;
  c.cin = cin; c.cout = cout;
  c.w = push(h->host, w->data, (size_t)w->numel);
  if (pack_tc && ks == 3 && conv3x3_tc_supported(cin, cin, cout)) {  // (a concatenated input splits at cin / 2, a multiple of 64 too)
    c.img = (h->host16.size() + 127) / 128 * 128;
    h->host16.resize(c.img + conv3x3_tc_image_elems(cin, cout));
    conv3x3_tc_pack_host(w->data, cin, cout, h->host16.data() + c.img);
    c.has_img = true;
  }
  if (pack_tc && ks == 3 && conv3x3_tma_supported(cin, cout)) {
    c.img2 = (h->host16.size() + 127) / 128 * 128;
    h->host16.resize(c.img2 + (size_t)9 * cout * cin);
    conv3x3_tma_pack_host(w->data, cin, cout, 9, h->host16.data() + c.img2);
    c.has_img2 = true;
  }
  const rvcdec_tensor *g = tm.get(bn + ".weight", cout, &rc), *b = tm.get(bn + ".bias", cout, &rc), *mu = tm.get(bn + ".running_mean", cout, &rc),
                      *var = tm.get(bn + ".running_var", cout, &rc);
  if (!g || !b || !mu || !var) return rc;
  c.scale = push(h->host, nullptr, cout);
  c.shift = push(h->host, nullptr, cout);
  for (int i = 0; i < cout; ++i) {
    const float sc = g->data[i] / std::sqrt(var->data[i] + BN_EPS);
    h->host[c.scale + i] = sc;
    h->host[c.shift + i] = b->data[i] - mu->data[i] * sc;
  }
  c.has_scale = c.has_shift = true;
  return 0;
}

int load_block(rvcrmvpe_handle* h, const TM& tm, const std::string& p, int cin, int cout, Block& blk) {
  int rc;
  if ((rc = load_conv_bn(h, tm, p + ".conv.0", p + ".conv.1", cin, cout, 3, blk.c1, true))) return rc;
  if ((rc = load_conv_bn(h, tm, p + ".conv.3", p + ".conv.4", cout, cout, 3, blk.c2, true))) return rc;
  blk.shortcut = cin != cout;  // RMVPE.py:47-51
  if (blk.shortcut) {
    const rvcdec_tensor *w = tm.get(p + ".shortcut.weight", (int64_t)cin * cout, &rc), *b = tm.get(p + ".shortcut.bias", cout, &rc);
    if (!w || !b) return rc;
    blk.sc.cin = cin; blk.sc.cout = cout;
    blk.sc.w = push(h->host, w->data, (size_t)w->numel);
    blk.sc.shift = push(h->host, b->data, cout);
    blk.sc.has_shift = true;
    if (conv3x3_tma_supported(cin, cout)) {  // the 1 x 1 projection through the same TMA kernel (one tap)
      blk.sc.img2 = (h->host16.size() + 127) / 128 * 128;
      h->host16.resize(blk.sc.img2 + (size_t)cout * cin);
      conv3x3_tma_pack_host(w->data, cin, cout, 1, h->host16.data() + blk.sc.img2);
      blk.sc.has_img2 = true;
    }
  }
  return 0;
}

int load_blocks(rvcrmvpe_handle* h, const TM& tm, const std::string& p, int cin, int cout, std::vector<Block>& out) {
  out.assign(BLOCKS, Block());
  for (int j = 0; j < BLOCKS; ++j) {
    int rc = load_block(h, tm, p + "." + std::to_string(j), j == 0 ? cin : cout, cout, out[j]);
    if (rc) return rc;
  }
  return 0;
}

}  // namespace

extern "C" {

const char* rvcrmvpe_last_error(void) { return g_merr.c_str(); }

int rvcrmvpe_create(int device, rvcrmvpe_handle** out) {
  if (!out) return mfail(RVCDEC_ERR_ARG, "null argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return mfail(RVCDEC_ERR_CUDA, "no CUDA device %d (the RMVPE path has no CPU fallback)", device);
  auto h = new rvcrmvpe_handle();
  h->device = device;
  *out = h;
  return 0;
}

void rvcrmvpe_destroy(rvcrmvpe_handle* h) {
  if (!h) return;
  if (h->dev || h->dev16) {
    MGuard g(h->device);
    cudaFree(h->dev);
    cudaFree(h->dev16);
  }
  delete h;
}

int64_t rvcrmvpe_last_launch_count(const rvcrmvpe_handle* h) { return h ? h->launches : 0; }

int rvcrmvpe_set_option(rvcrmvpe_handle* h, const char* key, int64_t value) {
  if (!h || !key) return mfail(RVCDEC_ERR_ARG, "null argument");
  if (!strcmp(key, "tc")) { h->tc = value != 0; return 0; }
  if (!strcmp(key, "keep_inter")) { h->keep_inter = value != 0; return 0; }
  if (!strcmp(key, "tma")) { h->tma = value != 0; return 0; }
  return mfail(RVCDEC_ERR_ARG, "unknown option %s", key);
}

int rvcrmvpe_load_weights(rvcrmvpe_handle* h, const rvcdec_tensor* tensors, int32_t n) {
  if (!h || !tensors) return mfail(RVCDEC_ERR_ARG, "null argument");
  TM tm;
  for (int i = 0; i < n; ++i) tm.m[tensors[i].name ? tensors[i].name : ""] = &tensors[i];
  h->host.clear();
  h->host16.clear();
  h->loaded = false;
  int rc = 0;
  {  // encoder input BatchNorm, one channel (RMVPE.py:119, :138)
    const rvcdec_tensor *g = tm.get("unet.encoder.bn.weight", 1, &rc), *b = tm.get("unet.encoder.bn.bias", 1, &rc),
                        *mu = tm.get("unet.encoder.bn.running_mean", 1, &rc), *var = tm.get("unet.encoder.bn.running_var", 1, &rc);
    if (!g || !b || !mu || !var) return rc;
    h->in_a = g->data[0] / std::sqrt(var->data[0] + BN_EPS);
    h->in_c = b->data[0] - mu->data[0] * h->in_a;
  }
  int cin = 1, cout = C_BASE;
  for (int i = 0; i < LEVELS; ++i) {
    if ((rc = load_blocks(h, tm, "unet.encoder.layers." + std::to_string(i) + ".conv", cin, cout, h->enc[i]))) return rc;
    cin = cout; cout *= 2;
  }
  // cin = 256 (encoder output channels), cout = 512 (Encoder.out_channel)
  for (int i = 0; i < INTERS; ++i)
    if ((rc = load_blocks(h, tm, "unet.intermediate.layers." + std::to_string(i) + ".conv", i == 0 ? cin : cout, cout, h->inter[i]))) return rc;
  int c = cout;
  for (int i = 0; i < LEVELS; ++i) {
    const std::string p = "unet.decoder.layers." + std::to_string(i);
    const int co = c / 2;
    if ((rc = load_conv_bn(h, tm, p + ".conv1.0", p + ".conv1.1", c, co, 3, h->up[i]))) return rc;  // ConvTranspose2d weight: (Cin, Cout, 3, 3)
    if (conv3x3_tma_supported(c, 4 * co)) {  // the transposed convolution as a four-tap, four-phase GEMM (rmvpe_conv_tma.cu)
      ConvW& u = h->up[i];
      u.img2 = (h->host16.size() + 127) / 128 * 128;
      h->host16.resize(u.img2 + (size_t)16 * co * c);
      convt2x_tma_pack_host(h->host.data() + u.w, c, co, h->host16.data() + u.img2);
      u.has_img2 = true;
    }
    if ((rc = load_blocks(h, tm, p + ".conv2", 2 * co, co, h->dec[i]))) return rc;
    c = co;
  }
  {  // cnn: Conv2d(16, 3, 3x3) with bias (RMVPE.py:320)
    const rvcdec_tensor *w = tm.get("cnn.weight", 3 * C_BASE * 9, &rc), *b = tm.get("cnn.bias", 3, &rc);
    if (!w || !b) return rc;
    h->cnn.cin = C_BASE; h->cnn.cout = 3;
    h->cnn.w = push(h->host, w->data, (size_t)w->numel);
    h->cnn.shift = push(h->host, b->data, 3);
    h->cnn.has_shift = true;
  }
  {  // BiGRU(384, 256) + Linear(512, 360)
    const int G = 3 * GRU_HID;
    h->wih = push(h->host, nullptr, (size_t)2 * G * GRU_IN);
    h->bih = push(h->host, nullptr, 2 * G);
    h->whhT = push(h->host, nullptr, (size_t)2 * GRU_HID * G);
    h->bhh = push(h->host, nullptr, 2 * G);
    for (int d = 0; d < 2; ++d) {
      const std::string sfx = d ? "_reverse" : "";
      const rvcdec_tensor *wi = tm.get("fc.0.gru.weight_ih_l0" + sfx, (int64_t)G * GRU_IN, &rc), *wh = tm.get("fc.0.gru.weight_hh_l0" + sfx, (int64_t)G * GRU_HID, &rc),
                          *bi = tm.get("fc.0.gru.bias_ih_l0" + sfx, G, &rc), *bh = tm.get("fc.0.gru.bias_hh_l0" + sfx, G, &rc);
      if (!wi || !wh || !bi || !bh) return rc;
      memcpy(h->host.data() + h->wih + (size_t)d * G * GRU_IN, wi->data, sizeof(float) * G * GRU_IN);
      memcpy(h->host.data() + h->bih + (size_t)d * G, bi->data, sizeof(float) * G);
      memcpy(h->host.data() + h->bhh + (size_t)d * G, bh->data, sizeof(float) * G);
      for (int j = 0; j < G; ++j)
        for (int k = 0; k < GRU_HID; ++k) h->host[h->whhT + ((size_t)d * GRU_HID + k) * G + j] = wh->data[(size_t)j * GRU_HID + k];
    }
    const rvcdec_tensor *fw = tm.get("fc.1.weight", (int64_t)N_CLASS * 2 * GRU_HID, &rc), *fb = tm.get("fc.1.bias", N_CLASS, &rc);
    if (!fw || !fb) return rc;
    h->fcw = push(h->host, fw->data, (size_t)fw->numel);
    h->fcb = push(h->host, fb->data, N_CLASS);
  }
  {  // mel front end: filter bank (from the host) and the windowed DFT table [cos rows | -sin rows] x 1024
    const rvcdec_tensor* mb = tm.get("mel_basis", (int64_t)N_MELS * NBINS, &rc);
    if (!mb) return rc;
    h->basis = push(h->host, mb->data, (size_t)mb->numel);
    h->dft = push(h->host, nullptr, (size_t)2 * NBINS * N_FFT);
    const double PI = 3.14159265358979323846;
    for (int nb = 0; nb < NBINS; ++nb)
      for (int k = 0; k < N_FFT; ++k) {
        const double win = 0.5 - 0.5 * std::cos(2.0 * PI * k / N_FFT);   // torch.hann_window(1024) (periodic)
        const double ang = 2.0 * PI * (double)((long long)nb * k % N_FFT) / N_FFT;
        h->host[h->dft + (size_t)nb * N_FFT + k] = (float)(win * std::cos(ang));
        h->host[h->dft + (size_t)(NBINS + nb) * N_FFT + k] = (float)(-win * std::sin(ang));
      }
  }
  MGuard g(h->device);
  MCK(g.err);
  if (h->dev) { MCK(cudaFree(h->dev)); h->dev = nullptr; }
  MCK(cudaMalloc((void**)&h->dev, h->host.size() * sizeof(float)));
  MCK(cudaMemcpy(h->dev, h->host.data(), h->host.size() * sizeof(float), cudaMemcpyHostToDevice));
  if (h->dev16) { MCK(cudaFree(h->dev16)); h->dev16 = nullptr; }
  if (!h->host16.empty()) {
    MCK(cudaMalloc((void**)&h->dev16, h->host16.size() * sizeof(__half)));
    MCK(cudaMemcpy(h->dev16, h->host16.data(), h->host16.size() * sizeof(__half), cudaMemcpyHostToDevice));
  }
  std::vector<__half>().swap(h->host16);
  h->loaded = true;
  return 0;
}

}  // extern "C"

namespace {

struct MWs {
  float* buf[4];          // rotating activations, each B x 16 x T x 128 floats (the largest tensor of the U-Net)
  float* skip[LEVELS];    // encoder outputs kept for the decoder's concatenation
  float *feat, *gi, *gout, *inter, *part;
  __half *t16, *y16a, *y16b;
  size_t bytes;
};

MWs carve_m(void* base, int B, int T) {
  MWs w{};
  size_t off = 0;
  auto take = [&](size_t nfloat) { float* p = base ? (float*)((uint8_t*)base + off) : nullptr; off += up256(nfloat * 4); return p; };
  const size_t big = (size_t)B * C_BASE * T * N_MELS;
  for (int i = 0; i < 4; ++i) w.buf[i] = take(big);
  for (int i = 0; i < LEVELS; ++i) w.skip[i] = take(big >> i);
  w.feat = take((size_t)B * T * GRU_IN);
  w.gi = take((size_t)B * T * 6 * GRU_HID);
  w.gout = take((size_t)B * T * 2 * GRU_HID);
  w.inter = take((size_t)B * 512 * (T / 32) * 4);
  {  // three channels-last fp16 buffers, each large enough for the widest convolution input that takes the TMA path
    const size_t halves = (size_t)B * (T + 2) * 130 * 32;
    w.t16 = (__half*)take(halves / 2); w.y16a = (__half*)take(halves / 2); w.y16b = (__half*)take(halves / 2);
  }
  w.part = take((size_t)16 * B * 256 * T);  // split-K partial sums: at most 16 splits of the largest layer that may split (128 ch x T/8 x 16)   // copy of the intermediate blocks' output (test hook, option keep_inter)
  w.bytes = off;
  return w;
}

// ConvBlockRes.forward (RMVPE.py:53-57): dst = relu(bn(conv(relu(bn(conv(x)))))) + (shortcut(x) | x); x = cat(in0, in1)
int run_block(rvcrmvpe_handle* h, const Block& k, const float* in0, const float* in1, int C0, int B, int H, int W, float* t1, float* t2,
              float* dst, cudaStream_t s) {
  const float* P = h->dev;
  if (h->tc && h->tma && k.c1.has_img2 && k.c2.has_img2 && (in1 == nullptr || C0 % 8 == 0) &&
      (size_t)B * (H + 2) * (W + 2) * (size_t)std::max(k.c1.cin, k.c1.cout) <= (size_t)B * (h->T_ws + 2) * 130 * 32) {
    // both convolutions as TMA-fed implicit GEMMs over channels-last fp16 copies (rmvpe_conv_tma.cu)
    const __half* x16;
    if (in1 == nullptr && h->y16_of == in0) {
      x16 = h->y16[h->y16_cur];  // left by the previous block's epilogue
    } else {
      h->y16_cur ^= 1;
      MCK(launch_rmvpe_to_nhwc16(in0, in1, in1 ? C0 : k.c1.cin, k.c1.cin, B, H, W, h->y16[h->y16_cur], s)); ++h->launches;
      x16 = h->y16[h->y16_cur];
    }
    int nl = 0;
    cudaError_t e = launch_conv3x3_tma(x16, h->dev16 + k.c1.img2, P + k.c1.scale, P + k.c1.shift, nullptr, nullptr, h->t16, B, k.c1.cin, k.c1.cout, H, W,
                                       1, 9, 0, h->part, h->part_floats, &nl, s);
    if (e == cudaSuccess) {
      h->launches += nl;
      const float* res = in0;
      if (k.shortcut && k.sc.has_img2) {
        MCK(launch_conv3x3_tma(x16, h->dev16 + k.sc.img2, nullptr, P + k.sc.shift, nullptr, t2, nullptr, B, k.sc.cin, k.sc.cout, H, W, 0, 1, 0, nullptr, 0,
                               &nl, s));
        h->launches += nl;
        res = t2;
      } else if (k.shortcut) {
        Conv2dParams q{};
        q.in0 = in0; q.in1 = in1; q.C0 = C0; q.Cin = k.sc.cin; q.w = P + k.sc.w; q.shift = P + k.sc.shift;
        q.out = t2; q.B = B; q.Cout = k.sc.cout; q.H = H; q.W = W;
        MCK(launch_conv2d(q, 1, s)); ++h->launches;
        res = t2;
      }
      __half* y16 = h->y16[h->y16_cur ^ 1];
      MCK(launch_conv3x3_tma(h->t16, h->dev16 + k.c2.img2, P + k.c2.scale, P + k.c2.shift, res, dst, y16, B, k.c2.cin, k.c2.cout, H, W, 1, 9, 0, h->part,
                             h->part_floats, &nl, s));
      h->launches += nl;
      h->y16_cur ^= 1;
      h->y16_of = dst;
      return 0;
    }
    if (e != cudaErrorNotSupported) MCK(e);
    h->y16_of = nullptr;  // no tensor maps on this driver: the gathered kernels below
  }
  h->y16_of = nullptr;
  Conv2dParams p{};
  p.in0 = in0; p.in1 = in1; p.C0 = C0; p.Cin = k.c1.cin; p.w = P + k.c1.w; p.scale = P + k.c1.scale; p.shift = P + k.c1.shift;
  p.out = t1; p.B = B; p.Cout = k.c1.cout; p.H = H; p.W = W; p.relu = 1;
  auto conv3 = [&](const Conv2dParams& q, const ConvW& cw) -> cudaError_t {
    ++h->launches;
    if (h->tc && cw.has_img && conv3x3_tc_supported(q.Cin, q.C0, q.Cout)) {
      ConvTcParams t{q.in0, q.in1, q.C0, q.Cin, h->dev16 + cw.img, q.scale, q.shift, q.res, q.out, q.B, q.Cout, q.H, q.W, q.relu, nullptr, 1};
      const int sp = conv3x3_tc_splits(q.Cin, q.Cout, q.H * q.W, q.B);
      if (sp > 1 && h->part && (size_t)sp * q.B * q.Cout * q.H * q.W <= h->part_floats) {
        t.part = h->part; t.ksplit = sp;
        ++h->launches;
      }
      return launch_conv3x3_tc(t, s);
    }
    return launch_conv2d(q, 3, s);
  };
  MCK(conv3(p, k.c1));
  const float* res = in0;  // identity shortcut: only when the input is one tensor of Cout channels
  if (k.shortcut) {
    Conv2dParams q{};
    q.in0 = in0; q.in1 = in1; q.C0 = C0; q.Cin = k.sc.cin; q.w = P + k.sc.w; q.shift = P + k.sc.shift;
    q.out = t2; q.B = B; q.Cout = k.sc.cout; q.H = H; q.W = W;
    MCK(launch_conv2d(q, 1, s)); ++h->launches;
    res = t2;
  }
  Conv2dParams r{};
  r.in0 = t1; r.C0 = k.c2.cin; r.Cin = k.c2.cin; r.w = P + k.c2.w; r.scale = P + k.c2.scale; r.shift = P + k.c2.shift; r.res = res;
  r.out = dst; r.B = B; r.Cout = k.c2.cout; r.H = H; r.W = W; r.relu = 1;
  MCK(conv3(r, k.c2));
  return 0;
}

// four blocks in a row; the result lands in `dst`.  x0 (one tensor, or the pair in0 | in1) is read by the first block only.
int run_blocks(rvcrmvpe_handle* h, const std::vector<Block>& blocks, const float* in0, const float* in1, int C0, int B, int H, int W,
               float* const scratch[3], float* dst, cudaStream_t s) {
  // scratch[0], scratch[1]: block internals; scratch[2] and dst alternate as block outputs so that the last one is dst
  const float* x0 = in0; const float* x1 = in1; int c0 = C0;
  for (size_t j = 0; j < blocks.size(); ++j) {
    float* o = ((blocks.size() - 1 - j) % 2 == 0) ? dst : scratch[2];
    int rc = run_block(h, blocks[j], x0, x1, c0, B, H, W, scratch[0], scratch[1], o, s);
    if (rc) return rc;
    x0 = o; x1 = nullptr; c0 = blocks[j].c2.cout;
  }
  return 0;
}

}  // namespace

extern "C" {

size_t rvcrmvpe_workspace_bytes(const rvcrmvpe_handle* h, int32_t B, int32_t T) {
  if (!h || B < 1 || T < 32 || T % 32) return 0;
  return carve_m(nullptr, B, T).bytes;
}

int rvcrmvpe_forward(rvcrmvpe_handle* h, const float* mel, float* hidden, void* workspace, size_t workspace_bytes, int32_t B, int32_t T,
                     void* stream) {
  if (!h || !mel || !hidden || !workspace) return mfail(RVCDEC_ERR_ARG, "null argument");
  if (!h->loaded) return mfail(RVCDEC_ERR_STATE, "rvcrmvpe_load_weights has not succeeded on this handle");
  if (B < 1 || T < 32 || T % 32) return mfail(RVCDEC_ERR_ARG, "T must be a positive multiple of 32 (mel2hidden pads to it, RMVPE.py:451-453); got %d", T);
  const size_t need = rvcrmvpe_workspace_bytes(h, B, T);
  if (workspace_bytes < need) return mfail(RVCDEC_ERR_WORKSPACE, "workspace of %zu bytes is smaller than the %zu needed for B=%d, T=%d", workspace_bytes, need, B, T);
  MGuard g(h->device);
  MCK(g.err);
  cudaStream_t s = (cudaStream_t)stream;
  MWs w = carve_m(workspace, B, T);
  const float* P = h->dev;
  h->launches = 0;
  h->part = w.part;
  h->T_ws = T;
  h->t16 = w.t16; h->y16[0] = w.y16a; h->y16[1] = w.y16b; h->y16_cur = 0; h->y16_of = nullptr;
  h->part_floats = (size_t)16 * B * 256 * T;
  int rc;
  // mel (B, 128, T) -> (B, 1, T, 128), encoder input BatchNorm
  MCK(launch_mel_in(mel, w.buf[0], B, T, h->in_a, h->in_c, s)); ++h->launches;
  float* cur = w.buf[0];
  int H = T, W = N_MELS, C = 1;
  // ---- encoder (RMVPE.py:136-142): 4 blocks -> skip, AvgPool2d(2) ----
  for (int i = 0; i < LEVELS; ++i) {
    float* scratch[3] = {nullptr, nullptr, nullptr};
    int k = 0;
    for (int j = 0; j < 4; ++j) if (w.buf[j] != cur) scratch[k++] = w.buf[j];
    if ((rc = run_blocks(h, h->enc[i], cur, nullptr, C, B, H, W, scratch, w.skip[i], s))) return rc;
    C = h->enc[i].back().c2.cout;
    MCK(launch_avgpool2(w.skip[i], scratch[0], (size_t)B * C, H, W, s)); ++h->launches;
    h->y16_of = nullptr;  // scratch[0] now holds another tensor
    cur = scratch[0];
    H /= 2; W /= 2;
  }
  // ---- intermediate (RMVPE.py:169-172) ----
  for (int i = 0; i < INTERS; ++i) {
    float* scratch[3]; int k = 0; float* dst = nullptr;
    for (int j = 0; j < 4; ++j) if (w.buf[j] != cur) { if (k < 3) scratch[k++] = w.buf[j]; }
    dst = scratch[2];  // run_blocks alternates scratch[2] / dst: give it a fourth buffer by reusing the rotation below
    // three free buffers: two for block internals, one that alternates with `cur`'s old place is not possible (x0 is read
    // by block 0 only, afterwards its buffer is free): blocks 1.. may write where the level's input was
    float* sc3[3] = {scratch[0], scratch[1], cur};
    if ((rc = run_block(h, h->inter[i][0], cur, nullptr, C, B, H, W, scratch[0], scratch[1], dst, s))) return rc;
    C = h->inter[i][0].c2.cout;
    const float* x = dst; float* o = cur;
    for (size_t j = 1; j < h->inter[i].size(); ++j) {
      if ((rc = run_block(h, h->inter[i][j], x, nullptr, C, B, H, W, sc3[0], sc3[1], o, s))) return rc;
      float* nx = const_cast<float*>(x); x = o; o = nx;
    }
    cur = const_cast<float*>(x);
  }
  if (h->keep_inter) MCK(cudaMemcpyAsync(w.inter, cur, (size_t)B * C * H * W * sizeof(float), cudaMemcpyDeviceToDevice, s));
  // ---- decoder (RMVPE.py:209-214, :240-243): ConvTranspose2d + BN + ReLU, cat with the skip, 4 blocks ----
  for (int i = 0; i < LEVELS; ++i) {
    float* fr[3]; int k = 0;
    for (int j = 0; j < 4; ++j) if (w.buf[j] != cur) fr[k++] = w.buf[j];
    const ConvW& u = h->up[i];
    bool up_done = false;
    if (h->tc && h->tma && u.has_img2 && (size_t)(H + 2) * (W + 2) * u.cin <= (size_t)(T + 2) * 130 * 32) {
      const __half* x16;
      if (h->y16_of == cur) {
        x16 = h->y16[h->y16_cur];
      } else {
        h->y16_cur ^= 1;
        MCK(launch_rmvpe_to_nhwc16(cur, nullptr, u.cin, u.cin, B, H, W, h->y16[h->y16_cur], s)); ++h->launches;
        x16 = h->y16[h->y16_cur];
      }
      int nl = 0;
      cudaError_t e = launch_conv3x3_tma(x16, h->dev16 + u.img2, P + u.scale, P + u.shift, nullptr, fr[0], nullptr, B, u.cin, 4 * u.cout, H, W, 1, 4,
                                         u.cout, nullptr, 0, &nl, s);
      if (e == cudaSuccess) { h->launches += nl; up_done = true; }
      else if (e != cudaErrorNotSupported) MCK(e);
    }
    if (!up_done) { MCK(launch_convt2x(cur, P + u.w, P + u.scale, P + u.shift, fr[0], B, u.cin, u.cout, H, W, s)); ++h->launches; }
    h->y16_of = nullptr;
    H *= 2; W *= 2; C = u.cout;
    // block 0 reads cat(fr[0], skip); internals fr[1], fr[2]; output -> cur (its old content, the level's input, is dead)
    if ((rc = run_block(h, h->dec[i][0], fr[0], w.skip[LEVELS - 1 - i], C, B, H, W, fr[1], fr[2], cur, s))) return rc;
    const float* x = cur; float* o = fr[0];
    for (size_t j = 1; j < h->dec[i].size(); ++j) {
      if ((rc = run_block(h, h->dec[i][j], x, nullptr, C, B, H, W, fr[1], fr[2], o, s))) return rc;
      float* nx = const_cast<float*>(x); x = o; o = nx;
    }
    cur = const_cast<float*>(x);
  }
  // ---- cnn (16 -> 3) stored as (B, T, 3 * 128) (RMVPE.py:337) ----
  {
    Conv2dParams p{};
    p.in0 = cur; p.C0 = C_BASE; p.Cin = C_BASE; p.w = P + h->cnn.w; p.shift = P + h->cnn.shift; p.out = w.feat;
    p.B = B; p.Cout = 3; p.H = H; p.W = W; p.out_btcw = 1;
    MCK(launch_conv2d(p, 3, s)); ++h->launches;
  }
  // ---- BiGRU + Linear + sigmoid (RMVPE.py:322-327) ----
  MCK(launch_sgemm_nt(w.feat, GRU_IN, 0, P + h->wih, P + h->bih, w.gi, 6 * GRU_HID, 0, 1, B * T, 6 * GRU_HID, GRU_IN, 0, s)); ++h->launches;
  MCK(launch_gru(w.gi, P + h->whhT, P + h->bhh, w.gout, B, T, s)); ++h->launches;
  MCK(launch_sgemm_nt(w.gout, 2 * GRU_HID, 0, P + h->fcw, P + h->fcb, hidden, N_CLASS, 0, 1, B * T, N_CLASS, 2 * GRU_HID, 1, s)); ++h->launches;
  return 0;
}

size_t rvcrmvpe_mel_workspace_bytes(const rvcrmvpe_handle* h, int32_t B, int32_t n) {
  if (!h || B < 1 || n < 1) return 0;
  const size_t frames = 1 + (size_t)n / HOP;
  return up256((size_t)B * ((size_t)n + N_FFT) * 4) + up256((size_t)B * frames * 2 * NBINS * 4);
}

int rvcrmvpe_mel(rvcrmvpe_handle* h, const float* audio, float* mel, void* workspace, size_t workspace_bytes, int32_t B, int32_t n,
                 void* stream) {
  if (!h || !audio || !mel || !workspace) return mfail(RVCDEC_ERR_ARG, "null argument");
  if (!h->loaded) return mfail(RVCDEC_ERR_STATE, "rvcrmvpe_load_weights has not succeeded on this handle");
  if (B < 1 || n <= N_FFT / 2) return mfail(RVCDEC_ERR_ARG, "need more than %d samples per item (reflect padding of torch.stft)", N_FFT / 2);
  if (workspace_bytes < rvcrmvpe_mel_workspace_bytes(h, B, n)) return mfail(RVCDEC_ERR_WORKSPACE, "mel workspace too small");
  MGuard g(h->device);
  MCK(g.err);
  cudaStream_t s = (cudaStream_t)stream;
  const int frames = 1 + n / HOP;
  const size_t np = (size_t)n + N_FFT;
  float* padded = (float*)workspace;
  float* spec = (float*)((uint8_t*)workspace + up256((size_t)B * np * 4));
  MCK(launch_reflect_pad(audio, padded, B, n, N_FFT / 2, s));
  // frame f = padded[f * 160 .. f * 160 + 1024): the DFT as one product with the windowed cos / sin table
  MCK(launch_sgemm_nt(padded, HOP, np, h->dev + h->dft, nullptr, spec, 2 * NBINS, (size_t)frames * 2 * NBINS, B, frames, 2 * NBINS, N_FFT, 0, s));
  MCK(launch_mel_log(spec, h->dev + h->basis, mel, B, frames, NBINS, 1e-5f, s));
  h->launches = 3;
  return 0;
}

int rvcrmvpe_decode(rvcrmvpe_handle* h, const float* hidden, float thred, float* f0, int32_t T, void* stream) {
  if (!h || !hidden || !f0 || T < 1) return mfail(RVCDEC_ERR_ARG, "bad argument");
  MGuard g(h->device);
  MCK(g.err);
  MCK(launch_rmvpe_decode(hidden, thred, f0, T, (cudaStream_t)stream));
  return 0;
}

int64_t rvcrmvpe_debug_read(rvcrmvpe_handle* h, const char* name, float* dst, int64_t dst_numel, void* workspace, int32_t B, int32_t T,
                            void* stream) {
  if (!h || !name || !dst || !workspace) return mfail(RVCDEC_ERR_ARG, "null argument");
  MGuard g(h->device);
  MCK(g.err);
  MWs w = carve_m(workspace, B, T);
  const std::string nm = name;
  const float* src = nullptr;
  int64_t n = 0;
  if (nm == "cnn") { src = w.feat; n = (int64_t)B * T * GRU_IN; }
  else if (nm == "gru") { src = w.gout; n = (int64_t)B * T * 2 * GRU_HID; }
  else if (nm == "skip4") { src = w.skip[4]; n = (int64_t)B * 256 * (T / 16) * 8; }
  else if (nm == "inter") {
    if (!h->keep_inter) return mfail(RVCDEC_ERR_ARG, "inter is kept only with the keep_inter option");
    src = w.inter; n = (int64_t)B * 512 * (T / 32) * 4;
  }
  else return mfail(RVCDEC_ERR_ARG, "unknown intermediate '%s'", name);
  if (dst_numel < n) return mfail(RVCDEC_ERR_ARG, "destination too small");
  MCK(cudaMemcpyAsync(dst, src, (size_t)n * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return n;
}

}  // extern "C"
