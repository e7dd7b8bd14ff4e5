/*
 * rvcsyn.h - C ABI of the whole Synthesizer.infer on the B200 (librvcdec.so), SURVEY 8f row 1.
 *
 * Widens the drop-in boundary of rvcdec.h by the step immediately in front of the decoder: everything
 * rvc/lib/algorithm/synthesizers.py:205-243 computes for an f0 (or no-f0) model with rate = None -
 *   g = emb_g(sid); m_p, logs_p, x_mask = enc_p(phone, pitch, lengths)            (encoders.py:88-144, attentions.py:6-240)
 *   z_p = (m_p + exp(logs_p) * randn * 0.66666) * x_mask                           (synthesizers.py:228)
 *   z = flow(z_p, x_mask, g, reverse=True)                                         (residuals.py:103-261, modules.py:5-117)
 *   o = dec(z * x_mask, nsff0, g)                                                  (rvcdec_forward)
 * - behind one call.  The front end runs in fp32 (it is ~3 % of the FLOPs of net_g.infer) on hand-written CUDA kernels:
 * the fp32 convolution back-end of the decoder for every Linear / Conv1d, a windowed relative-position attention kernel
 * with online softmax, masked LayerNorm, the WaveNet gate, z_p sampling; the flow's Flip layers are folded into the
 * coupling layers' weights at load time.  Conventions (status codes, device pointers, no CPU fallback) as in rvcdec.h.
 */
#ifndef RVCSYN_H_
#define RVCSYN_H_

#include "rvcdec.h"

#ifdef __cplusplus
extern "C" {
#endif

/* The constructor values of Synthesizer (synthesizers.py:40-66) that the inference path uses. */
typedef struct {
  int32_t inter_channels;    /* 192 */
  int32_t hidden_channels;   /* 192 */
  int32_t filter_channels;   /* 768 */
  int32_t n_heads;           /* 2   (hidden_channels / n_heads must be 96) */
  int32_t n_layers;          /* 6 */
  int32_t kernel_size;       /* 3   (FFN convolutions) */
  int32_t text_dim;          /* text_enc_hidden_dim: 768 (v2) or 256 (v1) */
  int32_t spk_embed_dim;     /* rows of emb_g (infer.py:472) */
  int32_t gin_channels;      /* 256 */
  int32_t use_f0;            /* 1 = pitch embedding + NSF decoder, 0 = no-f0 model */
  int32_t window_size;       /* 10  (encoders.py:33) */
  int32_t flow_kernel;       /* 5   (synthesizers.py:146-152) */
  int32_t flow_layers;       /* 3 */
  int32_t n_flows;           /* 4 */
  rvcdec_config dec;         /* the decoder's configuration */
} rvcsyn_config;

typedef struct rvcsyn_handle rvcsyn_handle;

int rvcsyn_create(const rvcsyn_config* cfg, int device, rvcsyn_handle** out);

/* The Synthesizer's state dict (host fp32 tensors, the reference's names: "enc_p.emb_phone.weight",
 * "enc_p.encoder.attn_layers.0.conv_q.weight", "flow.flows.0.enc.in_layers.0.weight_g" / "...parametrizations.weight.original0",
 * "emb_g.weight", "dec.*" ...).  enc_q.* is not needed (infer.py:484 deletes it) and ignored. */
int rvcsyn_load_weights(rvcsyn_handle* h, const rvcdec_tensor* tensors, int32_t n);

/* Device scratch for one call over B items of T frames (front end + decoder). */
size_t rvcsyn_workspace_bytes(const rvcsyn_handle* h, int32_t B, int32_t T, int32_t mode);

/* net_g.infer(phone, phone_lengths, pitch, nsff0, sid) -> o.  All pointers DEVICE:
 *   phone     (B, T, text_dim) fp32        lengths (B) int32         pitch (B, T) int64 (NULL for no-f0 models)
 *   nsff0     (B, T) fp32 Hz               sid (B) int64
 *   noise_zp  (B, inter_channels, T) fp32 standard-normal draws for z_p (what torch.randn_like(m_p) yields), or NULL to
 *             draw them in-kernel (Philox keyed by seed)
 *   noise_src (B, T*upp) fp32 draws for the harmonic source, or NULL (see rvcdec_forward)
 *   out       (B, T*upp) fp32              z_out  optional (B, inter_channels, T): receives z * x_mask
 * The batch must fit the workspace in one pass (rvcsyn_workspace_bytes); asynchronous on `stream`. */
int rvcsyn_infer(rvcsyn_handle* h, const float* phone, const int32_t* lengths, const int64_t* pitch, const float* nsff0,
                 const int64_t* sid, const float* noise_zp, const float* noise_src, uint64_t seed, float* out, float* z_out,
                 void* workspace, size_t workspace_bytes, int32_t B, int32_t T, int32_t mode, void* stream);

/* Tuning knobs.  "front_fp32" (0 / 1, default 0): in the fast modes (mode != 0) everything before the decoder - q|k|v, o,
 * FFN, proj, the coupling layers' pre / WaveNet / post convolutions - runs on the tensor-core back-end from 16-bit operand
 * copies (fp16; bf16 in mode 3) with fp32 accumulation; 1 keeps that part on the fp32 back-end whatever the mode.  mode 0 is
 * fp32 throughout.  "att_mma" (0 / 1, default 1): in the fast modes the windowed relative-position attention runs on the tensor
 * cores too (scores and P V as tcgen05 GEMMs, online softmax in fp32); 0 = the fp32 attention kernel.  "graphs" (0 / 1, default
 * 1) with "graph_max_frames" (default 16384): calls with injected z_p noise and B * T up to that many frames copy their inputs to
 * fixed places in the workspace and replay the front end's ~130 launches as one CUDA graph per (workspace, B, T, mode). */
int rvcsyn_set_option(rvcsyn_handle* h, const char* key, int64_t value);

/* ---- test hooks ---- stop after: 1 embedding, 2 / 3 / 4 / 5 q|k|v / attention / first LayerNorm / end of encoder layer 0,
 * 6 encoder (stats), 7 z_p, 8 first coupling layer, 9 flow; 0 = run everything.  rvcsyn_debug_read copies a named
 * intermediate of the last call ("x", "qkv", "att", "stats", "xf", "hh") as plain (B, C, T) fp32 into a device buffer. */
int rvcsyn_debug_set_stop(rvcsyn_handle* h, int32_t stop);
int64_t rvcsyn_debug_read(rvcsyn_handle* h, const char* name, float* dst, int64_t dst_numel, void* workspace, int32_t B, int32_t T,
                          void* stream);

/* The decoder handle this synthesizer owns (options, profiling: rvcdec_set_option / rvcdec_profile_*). */
rvcdec_handle* rvcsyn_decoder(rvcsyn_handle* h);
int64_t rvcsyn_last_launch_count(const rvcsyn_handle* h);
void rvcsyn_destroy(rvcsyn_handle* h);
const char* rvcsyn_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* RVCSYN_H_ */
