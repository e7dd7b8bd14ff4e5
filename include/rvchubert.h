/* rvchubert.h - C ABI of the B200 HuBERT / ContentVec content embedder (second half: the features feeding
 * the synthesizer).
 *
 * Replaces, for CUDA tensors, the reference's HubertModelWithFinalProj (rvc/lib/utils.py:31-34), i.e. transformers'
 * HubertModel (models/hubert/modeling_hubert.py; a third-party dependency of the reference) with the published base
 * configuration the reference's embedders ship with (contentvec / chinese- / japanese- / korean-hubert-base config.json =
 * HubertConfig defaults), as called by rvc/infer/pipeline.py:326-334: one waveform batch, no attention mask, eval mode.
 *   rvchubert_forward      model(feats)["last_hidden_state"]: 7-layer conv feature extractor (GroupNorm on the first layer,
 *                          GELU), LayerNorm + Linear 512 -> 768, weight-normed grouped positional conv + GELU, LayerNorm,
 *                          12 post-LN transformer layers (12 heads, FFN 3072, erf-GELU)
 *   rvchubert_final_proj   model.final_proj (Linear 768 -> 256; v1 voice models, pipeline.py:332-334)
 * fp32 parameters and outputs; the GEMM operands are fp16 in the default tensor-core mode (rvchubert_set_option).  Same conventions as rvcdec.h: plain C types, device pointers of the handle's device, 0 on success /
 * negative rvcdec_status, rvchubert_last_error() for the text.
 */
#ifndef RVCHUBERT_H_
#define RVCHUBERT_H_

#include <stddef.h>
#include <stdint.h>
#include "rvcdec.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rvchubert_handle rvchubert_handle;

int rvchubert_create(int device, rvchubert_handle** out);

/* HubertModelWithFinalProj's state dict under transformers' names ("feature_extractor.conv_layers.0.conv.weight",
 * "feature_projection.projection.weight", "encoder.pos_conv_embed.conv.parametrizations.weight.original0" / "...original1"
 * (or the older "...conv.weight_g" / "...conv.weight_v"), "encoder.layers.11.feed_forward.output_dense.bias", ...).
 * "final_proj.weight" / "final_proj.bias" are optional (rvchubert_final_proj fails without them); "masked_spec_embed"
 * is ignored (used only in training). */
int rvchubert_load_weights(rvchubert_handle* h, const rvcdec_tensor* tensors, int32_t n);

/* frames the feature extractor yields for n_samples of 16 kHz audio (HubertModel._get_feat_extract_output_lengths):
 * ((n - 10) / 5 + 1, then five times (L - 3) / 2 + 1 ... ) ~ n / 320; 0 when the clip is shorter than one frame (400 samples). */
int32_t rvchubert_frames(int32_t n_samples);

size_t rvchubert_workspace_bytes(const rvchubert_handle* h, int32_t B, int32_t n_samples);

/* audio (B, n_samples) fp32 -> last_hidden_state (B, rvchubert_frames(n_samples), 768) fp32. */
int rvchubert_forward(rvchubert_handle* h, const float* audio, float* hidden, void* workspace, size_t workspace_bytes, int32_t B,
                      int32_t n_samples, void* stream);

/* x (rows, 768) fp32 -> (rows, 256) fp32. */
int rvchubert_final_proj(rvchubert_handle* h, const float* x, float* out, int64_t rows, void* stream);

/* test hook: copies a named intermediate of the last rvchubert_forward ("features" = conv feature extractor output
 * (B, T, 512), "embedded" = encoder input after the positional conv and LayerNorm (B, T, 768), "layer0" = output of the first
 * transformer layer (B, T, 768); "layer0" is kept only when the debug flag was set through rvchubert_set_option) into a
 * device buffer; returns the number of floats or < 0. */
int64_t rvchubert_debug_read(rvchubert_handle* h, const char* name, float* dst, int64_t dst_numel, void* workspace, int32_t B,
                             int32_t n_samples, void* stream);

/* options: "tc" (default 1: the linear layers and strided convolutions as fp16-operand, fp32-accumulating tensor-core GEMMs,
 * ~6e-3 max error on the O(1) features; 0: fp32 CUDA-core GEMMs, 1e-5), "tma" (default 1; with tc: GEMM operands through
 * 2-D tensor maps, 0: gathered with cp.async - same arithmetic), "att_tc" (default 1; with tc: attention on the tensor
 * cores from the fp16 q | k | v rows; 0: the fp32 attention kernel), "keep_layer0" (0 / 1, default 0: copy the first
 * transformer layer's output aside for rvchubert_debug_read). */
int rvchubert_set_option(rvchubert_handle* h, const char* key, int64_t value);

int64_t rvchubert_last_launch_count(const rvchubert_handle* h);
void rvchubert_destroy(rvchubert_handle* h);
const char* rvchubert_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* RVCHUBERT_H_ */
