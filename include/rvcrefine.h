/*
 * rvcrefine.h - C ABI of the sibling decoder RefineGANGenerator on the B200 (librvcdec.so), SURVEY 8f row 3.
 *
 * rvc/lib/algorithm/generators/refinegan.py:268-451, the decoder Synthesizer builds for vocoder "RefineGAN"
 * (synthesizers.py:98-106): forward(mel, f0, g) -> waveform.  fp32 throughout (parity mode): every convolution runs on
 * the decoder's fp32 back-end (conv_f32.cu); the harmonic source on the linearly interpolated f0, the sinc / Kaiser
 * decimation, the linear upsampling + skip concatenation and the AdaIN noise layers are kernels of refine_kernels.cu.
 * Conventions (status codes, device pointers, no CPU fallback) as in rvcdec.h.
 */
#ifndef RVCREFINE_H_
#define RVCREFINE_H_

#include "rvcdec.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Constructor values of RefineGANGenerator (refinegan.py:290-302) that shape the network. */
typedef struct {
  int32_t sample_rate;                        /* 48000 */
  int32_t num_upsamples;                      /* 4 */
  int32_t upsample_rates[RVCDEC_MAX_UPS];     /* 12,10,2,2 (even rates; the downsample path uses them in reverse) */
  int32_t num_mels;                           /* 192 (= inter_channels) */
  int32_t start_channels;                     /* 16 */
  int32_t upsample_initial_channel;           /* 512 */
  int32_t gin_channels;                       /* 256, 0 = no cond */
} rvcrefine_config;

typedef struct rvcrefine_handle rvcrefine_handle;

int rvcrefine_create(const rvcrefine_config* cfg, int device, rvcrefine_handle** out);

/* The module's state dict under the reference's names ("pre_conv.weight_g" / "...parametrizations.weight.original0",
 * "upsample_conv_blocks.2.blocks.1.1.convs1.0.bias", "upsample_conv_blocks.0.blocks.0.0.weight" (AdaIN), "m_source.merge.0.weight",
 * "conv_post...") plus one tensor "resample_taps.<i>" per downsample stage i: the torchaudio sinc_interp_kaiser filter of
 * that stage's rate (2 * width + rate taps; refinegan.py:424-433), computed by the host side (rvc_b200.refinegan). */
int rvcrefine_load_weights(rvcrefine_handle* h, const rvcdec_tensor* tensors, int32_t n);

size_t rvcrefine_workspace_bytes(const rvcrefine_handle* h, int32_t B, int32_t T);

/* forward(mel, f0, g) -> waveform, all pointers DEVICE fp32:
 *   mel (B, num_mels, T)    f0 (B, T) Hz    g (B, gin_channels) or NULL    out (B, T*upp)
 *   noise  every standard-normal draw of the reference in its order, or NULL (in-kernel Philox keyed by seed):
 *          [source (B, T*upp)] then for stage i = 0.., parallel block j = 0..2: [AdaIN in front (B, C_i, L_i)] [AdaIN behind (B, C_i, L_i)]
 *          with C_i = upsample_initial_channel >> (i+1), L_i = T * rates[0] * .. * rates[i]   (refinegan.py:257, :98-99) */
int rvcrefine_forward(rvcrefine_handle* h, const float* mel, const float* f0, const float* g, const float* noise, uint64_t seed,
                      float* out, void* workspace, size_t workspace_bytes, int32_t B, int32_t T, void* stream);

int64_t rvcrefine_noise_floats(const rvcrefine_handle* h, int32_t B, int32_t T);  /* size of `noise` */
int64_t rvcrefine_last_launch_count(const rvcrefine_handle* h);
void rvcrefine_destroy(rvcrefine_handle* h);
const char* rvcrefine_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* RVCREFINE_H_ */
