/* rvcrmvpe.h - C ABI of the B200 RMVPE f0 predictor (first half: the f0 front end feeding the decoder).
 *
 * Replaces, for CUDA tensors, the reference's rvc/lib/predictors/RMVPE.py:
 *   rvcrmvpe_mel       MelSpectrogram.forward (RMVPE.py:388-417, keyshift 0, speed 1, center=True): STFT magnitude (n_fft 1024,
 *                      hop 160, Hann) -> mel filter bank -> log(clamp 1e-5)
 *   rvcrmvpe_forward   E2E.forward (RMVPE.py:335-339) for E2E(4, 1, (2, 2)) (RMVPE.py:431): DeepUnet (5 encoder levels of 4
 *                      ConvBlockRes + AvgPool, 4 intermediate blocks, 5 decoder levels of ConvTranspose2d + skip concat +
 *                      4 ConvBlockRes; BatchNorm folded at load), Conv2d 16 -> 3, BiGRU(384, 256), Linear(512, 360), sigmoid
 *   rvcrmvpe_decode    RMVPE0Predictor.decode / to_local_average_cents (RMVPE.py:484-541)
 * mel2hidden's reflect padding of the frame axis to a multiple of 32 (RMVPE.py:445-482) is host-side (rvc_b200.rmvpe).
 * fp32 parameters and activations; the deep levels' 3 x 3 convolutions take fp16 operands on the tensor cores by default
 * (rvcrmvpe_set_option "tc").  Same conventions as rvcdec.h: plain C
 * types, device pointers of the handle's device, 0 on success / negative rvcdec_status, rvcrmvpe_last_error() for the text.
 */
#ifndef RVCRMVPE_H_
#define RVCRMVPE_H_

#include <stddef.h>
#include <stdint.h>
#include "rvcdec.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rvcrmvpe_handle rvcrmvpe_handle;

int rvcrmvpe_create(int device, rvcrmvpe_handle** out);

/* E2E's state dict under the reference's names ("unet.encoder.bn.weight", "unet.encoder.layers.0.conv.0.conv.0.weight",
 * "...conv.1.running_mean", "...shortcut.weight", "unet.decoder.layers.4.conv1.0.weight", "cnn.weight", "fc.0.gru.weight_ih_l0",
 * "fc.0.gru.weight_hh_l0_reverse", "fc.1.weight", ...; "num_batches_tracked" entries are ignored) plus "mel_basis": the
 * (128, 513) filter bank MelSpectrogram registers as a buffer (RMVPE.py:377-378; librosa.filters.mel, computed by the host). */
int rvcrmvpe_load_weights(rvcrmvpe_handle* h, const rvcdec_tensor* tensors, int32_t n);

/* audio (B, n_samples) fp32 at 16 kHz -> mel (B, 128, 1 + n_samples / 160) fp32.  workspace: rvcrmvpe_mel_workspace_bytes. */
size_t rvcrmvpe_mel_workspace_bytes(const rvcrmvpe_handle* h, int32_t B, int32_t n_samples);
int rvcrmvpe_mel(rvcrmvpe_handle* h, const float* audio, float* mel, void* workspace, size_t workspace_bytes, int32_t B,
                 int32_t n_samples, void* stream);

/* mel (B, 128, T) fp32 with T % 32 == 0 -> salience (B, T, 360) fp32. */
size_t rvcrmvpe_workspace_bytes(const rvcrmvpe_handle* h, int32_t B, int32_t T);
int rvcrmvpe_forward(rvcrmvpe_handle* h, const float* mel, float* hidden, void* workspace, size_t workspace_bytes, int32_t B,
                     int32_t T, void* stream);

/* salience (T, 360) fp32 -> f0 (T) fp32 in Hz (0 = unvoiced: maximum salience <= thred). */
int rvcrmvpe_decode(rvcrmvpe_handle* h, const float* hidden, float thred, float* f0, int32_t T, void* stream);

/* test hook: copies a named intermediate of the last rvcrmvpe_forward ("skip4" = last encoder level before pooling (B, 256, T/16, 8), "cnn" =
 * (B, T, 384), "gru" = (B, T, 512), "inter" = output of the intermediate blocks (B, 512, T/32, 4), kept only with the option keep_inter) into a device buffer; returns the number of floats or < 0. */
int64_t rvcrmvpe_debug_read(rvcrmvpe_handle* h, const char* name, float* dst, int64_t dst_numel, void* workspace, int32_t B, int32_t T,
                            void* stream);

/* options: "tc" (default 1: the 3 x 3 convolutions with Cin % 64 == 0 and Cout % 128 == 0 - encoder levels 3-4, the intermediate
 * blocks, decoder levels 0-1 - as fp16-operand, fp32-accumulating implicit GEMMs on the tensor cores; 0: every layer in fp32),
 * "tma" (default 1; with tc: the convolutions with Cin % 64 == 0 read a channels-last fp16 copy of their input through tensor
 * maps instead of gathering it per pixel - same arithmetic), "keep_inter" (0 / 1, default 0: test hook, see rvcrmvpe_debug_read). */
int rvcrmvpe_set_option(rvcrmvpe_handle* h, const char* key, int64_t value);

int64_t rvcrmvpe_last_launch_count(const rvcrmvpe_handle* h);
void rvcrmvpe_destroy(rvcrmvpe_handle* h);
const char* rvcrmvpe_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* RVCRMVPE_H_ */
