/*
 * rvcdec.h - C ABI of the B200-native NSF-HiFiGAN decoder (librvcdec.so).
 *
 * This is the drop-in boundary for ONE path of the reference
 * (Acelogic/Retrieval-based-Voice-Conversion-MLX): HiFiGANNSFGenerator.forward,
 * rvc/lib/algorithm/generators/hifigan_nsf.py:173-212, as called from
 * Synthesizer.infer (rvc/lib/algorithm/synthesizers.py:237-241).  The reference
 * has no FFI for this path - its boundary is a Python class - so the entry
 * points below are what a ctypes stub placed in the reference's
 * rvc/lib/algorithm/generators/ would bind.  Plain C types
 * only; every pointer documented as "device" is a CUDA device pointer of the
 * device the handle was created on, every pointer documented as "host" is
 * ordinary (preferably pinned) host memory.
 *
 * All functions return 0 on success and a negative rvcdec_status otherwise;
 * nothing throws across this boundary.  rvcdec_last_error() returns a
 * thread-local human-readable message for the last failure.
 * There is no CPU fallback: every compute entry point fails with
 * RVCDEC_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef RVCDEC_H_
#define RVCDEC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RVCDEC_MAX_UPS 8
#define RVCDEC_MAX_KERNELS 8
#define RVCDEC_MAX_DILATIONS 8

typedef enum {
  RVCDEC_OK = 0,
  RVCDEC_ERR_ARG = -1,         /* bad argument / unsupported configuration */
  RVCDEC_ERR_CUDA = -2,        /* CUDA runtime error (message has the detail) */
  RVCDEC_ERR_WEIGHTS = -3,     /* missing / mis-shaped tensor in load_weights */
  RVCDEC_ERR_WORKSPACE = -4,   /* workspace too small for even one batch item */
  RVCDEC_ERR_STATE = -5        /* forward before load_weights, etc. */
} rvcdec_status;

/* Arithmetic of the convolution stack.
 *   FP32: CUDA-core FFMA kernels, fp32 everywhere (parity mode, <= 2e-3 max-abs).
 *   FAST: tcgen05 tensor-core implicit GEMMs, 16-bit operands, fp32 accumulate,
 *         fp32 residual stream (>= 0.999 multi-resolution log-mel correlation). */
typedef enum {
  RVCDEC_MODE_FP32 = 0,
  RVCDEC_MODE_FAST = 1,       /* bf16 operands up to and including stage 1, fp16 after (the default fast plan) */
  RVCDEC_MODE_FAST_F16 = 2,   /* fp16 operands everywhere (experiments; range-unsafe for arbitrary checkpoints) */
  RVCDEC_MODE_FAST_BF16 = 3   /* bf16 operands everywhere (experiments; misses the fast-mode bar, SURVEY 8d probe) */
} rvcdec_mode;

/* The constructor values of the reference class, hifigan_nsf.py:75-86
 * (checkpointing is training-only and has no counterpart here). */
typedef struct {
  int32_t initial_channel;                 /* 192 */
  int32_t upsample_initial_channel;        /* 512 */
  int32_t gin_channels;                    /* 256, 0 = no conditioning conv */
  int32_t sr;                              /* 48000 */
  int32_t num_upsamples;                   /* 4 */
  int32_t upsample_rates[RVCDEC_MAX_UPS];          /* 12,10,2,2 (odd rates need kernel = 2 x rate) */
  int32_t upsample_kernel_sizes[RVCDEC_MAX_UPS];   /* 24,20,4,4 */
  int32_t num_kernels;                     /* 3 */
  int32_t resblock_kernel_sizes[RVCDEC_MAX_KERNELS];   /* 3,7,11 (odd) */
  int32_t num_dilations;                   /* 3 */
  int32_t resblock_dilation_sizes[RVCDEC_MAX_KERNELS][RVCDEC_MAX_DILATIONS]; /* 1,3,5 */
  int32_t harmonic_num;                    /* overtones of the harmonic source: 0 = HiFiGANNSFGenerator (hifigan_nsf.py:93);
                                              8 = the sibling HiFiGANMRFGenerator (hifigan_mrf.py:223-374 as built at
                                              synthesizers.py:86-97): 9 sines with random initial phases, Linear(9, 1) merge,
                                              and a conv_post with a bias */
} rvcdec_config;

/* One fp32 tensor of the decoder's state_dict, in HOST memory, named exactly as
 * the reference names it without the "dec." prefix, e.g. "conv_pre.weight",
 * "ups.0.parametrizations.weight.original0" or the legacy "ups.0.weight_g"
 * (rvc/train/utils.py:49-55), "resblocks.3.convs1.2.bias", "m_source.l_linear.weight". */
typedef struct {
  const char* name;
  const float* data;   /* host */
  int64_t numel;
} rvcdec_tensor;

typedef struct rvcdec_handle rvcdec_handle;

/* Create a decoder bound to CUDA device `device`.  Replaces
 * HiFiGANNSFGenerator.__init__ (hifigan_nsf.py:75-171). */
int rvcdec_create(const rvcdec_config* cfg, int device, rvcdec_handle** out);

/* Fold weight-norm (w = v * g/||v||, hifigan_nsf.py:120, residuals.py:14), repack
 * every layer into the kernels' layouts and upload.  Replaces
 * load_state_dict + the per-forward weight-norm recomputation of the reference
 * (rvc/infer/infer.py:479-488).  May be called again after the weights change. */
int rvcdec_load_weights(rvcdec_handle* h, const rvcdec_tensor* tensors, int32_t n);

/* Device scratch bytes for one pass over B items of T frames. Linear in B. */
size_t rvcdec_workspace_bytes(const rvcdec_handle* h, int32_t B, int32_t T, int32_t mode);

/* forward(x, f0, g) -> waveform, hifigan_nsf.py:173-212, all pointers DEVICE fp32:
 *   x     (B, initial_channel, T) contiguous
 *   f0    (B, T) Hz, 0 = unvoiced
 *   g     (B, gin_channels) or NULL
 *   noise (B, T*upp) standard-normal draws to inject (what the reference draws at
 *         hifigan.py:223), or NULL to draw them in-kernel from Philox4x32-10(seed)
 *   out   (B, T*upp)  (the reference's (B,1,T*upp))
 * If workspace_bytes is smaller than rvcdec_workspace_bytes(B) the batch is
 * processed in as many equal passes as needed.  Asynchronous on `stream`
 * (a cudaStream_t passed as void*). */
int rvcdec_forward(rvcdec_handle* h, const float* x, const float* f0, const float* g,
                   const float* noise, uint64_t seed, float* out,
                   void* workspace, size_t workspace_bytes,
                   int32_t B, int32_t T, int32_t mode, void* stream);

/* Sources with overtones (harmonic_num > 0): `noise` of rvcdec_forward is then (B, T*upp, harmonic_num+1) and the random
 * initial phases the reference draws with torch.rand(B, harmonic_num+1) (hifigan_mrf.py:129-134; column 0 is forced to 0)
 * are taken from `phases` (DEVICE, (B, harmonic_num+1), cycles in [0,1)) by every following rvcdec_forward on this handle;
 * NULL (the default) = drawn in-kernel from the Philox key. */
int rvcdec_set_initial_phases(rvcdec_handle* h, const float* phases);

/* Same call with HOST buffers (pinned memory recommended): copies inputs to the
 * device, decodes, copies the waveform back, and returns after the copy has
 * completed.  `noise` may be NULL (in-kernel draws).  Scratch is owned by the
 * handle and capped at `max_device_bytes` (0 = automatic: 60 % of the free device memory, and a large batch is
 * cut into at least three passes so that the copies of one pass overlap the kernels of its neighbours: inputs
 * and waveform are double-buffered on two copy streams). This is the call
 * the reference's pipeline makes per chunk (rvc/infer/pipeline.py:366-371:
 * net_g.infer(...)[0][0,0].data.cpu().float().numpy()). */
int rvcdec_forward_host(rvcdec_handle* h, const float* x, const float* f0, const float* g,
                        const float* noise, uint64_t seed, float* out,
                        int32_t B, int32_t T, int32_t mode, size_t max_device_bytes);

/* Tuning knobs of the fast mode (results stay within the fast-mode tolerance for every setting):
 *   "fused_mrf"  0 = every ResBlock convolution is its own kernel on every stage;
 *                1 = (default) the 64/32-channel stages run the fused ResBlock-chain kernel, one launch
 *                    per ResBlock (residuals.py:71-80 kept on chip: residual in TMEM, operands in smem);
 *                2 = same kernel, one launch per stage (all ResBlocks of the MRF over one tile).
 *   "zero_stuff" 1 = (default) a stride-2 upsample that feeds a fused-MRF stage runs as an ordinary convolution
 *                    over a zero-stuffed copy of its input (written by the previous stage's last kernel);
 *                0 = polyphase form everywhere.
 *   "poly2"      1 = a stride-2 upsample that feeds a fused-MRF stage runs in true polyphase form: dense 16-bit input, GEMM
 *                    columns ordered (slice, phase, channel) so that an accumulator row holds output rows 2t and 2t+1 of a slice,
 *                    warp-transposed coalesced fp32 epilogue with the noise conv fused; no zeros are read or multiplied.
 *                    0 = (default) the zero-stuffed form below - the two measure the same end to end (profiles/r2e_*).
 *   "ups_pm"     1 = (default) the rate-10 / 12 upsamples (hifigan_nsf.py:198) use their phase-major weight image: GEMM
 *                    columns ordered (16-channel group, half of the phases, slice, phase, channel), warp-transposed
 *                    coalesced epilogue; on a per-conv MRF stage whose noise conv has 8 taps (stage 1) the epilogue
 *                    also adds noise_convs[i](har) (:199) and writes the ResBlocks' activated 16-bit operand, which
 *                    replaces the noise_add pass.  0 = one phase per N tile with the scatter epilogue.
 *   "no_source"  1 = the decoder has no harmonic source and no noise convs - the reference's sibling HiFiGANGenerator
 *                    (rvc/lib/algorithm/generators/hifigan.py:72-97, the no-f0 models): the source kernel and every
 *                    x + noise_convs[i](har) are skipped (f0 / noise arguments are ignored); 0 = (default) NSF decoder.
 *   "conv_epi_warps" / "conv_epi_warps_small"  16 (default) or 8 epilogue warps in the 128-wide / the 64- and 32-wide
 *                    conv_mma tilings; 16 warps work on 16-column items, 8 warps on 32-column items (profiles/r1bq, r1br).
 *   "pm_a_stages" 2 (default) or 3 activation stages of the rate-10 phase-major tile (no measurable difference).
 *   "fused_post" 1 = (default) conv_post + tanh (hifigan_nsf.py:209-210) are computed in the epilogue of the last
 *                    stage's final fused-MRF launch (32-channel last stage, 7 taps, "fused_mrf" 1); 0 = own kernel.
 *   "fused_epi_warps" 8 (default) or 16 epilogue warps in the fused ResBlock kernel.
 *   "pair_mrf"   1 = the 128-channel stage runs one launch per (conv1, conv2) pair (mrf_pair.cu: fp32 residual in
 *                    TMEM, operand tile in smem, half the HBM traffic); 0 = (default) one launch per conv - the two
 *                    measure the same on B200 (profiles/r1z_*), see DESIGN.md section 3.
 *   "pair_max_k" N > 0 = on the 128-channel stage, ResBlocks whose kernel size is <= N run one fused launch per (conv1,
 *                    conv2) pair (mrf_pair.cu) while the others keep one launch per conv; 0 = none (see DESIGN.md 3 for the
 *                    measured default).
 *   "pair_noise" 1 = with "pair_mrf", the stage's short noise conv rides in its upsample's epilogue; 0 = (default).
 *   "fused_split_k" fused ResBlock chains as two launches (dilations 1,3 -> fp32 -> dilation 5), which halves the rows
 *                    recomputed at tile edges: N > 0 = chains with kernel size >= N; 0 = (default) chains whose recompute
 *                    exceeds a quarter of the tile (the k = 11 chain of the 64-channel stage); -1 = never.
 *   "max_ntile"  128 (default) or 256: widest GEMM N tile of the tensor-core weight images; takes effect at the next
 *                    rvcdec_load_weights.  128 keeps the accumulators double-buffered in TMEM (profiles/r1ab_*).
 *   "a_stages"   2 (default) or 3 activation stages in the 128-wide tiling (3 measured slower, profiles/r1af_*).
 *   "graphs"     1 = (default) the launch sequence of a pass is captured into a CUDA graph the first time a (buffers,
 *                    B, T, mode) combination is seen and replayed afterwards (the Philox key is read from device memory,
 *                    so one graph serves every seed); calls that need per-launch host side effects (profiling, debug
 *                    stops) launch directly.  0 = always launch directly.  Results are bit-identical either way.
 *   "mrf_items"  0 (default) = every launch of a per-conv MRF stage covers all items; N > 0 = launches run over groups
 *                    of items whose 16-bit hand-off tensor stays under N MiB (bit-identical results; measured slower on
 *                    30 s chunks, profiles/r1ag_*). */
int rvcdec_set_option(rvcdec_handle* h, const char* key, int64_t value);

/* Number of kernels launched by the last forward on this handle. */
int64_t rvcdec_last_launch_count(const rvcdec_handle* h);

/* Per-layer timing for bench.py.  While enabled, every forward records a CUDA event on the
 * launching stream in front of each kernel group; rvcdec_profile_read (which synchronises on the
 * last event) returns the accumulated device time, algorithmic FLOPs (2 x MAC, biases and
 * zero-padded polyphase taps excluded) and algorithmic HBM bytes (each operand/result once) per
 * layer group since the last enable: "source", "prep", "conv_pre", "ups{i}", "noise_add{i}",
 * "mrf{i}", "conv_post".  Returns the number of entries written. */
typedef struct {
  char name[32];
  double ms;
  double flops;
  double bytes;
  int64_t spans;
} rvcdec_profile_entry;
int rvcdec_profile_enable(rvcdec_handle* h, int32_t on);
int32_t rvcdec_profile_read(rvcdec_handle* h, rvcdec_profile_entry* out, int32_t max_entries);

/* ---- test hooks (used by tests/ only) ------------------------------------- */
/* Make the next forwards stop early: 0 = run everything, 1 = after the harmonic source,
 * 2 = after conv_pre, 10+i = after stage i's upsample + noise_conv add, 20+i = after stage i.
 * rvcdec_debug_read then copies a named intermediate of the LAST pass, converted to plain
 * (B, C, L) fp32, into a DEVICE buffer.  Names: "har" (B, T*upp); "pre" = leaky_relu(conv_pre
 * + cond, 0.1); "stage{i}.in" = ups + noise_conv output; "stage{i}.out" = leaky_relu(stage
 * output, 0.1) for i < last, the raw stage output for the last stage.
 * Returns the number of floats written, or <0. */
int rvcdec_debug_set_stop(rvcdec_handle* h, int32_t stop_after);
/* Counters / host-side timings of the handle: "graph_hits", "graph_captures", "graphs_cached" (CUDA-graph cache), and
 * for the last rvcdec_forward_host call "host_setup_s", "host_enqueue_s", "host_wait_s".  -1 for an unknown name. */
double rvcdec_debug_stat(const rvcdec_handle* h, const char* name);
int64_t rvcdec_debug_read(rvcdec_handle* h, const char* name, float* dst, int64_t dst_numel,
                          void* workspace, int32_t B, int32_t T, int32_t mode, void* stream);
/* Instrumented twin of the fused ResBlock kernel: after rvcdec_set_option(h, "fused_timing", 1) every
 * fused launch (up to 32) records, per SM, clock64 spans {MMA warp: wait operand rows, wait acc1 drained,
 * wait weights, total, -,-,-,-; epilogue warp: wait conv1 accumulator, wait conv2 accumulator, stage-input
 * load, total, ...} (16 int64 per SM).  Returns the number of launches recorded. */
int64_t rvcdec_debug_fused_timing(rvcdec_handle* h, int64_t* dst, int64_t dst_numel);
/* Copy a packed host-side weight image for CPU tests (no GPU needed). Layer = reference module
 * name ("conv_pre", "ups.0", "resblocks.4.convs1.2").  kind 0 = fp32 [tap][ci][n] image, 1 = fp16
 * tensor-core image, 2 = tap shifts (int32), 3 = tensor-core tap masks (uint32 per N tile),
 * 4 = bf16 tensor-core image, 5 = int32[8] {Cin, N, cout, up, ntaps, halo, n_tile, kc}, 6 = bias.
 * dst NULL returns the size in bytes. */
/* Fold + repack on the host only (no CUDA call), so the images can be inspected without a GPU. */
int rvcdec_debug_pack_host(rvcdec_handle* h, const rvcdec_tensor* tensors, int32_t n);
int64_t rvcdec_debug_packed(const rvcdec_handle* h, const char* layer, int32_t kind, void* dst, int64_t dst_bytes);

/* ---- allocation-free harmonic source -------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * har[b][s] = tanh(lin_w * SineGenerator(f0)[b][s] + lin_b), the quantity the "har" tap of rvcdec_forward exposes (one
 * sine, no overtones), computed with the frame-level fp64 phase prefix in a STATIC 512 KB device array inside the library
 * instead of a workspace: the call allocates nothing (no cudaMalloc / cudaFree, no handle, no graph) and only enqueues
 * two kernels on `stream`.  Limit: B * (T + 1) <= 65536; larger requests return RVCDEC_ERR_ARG with a message.
 * All pointers DEVICE, on the CURRENT device: f0 (B, T) Hz; noise (B, T*upp) standard-normal draws or NULL (in-kernel
 * Philox keyed by seed and the sample index); har (B, T*upp) out.  sr / upp / lin_w / lin_b: sampling rate, samples per
 * frame, and m_source.l_linear's weight and bias.
 * Every GPU that loads the library holds the array until exit; every call shares it and it keeps the last call's
 * contents (a call writes each entry before it reads it).  Calls that overlap on the GPU race on it: serialise them. */
int rvcdec_source_noalloc(const float* f0, const float* noise, uint64_t seed, float* har, int32_t B, int32_t T, int32_t upp,
                          float sr, float lin_w, float lin_b, void* stream);

void rvcdec_destroy(rvcdec_handle* h);
const char* rvcdec_last_error(void);
const char* rvcdec_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RVCDEC_H_ */
