// Shared definitions for the NSF-HiFiGAN decoder kernels (sm_100a).
//
// HBM layout of every activation tensor ("blocked-8, time-major"):
//   element (b, c, t) of a (B, C, L) tensor lives at
//     ((b * (C/8) + c/8) * Lp + PADL + t) * 8 + (c % 8)
//   with Lp = PADL + roundup(L, 512) + PADR rows.  A "row" is 8 consecutive
//   channels of one time step (16 B in 16-bit, 32 B in fp32); the rows of one
//   channel-slice are contiguous in time.  That makes
//     * a conv tap a pure row offset (any dilation, no alignment constraint),
//     * one channel-slice of a time tile a single contiguous run that a 1-D TMA
//       bulk copy (cp.async.bulk) moves into shared memory,
//     * the shared-memory image exactly the no-swizzle K-major UMMA operand
//       layout (8x16B core matrices, SBO = 128 B, LBO = slice stride),
//     * epilogue stores coalesced: 32 lanes = 32 consecutive rows.
//   Rows [-PADL, 0) and [L, Lp-PADL) of 16-bit operand tensors are kept zero: they
//   are the convolutions' zero padding (commons.py:19-27 get_padding).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rvcdec {

constexpr int PADL = 32;
constexpr int PADR = 32;
constexpr int ROW_ROUND = 512;
constexpr int MAX_TAPS = 12;
constexpr int MAX_NTILES = 64;

__host__ __device__ inline int padded_rows(int L) { return PADL + ((L + ROW_ROUND - 1) / ROW_ROUND) * ROW_ROUND + PADR; }

enum ActFmt : int { ACT_F16 = 0, ACT_BF16 = 1, ACT_F32 = 2 };

__host__ __device__ inline int act_bytes(int fmt) { return fmt == ACT_F32 ? 4 : 2; }

// Epilogue of every convolution kernel (both back-ends):
//   v = acc + bias[b*bias_bstride + n]
//   if (res)    v += res[..]
//   v *= scale
//   if (sum_in) v += sum_in[..]
//   if (out_f32) out_f32[..] = v
//   if (out_act) out_act[..] = cvt(leaky_relu(v, slope))
// Output column n of input row t maps to channel n % cout and time t*up + n / cout
// (up > 1 is the polyphase transposed convolution: the phase is folded into N).
struct Epilogue {
  const float* bias;
  const float* res;
  const float* sum_in;
  float* out_f32;
  void* out_act;
  float scale;
  float slope;
  int bias_bstride;
  int act_fmt;   // format of out_act
  int up;
  int cout;      // physical channels of the output tensors
  int L_out;     // valid rows of the output tensors
  int Lp_out;    // padded rows of the output tensors
  // optional fused noise_convs[i] (hifigan_nsf.py:199): v += nz_bias[co] + sum_j har[b][row*stride - pad + j] * nz_w[j][co]
  // (applied before res / scale; used by the upsample layers whose noise conv has <= 8 taps)
  const float* nz_har;   // (B, nz_src) or null
  const float* nz_w;     // [k][cout]
  const float* nz_bias;  // [cout]
  int nz_k, nz_stride, nz_pad, nz_src;
  // optional zero-stuffed 16-bit output: out_act row t goes to row t*act_zs of a tensor with Lp_act padded
  // rows, rows t*act_zs+1 .. are written as zeros.  A following stride-2 transposed convolution then is an
  // ordinary convolution over that tensor (ConvTranspose1d == conv over the zero-stuffed input with the
  // flipped kernel), which gives it the coalesced "output row == accumulator row" epilogue.
  int act_zs;   // 0 / 1 = dense
  int Lp_act;   // padded rows of out_act when act_zs > 1
  // 1 = phase-major polyphase upsample (tensor-core back-end only): the GEMM columns are ordered
  // (channel group of 16, half of the phases, slice of the group, phase within the half, channel), see
  // conv_mma.cu "phase-major"; requires an even `up` and cout % 16 == 0.  A fused noise conv must then have
  // stride 4, 8 taps, padding 2.
  int pm;
  // optional per-item valid length (rows >= lengths[b] are written as zeros: the x * x_mask of the reference's masked
  // sequence models, encoders.py:139; honoured by the fp32 back-end only)
  const int* lengths;
};

// One convolution as an implicit GEMM:  D[t, n] = sum_tap sum_ci A[t + shift[tap], ci] * W[tap][ci][n]
struct ConvParams {
  const void* act_in;      // blocked-8 input, format act_fmt_in
  const void* w;           // packed weights (layout depends on the back-end)
  int act_fmt_in;
  int B, Cin, L_in, Lp_in;
  int N_total;             // total GEMM columns (= up * cout)
  int ntaps;
  int halo;                // max |shift|
  int shift[MAX_TAPS];
  uint32_t tapmask[MAX_NTILES];  // per N tile: bit j set = tap j has non-zero weights
  int m_tiles, n_tiles;    // tiles per batch item along time, tiles along N
  int a_stages;            // activation ring depth of the 128-wide tensor-core tiling (3; anything else = 2)
  int epi_warps_small;     // same for the 64- / 32-wide tilings (zero-stuffed stride-2 upsamples, per-conv stages 2-3)
  int epi_warps;           // 16 = the 128-wide tiling runs 16 epilogue warps over 16-column items; anything else = 8 / 32
  int lean;                // 1 = ordinary convolutions of the 128-wide tiling may run the LEAN instantiation (option "conv_lean")
  int in_slices_phys;      // 8-channel slices per item of the tensor act_in points into (0 = Cin / 8); lets a
                           // convolution read a channel sub-range of a wider tensor (coupling layers, residuals.py:239)
  Epilogue epi;
};

__device__ __forceinline__ float lrelu(float v, float slope) { return v > 0.f ? v : v * slope; }

// Apply the epilogue to NV (4 or 8) consecutive GEMM columns n0.. of input row t_in of item b.
// MASK = true also honours Epilogue::lengths (fp32 back-end only: the tensor-core kernels are compiled without it, the
// extra live pointer cost the 128-wide tiling 16 B of spills and 2 % of its time, A/B profiles/r2n_*)
template <int NV, bool MASK = false>
__device__ __forceinline__ void epilogue_store(const Epilogue& e, int b, int t_in, int n0, float (&v)[NV]) {
  const int r = n0 / e.cout;
  const int co = n0 - r * e.cout;
  const int row = t_in * e.up + r;
  if (row >= e.L_out) return;
  const float* bp = e.bias + (size_t)b * e.bias_bstride + n0;
#pragma unroll
  for (int i = 0; i < NV; i += 4) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(bp + i));
    v[i] += q.x; v[i + 1] += q.y; v[i + 2] += q.z; v[i + 3] += q.w;
  }
  if (e.nz_har) {
#pragma unroll
    for (int i = 0; i < NV; i += 4) {
      const float4 q = __ldg(reinterpret_cast<const float4*>(e.nz_bias + co + i));
      v[i] += q.x; v[i + 1] += q.y; v[i + 2] += q.z; v[i + 3] += q.w;
    }
    const float* hb = e.nz_har + (size_t)b * e.nz_src;
    const int s0 = row * e.nz_stride - e.nz_pad;
    for (int j = 0; j < e.nz_k; ++j) {
      const int sidx = s0 + j;
      if (sidx < 0 || sidx >= e.nz_src) continue;
      const float h = __ldg(hb + sidx);
#pragma unroll
      for (int i = 0; i < NV; i += 4) {
        const float4 q = __ldg(reinterpret_cast<const float4*>(e.nz_w + (size_t)j * e.cout + co + i));
        v[i] = fmaf(h, q.x, v[i]); v[i + 1] = fmaf(h, q.y, v[i + 1]); v[i + 2] = fmaf(h, q.z, v[i + 2]); v[i + 3] = fmaf(h, q.w, v[i + 3]);
      }
    }
  }
  const size_t off = (((size_t)b * (e.cout >> 3) + (co >> 3)) * e.Lp_out + PADL + row) * 8 + (co & 7);
  if (e.res) {
#pragma unroll
    for (int i = 0; i < NV; i += 4) {
      const float4 q = *reinterpret_cast<const float4*>(e.res + off + i);
      v[i] += q.x; v[i + 1] += q.y; v[i + 2] += q.z; v[i + 3] += q.w;
    }
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] *= e.scale;
  if (e.sum_in) {
#pragma unroll
    for (int i = 0; i < NV; i += 4) {
      const float4 q = *reinterpret_cast<const float4*>(e.sum_in + off + i);
      v[i] += q.x; v[i + 1] += q.y; v[i + 2] += q.z; v[i + 3] += q.w;
    }
  }
  if constexpr (MASK) {
    if (e.lengths && row >= __ldg(e.lengths + b)) {
#pragma unroll
      for (int i = 0; i < NV; ++i) v[i] = 0.f;
    }
  }
  if (e.out_f32) {
#pragma unroll
    for (int i = 0; i < NV; i += 4)
      *reinterpret_cast<float4*>(e.out_f32 + off + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
  }
  if (e.out_act) {
    if (e.act_fmt == ACT_F32) {
      float* o = reinterpret_cast<float*>(e.out_act) + off;
#pragma unroll
      for (int i = 0; i < NV; i += 4)
        *reinterpret_cast<float4*>(o + i) = make_float4(lrelu(v[i], e.slope), lrelu(v[i + 1], e.slope),
                                                        lrelu(v[i + 2], e.slope), lrelu(v[i + 3], e.slope));
    } else {
      uint32_t pk[NV / 2];
      if (e.act_fmt == ACT_F16) {
#pragma unroll
        for (int i = 0; i < NV; i += 2) {
          __half2 h = __floats2half2_rn(lrelu(v[i], e.slope), lrelu(v[i + 1], e.slope));
          pk[i / 2] = *reinterpret_cast<uint32_t*>(&h);
        }
      } else {
#pragma unroll
        for (int i = 0; i < NV; i += 2) {
          __nv_bfloat162 h = __floats2bfloat162_rn(lrelu(v[i], e.slope), lrelu(v[i + 1], e.slope));
          pk[i / 2] = *reinterpret_cast<uint32_t*>(&h);
        }
      }
      uint16_t* o = reinterpret_cast<uint16_t*>(e.out_act) + off;
      if (NV == 8) *reinterpret_cast<uint4*>(o) = make_uint4(pk[0], pk[1], pk[NV / 2 - 2], pk[NV / 2 - 1]);
      else *reinterpret_cast<uint2*>(o) = make_uint2(pk[0], pk[1]);
    }
  }
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is per (function, device): set it once, not on every launch
// (59 launches per single-utterance decode each paid a driver call for it).  `done` is a per-call-site bit mask
// by device ordinal; a lost race only repeats the idempotent call.
inline cudaError_t set_smem_once(const void* func, int bytes, unsigned long long& done) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && ((done >> dev) & 1ull)) return cudaSuccess;
  e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess && dev < 64) done |= 1ull << dev;
  return e;
}

// ---- launchers implemented in the .cu files --------------------------------
cudaError_t launch_conv_f32(const ConvParams& p, cudaStream_t s);
// returns cudaErrorInvalidValue when no instantiation fits (N tile / K chunk)
cudaError_t launch_conv_mma(const ConvParams& p, int n_tile, int kc, int num_sms, cudaStream_t s);
bool conv_mma_supported(int n_tile, int kc);

// ---- fused ResBlock chain (mrf_fused.cu) ------------------------------------
// One ResBlock (residuals.py:71-80) = nd x [ xt = conv1_d(lrelu(x)); xt = conv2(lrelu(xt)); x = xt + x ].
// The kernel keeps x (fp32) in TMEM for the whole chain and the two 16-bit operand tiles in shared
// memory; only the stage input is read from HBM and only the MRF sum / stage output is written.
constexpr int FUSED_MAX_D = 4;   // dilations per ResBlock
constexpr int FUSED_MAX_RB = 4;  // ResBlocks per launch
struct FusedConv {
  const void* w;      // tensor-core image of the conv: [tap][C/8][C][8] 16-bit, one tap block = C*C*2 bytes
  const float* bias;  // conv1: its bias; conv2: CUMULATIVE bias b2[0]+..+b2[m] (x stays un-biased in TMEM)
  int ntaps, dil;
};
struct FusedRB {
  FusedConv c1[FUSED_MAX_D], c2[FUSED_MAX_D];
  int sum_in;   // 1: add the running MRF sum (p.sum) to this ResBlock's output
  int is_last;  // 1: this ResBlock completes the MRF average -> write out_f32 / out_act; 0: write p.sum
};
struct FusedParams {
  const float* x_in;  // stage input, fp32 blocked-8 (B, C, L)
  float* sum;         // MRF accumulator, fp32 blocked-8
  float* out_f32;     // stage output (average), fp32 blocked-8, or null
  void* out_act;      // leaky_relu(stage output, out_slope) in out_fmt, blocked-8, or null
  int fmt;            // operand format of this stage (ACT_F16 / ACT_BF16)
  int out_fmt;
  int out_zs;         // > 1: out_act is zero-stuffed (row t -> t*out_zs, see Epilogue::act_zs), Lp_out_act padded rows
  int Lp_out_act;
  float scale;        // 1 / num_kernels
  float slope;        // 0.1 inside the ResBlocks
  float out_slope;
  int B, L, Lp, C;
  int nd, n_rb;
  int margin;         // rows at each tile edge that are recomputed by the neighbouring tile
  int tiles_per_item;
  int epi_warps;      // 8 or 16 epilogue warps (anything else = 16)
  long long* timing;  // null, or [gridDim.x][16] clock64 spans (instrumented kernel twin)
  // optional fused tail (last stage, C = 32, 7 taps): post_out[b][t] = tanh(conv_post(leaky_relu(avg, post_slope)))
  // (hifigan_nsf.py:209-210) computed from the last ResBlock's epilogue instead of writing the fp32 average to HBM;
  // `margin` must then include the 3 extra rows per side that conv_post reads.
  const float* post_w;  // [7][C], or null
  float post_wc[7 * 32]; // the same weights in the kernel's constant bank (filled by the launcher): the tail's FMAs read
                         // them as constant operands instead of taking shared-memory bandwidth from the MMAs
  float* post_out;      // (B, L)
  float post_slope;
  float post_bias;      // conv_post bias (0 for the NSF decoder, hifigan_nsf.py:164; the MRF sibling has one, hifigan_mrf.py:327)
  FusedRB rb[FUSED_MAX_RB];
};
bool mrf_fused_supported(int C, int max_taps, int max_halo);
bool mrf_fused_post_supported(int C, int post_k);
int mrf_fused_tile_rows(int C);
cudaError_t launch_mrf_fused(const FusedParams& p, int num_sms, cudaStream_t s);

// ---- fused ResBlock pair for the 128-channel stage (mrf_pair.cu) --------------
//   v = x + conv2(lrelu(conv1(lrelu(x)) + b1)) + b2 ;  o = v * scale (+ sum_in) ;  out_f32 = o ; out_act = cvt(lrelu(o, out_slope))
struct PairParams {
  const float* x_in;          // fp32 blocked-8 (B, 128, L); must not alias out_f32 (tiles read their neighbours' rows)
  const void* w1;             // conv_mma weight images, N tile 128 / K chunk 64: [tap][chunk][slice 8][n 128][8] 16-bit
  const void* w2;
  const void* w1h;            // the same weights split by output-channel half: [tap][half 2][slice 16][n 64][8] (mrf_pair2.cu)
  const void* w2h;
  const float* b1;
  const float* b2;
  int k1, d1, k2, d2;         // taps / dilation of conv1, conv2
  int fmt;                    // operand format (ACT_F16 / ACT_BF16)
  float slope;                // 0.1 inside the ResBlocks
  int B, L, Lp;
  int margin;                 // halo(conv1) + halo(conv2): rows per tile edge recomputed by the neighbouring tile
  int tiles_per_item;
  float scale;                // 1 (inner pairs) or 1 / num_kernels (last pair of a ResBlock)
  const float* sum_in;        // running MRF sum to add, or null
  float* out_f32;             // residual stream for the next pair / running MRF sum / stage output, or null
  void* out_act;              // leaky_relu(o, out_slope) as the next stage's 16-bit operand, or null
  int out_fmt, out_zs, Lp_out_act;
  float out_slope;
  long long* timing;          // null, or [gridDim.x][16] clock64 spans (instrumented twin)
};
bool mrf_pair_supported(int C, int halo1, int halo2);
cudaError_t launch_mrf_pair(const PairParams& p, int num_sms, cudaStream_t s);
// the same pair on CTA pairs (cta_group::2, two tiles in flight; mrf_pair2.cu)
bool mrf_pair2_supported(int C, int halo1, int halo2);
cudaError_t launch_mrf_pair2(const PairParams& p, int num_sms, cudaStream_t s);

struct SourceParams {
  const float* f0;      // (B, T)
  const float* noise;   // (B, T*upp) or null
  float* har;           // (B, T*upp)
  const uint64_t* seed_dev;  // non-null: the key is read from device memory (launch sequences replayed as CUDA graphs)
  uint64_t seed;     // Philox key of the whole call
  uint64_t item0;    // global index of item 0 of this launch: draw (item, sample) is a function of (seed, item0 + b, sample)
                     // only, so the waveform does not depend on how a batch is cut into passes
  int B, T, upp;
  float sr, lin_w, lin_b;
  // sources with overtones (HiFiGANMRFGenerator, hifigan_mrf.py:86-176): H = harmonic_num + 1 sines, harmonic k at (k+1) f0
  // with initial phase rand_ini[b][k] (cycles; null = in-kernel uniform draws, 0 for k = 0), noise (B, T*upp, H), merged by
  // Linear(H, 1) = lin_wv . v + lin_b.  H = 1 is the NSF source (lin_wv[0] = lin_w).
  int H;
  float lin_wv[16];
  const float* rand_ini;
};
cudaError_t launch_source(const SourceParams& p, cudaStream_t s);
cudaError_t launch_set_u64(uint64_t* dst, uint64_t v, cudaStream_t s);

// biasB[b][n] = bias_pre[n] + (g ? bias_cond[n] + sum_c Wc[n][c] g[b][c] : 0)
cudaError_t launch_cond(const float* g, const float* Wc, const float* bias_cond, const float* bias_pre,
                        float* biasB, int B, int gin, int N, cudaStream_t s);
// (B, C, T) fp32 contiguous -> blocked-8 act tensor (format fmt)
cudaError_t launch_pack_input(const float* x, void* act, int fmt, int B, int C, int T, int Lp, cudaStream_t s);
// zero the pad rows [-PADL,0) and [L, Lp-PADL) of up to 3 blocked-8 tensors of the same geometry
cudaError_t launch_zero_pads(void* a0, void* a1, void* a2, int elem_bytes, int B, int C, int L, int Lp, cudaStream_t s);
// x += noise_conv(har) (+bias); writes x (fp32 blocked) and act = lrelu(x, slope)
struct NoiseAddParams {
  const float* har;     // (B, n_src)
  const float* w;       // [k][C] transposed noise_conv weight
  const float* bias;    // [C]
  float* x;             // fp32 blocked (B, C, L)
  void* act;            // blocked act out
  int act_fmt;
  int B, C, L, Lp, n_src, k, stride, pad;
  float slope;
};
cudaError_t launch_noise_add(const NoiseAddParams& p, cudaStream_t s);
// out[b][t] = tanh(sum_j sum_c w[j][c] * lrelu(x[b][c][t+j-pad], slope))
cudaError_t launch_post(const float* x, const float* w, float* out, int B, int C, int L, int Lp, int k, float slope,
                        cudaStream_t s, float bias = 0.f);
// blocked-8 (fp32 or 16-bit) -> plain (B, C, L) fp32, for tests
cudaError_t launch_unpack(const void* src, int fmt, float* dst, int B, int C, int L, int Lp, cudaStream_t s, int row_stride = 1);

}  // namespace rvcdec
