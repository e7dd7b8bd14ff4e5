// Kernels of the Synthesizer front end (everything Synthesizer.infer computes before the decoder,
// rvc/lib/algorithm/synthesizers.py:226-236): phone / pitch embedding, windowed relative-position multi-head
// attention, LayerNorm over channels, WaveNet gate, z_p sampling.  fp32 throughout (this part is ~3 % of the FLOPs of
// net_g.infer; parity first), activations in the decoder's blocked-8 time-major layout so that the convolutions run on
// the same back-end (conv_f32.cu) and the result lands where the decoder reads its input.
#include <algorithm>
#include <cstdio>
#include "common.cuh"
#include "philox.cuh"

namespace rvcdec {

__device__ __forceinline__ size_t blk_off(int b, int nsl, int sl, int Lp, int t) {
  return (((size_t)b * nsl + sl) * Lp + PADL + t) * 8;
}

// (B, T, C) row-major fp32 (phone features, encoders.py:130) -> blocked-8
// 8 fp32 values -> one 16-byte row of the 16-bit operand copy (same element offset as the fp32 tensor): the
// tensor-core front end (mode != 0) feeds every convolution from such copies
__device__ __forceinline__ void store16(void* base16, size_t elem_off, const float (&v)[8], int fmt) {
  uint32_t pk[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (fmt == ACT_BF16) {
      __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      pk[i] = *reinterpret_cast<uint32_t*>(&h);
    } else {
      __half2 h = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
      pk[i] = *reinterpret_cast<uint32_t*>(&h);
    }
  }
  *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(base16) + elem_off) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
}

__global__ void pack_btc_kernel(const float* __restrict__ x, float* __restrict__ out, int B, int T, int C, int Lp) {
  const int nsl = C >> 3;
  const size_t total = (size_t)B * nsl * T;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int sl = (int)(i % nsl);
    const size_t bt = i / nsl;
    const int t = (int)(bt % T);
    const int b = (int)(bt / T);
    const float4* src = reinterpret_cast<const float4*>(x + ((size_t)b * T + t) * C + sl * 8);
    float4* dst = reinterpret_cast<float4*>(out + blk_off(b, nsl, sl, Lp, t));
    dst[0] = __ldg(src);
    dst[1] = __ldg(src + 1);
  }
}
cudaError_t launch_pack_btc(const float* x, float* out, int B, int T, int C, int Lp, cudaStream_t s) {
  const size_t total = (size_t)B * (C / 8) * T;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 16));
  pack_btc_kernel<<<blocks, 256, 0, s>>>(x, out, B, T, C, Lp);
  return cudaGetLastError();
}

// x = leaky_relu((x + emb_pitch[pitch]) * sqrt(H), 0.1) * mask      (encoders.py:130-139; x holds emb_phone(phone))
__global__ void embed_post_kernel(float* __restrict__ x, const long long* __restrict__ pitch, const float* __restrict__ emb_pitch,
                                  const int* __restrict__ lengths, int B, int T, int H, int Lp, float scale, void* x16, int fmt) {
  const int nsl = H >> 3;
  const size_t total = (size_t)B * nsl * T;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    const size_t bs = i / T;
    const int sl = (int)(bs % nsl);
    const int b = (int)(bs / nsl);
    float* px = x + blk_off(b, nsl, sl, Lp, t);
    float v[8];
    *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(px);
    *reinterpret_cast<float4*>(v + 4) = *reinterpret_cast<const float4*>(px + 4);
    const bool ok = t < lengths[b];
    if (pitch && emb_pitch) {
      const long long pi = pitch[(size_t)b * T + t];
      const float* e = emb_pitch + (size_t)pi * H + sl * 8;
#pragma unroll
      for (int c = 0; c < 8; ++c) v[c] += __ldg(e + c);
    }
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float s = v[c] * scale;
      v[c] = ok ? (s > 0.f ? s : 0.1f * s) : 0.f;
    }
    *reinterpret_cast<float4*>(px) = *reinterpret_cast<float4*>(v);
    *reinterpret_cast<float4*>(px + 4) = *reinterpret_cast<float4*>(v + 4);
    if (x16) store16(x16, blk_off(b, nsl, sl, Lp, t), v, fmt);
  }
}
cudaError_t launch_embed_post(float* x, const long long* pitch, const float* emb_pitch, const int* lengths, int B, int T, int H,
                              int Lp, float scale, void* x16, int fmt, cudaStream_t s) {
  const size_t total = (size_t)B * (H / 8) * T;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 16));
  embed_post_kernel<<<blocks, 256, 0, s>>>(x, pitch, emb_pitch, lengths, B, T, H, Lp, scale, x16, fmt);
  return cudaGetLastError();
}

// out[b][:] = table[idx[b]][:]       (emb_g lookup, synthesizers.py:226)
__global__ void gather_rows_kernel(const float* __restrict__ table, const long long* __restrict__ idx, float* __restrict__ out, int B,
                                   int D, int rows) {
  const int b = blockIdx.x;
  long long r = idx[b];
  if (r < 0) r = 0;
  if (r >= rows) r = rows - 1;
  for (int d = threadIdx.x; d < D; d += blockDim.x) out[(size_t)b * D + d] = table[(size_t)r * D + d];
}
cudaError_t launch_gather_rows(const float* table, const long long* idx, float* out, int B, int D, int rows, cudaStream_t s) {
  gather_rows_kernel<<<B, 128, 0, s>>>(table, idx, out, B, D, rows);
  return cudaGetLastError();
}

// LayerNorm over channels per time step (normalization.py:19-26), masked: y = LN(x) for t < len, 0 otherwise.
// One warp per time step would waste lanes at C = 192; a thread per (b, t) reads its row's slices at a stride and the
// warp's 32 consecutive time steps make every access a contiguous 1 KB run.
__global__ void layernorm_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ gamma,
                                 const float* __restrict__ beta, const int* __restrict__ lengths, int B, int T, int C, int Lp, float eps,
                                 void* y16, int fmt) {
  const int nsl = C >> 3;
  const size_t total = (size_t)B * T;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    const int b = (int)(i / T);
    const bool ok = t < lengths[b];
    float sum = 0.f;
    for (int sl = 0; sl < nsl; ++sl) {
      const float* px = x + blk_off(b, nsl, sl, Lp, t);
      const float4 a = *reinterpret_cast<const float4*>(px), c = *reinterpret_cast<const float4*>(px + 4);
      sum += (a.x + a.y) + (a.z + a.w) + (c.x + c.y) + (c.z + c.w);
    }
    const float mean = sum / (float)C;
    float var = 0.f;
    for (int sl = 0; sl < nsl; ++sl) {
      const float* px = x + blk_off(b, nsl, sl, Lp, t);
      const float4 a = *reinterpret_cast<const float4*>(px), c = *reinterpret_cast<const float4*>(px + 4);
      const float d0 = a.x - mean, d1 = a.y - mean, d2 = a.z - mean, d3 = a.w - mean, d4 = c.x - mean, d5 = c.y - mean,
                  d6 = c.z - mean, d7 = c.w - mean;
      var += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3) + (d4 * d4 + d5 * d5) + (d6 * d6 + d7 * d7);
    }
    const float rstd = rsqrtf(var / (float)C + eps);
    for (int sl = 0; sl < nsl; ++sl) {
      const size_t o = blk_off(b, nsl, sl, Lp, t);
      float v[8];
      *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(x + o);
      *reinterpret_cast<float4*>(v + 4) = *reinterpret_cast<const float4*>(x + o + 4);
#pragma unroll
      for (int c = 0; c < 8; ++c) v[c] = ok ? (v[c] - mean) * rstd * __ldg(gamma + sl * 8 + c) + __ldg(beta + sl * 8 + c) : 0.f;
      *reinterpret_cast<float4*>(y + o) = *reinterpret_cast<float4*>(v);
      *reinterpret_cast<float4*>(y + o + 4) = *reinterpret_cast<float4*>(v + 4);
      if (y16) store16(y16, o, v, fmt);
    }
  }
}
cudaError_t launch_layernorm(const float* x, float* y, const float* gamma, const float* beta, const int* lengths, int B, int T, int C,
                             int Lp, float eps, void* y16, int fmt, cudaStream_t s) {
  const size_t total = (size_t)B * T;
  const int blocks = (int)std::min<size_t>((total + 127) / 128, (size_t)(148 * 16));
  layernorm_kernel<<<blocks, 128, 0, s>>>(x, y, gamma, beta, lengths, B, T, C, Lp, eps, y16, fmt);
  return cudaGetLastError();
}

// acts = tanh(a[:, :H]) * sigmoid(a[:, H:])     (commons.py:88-103; the conditioning g_l is part of a's per-item bias)
__global__ void gate_kernel(const float* __restrict__ a, float* __restrict__ acts, int B, int T, int H, int Lp, void* acts16, int fmt) {
  const int nsl = H >> 3;
  const size_t total = (size_t)B * nsl * T;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    const size_t bs = i / T;
    const int sl = (int)(bs % nsl);
    const int b = (int)(bs / nsl);
    const float* pt = a + blk_off(b, 2 * nsl, sl, Lp, t);
    const float* ps = a + blk_off(b, 2 * nsl, nsl + sl, Lp, t);
    float tv[8], sv[8];
    *reinterpret_cast<float4*>(tv) = *reinterpret_cast<const float4*>(pt);
    *reinterpret_cast<float4*>(tv + 4) = *reinterpret_cast<const float4*>(pt + 4);
    *reinterpret_cast<float4*>(sv) = *reinterpret_cast<const float4*>(ps);
    *reinterpret_cast<float4*>(sv + 4) = *reinterpret_cast<const float4*>(ps + 4);
#pragma unroll
    for (int c = 0; c < 8; ++c) tv[c] = tanhf(tv[c]) * (1.f / (1.f + expf(-sv[c])));
    float* po = acts + blk_off(b, nsl, sl, Lp, t);
    *reinterpret_cast<float4*>(po) = *reinterpret_cast<float4*>(tv);
    *reinterpret_cast<float4*>(po + 4) = *reinterpret_cast<float4*>(tv + 4);
    if (acts16) store16(acts16, blk_off(b, nsl, sl, Lp, t), tv, fmt);
  }
}
cudaError_t launch_gate(const float* a, float* acts, int B, int T, int H, int Lp, void* acts16, int fmt, cudaStream_t s) {
  const size_t total = (size_t)B * (H / 8) * T;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 16));
  gate_kernel<<<blocks, 256, 0, s>>>(a, acts, B, T, H, Lp, acts16, fmt);
  return cudaGetLastError();
}

// z_p = (m + exp(logs) * noise * 0.66666) * mask       (synthesizers.py:228); stats = [m | logs] (2C channels, masked)
// noise: (B, C, T) fp32 draws to inject, or null -> in-kernel Philox keyed by (seed, global item, channel, frame)
__global__ void zp_kernel(const float* __restrict__ stats, const float* __restrict__ noise, float* __restrict__ zp,
                          const int* __restrict__ lengths, uint64_t seed, uint64_t item0, int B, int T, int C, int Lp, void* zp16, int fmt) {
  const int nsl = C >> 3;
  const size_t total = (size_t)B * nsl * T;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    const size_t bs = i / T;
    const int sl = (int)(bs % nsl);
    const int b = (int)(bs / nsl);
    const float* pm = stats + blk_off(b, 2 * nsl, sl, Lp, t);
    const float* pl = stats + blk_off(b, 2 * nsl, nsl + sl, Lp, t);
    float m[8], l[8];
    *reinterpret_cast<float4*>(m) = *reinterpret_cast<const float4*>(pm);
    *reinterpret_cast<float4*>(m + 4) = *reinterpret_cast<const float4*>(pm + 4);
    *reinterpret_cast<float4*>(l) = *reinterpret_cast<const float4*>(pl);
    *reinterpret_cast<float4*>(l + 4) = *reinterpret_cast<const float4*>(pl + 4);
    const bool ok = t < lengths[b];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const int ch = sl * 8 + c;
      const float z = noise ? __ldg(noise + ((size_t)b * C + ch) * T + t)
                            : philox_normal(seed ^ 0x5A17ED5EEDull, ((item0 + (uint64_t)b) * (uint64_t)C + (uint64_t)ch) * (uint64_t)T + (uint64_t)t);
      m[c] = ok ? m[c] + expf(l[c]) * z * 0.66666f : 0.f;
    }
    float* po = zp + blk_off(b, nsl, sl, Lp, t);
    *reinterpret_cast<float4*>(po) = *reinterpret_cast<float4*>(m);
    *reinterpret_cast<float4*>(po + 4) = *reinterpret_cast<float4*>(m + 4);
    if (zp16) store16(zp16, blk_off(b, nsl, sl, Lp, t), m, fmt);
  }
}
cudaError_t launch_zp(const float* stats, const float* noise, float* zp, const int* lengths, uint64_t seed, uint64_t item0, int B,
                      int T, int C, int Lp, void* zp16, int fmt, cudaStream_t s) {
  const size_t total = (size_t)B * (C / 8) * T;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 16));
  zp_kernel<<<blocks, 256, 0, s>>>(stats, noise, zp, lengths, seed, item0, B, T, C, Lp, zp16, fmt);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// Windowed relative-position self-attention (attentions.py:89-125), one (query tile, head, item) per CTA, online softmax:
//   score[t, s] = q_t . k_s / sqrt(kc) + (|s - t| <= w ? q_t . E_k[s - t + w] / sqrt(kc) : 0);   keys s >= len -> -1e4
//   out[t]      = sum_s p[t, s] v_s + sum_{|s - t| <= w} p[t, s] E_v[s - t + w]
// qkv: blocked-8 fp32 with 3*C channels ([q | k | v], each C = heads * KC).  KC = 96, window <= 10.
// 256 threads: thread (ty = tid / 16, tx = tid % 16) owns score rows 4*ty .. 4*ty+3 x key columns 4*tx .. 4*tx+3 and
// output rows 4*ty .. x channels tx + 16*j.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int ATT_BQ = 64, ATT_BK = 64, ATT_KC = 96, ATT_W = 10, ATT_NR = 2 * ATT_W + 1;
constexpr int ATT_QS = ATT_KC + 4;   // row stride of Qs (float4-aligned)
constexpr int ATT_KT = ATT_BK + 4;   // row stride of Kt[k][s]
constexpr int ATT_PS = ATT_BK + 1;   // row stride of Ps
constexpr int ATT_SMEM_FLOATS = ATT_BQ * ATT_QS + ATT_KC * ATT_KT + ATT_BK * ATT_KC + ATT_BQ * ATT_PS + ATT_BQ * ATT_NR +
                                2 * ATT_NR * ATT_KC;

__global__ void __launch_bounds__(256) attention_kernel(const float* __restrict__ qkv, float* __restrict__ out,
                                                        const float* __restrict__ emb_rel_k, const float* __restrict__ emb_rel_v,
                                                        const int* __restrict__ lengths, int B, int T, int heads, int Lp, float qscale) {
  extern __shared__ float att_smem[];
  float* Qs = att_smem;                        // [BQ][QS]   queries (pre-scaled)
  float* Kt = Qs + ATT_BQ * ATT_QS;            // [KC][KT]   keys, transposed
  float* Vs = Kt + ATT_KC * ATT_KT;            // [BK][KC]
  float* Ps = Vs + ATT_BK * ATT_KC;            // [BQ][PS]   probabilities of the current tile
  float* Rs = Ps + ATT_BQ * ATT_PS;            // [BQ][NR]   q_t . E_k[r]
  float* Ek = Rs + ATT_BQ * ATT_NR;            // [NR][KC]
  float* Ev = Ek + ATT_NR * ATT_KC;            // [NR][KC]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int q0 = blockIdx.x * ATT_BQ, h = blockIdx.y, b = blockIdx.z;
  const int C = heads * ATT_KC, nsl3 = 3 * (C >> 3), sl_per_head = ATT_KC >> 3;
  const int len = min(lengths[b], T);

  for (int i = tid; i < ATT_NR * ATT_KC; i += 256) { Ek[i] = emb_rel_k[i]; Ev[i] = emb_rel_v[i]; }
  // queries of this tile (rows beyond T read as zeros)
  for (int i = tid; i < ATT_BQ * sl_per_head; i += 256) {
    const int r = i % ATT_BQ, sl = i / ATT_BQ;
    const int t = q0 + r;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a;
    if (t < T) {
      const float* p = qkv + blk_off(b, nsl3, h * sl_per_head + sl, Lp, t);
      a = *reinterpret_cast<const float4*>(p); c = *reinterpret_cast<const float4*>(p + 4);
    }
    float* d = Qs + r * ATT_QS + sl * 8;
    d[0] = a.x * qscale; d[1] = a.y * qscale; d[2] = a.z * qscale; d[3] = a.w * qscale;
    d[4] = c.x * qscale; d[5] = c.y * qscale; d[6] = c.z * qscale; d[7] = c.w * qscale;
  }
  __syncthreads();
  for (int i = tid; i < ATT_BQ * ATT_NR; i += 256) {
    const int r = i / ATT_NR, j = i % ATT_NR;
    float acc = 0.f;
    for (int k = 0; k < ATT_KC; ++k) acc = fmaf(Qs[r * ATT_QS + k], Ek[j * ATT_KC + k], acc);
    Rs[i] = acc;
  }

  float m_run[4], l_run[4], o[4][6];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m_run[i] = -3.0e38f; l_run[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) o[i][j] = 0.f;
  }

  for (int s0 = 0; s0 < T; s0 += ATT_BK) {
    __syncthreads();  // previous tile's readers of Kt / Vs / Ps are done (also orders Rs on the first pass)
    for (int i = tid; i < ATT_BK * sl_per_head; i += 256) {
      const int r = i % ATT_BK, sl = i / ATT_BK;
      const int s = s0 + r;
      float4 ka = make_float4(0.f, 0.f, 0.f, 0.f), kb = ka, va = ka, vb = ka;
      if (s < T) {
        const float* pk = qkv + blk_off(b, nsl3, (heads + h) * sl_per_head + sl, Lp, s);
        const float* pv = qkv + blk_off(b, nsl3, (2 * heads + h) * sl_per_head + sl, Lp, s);
        ka = *reinterpret_cast<const float4*>(pk); kb = *reinterpret_cast<const float4*>(pk + 4);
        va = *reinterpret_cast<const float4*>(pv); vb = *reinterpret_cast<const float4*>(pv + 4);
      }
      const int k0 = sl * 8;
      Kt[(k0 + 0) * ATT_KT + r] = ka.x; Kt[(k0 + 1) * ATT_KT + r] = ka.y; Kt[(k0 + 2) * ATT_KT + r] = ka.z; Kt[(k0 + 3) * ATT_KT + r] = ka.w;
      Kt[(k0 + 4) * ATT_KT + r] = kb.x; Kt[(k0 + 5) * ATT_KT + r] = kb.y; Kt[(k0 + 6) * ATT_KT + r] = kb.z; Kt[(k0 + 7) * ATT_KT + r] = kb.w;
      float* dv = Vs + r * ATT_KC + k0;
      *reinterpret_cast<float4*>(dv) = va; *reinterpret_cast<float4*>(dv + 4) = vb;
    }
    __syncthreads();
    // S = Q K^T (4 x 4 per thread)
    float sc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) sc[i][j] = 0.f;
#pragma unroll 4
    for (int k = 0; k < ATT_KC; ++k) {
      const float4 kv = *reinterpret_cast<const float4*>(Kt + k * ATT_KT + 4 * tx);
      float qv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) qv[i] = Qs[(4 * ty + i) * ATT_QS + k];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        sc[i][0] = fmaf(qv[i], kv.x, sc[i][0]); sc[i][1] = fmaf(qv[i], kv.y, sc[i][1]);
        sc[i][2] = fmaf(qv[i], kv.z, sc[i][2]); sc[i][3] = fmaf(qv[i], kv.w, sc[i][3]);
      }
    }
    // relative keys, masks, online softmax
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = 4 * ty + i, t = q0 + r;
      float mx = -3.0e38f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int s = s0 + 4 * tx + j;
        const int d = s - t;
        float v = sc[i][j];
        if (d >= -ATT_W && d <= ATT_W) v += Rs[r * ATT_NR + d + ATT_W];
        if (s >= len || t >= len) v = -1e4f;          // masked_fill(mask == 0, -1e4), attentions.py:107
        if (s >= T) v = -3.0e38f;                      // beyond the tensor: not part of the softmax at all
        sc[i][j] = v;
        mx = fmaxf(mx, v);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));  // the 16 lanes of a row
      const float m_new = fmaxf(m_run[i], mx);
      const float corr = __expf(m_run[i] - m_new);
      float ps = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float p = (sc[i][j] <= -1.0e38f) ? 0.f : __expf(sc[i][j] - m_new);
        Ps[r * ATT_PS + 4 * tx + j] = p;
        ps += p;
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, off);
      l_run[i] = l_run[i] * corr + ps;
      m_run[i] = m_new;
#pragma unroll
      for (int j = 0; j < 6; ++j) o[i][j] *= corr;
    }
    __syncthreads();
    // O += P V
#pragma unroll 2
    for (int s = 0; s < ATT_BK; ++s) {
      float vv[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) vv[j] = Vs[s * ATT_KC + tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float p = Ps[(4 * ty + i) * ATT_PS + s];
#pragma unroll
        for (int j = 0; j < 6; ++j) o[i][j] = fmaf(p, vv[j], o[i][j]);
      }
    }
    // relative values: only key tiles that touch the band |s - t| <= w of this query tile
    if (s0 + ATT_BK - 1 >= q0 - ATT_W && s0 <= q0 + ATT_BQ - 1 + ATT_W) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = 4 * ty + i, t = q0 + r;
        for (int d = -ATT_W; d <= ATT_W; ++d) {
          const int sl = t + d - s0;
          if (sl < 0 || sl >= ATT_BK) continue;
          const float p = Ps[r * ATT_PS + sl];
#pragma unroll
          for (int j = 0; j < 6; ++j) o[i][j] = fmaf(p, Ev[(d + ATT_W) * ATT_KC + tx + 16 * j], o[i][j]);
        }
      }
    }
  }
  // normalise and store: channel h*KC + tx + 16 j of row t
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = q0 + 4 * ty + i;
    if (t >= T) continue;
    const float inv = 1.f / l_run[i];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const int ch = h * ATT_KC + tx + 16 * j;
      out[blk_off(b, C >> 3, ch >> 3, Lp, t) + (ch & 7)] = o[i][j] * inv;
    }
  }
}

cudaError_t launch_attention(const float* qkv, float* out, const float* emb_rel_k, const float* emb_rel_v, const int* lengths, int B,
                             int T, int heads, int kc, int window, int Lp, cudaStream_t s) {
  if (kc != ATT_KC || window != ATT_W) return cudaErrorInvalidValue;
  static unsigned long long done = 0;
  const int smem = ATT_SMEM_FLOATS * (int)sizeof(float);
  cudaError_t e = set_smem_once(reinterpret_cast<const void*>(attention_kernel), smem, done);
  if (e != cudaSuccess) return e;
  dim3 grid((T + ATT_BQ - 1) / ATT_BQ, heads, B);
  attention_kernel<<<grid, 256, smem, s>>>(qkv, out, emb_rel_k, emb_rel_v, lengths, B, T, heads, Lp, 1.f / sqrtf((float)kc));
  return cudaGetLastError();
}

// fp32 blocked-8 tensor -> 16-bit operand copy (attention output, which the o-projection reads)
__global__ void cast16_kernel(const float* __restrict__ x, void* __restrict__ y16, int fmt, int B, int T, int C, int Lp) {
  const int nsl = C >> 3;
  const size_t total = (size_t)B * nsl * T;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    const size_t bs = i / T;
    const size_t o = blk_off((int)(bs / nsl), nsl, (int)(bs % nsl), Lp, t);
    float v[8];
    *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(x + o);
    *reinterpret_cast<float4*>(v + 4) = *reinterpret_cast<const float4*>(x + o + 4);
    store16(y16, o, v, fmt);
  }
}
cudaError_t launch_cast16(const float* x, void* y16, int fmt, int B, int T, int C, int Lp, cudaStream_t s) {
  const size_t total = (size_t)B * (C / 8) * T;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 16));
  cast16_kernel<<<blocks, 256, 0, s>>>(x, y16, fmt, B, T, C, Lp);
  return cudaGetLastError();
}

// x * x_mask after a tensor-core convolution (whose epilogue knows nothing of lengths): rows t >= lengths[b] of the fp32
// tensor and / or its 16-bit copy (each with its own slices-per-item count: a coupling layer updates one channel half of z)
// are zeroed.  One CTA per (item, slice); a full-length item costs one comparison.
__global__ void mask_rows_kernel(float* __restrict__ x32, int phys32, void* __restrict__ x16, int phys16, const int* __restrict__ lengths,
                                 int T, int C, int Lp) {
  const int nsl = C >> 3;
  const int b = blockIdx.x / nsl, sl = blockIdx.x % nsl;
  const int len = lengths[b];
  for (int t = len + (int)threadIdx.x; t < T; t += blockDim.x) {
    if (x32) {
      float* p = x32 + blk_off(b, phys32, sl, Lp, t);
      *reinterpret_cast<float4*>(p) = make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4*>(p + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (x16) *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(x16) + blk_off(b, phys16, sl, Lp, t)) = make_uint4(0, 0, 0, 0);
  }
}
cudaError_t launch_mask_rows(float* x32, int phys32, void* x16, int phys16, const int* lengths, int B, int T, int C, int Lp, cudaStream_t s) {
  mask_rows_kernel<<<B * (C / 8), 128, 0, s>>>(x32, phys32 ? phys32 : C / 8, x16, phys16 ? phys16 : C / 8, lengths, T, C, Lp);
  return cudaGetLastError();
}

}  // namespace rvcdec
