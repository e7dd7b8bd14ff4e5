// Kernels of the HuBERT / ContentVec content embedder (include/rvchubert.h; the reference's HubertModelWithFinalProj,
// rvc/lib/utils.py:31-34 = transformers' HubertModel, models/hubert/modeling_hubert.py).  fp32, time-major activations
// (B, T, C): every linear layer and every strided convolution of the feature extractor is one GEMM over rows (a stride-s,
// k-tap Conv1d reads k * C contiguous floats per output row with a row pitch of s * C), so the whole model is
//   gemm128 (bias / GELU / residual epilogues), layer norm, the grouped positional convolution, attention,
// plus the first conv layer (one input channel) with its GroupNorm statistics from a 10 x 10 autocorrelation of the audio.
#include <cuda_fp16.h>

#include <cstdint>

#include "common.cuh"

namespace rvcdec {

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }

// ---------------------------------------------------------------------------------------------------------------------
// C[m, n] = epi(sum_k A[m * lda + k] * W[n * K + k] + bias[n]) (+ R[m * ldc + n]);  epi: 0 none, 1 erf-GELU.
// 128 x 128 tile, 16-deep steps through two shared-memory buffers, 8 x 8 outputs per thread (two 4-wide halves per
// direction so that every shared-memory read is one conflict-free 16-byte load).  Requires K % 16 == 0, N % 128 == 0,
// lda % 4 == 0.  blockIdx.z: batch (A + z * strideA, C / R + z * strideC).
// ---------------------------------------------------------------------------------------------------------------------
constexpr int GT = 128, GK = 16, GP = GT + 4;

template <int ACT, bool RES>
__global__ void __launch_bounds__(256, 2) gemm128_kernel(const float* __restrict__ A, size_t lda, size_t strideA, const float* __restrict__ W,
                                                         const float* __restrict__ bias, const float* __restrict__ R, float* __restrict__ C,
                                                         size_t ldc, size_t strideC, int M, int N, int K) {
  __shared__ __align__(16) float As[2][GK][GP];
  __shared__ __align__(16) float Bs[2][GK][GP];
  A += blockIdx.z * strideA;
  C += blockIdx.z * strideC;
  const int m0 = blockIdx.y * GT, n0 = blockIdx.x * GT;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int lrow0 = tid >> 2, lkq = tid & 3;  // this thread's two rows of each operand tile: lrow0, lrow0 + 64
  const bool va0 = m0 + lrow0 < M, va1 = m0 + lrow0 + 64 < M;
  const float* a0p = A + (size_t)(m0 + lrow0) * lda + lkq * 4;
  const float* a1p = A + (size_t)(m0 + lrow0 + 64) * lda + lkq * 4;
  const float* b0p = W + (size_t)(n0 + lrow0) * K + lkq * 4;
  const float* b1p = W + (size_t)(n0 + lrow0 + 64) * K + lkq * 4;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 ra0 = va0 ? __ldg((const float4*)a0p) : z4, ra1 = va1 ? __ldg((const float4*)a1p) : z4;
  float4 rb0 = __ldg((const float4*)b0p), rb1 = __ldg((const float4*)b1p);
  auto stash = [&](int buf) {
    As[buf][lkq * 4 + 0][lrow0] = ra0.x; As[buf][lkq * 4 + 1][lrow0] = ra0.y; As[buf][lkq * 4 + 2][lrow0] = ra0.z; As[buf][lkq * 4 + 3][lrow0] = ra0.w;
    As[buf][lkq * 4 + 0][lrow0 + 64] = ra1.x; As[buf][lkq * 4 + 1][lrow0 + 64] = ra1.y; As[buf][lkq * 4 + 2][lrow0 + 64] = ra1.z; As[buf][lkq * 4 + 3][lrow0 + 64] = ra1.w;
    Bs[buf][lkq * 4 + 0][lrow0] = rb0.x; Bs[buf][lkq * 4 + 1][lrow0] = rb0.y; Bs[buf][lkq * 4 + 2][lrow0] = rb0.z; Bs[buf][lkq * 4 + 3][lrow0] = rb0.w;
    Bs[buf][lkq * 4 + 0][lrow0 + 64] = rb1.x; Bs[buf][lkq * 4 + 1][lrow0 + 64] = rb1.y; Bs[buf][lkq * 4 + 2][lrow0 + 64] = rb1.z; Bs[buf][lkq * 4 + 3][lrow0 + 64] = rb1.w;
  };
  stash(0);
  __syncthreads();
  const int nk = K / GK;
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) {
      const int ko = (kt + 1) * GK;
      ra0 = va0 ? __ldg((const float4*)(a0p + ko)) : z4;
      ra1 = va1 ? __ldg((const float4*)(a1p + ko)) : z4;
      rb0 = __ldg((const float4*)(b0p + ko));
      rb1 = __ldg((const float4*)(b1p + ko));
    }
#pragma unroll
    for (int k = 0; k < GK; ++k) {
      const float4 al = *(const float4*)&As[buf][k][ty * 4], ah = *(const float4*)&As[buf][k][64 + ty * 4];
      const float4 bl = *(const float4*)&Bs[buf][k][tx * 4], bh = *(const float4*)&Bs[buf][k][64 + tx * 4];
      const float a[8] = {al.x, al.y, al.z, al.w, ah.x, ah.y, ah.z, ah.w};
      const float b[8] = {bl.x, bl.y, bl.z, bl.w, bh.x, bh.y, bh.z, bh.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kt + 1 < nk) stash(buf ^ 1);
    __syncthreads();
  }
  const float* Rz = RES ? R + blockIdx.z * strideC : nullptr;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + i - 4);
    if (m >= M) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int n = n0 + jh * 64 + tx * 4;
      float4 v = make_float4(acc[i][jh * 4], acc[i][jh * 4 + 1], acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]);
      if (bias) {
        const float4 bb = __ldg((const float4*)(bias + n));
        v.x += bb.x; v.y += bb.y; v.z += bb.z; v.w += bb.w;
      }
      if (ACT == 1) { v.x = gelu_erf(v.x); v.y = gelu_erf(v.y); v.z = gelu_erf(v.z); v.w = gelu_erf(v.w); }
      if (RES) {
        const float4 rr = __ldg((const float4*)(Rz + (size_t)m * ldc + n));
        v.x += rr.x; v.y += rr.y; v.z += rr.z; v.w += rr.w;
      }
      *(float4*)(C + (size_t)m * ldc + n) = v;
    }
  }
}

cudaError_t launch_gemm128(const float* A, size_t lda, size_t strideA, const float* W, const float* bias, const float* R, float* C, size_t ldc,
                           size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s) {
  if (K % GK || N % GT || lda % 4 || ldc % 4 || M <= 0) return cudaErrorInvalidValue;
  dim3 grid(N / GT, (M + GT - 1) / GT, batch);
  if (R) {
    if (act) gemm128_kernel<1, true><<<grid, 256, 0, s>>>(A, lda, strideA, W, bias, R, C, ldc, strideC, M, N, K);
    else gemm128_kernel<0, true><<<grid, 256, 0, s>>>(A, lda, strideA, W, bias, R, C, ldc, strideC, M, N, K);
  } else {
    if (act) gemm128_kernel<1, false><<<grid, 256, 0, s>>>(A, lda, strideA, W, bias, R, C, ldc, strideC, M, N, K);
    else gemm128_kernel<0, false><<<grid, 256, 0, s>>>(A, lda, strideA, W, bias, R, C, ldc, strideC, M, N, K);
  }
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// out[r] = LayerNorm(a[r] (+ b[r])) * gamma + beta over C channels, one warp per row (two-pass mean / variance in fp32).
// ---------------------------------------------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(256) hubert_ln_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, float* __restrict__ out, __half* __restrict__ out16,
                                                        size_t rows, float eps) {
  constexpr int PER = C / 32;
  const size_t r = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= rows) return;
  float v[PER];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int c = lane + 32 * i;
    v[i] = a[r * C + c] + (b ? b[r * C + c] : 0.f);
    sum += v[i];
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum * (1.f / C);
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < PER; ++i) { const float d = v[i] - mean; sq = fmaf(d, d, sq); }
#pragma unroll
  for (int o = 16; o; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float inv = rsqrtf(sq * (1.f / C) + eps);
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int c = lane + 32 * i;
    const float y = (v[i] - mean) * inv * __ldg(gamma + c) + __ldg(beta + c);
    if (out) out[r * C + c] = y;
    if (out16) out16[r * C + c] = __float2half_rn(y);
  }
}
cudaError_t launch_hubert_ln(const float* a, const float* b, const float* gamma, const float* beta, float* out, __half* out16, size_t rows, int C,
                             float eps, cudaStream_t s) {
  const unsigned grid = (unsigned)((rows + 7) / 8);
  if (C == 768) hubert_ln_kernel<768><<<grid, 256, 0, s>>>(a, b, gamma, beta, out, out16, rows, eps);
  else if (C == 512) hubert_ln_kernel<512><<<grid, 256, 0, s>>>(a, b, gamma, beta, out, out16, rows, eps);
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// First conv layer: Conv1d(1, 512, k = 10, stride 5, no bias) -> GroupNorm(512 groups: per channel over time) -> GELU
// (HubertGroupNormConvLayer, modeling_hubert.py:154-175).  The conv output is linear in the audio, so the per-channel mean
// and mean square over time follow from S[j] = sum_t a[5 t + j] and R[j][j'] = sum_t a[5 t + j] a[5 t + j'] (65 numbers per
// item, accumulated in fp64): mean_c = w_c . S / L, E[y^2]_c = w_c^T R w_c / L.  stats layout per item: S[10], then R as
// the 55 entries j <= j' row by row.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int F0K = 10, F0S = 5, F0NS = F0K + F0K * (F0K + 1) / 2;

__global__ void __launch_bounds__(256) hubert_feat0_stats_kernel(const float* __restrict__ audio, size_t n, int L0, double* __restrict__ stats) {
  const float* a = audio + (size_t)blockIdx.y * n;
  float acc[F0NS];
#pragma unroll
  for (int i = 0; i < F0NS; ++i) acc[i] = 0.f;
  const int t0 = blockIdx.x * 256 * 16;
  for (int it = 0; it < 16; ++it) {
    const int t = t0 + it * 256 + threadIdx.x;
    if (t >= L0) break;
    float x[F0K];
#pragma unroll
    for (int j = 0; j < F0K; ++j) x[j] = __ldg(a + (size_t)t * F0S + j);
    int q = F0K;
#pragma unroll
    for (int j = 0; j < F0K; ++j) {
      acc[j] += x[j];
#pragma unroll
      for (int k = j; k < F0K; ++k) { acc[q] = fmaf(x[j], x[k], acc[q]); ++q; }
    }
  }
  __shared__ double red[8][F0NS];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < F0NS; ++i) {
    double v = (double)acc[i];
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[wid][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < F0NS) {
    double v = 0.0;
    for (int w = 0; w < 8; ++w) v += red[w][threadIdx.x];
    atomicAdd(stats + (size_t)blockIdx.y * F0NS + threadIdx.x, v);
  }
}

// per (item, channel): scale = gamma / sqrt(var + eps), shift = beta - mean * scale
__global__ void hubert_feat0_affine_kernel(const double* __restrict__ stats, const float* __restrict__ w, const float* __restrict__ gamma,
                                           const float* __restrict__ beta, float* __restrict__ scale, float* __restrict__ shift, int L0, int Cout,
                                           float eps) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
  if (c >= Cout) return;
  const double* st = stats + (size_t)b * F0NS;
  double wv[F0K];
  for (int j = 0; j < F0K; ++j) wv[j] = (double)w[c * F0K + j];
  double m = 0.0, e2 = 0.0;
  int q = F0K;
  for (int j = 0; j < F0K; ++j) {
    m += wv[j] * st[j];
    for (int k = j; k < F0K; ++k) e2 += (k == j ? 1.0 : 2.0) * wv[j] * wv[k] * st[q++];
  }
  m /= L0; e2 /= L0;
  const double var = fmax(e2 - m * m, 0.0);
  const double sc = (double)gamma[c] / sqrt(var + (double)eps);
  scale[(size_t)b * Cout + c] = (float)sc;
  shift[(size_t)b * Cout + c] = (float)((double)beta[c] - m * sc);
}

// out[b][t][c] = gelu(scale[b][c] * sum_j w[c][j] a[b][5 t + j] + shift[b][c]);  32 frames x 512 channels per CTA
__global__ void __launch_bounds__(256) hubert_feat0_kernel(const float* __restrict__ audio, size_t n, const float* __restrict__ w,
                                                           const float* __restrict__ scale, const float* __restrict__ shift, float* __restrict__ out,
                                                           __half* __restrict__ out16, int L0, int Lstride, int Cout) {
  constexpr int TT = 32;
  __shared__ float as[TT * F0S + F0K];
  const int b = blockIdx.y, t0 = blockIdx.x * TT;
  const float* a = audio + (size_t)b * n;
  for (int i = threadIdx.x; i < TT * F0S + F0K; i += 256) {
    const size_t p = (size_t)t0 * F0S + i;
    as[i] = p < n ? __ldg(a + p) : 0.f;
  }
  __syncthreads();
  for (int c = 2 * threadIdx.x; c < Cout; c += 512) {  // two adjacent channels per thread: 4- / 8-byte stores, a warp writes 128 / 256 B
    float w0[F0K], w1[F0K];
#pragma unroll
    for (int j = 0; j < F0K; ++j) { w0[j] = __ldg(w + c * F0K + j); w1[j] = __ldg(w + (c + 1) * F0K + j); }
    const float2 sc = *reinterpret_cast<const float2*>(scale + (size_t)b * Cout + c), sh = *reinterpret_cast<const float2*>(shift + (size_t)b * Cout + c);
    for (int tt = 0; tt < TT && t0 + tt < L0; ++tt) {
      float y0 = 0.f, y1 = 0.f;
#pragma unroll
      for (int j = 0; j < F0K; ++j) {
        const float av = as[tt * F0S + j];
        y0 = fmaf(w0[j], av, y0);
        y1 = fmaf(w1[j], av, y1);
      }
      const float g0 = gelu_erf(fmaf(y0, sc.x, sh.x)), g1 = gelu_erf(fmaf(y1, sc.y, sh.y));
      const size_t o = ((size_t)b * Lstride + t0 + tt) * Cout + c;
      if (out) *reinterpret_cast<float2*>(out + o) = make_float2(g0, g1);
      if (out16) *reinterpret_cast<__half2*>(out16 + o) = __floats2half2_rn(g0, g1);
    }
  }
}

cudaError_t launch_hubert_feat0(const float* audio, size_t n, const float* w, const float* gamma, const float* beta, double* stats, float* scale,
                                float* shift, float* out, __half* out16, int B, int L0, int Lstride, int Cout, float eps, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(stats, 0, (size_t)B * F0NS * sizeof(double), s);
  if (e != cudaSuccess) return e;
  hubert_feat0_stats_kernel<<<dim3((L0 + 4095) / 4096, B), 256, 0, s>>>(audio, n, L0, stats);
  hubert_feat0_affine_kernel<<<dim3((Cout + 127) / 128, B), 128, 0, s>>>(stats, w, gamma, beta, scale, shift, L0, Cout, eps);
  hubert_feat0_kernel<<<dim3((L0 + 31) / 32, B), 256, 0, s>>>(audio, n, w, scale, shift, out, out16, L0, Lstride, Cout);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// Positional convolution: Conv1d(768, 768, k = 128, padding 64, 16 groups of 48) + bias, last of the T + 1 outputs dropped,
// GELU (HubertPositionalConvEmbedding, modeling_hubert.py:45-92).  out[b][t][g 48 + co] = gelu(bias + sum_{j, ci}
// w[g][ci][j][co] x[b][t + j - 64][g 48 + ci]).  One CTA: 64 frames of one group; 192 threads = 48 output channels x 4
// frame quarters, 16 frames each, sliding a 23-frame register window over 8 taps at a time.  The weight image
// (g, ci, j, co) is packed (and weight-normalised) at load so that a warp reads consecutive co.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int PC_G = 48, PC_K = 128, PC_T = 64, PC_ROW = PC_T + PC_K - 1 + 2;  // 193: odd pitch, conflict-free fill

__global__ void __launch_bounds__(192) hubert_posconv_kernel(const float* __restrict__ x, const float* __restrict__ wp, const float* __restrict__ bias,
                                                             float* __restrict__ out, int T, int C) {
  __shared__ float xs[PC_G][PC_ROW];
  const int t0 = blockIdx.x * PC_T, g = blockIdx.y, b = blockIdx.z;
  const float* xb = x + (size_t)b * T * C + g * PC_G;
  for (int i = threadIdx.x; i < PC_G * (PC_T + PC_K - 1); i += 192) {
    const int ci = i % PC_G, tt = i / PC_G;
    const int t = t0 - PC_K / 2 + tt;
    xs[ci][tt] = (t >= 0 && t < T) ? __ldg(xb + (size_t)t * C + ci) : 0.f;
  }
  __syncthreads();
  const int co = threadIdx.x % PC_G, tq = threadIdx.x / PC_G;
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  const float* wg = wp + (size_t)g * PC_G * PC_K * PC_G + co;
  for (int ci = 0; ci < PC_G; ++ci) {
    const float* xr = &xs[ci][tq * 16];
    const float* wr = wg + (size_t)ci * PC_K * PC_G;
    for (int jb = 0; jb < PC_K; jb += 8) {
      float win[23], wv[8];
#pragma unroll
      for (int i = 0; i < 23; ++i) win[i] = xr[jb + i];
#pragma unroll
      for (int j = 0; j < 8; ++j) wv[j] = __ldg(wr + (size_t)(jb + j) * PC_G);
#pragma unroll
      for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fmaf(win[i + j], wv[j], acc[i]);
    }
  }
  const float bb = __ldg(bias + g * PC_G + co);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int t = t0 + tq * 16 + i;
    if (t < T) out[((size_t)b * T + t) * C + g * PC_G + co] = gelu_erf(acc[i] + bb);
  }
}
cudaError_t launch_hubert_posconv(const float* x, const float* wp, const float* bias, float* out, int B, int T, int C, cudaStream_t s) {
  if (C % PC_G) return cudaErrorInvalidValue;
  hubert_posconv_kernel<<<dim3((T + PC_T - 1) / PC_T, C / PC_G, B), 192, 0, s>>>(x, wp, bias, out, T, C);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// Self-attention without mask (HubertAttention + eager_attention_forward, modeling_hubert.py:237-345): qkv (B, T, 3 * 768)
// = [q | k | v] rows from one GEMM, 12 heads of 64; ctx[b][t][h 64 + d] = softmax_k(q . k / 8) v.  One CTA: 64 queries of
// one (item, head); keys in tiles of 64 with the online softmax (running maximum / sum per query, output rescaled when the
// maximum moves); 256 threads, each a 4 x 4 block of the score tile and of the 64 x 64 output.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int AT = 64, AD = 64, AP = AT + 4;
constexpr size_t HUBERT_ATT_SMEM = (size_t)4 * AD * AP * sizeof(float);

__global__ void __launch_bounds__(256) hubert_attention_kernel(const float* __restrict__ qkv, float* __restrict__ ctx, __half* __restrict__ ctx16, int T, int heads,
                                                               float scale) {
  extern __shared__ __align__(16) float sm[];
  float(*Qs)[AP] = (float(*)[AP])sm;                    // [d][q], pre-scaled
  float(*Ks)[AP] = (float(*)[AP])(sm + AD * AP);        // [d][k]
  float(*Vs)[AP] = (float(*)[AP])(sm + 2 * AD * AP);    // [k][d]
  float(*Ps)[AP] = (float(*)[AP])(sm + 3 * AD * AP);    // [k][q]
  const int q0 = blockIdx.x * AT, h = blockIdx.y, b = blockIdx.z;
  const int C = heads * AD, C3 = 3 * C;
  const float* base = qkv + (size_t)b * T * C3 + h * AD;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  // loads: 64 rows x 16 float4; thread -> rows lr, lr + 16, ...; quad lq
  const int lr = tid >> 4, lq = tid & 15;
  for (int r = lr; r < AT; r += 16) {
    const int t = q0 + r;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < T) v = __ldg((const float4*)(base + (size_t)t * C3 + lq * 4));
    Qs[lq * 4 + 0][r] = v.x * scale; Qs[lq * 4 + 1][r] = v.y * scale; Qs[lq * 4 + 2][r] = v.z * scale; Qs[lq * 4 + 3][r] = v.w * scale;
  }
  float o[4][4], mrow[4], lrow[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    mrow[i] = -INFINITY; lrow[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) o[i][j] = 0.f;
  }
  for (int k0 = 0; k0 < T; k0 += AT) {
    __syncthreads();  // previous tile's P V is done (and Q is in place)
    for (int r = lr; r < AT; r += 16) {
      const int t = k0 + r;
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (t < T) {
        kv = __ldg((const float4*)(base + (size_t)t * C3 + C + lq * 4));
        vv = __ldg((const float4*)(base + (size_t)t * C3 + 2 * C + lq * 4));
      }
      Ks[lq * 4 + 0][r] = kv.x; Ks[lq * 4 + 1][r] = kv.y; Ks[lq * 4 + 2][r] = kv.z; Ks[lq * 4 + 3][r] = kv.w;
      *(float4*)&Vs[r][lq * 4] = vv;
    }
    __syncthreads();
    float sc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) sc[i][j] = 0.f;
#pragma unroll 8
    for (int d = 0; d < AD; ++d) {
      const float4 a = *(const float4*)&Qs[d][ty * 4], kk = *(const float4*)&Ks[d][tx * 4];
      const float av[4] = {a.x, a.y, a.z, a.w}, kv[4] = {kk.x, kk.y, kk.z, kk.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) sc[i][j] = fmaf(av[i], kv[j], sc[i][j]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (k0 + tx * 4 + j >= T) {
#pragma unroll
        for (int i = 0; i < 4; ++i) sc[i][j] = -INFINITY;
      }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float mx = fmaxf(fmaxf(sc[i][0], sc[i][1]), fmaxf(sc[i][2], sc[i][3]));
#pragma unroll
      for (int off = 8; off; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float mnew = fmaxf(mrow[i], mx);          // finite: the first tile always holds a valid key
      const float corr = expf(mrow[i] - mnew);        // exp(-inf) = 0 on the first tile
      float ps = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) { sc[i][j] = expf(sc[i][j] - mnew); ps += sc[i][j]; }
#pragma unroll
      for (int off = 8; off; off >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, off);
      lrow[i] = lrow[i] * corr + ps;
      mrow[i] = mnew;
#pragma unroll
      for (int j = 0; j < 4; ++j) o[i][j] *= corr;
#pragma unroll
      for (int j = 0; j < 4; ++j) Ps[tx * 4 + j][ty * 4 + i] = sc[i][j];
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < AT; ++k) {
      const float4 p = *(const float4*)&Ps[k][ty * 4], vv = *(const float4*)&Vs[k][tx * 4];
      const float pv[4] = {p.x, p.y, p.z, p.w}, vd[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) o[i][j] = fmaf(pv[i], vd[j], o[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = q0 + ty * 4 + i;
    if (t >= T) continue;
    const float inv = 1.f / lrow[i];
    const size_t oo = ((size_t)b * T + t) * C + h * AD + tx * 4;
    if (ctx) *(float4*)(ctx + oo) = make_float4(o[i][0] * inv, o[i][1] * inv, o[i][2] * inv, o[i][3] * inv);
    if (ctx16) {
      __half2 h0 = __floats2half2_rn(o[i][0] * inv, o[i][1] * inv), h1 = __floats2half2_rn(o[i][2] * inv, o[i][3] * inv);
      *(uint2*)(ctx16 + oo) = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
    }
  }
}
cudaError_t launch_hubert_attention(const float* qkv, float* ctx, __half* ctx16, int B, int T, int heads, cudaStream_t s) {
  static bool attr_set[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && !attr_set[dev]) {
    e = cudaFuncSetAttribute(hubert_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HUBERT_ATT_SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  hubert_attention_kernel<<<dim3((T + AT - 1) / AT, heads, B), 256, HUBERT_ATT_SMEM, s>>>(qkv, ctx, ctx16, T, heads, 0.125f);
  return cudaGetLastError();
}

}  // namespace rvcdec
