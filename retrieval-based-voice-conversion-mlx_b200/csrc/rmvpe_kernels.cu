// Kernels of the RMVPE f0 predictor (rvc/lib/predictors/RMVPE.py): fp32, plain NCHW activations (B, C, H = frames, W = mel
// bins), everything the decoder's back-ends do not have - 2-D convolutions with folded BatchNorm, stride-2 transposed
// convolution, average pooling, a small SGEMM (STFT as a DFT matrix product, GRU input projection, the final Linear), the
// GRU recurrence, mel / log and the salience -> f0 decode.  Parity first: none of this is tuned beyond sane tiling.
#include <algorithm>
#include <cstdio>
#include "common.cuh"

namespace rvcdec {

// ---------------------------------------------------------------------------------------------------------------------
// conv2d, kernel KS x KS (3 or 1), stride 1, padding KS / 2:
//   out[b, co, y, x] = act(scale[co] * sum_{ci, ky, kx} in[b, ci, y + ky - p, x + kx - p] w[co, ci, ky, kx] + shift[co]) + res
// The input may be the channel concatenation of two tensors (in0: C0 channels, in1: Cin - C0), as the decoder blocks read
// torch.cat((x, skip), 1) (RMVPE.py:211).  One thread = one pixel x 16 output channels; a CTA covers th x tw pixels
// (tw = min(32, W), th * tw = 256) and stages 8 input channels and their weights per step in shared memory.
// out_btcw: store as (B, H, Cout, W) instead of (B, Cout, H, W) (the transpose(1, 2).flatten(-2) of E2E.forward).
// ---------------------------------------------------------------------------------------------------------------------
struct Conv2dParams {
  const float* in0; const float* in1; int C0, Cin;
  const float* w;       // [Cout][Cin][KS][KS]
  const float* scale;   // [Cout] or null (= 1)
  const float* shift;   // [Cout] or null (= 0)
  const float* res;     // (B, Cout, H, W) added after the activation, or null
  float* out;
  int B, Cout, H, W, relu, out_btcw, tw;
};

template <int KS>
__global__ void __launch_bounds__(256) conv2d_kernel(Conv2dParams p) {
  constexpr int CI = 8, CO = 16, PADK = KS / 2;
  __shared__ float tile[CI * 66 * 6 > CI * 10 * 34 ? CI * 66 * 6 : CI * 10 * 34];  // (th + 2) x (tw + 2) per channel
  __shared__ __align__(16) float wsm[CI * KS * KS * CO];
  const int tw = p.tw, th = 256 / tw;
  const int tx = threadIdx.x % tw, ty = threadIdx.x / tw;
  const int tiles_x = (p.W + tw - 1) / tw;
  const int x0 = (blockIdx.x % tiles_x) * tw, y0 = (blockIdx.x / tiles_x) * th;
  const int co0 = blockIdx.y * CO, b = blockIdx.z;
  const int sw = tw + 2 * PADK, sh = th + 2 * PADK;
  float acc[CO];
#pragma unroll
  for (int o = 0; o < CO; ++o) acc[o] = 0.f;
  // What a thread stages per 8-channel step does not depend on the step: (channel within the step, pixel offset) of its
  // tile elements and the offsets of its weights are computed once (the divisions by the run-time tile shape cost more
  // than the loads), and the NEXT step's values are fetched into registers while the current one is computed.
  constexpr int NT = (CI * 66 * 6 + 255) / 256, NW = (CI * KS * KS * CO + 255) / 256;
  const int n_tile = CI * sh * sw;
  int tpix[NT];   // pixel offset inside a channel plane, -1 = zero padding / outside, -2 = no element
  int tch[NT];
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int i = threadIdx.x + 256 * j;
    tpix[j] = -2; tch[j] = 0;
    if (i < n_tile) {
      const int c = i / (sh * sw), r = i - c * sh * sw;
      const int yy = y0 + r / sw - PADK, xx = x0 + r % sw - PADK;
      tch[j] = c;
      tpix[j] = (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W) ? yy * p.W + xx : -1;
    }
  }
  int woff[NW], wch[NW];
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const int i = threadIdx.x + 256 * j;
    woff[j] = -1; wch[j] = 0;
    if (i < CI * KS * KS * CO) {
      const int o = i % CO, t = (i / CO) % (KS * KS), c = i / (CO * KS * KS);
      wch[j] = c;
      if (co0 + o < p.Cout) woff[j] = ((co0 + o) * p.Cin + c) * KS * KS + t;
    }
  }
  const size_t plane = (size_t)p.H * p.W;
  const float* base0 = p.in0 + (size_t)b * p.C0 * plane;
  const float* base1 = p.in1 ? p.in1 + (size_t)b * (p.Cin - p.C0) * plane : nullptr;
  float tv[NT], wv_[NW];
  auto fetch = [&](int c0) {
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int ci = c0 + tch[j];
      tv[j] = 0.f;
      if (tpix[j] >= 0 && ci < p.Cin) tv[j] = __ldg((ci < p.C0 ? base0 + (size_t)ci * plane : base1 + (size_t)(ci - p.C0) * plane) + tpix[j]);
    }
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      wv_[j] = 0.f;
      if (woff[j] >= 0 && c0 + wch[j] < p.Cin) wv_[j] = __ldg(p.w + woff[j] + (size_t)c0 * KS * KS);
    }
  };
  fetch(0);
  for (int c0 = 0; c0 < p.Cin; c0 += CI) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < NT; ++j) if (tpix[j] != -2) tile[threadIdx.x + 256 * j] = tv[j];
#pragma unroll
    for (int j = 0; j < NW; ++j) if (threadIdx.x + 256 * j < CI * KS * KS * CO) wsm[threadIdx.x + 256 * j] = wv_[j];
    __syncthreads();
    if (c0 + CI < p.Cin) fetch(c0 + CI);
#pragma unroll 1
    for (int c = 0; c < CI; ++c) {
#pragma unroll
      for (int t = 0; t < KS * KS; ++t) {
        const float v = tile[c * sh * sw + (ty + t / KS) * sw + tx + t % KS];
        const float4* wv = reinterpret_cast<const float4*>(wsm + (c * KS * KS + t) * CO);
#pragma unroll
        for (int q = 0; q < CO / 4; ++q) {
          const float4 w4 = wv[q];
          acc[4 * q] = fmaf(v, w4.x, acc[4 * q]); acc[4 * q + 1] = fmaf(v, w4.y, acc[4 * q + 1]);
          acc[4 * q + 2] = fmaf(v, w4.z, acc[4 * q + 2]); acc[4 * q + 3] = fmaf(v, w4.w, acc[4 * q + 3]);
        }
      }
    }
  }
  const int y = y0 + ty, x = x0 + tx;
  if (y >= p.H || x >= p.W) return;
#pragma unroll
  for (int o = 0; o < CO; ++o) {
    const int co = co0 + o;
    if (co >= p.Cout) break;
    float v = acc[o] * (p.scale ? __ldg(p.scale + co) : 1.f) + (p.shift ? __ldg(p.shift + co) : 0.f);
    if (p.relu) v = fmaxf(v, 0.f);
    if (p.res) v += p.res[(((size_t)b * p.Cout + co) * p.H + y) * p.W + x];
    const size_t oi = p.out_btcw ? (((size_t)b * p.H + y) * p.Cout + co) * p.W + x : (((size_t)b * p.Cout + co) * p.H + y) * p.W + x;
    p.out[oi] = v;
  }
}

cudaError_t launch_conv2d(const Conv2dParams& p0, int ks, cudaStream_t s) {
  Conv2dParams p = p0;
  p.tw = p.W >= 32 ? 32 : (p.W >= 16 ? 16 : (p.W >= 8 ? 8 : 4));
  const int th = 256 / p.tw;
  dim3 grid(((p.W + p.tw - 1) / p.tw) * ((p.H + th - 1) / th), (p.Cout + 15) / 16, p.B);
  if (ks == 3) conv2d_kernel<3><<<grid, 256, 0, s>>>(p);
  else if (ks == 1) conv2d_kernel<1><<<grid, 256, 0, s>>>(p);
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// ConvTranspose2d(k = 3, stride 2, padding 1, output_padding 1) + folded BatchNorm + ReLU (RMVPE.py:190-202):
//   out[b, co, Y, X] = relu(scale * sum_{ci, ky, kx : (Y + 1 - ky), (X + 1 - kx) even} in[b, ci, (Y+1-ky)/2, (X+1-kx)/2] w[ci, co, ky, kx] + shift)
// weight layout (Cin, Cout, 3, 3) as torch stores it.  One thread = one output pixel x 16 output channels.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) convt2x_kernel(const float* __restrict__ in, const float* __restrict__ w, const float* __restrict__ scale,
                                                      const float* __restrict__ shift, float* __restrict__ out, int B, int Cin, int Cout,
                                                      int H, int W) {
  constexpr int CI = 8, CO = 16;
  __shared__ __align__(16) float wsm[CI * 9 * CO];
  const int HO = 2 * H, WO = 2 * W;
  const size_t pix = blockIdx.x * (size_t)256 + threadIdx.x;
  const bool live = pix < (size_t)HO * WO;
  const int Y = live ? (int)(pix / WO) : 0, X = live ? (int)(pix % WO) : 0;
  const int co0 = blockIdx.y * CO, b = blockIdx.z;
  float acc[CO];
#pragma unroll
  for (int o = 0; o < CO; ++o) acc[o] = 0.f;
  for (int c0 = 0; c0 < Cin; c0 += CI) {
    __syncthreads();
    for (int i = threadIdx.x; i < CI * 9 * CO; i += 256) {
      const int o = i % CO, t = (i / CO) % 9, c = i / (CO * 9);
      const int co = co0 + o, ci = c0 + c;
      wsm[i] = (co < Cout && ci < Cin) ? __ldg(w + ((size_t)ci * Cout + co) * 9 + t) : 0.f;
    }
    __syncthreads();
    if (!live) continue;
    for (int c = 0; c < CI && c0 + c < Cin; ++c) {
      const float* src = in + ((size_t)b * Cin + c0 + c) * H * W;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        const int ny = Y + 1 - ky;
        if (ny < 0 || (ny & 1) || (ny >> 1) >= H) continue;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const int nx = X + 1 - kx;
          if (nx < 0 || (nx & 1) || (nx >> 1) >= W) continue;
          const float v = __ldg(src + (size_t)(ny >> 1) * W + (nx >> 1));
          const float* wv = wsm + (c * 9 + ky * 3 + kx) * CO;
#pragma unroll
          for (int o = 0; o < CO; ++o) acc[o] = fmaf(v, wv[o], acc[o]);
        }
      }
    }
  }
  if (!live) return;
#pragma unroll
  for (int o = 0; o < CO; ++o) {
    const int co = co0 + o;
    if (co >= Cout) break;
    out[(((size_t)b * Cout + co) * HO + Y) * WO + X] = fmaxf(acc[o] * __ldg(scale + co) + __ldg(shift + co), 0.f);
  }
}
cudaError_t launch_convt2x(const float* in, const float* w, const float* scale, const float* shift, float* out, int B, int Cin, int Cout,
                           int H, int W, cudaStream_t s) {
  dim3 grid((unsigned)(((size_t)4 * H * W + 255) / 256), (Cout + 15) / 16, B);
  convt2x_kernel<<<grid, 256, 0, s>>>(in, w, scale, shift, out, B, Cin, Cout, H, W);
  return cudaGetLastError();
}

// AvgPool2d(2) (RMVPE.py:81)
__global__ void avgpool2_kernel(const float* __restrict__ in, float* __restrict__ out, size_t planes, int H, int W) {
  const int HO = H / 2, WO = W / 2;
  const size_t total = planes * HO * WO;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % WO), y = (int)((i / WO) % HO);
    const size_t pl = i / ((size_t)WO * HO);
    const float* s = in + (pl * H + 2 * y) * W + 2 * x;
    out[i] = 0.25f * ((s[0] + s[1]) + (s[W] + s[W + 1]));
  }
}
cudaError_t launch_avgpool2(const float* in, float* out, size_t planes, int H, int W, cudaStream_t s) {
  const size_t total = planes * (H / 2) * (W / 2);
  avgpool2_kernel<<<(int)std::min<size_t>((total + 255) / 256, 148 * 16), 256, 0, s>>>(in, out, planes, H, W);
  return cudaGetLastError();
}

// mel (B, 128, T) -> x (B, 1, T, 128) with the encoder's input BatchNorm (one channel: x * a + c) (RMVPE.py:138, :336)
__global__ void mel_in_kernel(const float* __restrict__ mel, float* __restrict__ x, int B, int T, float a, float c) {
  const size_t total = (size_t)B * T * 128;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int m = (int)(i % 128), t = (int)((i / 128) % T), b = (int)(i / ((size_t)128 * T));
    x[i] = mel[((size_t)b * 128 + m) * T + t] * a + c;
  }
}
cudaError_t launch_mel_in(const float* mel, float* x, int B, int T, float a, float c, cudaStream_t s) {
  const size_t total = (size_t)B * T * 128;
  mel_in_kernel<<<(int)std::min<size_t>((total + 255) / 256, 148 * 16), 256, 0, s>>>(mel, x, B, T, a, c);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// C[m, n] = act(sum_k A[m * lda + k] * Bm[n * K + k] + bias[n])     (A rows may overlap: lda < K is the STFT's frame hop)
// 64 x 64 tile, 16-deep steps, 4 x 4 outputs per thread.  act: 0 none, 1 sigmoid.
// batch: blockIdx.z selects A + z * strideA, C + z * strideC (Bm shared).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sgemm_nt_kernel(const float* __restrict__ A, size_t lda, size_t strideA, const float* __restrict__ Bm,
                                                       const float* __restrict__ bias, float* __restrict__ C, size_t ldc, size_t strideC,
                                                       int M, int N, int K, int act) {
  __shared__ float As[16][64 + 1], Bs[16][64 + 1];
  A += blockIdx.z * strideA;
  C += blockIdx.z * strideC;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      const int r = i >> 4, k = i & 15;
      As[k][r] = (m0 + r < M && k0 + k < K) ? __ldg(A + (size_t)(m0 + r) * lda + k0 + k) : 0.f;
      Bs[k][r] = (n0 + r < N && k0 + k < K) ? __ldg(Bm + (size_t)(n0 + r) * K + k0 + k) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float a[4], bb[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[k][ty * 4 + i]; bb[i] = Bs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j] + (bias ? __ldg(bias + n) : 0.f);
      if (act == 1) v = 1.f / (1.f + expf(-v));
      C[(size_t)m * ldc + n] = v;
    }
  }
}
cudaError_t launch_sgemm_nt(const float* A, size_t lda, size_t strideA, const float* Bm, const float* bias, float* C, size_t ldc,
                            size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s) {
  dim3 grid((N + 63) / 64, (M + 63) / 64, batch);
  sgemm_nt_kernel<<<grid, 256, 0, s>>>(A, lda, strideA, Bm, bias, C, ldc, strideC, M, N, K, act);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// GRU recurrence (nn.GRU, one layer, both directions; RMVPE.py:553-564).  gi = x W_ih^T + b_ih is precomputed by the SGEMM:
// (B, T, 2 * 3H) = [forward r | z | n | reverse r | z | n].  One CTA per (direction, item), 3H = 768 threads: thread j
// forms row j of W_hh h + b_hh from the TRANSPOSED recurrent weights whhT[dir][k][j] (coalesced, L2-resident: 786 KB per
// direction), the first H threads then apply the gates.  out (B, T, 2H) = [forward h | reverse h].
// ---------------------------------------------------------------------------------------------------------------------
constexpr int GRU_H = 256;
__global__ void __launch_bounds__(3 * GRU_H) gru_kernel(const float* __restrict__ gi, const float* __restrict__ whhT, const float* __restrict__ bhh,
                                                        float* __restrict__ out, int B, int T) {
  __shared__ float h[GRU_H];
  __shared__ float gh[3 * GRU_H];
  const int dir = blockIdx.x, b = blockIdx.y, j = threadIdx.x;
  const float* w = whhT + (size_t)dir * GRU_H * 3 * GRU_H;
  const float bj = bhh[dir * 3 * GRU_H + j];
  if (j < GRU_H) h[j] = 0.f;
  __syncthreads();
  for (int step = 0; step < T; ++step) {
    const int t = dir ? T - 1 - step : step;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll 8
    for (int k = 0; k < GRU_H; k += 4) {
      a0 = fmaf(__ldg(w + (size_t)k * 3 * GRU_H + j), h[k], a0);
      a1 = fmaf(__ldg(w + (size_t)(k + 1) * 3 * GRU_H + j), h[k + 1], a1);
      a2 = fmaf(__ldg(w + (size_t)(k + 2) * 3 * GRU_H + j), h[k + 2], a2);
      a3 = fmaf(__ldg(w + (size_t)(k + 3) * 3 * GRU_H + j), h[k + 3], a3);
    }
    gh[j] = (a0 + a1) + (a2 + a3) + bj;
    __syncthreads();
    if (j < GRU_H) {
      const float* g = gi + ((size_t)b * T + t) * 6 * GRU_H + dir * 3 * GRU_H;
      const float r = 1.f / (1.f + expf(-(g[j] + gh[j])));
      const float z = 1.f / (1.f + expf(-(g[GRU_H + j] + gh[GRU_H + j])));
      const float n = tanhf(g[2 * GRU_H + j] + r * gh[2 * GRU_H + j]);
      const float hn = (1.f - z) * n + z * h[j];
      out[((size_t)b * T + t) * 2 * GRU_H + dir * GRU_H + j] = hn;
      h[j] = hn;  // only this thread reads h[j] between the two barriers of a step
    }
    __syncthreads();
  }
}
cudaError_t launch_gru_simple(const float* gi, const float* whhT, const float* bhh, float* out, int B, int T, cudaStream_t s) {
  gru_kernel<<<dim3(2, B), 3 * GRU_H, 0, s>>>(gi, whhT, bhh, out, B, T);
  return cudaGetLastError();
}

// The same recurrence on a CLUSTER of 8 CTAs per (direction, item) with the recurrent weights resident in REGISTERS:
// CTA c owns hidden units [32 c, 32 c + 32), i.e. 96 of the 768 gate rows; its 384 threads hold 64 weights each (thread =
// one row x one quarter of k; a warp = 32 rows of one gate and one quarter, so the hidden vector is read with broadcast
// 16-byte loads: 16 per thread and step).  A first version kept the weights in shared memory (4 CTAs x 192 KB) and was
// bound by streaming them through the LSU every step (3072 wavefronts, 2.3 us per step, profiles/r2av_*).  Per step:
// 64 FMAs per thread, the four quarter sums meet in shared memory, the CTA's first warp applies the gates and writes the 32
// new h values into the h buffers of all eight CTAs with st.async: the store carries its own completion (complete_tx on an
// mbarrier of the receiving CTA, which expects 256 x 4 bytes per step), so a step needs no cluster barrier - every thread
// waits on its CTA's local mbarrier for the next h (h is double-buffered; a CTA sends h(s+1) only after it received all
// of h(s), i.e. after every CTA finished reading the buffer that h(s+1) overwrites).
constexpr int GRU_CL = 8, GRU_UNITS = GRU_H / GRU_CL, GRU_ROWS = 3 * GRU_UNITS, GRU_Q = 4, GRU_KQ = GRU_H / GRU_Q, GRU_CT = GRU_Q * GRU_ROWS;
static_assert(GRU_UNITS == 32 && GRU_CT == 384, "one warp per (gate, quarter)");
__global__ void __cluster_dims__(GRU_CL, 1, 1) __launch_bounds__(GRU_CT, 1)
gru_cluster_kernel(const float* __restrict__ gi, const float* __restrict__ whhT, const float* __restrict__ bhh, float* __restrict__ out, int B, int T) {
  __shared__ __align__(16) float hbuf[2][GRU_H];
  __shared__ float part[GRU_Q][GRU_ROWS];
  __shared__ __align__(8) unsigned long long hbar[2];
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int dir = blockIdx.y, b = blockIdx.z, t_ = threadIdx.x;
  const int warp = t_ >> 5, lane = t_ & 31;
  const int gate = warp % 3, q = warp / 3, row = gate * GRU_UNITS + lane;      // local row
  const int grow = gate * GRU_H + GRU_UNITS * (int)rank + lane;                 // row of W_hh
  float wreg[GRU_KQ];
  {
    const float* wg = whhT + ((size_t)dir * GRU_H + q * GRU_KQ) * 3 * GRU_H + grow;
#pragma unroll
    for (int j = 0; j < GRU_KQ; ++j) wreg[j] = __ldg(wg + (size_t)j * 3 * GRU_H);
  }
  for (int i = t_; i < 2 * GRU_H; i += GRU_CT) (&hbuf[0][0])[i] = 0.f;
  const uint32_t bar_local = (uint32_t)__cvta_generic_to_shared(&hbar[0]);
  if (t_ == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_local) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_local + 8) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const int ug = GRU_UNITS * (int)rank + lane;  // gate threads: warp 0
  float br = 0.f, bz = 0.f, bn = 0.f;
  if (warp == 0) {
    const float* bb = bhh + dir * 3 * GRU_H;
    br = bb[ug]; bz = bb[GRU_H + ug]; bn = bb[2 * GRU_H + ug];
  }
  uint32_t hdst[GRU_CL], bdst[GRU_CL];  // this CTA's h buffer / mbarriers as seen in each CTA of the cluster
  {
    const uint32_t local = (uint32_t)__cvta_generic_to_shared(&hbuf[0][0]);
#pragma unroll
    for (int c = 0; c < GRU_CL; ++c) {
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(hdst[c]) : "r"(local), "r"(c));
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(bdst[c]) : "r"(bar_local), "r"(c));
    }
  }
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  float g_r = 0.f, g_z = 0.f, g_n = 0.f;
  auto load_gi = [&](int step) {
    const int t = dir ? T - 1 - step : step;
    const float* g = gi + ((size_t)b * T + t) * 6 * GRU_H + dir * 3 * GRU_H;
    g_r = __ldg(g + ug); g_z = __ldg(g + GRU_H + ug); g_n = __ldg(g + 2 * GRU_H + ug);
  };
  if (warp == 0 && T > 0) load_gi(0);
  for (int step = 0; step < T; ++step) {
    const int cur = step & 1, nxt = cur ^ 1;
    const bool more = step + 1 < T;
    // arm the barrier that collects h(step + 1): one local arrival + 256 floats from the eight CTAs' gate warps
    if (t_ == 0 && more)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_local + 8 * nxt), "r"(GRU_H * 4) : "memory");
    const float* h = hbuf[cur];
    const float4* hp = reinterpret_cast<const float4*>(h + q * GRU_KQ);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int k = 0; k < GRU_KQ / 4; ++k) {
      const float4 hv = hp[k];
      a0 = fmaf(wreg[4 * k + 0], hv.x, a0);
      a1 = fmaf(wreg[4 * k + 1], hv.y, a1);
      a2 = fmaf(wreg[4 * k + 2], hv.z, a2);
      a3 = fmaf(wreg[4 * k + 3], hv.w, a3);
    }
    part[q][row] = (a0 + a1) + (a2 + a3);
    __syncthreads();
    if (warp == 0) {
      const float cr = g_r, cz = g_z, cn = g_n;
      if (more) load_gi(step + 1);  // in flight while the others' h values arrive
      const float hr = ((part[0][lane] + part[1][lane]) + (part[2][lane] + part[3][lane])) + br;
      const float hz = ((part[0][GRU_UNITS + lane] + part[1][GRU_UNITS + lane]) + (part[2][GRU_UNITS + lane] + part[3][GRU_UNITS + lane])) + bz;
      const float hn = ((part[0][2 * GRU_UNITS + lane] + part[1][2 * GRU_UNITS + lane]) +
                        (part[2][2 * GRU_UNITS + lane] + part[3][2 * GRU_UNITS + lane])) + bn;
      const float r = 1.f / (1.f + expf(-(cr + hr)));
      const float z = 1.f / (1.f + expf(-(cz + hz)));
      const float n = tanhf(cn + r * hn);
      const float hnew = (1.f - z) * n + z * h[ug];
      const int t = dir ? T - 1 - step : step;
      out[((size_t)b * T + t) * 2 * GRU_H + dir * GRU_H + ug] = hnew;
      if (more) {
        const uint32_t off = (uint32_t)((nxt * GRU_H + ug) * 4);
#pragma unroll
        for (int c = 0; c < GRU_CL; ++c)
          asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(hdst[c] + off),
                       "r"(__float_as_uint(hnew)), "r"(bdst[c] + 8 * nxt)
                       : "memory");
      }
    }
    if (more) {  // h(step + 1) complete in this CTA's buffer: the (step / 2)-th use of barrier nxt
      const uint32_t parity = (uint32_t)((step >> 1) & 1);
      uint32_t done = 0;
      for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar_local + 8 * nxt), "r"(parity)
            : "memory");
        if (spin > (1u << 26)) __trap();  // a protocol bug must not hang the GPU
      }
    }
  }
  // no CTA may exit while others still send into its shared memory
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

cudaError_t launch_gru(const float* gi, const float* whhT, const float* bhh, float* out, int B, int T, cudaStream_t s) {
  gru_cluster_kernel<<<dim3(GRU_CL, 2, B), GRU_CT, 0, s>>>(gi, whhT, bhh, out, B, T);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// Mel spectrogram pieces (RMVPE.py:388-417).  reflect_pad: torch.stft(center=True) pads n_fft / 2 samples per side by
// reflection; the DFT itself is launch_sgemm_nt with the windowed cos / sin table; mel_log: |X| -> basis -> log(clamp).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void reflect_pad_kernel(const float* __restrict__ audio, float* __restrict__ padded, int B, int n, int pad) {
  const size_t np = (size_t)n + 2 * pad, total = (size_t)B * np;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / np);
    long long j = (long long)(i - (size_t)b * np) - pad;
    if (j < 0) j = -j;
    if (j >= n) j = 2LL * (n - 1) - j;
    padded[i] = audio[(size_t)b * n + j];
  }
}
cudaError_t launch_reflect_pad(const float* audio, float* padded, int B, int n, int pad, cudaStream_t s) {
  const size_t total = (size_t)B * ((size_t)n + 2 * pad);
  reflect_pad_kernel<<<(int)std::min<size_t>((total + 255) / 256, 148 * 16), 256, 0, s>>>(audio, padded, B, n, pad);
  return cudaGetLastError();
}

// spec (B, frames, 2 * NB) = [re 0..NB-1 | im 0..NB-1] -> mel (B, 128, frames); one CTA per (frame, item)
__global__ void __launch_bounds__(128) mel_log_kernel(const float* __restrict__ spec, const float* __restrict__ basis, float* __restrict__ mel,
                                                      int frames, int NB, float clampv) {
  extern __shared__ float mag[];
  const int f = blockIdx.x, b = blockIdx.y;
  const float* sp = spec + ((size_t)b * frames + f) * 2 * NB;
  for (int k = threadIdx.x; k < NB; k += 128) mag[k] = sqrtf(sp[k] * sp[k] + sp[NB + k] * sp[NB + k]);
  __syncthreads();
  const float* bw = basis + (size_t)threadIdx.x * NB;
  float acc = 0.f;
  for (int k = 0; k < NB; ++k) acc = fmaf(__ldg(bw + k), mag[k], acc);
  mel[((size_t)b * 128 + threadIdx.x) * frames + f] = logf(fmaxf(acc, clampv));
}
cudaError_t launch_mel_log(const float* spec, const float* basis, float* mel, int B, int frames, int NB, float clampv, cudaStream_t s) {
  mel_log_kernel<<<dim3(frames, B), 128, NB * sizeof(float), s>>>(spec, basis, mel, frames, NB, clampv);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// decode (RMVPE.py:484-541): per frame argmax over the 360 bins, salience-weighted mean of the cents of the 9 bins around
// it, 0 when the maximum is <= thred; f0 = 10 * 2^(cents / 1200) with 10 -> 0.  One warp per frame; fp64 for the mean as
// the reference computes it in numpy float64 from the float32 salience.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) rmvpe_decode_kernel(const float* __restrict__ sal, float thred, float* __restrict__ f0, int T) {
  const int t = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (t >= T) return;
  const float* s = sal + (size_t)t * 360;
  float best = -1.f;
  int bi = 0;
  for (int k = lane; k < 360; k += 32) {
    const float v = s[k];
    if (v > best) { best = v; bi = k; }   // first maximum within this lane's strided bins
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }  // np.argmax: the first maximum
  }
  if (lane == 0) {
    double ps = 0.0, ws = 0.0;
    for (int d = -4; d <= 4; ++d) {
      const int k = bi + d;
      if (k < 0 || k >= 360) continue;  // the zero padding of salience and cents_mapping
      const double v = (double)s[k];
      ps += v * (20.0 * k + 1997.3794084376191);
      ws += v;
    }
    double cents = ps / ws;
    if (!(best > thred)) cents = 0.0;
    double f = 10.0 * exp2(cents / 1200.0);
    if (f == 10.0) f = 0.0;
    f0[t] = (float)f;
  }
}
cudaError_t launch_rmvpe_decode(const float* sal, float thred, float* f0, int T, cudaStream_t s) {
  rmvpe_decode_kernel<<<(T + 3) / 4, 128, 0, s>>>(sal, thred, f0, T);
  return cudaGetLastError();
}

}  // namespace rvcdec
