// Small HBM-bound kernels around the convolution stack: harmonic source, speaker
// conditioning bias, input packing, pad zeroing, noise_conv add, conv_post + tanh.
#include <algorithm>
#include <cstdio>
#include "common.cuh"
#include "philox.cuh"

namespace rvcdec {

// ---------------------------------------------------------------------------
// Harmonic source: SineGenerator.forward (hifigan.py:197-228) + SourceModuleHnNSF
// merge (hifigan_nsf.py:48-52) in one kernel.
//   inc[l][k]  = f0[l]/sr * (k+1)
//   rem[l]     = fmod(inc[l][upp-1] + 0.5, 1) - 0.5          (l < T-1)
//   C[l]       = sum_{j<l} rem[j]                            (frame-level scan, fp64)
//   sine       = 0.1 * sin(2*pi*(inc[l][k] + C[l]))
//   out        = uv ? sine + 0.003*z : (0.1/3)*z ;  har = tanh(w*out + b)
// Each CTA owns FPB consecutive frames of one item.  The scan is only T (<= a few
// thousand) elements long, so every CTA re-reduces the prefix of its own span in fp64
// instead of running a multi-pass device scan: no inter-CTA dependency, one launch.
// ---------------------------------------------------------------------------
constexpr int SRC_FPB = 32;
constexpr int SRC_THREADS = 256;

__global__ void __launch_bounds__(SRC_THREADS) source_kernel(SourceParams p) {
  __shared__ double red[SRC_THREADS / 32];
  __shared__ double s_prefix;
  __shared__ double s_C[SRC_FPB];
  __shared__ double s_finc[SRC_FPB];
  __shared__ float s_f0[SRC_FPB];
  const int b = blockIdx.y;
  const int l0 = blockIdx.x * SRC_FPB;
  const uint64_t seed = p.seed_dev ? __ldg(p.seed_dev) : p.seed;
  const float* f0 = p.f0 + (size_t)b * p.T;
  const double inv_sr = 1.0 / (double)p.sr;
  const double dupp = (double)p.upp;

  // (1) prefix of the wrapped per-frame phase advance over frames [0, l0)
  double part = 0.0;
  for (int j = threadIdx.x; j < l0; j += SRC_THREADS) {
    const double adv = (double)f0[j] * inv_sr * dupp + 0.5;
    part += (adv - floor(adv)) - 0.5;  // f0 >= 0 so fmod(adv, 1) == adv - floor(adv)
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < SRC_THREADS / 32; ++i) t += red[i];
    s_prefix = t;
  }
  __syncthreads();
  // (2) scan inside the span (<= 32 frames: one thread per frame, serial partial sums)
  if (threadIdx.x < SRC_FPB) {
    const int l = l0 + threadIdx.x;
    double c = s_prefix;
    for (int j = l0; j < l && j < p.T; ++j) {
      const double adv = (double)f0[j] * inv_sr * dupp + 0.5;
      c += (adv - floor(adv)) - 0.5;
    }
    s_C[threadIdx.x] = c - floor(c);  // phase only matters modulo 1
    const float f = (l < p.T) ? f0[l] : 0.f;
    s_f0[threadIdx.x] = f;
    s_finc[threadIdx.x] = (double)f * inv_sr;
  }
  __syncthreads();
  // (3) samples
  const int nframes = min(SRC_FPB, p.T - l0);
  const int nsamp = nframes * p.upp;
  const size_t base = (size_t)b * p.T * p.upp + (size_t)l0 * p.upp;
  // Philox counter of sample 0 of this span: indexed by the GLOBAL item (item0 + b), not by the position in the launch
  const uint64_t pbase = (p.item0 + (uint64_t)b) * (uint64_t)p.T * (uint64_t)p.upp + (uint64_t)l0 * (uint64_t)p.upp;
  for (int i = threadIdx.x; i < nsamp; i += SRC_THREADS) {
    const int fr = i / p.upp;
    const int k = i - fr * p.upp;
    const double ph = s_finc[fr] * (double)(k + 1) + s_C[fr];  // fundamental phase in cycles (closed form inside a frame)
    const bool voiced = s_f0[fr] > 0.f;
    if (p.H <= 1) {
      const float frac = (float)(ph - floor(ph));
      const float sine = 0.1f * sinpif(2.0f * frac);
      const float z = p.noise ? __ldg(p.noise + base + i) : philox_normal(seed, pbase + (uint64_t)i);
      const float v = voiced ? (sine + 0.003f * z) : ((0.1f / 3.0f) * z);
      p.har[base + i] = tanhf(p.lin_w * v + p.lin_b);
    } else {
      // overtone k2 runs at (k2 + 1) times the fundamental's phase plus its own initial phase (hifigan_mrf.py:124-142:
      // per-sample cumsum of (f0 (k2+1) / sr) % 1 with rand_ini added to sample 0)
      float acc = p.lin_b;
      for (int k2 = 0; k2 < p.H; ++k2) {
        float ini = 0.f;
        if (k2 > 0) {
          if (p.rand_ini) ini = __ldg(p.rand_ini + (size_t)b * p.H + k2);
          else {
            const uint64_t ci = (1ull << 62) + (p.item0 + (uint64_t)b) * (uint64_t)p.H + (uint64_t)k2;
            ini = (float)philox4x32_10((uint32_t)ci, (uint32_t)(ci >> 32), (uint32_t)seed, (uint32_t)(seed >> 32)).x * 2.3283064365386963e-10f;
          }
        }
        const double phk = ph * (double)(k2 + 1) + (double)ini;
        const float frac = (float)(phk - floor(phk));
        const float sine = 0.1f * sinpif(2.0f * frac);
        const uint64_t ni = (uint64_t)(base + i) * (uint64_t)p.H + (uint64_t)k2;
        const float z = p.noise ? __ldg(p.noise + ni) : philox_normal(seed, (pbase + (uint64_t)i) * (uint64_t)p.H + (uint64_t)k2);
        const float v = voiced ? (sine + 0.003f * z) : ((0.1f / 3.0f) * z);
        acc = fmaf(p.lin_wv[k2], v, acc);
      }
      p.har[base + i] = tanhf(acc);
    }
  }
}

__global__ void set_u64_kernel(uint64_t* dst, uint64_t v) { *dst = v; }
cudaError_t launch_set_u64(uint64_t* dst, uint64_t v, cudaStream_t s) {
  set_u64_kernel<<<1, 1, 0, s>>>(dst, v);
  return cudaGetLastError();
}

cudaError_t launch_source(const SourceParams& p, cudaStream_t s) {
  dim3 grid((p.T + SRC_FPB - 1) / SRC_FPB, p.B);
  source_kernel<<<grid, SRC_THREADS, 0, s>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// cond(g) folded into a per-item bias for conv_pre (hifigan_nsf.py:179-182).
// ---------------------------------------------------------------------------
__global__ void cond_kernel(const float* __restrict__ g, const float* __restrict__ Wc, const float* __restrict__ bc,
                            const float* __restrict__ bpre, float* __restrict__ biasB, int B, int gin, int N) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (n >= N) return;
  float acc = bpre[n];
  if (g) {
    acc += bc[n];
    const float* w = Wc + (size_t)n * gin;
    const float* gb = g + (size_t)b * gin;
    for (int c = 0; c < gin; ++c) acc = fmaf(w[c], gb[c], acc);
  }
  biasB[(size_t)b * N + n] = acc;
}

cudaError_t launch_cond(const float* g, const float* Wc, const float* bias_cond, const float* bias_pre, float* biasB,
                        int B, int gin, int N, cudaStream_t s) {
  dim3 grid((N + 127) / 128, B);
  cond_kernel<<<grid, 128, 0, s>>>(g, Wc, bias_cond, bias_pre, biasB, B, gin, N);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T cvt_act(float v);
template <> __device__ __forceinline__ float cvt_act<float>(float v) { return v; }
template <> __device__ __forceinline__ __half cvt_act<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 cvt_act<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <typename T>
__global__ void pack_input_kernel(const float* __restrict__ x, T* __restrict__ act, int B, int C, int T_, int Lp) {
  const size_t total = (size_t)B * (C / 8) * T_;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % T_);
    const size_t bs = i / T_;
    const int sl = (int)(bs % (C / 8));
    const int b = (int)(bs / (C / 8));
    const float* src = x + ((size_t)b * C + sl * 8) * T_ + t;
    T* dst = act + ((bs * Lp) + PADL + t) * 8;
#pragma unroll
    for (int c = 0; c < 8; ++c) dst[c] = cvt_act<T>(src[(size_t)c * T_]);
  }
}

cudaError_t launch_pack_input(const float* x, void* act, int fmt, int B, int C, int T, int Lp, cudaStream_t s) {
  const size_t total = (size_t)B * (C / 8) * T;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 16));
  if (fmt == ACT_F32) pack_input_kernel<float><<<blocks, 256, 0, s>>>(x, (float*)act, B, C, T, Lp);
  else if (fmt == ACT_F16) pack_input_kernel<__half><<<blocks, 256, 0, s>>>(x, (__half*)act, B, C, T, Lp);
  else pack_input_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>(x, (__nv_bfloat16*)act, B, C, T, Lp);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
__global__ void zero_pads_kernel(uint4* a0, uint4* a1, uint4* a2, int chunks_per_row, int L, int Lp) {
  // one block per (b*slice, tensor); pad rows = [0,PADL) U [PADL+L, Lp)
  uint4* base = (blockIdx.y == 0 ? a0 : (blockIdx.y == 1 ? a1 : a2));
  if (!base) return;
  const int right = Lp - PADL - L;
  const int total = (PADL + right) * chunks_per_row;
  uint4* sl = base + (size_t)blockIdx.x * Lp * chunks_per_row;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int r = i / chunks_per_row;
    const int c = i - r * chunks_per_row;
    const int row = (r < PADL) ? r : (L + r);  // r >= PADL: PADL + L + (r - PADL)
    sl[(size_t)row * chunks_per_row + c] = make_uint4(0, 0, 0, 0);
  }
}

cudaError_t launch_zero_pads(void* a0, void* a1, void* a2, int elem_bytes, int B, int C, int L, int Lp,
                             cudaStream_t s) {
  const int chunks = elem_bytes / 2;  // 16-byte chunks per 8-channel row
  dim3 grid(B * (C / 8), 3);
  zero_pads_kernel<<<grid, 256, 0, s>>>((uint4*)a0, (uint4*)a1, (uint4*)a2, chunks, L, Lp);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// x = ups(x) [already in p.x] + noise_convs[i](har)  (hifigan_nsf.py:199), then the
// activated copy lrelu(x, 0.1) every ResBlock's first conv consumes (residuals.py:74).
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) noise_add_kernel(NoiseAddParams p) {
  const int nsl = p.C / 8;
  const size_t total = (size_t)p.B * nsl * p.L;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int row = (int)(i % p.L);
    const size_t bs = i / p.L;
    const int sl = (int)(bs % nsl);
    const int b = (int)(bs / nsl);
    const size_t off = (bs * p.Lp + PADL + row) * 8;
    float v[8];
    {
      const float4 q0 = *reinterpret_cast<const float4*>(p.x + off);
      const float4 q1 = *reinterpret_cast<const float4*>(p.x + off + 4);
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + sl * 8));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + sl * 8 + 4));
      v[0] = q0.x + b0.x; v[1] = q0.y + b0.y; v[2] = q0.z + b0.z; v[3] = q0.w + b0.w;
      v[4] = q1.x + b1.x; v[5] = q1.y + b1.y; v[6] = q1.z + b1.z; v[7] = q1.w + b1.w;
    }
    const float* hb = p.har + (size_t)b * p.n_src;
    const int s0 = row * p.stride - p.pad;
    for (int j = 0; j < p.k; ++j) {
      const int sidx = s0 + j;
      if (sidx < 0 || sidx >= p.n_src) continue;
      const float h = __ldg(hb + sidx);
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(p.w + (size_t)j * p.C + sl * 8));
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(p.w + (size_t)j * p.C + sl * 8 + 4));
      v[0] = fmaf(h, w0.x, v[0]); v[1] = fmaf(h, w0.y, v[1]); v[2] = fmaf(h, w0.z, v[2]); v[3] = fmaf(h, w0.w, v[3]);
      v[4] = fmaf(h, w1.x, v[4]); v[5] = fmaf(h, w1.y, v[5]); v[6] = fmaf(h, w1.z, v[6]); v[7] = fmaf(h, w1.w, v[7]);
    }
    *reinterpret_cast<float4*>(p.x + off) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p.x + off + 4) = make_float4(v[4], v[5], v[6], v[7]);
    T* a = reinterpret_cast<T*>(p.act) + off;
    if (sizeof(T) == 4) {
      *reinterpret_cast<float4*>(a) = make_float4(lrelu(v[0], p.slope), lrelu(v[1], p.slope), lrelu(v[2], p.slope), lrelu(v[3], p.slope));
      *reinterpret_cast<float4*>(a + 4) = make_float4(lrelu(v[4], p.slope), lrelu(v[5], p.slope), lrelu(v[6], p.slope), lrelu(v[7], p.slope));
    } else {
      T tmp[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) tmp[c] = cvt_act<T>(lrelu(v[c], p.slope));
      *reinterpret_cast<uint4*>(a) = *reinterpret_cast<uint4*>(tmp);
    }
  }
}

// Long noise convs (stage 0: 80 taps, stride 40, 256 channels): weights and the har span of a row tile
// are staged in shared memory and every thread keeps 8 rows x 8 channels in registers, so one weight
// fetch feeds 8 rows (the generic kernel above re-read the weights from L1 for every row: 30 ms at the
// bench shape; this one is FMA-bound at ~3 ms).
constexpr int NZB_ROWS_PER_THREAD = 8;
template <typename T>
__global__ void __launch_bounds__(256) noise_add_big_kernel(NoiseAddParams p, int tiles_per_item, int rows_per_tile) {
  extern __shared__ float nz_smem[];
  float* sw = nz_smem;                 // [k][C]
  float* sh = nz_smem + p.k * p.C;     // har span of the tile
  const int nsl = p.C / 8;
  const int rgroups = 256 / nsl;       // row groups per block
  const int sl = threadIdx.x % nsl;
  const int rg = threadIdx.x / nsl;
  for (int i = threadIdx.x; i < p.k * p.C; i += 256) sw[i] = p.w[i];
  const int span = (rows_per_tile - 1) * p.stride + p.k;
  const int total = p.B * tiles_per_item;
  for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
    const int b = tile / tiles_per_item;
    const int row0 = (tile - b * tiles_per_item) * rows_per_tile;
    const float* hb = p.har + (size_t)b * p.n_src;
    const int s_base = row0 * p.stride - p.pad;
    __syncthreads();  // previous tile's readers done (also covers the weight staging on the first pass)
    for (int i = threadIdx.x; i < span; i += 256) {
      const int sidx = s_base + i;
      sh[i] = (sidx >= 0 && sidx < p.n_src) ? __ldg(hb + sidx) : 0.f;
    }
    __syncthreads();
    if (rg < rgroups) {
      float acc[NZB_ROWS_PER_THREAD][8];
#pragma unroll
      for (int r = 0; r < NZB_ROWS_PER_THREAD; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[r][c] = 0.f;
      const float* hrow = sh + rg * NZB_ROWS_PER_THREAD * p.stride;
      for (int j = 0; j < p.k; ++j) {
        const float4 w0 = *reinterpret_cast<const float4*>(sw + j * p.C + sl * 8);
        const float4 w1 = *reinterpret_cast<const float4*>(sw + j * p.C + sl * 8 + 4);
#pragma unroll
        for (int r = 0; r < NZB_ROWS_PER_THREAD; ++r) {
          const float hv = hrow[r * p.stride + j];
          acc[r][0] = fmaf(hv, w0.x, acc[r][0]); acc[r][1] = fmaf(hv, w0.y, acc[r][1]);
          acc[r][2] = fmaf(hv, w0.z, acc[r][2]); acc[r][3] = fmaf(hv, w0.w, acc[r][3]);
          acc[r][4] = fmaf(hv, w1.x, acc[r][4]); acc[r][5] = fmaf(hv, w1.y, acc[r][5]);
          acc[r][6] = fmaf(hv, w1.z, acc[r][6]); acc[r][7] = fmaf(hv, w1.w, acc[r][7]);
        }
      }
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + sl * 8));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + sl * 8 + 4));
#pragma unroll
      for (int r = 0; r < NZB_ROWS_PER_THREAD; ++r) {
        const int row = row0 + rg * NZB_ROWS_PER_THREAD + r;
        if (row >= p.L) break;
        const size_t off = (((size_t)b * nsl + sl) * p.Lp + PADL + row) * 8;
        const float4 q0 = *reinterpret_cast<const float4*>(p.x + off);
        const float4 q1 = *reinterpret_cast<const float4*>(p.x + off + 4);
        float v[8] = {q0.x + b0.x + acc[r][0], q0.y + b0.y + acc[r][1], q0.z + b0.z + acc[r][2], q0.w + b0.w + acc[r][3],
                      q1.x + b1.x + acc[r][4], q1.y + b1.y + acc[r][5], q1.z + b1.z + acc[r][6], q1.w + b1.w + acc[r][7]};
        *reinterpret_cast<float4*>(p.x + off) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(p.x + off + 4) = make_float4(v[4], v[5], v[6], v[7]);
        T* a = reinterpret_cast<T*>(p.act) + off;
        if (sizeof(T) == 4) {
          *reinterpret_cast<float4*>(a) = make_float4(lrelu(v[0], p.slope), lrelu(v[1], p.slope), lrelu(v[2], p.slope), lrelu(v[3], p.slope));
          *reinterpret_cast<float4*>(a + 4) = make_float4(lrelu(v[4], p.slope), lrelu(v[5], p.slope), lrelu(v[6], p.slope), lrelu(v[7], p.slope));
        } else {
          T tmp[8];
#pragma unroll
          for (int c = 0; c < 8; ++c) tmp[c] = cvt_act<T>(lrelu(v[c], p.slope));
          *reinterpret_cast<uint4*>(a) = *reinterpret_cast<uint4*>(tmp);
        }
      }
    }
  }
}

template <typename T>
static cudaError_t launch_noise_add_big(const NoiseAddParams& p, cudaStream_t s) {
  const int nsl = p.C / 8;
  const int rows_per_tile = (256 / nsl) * NZB_ROWS_PER_THREAD;
  const int span = (rows_per_tile - 1) * p.stride + p.k;
  const size_t smem = ((size_t)p.k * p.C + span) * sizeof(float);
  auto kern = noise_add_big_kernel<T>;
  // (the size depends on the layer, so the attribute is raised to the largest request seen per device)
  static int granted[64] = {0};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 64 || granted[dev] < (int)smem) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (dev < 64) granted[dev] = (int)smem;
  }
  const int tiles_per_item = (p.L + rows_per_tile - 1) / rows_per_tile;
  const long long total = (long long)p.B * tiles_per_item;
  const int blocks = (int)std::min<long long>(total, 148 * 2);
  kern<<<blocks, 256, smem, s>>>(p, tiles_per_item, rows_per_tile);
  return cudaGetLastError();
}

cudaError_t launch_noise_add(const NoiseAddParams& p, cudaStream_t s) {
  // long kernels: smem-staged register-tiled variant (needs C/8 to divide 256 and the image to fit)
  const int nsl = p.C / 8;
  if (p.k >= 16 && nsl <= 256 && 256 % nsl == 0 &&
      ((size_t)p.k * p.C + (size_t)((256 / nsl) * NZB_ROWS_PER_THREAD - 1) * p.stride + p.k) * 4 <= 200 * 1024) {
    if (p.act_fmt == ACT_F32) return launch_noise_add_big<float>(p, s);
    if (p.act_fmt == ACT_F16) return launch_noise_add_big<__half>(p, s);
    return launch_noise_add_big<__nv_bfloat16>(p, s);
  }
  const size_t total = (size_t)p.B * (p.C / 8) * p.L;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 32));
  if (p.act_fmt == ACT_F32) noise_add_kernel<float><<<blocks, 256, 0, s>>>(p);
  else if (p.act_fmt == ACT_F16) noise_add_kernel<__half><<<blocks, 256, 0, s>>>(p);
  else noise_add_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Tail: leaky_relu(x) with the DEFAULT slope 0.01, conv_post (C->1, k=7, no bias), tanh
// (hifigan_nsf.py:209-210).  One thread per output sample, fp32 throughout.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) post_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                   float* __restrict__ out, int B, int C, int L, int Lp, int k,
                                                   float slope, float bias) {
  extern __shared__ float sw[];  // [k][C]
  for (int i = threadIdx.x; i < k * C; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  const int nsl = C / 8;
  const int pad = (k - 1) / 2;
  const size_t total = (size_t)B * L;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % L);
    const int b = (int)(i / L);
    float acc = 0.f;
    for (int sl = 0; sl < nsl; ++sl) {
      const float* xs = x + (((size_t)b * nsl + sl) * Lp + PADL) * 8;
      for (int j = 0; j < k; ++j) {
        const int tt = t + j - pad;
        if (tt < 0 || tt >= L) continue;
        const float4 q0 = *reinterpret_cast<const float4*>(xs + (size_t)tt * 8);
        const float4 q1 = *reinterpret_cast<const float4*>(xs + (size_t)tt * 8 + 4);
        const float* wj = sw + j * C + sl * 8;
        acc = fmaf(lrelu(q0.x, slope), wj[0], acc); acc = fmaf(lrelu(q0.y, slope), wj[1], acc);
        acc = fmaf(lrelu(q0.z, slope), wj[2], acc); acc = fmaf(lrelu(q0.w, slope), wj[3], acc);
        acc = fmaf(lrelu(q1.x, slope), wj[4], acc); acc = fmaf(lrelu(q1.y, slope), wj[5], acc);
        acc = fmaf(lrelu(q1.z, slope), wj[6], acc); acc = fmaf(lrelu(q1.w, slope), wj[7], acc);
      }
    }
    out[i] = tanhf(acc + bias);
  }
}

// k = 7 specialisation: every thread produces 4 consecutive samples from a 10-row window, so each
// input row is fetched 2.5x instead of 7x per output (the generic kernel above is L1-bandwidth bound).
__global__ void __launch_bounds__(256) post7_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                    float* __restrict__ out, int B, int C, int L, int Lp, float slope, float bias) {
  extern __shared__ float sw[];  // [7][C]
  for (int i = threadIdx.x; i < 7 * C; i += blockDim.x) sw[i] = w[i];
  __syncthreads();
  const int nsl = C / 8;
  const int groups_per_item = (L + 3) / 4;
  const size_t total = (size_t)B * groups_per_item;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int gq = (int)(i % groups_per_item);
    const int b = (int)(i / groups_per_item);
    const int t0 = gq * 4;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int sl = 0; sl < nsl; ++sl) {
      const float* xs = x + (((size_t)b * nsl + sl) * Lp + PADL) * 8;
      const float* ws = sw + sl * 8;
#pragma unroll
      for (int r = 0; r < 10; ++r) {
        const int tt = t0 + r - 3;
        if (tt < 0 || tt >= L) continue;  // zero padding of conv_post
        const float4 q0 = *reinterpret_cast<const float4*>(xs + (size_t)tt * 8);
        const float4 q1 = *reinterpret_cast<const float4*>(xs + (size_t)tt * 8 + 4);
        const float v[8] = {lrelu(q0.x, slope), lrelu(q0.y, slope), lrelu(q0.z, slope), lrelu(q0.w, slope),
                            lrelu(q1.x, slope), lrelu(q1.y, slope), lrelu(q1.z, slope), lrelu(q1.w, slope)};
#pragma unroll
        for (int o = 0; o < 4; ++o) {
          const int j = r - o;  // tap index of row r for output t0 + o
          if (j < 0 || j >= 7) continue;
          const float4 w0 = *reinterpret_cast<const float4*>(ws + j * C);
          const float4 w1 = *reinterpret_cast<const float4*>(ws + j * C + 4);
          acc[o] = fmaf(v[0], w0.x, acc[o]); acc[o] = fmaf(v[1], w0.y, acc[o]); acc[o] = fmaf(v[2], w0.z, acc[o]);
          acc[o] = fmaf(v[3], w0.w, acc[o]); acc[o] = fmaf(v[4], w1.x, acc[o]); acc[o] = fmaf(v[5], w1.y, acc[o]);
          acc[o] = fmaf(v[6], w1.z, acc[o]); acc[o] = fmaf(v[7], w1.w, acc[o]);
        }
      }
    }
    float* o = out + (size_t)b * L + t0;
    if (t0 + 3 < L && ((reinterpret_cast<uintptr_t>(o) & 15u) == 0)) {
      *reinterpret_cast<float4*>(o) = make_float4(tanhf(acc[0] + bias), tanhf(acc[1] + bias), tanhf(acc[2] + bias), tanhf(acc[3] + bias));
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (t0 + q < L) o[q] = tanhf(acc[q] + bias);
    }
  }
}

cudaError_t launch_post(const float* x, const float* w, float* out, int B, int C, int L, int Lp, int k, float slope,
                        cudaStream_t s, float bias) {
  if (k == 7) {
    const size_t total = (size_t)B * ((L + 3) / 4);
    const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 32));
    post7_kernel<<<blocks, 256, (size_t)7 * C * sizeof(float), s>>>(x, w, out, B, C, L, Lp, slope, bias);
    return cudaGetLastError();
  }
  const size_t total = (size_t)B * L;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 32));
  post_kernel<<<blocks, 256, (size_t)k * C * sizeof(float), s>>>(x, w, out, B, C, L, Lp, k, slope, bias);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
template <typename T>
__global__ void unpack_kernel(const T* __restrict__ src, float* __restrict__ dst, int B, int C, int L, int Lp, int rs) {
  const size_t total = (size_t)B * C * L;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % L);
    const size_t bc = i / L;
    const int c = (int)(bc % C);
    const int b = (int)(bc / C);
    dst[i] = (float)src[(((size_t)b * (C / 8) + (c >> 3)) * Lp + PADL + (size_t)t * rs) * 8 + (c & 7)];
  }
}

cudaError_t launch_unpack(const void* src, int fmt, float* dst, int B, int C, int L, int Lp, cudaStream_t s, int rs) {
  const size_t total = (size_t)B * C * L;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 32));
  if (fmt == ACT_F32) unpack_kernel<float><<<blocks, 256, 0, s>>>((const float*)src, dst, B, C, L, Lp, rs);
  else if (fmt == ACT_F16) unpack_kernel<__half><<<blocks, 256, 0, s>>>((const __half*)src, dst, B, C, L, Lp, rs);
  else unpack_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>((const __nv_bfloat16*)src, dst, B, C, L, Lp, rs);
  return cudaGetLastError();
}

}  // namespace rvcdec
