// Kernels of the sibling decoder RefineGANGenerator (rvc/lib/algorithm/generators/refinegan.py:268-451) that the NSF
// decoder does not have: the harmonic source on a LINEARLY interpolated f0 (a real per-sample phase integral), the
// sinc / Kaiser decimation of the downsample path, linear upsampling + skip concatenation, and the AdaIN noise layers.
// fp32, blocked-8 time-major activations; every convolution of the network runs on conv_f32.cu.
#include <algorithm>
#include <cstdio>
#include "common.cuh"
#include "philox.cuh"

namespace rvcdec {

__device__ __forceinline__ size_t rblk(int b, int nsl, int sl, int Lp, int t) { return (((size_t)b * nsl + sl) * Lp + PADL + t) * 8; }

// ---------------------------------------------------------------------------------------------------------------------
// Harmonic source (refinegan.py:180-263 with harmonic_num = 0, f0 interpolated at :419).
// F.interpolate(mode="linear", align_corners=False) is piecewise linear between frame centres: piece 0 = the first upp/2
// samples (constant f0[0]), piece p = 1 .. T-1 the upp samples between centres p-1 and p, piece T the last upp/2 samples
// (constant f0[T-1]).  The per-sample cumsum of f / sr (refinegan.py:232-238) is therefore
//     phase(s) = before[p] + (j+1) fa + (fb - fa) (j+1)^2 / (2 upp)        (j = index inside the piece)
// with before[p] = sum of the whole pieces in front, a prefix over T+1 values (fp64) that one kernel computes per item.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void refine_prefix_kernel(const float* __restrict__ f0, double* __restrict__ before, int T, int upp, double inv_sr) {
  // one CTA per item; T is a few thousand at most: a warp-strided two-level scan would be overkill
  const int b = blockIdx.x;
  const float* f = f0 + (size_t)b * T;
  double* o = before + (size_t)b * (T + 1);
  if (threadIdx.x == 0) {
    double acc = 0.0;
    o[0] = 0.0;
    acc += (double)f[0] * inv_sr * (upp / 2);
    o[1 > T ? T : 1] = acc - floor(acc);
    for (int p = 1; p < T; ++p) {
      acc += (double)upp * ((double)f[p - 1] + (double)f[p]) * 0.5 * inv_sr;
      acc -= floor(acc);  // only the phase modulo 1 matters
      o[p + 1] = acc;
    }
  }
}

__global__ void refine_source_kernel(const float* __restrict__ f0, const double* __restrict__ before, const float* __restrict__ noise,
                                     float* __restrict__ har8, uint64_t seed, uint64_t item0, int B, int T, int upp, int Lp,
                                     float inv_sr, float merge_w) {
  const size_t n = (size_t)T * upp;
  const size_t total = (size_t)B * n;
  const int half = upp / 2;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / n);
    const int s = (int)(i - (size_t)b * n);
    const float* f = f0 + (size_t)b * T;
    int p, j;
    if (s < half) { p = 0; j = s; }
    else { p = (s - half) / upp + 1; j = (s - half) - (p - 1) * upp; }
    const int ia = p == 0 ? 0 : p - 1, ib = p >= T ? T - 1 : p;
    const float f_a = f[ia], f_b = f[ib];
    const double fa = (double)f_a * inv_sr, fb = (double)f_b * inv_sr;
    const double jj = (double)(j + 1);
    const double ph = before[(size_t)b * (T + 1) + p] + jj * fa + (fb - fa) * jj * jj / (2.0 * (double)upp);
    const float frac = (float)(ph - floor(ph));
    const float sine = 0.1f * sinpif(2.0f * frac);
    // interpolated f0 > 0 (refinegan.py:221-224 on the interpolated contour): both weights are strictly positive inside a piece
    const bool voiced = (f_a > 0.f) || (f_b > 0.f);
    const float z = noise ? __ldg(noise + i) : philox_normal(seed, (item0 + (uint64_t)b) * (uint64_t)n + (uint64_t)s);
    const float v = voiced ? (sine + 0.003f * z) : ((0.1f / 3.0f) * z);
    float* o = har8 + rblk(b, 1, 0, Lp, s);
    *reinterpret_cast<float4*>(o) = make_float4(tanhf(merge_w * v), 0.f, 0.f, 0.f);
    *reinterpret_cast<float4*>(o + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

cudaError_t launch_refine_source(const float* f0, double* before, const float* noise, float* har8, uint64_t seed, uint64_t item0,
                                 int B, int T, int upp, int Lp, float sr, float merge_w, cudaStream_t s) {
  refine_prefix_kernel<<<B, 32, 0, s>>>(f0, before, T, upp, 1.0 / (double)sr);
  const size_t total = (size_t)B * T * upp;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 16));
  refine_source_kernel<<<blocks, 256, 0, s>>>(f0, before, noise, har8, seed, item0, B, T, upp, Lp, 1.f / sr, merge_w);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// Decimation by `rate` with torchaudio's sinc_interp_kaiser filter (refinegan.py:424-433): one FIR of 2*width + rate taps,
// the same for every channel, stride `rate`, input zero-padded by `width` in front:  y[o] = sum_j x[o*rate + j - width] k[j].
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) resample_kernel(const float* __restrict__ x, const float* __restrict__ taps, float* __restrict__ y,
                                                       int B, int C, int L_in, int Lp_in, int L_out, int Lp_out, int rate, int width,
                                                       int ntaps) {
  extern __shared__ float s_taps[];
  for (int i = threadIdx.x; i < ntaps; i += blockDim.x) s_taps[i] = taps[i];
  __syncthreads();
  const int nsl = C >> 3;
  const size_t total = (size_t)B * nsl * L_out;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int o = (int)(i % L_out);
    const size_t bs = i / L_out;
    const int sl = (int)(bs % nsl);
    const int b = (int)(bs / nsl);
    const float* xs = x + rblk(b, nsl, sl, Lp_in, 0);
    const int r0 = o * rate - width;
    const int j0 = r0 < 0 ? -r0 : 0;
    const int j1 = min(ntaps, L_in - r0);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int j = j0; j < j1; ++j) {
      const float k = s_taps[j];
      const float* px = xs + (size_t)(r0 + j) * 8;
      const float4 a = *reinterpret_cast<const float4*>(px), c = *reinterpret_cast<const float4*>(px + 4);
      acc[0] = fmaf(a.x, k, acc[0]); acc[1] = fmaf(a.y, k, acc[1]); acc[2] = fmaf(a.z, k, acc[2]); acc[3] = fmaf(a.w, k, acc[3]);
      acc[4] = fmaf(c.x, k, acc[4]); acc[5] = fmaf(c.y, k, acc[5]); acc[6] = fmaf(c.z, k, acc[6]); acc[7] = fmaf(c.w, k, acc[7]);
    }
    float* po = y + rblk(b, nsl, sl, Lp_out, o);
    *reinterpret_cast<float4*>(po) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(po + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}
cudaError_t launch_resample(const float* x, const float* taps, float* y, int B, int C, int L_in, int Lp_in, int L_out, int Lp_out,
                            int rate, int width, int ntaps, cudaStream_t s) {
  const size_t total = (size_t)B * (C / 8) * L_out;
  const int blocks = (int)std::min<size_t>((total + 127) / 128, (size_t)(148 * 32));
  resample_kernel<<<blocks, 128, (size_t)ntaps * sizeof(float), s>>>(x, taps, y, B, C, L_in, Lp_in, L_out, Lp_out, rate, width, ntaps);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// x = cat([Upsample(rate, linear)(leaky_relu(x, slope)), skip])   (refinegan.py:444-452).  skip is stored activated already.
// out has C + Cs channels at L * rate rows.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void upsample_cat_kernel(const float* __restrict__ x, const float* __restrict__ skip, float* __restrict__ out, int B, int C,
                                    int Cs, int L, int Lp, int rate, int Lp_out, float slope) {
  const int nsl = C >> 3, nss = Cs >> 3, nso = nsl + nss;
  const int Lo = L * rate;
  const size_t total = (size_t)B * nso * Lo;
  const float inv = 1.0f / (float)rate;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int s = (int)(i % Lo);
    const size_t bs = i / Lo;
    const int sl = (int)(bs % nso);
    const int b = (int)(bs / nso);
    float v[8];
    if (sl < nsl) {
      // torch upsample_linear1d, align_corners=False: src = scale * (dst + 0.5) - 0.5, clamped at 0
      float src = inv * ((float)s + 0.5f) - 0.5f;
      if (src < 0.f) src = 0.f;
      const int i0 = min((int)src, L - 1);
      const int i1 = min(i0 + 1, L - 1);
      const float l1 = src - (float)i0, l0 = 1.f - l1;
      const float* p0 = x + rblk(b, nsl, sl, Lp, i0);
      const float* p1 = x + rblk(b, nsl, sl, Lp, i1);
      float a[8], c[8];
      *reinterpret_cast<float4*>(a) = *reinterpret_cast<const float4*>(p0);
      *reinterpret_cast<float4*>(a + 4) = *reinterpret_cast<const float4*>(p0 + 4);
      *reinterpret_cast<float4*>(c) = *reinterpret_cast<const float4*>(p1);
      *reinterpret_cast<float4*>(c + 4) = *reinterpret_cast<const float4*>(p1 + 4);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = l0 * lrelu(a[k], slope) + l1 * lrelu(c[k], slope);
    } else {
      const float* ps = skip + rblk(b, nss, sl - nsl, Lp_out, s);
      *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(ps);
      *reinterpret_cast<float4*>(v + 4) = *reinterpret_cast<const float4*>(ps + 4);
    }
    float* po = out + rblk(b, nso, sl, Lp_out, s);
    *reinterpret_cast<float4*>(po) = *reinterpret_cast<float4*>(v);
    *reinterpret_cast<float4*>(po + 4) = *reinterpret_cast<float4*>(v + 4);
  }
}
cudaError_t launch_upsample_cat(const float* x, const float* skip, float* out, int B, int C, int Cs, int L, int Lp, int rate,
                                int Lp_out, float slope, cudaStream_t s) {
  const size_t total = (size_t)B * ((C + Cs) / 8) * L * rate;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 32));
  upsample_cat_kernel<<<blocks, 256, 0, s>>>(x, skip, out, B, C, Cs, L, Lp, rate, Lp_out, slope);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// AdaIN (refinegan.py:88-100): y = leaky_relu(x + randn * weight[c], slope).
//   mode 0 (in front of a ResBlock):  res = y,  act = leaky_relu(y, slope)  (the ResBlock's first activation, :79)
//   mode 1 (behind it, + the mean over the parallel blocks, :158):  acc = (first ? 0 : acc) + y * scale
// noise: (B, C, L) plain fp32 draws to inject, or null -> in-kernel Philox at counter offset `ctr0`.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void adain_kernel(const float* __restrict__ x, const float* __restrict__ noise, const float* __restrict__ w, float* __restrict__ res,
                             float* __restrict__ act, float* __restrict__ acc, int mode, int first, float scale, float slope, uint64_t seed,
                             uint64_t ctr0, int B, int C, int L, int Lp) {
  const int nsl = C >> 3;
  const size_t total = (size_t)B * nsl * L;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % L);
    const size_t bs = i / L;
    const int sl = (int)(bs % nsl);
    const int b = (int)(bs / nsl);
    const size_t o = rblk(b, nsl, sl, Lp, t);
    float v[8];
    *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(x + o);
    *reinterpret_cast<float4*>(v + 4) = *reinterpret_cast<const float4*>(x + o + 4);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int ch = sl * 8 + k;
      const size_t ni = ((size_t)b * C + ch) * L + t;
      const float z = noise ? __ldg(noise + ni) : philox_normal(seed, ctr0 + (uint64_t)ni);
      v[k] = lrelu(v[k] + z * __ldg(w + ch), slope);
    }
    if (mode == 0) {
      *reinterpret_cast<float4*>(res + o) = *reinterpret_cast<float4*>(v);
      *reinterpret_cast<float4*>(res + o + 4) = *reinterpret_cast<float4*>(v + 4);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = lrelu(v[k], slope);
      *reinterpret_cast<float4*>(act + o) = *reinterpret_cast<float4*>(v);
      *reinterpret_cast<float4*>(act + o + 4) = *reinterpret_cast<float4*>(v + 4);
    } else {
      float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (!first) {
        *reinterpret_cast<float4*>(a) = *reinterpret_cast<const float4*>(acc + o);
        *reinterpret_cast<float4*>(a + 4) = *reinterpret_cast<const float4*>(acc + o + 4);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] += v[k] * scale;
      *reinterpret_cast<float4*>(acc + o) = *reinterpret_cast<float4*>(a);
      *reinterpret_cast<float4*>(acc + o + 4) = *reinterpret_cast<float4*>(a + 4);
    }
  }
}
cudaError_t launch_adain(const float* x, const float* noise, const float* w, float* res, float* act, float* acc, int mode, int first,
                         float scale, float slope, uint64_t seed, uint64_t ctr0, int B, int C, int L, int Lp, cudaStream_t s) {
  const size_t total = (size_t)B * (C / 8) * L;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)(148 * 32));
  adain_kernel<<<blocks, 256, 0, s>>>(x, noise, w, res, act, acc, mode, first, scale, slope, seed, ctr0, B, C, L, Lp);
  return cudaGetLastError();
}

}  // namespace rvcdec
