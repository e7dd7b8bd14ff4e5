// fp32 parity back-end: every convolution of the decoder as a CUDA-core FFMA implicit
// GEMM over the blocked-8 layout.  Same ConvParams / Epilogue contract as the tensor-core
// back-end (conv_mma.cu); this one exists so that the whole decoder can be checked against
// the reference at fp32 accuracy (max-abs <= 2e-3, BASELINE.json north_star).
//
// CTA tile: 128 time rows x 64 GEMM columns, 256 threads, 8x4 accumulators per thread.
// K loop: one 8-channel slice at a time; the slice's rows (+halo) are staged transposed in
// shared memory ([channel][row], so tap-shifted reads are conflict-free broadcasts) together
// with that slice's weights for every tap.
#include "common.cuh"

namespace rvcdec {

constexpr int F32_TM = 128;
constexpr int F32_TN = 64;
constexpr int F32_HALO = 32;
constexpr int F32_ROWS = F32_TM + 2 * F32_HALO;

__global__ void __launch_bounds__(256) conv_f32_kernel(const __grid_constant__ ConvParams p) {
  __shared__ float As[8][F32_ROWS];
  __shared__ __align__(16) float Ws[MAX_TAPS][8][F32_TN];

  const int tid = threadIdx.x;
  const int tx = tid & 15;   // column group: cols tx*4 .. tx*4+3
  const int ty = tid >> 4;   // row group:    rows ty*8 .. ty*8+7
  int blk = blockIdx.x;
  const int n_tile = blk % p.n_tiles; blk /= p.n_tiles;
  const int m_tile = blk % p.m_tiles;
  const int b = blk / p.m_tiles;
  const int t0 = m_tile * F32_TM;
  const int n0 = n_tile * F32_TN;
  const uint32_t mask = p.tapmask[n_tile];
  const int nsl = p.Cin >> 3;
  const int nsl_phys = p.in_slices_phys ? p.in_slices_phys : nsl;
  const int rows = F32_TM + 2 * p.halo;
  const float* act = reinterpret_cast<const float*>(p.act_in);
  const float* W = reinterpret_cast<const float*>(p.w);

  float acc[8][4];
#pragma unroll
  for (int r = 0; r < 8; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;

  for (int cs = 0; cs < nsl; ++cs) {
    // stage activations: rows t0-halo .. t0+TM+halo of slice cs, transposed
    const float* src = act + (((size_t)b * nsl_phys + cs) * p.Lp_in + PADL + t0 - p.halo) * 8;
    for (int i = tid; i < rows * 2; i += 256) {
      const int r = i >> 1, h = i & 1;
      const float4 q = *reinterpret_cast<const float4*>(src + (size_t)r * 8 + h * 4);
      As[h * 4 + 0][r] = q.x; As[h * 4 + 1][r] = q.y; As[h * 4 + 2][r] = q.z; As[h * 4 + 3][r] = q.w;
    }
    // stage weights: [tap][ci][64 cols]
    for (int i = tid; i < p.ntaps * 8 * (F32_TN / 4); i += 256) {
      const int n4 = i % (F32_TN / 4);
      const int ci = (i / (F32_TN / 4)) & 7;
      const int tap = i / (8 * (F32_TN / 4));
      float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n0 + n4 * 4 < p.N_total)
        q = __ldg(reinterpret_cast<const float4*>(W + ((size_t)tap * p.Cin + cs * 8 + ci) * p.N_total + n0 + n4 * 4));
      *reinterpret_cast<float4*>(&Ws[tap][ci][n4 * 4]) = q;
    }
    __syncthreads();
    for (int tap = 0; tap < p.ntaps; ++tap) {
      if (!((mask >> tap) & 1u)) continue;
      const int sh = p.halo + p.shift[tap] + ty * 8;
#pragma unroll
      for (int ci = 0; ci < 8; ++ci) {
        float a[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) a[r] = As[ci][sh + r];
        const float4 w = *reinterpret_cast<const float4*>(&Ws[tap][ci][tx * 4]);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          acc[r][0] = fmaf(a[r], w.x, acc[r][0]);
          acc[r][1] = fmaf(a[r], w.y, acc[r][1]);
          acc[r][2] = fmaf(a[r], w.z, acc[r][2]);
          acc[r][3] = fmaf(a[r], w.w, acc[r][3]);
        }
      }
    }
    __syncthreads();
  }

  const int n = n0 + tx * 4;
  if (n >= p.N_total) return;
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const int t = t0 + ty * 8 + r;
    if (t < p.L_in) epilogue_store<4, true>(p.epi, b, t, n, acc[r]);
  }
}

cudaError_t launch_conv_f32(const ConvParams& p, cudaStream_t s) {
  if (p.halo > F32_HALO || p.ntaps > MAX_TAPS) return cudaErrorInvalidValue;
  ConvParams q = p;
  q.m_tiles = (p.L_in + F32_TM - 1) / F32_TM;
  q.n_tiles = (p.N_total + F32_TN - 1) / F32_TN;
  const long long grid = (long long)q.B * q.m_tiles * q.n_tiles;
  if (grid <= 0 || grid > 0x7fffffffLL) return cudaErrorInvalidValue;
  conv_f32_kernel<<<(unsigned)grid, 256, 0, s>>>(q);
  return cudaGetLastError();
}

}  // namespace rvcdec
