// 3 x 3 convolutions of the RMVPE U-Net on the tensor cores (include/rvcrmvpe.h, option "tc"): the twin of conv2d_kernel<3>
// (rmvpe_kernels.cu) for Cin % 16 == 0 and Cout = 32, 64 or a multiple of 128 - every ConvBlockRes convolution from level 1 down.  On the CUDA cores each of them took ~0.2 ms at 60 s of audio (1.8 GMAC whatever the
// level) and the deep ones 0.7 ms whatever the clip length (a 512-channel convolution over a 188 x 4 map on 96 small CTAs).
//
// Implicit GEMM per item: rows = pixels (128 per CTA), columns = output channels (BN = min(Cout, 128) per CTA),
// k = tap * Cin + ci in steps of 64 (zero-padded to a multiple of 64: 9 * 16 = 144 -> 192).  Activations stay (B, C, H, W)
// fp32 between layers: eight producer warps read them (one pixel per thread, so a warp reads one coalesced 128-byte line
// per channel), convert to fp16 and store the A tile in the no-swizzle K-major layout ([8-channel chunk][pixel][8]; a
// chunk never straddles a tap because Cin % 8 == 0), two groups of four warps alternating steps; the weights come as
// pre-packed fp16 tile images (one bulk copy per step); one warp issues tcgen05.mma into TMEM; the first four warps then
// read the accumulator (thread = pixel, so every store of an output channel is a coalesced line in (B, C, H, W)) and apply
// the folded BatchNorm, ReLU and the residual.  The input may be the concatenation of two tensors (decoder blocks).
#include <cuda_fp16.h>

#include <cstdint>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

namespace ctc {
constexpr int BM = 128, BKK = 64;
constexpr int A_TILE = BM * BKK * 2;          // 16 KB
constexpr int CH = BM * 16;                   // bytes between 8-element K chunks of the A tile
constexpr int THREADS = 320;                  // warps 0-7 producers, 8 weight loader, 9 MMA issuer
// BN = 128 (the deep levels: few CTAs, 18-144 K steps each): one CTA per SM, four stages, all 64 loads of a step in flight per
// thread.  BN < 128 (the shallow levels: thousands of CTAs of 3-9 steps): two CTAs per SM, three stages, loads in two batches
// of 32 to stay under 102 registers.
template <int BN> struct Geo {
  static constexpr bool DEEP = BN == 128;
  static constexpr int STAGES = DEEP ? 4 : 3, BATCHES = DEEP ? 1 : 2, MIN_CTAS = DEEP ? 1 : 2;
  static constexpr int MT = 1;  // pixel tiles per CTA, each with its own BN accumulator columns (128 / BN tiles per CTA were
                                // measured no faster: the producers' instruction stream, not the set-up, bounds the shallow levels)
  static constexpr int STAGE = A_TILE + BN * BKK * 2;
  static constexpr int SMEM = 128 + STAGES * STAGE + 8 * (2 * STAGES + 1) + 16;
};
inline int tile_n(int Cout) { return Cout >= 128 ? 128 : Cout; }
}  // namespace ctc

struct ConvTcParams {
  const float* in0; const float* in1; int C0, Cin;
  const __half* wimg;   // [Cout / BN][nk][8 chunks][BN rows][8]: k = tap * Cin + ci, zero beyond 9 * Cin
  const float* scale; const float* shift; const float* res;
  float* out;
  int B, Cout, H, W, relu;
  float* part;   // split-K: raw partial sums [split][B][Cout][H W] (conv3x3_tc_finish_kernel reduces them), else null
  int ksplit;    // number of K splits (1 = none)
};

template <int BN>
__global__ void __launch_bounds__(ctc::THREADS, ctc::Geo<BN>::MIN_CTAS) conv3x3_tc_kernel(const ConvTcParams p) {
  using namespace ctc;
  constexpr int STAGES = Geo<BN>::STAGES, NB = Geo<BN>::BATCHES, CPB = 8 / NB;  // chunks per batch
  constexpr int STAGE = Geo<BN>::STAGE, W_TILE = BN * BKK * 2, WCH = BN * 16;
  constexpr int OFF_BAR = STAGES * STAGE;
  constexpr int MT = Geo<BN>::MT;
  constexpr uint32_t TM_COLS = MT * BN < 32 ? 32 : MT * BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sg = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const uint32_t sb = smem_u32(sg);
  const uint32_t full0 = sb + OFF_BAR, empty0 = full0 + 8 * STAGES, acc_full = empty0 + 8 * STAGES, tmem_slot = acc_full + 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sg + OFF_BAR + 8 * (2 * STAGES + 1));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  static_assert(MT == 1, "one pixel tile per CTA");
  const int n_tile = blockIdx.x, m0 = blockIdx.y * BM, b = blockIdx.z / p.ksplit, sp = blockIdx.z - b * p.ksplit;
  const int HW = p.H * p.W, K9 = 9 * p.Cin, nk = (K9 + BKK - 1) / BKK;
  const int kbeg = (int)((long long)nk * sp / p.ksplit), nq = (int)((long long)nk * (sp + 1) / p.ksplit) - kbeg;  // this CTA's K steps
  constexpr int ntile = 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full0 + 8 * s, 5); /* four producer warps + the weight loader's expect_tx */ mbar_init(empty0 + 8 * s, 1); }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 9) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot_ptr;

  if (warp < 8) {
    // ===================== producers: im2col rows, fp32 -> fp16 =====================
    const int grp = warp >> 2, r = (warp & 3) * 32 + lane;
    const int C1 = p.Cin - p.C0;
    for (int q = grp; q < nq; q += 2) {
      const int tile = 0, kt = kbeg + q, s = q % STAGES;
      const int m = m0 + tile * BM + r;
      const bool valid = m < HW;
      const int y = valid ? m / p.W : 0, x = valid ? m - y * p.W : 0;
      uint8_t* dst = sg + s * STAGE + r * 16;
      int tap = kt * BKK / p.Cin, ci = kt * BKK - tap * p.Cin;  // of the step's first chunk; chunks advance by 8 channels
#pragma unroll 1
      for (int hb = 0; hb < NB; ++hb) {
        float v[CPB * 8];
#pragma unroll
        for (int c4 = 0; c4 < CPB; ++c4) {
          const int dy = (tap * 11) >> 5, dx = tap - 3 * dy;  // tap / 3, tap % 3 for tap < 9
          const int yy = y + dy - 1, xx = x + dx - 1;
          const bool inb = valid && tap < 9 && (unsigned)yy < (unsigned)p.H && (unsigned)xx < (unsigned)p.W;
          const float* src = (ci < p.C0 ? p.in0 + ((size_t)b * p.C0 + ci) * HW : p.in1 + ((size_t)b * C1 + (ci - p.C0)) * HW) +
                             (inb ? yy * p.W + xx : 0);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[c4 * 8 + e] = inb ? __ldg(src + (size_t)e * HW) : 0.f;
          ci += 8;
          if (ci >= p.Cin) { ci = 0; ++tap; }
        }
        if (hb == 0 && q >= STAGES) mbar_wait(empty0 + 8 * s, ((q / STAGES) - 1) & 1);
#pragma unroll
        for (int c4 = 0; c4 < CPB; ++c4) {
          __half2 h0 = __floats2half2_rn(v[8 * c4], v[8 * c4 + 1]), h1 = __floats2half2_rn(v[8 * c4 + 2], v[8 * c4 + 3]);
          __half2 h2 = __floats2half2_rn(v[8 * c4 + 4], v[8 * c4 + 5]), h3 = __floats2half2_rn(v[8 * c4 + 6], v[8 * c4 + 7]);
          *reinterpret_cast<uint4*>(dst + (hb * CPB + c4) * CH) = make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1),
                                                                          *reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
        }
      }
      fence_proxy_async();  // generic-proxy stores, read by the MMA through the async proxy
      __syncwarp();
      if (lane == 0) mbar_arrive(full0 + 8 * s);  // one arrival per warp
    }
    if (warp < 4) {
      // ===================== epilogue: thread = pixel = TMEM lane =====================
      mbar_wait(acc_full, 0);
      tc_fence_after();
      const uint32_t tl = tm + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
      for (int tile = 0; tile < ntile; ++tile) {
        const int m = m0 + tile * BM + r;
        const bool valid = m < HW;
        auto finish = [&](int co, uint32_t a) {
          if (p.part) {  // split-K: the raw partial sum; scale / shift / ReLU / residual wait for the reduction
            p.part[(((size_t)sp * p.B + b) * p.Cout + co) * HW + m] = __uint_as_float(a);
            return;
          }
          float o = __uint_as_float(a) * (p.scale ? __ldg(p.scale + co) : 1.f) + (p.shift ? __ldg(p.shift + co) : 0.f);
          if (p.relu) o = fmaxf(o, 0.f);
          const size_t oi = ((size_t)b * p.Cout + co) * HW + m;
          if (p.res) o += p.res[oi];
          p.out[oi] = o;
        };
        if constexpr (BN >= 32) {
#pragma unroll 1
          for (int c = 0; c < BN / 32; ++c) {
            uint32_t acc[32];
            tc_ld32(tl + tile * BN + c * 32, acc);
            if (!valid) continue;
#pragma unroll
            for (int i = 0; i < 32; ++i) finish(n_tile * BN + c * 32 + i, acc[i]);
          }
        } else {
          uint32_t acc[16];
          tc_ld16(tl + tile * BN, acc);
          if (valid) {
#pragma unroll
            for (int i = 0; i < 16; ++i) finish(n_tile * BN + i, acc[i]);
          }
        }
      }
      tc_fence_before();
    }
  } else if (warp == 8) {
    // ===================== weight loader: one tile image per step =====================
    if (lane == 0) {
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.wimg) + (size_t)n_tile * nk * W_TILE;
      for (int q = 0; q < nq; ++q) {
        const int s = q % STAGES, kt = kbeg + q;
        if (q >= STAGES) mbar_wait(empty0 + 8 * s, ((q / STAGES) - 1) & 1);
        mbar_expect_tx(full0 + 8 * s, W_TILE);
        bulk_g2s(sb + s * STAGE + A_TILE, wsrc + (size_t)kt * W_TILE, W_TILE, full0 + 8 * s);
      }
    }
    __syncwarp();
  } else {
    // ===================== MMA issuer =====================
    const uint32_t idesc = make_idesc(ACT_F16, BN);
    for (int q = 0; q < nq; ++q) {
      const int s = q % STAGES, tile = 0;
      mbar_wait(full0 + 8 * s, (q / STAGES) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = sb + s * STAGE, sw = sa + A_TILE;
#pragma unroll
        for (int ks = 0; ks < BKK / 16; ++ks)
          tc_mma_f16(tm + tile * BN, make_smem_desc(sa + 2 * ks * CH, CH, 128), make_smem_desc(sw + 2 * ks * WCH, WCH, 128), idesc,
                     (q | ks) ? 1u : 0u);
        tc_commit(empty0 + 8 * s);
        if (q == nq - 1) tc_commit(acc_full);
      }
      __syncwarp();
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(TM_COLS) : "memory");
  }
}

// (Cout = 16 compiles and is correct, but is no faster than the CUDA-core kernel - 216 against 214 us at 6016 frames: with nine
// taps of 16 channels a thread-gathered im2col row costs as many instructions as the FMAs it replaces - so level 0 stays there.)
bool conv3x3_tc_supported(int Cin, int C0, int Cout) {
  return Cin % 16 == 0 && C0 % 8 == 0 && (Cout == 32 || Cout == 64 || Cout % 128 == 0);
}
// halves of the packed image of one layer
size_t conv3x3_tc_image_elems(int Cin, int Cout) { return (size_t)Cout * ((9 * Cin + ctc::BKK - 1) / ctc::BKK) * ctc::BKK; }

// w (Cout, Cin, 3, 3) fp32 on the host -> the tile images the kernel copies: [Cout / BN][nk][8][BN][8], k = tap * Cin + ci
void conv3x3_tc_pack_host(const float* w, int Cin, int Cout, __half* img) {
  using namespace ctc;
  const int BN = tile_n(Cout), K9 = 9 * Cin, nk = (K9 + BKK - 1) / BKK;
  for (int nt = 0; nt < Cout / BN; ++nt)
    for (int kt = 0; kt < nk; ++kt)
      for (int c = 0; c < 8; ++c)
        for (int n = 0; n < BN; ++n)
          for (int e = 0; e < 8; ++e) {
            const int k = kt * BKK + c * 8 + e;
            float v = 0.f;
            if (k < K9) {
              const int tap = k / Cin, ci = k % Cin;
              v = w[((size_t)(nt * BN + n) * Cin + ci) * 9 + tap];
            }
            img[((((size_t)nt * nk + kt) * 8 + c) * BN + n) * 8 + e] = __float2half_rn(v);
          }
}

// split-K reduction in a fixed order (bit-reproducible), then the epilogue of the unsplit kernel
__global__ void conv3x3_tc_finish_kernel(const float* __restrict__ part, int ksplit, const float* __restrict__ scale, const float* __restrict__ shift,
                                         const float* __restrict__ res, float* __restrict__ out, size_t n, int Cout, int HW, int relu) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float a = 0.f;
  for (int s = 0; s < ksplit; ++s) a += part[(size_t)s * n + i];
  const int co = (int)((i / HW) % Cout);
  float o = a * (scale ? __ldg(scale + co) : 1.f) + (shift ? __ldg(shift + co) : 0.f);
  if (relu) o = fmaxf(o, 0.f);
  if (res) o += res[i];
  out[i] = o;
}

// number of K splits for a layer: only where the grid would leave most SMs idle (the deep levels: 4-24 CTAs of 36-144 steps)
int conv3x3_tc_splits(int Cin, int Cout, int HW, int B) {
  using namespace ctc;
  const int BN = tile_n(Cout), nk = (9 * Cin + BKK - 1) / BKK;
  const long long ctas = (long long)(Cout / BN) * ((HW + BM - 1) / BM) * B;
  if (BN != 128 || ctas >= 74 || nk < 8) return 1;
  int sp = (int)((148 + ctas - 1) / ctas);
  if (sp > nk / 4) sp = nk / 4;
  if (sp > 16) sp = 16;
  return sp < 2 ? 1 : sp;
}

template <int BN>
static cudaError_t launch_conv3x3_tc_n(const ConvTcParams& p, cudaStream_t s) {
  using namespace ctc;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && !attr_set[dev]) {
    e = cudaFuncSetAttribute(conv3x3_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Geo<BN>::SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  conv3x3_tc_kernel<BN><<<dim3(p.Cout / BN, (p.H * p.W + BM - 1) / BM, p.B * p.ksplit), THREADS, Geo<BN>::SMEM, s>>>(p);
  if (p.part) {
    const size_t n = (size_t)p.B * p.Cout * p.H * p.W;
    conv3x3_tc_finish_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(p.part, p.ksplit, p.scale, p.shift, p.res, p.out, n, p.Cout, p.H * p.W, p.relu);
  }
  return cudaGetLastError();
}

cudaError_t launch_conv3x3_tc(const ConvTcParams& p, cudaStream_t s) {
  if (!conv3x3_tc_supported(p.Cin, p.C0, p.Cout) || p.ksplit < 1 || (p.ksplit > 1 && !p.part)) return cudaErrorInvalidValue;
  switch (ctc::tile_n(p.Cout)) {
    case 128: return launch_conv3x3_tc_n<128>(p, s);
    case 64: return launch_conv3x3_tc_n<64>(p, s);
    case 32: return launch_conv3x3_tc_n<32>(p, s);
    case 16: return launch_conv3x3_tc_n<16>(p, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace rvcdec
