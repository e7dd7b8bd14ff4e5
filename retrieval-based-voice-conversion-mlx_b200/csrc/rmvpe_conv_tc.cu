// 3 x 3 convolutions of the RMVPE U-Net's deep levels on the tensor cores (include/rvcrmvpe.h, option "tc"): the twin of
// conv2d_kernel<3> (rmvpe_kernels.cu) for Cin % 64 == 0 and Cout % 128 == 0 (encoder levels 3-4, the intermediate blocks,
// decoder levels 0-1: 0.7 ms per convolution on the CUDA cores whatever the clip length, because a 512-channel 3 x 3
// convolution over a 188 x 4 map is 1.8 GMAC on 96 small CTAs).
//
// Implicit GEMM per item: rows = pixels (128 per CTA), columns = output channels (128 per CTA), K = 9 taps x Cin in steps of
// 64 (one tap, 64 input channels).  Activations stay (B, C, H, W) fp32 between layers: eight producer warps read them
// (one pixel per thread, one coalesced 128-byte line per warp and channel), convert to fp16 and store the A tile in the
// no-swizzle K-major layout ([8-channel chunk][pixel][8]), two groups of four warps alternating steps; the weights come as
// pre-packed fp16 tile images (one 16 KB bulk copy per step); one warp issues tcgen05.mma into TMEM; the first four warps
// then read the accumulator (thread = pixel, so every store of an output channel is a coalesced line in (B, C, H, W)) and
// apply the folded BatchNorm, ReLU and the residual.  The input may be the concatenation of two tensors (decoder blocks).
#include <cuda_fp16.h>

#include <cstdint>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

namespace ctc {
constexpr int BM = 128, BN = 128, BKK = 64, STAGES = 4;
constexpr int TILE = BM * BKK * 2;            // 16 KB per operand per stage
constexpr int CH = BM * 16;                   // bytes between 8-element K chunks of a tile
constexpr int THREADS = 320;                  // warps 0-7 producers, 8 weight loader, 9 MMA issuer
constexpr int OFF_BAR = STAGES * 2 * TILE;
constexpr int SMEM = 128 + OFF_BAR + 8 * (2 * STAGES + 1) + 16;
}  // namespace ctc

struct ConvTcParams {
  const float* in0; const float* in1; int C0, Cin;
  const __half* wimg;   // [Cout / 128][9 * Cin / 64][8 chunks][128 rows][8]: k = tap * Cin + ci
  const float* scale; const float* shift; const float* res;
  float* out;
  int B, Cout, H, W, relu;
};

__global__ void __launch_bounds__(ctc::THREADS, 1) conv3x3_tc_kernel(const ConvTcParams p) {
  using namespace ctc;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sg = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const uint32_t sb = smem_u32(sg);
  const uint32_t full0 = sb + OFF_BAR, empty0 = full0 + 8 * STAGES, acc_full = empty0 + 8 * STAGES, tmem_slot = acc_full + 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sg + OFF_BAR + 8 * (2 * STAGES + 1));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x, m0 = blockIdx.y * BM, b = blockIdx.z;
  const int HW = p.H * p.W, cpt = p.Cin / BKK, nk = 9 * cpt;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full0 + 8 * s, 129); mbar_init(empty0 + 8 * s, 1); }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 9) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot_ptr;

  if (warp < 8) {
    // ===================== producers: im2col rows of one tap and 64 channels, fp32 -> fp16 =====================
    const int grp = warp >> 2, r = (warp & 3) * 32 + lane, m = m0 + r;
    const bool valid = m < HW;
    const int y = valid ? m / p.W : 0, x = valid ? m - y * p.W : 0;
    const int C1 = p.Cin - p.C0;
    for (int kt = grp; kt < nk; kt += 2) {
      const int s = kt % STAGES;
      const int tap = kt / cpt, ci0 = (kt - tap * cpt) * BKK;
      const int yy = y + tap / 3 - 1, xx = x + tap % 3 - 1;
      const bool inb = valid && (unsigned)yy < (unsigned)p.H && (unsigned)xx < (unsigned)p.W;
      const float* src = (ci0 < p.C0 ? p.in0 + ((size_t)b * p.C0 + ci0) * HW : p.in1 + ((size_t)b * C1 + (ci0 - p.C0)) * HW) + (inb ? yy * p.W + xx : 0);
      float v[BKK];
#pragma unroll
      for (int i = 0; i < BKK; ++i) v[i] = inb ? __ldg(src + (size_t)i * HW) : 0.f;
      if (kt >= STAGES) mbar_wait(empty0 + 8 * s, ((kt / STAGES) - 1) & 1);
      uint8_t* dst = sg + s * 2 * TILE + r * 16;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        __half2 h0 = __floats2half2_rn(v[8 * c], v[8 * c + 1]), h1 = __floats2half2_rn(v[8 * c + 2], v[8 * c + 3]);
        __half2 h2 = __floats2half2_rn(v[8 * c + 4], v[8 * c + 5]), h3 = __floats2half2_rn(v[8 * c + 6], v[8 * c + 7]);
        *reinterpret_cast<uint4*>(dst + c * CH) = make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1),
                                                             *reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
      }
      fence_proxy_async();  // generic-proxy stores, read by the MMA through the async proxy
      mbar_arrive(full0 + 8 * s);
    }
    if (warp < 4) {
      // ===================== epilogue: thread = pixel = TMEM lane =====================
      mbar_wait(acc_full, 0);
      tc_fence_after();
      const uint32_t tl = tm + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t acc[32];
        tc_ld32(tl + c * 32, acc);
        if (!valid) continue;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int co = n_tile * BN + c * 32 + i;
          float o = __uint_as_float(acc[i]) * (p.scale ? __ldg(p.scale + co) : 1.f) + (p.shift ? __ldg(p.shift + co) : 0.f);
          if (p.relu) o = fmaxf(o, 0.f);
          const size_t oi = ((size_t)b * p.Cout + co) * HW + m;
          if (p.res) o += p.res[oi];
          p.out[oi] = o;
        }
      }
      tc_fence_before();
    }
  } else if (warp == 8) {
    // ===================== weight loader: one 16 KB tile image per step =====================
    if (lane == 0) {
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.wimg) + (size_t)n_tile * nk * TILE;
      for (int kt = 0; kt < nk; ++kt) {
        const int s = kt % STAGES;
        if (kt >= STAGES) mbar_wait(empty0 + 8 * s, ((kt / STAGES) - 1) & 1);
        mbar_expect_tx(full0 + 8 * s, TILE);
        bulk_g2s(sb + s * 2 * TILE + TILE, wsrc + (size_t)kt * TILE, TILE, full0 + 8 * s);
      }
    }
    __syncwarp();
  } else {
    // ===================== MMA issuer =====================
    const uint32_t idesc = make_idesc(ACT_F16, BN);
    for (int kt = 0; kt < nk; ++kt) {
      const int s = kt % STAGES;
      mbar_wait(full0 + 8 * s, (kt / STAGES) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = sb + s * 2 * TILE, sw = sa + TILE;
#pragma unroll
        for (int ks = 0; ks < BKK / 16; ++ks)
          tc_mma_f16(tm, make_smem_desc(sa + 2 * ks * CH, CH, 128), make_smem_desc(sw + 2 * ks * CH, CH, 128), idesc, (kt | ks) ? 1u : 0u);
        tc_commit(empty0 + 8 * s);
        if (kt == nk - 1) tc_commit(acc_full);
      }
      __syncwarp();
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(128u) : "memory");
  }
}

bool conv3x3_tc_supported(int Cin, int C0, int Cout) { return Cin % ctc::BKK == 0 && C0 % ctc::BKK == 0 && Cout % ctc::BN == 0; }
// halves of the packed image of one layer
size_t conv3x3_tc_image_elems(int Cin, int Cout) { return (size_t)Cout * Cin * 9; }

// w (Cout, Cin, 3, 3) fp32 on the host -> the tile images the kernel copies: [Cout / 128][9 Cin / 64][8][128][8], k = tap * Cin + ci
void conv3x3_tc_pack_host(const float* w, int Cin, int Cout, __half* img) {
  using namespace ctc;
  const int nk = 9 * Cin / BKK;
  for (int nt = 0; nt < Cout / BN; ++nt)
    for (int kt = 0; kt < nk; ++kt)
      for (int c = 0; c < 8; ++c)
        for (int n = 0; n < BN; ++n)
          for (int e = 0; e < 8; ++e) {
            const int k = kt * BKK + c * 8 + e, tap = k / Cin, ci = k % Cin;
            img[((((size_t)nt * nk + kt) * 8 + c) * BN + n) * 8 + e] = __float2half_rn(w[((size_t)(nt * BN + n) * Cin + ci) * 9 + tap]);
          }
}

cudaError_t launch_conv3x3_tc(const ConvTcParams& p, cudaStream_t s) {
  using namespace ctc;
  if (!conv3x3_tc_supported(p.Cin, p.C0, p.Cout)) return cudaErrorInvalidValue;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && !attr_set[dev]) {
    e = cudaFuncSetAttribute(conv3x3_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  conv3x3_tc_kernel<<<dim3(p.Cout / BN, (p.H * p.W + BM - 1) / BM, p.B), THREADS, SMEM, s>>>(p);
  return cudaGetLastError();
}

}  // namespace rvcdec
