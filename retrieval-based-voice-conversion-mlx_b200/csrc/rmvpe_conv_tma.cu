// 3 x 3 convolutions of the RMVPE U-Net as TMA-fed implicit GEMMs (include/rvcrmvpe.h, options "tc" + "tma"): the successor of
// rmvpe_conv_tc.cu, whose producer warps gather im2col rows pixel by pixel (~1 us per K step).
//
// The convolution input is kept as a second copy of the activation: fp16, channels-last, with a one-pixel zero border,
//   in16[b][y + 1][x + 1][c]   ((H + 2) x (W + 2) pixels per item, rows of C halves),
// so that in the flattened padded raster r = (b (H + 2) + y + 1)(W + 2) + x + 1 the nine taps are nine ROW SHIFTS,
// r + (dy - 1)(W + 2) + (dx - 1), and the A tile of (128 consecutive raster rows, one tap, 64 channels) is one box of a 2-D
// tensor map {C, rows} (rows before the first / after the last pixel come back as zeros).  The weights are a 2-D map over
// [tap][Cout][Cin].  One thread issues two cp.async.bulk.tensor loads per K step, one thread the tcgen05.mma (128-byte swizzle,
// as in gemm_tma.cu); the epilogue (thread = raster row) applies the folded BatchNorm / ReLU / residual and writes
//   out32: the (B, C, H, W) fp32 activation the rest of the model reads (interior pixels only), and / or
//   out16: the padded channels-last fp16 copy for the next convolution (zeros on the border rows it owns).
// Border rows cost (H + 2)(W + 2) / (H W) extra MMA rows (1.5 x at the 4-wide deepest level, 1.03 x at level 1).
// rmvpe_to_nhwc16_kernel builds the copy for tensors that do not come out of such an epilogue (pooling, transposed conv + skip).
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdint>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

namespace ctm {
constexpr int BM = 128;
constexpr int THREADS = 192;         // warps 0-3 epilogue, 4 TMA producer, 5 MMA issuer
// KB = channels per K step = min(Cin, 64): rows of 128 / 64 / 32 bytes, i.e. the 128- / 64- / 32-byte swizzle of TMA and UMMA
// A CTA of the shallow levels has only nine K steps (profiles/r2bo_*: ~8 us per CTA, the epilogue warps wait for the accumulator
// 37 % of the time); keeping all nine steps in flight at once changed nothing measurable, and four pixel tiles per CTA made it
// slower (5 s clip 2.8 -> 3.5 ms): the shallow levels want MORE CTAs in flight, not fewer set-ups - six per SM instead of four
// (56 registers, 30 KB of stages each) took 5 min of audio from 60.8 to 57.4 ms; eight do not fit the registers of tcgen05.ld.
template <int BN, int KB> struct Geo {
  static constexpr int STAGES = KB == 16 ? 6 : (KB == 32 ? 3 : 4);  // sized so that six CTAs of the shallow variants share an SM
  static constexpr int A_TILE = BM * KB * 2, W_TILE = BN * KB * 2, STAGE = A_TILE + (W_TILE + 1023) / 1024 * 1024;
  static constexpr int SMEM = 1024 + STAGES * STAGE + 8 * (2 * STAGES + 1) + 16;
  static constexpr int MIN_CTAS = (KB == 64 && BN == 128) ? 1 : (KB == 64 ? 2 : 6);
};
// K-major operand tile written by TMA with the swizzle that matches its row length: 8-row atoms, 8 * row bytes apart
template <int KB> __device__ __forceinline__ uint64_t sw_desc(uint32_t addr) {
  constexpr uint64_t layout = KB == 64 ? 2 : (KB == 32 ? 4 : 6);  // SWIZZLE_128B / 64B / 32B
  constexpr uint64_t sbo = 8 * KB * 2;
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((sbo >> 4) << 32) | ((uint64_t)1 << 46) | (layout << 61);
}
__device__ __forceinline__ void tma2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
               "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar)
               : "memory");
}
}  // namespace ctm

struct ConvTmaParams {
  CUtensorMap tmX, tmW;
  const float* scale; const float* shift; const float* res;  // res: (B, Cout, H, W) fp32 or null
  float* out32;      // (B, Cout, H, W) or null
  __half* out16;     // padded channels-last (B, H + 2, W + 2, Cout) or null
  int B, Cin, Cout, H, W, relu;
  int taps;      // 9 (3 x 3), 1 (the 1 x 1 shortcut projection: centre tap only) or 4 (stride-2 transposed convolution, see up_cout)
  int up_cout;   // taps == 4: Cout = 4 phases x up_cout channels; column (2 py + px) up_cout + co of input pixel (y, x) is output
                 // pixel (2 y + py, 2 x + px) of channel co, fed by the input pixels (y + oy, x + ox), tap = 2 oy + ox
  float* part;   // split-K: raw partial sums [split][B][Cout][H W]; conv3x3_tma_finish_kernel reduces them and runs the epilogue
  int ksplit;
};

template <int BN, int KB>
__global__ void __launch_bounds__(ctm::THREADS, ctm::Geo<BN, KB>::MIN_CTAS) conv3x3_tma_kernel(const __grid_constant__ ConvTmaParams p) {
  using namespace ctm;
  using G = Geo<BN, KB>;
  constexpr int STAGES = G::STAGES, STAGE = G::STAGE, A_TILE = G::A_TILE, W_TILE = G::W_TILE, OFF_BAR = STAGES * STAGE;
  constexpr uint32_t TM_COLS = BN < 32 ? 32 : BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sg = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sb = smem_u32(sg);
  const uint32_t full0 = sb + OFF_BAR, empty0 = full0 + 8 * STAGES, acc_full = empty0 + 8 * STAGES, tmem_slot = acc_full + 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sg + OFF_BAR + 8 * (2 * STAGES + 1));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, r0 = blockIdx.y * BM;
  const int Hp = p.H + 2, Wp = p.W + 2, kcs = p.Cin / KB, nk_all = p.taps * kcs, sp = blockIdx.z;
  const int kbeg = (int)((long long)nk_all * sp / p.ksplit), nk = (int)((long long)nk_all * (sp + 1) / p.ksplit) - kbeg;  // this CTA's K steps

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot_ptr;

  if (warp == 4) {
    if (lane == 0) {
      for (int q = 0; q < nk; ++q) {
        const int s = q % STAGES, tap = (kbeg + q) / kcs, kc = (kbeg + q) - tap * kcs;
        const int dy = p.taps == 1 ? 1 : (tap * 11) >> 5, dx = p.taps == 1 ? 1 : tap - 3 * dy;  // tap / 3, tap % 3
        if (q >= STAGES) mbar_wait(empty0 + 8 * s, ((q / STAGES) - 1) & 1);
        const uint32_t sa = sb + s * STAGE, sw = sa + A_TILE;
        mbar_expect_tx(full0 + 8 * s, A_TILE + W_TILE);
        const int shift = p.taps == 4 ? (tap >> 1) * Wp + (tap & 1) : (dy - 1) * Wp + (dx - 1);
        tma2d(sa, &p.tmX, kc * KB, r0 + shift, full0 + 8 * s);
        tma2d(sw, &p.tmW, kc * KB, tap * p.Cout + n0, full0 + 8 * s);
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    const uint32_t idesc = make_idesc(ACT_F16, BN);
    for (int q = 0; q < nk; ++q) {
      const int s = q % STAGES;
      mbar_wait(full0 + 8 * s, (q / STAGES) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = sb + s * STAGE, sw = sa + A_TILE;
#pragma unroll
        for (int ks = 0; ks < KB / 16; ++ks) tc_mma_f16(tm, sw_desc<KB>(sa + ks * 32), sw_desc<KB>(sw + ks * 32), idesc, (q | ks) ? 1u : 0u);
        tc_commit(empty0 + 8 * s);
        if (q == nk - 1) tc_commit(acc_full);
      }
      __syncwarp();
    }
    tc_fence_before();
  } else {
    // ===================== epilogue: thread = padded raster row =====================
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const long long R = (long long)p.B * Hp * Wp;
    const int r = r0 + warp * 32 + lane;
    const bool in_range = r < R;
    const int b = in_range ? r / (Hp * Wp) : 0, rem = r - b * Hp * Wp, yp = rem / Wp, xp = rem - yp * Wp;
    const bool inside = in_range && yp >= 1 && yp <= p.H && xp >= 1 && xp <= p.W;
    const int HW = p.H * p.W, m = (yp - 1) * p.W + (xp - 1);
    const uint32_t tl = tm + ((uint32_t)(warp * 32) << 16);
    auto chunk = [&](int c0, const uint32_t* acc, int n) {  // n = 16 or 32 channels starting at n0 + c0
      if (p.taps == 4) {  // transposed convolution: scatter the four phases into the (2 H, 2 W) fp32 tensor
        if (inside) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (i >= n) break;
            const int col = n0 + c0 + i, ph = col / p.up_cout, co = col - ph * p.up_cout;
            float o = __uint_as_float(acc[i]) * (p.scale ? __ldg(p.scale + co) : 1.f) + (p.shift ? __ldg(p.shift + co) : 0.f);
            if (p.relu) o = fmaxf(o, 0.f);
            p.out32[(((size_t)b * p.up_cout + co) * (2 * p.H) + 2 * (yp - 1) + (ph >> 1)) * (2 * p.W) + 2 * (xp - 1) + (ph & 1)] = o;
          }
        }
        return;
      }
      if (p.part) {
        if (inside) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (i >= n) break;
            p.part[(((size_t)sp * p.B + b) * p.Cout + n0 + c0 + i) * HW + m] = __uint_as_float(acc[i]);
          }
        }
        return;
      }
      float v[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (i >= n) break;
        const int co = n0 + c0 + i;
        float o = __uint_as_float(acc[i]) * (p.scale ? __ldg(p.scale + co) : 1.f) + (p.shift ? __ldg(p.shift + co) : 0.f);
        if (p.relu) o = fmaxf(o, 0.f);
        if (inside) {
          const size_t oi = ((size_t)b * p.Cout + co) * HW + m;
          if (p.res) o += p.res[oi];
          if (p.out32) p.out32[oi] = o;
        }
        v[i] = inside ? o : 0.f;
      }
      if (p.out16 && in_range) {
        __half* d = p.out16 + (size_t)r * p.Cout + n0 + c0;
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          if (i >= n) break;
          __half2 h0 = __floats2half2_rn(v[i], v[i + 1]), h1 = __floats2half2_rn(v[i + 2], v[i + 3]);
          __half2 h2 = __floats2half2_rn(v[i + 4], v[i + 5]), h3 = __floats2half2_rn(v[i + 6], v[i + 7]);
          *reinterpret_cast<uint4*>(d + i) = make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1),
                                                        *reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
        }
      }
    };
    if constexpr (BN >= 32) {
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t acc[32];
        tc_ld32(tl + c * 32, acc);
        chunk(c * 32, acc, 32);
      }
    } else {
      uint32_t acc[32];
      uint32_t a16[16];
      tc_ld16(tl, a16);
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = a16[i];
      chunk(0, acc, 16);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(TM_COLS) : "memory");
  }
}

// (B, C0 [+ C1], H, W) fp32 -> padded channels-last fp16 (B, H + 2, W + 2, C0 + C1), zero border.  One thread = one padded pixel x
// 8 channels (reads coalesced along x, one 16-byte store).
__global__ void rmvpe_to_nhwc16_kernel(const float* __restrict__ in0, const float* __restrict__ in1, int C0, int C, int B, int H, int W,
                                       __half* __restrict__ out) {
  const int Hp = H + 2, Wp = W + 2, groups = C / 8;
  const long long total = (long long)B * Hp * Wp * groups;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  // pixel fastest within a channel group keeps the reads of a warp on consecutive x
  const long long pix = i % ((long long)B * Hp * Wp);
  const int g = (int)(i / ((long long)B * Hp * Wp));
  const int b = (int)(pix / (Hp * Wp)), rem = (int)(pix - (long long)b * Hp * Wp), yp = rem / Wp, xp = rem - yp * Wp;
  uint4 u = make_uint4(0, 0, 0, 0);
  if (yp >= 1 && yp <= H && xp >= 1 && xp <= W) {
    const int c = g * 8, HW = H * W;
    const float* src = (c < C0 ? in0 + ((size_t)b * C0 + c) * HW : in1 + ((size_t)b * (C - C0) + (c - C0)) * HW) + (yp - 1) * W + (xp - 1);
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = __ldg(src + (size_t)e * HW);
    __half2 h0 = __floats2half2_rn(v[0], v[1]), h1 = __floats2half2_rn(v[2], v[3]), h2 = __floats2half2_rn(v[4], v[5]), h3 = __floats2half2_rn(v[6], v[7]);
    u = make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1), *reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
  }
  *reinterpret_cast<uint4*>(out + (size_t)pix * C + g * 8) = u;
}
cudaError_t launch_rmvpe_to_nhwc16(const float* in0, const float* in1, int C0, int C, int B, int H, int W, __half* out, cudaStream_t s) {
  if (C % 8 || C0 % 8) return cudaErrorInvalidValue;
  const long long total = (long long)B * (H + 2) * (W + 2) * (C / 8);
  rmvpe_to_nhwc16_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(in0, in1, C0, C, B, H, W, out);
  return cudaGetLastError();
}

// split-K reduction in a fixed order + the epilogue: one thread = one padded pixel x 8 output channels
__global__ void conv3x3_tma_finish_kernel(const float* __restrict__ part, int ksplit, const float* __restrict__ scale, const float* __restrict__ shift,
                                          const float* __restrict__ res, float* __restrict__ out32, __half* __restrict__ out16, int B, int Cout, int H,
                                          int W, int relu) {
  const int Hp = H + 2, Wp = W + 2, groups = Cout / 8, HW = H * W;
  const long long npix = (long long)B * Hp * Wp;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix * groups) return;
  const long long pix = i % npix;
  const int g = (int)(i / npix);
  const int b = (int)(pix / (Hp * Wp)), rem = (int)(pix - (long long)b * Hp * Wp), yp = rem / Wp, xp = rem - yp * Wp;
  float v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) v[e] = 0.f;
  if (yp >= 1 && yp <= H && xp >= 1 && xp <= W) {
    const int m = (yp - 1) * W + (xp - 1);
    const size_t n = (size_t)B * Cout * HW;
    const size_t oi0 = ((size_t)b * Cout + g * 8) * HW + m;
    float a8[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) a8[e] = 0.f;
    for (int s = 0; s < ksplit; ++s) {  // splits outermost: the eight channels' loads of a split are in flight together
      const float* ps = part + (size_t)s * n + oi0;
#pragma unroll
      for (int e = 0; e < 8; ++e) a8[e] += __ldg(ps + (size_t)e * HW);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int co = g * 8 + e;
      const size_t oi = oi0 + (size_t)e * HW;
      const float a = a8[e];
      float o = a * (scale ? __ldg(scale + co) : 1.f) + (shift ? __ldg(shift + co) : 0.f);
      if (relu) o = fmaxf(o, 0.f);
      if (res) o += res[oi];
      if (out32) out32[oi] = o;
      v[e] = o;
    }
  }
  if (out16) {
    __half2 h0 = __floats2half2_rn(v[0], v[1]), h1 = __floats2half2_rn(v[2], v[3]), h2 = __floats2half2_rn(v[4], v[5]), h3 = __floats2half2_rn(v[6], v[7]);
    *reinterpret_cast<uint4*>(out16 + (size_t)pix * Cout + g * 8) =
        make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1), *reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
  }
}

bool conv3x3_tma_supported(int Cin, int Cout) {
  return (Cin == 16 || Cin == 32 || Cin % 64 == 0) && (Cout == 16 || Cout == 32 || Cout == 64 || Cout % 128 == 0);
}

// w (Cout, Cin, k, k) fp32 on the host, k * k = taps (9 or 1) -> [tap][Cout][Cin] fp16
void conv3x3_tma_pack_host(const float* w, int Cin, int Cout, int taps, __half* img) {
  for (int tap = 0; tap < taps; ++tap)
    for (int co = 0; co < Cout; ++co)
      for (int ci = 0; ci < Cin; ++ci) img[((size_t)tap * Cout + co) * Cin + ci] = __float2half_rn(w[((size_t)co * Cin + ci) * taps + tap]);
}

// ConvTranspose2d weight (Cin, Cout, 3, 3), stride 2, padding 1 -> [tap 4][phase 4 x Cout][Cin] fp16: tap = 2 oy + ox reads input pixel
// (y + oy, x + ox); phase = 2 py + px writes output pixel (2 y + py, 2 x + px); kernel index k = p + 1 - 2 o (zero block if k < 0)
void convt2x_tma_pack_host(const float* w, int Cin, int Cout, __half* img) {
  for (int tap = 0; tap < 4; ++tap)
    for (int ph = 0; ph < 4; ++ph) {
      const int ky = (ph >> 1) + 1 - 2 * (tap >> 1), kx = (ph & 1) + 1 - 2 * (tap & 1);
      for (int co = 0; co < Cout; ++co)
        for (int ci = 0; ci < Cin; ++ci)
          img[(((size_t)tap * 4 + ph) * Cout + co) * Cin + ci] =
              __float2half_rn((ky < 0 || kx < 0) ? 0.f : w[(((size_t)ci * Cout + co) * 3 + ky) * 3 + kx]);
    }
}

namespace {
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeFn conv_encode_fn() {
  static EncodeFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) f = nullptr;
    return (EncodeFn)f;
  }();
  return fn;
}
bool map2d(const void* ptr, uint64_t inner, uint64_t rows, uint32_t kb, uint32_t box_rows, CUtensorMap* out) {
  EncodeFn fn = conv_encode_fn();
  if (!fn) return false;
  const cuuint64_t dims[2] = {inner, rows}, strides[1] = {inner * 2};
  const cuuint32_t box[2] = {kb, box_rows}, es[2] = {1, 1};
  const CUtensorMapSwizzle sw = kb == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (kb == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  return fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
template <int BN, int KB>
cudaError_t launch_n(ConvTmaParams& p, cudaStream_t s) {
  using namespace ctm;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && !attr_set[dev]) {
    e = cudaFuncSetAttribute(conv3x3_tma_kernel<BN, KB>, cudaFuncAttributeMaxDynamicSharedMemorySize, Geo<BN, KB>::SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  const long long R = (long long)p.B * (p.H + 2) * (p.W + 2);
  conv3x3_tma_kernel<BN, KB><<<dim3(p.Cout / BN, (unsigned)((R + BM - 1) / BM), p.ksplit), THREADS, Geo<BN, KB>::SMEM, s>>>(p);
  return cudaGetLastError();
}
}  // namespace

// in16: padded channels-last input (B, H + 2, W + 2, Cin); wimg: [9][Cout][Cin] fp16.  cudaErrorNotSupported: no tensor map -> caller falls back.
// part / part_floats: scratch for split-K (used when the grid would leave most SMs idle: the deep levels), or null.
cudaError_t launch_conv3x3_tma(const __half* in16, const __half* wimg, const float* scale, const float* shift, const float* res, float* out32,
                               __half* out16, int B, int Cin, int Cout, int H, int W, int relu, int taps, int up_cout, float* part,
                               size_t part_floats, int* launches, cudaStream_t s) {
  if (!conv3x3_tma_supported(Cin, Cout) || (!out32 && !out16) || (taps != 9 && taps != 1 && taps != 4)) return cudaErrorInvalidValue;
  if (taps == 4 && (Cout != 4 * up_cout || !out32 || out16 || res)) return cudaErrorInvalidValue;
  if (taps == 4) part = nullptr;
  ConvTmaParams p;
  const int BN = Cout >= 128 ? 128 : Cout;
  const long long R = (long long)B * (H + 2) * (W + 2);
  const long long ctas = (long long)(Cout / BN) * ((R + ctm::BM - 1) / ctm::BM);
  const int KBv = Cin >= 64 ? 64 : Cin;
  const int nk = taps * Cin / KBv;
  int sp = 1;
  if (part && ctas < 74 && nk >= 8) {
    sp = (int)((148 + ctas - 1) / ctas);
    if (sp > nk / 4) sp = nk / 4;
    if (sp > 16) sp = 16;
    if (sp < 2 || (size_t)sp * B * Cout * H * W > part_floats) sp = 1;
  }
  if (!map2d(in16, (uint64_t)Cin, (uint64_t)B * (H + 2) * (W + 2), (uint32_t)KBv, ctm::BM, &p.tmX) ||
      !map2d(wimg, (uint64_t)Cin, (uint64_t)taps * Cout, (uint32_t)KBv, (uint32_t)BN, &p.tmW))
    return cudaErrorNotSupported;
  p.scale = scale; p.shift = shift; p.res = res; p.out32 = out32; p.out16 = out16;
  p.B = B; p.Cin = Cin; p.Cout = Cout; p.H = H; p.W = W; p.relu = relu; p.taps = taps; p.up_cout = up_cout;
  p.part = sp > 1 ? part : nullptr; p.ksplit = sp;
  cudaError_t e = cudaErrorInvalidValue;
  switch (BN * 100 + KBv) {
    case 12864: e = launch_n<128, 64>(p, s); break;
    case 6464: e = launch_n<64, 64>(p, s); break;
    case 6432: e = launch_n<64, 32>(p, s); break;
    case 3264: e = launch_n<32, 64>(p, s); break;
    case 3232: e = launch_n<32, 32>(p, s); break;
    case 3216: e = launch_n<32, 16>(p, s); break;
    case 1632: e = launch_n<16, 32>(p, s); break;
    case 1616: e = launch_n<16, 16>(p, s); break;
  }
  if (launches) *launches = 1;
  if (e != cudaSuccess || sp == 1) return e;
  const long long total = R * (Cout / 8);
  conv3x3_tma_finish_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(part, sp, scale, shift, res, out32, out16, B, Cout, H, W, relu);
  if (launches) *launches = 2;
  return cudaGetLastError();
}

}  // namespace rvcdec
