// Tensor-core GEMM of the HuBERT embedder's fast mode (include/rvchubert.h, option "tc"):
//   C[m, n] = epi(sum_k A[m * lda + k] * W[n * K + k] + bias[n]) (+ R[m * ldc + n]),  A, W fp16, accumulation fp32 in TMEM.
// A is row-major with an arbitrary row pitch (the feature extractor's strided convolutions read k * 512 halves per frame at
// a pitch of s * 512, hubert_kernels.cu), which rules out one bulk copy per tile: four producer warps gather the 128 x 64
// tiles with 16-byte cp.async into the no-swizzle K-major core-matrix layout ([k chunk of 8][row][8 halves], the layout of
// conv_mma.cu's operands; descriptor SBO = 128 B, LBO = the chunk stride, padded by 16 B so that the eight lanes that copy
// one row's 128 contiguous bytes land in eight different bank groups), signalling an mbarrier per stage
// (cp.async.mbarrier.arrive.noinc); a fifth warp issues tcgen05.mma (M 128 x N 128 x K 16, four per stage) and frees each
// stage with tcgen05.commit; the producer warps then read the accumulator (tcgen05.ld, one row per thread) and apply
// bias / erf-GELU / residual, writing fp32 and / or fp16.  Three 32 KB stages, so two CTAs share an SM and one's epilogue
// overlaps the other's main loop.  Requires K % 64 == 0, N % 128 == 0, lda % 8 == 0.
#include <cuda_fp16.h>

#include <cstdint>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

namespace {
constexpr int TG_M = 128, TG_N = 128, TG_K = 64, TG_STAGES = 3;
constexpr int TG_CHUNK_STRIDE = TG_M * 16 + 16;   // bytes between consecutive 8-element K chunks of a tile (padded, see above)
constexpr int TG_TILE = 16640;                    // per operand per stage: 8 chunks x 2064 B, rounded up to 128 B
constexpr int TG_THREADS = 160;
constexpr int TG_SMEM = 128 + TG_STAGES * 2 * TG_TILE + 64 + 16;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ float gelu_erf_tc(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }
}  // namespace

struct GemmTcParams {
  const __half* A; size_t lda, strideA;
  const __half* W;
  const float* bias; const float* R;
  float* C32; __half* C16; size_t ldc, strideC;
  int M, N, K, act;
};

__global__ void __launch_bounds__(TG_THREADS, 2) gemm_tc_kernel(const GemmTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sg = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const uint32_t sb = smem_u32(sg);
  const uint32_t OFF_BAR = TG_STAGES * 2 * TG_TILE;
  const uint32_t full0 = sb + OFF_BAR, empty0 = full0 + 8 * TG_STAGES, acc_full = empty0 + 8 * TG_STAGES, tmem_slot = acc_full + 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sg + OFF_BAR + 8 * (2 * TG_STAGES + 1));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * TG_M, n0 = blockIdx.x * TG_N;
  const __half* A = p.A + blockIdx.z * p.strideA;
  const int nk = p.K / TG_K;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TG_STAGES; ++s) { mbar_init(full0 + 8 * s, 128); mbar_init(empty0 + 8 * s, 1); }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot_ptr;

  if (warp < 4) {
    // ===================== producers: gather the A and W tiles of every K step =====================
    // one warp instruction covers 4 rows x 8 chunks: a quarter warp (the unit a 16-byte-per-lane access is processed in)
    // reads one row's 128 contiguous bytes.  (With 8 rows x 4 chunks per instruction every quarter touched 8 half-used
    // sectors and the L1 tag stage, not the tensor pipe, set the pace: profiles/r2ar_*.)  This thread: chunk l % 8 of
    // rows warp * 32 + j * 4 + l / 8, j < 8.
    const int ch = lane & 7, rs = lane >> 3;
    const int row0 = warp * 32 + rs;
    const __half* ap = A + (size_t)(m0 + row0) * p.lda + ch * 8;
    const __half* wp = p.W + (size_t)(n0 + row0) * p.K + ch * 8;
    const size_t a4 = 4 * p.lda, w4 = (size_t)4 * p.K;
    const uint32_t d0 = (uint32_t)(ch * TG_CHUNK_STRIDE + row0 * 16);
    uint32_t amask = 0;  // bit j: row j of this thread exists
#pragma unroll
    for (int j = 0; j < 8; ++j) amask |= (m0 + row0 + j * 4 < p.M) ? (1u << j) : 0u;
    for (int kt = 0; kt < nk; ++kt) {
      const int s = kt % TG_STAGES;
      if (kt >= TG_STAGES) mbar_wait(empty0 + 8 * s, ((kt / TG_STAGES) - 1) & 1);
      const uint32_t sa = sb + s * 2 * TG_TILE + d0, sw = sa + TG_TILE;
      const int ko = kt * TG_K;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const bool ok = (amask >> j) & 1u;
        cp_async16(sa + j * 64, ok ? ap + j * a4 + ko : A, ok ? 16u : 0u);
        cp_async16(sw + j * 64, wp + j * w4 + ko, 16u);
      }
      cp_async_arrive_noinc(full0 + 8 * s);
    }
    // ===================== epilogue: one accumulator row per thread =====================
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int row = m0 + warp * 32 + lane;
    const bool ok = row < p.M;
    const size_t zoff = blockIdx.z * p.strideC;
    const uint32_t tl = tm + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < TG_N / 32; ++c) {
      uint32_t r[32];
      tc_ld32(tl + c * 32, r);
      if (!ok) continue;
      const int n = n0 + c * 32;
      const size_t o = zoff + (size_t)row * p.ldc + n;
      float v[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
      if (p.bias) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n + i));
          v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
        }
      }
      if (p.act == 1) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = gelu_erf_tc(v[i]);
      }
      if (p.R) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(p.R + o + i));
          v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
        }
      }
      if (p.C32) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(p.C32 + o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
      }
      if (p.C16) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          __half2 h0 = __floats2half2_rn(v[i], v[i + 1]), h1 = __floats2half2_rn(v[i + 2], v[i + 3]);
          __half2 h2 = __floats2half2_rn(v[i + 4], v[i + 5]), h3 = __floats2half2_rn(v[i + 6], v[i + 7]);
          uint4 u;
          u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
          u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
          *reinterpret_cast<uint4*>(p.C16 + o + i) = u;
        }
      }
    }
    tc_fence_before();
  } else {
    // ===================== MMA issuer =====================
    const uint32_t idesc = make_idesc(ACT_F16, TG_N);
    for (int kt = 0; kt < nk; ++kt) {
      const int s = kt % TG_STAGES;
      mbar_wait(full0 + 8 * s, (kt / TG_STAGES) & 1);
      fence_proxy_async();  // the tiles were written through the generic proxy (cp.async); the MMA reads through the async proxy
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = sb + s * 2 * TG_TILE, sw = sa + TG_TILE;
#pragma unroll
        for (int ks = 0; ks < TG_K / 16; ++ks)
          tc_mma_f16(tm, make_smem_desc(sa + 2 * ks * TG_CHUNK_STRIDE, TG_CHUNK_STRIDE, 128),
                     make_smem_desc(sw + 2 * ks * TG_CHUNK_STRIDE, TG_CHUNK_STRIDE, 128), idesc, (kt | ks) ? 1u : 0u);
        tc_commit(empty0 + 8 * s);
        if (kt == nk - 1) tc_commit(acc_full);
      }
      __syncwarp();
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(128u) : "memory");
  }
}

cudaError_t launch_gemm_tc(const __half* A, size_t lda, size_t strideA, const __half* W, const float* bias, const float* R, float* C32, __half* C16,
                           size_t ldc, size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s) {
  if (K % TG_K || N % TG_N || lda % 8 || ldc % 8 || M <= 0 || (!C32 && !C16)) return cudaErrorInvalidValue;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && !attr_set[dev]) {
    e = cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TG_SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  GemmTcParams p{A, lda, strideA, W, bias, R, C32, C16, ldc, strideC, M, N, K, act};
  gemm_tc_kernel<<<dim3(N / TG_N, (M + TG_M - 1) / TG_M, batch), TG_THREADS, TG_SMEM, s>>>(p);
  return cudaGetLastError();
}

}  // namespace rvcdec
