// TMA-fed tensor-core GEMM of the HuBERT embedder's fast mode (same contract as gemm_tc.cu):
//   C[m, n] = epi(sum_k A[m * lda + k] * W[n * K + k] + bias[n]) (+ R[m * ldc + n]),  A, W fp16, accumulation fp32 in TMEM.
// gemm_tc.cu gathers its tiles with cp.async and runs at ~1.2 us per K step (profiles/r2ar_*): 2048 16-byte requests, a
// proxy fence and 128 barrier arrivals per stage.  Here both operands are described by 2-D tensor maps (row pitch = lda, so
// the overlapping rows of the strided convolutions are just a strided view) and ONE thread issues two
// cp.async.bulk.tensor loads per stage: 128 rows x 64 halves with the 128-byte swizzle, which is the canonical K-major
// operand layout of tcgen05.mma (descriptor: swizzle mode 2, stride between 8-row groups 1024 B; a K = 16 step advances the
// start address by 32 B).  Rows beyond M are zero-filled by the TMA unit.  Warps: 0-3 epilogue (one accumulator row per
// thread, as in gemm_tc.cu), 4 = TMA producer, 5 = MMA issuer; three 32 KB stages, two CTAs per SM.
// Batched calls (the feature extractor with B > 1) fold the items into the row coordinate, which needs strideA % lda == 0.
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdint>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

cudaError_t launch_gemm_tc(const __half* A, size_t lda, size_t strideA, const __half* W, const float* bias, const float* R, float* C32, __half* C16,
                           size_t ldc, size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s);

namespace {
constexpr int TM_M = 128, TM_N = 128, TM_K = 64, TM_STAGES = 3;
constexpr int TM_TILE = TM_M * TM_K * 2;  // 16 KB, rows of 128 B
constexpr int TM_THREADS = 192;
constexpr int TM_SMEM = 1024 + TM_STAGES * 2 * TM_TILE + 64 + 16;

__device__ __forceinline__ float gelu_erf_tm(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }
// K-major operand tile written by TMA with CU_TENSOR_MAP_SWIZZLE_128B: 8-row x 128-byte atoms, 1024 B apart
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
               "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar)
               : "memory");
}
}  // namespace

struct GemmTmaParams {
  CUtensorMap tmA, tmW;
  const float* bias; const float* R;
  float* C32; __half* C16; size_t ldc, strideC;
  int M, N, K, act, rows_per_item;  // rows_per_item: strideA / lda (row coordinate of item z = z * rows_per_item)
};

template <int BN>
__global__ void __launch_bounds__(TM_THREADS, 2) gemm_tma_kernel(const __grid_constant__ GemmTmaParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sg = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sb = smem_u32(sg);
  const uint32_t OFF_BAR = TM_STAGES * 2 * TM_TILE;
  const uint32_t full0 = sb + OFF_BAR, empty0 = full0 + 8 * TM_STAGES, acc_full = empty0 + 8 * TM_STAGES, tmem_slot = acc_full + 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sg + OFF_BAR + 8 * (2 * TM_STAGES + 1));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * TM_M, n0 = blockIdx.x * BN;
  const int nk = p.K / TM_K;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TM_STAGES; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot_ptr;

  if (warp == 4) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      const int row0 = (int)blockIdx.z * p.rows_per_item + m0;
      for (int kt = 0; kt < nk; ++kt) {
        const int s = kt % TM_STAGES;
        if (kt >= TM_STAGES) mbar_wait(empty0 + 8 * s, ((kt / TM_STAGES) - 1) & 1);
        const uint32_t sa = sb + s * 2 * TM_TILE, sw = sa + TM_TILE;
        mbar_expect_tx(full0 + 8 * s, TM_TILE + BN * TM_K * 2);
        tma_load_2d(sa, &p.tmA, kt * TM_K, row0, full0 + 8 * s);
        tma_load_2d(sw, &p.tmW, kt * TM_K, n0, full0 + 8 * s);
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = make_idesc(ACT_F16, BN);
    for (int kt = 0; kt < nk; ++kt) {
      const int s = kt % TM_STAGES;
      mbar_wait(full0 + 8 * s, (kt / TM_STAGES) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = sb + s * 2 * TM_TILE, sw = sa + TM_TILE;
#pragma unroll
        for (int ks = 0; ks < TM_K / 16; ++ks)
          tc_mma_f16(tm, make_sw128_desc(sa + ks * 32), make_sw128_desc(sw + ks * 32), idesc, (kt | ks) ? 1u : 0u);
        tc_commit(empty0 + 8 * s);
        if (kt == nk - 1) tc_commit(acc_full);
      }
      __syncwarp();
    }
    tc_fence_before();
  } else {
    // ===================== epilogue: one accumulator row per thread =====================
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int row = m0 + warp * 32 + lane;
    const bool ok = row < p.M;
    const size_t zoff = blockIdx.z * p.strideC;
    const uint32_t tl = tm + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
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
        for (int i = 0; i < 32; ++i) v[i] = gelu_erf_tm(v[i]);
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
  }
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(128u) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Positional convolution (Conv1d 768 -> 768, k = 128, padding 64, 16 groups of 48, + bias, GELU; the last of the T + 1 outputs
// dropped) as 16 grouped GEMMs on the tensor cores: out[b, t, g 48 + co] = gelu(bias + sum_{j < 128} sum_{ci < 48}
// x[b, t + j - 64, g 48 + ci] w[g][j][co][ci]).  One CTA: 128 frames of one group; per tap one TMA box of the fp16 input
// (3-D map {768, T, B}: frames before 0 / after T - 1 and channels past 768 are zero-filled by the TMA unit; the box is 64
// channels wide, of which the MMAs use 48 = three K = 16 steps) and one 48 x 64 box of the packed weights; N = 48.
// ---------------------------------------------------------------------------------------------------------------------
namespace {
constexpr int PC_M = 128, PC_N = 48, PC_TAPS = 128, PC_STAGES = 4;
constexpr int PC_A = PC_M * 128, PC_W = PC_N * 128, PC_STAGE = PC_A + 6144 + 2048;  // W tile padded to keep stages 1024-aligned
static_assert(PC_W == 6144 && PC_STAGE % 1024 == 0, "stage layout");
constexpr int PC_THREADS = 192;
constexpr int PC_SMEM = 1024 + PC_STAGES * PC_STAGE + 64 + 16;
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
               "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
               : "memory");
}
}  // namespace

struct PosconvTmaParams {
  CUtensorMap tmX, tmW;
  const float* bias;
  float* out;
  int T, C;
};

__global__ void __launch_bounds__(PC_THREADS, 2) posconv_tma_kernel(const __grid_constant__ PosconvTmaParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sg = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sb = smem_u32(sg);
  const uint32_t OFF_BAR = PC_STAGES * PC_STAGE;
  const uint32_t full0 = sb + OFF_BAR, empty0 = full0 + 8 * PC_STAGES, acc_full = empty0 + 8 * PC_STAGES, tmem_slot = acc_full + 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sg + OFF_BAR + 8 * (2 * PC_STAGES + 1));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t0 = blockIdx.x * PC_M, g = blockIdx.y, b = blockIdx.z;

  if (threadIdx.x == 0) {
    for (int s = 0; s < PC_STAGES; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(64u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot_ptr;

  if (warp == 4) {
    if (lane == 0) {
      for (int j = 0; j < PC_TAPS; ++j) {
        const int s = j % PC_STAGES;
        if (j >= PC_STAGES) mbar_wait(empty0 + 8 * s, ((j / PC_STAGES) - 1) & 1);
        const uint32_t sa = sb + s * PC_STAGE, sw = sa + PC_A;
        mbar_expect_tx(full0 + 8 * s, PC_A + PC_W);
        tma_load_3d(sa, &p.tmX, g * PC_N, t0 + j - PC_TAPS / 2, b, full0 + 8 * s);
        tma_load_2d(sw, &p.tmW, 0, (g * PC_TAPS + j) * PC_N, full0 + 8 * s);
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    const uint32_t idesc = make_idesc(ACT_F16, PC_N);
    for (int j = 0; j < PC_TAPS; ++j) {
      const int s = j % PC_STAGES;
      mbar_wait(full0 + 8 * s, (j / PC_STAGES) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = sb + s * PC_STAGE, sw = sa + PC_A;
#pragma unroll
        for (int ks = 0; ks < PC_N / 16; ++ks)  // 48 of the box's 64 channels
          tc_mma_f16(tm, make_sw128_desc(sa + ks * 32), make_sw128_desc(sw + ks * 32), idesc, (j | ks) ? 1u : 0u);
        tc_commit(empty0 + 8 * s);
        if (j == PC_TAPS - 1) tc_commit(acc_full);
      }
      __syncwarp();
    }
    tc_fence_before();
  } else {
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int t = t0 + warp * 32 + lane;
    const uint32_t tl = tm + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < PC_N / 16; ++c) {
      uint32_t r[16];
      tc_ld16(tl + c * 16, r);
      if (t >= p.T) continue;
      const int n = g * PC_N + c * 16;
      float* o = p.out + ((size_t)b * p.T + t) * p.C + n;
#pragma unroll
      for (int i = 0; i < 16; i += 4) {
        const float4 bb = __ldg(reinterpret_cast<const float4*>(p.bias + n + i));
        *reinterpret_cast<float4*>(o + i) = make_float4(gelu_erf_tm(__uint_as_float(r[i]) + bb.x), gelu_erf_tm(__uint_as_float(r[i + 1]) + bb.y),
                                                        gelu_erf_tm(__uint_as_float(r[i + 2]) + bb.z), gelu_erf_tm(__uint_as_float(r[i + 3]) + bb.w));
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(64u) : "memory");
  }
}

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) f = nullptr;
    return (EncodeTiledFn)f;
  }();
  return fn;
}
struct MapKey {
  const void* p; uint64_t pitch, rows, k;
  bool operator==(const MapKey& o) const { return p == o.p && pitch == o.pitch && rows == o.rows && k == o.k; }
};
struct MapHash {
  size_t operator()(const MapKey& k) const { return (size_t)((uintptr_t)k.p * 1315423911u ^ k.pitch * 2654435761u ^ k.rows * 97 ^ k.k); }
};
// fp16 matrix of `rows` rows of `k` halves, `pitch` halves apart (pitch < k = overlapping rows); box 64 x 128, 128-byte swizzle
bool tensor_map(const void* ptr, uint64_t pitch, uint64_t rows, uint64_t k, uint32_t box_rows, CUtensorMap* out) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapHash> cache;
  const MapKey key{ptr, pitch, rows, k * 1024 + box_rows};
  std::lock_guard<std::mutex> g(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return true; }
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t dims[2] = {k, rows}, strides[1] = {pitch * 2};
  const cuuint32_t box[2] = {TM_K, box_rows}, es[2] = {1, 1};
  CUtensorMap m;
  if (fn(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return true;
}
}  // namespace

// Same arguments as launch_gemm_tc; falls back to it when a tensor map cannot describe the operand.
cudaError_t launch_gemm_tma(const __half* A, size_t lda, size_t strideA, const __half* W, const float* bias, const float* R, float* C32, __half* C16,
                            size_t ldc, size_t strideC, int batch, int M, int N, int K, int act, cudaStream_t s) {
  if (K % TM_K || N % TM_N || lda % 8 || ldc % 8 || M <= 0 || (!C32 && !C16)) return cudaErrorInvalidValue;
  GemmTmaParams p;
  const bool foldable = batch == 1 || (strideA % lda == 0);
  const uint64_t rows = (uint64_t)(batch - 1) * (foldable && batch > 1 ? strideA / lda : 0) + (uint64_t)M;
  // 64-column tiles where 128-column ones would leave half the SMs without a CTA (N = 768 at M = 1499: 72 CTAs)
  const int bn = ((long long)(N / TM_N) * ((M + TM_M - 1) / TM_M) * batch < 100) ? 64 : TM_N;
  if (!foldable || !tensor_map(A, lda, rows, (uint64_t)K, TM_M, &p.tmA) || !tensor_map(W, (uint64_t)K, (uint64_t)N, (uint64_t)K, (uint32_t)bn, &p.tmW))
    return launch_gemm_tc(A, lda, strideA, W, bias, R, C32, C16, ldc, strideC, batch, M, N, K, act, s);
  static bool attr_set[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && !attr_set[dev]) {
    e = cudaFuncSetAttribute(gemm_tma_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, TM_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, TM_SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  p.bias = bias; p.R = R; p.C32 = C32; p.C16 = C16; p.ldc = ldc; p.strideC = strideC;
  p.M = M; p.N = N; p.K = K; p.act = act; p.rows_per_item = batch > 1 ? (int)(strideA / lda) : 0;
  if (bn == 64) gemm_tma_kernel<64><<<dim3(N / 64, (M + TM_M - 1) / TM_M, batch), TM_THREADS, TM_SMEM, s>>>(p);
  else gemm_tma_kernel<128><<<dim3(N / TM_N, (M + TM_M - 1) / TM_M, batch), TM_THREADS, TM_SMEM, s>>>(p);
  return cudaGetLastError();
}

// x16 (B, T, 768) fp16, wimg [16 groups][128 taps][48 co][64 ci] fp16 (ci >= 48 unused) -> out (B, T, 768) fp32 = gelu(conv + bias).
// Returns cudaErrorNotSupported when the tensor maps cannot be built (the caller then runs hubert_posconv_kernel).
cudaError_t launch_posconv_tma(const __half* x16, const __half* wimg, const float* bias, float* out, int B, int T, int C, cudaStream_t s) {
  if (C != 16 * PC_N) return cudaErrorInvalidValue;
  EncodeTiledFn fn = encode_fn();
  if (!fn) return cudaErrorNotSupported;
  PosconvTmaParams p;
  {
    const cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)T, (cuuint64_t)B}, strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)T * C * 2};
    const cuuint32_t box[3] = {64, PC_M, 1}, es[3] = {1, 1, 1};
    if (fn(&p.tmX, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<__half*>(x16), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return cudaErrorNotSupported;
    const cuuint64_t wd[2] = {64, (cuuint64_t)16 * PC_TAPS * PC_N}, ws[1] = {128};
    const cuuint32_t wb[2] = {64, PC_N}, we[2] = {1, 1};
    if (fn(&p.tmW, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half*>(wimg), wd, ws, wb, we, CU_TENSOR_MAP_INTERLEAVE_NONE,
           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return cudaErrorNotSupported;
  }
  static bool attr_set[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && !attr_set[dev]) {
    e = cudaFuncSetAttribute(posconv_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PC_SMEM);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  p.bias = bias; p.out = out; p.T = T; p.C = C;
  posconv_tma_kernel<<<dim3((T + PC_M - 1) / PC_M, 16, B), PC_THREADS, PC_SMEM, s>>>(p);
  return cudaGetLastError();
}

}  // namespace rvcdec
