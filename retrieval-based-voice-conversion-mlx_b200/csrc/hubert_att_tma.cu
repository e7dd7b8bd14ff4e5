// Self-attention of the HuBERT encoder layers on the tensor cores, TMA-fed: hubert_att_tc.cu with the cp.async gathers replaced by
// cp.async.bulk.tensor boxes of a 3-D map {2304, T, B} over the fp16 q | k | v rows (128-byte swizzle; frames past T come back as
// zeros).  Q and K are K-major operands, V is the MN-major B operand of P V in the layout the TMA unit writes (keys x 64
// channels, 128-byte rows); P stays a no-swizzle tile written by the softmax threads.
//   qkv16 (B * T, 3 * 768) fp16 rows [q | k | v] (gemm_tc.cu's output), 12 heads of 64;
//   ctx16[b, t, h 64 + d] = softmax_s(q_t . k_s / 8) v_s   as fp16 rows (the A operand of the out_proj GEMM).
// One CTA per (128-query tile, head, item), flash-attention style over 128-key tiles:
//   TMEM  S[2] 128 x 128 fp32 (double-buffered scores), O 128 x 64 fp32 (running output)
//   smem  Q, K[2], V[2] tiles gathered with 16-byte cp.async from the row-major tensor into [8-channel chunk][row][8]
//         (Q and K: K-major operands; V: the MN-major B operand of P V), P[2] probabilities as the fp16 A operand
// Warps: 0 = producer, 1 = MMA issuer, 2-5 = softmax (one query row per thread = its TMEM lane); exp2 domain; the running
// output is rescaled in TMEM only when a row maximum of the warp moves.
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

namespace hatm {
constexpr int BQ = 128, BK = 128, KC = 64;
constexpr int NCH = KC / 8;                // 8-channel chunks per head
constexpr int T_SLICE = 128 * 16;          // bytes per key block of the P tile (128 query rows x 8 keys)
constexpr int QKV_TILE = 128 * 128;        // 128 rows of 128 B (64 channels), 128-byte swizzle
constexpr int P_TILE = (BK / 8) * T_SLICE; // 32 KB: [key block 16][query row 128][8 keys]
constexpr int OFF_Q = 0, OFF_K = OFF_Q + QKV_TILE, OFF_V = OFF_K + 2 * QKV_TILE, OFF_P = OFF_V + 2 * QKV_TILE, OFF_BAR = OFF_P + 2 * P_TILE;
constexpr int NUM_BARS = 11;
constexpr int SMEM_BYTES = 1024 + OFF_BAR + NUM_BARS * 8 + 16;
constexpr int THREADS = 192;
constexpr int TM_S = 0, TM_O = 256;  // TMEM columns
}  // namespace hatt

namespace {
__device__ __forceinline__ uint64_t msw128(uint32_t addr) {  // K-major, 128-byte swizzle (gemm_tma.cu); LBO = SBO = 1024 also serves the MN-major V tile
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)(1024 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ void mtma3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
               "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
               : "memory");
}
__device__ __forceinline__ uint32_t midesc_bmn(int n) { return make_idesc(ACT_F16, n) | (1u << 16); }
__device__ __forceinline__ float mex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint64_t mu2(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ uint64_t mf2(float lo, float hi) { return mu2(__float_as_uint(lo), __float_as_uint(hi)); }
__device__ __forceinline__ uint32_t mpack2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
}  // namespace

__global__ void __launch_bounds__(hatm::THREADS, 1) hubert_att_tma_kernel(const __grid_constant__ CUtensorMap tmQKV, __half* __restrict__ ctx16, int T,
                                                                          int heads) {
  using namespace hatm;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sb = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sg = smem_raw + (sb - smem_u32(smem_raw));
  const uint32_t bars = sb + OFF_BAR;
  const uint32_t q_full = bars, k_full = bars + 8, v_full = bars + 24, s_full = bars + 40, o_done = bars + 56, p_ready = bars + 72;
  const uint32_t tmem_slot = bars + NUM_BARS * 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sg + OFF_BAR + NUM_BARS * 8);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, h = blockIdx.y, b = blockIdx.z;
  const int C = heads * KC;
  const int n_kt = (T + BK - 1) / BK;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(k_full + 8 * i, 1); mbar_init(v_full + 8 * i, 1); mbar_init(s_full + 8 * i, 1); mbar_init(o_done + 8 * i, 1);
      mbar_init(p_ready + 8 * i, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== producer: Q once; K / V tiles double-buffered =====================
    if (lane == 0) {
      mbar_expect_tx(q_full, QKV_TILE);
      mtma3d(sb + OFF_Q, &tmQKV, h * KC, q0, b, q_full);
      for (int j = 0; j < n_kt; ++j) {
        const int st = j & 1;
        if (j >= 2) mbar_wait(s_full + 8 * st, ((j - 2) >> 1) & 1);  // K stage free: S(j-2) has completed
        mbar_expect_tx(k_full + 8 * st, QKV_TILE);
        mtma3d(sb + OFF_K + st * QKV_TILE, &tmQKV, C + h * KC, j * BK, b, k_full + 8 * st);
        if (j >= 2) mbar_wait(o_done + 8 * st, ((j - 2) >> 1) & 1);  // V stage free: P V(j-2) has completed
        mbar_expect_tx(v_full + 8 * st, QKV_TILE);
        mtma3d(sb + OFF_V + st * QKV_TILE, &tmQKV, 2 * C + h * KC, j * BK, b, v_full + 8 * st);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer (one elected lane) =====================
    if (elect_one()) {
      const uint32_t idS = make_idesc(ACT_F16, BK), idO = midesc_bmn(KC);
      mbar_wait(q_full, 0);
      for (int j = 0; j <= n_kt; ++j) {
        if (j < n_kt) {  // S(j) = Q K(j)^T; its buffer was read out by softmax(j - 2), which P V(j - 2) waited for
          const int st = j & 1;
          mbar_wait(k_full + 8 * st, (j >> 1) & 1);
          tc_fence_after();
          for (int ks = 0; ks < KC / 16; ++ks)
            tc_mma_f16(tm + TM_S + st * BK, msw128(sb + OFF_Q + ks * 32), msw128(sb + OFF_K + st * QKV_TILE + ks * 32), idS, ks ? 1u : 0u);
          tc_commit(s_full + 8 * st);
        }
        if (j >= 1) {  // O += P(j-1) V(j-1)
          const int i = j - 1, st = i & 1;
          mbar_wait(p_ready + 8 * st, (i >> 1) & 1);
          mbar_wait(v_full + 8 * st, (i >> 1) & 1);
          tc_fence_after();
          for (int ks = 0; ks < BK / 16; ++ks)  // 16 keys = two 8-row swizzle atoms of the V tile = 2048 B
            tc_mma_f16(tm + TM_O, make_smem_desc(sb + OFF_P + st * P_TILE + 2 * ks * T_SLICE, T_SLICE, 128),
                       msw128(sb + OFF_V + st * QKV_TILE + ks * 2048), idO, (i | ks) ? 1u : 0u);
          tc_commit(o_done + 8 * st);
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== softmax warps: thread = query row = TMEM lane =====================
    const int qd = warp & 3, r = qd * 32 + lane, t = q0 + r;
    const uint32_t tl = tm + ((uint32_t)(qd * 32) << 16);
    const float c2 = 1.4426950408889634f * 0.125f;  // scores / sqrt(64) in the log2 domain
    float m_run = -3.0e38f, l_run = 0.f;
    for (int j = 0; j < n_kt; ++j) {
      const int st = j & 1, s0 = j * BK;
      mbar_wait(s_full + 8 * st, (j >> 1) & 1);
      tc_fence_after();
      // Full key tiles (all but the last of a clip) take the short path: the maximum of the RAW scores (the scale is positive),
      // then exponent arguments and row sums in packed fp32x2 arithmetic - the softmax warps, not the tensor pipe, set this
      // kernel's pace, and their instruction count per score drops from ~4.5 to ~2.5.
      const bool full = s0 + BK <= T;
      float mx = -3.0e38f;
#pragma unroll 1
      for (int c = 0; c < BK / 32; ++c) {
        uint32_t sc[32];
        tc_ld32(tl + TM_S + st * BK + c * 32, sc);
        if (full) {
#pragma unroll
          for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(sc[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (s0 + c * 32 + i < T) mx = fmaxf(mx, __uint_as_float(sc[i]));
        }
      }
      mx *= c2;
      const float m_new = fmaxf(m_run, mx);
      const float corr = mex2(m_run - m_new);
      if (j >= 2) mbar_wait(o_done + 8 * st, ((j - 2) >> 1) & 1);  // the P buffer of this stage was last read by P V(j - 2)
      uint8_t* Pt = sg + OFF_P + st * P_TILE + r * 16;
      float ps = 0.f;
      const uint64_t c2p = mf2(c2, c2), mnp = mf2(-m_new, -m_new);
      uint64_t ps2 = 0ull;
#pragma unroll 1
      for (int c = 0; c < BK / 32; ++c) {
        uint32_t sc[32];
        tc_ld32(tl + TM_S + st * BK + c * 32, sc);
        float pv[32];
        if (full) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            uint64_t t;
            asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(t) : "l"(mu2(sc[i], sc[i + 1])), "l"(c2p), "l"(mnp));
            const float e0 = mex2(__uint_as_float((uint32_t)t)), e1 = mex2(__uint_as_float((uint32_t)(t >> 32)));
            pv[i] = e0; pv[i + 1] = e1;
            asm("add.rn.f32x2 %0, %0, %1;" : "+l"(ps2) : "l"(mf2(e0, e1)));
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float e = (s0 + c * 32 + i < T) ? mex2(fmaf(__uint_as_float(sc[i]), c2, -m_new)) : 0.f;
            pv[i] = e;
            ps += e;
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
          *reinterpret_cast<uint4*>(Pt + (c * 4 + k) * T_SLICE) = make_uint4(mpack2(pv[8 * k], pv[8 * k + 1]), mpack2(pv[8 * k + 2], pv[8 * k + 3]),
                                                                             mpack2(pv[8 * k + 4], pv[8 * k + 5]), mpack2(pv[8 * k + 6], pv[8 * k + 7]));
      }
      ps += __uint_as_float((uint32_t)ps2) + __uint_as_float((uint32_t)(ps2 >> 32));
      l_run = l_run * corr + ps;
      m_run = m_new;
      // rescale the running output only if a maximum of this warp's rows moved (P V(j - 1) must have landed first)
      if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {
        mbar_wait(o_done + 8 * ((j - 1) & 1), ((j - 1) >> 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < KC / 8; ++c) {
          uint32_t o8[8];
          tc_ld8(tl + TM_O + c * 8, o8);
#pragma unroll
          for (int i = 0; i < 8; ++i) o8[i] = __float_as_uint(__uint_as_float(o8[i]) * corr);
          tc_st8(tl + TM_O + c * 8, o8);
        }
        tc_wait_st();
      }
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_ready + 8 * st);
    }
    {  // normalise and store channels h * 64 .. of row t
      const int last = n_kt - 1;
      mbar_wait(o_done + 8 * (last & 1), (last >> 1) & 1);
      tc_fence_after();
      const float inv = 1.f / l_run;
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        uint32_t o8[8];
        tc_ld8(tl + TM_O + c * 8, o8);
        if (t < T) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(o8[i]) * inv;
          *reinterpret_cast<uint4*>(ctx16 + ((size_t)b * T + t) * C + h * KC + c * 8) =
              make_uint4(mpack2(v[0], v[1]), mpack2(v[2], v[3]), mpack2(v[4], v[5]), mpack2(v[6], v[7]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512u) : "memory");
  }
}

// cudaErrorNotSupported when the tensor map cannot be built: the caller runs hubert_att_tc.cu instead
cudaError_t launch_hubert_att_tma(const __half* qkv16, __half* ctx16, int B, int T, int heads, cudaStream_t s) {
  using namespace hatm;
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) f = nullptr;
    return (EncodeFn)f;
  }();
  if (!fn) return cudaErrorNotSupported;
  const int C3 = 3 * heads * KC;
  CUtensorMap tm;
  const cuuint64_t dims[3] = {(cuuint64_t)C3, (cuuint64_t)T, (cuuint64_t)B}, strides[2] = {(cuuint64_t)C3 * 2, (cuuint64_t)T * C3 * 2};
  const cuuint32_t box[3] = {KC, 128, 1}, es[3] = {1, 1, 1};
  if (fn(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<__half*>(qkv16), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return cudaErrorNotSupported;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && !attr_set[dev]) {
    e = cudaFuncSetAttribute(hubert_att_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  hubert_att_tma_kernel<<<dim3((T + BQ - 1) / BQ, heads, B), THREADS, SMEM_BYTES, s>>>(tm, ctx16, T, heads);
  return cudaGetLastError();
}

}  // namespace rvcdec
