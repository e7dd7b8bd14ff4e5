// Windowed relative-position self-attention of the TextEncoder (attentions.py:89-125) on the tensor cores: the fast-mode
// twin of attention_kernel (synth_front.cu).  Scores run over ALL keys of an item (the window only limits the relative-
// position terms), so at T = 3000 frames the two batched GEMMs are 7 GFLOP per item, head and layer - on CUDA cores that
// was 80 % of the front end's time once its convolutions ran on the tensor cores (profiles/r2ae_synth_launches).
//
//   S[t, s] = (q_t . k_s + [|s - t| <= w] q_t . Ek[s - t + w]) / sqrt(96);   keys s >= len (or a query t >= len): -1e4
//   O[t]    = sum_s p[t, s] v_s + sum_{|s - t| <= w} p[t, s] Ev[s - t + w]
//
// One CTA per (128-query tile, head, item), flash-attention style over 128-key tiles:
//   TMEM  S[2] 128 x 128 fp32 (double-buffered scores), O 128 x 96 fp32 (running output), R 128 x 32 (q . Ek[j], j < 21)
//   smem  Q, K[2], V[2] tiles straight out of the 16-bit q|k|v tensor (blocked-8: Q and K are K-major operands, V is the
//         MN-major B operand of P V), P[2] probabilities as the 16-bit A operand, Pb[2] the 21 band probabilities per
//         query (A operand of the relative-value term: one more K = 32 MMA into the same accumulator, so the online-
//         softmax rescaling covers it), Ek / Ev as 16-bit B operands
// Warps: 0 = bulk-copy producer, 1 = MMA issuer, 2-5 = softmax (one query row per thread = its TMEM lane).
// The running output is rescaled in TMEM only when a row maximum of the warp moves (vote).
#include <algorithm>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

namespace attm {
constexpr int BQ = 128, BK = 128, KC = 96, W = 10, NR = 2 * W + 1;
constexpr int NSLH = KC / 8;               // slices per head
constexpr int T_SLICE = 128 * 16;          // bytes per slice of a 128-row 16-bit tile
constexpr int QKV_TILE = NSLH * T_SLICE;   // 24 KB
constexpr int P_TILE = (BK / 8) * T_SLICE; // 32 KB: [key block 16][query row 128][8 keys]
constexpr int PB_TILE = 4 * T_SLICE;       // 8 KB:  [band block 4][query row 128][8] (21 of 32 used)
constexpr int EK_BYTES = 32 * KC * 2;      // Ek as B of R = Q Ek^T: N = 32 rows (21 used) x K = 96, K-major: [slice 12][n 32][8]
constexpr int EV_BYTES = KC * 32 * 2;      // Ev as B of O += Pb Ev: N = 96 x K = 32 band, K-major: [band block 4][n 96][8]
constexpr int OFF_Q = 0, OFF_K = OFF_Q + QKV_TILE, OFF_V = OFF_K + 2 * QKV_TILE, OFF_P = OFF_V + 2 * QKV_TILE,
              OFF_PB = OFF_P + 2 * P_TILE, OFF_EK = OFF_PB + 2 * PB_TILE, OFF_EV = OFF_EK + EK_BYTES, OFF_RS = OFF_EV + EV_BYTES,
              OFF_BAR = OFF_RS + 128 * NR * 4;
constexpr int NUM_BARS = 12;
constexpr int SMEM_BYTES = 128 + OFF_BAR + NUM_BARS * 8 + 16;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
constexpr int THREADS = 192;
constexpr int TM_S = 0, TM_O = 256, TM_R = 384;  // TMEM columns
}  // namespace attm

struct AttMmaParams {
  const void* qkv16;     // blocked-8 16-bit, 3 * heads * 96 channels: [q | k | v]; pad rows zeroed by the caller
  float* out;            // blocked-8 fp32, heads * 96 channels
  void* out16;           // the same as 16-bit operand copy (conv_o reads it), or null
  const void* ek16;      // packed Ek image (EK_BYTES), format fmt
  const void* ev16;      // packed Ev image (EV_BYTES)
  const int* lengths;
  int B, T, heads, Lp, fmt;
};

namespace {
// instruction descriptor with B MN-major (bit 16): V tiles are [channel slice][key row][8 channels], i.e. the canonical
// no-swizzle MN-major layout ((8 ch, m) , (8 keys, k)) : ((1, SBO = slice stride), (16 B, LBO = 128 B))
__device__ __forceinline__ uint32_t idesc_bmn(int fmt, int n) { return make_idesc(fmt, n) | (1u << 16); }
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t apack2(float a, float b, int fmt) {
  if (fmt == ACT_BF16) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
}  // namespace

__global__ void __launch_bounds__(attm::THREADS, 1) att_mma_kernel(const __grid_constant__ AttMmaParams p) {
  using namespace attm;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sb = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* sg = smem_raw + (sb - smem_u32(smem_raw));
  const uint32_t bars = sb + OFF_BAR;
  const uint32_t q_full = bars, r_full = bars + 8, k_full = bars + 16, v_full = bars + 32, s_full = bars + 48, o_done = bars + 64,
                 p_ready = bars + 80;
  const uint32_t tmem_slot = bars + NUM_BARS * 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sg + OFF_BAR + NUM_BARS * 8);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, h = blockIdx.y, b = blockIdx.z;
  const int nsl3 = 3 * p.heads * NSLH;
  const int n_kt = (p.T + BK - 1) / BK;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1); mbar_init(r_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(k_full + 8 * i, 1); mbar_init(v_full + 8 * i, 1); mbar_init(s_full + 8 * i, 1); mbar_init(o_done + 8 * i, 1);
      mbar_init(p_ready + 8 * i, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // band tiles start as zeros (columns 21..31 of every row stay zero)
  for (int i = threadIdx.x; i < 2 * PB_TILE / 16; i += THREADS) reinterpret_cast<uint4*>(sg + OFF_PB)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async();
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_slot_ptr;
  const uint8_t* qkv = reinterpret_cast<const uint8_t*>(p.qkv16);
  // byte address of row t of slice sl (of the q / k / v part `which`) of this item and head
  auto row_ptr = [&](int which, int sl, int t) {
    return qkv + ((((size_t)b * nsl3 + (size_t)(which * p.heads + h) * NSLH + sl) * p.Lp) + PADL + t) * 16;
  };

  if (warp == 0) {
    // ===================== producer: Q, Ek, Ev once; K / V tiles double-buffered =====================
    // (rows beyond T up to the tile end are the tensor's zeroed pad rows: Lp covers T rounded up to 512.  Their scores
    //  are masked out and their probabilities are exact zeros.)
    if (lane == 0) {
      mbar_expect_tx(q_full, QKV_TILE + EK_BYTES + EV_BYTES);
      for (int sl = 0; sl < NSLH; ++sl) bulk_g2s(sb + OFF_Q + sl * T_SLICE, row_ptr(0, sl, q0), T_SLICE, q_full);
      bulk_g2s(sb + OFF_EK, p.ek16, EK_BYTES, q_full);
      bulk_g2s(sb + OFF_EV, p.ev16, EV_BYTES, q_full);
    }
    for (int j = 0; j < n_kt; ++j) {
      const int st = j & 1;
      if (j >= 2) mbar_wait(s_full + 8 * st, ((j - 2) >> 1) & 1);  // K stage free: S(j-2) has completed
      if (lane == 0) {
        mbar_expect_tx(k_full + 8 * st, QKV_TILE);
        for (int sl = 0; sl < NSLH; ++sl)
          bulk_g2s(sb + OFF_K + st * QKV_TILE + sl * T_SLICE, row_ptr(1, sl, j * BK), T_SLICE, k_full + 8 * st);
      }
      if (j >= 2) mbar_wait(o_done + 8 * st, ((j - 2) >> 1) & 1);  // V stage free: P V(j-2) has completed
      if (lane == 0) {
        mbar_expect_tx(v_full + 8 * st, QKV_TILE);
        for (int sl = 0; sl < NSLH; ++sl)
          bulk_g2s(sb + OFF_V + st * QKV_TILE + sl * T_SLICE, row_ptr(2, sl, j * BK), T_SLICE, v_full + 8 * st);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one elected lane) =====================
    if (elect_one()) {
      const uint32_t idS = make_idesc(p.fmt, BK), idR = make_idesc(p.fmt, 32), idO = idesc_bmn(p.fmt, KC), idB = make_idesc(p.fmt, KC);
      mbar_wait(q_full, 0);
      tc_fence_after();
      // R[t][j] = q_t . Ek[j]  (N = 32, 21 used)
      for (int ks = 0; ks < KC / 16; ++ks)
        tc_mma_f16(tm + TM_R, make_smem_desc(sb + OFF_Q + 2 * ks * T_SLICE, T_SLICE, 128),
                   make_smem_desc(sb + OFF_EK + 2 * ks * (32 * 16), 32 * 16, 128), idR, ks ? 1u : 0u);
      tc_commit(r_full);
      for (int j = 0; j <= n_kt; ++j) {
        if (j < n_kt) {  // S(j) = Q K(j)^T; its buffer was read out by softmax(j - 2), which P V(j - 2) waited for
          const int st = j & 1;
          mbar_wait(k_full + 8 * st, (j >> 1) & 1);
          tc_fence_after();
          for (int ks = 0; ks < KC / 16; ++ks)
            tc_mma_f16(tm + TM_S + st * BK, make_smem_desc(sb + OFF_Q + 2 * ks * T_SLICE, T_SLICE, 128),
                       make_smem_desc(sb + OFF_K + st * QKV_TILE + 2 * ks * T_SLICE, T_SLICE, 128), idS, ks ? 1u : 0u);
          tc_commit(s_full + 8 * st);
        }
        if (j >= 1) {  // O += P(j-1) V(j-1) (+ band probabilities x Ev where the key tile touches the band)
          const int i = j - 1, st = i & 1;
          mbar_wait(p_ready + 8 * st, (i >> 1) & 1);
          mbar_wait(v_full + 8 * st, (i >> 1) & 1);
          tc_fence_after();
          for (int ks = 0; ks < BK / 16; ++ks)
            tc_mma_f16(tm + TM_O, make_smem_desc(sb + OFF_P + st * P_TILE + 2 * ks * T_SLICE, T_SLICE, 128),
                       make_smem_desc(sb + OFF_V + st * QKV_TILE + ks * 256, 128, T_SLICE), idO, (i | ks) ? 1u : 0u);
          if (i * BK + BK - 1 >= q0 - W && i * BK <= q0 + BQ - 1 + W)
            for (int ks = 0; ks < 2; ++ks)
              tc_mma_f16(tm + TM_O, make_smem_desc(sb + OFF_PB + st * PB_TILE + 2 * ks * T_SLICE, T_SLICE, 128),
                         make_smem_desc(sb + OFF_EV + 2 * ks * (KC * 16), KC * 16, 128), idB, 1u);
          tc_commit(o_done + 8 * st);
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== softmax warps: thread = query row = TMEM lane =====================
    const int qd = warp & 3, r = qd * 32 + lane, t = q0 + r;
    const uint32_t tl = tm + ((uint32_t)(qd * 32) << 16);
    const int len = min(p.lengths[b], p.T);
    const float c2 = 1.4426950408889634f / sqrtf((float)KC);  // scores in the log2 domain
    const float masked = -1e4f * 1.4426950408889634f;          // masked_fill(mask == 0, -1e4), attentions.py:107
    float* Rs = reinterpret_cast<float*>(sg + OFF_RS) + r * NR;
    const bool row_dead = t >= len;
    {
      mbar_wait(r_full, 0);
      tc_fence_after();
      uint32_t rr[32];
      tc_ld32(tl + TM_R, rr);
#pragma unroll
      for (int i = 0; i < NR; ++i) Rs[i] = __uint_as_float(rr[i]);
    }
    float m_run = -3.0e38f, l_run = 0.f;
    for (int j = 0; j < n_kt; ++j) {
      const int st = j & 1, s0 = j * BK;
      const bool band_tile = s0 + BK - 1 >= q0 - W && s0 <= q0 + BQ - 1 + W;
      mbar_wait(s_full + 8 * st, (j >> 1) & 1);
      tc_fence_after();
      // pass 1: row maximum of this tile
      float mx = -3.0e38f;
#pragma unroll 1
      for (int c = 0; c < BK / 32; ++c) {
        uint32_t sc[32];
        tc_ld32(tl + TM_S + st * BK + c * 32, sc);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int s = s0 + c * 32 + i, d = s - t;
          float v = __uint_as_float(sc[i]);
          if (band_tile && (unsigned)(d + W) <= (unsigned)(2 * W)) v += Rs[d + W];
          v *= c2;
          if (s >= len || row_dead) v = masked;
          if (s < p.T) mx = fmaxf(mx, v);
        }
      }
      const float m_new = fmaxf(m_run, mx);
      const float corr = ex2(m_run - m_new);
      // the P / Pb buffers of this stage were last read by P V(j - 2)
      if (j >= 2) mbar_wait(o_done + 8 * st, ((j - 2) >> 1) & 1);
      uint8_t* Pt = sg + OFF_P + st * P_TILE + r * 16;
      uint8_t* Pb = sg + OFF_PB + st * PB_TILE + r * 16;
      if (band_tile) {
#pragma unroll
        for (int k = 0; k < 4; ++k) *reinterpret_cast<uint4*>(Pb + k * T_SLICE) = make_uint4(0, 0, 0, 0);
      }
      // pass 2: probabilities -> 16-bit A operand (and the band copy), row sum
      float ps = 0.f;
#pragma unroll 1
      for (int c = 0; c < BK / 32; ++c) {
        uint32_t sc[32];
        tc_ld32(tl + TM_S + st * BK + c * 32, sc);
        float pv[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int s = s0 + c * 32 + i, d = s - t;
          float v = __uint_as_float(sc[i]);
          if (band_tile && (unsigned)(d + W) <= (unsigned)(2 * W)) v += Rs[d + W];
          v *= c2;
          if (s >= len || row_dead) v = masked;
          const float e = (s < p.T) ? ex2(v - m_new) : 0.f;
          pv[i] = e;
          ps += e;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
          *reinterpret_cast<uint4*>(Pt + (c * 4 + k) * T_SLICE) =
              make_uint4(apack2(pv[8 * k], pv[8 * k + 1], p.fmt), apack2(pv[8 * k + 2], pv[8 * k + 3], p.fmt),
                         apack2(pv[8 * k + 4], pv[8 * k + 5], p.fmt), apack2(pv[8 * k + 6], pv[8 * k + 7], p.fmt));
        if (band_tile) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int bi = s0 + c * 32 + i - t + W;
            if ((unsigned)bi <= (unsigned)(2 * W)) {
              const uint32_t pk = apack2(pv[i], 0.f, p.fmt);
              *reinterpret_cast<uint16_t*>(Pb + (bi >> 3) * T_SLICE + (bi & 7) * 2) = (uint16_t)(pk & 0xffffu);
            }
          }
        }
      }
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
    // ---- normalise and store: channels h*96 .. of row t ----
    {
      const int last = n_kt - 1;
      mbar_wait(o_done + 8 * (last & 1), (last >> 1) & 1);
      tc_fence_after();
      const float inv = 1.f / l_run;
      const int nslo = p.heads * NSLH;
#pragma unroll 1
      for (int c = 0; c < NSLH; ++c) {
        uint32_t o8[8];
        tc_ld8(tl + TM_O + c * 8, o8);
        if (t < p.T) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(o8[i]) * inv;
          const size_t off = ((((size_t)b * nslo + (size_t)h * NSLH + c) * p.Lp) + PADL + t) * 8;
          *reinterpret_cast<float4*>(p.out + off) = make_float4(v[0], v[1], v[2], v[3]);
          *reinterpret_cast<float4*>(p.out + off + 4) = make_float4(v[4], v[5], v[6], v[7]);
          if (p.out16)
            *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out16) + off) =
                make_uint4(apack2(v[0], v[1], p.fmt), apack2(v[2], v[3], p.fmt), apack2(v[4], v[5], p.fmt), apack2(v[6], v[7], p.fmt));
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

// Ek / Ev ((2w+1) x 96 fp32) -> the 16-bit B operand images the kernel copies into shared memory
//   ek_img: [slice 12][n 32][8]     ek_img[sl][n][e] = Ek[n][sl*8 + e]   (n < 21, else 0)
//   ev_img: [band block 4][n 96][8] ev_img[kb][n][e] = Ev[kb*8 + e][n]   (band index < 21, else 0)
__global__ void att_pack_rel_kernel(const float* __restrict__ ek, const float* __restrict__ ev, uint16_t* __restrict__ ek_img,
                                    uint16_t* __restrict__ ev_img, int fmt) {
  using namespace attm;
  for (int i = threadIdx.x; i < 32 * KC; i += blockDim.x) {
    const int e = i & 7, n = (i >> 3) & 31, sl = i >> 8;
    const float v = n < NR ? ek[n * KC + sl * 8 + e] : 0.f;
    ek_img[i] = (uint16_t)(apack2(v, 0.f, fmt) & 0xffffu);
  }
  for (int i = threadIdx.x; i < KC * 32; i += blockDim.x) {
    const int e = i & 7, n = (i >> 3) % KC, kb = (i >> 3) / KC;
    const int bi = kb * 8 + e;
    const float v = bi < NR ? ev[bi * KC + n] : 0.f;
    ev_img[i] = (uint16_t)(apack2(v, 0.f, fmt) & 0xffffu);
  }
}
// both images of one layer: ek_img then ev_img, att_mma_rel_bytes() each
cudaError_t launch_att_pack_rel(const float* ek, const float* ev, void* ek_img, void* ev_img, int fmt, cudaStream_t s) {
  att_pack_rel_kernel<<<1, 256, 0, s>>>(ek, ev, (uint16_t*)ek_img, (uint16_t*)ev_img, fmt);
  return cudaGetLastError();
}
int att_mma_rel_bytes() { return attm::EK_BYTES; }

cudaError_t launch_attention_mma(const void* qkv16, float* out, void* out16, const void* ek_img, const void* ev_img, const int* lengths,
                                 int B, int T, int heads, int kc, int window, int Lp, int fmt, cudaStream_t s) {
  using namespace attm;
  if (kc != KC || window != W) return cudaErrorInvalidValue;
  static unsigned long long done = 0;
  cudaError_t e = set_smem_once(reinterpret_cast<const void*>(att_mma_kernel), SMEM_BYTES, done);
  if (e != cudaSuccess) return e;
  AttMmaParams p{qkv16, out, out16, ek_img, ev_img, lengths, B, T, heads, Lp, fmt};
  dim3 grid((T + BQ - 1) / BQ, heads, B);
  att_mma_kernel<<<grid, THREADS, SMEM_BYTES, s>>>(p);
  return cudaGetLastError();
}

}  // namespace rvcdec
