// Fused ResBlock pair for the 128-channel stage on CTA PAIRS (cta_group::2), two tiles in flight:
//
//     x_out = x + conv2(lrelu(conv1_d(lrelu(x)) + b1)) + b2          (residuals.py:71-80, one iteration of the loop)
//
// mrf_pair.cu (one CTA, 256-row tile) halves the HBM traffic of the per-conv path and still measures the same: its TMEM
// and shared memory are full with ONE tile, so the tensor pipe idles through every epilogue.  Here a cluster
// of two CTAs shares one 256-row tile - CTA r owns rows [128 r, 128 r + 128) - and one tcgen05.mma.cta_group::2
// (M = 256) covers both halves while each CTA streams only HALF of every weight block (N/2 = 64 output channels).
// Per CTA that frees half of TMEM and half of the operand memory, which go to a SECOND tile in flight:
//
//   TMEM   per buffer p in {0,1}:  x[p] 128 columns (residual; conv2 accumulates into it), acc1[p] 128 columns
//   smem   A[p]   (128 + 2*32 pad rows) x 128 ch, blocked-8 K-major 16-bit; conv2's operand overwrites conv1's in place
//          W ring 7 x 16 KB: this CTA's half (64 of 128 output channels) of one tap's weight block, one bulk copy out
//                 of the half-split weight image (a stage feeds 8 MMAs = 512 clk).  Fetching the halves out of the
//                 conv_mma image as 1 KB pieces starved the tensor pipe: 27-33 % of its time went into weight waits.
//
// Schedule (j = tiles of this cluster in order, buffer p = j & 1):
//   tensor pipe   C1(0) C1(1) C2(0) C2(1) C1(2) C1(3) C2(2) ...
//   epilogue      L(0) L(1) | M(0) M(1) F(0)+L(2) F(1)+L(3) | M(2) ...
// so every epilogue phase (L = load x: TMEM residual + lrelu operand rows; M = conv1 accumulator -> conv2 operand;
// F = residual out of TMEM -> HBM) runs while the tensor pipe works on the OTHER buffer.
//
// Tiles: every CTA owns an independent 128-row tile; only the MMA instruction and the weight halves are shared by
// the pair.  conv1's input is the stage's fp32 stream in HBM, so a tile's pad rows (32 per side) are filled with the
// REAL neighbouring rows (zeros outside the sequence): conv1 is exact on all 128 rows and only conv2's halo
// (margin = (k2 - 1) / 2 * d2 <= 5 rows per side) is recomputed by the neighbouring tile - 92-98 % useful rows against
// 77-97 % for the 256-row tiles of mrf_pair.cu, whose margin is halo(conv1) + halo(conv2).  No row crosses from one CTA
// to the other, so nothing is exchanged through distributed shared memory.
//
// Cross-CTA protocol (validated in isolation by tools/umma_bench3.cu): epilogue warps of BOTH CTAs arrive on the
// LEADER's barriers (mapa + mbarrier.arrive on the shared::cluster address; CTA-scope release / acquire as in CUTLASS's
// ClusterBarrier - the cluster-scope forms compile to MEMBAR.ALL.GPU on the arriving and CCTL.IVALL on the waiting side
// and made the first version of this kernel 2.4x slower than the per-conv path); MMA completion is multicast to the barrier of the same name in both CTAs (tcgen05.commit ... multicast::cluster); the
// peer's weight stages are announced to the leader by a relay warp (a bulk copy can only signal a barrier of the CTA
// it lands in).
//
// Warps (576 threads per CTA): 0 = weight producer, 1 = MMA issuer (leader) / weight relay (peer), 2-17 = epilogue
// (4 column groups of 32 channels x 4 TMEM lane quarters).
#include <algorithm>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

namespace pair2 {
constexpr int EW = 16;
constexpr int THREADS = 32 * (EW + 2);
constexpr int C = 128;
constexpr int PAD = 32;
constexpr int ROWS = 128 + 2 * PAD;
constexpr int NSL = C / 8;
constexpr int A_SLICE = ROWS * 16;
constexpr int A_BUF = NSL * A_SLICE;       // 48 KB
constexpr int IMG_TAP = C * C * 2;         // one tap of the half-split image: [half 2][slice 16][n 64][8]
constexpr int BH_SLICE = (C / 2) * 16;     // this CTA's half of a K block (64 output channels)
constexpr int W_STAGE = (C / 8) * BH_SLICE;   // 16 KB: this CTA's half of one tap
constexpr int S = 7;
constexpr int MT = 128;                    // rows of a tile (one per CTA)
constexpr int NUM_BARS = 2 * S + 8;
constexpr int SMEM_BYTES = 128 + 2 * A_BUF + S * W_STAGE + NUM_BARS * 8 + 16 + 2 * C * 4;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
constexpr int NC = C / (EW / 4);  // 32 accumulator columns per epilogue warp
constexpr int NS8 = NC / 8;
}  // namespace pair2

namespace {
__device__ __forceinline__ uint32_t quni(uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); }
__device__ __forceinline__ uint32_t qpack2(float a, float b, bool bf16) {
  if (!bf16) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
// packed fp32x2 arithmetic (FADD2 / FMUL2): halves the epilogues' instruction count, which is what bounds them
__device__ __forceinline__ uint64_t q2_pack(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ uint64_t q2_packf(float lo, float hi) { return q2_pack(__float_as_uint(lo), __float_as_uint(hi)); }
__device__ __forceinline__ float2 q2_unpack(uint64_t v) {
  uint32_t lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
  return make_float2(__uint_as_float(lo), __uint_as_float(hi));
}
__device__ __forceinline__ uint64_t q2_add(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t q2_mul(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// 16-bit pair of leaky_relu(v) = max(v, slope * v), 0 < slope < 1
__device__ __forceinline__ uint32_t q2_lrelu_pack(uint64_t v, uint64_t slope2, bool bf16) {
  const float2 a = q2_unpack(v), m = q2_unpack(q2_mul(v, slope2));
  return qpack2(fmaxf(a.x, m.x), fmaxf(a.y, m.y), bf16);
}
}  // namespace

#define PT2(slot_, ...)                                    \
  do {                                                     \
    if (TIMING) {                                          \
      const long long _t = clock64();                      \
      __VA_ARGS__;                                         \
      tacc[slot_] += clock64() - _t;                       \
    } else {                                               \
      __VA_ARGS__;                                         \
    }                                                      \
  } while (0)

template <bool BF16, bool TIMING>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(pair2::THREADS, 1) mrf_pair2_kernel(const __grid_constant__ PairParams p) {
  using namespace pair2;
  long long tacc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  const long long t_begin = TIMING ? clock64() : 0;
  extern __shared__ uint8_t smem_raw[];
  // (the dynamic shared window starts at the same offset in both CTAs, so aligned offsets agree across the pair)
  const uint32_t smem_base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t a_smem = smem_base;  // A[0], A[1]
  const uint32_t w_smem = a_smem + 2 * A_BUF;
  const uint32_t bars = w_smem + S * W_STAGE;
  const uint32_t w_full = bars, w_empty = w_full + 8 * S, in_ready = w_empty + 8 * S, mid_ready = in_ready + 16,
                 acc1_full = mid_ready + 16, x_full = acc1_full + 16;
  const uint32_t tmem_slot = x_full + 16;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));
  // b1 | b2 in shared memory: the epilogues re-read them for every tile, and the x stream evicts them from L1
  float* bias_s = reinterpret_cast<float*>(smem_gen + (tmem_slot + 16 - smem_base));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();

  // operand tiles start as zeros: the OUTER pad rows are the convolutions' zero padding at the tile edges
  for (int i = threadIdx.x; i < 2 * A_BUF / 16; i += THREADS) reinterpret_cast<uint4*>(smem_gen)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x < 2 * C) bias_s[threadIdx.x] = threadIdx.x < C ? p.b1[threadIdx.x] : p.b2[threadIdx.x - C];
  if (threadIdx.x == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(w_full + 8 * i, rank == 0 ? 2 : 1);  // leader: own bulk copies + the relay's word for the peer's
      mbar_init(w_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(in_ready + 8 * i, 2 * EW);   // epilogue warps of both CTAs (only the leader's copy is used)
      mbar_init(mid_ready + 8 * i, 2 * EW);
      mbar_init(acc1_full + 8 * i, 1);
      mbar_init(x_full + 8 * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's barriers are initialised and its operand tiles zeroed before anything remote happens
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  const int total_work = p.B * p.tiles_per_item;  // tiles; cluster step j works on tiles 2 (cid + j ncl) + rank
  const int total_pairs = (total_work + 1) >> 1;
  const int ncl = (int)(gridDim.x >> 1), cid = (int)(blockIdx.x >> 1);
  const int n_tiles = cid < total_pairs ? (total_pairs - cid + ncl - 1) / ncl : 0;  // steps of this cluster (same in both CTAs)

  if (warp == 0) {
    // ===================== weight producer (both CTAs: this CTA's 64 output channels of every block) =====================
    uint32_t slot = 0, ph = 0;
    const int margin_p = p.margin, valid_p = MT - 2 * p.margin;
    // L2 prefetch of everything the epilogue warps will load for tile step j (x rows incl. pads, running MRF sum):
    // one bulk prefetch per 8-channel slice (rows of a slice are contiguous), lanes 0-15 one slice each.  Issued two
    // steps ahead; the epilogue's loads then hit L2.  (Per-row prefetch instructions in the epilogue warps cost
    // ~10 LSU instructions per thread and tile.)
    auto prefetch_step = [&](int j) {
      const int work = 2 * (cid + j * ncl) + (int)rank;
      if (work >= total_work || lane >= NSL) return;
      const int bb = work / p.tiles_per_item, tile = work - bb * p.tiles_per_item;
      const int t0 = tile * valid_p - margin_p;
      const size_t row0 = ((size_t)bb * NSL + (size_t)lane) * p.Lp + PADL;
      int a = t0 - PAD, b = t0 + MT + PAD;
      a = a < 0 ? 0 : a; b = b > p.L ? p.L : b;
      if (b > a) bulk_prefetch_l2(p.x_in + (row0 + a) * 8, (uint32_t)(b - a) * 32u);
      if (p.sum_in) {
        int sa = t0 + margin_p, sb = t0 + MT - margin_p;
        sa = sa < 0 ? 0 : sa; sb = sb > p.L ? p.L : sb;
        if (sb > sa) bulk_prefetch_l2(p.sum_in + (row0 + sa) * 8, (uint32_t)(sb - sa) * 32u);
      }
    };
    for (int jp = 0; jp < n_tiles; jp += 2) {
      const int cnt = (n_tiles - jp) < 2 ? (n_tiles - jp) : 2;
      if (jp == 0) { prefetch_step(0); if (n_tiles > 1) prefetch_step(1); }
      if (jp + 2 < n_tiles) prefetch_step(jp + 2);
      if (jp + 3 < n_tiles) prefetch_step(jp + 3);
#pragma unroll 1
      for (int which = 0; which < 2; ++which) {
        const uint8_t* src = reinterpret_cast<const uint8_t*>(which ? p.w2h : p.w1h) + rank * W_STAGE;
        const int nblk = which ? p.k2 : p.k1;
        for (int c = 0; c < cnt; ++c)
          for (int i = 0; i < nblk; ++i) {
            PT2(0, mbar_wait(w_empty + 8 * slot, ph ^ 1u));
            if (elect_one()) {
              mbar_expect_tx(w_full + 8 * slot, W_STAGE);
              bulk_g2s(w_smem + slot * W_STAGE, src + (size_t)i * IMG_TAP, W_STAGE, w_full + 8 * slot);
            }
            __syncwarp();
            if (++slot == S) { slot = 0; ph ^= 1u; }
          }
      }
    }
  } else if (warp == 1 && rank != 0) {
    // ===================== weight relay (peer): "my half of stage s has landed" -> the leader's w_full[s] =====================
    const uint32_t w_full_leader = mapa_u32(w_full, 0);
    uint32_t slot = 0, ph = 0;
    for (int jp = 0; jp < n_tiles; jp += 2) {
      const int cnt = (n_tiles - jp) < 2 ? (n_tiles - jp) : 2;
      const int nblk = cnt * (p.k1 + p.k2);
      for (int i = 0; i < nblk; ++i) {
        PT2(0, mbar_wait(w_full + 8 * slot, ph));
        if (lane == 0) mbar_arrive_remote(w_full_leader + 8 * slot);
        __syncwarp();
        if (++slot == S) { slot = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA; converged warp, elected lane issues) =====================
    const uint32_t idesc = make_idesc_m(p.fmt, C, 256);
    const uint64_t a_desc0 = make_smem_desc(a_smem, A_SLICE, 128) + (uint64_t)PAD;
    const uint64_t w_desc = make_smem_desc(w_smem, BH_SLICE, 128);
    const uint32_t leader = elect_one() ? 1u : 0u;
    const uint32_t a_hi = (uint32_t)(a_desc0 >> 32), w_lo = (uint32_t)w_desc, w_hi = (uint32_t)(w_desc >> 32);
    uint32_t slot = 0, ph = 0;
    for (int jp = 0; jp < n_tiles; jp += 2) {
      const int cnt = (n_tiles - jp) < 2 ? (n_tiles - jp) : 2;
      const uint32_t par = (uint32_t)(jp >> 1) & 1u;
#pragma unroll 1
      for (int which = 0; which < 2; ++which) {
        const int ntaps = (int)quni((uint32_t)(which ? p.k2 : p.k1)), dil = (int)quni((uint32_t)(which ? p.d2 : p.d1));
#pragma unroll 1
        for (int c = 0; c < cnt; ++c) {
          PT2(which, mbar_wait((which ? mid_ready : in_ready) + 8 * c, par));
          tc_fence_after();
          const uint32_t d_tmem = quni(tmem_base + (uint32_t)(c * 2 * C) + (which ? 0u : (uint32_t)C));
          uint32_t a_tap = quni((uint32_t)a_desc0 + (uint32_t)((c * A_BUF) >> 4) + (uint32_t)(-((ntaps - 1) / 2) * dil));
          uint32_t fresh = which ? 1u : 0u;  // conv1 overwrites acc1 with its first MMA; conv2 accumulates into x
#pragma unroll 1
          for (int tap = 0; tap < ntaps; ++tap) {
            PT2(2, mbar_wait(w_full + 8 * slot, ph));
            tc_fence_after();
            // (the shuffles tell ptxas that these are warp-uniform: the 8 MMAs below then issue from uniform registers
            //  instead of one ELECT / vote loop per instruction)
            const uint32_t b_lo = quni(w_lo + slot * (uint32_t)(W_STAGE >> 4));
            const uint32_t a_t = quni(a_tap), fr = quni(fresh);
#pragma unroll
            for (int k16 = 0; k16 < C / 16; ++k16)
              tc2_mma_f16_pred_lh(d_tmem, a_t + (uint32_t)((2 * k16 * A_SLICE) >> 4), a_hi, b_lo + (uint32_t)((2 * k16 * BH_SLICE) >> 4), w_hi,
                                  idesc, k16 == 0 ? fr : 1u, leader);
            fresh = 1u;
            tc2_commit_pred(quni(w_empty + 8 * slot), leader);
            if (++slot == S) { slot = 0; ph ^= 1u; }
            a_tap += (uint32_t)dil;
          }
          tc2_commit_pred((which ? x_full : acc1_full) + 8 * c, leader);
        }
      }
    }
  } else {
    // ===================== epilogue warps (both CTAs) =====================
    const int q = warp & 3;           // TMEM lane quarter this warp may access
    const int cg = (warp - 2) >> 2;   // column group (32 channels = 4 slices)
    const int r_in_blk = q * 32 + lane;
    const int sl0 = cg * NS8;
    const int margin = p.margin;      // conv2's halo: rows per tile edge recomputed by the neighbouring tile
    const int valid_rows = MT - 2 * margin;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(cg * NC);
    const uint64_t slope2 = q2_packf(p.slope, p.slope);
    const uint32_t in_ready_l = mapa_u32(in_ready, 0), mid_ready_l = mapa_u32(mid_ready, 0);
    // pad rows (conv1's halo, real neighbouring rows): warp (q, cg) also fills 32 rows x 2 slices of them -
    // q >> 1 = side (left / right), q & 1 = which two of the column group's four slices
    const int pad_side = q >> 1, pad_sl = sl0 + (q & 1) * 2;
    const int pad_dt = pad_side ? MT + lane : lane - PAD;             // row of this lane relative to the tile start
    const uint32_t pad_row_off = (uint32_t)(pad_side ? PAD + MT + lane : lane) * 16;
    const uint32_t own_row_off = (uint32_t)(PAD + r_in_blk) * 16;

    auto publish = [&](uint32_t cluster_bar) {
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(cluster_bar);
    };
    auto store_row = [&](int pbuf, const uint32_t (&pk)[NC / 2]) {
#pragma unroll
      for (int s8 = 0; s8 < NS8; ++s8) {
        const uint32_t dst = a_smem + (uint32_t)(pbuf * A_BUF) + (uint32_t)(sl0 + s8) * A_SLICE + own_row_off;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk[4 * s8]), "r"(pk[4 * s8 + 1]), "r"(pk[4 * s8 + 2]),
                     "r"(pk[4 * s8 + 3])
                     : "memory");
      }
    };
    // (kept small: two current + two next units live in registers; item_base is recomputed where it is used)
    struct Unit { int t0, item; bool interior, real; };
    auto make_unit = [&](int j) {
      Unit u;
      const int work = 2 * (cid + j * ncl) + (int)rank;
      u.real = work < total_work;  // the last step of an odd tile count has no tile for the peer: zeros in, nothing out
      const int wk = u.real ? work : total_work - 1;
      const int bb = wk / p.tiles_per_item;
      const int tile = wk - bb * p.tiles_per_item;
      u.t0 = tile * valid_rows - margin;
      u.item = bb;
      u.interior = u.real && (u.t0 - PAD >= 0) && (u.t0 + MT + PAD <= p.L);  // no row of tile + pads needs a bounds check
      return u;
    };
    auto item_base_of = [&](const Unit& u) { return (size_t)u.item * NSL * p.Lp; };
    // the tile's rows of this warp's 4 slices + its share of the pad rows
    auto load_x = [&](float4 (&dst)[NC / 4], float4 (&pad)[4], const Unit& u) {
      const int t = u.t0 + r_in_blk, tp = u.t0 + pad_dt;
      const bool ok = u.interior || (u.real && t >= 0 && t < p.L);
      const bool okp = u.interior || (u.real && tp >= 0 && tp < p.L);
      const size_t item_base = item_base_of(u);
#pragma unroll
      for (int s8 = 0; s8 < NS8; ++s8) {
        dst[2 * s8] = make_float4(0.f, 0.f, 0.f, 0.f);
        dst[2 * s8 + 1] = dst[2 * s8];
        if (ok) {
          const float* src = p.x_in + ((item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8;
          dst[2 * s8] = *reinterpret_cast<const float4*>(src);
          dst[2 * s8 + 1] = *reinterpret_cast<const float4*>(src + 4);
        }
      }
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        pad[2 * s] = make_float4(0.f, 0.f, 0.f, 0.f);
        pad[2 * s + 1] = pad[2 * s];
        if (okp) {
          const float* src = p.x_in + ((item_base + (size_t)(pad_sl + s) * p.Lp) + PADL + tp) * 8;
          pad[2 * s] = *reinterpret_cast<const float4*>(src);
          pad[2 * s + 1] = *reinterpret_cast<const float4*>(src + 4);
        }
      }
    };
    // L: x rows (registers) -> TMEM residual x[pbuf] + lrelu(x) operand rows (tile and pads), then tell the leader
    auto phase_L = [&](const float4 (&cur)[NC / 4], const float4 (&pad)[4], int pbuf) {
#pragma unroll
      for (int s8 = 0; s8 < NS8; ++s8) {
        uint32_t r8[8] = {__float_as_uint(cur[2 * s8].x), __float_as_uint(cur[2 * s8].y), __float_as_uint(cur[2 * s8].z),
                          __float_as_uint(cur[2 * s8].w), __float_as_uint(cur[2 * s8 + 1].x), __float_as_uint(cur[2 * s8 + 1].y),
                          __float_as_uint(cur[2 * s8 + 1].z), __float_as_uint(cur[2 * s8 + 1].w)};
        tc_st8(t_lane + (uint32_t)(pbuf * 2 * C) + s8 * 8, r8);
      }
      uint32_t pk[NC / 2];
#pragma unroll
      for (int i = 0; i < NC / 4; ++i) {
        pk[2 * i] = q2_lrelu_pack(q2_packf(cur[i].x, cur[i].y), slope2, BF16);
        pk[2 * i + 1] = q2_lrelu_pack(q2_packf(cur[i].z, cur[i].w), slope2, BF16);
      }
      store_row(pbuf, pk);
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const uint32_t dst = a_smem + (uint32_t)(pbuf * A_BUF) + (uint32_t)(pad_sl + s) * A_SLICE + pad_row_off;
        const float4 a = pad[2 * s], b = pad[2 * s + 1];
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(q2_lrelu_pack(q2_packf(a.x, a.y), slope2, BF16)),
                     "r"(q2_lrelu_pack(q2_packf(a.z, a.w), slope2, BF16)), "r"(q2_lrelu_pack(q2_packf(b.x, b.y), slope2, BF16)),
                     "r"(q2_lrelu_pack(q2_packf(b.z, b.w), slope2, BF16))
                     : "memory");
      }
      tc_wait_st();
      publish(in_ready_l + 8 * pbuf);
    };

    // prologue: the first two tiles have nothing to hide behind
    Unit ucur[2], unext[2];
    {
      const long long t_p = TIMING ? clock64() : 0;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        ucur[j] = make_unit(j);
        unext[j] = ucur[j];
        if (j < n_tiles) {
          float4 cur[NC / 4], pad[4];
          load_x(cur, pad, ucur[j]);
          phase_L(cur, pad, j);
        }
      }
      if (TIMING) tacc[4] += clock64() - t_p;
    }
    const uint64_t scale2 = q2_packf(p.scale, p.scale);
    for (int jp = 0; jp < n_tiles; jp += 2) {
      const int cnt = (n_tiles - jp) < 2 ? (n_tiles - jp) : 2;
      const uint32_t par = (uint32_t)(jp >> 1) & 1u;
      // ---- M: conv1 accumulator -> conv2 operand rows, zero outside the sequence (conv2's zero padding) ----
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        if (c >= cnt) continue;
        const Unit& u = ucur[c];
        const int t = u.t0 + r_in_blk;
        PT2(0, mbar_wait(acc1_full + 8 * c, par));
        tc_fence_after();
        const long long t_m = TIMING ? clock64() : 0;
        uint32_t rr[NC];
        PT2(8, tc_ld32(t_lane + (uint32_t)(c * 2 * C + C), rr));
        uint32_t pk[NC / 2];
        const ulonglong2* bp = reinterpret_cast<const ulonglong2*>(bias_s + cg * NC);
#pragma unroll
        for (int i = 0; i < NC / 4; ++i) {
          const ulonglong2 bb = bp[i];
          pk[2 * i] = q2_lrelu_pack(q2_add(q2_pack(rr[4 * i], rr[4 * i + 1]), bb.x), slope2, BF16);
          pk[2 * i + 1] = q2_lrelu_pack(q2_add(q2_pack(rr[4 * i + 2], rr[4 * i + 3]), bb.y), slope2, BF16);
        }
        if (!u.interior && (t < 0 || t >= p.L)) {
#pragma unroll
          for (int i = 0; i < NC / 2; ++i) pk[i] = 0u;
        }
        store_row(c, pk);
        PT2(9, publish(mid_ready_l + 8 * c));
        if (jp + c + 2 < n_tiles) unext[c] = make_unit(jp + c + 2);
        if (TIMING) tacc[1] += clock64() - t_m;
      }
      // ---- F: x[pbuf] (= x + conv2 without b2) -> output, and L of the tile that takes this buffer next ----
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        if (c >= cnt) continue;
        const Unit u = ucur[c];
        const int t = u.t0 + r_in_blk;
        const bool has_next = jp + c + 2 < n_tiles;
        float4 xn[NC / 4], xp[4];
        if (has_next) PT2(10, load_x(xn, xp, unext[c]));  // in flight while conv2 finishes
        PT2(2, mbar_wait(x_full + 8 * c, par));
        tc_fence_after();
        const long long t_f = TIMING ? clock64() : 0;
        uint32_t rr[NC];
        tc_ld32(t_lane + (uint32_t)(c * 2 * C), rr);
        // the next tile's rows first (the tensor pipe waits for them), this tile's stores behind them
        if (has_next) PT2(5, phase_L(xn, xp, c));
        ucur[c] = unext[c];
        const long long t_o = TIMING ? clock64() : 0;
        const bool ok = u.real && (u.interior || ((t >= 0) && (t < p.L)));
        if (ok && r_in_blk >= margin && r_in_blk < MT - margin) {
          const size_t item_base = item_base_of(u);
#pragma unroll
          for (int s8 = 0; s8 < NS8; ++s8) {
            const size_t off = ((item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8;
            const ulonglong2 c0 = *reinterpret_cast<const ulonglong2*>(bias_s + C + cg * NC + s8 * 8);
            const ulonglong2 c1 = *reinterpret_cast<const ulonglong2*>(bias_s + C + cg * NC + s8 * 8 + 4);
            uint64_t o2[4];
            o2[0] = q2_mul(q2_add(q2_pack(rr[8 * s8 + 0], rr[8 * s8 + 1]), c0.x), scale2);
            o2[1] = q2_mul(q2_add(q2_pack(rr[8 * s8 + 2], rr[8 * s8 + 3]), c0.y), scale2);
            o2[2] = q2_mul(q2_add(q2_pack(rr[8 * s8 + 4], rr[8 * s8 + 5]), c1.x), scale2);
            o2[3] = q2_mul(q2_add(q2_pack(rr[8 * s8 + 6], rr[8 * s8 + 7]), c1.y), scale2);
            if (p.sum_in) {
              const ulonglong2 s0 = *reinterpret_cast<const ulonglong2*>(p.sum_in + off);
              const ulonglong2 s1 = *reinterpret_cast<const ulonglong2*>(p.sum_in + off + 4);
              o2[0] = q2_add(o2[0], s0.x); o2[1] = q2_add(o2[1], s0.y);
              o2[2] = q2_add(o2[2], s1.x); o2[3] = q2_add(o2[3], s1.y);
            }
            if (p.out_f32) {
              *reinterpret_cast<ulonglong2*>(p.out_f32 + off) = make_ulonglong2(o2[0], o2[1]);
              *reinterpret_cast<ulonglong2*>(p.out_f32 + off + 4) = make_ulonglong2(o2[2], o2[3]);
            }
            if (p.out_act) {
              const uint64_t os2 = q2_packf(p.out_slope, p.out_slope);
              uint32_t pk[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) pk[i] = q2_lrelu_pack(o2[i], os2, p.out_fmt == ACT_BF16);
              if (p.out_zs > 1) {
                uint16_t* oa = reinterpret_cast<uint16_t*>(p.out_act) +
                               (((size_t)u.item * NSL + (size_t)(sl0 + s8)) * p.Lp_out_act + PADL + (size_t)t * p.out_zs) * 8;
                *reinterpret_cast<uint4*>(oa) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                for (int z = 1; z < p.out_zs; ++z) *reinterpret_cast<uint4*>(oa + z * 8) = make_uint4(0, 0, 0, 0);
              } else {
                *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out_act) + off) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              }
            }
          }
        }
        if (TIMING) { tacc[11] += clock64() - t_o; tacc[3] += clock64() - t_f; }
      }
    }
  }

  if (TIMING && p.timing && (threadIdx.x == 0 || threadIdx.x == 32 || threadIdx.x == 64)) {
    // warp 1 (leader: MMA issuer) [0] wait conv1 operand [1] wait conv2 operand [2] wait weights [5] total;
    // (peer: relay) [0] wait own weight stage; epilogue warp 2: [8] wait acc1 [9] M [10] wait x [11] F + L [12] prologue
    // [13] total [6] L inside F; producer: [7] wait free weight stage
    long long* o = p.timing + (size_t)blockIdx.x * 16;
    if (threadIdx.x == 32) { o[0] = tacc[0]; o[1] = tacc[1]; o[2] = tacc[2]; o[5] = clock64() - t_begin; }
    if (threadIdx.x == 64) { o[8] = tacc[0]; o[9] = tacc[1]; o[10] = tacc[2]; o[11] = tacc[3]; o[12] = tacc[4]; o[13] = clock64() - t_begin; o[6] = tacc[5];
                             o[3] = tacc[10]; o[4] = tacc[8]; o[14] = tacc[9]; o[15] = tacc[11]; }
    if (threadIdx.x == 0) o[7] = tacc[0];
  }
  // nothing of this CTA (barriers, operand tiles, TMEM) may go away while the peer can still signal or read it
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ------------------------------ host side ----------------------------------
bool mrf_pair2_supported(int C, int halo1, int halo2) {
  return C == pair2::C && halo1 <= pair2::PAD && halo2 <= pair2::PAD && pair2::MT - 2 * halo2 >= 64;
}

cudaError_t launch_mrf_pair2(const PairParams& p, int num_sms, cudaStream_t s) {
  using namespace pair2;
  auto kern = p.timing ? (p.fmt == ACT_BF16 ? mrf_pair2_kernel<true, true> : mrf_pair2_kernel<false, true>)
                       : (p.fmt == ACT_BF16 ? mrf_pair2_kernel<true, false> : mrf_pair2_kernel<false, false>);
  static unsigned long long done[4] = {0, 0, 0, 0};
  cudaError_t e = set_smem_once(reinterpret_cast<const void*>(kern), SMEM_BYTES, done[(p.timing ? 2 : 0) + (p.fmt == ACT_BF16 ? 1 : 0)]);
  if (e != cudaSuccess) return e;
  const int h1 = (p.k1 - 1) / 2 * p.d1, h2 = (p.k2 - 1) / 2 * p.d2;
  if (!p.w1h || !p.w2h || p.k1 < 1 || p.k2 < 1 || h1 > PAD || h2 > PAD || MT - 2 * h2 < 64) return cudaErrorInvalidValue;
  PairParams q = p;
  q.margin = h2;  // conv1's halo comes out of the pad rows (real neighbouring rows), only conv2's is recomputed
  const int valid = MT - 2 * h2;
  q.tiles_per_item = (p.L + valid - 1) / valid;
  const long long total = (long long)q.B * q.tiles_per_item;
  if (total <= 0 || total > 0x7fffffffLL) return cudaErrorInvalidValue;
  long long clusters = num_sms / 2;
  if ((total + 1) / 2 < clusters) clusters = (total + 1) / 2;
  kern<<<(unsigned)(2 * clusters), THREADS, SMEM_BYTES, s>>>(q);
  return cudaGetLastError();
}

}  // namespace rvcdec
