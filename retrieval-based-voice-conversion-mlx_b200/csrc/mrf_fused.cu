// Fused multi-receptive-field stage for the low-channel, high-rate stages (C = 64, 32):
// a whole ResBlock chain (residuals.py:71-80; 3 x [lrelu -> dilated conv1 -> lrelu -> conv2 -> +x])
// runs inside one CTA per time tile, with the MRF average (hifigan_nsf.py:200-207) folded into the
// last write.  HBM sees the stage input once per ResBlock and the running MRF sum; the ~7 tensors
// per conv pair the un-fused path streams through HBM (SURVEY 8a "activation volume") stay on chip:
//
//   TMEM   x[MB][128 x C] fp32   the residual stream; conv2 ACCUMULATES into it (the "+x" is free)
//          acc1[2][128 x C] fp32 conv1 accumulators, double-buffered per 128-row block
//   smem   A0, A1                16-bit operand tiles (MB*128 + 2*32 pad rows) x C, blocked-8 K-major
//                                = the no-swizzle UMMA layout, a tap / dilation is a row offset
//          W ring                S tap blocks (C x C 16-bit); a conv's k taps stay resident while the
//                                MB row blocks pass over them, so L2 -> smem weight traffic is per tile
//
// Tiles overlap by `margin` rows per side (the chain's receptive field); every conv is computed on
// all MB*128 rows and the rows whose inputs fell outside the tile are simply never stored.
//
// Warp roles (320 threads): warp 0 = TMA weight producer, warp 1 = tcgen05.mma issuer,
// warps 2-9 = epilogue (TMEM -> bias/lrelu/convert -> smem operand; stage input load; final store).
#include <algorithm>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

// EW = epilogue warps (8 or 16): EW / 4 column groups x 4 TMEM lane quarters.  Warps: 0 = weight producer,
// 1 = MMA issuer A, 2 .. EW+1 = epilogue, EW+2 = MMA issuer B ((EW + 2) % 4 = 2: another scheduler than warp 1).
__host__ __device__ constexpr int fused_threads(int ew) { return 32 * (ew + 3); }

template <int C, int MB, int S>
struct FusedCfg {
  static constexpr int MT = MB * 128;
  static constexpr int PAD = 32;
  static constexpr int ROWS = MT + 2 * PAD;
  static constexpr int NSL = C / 8;
  static constexpr int A_SLICE = ROWS * 16;
  static constexpr int A_BUF = NSL * A_SLICE;
  static constexpr int W_SLOT = C * C * 2;
  static constexpr int B_SLICE = C * 16;
  static constexpr int X_COLS = MB * C;
  static constexpr int ACC1_COL0 = X_COLS;
  static constexpr int USED_COLS = X_COLS + 2 * C;
  static constexpr int TMEM_COLS = USED_COLS <= 128 ? 128 : USED_COLS <= 256 ? 256 : 512;
  static constexpr int NUM_BARS = 2 * S + 4 + 3 * MB + 1;
  // fused conv_post (last stage only, C = 32): ring of 3 blocks x 2 column halves x 7 taps x 128 rows of partial
  // dot products + the [7][C] conv_post weights
  static constexpr int POST_TAPS = 7;
  // partial sums: double-buffered block [2][4 column groups][7][128] + the last 8 rows of the previous blocks [3][4][7][8]
  static constexpr int POST_MAIN = 2 * 4 * POST_TAPS * 128;
  static constexpr int POST_TAIL = 3 * 4 * POST_TAPS * 8;
  static constexpr int POST_BYTES = (C == 32) ? (POST_MAIN + POST_TAIL) * 4 : 0;
  static constexpr int SMEM_BYTES = 128 + 2 * A_BUF + S * W_SLOT + NUM_BARS * 8 + 16 + 16 + POST_BYTES;
  static_assert(USED_COLS <= 512, "TMEM budget");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
  static_assert(C == 32 || C == 64, "column split below assumes 8, 16 or 32 columns per epilogue warp");
};

__device__ __forceinline__ uint32_t uni(uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); }

__device__ __forceinline__ uint32_t pack2(float a, float b, int fmt) {
  if (fmt == ACT_F16) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// packed fp32x2 arithmetic (sm_100 FADD2 / FMUL2): halves the epilogue's instruction count
__device__ __forceinline__ uint64_t f2_pack_u(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) { return f2_pack_u(__float_as_uint(lo), __float_as_uint(hi)); }
__device__ __forceinline__ uint64_t f2_splat(float v) { return f2_pack(v, v); }
__device__ __forceinline__ float2 f2_unpack(uint64_t v) {
  uint32_t lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
  return make_float2(__uint_as_float(lo), __uint_as_float(hi));
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t f2_mul(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// 16-bit pair of leaky_relu(v) = max(v, slope * v), 0 < slope < 1
template <bool BF16>
__device__ __forceinline__ uint32_t lrelu_pack(uint64_t v, uint64_t slope2) {
  const float2 a = f2_unpack(v), m = f2_unpack(f2_mul(v, slope2));
  return pack2(fmaxf(a.x, m.x), fmaxf(a.y, m.y), BF16 ? ACT_BF16 : ACT_F16);
}

template <int NC>
__device__ __forceinline__ void tmem_load_cols(uint32_t taddr, uint32_t (&r)[NC]);
template <>
__device__ __forceinline__ void tmem_load_cols<8>(uint32_t taddr, uint32_t (&r)[8]) { tc_ld8(taddr, r); }
template <>
__device__ __forceinline__ void tmem_load_cols<16>(uint32_t taddr, uint32_t (&r)[16]) { tc_ld16(taddr, r); }
template <>
__device__ __forceinline__ void tmem_load_cols<32>(uint32_t taddr, uint32_t (&r)[32]) { tc_ld32(taddr, r); }

// TIMING = true compiles the instrumented twin used by tools/fused_timing.py: the MMA warp and one
// epilogue warp accumulate clock64() spans around every wait into p.timing[blockIdx.x][16].
#define TWAIT(slot_, ...)                                  \
  do {                                                     \
    if (TIMING) {                                          \
      const long long _t = clock64();                      \
      __VA_ARGS__;                                         \
      tacc[slot_] += clock64() - _t;                       \
    } else {                                               \
      __VA_ARGS__;                                         \
    }                                                      \
  } while (0)

template <int C, int MB, int S, int EW, bool BF16, bool TIMING>
__global__ void __launch_bounds__(fused_threads(EW), 1) mrf_fused_kernel(const __grid_constant__ FusedParams p) {
  using Cfg = FusedCfg<C, MB, S>;
  constexpr int FUSED_THREADS = fused_threads(EW);
  constexpr int FUSED_EPI_WARPS = EW;
  constexpr int FUSED_MMA_WARP_B = EW + 2;
  static_assert(EW == 8 || EW == 16, "epilogue warps");
  long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long t_taps = 0;  // MMA warp: time inside the tap loops (issue + queue back-pressure)
  const long long t_begin = TIMING ? clock64() : 0;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t a0_smem = smem_base;
  const uint32_t a1_smem = a0_smem + Cfg::A_BUF;
  const uint32_t w_smem = a1_smem + Cfg::A_BUF;
  const uint32_t bars = w_smem + S * Cfg::W_SLOT;
  const uint32_t w_full = bars, w_empty = w_full + 8 * S, acc1_full = w_empty + 8 * S, acc1_empty = acc1_full + 16,
                 x_full = acc1_empty + 16, a0_ready = x_full + 8 * MB, a1_ready = a0_ready + 8 * MB;
  const uint32_t dummy_bar = a1_ready + 8 * MB;  // sink for commits that release nothing
  const uint32_t tmem_slot = dummy_bar + 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));
  const uint32_t post_s = (tmem_slot + 16 + 15) & ~15u;  // partial-sum ring of the fused conv_post
  float* post_ring = reinterpret_cast<float*>(smem_gen + (post_s - smem_base));
  float* post_tail = post_ring + Cfg::POST_MAIN;
  const bool do_post_any = Cfg::POST_BYTES > 0 && p.post_w != nullptr;

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);  // provably warp-uniform role index
  const int lane = threadIdx.x & 31;

  // operand tiles start as zeros: the pad rows are the convolutions' zero padding at the tile edges
  for (int i = threadIdx.x; i < 2 * Cfg::A_BUF / 16; i += FUSED_THREADS)
    reinterpret_cast<uint4*>(smem_gen)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(w_full + 8 * i, 1); mbar_init(w_empty + 8 * i, MB > 1 ? 2 : 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(acc1_full + 8 * i, 1); mbar_init(acc1_empty + 8 * i, FUSED_EPI_WARPS); }
    for (int i = 0; i < MB; ++i) {
      mbar_init(x_full + 8 * i, 1);
      mbar_init(a0_ready + 8 * i, FUSED_EPI_WARPS);
      mbar_init(a1_ready + 8 * i, FUSED_EPI_WARPS);
    }
    mbar_init(dummy_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "r"((uint32_t)Cfg::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  const int nd = p.nd, n_rb = p.n_rb;
  const int total_work = p.B * p.tiles_per_item;

  if (warp == 0) {
    // ===================== weight producer =====================
    uint32_t slot = 0, ph = 0;
    for (int work = blockIdx.x; work < total_work; work += gridDim.x) {
      for (int rb = 0; rb < n_rb; ++rb) {
        for (int m = 0; m < nd; ++m) {
#pragma unroll 1
          for (int which = 0; which < 2; ++which) {
            const FusedConv& cv = which ? p.rb[rb].c2[m] : p.rb[rb].c1[m];
            const uint8_t* src = reinterpret_cast<const uint8_t*>(cv.w);
            const int ntaps = cv.ntaps;
            for (int tap = 0; tap < ntaps; ++tap) {
              mbar_wait(w_empty + 8 * slot, ph ^ 1u);
              if (elect_one()) {
                mbar_expect_tx(w_full + 8 * slot, Cfg::W_SLOT);
                bulk_g2s(w_smem + slot * Cfg::W_SLOT, src + (size_t)tap * Cfg::W_SLOT, Cfg::W_SLOT, w_full + 8 * slot);
              }
              __syncwarp();
              if (++slot == S) { slot = 0; ph ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == 1 || warp == FUSED_MMA_WARP_B) {
    // ===================== MMA issuers (two converged warps on different schedulers) =====================
    // The issue path (barrier probes + ~9 uniform instructions per MMA) costs about as much time as the
    // tensor pipe needs to execute an N = 64 / 32 MMA (timing twin, r1g: 48 clk per MMA in the tap loop +
    // ~600 clk per 128-row block), so one issuing thread cannot keep the pipe busy.  Blocks alternate
    // between two issuers; all ordering between them goes through the operand / accumulator barriers.
    const uint32_t issuer = (warp == 1) ? 0u : 1u;
    const uint32_t idesc = make_idesc(p.fmt, C);
    const uint64_t a0_desc = make_smem_desc(a0_smem, Cfg::A_SLICE, 128) + (uint64_t)Cfg::PAD;
    const uint64_t a1_desc = make_smem_desc(a1_smem, Cfg::A_SLICE, 128) + (uint64_t)Cfg::PAD;
    const uint64_t w_desc = make_smem_desc(w_smem, Cfg::B_SLICE, 128);
    const uint32_t leader = elect_one() ? 1u : 0u;
    uint32_t slot0 = 0, ph0 = 0;  // ring position of tap 0 of the current conv
    uint32_t cnt[2] = {0, 0};     // conv1 / conv2 blocks issued so far (by both issuers); parity = owner
    uint32_t seq = 0;             // (tile, ResBlock) pairs processed so far
    for (int work = blockIdx.x; work < total_work; work += gridDim.x) {
      for (int rb = 0; rb < n_rb; ++rb, ++seq) {
        for (int m = 0; m < nd; ++m) {
          const uint32_t par = (seq * (uint32_t)nd + (uint32_t)m) & 1u;
#pragma unroll 1
          for (int which = 0; which < 2; ++which) {
            const FusedConv& cv = which ? p.rb[rb].c2[m] : p.rb[rb].c1[m];
            const int ntaps = cv.ntaps, dil = cv.dil;
            const int shift0 = -((ntaps - 1) / 2) * dil;
            const uint32_t in_ready = which ? a1_ready : a0_ready;
            const uint64_t in_desc = which ? a1_desc : a0_desc;
            const uint32_t cnt0 = cnt[which];
            // blocks of this conv owned by this issuer: mb with ((cnt0 + mb) & 1) == issuer
            const int first_own = ((cnt0 & 1u) == issuer) ? 0 : 1;
            const int last_own = (((cnt0 + MB - 1) & 1u) == issuer) ? MB - 1 : MB - 2;
#pragma unroll 1
            for (int mb = first_own; mb < MB; mb += 2) {
              const uint32_t bcnt = cnt0 + (uint32_t)mb;
              const uint32_t buf = issuer;  // conv1 accumulator buffer == owner (bcnt & 1)
              uint32_t d_tmem = which ? (tmem_base + mb * C) : (tmem_base + Cfg::ACC1_COL0 + buf * C);
              // ---- every barrier this block depends on, probed by one lane each, one vote ----
              {
                uint32_t my_bar = 0, my_par = 0;
                bool mine = false;
                if (lane == 0 && mb > 0) { mine = true; my_bar = in_ready + 8 * (mb - 1); my_par = par; }
                if (lane == 1) { mine = true; my_bar = in_ready + 8 * mb; my_par = par; }
                if (lane == 2 && mb + 1 < MB) { mine = true; my_bar = in_ready + 8 * (mb + 1); my_par = par; }
                if (lane == 3 && !which) { mine = true; my_bar = acc1_empty + 8 * buf; my_par = (((bcnt >> 1) & 1u) ^ 1u); }
                if (mb == first_own && lane >= 4 && lane < 4 + ntaps) {  // weights of this conv: lane 4+t waits for tap t
                  uint32_t sl = slot0 + (uint32_t)(lane - 4), pz = ph0;
                  if (sl >= S) { sl -= S; pz ^= 1u; }
                  mine = true; my_bar = w_full + 8 * sl; my_par = pz;
                }
                const long long t_w = TIMING ? clock64() : 0;
                for (uint32_t spin = 0;; ++spin) {  // uniform exit (vote), so the warp stays provably converged
                  const uint32_t ok = mine ? mbar_try(my_bar, my_par) : 1u;
                  if (__all_sync(0xffffffffu, ok)) break;
                  if (spin > (1u << 24)) { printf("rvcdec mrf_fused: barrier timeout (block %d issuer %u)\n", (int)blockIdx.x, issuer); __trap(); }
                }
                if (TIMING) tacc[0] += clock64() - t_w;
              }
              tc_fence_after();
              // Warp-uniform tap loop; the MMAs / commits are predicated on the elected lane.
              // (the shuffles only tell ptxas that these loop inputs are warp-uniform, so that the loop
              //  below runs on the uniform datapath).  32-bit descriptor arithmetic: only the
              // start-address field (low word) changes.
              uint32_t a_lo = uni((uint32_t)in_desc + (uint32_t)(mb * 128 + shift0));
              const uint32_t a_hi = (uint32_t)(in_desc >> 32), w_lo = (uint32_t)w_desc, w_hi = (uint32_t)(w_desc >> 32);
              uint32_t slot = uni(slot0);
              // this issuer's last block of the conv releases each weight slot (w_empty counts both issuers);
              // its other blocks "release" a dummy barrier so that the commit stays unconditional
              const uint32_t rel_base = uni((mb == last_own) ? w_empty : dummy_bar);
              const uint32_t rel_step = uni((mb == last_own) ? 8u : 0u);
              uint32_t acc = uni(which ? 1u : 0u);
              d_tmem = uni(d_tmem);
              const int ntaps_u = (int)uni((uint32_t)ntaps), dil_u = (int)uni((uint32_t)dil);
              const long long t_loop = TIMING ? clock64() : 0;
#pragma unroll 1
              for (int tap = 0; tap < ntaps_u; ++tap) {
                const uint32_t b_lo = w_lo + slot * (uint32_t)(Cfg::W_SLOT >> 4);
#pragma unroll
                for (int ks = 0; ks < C / 16; ++ks)
                  tc_mma_f16_pred_lh(d_tmem, a_lo + (uint32_t)((2 * ks * Cfg::A_SLICE) >> 4), a_hi,
                                     b_lo + (uint32_t)((2 * ks * Cfg::B_SLICE) >> 4), w_hi, idesc, ks == 0 ? acc : 1u, leader);
                acc = 1u;
                tc_commit_pred(rel_base + rel_step * slot, leader);
                a_lo += (uint32_t)dil_u;
                if (++slot == S) slot = 0;
              }
              if (TIMING) t_taps += clock64() - t_loop;
              tc_commit_pred(which ? (x_full + 8 * mb) : (acc1_full + 8 * buf), leader);
            }
            cnt[which] = cnt0 + MB;
            slot0 += (uint32_t)ntaps;
            if (slot0 >= S) { slot0 -= S; ph0 ^= 1u; }
          }
        }
      }
    }
  } else {
    // ===================== epilogue warps =====================
    constexpr int CG = EW / 4;       // column groups
    constexpr int NC = C / CG;       // accumulator columns handled by this warp
    constexpr int NS8 = NC / 8;      // = 8-channel slices handled by this warp
    const int q = warp & 3;          // TMEM lane quarter this warp may access
    const int hh = (warp - 2) >> 2;  // column group
    const int r_in_blk = q * 32 + lane;
    const int sl0 = hh * NS8;        // first slice of this warp
    const uint64_t slope2 = f2_splat(p.slope);
    const int valid_rows = Cfg::MT - 2 * p.margin;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    uint32_t a1cnt = 0, seq = 0;

    // make this warp's smem / TMEM writes visible to the MMA (async proxy) and signal `bar`
    auto publish = [&](uint32_t bar) {
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar);
    };
    // 16-bit row of this warp's slices -> operand tile
    auto store_row = [&](uint32_t a_smem, int r, const uint32_t (&pk)[NC / 2]) {
#pragma unroll
      for (int s8 = 0; s8 < NS8; ++s8) {
        const uint32_t dst = a_smem + (uint32_t)(sl0 + s8) * Cfg::A_SLICE + (uint32_t)(Cfg::PAD + r) * 16;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk[4 * s8]), "r"(pk[4 * s8 + 1]),
                     "r"(pk[4 * s8 + 2]), "r"(pk[4 * s8 + 3])
                     : "memory");
      }
    };
    auto load_x = [&](float4 (&dst)[NC / 4], size_t item_base, int t, bool ok) {
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
    };

    // A unit = one ResBlock over one tile.  The stage input of unit n+1 is loaded block by block
    // inside the final epilogue of unit n (as soon as a block's residual has been read out of TMEM its
    // columns and operand rows are free), so the MMA warp finds the first blocks of the next unit ready
    // while the last block of this one is still in the tensor pipe.
    struct Unit { int work, rb, t0, item; size_t item_base; bool interior, valid; };
    auto make_unit = [&](int work, int rb) {
      Unit u;
      u.work = work; u.rb = rb; u.valid = work < total_work;
      const int bb = work / p.tiles_per_item;
      const int tile = work - bb * p.tiles_per_item;
      u.t0 = tile * valid_rows - p.margin;
      u.item = bb;
      u.item_base = (size_t)bb * Cfg::NSL * p.Lp;
      u.interior = (u.t0 >= 0) && (u.t0 + Cfg::MT <= p.L);  // no row of this tile needs zero padding
      return u;
    };
    auto load_unit_x = [&](float4 (&dst)[NC / 4], const Unit& u, int r) {
      const int t = u.t0 + r;
      load_x(dst, u.item_base, t, u.interior || (t >= 0 && t < p.L));
    };
    // L2 prefetch of what the final epilogue of this unit will read for row r: the running MRF sum and the
    // next unit's stage input.  Issued one convolution ahead (from the last conv1 epilogue), so the
    // final epilogue's loads hit L2 instead of paying an HBM round trip per block on the critical path
    // between two units (timing twin r1k: ~15k clk per unit were spent there).
    auto prefetch_final = [&](const Unit& u, const Unit& nx, int r, bool sum_in) {
      const int t = u.t0 + r;
      if (sum_in && r >= p.margin && r < Cfg::MT - p.margin && t >= 0 && t < p.L) {
#pragma unroll
        for (int s8 = 0; s8 < NS8; ++s8)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(p.sum + ((u.item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8));
      }
      const int tn = nx.t0 + r;
      if (nx.valid && tn >= 0 && tn < p.L) {
#pragma unroll
        for (int s8 = 0; s8 < NS8; ++s8)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(p.x_in + ((nx.item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + tn) * 8));
      }
    };
    // x block (registers) -> TMEM residual + lrelu(x) operand rows, then signal the MMA warp
    auto store_unit_x = [&](const float4 (&cur)[NC / 4], int mb) {
      const int r = mb * 128 + r_in_blk;
      uint32_t pk[NC / 2];
#pragma unroll
      for (int i = 0; i < NC / 4; ++i) {
        pk[2 * i] = lrelu_pack<BF16>(f2_pack(cur[i].x, cur[i].y), slope2);
        pk[2 * i + 1] = lrelu_pack<BF16>(f2_pack(cur[i].z, cur[i].w), slope2);
      }
#pragma unroll
      for (int s8 = 0; s8 < NS8; ++s8) {
        uint32_t r8[8] = {__float_as_uint(cur[2 * s8].x), __float_as_uint(cur[2 * s8].y), __float_as_uint(cur[2 * s8].z),
                          __float_as_uint(cur[2 * s8].w), __float_as_uint(cur[2 * s8 + 1].x), __float_as_uint(cur[2 * s8 + 1].y),
                          __float_as_uint(cur[2 * s8 + 1].z), __float_as_uint(cur[2 * s8 + 1].w)};
        tc_st8(t_lane + mb * C + hh * NC + s8 * 8, r8);
      }
      store_row(a0_smem, r, pk);  // rows outside the sequence were loaded as zeros
      tc_wait_st();
      publish(a0_ready + 8 * mb);
    };

    Unit u = make_unit(blockIdx.x, 0);
    if (u.valid) {  // the first unit of this CTA has nothing to hide behind
      const long long t_xl = TIMING ? clock64() : 0;
      float4 cur[NC / 4], nxt[NC / 4];
      load_unit_x(cur, u, r_in_blk);
#pragma unroll 1
      for (int mb = 0; mb < MB; ++mb) {
        if (mb + 1 < MB) load_unit_x(nxt, u, (mb + 1) * 128 + r_in_blk);
        store_unit_x(cur, mb);
#pragma unroll
        for (int i = 0; i < NC / 4; ++i) cur[i] = nxt[i];
      }
      if (TIMING) tacc[2] += clock64() - t_xl;
    }
    while (u.valid) {
      Unit nx = (u.rb + 1 < n_rb) ? make_unit(u.work, u.rb + 1) : make_unit(u.work + (int)gridDim.x, 0);
      const FusedRB& R = p.rb[u.rb];
      const int t0 = u.t0;
      const size_t item_base = u.item_base;
      const bool interior = u.interior;
      for (int m = 0; m < nd; ++m) {
        const uint32_t par = (seq * (uint32_t)nd + (uint32_t)m) & 1u;
        // ---- conv1 epilogue: A1 = lrelu(acc1 + b1) ----
        {
          uint64_t b1[NC / 2];  // bias as fp32x2 pairs in registers
          const uint64_t* g1 = reinterpret_cast<const uint64_t*>(R.c1[m].bias + hh * NC);
#pragma unroll
          for (int i = 0; i < NC / 2; ++i) b1[i] = __ldg(g1 + i);
#pragma unroll 1
          for (int mb = 0; mb < MB; ++mb) {
            const int r = mb * 128 + r_in_blk;
            const uint32_t buf = a1cnt & 1u;
            TWAIT(0, mbar_wait(acc1_full + 8 * buf, (a1cnt >> 1) & 1u));
            tc_fence_after();
            uint32_t rr[NC];
            TWAIT(3, tmem_load_cols<NC>(t_lane + Cfg::ACC1_COL0 + buf * C + hh * NC, rr));
            const long long t_c = TIMING ? clock64() : 0;
            uint32_t pk[NC / 2];
#pragma unroll
            for (int i = 0; i < NC / 2; ++i) pk[i] = lrelu_pack<BF16>(f2_add(f2_pack_u(rr[2 * i], rr[2 * i + 1]), b1[i]), slope2);
            if (!interior) {
              const int t = t0 + r;
              if (t < 0 || t >= p.L) {
#pragma unroll
                for (int i = 0; i < NC / 2; ++i) pk[i] = 0u;
              }
            }
            store_row(a1_smem, r, pk);
            if (TIMING) tacc[4] += clock64() - t_c;
            TWAIT(5, fence_proxy_async(); tc_fence_before(); __syncwarp();
                  if (lane == 0) { mbar_arrive(a1_ready + 8 * mb); mbar_arrive(acc1_empty + 8 * buf); });
            ++a1cnt;
            if (m == nd - 1) prefetch_final(u, nx, r, R.sum_in != 0);
          }
        }
        // ---- conv2 epilogue: x (TMEM) now holds x + conv2(...) without the biases ----
        uint64_t b2[NC / 2];
        {
          const uint64_t* g2 = reinterpret_cast<const uint64_t*>(R.c2[m].bias + hh * NC);
#pragma unroll
          for (int i = 0; i < NC / 2; ++i) b2[i] = __ldg(g2 + i);
        }
        if (m < nd - 1) {
#pragma unroll 1
          for (int mb = 0; mb < MB; ++mb) {
            const int r = mb * 128 + r_in_blk;
            const int t = t0 + r;
            TWAIT(1, mbar_wait(x_full + 8 * mb, par));
            tc_fence_after();
            uint32_t rr[NC];
            tmem_load_cols<NC>(t_lane + mb * C + hh * NC, rr);
            uint32_t pk[NC / 2];
#pragma unroll
            for (int i = 0; i < NC / 2; ++i) pk[i] = lrelu_pack<BF16>(f2_add(f2_pack_u(rr[2 * i], rr[2 * i + 1]), b2[i]), slope2);
            if (!interior && (t < 0 || t >= p.L)) {
#pragma unroll
              for (int i = 0; i < NC / 2; ++i) pk[i] = 0u;
            }
            store_row(a0_smem, r, pk);
            publish(a0_ready + 8 * mb);
          }
        } else {
          // ---- ResBlock output -> MRF sum / stage output, interleaved with the next unit's input ----
#pragma unroll 1
          for (int mb = 0; mb < MB; ++mb) {
            const int r = mb * 128 + r_in_blk;
            const int t = t0 + r;
            const bool ok = interior || ((t >= 0) && (t < p.L));
            // fused conv_post consumes rows up to 3 beyond the rows this tile writes (p.margin includes them)
            const bool do_post = do_post_any && R.is_last;
            const int extra = do_post ? (Cfg::POST_TAPS - 1) / 2 : 0;
            const bool store_out = ok && r >= p.margin - extra && r < Cfg::MT - p.margin + extra;
            float4 xn[NC / 4];
            if (nx.valid) load_unit_x(xn, nx, r);  // in flight while we wait for the accumulator
            float4 sm[NC / 4];                      // running MRF sum of this row: all loads before any store
            if (R.sum_in && store_out) {
#pragma unroll
              for (int s8 = 0; s8 < NS8; ++s8) {
                const float* sp = p.sum + ((item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8;
                sm[2 * s8] = *reinterpret_cast<const float4*>(sp);
                sm[2 * s8 + 1] = *reinterpret_cast<const float4*>(sp + 4);
              }
            }
            TWAIT(1, mbar_wait(x_full + 8 * mb, par));
            tc_fence_after();
            uint32_t rr[NC];
            tmem_load_cols<NC>(t_lane + mb * C + hh * NC, rr);
            // The next unit's operand rows are published BEFORE this block's global stores are issued: the proxy
            // fence in publish() waits for the thread's outstanding global stores (timing twin of mrf_pair.cu, r1q/r1s).
            if (nx.valid) TWAIT(6, store_unit_x(xn, mb));
            const long long t_out = TIMING ? clock64() : 0;
            float pj[Cfg::POST_TAPS];  // conv_post partial dot products of this row over this warp's channels
#pragma unroll
            for (int j = 0; j < Cfg::POST_TAPS; ++j) pj[j] = 0.f;
            if (store_out) {
#pragma unroll
              for (int s8 = 0; s8 < NS8; ++s8) {
                const size_t off = ((item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8;
                float o[8];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const float2 bb = f2_unpack(b2[4 * s8 + i]);
                  o[2 * i] = (__uint_as_float(rr[8 * s8 + 2 * i]) + bb.x) * p.scale;
                  o[2 * i + 1] = (__uint_as_float(rr[8 * s8 + 2 * i + 1]) + bb.y) * p.scale;
                }
                if (R.sum_in) {
                  const float4 s0 = sm[2 * s8], s1 = sm[2 * s8 + 1];
                  o[0] += s0.x; o[1] += s0.y; o[2] += s0.z; o[3] += s0.w;
                  o[4] += s1.x; o[5] += s1.y; o[6] += s1.z; o[7] += s1.w;
                }
                if (!R.is_last) {
                  *reinterpret_cast<float4*>(p.sum + off) = make_float4(o[0], o[1], o[2], o[3]);
                  *reinterpret_cast<float4*>(p.sum + off + 4) = make_float4(o[4], o[5], o[6], o[7]);
                } else if (do_post) {
                  // leaky_relu(x, 0.01) . conv_post weights (hifigan_nsf.py:209-210): tap j of this row feeds sample t - j + 3
#pragma unroll
                  for (int i = 0; i < 8; ++i) o[i] = fmaxf(o[i], o[i] * p.post_slope);
#pragma unroll
                  for (int j = 0; j < Cfg::POST_TAPS; ++j) {
                    const float* wc = p.post_wc + (j * C + (sl0 + s8) * 8) % (7 * 32);  // constant bank (C == 32 here)
                    float a = pj[j];
#pragma unroll
                    for (int i = 0; i < 8; ++i) a = fmaf(o[i], wc[i], a);
                    pj[j] = a;
                  }
                } else {
                  if (p.out_f32) {
                    *reinterpret_cast<float4*>(p.out_f32 + off) = make_float4(o[0], o[1], o[2], o[3]);
                    *reinterpret_cast<float4*>(p.out_f32 + off + 4) = make_float4(o[4], o[5], o[6], o[7]);
                  }
                  if (p.out_act) {
                    uint32_t pk[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                      pk[i] = pack2(fmaxf(o[2 * i], o[2 * i] * p.out_slope), fmaxf(o[2 * i + 1], o[2 * i + 1] * p.out_slope), p.out_fmt);
                    if (p.out_zs > 1) {
                      uint16_t* oa = reinterpret_cast<uint16_t*>(p.out_act) +
                                     (((size_t)(item_base / p.Lp) + (size_t)(sl0 + s8)) * p.Lp_out_act + PADL + (size_t)t * p.out_zs) * 8;
                      *reinterpret_cast<uint4*>(oa) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                      for (int z = 1; z < p.out_zs; ++z) *reinterpret_cast<uint4*>(oa + z * 8) = make_uint4(0, 0, 0, 0);
                    } else {
                      *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out_act) + off) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    }
                  }
                }
              }
            }
            if (TIMING) tacc[7] += clock64() - t_out;
            if (do_post) {
              // stage the partial sums of this block, then (column-half 0 warps) finish the samples whose 7-row
              // windows are complete: tile rows [mb*128 - 3, mb*128 + 125).  The ring keeps 3 blocks, so the one
              // barrier per block also orders these reads against the overwrite two blocks later.
              float* ring = post_ring + (((mb & 1) * 4 + hh) * Cfg::POST_TAPS) * 128 + r_in_blk;
#pragma unroll
              for (int j = 0; j < Cfg::POST_TAPS; ++j) ring[j * 128] = pj[j];
              if (r_in_blk >= 120) {  // the next block's first samples read these rows after `main` has moved on
                float* tl = post_tail + (((mb % 3) * 4 + hh) * Cfg::POST_TAPS) * 8 + (r_in_blk - 120);
#pragma unroll
                for (int j = 0; j < Cfg::POST_TAPS; ++j) tl[j * 8] = pj[j];
              }
              asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");
              if (hh == 0) {
                const int ro = r - extra;
                const int to = t0 + ro;
                if (ro >= p.margin && ro < Cfg::MT - p.margin && to >= 0 && to < p.L) {
                  float acc = 0.f;
#pragma unroll
                  for (int j = 0; j < Cfg::POST_TAPS; ++j) {
                    const int src = ro + j - extra - mb * 128;  // block-relative row whose tap-j product feeds sample ro
                    const float* q = (src >= 0) ? post_ring + (((mb & 1) * 4) * Cfg::POST_TAPS + j) * 128 + src
                                                : post_tail + ((((mb + 2) % 3) * 4) * Cfg::POST_TAPS + j) * 8 + (src + 8);
                    const int cg_stride = (src >= 0) ? Cfg::POST_TAPS * 128 : Cfg::POST_TAPS * 8;
#pragma unroll
                    for (int c = 0; c < CG; ++c) acc += q[c * cg_stride];
                  }
                  p.post_out[(size_t)u.item * p.L + to] = tanhf(acc + p.post_bias);
                }
              }
            }
          }
        }
      }
      u = nx;
      ++seq;
    }
  }

  if (TIMING && p.timing && (threadIdx.x == 32 || threadIdx.x == 64)) {
    // [0..3] MMA warp: wait operand rows, wait acc1 drained, wait weights, total; [8..11] epilogue warp 2:
    // wait acc1_full, wait x_full, stage-input load, total
    long long* o = p.timing + (size_t)blockIdx.x * 16 + (threadIdx.x == 32 ? 0 : 8);
    if (threadIdx.x == 32) {
      o[0] = tacc[0]; o[2] = 0;  // o[1] is written by the epilogue warp (final epilogue: output section)
      o[3] = clock64() - t_begin;
      // operand-row waits by position: conv1 block 0, conv1 blocks >= 1, conv2 block 0, conv2 blocks >= 1
      o[4] = 0; o[5] = 0; o[6] = t_taps; o[7] = 0;
    } else {
      o[0] = tacc[0]; o[1] = tacc[1]; o[2] = tacc[2]; o[3] = clock64() - t_begin;
      o[4] = tacc[3]; o[5] = tacc[4]; o[6] = tacc[5]; o[7] = tacc[6];
      (p.timing + (size_t)blockIdx.x * 16)[1] = tacc[7];
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)Cfg::TMEM_COLS)
                 : "memory");
  }
}

// ------------------------------ host side ----------------------------------
template <int C, int MB, int S, int EW>
static cudaError_t launch_fused_inst(const FusedParams& p, int num_sms, cudaStream_t s) {
  using Cfg = FusedCfg<C, MB, S>;
  auto kern = p.timing ? (p.fmt == ACT_BF16 ? mrf_fused_kernel<C, MB, S, EW, true, true> : mrf_fused_kernel<C, MB, S, EW, false, true>)
                       : (p.fmt == ACT_BF16 ? mrf_fused_kernel<C, MB, S, EW, true, false> : mrf_fused_kernel<C, MB, S, EW, false, false>);
  static unsigned long long done[4] = {0, 0, 0, 0};  // one per (format, timing) twin of this instantiation
  cudaError_t e = set_smem_once(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES, done[(p.timing ? 2 : 0) + (p.fmt == ACT_BF16 ? 1 : 0)]);
  if (e != cudaSuccess) return e;
  FusedParams q = p;
  const int valid = Cfg::MT - 2 * p.margin;
  // (one ResBlock per launch only: the tail reads the running MRF sum 3 rows beyond the rows this tile owns, which
  //  must come from an earlier launch, not from a neighbouring CTA of this one)
  if (p.post_w && (Cfg::POST_BYTES == 0 || p.margin < 3 || !p.post_out || p.n_rb != 1)) return cudaErrorInvalidValue;
  if (valid < 128 || p.nd < 1 || p.nd > FUSED_MAX_D || p.n_rb < 1 || p.n_rb > FUSED_MAX_RB) return cudaErrorInvalidValue;
  for (int r = 0; r < p.n_rb; ++r)
    for (int m = 0; m < p.nd; ++m) {
      const FusedConv* cs[2] = {&p.rb[r].c1[m], &p.rb[r].c2[m]};
      for (const FusedConv* c : cs)
        if (c->ntaps < 1 || c->ntaps > S || (c->ntaps - 1) / 2 * c->dil > Cfg::PAD) return cudaErrorInvalidValue;
    }
  q.tiles_per_item = (p.L + valid - 1) / valid;
  const long long total = (long long)q.B * q.tiles_per_item;
  if (total <= 0 || total > 0x7fffffffLL) return cudaErrorInvalidValue;
  const int grid = (int)(total < num_sms ? total : num_sms);
  kern<<<grid, fused_threads(EW), Cfg::SMEM_BYTES, s>>>(q);
  return cudaGetLastError();
}

// (C, MB, S): C = 64 -> 3 x 128-row blocks, 13 tap slots (221 KB smem); C = 32 -> 8 blocks, 24 slots (188 KB)
bool mrf_fused_supported(int C, int max_taps, int max_halo) {
  if (max_halo > 32) return false;
  if (C == 64) return max_taps <= 13;
  if (C == 32) return max_taps <= 24;
  return false;
}
bool mrf_fused_post_supported(int C, int post_k) { return C == 32 && post_k == 7; }
int mrf_fused_tile_rows(int C) { return C == 64 ? 3 * 128 : C == 32 ? 8 * 128 : 0; }

cudaError_t launch_mrf_fused(const FusedParams& p, int num_sms, cudaStream_t s) {
  if (p.C == 64) return p.epi_warps == 8 ? launch_fused_inst<64, 3, 13, 8>(p, num_sms, s) : launch_fused_inst<64, 3, 13, 16>(p, num_sms, s);
  if (p.C == 32) return p.epi_warps == 8 ? launch_fused_inst<32, 8, 24, 8>(p, num_sms, s) : launch_fused_inst<32, 8, 24, 16>(p, num_sms, s);
  return cudaErrorInvalidValue;
}

}  // namespace rvcdec
