// Fused ResBlock pair for the 128-channel stage (residuals.py:71-80, one iteration of the loop):
//
//     x_out = x + conv2(lrelu(conv1_d(lrelu(x)) + b1)) + b2
//
// The per-conv path (conv_mma.cu) streams ~6 tensors per pair through HBM at this width (16-bit operand in /
// out of conv1, operand + fp32 residual in, fp32 + operand out of conv2) and ncu shows the conv2 launches
// pinned at the HBM roof (r1l: 65 % dram, 17-40 % tensor).  Here a pair reads the fp32 stream once and
// writes it once; the two operand tiles live in shared memory and the residual in TMEM:
//
//   TMEM   x[2][128 rows x 128 ch] fp32   residual; conv2 accumulates into it ("+x" is free)
//          acc1[2][128 x 128]      fp32   conv1 accumulators                                  (= all 512 columns)
//   smem   A        (256 + 2*32 pad rows) x 128 ch, blocked-8 K-major 16-bit (no-swizzle UMMA layout); ONE tile
//                   serves both convs (see below)
//          W ring   4 x 16 KB: one (tap, 64-channel chunk) block of the conv_mma weight image per stage
//          out      16 x 4 KB: per-epilogue-warp staging of the fp32 output rows, drained by bulk async stores
//                   (cp.async.bulk shared -> global).  Round 1 stored the output with STG from registers: every CTA
//                   reaches its store phase at the same time, the 128 KB-per-SM bursts run at the HBM rate and the
//                   epilogue warps sat in them for ~8k clk per tile (timing twin r1x: fin:stg) while the tensor pipe
//                   waited for the NEXT tile's conv1 epilogue from the same warps.  The TMA engine now drains the
//                   staging in the background; a warp only waits for its previous drain when it needs the staging again.
//
// A 128 x 128 tap block is 32 KB, so (unlike mrf_fused.cu) a conv's taps cannot stay resident: the loop is
// tap-outer / row-block-inner, each weight block feeds 8 MMAs (2 row blocks x 4 K steps) and both
// accumulators of a conv complete together.  Consequence: when conv1's epilogue runs, every MMA that reads
// lrelu(x) has retired, so conv2's operand overwrites it in place - and the 80 KB a second tile would take
// go to the weight ring instead (128 KB in flight; with 3 stages the ring could not cover the L2 latency at
// the 32 B/clk the tensor pipe consumes weights).  Tiles overlap by margin = halo(conv1) + halo(conv2) rows.
//
// Warps (576 threads): 0 = weight producer (TMA bulk copies), 1 = MMA issuer, 2-17 = epilogue
// (4 column groups of 32 channels x 4 TMEM lane quarters).
#include <algorithm>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

namespace pair {
constexpr int EW = 16;
constexpr int THREADS = 32 * (EW + 2);
constexpr int C = 128;
constexpr int MB = 2;
constexpr int MT = MB * 128;
constexpr int PAD = 32;
constexpr int ROWS = MT + 2 * PAD;
constexpr int NSL = C / 8;
constexpr int A_SLICE = ROWS * 16;
constexpr int A_BUF = NSL * A_SLICE;
constexpr int KC = 64;
constexpr int B_SLICE = C * 16;
constexpr int W_STAGE = (KC / 8) * B_SLICE;
constexpr int S = 4;
constexpr int ACC1_COL0 = MB * C;
constexpr int NUM_BARS = 2 * S + 3 * MB + 2;
constexpr int STG_WARP = 4 * 32 * 32;  // 4 slices x 32 rows x 8 fp32: the output rows of one epilogue warp and row block
constexpr int STG_BYTES = EW * STG_WARP;
constexpr int SMEM_BYTES = 128 + A_BUF + S * W_STAGE + STG_BYTES + NUM_BARS * 8 + 16;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
static_assert(ACC1_COL0 + MB * C <= 512, "TMEM budget");
constexpr int NC = C / (EW / 4);  // 32 accumulator columns per epilogue warp
constexpr int NS8 = NC / 8;
}  // namespace pair

namespace {
__device__ __forceinline__ uint32_t puni(uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); }
__device__ __forceinline__ uint32_t ppack2(float a, float b, bool bf16) {
  if (!bf16) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float plrelu(float v, float slope) { return fmaxf(v, v * slope); }
// shared -> global bulk async copy (TMA engine), tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the staging may be overwritten once every committed group has READ its source
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
}  // namespace

// TIMING = true: instrumented twin (tools/pair_timing.py); the MMA warp and epilogue warp 2 accumulate clock64
// spans into p.timing[blockIdx.x][16].
#define PT(slot_, ...)                                     \
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
__global__ void __launch_bounds__(pair::THREADS, 1) mrf_pair_kernel(const __grid_constant__ PairParams p) {
  using namespace pair;
  long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long t_begin = TIMING ? clock64() : 0;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 127u) & ~127u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t a0_smem = smem_base;
  const uint32_t a1_smem = a0_smem;  // conv2's operand overwrites conv1's in place (tap-outer order, see the header)
  const uint32_t w_smem = a0_smem + A_BUF;
  const uint32_t stg_smem = w_smem + S * W_STAGE;
  const uint32_t bars = stg_smem + STG_BYTES;
  const uint32_t w_full = bars, w_empty = w_full + 8 * S, a0_ready = w_empty + 8 * S, a1_ready = a0_ready + 8 * MB,
                 x_ready = a1_ready + 8 * MB, acc1_full = x_ready + 8 * MB, x_full = acc1_full + 8;
  const uint32_t tmem_slot = x_full + 8;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  // operand tiles start as zeros: the pad rows are the convolutions' zero padding at the tile edges
  for (int i = threadIdx.x; i < A_BUF / 16; i += THREADS) reinterpret_cast<uint4*>(smem_gen)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(w_full + 8 * i, 1); mbar_init(w_empty + 8 * i, 1); }
    for (int i = 0; i < MB; ++i) { mbar_init(a0_ready + 8 * i, EW); mbar_init(a1_ready + 8 * i, EW); mbar_init(x_ready + 8 * i, EW); }
    mbar_init(acc1_full, 1);
    mbar_init(x_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  const int total_work = p.B * p.tiles_per_item;

  if (warp == 0) {
    // ===================== weight producer =====================
    uint32_t slot = 0, ph = 0;
    for (int work = blockIdx.x; work < total_work; work += gridDim.x) {
#pragma unroll 1
      for (int which = 0; which < 2; ++which) {
        const uint8_t* src = reinterpret_cast<const uint8_t*>(which ? p.w2 : p.w1);
        const int nblk = (which ? p.k2 : p.k1) * (C / KC);
        for (int i = 0; i < nblk; ++i) {
          mbar_wait(w_empty + 8 * slot, ph ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(w_full + 8 * slot, W_STAGE);
            bulk_g2s(w_smem + slot * W_STAGE, src + (size_t)i * W_STAGE, W_STAGE, w_full + 8 * slot);
          }
          __syncwarp();
          if (++slot == S) { slot = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (converged warp, elected lane issues) =====================
    const uint32_t idesc = make_idesc(p.fmt, C);
    const uint64_t a0_desc = make_smem_desc(a0_smem, A_SLICE, 128) + (uint64_t)PAD;
    const uint64_t a1_desc = make_smem_desc(a1_smem, A_SLICE, 128) + (uint64_t)PAD;
    const uint64_t w_desc = make_smem_desc(w_smem, B_SLICE, 128);
    const uint32_t leader = elect_one() ? 1u : 0u;
    const uint32_t a_hi = (uint32_t)(a0_desc >> 32), w_lo = (uint32_t)w_desc, w_hi = (uint32_t)(w_desc >> 32);
    uint32_t slot = 0, ph = 0, seq = 0;
    for (int work = blockIdx.x; work < total_work; work += gridDim.x, ++seq) {
      const uint32_t par = seq & 1u;
#pragma unroll 1
      for (int which = 0; which < 2; ++which) {
        const int ntaps = (int)puni((uint32_t)(which ? p.k2 : p.k1)), dil = (int)puni((uint32_t)(which ? p.d2 : p.d1));
        const uint32_t in_ready = which ? a1_ready : a0_ready;
        // conv2 accumulates into x: besides its operand it needs the residual of THIS unit in TMEM (stored by the
        // epilogue warps behind the previous unit's output, while conv1 runs)
        PT(which, for (int b = 0; b < MB; ++b) { mbar_wait(in_ready + 8 * b, par); if (which) mbar_wait(x_ready + 8 * b, par); });
        tc_fence_after();
        const long long t_loop = TIMING ? clock64() : 0;
        const uint32_t d_base = puni(tmem_base + (which ? 0u : (uint32_t)ACC1_COL0));
        uint32_t a_tap = puni((uint32_t)(which ? a1_desc : a0_desc) + (uint32_t)(-((ntaps - 1) / 2) * dil));
        uint32_t fresh = which ? 1u : 0u;  // conv1 overwrites acc1 with its first MMA; conv2 accumulates into x
#pragma unroll 1
        for (int tap = 0; tap < ntaps; ++tap) {
#pragma unroll 1
          for (int kc = 0; kc < C / KC; ++kc) {
            PT(2, mbar_wait(w_full + 8 * slot, ph));
            tc_fence_after();
            const uint32_t b_lo = w_lo + slot * (uint32_t)(W_STAGE >> 4);
            const uint32_t a_kc = a_tap + (uint32_t)((kc * (KC / 8) * A_SLICE) >> 4);
#pragma unroll
            for (int b = 0; b < MB; ++b)
#pragma unroll
              for (int ks = 0; ks < KC / 16; ++ks)
                tc_mma_f16_pred_lh(d_base + b * C, a_kc + (uint32_t)(b * 128) + (uint32_t)((2 * ks * A_SLICE) >> 4), a_hi,
                                   b_lo + (uint32_t)((2 * ks * B_SLICE) >> 4), w_hi, idesc, ks == 0 ? fresh : 1u, leader);
            fresh = 1u;
            tc_commit_pred(w_empty + 8 * slot, leader);
            if (++slot == S) { slot = 0; ph ^= 1u; }
          }
          a_tap += (uint32_t)dil;
        }
        tc_commit_pred(which ? x_full : acc1_full, leader);
        if (TIMING) tacc[3] += clock64() - t_loop;
      }
    }
  } else {
    // ===================== epilogue warps =====================
    const int q = warp & 3;           // TMEM lane quarter this warp may access
    const int cg = (warp - 2) >> 2;   // column group (32 channels = 4 slices)
    const int r_in_blk = q * 32 + lane;
    const int sl0 = cg * NS8;
    const int valid_rows = MT - 2 * p.margin;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(cg * NC);
    const float slope = p.slope;

    auto publish = [&](uint32_t bar) {
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar);
    };
    auto store_row = [&](uint32_t a_smem, int r, const uint32_t (&pk)[NC / 2]) {
#pragma unroll
      for (int s8 = 0; s8 < NS8; ++s8) {
        const uint32_t dst = a_smem + (uint32_t)(sl0 + s8) * A_SLICE + (uint32_t)(PAD + r) * 16;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk[4 * s8]), "r"(pk[4 * s8 + 1]),
                     "r"(pk[4 * s8 + 2]), "r"(pk[4 * s8 + 3])
                     : "memory");
      }
    };
    struct Unit { int work, t0, item; size_t item_base; bool interior, valid; };
    auto make_unit = [&](int work) {
      Unit u;
      u.work = work; u.valid = work < total_work;
      const int bb = work / p.tiles_per_item;
      const int tile = work - bb * p.tiles_per_item;
      u.t0 = tile * valid_rows - p.margin;
      u.item = bb;
      u.item_base = (size_t)bb * NSL * p.Lp;
      u.interior = (u.t0 >= 0) && (u.t0 + MT <= p.L);
      return u;
    };
    auto load_x = [&](float4 (&dst)[NC / 4], const Unit& u, int r) {
      const int t = u.t0 + r;
      const bool ok = u.interior || (t >= 0 && t < p.L);
#pragma unroll
      for (int s8 = 0; s8 < NS8; ++s8) {
        dst[2 * s8] = make_float4(0.f, 0.f, 0.f, 0.f);
        dst[2 * s8 + 1] = dst[2 * s8];
        if (ok) {
          const float* src = p.x_in + ((u.item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8;
          dst[2 * s8] = *reinterpret_cast<const float4*>(src);
          dst[2 * s8 + 1] = *reinterpret_cast<const float4*>(src + 4);
        }
      }
    };
    auto prefetch_x = [&](const Unit& u, int r) {
      const int t = u.t0 + r;
      if (u.valid && t >= 0 && t < p.L) {
#pragma unroll
        for (int s8 = 0; s8 < NS8; ++s8)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(p.x_in + ((u.item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8));
      }
    };
    // x block (registers) -> lrelu(x) operand rows, then signal the MMA warp: conv1 can start
    auto store_x_operand = [&](const float4 (&cur)[NC / 4], int b) {
      const int r = b * 128 + r_in_blk;
      uint32_t pk[NC / 2];
#pragma unroll
      for (int i = 0; i < NC / 4; ++i) {
        pk[2 * i] = ppack2(plrelu(cur[i].x, slope), plrelu(cur[i].y, slope), BF16);
        pk[2 * i + 1] = ppack2(plrelu(cur[i].z, slope), plrelu(cur[i].w, slope), BF16);
      }
      store_row(a0_smem, r, pk);
      publish(a0_ready + 8 * b);
    };
    // x block (registers) -> TMEM residual, then signal the MMA warp: conv2 may accumulate into it
    auto store_x_residual = [&](const float4 (&cur)[NC / 4], int b) {
#pragma unroll
      for (int s8 = 0; s8 < NS8; ++s8) {
        uint32_t r8[8] = {__float_as_uint(cur[2 * s8].x), __float_as_uint(cur[2 * s8].y), __float_as_uint(cur[2 * s8].z),
                          __float_as_uint(cur[2 * s8].w), __float_as_uint(cur[2 * s8 + 1].x), __float_as_uint(cur[2 * s8 + 1].y),
                          __float_as_uint(cur[2 * s8 + 1].z), __float_as_uint(cur[2 * s8 + 1].w)};
        tc_st8(t_lane + b * C + s8 * 8, r8);
      }
      tc_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(x_ready + 8 * b);
    };

    Unit u = make_unit(blockIdx.x);
    if (u.valid) {
      const long long t_x = TIMING ? clock64() : 0;
#pragma unroll 1
      for (int b = 0; b < MB; ++b) {
        float4 cur[NC / 4];
        load_x(cur, u, b * 128 + r_in_blk);
        store_x_operand(cur, b);
        store_x_residual(cur, b);
      }
      if (TIMING) tacc[4] += clock64() - t_x;
    }
    uint32_t seq = 0;
    while (u.valid) {
      const Unit nx = make_unit(u.work + (int)gridDim.x);
      const uint32_t par = seq & 1u;
      const int t0 = u.t0;
      // ---- conv1 epilogue: A1 = lrelu(acc1 + b1), zero outside the sequence (conv2's zero padding) ----
      PT(0, mbar_wait(acc1_full, par));
      tc_fence_after();
      const long long t_e1 = TIMING ? clock64() : 0;
#pragma unroll 1
      for (int b = 0; b < MB; ++b) {
        const int r = b * 128 + r_in_blk;
        const int t = t0 + r;
        uint32_t rr[NC];
        tc_ld32(t_lane + ACC1_COL0 + b * C, rr);
        uint32_t pk[NC / 2];
        const float4* bp = reinterpret_cast<const float4*>(p.b1 + cg * NC);
#pragma unroll
        for (int i = 0; i < NC / 4; ++i) {
          const float4 bb = __ldg(bp + i);
          pk[2 * i] = ppack2(plrelu(__uint_as_float(rr[4 * i]) + bb.x, slope), plrelu(__uint_as_float(rr[4 * i + 1]) + bb.y, slope), BF16);
          pk[2 * i + 1] = ppack2(plrelu(__uint_as_float(rr[4 * i + 2]) + bb.z, slope), plrelu(__uint_as_float(rr[4 * i + 3]) + bb.w, slope), BF16);
        }
        if (!u.interior && (t < 0 || t >= p.L)) {
#pragma unroll
          for (int i = 0; i < NC / 2; ++i) pk[i] = 0u;
        }
        store_row(a1_smem, r, pk);
        publish(a1_ready + 8 * b);
      }
      // L2 prefetch of what the final epilogue below loads (behind the publishes: a proxy fence waits for
      // outstanding global accesses of the thread)
#pragma unroll 1
      for (int b = 0; b < MB; ++b) {
        prefetch_x(nx, b * 128 + r_in_blk);
        const int r = b * 128 + r_in_blk, t = t0 + r;
        if (p.sum_in && r >= p.margin && r < MT - p.margin && t >= 0 && t < p.L) {
#pragma unroll
          for (int s8 = 0; s8 < NS8; ++s8)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.sum_in + ((u.item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8));
        }
      }
      if (TIMING) tacc[1] += clock64() - t_e1;
      // ---- next unit's operand first (critical path), then this unit's output (behind the next conv1) ----
      // All CTAs run the same schedule, so their global-memory phases coincide and each 64 KB-per-SM burst of
      // stores is HBM-bound while it lasts (timing twin r1s/r1v: ~3.8k clk per block).  The tensor pipe must not
      // wait for that: as soon as conv2 has retired, the next unit's lrelu(x) rows (L2-prefetched) are published
      // and conv1 of the next unit starts; only then is x read out of TMEM and stored, and the next unit's
      // residual written into TMEM (conv2 of the next unit waits for it on x_ready).
      const long long t_e2 = TIMING ? clock64() : 0;
      // next unit's operand rows, packed in registers while conv2 is still running
      uint32_t pkn[MB][NC / 2];
      if (nx.valid) {
#pragma unroll
        for (int b = 0; b < MB; ++b) {
          float4 xn[NC / 4];
          load_x(xn, nx, b * 128 + r_in_blk);
#pragma unroll
          for (int i = 0; i < NC / 4; ++i) {
            pkn[b][2 * i] = ppack2(plrelu(xn[i].x, slope), plrelu(xn[i].y, slope), BF16);
            pkn[b][2 * i + 1] = ppack2(plrelu(xn[i].z, slope), plrelu(xn[i].w, slope), BF16);
          }
        }
      }
      PT(2, mbar_wait(x_full, par));
      tc_fence_after();
      if (nx.valid) {
        const long long t_p = TIMING ? clock64() : 0;
#pragma unroll
        for (int b = 0; b < MB; ++b) {
          store_row(a0_smem, b * 128 + r_in_blk, pkn[b]);
          publish(a0_ready + 8 * b);
        }
        if (TIMING) tacc[6] += clock64() - t_p;
      }
#pragma unroll 1
      for (int b = 0; b < MB; ++b) {
        const int r = b * 128 + r_in_blk;
        const int t = t0 + r;
        const bool ok = u.interior || ((t >= 0) && (t < p.L));
        const bool store_out = ok && r >= p.margin && r < MT - p.margin;
        float4 xn[NC / 4];
        if (nx.valid) load_x(xn, nx, r);  // second read of these rows (L1 / L2 hit)
        uint32_t rr[NC];
        PT(5, tc_ld32(t_lane + b * C, rr));
        // the next unit's residual goes into TMEM before this unit's stores are issued (conv2 of the next unit
        // waits for it; the stores may take several thousand clocks)
        if (nx.valid) store_x_residual(xn, b);
        const long long t_st = TIMING ? clock64() : 0;
        {
          const int tq = t - lane;  // time of lane 0's row (the warp's 32 rows are consecutive)
          // rows of this warp that are stored: every condition is an interval in the row index, so they are contiguous
          const uint32_t smask = __ballot_sync(0xffffffffu, store_out);
          const int lo = smask ? (__ffs((int)smask) - 1) : 0, cnt = __popc(smask);
          const uint32_t my_stg = stg_smem + (uint32_t)(warp - 2) * STG_WARP;
          if (p.out_f32) {  // the previous drain of this staging area must have read it
            if (lane == 0) bulk_wait_read();
            __syncwarp();
          }
#pragma unroll
          for (int s8 = 0; s8 < NS8; ++s8) {
            const size_t off = ((u.item_base + (size_t)(sl0 + s8) * p.Lp) + PADL + t) * 8;
            float o[8];
            if (store_out) {
              const float4 c0 = __ldg(reinterpret_cast<const float4*>(p.b2 + cg * NC + s8 * 8));
              const float4 c1 = __ldg(reinterpret_cast<const float4*>(p.b2 + cg * NC + s8 * 8 + 4));
              o[0] = (__uint_as_float(rr[8 * s8 + 0]) + c0.x) * p.scale; o[1] = (__uint_as_float(rr[8 * s8 + 1]) + c0.y) * p.scale;
              o[2] = (__uint_as_float(rr[8 * s8 + 2]) + c0.z) * p.scale; o[3] = (__uint_as_float(rr[8 * s8 + 3]) + c0.w) * p.scale;
              o[4] = (__uint_as_float(rr[8 * s8 + 4]) + c1.x) * p.scale; o[5] = (__uint_as_float(rr[8 * s8 + 5]) + c1.y) * p.scale;
              o[6] = (__uint_as_float(rr[8 * s8 + 6]) + c1.z) * p.scale; o[7] = (__uint_as_float(rr[8 * s8 + 7]) + c1.w) * p.scale;
              if (p.sum_in) {
                const float4 s0 = *reinterpret_cast<const float4*>(p.sum_in + off);
                const float4 s1 = *reinterpret_cast<const float4*>(p.sum_in + off + 4);
                o[0] += s0.x; o[1] += s0.y; o[2] += s0.z; o[3] += s0.w;
                o[4] += s1.x; o[5] += s1.y; o[6] += s1.z; o[7] += s1.w;
              }
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i) o[i] = 0.f;
            }
            if (p.out_f32) {  // rows in the global layout: [row][8 fp32] per slice
              const uint32_t dst = my_stg + (uint32_t)(s8 * 1024 + lane * 32);
              asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "f"(o[0]), "f"(o[1]), "f"(o[2]), "f"(o[3]) : "memory");
              asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 16), "f"(o[4]), "f"(o[5]), "f"(o[6]), "f"(o[7]) : "memory");
            }
            if (p.out_act && store_out) {
              uint32_t pk[4];
#pragma unroll
              for (int i = 0; i < 4; ++i)
                pk[i] = ppack2(plrelu(o[2 * i], p.out_slope), plrelu(o[2 * i + 1], p.out_slope), p.out_fmt == ACT_BF16);
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
          if (p.out_f32) {
            fence_proxy_async();  // generic-proxy writes of the staging -> visible to the bulk copy engine
            __syncwarp();
            if (lane == 0 && cnt > 0) {
#pragma unroll
              for (int s8 = 0; s8 < NS8; ++s8) {
                // (pointer arithmetic in ptrdiff_t: tq may be negative at a sequence start, lo moves it to a valid row)
                float* g_dst = p.out_f32 + ((ptrdiff_t)((u.item_base + (size_t)(sl0 + s8) * p.Lp) + PADL) + tq + lo) * 8;
                bulk_s2g(g_dst, my_stg + (uint32_t)(s8 * 1024 + lo * 32), (uint32_t)cnt * 32u);
              }
              bulk_commit();
            }
          }
        }
        if (TIMING) tacc[7] += clock64() - t_st;
      }
      if (TIMING) tacc[3] += clock64() - t_e2;
      u = nx;
      ++seq;
    }
    if (lane == 0) bulk_wait_all();  // the staging must outlive its drains
  }
  if (TIMING && p.timing && (threadIdx.x == 32 || threadIdx.x == 64)) {
    // MMA warp: [0] wait operand (conv1), [1] wait operand (conv2), [2] wait weights, [3] tap loops incl. [2], [5] total
    // epilogue warp 2: [8] wait acc1, [9] conv1 epilogue, [10] wait x, [11] final epilogue incl. [10], [12] first x load, [13] total
    long long* o = p.timing + (size_t)blockIdx.x * 16 + (threadIdx.x == 32 ? 0 : 8);
    for (int i = 0; i < 5; ++i) o[i] = tacc[i];
    o[5] = clock64() - t_begin;
    if (threadIdx.x == 64) { o[-2] = tacc[5]; o[-1] = tacc[6]; o[6] = tacc[7]; }  // final: tmem ld, x store + publish, global stores
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ------------------------------ host side ----------------------------------
bool mrf_pair_supported(int C, int halo1, int halo2) {
  return C == pair::C && halo1 <= pair::PAD && halo2 <= pair::PAD && pair::MT - 2 * (halo1 + halo2) >= 128;
}

cudaError_t launch_mrf_pair(const PairParams& p, int num_sms, cudaStream_t s) {
  using namespace pair;
  auto kern = p.timing ? (p.fmt == ACT_BF16 ? mrf_pair_kernel<true, true> : mrf_pair_kernel<false, true>)
                       : (p.fmt == ACT_BF16 ? mrf_pair_kernel<true, false> : mrf_pair_kernel<false, false>);
  static unsigned long long done[4] = {0, 0, 0, 0};
  cudaError_t e = set_smem_once(reinterpret_cast<const void*>(kern), SMEM_BYTES, done[(p.timing ? 2 : 0) + (p.fmt == ACT_BF16 ? 1 : 0)]);
  if (e != cudaSuccess) return e;
  const int valid = MT - 2 * p.margin;
  if (valid < 128 || p.k1 < 1 || p.k2 < 1 || (p.k1 - 1) / 2 * p.d1 > PAD || (p.k2 - 1) / 2 * p.d2 > PAD ||
      p.margin < (p.k1 - 1) / 2 * p.d1 + (p.k2 - 1) / 2 * p.d2)
    return cudaErrorInvalidValue;
  PairParams q = p;
  q.tiles_per_item = (p.L + valid - 1) / valid;
  const long long total = (long long)q.B * q.tiles_per_item;
  if (total <= 0 || total > 0x7fffffffLL) return cudaErrorInvalidValue;
  const int grid = (int)(total < num_sms ? total : num_sms);
  kern<<<grid, THREADS, SMEM_BYTES, s>>>(q);
  return cudaGetLastError();
}

}  // namespace rvcdec
