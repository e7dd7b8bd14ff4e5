// Tensor-core back-end: every convolution of the decoder as a tcgen05 implicit GEMM (sm_100a).
//
//   D[t, n] = sum_tap sum_ci A[t + shift[tap], ci] * W[tap][ci][n]      (fp32 accumulate in TMEM)
//
// CTA = 18 warps (10 with EW = 8), warp-specialised, persistent over (item, time-tile, N-tile):
//   warp 0   TMA producer   1-D bulk copies (cp.async.bulk, UBLKCP) of activation slices and
//                           weight blocks into shared memory, completion on mbarriers
//   warp 1   MMA issuer     one thread issues tcgen05.mma (kind::f16, M=128, N=N_TILE, K=16),
//                           tcgen05.commit releases smem stages / publishes the accumulator
//   warps 2-17 epilogue     tcgen05.ld accumulator -> registers -> fused epilogue (bias, residual,
//                           MRF average, leaky_relu, 16-bit convert) -> coalesced global stores;
//                           16 warps over 16-column items (EW = 16, default) or 8 over 32-column items;
//                           every global load of an item is issued before the (volatile) TMEM load
//   PM = 1                  phase-major polyphase upsample tiles with a warp-transposed epilogue (see below)
//
// Shared-memory operand layout: NO swizzle, K-major "interleaved" core matrices.  The blocked-8
// HBM layout ([slice][row][8 ch]) copied verbatim IS that layout: 8 rows x 16 B core matrices,
// consecutive 8-row groups 128 B apart (SBO), the two K halves of an MMA one slice apart (LBO).
// A convolution tap is then a +/- n*16 B change of the descriptor start address - any dilation,
// no 1024 B swizzle-atom alignment to respect - and the tile (+halo) is loaded once per K chunk
// and reused by every tap.
//
// Time tile = MB x 128 rows (MB accumulators side by side in TMEM) so that one weight block
// fetched from L2 feeds MB MMAs; accumulators are double-buffered in TMEM when they fit
// (2*MB*N_TILE <= 512 columns) so the epilogue of tile i overlaps the MMAs of tile i+1.
#include <algorithm>
#include <cstdio>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace rvcdec {

// packed fp32x2 arithmetic (sm_100 FADD2 / FFMA2): halves the phase-major epilogue's FP instruction count
__device__ __forceinline__ uint64_t f2_pack_u(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ uint64_t f2_splat(float v) { return f2_pack_u(__float_as_uint(v), __float_as_uint(v)); }
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

// ------------------------------ kernel -------------------------------------
template <int N_TILE, int MB, int KC, int B_STAGES, int A_STAGES = 2, int PM = 0, int EW = 8>
struct MmaCfg {
  static constexpr int MT = MB * 128;
  // PM = 1: phase-major polyphase upsample tile (see the PM epilogue below): N_TILE = PH phases x 2 slices x 8
  // channels; every epilogue warp owns a staging area of 32 x (2*PH*32 + 16) bytes for the warp-level transpose
  // PM = 2: stride-2 polyphase upsample tile, columns (slice, phase, channel): every epilogue warp transposes its 32 rows x
  // 64 B through a staging area of 32 x 80 B (16 B of padding per lane keeps the quarter-warp stores conflict-free)
  static constexpr int PH = N_TILE / 16;
  static constexpr int STG_LANE = PM == 2 ? 80 : PH * 32 + 16;
  static constexpr int STG_WARP = PM ? 32 * STG_LANE : 0;
  static constexpr int STG_BYTES = 16 * STG_WARP;
  static constexpr int CONST_BYTES = PM == 1 ? (256 + 8 * 128) * 4 : 0;  // per-channel bias (<= 256) + noise-conv taps [8][<= 128]
  static constexpr int EPI_WARPS = PM ? 16 : EW;        // 8, or 16 (items of 16 columns: twice the warps to hide the epilogue's latencies)
  static constexpr int CW = EPI_WARPS == 16 ? 16 : 32;  // accumulator columns per epilogue work item (PM != 1 paths)
  static constexpr int THREADS = 64 + EPI_WARPS * 32;   // producer warp + MMA warp + epilogue warps
  static_assert(EW == 8 || EW == 16, "8 or 16 epilogue warps");
  static constexpr int ACC_STAGES = (2 * MB * N_TILE <= 512) ? 2 : 1;
  static constexpr int ACC_COLS = ACC_STAGES * MB * N_TILE;
  static constexpr int TMEM_COLS = ACC_COLS <= 32 ? 32 : ACC_COLS <= 64 ? 64 : ACC_COLS <= 128 ? 128 : ACC_COLS <= 256 ? 256 : 512;
  static constexpr int A_HALO = PM ? 8 : PADL;  // rows of halo an activation stage has room for (upsamples shift by +-1)
  static constexpr int A_ROWS = MT + 2 * A_HALO;
  static constexpr int A_SLICE_BYTES = A_ROWS * 16;
  static constexpr int A_BUF_BYTES = (KC / 8) * A_SLICE_BYTES;
  static constexpr int B_SLICE_BYTES = N_TILE * 16;
  static constexpr int B_STAGE_BYTES = (KC / 8) * B_SLICE_BYTES;
  static constexpr int NUM_BARS = 2 * A_STAGES + 2 * B_STAGES + 2 * ACC_STAGES;
  static constexpr int SMEM_BYTES = 128 /*align slack*/ + A_STAGES * A_BUF_BYTES + B_STAGES * B_STAGE_BYTES + STG_BYTES + CONST_BYTES + NUM_BARS * 8 + 16;
  static_assert(ACC_COLS <= 512, "accumulators exceed TMEM");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

constexpr int MMA_THREADS = 320;  // producer warp + MMA warp + 2 epilogue warpgroups

// LEAN = 1: ordinary convolutions only (up == 1, no fused noise conv, dense 16-bit output) - the 36 ResBlock launches of
// the 256- / 128-channel stages.  The epilogue warps run at the 96-register cap of an 18-warp CTA, and the registers the
// noise-conv / zero-stuffing / phase-scatter code keeps alive cost the general kernel 24 B of spills per thread.
template <int N_TILE, int MB, int KC, int B_STAGES, int A_STAGES, int PM, int EW, int LEAN>
__global__ void __launch_bounds__((PM || EW == 16) ? 576 : MMA_THREADS, 1) conv_mma_kernel(const __grid_constant__ ConvParams p) {
  using Cfg = MmaCfg<N_TILE, MB, KC, B_STAGES, A_STAGES, PM, EW>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 127u) & ~127u;
  const uint32_t a_smem = smem_base;
  const uint32_t b_smem = a_smem + A_STAGES * Cfg::A_BUF_BYTES;
  const uint32_t stg_smem = b_smem + B_STAGES * Cfg::B_STAGE_BYTES;
  const uint32_t const_smem = stg_smem + Cfg::STG_BYTES;
  const uint32_t bars = const_smem + Cfg::CONST_BYTES;
  const uint32_t a_full = bars, a_empty = bars + 8 * A_STAGES, b_full = bars + 16 * A_STAGES, b_empty = b_full + 8 * B_STAGES,
                 acc_full = b_empty + 8 * B_STAGES, acc_empty = acc_full + 8 * Cfg::ACC_STAGES;
  const uint32_t tmem_slot = acc_empty + 8 * Cfg::ACC_STAGES;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);  // provably warp-uniform role index
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < A_STAGES; ++i) { mbar_init(a_full + 8 * i, 1); mbar_init(a_empty + 8 * i, 1); }
    for (int i = 0; i < B_STAGES; ++i) { mbar_init(b_full + 8 * i, 1); mbar_init(b_empty + 8 * i, 1); }
    for (int i = 0; i < Cfg::ACC_STAGES; ++i) { mbar_init(acc_full + 8 * i, 1); mbar_init(acc_empty + 8 * i, Cfg::THREADS - 64); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "r"((uint32_t)Cfg::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if constexpr (PM == 1) {
    // per-channel bias (+ noise-conv bias) and the noise-conv taps live in shared memory: the strided har reads of
    // the epilogue sweep the small L1 (most of it is carved out as shared memory), so uniform global loads of these
    // constants kept missing it (profiles/r1bc: the first use of each was a long-scoreboard stall)
    float* cst = reinterpret_cast<float*>(smem_gen + (const_smem - smem_base));
    const Epilogue& e = p.epi;
    const int PHc = Cfg::PH;
    for (int c = threadIdx.x; c < e.cout; c += Cfg::THREADS) {
      const int n_new = (c >> 4) * 2 * N_TILE + (((c >> 3) & 1) * PHc) * 8 + (c & 7);
      cst[c] = __ldg(e.bias + n_new) + (e.nz_har ? __ldg(e.nz_bias + c) : 0.f);
    }
    if (e.nz_har)
      for (int i = threadIdx.x; i < 8 * e.cout; i += Cfg::THREADS) cst[256 + i] = __ldg(e.nz_w + i);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  const int nkc = p.Cin / KC;
  const int nsl = p.Cin >> 3;
  const int tiles_per_item = p.m_tiles * p.n_tiles;
  const int total_tiles = p.B * tiles_per_item;
  const int rows = Cfg::MT + 2 * p.halo;

  if (warp == 0) {
    // ===================== TMA producer (converged warp, one elected lane issues) =====================
    uint32_t a_stage = 0, a_phase = 0, b_stage = 0, b_phase = 0;
    const uint8_t* act = reinterpret_cast<const uint8_t*>(p.act_in);
    const uint8_t* wgt = reinterpret_cast<const uint8_t*>(p.w);
    const int n_tiles = p.n_tiles, ntaps = p.ntaps;
    const uint32_t a_bytes = (uint32_t)rows * 16;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int b = tile / tiles_per_item;
      const int rem = tile - b * tiles_per_item;
      const int m_tile = rem / n_tiles;
      const int n_tile = rem - m_tile * n_tiles;
      const int t0 = m_tile * Cfg::MT;
      const uint32_t mask = p.tapmask[n_tile];
      const uint8_t* a_src = act + (((size_t)b * (p.in_slices_phys ? p.in_slices_phys : nsl)) * p.Lp_in + PADL + t0 - p.halo) * 16;
      const uint8_t* w_src = wgt + (size_t)n_tile * ntaps * nkc * Cfg::B_STAGE_BYTES;
      for (int kc = 0; kc < nkc; ++kc) {
        mbar_wait(a_empty + 8 * a_stage, a_phase ^ 1u);
        if (elect_one()) {
          mbar_expect_tx(a_full + 8 * a_stage, (uint32_t)(KC / 8) * a_bytes);
#pragma unroll
          for (int sl = 0; sl < KC / 8; ++sl)
            bulk_g2s(a_smem + a_stage * Cfg::A_BUF_BYTES + sl * Cfg::A_SLICE_BYTES,
                     a_src + (size_t)(kc * (KC / 8) + sl) * p.Lp_in * 16, a_bytes, a_full + 8 * a_stage);
        }
        __syncwarp();
        if (++a_stage == A_STAGES) { a_stage = 0; a_phase ^= 1u; }
        for (int tap = 0; tap < ntaps; ++tap) {
          if (!((mask >> tap) & 1u)) continue;
          mbar_wait(b_empty + 8 * b_stage, b_phase ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(b_full + 8 * b_stage, Cfg::B_STAGE_BYTES);
            bulk_g2s(b_smem + b_stage * Cfg::B_STAGE_BYTES, w_src + (size_t)(tap * nkc + kc) * Cfg::B_STAGE_BYTES,
                     Cfg::B_STAGE_BYTES, b_full + 8 * b_stage);
          }
          __syncwarp();
          if (++b_stage == B_STAGES) { b_stage = 0; b_phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp runs the loop converged (so every descriptor stays in uniform registers and
    // ptxas emits no per-MMA "elect" waterfall); one elected lane issues the MMAs and the commits.
    // ncu r1c: the former `if (lane == 0)` form spent ~120 clk of issue overhead per MMA and was the
    // bottleneck of every convolution (the tensor pipe needs one N=128 MMA per 64 clk).
    uint32_t a_stage = 0, a_phase = 0, b_stage = 0, b_phase = 0, acc_stage = 0, acc_phase = 0;
    const uint32_t idesc = make_idesc(p.act_fmt_in, N_TILE);
    const uint64_t a_desc0 = make_smem_desc(a_smem, Cfg::A_SLICE_BYTES, 128);
    const uint64_t b_desc0 = make_smem_desc(b_smem, Cfg::B_SLICE_BYTES, 128);
    const int n_tiles = p.n_tiles, ntaps = p.ntaps;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const uint32_t mask = p.tapmask[tile % n_tiles];
      mbar_wait(acc_empty + 8 * acc_stage, acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_base = tmem_base + acc_stage * (MB * N_TILE);
      uint32_t fresh = 0u;  // 0 for the very first MMA of the tile (overwrite), 1 afterwards
      for (int kc = 0; kc < nkc; ++kc) {
        mbar_wait(a_full + 8 * a_stage, a_phase);
        const uint64_t a_kc = a_desc0 + (uint64_t)((a_stage * Cfg::A_BUF_BYTES) >> 4) + (uint64_t)p.halo;
        for (int tap = 0; tap < ntaps; ++tap) {
          if (!((mask >> tap) & 1u)) continue;
          mbar_wait(b_full + 8 * b_stage, b_phase);
          tc_fence_after();
          if (elect_one()) {
            // one row of the blocked layout is 16 B, so a tap shift is +shift in descriptor units
            const uint64_t a_tap = a_kc + (uint64_t)(int64_t)p.shift[tap];
            const uint64_t b_st = b_desc0 + (uint64_t)((b_stage * Cfg::B_STAGE_BYTES) >> 4);
#pragma unroll
            for (int mb = 0; mb < MB; ++mb) {
#pragma unroll
              for (int ks = 0; ks < KC / 16; ++ks) {
                tc_mma_f16(d_base + mb * N_TILE, a_tap + (uint64_t)((mb * 128 * 16 + 2 * ks * Cfg::A_SLICE_BYTES) >> 4),
                           b_st + (uint64_t)((2 * ks * Cfg::B_SLICE_BYTES) >> 4), idesc, (ks == 0) ? fresh : 1u);
              }
            }
            tc_commit(b_empty + 8 * b_stage);
          }
          __syncwarp();
          fresh = 1u;
          if (++b_stage == B_STAGES) { b_stage = 0; b_phase ^= 1u; }
        }
        if (elect_one()) tc_commit(a_empty + 8 * a_stage);
        __syncwarp();
        if (++a_stage == A_STAGES) { a_stage = 0; a_phase ^= 1u; }
      }
      if (elect_one()) tc_commit(acc_full + 8 * acc_stage);
      __syncwarp();
      if (++acc_stage == Cfg::ACC_STAGES) { acc_stage = 0; acc_phase ^= 1u; }
    }
  } else {
    // ===================== epilogue warps (2 warpgroups) =====================
    const int q = warp & 3;          // TMEM lane quarter this warp may access
    const int wg = (warp - 2) >> 2;  // work items (mb, CW-column chunk) are dealt to the warpgroups in turn
    constexpr int CW = Cfg::CW;
    constexpr int WGS = Cfg::EPI_WARPS / 4;
    constexpr int CH = N_TILE / CW;
    constexpr int ITEMS = MB * CH;
    const Epilogue& e = p.epi;
    const bool fast = LEAN ? true : (e.up == 1);
    const size_t slice_stride = (size_t)e.Lp_out * 8;
    uint32_t acc_it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int b = tile / tiles_per_item;
      const int rem = tile - b * tiles_per_item;
      const int m_tile = rem / p.n_tiles;
      const int n_tile = rem - m_tile * p.n_tiles;
      const int t0 = m_tile * Cfg::MT;
      const uint32_t as = acc_it % Cfg::ACC_STAGES;
      // phase-major upsample: this row's har samples are requested before the accumulator wait (the epilogue has few
      // warps, so every exposed global-load latency counts).  A/B (profiles/r1be-r1bh): plain per-lane 8 B loads (lane
      // stride up*16 B) beat staging the warp's contiguous window through shared memory - the L1 data pipe, which
      // also carries the transpose, is the unit that saturates (80 %).
      [[maybe_unused]] float hv[PM == 1 ? 4 * (N_TILE / 16) + 4 : 1];
      if constexpr (PM == 1) {
        static_assert(MB == 2, "phase-major epilogue: one item per warp and tile");
        constexpr int PH = Cfg::PH, U = 2 * PH, NZS = 4, NZK = 8, NH = NZS * PH + NZK - NZS;
        if (e.nz_har) {
          const int t_in = t0 + (wg & 1) * 128 + q * 32 + lane;
          const float* hb = e.nz_har + (size_t)b * e.nz_src;
          const int s0 = (t_in * U + (n_tile & 1) * PH) * NZS - (NZK - NZS) / 2;  // even
          constexpr int NR4 = (NH + 2 + 3) / 4;
          if (t_in < p.L_in && s0 >= 2 && s0 - 2 + NR4 * 4 <= e.nz_src) {
            // s0 - 2 is a multiple of 4 samples: 16 B loads (7 instead of 12 requests per lane on the L1 data pipe)
            const float4* h4p = reinterpret_cast<const float4*>(hb + s0 - 2);
            float raw[NR4 * 4];
#pragma unroll
            for (int j = 0; j < NR4; ++j) {
              const float4 h4 = __ldg(h4p + j);
              raw[4 * j] = h4.x; raw[4 * j + 1] = h4.y; raw[4 * j + 2] = h4.z; raw[4 * j + 3] = h4.w;
            }
#pragma unroll
            for (int j = 0; j < NH; ++j) hv[j] = raw[j + 2];
          } else {
#pragma unroll
            for (int j = 0; j < NH / 2; ++j) {
              const int sidx = s0 + 2 * j;
              float2 h2 = make_float2(0.f, 0.f);
              if (t_in < p.L_in && sidx >= 0 && sidx < e.nz_src) h2 = __ldg(reinterpret_cast<const float2*>(hb + sidx));
              hv[2 * j] = h2.x; hv[2 * j + 1] = h2.y;
            }
          }
        }
      }
      mbar_wait(acc_full + 8 * as, (acc_it / Cfg::ACC_STAGES) & 1u);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + ((uint32_t)(q * 32) << 16) + as * (MB * N_TILE);
      if constexpr (PM == 2) {
        // ---- stride-2 polyphase upsample (ConvTranspose1d k = 4, stride 2; hifigan_nsf.py:198 [+ :199]) ----
        // GEMM columns are ordered (slice, phase, channel): the 16 columns of a work item are output rows 2t and 2t+1
        // of one slice for this thread's input row t, i.e. 64 contiguous bytes of fp32, and a warp owns a 2 KB run.
        // No zeros are read or multiplied (the zero-stuffed form did both) and the previous stage writes a dense 16-bit
        // tensor.  Stored straight from the registers, the 64 B lane stride makes every STG.128 touch 16 cache lines
        // (measured r2d: 26 ms against 17 ms for the zero-stuffed form), so the warp transposes the run through its
        // staging area and every store instruction writes 512 contiguous bytes.  fp32 output only: the fused MRF stage
        // that follows reads nothing else.
        static_assert(Cfg::CW == 16, "stride-2 polyphase epilogue works on 16-column items");
        constexpr int CH2 = N_TILE / 16;
        constexpr int ITEMS2 = MB * CH2;
        constexpr int WGS2 = Cfg::EPI_WARPS / 4;
        const uint32_t my_stg = stg_smem + (uint32_t)(warp - 2) * Cfg::STG_WARP;
#pragma unroll 1
        for (int item = wg; item < ITEMS2; item += WGS2) {
          const int mb = item / CH2;
          const int c0 = (item - mb * CH2) * 16;
          const int tb = t0 + mb * 128 + q * 32;
          const int t_in = tb + lane;
          const bool valid = t_in < p.L_in;
          const int n0 = n_tile * N_TILE + c0;
          const int co = (n0 >> 4) * 8;
          // har samples under the two output rows' noise convs (<= 8 taps), requested before the accumulator wait
          float hp[2][8];
          if (e.nz_har) {
            const float* hb = e.nz_har + (size_t)b * e.nz_src;
#pragma unroll
            for (int ph = 0; ph < 2; ++ph) {
              const int s0 = (2 * t_in + ph) * e.nz_stride - e.nz_pad;
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const int sidx = s0 + j;
                hp[ph][j] = (valid && j < e.nz_k && sidx >= 0 && sidx < e.nz_src) ? __ldg(hb + sidx) : 0.f;
              }
            }
          }
          float4 bb[4];
          {
            const float4* bp = reinterpret_cast<const float4*>(e.bias + n0);
#pragma unroll
            for (int i = 0; i < 4; ++i) bb[i] = __ldg(bp + i);
          }
          uint32_t r[16];
          tc_ld16(t_acc + mb * N_TILE + c0, r);
#pragma unroll
          for (int ph = 0; ph < 2; ++ph) {
            float v[8];
            const float4 b0 = bb[2 * ph], b1 = bb[2 * ph + 1];
            v[0] = __uint_as_float(r[ph * 8 + 0]) + b0.x; v[1] = __uint_as_float(r[ph * 8 + 1]) + b0.y;
            v[2] = __uint_as_float(r[ph * 8 + 2]) + b0.z; v[3] = __uint_as_float(r[ph * 8 + 3]) + b0.w;
            v[4] = __uint_as_float(r[ph * 8 + 4]) + b1.x; v[5] = __uint_as_float(r[ph * 8 + 5]) + b1.y;
            v[6] = __uint_as_float(r[ph * 8 + 6]) + b1.z; v[7] = __uint_as_float(r[ph * 8 + 7]) + b1.w;
            if (e.nz_har) {
              const float4 nb0 = __ldg(reinterpret_cast<const float4*>(e.nz_bias + co));
              const float4 nb1 = __ldg(reinterpret_cast<const float4*>(e.nz_bias + co + 4));
              v[0] += nb0.x; v[1] += nb0.y; v[2] += nb0.z; v[3] += nb0.w;
              v[4] += nb1.x; v[5] += nb1.y; v[6] += nb1.z; v[7] += nb1.w;
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                if (j < e.nz_k) {
                  const float hj = hp[ph][j];
                  const float4 w0 = __ldg(reinterpret_cast<const float4*>(e.nz_w + (size_t)j * e.cout + co));
                  const float4 w1 = __ldg(reinterpret_cast<const float4*>(e.nz_w + (size_t)j * e.cout + co + 4));
                  v[0] = fmaf(hj, w0.x, v[0]); v[1] = fmaf(hj, w0.y, v[1]); v[2] = fmaf(hj, w0.z, v[2]); v[3] = fmaf(hj, w0.w, v[3]);
                  v[4] = fmaf(hj, w1.x, v[4]); v[5] = fmaf(hj, w1.y, v[5]); v[6] = fmaf(hj, w1.z, v[6]); v[7] = fmaf(hj, w1.w, v[7]);
                }
              }
            }
            const uint32_t dst = my_stg + (uint32_t)(lane * Cfg::STG_LANE + ph * 32);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 16), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
          }
          __syncwarp();
          // staging -> global: piece pp = 32 k + lane of the warp's 128 16-byte pieces belongs to thread pp / 4
          float* run0 = e.out_f32 + (((size_t)b * (e.cout >> 3) + (n0 >> 4)) * e.Lp_out + PADL + 2 * (size_t)tb) * 8;
          const int rows_ok = min(32, p.L_in - tb);  // threads [0, rows_ok) hold real rows
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int pp = 32 * k + lane;
            float4 f;
            const uint32_t src = my_stg + (uint32_t)((pp >> 2) * Cfg::STG_LANE + (pp & 3) * 16);
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f.x), "=f"(f.y), "=f"(f.z), "=f"(f.w) : "r"(src) : "memory");
            if ((pp >> 2) < rows_ok) *reinterpret_cast<float4*>(run0 + pp * 4) = f;
          }
          __syncwarp();
        }
      } else if constexpr (PM == 1) {
        // ---- phase-major polyphase upsample (ConvTranspose1d, hifigan_nsf.py:198) [+ noise_convs[i](har), :199] ----
        // N tile n_tile = (channel group cg of 16) * 2 + half; its columns are ordered (slice sl of the group, phase rr
        // of this half, channel c): column (sl*PH + rr)*8 + c is output row t_in*up + half*PH + rr of slice 2*cg + sl.
        // A thread (= input row) therefore owns PH consecutive output rows of a slice (PH*32 B in fp32), and a warp
        // owns 32 such runs.  The warp transposes them through its shared-memory staging area so that every global
        // store instruction writes consecutive 16 B (fp32) / 8 B (16-bit) pieces: the former one-phase-per-tile
        // scatter wrote 32 B pieces 32*up bytes apart and ran at 1.3 TB/s (profiles/r1at id 20).
        // 16 epilogue warps: (TMEM lane quarter q) x (row block mb) x (slice sl); the epilogue is a long chain of
        // dependent loads and FMAs per warp, so the warps are what hides its latency (profiles/r1bb: 8 warps with
        // both slices each ran at ~10 clk per instruction).
        constexpr int PH = Cfg::PH;
        constexpr int NC = PH * 8;       // columns (= floats per thread) of one slice
        constexpr int NZS = 4, NZK = 8;  // the fused noise conv: stride 4, 8 taps, pad 2 (checked by the host)
        constexpr int U = 2 * PH;        // upsample rate == 16 B pieces per staged run
        constexpr int RUNS = 32 / U;     // runs one store instruction covers (lanes >= RUNS*U idle)
        constexpr int ITER = (32 + RUNS - 1) / RUNS;
        constexpr int SL4 = Cfg::STG_LANE / 4;
        const int cg = n_tile >> 1, half = n_tile & 1;
        const int sl = (warp - 2) >> 3;
        const int sub = lane / U, pp = lane - sub * U;  // this lane's run / piece in the store loop
        float* stg = reinterpret_cast<float*>(smem_gen + (stg_smem - smem_base) + (size_t)(warp - 2) * Cfg::STG_WARP);
        const float* cst = reinterpret_cast<const float*>(smem_gen + (const_smem - smem_base));
        const int co0 = cg * 16 + sl * 8;
        {
          const int mb = wg & 1;
          const int tb = t0 + mb * 128 + q * 32;
          uint64_t a2[NC / 2];  // packed fp32 pairs (FADD2 / FFMA2)
          {
            const uint32_t t_col = t_acc + mb * N_TILE + sl * NC;
            uint32_t r[32];
            tc_ld32(t_col, r);
#pragma unroll
            for (int i = 0; i < 16; ++i) a2[i] = f2_pack_u(r[2 * i], r[2 * i + 1]);
            if constexpr (NC - 32 == 8) {
              uint32_t r8[8];
              tc_ld8(t_col + 32, r8);
#pragma unroll
              for (int i = 0; i < 4; ++i) a2[16 + i] = f2_pack_u(r8[2 * i], r8[2 * i + 1]);
            } else {
              uint32_t r16[16];
              tc_ld16(t_col + 32, r16);
#pragma unroll
              for (int i = 0; i < 8; ++i) a2[16 + i] = f2_pack_u(r16[2 * i], r16[2 * i + 1]);
            }
            static_assert(NC == 40 || NC == 48, "phase-major tiles: 5 or 6 phases per half");
          }
          {  // per-channel bias (upsample + noise conv), the same for every phase
            const ulonglong2 bq0 = *reinterpret_cast<const ulonglong2*>(cst + co0), bq1 = *reinterpret_cast<const ulonglong2*>(cst + co0 + 4);
#pragma unroll
            for (int rr = 0; rr < PH; ++rr) {
              a2[rr * 4 + 0] = f2_add(a2[rr * 4 + 0], bq0.x); a2[rr * 4 + 1] = f2_add(a2[rr * 4 + 1], bq0.y);
              a2[rr * 4 + 2] = f2_add(a2[rr * 4 + 2], bq1.x); a2[rr * 4 + 3] = f2_add(a2[rr * 4 + 3], bq1.y);
            }
          }
          if (e.nz_har) {
#pragma unroll
            for (int j = 0; j < NZK; ++j) {
              const float* wj = cst + 256 + j * e.cout + co0;
              const ulonglong2 wq0 = *reinterpret_cast<const ulonglong2*>(wj), wq1 = *reinterpret_cast<const ulonglong2*>(wj + 4);
#pragma unroll
              for (int rr = 0; rr < PH; ++rr) {
                const uint64_t hh = f2_splat(hv[NZS * rr + j]);
                a2[rr * 4 + 0] = f2_fma(hh, wq0.x, a2[rr * 4 + 0]); a2[rr * 4 + 1] = f2_fma(hh, wq0.y, a2[rr * 4 + 1]);
                a2[rr * 4 + 2] = f2_fma(hh, wq1.x, a2[rr * 4 + 2]); a2[rr * 4 + 3] = f2_fma(hh, wq1.y, a2[rr * 4 + 3]);
              }
            }
          }
          // registers -> staging (lane stride STG_LANE = 11 / 13 x 16 B: conflict-free quarter-warp stores)
          {
            ulonglong2* mine = reinterpret_cast<ulonglong2*>(stg + lane * SL4);
#pragma unroll
            for (int i = 0; i < NC / 4; ++i) mine[i] = make_ulonglong2(a2[2 * i], a2[2 * i + 1]);
          }
          __syncwarp();
          // staging -> global: lane (sub, pp) takes 16 B piece pp of run it*RUNS + sub, so one store instruction writes
          // RUNS whole runs (PH*32 B each, 2*PH*32 B apart: the other half of the phases belongs to the neighbouring
          // N tile) and every address is the previous one plus a compile-time constant
          const size_t row0 = ((size_t)b * (e.cout >> 3) + cg * 2 + sl) * e.Lp_out + PADL + (size_t)half * PH;
          const size_t el0 = (row0 + (size_t)(tb + sub) * U) * 8 + pp * 4;
          const int lim = (sub < RUNS) ? min(32, p.L_in - tb) - sub : 0;  // run it*RUNS + sub exists iff it*RUNS < lim
          const float* src = stg + sub * SL4 + pp * 4;
          float4 f[ITER];
#pragma unroll
          for (int it = 0; it < ITER; ++it) f[it] = *reinterpret_cast<const float4*>(src + it * RUNS * SL4);
          float* of = e.out_f32 ? e.out_f32 + el0 : nullptr;
          uint16_t* oa = e.out_act ? reinterpret_cast<uint16_t*>(e.out_act) + el0 : nullptr;
          const bool f16 = e.act_fmt == ACT_F16;
#pragma unroll
          for (int it = 0; it < ITER; ++it) {
            const bool ok = it * RUNS < lim;
            const float4 ff = f[it];
            if (of && ok) *reinterpret_cast<float4*>(of + it * RUNS * U * 8) = ff;
            uint2 pk;
            const float m0 = fmaxf(ff.x, ff.x * e.slope), m1 = fmaxf(ff.y, ff.y * e.slope);
            const float m2 = fmaxf(ff.z, ff.z * e.slope), m3 = fmaxf(ff.w, ff.w * e.slope);
            if (f16) {
              __half2 h0 = __floats2half2_rn(m0, m1), h1 = __floats2half2_rn(m2, m3);
              pk.x = *reinterpret_cast<uint32_t*>(&h0); pk.y = *reinterpret_cast<uint32_t*>(&h1);
            } else {
              __nv_bfloat162 h0 = __floats2bfloat162_rn(m0, m1), h1 = __floats2bfloat162_rn(m2, m3);
              pk.x = *reinterpret_cast<uint32_t*>(&h0); pk.y = *reinterpret_cast<uint32_t*>(&h1);
            }
            if (oa && ok) *reinterpret_cast<uint2*>(oa + it * RUNS * U * 8) = pk;
          }
          __syncwarp();
        }
      } else {
#pragma unroll 1
      for (int item = wg; item < ITEMS; item += WGS) {
        const int mb = item / CH;
        const int c0 = (item - mb * CH) * CW;
        const int t_in = t0 + mb * 128 + q * 32 + lane;
        const bool valid = t_in < p.L_in;
        if (fast) {
          uint32_t r[CW];
          // ---- lean path (every ordinary conv): up == 1, output row == input row ----
          const int n0 = n_tile * N_TILE + c0;
          const size_t off = (((size_t)b * (e.cout >> 3) + (n0 >> 3)) * e.Lp_out + PADL + t_in) * 8;
          float4 rr[CW / 4], sm[CW / 4];
          if (e.res && valid) {  // issue the residual loads before waiting on TMEM
#pragma unroll
            for (int s8 = 0; s8 < CW / 8; ++s8) {
              rr[2 * s8] = *reinterpret_cast<const float4*>(e.res + off + s8 * slice_stride);
              rr[2 * s8 + 1] = *reinterpret_cast<const float4*>(e.res + off + s8 * slice_stride + 4);
            }
          }
          if (e.sum_in && valid) {
#pragma unroll
            for (int s8 = 0; s8 < CW / 8; ++s8) {
              sm[2 * s8] = *reinterpret_cast<const float4*>(e.sum_in + off + s8 * slice_stride);
              sm[2 * s8 + 1] = *reinterpret_cast<const float4*>(e.sum_in + off + s8 * slice_stride + 4);
            }
          }
          [[maybe_unused]] float hv[LEAN ? 1 : 8];  // har samples under this row's (<= 8-tap) noise conv, loaded before the TMEM wait
          if constexpr (!LEAN) {
            if (e.nz_har) {
              const float* hb = e.nz_har + (size_t)b * e.nz_src;
              const int s0 = t_in * e.nz_stride - e.nz_pad;
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const int sidx = s0 + j;
                hv[j] = (valid && j < e.nz_k && sidx >= 0 && sidx < e.nz_src) ? __ldg(hb + sidx) : 0.f;
              }
            }
          }
          // bias of the 32 columns, requested before the TMEM wait as well (the volatile TMEM load pins everything
          // after it: profiles/r1bl, the first FADD of every 8-column group sat on its own bias load)
          float4 bb[CW / 4];
          {
            const float4* bp = reinterpret_cast<const float4*>(e.bias + (size_t)b * e.bias_bstride + n0);
#pragma unroll
            for (int i = 0; i < CW / 4; ++i) bb[i] = __ldg(bp + i);
          }
          if constexpr (CW == 32) tc_ld32(t_acc + mb * N_TILE + c0, r); else tc_ld16(t_acc + mb * N_TILE + c0, r);
          if (valid) {
#pragma unroll
            for (int s8 = 0; s8 < CW / 8; ++s8) {
              float v[8];
              const float4 b0 = bb[2 * s8], b1 = bb[2 * s8 + 1];
              v[0] = __uint_as_float(r[s8 * 8 + 0]) + b0.x; v[1] = __uint_as_float(r[s8 * 8 + 1]) + b0.y;
              v[2] = __uint_as_float(r[s8 * 8 + 2]) + b0.z; v[3] = __uint_as_float(r[s8 * 8 + 3]) + b0.w;
              v[4] = __uint_as_float(r[s8 * 8 + 4]) + b1.x; v[5] = __uint_as_float(r[s8 * 8 + 5]) + b1.y;
              v[6] = __uint_as_float(r[s8 * 8 + 6]) + b1.z; v[7] = __uint_as_float(r[s8 * 8 + 7]) + b1.w;
              if (e.res) {
                v[0] += rr[2 * s8].x; v[1] += rr[2 * s8].y; v[2] += rr[2 * s8].z; v[3] += rr[2 * s8].w;
                v[4] += rr[2 * s8 + 1].x; v[5] += rr[2 * s8 + 1].y; v[6] += rr[2 * s8 + 1].z; v[7] += rr[2 * s8 + 1].w;
              }
              if constexpr (!LEAN) { if (e.nz_har) {  // fused noise_convs[i](har_source), hifigan_nsf.py:199 (har values hoisted above)
                const int co = n0 + s8 * 8;
                const float4 nb0 = __ldg(reinterpret_cast<const float4*>(e.nz_bias + co));
                const float4 nb1 = __ldg(reinterpret_cast<const float4*>(e.nz_bias + co + 4));
                v[0] += nb0.x; v[1] += nb0.y; v[2] += nb0.z; v[3] += nb0.w;
                v[4] += nb1.x; v[5] += nb1.y; v[6] += nb1.z; v[7] += nb1.w;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  if (j < e.nz_k) {
                    const float4 w0 = __ldg(reinterpret_cast<const float4*>(e.nz_w + (size_t)j * e.cout + co));
                    const float4 w1 = __ldg(reinterpret_cast<const float4*>(e.nz_w + (size_t)j * e.cout + co + 4));
                    v[0] = fmaf(hv[j], w0.x, v[0]); v[1] = fmaf(hv[j], w0.y, v[1]); v[2] = fmaf(hv[j], w0.z, v[2]); v[3] = fmaf(hv[j], w0.w, v[3]);
                    v[4] = fmaf(hv[j], w1.x, v[4]); v[5] = fmaf(hv[j], w1.y, v[5]); v[6] = fmaf(hv[j], w1.z, v[6]); v[7] = fmaf(hv[j], w1.w, v[7]);
                  }
                }
              } }
              if (e.scale != 1.f) {
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] *= e.scale;
              }
              if (e.sum_in) {
                v[0] += sm[2 * s8].x; v[1] += sm[2 * s8].y; v[2] += sm[2 * s8].z; v[3] += sm[2 * s8].w;
                v[4] += sm[2 * s8 + 1].x; v[5] += sm[2 * s8 + 1].y; v[6] += sm[2 * s8 + 1].z; v[7] += sm[2 * s8 + 1].w;
              }
              const size_t o = off + s8 * slice_stride;
              if (e.out_f32) {
                *reinterpret_cast<float4*>(e.out_f32 + o) = make_float4(v[0], v[1], v[2], v[3]);
                *reinterpret_cast<float4*>(e.out_f32 + o + 4) = make_float4(v[4], v[5], v[6], v[7]);
              }
              if (e.out_act) {
                // leaky_relu(v) = max(v, slope*v) for 0 < slope < 1
                uint32_t pk[4];
                if (e.act_fmt == ACT_F16) {
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    __half2 hh = __floats2half2_rn(fmaxf(v[2 * i], v[2 * i] * e.slope), fmaxf(v[2 * i + 1], v[2 * i + 1] * e.slope));
                    pk[i] = *reinterpret_cast<uint32_t*>(&hh);
                  }
                } else {
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    __nv_bfloat162 hh = __floats2bfloat162_rn(fmaxf(v[2 * i], v[2 * i] * e.slope), fmaxf(v[2 * i + 1], v[2 * i + 1] * e.slope));
                    pk[i] = *reinterpret_cast<uint32_t*>(&hh);
                  }
                }
                if (!LEAN && e.act_zs > 1) {  // zero-stuffed rows for the next stage's stride-act_zs transposed conv
                  uint16_t* oa = reinterpret_cast<uint16_t*>(e.out_act) +
                                 ((((size_t)b * (e.cout >> 3) + (n0 >> 3) + s8) * e.Lp_act + PADL) + (size_t)t_in * e.act_zs) * 8;
                  *reinterpret_cast<uint4*>(oa) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                  for (int z = 1; z < e.act_zs; ++z) *reinterpret_cast<uint4*>(oa + z * 8) = make_uint4(0, 0, 0, 0);
                } else {
                  *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(e.out_act) + o) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                }
              }
            }
          }
        } else if constexpr (!LEAN) {
          // ---- generic path (polyphase upsample: phase scatter) ----
          // 8 columns per TMEM load so the accumulator registers are indexed statically (a 32-wide
          // array indexed by the s8 loop variable lived in local memory: STL/LDL in the r1c capture).
#pragma unroll 1
          for (int s8 = 0; s8 < CW / 8; ++s8) {
            uint32_t r8[8];
            tc_ld8(t_acc + mb * N_TILE + c0 + s8 * 8, r8);
            if (valid) {
              float v[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r8[i]);
              epilogue_store<8>(e, b, t_in, n_tile * N_TILE + c0 + s8 * 8, v);
            }
          }
        }
      }
      }
      tc_fence_before();
      mbar_arrive(acc_empty + 8 * as);
      ++acc_it;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"((uint32_t)Cfg::TMEM_COLS)
                 : "memory");
  }
}

template <int N_TILE, int MB, int KC, int B_STAGES, int A_STAGES = 2, int PM = 0, int EW = 8, int LEAN = 0>
static cudaError_t launch_inst(const ConvParams& p, int num_sms, cudaStream_t s) {
  using Cfg = MmaCfg<N_TILE, MB, KC, B_STAGES, A_STAGES, PM, EW>;
  auto kern = conv_mma_kernel<N_TILE, MB, KC, B_STAGES, A_STAGES, PM, EW, LEAN>;
  if (LEAN && (p.epi.up != 1 || p.epi.nz_har || p.epi.act_zs > 1)) return cudaErrorInvalidValue;
  if (PM != p.epi.pm || p.halo > Cfg::A_HALO) return cudaErrorInvalidValue;
  if (PM == 2 && (p.epi.up != 2 || p.epi.act_zs > 1 || p.epi.res || p.epi.sum_in || p.epi.out_act || !p.epi.out_f32 || p.epi.bias_bstride ||
                  (p.epi.nz_har && p.epi.nz_k > 8)))
    return cudaErrorInvalidValue;
  {
    static unsigned long long done = 0;  // one per instantiation
    cudaError_t e = set_smem_once(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES, done);
    if (e != cudaSuccess) return e;
  }
  if (p.Cin % KC != 0 || p.N_total % N_TILE != 0 || p.halo > PADL || p.ntaps > MAX_TAPS) return cudaErrorInvalidValue;
  ConvParams q = p;
  q.m_tiles = (p.L_in + Cfg::MT - 1) / Cfg::MT;
  q.n_tiles = p.N_total / N_TILE;
  if (q.n_tiles > MAX_NTILES) return cudaErrorInvalidValue;
  const long long total = (long long)q.B * q.m_tiles * q.n_tiles;
  if (total <= 0 || total > 0x7fffffffLL) return cudaErrorInvalidValue;
  const int grid = (int)(total < num_sms ? total : num_sms);
  kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, s>>>(q);
  return cudaGetLastError();
}

bool conv_mma_supported(int n_tile, int kc) {
  return (n_tile == 256 && kc == 64) || (n_tile == 128 && kc == 64) || (n_tile == 64 && kc == 64) ||
         (n_tile == 32 && kc == 32) || (n_tile == 80 && kc == 64) || (n_tile == 96 && kc == 64);
}

cudaError_t launch_conv_mma(const ConvParams& p, int n_tile, int kc, int num_sms, cudaStream_t s) {
  // phase-major upsample tiles (rate 10 / 12): 5 / 6 phases x 16 channels, accumulators double-buffered
  // (three activation stages for the rate-10 tile: an upsample has only 2 taps of MMAs per 64-channel chunk, so with
  // two stages the MMA warp spent more than half of its time waiting for activation tiles, profiles/r1bg)
  if (n_tile == 80 && kc == 64) return p.a_stages == 2 ? launch_inst<80, 2, 64, 5, 2, 1>(p, num_sms, s) : launch_inst<80, 2, 64, 3, 3, 1>(p, num_sms, s);
  if (n_tile == 96 && kc == 64) return launch_inst<96, 2, 64, 4, 2, 1>(p, num_sms, s);
  // stride-2 polyphase upsamples (columns (slice, phase, channel), warp-transposed coalesced epilogue)
  if (p.epi.pm == 2) {
    if (n_tile == 128 && kc == 64) return launch_inst<128, 2, 64, 6, 2, 2, 16>(p, num_sms, s);
    if (n_tile == 64 && kc == 64) return launch_inst<64, 4, 64, 4, 2, 2, 16>(p, num_sms, s);
    return cudaErrorInvalidValue;
  }
  if (n_tile == 256 && kc == 64) return launch_inst<256, 2, 64, 4>(p, num_sms, s);
  // 128-wide tiles: option a_stages = 3 adds a third 40 KB activation stage.  Measured (profiles/r1af_*): the
  // 128-channel stage gets 7 % SLOWER with it (229 vs 214 ms), so two stages stay the default.
  if (n_tile == 128 && kc == 64) {
    if (p.a_stages == 3) return launch_inst<128, 2, 64, 6, 3>(p, num_sms, s);
    if (p.epi_warps == 16 && p.lean && p.epi.up == 1 && !p.epi.nz_har && p.epi.act_zs <= 1) return launch_inst<128, 2, 64, 6, 2, 0, 16, 1>(p, num_sms, s);
    return p.epi_warps == 16 ? launch_inst<128, 2, 64, 6, 2, 0, 16>(p, num_sms, s) : launch_inst<128, 2, 64, 6, 2>(p, num_sms, s);
  }
  if (n_tile == 64 && kc == 64) return p.epi_warps_small == 16 ? launch_inst<64, 4, 64, 4, 2, 0, 16>(p, num_sms, s) : launch_inst<64, 4, 64, 4>(p, num_sms, s);
  if (n_tile == 32 && kc == 32) return p.epi_warps_small == 16 ? launch_inst<32, 4, 32, 4, 2, 0, 16>(p, num_sms, s) : launch_inst<32, 4, 32, 4>(p, num_sms, s);
  return cudaErrorInvalidValue;
}

}  // namespace rvcdec
