// Micro-benchmark 3: tcgen05.mma with cta_group::2 (M = 256 over a CTA pair, K = 16, kind::f16, both operands in
// shared memory, no-swizzle K-major).  Two questions before any kernel is rebuilt around CTA pairs:
//   (1) correctness of the operand split as this repo would use it: every CTA supplies its own 128 rows of A and
//       HALF of B (N/2 rows), the accumulator of CTA r is rows [128 r, 128 r + 128) x all N columns;  the handshake
//       (remote mbarrier arrive from the peer, multicast commit back) is the one the kernels would need;
//   (2) the issue rate per MMA as a function of N: in cta_group::1 it is max(45.7, 32 + N/4) clk
//       (profiles/r1c_umma_issue_rate_vs_N.csv) - does halving the B bytes per SM move it for N = 64 / 32?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_bench3 umma_bench3.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | ((uint64_t)1 << 46);
}
__device__ __forceinline__ void mma2(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit2(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ uint32_t cta_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void remote_arrive(uint32_t bar, uint32_t rank) {
  asm volatile("{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\t"
               "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(bar), "r"(rank)
               : "memory");
}
__device__ __forceinline__ void wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (long long spin = 0; !done; ++spin) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done)
                 : "r"(bar), "r"(parity)
                 : "memory");
    if (spin > (1LL << 26)) { printf("umma_bench3: barrier timeout (block %d)\n", (int)blockIdx.x); __trap(); }
  }
}

constexpr int A_ROWS = 128 + 64;
constexpr int A_SLICE = A_ROWS * 16;   // bytes per 8-element K block of the A tile
constexpr int KB = 8;                  // K blocks held in shared memory (K = 64)
constexpr int B_BYTES = KB * 128 * 16; // room for N/2 <= 128 rows per K block

// mode 0: timing (operands zero); mode 1: correctness (operands from global, D written back)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
bench2(int N, int iters, int mode, const __half* A, const __half* B, float* D, long long* out, int commit_taps = 0) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar_done, bar_ready, bar_sink;
  __shared__ uint32_t tmem_slot;
  const uint32_t rank = cta_rank();
  const uint32_t a_s = smem_u32(smem), b_s = a_s + KB * A_SLICE;
  const int NH = N / 2;
  for (int i = threadIdx.x; i < (KB * A_SLICE + B_BYTES) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
  __syncthreads();
  if (mode == 1) {
    // A: [2][128][64] halves (CTA, row, k);  B: [N][64] (n, k): this CTA takes rows [rank*NH, rank*NH + NH)
    for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) {
      const int r = i / 64, k = i % 64;
      *(__half*)(smem + (k / 8) * A_SLICE + r * 16 + (k % 8) * 2) = A[((size_t)rank * 128 + r) * 64 + k];
    }
    for (int i = threadIdx.x; i < NH * 64; i += blockDim.x) {
      const int n = i / 64, k = i % 64;
      *(__half*)(smem + KB * A_SLICE + (k / 8) * (NH * 16) + n * 16 + (k % 8) * 2) = B[((size_t)rank * NH + n) * 64 + k];
    }
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar_done)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 2;" ::"r"(smem_u32(&bar_ready)));  // one arrive per CTA
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar_sink)));   // commits nobody waits for
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  cluster_sync();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tm = tmem_slot;
  // both CTAs tell the leader that their operands are in place (the kernels' "operand rows ready" handshake)
  if (threadIdx.x == 32) {
    asm volatile("fence.proxy.async.shared::cta;");
    remote_arrive(smem_u32(&bar_ready), 0);
  }
  if (rank == 0 && threadIdx.x == 0) {
    wait_cluster(smem_u32(&bar_ready), 0);
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
    const long long t0 = clock64();
    int n = 0;
    if (mode == 1) {
      for (int ks = 0; ks < 4; ++ks) {
        mma2(tm, make_desc(a_s + 2 * ks * A_SLICE, A_SLICE, 128), make_desc(b_s + 2 * ks * NH * 16, NH * 16, 128), idesc, ks ? 1u : 0u);
        ++n;
      }
    } else {
      for (int it = 0; it < iters; ++it)
        for (int tap = 0; tap < 8; ++tap) {
          const uint32_t sh = (uint32_t)(tap * 3) * 16;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            mma2(tm + (uint32_t)((it & 1) * N), make_desc(a_s + sh + 2 * ks * A_SLICE, A_SLICE, 128),
                 make_desc(b_s + 2 * ks * NH * 16, NH * 16, 128), idesc, (tap | ks) ? 1u : 0u);
            ++n;
          }
          // the kernels release a weight stage with one multicast commit per (tap, K chunk) = every 4 MMAs
          if (commit_taps && (tap % commit_taps) == commit_taps - 1) commit2(smem_u32(&bar_sink));
        }
    }
    commit2(smem_u32(&bar_done));
    wait_cluster(smem_u32(&bar_done), 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = n; }
  }
  if (mode == 1) {
    // every thread waits for the multicast commit in ITS OWN CTA, then reads its TMEM lane
    wait_cluster(smem_u32(&bar_done), 0);
    __syncwarp();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c0 = 0; c0 < N; c0 += 8) {
      uint32_t r[8];
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                   : "r"(tm + ((uint32_t)(w * 32) << 16) + (uint32_t)c0)
                   : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int i = 0; i < 8; ++i) D[((size_t)(blockIdx.x & 1) * 128 + w * 32 + lane) * N + c0 + i] = __uint_as_float(r[i]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  cluster_sync();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tm));
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  const int smem = KB * A_SLICE + B_BYTES;
  cudaFuncSetAttribute(bench2, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  // ---- (1) correctness: D_r[row][n] = sum_k A_r[row][k] * B[n][k] ----
  int bad_total = 0;
  for (int N : {32, 64, 128, 256}) {
    std::vector<__half> hA(2 * 128 * 64), hB((size_t)N * 64);
    std::vector<float> fA(hA.size()), fB(hB.size());
    srand(1234 + N);
    for (size_t i = 0; i < hA.size(); ++i) { fA[i] = (float)(rand() % 7 - 3); hA[i] = __float2half(fA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { fB[i] = (float)(rand() % 5 - 2); hB[i] = __float2half(fB[i]); }
    __half *dA, *dB;
    float* dD;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 2 * 128 * (size_t)N * 4);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0xff, 2 * 128 * (size_t)N * 4);
    bench2<<<2, 128, smem>>>(N, 1, 1, dA, dB, dD, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("correctness N=%d: error %s\n", N, cudaGetErrorString(e)); return 1; }
    std::vector<float> hD(2 * 128 * (size_t)N);
    cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int c = 0; c < 2; ++c)
      for (int r = 0; r < 128; ++r)
        for (int n = 0; n < N; ++n) {
          float ref = 0.f;
          for (int k = 0; k < 64; ++k) ref += fA[((size_t)c * 128 + r) * 64 + k] * fB[(size_t)n * 64 + k];
          const float got = hD[((size_t)c * 128 + r) * N + n];
          if (got != ref && bad++ < 4) printf("  mismatch N=%d cta=%d row=%d col=%d got=%g want=%g\n", N, c, r, n, got, ref);
        }
    printf("correctness,N=%d,mismatches=%d\n", N, bad);
    bad_total += bad;
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
  }
  // ---- (2) issue rate ----
  printf("N,grid,cycles_per_mma,floor_256N_over_512,mac_per_clk_per_sm\n");
  for (int grid : {2, 148})
    for (int N : {32, 48, 64, 96, 128, 192, 256}) {
      bench2<<<grid, 128, smem>>>(N, 64, 0, nullptr, nullptr, nullptr, d);
      bench2<<<grid, 128, smem>>>(N, 256, 0, nullptr, nullptr, nullptr, d);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      long long h[2];
      cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      const double cpm = (double)h[0] / (double)h[1];
      printf("%d,%d,%.2f,%.1f,%.0f\n", N, grid, cpm, 256.0 * N / 512.0, 128.0 * N * 16 / cpm);
    }
  // ---- (3) does a multicast commit every `commit_taps` x 4 MMAs cost issue rate? ----
  printf("N,grid,commit_every_mmas,cycles_per_mma\n");
  for (int N : {64, 128})
    for (int ct : {0, 1, 2, 4}) {
      bench2<<<148, 128, smem>>>(N, 64, 0, nullptr, nullptr, nullptr, d, ct);
      bench2<<<148, 128, smem>>>(N, 256, 0, nullptr, nullptr, nullptr, d, ct);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      long long h[2];
      cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      printf("%d,148,%d,%.2f\n", N, ct * 4, (double)h[0] / (double)h[1]);
    }
  return bad_total ? 2 : 0;
}
