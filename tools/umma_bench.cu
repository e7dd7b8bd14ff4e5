// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, M=128, K=16, cta_group::1) as a function of N
// with both operands in shared memory (no-swizzle K-major, the layout conv_mma.cu / mrf_fused.cu use).
// Answers: is the small-N (32/64-channel) implicit GEMM bound by the tensor pipe or by the
// shared-memory operand fetch?   Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_bench umma_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | ((uint64_t)1 << 46);
}
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

constexpr int A_ROWS = 128 + 64;
constexpr int A_SLICE = A_ROWS * 16;

__global__ void __launch_bounds__(128, 1) bench(int N, int iters, int shift_mode, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const uint32_t a_s = smem_u32(smem), b_s = a_s + 8 * A_SLICE;
  for (int i = threadIdx.x; i < (8 * A_SLICE + 8 * 256 * 16) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tm = tmem_slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const long long t0 = clock64();
    int n = 0;
    for (int it = 0; it < iters; ++it) {
      for (int tap = 0; tap < 8; ++tap) {
        const uint32_t sh = shift_mode ? (uint32_t)(tap * 3) * 16 : 0;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t ad = make_desc(a_s + sh + 2 * ks * A_SLICE, A_SLICE, 128);
          const uint64_t bd = make_desc(b_s + 2 * ks * N * 16, N * 16, 128);
          mma(tm + (uint32_t)((it & 1) * N), ad, bd, idesc, (tap | ks) ? 1u : 0u);
          ++n;
        }
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
    const long long t1 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = n; }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  const int smem = 8 * A_SLICE + 8 * 256 * 16;
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  printf("N,shifted_A,grid,cycles_per_mma,floor_MN_over_256,mac_per_clk_per_sm\n");
  for (int grid : {1, 148})
    for (int shift : {0, 1})
      for (int N : {16, 32, 48, 64, 96, 128, 192, 256}) {
        bench<<<grid, 128, smem>>>(N, 64, shift, d);
        bench<<<grid, 128, smem>>>(N, 256, shift, d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        long long h[2];
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        const double cpm = (double)h[0] / (double)h[1];
        printf("%d,%d,%d,%.2f,%.1f,%.0f\n", N, shift, grid, cpm, 128.0 * N / 256.0, 128.0 * N * 16 / cpm);
      }
  return 0;
}
