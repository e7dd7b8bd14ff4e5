// How fast can 16 warps of one CTA per SM push the fused kernels' epilogue store pattern?
//   pattern 0: blocked-8 fp32 rows as the kernels do it today: lane = row, two STG.128 per 32 B row-slice (32 B lane stride)
//   pattern 1: lane pairs exchange halves (shfl.xor 1): every STG.128 writes whole 32 B sectors (even rows, then odd rows)
//   pattern 2: lane l writes half (l & 1) of row (l >> 1): every STG.128 writes 512 contiguous bytes (needs indexed shuffles)
//   pattern 3: reference: plain contiguous float4 per lane (no layout constraint)
// Prints clock cycles per "block" (128 rows x 128 channels fp32 = 64 KB per SM) and the aggregate GB/s.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/stg_bench tools/stg_bench.cu && tools/stg_bench
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(512, 1) k(float* out, size_t slice_stride /*floats*/, int iters, int pattern, long long* cyc) {
  extern __shared__ char pad[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, cg = warp >> 2;  // lane quarter, column group of 4 slices
  const int r_in_blk = q * 32 + lane;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const size_t row0 = ((size_t)(blockIdx.x * iters + it)) * 128;
    float4 lo = make_float4(it, lane, warp, 1.f), hi = make_float4(it, lane, warp, 2.f);
#pragma unroll
    for (int s8 = 0; s8 < 4; ++s8) {
      float* sl = out + (size_t)(cg * 4 + s8) * slice_stride + row0 * 8;
      if (pattern == 0) {
        float* p = sl + (size_t)r_in_blk * 8;
        *reinterpret_cast<float4*>(p) = lo;
        *reinterpret_cast<float4*>(p + 4) = hi;
      } else if (pattern == 1) {
        const bool odd = lane & 1;
        float4 send = odd ? lo : hi, recv;
        recv.x = __shfl_xor_sync(0xffffffffu, send.x, 1); recv.y = __shfl_xor_sync(0xffffffffu, send.y, 1);
        recv.z = __shfl_xor_sync(0xffffffffu, send.z, 1); recv.w = __shfl_xor_sync(0xffffffffu, send.w, 1);
        float* pe = sl + (size_t)(r_in_blk & ~1) * 8 + (odd ? 4 : 0);  // even row of the pair
        *reinterpret_cast<float4*>(pe) = odd ? recv : lo;
        *reinterpret_cast<float4*>(pe + 8) = odd ? hi : recv;          // odd row of the pair
      } else if (pattern == 2) {
        const int src_a = lane >> 1, src_b = 16 + (lane >> 1);
        const bool h = lane & 1;
        float4 a, b;
        a.x = __shfl_sync(0xffffffffu, h ? hi.x : lo.x, src_a);  // NB: the selected half must be the SOURCE lane's: both are sent below
        float4 alo, ahi, blo, bhi;
        alo.x = __shfl_sync(0xffffffffu, lo.x, src_a); alo.y = __shfl_sync(0xffffffffu, lo.y, src_a); alo.z = __shfl_sync(0xffffffffu, lo.z, src_a); alo.w = __shfl_sync(0xffffffffu, lo.w, src_a);
        ahi.x = __shfl_sync(0xffffffffu, hi.x, src_a); ahi.y = __shfl_sync(0xffffffffu, hi.y, src_a); ahi.z = __shfl_sync(0xffffffffu, hi.z, src_a); ahi.w = __shfl_sync(0xffffffffu, hi.w, src_a);
        blo.x = __shfl_sync(0xffffffffu, lo.x, src_b); blo.y = __shfl_sync(0xffffffffu, lo.y, src_b); blo.z = __shfl_sync(0xffffffffu, lo.z, src_b); blo.w = __shfl_sync(0xffffffffu, lo.w, src_b);
        bhi.x = __shfl_sync(0xffffffffu, hi.x, src_b); bhi.y = __shfl_sync(0xffffffffu, hi.y, src_b); bhi.z = __shfl_sync(0xffffffffu, hi.z, src_b); bhi.w = __shfl_sync(0xffffffffu, hi.w, src_b);
        a = h ? ahi : alo; b = h ? bhi : blo;
        float* p = sl + (size_t)(q * 32) * 8 + (size_t)lane * 4;
        *reinterpret_cast<float4*>(p) = a;          // rows q*32 + 0..15
        *reinterpret_cast<float4*>(p + 128) = b;    // rows q*32 + 16..31
      } else {
        float* p = sl + (size_t)(q * 32) * 8 + (size_t)lane * 4;
        *reinterpret_cast<float4*>(p) = lo;
        *reinterpret_cast<float4*>(p + 128) = hi;
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  const int iters = 64, grid = 148;
  const size_t rows = (size_t)grid * iters * 128, slice_stride = rows * 8;
  float* out; long long* cyc;
  cudaMalloc(&out, slice_stride * 16 * sizeof(float));
  cudaMalloc(&cyc, grid * sizeof(long long));
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int pattern = 0; pattern < 4; ++pattern) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0);
      k<<<grid, 512, 200 * 1024>>>(out, slice_stride, iters, pattern, cyc);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      double mean = 0; for (int i = 0; i < grid; ++i) mean += h[i]; mean /= grid;
      if (rep) printf("pattern %d: %.0f clk per 64 KB block (issue side), %.1f GB/s aggregate, %s\n", pattern, mean / iters,
                      slice_stride * 16 * 4 / (ms * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
