// umma_rate.cu — microbenchmark: how fast can ONE thread issue tcgen05.mma (M=128, N, K=16 bf16, SS operands)
// and how fast does the tensor pipe retire them, as a function of N and of how many accumulators the
// stream rotates over. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o umma_rate umma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../cnn-itmo_b200/csrc/sm100_ptx.cuh"

template <int N, int ROT, int SWZ>
__global__ void __launch_bounds__(128, 1) rate_kernel(int iters, long long* out, int sbo_bytes, int a_off, int jstride, int rnd) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (sm100::smem_u32(raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);
  if (threadIdx.x == 0) { sm100::mbar_init(&bar, 1); sm100::fence_barrier_init(); }
  if (warp == 1) sm100::tmem_alloc(&tmem_slot, 512);
  // zero the operand area (values irrelevant for timing, but avoid NaN slow paths)
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(raw)[i] = rnd ? ((i * 2654435761u) & 0x3FFF3FFFu) | 0x3C003C00u * ((i >> 3) & 1) : 0u;  // bf16 pairs of mixed sign/magnitude
  sm100::fence_proxy_async_smem();
  sm100::tc_fence_before_sync();
  __syncthreads();
  sm100::tc_fence_after_sync();
  const uint32_t tmem = tmem_slot;
  if (warp == 0) {
    if (sm100::elect_one()) {
      constexpr uint32_t idesc = sm100::make_idesc(sm100::kFmtBF16, 128, N);
      const uint32_t a_addr = base, b_addr = base + 32 * 1024;
      const uint64_t da0 = sbo_bytes ? sm100::make_smem_desc_kmajor_sbo<SWZ>(a_addr + a_off, sbo_bytes) : sm100::make_smem_desc_kmajor<SWZ>(a_addr);
      const uint64_t db0 = sm100::make_smem_desc_kmajor<SWZ>(b_addr);
      unsigned long long g0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
      long long t0 = clock64();
      for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          sm100::umma_ss<false>(tmem + uint32_t((u % ROT) * N), da0 + 2 * (u % (SWZ / 32)) + uint32_t(((u % ROT) * jstride) >> 4), db0 + 2 * (u % (SWZ / 32)), idesc, 1u);
        }
      }
      long long t1 = clock64();
      sm100::umma_commit(&bar);
      sm100::mbar_wait(&bar, 0);
      long long t2 = clock64();
      unsigned long long g1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
      if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; out[2] = (long long)(g1 - g0); }
    }
  }
  sm100::tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) { sm100::tc_fence_after_sync(); sm100::tmem_dealloc(tmem, 512); }
}


// The conv kernel's MMA stream for one halo tile: 9 taps x KSTEPS x NM MMAs over distinct A / B addresses.
// ORDER 0: tap, k, j (k outer)   1: tap, j, k (M block outer)   2: k, tap... not used
template <int N, int NM, int SWZ, int ORDER>
__global__ void __launch_bounds__(128, 1) conv_pattern_kernel(int iters, long long* out, int b_same, int extras) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (sm100::smem_u32(raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar, bar2, bar3;
  __shared__ uint32_t tmem_slot;
  const int warp = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);
  if (threadIdx.x == 0) { sm100::mbar_init(&bar, 1); sm100::mbar_init(&bar2, 1); sm100::mbar_init(&bar3, 1); sm100::fence_barrier_init(); }
  if (warp == 1) sm100::tmem_alloc(&tmem_slot, 512);
  for (int i = threadIdx.x; i < (200 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(raw)[i] = 0;
  sm100::fence_proxy_async_smem();
  sm100::tc_fence_before_sync();
  __syncthreads();
  sm100::tc_fence_after_sync();
  const uint32_t tmem = tmem_slot;
  constexpr int KS = SWZ / 32;
  constexpr int halo_w = 8 * NM + 2;
  if (warp == 0) {
    if (sm100::elect_one()) {
      constexpr uint32_t idesc = sm100::make_idesc(sm100::kFmtBF16, 128, N);
      const uint32_t a_hi = sm100::smem_desc_hi<SWZ>(halo_w * SWZ), b_hi = sm100::smem_desc_hi<SWZ>(8 * SWZ);
      const uint32_t a_lo0 = sm100::smem_desc_lo(base), b_lo0 = sm100::smem_desc_lo(base + 48 * 1024);
      unsigned long long g0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
      long long t0 = clock64();
      for (int it = 0; it < iters; ++it) {
        if (extras & 1) sm100::tc_fence_after_sync();
        if (extras & 2) sm100::mbar_wait(&bar3, 1);   // phase 1 of a fresh barrier: already complete
        uint32_t b_lo = b_lo0;
        uint32_t a_row = a_lo0;
        for (int ty = 0; ty < 3; ++ty, a_row += (halo_w * SWZ) >> 4) {
          uint32_t a_tap = a_row;
          for (int tx = 0; tx < 3; ++tx, a_tap += SWZ >> 4) {
            if (ORDER == 0) {
#pragma unroll
              for (int k = 0; k < KS; ++k)
#pragma unroll
                for (int j = 0; j < NM; ++j)
                  sm100::umma_ss<false>(tmem + j * N, (uint64_t(a_hi) << 32) | (a_tap + ((j * 8 * SWZ + k * 32) >> 4)),
                                        (uint64_t(b_hi) << 32) | (b_lo + 2 * k), idesc, 1u);
            } else {
#pragma unroll
              for (int j = 0; j < NM; ++j)
#pragma unroll
                for (int k = 0; k < KS; ++k)
                  sm100::umma_ss<false>(tmem + j * N, (uint64_t(a_hi) << 32) | (a_tap + ((j * 8 * SWZ + k * 32) >> 4)),
                                        (uint64_t(b_hi) << 32) | (b_lo + 2 * k), idesc, 1u);
            }
            if (!b_same) b_lo += (N * SWZ) >> 4;
          }
        }
        if (extras & 4) sm100::umma_commit(&bar2);
        if (extras & 8) { sm100::umma_commit(&bar2); sm100::umma_commit(&bar2); }
      }
      long long t1 = clock64();
      sm100::umma_commit(&bar);
      sm100::mbar_wait(&bar, 0);
      long long t2 = clock64();
      unsigned long long g1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
      if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; out[2] = (long long)(g1 - g0); }
    }
  }
  sm100::tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) { sm100::tc_fence_after_sync(); sm100::tmem_dealloc(tmem, 512); }
}

template <int N, int NM, int SWZ, int ORDER>
void run_pattern(int b_same, int extras = 0) {
  long long* d; cudaMalloc(&d, 32);
  const int iters = 3000;
  auto k = conv_pattern_kernel<N, NM, SWZ, ORDER>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024);
  k<<<148, 128, 210 * 1024>>>(10, d, b_same, extras);
  k<<<148, 128, 210 * 1024>>>(iters, d, b_same, extras);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[3]; cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
  const double n = 9.0 * (SWZ / 32) * NM * iters;
  printf("pattern N=%3d nm=%d swz=%3d order=%d extras=%d: issue %.1f retire %.1f clk/MMA, %.0f MHz %s\n", N, NM, SWZ, ORDER, extras,
         h[0] / n, h[1] / n, h[1] * 1e3 / double(h[2]), e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(d);
}

template <int N, int ROT, int SWZ>
void run(int grid, int iters = 2000, int sbo = 0, int a_off = 0, int jstride = 0, int rnd = 0) {
  long long* d; cudaMalloc(&d, 32);
  auto k = rate_kernel<N, ROT, SWZ>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  k<<<grid, 128, 80 * 1024>>>(10, d, sbo, a_off, jstride, rnd);
  k<<<grid, 128, 80 * 1024>>>(iters, d, sbo, a_off, jstride, rnd);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[3]; cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
  const double n = 16.0 * iters;
  printf("%s N=%3d rot=%d swz=%3d grid=%3d sbo=%d off=%d js=%d: issue %.1f clk/MMA, retire %.1f clk/MMA (floor %d) SM clock %.0f MHz over %.1f ms %s\n", rnd ? "rand" : "zero", N, ROT, SWZ, grid, sbo, a_off, jstride, h[0] / n, h[1] / n,
         N / 2, h[1] * 1e3 / double(h[2]), h[2] / 1e6, e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  for (int ex : {0, 1, 2, 4, 8, 15}) run_pattern<64, 4, 64, 0>(0, ex);
  return 0;
}
