// exchange.cuh — the gradient exchange of data-parallel training as ONE kernel over NVLink peer memory
// (reduce-scatter -> optimizer on the local 1/N shard -> all-gather of the updated parameters).
//
// Every rank keeps its flat gradient and parameter buffers (11,159,011 floats each, itmo_train_layout) in a cudaMalloc'd
// WINDOW that the other ranks of the box map through CUDA IPC (one process per GPU, NVSwitch gives every GPU full
// bandwidth to every peer). exchange_step_kernel then does, for the 1/N shard of the parameters rank r owns:
//     g[i]  = sum over ranks q = 0 .. N-1, IN RANK ORDER, of grads_q[i]          (peer loads  = the reduce-scatter)
//     p[i] <- optimizer(p[i], g[i], state_r[i])                                   (RMSprop / Adam, train_ops.cuh's math)
//     params_q[i] = p[i] for every rank q                                         (peer stores = the all-gather)
// between two barriers over flags in the peers' windows: A = "every rank's backward has finished", B = "every rank's
// parameter writes have landed". No NCCL call, no intermediate buffer, no separate optimizer launch; the summation order
// is fixed, so every element is computed exactly once, identically for all ranks, and an N-rank step equals one rank
// accumulating the N shards in rank order bit for bit. Only the owner keeps the optimizer state of a shard live.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "train_ops.cuh"

namespace itmo {

constexpr int kMaxRanks = 8;
constexpr int kFlagWords = 64;   // per rank: [0..7] barrier A arrivals by rank, [8..15] barrier B arrivals

struct ExchangeArgs {
  float* grads[kMaxRanks];      // every rank's gradient buffer (own = local pointer, others = IPC mappings)
  float* params[kMaxRanks];
  unsigned* flags[kMaxRanks];
  float* state0;                // local optimizer state (RMSprop accumulator / Adam m)
  float* state1;                // Adam v
  unsigned long long n;         // floats per buffer
  unsigned long long begin, end;  // this rank's shard [begin, end), multiples of 4 except the last end
  int rank, world;
  unsigned epoch;               // barrier values of this step: A = 2 * epoch + 1, B = 2 * epoch + 2
  int adam;
  float lr, b1, b2, eps, gscale;
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// peer gradient loads must not be served from this SM's L1 (it may still hold the previous step's lines)
__device__ __forceinline__ float4 ld_peer4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_peer1(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// Signal "rank reached barrier `slot` of this step" into every peer's window, then wait for all of them (first `world`
// threads of block 0; bounded spin: a missing peer traps the kernel instead of hanging the box).
__device__ __forceinline__ void cross_rank_barrier(const ExchangeArgs& a, int slot, unsigned value) {
  if (blockIdx.x == 0 && int(threadIdx.x) < a.world) {
    const int q = threadIdx.x;
    st_release_sys(a.flags[q] + slot * kMaxRanks + a.rank, value);
    const long long t0 = clock64();
    while (int(ld_acquire_sys(a.flags[a.rank] + slot * kMaxRanks + q) - value) < 0) {
      if (clock64() - t0 > 8000000000ll) {
        printf("itmo: exchange barrier %d timed out on rank %d waiting for rank %d\n", slot, a.rank, q);
        __trap();
      }
    }
  }
}

__device__ __forceinline__ float opt_update(const ExchangeArgs& a, float p, float g, float& s0, float& s1) {
  return opt_element(a.adam != 0, a.lr, a.b1, a.b2, a.eps, a.gscale, p, g, s0, s1);   // train_ops.cuh; a.lr = lr_t for Adam
}

// Cooperative launch (all CTAs resident): grid.sync() orders the phases inside one rank, the flag barriers across ranks.
__global__ void __launch_bounds__(256) exchange_step_kernel(const ExchangeArgs a) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  // ---- barrier A: all ranks' gradient buffers are final (each rank's backward precedes this kernel in stream order)
  cross_rank_barrier(a, 0, 2 * a.epoch + 1);
  grid.sync();
  // ---- reduce-scatter + optimizer + all-gather on this rank's shard, 4 floats per thread and iteration
  const unsigned long long n4 = (a.end - a.begin) / 4;
  for (unsigned long long v = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; v < n4;
       v += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long i = a.begin + 4 * v;
    float4 g = ld_peer4(a.grads[0] + i);
    for (int q = 1; q < a.world; ++q) {              // rank order: the sum is the same on every rank and every run
      const float4 h = ld_peer4(a.grads[q] + i);
      g.x += h.x;
      g.y += h.y;
      g.z += h.z;
      g.w += h.w;
    }
    float4 p = *reinterpret_cast<const float4*>(a.params[a.rank] + i);
    float4 s0 = *reinterpret_cast<const float4*>(a.state0 + i);
    // Adam's second state half starts n floats into the state buffer and n is odd: scalar accesses
    float4 s1 = a.adam ? make_float4(a.state1[i], a.state1[i + 1], a.state1[i + 2], a.state1[i + 3])
                       : make_float4(0.f, 0.f, 0.f, 0.f);
    p.x = opt_update(a, p.x, g.x, s0.x, s1.x);
    p.y = opt_update(a, p.y, g.y, s0.y, s1.y);
    p.z = opt_update(a, p.z, g.z, s0.z, s1.z);
    p.w = opt_update(a, p.w, g.w, s0.w, s1.w);
    *reinterpret_cast<float4*>(a.state0 + i) = s0;
    if (a.adam) {
      a.state1[i] = s1.x;
      a.state1[i + 1] = s1.y;
      a.state1[i + 2] = s1.z;
      a.state1[i + 3] = s1.w;
    }
    for (int q = 0; q < a.world; ++q) *reinterpret_cast<float4*>(a.params[q] + i) = p;
  }
  // tail of the last shard (n is not a multiple of 4)
  const unsigned long long tail0 = a.begin + 4 * n4;
  for (unsigned long long i = tail0 + blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < a.end;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    float g = ld_peer1(a.grads[0] + i);
    for (int q = 1; q < a.world; ++q) g += ld_peer1(a.grads[q] + i);
    float s0 = a.state0[i], s1 = a.adam ? a.state1[i] : 0.f;
    const float p = opt_update(a, a.params[a.rank][i], g, s0, s1);
    a.state0[i] = s0;
    if (a.adam) a.state1[i] = s1;
    for (int q = 0; q < a.world; ++q) a.params[q][i] = p;
  }
  // ---- barrier B: this rank's peer stores are visible system-wide before it says so; it leaves only when every rank
  // has said so (so its own parameter buffer is complete and nobody still reads its gradients)
  __threadfence_system();
  grid.sync();
  cross_rank_barrier(a, 1, 2 * a.epoch + 2);
}

}  // namespace itmo
