// wgrad_igemm.cuh — weight-gradient implicit GEMM of the training step (main.py:126-132, model.py:281) on sm_100a.
//
// For a 3x3 'same' convolution (ConvBN, model.py:195-196)
//     dW[ky][kx][ci][co] = sum over (n, y, x) of  X[n, y+ky-1, x+kx-1, ci] * G[n, y, x, co]
// and for the 2x2 stride-2 transposed convolution (ConvBNTranspose, model.py:199-200), per kernel row a,
//     dW[a][b][co][ci]   = sum over (n, i, j) of  G[n, 2i+a, 2j+b, co] * X[n, i, j, ci]
// Both are GEMMs whose reduction (K) dimension is the PIXEL index: D[m][n] = sum_p A[p][m] * B[p][n]. In NHWC
// memory the channels of a pixel are contiguous, so both operands are "MN-major" for the tensor core: the kernel
// feeds tcgen05.mma straight from the TMA-written pixel tiles with MN-major shared-memory descriptors (SWIZZLE_128B:
// one 128-byte row = 64 bf16 channels of one pixel, 8 pixels = one swizzle atom), no transposition pass.
//
//   A (M side, 128 rows) : the tensor that is read SHIFTED per tap (X for the 3x3 conv; for the transposed conv the
//                          gradient viewed as [n, H, W, (b,co)] has a single tap). One TMA box per 64 channels brings a
//                          16 x (8+2*pad) pixel tile; the taps of one kernel row are start-address shifts of whole
//                          pixels inside that tile (zero fill outside the image = 'same' padding).
//   B (N side, BN columns): the unshifted tensor, 16 x 8 pixel tile.
//   K block              : 128 pixels (16 rows x 8 columns) = 8 MMAs of K = 16 (two pixel rows each).
//
// A CTA owns one output tile (128 A-channels x BN B-channels x the TG taps of one kernel row: TG accumulators of
// BN fp32 columns in TMEM) and one slice of the pixel range (split-K over CTAs so that tiles x slices ~ SM count).
// Deterministic split-K: with more than one slice every CTA writes its partial tile to a workspace
// ws[tile][slice][tap][row][col] and wgrad_reduce_kernel adds the slices IN SLICE ORDER, then adds the total to the
// fp32 gradient tensor once; with one slice the CTA owns its part of the gradient and adds to it directly. No atomics:
// two runs give bit-identical gradients, and the gradient buffer receives one addition per launch (micro-batch), so
// it accumulates across micro-batches until the optimizer step zeroes it.
#pragma once
#include <cuda_bf16.h>
#include "sm100_ptx.cuh"

namespace itmo {

constexpr int kWgThreads = 192;   // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..5 epilogue
constexpr int kWgMaxStages = 4;
constexpr int kWgBoxPixB = 16 * 8;           // pixels of a B box
constexpr int kWgRowBytes = 128;             // 64 bf16 channels

struct WgradParams {
  CUtensorMap tmA;   // dims (Ca, W, H, N), box (64, 8 + 2*pad, 16, 1), SWIZZLE_128B
  CUtensorMap tmB;   // dims (Cb, W, H, N), box (64, 8, 16, 1), SWIZZLE_128B
  int H, W, NB;      // spatial size / batch of the (common) pixel grid
  int pad;           // 1: 3x3 conv (taps shift A by -1..1), 0: single tap
  int tg;            // taps per tile = accumulators: 3 (one kernel row), 9 (the whole 3x3 kernel, narrow layers) or 1
  int n_groups;      // tiles along the kernel rows: 3 when tg == 3, else 1
  int a_rows;        // pixel rows of an A box: 16, or 18 when the CTA covers all three kernel rows (tg == 9)
  int m_blocks;      // ceil(Ca / 128)
  int n_blocks;      // Cb / BN
  int a_boxes;       // 64-channel boxes of A per stage (1 or 2)
  int b_boxes;       // 64-channel boxes of B per stage (ceil(BN / 64))
  int a_box_bytes;   // 16 * (8 + 2*pad) * 128
  int stage_bytes;
  int n_stages;
  int ksplit;        // pixel-range slices
  int ca;            // valid A channels (rows written)
  int m64;           // 1: UMMA M = 64 (ca <= 64): half the A operand bytes per MMA at the same MMA duration
  float* ws;         // split-K workspace [tiles][ksplit][tg][128][BN] fp32 (ksplit > 1), else null
  float* dw;         // [n_groups*tg][ld_m][ld_n] fp32, accumulated
  int ld_m, ld_n;    // rows / columns of one tap's matrix
  int m_off, n_off;  // where this launch's (A-channel, B-channel) block starts inside it
};

__device__ __forceinline__ uint32_t wg_desc_lo(uint32_t smem_addr, uint32_t lbo_bytes) {
  return ((smem_addr & 0x3FFFFu) >> 4) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
__device__ __forceinline__ uint32_t wg_desc_hi(uint32_t sbo_bytes) {
  return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);   // version 1, SWIZZLE_128B
}

template <int BN>
__global__ void __launch_bounds__(kWgThreads, 1) wgrad_igemm_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = sm100::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint8_t* aux = smem + p.n_stages * p.stage_bytes + (p.a_boxes == 1 ? p.a_box_bytes : 0);   // slack: see run_wgrad
  uint64_t* full = reinterpret_cast<uint64_t*>(aux);
  uint64_t* empty = full + kWgMaxStages;
  uint64_t* tmem_full = empty + kWgMaxStages;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp_idx = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  // tile / slice of this CTA
  const int split = blockIdx.x % p.ksplit;
  int tile = blockIdx.x / p.ksplit;
  const int nb = tile % p.n_blocks;
  tile /= p.n_blocks;
  const int mb = tile % p.m_blocks;
  const int g = tile / p.m_blocks;

  const int tiles_x = p.W / 8, tiles_y = p.H / 16;
  const long long pb_total = (long long)p.NB * tiles_y * tiles_x;
  const int pb_beg = int(pb_total * split / p.ksplit);
  const int pb_end = int(pb_total * (split + 1) / p.ksplit);
  const int halo_w = 8 + 2 * p.pad;
  uint32_t tmem_cols = 32;
  while (tmem_cols < uint32_t(p.tg * BN)) tmem_cols <<= 1;

  if (warp_idx == 0 && lane == 0) {
    sm100::tma_prefetch_desc(&p.tmA);
    sm100::tma_prefetch_desc(&p.tmB);
    for (int i = 0; i < kWgMaxStages; ++i) {
      sm100::mbar_init(&full[i], 1);
      sm100::mbar_init(&empty[i], 1);
    }
    sm100::mbar_init(tmem_full, 1);
    sm100::fence_barrier_init();
  }
  if (warp_idx == 1) sm100::tmem_alloc(tmem_ptr_smem, tmem_cols);
  sm100::tc_fence_before_sync();
  __syncthreads();
  sm100::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp_idx == 0) {
    if (sm100::elect_one()) {
      // bytes the TMA delivers per stage (a_box_bytes is the 1 KB-aligned SLOT of a box, its data can be smaller)
      const uint32_t bytes = uint32_t(p.a_boxes * p.a_rows * halo_w * kWgRowBytes + p.b_boxes * kWgBoxPixB * kWgRowBytes);
      int s = 0;
      uint32_t ph = 0;
      for (int pb = pb_beg; pb < pb_end; ++pb) {
        const int tx = pb % tiles_x;
        const int t2 = pb / tiles_x;
        const int ty = t2 % tiles_y;
        const int img = t2 / tiles_y;
        sm100::mbar_wait(&empty[s], ph ^ 1);
        uint8_t* st = smem + s * p.stage_bytes;
        sm100::mbar_arrive_expect_tx(&full[s], bytes);
        for (int j = 0; j < p.a_boxes; ++j)
          sm100::tma_load_4d(st + j * p.a_box_bytes, &p.tmA, &full[s], mb * 128 + j * 64, tx * 8 - p.pad,
                             ty * 16 + (p.tg == 9 ? 0 : g) - p.pad, img, sm100::kEvictNormal);
        uint8_t* sb = st + p.a_boxes * p.a_box_bytes;
        for (int j = 0; j < p.b_boxes; ++j)
          sm100::tma_load_4d(sb + j * kWgBoxPixB * kWgRowBytes, &p.tmB, &full[s], nb * BN + j * 64, tx * 8, ty * 16, img,
                             sm100::kEvictNormal);
        if (++s == p.n_stages) {
          s = 0;
          ph ^= 1;
        }
      }
    }
  } else if (warp_idx == 1) {
    if (sm100::elect_one()) {
      // both operands MN-major (bits 15, 16)
      // M = 64 for the narrow layers: an MMA lasts max(M,128)*N/256 clocks either way, but its A tile is 2 KB
      // instead of 4 KB of shared-memory reads (the 32/64-channel layers were bound by exactly those reads)
      const uint32_t idesc = sm100::make_idesc(sm100::kFmtBF16, p.m64 ? 64 : 128, BN) | (1u << 15) | (1u << 16);
      const uint32_t a_hi = wg_desc_hi(uint32_t(halo_w) * kWgRowBytes);   // next 8 pixels of K: one tile row down
      const uint32_t b_hi = wg_desc_hi(8u * kWgRowBytes);
      const uint32_t a_lbo = uint32_t(p.a_box_bytes);                      // next 64 channels: the next box
      const uint32_t b_lbo = uint32_t(kWgBoxPixB * kWgRowBytes);
      int s = 0;
      uint32_t ph = 0;
      for (int pb = pb_beg; pb < pb_end; ++pb) {
        sm100::mbar_wait(&full[s], ph);
        sm100::tc_fence_after_sync();
        const uint32_t a_addr = sm100::smem_u32(smem + s * p.stage_bytes);
        const uint32_t b_addr = a_addr + uint32_t(p.a_boxes * p.a_box_bytes);
        const uint32_t a_lo = wg_desc_lo(a_addr, a_lbo);
        const uint32_t b_lo = wg_desc_lo(b_addr, b_lbo);
        for (int t = 0; t < p.tg; ++t) {
          // tap -> whole-pixel shift inside the halo tile: (ky, kx) when the CTA holds all nine taps, else kx only
          const int tshift = p.tg == 9 ? (t / 3) * halo_w + (t % 3) : t;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t da =
                (uint64_t(a_hi) << 32) | uint64_t(a_lo + ((uint32_t((2 * ks * halo_w + tshift) * kWgRowBytes)) >> 4));
            const uint64_t db = (uint64_t(b_hi) << 32) | uint64_t(b_lo + ((uint32_t(2 * ks * 8 * kWgRowBytes)) >> 4));
            sm100::umma_ss<false>(tmem_base + uint32_t(t * BN), da, db, idesc, (pb != pb_beg || ks != 0) ? 1u : 0u);
          }
        }
        sm100::umma_commit(&empty[s]);
        if (++s == p.n_stages) {
          s = 0;
          ph ^= 1;
        }
      }
      sm100::umma_commit(tmem_full);
    }
  } else if (pb_end > pb_beg) {
    // epilogue: TMEM lane = A channel (row of dW), column = tap * BN + B channel
    const int q = warp_idx & 3;
    sm100::mbar_wait_relaxed(tmem_full, 0, 500);
    sm100::tc_fence_after_sync();
    // M = 128: accumulator row r lives in TMEM lane r. M = 64: row r lives in lane (r % 16) + 32 * (r / 16), i.e.
    // every warp's lane quarter holds 16 rows in its first 16 lanes
    const int row = p.m64 ? (lane < 16 ? 16 * q + lane : p.ca) : mb * 128 + 32 * q + lane;
    const uint32_t taddr = tmem_base + (uint32_t(32 * q) << 16);
    for (int t = 0; t < p.tg; ++t) {
      const int tap = g * p.tg + t;
#pragma unroll 1
      for (int chunk = 0; chunk < BN / 32; ++chunk) {
        uint32_t v[32];
        sm100::tmem_ld_32x32b_x32(taddr + uint32_t(t * BN + chunk * 32), v);
        sm100::tmem_ld_wait();
        if (row < p.ca) {
          if (p.ws != nullptr) {
            // partial tile of this slice; row index inside the 128-row tile
            const int rl = row - mb * 128;
            float4* dst = reinterpret_cast<float4*>(
                p.ws + ((size_t(blockIdx.x) * p.tg + t) * 128 + rl) * BN + chunk * 32);   // blockIdx.x = tile * ksplit + split
#pragma unroll
            for (int i = 0; i < 8; ++i)
              dst[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]),
                                   __uint_as_float(v[4 * i + 3]));
          } else {
            // single slice: this CTA is the only writer of its part of the gradient
            float4* dst = reinterpret_cast<float4*>(p.dw + (size_t(tap) * p.ld_m + p.m_off + row) * p.ld_n + p.n_off +
                                                    nb * BN + chunk * 32);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float4 o = dst[i];
              o.x += __uint_as_float(v[4 * i]);
              o.y += __uint_as_float(v[4 * i + 1]);
              o.z += __uint_as_float(v[4 * i + 2]);
              o.w += __uint_as_float(v[4 * i + 3]);
              dst[i] = o;
            }
          }
        }
      }
    }
  }
  sm100::tc_fence_before_sync();
  __syncthreads();
  if (warp_idx == 1) {
    sm100::tc_fence_after_sync();
    sm100::tmem_dealloc(tmem_base, tmem_cols);
  }
}

// Second half of the deterministic split-K: dw += sum over slices (in slice order) of the partial tiles.
// One thread = 4 consecutive columns of one (tile, tap, row).
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ ws, int n_tiles, int ksplit, int tg,
                                                           int bn, int m_blocks, int n_blocks, int ca, float* __restrict__ dw,
                                                           int ld_m, int ld_n, int m_off, int n_off) {
  const int c4 = bn >> 2;
  const size_t total = size_t(n_tiles) * tg * 128 * c4;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const int col = int(i % c4) * 4;
    size_t r = i / c4;
    const int rl = int(r % 128);
    r /= 128;
    const int t = int(r % tg);
    const int tile = int(r / tg);
    const int nb = tile % n_blocks;
    const int mb = (tile / n_blocks) % m_blocks;
    const int g = tile / (n_blocks * m_blocks);
    const int row = mb * 128 + rl;
    if (row >= ca) continue;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* src = ws + ((size_t(tile) * ksplit * tg + t) * 128 + rl) * bn + col;
    const size_t sp_stride = size_t(tg) * 128 * bn;
    int sp = 0;
    for (; sp + 4 <= ksplit; sp += 4) {   // four independent loads in flight, added in slice order
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const float4*>(src + (sp + u) * sp_stride);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        acc.x += v[u].x;
        acc.y += v[u].y;
        acc.z += v[u].z;
        acc.w += v[u].w;
      }
    }
    for (; sp < ksplit; ++sp) {
      const float4 v = *reinterpret_cast<const float4*>(src + sp * sp_stride);
      acc.x += v.x;
      acc.y += v.y;
      acc.z += v.z;
      acc.w += v.w;
    }
    float4* dst = reinterpret_cast<float4*>(dw + (size_t(g * tg + t) * ld_m + m_off + row) * ld_n + n_off + nb * bn + col);
    float4 o = *dst;
    o.x += acc.x;
    o.y += acc.y;
    o.z += acc.z;
    o.w += acc.w;
    *dst = o;
  }
}

}  // namespace itmo
