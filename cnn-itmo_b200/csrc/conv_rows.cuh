// conv_rows.cuh — 3x3 ConvBN layers with 64 output channels (conv2d_3 32 -> 64, conv2d_4 64 -> 64 + pool, conv2d_14 96 -> 64
// + head; model.py:214-216, 274-276) with the three kernel ROWS folded into the N dimension of the MMA.
//
// conv_igemm.cuh issues one M128 x N64 MMA per (tap, 16 input channels): 32 clocks of tensor-core math against 6 KB of
// shared-memory operand reads (4 KB of pixels + 2 KB of weights) = 48 clocks of the 128 B/clock shared-memory pipe —
// those layers sit on that pipe (ncu smem_tc_wavefront 73 - 90 %, tensor pipe 49 - 72 %). The pixel operand of an MMA is
// the same for the three taps of a kernel COLUMN applied to one input row: input row r contributes W[ky][kx] x in(r, x + kx - 1)
// to output row r + 1 - ky. So here
//   * a tile is 4 image rows x 128 pixels; an M block is ONE image row of 128 pixels: TMEM lane = pixel column, whatever
//     the row, and the accumulators of the tile's four output rows lie side by side in TMEM (row y at columns (3 - y) * 64);
//   * per (input row, kx, 16 channels) ONE MMA of N = 192 multiplies the row's 128 pixels by [W[0][kx] | W[1][kx] | W[2][kx]]
//     and accumulates into three ADJACENT output rows' accumulators (r + 1, r, r - 1: ascending columns) — 96 clocks of
//     math for 10 KB of operand reads (80 clocks): the tensor pipe is the bound again. The two halo rows and the tile's
//     first / last rows use the N = 64 / 128 sub-ranges of the same weight tile; per (kx, 16 channels) the six input rows
//     cost 48 + 64 + 96 + 96 + 64 + 48 = 416 clocks for 4 x 96 = 384 of useful math, against 4 x 3 x 48 = 576 before.
//   * an output row's accumulator is first written by the first MMA of the input row of the same index (its ky = 0 group):
//     that one MMA is split into an overwriting N = 64 part and an accumulating rest.
// The weight tile [ky][cout][32 channels] of a (channel block, kx) is a 4-D TMA view of the ordinary packed weights
// ([tap][cout][cin]) — no second copy; all tiles stay resident. Halo tile: 6 rows x 130 pixels x 64 bytes per channel block
// (SWIZZLE_64B), zero-filled outside the image by the TMA. The epilogue is conv_igemm.cuh's epilogue_fast with the
// kEpiRows pixel mapping. Results are bit-identical to conv_igemm_kernel's for the same layer only up to the order of the
// fp32 accumulation inside the tensor core (same products, same accumulator per output element, different MMA order).
#pragma once
#include "conv_igemm.cuh"

namespace itmo {

constexpr int kRowsNM = 4;                 // output rows per tile = M blocks = accumulators per TMEM buffer
constexpr int kRowsTileW = 128;
constexpr int kRowsHaloW = kRowsTileW + 2, kRowsHaloH = kRowsNM + 2;
constexpr int kRowsSwz = 64, kRowsBK = 32, kRowsCout = 64;
constexpr int kRowsABytes = kRowsSwz * kRowsHaloW * kRowsHaloH;            // 49,920: one halo tile of one channel block
constexpr int kRowsAStage = (kRowsABytes + 1023) / 1024 * 1024;
constexpr int kRowsBTile = 3 * kRowsCout * kRowsSwz;                       // 12,288: [ky][cout][32 ch] of one (channel block, kx)

// The MMAs of one channel block of a tile: six halo rows x three kernel columns x two K steps.
template <uint32_t FMT>
__device__ __forceinline__ void issue_kb_rows(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                              bool first_kb) {
#pragma unroll
  for (int hr = 0; hr < kRowsHaloH; ++hr) {            // halo row hr = input row hr - 1 of the tile
    constexpr int kNmM1 = kRowsNM - 1;
    const int ky_lo = hr > kNmM1 ? hr - kNmM1 : 0, ky_hi = hr < 2 ? hr : 2;    // output row = hr - ky must be 0 .. 3
    const int ng = ky_hi - ky_lo + 1;
    const uint32_t d = tmem_d + uint32_t((kNmM1 - (hr - ky_lo)) * kRowsCout);  // columns of output row hr - ky_lo
    const uint32_t idesc = sm100::make_idesc(FMT, kTileM, uint32_t(ng * kRowsCout));
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const uint64_t da = (uint64_t(a_hi) << 32) | uint64_t(a_lo + (uint32_t((hr * kRowsHaloW + dx) * kRowsSwz + k * 32) >> 4));
        const uint32_t b0 = b_lo + (uint32_t(dx * kRowsBTile + ky_lo * kRowsCout * kRowsSwz + k * 32) >> 4);
        const uint64_t db = (uint64_t(b_hi) << 32) | uint64_t(b0);
        if (dx == 0 && k == 0 && hr <= kNmM1) {
          // the first MMA that touches output row hr (its ky = 0 group, the lowest columns of this MMA)
          if (first_kb) {
            sm100::umma_ss<false>(d, da, db, sm100::make_idesc(FMT, kTileM, kRowsCout), 0u);
            if (ng > 1) {
              const uint64_t db1 = (uint64_t(b_hi) << 32) | uint64_t(b0 + (uint32_t(kRowsCout * kRowsSwz) >> 4));
              sm100::umma_ss<false>(d + kRowsCout, da, db1, sm100::make_idesc(FMT, kTileM, uint32_t((ng - 1) * kRowsCout)), 1u);
            }
            continue;
          }
        }
        sm100::umma_ss<false>(d, da, db, idesc, 1u);
      }
    }
  }
}

// Epilogue of a row-tile ConvBN layer followed by MaxPooling2D (conv2d_4, model.py:215-216): the two output rows of a pooling
// pair are two M blocks on the SAME TMEM lanes, so the vertical maximum is taken in registers (the first row's packed
// values are kept while the second is computed) and the horizontal one with a lane-pair shuffle; even lanes stage the
// pooled pixel. Warp set ws owns output rows 3 - 2 ws and 2 - 2 ws of the tile (M blocks 2 ws, 2 ws + 1); a warp holds 32
// consecutive pixels of a row. Staging per warp: 32 full-resolution pixel rows + 16 pooled ones of 128 bytes, SWIZZLE_128B.
constexpr int kRowsPoolStageB = 48 * 128;
template <typename Elem>
__device__ __forceinline__ void epilogue_rows_pool(const ConvParams& p, uint32_t tmem_base, uint64_t* tmem_full_bar,
                                                   uint64_t* tmem_empty_bar, uint8_t* epi_stage, int warp_idx, int lane) {
  constexpr bool kF16 = ElemTraits<Elem>::kFmt == sm100::kFmtF16;
  const int ew = warp_idx - 4, q = warp_idx & 3, ws = ew >> 2;
  const int tiles_x = p.W / kRowsTileW, tiles_y = p.H / kRowsNM;
  const int total_tiles = p.NB * tiles_y * tiles_x;
  uint8_t* const st_out = epi_stage + ew * kRowsPoolStageB;
  uint8_t* const st_pool = st_out + 32 * 128;
  const int sw_out = lane & 7;
  const bool count_sat = kF16 && p.sat_count != nullptr;
  int acc = 0;
  uint32_t acc_phase = 0;
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    int m = tile;
    const int tx = m % tiles_x;
    m /= tiles_x;
    const int ty = m % tiles_y;
    const int img = m / tiles_y;
    sm100::mbar_wait_relaxed(&tmem_full_bar[acc], acc_phase, 200);
    sm100::tc_fence_after_sync();
    const uint32_t taddr = tmem_base + (uint32_t(32 * q) << 16) + uint32_t(acc) * uint32_t(kRowsNM * kRowsCout);
    const int x0 = tx * kRowsTileW + 32 * q;
    uint32_t keep[32];
#pragma unroll 1
    for (int jj = 0; jj < 2; ++jj) {
      const int j = 2 * ws + jj;                      // M block; output row 3 - j of the tile
      // one 32-channel half at a time and a ROLLED row loop: with both halves' accumulators, this row's packed values and the
      // kept row live at once (and both rows' code interleaved) the epilogue spilled 496 bytes under the 384-thread register
      // bound and ran slower than conv_igemm_kernel's
      uint32_t pk[32];
      uint32_t m2 = 0;
      if (lane == 0) sm100::tma_store_wait_read<0>();   // the bulk stores that last read the staging tiles are done with them
      __syncwarp();
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t v[32];
        sm100::tmem_ld_32x32b_x32(taddr + uint32_t(j * kRowsCout + h * 32), v);
        sm100::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float r0, r1, f0, f1;
          relu_bias2(v[i], v[i + 1], cpair(p.c_bias, h * 32 + i), r0, r1);
          f2_unpack(f2_fma(f2_pack(r0, r1), cpair(p.c_scale, h * 32 + i), cpair(p.c_shift, h * 32 + i)), f0, f1);
          pk[h * 16 + (i >> 1)] = pack2<Elem>(f0, f1);
        }
      }
      if constexpr (kF16) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const __half2 a = __habs2(*reinterpret_cast<const __half2*>(&pk[i]));
          const __half2 b = __habs2(*reinterpret_cast<const __half2*>(&pk[i + 1]));
          const __half2 mm = __hmax2(__hmax2(*reinterpret_cast<const __half2*>(&m2), a), b);
          m2 = *reinterpret_cast<const uint32_t*>(&mm);
        }
        m2 &= 0x7FFF7FFFu;
        if (count_sat && ((m2 & 0xFFFFu) >= 0x7BFFu || (m2 >> 16) >= 0x7BFFu)) atomicAdd(p.sat_count, 1ull);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k)
        *reinterpret_cast<uint4*>(st_out + lane * 128 + ((k ^ sw_out) << 4)) = make_uint4(pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
      if (jj == 0) {
#pragma unroll
        for (int i = 0; i < 32; ++i) keep[i] = pk[i];
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          uint32_t mv = max2<Elem>(keep[i], pk[i]);                                   // the two rows of the pair
          mv = max2<Elem>(mv, __shfl_xor_sync(0xFFFFFFFFu, mv, 1));                    // the two columns
          pk[i] = mv;
        }
        if ((lane & 1) == 0) {
          const int pr = lane >> 1;
#pragma unroll
          for (int k = 0; k < 8; ++k)
            *reinterpret_cast<uint4*>(st_pool + pr * 128 + ((k ^ (pr & 7)) << 4)) = make_uint4(pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
        }
      }
      sm100::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        const int y = ty * kRowsNM + (kRowsNM - 1 - j);
        sm100::tma_store_4d(&p.tmOut[0], st_out, 0, x0, y, img);
        if (jj == 1) sm100::tma_store_4d(&p.tmPool, st_pool, 0, x0 >> 1, y >> 1, img);
        sm100::tma_store_commit();
      }
    }
    sm100::tc_fence_before_sync();
    __syncwarp();
    if (lane == 0) sm100::mbar_arrive(&tmem_empty_bar[acc]);
    acc ^= 1;
    if (acc == 0) acc_phase ^= 1;
  }
  if (lane == 0) sm100::tma_store_wait_all<0>();
}

template <typename Elem, int EPI>
__global__ void __launch_bounds__(kNumThreads, 1) conv_rows_kernel(const __grid_constant__ ConvParams p) {
  static_assert((EPI & kEpiRows) != 0 && (EPI & kEpiConstAffine) != 0 && (EPI & kEpiPair) == 0, "row-tile epilogue variants");
  constexpr int BLOCK_N = kRowsCout;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = sm100::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

  int kb_total = 0;
  for (int s = 0; s < p.n_seg; ++s) kb_total += p.seg_kb[s];
  const int n_a = p.n_a_stages;
  uint8_t* smem_a = smem;                                       // n_a halo tiles
  uint8_t* smem_b = smem_a + n_a * kRowsAStage;                 // kb_total x 3 weight tiles, resident
  uint8_t* epi_stage = smem_b + kb_total * 3 * kRowsBTile;      // TMA-store staging tiles (1 KB aligned: all sizes are)
  uint8_t* aux = epi_stage + p.epi_bytes;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(aux);          // [kMaxAStages]
  uint64_t* a_empty = a_full + kMaxAStages;
  uint64_t* b_full = a_empty + kMaxAStages;                     // [1]
  uint64_t* tmem_full_bar = b_full + 1;                         // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;                 // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  float* s_head = reinterpret_cast<float*>(aux + 512);          // [3*64 folded weights][3 folded biases]

  const int warp_idx = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int tiles_x = p.W / kRowsTileW, tiles_y = p.H / kRowsNM;
  const int total_tiles = p.NB * tiles_y * tiles_x;
  constexpr uint32_t kBufCols = kRowsNM * BLOCK_N;              // 256 columns per accumulator buffer, two buffers

  // ---------------------------------------------------------------- one-time setup
  if (warp_idx == 0 && lane == 0) {
    for (int s = 0; s < 2; ++s) sm100::tma_prefetch_desc(&p.tmA[s]);
    sm100::tma_prefetch_desc(&p.tmB);
    if (p.tma_store) {
      sm100::tma_prefetch_desc(&p.tmOut[0]);
      sm100::tma_prefetch_desc(&p.tmPool);
    }
  }
  if (warp_idx == 1 && lane == 0) {
    for (int i = 0; i < kMaxAStages; ++i) {
      sm100::mbar_init(&a_full[i], 1);
      sm100::mbar_init(&a_empty[i], 1);
    }
    sm100::mbar_init(b_full, 1);
    for (int i = 0; i < 2; ++i) {
      sm100::mbar_init(&tmem_full_bar[i], 1);
      sm100::mbar_init(&tmem_empty_bar[i], kNumEpiWarps);
    }
    sm100::fence_barrier_init();
  }
  if (warp_idx == 2) sm100::tmem_alloc(tmem_ptr_smem, 2 * kBufCols);
  if constexpr ((EPI & 15) == kEpiHead) {
    // conv2d_14's BN affine folded into the head's weights, exactly as conv_igemm_kernel's setup does
    for (int i = threadIdx.x; i < 3 * 64 + 3; i += blockDim.x) {
      if (i < 3 * 64) {
        s_head[i] = p.head_w[i] * p.c_scale[i & 63];
      } else {
        const int k = i - 3 * 64;
        float acc = p.head_b[k];
        for (int ch = 0; ch < 64; ++ch) acc = fmaf(p.head_w[k * 64 + ch], p.c_shift[ch], acc);
        s_head[i] = acc;
      }
    }
  }
  sm100::tc_fence_before_sync();
  __syncthreads();
  sm100::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // ---------------------------------------------------------------- warp roles
  if (warp_idx == 0) {
    // ===== TMA producer =====
    if (sm100::elect_one()) {
      // every weight tile once: (channel block, kx) -> [ky][cout][32 channels]
      sm100::mbar_arrive_expect_tx(b_full, uint32_t(kb_total * 3 * kRowsBTile));
      for (int kb = 0; kb < kb_total; ++kb)
        for (int dx = 0; dx < 3; ++dx)
          sm100::tma_load_4d(smem_b + (kb * 3 + dx) * kRowsBTile, &p.tmB, b_full, kb * kRowsBK, 0, dx, 0, sm100::kEvictLast);
      int sa = 0, a_loads = 0;
      uint32_t pa = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int m = tile;
        const int tx = m % tiles_x;
        m /= tiles_x;
        const int ty = m % tiles_y;
        const int img = m / tiles_y;
        const int x0 = tx * kRowsTileW - 1, y0 = ty * kRowsNM - 1;
        for (int s = 0; s < p.n_seg; ++s) {
          const CUtensorMap* tmA = &p.tmA[p.seg_map[s]];
          for (int kb = 0; kb < p.seg_kb[s]; ++kb) {
            sm100::mbar_wait(&a_empty[sa], pa ^ 1);
            if ((p.dbg & 16) && a_loads >= n_a) {   // timing experiment (results wrong): no activation traffic after the ring filled
              sm100::mbar_arrive(&a_full[sa]);
            } else {
              sm100::mbar_arrive_expect_tx(&a_full[sa], uint32_t(kRowsABytes));
              sm100::tma_load_4d(smem_a + sa * kRowsAStage, tmA, &a_full[sa], kb * kRowsBK, x0, y0, img, sm100::kEvictNormal);
            }
            ++a_loads;
            if (++sa == n_a) {
              sa = 0;
              pa ^= 1;
            }
          }
        }
      }
    }
  } else if (warp_idx == 1) {
    // ===== MMA issuer =====
    if (sm100::elect_one()) {
      const uint32_t a_hi = sm100::smem_desc_hi<kRowsSwz>(8u * kRowsSwz);   // 128 consecutive pixels of one halo row
      const uint32_t b_hi = sm100::smem_desc_hi<kRowsSwz>(8u * kRowsSwz);
      const uint32_t b_lo0 = sm100::smem_desc_lo(sm100::smem_u32(smem_b));
      int sa = 0;
      uint32_t pa = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      sm100::mbar_wait(b_full, 0);
      sm100::tc_fence_after_sync();
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        sm100::mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
        sm100::tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + uint32_t(acc) * kBufCols;
        for (int kb = 0; kb < kb_total; ++kb) {
          sm100::mbar_wait(&a_full[sa], pa);
          sm100::tc_fence_after_sync();
          const uint32_t a_lo = sm100::smem_desc_lo(sm100::smem_u32(smem_a + sa * kRowsAStage));
          if (!(p.dbg & 64))   // 64: timing experiment without the MMAs
            issue_kb_rows<ElemTraits<Elem>::kFmt>(tmem_d, a_lo, a_hi, b_lo0 + uint32_t(kb) * (uint32_t(3 * kRowsBTile) >> 4), b_hi,
                                                  kb == 0);
          sm100::umma_commit(&a_empty[sa]);
          if (++sa == n_a) {
            sa = 0;
            pa ^= 1;
          }
        }
        sm100::umma_commit(&tmem_full_bar[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp_idx >= 4) {
    // ===== epilogue warps (conv_igemm.cuh), row-tile pixel mapping =====
    if constexpr ((EPI & 15) == kEpiConvPool)
      epilogue_rows_pool<Elem>(p, tmem_base, tmem_full_bar, tmem_empty_bar, epi_stage, warp_idx, lane);
    else
      epilogue_fast<Elem, BLOCK_N, EPI>(p, tmem_base, tmem_full_bar, tmem_empty_bar, epi_stage, nullptr, nullptr, nullptr, s_head,
                                        warp_idx, lane);
  }

  // ---------------------------------------------------------------- teardown
  sm100::tc_fence_before_sync();
  __syncthreads();
  if (warp_idx == 2) {
    sm100::tc_fence_after_sync();
    sm100::tmem_dealloc(tmem_base, 2 * kBufCols);
  }
}

}  // namespace itmo
