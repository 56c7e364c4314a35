// stem_fused.cuh — conv2d_1 (3 -> 32, model.py:208) and conv2d_2 (32 -> 32 + MaxPooling2D, model.py:209-210) as ONE kernel
// for the 16-bit inference modes: the first layer's output never goes to HBM.
//
// Separately the two layers cost 1.9 + 2.9 ms per 256 images: conv2d_1 is bound by the instruction issue of its im2col,
// conv2d_2 by the shared-memory operand reads of its N = 32 MMAs (40 clocks of reads for 16 clocks of math), and between
// them 16.8 MB per image are written and read back. The two limits are different resources, so one CTA does both:
//
//   builders (4 warps)     raw uint8 halo rows -> im2col rows of conv2d_1 (rgb_conv_tc.cuh's byte -> fp16 PRMT trick) for
//                          the 18 x 34 pixels conv2d_2's 16 x 32 tile needs: 5 M blocks of 128 rows, ring of 5 tiles
//   MMA thread             stage 1: 4 tcgen05.mma (M128 N32 K16: x against w_hi, then against w_lo) per M block into TMEM;
//                          stage 2: conv2d_2's 72 MMAs per tile (conv_igemm.cuh's issue_tap) — its A operand is the halo
//                          tile stage 1's epilogue wrote. Per iteration: stage 1 of tile k, then stage 2 of tile k - 1, so
//                          stage 1's epilogue runs under stage 2's MMAs and the builders are a whole tile ahead
//   stage-1 epilogue (4 w) TMEM -> y = max(acc / 255 + b, 0) s + t -> fp16 -> the 64-byte pixel row of conv2d_2's halo
//                          tile, in exactly the SWIZZLE_64B layout a TMA load of conv2d_1's output would have produced;
//                          pixels outside the image are written as ZERO (conv2d_2's 'same' padding pads conv2d_1's
//                          OUTPUT, not its input)
//   stage-2 epilogue (8 w) conv_igemm.cuh's epilogue_fast<.., 32, conv + pool, constant-bank affine>, unchanged: staging
//                          tile, 2x2 max-pool, TMA stores of the skip tensor and the pooled tensor
//
// The halo pixels (18 x 34 for a 16 x 32 tile) are recomputed by neighbouring tiles: 20 % more first-layer MMAs, which
// are 1/5 of the kernel's MMA time. Numerics are those of the two separate kernels (same operand split, same affine
// sequence, same packing), so the result is bit-identical to running them one after the other.
#pragma once
#include "conv_igemm.cuh"
#include "rgb_conv_tc.cuh"

namespace itmo {

constexpr int kStemThreads = 640;       // 20 warps: 0 loader, 1 MMA, 2 TMEM, 3 idle, 4-11 stage-2 epilogue, 12-15 stage-1 epilogue, 16-19 builders
constexpr int kStemHaloW = 34, kStemHaloH = 18, kStemHaloPix = kStemHaloW * kStemHaloH;   // 612
constexpr int kStemBlocks = 5;          // ceil(612 / 128)
constexpr int kStemA1Bufs = 5;          // a whole tile's im2col blocks: the builders run a full tile ahead of the MMAs
constexpr int kStemRawRows = 20, kStemRawWords = 28;
constexpr int kStemA2Bytes = 40 * 1024, kStemA1Bytes = 8 * 1024, kStemB1Bytes = 4 * 1024, kStemB2Bytes = 9 * 32 * 64;
constexpr int kStemEpiBytes = kNumEpiWarps * 40 * 64;
constexpr int kStemRawBytes = kStemRawRows * kStemRawWords * 4;
constexpr int kStemSmemBytes = 1024 + 2 * kStemA2Bytes + kStemA1Bufs * kStemA1Bytes + kStemB1Bytes + kStemB2Bytes + 2048 /*pad*/ +
                               kStemEpiBytes + 2 * kStemRawBytes + 1024 /*barriers*/;

struct StemParams {
  ConvParams c2;            // conv2d_2 exactly as run_conv would launch it (tmB, tmOut[0], tmPool, c_bias / c_scale / c_shift, nm = 4 ...)
  const uint8_t* in;        // [NB, H, W, 3] uint8
  const float* w1_hwio;     // conv2d_1 kernel [27][32] fp32
  RgbAffine a1;             // conv2d_1 bias / BN scale / shift by value
  float in_scale;           // 1 / 255
};

template <typename Elem>
__global__ void __launch_bounds__(kStemThreads, 1) stem_fused_kernel(const __grid_constant__ StemParams sp) {
  const ConvParams& p = sp.c2;
  constexpr int BLOCK_N = 32, SWZ = 64, NM = 4, KSTEPS = 2;
  constexpr int EPI2 = kEpiConvPool | kEpiConstAffine;
  extern __shared__ uint8_t smem_raw_[];
  const uint32_t raw_addr = sm100::smem_u32(smem_raw_);
  uint8_t* smem = smem_raw_ + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint8_t* s_a2 = smem;                                           // [2][40 KB] conv2d_2 halo tiles
  uint8_t* s_a1 = s_a2 + 2 * kStemA2Bytes;                        // [5][8 KB] conv2d_1 im2col tiles: 128 rows x 64 B (32 K slots), SWIZZLE_64B
  uint8_t* s_b1 = s_a1 + kStemA1Bufs * kStemA1Bytes;              // conv2d_1 weights: [w_hi | w_lo] tiles of [32 rows][64 B], SWIZZLE_64B
  uint8_t* s_b2 = s_b1 + kStemB1Bytes;                            // conv2d_2 weights [9][32][64 B]
  uint8_t* epi_stage = smem + ((s_b2 + kStemB2Bytes - smem + 1023) / 1024) * 1024;   // 1 KB aligned staging tiles
  uint32_t* s_rawin = reinterpret_cast<uint32_t*>(epi_stage + kStemEpiBytes);   // [2][20][28] raw input words
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(s_rawin) + 2 * kStemRawBytes);
  uint64_t* b2_full = bars;                    // [1]
  uint64_t* a1_full = bars + 1;                // [3]
  uint64_t* a1_empty = a1_full + kStemA1Bufs;  // [3]
  uint64_t* d1_full = a1_empty + kStemA1Bufs;  // [5]
  uint64_t* d1_empty = d1_full + kStemBlocks;  // [5]
  uint64_t* a2_full = d1_empty + kStemBlocks;  // [2]
  uint64_t* a2_empty = a2_full + 2;            // [2]
  uint64_t* tmem_full_bar = a2_empty + 2;      // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;  // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp_idx = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int tiles_x = p.W / (kBlkW * NM), tiles_y = p.H / kBlkH;
  const int total_tiles = p.NB * tiles_y * tiles_x;
  constexpr uint32_t kStage2Cols = 2 * NM * BLOCK_N;             // 256: two accumulator buffers of conv2d_2
  constexpr uint32_t kTmemCols = 512;

  // ---------------------------------------------------------------- one-time setup
  if (warp_idx == 0 && lane == 0) {
    sm100::tma_prefetch_desc(&p.tmB);
    sm100::tma_prefetch_desc(&p.tmOut[0]);
    sm100::tma_prefetch_desc(&p.tmPool);
  }
  if (warp_idx == 1 && lane == 0) {
    sm100::mbar_init(b2_full, 1);
    for (int i = 0; i < kStemA1Bufs; ++i) {
      sm100::mbar_init(&a1_full[i], 4);      // one arrival per builder warp
      sm100::mbar_init(&a1_empty[i], 1);
    }
    for (int i = 0; i < kStemBlocks; ++i) {
      sm100::mbar_init(&d1_full[i], 1);
      sm100::mbar_init(&d1_empty[i], 4);     // one arrival per stage-1 epilogue warp
    }
    for (int i = 0; i < 2; ++i) {
      sm100::mbar_init(&a2_full[i], 4);
      sm100::mbar_init(&a2_empty[i], 1);
      sm100::mbar_init(&tmem_full_bar[i], 1);
      sm100::mbar_init(&tmem_empty_bar[i], kNumEpiWarps);
    }
    sm100::fence_barrier_init();
  }
  if (warp_idx == 2) sm100::tmem_alloc(tmem_ptr_smem, kTmemCols);
  if (warp_idx == 3) {
    // conv2d_1 weights -> two K-major swizzled B tiles (w_hi, then w_lo = the 16-bit rounding of the remainder): row n,
    // 32 K slots each, slot -> tap by rgb_k_tap. The im2col row is stored ONCE (64 B) and multiplied by both tiles —
    // the K order hi(0-15), hi(16-31), lo(0-15), lo(16-31) of rgb_conv_tc.cuh's 128-byte rows, at half the row's
    // shared-memory store traffic (the kernel is bound by the shared-memory pipe: operand reads + LSU wavefronts).
    const int n = lane;
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
      uint32_t pk[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k0 = (c8 & 3) * 8 + 2 * i;
        const int t0 = rgb_k_tap(k0), t1 = rgb_k_tap(k0 + 1);
        float a = t0 >= 0 ? sp.w1_hwio[t0 * 32 + n] : 0.f;
        float b = t1 >= 0 ? sp.w1_hwio[t1 * 32 + n] : 0.f;
        if (c8 >= 4) {
          a -= float(Elem(a));
          b -= float(Elem(b));
        }
        pk[i] = pack2<Elem>(a, b);
      }
      *reinterpret_cast<uint4*>(s_b1 + (c8 >> 2) * 2048 + n * 64 + (((c8 & 3) ^ ((n >> 1) & 3)) * 16)) =
          make_uint4(pk[0], pk[1], pk[2], pk[3]);
    }
    sm100::fence_proxy_async_smem();
  }
  sm100::tc_fence_before_sync();
  __syncthreads();
  sm100::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_d1 = tmem_base + kStage2Cols;             // stage-1 accumulators: 5 x 32 columns

  if (warp_idx == 0) {
    // ===== conv2d_2 weights, once =====
    if (sm100::elect_one()) {
      sm100::mbar_arrive_expect_tx(b2_full, kStemB2Bytes);
      for (int tap = 0; tap < 9; ++tap)
        sm100::tma_load_3d(s_b2 + tap * (BLOCK_N * SWZ), &p.tmB, b2_full, 0, 0, tap, sm100::kEvictLast);
    }
  } else if (warp_idx == 1) {
    // ===== MMA issuer =====
    if (sm100::elect_one()) {
      constexpr uint32_t idesc = sm100::make_idesc(ElemTraits<Elem>::kFmt, kTileM, BLOCK_N);
      const uint32_t a2_hi = sm100::smem_desc_hi<SWZ>(uint32_t(kStemHaloW) * SWZ);   // 8-pixel groups one halo row apart
      const uint32_t b2_hi = sm100::smem_desc_hi<SWZ>(8u * SWZ);
      const uint32_t b2_lo0 = sm100::smem_desc_lo(sm100::smem_u32(s_b2));
      const uint32_t b1_addr = sm100::smem_u32(s_b1);
      sm100::mbar_wait(b2_full, 0);
      sm100::tc_fence_after_sync();
      int ab = 0;                 // A1 ring position
      uint32_t ab_phase = 0;
      uint32_t d1_phase = 0;      // phase of d1_empty[*] to wait for (all blocks advance together, once per tile)
      int acc = 0;
      uint32_t acc_phase = 0;
      // one stage-1 block: 4 MMAs into its 32 TMEM columns
      auto stage1_block = [&](int b) {
        sm100::mbar_wait(&a1_full[ab], ab_phase);
        sm100::mbar_wait(&d1_empty[b], d1_phase ^ 1);
        sm100::tc_fence_after_sync();
        const uint32_t a_addr = sm100::smem_u32(s_a1 + ab * kStemA1Bytes);
#pragma unroll
        for (int k = 0; k < 4; ++k)   // k >> 1: w_hi / w_lo tile; k & 1: K step inside the 64-byte row
          sm100::umma_ss<false>(tmem_d1 + uint32_t(b * 32), sm100::make_smem_desc_kmajor<64>(a_addr + (k & 1) * 32),
                                sm100::make_smem_desc_kmajor<64>(b1_addr + (k >> 1) * 2048 + (k & 1) * 32), idesc, k);
        sm100::umma_commit(&a1_empty[ab]);
        sm100::umma_commit(&d1_full[b]);
        if (++ab == kStemA1Bufs) {
          ab = 0;
          ab_phase ^= 1;
        }
      };
      int k = 0;   // tiles of this CTA issued so far (stage 1)
      for (int tile = blockIdx.x; ; tile += gridDim.x, ++k) {
        const bool have1 = tile < total_tiles;      // stage 1 of tile k
        const bool have2 = k >= 1;                  // stage 2 of tile k - 1
        if (!have1 && !have2) break;
        // stage 1 of tile k first (its im2col tiles were built during the previous iteration): its epilogue — TMEM ->
        // conv2d_2's halo tile — then runs under the 72 MMAs of stage 2 of tile k - 1 issued right behind. (Interleaving
        // the stage-1 blocks between the taps left the last block's epilogue exposed after every tile: no faster than the
        // two separate kernels.)
        if (have1) {
          for (int b = 0; b < kStemBlocks; ++b) stage1_block(b);
          d1_phase ^= 1;
        }
        if (have2) {
          const int buf = (k - 1) & 1;
          sm100::mbar_wait(&a2_full[buf], uint32_t((k - 1) >> 1) & 1u);
          sm100::mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
          sm100::tc_fence_after_sync();
          const uint32_t tmem_d = tmem_base + uint32_t(acc) * (NM * BLOCK_N);
          const uint32_t a_lo = sm100::smem_desc_lo(sm100::smem_u32(s_a2 + buf * kStemA2Bytes));
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const int ty = tap / 3, tx = tap - ty * 3;
            issue_tap<false, NM, KSTEPS, SWZ, BLOCK_N>(tmem_d, a_lo + (uint32_t((ty * kStemHaloW + tx) * SWZ) >> 4), a2_hi,
                                                       b2_lo0 + uint32_t(tap) * ((BLOCK_N * SWZ) >> 4), b2_hi, idesc,
                                                       tap == 0 ? 0u : 1u);
          }
          sm100::umma_commit(&a2_empty[buf]);
          sm100::umma_commit(&tmem_full_bar[acc]);
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1;
        }
        if (!have1) break;
      }
    }
  } else if (warp_idx >= 4 && warp_idx < 12) {
    // ===== stage-2 epilogue: conv2d_2's own (TMEM -> affine -> pack -> staging -> pool -> TMA stores) =====
    epilogue_fast<Elem, BLOCK_N, EPI2>(p, tmem_base, tmem_full_bar, tmem_empty_bar, epi_stage, nullptr, nullptr, nullptr, nullptr,
                                       warp_idx, lane);
  } else if (warp_idx >= 12 && warp_idx < 16) {
    // ===== stage-1 epilogue: conv2d_1's accumulators -> conv2d_2's halo tile in shared memory =====
    const int q = warp_idx & 3;
    const uint64_t in2 = f2_pack(sp.in_scale, sp.in_scale);
    uint32_t d1_phase = 0;
    int k = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++k) {
      int m = tile;
      const int tx = m % tiles_x;
      m /= tiles_x;
      const int ty = m % tiles_y;
      const int x0 = tx * (kBlkW * NM) - 1, y0 = ty * kBlkH - 1;      // image position of halo pixel (0, 0)
      const int buf = k & 1;
      sm100::mbar_wait(&a2_empty[buf], (uint32_t(k >> 1) & 1u) ^ 1u);  // stage 2 of tile k - 2 has consumed this buffer
      uint8_t* a2 = s_a2 + buf * kStemA2Bytes;
      for (int b = 0; b < kStemBlocks; ++b) {
        sm100::mbar_wait(&d1_full[b], d1_phase);
        sm100::tc_fence_after_sync();
        uint32_t v[32];
        sm100::tmem_ld_32x32b_x32(tmem_d1 + (uint32_t(32 * q) << 16) + uint32_t(b * 32), v);
        sm100::tmem_ld_wait();
        sm100::tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) sm100::mbar_arrive(&d1_empty[b]);
        const int pix = b * 128 + q * 32 + lane;
        if (pix < kStemHaloPix) {
          const int hy = pix / kStemHaloW, hx = pix - hy * kStemHaloW;
          const bool inside = (y0 + hy) >= 0 && (y0 + hy) < p.H && (x0 + hx) >= 0 && (x0 + hx) < p.W;
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float u0, u1, f0, f1;
            f2_unpack(f2_fma(f2_pack(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])), in2, cpair(sp.a1.bias, 2 * i)), u0, u1);
            f2_unpack(f2_fma(f2_pack(fmaxf(u0, 0.f), fmaxf(u1, 0.f)), cpair(sp.a1.scale, 2 * i), cpair(sp.a1.shift, 2 * i)), f0, f1);
            pk[i] = inside ? pack2<Elem>(f0, f1) : 0u;                 // conv2d_2 pads conv2d_1's OUTPUT with zeros
          }
          const int sw = (pix >> 1) & 3;                                // SWIZZLE_64B: chunk ^= address bits [7, 9)
#pragma unroll
          for (int c4 = 0; c4 < 4; ++c4)
            *reinterpret_cast<uint4*>(a2 + pix * 64 + ((c4 ^ sw) << 4)) = make_uint4(pk[4 * c4], pk[4 * c4 + 1], pk[4 * c4 + 2], pk[4 * c4 + 3]);
        }
      }
      d1_phase ^= 1;
      sm100::fence_proxy_async_smem();       // generic-proxy writes -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) sm100::mbar_arrive(&a2_full[buf]);
    }
  } else if (warp_idx >= 16) {
    // ===== builders: raw uint8 halo -> im2col rows of conv2d_1 =====
    const int bt = threadIdx.x - 16 * 32;    // 0..127 = row of the im2col tile
    constexpr int kWords = kStemRawRows * kStemRawWords;   // 560
    // raw word w of halo row rr of a tile: image row y0 - 2 + rr (conv2d_2 halo - 1, conv2d_1 halo - 1), image bytes
    // [3 x0 - 8 + 4 w, +4) with x0 the tile's first output column (3 x0 is a multiple of 4: whole words, aligned)
    auto fetch = [&](int tile, int idx) -> uint32_t {
      if (idx >= kWords || tile >= total_tiles) return 0u;
      int m = tile;
      const int tx = m % tiles_x;
      m /= tiles_x;
      const int ty = m % tiles_y;
      const int img = m / tiles_y;
      const int rr = idx / kStemRawWords, w = idx - rr * kStemRawWords;
      const int yy = ty * kBlkH - 2 + rr;
      const int off = tx * (kBlkW * NM) * 3 - 8 + 4 * w;
      if (yy < 0 || yy >= p.H || off < 0 || off >= 3 * p.W) return 0u;     // conv2d_1's 'same' padding
      return __ldg(reinterpret_cast<const uint32_t*>(sp.in + (size_t(img) * p.H + yy) * p.W * 3 + off));
    };
    uint32_t nw[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) nw[j] = fetch(blockIdx.x, bt + 128 * j);
    int ab = 0;
    uint32_t ab_phase = 0;
    int k = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++k) {
      uint32_t* raw = s_rawin + (k & 1) * (kStemRawBytes / 4);
#pragma unroll
      for (int j = 0; j < 5; ++j)
        if (bt + 128 * j < kWords) raw[bt + 128 * j] = nw[j];
      asm volatile("bar.sync 1, 128;" ::: "memory");     // the four builder warps: raw rows complete (and the buffer of
                                                          // tile k - 2 is no longer read: every warp passed its last block)
#pragma unroll
      for (int j = 0; j < 5; ++j) nw[j] = fetch(tile + gridDim.x, bt + 128 * j);   // next tile's words in flight
      for (int b = 0; b < kStemBlocks; ++b) {
        sm100::mbar_wait(&a1_empty[ab], ab_phase ^ 1);
        const int pix = b * 128 + bt;
        if (pix < kStemHaloPix) {
          const int hy = pix / kStemHaloW, hx = pix - hy * kStemHaloW;
          // 3x3 window of conv2d_1 around halo pixel (hy, hx): raw rows hy .. hy + 2; halo pixel hx is image column
          // x0 - 1 + hx, its left neighbour starts at image byte 3 (x0 - 2 + hx) = raw byte 3 hx + 2 (raw byte 0 is image
          // byte 3 x0 - 8): nine consecutive bytes from there
          const int sb = 3 * hx + 2;
          const int w0 = sb >> 2;
          const uint32_t sh = uint32_t(sb & 3) * 8u;
          uint32_t pr[16];
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
            const uint32_t* row = raw + (hy + dy) * kStemRawWords + w0;
            const uint32_t W0 = row[0], W1 = row[1], W2 = row[2];
            const uint32_t q0 = __funnelshift_r(W0, W1, sh), q1 = __funnelshift_r(W1, W2, sh), q2 = W2 >> sh;
            pr[dy * 5 + 0] = bytes_to_pair<Elem>(q0, 0x4140u);
            pr[dy * 5 + 1] = bytes_to_pair<Elem>(q0, 0x4342u);
            pr[dy * 5 + 2] = bytes_to_pair<Elem>(q1, 0x4140u);
            pr[dy * 5 + 3] = bytes_to_pair<Elem>(q1, 0x4342u);
            pr[dy * 5 + 4] = bytes_to_pair<Elem>(q2, 0x4140u);
          }
          pr[15] = 0u;
          uint8_t* a1 = s_a1 + ab * kStemA1Bytes + bt * 64;
          const int sw = (bt >> 1) & 3;
#pragma unroll
          for (int c4 = 0; c4 < 4; ++c4)
            *reinterpret_cast<uint4*>(a1 + ((c4 ^ sw) * 16)) = make_uint4(pr[4 * c4], pr[4 * c4 + 1], pr[4 * c4 + 2], pr[4 * c4 + 3]);
        }
        sm100::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) sm100::mbar_arrive(&a1_full[ab]);
        if (++ab == kStemA1Bufs) {
          ab = 0;
          ab_phase ^= 1;
        }
      }
    }
  }

  // ---------------------------------------------------------------- teardown
  sm100::tc_fence_before_sync();
  __syncthreads();
  if (warp_idx == 2) {
    sm100::tc_fence_after_sync();
    sm100::tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace itmo
