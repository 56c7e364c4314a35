// rgb_conv_tc.cuh — the first ConvBN of the U-Net (conv2d_1, 3 -> 32 channels, model.py:208) on the
// tensor cores, for the 16-bit modes.
//
// K = 3*3*3 = 27 is far too thin for the TMA-fed implicit GEMM (a 6-byte pixel is not a TMA box) and a
// CUDA-core FMA loop needs 864 FMAs per pixel, ~8x over the layer's HBM time. Here the 128 threads of a
// CTA build the im2col rows of a 4 x 32 pixel tile in shared memory themselves — one thread = one pixel =
// one 128-byte K-major row of 64 elements in the SWIZZLE_128B layout tcgen05 reads — and one thread issues
// four tcgen05.mma (M=128, N=32, K=16). The K axis is [27 taps, 5 zeros | the same 27 taps, 5 zeros] against
// weight rows [w_hi | w_lo] (w_hi = the 16-bit rounding of w, w_lo = the rounding of the remainder), so the
// weights enter with ~2x the significand of the storage type: this layer's rounding error would otherwise be
// amplified by all 18 layers behind it. Raw uint8 pixels are exact in bf16 / fp16, so the operand is the
// integer pixel value and the /255 of predict.py:59 is applied to the fp32 accumulator.
// Epilogue: TMEM -> registers -> (+bias, ReLU, BN affine) -> 64 contiguous bytes per pixel, 2 KB per warp.
// Several CTAs share an SM (23 KB of shared memory, 32 TMEM columns each) and hide each other's latencies.
#pragma once
#include "conv_igemm.cuh"

namespace itmo {

constexpr int kRgbTileH = 4, kRgbTileW = 32;    // 128 pixels = UMMA M
constexpr int kRgbHaloH = kRgbTileH + 2, kRgbHaloW = kRgbTileW + 2;

// in:  [NB,H,W,3] uint8 (raw pixels) or float (already /255); out: [NB,H,W,32] Elem
// w_hwio: [27][32] fp32 (Keras HWIO flattened), in_scale: 1/255 for uint8 input, 1 for float input
// bias / scale / shift of the 32 output channels handed over BY VALUE: kernel parameters live in the constant bank, so
// the epilogue's FFMAs take them as immediate-indexed constant operands — no shared-memory reads at all (the 24
// broadcast reads per pixel row were ~40 % of this kernel's L1 data-pipe wavefronts, its bottleneck).
struct RgbAffine {
  alignas(16) float bias[32];
  alignas(16) float scale[32];
  alignas(16) float shift[32];
};

// K slot s of the im2col row (0..31, twice: against w_hi and w_lo): kernel row dy = s / 10, j = s % 10 = 3 * dx + ch; the
// tenth slot of a kernel row and slots 30, 31 are padding (zero weights). Ten slots per kernel row keep the 16-bit pairs
// inside one kernel row, so a pair is two adjacent input bytes.
__device__ __forceinline__ int rgb_k_tap(int s) {
  const int dy = s / 10, j = s - dy * 10;
  return (dy < 3 && j < 9) ? dy * 9 + j : -1;
}

// Two raw bytes -> two exact 16-bit floats. fp16 without an integer conversion: 0x6400 | b is the fp16 number 1024 + b
// (exact for b < 256) and one packed subtraction removes the offset. `sel` picks the two bytes of q (PRMT selector):
// 0x4140 = bytes 0, 1; 0x4342 = bytes 2, 3.
template <typename Elem>
__device__ __forceinline__ uint32_t bytes_to_pair(uint32_t q, uint32_t sel) {
  if constexpr (ElemTraits<Elem>::kFmt == sm100::kFmtF16) {
    const uint32_t w = __byte_perm(q, 0x64646464u, sel);
    const uint32_t off = 0x64006400u;
    const __half2 r = __hsub2(*reinterpret_cast<const __half2*>(&w), *reinterpret_cast<const __half2*>(&off));
    return *reinterpret_cast<const uint32_t*>(&r);
  } else {
    // bf16 has 8 significand bits: no offset leaves 256 consecutive integers exact, so convert (bf16 is the opt-in mode)
    const uint32_t lo = (sel & 0xFu) * 8u, hi = ((sel >> 8) & 0xFu) * 8u;
    return pack2<Elem>(float((q >> lo) & 0xFFu), float((q >> hi) & 0xFFu));
  }
}

template <typename Elem, typename InT, bool kConstAffine = false>
__global__ void __launch_bounds__(128, 8) rgb_conv_tc_kernel(const InT* __restrict__ in, Elem* __restrict__ out,
                                                          const float* __restrict__ w_hwio,
                                                          const float* __restrict__ bias,
                                                          const float* __restrict__ scale,
                                                          const float* __restrict__ shift, int NB, int H, int W,
                                                          float in_scale, const __grid_constant__ RgbAffine aff) {
  __shared__ __align__(1024) uint8_t s_a[128 * 128];  // im2col tile, 128 rows x 128 B, SWIZZLE_128B
  __shared__ __align__(1024) uint8_t s_b[32 * 128];   // weights, 32 rows (cout) x 128 B (K), SWIZZLE_128B
  __shared__ float s_halo[kRgbHaloH][kRgbHaloW * 3 + 2];   // float input: halo values
  __shared__ uint32_t s_raw[kRgbHaloH][28];                 // uint8 input: halo rows as the raw words (26 used); byte b
                                                            // of a row = image-row byte x0*3 - 4 + b = halo value b - 1
  __shared__ __align__(16) float s_bias[32], s_scale[32], s_shift[32];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x;
  const int warp = tid >> 5;

  // ---- one-time setup: weights -> K-major swizzled B tile, epilogue vectors, barrier, TMEM
  if (tid < 32) {
    // row n = tid of B: B[n][k] = hi(w[k][n]) for k < 32, lo(w[k-32][n]) for k >= 32; taps >= 27 -> 0
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
      uint32_t pk[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k0 = (c8 & 3) * 8 + 2 * i;
        const int t0 = rgb_k_tap(k0), t1 = rgb_k_tap(k0 + 1);
        float a = t0 >= 0 ? w_hwio[t0 * 32 + tid] : 0.f;
        float b = t1 >= 0 ? w_hwio[t1 * 32 + tid] : 0.f;
        if (c8 >= 4) {  // remainder after the 16-bit rounding
          a -= float(Elem(a));
          b -= float(Elem(b));
        }
        pk[i] = pack2<Elem>(a, b);
      }
      const int pos = c8 ^ (tid & 7);
      *reinterpret_cast<uint4*>(s_b + tid * 128 + pos * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    }
    if constexpr (!kConstAffine) {
      s_bias[tid] = bias[tid];
      s_scale[tid] = scale[tid];
      s_shift[tid] = shift[tid];
    }
  }
  if (tid == 0) {
    sm100::mbar_init(&s_bar, 1);
    sm100::fence_barrier_init();
  }
  if (warp == 1) sm100::tmem_alloc(&s_tmem, 32);
  sm100::fence_proxy_async_smem();
  sm100::tc_fence_before_sync();
  __syncthreads();
  sm100::tc_fence_after_sync();
  const uint32_t tmem = s_tmem;

  const int tiles_x = W / kRgbTileW, tiles_y = H / kRgbTileH;
  const int total = NB * tiles_y * tiles_x;
  const int r = tid >> 5, c = tid & 31;     // this thread's pixel inside the tile (M row = tid)
  uint32_t phase = 0;
  constexpr uint32_t idesc = sm100::make_idesc(ElemTraits<Elem>::kFmt, 128, 32);

  // uint8 input: the halo rows are fetched as aligned 32-bit words (26 words cover the 102 bytes of a halo row),
  // two independent loads per thread, issued one tile ahead so that their latency hides behind the MMA and the
  // epilogue of the current tile. Needs a 4-byte aligned input pointer (W is a multiple of 32, so every row is).
  constexpr int kWordsPerRow = 26, kWords = kRgbHaloH * kWordsPerRow;   // 156
  const bool fast_u8 = sizeof(InT) == 1 && (reinterpret_cast<uintptr_t>(in) & 3) == 0;
  // Tile coordinates advance by gridDim.x tiles per iteration: the stride is decomposed ONCE into (images, tile rows,
  // tile columns) and the coordinates move by adds with carries — the per-tile divisions by run-time values (seven of
  // them, ~250 integer instructions per warp and tile) were 40 % of this kernel's issue slots, its bottleneck.
  const int per_img = tiles_x * tiles_y;
  const int st_img = int(gridDim.x) / per_img, st_rem = int(gridDim.x) - st_img * per_img;
  const int st_ty = st_rem / tiles_x, st_tx = st_rem - st_ty * tiles_x;
  int img = int(blockIdx.x) / per_img;
  int ty = (int(blockIdx.x) - img * per_img) / tiles_x;
  int tx = int(blockIdx.x) - img * per_img - ty * tiles_x;
  auto advance = [&](int& ax, int& ay, int& ai) {
    ax += st_tx;
    if (ax >= tiles_x) {
      ax -= tiles_x;
      ++ay;
    }
    ay += st_ty;
    if (ay >= tiles_y) {
      ay -= tiles_y;
      ++ai;
    }
    ai += st_img;
  };
  // this thread's two halo words: (halo row, word) are loop-invariant
  const int hy0 = tid / kWordsPerRow, k0 = tid - hy0 * kWordsPerRow;
  const int hy1 = (tid + 128) / kWordsPerRow, k1 = (tid + 128) - hy1 * kWordsPerRow;
  const bool has1 = tid + 128 < kWords;
  auto fetch_word = [&](int ax, int ay, int ai, int hy, int k) -> uint32_t {
    const int yy = ay * kRgbTileH + hy - 1;
    const int off = ax * (kRgbTileW * 3) - 4 + 4 * k;           // byte offset inside the image row
    if (yy < 0 || yy >= H || off < 0 || off >= 3 * W) return 0u;  // 'same' padding (whole words: 3*W % 4 == 0)
    return __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(in) +
                                                   (size_t(ai) * H + yy) * W * 3 + off));
  };
  uint32_t nw0 = 0, nw1 = 0;
  if (fast_u8 && int(blockIdx.x) < total) {
    nw0 = fetch_word(tx, ty, img, hy0, k0);
    if (has1) nw1 = fetch_word(tx, ty, img, hy1, k1);
  }
  int ntx = tx, nty = ty, nimg = img;
  advance(ntx, nty, nimg);

  for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
    const int x0 = tx * kRgbTileW, y0 = ty * kRgbTileH;

    // ---- halo tile -> shared memory as floats (zero outside the image = 'same' padding)
    if (fast_u8) {
      s_raw[hy0][k0] = nw0;
      if (has1) s_raw[hy1][k1] = nw1;
    } else {
      for (int i = tid; i < kRgbHaloH * kRgbHaloW * 3; i += 128) {
        const int hy = i / (kRgbHaloW * 3), rem = i - hy * (kRgbHaloW * 3);
        const int hx = rem / 3;
        const int yy = y0 + hy - 1, xx = x0 + hx - 1;
        const bool inside = yy >= 0 && yy < H && xx >= 0 && xx < W;
        if constexpr (sizeof(InT) == 1) {   // pointer not 4-byte aligned: byte loads into the same raw layout
          reinterpret_cast<uint8_t*>(&s_raw[hy][0])[rem + 1] =
              inside ? uint8_t(in[((size_t(img) * H + yy) * W + x0 - 1) * 3 + rem]) : uint8_t(0);
        } else {
          s_halo[hy][rem] = inside ? float(in[((size_t(img) * H + yy) * W + x0 - 1) * 3 + rem]) : 0.f;
        }
      }
    }
    __syncthreads();
    if (fast_u8 && tile + int(gridDim.x) < total) {   // next tile's words: in flight during the im2col, the MMA and the epilogue
      nw0 = fetch_word(ntx, nty, nimg, hy0, k0);
      if (has1) nw1 = fetch_word(ntx, nty, nimg, hy1, k1);
    }

    // ---- im2col: slot s = dy*10 + j, j = 9 consecutive halo bytes starting at pixel c (see rgb_k_tap)
    {
      uint32_t pr[16];   // the row's 16 pairs of 16-bit operands
      if constexpr (sizeof(InT) == 1) {
        // 9 consecutive bytes from raw byte 3c + 1: three word reads + funnel shifts, then five PRMT + five packed
        // subtractions per kernel row (bytes_to_pair) instead of 9 x (extract, I2F) + 5 packs
        const int sb = 3 * c + 1, w0 = sb >> 2;
        const uint32_t sh = uint32_t(sb & 3) * 8u;
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
          const uint32_t W0 = s_raw[r + dy][w0], W1 = s_raw[r + dy][w0 + 1], W2 = s_raw[r + dy][w0 + 2];
          const uint32_t q0 = __funnelshift_r(W0, W1, sh), q1 = __funnelshift_r(W1, W2, sh), q2 = W2 >> sh;
          pr[dy * 5 + 0] = bytes_to_pair<Elem>(q0, 0x4140u);
          pr[dy * 5 + 1] = bytes_to_pair<Elem>(q0, 0x4342u);
          pr[dy * 5 + 2] = bytes_to_pair<Elem>(q1, 0x4140u);
          pr[dy * 5 + 3] = bytes_to_pair<Elem>(q1, 0x4342u);
          pr[dy * 5 + 4] = bytes_to_pair<Elem>(q2, 0x4140u);   // (j = 8, padding slot: any finite value, its weight is 0)
        }
      } else {
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
          for (int jj = 0; jj < 4; ++jj)
            pr[dy * 5 + jj] = pack2<Elem>(s_halo[r + dy][c * 3 + 2 * jj], s_halo[r + dy][c * 3 + 2 * jj + 1]);
          pr[dy * 5 + 4] = pack2<Elem>(s_halo[r + dy][c * 3 + 8], 0.f);
        }
      }
      pr[15] = 0u;
      const int sw = tid & 7;
#pragma unroll
      for (int c4 = 0; c4 < 4; ++c4) {
        const uint4 v = make_uint4(pr[4 * c4], pr[4 * c4 + 1], pr[4 * c4 + 2], pr[4 * c4 + 3]);
        *reinterpret_cast<uint4*>(s_a + tid * 128 + ((c4 ^ sw) * 16)) = v;        // pairs with w_hi
        *reinterpret_cast<uint4*>(s_a + tid * 128 + (((c4 + 4) ^ sw) * 16)) = v;  // pairs with w_lo
      }
    }
    sm100::fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor core (async proxy)
    sm100::tc_fence_before_sync();     // orders the previous tile's tcgen05.ld before this tile's MMA
    __syncthreads();

    if (tid == 0) {
      sm100::tc_fence_after_sync();
      const uint32_t a_addr = sm100::smem_u32(s_a), b_addr = sm100::smem_u32(s_b);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        sm100::umma_ss<false>(tmem, sm100::make_smem_desc_kmajor<128>(a_addr + k * 32),
                              sm100::make_smem_desc_kmajor<128>(b_addr + k * 32), idesc, k);
      sm100::umma_commit(&s_bar);
    }
    sm100::mbar_wait(&s_bar, phase);
    phase ^= 1;
    sm100::tc_fence_after_sync();

    // ---- epilogue: one pixel per thread, 32 channels = 64 contiguous bytes
    uint32_t v[32];
    sm100::tmem_ld_32x32b_x32(tmem + (uint32_t(32 * warp) << 16), v);
    sm100::tmem_ld_wait();
    uint32_t pk[16];
    const uint64_t in2 = f2_pack(in_scale, in_scale);
#pragma unroll
    for (int i = 0; i < 16; ++i) {   // packed fp32 pairs: FFMA2, 2 x FMNMX, FFMA2 per channel pair
      uint64_t b2, sc2, sh2;
      if constexpr (kConstAffine) {
        b2 = cpair(aff.bias, 2 * i);
        sc2 = cpair(aff.scale, 2 * i);
        sh2 = cpair(aff.shift, 2 * i);
      } else {
        b2 = *reinterpret_cast<const uint64_t*>(s_bias + 2 * i);
        sc2 = *reinterpret_cast<const uint64_t*>(s_scale + 2 * i);
        sh2 = *reinterpret_cast<const uint64_t*>(s_shift + 2 * i);
      }
      float u0, u1, f0, f1;
      f2_unpack(f2_fma(f2_pack(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])), in2, b2), u0, u1);
      f2_unpack(f2_fma(f2_pack(fmaxf(u0, 0.f), fmaxf(u1, 0.f)), sc2, sh2), f0, f1);
      pk[i] = pack2<Elem>(f0, f1);
    }
    // The 32 pixels of a warp are 2 KB of contiguous output. Stage them in the (now idle) im2col tile with a
    // 16-byte-piece swizzle, then store 512 contiguous bytes per instruction instead of 32 pieces 64 B apart.
    uint8_t* stage = s_a + warp * 2048;
    const int lane = tid & 31;
#pragma unroll
    for (int i = 0; i < 4; ++i)
      *reinterpret_cast<uint4*>(stage + lane * 64 + ((i ^ ((lane >> 1) & 3)) * 16)) =
          make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
    __syncwarp();
    uint4* dst = reinterpret_cast<uint4*>(out + ((size_t(img) * H + y0 + r) * W + x0) * 32);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int u = k * 32 + lane, px = u >> 2;
      dst[u] = *reinterpret_cast<const uint4*>(stage + px * 64 + (((u & 3) ^ ((px >> 1) & 3)) * 16));
    }
    tx = ntx;
    ty = nty;
    img = nimg;
    advance(ntx, nty, nimg);
  }

  sm100::tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    sm100::tc_fence_after_sync();
    sm100::tmem_dealloc(tmem, 32);
  }
}

}  // namespace itmo
