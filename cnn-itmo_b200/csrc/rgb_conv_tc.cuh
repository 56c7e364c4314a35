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
  float bias[32], scale[32], shift[32];
};

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
        float a = k0 < 27 ? w_hwio[k0 * 32 + tid] : 0.f;
        float b = k0 + 1 < 27 ? w_hwio[(k0 + 1) * 32 + tid] : 0.f;
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
  auto fetch_word = [&](int tile, int idx) -> uint32_t {
    if (idx >= kWords || tile >= total) return 0u;
    const int tx = tile % tiles_x;
    const int ty = (tile / tiles_x) % tiles_y;
    const int img = tile / (tiles_x * tiles_y);
    const int hy = idx / kWordsPerRow, k = idx - hy * kWordsPerRow;
    const int yy = ty * kRgbTileH + hy - 1;
    const int off = tx * kRgbTileW * 3 - 4 + 4 * k;             // byte offset inside the image row
    if (yy < 0 || yy >= H || off < 0 || off >= 3 * W) return 0u;  // 'same' padding (whole words: 3*W % 4 == 0)
    return __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(in) +
                                                   (size_t(img) * H + yy) * W * 3 + off));
  };
  uint32_t nw0 = 0, nw1 = 0;
  if (fast_u8) {
    nw0 = fetch_word(blockIdx.x, tid);
    nw1 = fetch_word(blockIdx.x, tid + 128);
  }

  for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
    const int tx = tile % tiles_x;
    const int ty = (tile / tiles_x) % tiles_y;
    const int img = tile / (tiles_x * tiles_y);
    const int x0 = tx * kRgbTileW, y0 = ty * kRgbTileH;

    // ---- halo tile -> shared memory as floats (zero outside the image = 'same' padding)
    if (fast_u8) {
      if (tid < kWords) s_raw[tid / kWordsPerRow][tid % kWordsPerRow] = nw0;
      if (tid + 128 < kWords) s_raw[(tid + 128) / kWordsPerRow][(tid + 128) % kWordsPerRow] = nw1;
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
    if (fast_u8) {   // next tile's words: in flight during the im2col, the MMA and the epilogue
      nw0 = fetch_word(tile + gridDim.x, tid);
      nw1 = fetch_word(tile + gridDim.x, tid + 128);
    }

    // ---- im2col: k = dy*9 + dx*3 + ch = dy*9 + (9 consecutive halo values starting at pixel c)
    {
      float a[32];
      if constexpr (sizeof(InT) == 1) {
        // 9 consecutive bytes from raw byte 3c + 1: three word reads + funnel shifts instead of 9 scalar reads
        const int sb = 3 * c + 1, w0 = sb >> 2;
        const uint32_t sh = uint32_t(sb & 3) * 8u;
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
          const uint32_t W0 = s_raw[r + dy][w0], W1 = s_raw[r + dy][w0 + 1], W2 = s_raw[r + dy][w0 + 2];
          const uint32_t q0 = __funnelshift_r(W0, W1, sh), q1 = __funnelshift_r(W1, W2, sh), q2 = W2 >> sh;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            a[dy * 9 + j] = float((q0 >> (8 * j)) & 0xFFu);
            a[dy * 9 + 4 + j] = float((q1 >> (8 * j)) & 0xFFu);
          }
          a[dy * 9 + 8] = float(q2 & 0xFFu);
        }
      } else {
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
          for (int j = 0; j < 9; ++j) a[dy * 9 + j] = s_halo[r + dy][c * 3 + j];
      }
#pragma unroll
      for (int j = 27; j < 32; ++j) a[j] = 0.f;
      const int sw = tid & 7;
#pragma unroll
      for (int c4 = 0; c4 < 4; ++c4) {
        const uint4 v = make_uint4(pack2<Elem>(a[c4 * 8 + 0], a[c4 * 8 + 1]), pack2<Elem>(a[c4 * 8 + 2], a[c4 * 8 + 3]),
                                   pack2<Elem>(a[c4 * 8 + 4], a[c4 * 8 + 5]), pack2<Elem>(a[c4 * 8 + 6], a[c4 * 8 + 7]));
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
#pragma unroll
    for (int i = 0; i < 8; ++i) {   // 128-bit broadcast reads: a quarter of the shared-memory wavefronts of scalar ones
      float4 b4, sc4, sh4;
      if constexpr (kConstAffine) {
        b4 = make_float4(aff.bias[4 * i], aff.bias[4 * i + 1], aff.bias[4 * i + 2], aff.bias[4 * i + 3]);
        sc4 = make_float4(aff.scale[4 * i], aff.scale[4 * i + 1], aff.scale[4 * i + 2], aff.scale[4 * i + 3]);
        sh4 = make_float4(aff.shift[4 * i], aff.shift[4 * i + 1], aff.shift[4 * i + 2], aff.shift[4 * i + 3]);
      } else {
        b4 = reinterpret_cast<const float4*>(s_bias)[i];
        sc4 = reinterpret_cast<const float4*>(s_scale)[i];
        sh4 = reinterpret_cast<const float4*>(s_shift)[i];
      }
      const float f0 = fmaf(fmaxf(fmaf(__uint_as_float(v[4 * i]), in_scale, b4.x), 0.f), sc4.x, sh4.x);
      const float f1 = fmaf(fmaxf(fmaf(__uint_as_float(v[4 * i + 1]), in_scale, b4.y), 0.f), sc4.y, sh4.y);
      const float f2 = fmaf(fmaxf(fmaf(__uint_as_float(v[4 * i + 2]), in_scale, b4.z), 0.f), sc4.z, sh4.z);
      const float f3 = fmaf(fmaxf(fmaf(__uint_as_float(v[4 * i + 3]), in_scale, b4.w), 0.f), sc4.w, sh4.w);
      pk[2 * i] = pack2<Elem>(f0, f1);
      pk[2 * i + 1] = pack2<Elem>(f2, f3);
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
  }

  sm100::tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    sm100::tc_fence_after_sync();
    sm100::tmem_dealloc(tmem, 32);
  }
}

}  // namespace itmo
