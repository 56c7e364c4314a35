// conv_igemm.cuh — implicit-GEMM convolution kernel for the CNN-ITMO U-Net on sm_100a.
//
// One kernel template covers the two GEMM-shaped blocks of the reference graph
//   ConvBN          (model.py:195-196)  = BN(ReLU(Conv2D(f, 3, padding='same')(x)))
//   ConvBNTranspose (model.py:199-200)  = BN(ReLU(Conv2DTranspose(f, 2, strides=2)(x)))
// plus the fusions the hot path needs around them:
//   * concatenate([skip, up], axis=3) (model.py:246,251,256,261) folded into the operand loader:
//     the K loop walks two NHWC source tensors through two TMA tensor maps, no concat buffer;
//   * MaxPooling2D((2,2), strides=2) (model.py:210,215,220,227) fused into the epilogue
//     (both the full-resolution skip tensor and the pooled tensor are written);
//   * the 1x1 sigmoid head Conv2D(3, 1, activation='sigmoid') (model.py:276) and the
//     uint8 truncation of predict.py:64 fused into the epilogue of the last 3x3 conv.
//
// GEMM view (per output tile):  D[128 pixels, BLOCK_N couts] = sum over (tap, cin-block) of
//   A[128 pixels, BK cins] (an 8x16-pixel NHWC box shifted by the tap, TMA zero-fills the halo
//   => 'same' padding for free) x B[BLOCK_N couts, BK cins] (pre-packed K-major weights).
// Both operands land in shared memory in the canonical K-major swizzled layout via TMA, a single
// thread issues tcgen05.mma with the fp32 accumulator in TMEM (double buffered so the epilogue of
// tile i overlaps the main loop of tile i+1), four epilogue warps read TMEM with tcgen05.ld and
// apply  y = max(acc + bias, 0) * bn_scale + bn_shift  (BN sits AFTER ReLU in the reference).
#pragma once
#include <cuda_bf16.h>
#include <stdio.h>
#include "sm100_ptx.cuh"

namespace itmo {

constexpr int kTileH = 8;    // output tile: 8 rows x 16 cols = 128 pixels = UMMA M
constexpr int kTileW = 16;
constexpr int kTileM = 128;
constexpr int kMaxCout = 512;

enum ConvMode : int { kModeConv3x3 = 0, kModeConvT2x2 = 1 };
enum HeadDtype : int { kHeadNone = 0, kHeadF32 = 1, kHeadU8 = 2 };

struct ConvParams {
  CUtensorMap tmA0;  // source 0 (skip / only source): dims (C0, W, H, N), box (BK, 16, 8, 1)
  CUtensorMap tmA1;  // source 1 (upsampled) or a copy of tmA0 when unused
  CUtensorMap tmB;   // packed weights: dims (Cin_total, rows, taps), box (BK, BLOCK_N, 1)
  int H, W, NB;      // input spatial size and batch
  int kb0, kb1;      // number of BK-wide channel blocks taken from source 0 / source 1
  int taps;          // 9 for the 3x3 conv, 1 for the transposed conv
  int n_tiles;       // weight rows / BLOCK_N
  int cout;          // output channels (per 2x2 phase for the transposed conv)
  int mode;          // ConvMode
  const float* bias;      // [cout]
  const float* bn_scale;  // [cout]  gamma / sqrt(var + eps)
  const float* bn_shift;  // [cout]  beta - mean * scale
  void* out;              // NHWC activations (bf16 or fp32); may be null when the head is fused
  void* pooled;           // NHWC 2x2-max-pooled copy or null
  const float* head_w;    // [3][64] 1x1 head weights (row k = output channel k) or null
  const float* head_b;    // [3]
  void* head_out;         // [NB, H, W, 3] f32 or u8
  int head_dtype;         // HeadDtype
};

template <typename Elem>
struct ElemTraits;
template <>
struct ElemTraits<__nv_bfloat16> {
  static constexpr bool kTF32 = false;
  static constexpr uint32_t kFmt = sm100::kFmtBF16;
};
template <>
struct ElemTraits<float> {
  static constexpr bool kTF32 = true;
  static constexpr uint32_t kFmt = sm100::kFmtTF32;
};

template <typename Elem, int BLOCK_N, int SWZ>
struct ConvCfg {
  static constexpr int kBK = SWZ / int(sizeof(Elem));     // channels per K block
  static constexpr int kKSteps = SWZ / 32;                // one UMMA consumes 32 bytes of K
  static constexpr int kABytes = kTileM * SWZ;
  static constexpr int kBBytes = BLOCK_N * SWZ;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBudget = 196 * 1024;
  static constexpr int kStagesRaw = kBudget / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = (2 * BLOCK_N) < 32 ? 32 : (2 * BLOCK_N);
  // barriers + tmem ptr + epilogue params (bias, scale, shift) + head (3*64 + 4)
  static constexpr int kAuxBytes = 512 + 3 * kMaxCout * 4 + (3 * 64 + 4) * 4;
  static constexpr int kSmemBytes = 1024 /*align slack*/ + kStages * kStageBytes + kAuxBytes;
  static_assert(kStages >= 2, "pipeline needs at least two stages");
  static_assert((kTmemCols & (kTmemCols - 1)) == 0 && kTmemCols <= 512, "TMEM columns: power of two <= 512");
};

// Pack two floats into a bf16x2 register (round-to-nearest-even), low half = a.
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

template <typename Elem, int BLOCK_N, int SWZ>
__global__ void __launch_bounds__(256, 1) conv_igemm_kernel(const __grid_constant__ ConvParams p) {
  using Cfg = ConvCfg<Elem, BLOCK_N, SWZ>;
  constexpr int kStages = Cfg::kStages;
  constexpr bool kTF32 = ElemTraits<Elem>::kTF32;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = sm100::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

  uint8_t* smem_a = smem;                                   // kStages x [128 rows x SWZ bytes]
  uint8_t* smem_b = smem + kStages * Cfg::kABytes;          // kStages x [BLOCK_N rows x SWZ bytes]
  uint8_t* aux = smem + kStages * Cfg::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(aux);    // [kStages]
  uint64_t* empty_bar = full_bar + kStages;                 // [kStages]
  uint64_t* tmem_full_bar = empty_bar + kStages;            // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;             // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  float* s_bias = reinterpret_cast<float*>(aux + 512);
  float* s_scale = s_bias + kMaxCout;
  float* s_shift = s_scale + kMaxCout;
  float* s_head = s_shift + kMaxCout;                       // [3*64 weights][3 bias]

  const int warp_idx = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int tiles_x = p.W / kTileW;
  const int tiles_y = p.H / kTileH;
  const int m_tiles = p.NB * tiles_y * tiles_x;
  const int total_tiles = m_tiles * p.n_tiles;
  const int k_iters = (p.kb0 + p.kb1) * p.taps;

  // ---------------------------------------------------------------- one-time setup
  if (warp_idx == 0 && lane == 0) {
    sm100::tma_prefetch_desc(&p.tmA0);
    sm100::tma_prefetch_desc(&p.tmA1);
    sm100::tma_prefetch_desc(&p.tmB);
  }
  if (warp_idx == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      sm100::mbar_init(&full_bar[i], 1);
      sm100::mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      sm100::mbar_init(&tmem_full_bar[i], 1);
      sm100::mbar_init(&tmem_empty_bar[i], 4);  // one arrive per epilogue warp
    }
    sm100::fence_barrier_init();
  }
  if (warp_idx == 2) {
    sm100::tmem_alloc(tmem_ptr_smem, Cfg::kTmemCols);
  }
  for (int i = threadIdx.x; i < p.cout; i += blockDim.x) {
    s_bias[i] = p.bias[i];
    s_scale[i] = p.bn_scale[i];
    s_shift[i] = p.bn_shift[i];
  }
  if (p.head_dtype != kHeadNone) {
    for (int i = threadIdx.x; i < 3 * 64 + 3; i += blockDim.x)
      s_head[i] = (i < 3 * 64) ? p.head_w[i] : p.head_b[i - 3 * 64];
  }
  sm100::tc_fence_before_sync();
  __syncthreads();
  sm100::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // ---------------------------------------------------------------- warp roles
  if (warp_idx == 0) {
    // ===== TMA producer (one thread) =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n_tile = tile % p.n_tiles;
        int m = tile / p.n_tiles;
        const int tx = m % tiles_x;
        m /= tiles_x;
        const int ty = m % tiles_y;
        const int img = m / tiles_y;
        const int x0 = tx * kTileW, y0 = ty * kTileH;
        for (int kb = 0; kb < p.kb0 + p.kb1; ++kb) {
          const bool second = kb >= p.kb0;
          const CUtensorMap* tmA = second ? &p.tmA1 : &p.tmA0;
          const int c_off = (second ? kb - p.kb0 : kb) * Cfg::kBK;
          for (int tap = 0; tap < p.taps; ++tap) {
            const int dy = (p.taps == 9) ? tap / 3 - 1 : 0;
            const int dx = (p.taps == 9) ? tap % 3 - 1 : 0;
            sm100::mbar_wait(&empty_bar[stage], phase ^ 1);
            sm100::mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
            sm100::tma_load_4d(smem_a + stage * Cfg::kABytes, tmA, &full_bar[stage], c_off, x0 + dx, y0 + dy, img,
                               sm100::kEvictNormal);
            sm100::tma_load_3d(smem_b + stage * Cfg::kBBytes, &p.tmB, &full_bar[stage], kb * Cfg::kBK,
                               n_tile * BLOCK_N, tap, sm100::kEvictLast);
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp_idx == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc = sm100::make_idesc(ElemTraits<Elem>::kFmt, kTileM, BLOCK_N);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        sm100::mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
        sm100::tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
        for (int it = 0; it < k_iters; ++it) {
          sm100::mbar_wait(&full_bar[stage], phase);
          sm100::tc_fence_after_sync();
          const uint32_t a_addr = sm100::smem_u32(smem_a + stage * Cfg::kABytes);
          const uint32_t b_addr = sm100::smem_u32(smem_b + stage * Cfg::kBBytes);
#pragma unroll
          for (int k = 0; k < Cfg::kKSteps; ++k) {
            const uint64_t da = sm100::make_smem_desc_kmajor<SWZ>(a_addr + k * 32);
            const uint64_t db = sm100::make_smem_desc_kmajor<SWZ>(b_addr + k * 32);
            sm100::umma_ss<kTF32>(tmem_d, da, db, idesc, (it | k) != 0 ? 1u : 0u);
          }
          sm100::umma_commit(&empty_bar[stage]);  // frees this smem stage once the MMAs retire
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        sm100::umma_commit(&tmem_full_bar[acc]);  // accumulator complete -> epilogue
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp_idx >= 4) {
    // ===== epilogue warps: TMEM -> registers -> bias/ReLU/BN (-> pool / head) -> global =====
    const int e = warp_idx - 4;             // TMEM lane quarter (warp_idx % 4)
    const int r = 2 * e + (lane >> 4);      // pixel row inside the 8x16 tile
    const int c = lane & 15;                // pixel column inside the tile
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int n_tile = tile % p.n_tiles;
      int m = tile / p.n_tiles;
      const int tx = m % tiles_x;
      m /= tiles_x;
      const int ty = m % tiles_y;
      const int img = m / tiles_y;
      const int y = ty * kTileH + r, x = tx * kTileW + c;

      sm100::mbar_wait(&tmem_full_bar[acc], acc_phase);
      sm100::tc_fence_after_sync();
      const uint32_t taddr = tmem_base + (uint32_t(32 * e) << 16) + acc * BLOCK_N;

      float h0 = 0.f, h1 = 0.f, h2 = 0.f;
#pragma unroll 1
      for (int chunk = 0; chunk < BLOCK_N / 16; ++chunk) {
        uint32_t v[16];
        sm100::tmem_ld_32x32b_x16(taddr + chunk * 16, v);
        sm100::tmem_ld_wait();
        const int n_glob = n_tile * BLOCK_N + chunk * 16;  // row of the packed weight matrix
        int co0, oy, ox, Ho, Wo;
        if (p.mode == kModeConvT2x2) {
          const int ph = n_glob / p.cout;  // phase a*2+b of the 2x2 transposed kernel
          co0 = n_glob - ph * p.cout;
          oy = 2 * y + (ph >> 1);
          ox = 2 * x + (ph & 1);
          Ho = 2 * p.H;
          Wo = 2 * p.W;
        } else {
          co0 = n_glob;
          oy = y;
          ox = x;
          Ho = p.H;
          Wo = p.W;
        }
        float f[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float a = __uint_as_float(v[j]) + s_bias[co0 + j];
          f[j] = fmaf(fmaxf(a, 0.f), s_scale[co0 + j], s_shift[co0 + j]);
        }
        if (p.head_dtype != kHeadNone) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            h0 = fmaf(f[j], s_head[0 * 64 + co0 + j], h0);
            h1 = fmaf(f[j], s_head[1 * 64 + co0 + j], h1);
            h2 = fmaf(f[j], s_head[2 * 64 + co0 + j], h2);
          }
        }
        const size_t pix = (size_t(img) * Ho + oy) * Wo + ox;
        if constexpr (sizeof(Elem) == 2) {
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
          if (p.out != nullptr) {
            uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + pix * p.cout + co0);
            dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          }
          if (p.pooled != nullptr) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&pk[j]);
              uint32_t o = __shfl_xor_sync(0xFFFFFFFFu, pk[j], 1);
              a = __hmax2(a, *reinterpret_cast<__nv_bfloat162*>(&o));
              uint32_t au = *reinterpret_cast<uint32_t*>(&a);
              o = __shfl_xor_sync(0xFFFFFFFFu, au, 16);
              a = __hmax2(a, *reinterpret_cast<__nv_bfloat162*>(&o));
              pk[j] = *reinterpret_cast<uint32_t*>(&a);
            }
            if ((lane & 17) == 0) {
              const size_t ppix = (size_t(img) * (p.H / 2) + (y >> 1)) * (p.W / 2) + (x >> 1);
              uint4* dst =
                  reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.pooled) + ppix * p.cout + co0);
              dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
          }
        } else {
          if (p.out != nullptr) {
            float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + pix * p.cout + co0);
#pragma unroll
            for (int j = 0; j < 4; ++j) dst[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
          }
          if (p.pooled != nullptr) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              f[j] = fmaxf(f[j], __shfl_xor_sync(0xFFFFFFFFu, f[j], 1));
              f[j] = fmaxf(f[j], __shfl_xor_sync(0xFFFFFFFFu, f[j], 16));
            }
            if ((lane & 17) == 0) {
              const size_t ppix = (size_t(img) * (p.H / 2) + (y >> 1)) * (p.W / 2) + (x >> 1);
              float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.pooled) + ppix * p.cout + co0);
#pragma unroll
              for (int j = 0; j < 4; ++j) dst[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
            }
          }
        }
      }
      // accumulator drained: hand the TMEM buffer back to the MMA warp
      sm100::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) sm100::mbar_arrive(&tmem_empty_bar[acc]);

      if (p.head_dtype != kHeadNone) {
        // Conv2D(3, 1, activation='sigmoid') (model.py:276) on the post-BN fp32 values
        const float z0 = h0 + s_head[192], z1 = h1 + s_head[193], z2 = h2 + s_head[194];
        const float q0 = 1.f / (1.f + expf(-z0)), q1 = 1.f / (1.f + expf(-z1)), q2 = 1.f / (1.f + expf(-z2));
        const size_t pix = (size_t(img) * p.H + y) * p.W + x;
        if (p.head_dtype == kHeadF32) {
          float* dst = reinterpret_cast<float*>(p.head_out) + pix * 3;
          dst[0] = q0;
          dst[1] = q1;
          dst[2] = q2;
        } else {
          // predict.py:64  (pred * 255).astype('uint8')  -> fp32 multiply, truncation toward zero
          uint8_t* dst = reinterpret_cast<uint8_t*>(p.head_out) + pix * 3;
          dst[0] = static_cast<uint8_t>(static_cast<int>(q0 * 255.f));
          dst[1] = static_cast<uint8_t>(static_cast<int>(q1 * 255.f));
          dst[2] = static_cast<uint8_t>(static_cast<int>(q2 * 255.f));
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  // ---------------------------------------------------------------- teardown
  sm100::tc_fence_before_sync();
  __syncthreads();
  if (warp_idx == 2) {
    sm100::tc_fence_after_sync();
    sm100::tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

}  // namespace itmo
