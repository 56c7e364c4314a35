// conv_igemm.cuh — implicit-GEMM convolution kernel for the CNN-ITMO U-Net on sm_100a.
//
// One kernel template covers the two GEMM-shaped blocks of the reference graph
//   ConvBN          (model.py:195-196)  = BN(ReLU(Conv2D(f, 3, padding='same')(x)))
//   ConvBNTranspose (model.py:199-200)  = BN(ReLU(Conv2DTranspose(f, 2, strides=2)(x)))
// plus the fusions the hot path needs around them:
//   * concatenate([skip, up], axis=3) (model.py:246,251,256,261) folded into the operand loader:
//     the K loop walks a list of (source tensor, channel blocks) segments through TMA tensor maps,
//     no concat buffer. The same mechanism carries the 3xTF32 split of the fp32 mode
//     (x*w ~= x*w_hi + x*w_lo + x_lo*w_hi: three segments per source);
//   * MaxPooling2D((2,2), strides=2) (model.py:210,215,220,227) fused into the epilogue
//     (both the full-resolution skip tensor and the pooled tensor are written);
//   * the 1x1 sigmoid head Conv2D(3, 1, activation='sigmoid') (model.py:276) and the
//     uint8 truncation of predict.py:64 fused into the epilogue of the last 3x3 conv.
//
// GEMM view. A CTA tile is 16 rows x (8*NM) columns of output pixels = NM "M blocks" of
// 128 pixels (16 rows x 8 columns each, UMMA M = 128) that share every weight tile. For each
// block of BK input channels ONE TMA box brings the (16+2) x (8*NM+2) pixel halo tile into
// shared memory (TMA zero-fills outside the image => 'same' padding for free); the nine taps of
// the 3x3 kernel are nine shared-memory descriptors into that one halo tile (start address
// shifted by whole pixel rows, 8-pixel groups one halo row apart), so the activations cross
// L2->SMEM ~1.3x instead of 9x. Weight tiles [BLOCK_N couts x BK cins] per (channel block, tap)
// stream through their own ring, or — when the whole layer fits — are loaded once per CTA and
// stay resident. A single thread issues tcgen05.mma with the fp32 accumulators in TMEM (double
// buffered: the epilogue of tile i overlaps the main loop of tile i+1); eight epilogue warps
// read TMEM with tcgen05.ld and apply  y = max(acc + bias, 0) * bn_scale + bn_shift
// (BN sits AFTER ReLU in the reference, so it cannot be folded into the weights).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include "sm100_ptx.cuh"

namespace itmo {

constexpr int kBlkH = 16;    // M block: 16 rows x 8 cols = 128 pixels = UMMA M
constexpr int kBlkW = 8;
constexpr int kTileM = 128;
constexpr int kMaxCout = 512;
constexpr int kMaxSeg = 6;
constexpr int kMaxAStages = 8;
constexpr int kMaxBStages = 8;
constexpr int kNumThreads = 384;       // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4..11 epilogue
constexpr int kNumEpiWarps = 8;
constexpr int kSmemLimit = 227 * 1024;
constexpr int kAuxBytes = 512 + 3 * kMaxCout * 4 + (3 * 64 + 4) * 4;

enum ConvMode : int { kModeConv3x3 = 0, kModeConvT2x2 = 1 };
enum HeadDtype : int { kHeadNone = 0, kHeadF32 = 1, kHeadU8 = 2 };
// how fp32 results are stored (Elem = float): as computed, rounded to TF32 (so the tensor core's
// operand truncation is a no-op and the error is unbiased), or split into a TF32 head + residual.
enum StoreMode : int { kStorePlain = 0, kStoreTf32Rna = 1, kStoreSplit = 2 };
// Epilogue variant compiled into the kernel (template parameter EPI). 0 = the generic epilogue (every mode behind
// run-time flags: training statistics, split accumulators, fp32 storage, linear / identity, odd tile shapes). The
// others are the inference forward pass's four shapes as straight-line code for the 16-bit element types — ncu
// counted ~480 warp instructions per 32-column chunk in the generic loop, ~46 of them branches, against ~150 the
// arithmetic needs, and the HBM-bound layers were bound by exactly that issue stream (two epilogue warps per
// scheduler). | kEpiConstAffine: bias / scale / shift as constant-bank operands (cout <= 64).
// | kEpiPair: the kernel runs as CTA PAIRS (cluster of two CTAs on one TPC, tcgen05.mma.cta_group::2): the two CTAs
// work on two horizontally adjacent tiles as one M = 256 MMA. Each CTA loads its own halo tile but only HALF of every
// weight tile (the tensor core reads the other half from the peer's shared memory), so weight traffic L2 -> SMEM and
// the weight-operand shared-memory reads per MMA halve: N = 64 MMAs go from 6 KB to 5 KB of operand reads each
// (conv2d_14 is bound by exactly that), the N = 256 layers stream 4 KB instead of 8 KB of weights per MMA.
// | kEpiHeadConst (with kEpiHead): the head's folded weights as 64-bit constant-bank operands too (ConvParams::c_head) — no
// shared-memory reads left in conv2d_14's epilogue, whose MMAs already use ~90 % of the shared-memory pipe.
// Training step (bf16): kEpiStats = z = max(acc + b, 0) stored as is + the BatchNormalization batch statistics of the stored
// values (ConvParams::stats), 3x3 and transposed convolutions alike; kEpiIdentity = the accumulator stored as is (the
// data-gradient convolutions).
// | kEpiRows (conv_rows.cuh): the tile is nm image rows x 128 pixels, M block j = output row nm - 1 - j of the tile and
// a TMEM lane is a pixel column, so an epilogue warp holds 32 consecutive pixels of one row (store box 32 x 1).
enum EpiKind : int { kEpiGeneric = 0, kEpiConv = 1, kEpiConvPool = 2, kEpiConvT = 3, kEpiHead = 4, kEpiStats = 5, kEpiIdentity = 6, kEpiConstAffine = 16, kEpiPair = 32, kEpiHeadConst = 64, kEpiRows = 128 };

struct ConvParams {
  CUtensorMap tmA[4];  // activation sources: dims (C, W, H, N), box (BK, 8*NM+2*pad, 16+2*pad, 1)
  CUtensorMap tmB;     // packed weights: dims (K_total, rows, taps), box (BK, BLOCK_N, 1)
  // TMA-store epilogue (16-bit element types): every epilogue warp stages its 4 x 8 pixels x 64 (32) channels in a
  // swizzled shared-memory tile and one lane issues cp.async.bulk.tensor stores — no per-lane global stores, no
  // transposition shuffles on the L1 data pipe the tensor core's operand reads share.
  //   tmOut[0]: 3x3 conv: the output (C, W, H, N), box (64 | 32 channels, 8, 4, 1).
  //   tmOut[a], a = 0 / 1: transposed conv: output rows 2y + a viewed as ((b, co) = 2C, W, H, N) with a row stride of
  //             two output rows — the 2x2 pixel shuffle is then a plain tile store, N columns a*2C .. (a+1)*2C - 1.
  //   tmPool:   the 2x2-max-pooled copy (C, W/2, H/2, N), box (64 | 32, 4, 2, 1).
  CUtensorMap tmOut[2];
  CUtensorMap tmPool;
  int tma_store;       // 1: use the maps above
  int epi_bytes;       // shared memory of the staging tiles (0 when tma_store == 0)
  int epi_bufs;        // staging tiles per epilogue warp (1, or 2: the next group is staged while the TMA reads the last)
  int H, W, NB;        // input spatial size and batch
  int n_seg;           // K segments; segment s reads seg_kb[s] channel blocks from tmA[seg_map[s]]
  int seg_map[kMaxSeg];
  int seg_kb[kMaxSeg];
  int taps;            // 9 for the 3x3 conv, 1 for the transposed conv
  int pad;             // 1 for the 3x3 conv, 0 for the transposed conv
  int nm;              // M blocks per CTA tile (1, 2 or 4)
  int n_tiles;         // weight rows / BLOCK_N
  int cout;            // output channels (per 2x2 phase for the transposed conv)
  int mode;            // ConvMode
  int a_stage_bytes;   // halo tile bytes rounded up to 1024
  int n_a_stages, n_b_stages;
  int b_resident;      // 1: all (channel block, tap) weight tiles are loaded once and kept in shared memory
  int n_stationary;    // 1 (with b_resident, several N tiles): a CTA (pair) keeps ONE N tile — (block index) mod n_tiles —
                       // for all its M tiles, so that N tile's weights stay resident instead of being re-streamed from L2
                       // for every 128 pixels (the transposed convolutions were bound by exactly that: convT_3 moved
                       // 192 KB L2 -> SMEM per tile of which 128 KB weights). The grid is a multiple of n_tiles.
  int store_mode;      // StoreMode (Elem = float only)
  int acc_split;       // S >= 1 TMEM accumulators per M block, summed in fp32 registers by the epilogue. The
                       // tensor core adds into its fp32 accumulator with truncation, a bias that grows with the
                       // number of chained MMAs; the fp32 mode spreads the x*w_hi products round-robin over S-1
                       // accumulators and keeps the two small cross terms in the last one.
  int dbg;             // timing experiments only (results wrong): 16 no activation TMA, 32 no epilogue, 64 no MMAs
  int kb_hi;           // channel blocks of the leading (x*w_hi) part; the rest go to accumulator S-1
  int cout_log2;       // log2(cout) when cout is a power of two, else -1
  int identity_epi;    // 1 (with no_relu): bias = 0, scale = 1, shift = 0 — the epilogue stores the accumulator as is
  int no_relu;         // 1: y = (acc + bias) * scale + shift without the ReLU (data-gradient convolutions of the training step)
  float* stats;        // training forward (16-bit modes): BatchNormalization batch statistics (model.py:196) of the stored
                       // output z = ReLU(acc + bias). Every CTA writes ITS per-channel sum [cout] and sum of squares [cout]
                       // to stats[blockIdx.x][2*cout]; bn_finalize_kernel adds the CTAs' rows in index order. No atomics
                       // anywhere on the way (each epilogue warp owns a shared-memory row, rows are added in warp order), so
                       // the statistics — and everything downstream of them — are bit-reproducible run to run. The BN
                       // affine is then the identity and bn_scale / bn_shift are not read
  // Narrow 3x3 layers of a loaded context (cout == BLOCK_N <= 64, plain ConvBN epilogue): the affine vectors BY VALUE
  // (the fused head's weights stay in shared memory: as constant operands they measured 4 % slower on conv2d_14). Kernel parameters sit in the constant bank, so the epilogue's FADD / FFMA take them
  // as immediate-indexed constant operands: no shared-memory broadcast reads on the L1 data pipe these layers are
  // bound by (it is shared with the tensor core's operand reads).
  int const_affine;
  alignas(16) float c_bias[128];    // transposed conv with 64 output channels: the vectors twice (indexed by column & 127)
  alignas(16) float c_scale[128];
  alignas(16) float c_shift[128];
  alignas(16) float c_head[196];    // kEpiHeadConst: head weights [3][64] with conv2d_14's BN scale folded in, then the 3 folded biases
  // fp16 storage only: incremented once per (thread, 32-channel chunk) in which a value was clamped to +-65504 by the
  // saturating pack, so a checkpoint whose activations leave fp16's range cannot clip silently (itmo_saturation_count)
  unsigned long long* sat_count;
  const float* bias;      // [cout]
  const float* bn_scale;  // [cout]  gamma / sqrt(var + eps)
  const float* bn_shift;  // [cout]  beta - mean * scale
  void* out;              // NHWC activations; may be null when the head is fused
  void* out_lo;           // kStoreSplit: residual tensor (or null)
  void* pooled;           // NHWC 2x2-max-pooled copy or null
  void* pooled_lo;
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
struct ElemTraits<__half> {
  static constexpr bool kTF32 = false;
  static constexpr uint32_t kFmt = sm100::kFmtF16;
};
template <>
struct ElemTraits<float> {
  static constexpr bool kTF32 = true;
  static constexpr uint32_t kFmt = sm100::kFmtTF32;
};

// Pack two floats into a 16-bit pair (round-to-nearest-even), low half = a.
template <typename Elem>
__device__ __forceinline__ uint32_t pack2(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  // saturate instead of overflowing to inf (fp16 tops out at 65504): one F2FP.SATFINITE, no separate clamps (the
  // fminf / fmaxf pairs cost the HBM-bound layers 15-20 % in the fp16 mode)
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
// max over |halves| of two packed fp16 pairs (HMNMX2.XORSIGN with |.| operands): magnitude bits only are meaningful
__device__ __forceinline__ uint32_t absmax2_f16(uint32_t m, uint32_t v) {
  uint32_t r;
  asm("max.xorsign.abs.f16x2 %0, %1, %2;" : "=r"(r) : "r"(m), "r"(v));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) { return pack2<__nv_bfloat16>(a, b); }

template <typename Elem>
__device__ __forceinline__ uint32_t max2(uint32_t a, uint32_t b);
template <>
__device__ __forceinline__ uint32_t max2<__nv_bfloat16>(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
template <>
__device__ __forceinline__ uint32_t max2<__half>(uint32_t a, uint32_t b) {
  __half2 r = __hmax2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

__device__ __forceinline__ float tf32_rna(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}
__device__ __forceinline__ float tf32_trunc(float v) { return __uint_as_float(__float_as_uint(v) & 0xFFFFE000u); }

// fp32 vector store of 4 consecutive channels under a StoreMode
__device__ __forceinline__ void store4_f32(float* dst, float* dst_lo, int mode, float a, float b, float c, float d) {
  if (mode == kStoreTf32Rna) {
    *reinterpret_cast<float4*>(dst) = make_float4(tf32_rna(a), tf32_rna(b), tf32_rna(c), tf32_rna(d));
  } else {
    *reinterpret_cast<float4*>(dst) = make_float4(a, b, c, d);
    if (mode == kStoreSplit && dst_lo != nullptr)
      *reinterpret_cast<float4*>(dst_lo) =
          make_float4(a - tf32_trunc(a), b - tf32_trunc(b), c - tf32_trunc(c), d - tf32_trunc(d));
  }
}

// Sum over the 32 lanes of a warp of 32 per-lane values at once: after five exchange rounds lane j holds
// sum over lanes of v[j] (31 shuffles instead of 32 x 5). v is clobbered.
__device__ __forceinline__ float warp_transpose_sum32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = upper ? v[i] : v[i + off];
      const float keep = upper ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, off);
    }
  }
  return v[0];
}

// 4x4 transposition of 16-byte pieces inside each quad of lanes: on entry lane j (= lane & 3) holds pieces
// t[4k..4k+3], k = 0..3, of ITS row; on exit t[4k..4k+3] is piece j of the row lane (quad base + k) came with.
__device__ __forceinline__ void quad_transpose_16B(uint32_t (&t)[16], int lane) {
#pragma unroll
  for (int bit = 1; bit <= 2; bit <<= 1) {
    const bool upper = (lane & bit) != 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (k & bit) continue;               // pairs (k, k + bit)
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const uint32_t send = upper ? t[4 * k + w] : t[4 * (k + bit) + w];
        const uint32_t got = __shfl_xor_sync(0xFFFFFFFFu, send, bit);
        if (upper) t[4 * k + w] = got; else t[4 * (k + bit) + w] = got;
      }
    }
  }
}

// Packed fp32 pairs (FADD2 / FFMA2 on sm_100: two lanes of fp32 arithmetic per issued instruction, same rounding as the
// scalar forms). The specialised epilogues are bound by their instruction issue, so the affine and the head use them.
__device__ __forceinline__ uint64_t f2_pack(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
// (r0, r1) = max(acc + bias, 0) for two adjacent channels
__device__ __forceinline__ void relu_bias2(uint32_t v0, uint32_t v1, uint64_t b01, float& r0, float& r1) {
  float u0, u1;
  f2_unpack(f2_add(f2_pack(__uint_as_float(v0), __uint_as_float(v1)), b01), u0, u1);
  r0 = fmaxf(u0, 0.f);
  r1 = fmaxf(u1, 0.f);
}
__device__ __forceinline__ void relu_bias2(uint32_t v0, uint32_t v1, float b0, float b1, float& r0, float& r1) {
  relu_bias2(v0, v1, f2_pack(b0, b1), r0, r1);
}
// adjacent pair of a by-value parameter vector as one 64-bit constant-bank operand
__device__ __forceinline__ uint64_t cpair(const float* v, int i) { return *reinterpret_cast<const uint64_t*>(v + i); }

// y = ReLU(acc + bias) * scale + shift on one 32-column accumulator chunk (kStats: training forward, identity
// affine; kLinear: no ReLU, the data-gradient convolutions). The vectors are read as 128-bit broadcasts.
template <bool kStats, bool kLinear>
__device__ __forceinline__ void epi_affine(const uint32_t (&v)[32], float (&f)[32], const float4 (&b4)[8],
                                           const float* __restrict__ ss, const float* __restrict__ st) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float a0 = __uint_as_float(v[4 * i]) + b4[i].x, a1 = __uint_as_float(v[4 * i + 1]) + b4[i].y;
    float a2 = __uint_as_float(v[4 * i + 2]) + b4[i].z, a3 = __uint_as_float(v[4 * i + 3]) + b4[i].w;
    if constexpr (!kLinear) {
      a0 = fmaxf(a0, 0.f);
      a1 = fmaxf(a1, 0.f);
      a2 = fmaxf(a2, 0.f);
      a3 = fmaxf(a3, 0.f);
    }
    f[4 * i] = a0;
    f[4 * i + 1] = a1;
    f[4 * i + 2] = a2;
    f[4 * i + 3] = a3;
  }
  if constexpr (!kStats) {
    float4 sc[8], sh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sc[i] = reinterpret_cast<const float4*>(ss)[i];
      sh[i] = reinterpret_cast<const float4*>(st)[i];
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      f[4 * i] = fmaf(f[4 * i], sc[i].x, sh[i].x);
      f[4 * i + 1] = fmaf(f[4 * i + 1], sc[i].y, sh[i].y);
      f[4 * i + 2] = fmaf(f[4 * i + 2], sc[i].z, sh[i].z);
      f[4 * i + 3] = fmaf(f[4 * i + 3], sc[i].w, sh[i].w);
    }
  }
}

// The same with the vectors read from the kernel parameters at compile-time offsets (constant-bank operands).
template <int CO0>
__device__ __forceinline__ void epi_affine_const(const uint32_t (&v)[32], float (&f)[32], const ConvParams& p) {
#pragma unroll
  for (int i = 0; i < 32; ++i)
    f[i] = fmaf(fmaxf(__uint_as_float(v[i]) + p.c_bias[CO0 + i], 0.f), p.c_scale[CO0 + i], p.c_shift[CO0 + i]);
}
template <typename Elem, int BLOCK_N, int SWZ>
struct ConvCfg {
  static constexpr int kBK = SWZ / int(sizeof(Elem));     // channels per K block
  static constexpr int kKSteps = SWZ / 32;                // one UMMA consumes 32 bytes of K
  static constexpr int kBBytes = BLOCK_N * SWZ;
  static constexpr int kChunk = 32;                       // accumulator columns per tcgen05.ld
};

// The NM * KSTEPS MMAs of one (channel block, tap). Their shared-memory descriptors differ from the tap's first one
// only by compile-time constants in the start-address field (units of 16 bytes), so each MMA costs the issuing
// thread two integer adds: at N <= 64 the tensor pipe retires an MMA every 40-48 clocks (tools/ubench/umma_rate.cu)
// and a longer issue sequence per MMA becomes the bottleneck of the whole kernel. K step outer, M block inner.
template <bool TF32, int NM, int KSTEPS, int SWZ, int BLOCK_N, bool PAIR = false>
__device__ __forceinline__ void issue_tap(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                          uint32_t idesc, uint32_t have) {
#pragma unroll
  for (int k = 0; k < KSTEPS; ++k) {
    const uint64_t db = (uint64_t(b_hi) << 32) | uint64_t(b_lo + uint32_t((k * 32) >> 4));
#pragma unroll
    for (int j = 0; j < NM; ++j) {
      const uint64_t da = (uint64_t(a_hi) << 32) | uint64_t(a_lo + uint32_t((j * kBlkW * SWZ + k * 32) >> 4));
      if constexpr (PAIR) sm100::umma_ss_pair(tmem_d + uint32_t(j * BLOCK_N), da, db, idesc, k == 0 ? have : 1u);
      else sm100::umma_ss<TF32>(tmem_d + uint32_t(j * BLOCK_N), da, db, idesc, k == 0 ? have : 1u);
    }
  }
}

// All taps of one channel block, fully unrolled for the tile shape the layer planner picks for this BLOCK_N
// (NM M blocks, 3x3 or single tap): every descriptor is "stage base + compile-time constant". The issuing thread's
// instructions between two MMAs are NOT hidden behind the tensor pipe beyond a few queued MMAs (measured: a 25-instruction
// tap prologue cost ~40 clocks per MMA at N = 64), so the stream has to be this dense.
template <bool TF32, int NM, int KSTEPS, int SWZ, int BLOCK_N, bool CONV, bool PAIR = false>
__device__ __forceinline__ void issue_kb(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_hi, uint32_t idesc,
                                         uint32_t have_first, bool resident, uint32_t b_res_lo, uint32_t b_lo0,
                                         uint64_t* b_full, uint64_t* b_empty, int& sb, uint32_t& pb, int n_b) {
  constexpr int kT = CONV ? 3 : 1;
  constexpr int kHaloW = kBlkW * NM + (CONV ? 2 : 0);
  constexpr uint32_t kB16 = uint32_t(BLOCK_N * SWZ / (PAIR ? 2 : 1)) >> 4;   // a CTA of a pair holds half of the rows
#pragma unroll
  for (int ty = 0; ty < kT; ++ty) {
#pragma unroll
    for (int tx = 0; tx < kT; ++tx) {
      const int tap = ty * kT + tx;
      uint32_t b_lo;
      if (resident) {
        b_lo = b_res_lo + uint32_t(tap) * kB16;
      } else {
        sm100::mbar_wait(&b_full[sb], pb);
        sm100::tc_fence_after_sync();
        b_lo = b_lo0 + uint32_t(sb) * kB16;
      }
      issue_tap<TF32, NM, KSTEPS, SWZ, BLOCK_N, PAIR>(tmem_d, a_lo + (uint32_t((ty * kHaloW + tx) * SWZ) >> 4), a_hi, b_lo, b_hi,
                                                       idesc, tap == 0 ? have_first : 1u);
      if (!resident) {
        if constexpr (PAIR) sm100::umma_commit_pair(&b_empty[sb]);
        else sm100::umma_commit(&b_empty[sb]);  // frees this weight stage once the MMAs retire
        if (++sb == n_b) {
          sb = 0;
          pb ^= 1;
        }
      }
    }
  }
}

// M blocks per CTA tile the planner uses for a given N tile (run_conv in itmo_api.cu); other values (narrow test
// images, the fp32 mode's split accumulators) take the generic issue loop.
__host__ __device__ constexpr int natural_nm(int block_n) { return block_n >= 256 ? 1 : block_n == 128 ? 2 : 4; }


// ------------------------------------------------------------------------------------------------------------------
// Specialised epilogues of the inference forward pass (EpiKind != kEpiGeneric), 16-bit element types, tile shape
// nm == natural_nm(BLOCK_N), one accumulator per M block. Everything the generic loop decides per chunk at run time
// is a template parameter here, the (M block, channel group, half) loops are fully unrolled, so column offsets — and
// with kEpiConstAffine the bias / scale / shift operands — are compile-time constants.
//   work item  = one M block (16 x 8 pixels) x one channel group (64 channels; 32 when BLOCK_N == 32); a warp handles
//                its TMEM lane quarter = 4 rows x 8 pixels of the block. The two warp sets split the items.
//   per item   : both tcgen05.ld of the group in flight together -> y = max(acc + b, 0) * s + t -> pack -> swizzled
//                staging tile (row = pixel) [-> 2x2 max-pool of the 4 x 8 pixels out of the staging tile] -> one lane
//                issues the cp.async.bulk.tensor store(s). With two staging tiles per warp the next item is staged
//                while the TMA engine still reads the previous one.
//   kEpiHead   : no activation store; the 64 -> 3 head (model.py:276), sigmoid and predict.py:64's truncation on the
//                fp32 post-BN registers.
template <typename Elem, int BLOCK_N, int EPI>
__device__ __forceinline__ void epilogue_fast(const ConvParams& p, uint32_t tmem_base, uint64_t* tmem_full_bar,
                                              uint64_t* tmem_empty_bar, uint8_t* epi_stage, const float* s_bias,
                                              const float* s_scale, const float* s_shift, const float* s_head,
                                              int warp_idx, int lane, float* s_stat = nullptr) {
  constexpr int kKind = EPI & 15;
  constexpr bool kConst = (EPI & kEpiConstAffine) != 0;
  constexpr bool kPool = kKind == kEpiConvPool;
  constexpr bool kHead = kKind == kEpiHead;
  constexpr bool kConvT = kKind == kEpiConvT;
  constexpr bool kStats = kKind == kEpiStats;        // runtime p.mode decides 3x3 / transposed addressing
  constexpr bool kIdent = kKind == kEpiIdentity;
  constexpr bool kF16 = ElemTraits<Elem>::kFmt == sm100::kFmtF16;
  const int nm = p.nm;                              // M blocks per tile: natural_nm(BLOCK_N) or half of it
  constexpr int kRowB = BLOCK_N >= 64 ? 128 : 64;   // bytes of one staged pixel row
  constexpr int kGroupCh = kRowB / 2;               // channels per work item
  constexpr int kGroups = BLOCK_N / kGroupCh;
  constexpr int kHalves = kGroupCh / 32;            // tcgen05.ld x32 per item
  constexpr int kStageB = 40 * kRowB;               // 32 pixel rows + 8 pooled rows
  constexpr bool kRows = (EPI & kEpiRows) != 0;
  const int kTileW = kRows ? 128 : kBlkW * nm;
  static_assert(!kRows || kKind == kEpiConv || kKind == kEpiHead, "row tiles: plain ConvBN and the fused head");
  static_assert(!kConst || kConvT || BLOCK_N <= 64, "constant-bank affine: cout <= 64");
  static_assert(!kHead || (BLOCK_N == 64 && kConst), "fused head: conv2d_14 only");
  static_assert(sizeof(Elem) == 2, "16-bit element types");
  static_assert(!kStats || ElemTraits<Elem>::kFmt == sm100::kFmtBF16, "training statistics: bf16 storage");

  const int ew = warp_idx - 4;            // 0..7
  const int q = warp_idx & 3;             // TMEM lane quarter this warp may read
  const int ws = ew >> 2;                 // warp set 0 / 1
  const int r = 4 * q + (lane >> 3);      // pixel row inside the 16x8 M block
  const int c = lane & 7;                 // pixel column inside the M block
  constexpr bool kPair = (EPI & kEpiPair) != 0;
  const uint32_t cta_rank = kPair ? sm100::cluster_ctarank() : 0u;
  const int tiles_x = p.W / kTileW;
  const int tiles_y = kRows ? p.H / nm : p.H / kBlkH;
  const int m_tiles = p.NB * tiles_y * tiles_x;
  const int m_units = kPair ? (m_tiles >> 1) : m_tiles;
  const int cta_unit = kPair ? int(blockIdx.x >> 1) : int(blockIdx.x);
  const int cta_units = kPair ? int(gridDim.x >> 1) : int(gridDim.x);
  // work units of this CTA (pair): (N tile, M unit) interleaved, or — n_stationary — M units of one fixed N tile
  const bool stat = p.n_stationary != 0;
  const int total_tiles = stat ? m_units : m_units * p.n_tiles;
  const int unit0 = stat ? cta_unit / p.n_tiles : cta_unit;
  const int unit_stride = stat ? cta_units / p.n_tiles : cta_units;
  const int n_fixed = cta_unit % p.n_tiles;
  // pair: the accumulators of both CTAs are handed back through the leader's barrier
  const uint32_t empty_remote[2] = {kPair ? sm100::mapa_u32(&tmem_empty_bar[0], 0) : 0u,
                                    kPair ? sm100::mapa_u32(&tmem_empty_bar[1], 0) : 0u};
  const uint32_t buf_cols = uint32_t(nm * BLOCK_N);
  uint8_t* const stage0 = epi_stage + ew * p.epi_bufs * kStageB;
  const int sw_out = kRowB == 128 ? (lane & 7) : ((lane >> 1) & 3);
  const bool count_sat = kF16 && p.sat_count != nullptr;
  int buf = 0;
  int acc = 0;
  uint32_t acc_phase = 0;
  for (int unit = unit0; unit < total_tiles; unit += unit_stride) {
    const int n_tile = stat ? n_fixed : unit % p.n_tiles;
    int m = stat ? unit : unit / p.n_tiles;
    if constexpr (kPair) m = 2 * m + int(cta_rank);
    const int tx = m % tiles_x;
    m /= tiles_x;
    const int ty = m % tiles_y;
    const int img = m / tiles_y;

    sm100::mbar_wait_relaxed(&tmem_full_bar[acc], acc_phase, 200);
    sm100::tc_fence_after_sync();
    const uint32_t taddr = tmem_base + (uint32_t(32 * q) << 16) + uint32_t(acc) * buf_cols;
    if (!(p.dbg & 32)) {   // 32: timing experiment without the epilogue work
    // M blocks of this warp set; NOT unrolled (j only moves the TMEM column and the x coordinate): the item body is
    // 150-500 instructions and the kernel already fills most of the instruction cache
#pragma unroll 1
    for (int j = (nm > 1 ? ws : 0); j < nm; j += (nm > 1 ? 2 : 1)) {
      if (nm == 1 && kGroups == 1 && ws != 0) break;   // a lone block with a single group stays with set 0
      uint64_t hh0 = 0ull, hh1 = 0ull, hh2 = 0ull;   // head sums, (even channels, odd channels) as packed fp32 pairs
#pragma unroll
      for (int g = 0; g < kGroups; ++g) {
        if (kGroups > 1 && nm == 1 && (g & 1) != ws) continue;
        uint32_t v[kHalves][32];
#pragma unroll
        for (int h = 0; h < kHalves; ++h)
          sm100::tmem_ld_32x32b_x32(taddr + uint32_t(j * BLOCK_N + g * kGroupCh + h * 32), v[h]);
        uint8_t* const st_out = stage0 + buf * kStageB;
        if constexpr (!kHead) {
          // the bulk store that last read this staging tile must be done with it
          if (lane == 0) {
            if (p.epi_bufs == 2) sm100::tma_store_wait_read<1>();
            else sm100::tma_store_wait_read<0>();
          }
          __syncwarp();
        }
        sm100::tmem_ld_wait();
        uint32_t m2 = 0;
#pragma unroll
        for (int h = 0; h < kHalves; ++h) {
          const int col = g * kGroupCh + h * 32;                      // compile-time after unrolling
          float f[32];
          if constexpr (kIdent) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[h][i]);
          } else if constexpr (kStats) {
            const int n_glob = n_tile * BLOCK_N + col;
            const int co0 = p.mode == kModeConvT2x2 ? (n_glob & (p.cout - 1)) : n_glob;
            const float4* bb = reinterpret_cast<const float4*>(s_bias + co0);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 b4 = bb[i];
              relu_bias2(v[h][4 * i], v[h][4 * i + 1], b4.x, b4.y, f[4 * i], f[4 * i + 1]);
              relu_bias2(v[h][4 * i + 2], v[h][4 * i + 3], b4.z, b4.w, f[4 * i + 2], f[4 * i + 3]);
            }
          } else if constexpr (kHead) {
            // the BN affine is folded into the head's weights (see the kernel's setup): only max(acc + b, 0) here
#pragma unroll
            for (int i = 0; i < 32; i += 2) relu_bias2(v[h][i], v[h][i + 1], cpair(p.c_bias, col + i), f[i], f[i + 1]);
          } else if constexpr (kConst) {
            const int cc = kConvT ? (col & 127) : col;                // transposed conv: cout == 128, or 64 stored twice
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float r0, r1;
              relu_bias2(v[h][i], v[h][i + 1], cpair(p.c_bias, cc + i), r0, r1);
              f2_unpack(f2_fma(f2_pack(r0, r1), cpair(p.c_scale, cc + i), cpair(p.c_shift, cc + i)), f[i], f[i + 1]);
            }
          } else {
            const int co0 = kConvT ? ((n_tile * BLOCK_N + col) & (p.cout - 1)) : (n_tile * BLOCK_N + col);
            const float4* bb = reinterpret_cast<const float4*>(s_bias + co0);
            const float4* sc = reinterpret_cast<const float4*>(s_scale + co0);
            const float4* sh = reinterpret_cast<const float4*>(s_shift + co0);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 b4 = bb[i], s4 = sc[i], t4 = sh[i];
              float r0, r1, r2, r3;
              relu_bias2(v[h][4 * i], v[h][4 * i + 1], b4.x, b4.y, r0, r1);
              relu_bias2(v[h][4 * i + 2], v[h][4 * i + 3], b4.z, b4.w, r2, r3);
              f2_unpack(f2_fma(f2_pack(r0, r1), f2_pack(s4.x, s4.y), f2_pack(t4.x, t4.y)), f[4 * i], f[4 * i + 1]);
              f2_unpack(f2_fma(f2_pack(r2, r3), f2_pack(s4.z, s4.w), f2_pack(t4.z, t4.w)), f[4 * i + 2], f[4 * i + 3]);
            }
          }
          if constexpr (kHead) {
            // even / odd channels accumulate in the two halves of a packed pair: 3 FFMA2 per channel pair
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              if constexpr ((EPI & kEpiHeadConst) != 0) {
                const uint64_t fa = f2_pack(f[i], f[i + 1]), fb = f2_pack(f[i + 2], f[i + 3]);
                hh0 = f2_fma(fb, cpair(p.c_head, 0 * 64 + col + i + 2), f2_fma(fa, cpair(p.c_head, 0 * 64 + col + i), hh0));
                hh1 = f2_fma(fb, cpair(p.c_head, 1 * 64 + col + i + 2), f2_fma(fa, cpair(p.c_head, 1 * 64 + col + i), hh1));
                hh2 = f2_fma(fb, cpair(p.c_head, 2 * 64 + col + i + 2), f2_fma(fa, cpair(p.c_head, 2 * 64 + col + i), hh2));
                continue;
              }
              const ulonglong2 w0 = *reinterpret_cast<const ulonglong2*>(s_head + 0 * 64 + col + i);
              const ulonglong2 w1 = *reinterpret_cast<const ulonglong2*>(s_head + 1 * 64 + col + i);
              const ulonglong2 w2 = *reinterpret_cast<const ulonglong2*>(s_head + 2 * 64 + col + i);
              const uint64_t fa = f2_pack(f[i], f[i + 1]), fb = f2_pack(f[i + 2], f[i + 3]);
              hh0 = f2_fma(fb, w0.y, f2_fma(fa, w0.x, hh0));
              hh1 = f2_fma(fb, w1.y, f2_fma(fa, w1.x, hh1));
              hh2 = f2_fma(fb, w2.y, f2_fma(fa, w2.x, hh2));
            }
          } else {
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) pk[i] = pack2<Elem>(f[2 * i], f[2 * i + 1]);
            if constexpr (kStats) {
              // statistics of the ROUNDED values (what bn_apply and the backward read back from HBM); each epilogue warp
              // owns a shared-memory row, plain adds in program order: bit-reproducible
              float sv[32], sq[32];
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                sv[2 * i] = __uint_as_float(pk[i] << 16);
                sv[2 * i + 1] = __uint_as_float(pk[i] & 0xFFFF0000u);
                sq[2 * i] = sv[2 * i] * sv[2 * i];
                sq[2 * i + 1] = sv[2 * i + 1] * sv[2 * i + 1];
              }
              const float tsum = warp_transpose_sum32(sv, lane);
              const float tsq = warp_transpose_sum32(sq, lane);
              const int n_glob = n_tile * BLOCK_N + col;
              const int co0 = p.mode == kModeConvT2x2 ? (n_glob & (p.cout - 1)) : n_glob;
              float* row = s_stat + ew * 2 * p.cout;
              row[co0 + lane] += tsum;
              row[p.cout + co0 + lane] += tsq;
            }
            if constexpr (kF16) {
              // |max| of the packed pairs, two words per instruction (three-input VHMNMX with |.| operands)
#pragma unroll
              for (int i = 0; i < 16; i += 2) {
                const __half2 a = __habs2(*reinterpret_cast<const __half2*>(&pk[i]));
                const __half2 b = __habs2(*reinterpret_cast<const __half2*>(&pk[i + 1]));
                const __half2 mm = __hmax2(__hmax2(*reinterpret_cast<const __half2*>(&m2), a), b);
                m2 = *reinterpret_cast<const uint32_t*>(&mm);
              }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
              *reinterpret_cast<uint4*>(st_out + lane * kRowB + (((h * 4 + k) ^ sw_out) << 4)) =
                  make_uint4(pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
          }
        }
        if constexpr (!kHead) {
          if constexpr (kF16) {
            // a value the saturating pack clamped reads back as +-65504 (0x7BFF)
            m2 &= 0x7FFF7FFFu;
            if (count_sat && ((m2 & 0xFFFFu) >= 0x7BFFu || (m2 >> 16) >= 0x7BFFu)) atomicAdd(p.sat_count, 1ull);
          }
          if constexpr (kPool) {
            // 2x2 max-pool of the warp's 4 x 8 pixels out of the staging tile: lane = (pooled pixel, 16-byte piece)
            __syncwarp();
            uint8_t* const st_pool = st_out + 32 * kRowB;
            const int pp = lane >> 2, k = lane & 3;
            const int p00 = (2 * (pp >> 2)) * 8 + 2 * (pp & 3);
            const int sw_p = kRowB == 128 ? (pp & 7) : ((pp >> 1) & 3);
#pragma unroll
            for (int h = 0; h < kHalves; ++h) {
              auto piece = [&](int px) {
                const int sw = kRowB == 128 ? (px & 7) : ((px >> 1) & 3);
                return *reinterpret_cast<const uint4*>(st_out + px * kRowB + (((h * 4 + k) ^ sw) << 4));
              };
              const uint4 a4 = piece(p00), b4v = piece(p00 + 1), c4 = piece(p00 + 8), d4 = piece(p00 + 9);
              uint4 m4;
              m4.x = max2<Elem>(max2<Elem>(a4.x, b4v.x), max2<Elem>(c4.x, d4.x));
              m4.y = max2<Elem>(max2<Elem>(a4.y, b4v.y), max2<Elem>(c4.y, d4.y));
              m4.z = max2<Elem>(max2<Elem>(a4.z, b4v.z), max2<Elem>(c4.z, d4.z));
              m4.w = max2<Elem>(max2<Elem>(a4.w, b4v.w), max2<Elem>(c4.w, d4.w));
              *reinterpret_cast<uint4*>(st_pool + pp * kRowB + (((h * 4 + k) ^ sw_p) << 4)) = m4;
            }
          }
          sm100::fence_proxy_async_smem();   // generic-proxy writes -> visible to the TMA (async proxy)
          __syncwarp();
          if (lane == 0) {
            const int ng = n_tile * BLOCK_N + g * kGroupCh;          // first column of the staged group
            const int x0 = kRows ? tx * kTileW + 32 * q : tx * kTileW + j * kBlkW;
            const int y0 = kRows ? ty * nm + (nm - 1 - j) : ty * kBlkH + 4 * q;
            if (kConvT || (kStats && p.mode == kModeConvT2x2)) {
              const int a_row = ng >> (p.cout_log2 + 1);               // kernel row a of this column group
              sm100::tma_store_4d(&p.tmOut[a_row], st_out, ng - (a_row << (p.cout_log2 + 1)), x0, y0, img);
            } else {
              sm100::tma_store_4d(&p.tmOut[0], st_out, ng, x0, y0, img);
            }
            if constexpr (kPool) sm100::tma_store_4d(&p.tmPool, st_out + 32 * kRowB, ng, x0 >> 1, y0 >> 1, img);
            sm100::tma_store_commit();
          }
          buf ^= p.epi_bufs - 1;
        }
      }
      if constexpr (kHead) {
        // Conv2D(3, 1, activation='sigmoid') (model.py:276) on the post-BN fp32 values
        const int y = kRows ? ty * nm + (nm - 1 - j) : ty * kBlkH + r;
        const int x = kRows ? tx * kTileW + 32 * q + lane : tx * kTileW + j * kBlkW + c;
        float e0, o0, e1, o1, e2, o2;
        f2_unpack(hh0, e0, o0);
        f2_unpack(hh1, e1, o1);
        f2_unpack(hh2, e2, o2);
        constexpr bool kHC = (EPI & kEpiHeadConst) != 0;
        const float z0 = (e0 + o0) + (kHC ? p.c_head[192] : s_head[192]), z1 = (e1 + o1) + (kHC ? p.c_head[193] : s_head[193]),
                    z2 = (e2 + o2) + (kHC ? p.c_head[194] : s_head[194]);
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
    }
    }
    // accumulators drained: hand the TMEM buffer back to the MMA warp
    sm100::tc_fence_before_sync();
    __syncwarp();
    if (lane == 0) {
      if constexpr (kPair) sm100::mbar_arrive_cluster(empty_remote[acc]);
      else sm100::mbar_arrive(&tmem_empty_bar[acc]);
    }
    acc ^= 1;
    if (acc == 0) acc_phase ^= 1;
  }
  if constexpr (!kHead) {
    if (lane == 0) sm100::tma_store_wait_all<0>();   // staging tiles stay valid until the TMA has read them
  }
}

template <typename Elem, int BLOCK_N, int SWZ, int EPI = kEpiGeneric>
__global__ void __launch_bounds__(kNumThreads, 1) conv_igemm_kernel(const __grid_constant__ ConvParams p) {
  using Cfg = ConvCfg<Elem, BLOCK_N, SWZ>;
  constexpr bool kTF32 = ElemTraits<Elem>::kTF32;
  static_assert(BLOCK_N % 32 == 0 && BLOCK_N <= 256, "BLOCK_N: multiple of 32 up to 256");
  constexpr bool kPair = (EPI & kEpiPair) != 0;                 // CTA pairs, see EpiKind
  static_assert(!kPair || (EPI & 15) != kEpiGeneric, "CTA pairs: specialised epilogues only");
  constexpr int kBStage = Cfg::kBBytes / (kPair ? 2 : 1);      // weight tile bytes in THIS CTA's shared memory
  const uint32_t cta_rank = kPair ? sm100::cluster_ctarank() : 0u;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = sm100::smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

  const int nm = p.nm;
  const int n_a = p.n_a_stages, n_b = p.n_b_stages;
  uint8_t* smem_a = smem;                                   // n_a x halo tile
  uint8_t* smem_b = smem + n_a * p.a_stage_bytes;           // n_b x [BLOCK_N rows x SWZ bytes]
  uint8_t* epi_stage = smem_b + n_b * kBStage;              // TMA-store staging: [kNumEpiWarps][40 rows][kEpiRowB], 1 KB aligned
  uint8_t* aux = epi_stage + p.epi_bytes;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(aux);      // [kMaxAStages]
  uint64_t* a_empty = a_full + kMaxAStages;
  uint64_t* b_full = a_empty + kMaxAStages;                 // [kMaxBStages]  (resident: b_full[0] only)
  uint64_t* b_empty = b_full + kMaxBStages;
  uint64_t* tmem_full_bar = b_empty + kMaxBStages;          // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;             // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  float* s_bias = reinterpret_cast<float*>(aux + 512);
  float* s_scale = s_bias + kMaxCout;
  float* s_shift = s_scale + kMaxCout;
  float* s_head = s_shift + kMaxCout;                       // [3*64 weights][3 bias]
  float* s_stat = reinterpret_cast<float*>(aux + kAuxBytes); // training forward: [kNumEpiWarps][2][cout] (see ConvParams::stats)

  // warp index made provably warp-uniform: with a uniform role branch and elect.sync the compiler issues
  // UTCHMMA / UBLKCP straight from uniform registers instead of a per-thread ELECT/BRA loop around every one
  const int warp_idx = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  const int tile_w = kBlkW * nm;
  const int tiles_x = p.W / tile_w;
  const int tiles_y = p.H / kBlkH;
  const int m_tiles = p.NB * tiles_y * tiles_x;
  // work units: one tile per CTA, or two horizontally adjacent tiles (same N tile) per CTA pair
  const int m_units = kPair ? (m_tiles >> 1) : m_tiles;
  const int cta_unit = kPair ? int(blockIdx.x >> 1) : int(blockIdx.x);
  const int cta_units = kPair ? int(gridDim.x >> 1) : int(gridDim.x);
  const bool stat = p.n_stationary != 0;        // see ConvParams::n_stationary
  const int total_tiles = stat ? m_units : m_units * p.n_tiles;
  const int unit0 = stat ? cta_unit / p.n_tiles : cta_unit;
  const int unit_stride = stat ? cta_units / p.n_tiles : cta_units;
  const int n_fixed = cta_unit % p.n_tiles;
  int kb_total = 0;
  for (int s = 0; s < p.n_seg; ++s) kb_total += p.seg_kb[s];
  const int halo_w = tile_w + 2 * p.pad;                    // halo pitch in pixels
  const int halo_h = kBlkH + 2 * p.pad;
  const int S = p.acc_split;
  const uint32_t buf_cols = uint32_t(S * nm * BLOCK_N);    // TMEM columns of one accumulator buffer
  const uint32_t tmem_cols_used = 2 * buf_cols;
  uint32_t tmem_cols = 32;
  while (tmem_cols < tmem_cols_used) tmem_cols <<= 1;

  // ---------------------------------------------------------------- one-time setup
  if (warp_idx == 0 && lane == 0) {
    for (int s = 0; s < 4; ++s) sm100::tma_prefetch_desc(&p.tmA[s]);
    sm100::tma_prefetch_desc(&p.tmB);
    if (p.tma_store) {
      sm100::tma_prefetch_desc(&p.tmOut[0]);
      sm100::tma_prefetch_desc(&p.tmOut[1]);
      sm100::tma_prefetch_desc(&p.tmPool);
    }
  }
  if (warp_idx == 1 && lane == 0) {
    // CTA pair: the "full" barriers that count are the leader's — ONE arrival (the leader's producer, which also
    // announces the bytes of both CTAs' loads; the peer's loads only complete bytes on it: a remote arrive per weight
    // tile cost the peer's producer ~600 clocks each and halved the streamed layers' speed); the leader's "TMEM
    // empty" barriers collect the epilogue warps of both CTAs
    constexpr uint32_t kCtas = kPair ? 2 : 1;
    for (int i = 0; i < kMaxAStages; ++i) {
      sm100::mbar_init(&a_full[i], 1);
      sm100::mbar_init(&a_empty[i], 1);
    }
    for (int i = 0; i < kMaxBStages; ++i) {
      sm100::mbar_init(&b_full[i], 1);
      sm100::mbar_init(&b_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      sm100::mbar_init(&tmem_full_bar[i], 1);
      sm100::mbar_init(&tmem_empty_bar[i], kNumEpiWarps * kCtas);  // one arrive per epilogue warp
    }
    sm100::fence_barrier_init();
  }
  if (warp_idx == 2) {
    if constexpr (kPair) sm100::tmem_alloc_pair(tmem_ptr_smem, tmem_cols);
    else sm100::tmem_alloc(tmem_ptr_smem, tmem_cols);
  }
  const bool with_stats = p.stats != nullptr;
  for (int i = threadIdx.x; i < p.cout; i += blockDim.x) {
    s_bias[i] = p.bias[i];
    s_scale[i] = with_stats ? 1.f : p.bn_scale[i];
    s_shift[i] = with_stats ? 0.f : p.bn_shift[i];
  }
  if (with_stats)
    for (int i = threadIdx.x; i < kNumEpiWarps * 2 * p.cout; i += blockDim.x) s_stat[i] = 0.f;
  if (p.head_dtype != kHeadNone) {
    if constexpr ((EPI & 15) == kEpiHead) {
      // specialised head: z_k = sum_c w_kc (r_c s_c + t_c) + b_k = sum_c (w_kc s_c) r_c + (b_k + sum_c w_kc t_c),
      // r = max(acc + bias, 0): the BN affine of conv2d_14 moves into the head's weights (fp32, once per CTA)
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
    } else {
      for (int i = threadIdx.x; i < 3 * 64 + 3; i += blockDim.x)
        s_head[i] = (i < 3 * 64) ? p.head_w[i] : p.head_b[i - 3 * 64];
    }
  }
  sm100::tc_fence_before_sync();
  if constexpr (kPair) sm100::cluster_sync();   // the peer's barriers must be initialised before anything signals them
  else __syncthreads();
  sm100::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // ---------------------------------------------------------------- warp roles
  if (warp_idx == 0) {
    // ===== TMA producer (one elected thread) =====
    if (sm100::elect_one() && !(p.dbg & 512)) {  // 512: timing experiment without the producer
      const uint32_t a_bytes = uint32_t(SWZ) * halo_w * halo_h;
      constexpr uint32_t kCtas = kPair ? 2 : 1;
      // pair: only the leader arms its barrier, with the bytes of both CTAs' loads. The peer's bytes may land before
      // that (the transaction count goes negative meanwhile); the phase cannot complete before the leader's arrival,
      // and the peer cannot be a whole ring ahead because its "empty" barriers are signalled by the leader's MMAs.
      auto arm = [&](uint64_t* bar, uint32_t bytes_all) {
        if (!kPair || cta_rank == 0) sm100::mbar_arrive_expect_tx(bar, bytes_all);
      };
      auto load_b = [&](uint8_t* dst, uint64_t* bar, int k0, int row0, int tap) {
        if constexpr (kPair)
          sm100::tma_load_3d_pair(dst, &p.tmB, sm100::mapa_u32(bar, 0), k0, row0 + int(cta_rank) * (BLOCK_N / 2), tap,
                                  sm100::kEvictLast);
        else
          sm100::tma_load_3d(dst, &p.tmB, bar, k0, row0, tap, sm100::kEvictLast);
      };
      if (p.b_resident) {
        // all weight tiles of the layer — of this CTA's N tile when there are several (n_stationary) — once
        arm(&b_full[0], uint32_t(kb_total * p.taps) * Cfg::kBBytes);
        for (int kb = 0; kb < kb_total; ++kb)
          for (int tap = 0; tap < p.taps; ++tap)
            load_b(smem_b + (kb * p.taps + tap) * kBStage, &b_full[0], kb * Cfg::kBK, (stat ? n_fixed : 0) * BLOCK_N, tap);
      }
      int sa = 0, sb = 0, a_loads = 0;
      uint32_t pa = 0, pb = 0;
      for (int unit = unit0; unit < total_tiles; unit += unit_stride) {
        const int n_tile = stat ? n_fixed : unit % p.n_tiles;
        int m = stat ? unit : unit / p.n_tiles;
        if constexpr (kPair) m = 2 * m + int(cta_rank);
        const int tx = m % tiles_x;
        m /= tiles_x;
        const int ty = m % tiles_y;
        const int img = m / tiles_y;
        const int x0 = tx * tile_w - p.pad, y0 = ty * kBlkH - p.pad;
        int kb_glob = 0;
        for (int s = 0; s < p.n_seg; ++s) {
          const CUtensorMap* tmA = &p.tmA[p.seg_map[s]];
          for (int kb = 0; kb < p.seg_kb[s]; ++kb, ++kb_glob) {
            sm100::mbar_wait(&a_empty[sa], pa ^ 1);
            if ((p.dbg & 16) && a_loads >= n_a) {   // timing experiment: no activation traffic after the ring filled
              if (!kPair || cta_rank == 0) sm100::mbar_arrive(&a_full[sa]);
            } else {
              arm(&a_full[sa], a_bytes * kCtas);
              if constexpr (kPair)
                sm100::tma_load_4d_pair(smem_a + sa * p.a_stage_bytes, tmA, sm100::mapa_u32(&a_full[sa], 0), kb * Cfg::kBK, x0,
                                        y0, img, sm100::kEvictNormal);
              else
                sm100::tma_load_4d(smem_a + sa * p.a_stage_bytes, tmA, &a_full[sa], kb * Cfg::kBK, x0, y0, img,
                                   sm100::kEvictNormal);
            }
            ++a_loads;
            if (++sa == n_a) {
              sa = 0;
              pa ^= 1;
            }
            if (!p.b_resident) {
              for (int tap = 0; tap < p.taps; ++tap) {
                sm100::mbar_wait(&b_empty[sb], pb ^ 1);
                arm(&b_full[sb], Cfg::kBBytes);
                load_b(smem_b + sb * kBStage, &b_full[sb], kb_glob * Cfg::kBK, n_tile * BLOCK_N, tap);
                if (++sb == n_b) {
                  sb = 0;
                  pb ^= 1;
                }
              }
            }
          }
        }
      }
    }
  } else if (warp_idx == 1) {
    // ===== MMA issuer (one elected thread) =====
    // (CTA pair: the leader issues the M = 256 MMAs for both CTAs; the peer's tensor core executes its half)
    if (sm100::elect_one() && cta_rank == 0) {
      constexpr uint32_t idesc = sm100::make_idesc(ElemTraits<Elem>::kFmt, kPair ? 2 * kTileM : kTileM, BLOCK_N);
      auto commit = [&](uint64_t* bar) {
        if constexpr (kPair) sm100::umma_commit_pair(bar);   // arrives on this barrier in both CTAs
        else sm100::umma_commit(bar);
      };
      // descriptor high words are constant; the low words (start address >> 4) advance by whole pixels / rows
      const uint32_t a_hi = sm100::smem_desc_hi<SWZ>(uint32_t(halo_w) * SWZ);  // 8-pixel groups one halo row apart
      const uint32_t b_hi = sm100::smem_desc_hi<SWZ>(8u * SWZ);
      const uint32_t row16 = (uint32_t(halo_w) * SWZ) >> 4;                    // one halo row, in descriptor units
      const uint32_t b_lo0 = sm100::smem_desc_lo(sm100::smem_u32(smem_b));
      const int taps_y = p.taps == 9 ? 3 : 1, taps_x = taps_y;
      int sa = 0, sb = 0;
      uint32_t pa = 0, pb = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      // the unrolled issue path exists for the natural tile and, for the 64-column / 128-byte layers (conv2d_4: with
      // two M blocks instead of four its weights stay resident), for half of it
      constexpr bool kHalfNm = BLOCK_N == 64 && SWZ == 128;
      const bool fast = S == 1 && (nm == natural_nm(BLOCK_N) || (kHalfNm && nm == natural_nm(BLOCK_N) / 2)) &&
                        !(p.dbg & (64 | 1024));  // 1024: force the generic loop
      const bool nowait = (p.dbg & 128) != 0;   // timing experiment: the issuing thread never waits on a barrier
      if (p.b_resident && !nowait) {
        sm100::mbar_wait(&b_full[0], 0);
        sm100::tc_fence_after_sync();
      }
      for (int unit = unit0; unit < total_tiles; unit += unit_stride) {
        if (!nowait) sm100::mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
        sm100::tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + uint32_t(acc) * buf_cols;
        uint32_t used = 0;   // accumulators that already hold a partial sum in this tile
        int rr = 0;
        uint32_t b_res_lo = b_lo0;  // resident weights: tiles follow each other in (channel block, tap) order
        for (int kb = 0; kb < kb_total; ++kb) {
          if (!nowait) sm100::mbar_wait(&a_full[sa], pa);
          sm100::tc_fence_after_sync();
          uint32_t a_row_lo = sm100::smem_desc_lo(sm100::smem_u32(smem_a + sa * p.a_stage_bytes));
          if (fast) {
            constexpr int kNM = natural_nm(BLOCK_N);
            if (kHalfNm && nm != kNM)
              issue_kb<kTF32, (kHalfNm ? kNM / 2 : kNM), Cfg::kKSteps, SWZ, BLOCK_N, true, kPair>(
                  tmem_d, a_row_lo, a_hi, b_hi, idesc, kb != 0, p.b_resident != 0, b_res_lo, b_lo0, b_full, b_empty, sb, pb, n_b);
            else if (p.taps == 9)
              issue_kb<kTF32, kNM, Cfg::kKSteps, SWZ, BLOCK_N, true, kPair>(tmem_d, a_row_lo, a_hi, b_hi, idesc, kb != 0, p.b_resident != 0,
                                                                     b_res_lo, b_lo0, b_full, b_empty, sb, pb, n_b);
            else
              issue_kb<kTF32, kNM, Cfg::kKSteps, SWZ, BLOCK_N, false, kPair>(tmem_d, a_row_lo, a_hi, b_hi, idesc, kb != 0, p.b_resident != 0,
                                                                      b_res_lo, b_lo0, b_full, b_empty, sb, pb, n_b);
            b_res_lo += uint32_t(p.taps) * (kBStage >> 4);
          } else
          for (int ty_ = 0; ty_ < taps_y; ++ty_, a_row_lo += row16) {
            uint32_t a_tap_lo = a_row_lo;   // halo row / column of the tap's first pixel
            for (int tx_ = 0; tx_ < taps_x; ++tx_, a_tap_lo += SWZ >> 4) {
              uint32_t b_lo;
              if (p.b_resident) {
                b_lo = b_res_lo;
                b_res_lo += kBStage >> 4;
              } else {
                sm100::mbar_wait(&b_full[sb], pb);
                sm100::tc_fence_after_sync();
                b_lo = b_lo0 + uint32_t(sb) * (kBStage >> 4);
              }
              // accumulator of this tap: always 0 unless the fp32 mode splits the sum (round-robin, no division:
              // everything in this loop is on the single issuing thread's critical path)
              int sacc = 0;
              if (S > 1) {
                if (kb < p.kb_hi) {
                  sacc = rr;
                  if (++rr == S - 1) rr = 0;
                } else {
                  sacc = S - 1;
                }
              }
              const uint32_t have = (used >> sacc) & 1u;
              used |= 1u << sacc;
              const uint32_t d0 = tmem_d + uint32_t(sacc * nm) * BLOCK_N;
              if (!(p.dbg & 64)) {  // 64: timing experiment without the MMAs (TMA + epilogue only)
                if (nm == 4) issue_tap<kTF32, 4, Cfg::kKSteps, SWZ, BLOCK_N, kPair>(d0, a_tap_lo, a_hi, b_lo, b_hi, idesc, have);
                else if (nm == 2) issue_tap<kTF32, 2, Cfg::kKSteps, SWZ, BLOCK_N, kPair>(d0, a_tap_lo, a_hi, b_lo, b_hi, idesc, have);
                else issue_tap<kTF32, 1, Cfg::kKSteps, SWZ, BLOCK_N, kPair>(d0, a_tap_lo, a_hi, b_lo, b_hi, idesc, have);
              }
              if (!p.b_resident) {
                commit(&b_empty[sb]);  // frees this weight stage once the MMAs retire
                if (++sb == n_b) {
                  sb = 0;
                  pb ^= 1;
                }
              }
            }
          }
          commit(&a_empty[sa]);      // halo tile consumed by all taps
          if (++sa == n_a) {
            sa = 0;
            pa ^= 1;
          }
        }
        commit(&tmem_full_bar[acc]);  // accumulators complete -> epilogue
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp_idx >= 4 && !(p.dbg & 256)) {  // 256: timing experiment without the epilogue warps
    // ===== epilogue warps: TMEM -> registers -> bias/ReLU/BN (-> pool / head) -> global =====
    if constexpr (EPI != kEpiGeneric && sizeof(Elem) == 2) {
      epilogue_fast<Elem, BLOCK_N, EPI>(p, tmem_base, tmem_full_bar, tmem_empty_bar, epi_stage, s_bias, s_scale, s_shift,
                                        s_head, warp_idx, lane, s_stat);
    } else {
    const int ew = warp_idx - 4;            // 0..7
    const int q = warp_idx & 3;             // TMEM lane quarter this warp may read
    const int ws = ew >> 2;                 // warp set 0 / 1: the two sets split the tile's work items
    const int r = 4 * q + (lane >> 3);      // pixel row inside the 16x8 M block
    const int c = lane & 7;                 // pixel column inside the M block
    constexpr int kChunk = Cfg::kChunk;
    constexpr int kChunksPerBlk = BLOCK_N / kChunk;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = unit0; tile < total_tiles; tile += unit_stride) {
      const int n_tile = stat ? n_fixed : tile % p.n_tiles;
      int m = stat ? tile : tile / p.n_tiles;
      const int tx = m % tiles_x;
      m /= tiles_x;
      const int ty = m % tiles_y;
      const int img = m / tiles_y;
      const int y = ty * kBlkH + r;

      if (p.dbg == 4) sm100::mbar_wait(&tmem_full_bar[acc], acc_phase);
      else sm100::mbar_wait_relaxed(&tmem_full_bar[acc], acc_phase, 200);
      sm100::tc_fence_after_sync();
      const uint32_t taddr = tmem_base + (uint32_t(32 * q) << 16) + uint32_t(acc) * buf_cols;

      // the two warp sets split the tile: by M block when there are several, else by column chunk; a fused
      // head needs every channel of a pixel in one thread, so a lone block then stays with set 0
      const bool split_by_chunk = nm == 1 && kChunksPerBlk > 1 && p.head_dtype == kHeadNone;
      // staged rows hold 64 channels = two chunks: with TMA stores the sets split a lone block by chunk PAIR
      const int split_shift = (p.tma_store && BLOCK_N >= 64) ? 1 : 0;
      for (int j = 0; j < ((p.dbg & 32) ? 0 : nm); ++j) {  // dbg 32: timing experiment without the epilogue
        if (nm > 1 ? (j & 1) != ws : (!split_by_chunk && ws != 0)) continue;
        const int x = tx * tile_w + j * kBlkW + c;
        float h0 = 0.f, h1 = 0.f, h2 = 0.f;
#pragma unroll 1
        for (int chunk = 0; chunk < kChunksPerBlk; ++chunk) {
          if (split_by_chunk && ((chunk >> split_shift) & 1) != ws) continue;
          uint32_t v[kChunk];
          sm100::tmem_ld_32x32b_x32(taddr + uint32_t(j * BLOCK_N + chunk * kChunk), v);
          // output coordinates and the bias vector while the TMEM load is in flight
          const int n_glob = n_tile * BLOCK_N + chunk * kChunk;  // row of the packed weight matrix
          int co0, oy, ox, Ho, Wo;
          if (p.mode == kModeConvT2x2) {
            // phase a*2+b of the 2x2 transposed kernel (shift when cout is a power of two: no division chain)
            const int ph = p.cout_log2 >= 0 ? (n_glob >> p.cout_log2) : n_glob / p.cout;
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
          // const_affine (host-checked: cout == BLOCK_N <= 64, 3x3 conv, one N tile): co0 = chunk * 32 is one of two
          // compile-time values. (Selecting by co0 for the narrow transposed convs as well was tried: the extra code
          // in the BLOCK_N = 256 instantiation and a different register allocation made conv2d_2 / conv2d_4 slower
          // than without the constant path, an A/B wash overall.)
          const bool use_const = BLOCK_N <= 64 && p.const_affine != 0;
          float4 b4[8];
          if (!use_const && !p.identity_epi) {
#pragma unroll
            for (int i = 0; i < 8; ++i) b4[i] = reinterpret_cast<const float4*>(s_bias + co0)[i];
          }
          sm100::tmem_ld_wait();
          for (int sa = 1; sa < S; ++sa) {  // split accumulators: sum the partial results in fp32 (rn)
            uint32_t u[kChunk];
            sm100::tmem_ld_32x32b_x32(taddr + uint32_t((sa * nm + j) * BLOCK_N + chunk * kChunk), u);
            sm100::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < kChunk; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) + __uint_as_float(u[i]));
          }
          // bias / ReLU / BN affine of the chunk; one straight-line variant per mode so that the broadcast
          // shared-memory reads are hoisted ahead of the arithmetic (run-time flags inside the unrolled loop
          // serialise load -> use -> load and leave the LSU latency exposed)
          float f[kChunk];
          if constexpr (BLOCK_N <= 64) {
            if (use_const) {
              if (chunk == 0) epi_affine_const<0>(v, f, p);
              else epi_affine_const<(BLOCK_N > 32 ? 32 : 0)>(v, f, p);
            }
          }
          if (!use_const) {
            if (with_stats) epi_affine<true, false>(v, f, b4, s_scale + co0, s_shift + co0);
            else if (p.no_relu && p.identity_epi) {   // data-gradient convolutions: nothing to add or scale
#pragma unroll
              for (int i = 0; i < kChunk; ++i) f[i] = __uint_as_float(v[i]);
            } else if (p.no_relu) epi_affine<false, true>(v, f, b4, s_scale + co0, s_shift + co0);
            else epi_affine<false, false>(v, f, b4, s_scale + co0, s_shift + co0);
          }
          {
            if (p.head_dtype != kHeadNone) {
#pragma unroll
              for (int i = 0; i < kChunk; i += 4) {
                const float4 w0 = *reinterpret_cast<const float4*>(s_head + 0 * 64 + co0 + i);
                const float4 w1 = *reinterpret_cast<const float4*>(s_head + 1 * 64 + co0 + i);
                const float4 w2 = *reinterpret_cast<const float4*>(s_head + 2 * 64 + co0 + i);
                h0 = fmaf(f[i + 3], w0.w, fmaf(f[i + 2], w0.z, fmaf(f[i + 1], w0.y, fmaf(f[i], w0.x, h0))));
                h1 = fmaf(f[i + 3], w1.w, fmaf(f[i + 2], w1.z, fmaf(f[i + 1], w1.y, fmaf(f[i], w1.x, h1))));
                h2 = fmaf(f[i + 3], w2.w, fmaf(f[i + 2], w2.z, fmaf(f[i + 1], w2.y, fmaf(f[i], w2.x, h2))));
              }
            }
          }
          const size_t pix = (size_t(img) * Ho + oy) * Wo + ox;
          const bool pool_writer = (lane & 9) == 0;          // even row, even column of the block
          const size_t ppix = (size_t(img) * (p.H / 2) + (y >> 1)) * (p.W / 2) + (x >> 1);
          if constexpr (sizeof(Elem) == 2) {
            uint32_t pk[kChunk / 2];
#pragma unroll
            for (int i = 0; i < kChunk / 2; ++i) pk[i] = pack2<Elem>(f[2 * i], f[2 * i + 1]);
            if constexpr (ElemTraits<Elem>::kFmt == sm100::kFmtF16) {
              // a value the saturating pack clamped reads back as +-65504 (0x7BFF): 16 packed |max| instead of 32 compares
              if (p.sat_count != nullptr && (p.out != nullptr || p.pooled != nullptr)) {
                uint32_t m2 = 0;
#pragma unroll
                for (int i = 0; i < kChunk / 2; ++i) m2 = absmax2_f16(m2, pk[i]);
                m2 &= 0x7FFF7FFFu;
                if ((m2 & 0xFFFFu) >= 0x7BFFu || (m2 >> 16) >= 0x7BFFu) atomicAdd(p.sat_count, 1ull);
              }
            }
            if constexpr (sizeof(Elem) == 2 && ElemTraits<Elem>::kFmt == sm100::kFmtBF16) {
              if (with_stats) {
                // statistics of the ROUNDED values (what bn_apply and the backward read back from HBM)
                float sv[kChunk], sq[kChunk];
#pragma unroll
                for (int i = 0; i < kChunk / 2; ++i) {
                  sv[2 * i] = __uint_as_float(pk[i] << 16);
                  sv[2 * i + 1] = __uint_as_float(pk[i] & 0xFFFF0000u);
                  sq[2 * i] = sv[2 * i] * sv[2 * i];
                  sq[2 * i + 1] = sv[2 * i + 1] * sv[2 * i + 1];
                }
                const float tsum = warp_transpose_sum32(sv, lane);
                const float tsq = warp_transpose_sum32(sq, lane);
                float* row = s_stat + ew * 2 * p.cout;     // this warp's own row: plain adds in program order
                row[co0 + lane] += tsum;
                row[p.cout + co0 + lane] += tsq;
              }
            }
            if (p.tma_store) {
              // ---- staged TMA store. Row of the staging tile = one pixel (lane), kEpiRowB bytes of channels; the 16-byte
              // pieces are XOR-swizzled exactly like the tensor map's SWIZZLE_128B / _64B mode, so the writes of a
              // quarter warp hit 8 different bank groups.
              constexpr int kRowB = BLOCK_N >= 64 ? 128 : 64;
              uint8_t* st_out = epi_stage + ew * (40 * kRowB);
              uint8_t* st_pool = st_out + 32 * kRowB;
              const int half = kRowB == 128 ? (chunk & 1) : 0;
              if (half == 0) {   // the previous group's bulk stores must have finished READING the staging tile
                if (lane == 0) sm100::tma_store_wait_read<0>();
                __syncwarp();
              }
              const int sw_out = kRowB == 128 ? (lane & 7) : ((lane >> 1) & 3);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                *reinterpret_cast<uint4*>(st_out + lane * kRowB + (((half * 4 + k) ^ sw_out) << 4)) =
                    make_uint4(pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
              if (p.pooled != nullptr) {
                // 2x2 max-pool of the warp's 4 x 8 pixels out of the staging tile: lane = (pooled pixel, 16-byte piece)
                __syncwarp();
                const int pp = lane >> 2, k = lane & 3;
                const int p00 = (2 * (pp >> 2)) * 8 + 2 * (pp & 3);
                uint4 m4;
                {
                  const int pa = p00, pb = p00 + 1, pc = p00 + 8, pd = p00 + 9;
                  auto piece = [&](int px) {
                    const int sw = kRowB == 128 ? (px & 7) : ((px >> 1) & 3);
                    return *reinterpret_cast<const uint4*>(st_out + px * kRowB + (((half * 4 + k) ^ sw) << 4));
                  };
                  const uint4 a4 = piece(pa), b4v = piece(pb), c4 = piece(pc), d4 = piece(pd);
                  m4.x = max2<Elem>(max2<Elem>(a4.x, b4v.x), max2<Elem>(c4.x, d4.x));
                  m4.y = max2<Elem>(max2<Elem>(a4.y, b4v.y), max2<Elem>(c4.y, d4.y));
                  m4.z = max2<Elem>(max2<Elem>(a4.z, b4v.z), max2<Elem>(c4.z, d4.z));
                  m4.w = max2<Elem>(max2<Elem>(a4.w, b4v.w), max2<Elem>(c4.w, d4.w));
                }
                const int sw_p = kRowB == 128 ? (pp & 7) : ((pp >> 1) & 3);
                *reinterpret_cast<uint4*>(st_pool + pp * kRowB + (((half * 4 + k) ^ sw_p) << 4)) = m4;
              }
              if (kRowB == 64 || half == 1) {
                sm100::fence_proxy_async_smem();   // generic-proxy writes -> visible to the TMA (async proxy)
                __syncwarp();
                if (lane == 0) {
                  const int ng = (n_tile * BLOCK_N + chunk * kChunk) & ~(kRowB / 2 - 1);   // first column of the staged group
                  const int x0 = tx * tile_w + j * kBlkW, y0 = ty * kBlkH + 4 * q;
                  if (p.out != nullptr) {
                    if (p.mode == kModeConvT2x2) {
                      const int a_row = ng >> (p.cout_log2 + 1);                       // kernel row a of this column group
                      sm100::tma_store_4d(&p.tmOut[a_row], st_out, ng - (a_row << (p.cout_log2 + 1)), x0, y0, img);
                    } else {
                      sm100::tma_store_4d(&p.tmOut[0], st_out, ng, x0, y0, img);
                    }
                  }
                  if (p.pooled != nullptr) sm100::tma_store_4d(&p.tmPool, st_pool, ng, x0 >> 1, y0 >> 1, img);
                  sm100::tma_store_commit();
                }
              }
            } else {
            // The four lanes of a quad hold four horizontally adjacent pixels (64 B of channels each). A 4x4
            // transposition of the 16-byte pieces inside the quad lets every store instruction write 64
            // contiguous bytes per pixel (8 lines per request instead of 32 half-written sectors' worth of lines).
            // After it lane j of a quad holds piece j of all four pixels, so the horizontal half of the 2x2
            // max-pool is in-thread and only the vertical half needs a shuffle (8 instead of 32 per chunk).
            uint32_t t[kChunk / 2];
#pragma unroll
            for (int i = 0; i < kChunk / 2; ++i) t[i] = pk[i];
            if (p.out != nullptr || p.pooled != nullptr) quad_transpose_16B(t, lane);
            const int jq = lane & 3;
            if (p.out != nullptr) {
              const int xstep = p.mode == kModeConvT2x2 ? 2 : 1;   // output pixels between two lanes of the quad
              Elem* base = reinterpret_cast<Elem*>(p.out) + co0 + 8 * jq;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const size_t pk_pix = size_t(ptrdiff_t(pix) + ptrdiff_t((k - jq) * xstep));
                *reinterpret_cast<uint4*>(base + pk_pix * p.cout) = make_uint4(t[4 * k], t[4 * k + 1], t[4 * k + 2], t[4 * k + 3]);
              }
            }
            if (p.pooled != nullptr) {
              uint32_t m[8];   // [0..3]: pixels 0,1 of the quad -> pooled column 0; [4..7]: pixels 2,3 -> column 1
#pragma unroll
              for (int w = 0; w < 4; ++w) {
                m[w] = max2<Elem>(t[w], t[4 + w]);
                m[4 + w] = max2<Elem>(t[8 + w], t[12 + w]);
              }
#pragma unroll
              for (int w = 0; w < 8; ++w) m[w] = max2<Elem>(m[w], __shfl_xor_sync(0xFFFFFFFFu, m[w], 8));  // row below
              if ((lane & 8) == 0) {                             // even row of the block writes
                const size_t pp0 = (size_t(img) * (p.H / 2) + (y >> 1)) * (p.W / 2) + ((x - jq) >> 1);
                Elem* pb = reinterpret_cast<Elem*>(p.pooled) + co0 + 8 * jq;
                *reinterpret_cast<uint4*>(pb + pp0 * p.cout) = make_uint4(m[0], m[1], m[2], m[3]);
                *reinterpret_cast<uint4*>(pb + (pp0 + 1) * p.cout) = make_uint4(m[4], m[5], m[6], m[7]);
              }
            }
            }
          } else {
            if (p.out != nullptr) {
              float* dst = reinterpret_cast<float*>(p.out) + pix * p.cout + co0;
              float* dst_lo = p.out_lo ? reinterpret_cast<float*>(p.out_lo) + pix * p.cout + co0 : nullptr;
#pragma unroll
              for (int i = 0; i < kChunk / 4; ++i)
                store4_f32(dst + 4 * i, dst_lo ? dst_lo + 4 * i : nullptr, p.store_mode, f[4 * i], f[4 * i + 1],
                           f[4 * i + 2], f[4 * i + 3]);
            }
            if (p.pooled != nullptr) {
#pragma unroll
              for (int i = 0; i < kChunk; ++i) {
                f[i] = fmaxf(f[i], __shfl_xor_sync(0xFFFFFFFFu, f[i], 1));
                f[i] = fmaxf(f[i], __shfl_xor_sync(0xFFFFFFFFu, f[i], 8));
              }
              if (pool_writer) {
                float* dst = reinterpret_cast<float*>(p.pooled) + ppix * p.cout + co0;
                float* dst_lo = p.pooled_lo ? reinterpret_cast<float*>(p.pooled_lo) + ppix * p.cout + co0 : nullptr;
#pragma unroll
                for (int i = 0; i < kChunk / 4; ++i)
                  store4_f32(dst + 4 * i, dst_lo ? dst_lo + 4 * i : nullptr, p.store_mode, f[4 * i], f[4 * i + 1],
                             f[4 * i + 2], f[4 * i + 3]);
              }
            }
          }
        }
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
      }
      // accumulators drained: hand the TMEM buffer back to the MMA warp
      sm100::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) sm100::mbar_arrive(&tmem_empty_bar[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (p.tma_store && lane == 0) sm100::tma_store_wait_all<0>();   // staging tiles stay valid until the TMA has read them
    }
  }

  // ---------------------------------------------------------------- teardown
  sm100::tc_fence_before_sync();
  if constexpr (kPair) sm100::cluster_sync();   // neither CTA may leave while the peer can still signal or read it
  else __syncthreads();
  if (with_stats) {
    float* dst = p.stats + size_t(blockIdx.x) * 2 * p.cout;
    for (int i = threadIdx.x; i < 2 * p.cout; i += blockDim.x) {
      float acc = 0.f;
#pragma unroll
      for (int w = 0; w < kNumEpiWarps; ++w) acc += s_stat[w * 2 * p.cout + i];   // fixed warp order
      dst[i] = acc;
    }
  }
  if (warp_idx == 2) {
    sm100::tc_fence_after_sync();
    if constexpr (kPair) sm100::tmem_dealloc_pair(tmem_base, tmem_cols);
    else sm100::tmem_dealloc(tmem_base, tmem_cols);
  }
}

}  // namespace itmo
