// frame_ops.cuh — HBM-bound kernels around the U-Net for frames larger than its native 512x512 input
// (BASELINE config 4: 3840x2160 SDR frames) and the optional inverse-Reinhard output stage.
//
//   tile_gather_kernel      frame [F,H,W,3] u8 -> 512x512 tiles (origins from the tile plan, all inside the frame)
//   tile_scatter_kernel     tile outputs -> frame, each tile contributing its "valid" rectangle; for uint8 output
//                           the per-frame reductions of m-files/inverse_Reinhard.m:5-14 are fused in
//                           (sum of log(max(X, realmin)), count of zero-luminance pixels)
//   inverse_reinhard_kernel m-files/inverse_Reinhard.m:2-4,16-23 per pixel, from the reduced statistics
//
// The reference has no tiling (fixed 512x512 Input, model.py:204,207). The plan keeps every tile inside the frame
// on the 16-pixel grid of the four 2x2 pools, 96 >= 92 pixels (the receptive-field reach of this graph, SURVEY
// section 8e) away from every interior cut, so that a stitched frame equals the fully-convolutional network
// applied to the whole frame, including the 'same' zero padding at the frame border.
#pragma once
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

namespace itmo {

constexpr int kTile = 512;
constexpr int kTileHalo = 96;
constexpr int kPixPerThread = 16;  // 48 bytes of uint8 RGB = three 16-byte vectors

struct TileRect {  // one tile of the plan
  int oy, ox;      // origin in the frame
  int y0, y1;      // valid output rows [y0, y1) in frame coordinates
  int x0, x1;
};

// 1-D plan: origins and valid ranges along one axis of length `extent` (multiple of 16, >= 512).
// Returns the number of tiles (or -1); writes up to cap entries.
inline int plan_axis(int extent, int* origin, int* lo, int* hi, int cap) {
  if (extent < kTile || extent % 16) return -1;
  int n = 0, covered = 0;
  while (covered < extent) {
    int o = covered == 0 ? 0 : covered - kTileHalo;
    if (o + kTile > extent) o = extent - kTile;
    const int v0 = covered;
    const int v1 = (o + kTile == extent) ? extent : o + kTile - kTileHalo;
    if (n < cap) {
      origin[n] = o;
      lo[n] = v0;
      hi[n] = v1;
    }
    ++n;
    covered = v1;
  }
  return n;
}

// tiles [nt,512,512,3] u8 <- frame rows; one thread moves 16 bytes
__global__ void __launch_bounds__(256) tile_gather_kernel(const uint8_t* __restrict__ frame, int W, uint8_t* __restrict__ tiles,
                                                          const TileRect* __restrict__ rects, int nt) {
  constexpr int kVecPerRow = kTile * 3 / 16;  // 96
  const size_t total = size_t(nt) * kTile * kVecPerRow;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const int v = int(i % kVecPerRow);
    const int row = int((i / kVecPerRow) % kTile);
    const int t = int(i / (size_t(kVecPerRow) * kTile));
    const TileRect r = rects[t];
    const uint4* src = reinterpret_cast<const uint4*>(frame + (size_t(r.oy + row) * W + r.ox) * 3) + v;
    reinterpret_cast<uint4*>(tiles)[i] = *src;
  }
}

struct FrameStats {      // per frame, inverse_Reinhard.m:5-14
  double log_sum;        // 0, or +inf when a pixel's term was infinite (as-written mode, I == 1); the finite terms are below
  unsigned long long zero_lum;  // PB1 = #(I == 0)
  unsigned long long log_sum_fx;  // sum over pixels of log(max(X, realmin)) in 40.24 fixed point, two's complement.
                                  // INTEGER atomics: addition is associative, so the frame statistic — and with it every
                                  // HDR pixel — is bit-reproducible whatever order the blocks finish in (each warp's /
                                  // block's own partial is a fixed-order double sum, rounded to 2^-24 once; |term| <= 745,
                                  // 2^33 pixels fit)
};
constexpr double kLogFx = 16777216.0;   // 2^24
__device__ __forceinline__ void stats_add(FrameStats* st, double lsum, unsigned long long zeros) {
  atomicAdd(&st->log_sum_fx, static_cast<unsigned long long>(__double2ll_rn(lsum * kLogFx)));
  atomicAdd(&st->zero_lum, zeros);
}
__device__ __forceinline__ double stats_log_sum(const FrameStats* st) {
  if (isinf(st->log_sum)) return st->log_sum;   // an infinite term was seen (see the kernels): the double sum would be inf
  return double(static_cast<long long>(st->log_sum_fx)) / kLogFx;
}

__device__ __forceinline__ double srgb_pow22(uint8_t v) { return pow(double(v) / 255.0, 2.2); }  // inverse_Reinhard.m:2
__device__ __forceinline__ double rgb2lum(double r, double g, double b) {  // HDR-Toolbox RGB2Lum (Rec.709), :3
  return 0.2126 * r + 0.7152 * g + 0.0722 * b;
}
// inverse_Reinhard.m:4 as written divides by (I - 1); the inverse of Reinhard.m:22 L = X/(1+X) is X = L/(1-L).
// The repaired form clamps I to 1 - 0.5/255 first: a saturated (255,255,255) pixel would otherwise give X = +inf, an
// infinite geometric mean and an all-zero frame — and saturated highlights are the core inverse-tone-mapping case.
constexpr double kLumMax = 1.0 - 0.5 / 255.0;
__device__ __forceinline__ double tone_inverse(double I, int as_written) {
  return as_written ? I / (I - 1.0) : fmin(I, kLumMax) / (1.0 - fmin(I, kLumMax));
}
// One pixel's term of the log-average (inverse_Reinhard.m:5-11). As written: log(max(X, realmin)) for every pixel.
// Repaired: zero-luminance pixels are LEFT OUT of the geometric mean (each would add log(realmin) = -708 and 1 % of
// black pixels would raise the scale ~1000x into the 2^32 clip); the mean is taken over the non-zero count.
__device__ __forceinline__ double log_term(double I, int as_written) {
  if (!as_written && I == 0.0) return 0.0;
  return log(fmax(tone_inverse(I, as_written), DBL_MIN));   // max(NaN, realmin) = realmin in MATLAB and in fmax
}

// out frame <- valid rectangles of the tiles. T = uint8_t (optionally with the fused reductions) or float.
template <typename T, bool STATS>
__global__ void __launch_bounds__(256) tile_scatter_kernel(const T* __restrict__ tiles, T* __restrict__ frame, int W,
                                                           const TileRect* __restrict__ rects, int nt, FrameStats* stats,
                                                           int as_written) {
  // work item = (tile, row of its valid rectangle, group of 16 pixels); valid x ranges are multiples of 16
  double lsum = 0.0;
  unsigned long long zeros = 0;
  const int t = blockIdx.y;
  const TileRect r = rects[t];
  const int groups = (r.x1 - r.x0) / kPixPerThread;
  const int items = (r.y1 - r.y0) * groups;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < items; i += gridDim.x * blockDim.x) {
    const int y = r.y0 + i / groups;
    const int x = r.x0 + (i % groups) * kPixPerThread;
    const T* src = tiles + ((size_t(t) * kTile + (y - r.oy)) * kTile + (x - r.ox)) * 3;
    T* dst = frame + (size_t(y) * W + x) * 3;
    constexpr int kVec = int(sizeof(T)) * 3 * kPixPerThread / 16;
    uint4 v[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) v[k] = reinterpret_cast<const uint4*>(src)[k];
#pragma unroll
    for (int k = 0; k < kVec; ++k) reinterpret_cast<uint4*>(dst)[k] = v[k];
    if constexpr (STATS) {
      const uint8_t* px = reinterpret_cast<const uint8_t*>(v);
#pragma unroll 4
      for (int p = 0; p < kPixPerThread; ++p) {
        const double I = rgb2lum(srgb_pow22(px[3 * p]), srgb_pow22(px[3 * p + 1]), srgb_pow22(px[3 * p + 2]));
        const double lt = log_term(I, as_written);
        if (isinf(lt)) stats->log_sum = lt;   // as written, I == 1 gives X = +inf: kept out of the fixed-point sum, flagged
        else lsum += lt;                      // (every writer stores the same value: order-free)
        zeros += (I == 0.0);
      }
    }
  }
  if constexpr (STATS) {
    for (int o = 16; o > 0; o >>= 1) {
      lsum += __shfl_xor_sync(0xFFFFFFFFu, lsum, o);
      zeros += __shfl_xor_sync(0xFFFFFFFFu, zeros, o);
    }
    __shared__ double s_l[8];
    __shared__ unsigned long long s_z[8];
    if ((threadIdx.x & 31) == 0) {
      s_l[threadIdx.x >> 5] = lsum;
      s_z[threadIdx.x >> 5] = zeros;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < int(blockDim.x >> 5); ++w) {
        lsum += s_l[w];
        zeros += s_z[w];
      }
      stats_add(stats, lsum, zeros);
    }
  }
}

// Statistics of a whole uint8 frame (standalone inverse Reinhard, not behind the tiled predict).
__global__ void __launch_bounds__(256) frame_stats_kernel(const uint8_t* __restrict__ frame, size_t n_pix, FrameStats* stats,
                                                          int as_written) {
  double lsum = 0.0;
  unsigned long long zeros = 0;
  for (size_t p = blockIdx.x * size_t(blockDim.x) + threadIdx.x; p < n_pix; p += size_t(gridDim.x) * blockDim.x) {
    const double I = rgb2lum(srgb_pow22(frame[3 * p]), srgb_pow22(frame[3 * p + 1]), srgb_pow22(frame[3 * p + 2]));
    const double lt = log_term(I, as_written);
    if (isinf(lt)) stats->log_sum = lt;
    else lsum += lt;
    zeros += (I == 0.0);
  }
  for (int o = 16; o > 0; o >>= 1) {
    lsum += __shfl_xor_sync(0xFFFFFFFFu, lsum, o);
    zeros += __shfl_xor_sync(0xFFFFFFFFu, zeros, o);
  }
  if ((threadIdx.x & 31) == 0) stats_add(stats, lsum, zeros);
}

// inverse_Reinhard.m:11-23 per pixel. as_written = 1 follows the script literally (IEEE semantics of its
// degenerate cases included: G_E = exp(P log G_X / PB1 - PB2 log a / PB1), division by the zero-luminance count).
// as_written = 0 is the repaired form: X = I'/(1-I') with I' = min(I, 1 - 0.5/255), E = key * X / G_X (geometric mean
// of the recovered luminance over the non-black pixels normalised to `key`, mirroring Reinhard.m:19-21), zero-luminance
// pixels stay black.
// OutT = float, or __half (ITMO_OUT_HDR_INTENDED_F16: values above fp16's range clamp to 65504).
template <typename OutT>
__global__ void __launch_bounds__(256) inverse_reinhard_kernel(const uint8_t* __restrict__ frame, OutT* __restrict__ hdr,
                                                               size_t n_pix, const FrameStats* __restrict__ stats,
                                                               int as_written, double key) {
  const double P = double(n_pix);
  const double nz = P - double(stats->zero_lum);
  // :11; the repaired form averages over the non-zero-luminance pixels (an all-black frame stays black)
  const double log_sum = stats_log_sum(stats);
  const double G_X = as_written ? exp(log_sum / P) : (nz > 0 ? exp(log_sum / nz) : 1.0);
  double scale;
  if (as_written) {
    const double PB1 = double(stats->zero_lum), PB2 = P - PB1;  // :13-14
    const double a = 0.18;
    const double G_E = exp(P * log(G_X) / PB1 - PB2 * log(a) / PB1);  // :16
    scale = G_E / a;                                                      // :17  E = G_E * X / a
  } else {
    scale = key / G_X;
  }
  for (size_t p = blockIdx.x * size_t(blockDim.x) + threadIdx.x; p < n_pix; p += size_t(gridDim.x) * blockDim.x) {
    const double r = srgb_pow22(frame[3 * p]), g = srgb_pow22(frame[3 * p + 1]), b = srgb_pow22(frame[3 * p + 2]);
    const double I = rgb2lum(r, g, b);
    double E = scale * tone_inverse(I, as_written);
    if (E != E) E = DBL_MIN;                 // :18  E(isnan(E)) = realmin
    if (E >= 4294967296.0) E = 4294967296.0;  // :19
    double q = E / I;                         // :22
    if (!as_written && I == 0.0) q = 0.0;
    if constexpr (sizeof(OutT) == 2) {
      hdr[3 * p] = __float2half_rn(fminf(float(r * q), 65504.f));
      hdr[3 * p + 1] = __float2half_rn(fminf(float(g * q), 65504.f));
      hdr[3 * p + 2] = __float2half_rn(fminf(float(b * q), 65504.f));
    } else {
      hdr[3 * p] = float(r * q);
      hdr[3 * p + 1] = float(g * q);
      hdr[3 * p + 2] = float(b * q);
    }
  }
}

}  // namespace itmo
