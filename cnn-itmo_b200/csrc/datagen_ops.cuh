// datagen_ops.cuh — training-data synthesis and augmentation on the GPU (BASELINE configs[3], [4]
// "synthetic SDR/HDR pairs"). The reference makes its data OFFLINE in MATLAB and augments it on the host in Keras:
//   m-files/virtual_camera.m:13-35   HDR -> SDR network input: exposure 0.18*2^v/G (G = geometric-mean luminance,
//                                    :14-19), camera curve (1+n) X^y / (n + X^y) clipped at 1 (:21-28), colour by the
//                                    luminance ratio (:31-33), imwrite (8-bit PNG, :37)
//   m-files/Reinhard.m:12-26         HDR -> training target: X = 0.18/G * Y, L = X/(1+X), colour by L/Y, imwrite
//   main.py:71-76 (+78-100)          ImageDataGenerator(rescale=1/255, rotation_range=90, horizontal_flip,
//                                    vertical_flip, zoom_range=0.2), the SAME seed for inputs and targets
//                                    => keras_preprocessing apply_affine_transform (scipy.ndimage.affine_transform,
//                                    order 1, mode 'nearest') followed by the flips
// All of it is per-pixel arithmetic plus one log-luminance reduction per image: HBM-bound kernels, 3 floats per thread.
// Precisions follow MATLAB's rules for the script (EXR data is single; single*double = single; pow on double(...)):
// the numpy restatement in oracle/datagen_ref.py does the same, so uint8 outputs agree bit for bit up to the last-ulp
// freedom of pow().
#pragma once
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

namespace itmo {

// ---- seeded HDR field (stands in for the EXR corpus the reference does not ship) --------------------------------------
__host__ __device__ __forceinline__ uint32_t dg_hash(uint32_t x, uint32_t seed) {
  uint32_t h = (x * 0x9E3779B1u) ^ seed;
  h ^= h >> 16;
  h *= 0x85EBCA6Bu;
  h ^= h >> 13;
  h *= 0xC2B2AE35u;
  h ^= h >> 16;
  return h;
}
// standard normal from two hashes of (index, stream): Box-Muller in double (restated in oracle/datagen_ref.py)
__device__ __forceinline__ double dg_normal(uint32_t idx, uint32_t seed, uint32_t stream) {
  const uint32_t h1 = dg_hash(idx, seed ^ (stream * 0x632BE5ABu + 0x1234567u));
  const uint32_t h2 = dg_hash(idx, seed ^ (stream * 0x632BE5ABu + 0x89ABCDEu));
  const double u1 = (double(h1) + 0.5) * (1.0 / 4294967296.0);
  const double u2 = (double(h2) + 0.5) * (1.0 / 4294967296.0);
  return sqrt(-2.0 * log(u1)) * cos(6.283185307179586476925 * u2);
}
// hdr[n,y,x,c] = exp(1.2 * base[n, y/8, x/8, c] + 0.15 * fine[n,y,x,c]): log-normal radiance with 8-pixel structure,
// ~5 decades of dynamic range like an HDR photograph. float32 [n,h,w,3].
__global__ void __launch_bounds__(256) synth_hdr_kernel(float* __restrict__ hdr, int n, int h, int w, uint32_t seed) {
  const size_t total = size_t(n) * h * w * 3;
  const int bw = w / 8 + 1, bh = h / 8 + 1;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const int c = int(i % 3);
    size_t r = i / 3;
    const int x = int(r % w);
    r /= w;
    const int y = int(r % h);
    const int img = int(r / h);
    const uint32_t bidx = uint32_t(((size_t(img) * bh + y / 8) * bw + x / 8) * 3 + c);
    const double base = dg_normal(bidx, seed, 1u);
    const double fine = dg_normal(uint32_t(i), seed, 2u);
    hdr[i] = float(exp(1.2 * base + 0.15 * fine));
  }
}

// ---- luminance and its per-image log-average (virtual_camera.m:12-19, Reinhard.m:12-19) -------------------------------
// RGB2Lum of the HDR Toolbox: Rec.709 weights on single data — the three products and two sums each rounded to single
__device__ __forceinline__ float dg_lum(float r, float g, float b) {
  return __fadd_rn(__fadd_rn(__fmul_rn(0.2126f, r), __fmul_rn(0.7152f, g)), __fmul_rn(0.0722f, b));
}
constexpr int kDgBlocksPerImage = 64;
// partial[img][blockIdx.x] = sum over this block's pixels of log(max(double(Y), realmin)); fixed order (no atomics)
__global__ void __launch_bounds__(256) loglum_partial_kernel(const float* __restrict__ hdr, int h, int w,
                                                             double* __restrict__ partial) {
  const int img = blockIdx.y;
  const size_t npix = size_t(h) * w;
  const float* src = hdr + size_t(img) * npix * 3;
  double acc = 0.0;
  for (size_t p = blockIdx.x * size_t(blockDim.x) + threadIdx.x; p < npix; p += size_t(gridDim.x) * blockDim.x) {
    const float Y = dg_lum(src[3 * p], src[3 * p + 1], src[3 * p + 2]);
    acc += log(fmax(double(Y), DBL_MIN));
  }
  __shared__ double s[256];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {   // fixed tree
    if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[size_t(img) * gridDim.x + blockIdx.x] = s[0];
}
// G[img] = exp(sum / (h*w))
__global__ void geomean_finalize_kernel(const double* __restrict__ partial, int blocks, double npix, int n,
                                        double* __restrict__ G) {
  const int img = blockIdx.x * blockDim.x + threadIdx.x;
  if (img >= n) return;
  double s = 0.0;
  for (int b = 0; b < blocks; ++b) s += partial[size_t(img) * blocks + b];
  G[img] = exp(s * (1.0 / npix));
}

// imwrite of a double image: uint8(round(255 * x)) with saturation; NaN -> 0
__device__ __forceinline__ float dg_png_quant(float v) {
  if (!(v == v)) return 0.f;
  const double q = floor(255.0 * double(fmin(fmax(v, 0.f), 1.f)) + 0.5);
  return float(q);
}

struct CameraParams {   // one virtual camera per image (virtual_camera.m:13,21-22)
  double v;             // exposure stops, U[-4, 4]
  double n;             // N(0.6, sqrt(0.1))
  double y;             // N(0.9, sqrt(0.1))
};

// mode 0: virtual_camera.m:20-33 -> SDR input. mode 1: Reinhard.m:20-26 -> tone-mapped target.
// OutT = uint8_t: the 8-bit PNG value; OutT = float: that value / 255 (what rescale=1/255 feeds the network, main.py:71)
template <typename OutT>
__global__ void __launch_bounds__(256) tonemap_kernel(const float* __restrict__ hdr, OutT* __restrict__ out, int h, int w,
                                                      const double* __restrict__ G, const CameraParams* __restrict__ cam,
                                                      int mode, double key) {
  const int img = blockIdx.y;
  const size_t npix = size_t(h) * w;
  const float* src = hdr + size_t(img) * npix * 3;
  OutT* dst = out + size_t(img) * npix * 3;
  double delta_t, cn = 0.0, cy = 1.0;
  if (mode == 0) {
    delta_t = key * exp2(cam[img].v) / G[img];   // :19
    cn = cam[img].n;
    cy = cam[img].y;
  } else {
    delta_t = key / G[img];                      // Reinhard.m:20
  }
  for (size_t p = blockIdx.x * size_t(blockDim.x) + threadIdx.x; p < npix; p += size_t(gridDim.x) * blockDim.x) {
    const float r = src[3 * p], g = src[3 * p + 1], b = src[3 * p + 2];
    const float Y = dg_lum(r, g, b);
    const float X = float(delta_t * double(Y));   // double * single = single in MATLAB
    float L;
    if (mode == 0) {
      const double xp = pow(double(X), cy);
      const double a = (1.0 + cn) * (xp / (cn + xp));   // :26
      L = float(fmin(1.0, a));                          // :27, stored back into the single array
    } else {
      L = float(double(X) / (1.0 + double(X)));         // Reinhard.m:22
    }
    const float ratio = __fdiv_rn(L, Y);                // single ./ single (NaN for a black pixel, like the script)
    const float o0 = dg_png_quant(__fmul_rn(r, ratio)), o1 = dg_png_quant(__fmul_rn(g, ratio)),
                o2 = dg_png_quant(__fmul_rn(b, ratio));
    if constexpr (sizeof(OutT) == 1) {
      dst[3 * p] = uint8_t(o0);
      dst[3 * p + 1] = uint8_t(o1);
      dst[3 * p + 2] = uint8_t(o2);
    } else {
      dst[3 * p] = __fdiv_rn(o0, 255.f);
      dst[3 * p + 1] = __fdiv_rn(o1, 255.f);
      dst[3 * p + 2] = __fdiv_rn(o2, 255.f);
    }
  }
}

// ---- augmentation (main.py:71-76) ---------------------------------------------------------------------------------
// out[r][c] = A[r'][c'] with (r', c') the flipped position and A = scipy.ndimage.affine_transform(in, M, offset,
// order=1, mode='nearest') = bilinear interpolation of `in` at clamp(M (r', c') + offset, 0, size-1). The host builds
// M / offset exactly like keras_preprocessing (rotation x zoom, transform_matrix_offset_center).
struct AffineParams {
  double m00, m01, m10, m11, off0, off1;   // input (row, col) = M * output (row, col) + off
  int flip_h, flip_v;
};
// in: [n,h,w,3] uint8 (0..255) or float; out float32 = interpolated value * scale (scale = 1/255: Keras rescale)
template <typename InT>
__global__ void __launch_bounds__(256) augment_kernel(const InT* __restrict__ in, float* __restrict__ out, int h, int w,
                                                      const AffineParams* __restrict__ params, float scale) {
  const int img = blockIdx.y;
  const AffineParams P = params[img];
  const size_t npix = size_t(h) * w;
  const InT* src = in + size_t(img) * npix * 3;
  float* dst = out + size_t(img) * npix * 3;
  for (size_t p = blockIdx.x * size_t(blockDim.x) + threadIdx.x; p < npix; p += size_t(gridDim.x) * blockDim.x) {
    const int r = int(p / w), c = int(p % w);
    const int rr = P.flip_v ? h - 1 - r : r, cc = P.flip_h ? w - 1 - c : c;
    double y = P.m00 * rr + P.m01 * cc + P.off0;
    double x = P.m10 * rr + P.m11 * cc + P.off1;
    y = fmin(fmax(y, 0.0), double(h - 1));   // mode='nearest'
    x = fmin(fmax(x, 0.0), double(w - 1));
    const int y0 = min(int(y), h - 2 < 0 ? 0 : h - 2), x0 = min(int(x), w - 2 < 0 ? 0 : w - 2);
    const double fy = y - y0, fx = x - x0;
    const int y1 = min(y0 + 1, h - 1), x1 = min(x0 + 1, w - 1);
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      const double v00 = double(src[(size_t(y0) * w + x0) * 3 + ch]), v01 = double(src[(size_t(y0) * w + x1) * 3 + ch]);
      const double v10 = double(src[(size_t(y1) * w + x0) * 3 + ch]), v11 = double(src[(size_t(y1) * w + x1) * 3 + ch]);
      const double v = (1.0 - fy) * ((1.0 - fx) * v00 + fx * v01) + fy * ((1.0 - fx) * v10 + fx * v11);
      dst[p * 3 + ch] = __fmul_rn(float(v), scale);
    }
  }
}

}  // namespace itmo
