// train_ops.cuh — the HBM-bound stages of the training step (main.py:126-132 + model.py:281) around the
// implicit-GEMM kernels: BatchNormalization in training mode (batch statistics over N,H,W of the ReLU output,
// model.py:196,200), Dropout(0.5) (model.py:226,239), MaxPooling2D forward/backward, the 1x1 sigmoid head with the
// 'mse' loss (model.py:276,281) and the optimizer update (RMSprop as compiled at model.py:281; the Adam variant of
// the comments at model.py:280,282). Activations and activation gradients are bf16 NHWC, parameters, statistics
// and parameter gradients fp32; every kernel moves 16 bytes (8 channels) per thread and access.
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

namespace itmo {

struct Bf8 {
  uint4 raw;
  __device__ __forceinline__ void get(float (&f)[8]) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 v = __bfloat1622float2(h[i]);
      f[2 * i] = v.x;
      f[2 * i + 1] = v.y;
    }
  }
  __device__ __forceinline__ void set(const float (&f)[8]) {
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  }
};
__device__ __forceinline__ Bf8 ld8(const __nv_bfloat16* p) {
  Bf8 v;
  v.raw = *reinterpret_cast<const uint4*>(p);
  return v;
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const Bf8& v) { *reinterpret_cast<uint4*>(p) = v.raw; }

// Dropout keep decision of element `e` (flat NHWC index, 32-bit wrap) under `seed`: murmur3 finaliser. The CPU oracle
// (oracle/unet_train_ref.py: dropout_keep_mask) restates exactly this so both sides drop the same units.
__device__ __forceinline__ bool drop_keep(uint32_t e, uint32_t seed) {
  uint32_t h = (e * 0x9E3779B1u) ^ seed;
  h ^= h >> 16;
  h *= 0x85EBCA6Bu;
  h ^= h >> 13;
  h *= 0xC2B2AE35u;
  h ^= h >> 16;
  return (h >> 15) & 1u;
}

// ---- deterministic reductions --------------------------------------------------------------------------------------
// No kernel of the training step uses floating-point atomics: every per-channel reduction is (1) a per-thread sum in
// program order, (2) a block reduction in thread-index order through shared memory, (3) one row of per-block partial
// sums in a workspace, (4) a finalize kernel that adds the rows in block-index order (in double). Grids are fixed
// functions of the shapes, so two runs of the same step give bit-identical gradients (tests/test_gpu_train.py::
// test_backward_is_bit_reproducible), and a gradient buffer receives exactly ONE addition per micro-batch and tensor,
// which makes "two micro-batches on one rank" == "one micro-batch on each of two ranks + all-reduce" bit for bit.

// Thread (pl, cg) of a 256-thread block holds v[8] for channels cg*8 .. cg*8+7 (groups = C/8 divides 256, lanes =
// 256 / groups). out_row[c] = sum over pl = 0 .. lanes-1, in that order. s_red: 256 * 8 floats.
__device__ __forceinline__ void block_reduce_channels(const float (&v)[8], float* s_red, int C, int lanes, int pl, int cg,
                                                      float* __restrict__ out_row) {
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 8; ++i) s_red[pl * C + cg * 8 + i] = v[i];
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float acc = 0.f;
    for (int l = 0; l < lanes; ++l) acc += s_red[l * C + c];
    out_row[c] = acc;
  }
}

// Finalize kernels: a block of 256 threads = 8 columns x 32 row groups. Row group g adds rows g, g+32, ... of the
// partial-sum workspace (in double), the 32 group sums are then added in group order: a fixed summation order with
// nrows/32 dependent loads per thread instead of nrows (a one-thread-per-column loop over ~600 rows cost ~25 us per
// launch, x 55 launches per micro-batch). The result is valid in the threads of row group 0 (threadIdx.x < 8).
__device__ __forceinline__ double reduce_rows_8x32(const float* __restrict__ partials, int nrows, int stride, int col,
                                                   bool col_ok, double* s_red /* [256] */) {
  const int g = threadIdx.x >> 3;
  double acc = 0.0;
  if (col_ok) {
    int b = g;
    for (; b + 96 < nrows; b += 128) {   // four independent loads in flight, added in row order
      const float v0 = partials[size_t(b) * stride + col], v1 = partials[size_t(b + 32) * stride + col];
      const float v2 = partials[size_t(b + 64) * stride + col], v3 = partials[size_t(b + 96) * stride + col];
      acc += double(v0);
      acc += double(v1);
      acc += double(v2);
      acc += double(v3);
    }
    for (; b < nrows; b += 32) acc += double(partials[size_t(b) * stride + col]);
  }
  __syncthreads();
  s_red[threadIdx.x] = acc;
  __syncthreads();
  double tot = 0.0;
  if (g == 0)
    for (int k = 0; k < 32; ++k) tot += s_red[k * 8 + (threadIdx.x & 7)];
  return tot;
}

// ---- BN training statistics: sum / sum of squares per channel of z = ReLU(conv + b) ------------------------------
// grid-stride over pixels; blockDim = 256; thread -> (pixel lane, 8-channel group).
// partials[blockIdx.x][0..C) = sum, [C..2C) = sumsq of this block's pixels.
__global__ void __launch_bounds__(256, 4) bn_stats_kernel(const __nv_bfloat16* __restrict__ z, size_t npix, int C,
                                                       float* __restrict__ partials) {
  __shared__ float s_red[256 * 8];
  const int groups = C >> 3;
  const int cg = threadIdx.x % groups;
  const int lanes = blockDim.x / groups;   // pixels per block iteration (C <= 512 -> groups <= 64 -> lanes >= 4)
  const int pl = threadIdx.x / groups;
  float a[8] = {}, q[8] = {};
  const size_t stride = size_t(gridDim.x) * lanes;
  for (size_t pix = size_t(blockIdx.x) * lanes + pl; pix < npix; pix += 4 * stride) {
    Bf8 v[4];   // four independent 16-byte loads in flight per thread
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (pix + u * stride < npix) v[u] = ld8(z + (pix + u * stride) * C + cg * 8);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (pix + u * stride >= npix) break;
      float f[8];
      v[u].get(f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        a[i] += f[i];
        q[i] = fmaf(f[i], f[i], q[i]);
      }
    }
  }
  float* row = partials + size_t(blockIdx.x) * 2 * C;
  block_reduce_channels(a, s_red, C, lanes, pl, cg, row);
  block_reduce_channels(q, s_red, C, lanes, pl, cg, row + C);
}

// mean / biased variance -> the affine bn_apply uses, the values backward needs, and the moving statistics
// (Keras BatchNormalization: momentum 0.99; the moving variance takes the sample variance var * M / (M - (1 + eps))).
// partials: [nblocks][2*C] per-block sums (bn_stats_kernel or the conv epilogue), added here in block order.
__global__ void __launch_bounds__(256) bn_finalize_kernel(const float* __restrict__ partials, int nblocks, int C, double M,
                                                          float eps, float momentum, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float* __restrict__ mean,
                                                          float* __restrict__ invstd, float* __restrict__ scale,
                                                          float* __restrict__ shift, float* __restrict__ moving_mean,
                                                          float* __restrict__ moving_var) {
  __shared__ double s_red[256];
  const int i = blockIdx.x * 8 + (threadIdx.x & 7);     // grid = ceil(C / 8)
  const double sum = reduce_rows_8x32(partials, nblocks, 2 * C, i, i < C, s_red);
  const double sq = reduce_rows_8x32(partials, nblocks, 2 * C, C + i, i < C, s_red);
  if (threadIdx.x >= 8 || i >= C) return;
  const double m = sum / M;
  double v = sq / M - m * m;
  if (v < 0) v = 0;
  const float is = float(1.0 / sqrt(v + double(eps)));
  mean[i] = float(m);
  invstd[i] = is;
  const float s = gamma[i] * is;
  scale[i] = s;
  shift[i] = beta[i] - float(m) * s;
  if (moving_mean) {
    moving_mean[i] = momentum * moving_mean[i] + (1.f - momentum) * float(m);
    const double vs = v * (M / (M - (1.0 + double(eps))));
    moving_var[i] = momentum * moving_var[i] + (1.f - momentum) * float(vs);
  }
}

// ---- BN apply (+ Dropout) (+ MaxPool): one thread = one 2x2 pixel window x 8 channels --------------------------------
// a = z*scale + shift; with dropout a = keep ? 2a : 0 (inverted dropout, rate 0.5); pooled = max over the window of a.
__global__ void __launch_bounds__(256) bn_apply_kernel(const __nv_bfloat16* __restrict__ z, __nv_bfloat16* __restrict__ a,
                                                       __nv_bfloat16* __restrict__ pooled, int NB, int H, int W, int C,
                                                       const float* __restrict__ scale, const float* __restrict__ shift,
                                                       uint32_t drop_seed, int dropout) {
  const int groups = C >> 3;
  const int Hp = H >> 1, Wp = W >> 1;
  const size_t total = size_t(NB) * Hp * Wp * groups;
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  // the channel group of a thread never changes when the grid stride is a multiple of the group count (always, for the
  // power-of-two widths of this network): scale / shift are then loaded once, not once per window
  const bool fixed_cg = stride % groups == 0;
  const bool small = total < (size_t(1) << 31);          // 32-bit index arithmetic (64-bit divisions are ~10x dearer)
  float sc[8], sh[8];
  if (fixed_cg) {
    const int cg0 = int((blockIdx.x * size_t(blockDim.x) + threadIdx.x) % groups);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sc[i] = scale[cg0 * 8 + i];
      sh[i] = shift[cg0 * 8 + i];
    }
  }
  for (size_t gid = blockIdx.x * size_t(blockDim.x) + threadIdx.x; gid < total; gid += stride) {
    int cg, xp, yp, img;
    if (small) {
      uint32_t w = uint32_t(gid);
      cg = int(w % uint32_t(groups));
      w /= uint32_t(groups);
      xp = int(w % uint32_t(Wp));
      w /= uint32_t(Wp);
      yp = int(w % uint32_t(Hp));
      img = int(w / uint32_t(Hp));
    } else {
      cg = int(gid % groups);
      size_t w = gid / groups;
      xp = int(w % Wp);
      w /= Wp;
      yp = int(w % Hp);
      img = int(w / Hp);
    }
    if (!fixed_cg) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        sc[i] = scale[cg * 8 + i];
        sh[i] = shift[cg * 8 + i];
      }
    }
    float mx[8];
    Bf8 zin[4];   // the four loads of the window first: all in flight before the first dependent store
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const size_t pix = (size_t(img) * H + 2 * yp + (k >> 1)) * W + 2 * xp + (k & 1);
      zin[k] = ld8(z + pix * C + cg * 8);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const size_t pix = (size_t(img) * H + 2 * yp + (k >> 1)) * W + 2 * xp + (k & 1);
      const size_t e0 = pix * C + cg * 8;
      float f[8];
      zin[k].get(f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        f[i] = fmaf(f[i], sc[i], sh[i]);
        if (dropout) f[i] = drop_keep(uint32_t(e0 + i), drop_seed) ? 2.f * f[i] : 0.f;
      }
      Bf8 o;
      o.set(f);
      st8(a + e0, o);
      o.get(f);   // pool over the stored (rounded) values: what the next layer and the backward see
#pragma unroll
      for (int i = 0; i < 8; ++i) mx[i] = k == 0 ? f[i] : fmaxf(mx[i], f[i]);
    }
    if (pooled) {
      Bf8 o;
      o.set(mx);
      st8(pooled + ((size_t(img) * Hp + yp) * Wp + xp) * C + cg * 8, o);
    }
  }
}

// ---- head: Conv2D(3, 1, sigmoid) (model.py:276) + mse (model.py:281), forward and backward in one pass -----------------
// A warp covers 4 pixels x 8 channel groups per iteration: lane = (pixel slot lane >> 3, channel group lane & 7), so every
// load / store instruction of the 64-channel tensors moves 512 contiguous bytes per warp and a thread keeps only ITS 8
// channels: 24 head weights + 24 weight-gradient accumulators in registers (the first version kept all 64 channels of a
// pixel per thread: 255 registers, 12 % occupancy, half-used sectors, 22 % of DRAM peak). The three pre-activations are
// reduced over the 8 lanes of a pixel with xor shuffles. Per block: partials[blockIdx.x][0..191] = dW[k][c] (index
// k*64 + c), [192..194] = db, [195] = sum (p-y)^2, [196] = #pixels with argmax_k p == argmax_k y (Keras 'accuracy'
// resolves to categorical_accuracy for this output); head_finalize_kernel adds the blocks in index order.
// d_a = (2/count) (p-y) p (1-p) W. y: target in [0,1], float32 [P,3]. pred (optional): float32 [P,3].
constexpr int kHeadPartials = 200;
__global__ void __launch_bounds__(256) head_train_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ y,
                                                         const float* __restrict__ wh,   // [3][64]
                                                         const float* __restrict__ bh, size_t npix, float inv_count,
                                                         __nv_bfloat16* __restrict__ d_a, float* __restrict__ pred,
                                                         float* __restrict__ partials) {
  __shared__ float s_w[8][8][27];    // [warp][channel group][24 dW + 3 db]
  __shared__ float s_m[8][2];        // [warp][loss, hits]
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cg = lane & 7, ps = lane >> 3;
  float w[3][8], bk[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    bk[k] = bh[k];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[k][i] = wh[k * 64 + cg * 8 + i];
  }
  float g[3][8] = {}, gb[3] = {};
  float loss = 0.f, hits = 0.f;
  // Each lane handles kU pixels per iteration (pixel groups a whole grid stride apart): kU independent 16-byte loads in
  // flight per thread. With one load per iteration the kernel was latency-bound at 1.7 TB/s (8 KB in flight per SM).
  constexpr int kU = 4;
  const size_t per_iter = size_t(gridDim.x) * 32;   // 8 warps x 4 pixels per block
  const size_t n_iter = (npix + per_iter * kU - 1) / (per_iter * kU);
  for (size_t it = 0; it < n_iter; ++it) {
    const size_t pix0 = it * per_iter * kU + size_t(blockIdx.x) * 32 + wid * 4 + ps;
    Bf8 raw[kU];
    float yv[kU][3];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const size_t pix = pix0 + u * per_iter;
      if (pix < npix) {
        raw[u] = ld8(a + pix * 64 + cg * 8);
        yv[u][0] = y[pix * 3];
        yv[u][1] = y[pix * 3 + 1];
        yv[u][2] = y[pix * 3 + 2];
      } else {
        raw[u].raw = make_uint4(0u, 0u, 0u, 0u);
        yv[u][0] = yv[u][1] = yv[u][2] = 0.f;
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const size_t pix = pix0 + u * per_iter;
      const bool ok = pix < npix;
      float av[8];
      raw[u].get(av);
      float z[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t = fmaf(av[i], w[k][i], t);
        t += __shfl_xor_sync(0xFFFFFFFFu, t, 1);
        t += __shfl_xor_sync(0xFFFFFFFFu, t, 2);
        t += __shfl_xor_sync(0xFFFFFFFFu, t, 4);
        z[k] = t + bk[k];
      }
      const float p0 = 1.f / (1.f + expf(-z[0])), p1 = 1.f / (1.f + expf(-z[1])), p2 = 1.f / (1.f + expf(-z[2]));
      if (ok) {
        const float y0 = yv[u][0], y1 = yv[u][1], y2 = yv[u][2];
        const float e0 = p0 - y0, e1 = p1 - y1, e2 = p2 - y2;
        const float dl0 = 2.f * inv_count * e0 * p0 * (1.f - p0);
        const float dl1 = 2.f * inv_count * e1 * p1 * (1.f - p1);
        const float dl2 = 2.f * inv_count * e2 * p2 * (1.f - p2);
        if (cg == 0) {
          loss += e0 * e0 + e1 * e1 + e2 * e2;
          const int ap = p0 >= p1 ? (p0 >= p2 ? 0 : 2) : (p1 >= p2 ? 1 : 2);   // first maximum, like argmax
          const int ay = y0 >= y1 ? (y0 >= y2 ? 0 : 2) : (y1 >= y2 ? 1 : 2);
          hits += ap == ay ? 1.f : 0.f;
          gb[0] += dl0;
          gb[1] += dl1;
          gb[2] += dl2;
        }
        if (pred && cg < 3) pred[pix * 3 + cg] = cg == 0 ? p0 : cg == 1 ? p1 : p2;
        float f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          f[i] = dl0 * w[0][i] + dl1 * w[1][i] + dl2 * w[2][i];
          g[0][i] = fmaf(dl0, av[i], g[0][i]);
          g[1][i] = fmaf(dl1, av[i], g[1][i]);
          g[2][i] = fmaf(dl2, av[i], g[2][i]);
        }
        Bf8 o;
        o.set(f);
        st8(d_a + pix * 64 + cg * 8, o);
      }
    }
  }
  // the 4 pixel slots of a warp (lanes cg, cg+8, cg+16, cg+24): fixed xor tree
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      g[k][i] += __shfl_xor_sync(0xFFFFFFFFu, g[k][i], 8);
      g[k][i] += __shfl_xor_sync(0xFFFFFFFFu, g[k][i], 16);
    }
    gb[k] += __shfl_xor_sync(0xFFFFFFFFu, gb[k], 8);
    gb[k] += __shfl_xor_sync(0xFFFFFFFFu, gb[k], 16);
  }
#pragma unroll
  for (int o = 8; o <= 16; o <<= 1) {
    loss += __shfl_xor_sync(0xFFFFFFFFu, loss, o);
    hits += __shfl_xor_sync(0xFFFFFFFFu, hits, o);
  }
  if (ps == 0) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
      for (int i = 0; i < 8; ++i) s_w[wid][cg][k * 8 + i] = g[k][i];
      s_w[wid][cg][24 + k] = gb[k];
    }
    if (cg == 0) {
      s_m[wid][0] = loss;
      s_m[wid][1] = hits;
    }
  }
  __syncthreads();
  float* row = partials + size_t(blockIdx.x) * kHeadPartials;
  const int t = threadIdx.x;
  if (t < 197) {
    float acc = 0.f;
    if (t < 192) {
      const int k = t >> 6, c = t & 63;
      for (int wv = 0; wv < 8; ++wv) acc += s_w[wv][c >> 3][k * 8 + (c & 7)];
    } else if (t < 195) {
      for (int wv = 0; wv < 8; ++wv) acc += s_w[wv][0][24 + (t - 192)];
    } else {
      for (int wv = 0; wv < 8; ++wv) acc += s_m[wv][t - 195];
    }
    row[t] = acc;
  }
}

// adds the blocks' rows in index order: head kernel / bias gradients (Keras layout [64][3]) and the loss / accuracy sums
__global__ void __launch_bounds__(256) head_finalize_kernel(const float* __restrict__ partials, int nblocks,
                                                            float* __restrict__ dwh, float* __restrict__ dbh,
                                                            double* __restrict__ acc) {
  __shared__ double s_red[256];
  const int t = blockIdx.x * 8 + (threadIdx.x & 7);     // grid = ceil(197 / 8) = 25
  const double s = reduce_rows_8x32(partials, nblocks, kHeadPartials, t, t < 197, s_red);
  if (threadIdx.x >= 8 || t >= 197) return;
  if (t < 192) dwh[(t & 63) * 3 + (t >> 6)] += float(s);
  else if (t < 195) dbh[t - 192] += float(s);
  else acc[t - 195] += s;
}

// ---- BN backward, pass 1: per-channel sums of dA' and dA' * zhat (dA' = dA through the dropout mask) ---------------
// partials[blockIdx.x][0..C) = sum dA', [C..2C) = sum dA' * (z - mean) * invstd over this block's pixels
__global__ void __launch_bounds__(256, 4) bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dA,
                                                            const __nv_bfloat16* __restrict__ z, size_t npix, int C,
                                                            const float* __restrict__ mean, const float* __restrict__ invstd,
                                                            uint32_t drop_seed, int dropout, float* __restrict__ partials) {
  __shared__ float s_red[256 * 8];
  const int groups = C >> 3;
  const int cg = threadIdx.x % groups;
  const int lanes = blockDim.x / groups;
  const int pl = threadIdx.x / groups;
  float mu[8], is[8], sa[8] = {}, sb[8] = {};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    mu[i] = mean[cg * 8 + i];
    is[i] = invstd[cg * 8 + i];
  }
  const size_t stride = size_t(gridDim.x) * lanes;
  for (size_t pix = size_t(blockIdx.x) * lanes + pl; pix < npix; pix += 2 * stride) {
    Bf8 vd[2], vz[2];
#pragma unroll
    for (int u = 0; u < 2; ++u)
      if (pix + u * stride < npix) {
        const size_t e0 = (pix + u * stride) * C + cg * 8;
        vd[u] = ld8(dA + e0);
        vz[u] = ld8(z + e0);
      }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (pix + u * stride >= npix) break;
      const size_t e0 = (pix + u * stride) * C + cg * 8;
      float d[8], f[8];
      vd[u].get(d);
      vz[u].get(f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float dv = d[i];
        if (dropout) dv = drop_keep(uint32_t(e0 + i), drop_seed) ? 2.f * dv : 0.f;
        sa[i] += dv;
        sb[i] = fmaf(dv, (f[i] - mu[i]) * is[i], sb[i]);
      }
    }
  }
  float* row = partials + size_t(blockIdx.x) * 2 * C;
  block_reduce_channels(sa, s_red, C, lanes, pl, cg, row);
  block_reduce_channels(sb, s_red, C, lanes, pl, cg, row + C);
}

// totals of pass 1 in block order -> sums[0..2C) (double, read by pass 2) and the BN parameter gradients:
// dgamma += sum dA' zhat, dbeta += sum dA'  (ONE addition per micro-batch into the gradient buffer)
__global__ void __launch_bounds__(256) bn_bwd_finalize_kernel(const float* __restrict__ partials, int nblocks, int C,
                                                              double* __restrict__ sums, float* __restrict__ dgamma,
                                                              float* __restrict__ dbeta) {
  __shared__ double s_red[256];
  const int i = blockIdx.x * 8 + (threadIdx.x & 7);     // grid = ceil(C / 8)
  const double s0 = reduce_rows_8x32(partials, nblocks, 2 * C, i, i < C, s_red);
  const double s1 = reduce_rows_8x32(partials, nblocks, 2 * C, C + i, i < C, s_red);
  if (threadIdx.x >= 8 || i >= C) return;
  sums[i] = s0;
  sums[C + i] = s1;
  dbeta[i] += float(s0);
  dgamma[i] += float(s1);
}

// ---- BN backward, pass 2 (+ ReLU backward): dpre = [z > 0] * gamma*invstd * (dA' - mean(dA') - zhat * mean(dA' zhat)) --
// also writes this block's share of the bias gradient sum(dpre) per channel (of the bf16-rounded dpre the GEMMs consume)
// to db_partials[blockIdx.x][C]; vec_partials_add_kernel adds the rows to the gradient buffer in block order.
__global__ void __launch_bounds__(256, 3) bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dA,
                                                           const __nv_bfloat16* __restrict__ z,
                                                           __nv_bfloat16* __restrict__ dpre, size_t npix, int C,
                                                           const float* __restrict__ gamma, const float* __restrict__ mean,
                                                           const float* __restrict__ invstd, const double* __restrict__ sums,
                                                           double M, uint32_t drop_seed, int dropout,
                                                           float* __restrict__ db_partials) {
  __shared__ float s_red[256 * 8];
  const int groups = C >> 3;
  const int cg = threadIdx.x % groups;
  const int lanes = blockDim.x / groups;
  const int pl = threadIdx.x / groups;
  // dz = gs * (dA' - m1 - (z - mu) * is * m2) = gs * dA' + kz * z + k0   with kz = -gs*is*m2, k0 = gs * (mu*is*m2 - m1)
  float gs[8], kz[8], k0[8], db[8] = {};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = cg * 8 + i;
    const float mu = mean[c], is = invstd[c];
    const float m1 = float(sums[c] / M), m2 = float(sums[C + c] / M);
    gs[i] = gamma[c] * is;
    kz[i] = -gs[i] * is * m2;
    k0[i] = gs[i] * (mu * is * m2 - m1);
  }
  const size_t stride = size_t(gridDim.x) * lanes;
  for (size_t pix = size_t(blockIdx.x) * lanes + pl; pix < npix; pix += 2 * stride) {
    Bf8 vd[2], vz[2];
#pragma unroll
    for (int u = 0; u < 2; ++u)
      if (pix + u * stride < npix) {
        const size_t e0 = (pix + u * stride) * C + cg * 8;
        vd[u] = ld8(dA + e0);
        vz[u] = ld8(z + e0);
      }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (pix + u * stride >= npix) break;
      const size_t e0 = (pix + u * stride) * C + cg * 8;
      float d[8], f[8], o[8];
      vd[u].get(d);
      vz[u].get(f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float dv = d[i];
        if (dropout) dv = drop_keep(uint32_t(e0 + i), drop_seed) ? 2.f * dv : 0.f;
        const float dz = fmaf(gs[i], dv, fmaf(kz[i], f[i], k0[i]));
        o[i] = f[i] > 0.f ? dz : 0.f;
      }
      Bf8 ov;
      ov.set(o);
      st8(dpre + e0, ov);
      ov.get(o);
#pragma unroll
      for (int i = 0; i < 8; ++i) db[i] += o[i];
    }
  }
  block_reduce_channels(db, s_red, C, lanes, pl, cg, db_partials + size_t(blockIdx.x) * C);
}

// dst[i] += sum over blocks (in index order) of partials[b][i]
__global__ void __launch_bounds__(256) vec_partials_add_kernel(const float* __restrict__ partials, int nblocks, int n,
                                                               float* __restrict__ dst) {
  __shared__ double s_red[256];
  const int i = blockIdx.x * 8 + (threadIdx.x & 7);     // grid = ceil(n / 8)
  const double s = reduce_rows_8x32(partials, nblocks, n, i, i < n, s_red);
  if (threadIdx.x < 8 && i < n) dst[i] += float(s);
}

// ---- MaxPool backward + skip add: dA = dSkip + route(dPooled -> first maximum of the 2x2 window of a) -----------------
__global__ void __launch_bounds__(256) pool_bwd_add_kernel(const __nv_bfloat16* __restrict__ d_skip,
                                                           const __nv_bfloat16* __restrict__ d_pooled,
                                                           const __nv_bfloat16* __restrict__ a,
                                                           __nv_bfloat16* __restrict__ dA, int NB, int H, int W, int C) {
  const int groups = C >> 3;
  const int Hp = H >> 1, Wp = W >> 1;
  const size_t total = size_t(NB) * Hp * Wp * groups;
  for (size_t gid = blockIdx.x * size_t(blockDim.x) + threadIdx.x; gid < total; gid += size_t(gridDim.x) * blockDim.x) {
    const int cg = int(gid % groups);
    size_t w = gid / groups;
    const int xp = int(w % Wp);
    w /= Wp;
    const int yp = int(w % Hp);
    const int img = int(w / Hp);
    float dp[8];
    ld8(d_pooled + ((size_t(img) * Hp + yp) * Wp + xp) * C + cg * 8).get(dp);
    float av[4][8];
    size_t e[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      e[k] = ((size_t(img) * H + 2 * yp + (k >> 1)) * W + 2 * xp + (k & 1)) * C + cg * 8;
      ld8(a + e[k]).get(av[k]);
    }
    int arg[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      int b = 0;
      float m = av[0][i];
#pragma unroll
      for (int k = 1; k < 4; ++k)
        if (av[k][i] > m) {
          m = av[k][i];
          b = k;
        }
      arg[i] = b;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float o[8];
      if (d_skip) ld8(d_skip + e[k]).get(o);
      else {
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = 0.f;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] += arg[i] == k ? dp[i] : 0.f;
      Bf8 ov;
      ov.set(o);
      st8(dA + e[k], ov);
    }
  }
}

// ---- first layer input as a 32-channel bf16 tensor (channels 3..31 zero) so its weight gradient runs on the same GEMM
template <typename InT>
__global__ void __launch_bounds__(256) rgb_to_c32_kernel(const InT* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                         size_t npix, float mul) {
  for (size_t gid = blockIdx.x * size_t(blockDim.x) + threadIdx.x; gid < npix * 4; gid += size_t(gridDim.x) * blockDim.x) {
    const size_t pix = gid >> 2;
    const int cg = int(gid & 3);
    float f[8] = {};
    if (cg == 0) {
      if constexpr (sizeof(InT) == 1) {
        f[0] = __fdiv_rn(float(in[pix * 3]), 255.f);
        f[1] = __fdiv_rn(float(in[pix * 3 + 1]), 255.f);
        f[2] = __fdiv_rn(float(in[pix * 3 + 2]), 255.f);
      } else {
        f[0] = float(in[pix * 3]) * mul;
        f[1] = float(in[pix * 3 + 1]) * mul;
        f[2] = float(in[pix * 3 + 2]) * mul;
      }
    }
    Bf8 o;
    o.set(f);
    st8(out + pix * 32 + cg * 8, o);
  }
}

// ---- optimizer over the flat parameter buffer -----------------------------------------------------------------------
// One element of the optimizer step, shared by the plain kernels below and by the fused exchange kernel (exchange.cuh);
// explicit round-to-nearest intrinsics so that every caller rounds identically whatever the compiler would contract.
// Keras RMSprop (model.py:281 'rmsprop': lr 1e-3, rho 0.9, eps 1e-7):  a = rho a + (1-rho) g^2;  p -= lr g / (sqrt(a) + eps)
// Keras Adam (model.py:280,282 comments: lr 1e-3 / 2e-3, beta_1 0.5):  lr_t = lr sqrt(1-b2^t)/(1-b1^t);
// m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= lr_t m / (sqrt(v) + eps)       (lr = lr_t for Adam; b1 = rho for RMSprop)
__device__ __forceinline__ float opt_element(bool adam, float lr, float b1, float b2, float eps, float gscale, float p,
                                             float g, float& s0, float& s1) {
  g = __fmul_rn(g, gscale);
  if (adam) {
    s0 = __fmaf_rn(b1, s0, __fmul_rn(1.f - b1, g));
    s1 = __fmaf_rn(b2, s1, __fmul_rn(__fmul_rn(1.f - b2, g), g));
    return __fsub_rn(p, __fdiv_rn(__fmul_rn(lr, s0), __fadd_rn(__fsqrt_rn(s1), eps)));
  }
  s0 = __fmaf_rn(b1, s0, __fmul_rn(__fmul_rn(1.f - b1, g), g));
  return __fsub_rn(p, __fdiv_rn(__fmul_rn(lr, g), __fadd_rn(__fsqrt_rn(s0), eps)));
}
__global__ void rmsprop_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ a, size_t n,
                               float lr, float rho, float eps, float gscale) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    float s0 = a[i], s1 = 0.f;
    p[i] = opt_element(false, lr, rho, 0.f, eps, gscale, p[i], g[i], s0, s1);
    a[i] = s0;
  }
}
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, size_t n, float lr_t, float b1, float b2, float eps, float gscale) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    float s0 = m[i], s1 = v[i];
    p[i] = opt_element(true, lr_t, b1, b2, eps, gscale, p[i], g[i], s0, s1);
    m[i] = s0;
    v[i] = s1;
  }
}

// ---- weights for the data-gradient convolution: dX = conv3x3(dpre, W') with W'[tap'][ci][co] = W[8 - tap'][ci][co] --
// Keras HWIO [tap][ci][co] already has co (the dgrad K) contiguous: only the tap order reverses. rows [r0, r0+nr) of ci.
__global__ void pack_dgrad_weights_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ dst, int cin, int cout) {
  const size_t total = size_t(9) * cin * cout;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const size_t tap = i / (size_t(cin) * cout);
    const size_t rem = i - tap * size_t(cin) * cout;
    dst[i] = __float2bfloat16_rn(w[(8 - tap) * size_t(cin) * cout + rem]);
  }
}

__global__ void fill_kernel(float* __restrict__ p, size_t n, float v) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) p[i] = v;
}

}  // namespace itmo
