// train_api.cuh — the training step of the U-Net (fit_generator's inner step, main.py:126-132,
// compiled at model.py:281 with 'rmsprop' + 'mse') behind the C ABI of include/itmo_b200.h. Included at the end of
// itmo_api.cu (same translation unit: it reuses run_conv / the tensor-map helpers).
//
// One itmo_train_backward() call = forward in training mode + backward of one micro-batch:
//   forward : conv (tcgen05 implicit GEMM, epilogue bias + ReLU) -> z   | bn_stats -> bn_finalize -> bn_apply
//             (+ Dropout 0.5 after conv2d_8 / conv2d_10, + 2x2 MaxPool) -> a (, pooled)       ... x 18 layers
//             head: 1x1 conv + sigmoid + mse, forward and backward fused (head_train_kernel)
//   backward: per layer, last to first: bn_bwd_reduce, bn_bwd_apply (+ ReLU mask, bias gradient) -> dpre,
//             wgrad (wgrad_igemm_kernel, MN-major tcgen05 GEMM over pixels, fp32 red.add into the flat gradient
//             buffer), dgrad (the forward implicit-GEMM kernel on tap-reversed weights, linear epilogue; the
//             transposed conv's data gradient is a 1x1 GEMM over the 2x2 phases read as strided views),
//             pool_bwd_add (max-pool routing + skip-connection gradient sum).
// Parameters, gradients and optimizer state live in flat fp32 buffers owned by the caller (torch tensors), so the
// data-parallel exchange is ONE all-reduce of the gradient buffer and the optimizer ONE launch.
#pragma once
#include "train_ops.cuh"
#include "wgrad_igemm.cuh"

namespace {

constexpr int kBnChannelsTotal = 7808 / 2;  // sum of cout over the 18 BN layers = 3904 (layers.txt: 7,808 non-trainable)

struct TrainLayout {
  size_t kernel[ITMO_NUM_LAYERS], bias[ITMO_NUM_LAYERS], gamma[ITMO_NUM_LAYERS], beta[ITMO_NUM_LAYERS];
  size_t kernel_count[ITMO_NUM_LAYERS];
  size_t bn[ITMO_NUM_LAYERS];  // offset of the layer's channels inside the moving-mean (and moving-variance) block
  size_t total = 0, bn_total = 0;
  TrainLayout() {
    size_t off = 0, boff = 0;
    for (int li = 0; li < ITMO_NUM_LAYERS; ++li) {
      const LayerDef& L = kLayers[li];
      const int cin = L.cin0 + L.cin1;
      const size_t kc = L.kind == kRgbConv ? size_t(27) * L.cout
                        : L.kind == kConv  ? size_t(9) * cin * L.cout
                        : L.kind == kConvT ? size_t(4) * cin * L.cout
                                           : size_t(64) * 3;
      kernel[li] = off;
      kernel_count[li] = kc;
      off += kc;
      bias[li] = off;
      off += L.cout;
      if (L.kind != kHead) {
        gamma[li] = off;
        off += L.cout;
        beta[li] = off;
        off += L.cout;
        bn[li] = boff;
        boff += L.cout;
      } else {
        gamma[li] = beta[li] = bn[li] = size_t(-1);
      }
    }
    total = off;
    bn_total = boff;
  }
};
const TrainLayout& train_layout() {
  static TrainLayout l;
  return l;
}

int ew_grid(size_t work_items) {
  return int(std::min<size_t>((work_items + 255) / 256, size_t(num_sms()) * 8));
}

// Grid of a grid-stride kernel: every SM filled with as many 256-thread blocks as its registers / shared memory allow
// (one full wave: a partial second wave would idle most of the machine for half the kernel).
template <typename K>
int full_wave_grid(K kern, size_t max_blocks) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, 0) != cudaSuccess || per_sm < 1) per_sm = 2;
  return int(std::max<size_t>(1, std::min<size_t>(size_t(num_sms()) * per_sm, max_blocks)));
}

struct WgradArgs {
  const void* a;   // shifted operand (M side)
  int ca_tensor;   // channels of the tensor behind `a`
  int ca;          // valid channels = rows written
  long long a_row_stride, a_img_stride;
  const void* b;   // N side
  int cb;
  int n, h, w;
  int conv3x3;     // 1: nine taps with 'same' padding on `a`; 0: one tap
  float* dw;
  int ld_m, ld_n, m_off, n_off;
  float* ws;          // split-K workspace (wgrad_igemm.cuh) and its capacity in bytes
  size_t ws_bytes;
};

constexpr int kPartialRows = 2048;   // rows of the per-block partial-sum workspaces (grids are clamped to this)
inline size_t wgrad_ws_bytes() { return size_t(num_sms()) * 9 * 128 * 32 * 4 * 2; }   // >= grid * tg * 128 * BN * 4 for every layer

template <int BN>
int launch_wgrad_t(const itmo::WgradParams& p, int grid, int smem_bytes, cudaStream_t stream) {
  auto kern = itmo::wgrad_igemm_kernel<BN>;
  static bool configured = false;
  if (!configured) {
    ITMO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, itmo::kSmemLimit));
    configured = true;
  }
  kern<<<grid, itmo::kWgThreads, smem_bytes, stream>>>(p);
  ITMO_CUDA(cudaGetLastError());
  return ITMO_OK;
}

int run_wgrad(const WgradArgs& a, cudaStream_t stream) {
  if (a.h % 16 || a.w % 8) return fail(ITMO_ERR_INVALID, "wgrad: spatial size %dx%d must be a multiple of 16x8", a.h, a.w);
  if (a.cb % 32 || a.cb > 512) return fail(ITMO_ERR_INVALID, "wgrad: cb %d unsupported", a.cb);
  if (a.ca_tensor % 8) return fail(ITMO_ERR_INVALID, "wgrad: ca %d unsupported", a.ca_tensor);
  const int bn = std::min(a.cb, 128);
  if (a.cb % bn) return fail(ITMO_ERR_INVALID, "wgrad: cb %d not a multiple of %d", a.cb, bn);
  itmo::WgradParams p;
  memset(&p, 0, sizeof(p));
  p.pad = a.conv3x3 ? 1 : 0;
  // Narrow 3x3 layers (<= 64 shifted channels, <= 32 gradient channels: conv2d_1, conv2d_2): one CTA holds all nine
  // taps (9 x 32 accumulator columns), so a pixel block's halo tile and gradient tile are loaded once instead of once
  // per kernel row — these layers were bound by the shared-memory traffic of those loads plus the MMA operand reads.
  static const int allow_tg9 = env_int("ITMO_WGRAD_TG9", 1);
  const bool tg9 = allow_tg9 && a.conv3x3 && bn <= 32 && a.ca <= 64;
  p.tg = a.conv3x3 ? (tg9 ? 9 : 3) : 1;
  p.n_groups = (a.conv3x3 && !tg9) ? 3 : 1;
  p.a_rows = tg9 ? 18 : 16;
  int rc = make_act_map(&p.tmA, a.a, ITMO_PREC_BF16, a.ca_tensor, a.w, a.h, a.n, 128, 8 + 2 * p.pad, p.a_rows,
                        a.a_row_stride, a.a_img_stride);
  if (rc) return rc;
  rc = make_act_map(&p.tmB, a.b, ITMO_PREC_BF16, a.cb, a.w, a.h, a.n, 128, 8, 16);
  if (rc) return rc;
  p.H = a.h;
  p.W = a.w;
  p.NB = a.n;
  p.m_blocks = (a.ca + 127) / 128;
  p.n_blocks = a.cb / bn;
  p.a_boxes = a.ca > 64 ? 2 : 1;
  static const int allow_m64 = env_int("ITMO_WGRAD_M64", 1);
  p.m64 = (allow_m64 && a.ca <= 64) ? 1 : 0;
  p.b_boxes = (bn + 63) / 64;
  p.a_box_bytes = (p.a_rows * (8 + 2 * p.pad) * itmo::kWgRowBytes + 1023) / 1024 * 1024;   // 1 KB-aligned slot (swizzle atom)
  p.stage_bytes = p.a_boxes * p.a_box_bytes + p.b_boxes * itmo::kWgBoxPixB * itmo::kWgRowBytes;
  // With one A box the MMA (always M = 128) still reads a second 64-channel group one box further on: its rows are
  // never stored, but the read must stay inside the allocation -> one box of slack after the last stage.
  const int slack = p.a_boxes == 1 ? p.a_box_bytes : 0;
  p.n_stages = std::min(itmo::kWgMaxStages, (itmo::kSmemLimit - 2048 - slack) / p.stage_bytes);
  if (p.n_stages < 2) return fail(ITMO_ERR_INVALID, "wgrad: stage does not fit");
  const int tiles = p.n_groups * p.m_blocks * p.n_blocks;
  const long long pb_total = (long long)a.n * (a.h / 16) * (a.w / 8);
  p.ksplit = int(std::max<long long>(1, std::min<long long>(pb_total, num_sms() / tiles)));
  p.ca = a.ca;
  p.dw = a.dw;
  p.ld_m = a.ld_m;
  p.ld_n = a.ld_n;
  p.m_off = a.m_off;
  p.n_off = a.n_off;
  const int smem = 1024 + p.n_stages * p.stage_bytes + slack + 256;
  const int grid = tiles * p.ksplit;
  if (p.ksplit > 1) {
    const size_t need = size_t(grid) * p.tg * 128 * bn * 4;
    if (!a.ws || a.ws_bytes < need)
      return fail(ITMO_ERR_INVALID, "wgrad: split-K workspace of %zu bytes needed, %zu given", need, a.ws_bytes);
    p.ws = a.ws;
  }
  int rc2;
  switch (bn) {
    case 32: rc2 = launch_wgrad_t<32>(p, grid, smem, stream); break;
    case 64: rc2 = launch_wgrad_t<64>(p, grid, smem, stream); break;
    case 128: rc2 = launch_wgrad_t<128>(p, grid, smem, stream); break;
    default: return fail(ITMO_ERR_INVALID, "wgrad: BN %d unsupported", bn);
  }
  if (rc2 || p.ksplit == 1) return rc2;
  // slices added in slice order, then ONE addition into the gradient (deterministic split-K)
  const size_t items = size_t(tiles) * p.tg * 128 * (bn / 4);
  itmo::wgrad_reduce_kernel<<<int(std::min<size_t>((items + 255) / 256, size_t(num_sms()) * 8)), 256, 0, stream>>>(
      p.ws, tiles, p.ksplit, p.tg, bn, p.m_blocks, p.n_blocks, p.ca, p.dw, p.ld_m, p.ld_n, p.m_off, p.n_off);
  ITMO_CUDA(cudaGetLastError());
  return ITMO_OK;
}

enum TrainPhase { kPhFwdConv = 0, kPhDgrad = 1, kPhWgrad = 2, kPhOther = 3, kPhCount = 4 };

}  // namespace

struct itmo_trainer {
  int device = 0, cap = 0, H = 0, W = 0;
  float* params = nullptr;   // caller-owned flat buffers (train_layout())
  float* grads = nullptr;
  float* opt_state = nullptr;  // 2 x total
  float* bn_moving = nullptr;  // [moving_mean (bn_total) | moving_variance (bn_total)]
  long long launches = 0;
  long long opt_steps = 0;
  bool weights_synced = false;

  __nv_bfloat16* wfwd[ITMO_NUM_LAYERS] = {};  // forward packed weights (layers 1..17)
  __nv_bfloat16* wdg[ITMO_NUM_LAYERS] = {};   // data-gradient packed weights (layers 1..17)
  __nv_bfloat16* z[ITMO_NUM_LAYERS] = {};     // ReLU(conv + b)
  __nv_bfloat16* a[ITMO_NUM_LAYERS] = {};     // BN (+dropout) output
  __nv_bfloat16* pooled[ITMO_NUM_LAYERS] = {};
  __nv_bfloat16* dskip[ITMO_NUM_LAYERS] = {};
  __nv_bfloat16* dbg_dA[ITMO_NUM_LAYERS] = {};   // ITMO_TRAIN_KEEP_GRADS: copy of each layer's dL/da (parity tests)
  __nv_bfloat16* dbg_dpre[ITMO_NUM_LAYERS] = {};
  __nv_bfloat16* x32 = nullptr;
  __nv_bfloat16* g[2] = {};     // ping-pong activation gradients
  __nv_bfloat16* gp = nullptr;  // dpre of the layer in flight (= gpb[k])
  // The weight-gradient GEMMs run on a side stream: they depend only on dpre and on forward activations, nothing on the
  // main stream's backward chain depends on them, and they are tensor / shared-memory bound while the BN-backward and
  // pooling kernels the main stream runs next are HBM bound — two kinds of kernel that share an SM well. dpre is double
  // buffered so the next layer's BN backward can write while this layer's wgrad still reads.
  __nv_bfloat16* gpb[2] = {};
  cudaStream_t s2 = nullptr;
  cudaEvent_t ev_gp[2] = {}, ev_wg[2] = {};
  bool wg_pending[2] = {false, false};
  __nv_bfloat16* gq = nullptr;  // pooled gradient
  float* bn_vec = nullptr;      // per layer: mean, invstd, scale, shift (4 x 512 each)
  double* stats = nullptr;      // per layer 2 x 512: totals of the BN-backward reductions
  float* part = nullptr;        // per-block partial sums [kPartialRows][2 x 512] (BN statistics / BN backward / head)
  float* part2 = nullptr;       // per-block partial sums of the bias gradient [kPartialRows][512]
  float* wg_ws = nullptr;       // split-K workspace of the weight-gradient GEMM
  float* consts = nullptr;      // [512 ones][512 zeros]
  double* metrics = nullptr;    // [loss sum, hits, pixels]
  uint32_t drop_seed[ITMO_NUM_LAYERS] = {};
  std::vector<void*> allocs;

  bool profiling = false;
  struct Stamp { int phase; cudaEvent_t beg, end; };
  std::vector<Stamp> stamps;
  double phase_ms[kPhCount] = {};
  long long phase_calls[kPhCount] = {};
};

namespace {

struct PhaseStamp {
  itmo_trainer* t;
  cudaStream_t s;
  cudaEvent_t end = nullptr;
  int phase, line;
  PhaseStamp(itmo_trainer* t_, int phase_, cudaStream_t s_, int line_ = 0) : t(t_), s(s_), phase(phase_), line(line_) {
    t->launches++;
    if (!t->profiling) return;
    cudaEvent_t beg;
    cudaEventCreate(&beg);
    cudaEventCreate(&end);
    cudaEventRecord(beg, s);
    t->stamps.push_back({phase, beg, end});
  }
  ~PhaseStamp() {
    if (end) cudaEventRecord(end, s);
    static const int dbg_sync = env_int("ITMO_TRAIN_SYNC", 0);   // debug: find the launch that faults
    if (dbg_sync) {
      cudaError_t e = cudaStreamSynchronize(s);
      fprintf(stderr, "itmo train: launch group %lld (phase %d, line %d): %s\n", t->launches, phase, line,
              cudaGetErrorString(e));
    }
  }
};

inline int lay_hout(const itmo_trainer* t, const LayerDef& L) { return (L.kind == kConvT ? 2 : 1) * (L.hin * t->H / 512); }
inline int lay_wout(const itmo_trainer* t, const LayerDef& L) { return (L.kind == kConvT ? 2 : 1) * (L.hin * t->W / 512); }
inline int lay_hin(const itmo_trainer* t, const LayerDef& L) { return L.hin * t->H / 512; }
inline int lay_win(const itmo_trainer* t, const LayerDef& L) { return L.hin * t->W / 512; }

template <typename T>
int tr_alloc(itmo_trainer* t, T** out, size_t count) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, std::max<size_t>(count * sizeof(T), 256));
  if (e != cudaSuccess) return fail(ITMO_ERR_NOMEM, "cudaMalloc(%zu bytes) failed: %s", count * sizeof(T), cudaGetErrorString(e));
  t->allocs.push_back(p);
  *out = (T*)p;
  return ITMO_OK;
}

float* bnv(itmo_trainer* t, int li, int which) { return t->bn_vec + (size_t(li) * 4 + which) * 512; }
double* bstats(itmo_trainer* t, int li) { return t->stats + size_t(li) * 1024; }

int train_forward(itmo_trainer* t, const void* x, int x_dtype, int nb, int flags, cudaStream_t s) {
  const TrainLayout& lay = train_layout();
  const float* ones = t->consts;
  const float* zeros = t->consts + 512;
  static const bool fuse_stats = env_int("ITMO_TRAIN_FUSE_STATS", 1) != 0;
  for (int li = 0; li < 18; ++li) {
    const LayerDef& L = kLayers[li];
    const int hi = lay_hin(t, L), wi = lay_win(t, L), ho = lay_hout(t, L), wo = lay_wout(t, L);
    int stat_rows = 0;   // rows of t->part the statistics of this layer occupy
    if (L.kind == kRgbConv) {
      PhaseStamp ps(t, kPhFwdConv, s, __LINE__);
      int rc = run_rgb_conv(ITMO_PREC_BF16, x, x_dtype, nb, hi, wi, t->params + lay.kernel[li], t->params + lay.bias[li],
                            ones, zeros, t->z[li], nullptr, s);
      if (rc) return rc;
    } else {
      ConvArgs a;
      memset(&a, 0, sizeof(a));
      a.precision = ITMO_PREC_BF16;
      a.mode = L.kind == kConvT ? itmo::kModeConvT2x2 : itmo::kModeConv3x3;
      a.src0 = L.src0_pooled ? t->pooled[L.src0] : t->a[L.src0];
      a.c0 = L.cin0;
      a.src1 = L.src1 >= 0 ? t->a[L.src1] : nullptr;
      a.c1 = L.cin1;
      a.n = nb;
      a.h = hi;
      a.w = wi;
      a.wpacked = t->wfwd[li];
      a.bias = t->params + lay.bias[li];
      a.scale = ones;
      a.shift = zeros;
      a.cout = L.cout;
      a.out = t->z[li];
      a.head_dtype = itmo::kHeadNone;
      a.stats = fuse_stats ? t->part : nullptr;   // BN batch statistics: per-CTA rows written by the conv epilogue
      a.stats_blocks = &stat_rows;
      PhaseStamp ps(t, kPhFwdConv, s, __LINE__);
      int rc = run_conv(a, s);
      if (rc) return rc;
    }
    const size_t npix = size_t(nb) * ho * wo;
    const bool dropout = (li == 7 || li == 9) && !(flags & ITMO_TRAIN_NO_DROPOUT);
    {
      PhaseStamp ps(t, kPhOther, s, __LINE__);
      static const int g_stats = full_wave_grid(itmo::bn_stats_kernel, kPartialRows);
      if (!(fuse_stats && L.kind != kRgbConv)) {
        stat_rows = int(std::min<size_t>(g_stats, (npix + 63) / 64));
        itmo::bn_stats_kernel<<<stat_rows, 256, 0, s>>>(t->z[li], npix, L.cout, t->part);
      }
      itmo::bn_finalize_kernel<<<(L.cout + 7) / 8, 256, 0, s>>>(
          t->part, stat_rows, L.cout, double(npix), 1e-3f, 0.99f, t->params + lay.gamma[li], t->params + lay.beta[li],
          bnv(t, li, 0), bnv(t, li, 1), bnv(t, li, 2), bnv(t, li, 3),
          (flags & ITMO_TRAIN_FREEZE_MOVING) ? nullptr : t->bn_moving + lay.bn[li],
          (flags & ITMO_TRAIN_FREEZE_MOVING) ? nullptr : t->bn_moving + lay.bn_total + lay.bn[li]);
      itmo::bn_apply_kernel<<<ew_grid(npix / 4 * (L.cout / 8)), 256, 0, s>>>(
          t->z[li], t->a[li], L.pool ? t->pooled[li] : nullptr, nb, ho, wo, L.cout, bnv(t, li, 2), bnv(t, li, 3),
          t->drop_seed[li], dropout ? 1 : 0);
      t->launches += 2;
    }
    ITMO_CUDA(cudaGetLastError());
  }
  return ITMO_OK;
}

int train_backward(itmo_trainer* t, const void* x, int x_dtype, const float* y, float* pred, int nb, double loss_count,
                   int flags, cudaStream_t s) {
  const TrainLayout& lay = train_layout();
  const float* ones = t->consts;
  const float* zeros = t->consts + 512;
  // head (layer 18): loss, accuracy, d a[17], head parameter gradients
  {
    const size_t npix = size_t(nb) * t->H * t->W;
    PhaseStamp ps(t, kPhOther, s, __LINE__);
    static const int g_head = full_wave_grid(itmo::head_train_kernel, kPartialRows);
    const int grid_h = int(std::min<size_t>(g_head, (npix + 31) / 32));
    itmo::head_train_kernel<<<grid_h, 256, 0, s>>>(t->a[17], y, (const float*)t->wfwd[18], t->params + lay.bias[18], npix,
                                                   float(1.0 / loss_count), t->g[0], pred, t->part);
    itmo::head_finalize_kernel<<<25, 256, 0, s>>>(t->part, grid_h, t->grads + lay.kernel[18], t->grads + lay.bias[18],
                                                 t->metrics);
    t->launches += 1;
    ITMO_CUDA(cudaGetLastError());
  }
  for (int li = 17; li >= 0; --li) {
    const LayerDef& L = kLayers[li];
    const int hi = lay_hin(t, L), wi = lay_win(t, L), ho = lay_hout(t, L), wo = lay_wout(t, L);
    const size_t npix = size_t(nb) * ho * wo;
    __nv_bfloat16* dA = t->g[(17 - li) & 1];
    __nv_bfloat16* dNext = t->g[(17 - li + 1) & 1];
    const bool dropout = (li == 7 || li == 9) && !(flags & ITMO_TRAIN_NO_DROPOUT);
    if (flags & ITMO_TRAIN_KEEP_GRADS) {
      if (!t->dbg_dA[li]) {
        int rc = tr_alloc(t, &t->dbg_dA[li], size_t(t->cap) * ho * wo * L.cout);
        if (rc == ITMO_OK) rc = tr_alloc(t, &t->dbg_dpre[li], size_t(t->cap) * ho * wo * L.cout);
        if (rc) return rc;
      }
      ITMO_CUDA(cudaMemcpyAsync(t->dbg_dA[li], dA, npix * L.cout * 2, cudaMemcpyDeviceToDevice, s));
    }
    // dpre buffer of this layer; the wgrad that last read it (two layers ago, on the side stream) must be done
    static const bool side_wgrad = env_int("ITMO_TRAIN_WGRAD_STREAM", 1) != 0;
    const int gb = side_wgrad ? ((17 - li) & 1) : 0;
    t->gp = t->gpb[gb];
    if (t->wg_pending[gb]) {
      ITMO_CUDA(cudaStreamWaitEvent(s, t->ev_wg[gb], 0));
      t->wg_pending[gb] = false;
    }
    cudaStream_t sw = side_wgrad ? t->s2 : s;   // stream of this layer's weight-gradient GEMMs
    {
      PhaseStamp ps(t, kPhOther, s, __LINE__);
      static const int g_red = full_wave_grid(itmo::bn_bwd_reduce_kernel, kPartialRows);
      static const int g_app = full_wave_grid(itmo::bn_bwd_apply_kernel, kPartialRows);
      const int grid_r = int(std::min<size_t>(g_red, (npix + 63) / 64));
      const int grid = int(std::min<size_t>(g_app, (npix + 63) / 64));
      itmo::bn_bwd_reduce_kernel<<<grid_r, 256, 0, s>>>(dA, t->z[li], npix, L.cout, bnv(t, li, 0), bnv(t, li, 1),
                                                      t->drop_seed[li], dropout ? 1 : 0, t->part);
      itmo::bn_bwd_finalize_kernel<<<(L.cout + 7) / 8, 256, 0, s>>>(t->part, grid_r, L.cout, bstats(t, li),
                                                                        t->grads + lay.gamma[li], t->grads + lay.beta[li]);
      itmo::bn_bwd_apply_kernel<<<grid, 256, 0, s>>>(dA, t->z[li], t->gp, npix, L.cout, t->params + lay.gamma[li],
                                                     bnv(t, li, 0), bnv(t, li, 1), bstats(t, li), double(npix),
                                                     t->drop_seed[li], dropout ? 1 : 0, t->part2);
      itmo::vec_partials_add_kernel<<<(L.cout + 7) / 8, 256, 0, s>>>(t->part2, grid, L.cout, t->grads + lay.bias[li]);
      t->launches += 3;
      ITMO_CUDA(cudaGetLastError());
    }
    if (flags & ITMO_TRAIN_KEEP_GRADS)
      ITMO_CUDA(cudaMemcpyAsync(t->dbg_dpre[li], t->gp, npix * L.cout * 2, cudaMemcpyDeviceToDevice, s));
    // ---- weight gradient (side stream: after dpre is complete)
    if (side_wgrad) {
      ITMO_CUDA(cudaEventRecord(t->ev_gp[gb], s));
      ITMO_CUDA(cudaStreamWaitEvent(sw, t->ev_gp[gb], 0));
    }
    if (L.kind == kRgbConv) {
      {
        PhaseStamp ps(t, kPhOther, sw, __LINE__);
        const size_t np = size_t(nb) * hi * wi;
        if (x_dtype == ITMO_DT_U8)
          itmo::rgb_to_c32_kernel<uint8_t><<<ew_grid(np * 4), 256, 0, sw>>>((const uint8_t*)x, t->x32, np, 1.f);
        else
          itmo::rgb_to_c32_kernel<float><<<ew_grid(np * 4), 256, 0, sw>>>((const float*)x, t->x32, np, 1.f);
        ITMO_CUDA(cudaGetLastError());
      }
      WgradArgs w;
      memset(&w, 0, sizeof(w));
      w.a = t->x32;
      w.ca_tensor = 32;
      w.ca = 3;
      w.b = t->gp;
      w.cb = L.cout;
      w.n = nb;
      w.h = hi;
      w.w = wi;
      w.conv3x3 = 1;
      w.dw = t->grads + lay.kernel[li];
      w.ld_m = 3;
      w.ld_n = L.cout;
      w.ws = t->wg_ws;
      w.ws_bytes = wgrad_ws_bytes();
      PhaseStamp ps(t, kPhWgrad, sw, __LINE__);
      int rc = run_wgrad(w, sw);
      if (rc) return rc;
    } else if (L.kind == kConv) {
      for (int src = 0; src < (L.cin1 > 0 ? 2 : 1); ++src) {
        WgradArgs w;
        memset(&w, 0, sizeof(w));
        w.a = src == 0 ? (L.src0_pooled ? t->pooled[L.src0] : t->a[L.src0]) : t->a[L.src1];
        w.ca_tensor = w.ca = src == 0 ? L.cin0 : L.cin1;
        w.b = t->gp;
        w.cb = L.cout;
        w.n = nb;
        w.h = hi;
        w.w = wi;
        w.conv3x3 = 1;
        w.dw = t->grads + lay.kernel[li];
        w.ld_m = L.cin0 + L.cin1;
        w.ld_n = L.cout;
        w.m_off = src == 0 ? 0 : L.cin0;
        w.ws = t->wg_ws;
        w.ws_bytes = wgrad_ws_bytes();
        PhaseStamp ps(t, kPhWgrad, sw, __LINE__);
        int rc = run_wgrad(w, sw);
        if (rc) return rc;
      }
    } else {  // kConvT: dW[a][b][co][ci] = sum G[n, 2i+a, 2j+b, co] X[n, i, j, ci]
      for (int ph = 0; ph < 2; ++ph) {
        WgradArgs w;
        memset(&w, 0, sizeof(w));
        w.a = t->gp + size_t(ph) * wo * L.cout;
        w.ca_tensor = w.ca = 2 * L.cout;
        w.a_row_stride = 2ll * wo * L.cout;
        w.a_img_stride = (long long)ho * wo * L.cout;
        w.b = t->a[L.src0];
        w.cb = L.cin0;
        w.n = nb;
        w.h = hi;
        w.w = wi;
        w.conv3x3 = 0;
        w.dw = t->grads + lay.kernel[li] + size_t(ph) * 2 * L.cout * L.cin0;
        w.ld_m = 2 * L.cout;
        w.ld_n = L.cin0;
        w.ws = t->wg_ws;
        w.ws_bytes = wgrad_ws_bytes();
        PhaseStamp ps(t, kPhWgrad, sw, __LINE__);
        int rc = run_wgrad(w, sw);
        if (rc) return rc;
      }
    }
    if (side_wgrad) {
      ITMO_CUDA(cudaEventRecord(t->ev_wg[gb], sw));
      t->wg_pending[gb] = true;
    }
    if (li == 0) break;
    // ---- data gradient
    if (L.kind == kConv) {
      for (int src = 0; src < (L.cin1 > 0 ? 2 : 1); ++src) {
        const int sl = src == 0 ? L.src0 : L.src1;
        __nv_bfloat16* dst;
        if (src == 0 && L.src0_pooled) dst = t->gq;
        else if (src == 0 && kLayers[sl].pool) dst = t->dskip[sl];
        else dst = dNext;
        ConvArgs a;
        memset(&a, 0, sizeof(a));
        a.precision = ITMO_PREC_BF16;
        a.mode = itmo::kModeConv3x3;
        a.src0 = t->gp;
        a.c0 = L.cout;
        a.n = nb;
        a.h = hi;
        a.w = wi;
        a.wpacked = t->wdg[li] + (src == 0 ? 0 : size_t(L.cin0) * L.cout);
        a.w_tap_stride_rows = L.cin0 + L.cin1;
        a.bias = zeros;
        a.scale = ones;
        a.shift = zeros;
        a.cout = src == 0 ? L.cin0 : L.cin1;
        a.out = dst;
        a.no_relu = 1;
        a.identity_epi = 1;   // zeros / ones above: plain store of the accumulator
        a.head_dtype = itmo::kHeadNone;
        {
          PhaseStamp ps(t, kPhDgrad, s, __LINE__);
          int rc = run_conv(a, s);
          if (rc) return rc;
        }
        if (src == 0 && L.src0_pooled) {
          const LayerDef& S = kLayers[sl];
          const int hs = lay_hout(t, S), ws = lay_wout(t, S);
          PhaseStamp ps(t, kPhOther, s, __LINE__);
          itmo::pool_bwd_add_kernel<<<ew_grid(size_t(nb) * (hs / 2) * (ws / 2) * (S.cout / 8)), 256, 0, s>>>(
              t->dskip[sl], t->gq, t->a[sl], dNext, nb, hs, ws, S.cout);
          ITMO_CUDA(cudaGetLastError());
        }
      }
    } else {  // kConvT: dX[i,j,ci] = sum_{a,b,co} G[2i+a, 2j+b, co] W[a][b][co][ci] -> 1x1 GEMM, K = 4*cout in two row views
      ConvArgs a;
      memset(&a, 0, sizeof(a));
      a.precision = ITMO_PREC_BF16;
      a.mode = itmo::kModeConv3x3;
      a.one_tap = 1;
      a.src0 = t->gp;
      a.c0 = 2 * L.cout;
      a.src1 = t->gp + size_t(wo) * L.cout;
      a.c1 = 2 * L.cout;
      a.row_stride[0] = a.row_stride[1] = 2ll * wo * L.cout;
      a.img_stride[0] = a.img_stride[1] = (long long)ho * wo * L.cout;
      a.n = nb;
      a.h = hi;
      a.w = wi;
      a.wpacked = t->wdg[li];
      a.bias = zeros;
      a.scale = ones;
      a.shift = zeros;
      a.cout = L.cin0;
      a.out = dNext;
      a.no_relu = 1;
      a.identity_epi = 1;   // zeros / ones above: plain store of the accumulator
      a.head_dtype = itmo::kHeadNone;
      PhaseStamp ps(t, kPhDgrad, s, __LINE__);
      int rc = run_conv(a, s);
      if (rc) return rc;
    }
  }
  // join: everything the caller launches next on `s` (the next micro-batch's forward overwrites the activations the
  // weight-gradient GEMMs read; update / exchange / zero_grad touch the gradient buffer) follows the side stream's work
  for (int gb = 0; gb < 2; ++gb)
    if (t->wg_pending[gb]) {
      ITMO_CUDA(cudaStreamWaitEvent(s, t->ev_wg[gb], 0));
      t->wg_pending[gb] = false;
    }
  return ITMO_OK;
}

}  // namespace

extern "C" {

long long itmo_train_param_count(void) { return (long long)train_layout().total; }
long long itmo_train_bn_count(void) { return (long long)train_layout().bn_total; }

int itmo_train_layout(int layer, int which, long long* offset, long long* count) {
  if (layer < 0 || layer >= ITMO_NUM_LAYERS || !offset || !count) return fail(ITMO_ERR_INVALID, "bad argument");
  const TrainLayout& lay = train_layout();
  const LayerDef& L = kLayers[layer];
  switch (which) {
    case 0: *offset = (long long)lay.kernel[layer]; *count = (long long)lay.kernel_count[layer]; return ITMO_OK;
    case 1: *offset = (long long)lay.bias[layer]; *count = L.cout; return ITMO_OK;
    case 2:
    case 3:
    case 4:
      if (L.kind == kHead) return fail(ITMO_ERR_INVALID, "the head has no BatchNormalization");
      *offset = (long long)(which == 2 ? lay.gamma[layer] : which == 3 ? lay.beta[layer] : lay.bn[layer]);
      *count = L.cout;
      return ITMO_OK;
  }
  return fail(ITMO_ERR_INVALID, "which must be 0..4");
}

void itmo_trainer_destroy(itmo_trainer* t) {
  if (!t) return;
  cudaSetDevice(t->device);
  cudaDeviceSynchronize();
  for (void* p : t->allocs) cudaFree(p);
  if (t->s2) cudaStreamDestroy(t->s2);
  for (int i = 0; i < 2; ++i) {
    if (t->ev_gp[i]) cudaEventDestroy(t->ev_gp[i]);
    if (t->ev_wg[i]) cudaEventDestroy(t->ev_wg[i]);
  }
  for (auto& st : t->stamps) {
    cudaEventDestroy(st.beg);
    cudaEventDestroy(st.end);
  }
  delete t;
}

int itmo_trainer_create(itmo_trainer** out, int device, int micro_batch, int h, int w, float* params_dev,
                        float* grads_dev, float* opt_state_dev, float* bn_moving_dev) {
  if (!out) return fail(ITMO_ERR_INVALID, "out is null");
  *out = nullptr;
  if (!params_dev || !grads_dev || !opt_state_dev || !bn_moving_dev) return fail(ITMO_ERR_INVALID, "null buffer");
  if (micro_batch <= 0) return fail(ITMO_ERR_INVALID, "micro_batch must be positive");
  if (h <= 0 || w <= 0 || h % 256 || w % 128)
    return fail(ITMO_ERR_INVALID, "image size %dx%d: height must be a multiple of 256 and width of 128", h, w);
  ITMO_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  ITMO_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(ITMO_ERR_CUDA, "device %d is sm_%d%d; libitmo_b200 only contains sm_100a code", device, prop.major,
                prop.minor);
  itmo_trainer* t = new itmo_trainer();
  t->device = device;
  t->cap = micro_batch;
  t->H = h;
  t->W = w;
  t->params = params_dev;
  t->grads = grads_dev;
  t->opt_state = opt_state_dev;
  t->bn_moving = bn_moving_dev;
  int rc = ITMO_OK;
  size_t gmax = 0, qmax = 0;
  for (int li = 0; li < 18 && rc == ITMO_OK; ++li) {
    const LayerDef& L = kLayers[li];
    const int cin = L.cin0 + L.cin1;
    const size_t out_elems = size_t(micro_batch) * lay_hout(t, L) * lay_wout(t, L) * L.cout;
    gmax = std::max(gmax, out_elems);
    rc = tr_alloc(t, &t->z[li], out_elems);
    if (rc == ITMO_OK) rc = tr_alloc(t, &t->a[li], out_elems);
    if (rc == ITMO_OK && L.pool) {
      rc = tr_alloc(t, &t->pooled[li], out_elems / 4);
      if (rc == ITMO_OK) rc = tr_alloc(t, &t->dskip[li], out_elems);
      qmax = std::max(qmax, out_elems / 4);
    }
    if (rc == ITMO_OK && L.kind != kRgbConv) {
      const size_t wn = size_t(L.kind == kConv ? 9 : 4) * cin * L.cout;
      rc = tr_alloc(t, &t->wfwd[li], wn);
      if (rc == ITMO_OK) rc = tr_alloc(t, &t->wdg[li], wn);
    }
  }
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->x32, size_t(micro_batch) * h * w * 32);
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->g[0], gmax);
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->g[1], gmax);
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->gpb[0], gmax);
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->gpb[1], gmax);
  t->gp = t->gpb[0];
  if (rc == ITMO_OK) {
    cudaError_t es = cudaStreamCreateWithFlags(&t->s2, cudaStreamNonBlocking);
    for (int i = 0; i < 2 && es == cudaSuccess; ++i) {
      es = cudaEventCreateWithFlags(&t->ev_gp[i], cudaEventDisableTiming);
      if (es == cudaSuccess) es = cudaEventCreateWithFlags(&t->ev_wg[i], cudaEventDisableTiming);
    }
    if (es != cudaSuccess) rc = fail(ITMO_ERR_CUDA, "side stream setup failed: %s", cudaGetErrorString(es));
  }
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->gq, qmax);
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->bn_vec, size_t(ITMO_NUM_LAYERS) * 4 * 512);
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->stats, size_t(ITMO_NUM_LAYERS) * 1024);
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->part, size_t(kPartialRows) * 1024);
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->part2, size_t(kPartialRows) * 512);
  if (rc == ITMO_OK) rc = tr_alloc(t, (char**)&t->wg_ws, wgrad_ws_bytes());
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->consts, size_t(1024));
  if (rc == ITMO_OK) rc = tr_alloc(t, &t->metrics, size_t(4));
  if (rc == ITMO_OK) rc = tr_alloc(t, (float**)&t->wfwd[18], size_t(192));  // head weights [3][64] fp32
  if (rc != ITMO_OK) {
    itmo_trainer_destroy(t);
    return rc;
  }
  itmo::fill_kernel<<<2, 256>>>(t->consts, 512, 1.f);
  itmo::fill_kernel<<<2, 256>>>(t->consts + 512, 512, 0.f);
  cudaMemset(t->metrics, 0, sizeof(double) * 4);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    itmo_trainer_destroy(t);
    return fail(ITMO_ERR_CUDA, "trainer setup failed: %s", cudaGetErrorString(e));
  }
  *out = t;
  return ITMO_OK;
}

/* Re-derives the tensor-core operand copies (bf16 K-major forward weights, tap-reversed data-gradient weights, the
 * transposed head) from the flat fp32 parameters. Called by itmo_train_update; call it after writing params directly. */
int itmo_trainer_sync_weights(itmo_trainer* t, void* stream) {
  if (!t) return fail(ITMO_ERR_INVALID, "null trainer");
  ITMO_CUDA(cudaSetDevice(t->device));
  cudaStream_t s = (cudaStream_t)stream;
  const TrainLayout& lay = train_layout();
  for (int li = 1; li < 18; ++li) {
    const LayerDef& L = kLayers[li];
    const int cin = L.cin0 + L.cin1;
    const float* k = t->params + lay.kernel[li];
    int rc = pack_weights(ITMO_PREC_BF16, L.kind == kConv ? itmo::kModeConv3x3 : itmo::kModeConvT2x2, k, t->wfwd[li], cin,
                          L.cout, s);
    if (rc) return rc;
    const size_t total = lay.kernel_count[li];
    const int grid = int(std::min<size_t>((total + 255) / 256, 4096));
    if (L.kind == kConv) {
      itmo::pack_dgrad_weights_kernel<<<grid, 256, 0, s>>>(k, t->wdg[li], cin, L.cout);
    } else {
      // Keras [a][b][co][ci] read as HWIO with one tap: K = (a,b,co), rows = ci
      pack_conv_weights_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(k, t->wdg[li], 1, 4 * L.cout, cin, ITMO_PREC_BF16, 1);
    }
    t->launches += 2;
  }
  transpose_head_kernel<<<1, 192, 0, s>>>(t->params + lay.kernel[18], (float*)t->wfwd[18]);
  t->launches += 1;
  ITMO_CUDA(cudaGetLastError());
  t->weights_synced = true;
  return ITMO_OK;
}

int itmo_train_zero_grad(itmo_trainer* t, void* stream) {
  if (!t) return fail(ITMO_ERR_INVALID, "null trainer");
  ITMO_CUDA(cudaSetDevice(t->device));
  ITMO_CUDA(cudaMemsetAsync(t->grads, 0, train_layout().total * 4, (cudaStream_t)stream));
  return ITMO_OK;
}

int itmo_train_backward(itmo_trainer* t, const void* x_dev, int x_dtype, const float* y_dev, float* pred_dev_or_null,
                        int n, double loss_count, unsigned seed, unsigned flags, void* stream) {
  if (!t || !x_dev || !y_dev) return fail(ITMO_ERR_INVALID, "null argument");
  if (n <= 0 || n > t->cap) return fail(ITMO_ERR_INVALID, "micro-batch %d outside 1..%d", n, t->cap);
  if (x_dtype != ITMO_DT_U8 && x_dtype != ITMO_DT_F32) return fail(ITMO_ERR_INVALID, "bad x_dtype");
  if (!(loss_count > 0)) return fail(ITMO_ERR_INVALID, "loss_count must be positive");
  if (!t->weights_synced) return fail(ITMO_ERR_STATE, "itmo_train_backward before itmo_trainer_sync_weights");
  ITMO_CUDA(cudaSetDevice(t->device));
  cudaStream_t s = (cudaStream_t)stream;
  for (int li = 0; li < ITMO_NUM_LAYERS; ++li) t->drop_seed[li] = seed * 0x9E3779B9u + uint32_t(li) * 0x85EBCA6Bu + 1u;
  int rc = train_forward(t, x_dev, x_dtype, n, int(flags), s);
  if (rc) return rc;
  return train_backward(t, x_dev, x_dtype, y_dev, pred_dev_or_null, n, loss_count, int(flags), s);
}

int itmo_train_dropout_seed(unsigned seed, int layer, unsigned* out) {
  if (!out) return fail(ITMO_ERR_INVALID, "null argument");
  *out = seed * 0x9E3779B9u + uint32_t(layer) * 0x85EBCA6Bu + 1u;
  return ITMO_OK;
}

int itmo_train_metrics(itmo_trainer* t, double* loss_sum, double* hits, int reset, void* stream) {
  if (!t) return fail(ITMO_ERR_INVALID, "null trainer");
  ITMO_CUDA(cudaSetDevice(t->device));
  cudaStream_t s = (cudaStream_t)stream;
  double h[2];
  ITMO_CUDA(cudaMemcpyAsync(h, t->metrics, sizeof(h), cudaMemcpyDeviceToHost, s));
  if (reset) ITMO_CUDA(cudaMemsetAsync(t->metrics, 0, sizeof(double) * 4, s));
  ITMO_CUDA(cudaStreamSynchronize(s));
  if (loss_sum) *loss_sum = h[0];
  if (hits) *hits = h[1];
  return ITMO_OK;
}

int itmo_train_update(itmo_trainer* t, int optimizer, float lr, float beta1_or_rho, float beta2, float eps,
                      float grad_scale, void* stream) {
  if (!t) return fail(ITMO_ERR_INVALID, "null trainer");
  ITMO_CUDA(cudaSetDevice(t->device));
  cudaStream_t s = (cudaStream_t)stream;
  const size_t n = train_layout().total;
  const int grid = num_sms() * 8;
  t->opt_steps++;
  if (optimizer == ITMO_OPT_RMSPROP) {
    itmo::rmsprop_kernel<<<grid, 256, 0, s>>>(t->params, t->grads, t->opt_state, n, lr, beta1_or_rho, eps, grad_scale);
  } else if (optimizer == ITMO_OPT_ADAM) {
    const double b1t = pow(double(beta1_or_rho), double(t->opt_steps)), b2t = pow(double(beta2), double(t->opt_steps));
    const float lr_t = float(double(lr) * sqrt(1.0 - b2t) / (1.0 - b1t));
    itmo::adam_kernel<<<grid, 256, 0, s>>>(t->params, t->grads, t->opt_state, t->opt_state + n, n, lr_t, beta1_or_rho,
                                           beta2, eps, grad_scale);
  } else {
    return fail(ITMO_ERR_INVALID, "unknown optimizer %d", optimizer);
  }
  ITMO_CUDA(cudaGetLastError());
  t->launches++;
  return itmo_trainer_sync_weights(t, stream);
}

int itmo_train_set_step(itmo_trainer* t, long long steps) {
  if (!t) return fail(ITMO_ERR_INVALID, "null trainer");
  if (steps < 0) return fail(ITMO_ERR_INVALID, "negative step count");
  t->opt_steps = steps;
  return ITMO_OK;
}

long long itmo_train_get_step(const itmo_trainer* t) { return t ? t->opt_steps : 0; }

int itmo_train_read_activation(itmo_trainer* t, int layer, int which, int n, float* dst_host, size_t cap_elems,
                               size_t* n_elems) {
  if (!t || !dst_host) return fail(ITMO_ERR_INVALID, "null argument");
  if (layer < 0 || layer >= 18) return fail(ITMO_ERR_INVALID, "layer out of range");
  const LayerDef& L = kLayers[layer];
  const __nv_bfloat16* src = which == 0   ? t->z[layer]
                             : which == 1 ? t->a[layer]
                             : which == 2 ? t->pooled[layer]
                             : which == 3 ? t->dbg_dA[layer]
                                          : t->dbg_dpre[layer];
  if (!src) return fail(ITMO_ERR_STATE, "tensor not materialised");
  size_t cnt = size_t(n) * lay_hout(t, L) * lay_wout(t, L) * L.cout;
  if (which == 2) cnt /= 4;
  if (n_elems) *n_elems = cnt;
  if (cnt > cap_elems) return fail(ITMO_ERR_INVALID, "destination too small");
  ITMO_CUDA(cudaSetDevice(t->device));
  ITMO_CUDA(cudaDeviceSynchronize());
  float* tmp = nullptr;
  ITMO_CUDA(cudaMalloc((void**)&tmp, cnt * 4));
  half_to_f32_kernel<__nv_bfloat16><<<1024, 256>>>(src, tmp, cnt);
  cudaError_t e = cudaMemcpy(dst_host, tmp, cnt * 4, cudaMemcpyDeviceToHost);
  cudaFree(tmp);
  if (e != cudaSuccess) return fail(ITMO_ERR_CUDA, "read failed: %s", cudaGetErrorString(e));
  return ITMO_OK;
}

int itmo_train_set_profiling(itmo_trainer* t, int enable) {
  if (!t) return fail(ITMO_ERR_INVALID, "null trainer");
  t->profiling = enable != 0;
  return ITMO_OK;
}

int itmo_train_phase_times(itmo_trainer* t, double* ms_sum, long long* calls, int reset) {
  if (!t) return fail(ITMO_ERR_INVALID, "null trainer");
  ITMO_CUDA(cudaSetDevice(t->device));
  for (auto& st : t->stamps) {
    ITMO_CUDA(cudaEventSynchronize(st.end));
    float ms = 0.f;
    ITMO_CUDA(cudaEventElapsedTime(&ms, st.beg, st.end));
    t->phase_ms[st.phase] += ms;
    t->phase_calls[st.phase]++;
    cudaEventDestroy(st.beg);
    cudaEventDestroy(st.end);
  }
  t->stamps.clear();
  for (int i = 0; i < kPhCount; ++i) {
    if (ms_sum) ms_sum[i] = t->phase_ms[i];
    if (calls) calls[i] = t->phase_calls[i];
    if (reset) {
      t->phase_ms[i] = 0;
      t->phase_calls[i] = 0;
    }
  }
  return ITMO_OK;
}

long long itmo_train_launch_count(const itmo_trainer* t) { return t ? t->launches : 0; }

/* Stateless weight gradient of one ConvBN layer (parity tests): x [n,h,w,cin] bf16, g [n,h,w,cout] bf16 (the gradient
 * at the convolution output), dw [3][3][cin][cout] fp32 in Keras HWIO layout, ACCUMULATED into. */
int itmo_wgrad3x3(const void* x, int cin, const void* g, int cout, int n, int h, int w, float* dw, void* stream) {
  if (!x || !g || !dw) return fail(ITMO_ERR_INVALID, "null argument");
  WgradArgs a;
  memset(&a, 0, sizeof(a));
  a.a = x;
  a.ca_tensor = a.ca = cin;
  a.b = g;
  a.cb = cout;
  a.n = n;
  a.h = h;
  a.w = w;
  a.conv3x3 = 1;
  a.dw = dw;
  a.ld_m = cin;
  a.ld_n = cout;
  cudaStream_t s = (cudaStream_t)stream;
  ITMO_CUDA(cudaMallocAsync((void**)&a.ws, wgrad_ws_bytes(), s));
  a.ws_bytes = wgrad_ws_bytes();
  const int rc = run_wgrad(a, s);
  cudaFreeAsync(a.ws, s);
  return rc;
}

/* Stateless data gradient of one ConvBN layer: g [n,h,w,cout] bf16, kernel fp32 HWIO [3][3][cin][cout] (device),
 * dx [n,h,w,cin] bf16 = full correlation of g with the kernel. */
int itmo_dgrad3x3(const void* g, int cout, int n, int h, int w, const float* kernel_hwio, int cin, void* dx,
                  void* stream) {
  if (!g || !kernel_hwio || !dx) return fail(ITMO_ERR_INVALID, "null argument");
  cudaStream_t s = (cudaStream_t)stream;
  __nv_bfloat16* wp = nullptr;
  float* cz = nullptr;
  const size_t total = size_t(9) * cin * cout;
  ITMO_CUDA(cudaMallocAsync((void**)&wp, total * 2, s));
  ITMO_CUDA(cudaMallocAsync((void**)&cz, 1024 * 4, s));
  itmo::pack_dgrad_weights_kernel<<<int(std::min<size_t>((total + 255) / 256, 4096)), 256, 0, s>>>(kernel_hwio, wp, cin, cout);
  itmo::fill_kernel<<<2, 256, 0, s>>>(cz, 512, 1.f);
  itmo::fill_kernel<<<2, 256, 0, s>>>(cz + 512, 512, 0.f);
  ConvArgs a;
  memset(&a, 0, sizeof(a));
  a.precision = ITMO_PREC_BF16;
  a.mode = itmo::kModeConv3x3;
  a.src0 = g;
  a.c0 = cout;
  a.n = n;
  a.h = h;
  a.w = w;
  a.wpacked = wp;
  a.bias = cz + 512;
  a.scale = cz;
  a.shift = cz + 512;
  a.cout = cin;
  a.out = dx;
  a.no_relu = 1;
  a.identity_epi = 1;   // zeros / ones above: plain store of the accumulator
  a.head_dtype = itmo::kHeadNone;
  int rc = run_conv(a, s);
  cudaFreeAsync(wp, s);
  cudaFreeAsync(cz, s);
  return rc;
}

}  // extern "C"
