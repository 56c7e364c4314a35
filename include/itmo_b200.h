/* itmo_b200.h — C ABI of libitmo_b200.so: the B200-native replacement for the one hot path of
 * LynxHack/CNN-ITMO, the Keras U-Net forward pass that predict.py runs.
 *
 * The reference has no FFI of its own: its boundary is the Keras Model object
 *   model.predict(X)            predict.py:62, main.py:58
 *   load_model(path)            predict.py:24
 *   U_net(...)                  model.py:204-289   (ConvBN :195-196, ConvBNTranspose :199-200)
 * The Python shim in cnn-itmo_b200/model.py mirrors that object API and binds these entry
 * points through ctypes. Signatures are plain C: pointers, ints, no torch
 * or C++ types. Every function returns 0 on success or a negative itmo_status; a human readable
 * message for the calling thread's last failure is available from itmo_last_error(). Nothing
 * throws across the ABI. Device pointers are owned by the caller; a context owns the packed
 * weights, the activation arena and its stream-ordered workspace. A context is bound to one
 * CUDA device and is not thread-safe.
 */
#ifndef ITMO_B200_H_
#define ITMO_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct itmo_ctx itmo_ctx;

enum itmo_status {
  ITMO_OK = 0,
  ITMO_ERR_INVALID = -1, /* bad argument (shape, dtype, null pointer) */
  ITMO_ERR_CUDA = -2,    /* CUDA runtime / driver error, see itmo_last_error() */
  ITMO_ERR_STATE = -3,   /* e.g. forward before weights were loaded */
  ITMO_ERR_NOMEM = -4
};

/* arithmetic mode of the tensor-core layers (fp32 accumulation in TMEM in both) */
enum itmo_precision {
  ITMO_PREC_TF32 = 0, /* fp32 activations in HBM rounded to TF32 (rna), tcgen05 kind::tf32, ONE product per MAC. A
                         reduced-precision mode: its bar is PSNR >= 45 dB (measured 50.7 dB, max-abs 2e-2 on the
                         calibrated checkpoint) — the north-star's "max abs <= 1e-3 in TF32/fp32 mode" is met by
                         ITMO_PREC_FP32 (3xTF32), not by this one */
  ITMO_PREC_BF16 = 1, /* bf16 activations in HBM, tcgen05 kind::f16 (bf16 operands) */
  ITMO_PREC_FP32 = 2, /* fp32 activations as TF32 head + residual, 3xTF32 products: x*w_hi + x*w_lo + x_lo*w_hi
                         (fp32-accurate; parity: max-abs <= 1e-3 against the fp32 reference) */
  ITMO_PREC_FP16 = 3  /* fp16 activations in HBM (saturating, counted: itmo_saturation_count), tcgen05 kind::f16 (fp16
                         operands): same speed as bf16 with 3 more mantissa bits (parity: PSNR >= 45 dB). The DEFAULT
                         of the Python shim, predict CLI and bench.py: bf16 reaches only ~32 dB on the calibrated
                         checkpoint and is opt-in */
};

enum itmo_dtype { ITMO_DT_U8 = 0, ITMO_DT_F32 = 1 };

enum itmo_flags {
  ITMO_FLAG_NONE = 0
};

#define ITMO_NUM_LAYERS 19 /* 14 ConvBN + 4 ConvBNTranspose + 1x1 head, in execution order */
#define ITMO_IMG 512       /* native input size, model.py:204, layers.txt:4 */

/* One layer's parameters in the layout Keras stores them in the HDF5 checkpoint
 * (/model_weights/<layer>/<layer>/{kernel:0,bias:0}, BN {gamma:0,beta:0,moving_mean:0,
 * moving_variance:0}), host memory, float32:
 *   3x3 conv   kernel [3][3][Cin][Cout]      (model.py:196)
 *   transposed kernel [2][2][Cout][Cin]      (model.py:200)
 *   head       kernel [1][1][64][3]          (model.py:276)  gamma..var = NULL
 */
typedef struct itmo_layer_weights {
  const float* kernel;
  const float* bias;
  const float* gamma;
  const float* beta;
  const float* moving_mean;
  const float* moving_variance;
} itmo_layer_weights;

/* Layer order (index -> Keras layer of layers.txt):
 *  0..9  conv2d_1..conv2d_10 (encoder + bottleneck), 10 conv2d_transpose_1, 11 conv2d_11,
 *  12 conv2d_transpose_2, 13 conv2d_12, 14 conv2d_transpose_3, 15 conv2d_13,
 *  16 conv2d_transpose_4, 17 conv2d_14, 18 conv2d_15 (1x1 sigmoid head). */

const char* itmo_last_error(void);
/* Library/ABI version, bumped when a signature changes. */
int itmo_abi_version(void);

/* Create a context on CUDA device `device`. `max_batch` is the number of images the activation
 * arena holds at once; itmo_forward() splits larger batches into arena-sized chunks. */
int itmo_create(itmo_ctx** out, int device, int precision, int max_batch);
void itmo_destroy(itmo_ctx* ctx);

/* Replaces load_model()'s weight loading (predict.py:24): repacks the 19 layers to K-major
 * tensor-core tiles and precomputes the BN affine  s = gamma/sqrt(var+eps), t = beta - mean*s. */
int itmo_load_weights(itmo_ctx* ctx, const itmo_layer_weights* layers, int n_layers, float bn_eps);

/* Replaces model.predict(X) (predict.py:62) on DEVICE buffers.
 *  in : [n,512,512,3] NHWC, ITMO_DT_U8 (raw pixels; the /255 of predict.py:59 is fused) or
 *       ITMO_DT_F32 (already in [0,1])
 *  out: [n,512,512,3] NHWC, ITMO_DT_F32 (sigmoid output) or ITMO_DT_U8 (trunc(p*255), predict.py:64)
 * Asynchronous on `stream` (a cudaStream_t, may be NULL for the default stream). */
int itmo_forward(itmo_ctx* ctx, const void* in_dev, int in_dtype, void* out_dev, int out_dtype, int n,
                 unsigned flags, void* stream);

/* Same call with HOST buffers: pinned staging, H2D, forward, D2H, synchronises before return.
 * This is the end-to-end path model.predict() uses. */
int itmo_predict_host(itmo_ctx* ctx, const void* in_host, int in_dtype, void* out_host, int out_dtype, int n,
                      unsigned flags);

/* CUDA-graph replay for itmo_predict_host: the 18 kernel launches of one arena chunk are captured once per (chunk size,
 * dtypes, staging slot) and replayed with one cudaGraphLaunch — for the reference's batch-1 loop (predict.py:50-62) the
 * launches are host-latency-bound. Graphs are dropped when weights are reloaded. Off by default at the ABI; the Python
 * shim turns it on. Has no effect while itmo_set_profiling is on. */
int itmo_set_graph_capture(itmo_ctx* ctx, int enable);

/* Pinned (page-locked) host memory for itmo_predict_host's in/out buffers: copies from/to it are asynchronous
 * and overlap the kernels. Free with itmo_host_free(). */
int itmo_host_alloc(void** out, size_t bytes);
void itmo_host_free(void* p);

/* Per-layer device timing. While enabled, every layer launch of itmo_forward / itmo_predict_host is bracketed by
 * a CUDA event pair on its stream. itmo_layer_times() waits for the recorded events and returns, per layer of
 * ITMO_NUM_LAYERS, the accumulated milliseconds and the number of launches since the last reset (either output
 * may be NULL). The fused head (layer 18) reports 0. */
int itmo_set_profiling(itmo_ctx* ctx, int enable);
int itmo_layer_times(itmo_ctx* ctx, double* ms_sum, long long* calls, int reset);

/* Number of kernels itmo launched on this context since creation (for bench.py's gpu_launches). */
long long itmo_launch_count(const itmo_ctx* ctx);

/* fp16 storage mode: activations are packed to fp16 with saturation (+-65504). This returns how many epilogue store
 * groups (one thread x 32 channels) clamped at least one value since context creation / the last reset, so a
 * checkpoint whose activations leave fp16's range cannot clip silently (always 0 in the other modes). Synchronises
 * the device. Model.predict / predict_u8 check it after every call and warn. */
int itmo_saturation_count(itmo_ctx* ctx, unsigned long long* count, int reset);

/* Read back one intermediate activation of the last forward chunk (debug / per-layer parity).
 * `layer` indexes ITMO_NUM_LAYERS; `which` 0 = the layer output, 1 = its 2x2 max-pooled copy.
 * Writes float32 NHWC to dst_host (converted from bf16 in BF16 mode); returns element count via
 * *n_elems. The fused head (layer 17's activation) is not materialised and returns ITMO_ERR_STATE. */
int itmo_read_activation(itmo_ctx* ctx, int layer, int which, float* dst_host, size_t cap_elems, size_t* n_elems);

/* ---- stateless per-layer operators (BASELINE config 2: per-layer sweep) -------------------------
 * All pointers are device pointers; activations NHWC in the element type of `precision`
 * (bf16 or fp32); kernel in Keras layout, float32, DEVICE memory; bias/scale/shift float32 [cout].
 * ConvBN (model.py:195-196): out = max(conv3x3_same(concat(src0,src1)) + bias, 0)*scale + shift,
 * optional 2x2 max-pooled copy (model.py:210), src1 may be NULL (c1 = 0). */
int itmo_conv3x3_bn(int precision, const void* src0, int c0, const void* src1, int c1, int n, int h, int w,
                    const float* kernel_hwio, const float* bias, const float* scale, const float* shift, int cout,
                    void* out, void* pooled_or_null, void* stream);
/* ConvBNTranspose (model.py:199-200): kernel [2][2][cout][cin], out is [n,2h,2w,cout]. */
int itmo_convT2x2_bn(int precision, const void* src, int cin, int n, int h, int w, const float* kernel_hwoi,
                     const float* bias, const float* scale, const float* shift, int cout, void* out, void* stream);
/* Last ConvBN with the 1x1 sigmoid head fused (model.py:262 + :276): cout is 64, head_kernel is the
 * Keras [1][1][64][3] kernel, out is [n,h,w,3] in out_dtype (ITMO_DT_F32 / ITMO_DT_U8). */
int itmo_conv3x3_bn_head(int precision, const void* src0, int c0, const void* src1, int c1, int n, int h, int w,
                         const float* kernel_hwio, const float* bias, const float* scale, const float* shift,
                         const float* head_kernel, const float* head_bias, void* out, int out_dtype, void* stream);
/* First layer (Cin = 3) from raw input, in_dtype ITMO_DT_U8 or ITMO_DT_F32. */
int itmo_conv3x3_bn_rgb(int precision, const void* in, int in_dtype, int n, int h, int w, const float* kernel_hwio,
                        const float* bias, const float* scale, const float* shift, int cout, void* out,
                        void* stream);

/* ---- frames larger than the native input (BASELINE config 4) and the inverse-Reinhard output stage ------------
 * The reference has no tiling (fixed 512x512 Input, model.py:204,207). A frame [h,w,3] (both sides multiples of 16,
 * >= 512) is cut into 512x512 tiles that all lie inside the frame, start on the 16-pixel grid of the four 2x2 pools
 * and overlap by >= 96 pixels (the receptive field of this graph reaches 92), so the stitched result equals the
 * fully-convolutional network on the whole frame. */
enum itmo_out_kind {
  ITMO_OUT_U8 = 0,            /* trunc(p*255) uint8, what predict.py:64 writes */
  ITMO_OUT_F32 = 1,           /* sigmoid output float32 */
  ITMO_OUT_HDR_INTENDED = 2,  /* float32 linear HDR: m-files/inverse_Reinhard.m with its sign error repaired */
  ITMO_OUT_HDR_AS_WRITTEN = 3, /* float32: m-files/inverse_Reinhard.m:2-23 literally (degenerate) */
  ITMO_OUT_HDR_INTENDED_F16 = 4 /* the repaired inverse as IEEE half floats (OpenEXR's HALF pixel type), clamped to 65504:
                                   half the bytes to bring back to the host; the arithmetic is the float64 one of kind 2 */
};
/* Tile plan along one axis: returns the tile count, fills origin / valid range [lo,hi) per tile (up to cap). */
int itmo_tile_axis_plan(int extent, int* origin, int* valid_lo, int* valid_hi, int cap);
/* frames_dev: [n_frames,h,w,3] uint8 on the device; out_dev: same shape, uint8, float32 or half by out_kind.
 * `key` is the geometric-mean luminance the repaired inverse normalises to (Reinhard.m:20 uses 0.18). */
int itmo_predict_frames(itmo_ctx* ctx, const void* frames_dev, int n_frames, int h, int w, void* out_dev, int out_kind,
                        double key, void* stream);
/* Replaces m-files/inverse_Reinhard.m on its own: uint8 CNN output frames -> float32 HDR.
 * stats_scratch_dev: 32 bytes per frame of device scratch. */
int itmo_inverse_reinhard(const void* frames_u8_dev, void* hdr_f32_dev, int n_frames, int h, int w, int as_written,
                          double key, void* stats_scratch_dev, void* stream);

/* ---- training step (BASELINE config 5) --------------------------------------------------------------
 * Replaces the step Keras runs inside model.fit_generator(...) (main.py:126-132) for the model compiled at model.py:281
 * (optimizer 'rmsprop', loss 'mse', metrics ['accuracy']): forward in training mode (BatchNormalization on batch
 * statistics, Dropout(0.5) at model.py:226,239), mse, backward, optimizer update, BN moving statistics.
 * bf16 activations / activation gradients with fp32 accumulation; fp32 parameters, gradients and optimizer state in
 * FLAT buffers owned by the caller (device memory), laid out per layer as kernel (Keras layout), bias, gamma, beta —
 * itmo_train_layout() gives the offsets — so a data-parallel run all-reduces ONE buffer (itmo_train_param_count()
 * floats = 11,159,011, layers.txt "Trainable params"). bn_moving = [moving_mean | moving_variance], itmo_train_bn_count()
 * floats each (2 x 3,904 = layers.txt "Non-trainable params: 7,808"). opt_state = 2 x itmo_train_param_count() floats. */
typedef struct itmo_trainer itmo_trainer;
enum itmo_train_flags {
  ITMO_TRAIN_NO_DROPOUT = 1,   /* Dropout layers act as identity (parity tests against a deterministic oracle) */
  ITMO_TRAIN_FREEZE_MOVING = 2, /* do not update moving_mean / moving_variance */
  ITMO_TRAIN_KEEP_GRADS = 4     /* keep a copy of every layer's activation gradients (itmo_train_read_activation 3 / 4) */
};
enum itmo_optimizer {
  ITMO_OPT_RMSPROP = 0, /* model.py:281; Keras defaults lr 1e-3, rho 0.9, eps 1e-7 */
  ITMO_OPT_ADAM = 1     /* model.py:280,282 (commented variants): beta_1 0.5 */
};
long long itmo_train_param_count(void);
long long itmo_train_bn_count(void);
/* which: 0 kernel, 1 bias, 2 gamma, 3 beta (offsets into params / grads), 4 = offset into one half of bn_moving */
int itmo_train_layout(int layer, int which, long long* offset, long long* count);
/* micro_batch images of h x w (h multiple of 256, w of 128; the reference trains at 512x512, main.py:96) fit the arena */
int itmo_trainer_create(itmo_trainer** out, int device, int micro_batch, int h, int w, float* params_dev,
                        float* grads_dev, float* opt_state_dev, float* bn_moving_dev);
void itmo_trainer_destroy(itmo_trainer* t);
/* Rebuild the tensor-core operand copies after the caller wrote params_dev (itmo_train_update does it itself). */
int itmo_trainer_sync_weights(itmo_trainer* t, void* stream);
int itmo_train_zero_grad(itmo_trainer* t, void* stream);
/* Forward + backward of one micro-batch, gradients ADDED into grads_dev.
 *  x: [n,h,w,3] uint8 (raw SDR, /255 fused) or float32 in [0,1]; y: [n,h,w,3] float32 target in [0,1];
 *  pred (optional): float32 [n,h,w,3] network output; loss_count: number of output elements of the WHOLE batch the mse
 *  averages over (all micro-batches of all ranks x h x w x 3), so summed gradients are the gradient of the batch mean;
 *  seed: dropout seed of this micro-batch. Asynchronous on `stream`. */
int itmo_train_backward(itmo_trainer* t, const void* x_dev, int x_dtype, const float* y_dev, float* pred_dev_or_null,
                        int n, double loss_count, unsigned seed, unsigned flags, void* stream);
/* The per-layer dropout seed itmo_train_backward derives from `seed` (layer 7 = conv2d_8/drop4, 9 = conv2d_10/drop5). */
int itmo_train_dropout_seed(unsigned seed, int layer, unsigned* out);
/* Sum of squared errors and number of 'accurate' pixels accumulated since the last reset (synchronises `stream`). */
int itmo_train_metrics(itmo_trainer* t, double* sq_err_sum, double* hits, int reset, void* stream);
/* Optimizer step on the flat buffers (gradients multiplied by grad_scale first), then itmo_trainer_sync_weights.
 * RMSprop: beta1_or_rho = rho, beta2 unused. Adam: bias-corrected with the trainer's step counter. */
int itmo_train_update(itmo_trainer* t, int optimizer, float lr, float beta1_or_rho, float beta2, float eps,
                      float grad_scale, void* stream);
/* Optimizer steps already taken (Keras' optimizer `iterations`): Adam's bias correction lr*sqrt(1-b2^t)/(1-b1^t) uses
 * t = steps + 1. itmo_train_update increments it; set it when resuming from a checkpoint's /optimizer_weights
 * (keras.models.load_model restores `iterations` the same way), so the first resumed step continues the schedule. */
int itmo_train_set_step(itmo_trainer* t, long long steps);
long long itmo_train_get_step(const itmo_trainer* t);
/* Debug / parity: activation of the last micro-batch as float32; which 0 = ReLU(conv+b), 1 = BN(+dropout) output, 2 = pooled,
 * 3 = dL/d(BN output), 4 = dL/d(conv output before ReLU) (both only after a backward with ITMO_TRAIN_KEEP_GRADS) */
int itmo_train_read_activation(itmo_trainer* t, int layer, int which, int n, float* dst_host, size_t cap_elems,
                               size_t* n_elems);
/* Device timing per phase: 0 forward convs, 1 data-gradient convs, 2 weight-gradient GEMMs, 3 everything else */
int itmo_train_set_profiling(itmo_trainer* t, int enable);
int itmo_train_phase_times(itmo_trainer* t, double* ms_sum4, long long* calls4, int reset);
long long itmo_train_launch_count(const itmo_trainer* t);
/* Stateless gradient operators of one ConvBN layer (bf16 NHWC device tensors, fp32 Keras-layout kernel on the device):
 * dw [3][3][cin][cout] += sum_p x[p + tap] g[p];   dx = full correlation of g with the kernel. */
int itmo_wgrad3x3(const void* x, int cin, const void* g, int cout, int n, int h, int w, float* dw, void* stream);
int itmo_dgrad3x3(const void* g, int cout, int n, int h, int w, const float* kernel_hwio, int cin, void* dx,
                  void* stream);

/* ---- training-data synthesis and augmentation (BASELINE configs[3], [4]: "synthetic SDR/HDR pairs") -----------
 * The reference prepares its data offline in MATLAB and augments on the host in Keras; these entry points run the same
 * arithmetic as CUDA kernels so a training / 4K run can be fed without leaving the GPU. All *_dev pointers are device
 * memory, images are NHWC [n,h,w,3]; parameter arrays are HOST memory, one entry per image. Asynchronous on `stream`
 * after an internal synchronisation for the parameter upload. */
/* Seeded log-normal HDR radiance field, float32 (stands in for the EXR corpus the reference does not ship). */
int itmo_synth_hdr_field(float* hdr_dev, int n, int h, int w, unsigned seed, void* stream);
/* One virtual camera per image (m-files/virtual_camera.m:13, 21-22): v ~ U[-4,4] stops, n ~ N(0.6, sqrt 0.1),
 * y ~ N(0.9, sqrt 0.1). */
typedef struct itmo_camera_params {
  double v, n, y;
} itmo_camera_params;
/* m-files/virtual_camera.m:12-37: HDR -> SDR network input. G = exp(mean log max(Y, realmin)), X = 0.18 * 2^v / G * Y,
 * a = (1+n) X^y / (n + X^y), min(1, a), colour by the luminance ratio, imwrite's 8-bit rounding.
 * out_dtype ITMO_DT_U8: the PNG's pixel values; ITMO_DT_F32: those / 255 (Keras rescale=1/255, main.py:71). */
int itmo_virtual_camera(const float* hdr_dev, int n, int h, int w, const itmo_camera_params* cams_host, void* sdr_dev,
                        int out_dtype, void* stream);
/* m-files/Reinhard.m:12-31: HDR -> tone-mapped training target. X = key / G * Y (key 0.18, :20), L = X / (1 + X) (:22),
 * colour by L / Y, imwrite's 8-bit rounding; out_dtype as above. */
int itmo_reinhard_target(const float* hdr_dev, int n, int h, int w, double key, void* out_dev, int out_dtype, void* stream);
/* main.py:71-76: ImageDataGenerator's random transform for one batch. Per image: input (row, col) = M * output (row, col)
 * + off — the matrix keras_preprocessing builds from rotation and zoom around its image centre — sampled bilinearly
 * with edge replication (scipy.ndimage.affine_transform order=1, mode='nearest'), then the horizontal / vertical flips.
 * Use the SAME params for the input batch and the target batch (the reference gives both generators seed=1,
 * main.py:78-100). in: uint8 or float32; out float32 = value * scale (1/255 = rescale). */
typedef struct itmo_affine_params {
  double m00, m01, m10, m11, off0, off1;
  int flip_h, flip_v;
} itmo_affine_params;
int itmo_augment(const void* in_dev, int in_dtype, int n, int h, int w, const itmo_affine_params* params_host,
                 float* out_dev, float scale, void* stream);

/* ---- fused gradient exchange over NVLink peer memory (replaces the ncclAllReduce + itmo_train_update pair
 * of a data-parallel step, main.py:126-132 run on N GPUs of one box) -------------------------------------------------------
 * ONE kernel per step and rank: reduce-scatter (peer loads of every rank's gradient shard, summed IN RANK ORDER) ->
 * optimizer on the local 1/N shard -> all-gather (peer stores of the updated parameters into every rank's buffer), between
 * two flag barriers in the peers' windows. One process per GPU; the windows are shared through CUDA IPC:
 *   1. every rank: itmo_exchange_create (allocates the window holding ITS flat gradient + parameter buffers, returns the
 *      IPC handle), then itmo_trainer_create with params_dev = itmo_exchange_params(), grads_dev = itmo_exchange_grads();
 *   2. the caller gathers all ranks' handles (any host channel; 64 bytes each) and calls itmo_exchange_open;
 *   3. per step, after the last itmo_train_backward: itmo_exchange_step on every rank (same arguments as
 *      itmo_train_update; also rebuilds the tensor-core operand copies). All ranks must call it the same number of times.
 * Every parameter is computed by exactly one rank and broadcast, so all ranks hold bit-identical parameters, and an
 * N-rank step equals one rank accumulating the N shards in rank order bit for bit. The owner alone keeps a shard's
 * optimizer state current (gather the shards before saving a checkpoint). world == 1 needs no open and equals
 * itmo_train_update. */
#define ITMO_IPC_HANDLE_BYTES 64
typedef struct itmo_exchange itmo_exchange;
int itmo_exchange_create(itmo_exchange** out, int device, void* ipc_handle_out /* ITMO_IPC_HANDLE_BYTES or NULL */);
float* itmo_exchange_grads(itmo_exchange* x);
float* itmo_exchange_params(itmo_exchange* x);
int itmo_exchange_open(itmo_exchange* x, int rank, int world, const void* all_handles /* [world][ITMO_IPC_HANDLE_BYTES] */);
int itmo_exchange_step(itmo_exchange* x, itmo_trainer* t, int optimizer, float lr, float beta1_or_rho, float beta2,
                       float eps, float grad_scale, void* stream);
void itmo_exchange_destroy(itmo_exchange* x);

/* ---- allocation-free variant (NO counterpart in the reference) ------------------
 * itmo_reinhard_target for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc): same kernels, same arguments, bit-identical output, but the per-image log-average scratch
 * (64 block partials + G per image, doubles) is a file-scope `static __device__` array of the library instead of two
 * cudaMallocAsync blocks. The call makes no cudaMalloc / cudaFree / cudaMallocAsync and calls no allocating library
 * routine; it runs on the CURRENT device, asynchronously on `stream`.
 *   limit: n <= ITMO_REINHARD_STATIC_MAX_IMAGES images per call (any h, w); a larger n returns ITMO_ERR_INVALID and
 *          launches nothing — split the batch or use itmo_reinhard_target.
 *   cost : every GPU that loads the library holds the array (256 * 65 * 8 = 133,120 bytes) until the process exits.
 *   races: ALL calls on a device share the one array — calls that overlap on the GPU (different streams, or two host
 *          threads on different streams) race on it; keep them on one stream or order them with events. The array keeps
 *          the previous call's contents; nothing in it has to start at zero (every element is written before it is read).
 *   The first call on each device resolves the array's device address (cudaGetSymbolAddress); make that call outside a
 *   stream capture. */
#define ITMO_REINHARD_STATIC_MAX_IMAGES 256
int itmo_reinhard_target_static(const float* hdr_dev, int n, int h, int w, double key, void* out_dev, int out_dtype,
                                void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ITMO_B200_H_ */
