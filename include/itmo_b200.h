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
  ITMO_PREC_TF32 = 0, /* fp32 activations in HBM rounded to TF32 (rna), tcgen05 kind::tf32 */
  ITMO_PREC_BF16 = 1, /* bf16 activations in HBM, tcgen05 kind::f16 (bf16 operands) */
  ITMO_PREC_FP32 = 2, /* fp32 activations as TF32 head + residual, 3xTF32 products: x*w_hi + x*w_lo + x_lo*w_hi
                         (fp32-accurate; parity: max-abs <= 1e-3 against the fp32 reference) */
  ITMO_PREC_FP16 = 3  /* fp16 activations in HBM (saturating), tcgen05 kind::f16 (fp16 operands): same speed as
                         bf16 with 3 more mantissa bits (parity: PSNR >= 45 dB) */
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
  ITMO_OUT_HDR_AS_WRITTEN = 3 /* float32: m-files/inverse_Reinhard.m:2-23 literally (degenerate) */
};
/* Tile plan along one axis: returns the tile count, fills origin / valid range [lo,hi) per tile (up to cap). */
int itmo_tile_axis_plan(int extent, int* origin, int* valid_lo, int* valid_hi, int cap);
/* frames_dev: [n_frames,h,w,3] uint8 on the device; out_dev: same shape, uint8 or float32 by out_kind.
 * `key` is the geometric-mean luminance the repaired inverse normalises to (Reinhard.m:20 uses 0.18). */
int itmo_predict_frames(itmo_ctx* ctx, const void* frames_dev, int n_frames, int h, int w, void* out_dev, int out_kind,
                        double key, void* stream);
/* Replaces m-files/inverse_Reinhard.m on its own: uint8 CNN output frames -> float32 HDR.
 * stats_scratch_dev: 32 bytes per frame of device scratch. */
int itmo_inverse_reinhard(const void* frames_u8_dev, void* hdr_f32_dev, int n_frames, int h, int w, int as_written,
                          double key, void* stats_scratch_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ITMO_B200_H_ */
