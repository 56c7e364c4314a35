"""Drop-in equivalents of the reference's model API for its one hot path.

    from cnn_itmo_b200.model import U_net, load_model          # reference: model.py:204, predict.py:9,24
    model = load_model('saved7-model-218-0.73.hdf5')            # predict.py:24
    pred = model.predict(X)                                     # predict.py:62  -> float32 [N,512,512,3]

``Model`` mirrors the subset of the Keras ``Model`` object the reference touches (predict, summary,
save / save_weights / load_weights, get_weights / set_weights, compile, fit_generator / train_on_batch)
with the same argument meaning and error behaviour (ValueError for a wrong input shape, OSError for an
unreadable checkpoint). All arithmetic runs in libitmo_b200.so on an sm_100a GPU; there is no CPU path.
"""
from __future__ import annotations

import ctypes
import math
import os
import warnings
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

from . import _lib, checkpoint, graph

# bf16 / fp16: 16-bit activations in HBM, tcgen05 kind::f16 (same speed; fp16 carries 3 more mantissa bits)
# tf32: fp32 activations rounded to TF32, tcgen05 kind::tf32
# fp32: 3xTF32 (x*w_hi + x*w_lo + x_lo*w_hi) — fp32-accurate, the strict-parity mode
_PRECISIONS = {'bf16': _lib.PREC_BF16, 'fp16': _lib.PREC_FP16, 'tf32': _lib.PREC_TF32, 'fp32': _lib.PREC_FP32}


def _default_precision() -> str:
    # fp16 is the default 16-bit mode: same tensor-core rate as bf16 with three more significand bits; it holds the
    # north-star's PSNR >= 45 dB on the calibrated checkpoint where bf16 measures ~32 dB. bf16 is opt-in.
    return os.environ.get('ITMO_PRECISION', 'fp16').lower()


def _he_normal(rng: np.random.Generator, shape, fan_in: int) -> np.ndarray:
    """Keras 2.2.4 he_normal = VarianceScaling(2, 'fan_in', 'normal'): truncated normal (|z| <= 2),
    stddev sqrt(2/fan_in)/0.8796 so the truncated distribution has the requested variance."""
    std = math.sqrt(2.0 / fan_in) / .87962566103423978
    out = rng.standard_normal(shape)
    bad = np.abs(out) > 2.0
    while bad.any():
        out[bad] = rng.standard_normal(int(bad.sum()))
        bad = np.abs(out) > 2.0
    return (out * std).astype(np.float32)


def initial_weights(seed: Optional[int] = None) -> checkpoint.Weights:
    """Fresh weights as U_net() creates them (model.py:196,200,276): he_normal kernels, zero biases,
    BN gamma=1 beta=0 mean=0 var=1, glorot_uniform head."""
    rng = np.random.default_rng(seed)
    ws: checkpoint.Weights = []
    for l in graph.LAYERS:
        ks = graph.kernel_shape(l)
        if l.kind == 'head':
            lim = math.sqrt(6.0 / (ks[2] + ks[3]))
            ws.append({'kernel': rng.uniform(-lim, lim, ks).astype(np.float32),
                       'bias': np.zeros(l.cout, np.float32)})
            continue
        fan_in = ks[0] * ks[1] * (ks[2] if l.kind == 'conv' else ks[3])
        ws.append({'kernel': _he_normal(rng, ks, fan_in), 'bias': np.zeros(l.cout, np.float32),
                   'gamma': np.ones(l.cout, np.float32), 'beta': np.zeros(l.cout, np.float32),
                   'moving_mean': np.zeros(l.cout, np.float32), 'moving_variance': np.ones(l.cout, np.float32)})
    return ws


FP16_MAX = 65504.0
FP16_MIN_NORMAL = 2.0 ** -14


def fp16_range_report(weights: checkpoint.Weights, bn_eps: float = graph.BN_EPS, sigmas: float = 1000.0) -> List[str]:
    """Load-time range check of the fp16 storage mode (the default). Every tensor the kernels keep in HBM is a
    post-BatchNormalization activation y = gamma * zhat + beta with zhat the standardised ReLU output, so its range
    follows from gamma / beta: flagged when |beta| + `sigmas` * |gamma| could leave fp16's +-65504. The first layer
    is bounded exactly (inputs are in [0,1]): |y| <= |s| * (sum |w| + |b|) + |t|. Weights are flagged when they
    overflow fp16 or when more than 1 % of a kernel's energy sits below fp16's smallest normal (2^-14), where
    precision is lost. Returns human-readable findings (empty = fine); the device-side counter
    (Model.saturation_count) catches what a static check cannot."""
    out: List[str] = []
    for l, w in zip(graph.LAYERS, weights):
        k = np.asarray(w['kernel'], np.float64)
        if np.abs(k).max(initial=0.0) > FP16_MAX:
            out.append(f'{l.name}: kernel magnitude {np.abs(k).max():.3g} exceeds the fp16 range')
        if l.kind in ('conv', 'convT') and l.index != 0:
            e = float((k ** 2).sum())
            small = float((k[np.abs(k) < FP16_MIN_NORMAL] ** 2).sum())
            if e > 0 and small / e > 0.01:
                out.append(f'{l.name}: {100 * small / e:.1f} % of the kernel energy is below fp16\'s smallest normal '
                           '(2^-14): reduced precision')
        if l.kind == 'head':
            continue
        g, b = np.asarray(w['gamma'], np.float64), np.asarray(w['beta'], np.float64)
        bound = float((np.abs(b) + sigmas * np.abs(g)).max())
        if l.index == 0:
            s = g / np.sqrt(np.asarray(w['moving_variance'], np.float64) + bn_eps)
            t = b - np.asarray(w['moving_mean'], np.float64) * s
            zmax = np.abs(k).reshape(-1, k.shape[-1]).sum(0) + np.abs(np.asarray(w['bias'], np.float64))
            bound = float((np.abs(s) * zmax + np.abs(t)).max())
        if bound > FP16_MAX:
            out.append(f'{l.name}: post-BatchNormalization activations may reach {bound:.3g} (> 65504): fp16 storage '
                       "could saturate; precision='tf32' / 'fp32' has fp32 range")
    return out


def pinned_empty(shape, dtype) -> np.ndarray:
    """numpy array in page-locked host memory (itmo_host_alloc): H2D/D2H copies from/to it are asynchronous, so
    ``predict`` / ``predict_u8`` overlap them with the kernels. The memory is released with the array."""
    import weakref
    lib = _lib.load()
    dtype = np.dtype(dtype)
    nbytes = int(np.prod(shape, dtype=np.int64)) * dtype.itemsize
    ptr = ctypes.c_void_p()
    _lib.check(lib.itmo_host_alloc(ctypes.byref(ptr), nbytes))
    buf = (ctypes.c_char * max(nbytes, 1)).from_address(ptr.value)
    arr = np.frombuffer(buf, dtype=dtype, count=nbytes // dtype.itemsize).reshape(shape)
    weakref.finalize(buf, lib.itmo_host_free, ptr)
    return arr


class Model:
    """The compiled U-Net ('ReinhardtPrediction', model.py:278) bound to one B200."""

    def __init__(self, weights: Optional[checkpoint.Weights] = None, input_size=(512, 512, 3),
                 precision: Optional[str] = None, device: int = 0, max_batch: int = 32,
                 bn_epsilon: float = graph.BN_EPS):
        if tuple(input_size) != (graph.IMG, graph.IMG, 3):
            raise ValueError(f'input_size {tuple(input_size)} is not supported: the kernels are built for the '
                             f'reference\'s native ({graph.IMG}, {graph.IMG}, 3) input (model.py:204); larger frames '
                             'go through cnn_itmo_b200.frames')
        precision = (precision or _default_precision()).lower()
        if precision not in _PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_PRECISIONS)}, got {precision!r}")
        self.name = graph.MODEL_NAME
        self.input_shape = (None,) + tuple(input_size)
        self.output_shape = (None, graph.IMG, graph.IMG, 3)
        self.precision = precision
        self.device = int(device)
        self.max_batch = int(max_batch)
        self.bn_epsilon = float(bn_epsilon)
        self.optimizer = None
        self.loss = None
        self.metrics: List[str] = []
        self.meta: Dict[str, Any] = {}
        self._weights = weights if weights is not None else initial_weights()
        self._ctx = ctypes.c_void_p()
        self._uploaded = False
        self._trainer = None
        self._trained_dirty = False     # the trainer's parameters are newer than self._weights
        self._range_checked = False

    # ---- device context -----------------------------------------------------------------------------
    def _ensure_ctx(self):
        lib = _lib.load()
        self._sync_trained()
        if not self._ctx:
            _lib.check(lib.itmo_create(ctypes.byref(self._ctx), self.device, _PRECISIONS[self.precision],
                                       self.max_batch))
            _lib.check(lib.itmo_set_graph_capture(self._ctx, int(self.use_cuda_graph)))
        if not self._uploaded:
            arr = (_lib.LayerWeights * graph.NUM_LAYERS)()
            keep = []
            for i, w in enumerate(self._weights):
                for field in ('kernel', 'bias', 'gamma', 'beta', 'moving_mean', 'moving_variance'):
                    if field in w:
                        a = np.ascontiguousarray(w[field], dtype=np.float32)
                        keep.append(a)
                        setattr(arr[i], field, a.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
            _lib.check(lib.itmo_load_weights(self._ctx, arr, graph.NUM_LAYERS, self.bn_epsilon))
            self._uploaded = True
        if self.precision == 'fp16' and not self._range_checked:
            self._range_checked = True
            for msg in fp16_range_report(self._weights, self.bn_epsilon):
                warnings.warn(msg, RuntimeWarning, stacklevel=3)
        return lib

    def saturation_count(self, reset: bool = False) -> int:
        """fp16 mode: epilogue store groups that clamped a value to +-65504 since the context was created / last reset
        (itmo_saturation_count; synchronises the device). 0 in every other mode."""
        if not self._ctx:
            return 0
        n = ctypes.c_ulonglong(0)
        _lib.check(_lib.load().itmo_saturation_count(self._ctx, ctypes.byref(n), int(reset)))
        return int(n.value)

    def _warn_if_saturated(self):
        if self.precision == 'fp16':
            n = self.saturation_count(reset=True)
            if n:
                warnings.warn(f'fp16 activations saturated in {n} store groups during this predict call: the result is '
                              "clipped; use precision='tf32' or 'fp32' for this checkpoint", RuntimeWarning, stacklevel=3)

    def close(self):
        self.__dict__.pop('_frame_stream_cache', None)   # frames.predict_frame_batches' streams and pinned buffers
        if getattr(self, '_ctx', None):
            _lib.load().itmo_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()
            self._uploaded = False

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    @property
    def ctx(self) -> ctypes.c_void_p:
        """The itmo_ctx* (for callers that drive itmo_forward on device buffers, e.g. bench.py)."""
        self._ensure_ctx()
        return self._ctx

    # ---- Keras Model surface ------------------------------------------------------------------------
    def summary(self, line_length=None, positions=None, print_fn: Optional[Callable[[str], None]] = None):
        graph.print_summary(print_fn or print)

    def count_params(self) -> int:
        return graph.count_params()['total']

    def compile(self, optimizer='rmsprop', loss='mse', metrics=None, **kwargs):
        """model.compile(optimizer='rmsprop', loss='mse', metrics=['accuracy']) (model.py:281)."""
        if loss not in ('mse', 'mean_squared_error'):
            raise ValueError(f'loss {loss!r} is not supported (the reference trains with mse)')
        self.optimizer, self.loss, self.metrics = optimizer, loss, list(metrics or [])

    def _check_input(self, x) -> np.ndarray:
        x = np.asarray(x)
        if x.ndim != 4 or tuple(x.shape[1:]) != (graph.IMG, graph.IMG, 3):
            raise ValueError('Error when checking input: expected input_1 to have shape '
                             f'({graph.IMG}, {graph.IMG}, 3) but got array with shape {tuple(x.shape[1:])}'
                             if x.ndim == 4 else
                             f'Error when checking input: expected input_1 to have 4 dimensions, but got array '
                             f'with shape {tuple(x.shape)}')
        return x

    def predict(self, x, batch_size: Optional[int] = None, verbose: int = 0, steps=None) -> np.ndarray:
        """model.predict(X) (predict.py:62): X [N,512,512,3] float in [0,1] (any float dtype, the reference
        passes float64) -> float32 [N,512,512,3]. Synchronous; host in, host out."""
        x = self._check_input(x)
        if x.dtype == np.uint8:
            raise ValueError('predict() takes normalised floats like the reference (predict.py:59); '
                             'use predict_u8() for raw 8-bit images')
        if x.shape[0] == 0:
            return np.empty((0, graph.IMG, graph.IMG, 3), dtype=np.float32)
        lib = self._ensure_ctx()
        xf = pinned_empty(x.shape, np.float32)
        np.copyto(xf, x, casting='unsafe')              # Keras casts the fed array to floatx
        out = pinned_empty((xf.shape[0], graph.IMG, graph.IMG, 3), np.float32)
        _lib.check(lib.itmo_predict_host(self._ctx, xf.ctypes.data, _lib.DT_F32, out.ctypes.data, _lib.DT_F32,
                                         xf.shape[0], 0))
        self._warn_if_saturated()
        return out

    def predict_u8(self, images_u8, out: Optional[np.ndarray] = None) -> np.ndarray:
        """predict.py:59-64 fused: raw uint8 [N,512,512,3] -> uint8 [N,512,512,3] = trunc(sigmoid(...)*255).
        ``out`` (optional) is a caller-owned C-contiguous uint8 array of the same shape to write into — pass
        ``pinned_empty`` buffers for both sides to keep the copies asynchronous and skip the per-call allocation."""
        x = self._check_input(images_u8)
        if x.dtype != np.uint8:
            raise ValueError('predict_u8() expects uint8 images')
        if x.shape[0] == 0:
            return np.empty_like(x)
        lib = self._ensure_ctx()
        x = np.ascontiguousarray(x)
        if out is None:
            out = pinned_empty(x.shape, np.uint8)
        elif out.dtype != np.uint8 or out.shape != x.shape or not out.flags.c_contiguous:
            raise ValueError('out must be a C-contiguous uint8 array with the input\'s shape')
        _lib.check(lib.itmo_predict_host(self._ctx, x.ctypes.data, _lib.DT_U8, out.ctypes.data, _lib.DT_U8,
                                         x.shape[0], 0))
        self._warn_if_saturated()
        return out

    def forward_device(self, in_ptr: int, in_dtype: str, out_ptr: int, out_dtype: str, n: int, stream: int = 0):
        """itmo_forward on caller-owned DEVICE buffers (data_ptr() of torch tensors). Asynchronous."""
        lib = self._ensure_ctx()
        dt = {'u8': _lib.DT_U8, 'f32': _lib.DT_F32}
        _lib.check(lib.itmo_forward(self._ctx, in_ptr, dt[in_dtype], out_ptr, dt[out_dtype], n, 0, stream or None))

    def set_profiling(self, enable: bool = True):
        """Bracket every layer launch with CUDA events (itmo_set_profiling); read with layer_times()."""
        lib = self._ensure_ctx()
        _lib.check(lib.itmo_set_profiling(self._ctx, int(bool(enable))))

    def layer_times(self, reset: bool = True):
        """[(milliseconds, launches)] per layer of graph.LAYERS accumulated since the last reset."""
        lib = self._ensure_ctx()
        ms = (ctypes.c_double * graph.NUM_LAYERS)()
        calls = (ctypes.c_longlong * graph.NUM_LAYERS)()
        _lib.check(lib.itmo_layer_times(self._ctx, ms, calls, int(reset)))
        return [(float(ms[i]), int(calls[i])) for i in range(graph.NUM_LAYERS)]

    def launch_count(self) -> int:
        return int(_lib.load().itmo_launch_count(self._ctx)) if self._ctx else 0

    def read_activation(self, layer: int, pooled: bool = False) -> np.ndarray:
        """Intermediate activation of the last forward chunk as float32 NHWC (per-layer parity tests)."""
        lib = self._ensure_ctx()
        l = graph.LAYERS[layer]
        ho = (2 * l.hin if l.kind == 'convT' else l.hin) // (2 if pooled else 1)
        buf = np.empty((self.max_batch, ho, ho, l.cout), dtype=np.float32)
        n = ctypes.c_size_t(0)
        _lib.check(lib.itmo_read_activation(self._ctx, layer, int(pooled),
                                            buf.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), buf.size,
                                            ctypes.byref(n)))
        return buf.reshape(-1)[:n.value].reshape(-1, ho, ho, l.cout)

    # ---- weights ------------------------------------------------------------------------------------
    def get_weights(self) -> List[np.ndarray]:
        """Flat list in Keras' model.get_weights() order (model.layers order; kernel, bias / gamma, beta,
        moving_mean, moving_variance per layer)."""
        self._sync_trained()
        out: List[np.ndarray] = []
        for k in graph.keras_layers():
            if k.weight_layer is None:
                continue
            w = self._weights[k.weight_layer]
            keys = checkpoint.BN_KEYS if k.cls == 'BatchNormalization' else ('kernel', 'bias')
            out += [w[key].copy() for key in keys]
        return out

    def set_weights(self, flat: Sequence[np.ndarray]):
        flat = list(flat)
        i = 0
        new = [dict(w) for w in self._weights]
        for k in graph.keras_layers():
            if k.weight_layer is None:
                continue
            keys = checkpoint.BN_KEYS if k.cls == 'BatchNormalization' else ('kernel', 'bias')
            for key in keys:
                if i >= len(flat):
                    raise ValueError('You called `set_weights(weights)` with a weight list that is too short')
                a = np.asarray(flat[i], dtype=np.float32)
                if a.shape != new[k.weight_layer][key].shape:
                    raise ValueError(f'Layer weight shape {new[k.weight_layer][key].shape} not compatible with '
                                     f'provided weight shape {a.shape}')
                new[k.weight_layer][key] = a.copy()
                i += 1
        if i != len(flat):
            raise ValueError(f'set_weights: expected {i} arrays, got {len(flat)}')
        self._install_weights(new)

    def _install_weights(self, new):
        """New host weights: the inference context re-uploads on next use and a live training engine takes them too."""
        self._weights = new
        self._uploaded = False
        self._range_checked = False
        self._trained_dirty = False
        if self._trainer is not None:
            self._trainer.set_weights(new)

    def load_weights(self, filepath: str, by_name: bool = False):
        new, meta = checkpoint.read_checkpoint(filepath)
        self.meta.update(meta)
        self._install_weights(new)

    def save_weights(self, filepath: str, overwrite: bool = True):
        if not overwrite and os.path.exists(filepath):
            raise OSError(f'{filepath} exists')
        self._sync_trained()
        checkpoint.write_checkpoint(filepath, self._weights, weights_only=True)

    def save(self, filepath: str, overwrite: bool = True, include_optimizer: bool = True):
        """model.save (what ModelCheckpoint calls every epoch, main.py:123-124): full Keras HDF5 layout."""
        if not overwrite and os.path.exists(filepath):
            raise OSError(f'{filepath} exists')
        opt_w, opt_name = getattr(self, '_optimizer_weights', None), getattr(self, '_optimizer_name', 'rmsprop')
        self._sync_trained()
        if self._trainer is not None:
            opt_w, opt_name = self._trainer.get_optimizer_weights(), self._trainer.optimizer
        # /optimizer_weights (RMSprop accumulators in trainable_weights order) so that training can resume from the
        # file, here or in Keras
        checkpoint.write_checkpoint(filepath, self._weights, weights_only=False,
                                    optimizer_weights=opt_w if include_optimizer else None, optimizer=opt_name)

    # ---- training (main.py:126-132) -----------------------------------------------------------------
    def _ensure_trainer(self, micro_batch: Optional[int] = None):
        """The training engine (train.Trainer over itmo_train_*), created on first use with the model's weights."""
        from . import train
        if self._trainer is None:
            if self.optimizer is None:
                raise RuntimeError('You must compile a model before training/testing. Use `model.compile(optimizer, loss)`.')
            opt = self.optimizer if isinstance(self.optimizer, str) else self.optimizer.get('name', 'rmsprop')
            kw = {} if isinstance(self.optimizer, str) else {k: v for k, v in self.optimizer.items() if k != 'name'}
            self._trainer = train.Trainer(self._weights, micro_batch=micro_batch or self.train_micro_batch,
                                          device=self.device, optimizer=opt, seed=self.train_seed, **kw)
            pending = getattr(self, '_optimizer_weights', None)
            if pending is not None and getattr(self, '_optimizer_name', opt) == self._trainer.optimizer:
                self._trainer.set_optimizer_weights(pending)   # resume from a checkpoint's /optimizer_weights
            self._optimizer_weights = None
        return self._trainer

    # predict() / predict_u8() replay a captured CUDA graph of the 18-launch forward (itmo_set_graph_capture); set to
    # False before the first call to launch the kernels one by one ($ITMO_CUDA_GRAPH=0 does the same)
    use_cuda_graph = os.environ.get('ITMO_CUDA_GRAPH', '1') != '0'
    train_micro_batch = 8
    train_seed: Optional[int] = None     # Dropout mask stream of the training engine; None = fresh entropy per engine

    def _sync_trained(self):
        """If train_on_batch ran since the last pull, bring the trainer's parameters (and BN moving statistics) back
        into the host copy that predict() / get_weights() / save*() use, and mark the inference context stale. One
        D2H of 11.2 M floats, only when something was trained in between (the reference's per-epoch makePrediction
        callback, main.py:36-63, and ModelCheckpoint both land here)."""
        if self._trained_dirty and self._trainer is not None:
            self._weights = self._trainer.get_weights()
            self._uploaded = False
            self._range_checked = False
        self._trained_dirty = False

    def train_on_batch(self, x, y, sample_weight=None, class_weight=None, prefetch=None):
        """One optimizer step on x [N,512,512,3] (floats in [0,1] like main.py:71's rescale=1./255, or raw uint8) and
        the target y [N,512,512,3]; returns [loss, accuracy] like Keras with metrics=['accuracy']. prefetch: the next
        step's (x, y), see Trainer.train_on_batch (fit_generator passes its look-ahead batch)."""
        x = self._check_input(x)
        y = np.asarray(y)
        if y.shape != x.shape:
            raise ValueError(f'Error when checking target: expected conv2d_15 to have shape {x.shape[1:]} but got '
                             f'array with shape {y.shape[1:]}')
        if prefetch is not None:
            prefetch = (self._check_input(prefetch[0]), np.asarray(prefetch[1]))
        out = self._ensure_trainer().train_on_batch(x, y, prefetch=prefetch)
        self._trained_dirty = True
        return out

    def fit_generator(self, generator, steps_per_epoch=None, epochs=1, verbose=1, callbacks=None,
                      validation_data=None, validation_steps=None, **kwargs):
        """model.fit_generator(train_generator, steps_per_epoch, epochs, validation_data, validation_steps, callbacks)
        (main.py:126-132). `generator` yields (x, y) batches (main.py:111's zip of two flows). Callbacks: objects with
        the keras.callbacks.Callback hooks (cnn_itmo_b200.callbacks has CSVLogger / ModelCheckpoint / LambdaCallback with
        main.py:69,121-124's arguments); on_epoch_end sees the TRAINED weights through model.predict / model.save. Returns {'loss': [...], 'acc': [...],
        'val_loss': [...], 'val_acc': [...]} per epoch like History.history."""
        if steps_per_epoch is None:
            raise ValueError('steps_per_epoch is required for a generator')
        history: Dict[str, List[float]] = {'loss': [], 'acc': []}
        it = iter(generator)
        callbacks = list(callbacks or [])
        for cb in callbacks:
            if hasattr(cb, 'set_model'):
                cb.set_model(self)
            if hasattr(cb, 'set_params'):
                cb.set_params({'epochs': int(epochs), 'steps': int(steps_per_epoch), 'verbose': verbose,
                               'do_validation': validation_data is not None,
                               'metrics': ['loss', 'acc'] + (['val_loss', 'val_acc'] if validation_data is not None else [])})
        for cb in callbacks:
            if hasattr(cb, 'on_train_begin'):
                cb.on_train_begin({})
        ahead, done_steps = None, 0
        for epoch in range(int(epochs)):
            for cb in callbacks:
                if hasattr(cb, 'on_epoch_begin'):
                    cb.on_epoch_begin(epoch, {})
            ls, ac = [], []
            # one batch of look-ahead: the next batch's first micro-batch is uploaded under this step's compute
            total_steps = int(epochs) * int(steps_per_epoch)
            for _ in range(int(steps_per_epoch)):
                if ahead is None:
                    ahead = next(it)
                x, y = ahead
                done_steps += 1
                ahead = next(it, None) if done_steps < total_steps else None
                l, a = self.train_on_batch(x, y, prefetch=ahead)
                ls.append(l)
                ac.append(a)
            logs = {'loss': float(np.mean(ls)), 'acc': float(np.mean(ac))}
            if validation_data is not None:
                vit = iter(validation_data)
                vl, va = [], []
                for _ in range(int(validation_steps or 1)):
                    vx, vy = next(vit)
                    l, a = self.test_on_batch(vx, vy)
                    vl.append(l)
                    va.append(a)
                logs['val_loss'], logs['val_acc'] = float(np.mean(vl)), float(np.mean(va))
            for k, v in logs.items():
                history.setdefault(k, []).append(v)
            if verbose:
                print(f'Epoch {epoch + 1}/{epochs} - ' + ' - '.join(f'{k}: {v:.4f}' for k, v in logs.items()))
            for cb in callbacks or []:
                if hasattr(cb, 'on_epoch_end'):
                    cb.on_epoch_end(epoch, logs)
        for cb in callbacks or []:
            if hasattr(cb, 'on_train_end'):
                cb.on_train_end({})
        self._sync_trained()
        return history

    def test_on_batch(self, x, y):
        """[loss, accuracy] in inference mode (moving statistics), like Keras' test_on_batch. The trained parameters are
        pulled once per training phase (the dirty flag), not once per validation step."""
        x = self._check_input(x)
        y = np.asarray(y, np.float32)
        pred = self.predict(np.asarray(x, np.float32) / (255.0 if np.asarray(x).dtype == np.uint8 else 1.0))
        loss = float(np.mean((pred - y) ** 2))
        acc = float(np.mean(pred.argmax(-1) == y.argmax(-1)))
        return [loss, acc]


def U_net(pretrained_weights=None, input_size=(512, 512, 3), precision: Optional[str] = None, device: int = 0,
          max_batch: int = 32, seed: Optional[int] = None) -> Model:
    """Builder with the reference's signature (model.py:204). Like the reference it IGNORES
    ``pretrained_weights`` (the load is commented out at model.py:285-286), compiles with
    rmsprop / mse / accuracy (model.py:281) and prints the summary (model.py:288)."""
    model = Model(initial_weights(seed), input_size=input_size, precision=precision, device=device,
                  max_batch=max_batch)
    model.compile(optimizer='rmsprop', loss='mse', metrics=['accuracy'])
    model.summary()
    return model


def load_model(filepath: str, custom_objects=None, compile: bool = True, precision: Optional[str] = None,
               device: int = 0, max_batch: int = 32) -> Model:
    """keras.models.load_model (predict.py:24): read a Keras full-model (or weights-only) HDF5 file."""
    weights, meta = checkpoint.read_checkpoint(filepath)
    model = Model(weights, precision=precision, device=device, max_batch=max_batch,
                  bn_epsilon=meta.get('bn_epsilon', graph.BN_EPS))
    model.meta = meta
    opt_name = 'rmsprop'
    tc = meta.get('training_config')
    if isinstance(tc, dict) and str(tc.get('optimizer_config', {}).get('class_name', '')).lower() == 'adam':
        opt_name = 'adam'
    if compile:
        model.compile(optimizer=opt_name, loss='mse', metrics=['accuracy'])
        if meta.get('optimizer_weights'):
            # like keras.models.load_model: the optimizer state travels with the model and is installed when the
            # training engine is first needed (Model._ensure_trainer)
            model._optimizer_weights = meta['optimizer_weights']
            model._optimizer_name = opt_name
    return model
