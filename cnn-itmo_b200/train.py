"""Training step of the U-Net behind the Keras surface the reference uses.

    model = U_net()                                   # model.py:204, compiled rmsprop / mse / accuracy (model.py:281)
    model.fit_generator(gen, steps_per_epoch=..., epochs=...)        # main.py:126-132
    loss, acc = model.train_on_batch(x, y)

``Trainer`` owns the flat fp32 parameter / gradient / optimizer-state buffers (torch CUDA tensors: PyTorch is used for
device allocation, streams and torch.distributed only) and drives libitmo_b200.so's itmo_train_* entry points; every
kernel of the step is in the library. Data parallel: one process per GPU, each rank runs its share of the batch
(micro-batched), then ONE NCCL all-reduce (sum) of the flat gradient buffer over NVLink, then the same optimizer step on
every rank. There is no CPU fallback.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib, graph

_OPTIMIZERS = {'rmsprop': _lib.OPT_RMSPROP, 'adam': _lib.OPT_ADAM}
# Keras defaults: RMSprop(lr=1e-3, rho=0.9, epsilon=1e-7) (model.py:281); Adam of the comments: beta_1=0.5 (model.py:280,282)
_OPT_DEFAULTS = {'rmsprop': dict(lr=1e-3, beta1=0.9, beta2=0.0, eps=1e-7),
                 'adam': dict(lr=1e-3, beta1=0.5, beta2=0.999, eps=1e-7)}
WHICH = {'kernel': 0, 'bias': 1, 'gamma': 2, 'beta': 3}


def param_layout() -> List[Dict[str, Tuple[int, int]]]:
    """Per layer {field: (offset, count)} inside the flat parameter buffer; 'bn' = offset inside one half of bn_moving."""
    lib = _lib.load()
    out = []
    for li, l in enumerate(graph.LAYERS):
        d = {}
        for name, which in (('kernel', 0), ('bias', 1), ('gamma', 2), ('beta', 3), ('bn', 4)):
            if l.kind == 'head' and which >= 2:
                continue
            off, cnt = ctypes.c_longlong(), ctypes.c_longlong()
            _lib.check(lib.itmo_train_layout(li, which, ctypes.byref(off), ctypes.byref(cnt)))
            d[name] = (off.value, cnt.value)
        out.append(d)
    return out


def world_info() -> Tuple[int, int]:
    """(world_size, rank) of the default process group, (1, 0) when torch.distributed is not initialised."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(), dist.get_rank()
    return 1, 0


def shard_range(n_global: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [begin, end) slice of a global batch that `rank` trains on (BASELINE configs[4]: 512 images split
    512 / 256 / 128 / 64 per GPU). The batch must divide evenly: every rank runs the same number of micro-batches, so
    the all-reduce is reached in lock step."""
    if n_global % world:
        raise ValueError(f'global batch {n_global} is not divisible by {world} ranks')
    per = n_global // world
    return rank * per, (rank + 1) * per


def global_loss_count(n_local: int, h: int, w: int, world: int) -> float:
    """Number of output elements the mse averages over for the WHOLE batch (Keras 'mse' = mean over N*H*W*3): every
    rank divides its squared-error gradient by this, so the SUM over ranks is the gradient of the global mean."""
    return float(n_local * world) * h * w * 3


def allreduce_flat(flat_tensor):
    """The one exchange step of data-parallel training: in-place SUM of a flat gradient tensor over all ranks
    (NCCL over NVLink for CUDA tensors, gloo for the CPU tests). No-op for a single process."""
    import torch.distributed as dist
    world, _ = world_info()
    if world > 1:
        dist.all_reduce(flat_tensor, op=dist.ReduceOp.SUM)
    return flat_tensor


def flatten_layers(layers: Sequence[Dict[str, np.ndarray]], layout=None) -> np.ndarray:
    """Per-layer {kernel, bias, gamma, beta} dicts -> the flat fp32 buffer in the library's layout (itmo_train_layout)."""
    layout = layout or param_layout()
    n = int(_lib.load().itmo_train_param_count())
    flat = np.zeros(n, np.float32)
    for li, lay in enumerate(layout):
        for name in ('kernel', 'bias', 'gamma', 'beta'):
            if name in lay:
                off, cnt = lay[name]
                flat[off:off + cnt] = np.asarray(layers[li][name], np.float32).reshape(-1)
    return flat


def keras_trainable_order() -> List[Tuple[int, str]]:
    """(layer index, field) of every trainable weight in the order of Keras' ``model.trainable_weights`` — layers as
    ``model.layers`` lists them, Conv2D / Conv2DTranspose [kernel, bias], BatchNormalization [gamma, beta]. This is
    the order of the RMSprop accumulators in a checkpoint's /optimizer_weights group (main.py:123-124)."""
    out: List[Tuple[int, str]] = []
    for k in graph.keras_layers():
        if k.weight_layer is None:
            continue
        out += [(k.weight_layer, f) for f in (('gamma', 'beta') if k.cls == 'BatchNormalization' else ('kernel', 'bias'))]
    return out


def optimizer_weights_from_flat(state: np.ndarray, layout, optimizer: str, iterations: int = 0) -> List[np.ndarray]:
    """Flat optimizer state [2 * n_params] (RMSprop: accumulators in the first half; Adam: m | v) -> the list Keras
    2.2.x keeps in ``optimizer.weights``: RMSprop = one accumulator per trainable weight; Adam = [iterations] + ms + vs."""
    n = state.size // 2
    order = keras_trainable_order()

    def half(h):
        arrs = []
        for li, f in order:
            off, cnt = layout[li][f]
            a = state[h * n + off:h * n + off + cnt].copy()
            arrs.append(a.reshape(graph.kernel_shape(graph.LAYERS[li])) if f == 'kernel' else a)
        return arrs
    if optimizer == 'rmsprop':
        return half(0)
    return [np.asarray(iterations, np.int64)] + half(0) + half(1)


def optimizer_weights_to_flat(weights: Sequence[np.ndarray], layout, optimizer: str, n_params: int):
    """Inverse of optimizer_weights_from_flat -> (flat state [2 * n_params], iterations)."""
    order = keras_trainable_order()
    want = len(order) if optimizer == 'rmsprop' else 1 + 2 * len(order)
    if len(weights) != want:
        raise ValueError(f'{optimizer} state of this model has {want} arrays, got {len(weights)}')
    state = np.zeros(2 * n_params, np.float32)
    it = 0
    ws = list(weights)
    if optimizer != 'rmsprop':
        it = int(np.asarray(ws.pop(0)).reshape(-1)[0])
    for h in range(1 if optimizer == 'rmsprop' else 2):
        for i, (li, f) in enumerate(order):
            off, cnt = layout[li][f]
            a = np.asarray(ws[h * len(order) + i], np.float32).reshape(-1)
            if a.size != cnt:
                raise ValueError(f'optimizer weight {i} ({graph.LAYERS[li].name}/{f}) has {a.size} elements, expected {cnt}')
            state[h * n_params + off:h * n_params + off + cnt] = a
    return state, it


class _DeviceArray:
    """A float32 device buffer owned by libitmo_b200.so, viewed by torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {'shape': (int(n),), 'typestr': '<f4', 'data': (int(ptr), False), 'version': 3,
                                         'strides': None}


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """[begin, end) of the flat parameter buffer that `rank` owns in the fused exchange (itmo_exchange_step computes the
    same): ceil(n / world) rounded up to 4 floats per rank, the last rank takes what is left."""
    per = ((n + world - 1) // world + 3) // 4 * 4
    b = min(n, per * rank)
    return b, min(n, b + per)


class Trainer:
    """Gradient exchange (`exchange`): 'nccl' = ONE all-reduce of the flat gradient buffer, then the optimizer on every
    rank; 'fused' = itmo_exchange_step: reduce-scatter -> optimizer on the local 1/N shard -> all-gather as ONE kernel
    over NVLink peer memory (the flat buffers live in an IPC-shared window; torch.distributed only carries the 64-byte
    handles at start-up). With 'fused' only the owner keeps a shard's optimizer state current: get_optimizer_weights
    gathers the shards first.

    Batch statistics note: BatchNormalization's batch mean / variance (and its backward) are taken over ONE
    micro-batch of this rank, not over the whole global batch — Keras on one device normalises over the whole batch it
    is fed (main.py:87 feeds 2 images). With N <= micro_batch on one rank the two agree exactly; larger batches are a
    gradient-accumulation approximation, like per-replica statistics in Keras' own multi-GPU data parallelism. The
    moving statistics move ONCE per optimizer step (from the step's last micro-batch), as in Keras.

    `seed`: base of the Dropout mask stream (mixed with the step, the micro-batch and the data-parallel rank, so ranks
    draw different masks); None draws one from os.urandom so that a resumed run does not replay the masks of the run
    it continues (RMSprop checkpoints do not carry an iteration count, Keras 2.2.x keeps only the accumulators)."""

    def __init__(self, weights: Sequence[Dict[str, np.ndarray]], micro_batch: int = 8, image_hw=(512, 512),
                 device: int = 0, optimizer: str = 'rmsprop', lr: Optional[float] = None, seed: Optional[int] = 0,
                 exchange: str = 'nccl', **opt_kwargs):
        import torch
        if not torch.cuda.is_available():
            raise _lib.ItmoError('the training step needs an sm_100a GPU: there is no CPU fallback')
        lib = _lib.load()
        optimizer = optimizer.lower()
        if optimizer not in _OPTIMIZERS:
            raise ValueError(f'optimizer must be one of {sorted(_OPTIMIZERS)}, got {optimizer!r}')
        self.optimizer = optimizer
        self.opt = dict(_OPT_DEFAULTS[optimizer])
        if lr is not None:
            self.opt['lr'] = float(lr)
        self.opt.update({k: float(v) for k, v in opt_kwargs.items() if k in self.opt})
        self.device = int(device)
        self.micro_batch = int(micro_batch)
        self.h, self.w = int(image_hw[0]), int(image_hw[1])
        self.layout = param_layout()
        self.n_params = int(lib.itmo_train_param_count())
        self.n_bn = int(lib.itmo_train_bn_count())
        dev = torch.device('cuda', self.device)
        self._torch = torch
        self._dev = dev
        if exchange not in ('nccl', 'fused'):
            raise ValueError("exchange must be 'nccl' or 'fused'")
        self.exchange = exchange
        self._x = ctypes.c_void_p()
        if exchange == 'fused':
            # the flat gradient / parameter buffers live in the library's IPC-shared window
            handle = ctypes.create_string_buffer(_lib.IPC_HANDLE_BYTES)
            with torch.cuda.device(dev):
                _lib.check(lib.itmo_exchange_create(ctypes.byref(self._x), self.device, handle))
                self.params = torch.as_tensor(_DeviceArray(lib.itmo_exchange_params(self._x), self.n_params), device=dev)
                self.grads = torch.as_tensor(_DeviceArray(lib.itmo_exchange_grads(self._x), self.n_params), device=dev)
            world, rank = world_info()
            if world > 1:
                import torch.distributed as dist
                handles = [None] * world
                dist.all_gather_object(handles, handle.raw)
                err = None
                try:
                    _lib.check(lib.itmo_exchange_open(self._x, rank, world, b''.join(handles)))
                except _lib.ItmoError as ex:     # every rank must learn about it: the others are about to wait for this one
                    err = ex
                ok = torch.tensor([0 if err else 1], dtype=torch.int32, device=dev)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN)
                if int(ok[0]) == 0:
                    self.params = self.grads = None
                    lib.itmo_exchange_destroy(self._x)
                    self._x = ctypes.c_void_p()
                    raise err or _lib.ItmoError('fused exchange: a peer rank could not map the windows (CUDA IPC)')
            else:
                _lib.check(lib.itmo_exchange_open(self._x, 0, 1, None))
        else:
            self.params = torch.zeros(self.n_params, dtype=torch.float32, device=dev)
            self.grads = torch.zeros(self.n_params, dtype=torch.float32, device=dev)
        self.opt_state = torch.zeros(2 * self.n_params, dtype=torch.float32, device=dev)
        self.bn_moving = torch.zeros(2 * self.n_bn, dtype=torch.float32, device=dev)
        self._h = ctypes.c_void_p()
        torch.cuda.synchronize(dev)
        _lib.check(lib.itmo_trainer_create(ctypes.byref(self._h), self.device, self.micro_batch, self.h, self.w,
                                           self.params.data_ptr(), self.grads.data_ptr(), self.opt_state.data_ptr(),
                                           self.bn_moving.data_ptr()))
        self.set_weights(weights)
        self.seed_base = int.from_bytes(os.urandom(4), 'little') if seed is None else int(seed)
        self._copy_stream = None
        self._staged = None         # (key, (xs, ys, event), host arrays) of a prefetched first micro-batch

    @property
    def step_index(self) -> int:
        """Optimizer steps taken so far = Keras' `optimizer.iterations`; the library's counter is the one source of truth
        (Adam's bias correction reads it)."""
        return int(_lib.load().itmo_train_get_step(self._h)) if self._h else 0

    @step_index.setter
    def step_index(self, value: int):
        _lib.check(_lib.load().itmo_train_set_step(self._h, int(value)))

    def micro_seed(self, step: int, i0: int) -> int:
        """Dropout seed of the micro-batch starting at local image i0 of optimizer step `step` on this rank."""
        _, rank = world_info()
        return (self.seed_base * 2654435761 + step * 7919 + i0 + rank * 104729) & 0xFFFFFFFF

    # ---- plumbing ---------------------------------------------------------------------------------------------
    def _stream(self) -> int:
        return self._torch.cuda.current_stream(self._dev).cuda_stream

    def close(self):
        self._staged = None                          # a prefetched micro-batch nobody came for
        if getattr(self, '_h', None):
            _lib.load().itmo_trainer_destroy(self._h)
            self._h = ctypes.c_void_p()
        if getattr(self, '_x', None):
            self.params = self.grads = None          # views of the window
            _lib.load().itmo_exchange_destroy(self._x)
            self._x = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def set_weights(self, weights: Sequence[Dict[str, np.ndarray]]):
        flat = np.zeros(self.n_params, np.float32)
        mov = np.zeros(2 * self.n_bn, np.float32)
        for li, lay in enumerate(self.layout):
            w = weights[li]
            for name in ('kernel', 'bias', 'gamma', 'beta'):
                if name in lay:
                    off, cnt = lay[name]
                    flat[off:off + cnt] = np.asarray(w[name], np.float32).reshape(-1)
            if 'bn' in lay:
                off, cnt = lay['bn']
                mov[off:off + cnt] = np.asarray(w['moving_mean'], np.float32)
                mov[self.n_bn + off:self.n_bn + off + cnt] = np.asarray(w['moving_variance'], np.float32)
        self.params.copy_(self._torch.from_numpy(flat))
        self.bn_moving.copy_(self._torch.from_numpy(mov))
        _lib.check(_lib.load().itmo_trainer_sync_weights(self._h, self._stream()))

    def _unflatten(self, flat: np.ndarray, with_moving: bool) -> List[Dict[str, np.ndarray]]:
        mov = self.bn_moving.cpu().numpy() if with_moving else None
        out = []
        for li, lay in enumerate(self.layout):
            l = graph.LAYERS[li]
            d = {}
            for name in ('kernel', 'bias', 'gamma', 'beta'):
                if name in lay:
                    off, cnt = lay[name]
                    a = flat[off:off + cnt].copy()
                    d[name] = a.reshape(graph.kernel_shape(l)) if name == 'kernel' else a
            if with_moving and 'bn' in lay:
                off, cnt = lay['bn']
                d['moving_mean'] = mov[off:off + cnt].copy()
                d['moving_variance'] = mov[self.n_bn + off:self.n_bn + off + cnt].copy()
            out.append(d)
        return out

    def get_weights(self) -> List[Dict[str, np.ndarray]]:
        """Per-layer dicts in the Keras layouts (what checkpoint.write_checkpoint / Model take)."""
        self._torch.cuda.synchronize(self._dev)
        return self._unflatten(self.params.cpu().numpy(), True)

    def get_optimizer_weights(self) -> List[np.ndarray]:
        """The optimizer state as Keras lists it in ``model.optimizer.weights`` (what model.save writes under
        /optimizer_weights, main.py:123-124)."""
        self._torch.cuda.synchronize(self._dev)
        self._gather_optimizer_shards()
        return optimizer_weights_from_flat(self.opt_state.cpu().numpy(), self.layout, self.optimizer, self.step_index)

    def _gather_optimizer_shards(self):
        """Fused exchange: every rank's optimizer state is current only on the shard it owns; make it whole everywhere
        (one all-reduce of the state with the foreign shards zeroed — checkpoint time only)."""
        world, rank = world_info()
        if self.exchange != 'fused' or world == 1:
            return
        import torch.distributed as dist
        b, e = shard_bounds(self.n_params, world, rank)
        for h in range(2):
            half = self.opt_state[h * self.n_params:(h + 1) * self.n_params]
            half[:b].zero_()
            half[e:].zero_()
        dist.all_reduce(self.opt_state)

    def set_optimizer_weights(self, weights: Sequence[np.ndarray]):
        """Resume: load the state a checkpoint's /optimizer_weights group holds."""
        state, it = optimizer_weights_to_flat(weights, self.layout, self.optimizer, self.n_params)
        self.opt_state.copy_(self._torch.from_numpy(state))
        if self.optimizer != 'rmsprop':
            self.step_index = it     # itmo_train_set_step: the resumed step continues Adam's bias-correction schedule

    def get_grads(self) -> List[Dict[str, np.ndarray]]:
        self._torch.cuda.synchronize(self._dev)
        return self._unflatten(self.grads.cpu().numpy(), False)

    # ---- the step ---------------------------------------------------------------------------------------------
    def zero_grad(self):
        _lib.check(_lib.load().itmo_train_zero_grad(self._h, self._stream()))

    def backward(self, x_dev, y_dev, loss_count: float, seed: int = 0, dropout: bool = True,
                 freeze_moving: bool = False, pred_dev=None, keep_grads: bool = False):
        """Forward (training mode) + backward of one micro-batch of DEVICE tensors: x uint8 or float32 [n,h,w,3],
        y float32 [n,h,w,3]; gradients are added into self.grads."""
        torch = self._torch
        n = int(x_dev.shape[0])
        if tuple(x_dev.shape[1:]) != (self.h, self.w, 3) or tuple(y_dev.shape) != tuple(x_dev.shape):
            raise ValueError(f'expected x, y of shape (n, {self.h}, {self.w}, 3), got {tuple(x_dev.shape)} / '
                             f'{tuple(y_dev.shape)}')
        if x_dev.dtype not in (torch.uint8, torch.float32) or y_dev.dtype != torch.float32:
            raise ValueError('x must be uint8 or float32 and y float32')
        flags = (0 if dropout else _lib.TRAIN_NO_DROPOUT) | (_lib.TRAIN_FREEZE_MOVING if freeze_moving else 0) | \
            (_lib.TRAIN_KEEP_GRADS if keep_grads else 0)
        _lib.check(_lib.load().itmo_train_backward(
            self._h, x_dev.data_ptr(), _lib.DT_U8 if x_dev.dtype == torch.uint8 else _lib.DT_F32, y_dev.data_ptr(),
            pred_dev.data_ptr() if pred_dev is not None else None, n, float(loss_count), int(seed) & 0xFFFFFFFF, flags,
            self._stream()))

    def allreduce_grads(self):
        """The one exchange step of data-parallel training: sum of the flat gradient buffer over all ranks
        (NCCL over NVLink when the process group's backend is nccl)."""
        allreduce_flat(self.grads)

    def update(self, grad_scale: float = 1.0):
        o = self.opt
        _lib.check(_lib.load().itmo_train_update(self._h, _OPTIMIZERS[self.optimizer], o['lr'], o['beta1'], o['beta2'],
                                                 o['eps'], float(grad_scale), self._stream()))

    def exchange_and_update(self, grad_scale: float = 1.0):
        """The exchange step + optimizer of one data-parallel step: all-reduce then update ('nccl'), or the one fused
        kernel over peer memory ('fused')."""
        if self.exchange == 'fused':
            o = self.opt
            _lib.check(_lib.load().itmo_exchange_step(self._x, self._h, _OPTIMIZERS[self.optimizer], o['lr'], o['beta1'],
                                                      o['beta2'], o['eps'], float(grad_scale), self._stream()))
        else:
            self.allreduce_grads()
            self.update(grad_scale)

    def metrics(self, reset: bool = True) -> Tuple[float, float]:
        sq, hits = ctypes.c_double(), ctypes.c_double()
        _lib.check(_lib.load().itmo_train_metrics(self._h, ctypes.byref(sq), ctypes.byref(hits), int(reset),
                                                  self._stream()))
        return sq.value, hits.value

    @staticmethod
    def _host_batch(x, y):
        x = np.ascontiguousarray(x if np.asarray(x).dtype == np.uint8 else np.asarray(x, np.float32))
        y = np.ascontiguousarray(np.asarray(y, np.float32))
        return x, y

    @staticmethod
    def _batch_key(x, y):
        # identity of the arrays AS PASSED (before any dtype conversion makes a copy)
        x, y = np.asarray(x), np.asarray(y)
        return (x.ctypes.data, y.ctypes.data, x.shape, x.dtype.str, y.dtype.str)

    def train_on_batch(self, x, y, seed: Optional[int] = None, dropout: bool = True, prefetch=None) -> List[float]:
        """One optimizer step on HOST arrays x [N,h,w,3] (uint8, or float in [0,1]) and y [N,h,w,3] (float target): this
        rank's share of the batch. Returns [loss, accuracy] of the whole (all-rank) batch, like Keras' train_on_batch
        with metrics=['accuracy']. The H2D copy of micro-batch i+1 runs on a side stream while micro-batch i computes
        (asynchronous only from pinned arrays — model.pinned_empty). `prefetch=(x_next, y_next)`: the NEXT step's batch
        (what fit_generator's look-ahead passes): its first micro-batch is uploaded under this step's last micro-batch,
        so no copy of the next call is exposed. The arrays must stay unmodified until that call, which must be given the
        same arrays (otherwise the staged copy is dropped and nothing is lost but the overlap)."""
        import torch
        import torch.distributed as dist
        key = self._batch_key(x, y)
        x, y = self._host_batch(x, y)
        n = x.shape[0]
        world, _ = world_info()
        n_global = n * world
        loss_count = global_loss_count(n, self.h, self.w, world)
        step = self.step_index if seed is None else seed
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self._dev)
        main = torch.cuda.current_stream(self._dev)
        starts = list(range(0, n, self.micro_batch))

        def upload(i0, hx=x, hy=y):
            # staged on the side stream; the tensors are handed to the compute stream through an event
            with torch.cuda.stream(self._copy_stream):
                xs = torch.from_numpy(hx[i0:i0 + self.micro_batch]).to(self._dev, non_blocking=True)
                ys = torch.from_numpy(hy[i0:i0 + self.micro_batch]).to(self._dev, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
            return xs, ys, ev

        staged, self._staged = self._staged, None
        self.zero_grad()
        nxt = staged[1] if staged is not None and staged[0] == key else upload(starts[0])
        for k, i0 in enumerate(starts):
            xs, ys, ev = nxt
            if k + 1 < len(starts):
                nxt = upload(starts[k + 1])
            elif prefetch is not None:
                px, py = self._host_batch(*prefetch)
                self._staged = (self._batch_key(*prefetch), upload(0, px, py), (prefetch, px, py))   # keeps the arrays alive
            main.wait_event(ev)
            # BN moving statistics move once per optimizer step (Keras updates them once per batch): last micro-batch only
            self.backward(xs, ys, loss_count, seed=self.micro_seed(step, i0), dropout=dropout,
                          freeze_moving=(k + 1 < len(starts)))
            xs.record_stream(main)
            ys.record_stream(main)
        self.exchange_and_update()
        sq, hits = self.metrics()
        if world > 1:
            t = torch.tensor([sq, hits], dtype=torch.float64, device=self._dev)
            dist.all_reduce(t)
            sq, hits = float(t[0]), float(t[1])
        return [sq / loss_count, hits / (n_global * self.h * self.w)]

    def read_activation(self, layer: int, which: str, n: int) -> np.ndarray:
        """z (ReLU output), a (BN/dropout output) or pooled tensor of the last micro-batch as float32 NHWC."""
        l = graph.LAYERS[layer]
        scale = 2 if l.kind == 'convT' else 1
        ho, wo = scale * l.hin * self.h // 512, scale * l.hin * self.w // 512
        idx = {'z': 0, 'a': 1, 'pooled': 2, 'dA': 3, 'dpre': 4}[which]
        if idx == 2:
            ho, wo = ho // 2, wo // 2
        buf = np.empty((n, ho, wo, l.cout), np.float32)
        cnt = ctypes.c_size_t(0)
        _lib.check(_lib.load().itmo_train_read_activation(self._h, layer, idx, n,
                                                          buf.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), buf.size,
                                                          ctypes.byref(cnt)))
        return buf

    def set_profiling(self, enable: bool = True):
        _lib.check(_lib.load().itmo_train_set_profiling(self._h, int(bool(enable))))

    def phase_times(self, reset: bool = True) -> Dict[str, Tuple[float, int]]:
        ms = (ctypes.c_double * 4)()
        calls = (ctypes.c_longlong * 4)()
        _lib.check(_lib.load().itmo_train_phase_times(self._h, ms, calls, int(reset)))
        return {k: (float(ms[i]), int(calls[i])) for i, k in enumerate(('fwd_conv', 'dgrad', 'wgrad', 'other'))}

    def launch_count(self) -> int:
        return int(_lib.load().itmo_train_launch_count(self._h))
