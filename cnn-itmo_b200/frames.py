"""Frames larger than the U-Net's native 512x512 input (BASELINE config 4: 3840x2160 SDR frames), the optional
inverse-Reinhard output stage (m-files/inverse_Reinhard.m) and the sharding of independent frames / image batches
over the GPUs of one box.

The reference has none of this: its Input is fixed at 512x512 (model.py:204,207) and predict.py:50-72 feeds one
image at a time to one device. Frames and images are independent, so the multi-GPU path has NO data-path
collective: every rank runs the whole network on its own slice; only the finished outputs are gathered.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib

TILE = 512
HALO = 96          # >= 92, the receptive-field reach of the graph, on the 16-pixel pool grid
OUT_KINDS = {'u8': _lib.OUT_U8, 'f32': _lib.OUT_F32, 'hdr': _lib.OUT_HDR_INTENDED,
             'hdr_as_written': _lib.OUT_HDR_AS_WRITTEN, 'hdr_f16': _lib.OUT_HDR_INTENDED_F16}
OUT_DTYPES = {'u8': 'uint8', 'f32': 'float32', 'hdr': 'float32', 'hdr_as_written': 'float32', 'hdr_f16': 'float16'}


def plan_axis(extent: int) -> List[Tuple[int, int, int]]:
    """[(origin, valid_lo, valid_hi)] of the tiles along one frame axis — the library's own plan
    (itmo_tile_axis_plan, a host function: no GPU needed). Every tile lies inside the frame, starts on the
    16-pixel grid and keeps HALO pixels between its valid range and every interior cut."""
    lib = _lib.load()
    cap = 64
    o, lo, hi = (ctypes.c_int * cap)(), (ctypes.c_int * cap)(), (ctypes.c_int * cap)()
    n = lib.itmo_tile_axis_plan(int(extent), o, lo, hi, cap)
    if n < 0:
        raise ValueError(f'frame extent {extent}: must be a multiple of 16 and at least {TILE}')
    return [(o[i], lo[i], hi[i]) for i in range(n)]


def plan_tiles(h: int, w: int) -> List[Tuple[int, int, int, int, int, int]]:
    """[(oy, ox, y0, y1, x0, x1)] row-major: tile origin and the rectangle of the frame it contributes."""
    return [(oy, ox, y0, y1, x0, x1) for (oy, y0, y1) in plan_axis(h) for (ox, x0, x1) in plan_axis(w)]


# ------------------------------------------------------------------------------------------------
# sharding of independent units (images, frames) over ranks
# ------------------------------------------------------------------------------------------------
def shard_slice(n: int, rank: int, world: int) -> slice:
    """Contiguous slice of n units for `rank` of `world` (sizes differ by at most one, earlier ranks larger)."""
    if not (0 <= rank < world):
        raise ValueError(f'rank {rank} outside world of {world}')
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return slice(start, start + base + (1 if rank < rem else 0))


def shard_round_robin(n: int, rank: int, world: int) -> List[int]:
    """Units rank, rank+world, ... — frames of a stream keep arriving in order on every GPU."""
    if not (0 <= rank < world):
        raise ValueError(f'rank {rank} outside world of {world}')
    return list(range(rank, n, world))


def gather_units(local: np.ndarray, indices: Sequence[int], n_total: int, group=None) -> Optional[np.ndarray]:
    """Collect every rank's finished outputs on rank 0 (returns the [n_total, ...] array there, None elsewhere).
    `local[k]` is the result for unit `indices[k]`. Host-side plumbing over torch.distributed (gloo or nccl);
    single-process when torch.distributed is not initialised."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        out = np.empty((n_total,) + local.shape[1:], dtype=local.dtype)
        out[list(indices)] = local
        return out
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    payload = (list(indices), local)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(payload, gathered, dst=0, group=group)
    if rank != 0:
        return None
    out = np.empty((n_total,) + local.shape[1:], dtype=local.dtype)
    seen = np.zeros(n_total, dtype=bool)
    for idx, arr in gathered:
        if len(idx):
            out[idx] = arr
            seen[idx] = True
    if not seen.all():
        raise RuntimeError(f'units {np.flatnonzero(~seen).tolist()} were not produced by any rank')
    return out


# ------------------------------------------------------------------------------------------------
# device path
# ------------------------------------------------------------------------------------------------
def _check_frames(frames_u8) -> np.ndarray:
    x = np.asarray(frames_u8)
    if x.ndim == 3:
        x = x[None]
    if x.ndim != 4 or x.shape[-1] != 3 or x.dtype != np.uint8:
        raise ValueError(f'expected uint8 frames [n,h,w,3], got {x.dtype} {tuple(x.shape)}')
    h, w = x.shape[1:3]
    if h < TILE or w < TILE or h % 16 or w % 16:
        raise ValueError(f'frame {h}x{w}: both sides must be multiples of 16 and at least {TILE}')
    return np.ascontiguousarray(x)


def predict_frames_device(model, frames_ptr: int, n: int, h: int, w: int, out_ptr: int, out: str = 'u8',
                          key: float = 0.18, stream: int = 0):
    """itmo_predict_frames on caller-owned DEVICE buffers; asynchronous on `stream`."""
    lib = model._ensure_ctx()
    _lib.check(lib.itmo_predict_frames(model._ctx, frames_ptr, n, h, w, out_ptr, OUT_KINDS[out], float(key),
                                       stream or None))


def predict_frames(model, frames_u8, out: str = 'u8', key: float = 0.18) -> np.ndarray:
    """uint8 frames [n,h,w,3] -> 'u8': what predict.py:64 would write for the whole frame; 'f32': the sigmoid
    output; 'hdr' / 'hdr_as_written': float32 linear HDR through the inverse-Reinhard stage; 'hdr_f16': the repaired
    inverse as half floats (OpenEXR's HALF pixel type, clamped to 65504): half the bytes to copy back."""
    import torch
    x = _check_frames(frames_u8)
    n, h, w, _ = x.shape
    dev = torch.device('cuda', model.device)
    xd = torch.from_numpy(x).to(dev)
    yd = torch.empty((n, h, w, 3), dtype=getattr(torch, OUT_DTYPES[out]), device=dev)
    with torch.cuda.device(dev):
        s = torch.cuda.current_stream()
        predict_frames_device(model, xd.data_ptr(), n, h, w, yd.data_ptr(), out, key, s.cuda_stream)
        s.synchronize()
    return yd.cpu().numpy()


def predict_frame_batches(model, batches, out: str = 'hdr', key: float = 0.18):
    """A stream of frame batches (a video): for every uint8 batch [n,h,w,3] of the iterable — all of one shape, in pinned
    host memory (model.pinned_empty) for the copies to be asynchronous — yields the result of predict_frames as a pinned
    host array. Three CUDA streams and double-buffered device / host buffers: the upload of batch k + 1 and the download of
    batch k - 1 run under the kernels of batch k, so a step costs max(copy, compute) instead of their sum (the float32 HDR
    frame is 4 x the input's bytes). The yielded array is one of two buffers: it is overwritten two batches later — use it
    (or copy it) before asking for the batch after the next; the buffers (and the three streams) stay with the model for the
    next call of the same shape and kind."""
    import torch
    from .model import pinned_empty
    dev = torch.device('cuda', model.device)
    # streams and buffers are kept on the model between calls (page-locking 2 x 100 MB costs more than a 4K step)
    cache = model.__dict__.setdefault('_frame_stream_cache', {})
    if 'streams' not in cache:
        cache['streams'] = tuple(torch.cuda.Stream(dev) for _ in range(3))
    s_in, s_cmp, s_out = cache['streams']
    for a_ in cache['streams']:          # an abandoned earlier stream of batches may still be in flight on the shared buffers
        for b_ in cache['streams']:
            if a_ is not b_:
                a_.wait_stream(b_)
    xd, yd, hy, ht = [None, None], [None, None], [None, None], [None, None]
    ev_in, ev_cmp, ev_out = [None, None], [None, None], [None, None]
    shape = None
    pending = None     # buffer index whose download is in flight and not yet yielded
    k = 0
    with torch.cuda.device(dev):
        for batch in batches:
            x = _check_frames(batch)
            if shape is None:
                shape = x.shape
                n, h, w, _ = shape
                bufs = cache.get((shape, out))
                if bufs is None:
                    for key_ in [k_ for k_ in cache if k_ != 'streams']:
                        del cache[key_]               # one shape / kind at a time: a 4K float32 pair is 250 MB of pinned memory
                    bufs = cache[(shape, out)] = {
                        'xd': [torch.empty(shape, dtype=torch.uint8, device=dev) for _ in range(2)],
                        'yd': [torch.empty(shape, dtype=getattr(torch, OUT_DTYPES[out]), device=dev) for _ in range(2)],
                        'hy': [pinned_empty(shape, np.dtype(OUT_DTYPES[out])) for _ in range(2)]}
                    bufs['ht'] = [torch.from_numpy(a_) for a_ in bufs['hy']]
                xd, yd, hy, ht = bufs['xd'], bufs['yd'], bufs['hy'], bufs['ht']
            elif x.shape != shape:
                raise ValueError(f'all batches of a stream must have one shape: {shape} then {x.shape}')
            b = k & 1
            if ev_cmp[b] is not None:
                s_in.wait_event(ev_cmp[b])            # the kernels of batch k - 2 have read xd[b]
            with torch.cuda.stream(s_in):
                xd[b].copy_(torch.from_numpy(x), non_blocking=True)
                ev_in[b] = torch.cuda.Event()
                ev_in[b].record(s_in)
            s_cmp.wait_event(ev_in[b])
            if ev_out[b] is not None:
                s_cmp.wait_event(ev_out[b])           # the download of batch k - 2 has read yd[b]
            predict_frames_device(model, xd[b].data_ptr(), n, h, w, yd[b].data_ptr(), out, key, s_cmp.cuda_stream)
            ev_cmp[b] = torch.cuda.Event()
            ev_cmp[b].record(s_cmp)
            if pending is not None:                    # batch k - 1: its download was queued last iteration
                ev_out[pending].synchronize()
                yield hy[pending]
            s_out.wait_event(ev_cmp[b])
            with torch.cuda.stream(s_out):
                ht[b].copy_(yd[b], non_blocking=True)
                ev_out[b] = torch.cuda.Event()
                ev_out[b].record(s_out)
            pending = b
            k += 1
        if pending is not None:
            ev_out[pending].synchronize()
            yield hy[pending]


def inverse_reinhard(frames_u8, as_written: bool = False, key: float = 0.18, device: int = 0) -> np.ndarray:
    """m-files/inverse_Reinhard.m on the GPU: uint8 CNN-output frames [n,h,w,3] -> float32 HDR."""
    import torch
    x = np.asarray(frames_u8)
    if x.ndim == 3:
        x = x[None]
    if x.ndim != 4 or x.shape[-1] != 3 or x.dtype != np.uint8:
        raise ValueError(f'expected uint8 frames [n,h,w,3], got {x.dtype} {tuple(x.shape)}')
    lib = _lib.load()
    dev = torch.device('cuda', device)
    xd = torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    yd = torch.empty(x.shape, dtype=torch.float32, device=dev)
    scratch = torch.empty((max(1, x.shape[0]), 4), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        s = torch.cuda.current_stream()
        _lib.check(lib.itmo_inverse_reinhard(xd.data_ptr(), yd.data_ptr(), x.shape[0], x.shape[1], x.shape[2],
                                             int(as_written), float(key), scratch.data_ptr(), s.cuda_stream))
        s.synchronize()
    return yd.cpu().numpy()


def predict_frames_sharded(model, frames_u8, out: str = 'u8', key: float = 0.18, group=None) -> Optional[np.ndarray]:
    """Frames round-robin over the ranks of the process group, each rank on its own GPU; rank 0 gets the result."""
    import torch.distributed as dist
    x = _check_frames(frames_u8)
    if dist.is_available() and dist.is_initialized():
        rank, world = dist.get_rank(group), dist.get_world_size(group)
    else:
        rank, world = 0, 1
    mine = shard_round_robin(x.shape[0], rank, world)
    local = predict_frames(model, x[mine], out, key) if mine else \
        np.empty((0,) + x.shape[1:], dtype=np.uint8 if out == 'u8' else np.float32)
    return gather_units(local, mine, x.shape[0], group)
