"""Training-data synthesis and augmentation on the GPU.

The reference prepares its training pairs offline in MATLAB — ``m-files/virtual_camera.m`` renders the SDR network
input from an HDR photograph, ``m-files/Reinhard.m`` the tone-mapped target — and augments them on the host with
Keras' ``ImageDataGenerator`` (main.py:71-76: rescale 1/255, rotation_range 90, horizontal / vertical flips,
zoom_range 0.2; the same seed for inputs and targets, main.py:78-100). Here the same arithmetic runs as CUDA kernels
of libitmo_b200.so (csrc/datagen_ops.cuh), so BASELINE configs[3] / [4] can be fed with "synthetic SDR/HDR pairs"
without leaving the device:

    x_u8, y_f32 = synthetic_pairs_device(n, 512, 512, seed)          # torch CUDA tensors
    xa, ya = augment_pair_device(x_u8, y_f32, params)                 # float32, ready for Trainer.backward

PyTorch is used for device allocation only. There is no CPU fallback.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib


# ---- host-side parameter sampling (what MATLAB's rand / normrnd and Keras' get_random_transform draw) ------------------
def sample_camera_params(n: int, seed: int, safe: bool = True) -> List[Dict[str, float]]:
    """One virtual camera per image (virtual_camera.m:13,22-23): v = 8*rand()-4, n = normrnd(0.6, sqrt(0.1)),
    y = normrnd(0.9, sqrt(0.1)). The script uses the draws as they come; a non-positive n makes its curve
    (1+n) X^y / (n + X^y) singular (3 % of the draws), so ``safe`` redraws n, y until both exceed 0.05."""
    rng = np.random.RandomState(seed)
    out = []
    for _ in range(n):
        v = 8.0 * rng.random_sample() - 4.0
        while True:
            nn = rng.normal(0.6, np.sqrt(0.1))
            yy = rng.normal(0.9, np.sqrt(0.1))
            if not safe or (nn > 0.05 and yy > 0.05):
                break
        out.append(dict(v=float(v), n=float(nn), y=float(yy)))
    return out


def sample_transform_params(n: int, seed: int, rotation_range: float = 90.0, zoom_range: float = 0.2,
                            horizontal_flip: bool = True, vertical_flip: bool = True) -> List[Dict[str, float]]:
    """ImageDataGenerator.get_random_transform for n images, draws in its order (theta, zx / zy, the two flips),
    main.py:71-76's settings by default."""
    rng = np.random.RandomState(seed)
    out = []
    for _ in range(n):
        theta = rng.uniform(-rotation_range, rotation_range) if rotation_range else 0.0
        zx, zy = rng.uniform(1 - zoom_range, 1 + zoom_range, 2) if zoom_range else (1.0, 1.0)
        fh = bool(rng.random_sample() < 0.5) and horizontal_flip
        fv = bool(rng.random_sample() < 0.5) and vertical_flip
        out.append(dict(theta=float(theta), zx=float(zx), zy=float(zy), flip_h=fh, flip_v=fv))
    return out


def affine_matrix(theta_deg: float, zx: float, zy: float, h: int, w: int) -> Tuple[np.ndarray, np.ndarray]:
    """The (2x2 matrix, offset) keras_preprocessing's apply_affine_transform hands to scipy.ndimage.affine_transform for a
    rotation by theta_deg and a zoom (zx, zy): rotation @ zoom, recentred on (h/2 + 0.5, w/2 + 0.5)."""
    m = np.eye(3)
    if theta_deg != 0:
        t = np.deg2rad(theta_deg)
        m = np.array([[np.cos(t), -np.sin(t), 0.0], [np.sin(t), np.cos(t), 0.0], [0.0, 0.0, 1.0]])
    if zx != 1 or zy != 1:
        m = m @ np.array([[zx, 0.0, 0.0], [0.0, zy, 0.0], [0.0, 0.0, 1.0]])
    ox, oy = float(h) / 2 + 0.5, float(w) / 2 + 0.5
    m = np.array([[1.0, 0.0, ox], [0.0, 1.0, oy], [0.0, 0.0, 1.0]]) @ m @ \
        np.array([[1.0, 0.0, -ox], [0.0, 1.0, -oy], [0.0, 0.0, 1.0]])
    return m[:2, :2], m[:2, 2]


def _affine_array(params: Sequence[Dict[str, float]], h: int, w: int):
    arr = (_lib.AffineParams * len(params))()
    for i, p in enumerate(params):
        M, off = affine_matrix(p.get('theta', 0.0), p.get('zx', 1.0), p.get('zy', 1.0), h, w)
        arr[i].m00, arr[i].m01, arr[i].m10, arr[i].m11 = M[0, 0], M[0, 1], M[1, 0], M[1, 1]
        arr[i].off0, arr[i].off1 = off[0], off[1]
        arr[i].flip_h, arr[i].flip_v = int(bool(p.get('flip_h', False))), int(bool(p.get('flip_v', False)))
    return arr


# ---- device entry points ----------------------------------------------------------------------------------------------
def _torch_dev(device: int):
    import torch
    if not torch.cuda.is_available():
        raise _lib.ItmoError('the data-synthesis kernels need an sm_100a GPU: there is no CPU fallback')
    return torch, torch.device('cuda', device)


def synth_hdr_field_device(n: int, h: int, w: int, seed: int, device: int = 0):
    """Seeded log-normal HDR radiance field, float32 CUDA tensor [n,h,w,3] (itmo_synth_hdr_field)."""
    torch, dev = _torch_dev(device)
    hdr = torch.empty((n, h, w, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().itmo_synth_hdr_field(hdr.data_ptr(), n, h, w, int(seed) & 0xFFFFFFFF,
                                                    torch.cuda.current_stream().cuda_stream))
    return hdr


def virtual_camera_device(hdr, cams: Sequence[Dict[str, float]], as_float: bool = False):
    """virtual_camera.m on a float32 CUDA tensor [n,h,w,3] -> uint8 SDR (or float32 /255)."""
    import torch
    n, h, w, _ = hdr.shape
    if len(cams) != n:
        raise ValueError(f'{n} images but {len(cams)} camera parameter sets')
    arr = (_lib.CameraParams * n)()
    for i, c in enumerate(cams):
        arr[i].v, arr[i].n, arr[i].y = float(c['v']), float(c['n']), float(c['y'])
    out = torch.empty((n, h, w, 3), dtype=torch.float32 if as_float else torch.uint8, device=hdr.device)
    with torch.cuda.device(hdr.device):
        _lib.check(_lib.load().itmo_virtual_camera(hdr.data_ptr(), n, h, w, arr, out.data_ptr(),
                                                   _lib.DT_F32 if as_float else _lib.DT_U8,
                                                   torch.cuda.current_stream().cuda_stream))
    return out


def reinhard_target_device(hdr, key: float = 0.18, as_float: bool = True):
    """Reinhard.m on a float32 CUDA tensor [n,h,w,3] -> float32 target in [0,1] (the 8-bit PNG value / 255) or uint8."""
    import torch
    n, h, w, _ = hdr.shape
    out = torch.empty((n, h, w, 3), dtype=torch.float32 if as_float else torch.uint8, device=hdr.device)
    with torch.cuda.device(hdr.device):
        _lib.check(_lib.load().itmo_reinhard_target(hdr.data_ptr(), n, h, w, float(key), out.data_ptr(),
                                                    _lib.DT_F32 if as_float else _lib.DT_U8,
                                                    torch.cuda.current_stream().cuda_stream))
    return out


def reinhard_target_static_into(hdr, out, key: float = 0.18):
    """reinhard_target_device without any allocation: the caller passes the output tensor ([n,h,w,3] float32 or uint8, same
    device) and the library's static device scratch holds the per-image log-average (itmo_reinhard_target_static; at most
    _lib.REINHARD_STATIC_MAX_IMAGES images per call; calls that overlap on the GPU race on that scratch). No counterpart
    in the reference."""
    import torch
    n, h, w, _ = hdr.shape
    if tuple(out.shape) != (n, h, w, 3) or out.dtype not in (torch.float32, torch.uint8) or out.device != hdr.device:
        raise ValueError('out must be [n,h,w,3] float32 or uint8 on the device of hdr')
    if hdr.dtype != torch.float32 or not hdr.is_contiguous() or not out.is_contiguous():
        raise ValueError('hdr must be contiguous float32, out contiguous')
    with torch.cuda.device(hdr.device):
        _lib.check(_lib.load().itmo_reinhard_target_static(hdr.data_ptr(), n, h, w, float(key), out.data_ptr(),
                                                           _lib.DT_F32 if out.dtype == torch.float32 else _lib.DT_U8,
                                                           torch.cuda.current_stream().cuda_stream))
    return out


def augment_device(x, params: Sequence[Dict[str, float]], scale: Optional[float] = None):
    """ImageDataGenerator's random transform (main.py:71-76) of a CUDA batch [n,h,w,3] (uint8 0..255, or float32) with one
    parameter set per image -> float32. scale defaults to 1/255 for uint8 input (rescale) and 1 for float input."""
    import torch
    n, h, w, _ = x.shape
    if len(params) != n:
        raise ValueError(f'{n} images but {len(params)} transform parameter sets')
    if x.dtype not in (torch.uint8, torch.float32):
        raise ValueError('augment_device takes uint8 or float32 images')
    if scale is None:
        scale = 1.0 / 255.0 if x.dtype == torch.uint8 else 1.0
    out = torch.empty((n, h, w, 3), dtype=torch.float32, device=x.device)
    arr = _affine_array(params, h, w)
    x = x.contiguous()
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().itmo_augment(x.data_ptr(), _lib.DT_U8 if x.dtype == torch.uint8 else _lib.DT_F32, n, h, w,
                                            arr, out.data_ptr(), float(np.float32(scale)),
                                            torch.cuda.current_stream().cuda_stream))
    return out


def augment_pair_device(x, y, params: Sequence[Dict[str, float]]):
    """The SAME random transform on the input batch and the target batch (main.py:78-100: both generators get seed=1)."""
    return augment_device(x, params), augment_device(y, params)


def synthetic_pairs_device(n: int, h: int, w: int, seed: int, device: int = 0, want_target: bool = True,
                           augment: bool = False):
    """n training pairs made on the device: HDR field -> (virtual-camera SDR uint8 [n,h,w,3], Reinhard target float32
    [n,h,w,3] or None). With ``augment`` both are passed through the same random transforms (float32 outputs)."""
    hdr = synth_hdr_field_device(n, h, w, seed, device)
    x = virtual_camera_device(hdr, sample_camera_params(n, seed + 1))
    y = reinhard_target_device(hdr) if want_target else None
    if augment:
        params = sample_transform_params(n, seed + 2)
        x = augment_device(x, params)
        if y is not None:
            y = augment_device(y, params)
    return x, y
