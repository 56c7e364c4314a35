"""CPU restatement of CNN-ITMO's data preparation and augmentation — TEST INFRASTRUCTURE ONLY.

Only tests/ (and bench.py's checker legs) may import this module; the product path (cnn-itmo_b200/) never does.

What is restated (reference file:line):
  * m-files/virtual_camera.m:12-37   HDR radiance -> SDR network input
  * m-files/Reinhard.m:12-31         HDR radiance -> tone-mapped training target
  * main.py:71-76, 78-100            ImageDataGenerator(rescale=1/255, rotation_range=90, horizontal_flip, vertical_flip,
                                     zoom_range=0.2) with the same seed for inputs and targets
The arithmetic of the augmentation lives in third-party code that is NOT in /root/reference: Keras 2.2.x delegates to
keras_preprocessing (1.0.x) ``apply_affine_transform``, which calls ``scipy.ndimage.affine_transform(order=1,
mode='nearest')`` per channel. scipy IS installed here, so ``augment_np`` is pinned against the real scipy call in
tests/test_datagen_cpu.py. MATLAB (and the HDR Toolbox's RGB2Lum, the Rec.709 luminance) cannot run here and the
reference ships no EXR inputs or golden PNGs for these scripts: the tone curves are restated from the script text with
MATLAB's precision rules (EXR data is single; single * double = single; pow on double(...)) — parity unpinned.
"""
from __future__ import annotations

from typing import Dict, Sequence, Tuple

import numpy as np

REALMIN = np.finfo(np.float64).tiny


# ---------------------------------------------------------------------------------------------------
# seeded HDR field (the device's counter-hash generator, cnn-itmo_b200/csrc/datagen_ops.cuh)
# ---------------------------------------------------------------------------------------------------
def _hash(x: np.ndarray, seed: int) -> np.ndarray:
    with np.errstate(over='ignore'):
        h = (x.astype(np.uint32) * np.uint32(0x9E3779B1)) ^ np.uint32(seed & 0xFFFFFFFF)
        h ^= h >> np.uint32(16)
        h *= np.uint32(0x85EBCA6B)
        h ^= h >> np.uint32(13)
        h *= np.uint32(0xC2B2AE35)
        h ^= h >> np.uint32(16)
    return h


def _normal(idx: np.ndarray, seed: int, stream: int) -> np.ndarray:
    s1 = (seed ^ ((stream * 0x632BE5AB + 0x1234567) & 0xFFFFFFFF)) & 0xFFFFFFFF
    s2 = (seed ^ ((stream * 0x632BE5AB + 0x89ABCDE) & 0xFFFFFFFF)) & 0xFFFFFFFF
    u1 = (_hash(idx, s1).astype(np.float64) + 0.5) / 4294967296.0
    u2 = (_hash(idx, s2).astype(np.float64) + 0.5) / 4294967296.0
    return np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)


def synth_hdr_field(n: int, h: int, w: int, seed: int) -> np.ndarray:
    """float32 [n,h,w,3] log-normal radiance: exp(1.2 * base(8x8 blocks) + 0.15 * fine)."""
    bh, bw = h // 8 + 1, w // 8 + 1
    img, y, x, c = np.meshgrid(np.arange(n), np.arange(h), np.arange(w), np.arange(3), indexing='ij')
    bidx = ((img * bh + y // 8) * bw + x // 8) * 3 + c
    fidx = ((img * h + y) * w + x) * 3 + c
    base = _normal(bidx.astype(np.uint64).astype(np.uint32), seed, 1)
    fine = _normal(fidx.astype(np.uint64).astype(np.uint32), seed, 2)
    return np.exp(1.2 * base + 0.15 * fine).astype(np.float32)


# ---------------------------------------------------------------------------------------------------
# virtual_camera.m / Reinhard.m
# ---------------------------------------------------------------------------------------------------
def rgb2lum(img_f32: np.ndarray) -> np.ndarray:
    """HDR Toolbox RGB2Lum on single data: Rec.709 weights, every product and sum rounded to single."""
    img = img_f32.astype(np.float32)
    return (np.float32(0.2126) * img[..., 0] + np.float32(0.7152) * img[..., 1]) + np.float32(0.0722) * img[..., 2]


def geomean_lum(Y: np.ndarray) -> float:
    """virtual_camera.m:14-19 / Reinhard.m:13-19: exp(mean(log(max(double(Y), realmin))))."""
    return float(np.exp(np.log(np.maximum(Y.astype(np.float64), REALMIN)).sum() / Y.size))


def _imwrite_u8(x: np.ndarray) -> np.ndarray:
    """imwrite of a floating image: uint8(round(255 x)) with saturation, NaN -> 0."""
    x = np.where(np.isnan(x), 0.0, x)
    return np.floor(255.0 * np.clip(x, 0.0, 1.0).astype(np.float64) + 0.5).astype(np.uint8)


def virtual_camera(hdr: np.ndarray, v: float, n: float, y: float, key: float = 0.18) -> np.ndarray:
    """One HDR image float32 [h,w,3] -> uint8 SDR [h,w,3] (virtual_camera.m:12-37)."""
    hdr = hdr.astype(np.float32)
    Y = rgb2lum(hdr)                                         # :12
    G = geomean_lum(Y)                                       # :14-19
    delta_t = key * 2.0 ** v / G                             # :20
    X = (np.float64(delta_t) * Y.astype(np.float64)).astype(np.float32)   # :21  double * single = single
    with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
        xp = X.astype(np.float64) ** y
        a = (1.0 + n) * (xp / (n + xp))                      # :26
        L = np.minimum(1.0, a).astype(np.float32)            # :27
        ratio = L / Y                                        # :32 single ./ single
        out = hdr * ratio[..., None]
    return _imwrite_u8(out)                                  # :37


def reinhard_target(hdr: np.ndarray, key: float = 0.18) -> np.ndarray:
    """One HDR image float32 [h,w,3] -> uint8 tone-mapped target (Reinhard.m:12-31)."""
    hdr = hdr.astype(np.float32)
    Y = rgb2lum(hdr)
    G = geomean_lum(Y)
    delta_t = key / G                                        # :20
    X = (np.float64(delta_t) * Y.astype(np.float64)).astype(np.float32)   # :21
    with np.errstate(divide='ignore', invalid='ignore'):
        L = (X.astype(np.float64) / (1.0 + X.astype(np.float64))).astype(np.float32)   # :22
        ratio = L / Y
        out = hdr * ratio[..., None]                         # :25
    return _imwrite_u8(out)


# ---------------------------------------------------------------------------------------------------
# main.py:71-76 augmentation
# ---------------------------------------------------------------------------------------------------
def keras_affine(theta_deg: float, zx: float, zy: float, h: int, w: int) -> Tuple[np.ndarray, np.ndarray]:
    """keras_preprocessing.image.apply_affine_transform's matrix for rotation + zoom (no shift / shear in
    main.py:71-76): transform = rotation @ zoom, recentred by transform_matrix_offset_center (o_x = h/2 + 0.5,
    o_y = w/2 + 0.5). Returns (2x2 matrix, offset) as scipy.ndimage.affine_transform takes them."""
    t = np.deg2rad(theta_deg)
    rot = np.array([[np.cos(t), -np.sin(t), 0], [np.sin(t), np.cos(t), 0], [0, 0, 1]])
    zoom = np.array([[zx, 0, 0], [0, zy, 0], [0, 0, 1]])
    m = np.dot(np.eye(3), rot) if theta_deg != 0 else np.eye(3)
    if zx != 1 or zy != 1:
        m = np.dot(m, zoom)
    o_x, o_y = float(h) / 2 + 0.5, float(w) / 2 + 0.5
    off = np.array([[1, 0, o_x], [0, 1, o_y], [0, 0, 1]])
    reset = np.array([[1, 0, -o_x], [0, 1, -o_y], [0, 0, 1]])
    m = np.dot(np.dot(off, m), reset)
    return m[:2, :2].copy(), m[:2, 2].copy()


def augment_np(img: np.ndarray, theta: float, zx: float, zy: float, flip_h: bool, flip_v: bool,
               scale: float = 1.0 / 255.0) -> np.ndarray:
    """Explicit restatement: bilinear sampling at clamp(M (r, c) + off), then the flips, then rescale. img [h,w,3]
    (uint8 or float) -> float32 [h,w,3]."""
    h, w = img.shape[:2]
    M, off = keras_affine(theta, zx, zy, h, w)
    r, c = np.meshgrid(np.arange(h, dtype=np.float64), np.arange(w, dtype=np.float64), indexing='ij')
    y = np.clip(M[0, 0] * r + M[0, 1] * c + off[0], 0, h - 1)
    x = np.clip(M[1, 0] * r + M[1, 1] * c + off[1], 0, w - 1)
    y0 = np.minimum(np.floor(y).astype(np.int64), max(h - 2, 0))
    x0 = np.minimum(np.floor(x).astype(np.int64), max(w - 2, 0))
    y1, x1 = np.minimum(y0 + 1, h - 1), np.minimum(x0 + 1, w - 1)
    fy, fx = (y - y0)[..., None], (x - x0)[..., None]
    f = img.astype(np.float64)
    out = (1 - fy) * ((1 - fx) * f[y0, x0] + fx * f[y0, x1]) + fy * ((1 - fx) * f[y1, x0] + fx * f[y1, x1])
    out = out.astype(np.float32)
    if flip_h:
        out = out[:, ::-1]
    if flip_v:
        out = out[::-1]
    return (out * np.float32(scale)).astype(np.float32)


def augment_scipy(img: np.ndarray, theta: float, zx: float, zy: float, flip_h: bool, flip_v: bool,
                  scale: float = 1.0 / 255.0) -> np.ndarray:
    """The same through the library call Keras makes: scipy.ndimage.affine_transform per channel (order=1,
    mode='nearest') on the float32 image, flip_axis, x *= rescale."""
    from scipy import ndimage
    h, w = img.shape[:2]
    M, off = keras_affine(theta, zx, zy, h, w)
    x = img.astype(np.float32)
    chans = [ndimage.affine_transform(x[..., ch], M, off, order=1, mode='nearest', cval=0.0) for ch in range(3)]
    out = np.stack(chans, axis=-1)
    if flip_h:
        out = out[:, ::-1]
    if flip_v:
        out = out[::-1]
    return (out * np.float32(scale)).astype(np.float32)


def random_transform_params(rng: np.random.RandomState, rotation_range=90.0, zoom_range=0.2) -> Dict[str, float]:
    """ImageDataGenerator.get_random_transform's draws, in its order: theta, (zx, zy), flip_horizontal, flip_vertical."""
    theta = rng.uniform(-rotation_range, rotation_range)
    zx, zy = rng.uniform(1 - zoom_range, 1 + zoom_range, 2)
    flip_h = bool(rng.random_sample() < 0.5)
    flip_v = bool(rng.random_sample() < 0.5)
    return dict(theta=float(theta), zx=float(zx), zy=float(zy), flip_h=flip_h, flip_v=flip_v)
