"""CPU restatement of CNN-ITMO's training step — TEST INFRASTRUCTURE ONLY (parity unpinned).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this module; the
product path (cnn-itmo_b200/) never does.

What is restated (reference file:line):
  * the step Keras runs inside model.fit_generator(...)              main.py:126-132
  * on the model compiled with optimizer='rmsprop', loss='mse',
    metrics=['accuracy']                                              model.py:281
  * forward in TRAINING mode: BatchNormalization uses the batch
    statistics over (N, H, W) of the ReLU output (BN sits after the
    activation)                                                       model.py:195-200
    Dropout(0.5) after conv2d_8 and conv2d_10 (inverted scaling)      model.py:226,239
  * Keras RMSprop defaults lr=1e-3, rho=0.9, eps=1e-7; the Adam variant of the comments (beta_1=0.5)  model.py:280,282
  * BN moving statistics: momentum 0.99; moving variance fed with var * M / (M - (1 + eps)) (Keras 2.2.x
    BatchNormalization.call)

The arithmetic of the reference lives in un-vendored Keras/TensorFlow 1.x, which cannot run here, and the
reference has no tests or golden vectors for training: numerics parity is UNPINNED. The restatement is torch-CPU autograd
over the same graph as oracle/unet_ref.py (whose topology is pinned against layers.txt); tests gate it with central
finite differences in float64.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np

from . import unet_ref as R

RMSPROP_DEFAULTS = dict(lr=1e-3, rho=0.9, eps=1e-7)
ADAM_COMMENT_DEFAULTS = dict(lr=1e-3, beta_1=0.5, beta_2=0.999, eps=1e-7)


def dropout_keep_mask(shape, seed: int) -> np.ndarray:
    """Keep mask of an NHWC tensor of `shape` under the 32-bit per-element hash the device uses
    (cnn-itmo_b200/csrc/train_ops.cuh: drop_keep). seed = the per-layer seed (layer_dropout_seed)."""
    n = int(np.prod(shape))
    with np.errstate(over='ignore'):
        e = np.arange(n, dtype=np.uint64).astype(np.uint32)
        h = (e * np.uint32(0x9E3779B1)) ^ np.uint32(seed & 0xFFFFFFFF)
        h ^= h >> np.uint32(16)
        h *= np.uint32(0x85EBCA6B)
        h ^= h >> np.uint32(13)
        h *= np.uint32(0xC2B2AE35)
        h ^= h >> np.uint32(16)
    return (((h >> np.uint32(15)) & np.uint32(1)) != 0).reshape(shape)


def layer_dropout_seed(seed: int, layer: int) -> int:
    """Per-layer seed derived from the micro-batch seed (itmo_train_dropout_seed)."""
    return (seed * 0x9E3779B9 + layer * 0x85EBCA6B + 1) & 0xFFFFFFFF


def _bf16_round_fn():
    """Autograd function that rounds the value to bf16 on the way forward and the gradient to bf16 on the way back:
    what storing a tensor (and, later, its gradient) as bf16 in HBM does."""
    import torch

    class RoundBf16(torch.autograd.Function):
        @staticmethod
        def forward(ctx, t):
            return t.to(torch.bfloat16).to(t.dtype)

        @staticmethod
        def backward(ctx, g):
            return g.to(torch.bfloat16).to(g.dtype)

    return RoundBf16.apply


def train_forward_backward(weights: Sequence[Dict[str, np.ndarray]], x_nhwc: np.ndarray, y_nhwc: np.ndarray,
                           drop_seed: Optional[int] = None, dtype=None, loss_count: Optional[float] = None,
                           return_acts: bool = False, emulate_bf16: bool = False):
    """One forward (training mode) + backward. Returns dict(loss=sum of squared errors / loss_count, sq_err_sum, acc,
    pred [N,H,W,3], grads = list of {kernel, bias, gamma, beta} in Keras layouts, batch_mean / batch_var per layer).
    drop_seed None = dropout disabled (identity), else the micro-batch seed."""
    import torch
    F = torch.nn.functional
    dtype = dtype or torch.float32
    x = torch.from_numpy(np.ascontiguousarray(np.asarray(x_nhwc, dtype=np.float64))).to(dtype).permute(0, 3, 1, 2)
    y = torch.from_numpy(np.ascontiguousarray(np.asarray(y_nhwc, dtype=np.float64))).to(dtype).permute(0, 3, 1, 2)
    params: List[Dict[str, 'torch.Tensor']] = []
    for s in R.LAYERS:
        w = weights[s.index]
        d = {'kernel': torch.from_numpy(np.ascontiguousarray(w['kernel'])).to(dtype).requires_grad_(True),
             'bias': torch.from_numpy(np.ascontiguousarray(w['bias'])).to(dtype).requires_grad_(True)}
        if s.bn_name:
            d['gamma'] = torch.from_numpy(np.ascontiguousarray(w['gamma'])).to(dtype).requires_grad_(True)
            d['beta'] = torch.from_numpy(np.ascontiguousarray(w['beta'])).to(dtype).requires_grad_(True)
        params.append(d)
    outs, pooled, zs, means, vars_ = {}, {}, {}, {}, {}
    # emulate_bf16: mimic where the device path stores bf16 (every conv input and its gradient, z, a, the tensor-core
    # layers' weights); everything else (accumulation, statistics, parameter gradients) stays in `dtype`.
    rnd = _bf16_round_fn() if emulate_bf16 else (lambda t: t)
    for s in R.LAYERS:
        p = params[s.index]
        if s.kind == 'head':
            k = p['kernel'].permute(3, 2, 0, 1)  # HWIO -> OIHW
            pred = torch.sigmoid(F.conv2d(rnd(outs[17]), k, p['bias']))
            continue
        if len(s.src) == 2:
            inp = torch.cat([rnd(outs[s.src[0]]), rnd(outs[s.src[1]])], dim=1)   # (skip, up): model.py:246
        elif s.src[0] < 0:
            inp = x
        else:
            inp = rnd(pooled[s.src[0]] if s.src_pooled else outs[s.src[0]])
        kern = p['kernel']
        if emulate_bf16 and s.index != 0:
            kern = kern + (kern.detach().to(torch.bfloat16).to(dtype) - kern.detach())   # bf16 value, identity gradient
        if s.kind == 'conv':
            z = F.conv2d(inp, kern.permute(3, 2, 0, 1), p['bias'], padding=1)
        else:
            z = F.conv_transpose2d(inp, kern.permute(3, 2, 0, 1), p['bias'], stride=2)  # [a,b,Co,Ci] -> [Ci,Co,a,b]
        z = rnd(torch.relu(z))
        m = z.mean(dim=(0, 2, 3))
        v = z.var(dim=(0, 2, 3), unbiased=False)
        a = p['gamma'][None, :, None, None] * (z - m[None, :, None, None]) / torch.sqrt(v[None, :, None, None] + R.BN_EPS) \
            + p['beta'][None, :, None, None]
        if s.dropout and drop_seed is not None:
            n, c, h, w_ = a.shape
            keep = dropout_keep_mask((n, h, w_, c), layer_dropout_seed(drop_seed, s.index))
            keep_t = torch.from_numpy(np.ascontiguousarray(keep.transpose(0, 3, 1, 2))).to(dtype)
            a = a * keep_t * (1.0 / (1.0 - R.DROPOUT_RATE))
        a = rnd(a)
        outs[s.index] = a
        zs[s.index] = z
        means[s.index] = m.detach().numpy().astype(np.float64)
        vars_[s.index] = v.detach().numpy().astype(np.float64)
        if s.pool:
            pooled[s.index] = F.max_pool2d(a, 2, 2)
    count = float(loss_count) if loss_count is not None else float(pred.numel())
    sq = ((pred - y) ** 2).sum()
    loss = sq / count
    loss.backward()
    acc = (pred.argmax(dim=1) == y.argmax(dim=1)).to(torch.float64).mean().item()
    grads = []
    for s in R.LAYERS:
        p = params[s.index]
        grads.append({k: v.grad.detach().numpy().astype(np.float64) for k, v in p.items()})
    res = dict(loss=float(loss.item()), sq_err_sum=float(sq.item()), acc=acc,
               pred=pred.detach().permute(0, 2, 3, 1).contiguous().numpy(), grads=grads, batch_mean=means,
               batch_var=vars_)
    if return_acts:
        nhwc = lambda t: t.detach().permute(0, 2, 3, 1).contiguous().numpy()
        res['z'] = {k: nhwc(v) for k, v in zs.items()}
        res['a'] = {k: nhwc(v) for k, v in outs.items()}
    return res


def loss_only(weights, x_nhwc, y_nhwc, dtype=None) -> float:
    """mse of the training-mode forward without dropout (for finite-difference checks)."""
    import torch
    with torch.enable_grad():
        return train_forward_backward(weights, x_nhwc, y_nhwc, None, dtype)['loss']


def rmsprop_step(p: np.ndarray, g: np.ndarray, a: np.ndarray, lr=1e-3, rho=0.9, eps=1e-7):
    """Keras RMSprop.get_updates: a = rho a + (1-rho) g^2; p = p - lr g / (sqrt(a) + eps). Returns (p, a)."""
    a = rho * a + (1.0 - rho) * g * g
    return p - lr * g / (np.sqrt(a) + eps), a


def adam_step(p, g, m, v, t: int, lr=1e-3, beta_1=0.5, beta_2=0.999, eps=1e-7):
    """Keras Adam.get_updates (t = 1 for the first step). Returns (p, m, v)."""
    lr_t = lr * np.sqrt(1.0 - beta_2 ** t) / (1.0 - beta_1 ** t)
    m = beta_1 * m + (1.0 - beta_1) * g
    v = beta_2 * v + (1.0 - beta_2) * g * g
    return p - lr_t * m / (np.sqrt(v) + eps), m, v


def moving_update(moving_mean, moving_var, batch_mean, batch_var, count: float, momentum=R.BN_MOMENTUM, eps=R.BN_EPS):
    """Keras 2.2.x BatchNormalization.call in training: moving = momentum*moving + (1-momentum)*batch, the variance
    after the sample-size correction var * M / (M - (1 + eps))."""
    vs = batch_var * (count / (count - (1.0 + eps)))
    return (momentum * moving_mean + (1 - momentum) * batch_mean, momentum * moving_var + (1 - momentum) * vs)


def synthetic_pair(n: int, h: int, w: int, seed: int = 0):
    """Synthetic SDR input / tone-mapped target pair with the statistics of the reference's data flow: a log-normal HDR
    field, SDR = clipped gamma-encoded exposure (virtual_camera.m:13-28 in spirit), target = Reinhard global operator
    L/(1+L) (Reinhard.m:19-26). x uint8 [n,h,w,3], y float32 [n,h,w,3] in [0,1]."""
    rng = np.random.default_rng(seed)
    base = rng.normal(0.0, 1.0, size=(n, h // 8 + 2, w // 8 + 2, 3))
    up = np.repeat(np.repeat(base, 8, axis=1), 8, axis=2)[:, :h, :w, :]
    hdr = np.exp(1.2 * up + 0.15 * rng.normal(size=(n, h, w, 3)))
    sdr = np.clip(hdr * 0.6, 0, 1) ** (1 / 2.2)
    x = np.floor(sdr * 255.0 + 0.5).astype(np.uint8)
    lum = hdr / (1.0 + hdr)
    return x, lum.astype(np.float32)
