"""CPU restatement of the CNN-ITMO U-Net hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module. The product path (``cnn-itmo_b200``) never does: it fails loudly when
the CUDA library is missing instead of falling back to anything in here.

PARITY UNPINNED (numerics): the reference's arithmetic lives in un-vendored, un-pinned Keras +
TensorFlow (reference README.md:11-15); neither can be installed here (no wheels, no network, TF1
does not run on Python 3.12) and the checkpoint ``saved7-model-218-0.73.hdf5`` named at
predict.py:24 is stripped from the reference tree (.MISSING_LARGE_BLOBS:1). The reference has no
tests or golden vectors for the path. What IS pinned: the topology / shape / parameter-count
contract of the reference's own ``layers.txt`` (tests/test_oracle_cpu.py regenerates that table from
this module and compares every row and the totals), and the two independent restatements below
(numpy float64 loops vs torch CPU) against each other.

Two restatements of the same graph:
  (A) numpy float64, written as explicit tap loops + einsum — slow, unambiguous, small shapes;
  (B) torch CPU (oneDNN) float32/float64 — fast enough for 512x512 batches; gated on agreeing
      with (A) in float64 to 1e-10.

Reference semantics encoded here (file:line in /root/reference):
  model.py:195-196  ConvBN          = BN(ReLU(Conv2D(f, 3, padding='same', he_normal)(x)))
  model.py:199-200  ConvBNTranspose = BN(ReLU(Conv2DTranspose(f, 2, strides=2, padding='valid')(x)))
  model.py:204-278  U_net wiring, skip order concatenate([skip, up], axis=3), Dropout(0.5) placement
  model.py:276      head Conv2D(3, 1, activation='sigmoid')
  model.py:281      compile(optimizer='rmsprop', loss='mse', metrics=['accuracy'])
  predict.py:59-64  u8/255 in float64 -> float32 predict -> (pred*255).astype('uint8') truncation
Keras 2.2.x defaults: Conv2D = cross-correlation, kernel HWIO, zero pad 1 for 'same'; Conv2DTranspose
kernel [kh,kw,Cout,Cin], out[2i+a,2j+b,co] = sum_ci x[i,j,ci] W[a,b,co,ci] + b[co]; BatchNormalization
axis=-1, eps=1e-3, momentum=0.99, weight order gamma,beta,moving_mean,moving_variance; MaxPooling2D
'valid'; Dropout inverted scaling (identity at inference); RMSprop lr=1e-3, rho=0.9, eps=1e-7.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

BN_EPS = 1e-3        # keras.layers.BatchNormalization default epsilon
BN_MOMENTUM = 0.99   # keras default momentum
DROPOUT_RATE = 0.5   # model.py:226,239
IMG = 512            # model.py:204


# --------------------------------------------------------------------------------------------
# graph description (execution order of the 19 weight layers; model.py:208-276)
# --------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class LayerSpec:
    index: int
    kind: str            # 'conv' | 'convT' | 'head'
    keras_name: str      # name in layers.txt
    bn_name: Optional[str]
    cin: Tuple[int, ...]  # channels of each source (skip first, up second)
    cout: int
    hin: int             # input spatial size at 512x512
    pool: bool           # followed by MaxPooling2D(2,2)
    dropout: bool        # followed by Dropout(0.5) (before the pool / skip)
    src: Tuple[int, ...]  # producing layer indices (-1 = network input)
    src_pooled: bool


def _specs() -> List[LayerSpec]:
    L = []
    enc = [(3, 32, 512), (32, 32, 512), (32, 64, 256), (64, 64, 256), (64, 128, 128), (128, 128, 128),
           (128, 256, 64), (256, 256, 64), (256, 512, 32), (512, 512, 32)]
    for i, (ci, co, h) in enumerate(enc):
        L.append(LayerSpec(i, 'conv', f'conv2d_{i + 1}', f'batch_normalization_{i + 1}', (ci,), co, h,
                           pool=i in (1, 3, 5, 7), dropout=i in (7, 9), src=(i - 1,),
                           src_pooled=i in (2, 4, 6, 8)))
    # decoder: (up layer, conv layer) x 4; skips are layers 7, 5, 3, 1
    dec = [(512, 512, 32, 7, 256), (512, 256, 64, 5, 128), (256, 128, 128, 3, 64), (128, 64, 256, 1, 32)]
    idx, bn, conv_no = 10, 11, 11
    prev = 9
    for t, (ci, co, h, skip, cskip) in enumerate(dec):
        L.append(LayerSpec(idx, 'convT', f'conv2d_transpose_{t + 1}', f'batch_normalization_{bn}', (ci,), co, h,
                           False, False, (prev,), False))
        L.append(LayerSpec(idx + 1, 'conv', f'conv2d_{conv_no}', f'batch_normalization_{bn + 1}', (cskip, co), co,
                           2 * h, False, False, (skip, idx), False))
        prev = idx + 1
        idx += 2
        bn += 2
        conv_no += 1
    L.append(LayerSpec(18, 'head', 'conv2d_15', None, (64,), 3, 512, False, False, (17,), False))
    return L


LAYERS: List[LayerSpec] = _specs()
NUM_LAYERS = len(LAYERS)


def kernel_shape(spec: LayerSpec) -> Tuple[int, int, int, int]:
    cin = sum(spec.cin)
    if spec.kind == 'conv':
        return (3, 3, cin, spec.cout)
    if spec.kind == 'convT':
        return (2, 2, spec.cout, cin)
    return (1, 1, cin, spec.cout)


def param_count() -> Dict[str, int]:
    """Total / trainable / non-trainable parameter counts (layers.txt:140-142)."""
    total = nontrain = 0
    for s in LAYERS:
        total += int(np.prod(kernel_shape(s))) + s.cout
        if s.bn_name:
            total += 4 * s.cout
            nontrain += 2 * s.cout
    return {'total': total, 'trainable': total - nontrain, 'non_trainable': nontrain}


# --------------------------------------------------------------------------------------------
# Keras model.summary() table (model.py:288) — regenerated to pin the oracle against layers.txt
# --------------------------------------------------------------------------------------------
def keras_layer_rows(input_size=(512, 512, 3)) -> List[Tuple[str, str, Tuple, int, List[str]]]:
    """Rows (name, class, output_shape, params, inbound) in Keras' summary order.

    Keras orders model.layers by depth from the output; within the U-Net this is creation order of
    each block except that the encoder skip tensors' consumers sit deeper than the decoder. The order
    below is the order layers.txt shows (it is reproduced by a topological sort by depth, ties by
    creation order).
    """
    H = input_size[0]
    C = input_size[2]
    nodes = []  # (name, cls, shape, params, inbound, creation index)

    def add(name, cls, shape, params, inbound):
        nodes.append([name, cls, shape, params, list(inbound)])
        return name

    act = [0]
    pool = [0]
    drop = [0]
    cat = [0]

    def block(spec: LayerSpec, inbound: str, h_out: int) -> str:
        cls = 'Conv2D' if spec.kind == 'conv' else 'Conv2DTranspose'
        n = add(spec.keras_name, cls, (None, h_out, h_out, spec.cout),
                int(np.prod(kernel_shape(spec))) + spec.cout, [inbound])
        act[0] += 1
        a = add(f'activation_{act[0]}', 'Activation', (None, h_out, h_out, spec.cout), 0, [n])
        return add(spec.bn_name, 'BatchNormalization', (None, h_out, h_out, spec.cout), 4 * spec.cout, [a])

    x = add('input_1', 'InputLayer', (None, H, H, C), 0, [])
    outs: Dict[int, str] = {}
    pooled: Dict[int, str] = {}
    for s in LAYERS:
        hin = s.hin * H // IMG
        if s.kind == 'head':
            add(s.keras_name, 'Conv2D', (None, hin, hin, s.cout), int(np.prod(kernel_shape(s))) + s.cout,
                [outs[s.src[0]]])
            continue
        if len(s.src) == 2:
            cat[0] += 1
            inbound = add(f'concatenate_{cat[0]}', 'Concatenate', (None, hin, hin, sum(s.cin)), 0,
                          [outs[s.src[0]], outs[s.src[1]]])
        elif s.src[0] < 0:
            inbound = x
        else:
            inbound = pooled[s.src[0]] if s.src_pooled else outs[s.src[0]]
        h_out = 2 * hin if s.kind == 'convT' else hin
        o = block(s, inbound, h_out)
        if s.dropout:
            drop[0] += 1
            o = add(f'dropout_{drop[0]}', 'Dropout', (None, h_out, h_out, s.cout), 0, [o])
        outs[s.index] = o
        if s.pool:
            pool[0] += 1
            pooled[s.index] = add(f'max_pooling2d_{pool[0]}', 'MaxPooling2D', (None, h_out // 2, h_out // 2, s.cout),
                                  0, [o])
    # Keras summary order: layers sorted by depth (max distance to the output), descending, ties by
    # creation order (keras.engine.network.Network._init_graph_network).
    idx = {n[0]: i for i, n in enumerate(nodes)}
    depth = {n[0]: 0 for n in nodes}
    for n in reversed(nodes):  # creation order is topological
        for inb in n[4]:
            depth[inb] = max(depth[inb], depth[n[0]] + 1)
    order = sorted(nodes, key=lambda n: (-depth[n[0]], idx[n[0]]))
    return [(n[0], n[1], n[2], n[3], n[4]) for n in order]


def keras_summary_text(input_size=(512, 512, 3)) -> str:
    """Text of model.summary() as Keras 2.2.x prints it (line_length 98, positions .33/.55/.67/1.)."""
    line_length = 98
    positions = [int(line_length * p) for p in (.33, .55, .67, 1.)]
    lines: List[str] = []

    def row(fields):
        line = ''
        for i, f in enumerate(fields):
            if i > 0:
                line = line[:-1] + ' '
            line += str(f)
            line = line[:positions[i]]
            line += ' ' * (positions[i] - len(line))
        lines.append(line)

    lines.append('_' * line_length)
    row(['Layer (type)', 'Output Shape', 'Param #', 'Connected to'])
    lines.append('=' * line_length)
    rows = keras_layer_rows(input_size)
    for i, (name, cls, shape, params, inbound) in enumerate(rows):
        conns = [f'{n}[0][0]' for n in inbound]
        row([f'{name} ({cls})', str(shape), params, conns[0] if conns else ''])
        for c in conns[1:]:
            row(['', '', '', c])
        lines.append(('=' if i == len(rows) - 1 else '_') * line_length)
    pc = param_count()
    lines.append('Total params: {:,}'.format(pc['total']))
    lines.append('Trainable params: {:,}'.format(pc['trainable']))
    lines.append('Non-trainable params: {:,}'.format(pc['non_trainable']))
    lines.append('_' * line_length)
    return '\n'.join(lines) + '\n'


# --------------------------------------------------------------------------------------------
# weights
# --------------------------------------------------------------------------------------------
def synthetic_weights(seed: int = 218, dtype=np.float32) -> List[Dict[str, np.ndarray]]:
    """Seeded stand-in for the stripped checkpoint: he-normal kernels (model.py:196,200), small biases and
    deliberately NON-trivial BatchNormalization statistics (negative gammas on some channels: this is
    what breaks a 'pool before BN' or 'fold BN into the next conv' shortcut). Head: glorot-uniform."""
    rng = np.random.default_rng(seed)
    ws = []
    for s in LAYERS:
        ks = kernel_shape(s)
        if s.kind == 'head':
            fan_in, fan_out = ks[2], ks[3]
            lim = math.sqrt(6.0 / (fan_in + fan_out))
            k = rng.uniform(-lim, lim, ks)
            ws.append({'kernel': k.astype(dtype), 'bias': rng.uniform(-0.1, 0.1, s.cout).astype(dtype)})
            continue
        fan_in = ks[0] * ks[1] * (ks[2] if s.kind == 'conv' else ks[3])
        k = rng.normal(0.0, math.sqrt(2.0 / fan_in), ks)
        gamma = rng.normal(1.0, 0.25, s.cout)
        gamma[rng.random(s.cout) < 0.1] *= -1.0
        ws.append({
            'kernel': k.astype(dtype),
            'bias': rng.uniform(-0.1, 0.1, s.cout).astype(dtype),
            'gamma': gamma.astype(dtype),
            'beta': rng.normal(0.0, 0.1, s.cout).astype(dtype),
            'moving_mean': rng.normal(0.3, 0.1, s.cout).astype(dtype),
            'moving_variance': rng.uniform(0.05, 1.0, s.cout).astype(dtype),
        })
    return ws


# --------------------------------------------------------------------------------------------
# (A) numpy float64 restatement
# --------------------------------------------------------------------------------------------
def conv2d_same_np(x: np.ndarray, k: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Keras Conv2D(padding='same', strides=1): out[n,y,x,o] = b[o] + sum_{i,j,c} in[n,y+i-ph,x+j-pw,c] k[i,j,c,o]."""
    n, h, w, c = x.shape
    kh, kw, ci, co = k.shape
    assert ci == c
    ph, pw = (kh - 1) // 2, (kw - 1) // 2
    xp = np.zeros((n, h + kh - 1, w + kw - 1, c), dtype=np.float64)
    xp[:, ph:ph + h, pw:pw + w, :] = x
    out = np.zeros((n, h, w, co), dtype=np.float64)
    for i in range(kh):
        for j in range(kw):
            out += np.einsum('nhwc,co->nhwo', xp[:, i:i + h, j:j + w, :], k[i, j].astype(np.float64))
    return out + b.astype(np.float64)


def conv2d_transpose_k2s2_np(x: np.ndarray, k: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Keras Conv2DTranspose(f, 2, strides=2, padding='valid'); kernel [2,2,Cout,Cin]."""
    n, h, w, c = x.shape
    assert k.shape[0] == 2 and k.shape[1] == 2 and k.shape[3] == c
    co = k.shape[2]
    out = np.zeros((n, 2 * h, 2 * w, co), dtype=np.float64)
    for a in range(2):
        for bb in range(2):
            out[:, a::2, bb::2, :] = np.einsum('nhwc,oc->nhwo', x.astype(np.float64), k[a, bb].astype(np.float64))
    return out + b.astype(np.float64)


def bn_infer_np(x, gamma, beta, mean, var, eps=BN_EPS):
    g, be, m, v = (a.astype(np.float64) for a in (gamma, beta, mean, var))
    return g * (x - m) / np.sqrt(v + eps) + be


def maxpool2_np(x):
    n, h, w, c = x.shape
    return x.reshape(n, h // 2, 2, w // 2, 2, c).max(axis=(2, 4))


def sigmoid_np(x):
    return 1.0 / (1.0 + np.exp(-x))


def block_np(spec: LayerSpec, w: Dict[str, np.ndarray], x: np.ndarray) -> np.ndarray:
    """ConvBN / ConvBNTranspose (model.py:195-200), inference mode."""
    if spec.kind == 'conv':
        z = conv2d_same_np(x, w['kernel'], w['bias'])
    else:
        z = conv2d_transpose_k2s2_np(x, w['kernel'], w['bias'])
    z = np.maximum(z, 0.0)
    return bn_infer_np(z, w['gamma'], w['beta'], w['moving_mean'], w['moving_variance'])


def forward_np(weights: Sequence[Dict[str, np.ndarray]], x: np.ndarray, return_all: bool = False):
    """U_net forward (model.py:204-278), inference mode, float64. x: [N,H,W,3] in [0,1], H,W % 16 == 0."""
    x = np.asarray(x, dtype=np.float32).astype(np.float64)  # Keras casts the fed array to float32 first
    outs: Dict[int, np.ndarray] = {}
    pooled: Dict[int, np.ndarray] = {}
    for s in LAYERS:
        w = weights[s.index]
        if s.kind == 'head':
            z = np.einsum('nhwc,co->nhwo', outs[17], w['kernel'][0, 0].astype(np.float64)) + w['bias'].astype(np.float64)
            outs[18] = sigmoid_np(z)
            continue
        if len(s.src) == 2:
            inp = np.concatenate([outs[s.src[0]], outs[s.src[1]]], axis=3)  # skip first, up second
        elif s.src[0] < 0:
            inp = x
        else:
            inp = pooled[s.src[0]] if s.src_pooled else outs[s.src[0]]
        outs[s.index] = block_np(s, w, inp)  # Dropout: identity at inference
        if s.pool:
            pooled[s.index] = maxpool2_np(outs[s.index])
    if return_all:
        return outs, pooled
    return outs[18]


# --------------------------------------------------------------------------------------------
# pre / post processing (predict.py:59-64)
# --------------------------------------------------------------------------------------------
def preprocess_u8(img_u8: np.ndarray) -> np.ndarray:
    """np.true_divide(np.asarray(img).astype(float), 255) (predict.py:59): float64 in [0,1]."""
    return np.true_divide(np.asarray(img_u8).astype(float), 255)


def postprocess_u8(pred_f32: np.ndarray) -> np.ndarray:
    """(pred * 255).astype('uint8') (predict.py:64): float32 multiply, truncation toward zero."""
    return (np.asarray(pred_f32, dtype=np.float32) * 255).astype('uint8')


def psnr(a: np.ndarray, b: np.ndarray, peak: float = 1.0) -> float:
    mse = float(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))
    return float('inf') if mse == 0 else 10.0 * math.log10(peak * peak / mse)


# --------------------------------------------------------------------------------------------
# (B) torch CPU restatement (oneDNN) — used for 512x512 batches and as the timed CPU baseline
# --------------------------------------------------------------------------------------------
def _torch():
    import torch
    return torch


def _round_bf16(t):
    torch = _torch()
    return t.to(torch.bfloat16).to(t.dtype)


def _trunc_tf32(t):
    """TF32 operand view of an fp32 tensor: the tensor core ignores the low 13 mantissa bits."""
    torch = _torch()
    if t.dtype != torch.float32:
        return t
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


class TorchWeights:
    """Weights converted once to torch tensors in conv-friendly (OIHW) layouts."""

    def __init__(self, weights: Sequence[Dict[str, np.ndarray]], dtype=None, emulate: Optional[str] = None):
        torch = _torch()
        dtype = dtype or torch.float32
        self.dtype = dtype
        self.emulate = emulate
        self.layers = []
        for s in LAYERS:
            w = weights[s.index]
            k = torch.from_numpy(np.ascontiguousarray(w['kernel'])).to(dtype)
            if s.kind in ('conv', 'head'):
                k = k.permute(3, 2, 0, 1).contiguous()        # HWIO -> OIHW
            else:
                k = k.permute(3, 2, 0, 1).contiguous()        # [kh,kw,Cout,Cin] -> [Cin,Cout,kh,kw]
            tensor_core_layer = s.kind != 'head' and s.index != 0
            if emulate == 'bf16' and tensor_core_layer:
                k = _round_bf16(k)
            elif emulate == 'tf32' and tensor_core_layer:
                k = _trunc_tf32(k)
            d = {'kernel': k, 'bias': torch.from_numpy(np.asarray(w['bias'])).to(dtype)}
            if s.bn_name:
                g = torch.from_numpy(np.asarray(w['gamma'])).to(dtype)
                be = torch.from_numpy(np.asarray(w['beta'])).to(dtype)
                m = torch.from_numpy(np.asarray(w['moving_mean'])).to(dtype)
                v = torch.from_numpy(np.asarray(w['moving_variance'])).to(dtype)
                d.update(gamma=g, beta=be, mean=m, var=v)
            self.layers.append(d)


def block_torch(spec: LayerSpec, w: Dict, x, emulate: Optional[str] = None):
    """ConvBN / ConvBNTranspose on NCHW torch tensors. With ``emulate`` the device's storage rounding is
    mimicked: 'bf16' = operands and the stored result are bf16-rounded (fp32 accumulate); 'tf32' = operands
    truncated to 10 mantissa bits. Layer 0 runs in plain fp32 on the device in both modes."""
    torch = _torch()
    F = torch.nn.functional
    tensor_core_layer = spec.index != 0
    if emulate == 'tf32' and tensor_core_layer:
        x = _trunc_tf32(x)
    if spec.kind == 'conv':
        z = F.conv2d(x, w['kernel'], w['bias'], padding=1)
    else:
        z = F.conv_transpose2d(x, w['kernel'], w['bias'], stride=2)
    z = torch.relu(z)
    if emulate is not None:
        # the device folds BN to one fused multiply-add: s = g/sqrt(v+eps), t = b - m*s
        s = w['gamma'] / torch.sqrt(w['var'] + BN_EPS)
        t = w['beta'] - w['mean'] * s
        y = z * s[None, :, None, None] + t[None, :, None, None]
    else:
        y = (w['gamma'][None, :, None, None] * (z - w['mean'][None, :, None, None])
             / torch.sqrt(w['var'][None, :, None, None] + BN_EPS) + w['beta'][None, :, None, None])
    return y


def forward_torch(tw: TorchWeights, x_nhwc: np.ndarray, return_all: bool = False):
    """U_net forward on torch CPU. x_nhwc: [N,H,W,3] float in [0,1]. Returns NHWC numpy (dtype of tw)."""
    torch = _torch()
    F = torch.nn.functional
    emulate = tw.emulate
    x = torch.from_numpy(np.ascontiguousarray(np.asarray(x_nhwc, dtype=np.float32))).to(tw.dtype)
    x = x.permute(0, 3, 1, 2).contiguous()
    outs: Dict[int, 'torch.Tensor'] = {}
    pooled: Dict[int, 'torch.Tensor'] = {}
    with torch.no_grad():
        for s in LAYERS:
            w = tw.layers[s.index]
            if s.kind == 'head':
                z = F.conv2d(outs[17], w['kernel'], w['bias'])
                outs[18] = torch.sigmoid(z)
                continue
            if len(s.src) == 2:
                inp = torch.cat([outs[s.src[0]], outs[s.src[1]]], dim=1)
            elif s.src[0] < 0:
                inp = x
            else:
                inp = pooled[s.src[0]] if s.src_pooled else outs[s.src[0]]
            y = block_torch(s, w, inp, emulate)
            if emulate == 'bf16' and s.index != 17:
                y = _round_bf16(y)  # activations are stored in HBM as bf16; layer 17 feeds the fused head in fp32
            outs[s.index] = y
            if s.pool:
                pooled[s.index] = F.max_pool2d(y, 2, 2)
    to_np = lambda t: t.permute(0, 2, 3, 1).contiguous().numpy()
    if return_all:
        return {k: to_np(v) for k, v in outs.items()}, {k: to_np(v) for k, v in pooled.items()}
    return to_np(outs[18])


def predict_like_reference(tw: TorchWeights, images_u8: np.ndarray, one_at_a_time: bool = True) -> np.ndarray:
    """predict.py:50-64 on the CPU: per image u8/255 -> model.predict -> trunc(pred*255) u8."""
    outs = []
    if one_at_a_time:
        for i in range(images_u8.shape[0]):
            X = np.asarray([preprocess_u8(images_u8[i])])
            outs.append(postprocess_u8(forward_torch(tw, X))[0])
        return np.stack(outs)
    return postprocess_u8(forward_torch(tw, preprocess_u8(images_u8)))


# --------------------------------------------------------------------------------------------
# calibrated synthetic checkpoint
# --------------------------------------------------------------------------------------------
def synthetic_images_u8(n: int, seed: int = 0, size: int = IMG) -> np.ndarray:
    """Seeded SDR-like test images [n,size,size,3] uint8: smooth illumination gradients, a few hard edges and
    sensor-like noise, with per-image exposure — closer to camera output than white noise, and they exercise
    saturated (0 / 255) regions like the reference's virtual_camera.m clipping does."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:size, 0:size].astype(np.float64) / size
    out = np.empty((n, size, size, 3), dtype=np.uint8)
    for i in range(n):
        img = np.zeros((size, size, 3))
        for c in range(3):
            fx, fy, ph = rng.uniform(0.5, 6.0), rng.uniform(0.5, 6.0), rng.uniform(0, 2 * np.pi)
            img[:, :, c] = 0.5 + 0.35 * np.sin(2 * np.pi * (fx * xx + fy * yy) + ph)
        for _ in range(4):  # rectangles with hard edges
            x0, y0 = rng.integers(0, size - 8, 2)
            w, h = rng.integers(8, size // 2, 2)
            img[y0:y0 + h, x0:x0 + w, :] = rng.uniform(0, 1, 3)
        img *= 2.0 ** rng.uniform(-1.5, 1.5)           # exposure
        img += rng.normal(0, 0.02, img.shape)
        out[i] = np.clip(img * 255.0, 0, 255).astype(np.uint8)
    return out


def calibrated_weights(seed: int = 218, calib_size: int = 128, calib_n: int = 4) -> List[Dict[str, np.ndarray]]:
    """``synthetic_weights`` with BatchNormalization moving statistics set the way training would leave them:
    equal to the statistics of each layer's post-ReLU activations on a calibration batch. Activations then
    stay O(1) through all 19 layers (a freshly drawn moving_mean/variance makes them explode, which would
    saturate the sigmoid head and hide errors)."""
    torch = _torch()
    F = torch.nn.functional
    ws = synthetic_weights(seed)
    x = torch.from_numpy(preprocess_u8(synthetic_images_u8(calib_n, seed=seed + 1, size=calib_size))).float()
    x = x.permute(0, 3, 1, 2).contiguous()
    outs: Dict[int, 'torch.Tensor'] = {}
    pooled: Dict[int, 'torch.Tensor'] = {}
    with torch.no_grad():
        for s in LAYERS:
            if s.kind == 'head':
                break
            w = ws[s.index]
            if len(s.src) == 2:
                inp = torch.cat([outs[s.src[0]], outs[s.src[1]]], dim=1)
            elif s.src[0] < 0:
                inp = x
            else:
                inp = pooled[s.src[0]] if s.src_pooled else outs[s.src[0]]
            k = torch.from_numpy(w['kernel']).permute(3, 2, 0, 1).contiguous()
            b = torch.from_numpy(w['bias'])
            z = F.conv2d(inp, k, b, padding=1) if s.kind == 'conv' else F.conv_transpose2d(inp, k, b, stride=2)
            z = torch.relu(z)
            mean = z.mean(dim=(0, 2, 3))
            var = z.var(dim=(0, 2, 3), unbiased=False)
            w['moving_mean'] = mean.numpy().astype(np.float32)
            w['moving_variance'] = np.maximum(var.numpy(), 1e-4).astype(np.float32)
            g, be = torch.from_numpy(w['gamma']), torch.from_numpy(w['beta'])
            y = (g[None, :, None, None] * (z - mean[None, :, None, None])
                 / torch.sqrt(torch.from_numpy(w['moving_variance'])[None, :, None, None] + BN_EPS)
                 + be[None, :, None, None])
            outs[s.index] = y
            if s.pool:
                pooled[s.index] = F.max_pool2d(y, 2, 2)
    return ws


# --------------------------------------------------------------------------------------------
# frames larger than 512x512 (BASELINE config 4) and m-files/inverse_Reinhard.m
# --------------------------------------------------------------------------------------------
TILE = 512
TILE_HALO = 96   # >= 92 = receptive-field reach of the graph (3x3 convs and 2x2 pools of model.py:208-262)


def plan_axis_ref(extent: int) -> List[Tuple[int, int, int]]:
    """Independent restatement of the tile plan along one axis: [(origin, valid_lo, valid_hi)].
    Tiles stay inside the frame (so the frame border sees the network's own 'same' zero padding), start on
    the 16-pixel grid of the four 2x2 pools, and keep TILE_HALO pixels between a valid range and an interior cut."""
    assert extent >= TILE and extent % 16 == 0
    out, covered = [], 0
    while covered < extent:
        o = 0 if covered == 0 else covered - TILE_HALO
        o = min(o, extent - TILE)
        hi = extent if o + TILE == extent else o + TILE - TILE_HALO
        out.append((o, covered, hi))
        covered = hi
    return out


def predict_tiled_torch(tw: 'TorchWeights', frame_f32: np.ndarray) -> np.ndarray:
    """Whole frame [H,W,3] through 512x512 tiles of the plan, stitched (CPU). The reference itself cannot do this
    (fixed 512x512 Input, model.py:204); the property tested is tiled == fully-convolutional whole frame."""
    h, w, _ = frame_f32.shape
    out = np.empty((h, w, 3), dtype=np.float32)
    for (oy, y0, y1) in plan_axis_ref(h):
        for (ox, x0, x1) in plan_axis_ref(w):
            t = forward_torch(tw, frame_f32[None, oy:oy + TILE, ox:ox + TILE])[0]
            out[y0:y1, x0:x1] = t[y0 - oy:y1 - oy, x0 - ox:x1 - ox]
    return out


def inverse_reinhard_np(img_u8: np.ndarray, as_written: bool = False, key: float = 0.18) -> np.ndarray:
    """m-files/inverse_Reinhard.m:2-23 in numpy float64 on one uint8 image [H,W,3] (the script reads the PNG that
    predict.py:70 wrote). Returns float64 [H,W,3].

    as_written=True follows the script line by line, IEEE semantics of its degenerate cases included:
      :2  ImgIn = (u8/255)^2.2          :3  I = RGB2Lum(ImgIn)  (HDR-Toolbox, Rec.709: .2126 .7152 .0722)
      :4  X = I/(I-1)                   (sign error: the inverse of Reinhard.m:22 L = X/(1+X) is X = L/(1-L))
      :5-11 G_X = exp(mean(log(max(X, realmin))))
      :12-16 G_E = exp(P log(G_X)/PB1 - PB2 log(a)/PB1), PB1 = #(I==0), PB2 = #(I~=0)  (divides by the count of
             zero-luminance pixels)     :17 E = G_E X / a     :18 NaN -> realmin     :19 clamp at 2^32
      :20-23 out_c = ImgIn_c * E / I
    as_written=False is the repaired form implemented as the product's default: X = I'/(1-I') with
    I' = min(I, 1 - 0.5/255) (a saturated pixel must not drive the geometric mean to infinity), E = key * X / G_X
    with G_X the geometric mean of X over the NON-ZERO-luminance pixels (the recovered luminance is normalised to
    `key`, mirroring Reinhard.m:19-21; black pixels are left out of the log-average instead of contributing
    log(realmin) each), zero-luminance pixels stay black instead of NaN."""
    realmin = np.finfo(np.float64).tiny
    img = (img_u8.astype(np.float64) / 255.0) ** 2.2
    I = 0.2126 * img[..., 0] + 0.7152 * img[..., 1] + 0.0722 * img[..., 2]
    with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
        Ic = np.minimum(I, 1.0 - 0.5 / 255.0)
        X = I / (I - 1.0) if as_written else Ic / (1.0 - Ic)
        Xc = np.where(np.isnan(X), realmin, np.maximum(X, realmin))   # MATLAB max(NaN, realmin) = realmin
        P = float(I.size)
        if as_written:
            G_X = np.exp(np.log(Xc).sum() / P)
        else:
            nz = I != 0
            G_X = np.exp(np.log(Xc[nz]).sum() / float(nz.sum())) if nz.any() else 1.0
        if as_written:
            PB1 = float((I == 0).sum())
            PB2 = P - PB1
            a = 0.18
            G_E = np.exp(np.float64(P) * np.log(G_X) / np.float64(PB1) - np.float64(PB2) * np.log(a) / np.float64(PB1))
            E = G_E / a * X
        else:
            E = key / G_X * X
        E = np.where(np.isnan(E), realmin, E)
        E = np.where(E >= 2.0 ** 32, 2.0 ** 32, E)
        q = E / I
        if not as_written:
            q = np.where(I == 0, 0.0, q)
        return img * q[..., None]
