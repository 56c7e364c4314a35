"""Static description of the CNN-ITMO U-Net graph (reference model.py:204-278) and its Keras summary.

The table below is the host-side mirror of ``kLayers`` in csrc/itmo_api.cu: the 19 weight-carrying
layers in execution order. Everything else in the Keras graph (Activation, BatchNormalization,
MaxPooling2D, Dropout, Concatenate) is fused into the kernels of these layers, but ``summary()``
still lists all 66 Keras layers exactly as ``model.summary()`` does (model.py:288, layers.txt).
"""
from __future__ import annotations

from collections import namedtuple
from typing import Callable, Dict, List, Tuple

Layer = namedtuple('Layer', 'index kind name bn cin cout hin pool dropout src src_pooled')

IMG = 512          # model.py:204 input_size=(512,512,3)
BN_EPS = 1e-3      # Keras BatchNormalization default
MODEL_NAME = 'ReinhardtPrediction'  # model.py:278


def _build() -> List[Layer]:
    L: List[Layer] = []
    # encoder + bottleneck: conv2d_1 .. conv2d_10 (model.py:208-238)
    widths = [(3, 32), (32, 32), (32, 64), (64, 64), (64, 128), (128, 128), (128, 256), (256, 256), (256, 512),
              (512, 512)]
    for i, (ci, co) in enumerate(widths):
        L.append(Layer(i, 'conv', f'conv2d_{i + 1}', f'batch_normalization_{i + 1}', (ci,), co, IMG >> (i // 2),
                       i in (1, 3, 5, 7), i in (7, 9), (i - 1,), i in (2, 4, 6, 8)))
    # decoder (model.py:245-262): ConvBNTranspose -> concatenate([skip, up]) -> ConvBN
    skips = [(7, 256), (5, 128), (3, 64), (1, 32)]
    up_in = 512
    for t, (skip, cskip) in enumerate(skips):
        up_out = (512, 256, 128, 64)[t]
        h = 32 << t
        ti = 10 + 2 * t
        L.append(Layer(ti, 'convT', f'conv2d_transpose_{t + 1}', f'batch_normalization_{11 + 2 * t}', (up_in,), up_out,
                       h, False, False, (ti - 1,), False))
        L.append(Layer(ti + 1, 'conv', f'conv2d_{11 + t}', f'batch_normalization_{12 + 2 * t}', (cskip, up_out),
                       up_out, 2 * h, False, False, (skip, ti), False))
        up_in = up_out
    L.append(Layer(18, 'head', 'conv2d_15', None, (64,), 3, IMG, False, False, (17,), False))  # model.py:276
    return L


LAYERS: List[Layer] = _build()
NUM_LAYERS = len(LAYERS)


def kernel_shape(l: Layer) -> Tuple[int, int, int, int]:
    """Keras kernel shape: Conv2D HWIO, Conv2DTranspose [kh,kw,Cout,Cin]."""
    cin = sum(l.cin)
    if l.kind == 'conv':
        return (3, 3, cin, l.cout)
    if l.kind == 'convT':
        return (2, 2, l.cout, cin)
    return (1, 1, cin, l.cout)


def count_params() -> Dict[str, int]:
    total = frozen = 0
    for l in LAYERS:
        ks = kernel_shape(l)
        total += ks[0] * ks[1] * ks[2] * ks[3] + l.cout
        if l.bn:
            total += 4 * l.cout
            frozen += 2 * l.cout  # moving_mean / moving_variance
    return {'total': total, 'trainable': total - frozen, 'non_trainable': frozen}


def forward_flops_per_image() -> float:
    """2*MACs of the 19 conv layers for one 512x512 image (165.37 GFLOP)."""
    f = 0.0
    for l in LAYERS:
        ks = kernel_shape(l)
        if l.kind == 'convT':
            f += 2.0 * (2 * l.hin) ** 2 * ks[2] * ks[3]
        else:
            f += 2.0 * l.hin ** 2 * ks[0] * ks[1] * ks[2] * ks[3]
    return f


# ----------------------------------------------------------------------------------------------
# Keras layer list (66 layers) in model.layers order, and the summary text
# ----------------------------------------------------------------------------------------------
KerasLayer = namedtuple('KerasLayer', 'name cls shape params inbound weight_layer')


def keras_layers() -> List[KerasLayer]:
    made: List[KerasLayer] = []
    counters = {'activation': 0, 'max_pooling2d': 0, 'dropout': 0, 'concatenate': 0}

    def mk(name, cls, shape, params, inbound, wl=None):
        made.append(KerasLayer(name, cls, shape, params, tuple(inbound), wl))
        return name

    def nxt(prefix):
        counters[prefix] += 1
        return f'{prefix}_{counters[prefix]}'

    mk('input_1', 'InputLayer', (None, IMG, IMG, 3), 0, ())
    tensor_of: Dict[int, str] = {}
    pooled_of: Dict[int, str] = {}
    for l in LAYERS:
        ks = kernel_shape(l)
        nparam = ks[0] * ks[1] * ks[2] * ks[3] + l.cout
        if l.kind == 'head':
            mk(l.name, 'Conv2D', (None, l.hin, l.hin, l.cout), nparam, (tensor_of[l.src[0]],), l.index)
            break
        if len(l.src) == 2:
            src = mk(nxt('concatenate'), 'Concatenate', (None, l.hin, l.hin, sum(l.cin)), 0,
                     (tensor_of[l.src[0]], tensor_of[l.src[1]]))
        elif l.src[0] < 0:
            src = 'input_1'
        else:
            src = pooled_of[l.src[0]] if l.src_pooled else tensor_of[l.src[0]]
        ho = 2 * l.hin if l.kind == 'convT' else l.hin
        shp = (None, ho, ho, l.cout)
        t = mk(l.name, 'Conv2D' if l.kind == 'conv' else 'Conv2DTranspose', shp, nparam, (src,), l.index)
        t = mk(nxt('activation'), 'Activation', shp, 0, (t,))
        t = mk(l.bn, 'BatchNormalization', shp, 4 * l.cout, (t,), l.index)
        if l.dropout:
            t = mk(nxt('dropout'), 'Dropout', shp, 0, (t,))
        tensor_of[l.index] = t
        if l.pool:
            pooled_of[l.index] = mk(nxt('max_pooling2d'), 'MaxPooling2D', (None, ho // 2, ho // 2, l.cout), 0, (t,))
    # Keras sorts model.layers by depth (longest path to the output), creation order breaking ties
    pos = {k.name: i for i, k in enumerate(made)}
    depth = {k.name: 0 for k in made}
    for k in reversed(made):
        for src in k.inbound:
            depth[src] = max(depth[src], depth[k.name] + 1)
    return sorted(made, key=lambda k: (-depth[k.name], pos[k.name]))


def summary_text(line_length: int = 98, positions=(.33, .55, .67, 1.)) -> str:
    """The text Keras 2.2.x ``model.summary()`` prints for this graph (compare reference layers.txt)."""
    cols = [int(line_length * p) for p in positions]
    out: List[str] = []

    def emit(fields):
        line = ''
        for i, field in enumerate(fields):
            if i:
                line = line[:-1] + ' '
            line = (line + str(field))[:cols[i]]
            line += ' ' * (cols[i] - len(line))
        out.append(line)

    out.append('_' * line_length)
    emit(['Layer (type)', 'Output Shape', 'Param #', 'Connected to'])
    out.append('=' * line_length)
    layers = keras_layers()
    for i, k in enumerate(layers):
        conn = [f'{s}[0][0]' for s in k.inbound]
        emit([f'{k.name} ({k.cls})', k.shape, k.params, conn[0] if conn else ''])
        for extra in conn[1:]:
            emit(['', '', '', extra])
        out.append(('=' if i + 1 == len(layers) else '_') * line_length)
    c = count_params()
    out.append(f"Total params: {c['total']:,}")
    out.append(f"Trainable params: {c['trainable']:,}")
    out.append(f"Non-trainable params: {c['non_trainable']:,}")
    out.append('_' * line_length)
    return '\n'.join(out) + '\n'


def print_summary(print_fn: Callable[[str], None] = print):
    for line in summary_text().rstrip('\n').split('\n'):
        print_fn(line)
