"""Keras-layout HDF5 checkpoints for the U-Net (what ``load_model`` reads at reference predict.py:24 and
``ModelCheckpoint`` writes at main.py:123-124).

File layout (Keras 2.2.x ``model.save``):
  /                      attrs: keras_version, backend, model_config (JSON), training_config (JSON)
  /model_weights         attrs: layer_names [S], backend, keras_version
  /model_weights/<layer> attrs: weight_names [S]   (one group per Keras layer, 66 of them, most empty)
  /model_weights/<layer>/<layer>/kernel:0, bias:0                       Conv2D / Conv2DTranspose
  /model_weights/<layer>/<layer>/gamma:0, beta:0, moving_mean:0, moving_variance:0   BatchNormalization
  /optimizer_weights     RMSprop accumulators (ignored for inference)
``model.save_weights`` files have the ``model_weights`` content at the root instead.

Layers are matched by TYPE ORDER, not by numeric suffix: the 19 kernel-carrying layers form a dependency
chain, so any topological order (``layer_names`` is one) lists them in execution order; every shape is
then checked against the graph table (== layers.txt). h5py is used when importable, else h5min.
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

from . import graph, h5min

try:  # pragma: no cover - h5py is absent in the build image
    import h5py  # type: ignore
except Exception:  # noqa: BLE001
    h5py = None

Weights = List[Dict[str, np.ndarray]]
BN_KEYS = ('gamma', 'beta', 'moving_mean', 'moving_variance')


def _open(path: str):
    if not os.path.exists(path):
        raise OSError(f'Unable to open file (unable to open file: name = {path!r}, errno = 2, '
                      f"error message = 'No such file or directory')")
    if h5py is not None:  # pragma: no cover
        return h5py.File(path, 'r')
    return h5min.File(path)


def _as_str(v) -> str:
    if isinstance(v, bytes):
        return v.decode('utf-8')
    if isinstance(v, np.ndarray) and v.shape == ():
        return _as_str(v[()])
    return str(v)


def _names(attr) -> List[str]:
    if attr is None:
        return []
    arr = np.asarray(attr)
    if arr.size == 0:
        return []
    return [_as_str(x) for x in arr.ravel()]


def read_checkpoint(path: str) -> Tuple[Weights, Dict[str, Any]]:
    """Returns (weights in execution order, meta). Raises OSError for a missing / corrupt file and ValueError
    when the file holds a different architecture."""
    f = _open(path)
    try:
        root = f['model_weights'] if 'model_weights' in f else f
        attrs = root.attrs
        if 'layer_names' in attrs:
            layer_names = _names(attrs['layer_names'])
        else:  # Keras chunks very long attribute lists: layer_names0, layer_names1, ...
            layer_names, i = [], 0
            while f'layer_names{i}' in attrs:
                layer_names += _names(attrs[f'layer_names{i}'])
                i += 1
        if not layer_names:
            raise ValueError(f'{path}: no layer_names attribute — not a Keras weights file')
        kernels: List[Dict[str, np.ndarray]] = []
        bns: List[Dict[str, np.ndarray]] = []
        for lname in layer_names:
            g = root[lname]
            wnames = _names(g.attrs['weight_names']) if 'weight_names' in g.attrs else []
            vals = {wn.split('/')[-1].split(':')[0]: np.asarray(g[wn].read() if hasattr(g[wn], 'read') else g[wn])
                    for wn in wnames}
            if 'kernel' in vals:
                kernels.append({'kernel': vals['kernel'], 'bias': vals.get('bias'), '_name': lname})
            elif 'gamma' in vals or 'moving_mean' in vals:
                bns.append({**{k: vals.get(k) for k in BN_KEYS}, '_name': lname})
        meta: Dict[str, Any] = {'path': path, 'layer_names': layer_names}
        top = f.attrs
        for key in ('keras_version', 'backend'):
            if key in top:
                meta[key] = _as_str(top[key])
        for key in ('model_config', 'training_config'):
            if key in top:
                try:
                    meta[key] = json.loads(_as_str(top[key]))
                except Exception:  # noqa: BLE001
                    meta[key] = _as_str(top[key])
        if 'optimizer_weights' in f:   # model.save(include_optimizer=True): the state ModelCheckpoint keeps (main.py:123-124)
            og = f['optimizer_weights']
            onames = _names(og.attrs['weight_names']) if 'weight_names' in og.attrs else []
            meta['optimizer_weight_names'] = onames
            meta['optimizer_weights'] = [np.asarray(og[n].read() if hasattr(og[n], 'read') else og[n]) for n in onames]
    finally:
        f.close()
    if len(kernels) != graph.NUM_LAYERS or len(bns) != graph.NUM_LAYERS - 1:
        raise ValueError(f'{path}: expected {graph.NUM_LAYERS} conv layers and {graph.NUM_LAYERS - 1} '
                         f'BatchNormalization layers, found {len(kernels)} and {len(bns)}')
    weights: Weights = []
    bn_iter = iter(bns)
    for l, k in zip(graph.LAYERS, kernels):
        want = graph.kernel_shape(l)
        if tuple(k['kernel'].shape) != want:
            raise ValueError(f"{path}: layer {k['_name']} kernel shape {tuple(k['kernel'].shape)} != {want} "
                             f'expected for {l.name}')
        if k['bias'] is None or tuple(k['bias'].shape) != (l.cout,):
            raise ValueError(f"{path}: layer {k['_name']} has no bias of shape ({l.cout},)")
        w = {'kernel': np.ascontiguousarray(k['kernel'], dtype=np.float32),
             'bias': np.ascontiguousarray(k['bias'], dtype=np.float32)}
        if l.bn:
            b = next(bn_iter)
            for key in BN_KEYS:
                if b[key] is None or tuple(b[key].shape) != (l.cout,):
                    raise ValueError(f"{path}: {b['_name']} {key} missing or not of shape ({l.cout},)")
                w[key] = np.ascontiguousarray(b[key], dtype=np.float32)
        weights.append(w)
    eps = _bn_epsilon_from_config(meta.get('model_config'))
    if eps is not None:
        meta['bn_epsilon'] = eps
    return weights, meta


def _bn_epsilon_from_config(cfg) -> Optional[float]:
    try:
        eps = {float(l['config']['epsilon']) for l in cfg['config']['layers'] if l['class_name'] == 'BatchNormalization'}
        return eps.pop() if len(eps) == 1 else None
    except Exception:  # noqa: BLE001
        return None


# ----------------------------------------------------------------------------------------------
# writing
# ----------------------------------------------------------------------------------------------
def model_config() -> Dict[str, Any]:
    """Functional-model config JSON the way Keras 2.2.4 serialises this graph (enough for load_model to
    rebuild it: class names, layer configs that differ from defaults, inbound nodes)."""
    layers = []
    for k in graph.keras_layers():
        cfg: Dict[str, Any] = {'name': k.name, 'trainable': True}
        if k.cls == 'InputLayer':
            cfg = {'name': k.name, 'batch_input_shape': [None, graph.IMG, graph.IMG, 3], 'dtype': 'float32',
                   'sparse': False}
        elif k.cls in ('Conv2D', 'Conv2DTranspose'):
            l = graph.LAYERS[k.weight_layer]
            ks = graph.kernel_shape(l)
            cfg.update(filters=l.cout, kernel_size=[ks[0], ks[1]],
                       strides=[2, 2] if k.cls == 'Conv2DTranspose' else [1, 1],
                       padding='valid' if l.kind != 'conv' else 'same', data_format='channels_last',
                       activation='sigmoid' if l.kind == 'head' else 'linear', use_bias=True,
                       kernel_initializer={'class_name': 'VarianceScaling',
                                           'config': ({'scale': 1.0, 'mode': 'fan_avg', 'distribution': 'uniform',
                                                       'seed': None} if l.kind == 'head' else
                                                      {'scale': 2.0, 'mode': 'fan_in', 'distribution': 'normal',
                                                       'seed': None})},
                       bias_initializer={'class_name': 'Zeros', 'config': {}})
            if k.cls == 'Conv2D':
                cfg['dilation_rate'] = [1, 1]
        elif k.cls == 'Activation':
            cfg['activation'] = 'relu'
        elif k.cls == 'BatchNormalization':
            cfg.update(axis=-1, momentum=0.99, epsilon=graph.BN_EPS, center=True, scale=True)
        elif k.cls == 'MaxPooling2D':
            cfg.update(pool_size=[2, 2], padding='valid', strides=[2, 2], data_format='channels_last')
        elif k.cls == 'Dropout':
            cfg.update(rate=0.5, noise_shape=None, seed=None)
        elif k.cls == 'Concatenate':
            cfg['axis'] = 3
        inbound = [[[s, 0, 0, {}] for s in k.inbound]] if k.inbound else []
        layers.append({'name': k.name, 'class_name': k.cls, 'config': cfg, 'inbound_nodes': inbound})
    return {'class_name': 'Model',
            'config': {'name': graph.MODEL_NAME, 'layers': layers, 'input_layers': [['input_1', 0, 0]],
                       'output_layers': [['conv2d_15', 0, 0]]}}


def training_config(optimizer: str = 'rmsprop') -> Dict[str, Any]:
    """model.compile(optimizer='rmsprop', loss='mse', metrics=['accuracy']) (model.py:281); 'adam' = the variant in
    the comments of model.py:280,282 (beta_1 = 0.5)."""
    opt = {'class_name': 'RMSprop', 'config': {'lr': 0.001, 'rho': 0.9, 'decay': 0.0, 'epsilon': 1e-07}}
    if optimizer == 'adam':
        opt = {'class_name': 'Adam', 'config': {'lr': 0.001, 'beta_1': 0.5, 'beta_2': 0.999, 'decay': 0.0,
                                                'epsilon': 1e-07, 'amsgrad': False}}
    return {'optimizer_config': opt,
            'loss': 'mse', 'metrics': ['accuracy'], 'sample_weight_mode': None, 'loss_weights': None}


def write_checkpoint(path: str, weights: Weights, weights_only: bool = False,
                     optimizer_weights: Optional[List[np.ndarray]] = None, optimizer: str = 'rmsprop'):
    """Write ``weights`` (execution order) in the Keras HDF5 layout. ``weights_only`` mimics
    ``model.save_weights`` (no model_config, groups at the root)."""
    if len(weights) != graph.NUM_LAYERS:
        raise ValueError(f'expected {graph.NUM_LAYERS} layers')
    w = h5min.Writer()
    prefix = '' if weights_only else 'model_weights/'
    klayers = graph.keras_layers()
    for k in klayers:
        g = w.group(prefix + k.name)
        names: List[str] = []
        if k.weight_layer is not None:
            lw = weights[k.weight_layer]
            keys = ('kernel', 'bias') if k.cls != 'BatchNormalization' else BN_KEYS
            for key in keys:
                wn = f'{k.name}/{key}:0'
                w.dataset(prefix + k.name + '/' + wn, np.asarray(lw[key], dtype=np.float32))
                names.append(wn)
        g.attrs['weight_names'] = (np.array([n.encode('utf-8') for n in names]) if names
                                   else np.zeros((0,), dtype=np.float64))
    top = w.group(prefix) if prefix else w.root
    top.attrs['layer_names'] = np.array([k.name.encode('utf-8') for k in klayers])
    top.attrs['backend'] = h5min.VlenStr(b'tensorflow')
    top.attrs['keras_version'] = h5min.VlenStr(b'2.2.4')
    if not weights_only:
        w.root.attrs['keras_version'] = h5min.VlenStr(b'2.2.4')
        w.root.attrs['backend'] = h5min.VlenStr(b'tensorflow')
        w.root.attrs['model_config'] = h5min.VlenStr(json.dumps(model_config()).encode('utf-8'))
        w.root.attrs['training_config'] = h5min.VlenStr(json.dumps(training_config(optimizer)).encode('utf-8'))
        if optimizer_weights is not None:
            og = w.group('optimizer_weights')
            names = []
            cls = 'Adam' if optimizer == 'adam' else 'RMSprop'
            for i, a in enumerate(optimizer_weights):
                # Keras 2.2.x names: K.zeros(...) variables of the optimizer's name scope, numbered in creation order
                n = f'training/{cls}/Variable{"_" + str(i) if i else ""}:0'
                a = np.asarray(a)
                w.dataset('optimizer_weights/' + n, a if a.dtype == np.int64 else a.astype(np.float32))
                names.append(n.encode('utf-8'))
            og.attrs['weight_names'] = np.array(names) if names else np.zeros((0,), dtype=np.float64)
    w.save(path)


# ----------------------------------------------------------------------------------------------
# locating the checkpoint predict.py:24 names
# ----------------------------------------------------------------------------------------------
REFERENCE_CHECKPOINT = 'saved7-model-218-0.73.hdf5'


def find_reference_checkpoint() -> Optional[str]:
    """$ITMO_CKPT, ./saved7-model-218-0.73.hdf5, repo root, baseline/_ref/ — in that order."""
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cands = [os.environ.get('ITMO_CKPT'), REFERENCE_CHECKPOINT, os.path.join(here, REFERENCE_CHECKPOINT),
             os.path.join(here, 'baseline', '_ref', REFERENCE_CHECKPOINT)]
    for c in cands:
        if c and os.path.isfile(c):
            return c
    return None
