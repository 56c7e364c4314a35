"""Host-side logic that needs no GPU: graph table, checkpoint layout, Model API error behaviour, and that the
C-ABI library loads and exports every symbol include/itmo_b200.h declares (no compute calls)."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

import cnn_itmo_b200 as pkg
from cnn_itmo_b200 import _lib, checkpoint, graph, model as model_mod
from oracle import unet_ref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_graph_table_matches_oracle_and_layers_txt(golden_dir):
    assert graph.summary_text() == R.keras_summary_text()
    assert graph.count_params() == R.param_count()
    tab = json.load(open(os.path.join(golden_dir, 'layers_table.json')))
    assert [k.name for k in graph.keras_layers()] == [r['name'] for r in tab['rows']]
    for a, b in zip(graph.LAYERS, R.LAYERS):
        assert (a.kind, a.name, a.bn, tuple(a.cin), a.cout, a.hin, a.pool, a.dropout, tuple(a.src), a.src_pooled) == \
               (b.kind, b.keras_name, b.bn_name, tuple(b.cin), b.cout, b.hin, b.pool, b.dropout, tuple(b.src),
                b.src_pooled)
    assert abs(graph.forward_flops_per_image() / 1e9 - 165.373) < 1e-3


def test_u_net_prints_summary_and_ignores_pretrained_weights(capsys):
    m = pkg.U_net(pretrained_weights='does-not-exist.hdf5')                  # ignored, as at model.py:285-286
    out = capsys.readouterr().out
    assert 'conv2d_transpose_4 (Conv2DTrans' in out and 'Total params: 11,166,819' in out
    assert m.name == 'ReinhardtPrediction' and m.loss == 'mse' and m.optimizer == 'rmsprop'
    assert m.metrics == ['accuracy'] and m.count_params() == 11166819
    with pytest.raises(ValueError):
        pkg.Model(input_size=(256, 256, 3))


def test_checkpoint_roundtrip_keras_layout(tmp_path):
    ws = R.synthetic_weights(5)
    p = str(tmp_path / 'saved7-model-01-0.50.hdf5')
    checkpoint.write_checkpoint(p, ws)
    got, meta = checkpoint.read_checkpoint(p)
    for a, b in zip(ws, got):
        assert set(a) == set(b)
        for k in a:
            np.testing.assert_array_equal(a[k], b[k])
    assert meta['keras_version'] == '2.2.4' and meta['backend'] == 'tensorflow'
    assert meta['bn_epsilon'] == pytest.approx(1e-3)
    assert meta['model_config']['config']['name'] == 'ReinhardtPrediction'
    assert meta['training_config']['optimizer_config']['class_name'] == 'RMSprop'
    assert len(meta['layer_names']) == 66
    # the on-disk paths are exactly Keras': /model_weights/<layer>/<layer>/kernel:0
    from cnn_itmo_b200 import h5min
    f = h5min.File(p)
    assert f['model_weights/conv2d_transpose_2/conv2d_transpose_2/kernel:0'].shape == (2, 2, 256, 512)
    assert f['model_weights/batch_normalization_18/batch_normalization_18/moving_variance:0'].shape == (64,)
    assert f['model_weights/max_pooling2d_1'].attrs['weight_names'].shape == (0,)
    # weights-only flavour (model.save_weights)
    p2 = str(tmp_path / 'w.h5')
    checkpoint.write_checkpoint(p2, ws, weights_only=True)
    got2, _ = checkpoint.read_checkpoint(p2)
    np.testing.assert_array_equal(got2[11]['kernel'], ws[11]['kernel'])


def _fake_layout():
    """A flat parameter layout with the library's field order (kernel, bias, gamma, beta per layer) built from the graph
    table — enough to exercise the optimizer-state ordering without the CUDA library."""
    layout, off = [], 0
    for l in graph.LAYERS:
        ks = graph.kernel_shape(l)
        d = {}
        for name, cnt in (('kernel', int(np.prod(ks))), ('bias', l.cout)) + ((('gamma', l.cout), ('beta', l.cout)) if l.bn else ()):
            d[name] = (off, cnt)
            off += cnt
        layout.append(d)
    return layout, off


@pytest.mark.parametrize('opt', ['rmsprop', 'adam'])
def test_optimizer_state_roundtrip_in_keras_order(tmp_path, opt):
    """model.save writes the optimizer state under /optimizer_weights in the order of Keras'
    trainable_weights (main.py:123-124 ModelCheckpoint); reading it back restores the flat state exactly."""
    from cnn_itmo_b200 import train, h5min
    layout, n = _fake_layout()
    assert n == graph.count_params()['trainable'] == 11159011
    order = train.keras_trainable_order()
    assert len(order) == 2 * graph.NUM_LAYERS + 2 * (graph.NUM_LAYERS - 1)
    assert order[:4] == [(0, 'kernel'), (0, 'bias'), (0, 'gamma'), (0, 'beta')]
    assert order[-2:] == [(18, 'kernel'), (18, 'bias')]
    rng = np.random.default_rng(3)
    state = rng.random(2 * n, dtype=np.float32)
    if opt == 'rmsprop':
        state[n:] = 0
    ows = train.optimizer_weights_from_flat(state, layout, opt, iterations=41)
    assert len(ows) == (len(order) if opt == 'rmsprop' else 1 + 2 * len(order))
    first = ows[0] if opt == 'rmsprop' else ows[1]
    assert first.shape == (3, 3, 3, 32)                       # accumulator of conv2d_1/kernel, Keras HWIO
    p = str(tmp_path / f'ck-{opt}.hdf5')
    ws = R.synthetic_weights(5)
    checkpoint.write_checkpoint(p, ws, optimizer_weights=ows, optimizer=opt)
    got, meta = checkpoint.read_checkpoint(p)
    np.testing.assert_array_equal(got[3]['kernel'], ws[3]['kernel'])
    cls = 'RMSprop' if opt == 'rmsprop' else 'Adam'
    assert meta['training_config']['optimizer_config']['class_name'] == cls
    assert meta['optimizer_weight_names'][0] == f'training/{cls}/Variable:0'
    assert meta['optimizer_weight_names'][5] == f'training/{cls}/Variable_5:0'
    back, it = train.optimizer_weights_to_flat(meta['optimizer_weights'], layout, opt, n)
    np.testing.assert_array_equal(back, state)
    assert it == (0 if opt == 'rmsprop' else 41)
    assert h5min.File(p)[f'optimizer_weights/training/{cls}/Variable_1:0'].shape == ((32,) if opt == 'rmsprop' else (3, 3, 3, 32))
    with pytest.raises(ValueError):
        train.optimizer_weights_to_flat(meta['optimizer_weights'][:-1], layout, opt, n)


def test_checkpoint_matching_is_by_type_order_not_suffix(tmp_path):
    """A model built second in a Keras session has conv2d_16..; matching must not depend on the numbers."""
    from cnn_itmo_b200 import h5min
    ws = R.synthetic_weights(6)
    w = h5min.Writer()
    names = []
    for k in graph.keras_layers():
        new = re.sub(r'(\d+)$', lambda m: str(int(m.group(1)) + 100), k.name)
        names.append(new)
        g = w.group('model_weights/' + new)
        wn = []
        if k.weight_layer is not None:
            keys = ('kernel', 'bias') if k.cls != 'BatchNormalization' else checkpoint.BN_KEYS
            for key in keys:
                w.dataset(f'model_weights/{new}/{new}/{key}:0', ws[k.weight_layer][key])
                wn.append(f'{new}/{key}:0'.encode())
        g.attrs['weight_names'] = np.array(wn) if wn else np.zeros((0,))
    w.group('model_weights').attrs['layer_names'] = np.array([n.encode() for n in names])
    p = str(tmp_path / 'renamed.h5')
    w.save(p)
    got, _ = checkpoint.read_checkpoint(p)
    np.testing.assert_array_equal(got[13]['kernel'], ws[13]['kernel'])
    np.testing.assert_array_equal(got[17]['moving_mean'], ws[17]['moving_mean'])


def test_load_model_errors(tmp_path):
    with pytest.raises(OSError):
        pkg.load_model(str(tmp_path / 'saved7-model-218-0.73.hdf5'))
    bad = tmp_path / 'bad.hdf5'
    bad.write_bytes(b'\x00' * 4096)
    with pytest.raises(OSError):
        pkg.load_model(str(bad))
    # a readable HDF5 file with the wrong architecture -> ValueError
    from cnn_itmo_b200 import h5min
    w = h5min.Writer()
    w.group('model_weights').attrs['layer_names'] = np.array([b'dense_1'])
    w.group('model_weights/dense_1').attrs['weight_names'] = np.array([b'dense_1/kernel:0'])
    w.dataset('model_weights/dense_1/dense_1/kernel:0', np.zeros((4, 4), np.float32))
    p = str(tmp_path / 'other.h5')
    w.save(p)
    with pytest.raises(ValueError):
        pkg.load_model(p)


def test_model_input_validation_and_weight_lists():
    m = pkg.Model(R.synthetic_weights(1))
    for shape in [(1, 256, 256, 3), (512, 512, 3), (1, 512, 512, 1)]:
        with pytest.raises(ValueError):
            m.predict(np.zeros(shape, np.float32))
    with pytest.raises(ValueError):
        m.predict(np.zeros((1, 512, 512, 3), np.uint8))
    flat = m.get_weights()
    assert len(flat) == 19 * 2 + 18 * 4 and sum(a.size for a in flat) == 11166819
    assert flat[0].shape == (3, 3, 3, 32) and flat[2].shape == (32,)          # conv2d_1 kernel, then BN_1 gamma
    flat2 = [a + 1 for a in flat]
    m.set_weights(flat2)
    np.testing.assert_array_equal(m.get_weights()[5], flat[5] + 1)
    with pytest.raises(ValueError):
        m.set_weights(flat[:-1])
    with pytest.raises(ValueError):
        pkg.Model(precision='fp8')


def test_library_loads_and_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, 'include', 'itmo_b200.h')).read()
    hdr = re.sub(r'/\*.*?\*/', '', hdr, flags=re.S)
    declared = set(re.findall(r'\b(itmo_[A-Za-z0-9_]+)\s*\(', hdr))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    lib = _lib.load()                                                         # raises if the .so is not built
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.itmo_abi_version() >= 2
    assert lib.itmo_last_error() is not None


def test_product_code_never_imports_the_oracle():
    pkg_dir = os.path.join(ROOT, 'cnn-itmo_b200')
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dirpath, f)).read()
                assert 'unet_ref' not in src and not re.search(r'^\s*(from|import)\s+oracle', src, flags=re.M), f


def test_bench_layer_table_with_and_without_the_fused_stem():
    """bench.py's per-layer roofline table: with conv2d_1 + conv2d_2 fused (no launch booked under conv2d_1) the first
    row says so and conv2d_2's row carries both layers' FLOPs; the JSON `layers` list must build in both cases."""
    import bench
    from cnn_itmo_b200 import graph
    peaks = {'tflops_sustained': 1381.5, 'hbm_gbs': 6554.0}
    keys = ('layer', 'ms_per_step', 'tflops', 'gbs', 'bound', 'frac_of_bound', 'fused_into')
    flops = {}
    for fused in (True, False):
        lt = {l.index: (5.0, 5) for l in graph.LAYERS}
        if fused:
            lt[0] = (0.0, 0)
        per, fl, ms = bench.layer_table(graph, lt, 5, 256, 'fp16', peaks)
        layers = [{k: p[k] for k in keys if k in p} for p in per]
        assert len(layers) == 18 and layers[0]['layer'] == 'conv2d_1'
        assert ('fused_into' in layers[0]) == fused
        assert sum(int(p.get('launches_per_step', 0)) for p in per[1:]) == 17
        flops[fused] = fl
    assert flops[True] > flops[False]          # the first layer's FLOPs are counted with the fused kernel
