"""The oracle (oracle/unet_ref.py) against everything the reference offers to pin it, and against itself.

The reference has no tests and no numeric golden vectors for the path: what it does
ship is layers.txt, the verbatim model.summary() of the graph. The oracle must regenerate that table
row by row (names, classes, shapes, parameter counts, connectivity, totals)."""
import json
import os

import numpy as np
import pytest

from oracle import unet_ref as R


def test_summary_matches_reference_layers_txt(golden_dir):
    """Known-answer test: oracle graph -> Keras summary == the reference's layers.txt (parsed fixture)."""
    tab = json.load(open(os.path.join(golden_dir, 'layers_table.json')))
    rows = R.keras_layer_rows()
    assert len(rows) == len(tab['rows']) == 66
    for (name, cls, shape, params, inbound), ref in zip(rows, tab['rows']):
        assert name == ref['name']
        assert cls.startswith(ref['cls_prefix'])          # layers.txt truncates '(BatchNormalization)' at col 32
        # the summary column truncates '(None, 128, 128, 128)' to 20 characters
        assert str(shape)[:20].strip() == ref['shape'], (name, shape, ref['shape'])
        assert params == ref['params'], name
        assert [f'{i}[0][0]' for i in inbound] == ref['inbound'], name
    assert R.param_count() == {'total': tab['totals']['total'], 'trainable': tab['totals']['trainable'],
                               'non_trainable': tab['totals']['non_trainable']}
    assert R.param_count() == {'total': 11166819, 'trainable': 11159011, 'non_trainable': 7808}  # layers.txt:140-142


@pytest.mark.skipif(not os.path.exists('/root/reference/layers.txt'), reason='reference tree not mounted')
def test_summary_text_is_byte_identical_to_layers_txt():
    ref = open('/root/reference/layers.txt').read().splitlines()
    mine = R.keras_summary_text().splitlines()
    assert [l.rstrip() for l in mine[:len(ref)]] == [l.rstrip() for l in ref]


def test_oracle_golden_small(golden_dir):
    """The oracle against its own committed outputs (tests/golden/make_golden.py): guards against drift."""
    g = np.load(os.path.join(golden_dir, 'oracle_small.npz'))
    ws = R.synthetic_weights(218)
    with np.errstate(over='ignore'):
        outs, pooled = R.forward_np(ws, g['x'], return_all=True)
    for key, got in (('y', outs[18]), ('l1', outs[1]), ('p1', pooled[1]), ('l10', outs[10]), ('l17', outs[17])):
        np.testing.assert_allclose(got, g[key], rtol=1e-9, atol=1e-9, err_msg=key)


def test_numpy_and_torch_restatements_agree_fp64():
    """(A) numpy float64 loops vs (B) torch CPU float64, every intermediate tensor."""
    import torch
    ws = R.synthetic_weights(3)
    x = np.random.default_rng(1).random((2, 32, 48, 3))
    with np.errstate(over='ignore'):
        a_outs, a_pool = R.forward_np(ws, x, return_all=True)
    b_outs, b_pool = R.forward_torch(R.TorchWeights(ws, dtype=torch.float64), x, return_all=True)
    for i in a_outs:
        scale = max(1.0, float(np.abs(a_outs[i]).max()))
        assert np.abs(a_outs[i] - b_outs[i]).max() <= 1e-10 * scale, i
    for i in a_pool:
        assert np.abs(a_pool[i] - b_pool[i]).max() <= 1e-10 * max(1.0, float(np.abs(a_pool[i]).max())), i


def test_conv_transpose_is_four_gemms_plus_pixel_shuffle():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((1, 3, 5, 4))
    k = rng.standard_normal((2, 2, 6, 4))
    b = rng.standard_normal(6)
    y = R.conv2d_transpose_k2s2_np(x, k, b)
    for i in range(3):
        for j in range(5):
            for a in range(2):
                for c in range(2):
                    np.testing.assert_allclose(y[0, 2 * i + a, 2 * j + c], k[a, c] @ x[0, i, j] + b, atol=1e-12)


def test_conv_same_zero_padding_and_cross_correlation():
    x = np.zeros((1, 4, 4, 1))
    x[0, 0, 0, 0] = 1.0
    k = np.arange(9, dtype=np.float64).reshape(3, 3, 1, 1)
    y = R.conv2d_same_np(x, k, np.zeros(1))
    # out[y,x] = sum_ij in[y+i-1,x+j-1] k[i,j]: the impulse at (0,0) shows k[1,1] at (0,0), k[0,0] at (1,1)
    assert y[0, 0, 0, 0] == 4 and y[0, 1, 1, 0] == 0 and y[0, 0, 1, 0] == 3 and y[0, 1, 0, 0] == 1


def test_bn_sits_after_relu_and_negative_gamma_matters():
    """model.py:195-196: BN(ReLU(conv)). With gamma < 0 pooling before the affine would pick the wrong element."""
    s = R.LAYERS[1]
    w = {k: v.copy() for k, v in R.synthetic_weights(0)[1].items()}
    w['gamma'][:] = -1.0
    x = np.random.default_rng(2).standard_normal((1, 8, 8, 32))
    y = R.block_np(s, w, x)
    z = np.maximum(R.conv2d_same_np(x, w['kernel'], w['bias']), 0)
    wrong = R.bn_infer_np(R.maxpool2_np(z), w['gamma'], w['beta'], w['moving_mean'], w['moving_variance'])
    assert np.abs(R.maxpool2_np(y) - wrong).max() > 1e-3


def test_pre_and_post_processing_follow_predict_py():
    u8 = np.array([[[0, 127, 255]]], dtype=np.uint8)
    f = R.preprocess_u8(u8)
    assert f.dtype == np.float64 and f[0, 0, 2] == 1.0 and f[0, 0, 1] == 127 / 255
    p = np.array([0.0, 0.999, 0.5, 1.0], dtype=np.float32)
    np.testing.assert_array_equal(R.postprocess_u8(p), np.array([0, 254, 127, 255], dtype=np.uint8))  # truncation


def test_calibrated_weights_keep_activations_order_one():
    import torch
    ws = R.calibrated_weights(218, calib_size=64, calib_n=2)
    x = R.preprocess_u8(R.synthetic_images_u8(1, seed=9, size=64))
    outs, _ = R.forward_torch(R.TorchWeights(ws), x, return_all=True)
    for i in range(18):
        assert 0.3 < outs[i].std() < 3.0, (i, outs[i].std())
    assert 0.05 < outs[18].std() < 0.5 and 0.0 < outs[18].min() and outs[18].max() < 1.0
