"""CPU tests of the training-step oracle (oracle/unet_train_ref.py) and the host-side training surface.

The reference has no tests for training and its arithmetic lives in un-vendored Keras/TF1, so the oracle
is gated the only way available: autograd gradients against central finite differences in float64, the optimizer
restatements against hand-computed Keras formulas, and the dropout hash against a scalar re-implementation.
"""
import ctypes

import numpy as np
import pytest

from oracle import unet_ref as R
from oracle import unet_train_ref as T

torch = pytest.importorskip('torch')


def _small_weights(seed=2):
    ws = R.synthetic_weights(seed)
    return [{k: np.asarray(v, np.float64) for k, v in w.items()} for w in ws]


def test_autograd_matches_finite_differences():
    ws = _small_weights()
    rng = np.random.default_rng(0)
    x = rng.uniform(0, 1, (2, 32, 32, 3))
    y = rng.uniform(0, 1, (2, 32, 32, 3))
    res = T.train_forward_backward(ws, x, y, None, dtype=torch.float64)
    eps = 1e-7   # small step: the graph is piecewise smooth (ReLU, max-pool), larger steps cross kinks
    checks = [(0, 'kernel', (1, 1, 2, 5)), (3, 'bias', (7,)), (5, 'gamma', (3,)), (9, 'beta', (100,)),
              (10, 'kernel', (1, 0, 17, 33)), (11, 'kernel', (0, 2, 300, 40)), (17, 'kernel', (2, 2, 90, 10)),
              (18, 'kernel', (0, 0, 13, 2)), (18, 'bias', (1,))]
    for li, name, idx in checks:
        orig = ws[li][name][idx]
        ws[li][name][idx] = orig + eps
        lp = T.loss_only(ws, x, y, torch.float64)
        ws[li][name][idx] = orig - eps
        lm = T.loss_only(ws, x, y, torch.float64)
        ws[li][name][idx] = orig
        fd = (lp - lm) / (2 * eps)
        g = res['grads'][li][name][idx]
        assert abs(fd - g) <= 1e-7 + 2e-3 * abs(g), (li, name, idx, fd, g)


def test_loss_and_accuracy_definitions():
    ws = _small_weights(4)
    rng = np.random.default_rng(1)
    x = rng.uniform(0, 1, (1, 16, 16, 3))
    y = rng.uniform(0, 1, (1, 16, 16, 3))
    res = T.train_forward_backward(ws, x, y, None, dtype=torch.float64)
    assert np.isclose(res['loss'], np.mean((res['pred'] - y) ** 2))        # Keras 'mse'
    assert np.isclose(res['acc'], np.mean(res['pred'].argmax(-1) == y.argmax(-1)))  # categorical_accuracy
    # loss_count rescales the loss (global-batch mean when a rank only holds a share)
    res2 = T.train_forward_backward(ws, x, y, None, dtype=torch.float64, loss_count=2 * y.size)
    assert np.isclose(res2['loss'], res['loss'] / 2)
    assert np.allclose(res2['grads'][5]['kernel'], res['grads'][5]['kernel'] / 2)


def _keep_scalar(e, seed):
    h = ((e * 0x9E3779B1) & 0xFFFFFFFF) ^ seed
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & 0xFFFFFFFF
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & 0xFFFFFFFF
    h ^= h >> 16
    return (h >> 15) & 1


def test_dropout_mask_hash():
    seed = T.layer_dropout_seed(1234, 7)
    m = T.dropout_keep_mask((2, 4, 4, 8), seed)
    flat = m.reshape(-1)
    for e in (0, 1, 17, 255):
        assert bool(flat[e]) == bool(_keep_scalar(e, seed))
    big = T.dropout_keep_mask((1, 64, 64, 256), seed)
    assert abs(big.mean() - 0.5) < 0.01     # rate 0.5 (model.py:226,239)
    assert (T.dropout_keep_mask((1, 64, 64, 256), T.layer_dropout_seed(1234, 9)) != big).mean() > 0.4


def test_dropout_changes_forward_and_scales_by_two():
    ws = _small_weights(5)
    rng = np.random.default_rng(2)
    x = rng.uniform(0, 1, (1, 16, 16, 3))
    y = rng.uniform(0, 1, (1, 16, 16, 3))
    a = T.train_forward_backward(ws, x, y, None, dtype=torch.float64, return_acts=True)
    b = T.train_forward_backward(ws, x, y, 77, dtype=torch.float64, return_acts=True)
    keep = T.dropout_keep_mask(a['a'][7].shape, T.layer_dropout_seed(77, 7))
    assert np.allclose(b['a'][7], a['a'][7] * keep * 2.0)
    assert not np.isclose(a['loss'], b['loss'])


def test_optimizer_restatements():
    p, g, a = np.array([1.0, -2.0]), np.array([0.5, -0.25]), np.array([0.0, 0.01])
    p1, a1 = T.rmsprop_step(p, g, a)
    assert np.allclose(a1, 0.9 * a + 0.1 * g * g)
    assert np.allclose(p1, p - 1e-3 * g / (np.sqrt(a1) + 1e-7))
    p2, m2, v2 = T.adam_step(p, g, np.zeros(2), np.zeros(2), 1)
    # first Adam step moves by ~lr * sign(g)
    assert np.allclose(p2, p - 1e-3 * np.sign(g), atol=1e-6)
    mm, mv = T.moving_update(np.zeros(1), np.ones(1), np.array([2.0]), np.array([4.0]), 1000.0)
    assert np.allclose(mm, 0.02) and np.allclose(mv, 0.99 + 0.01 * 4.0 * 1000.0 / (1000.0 - 1.001))


def test_trainer_abi_layout_without_gpu():
    """The flat parameter layout is queryable without a GPU and tiles the trainable parameters exactly."""
    from cnn_itmo_b200 import _lib, train
    lib = _lib.load()
    lay = train.param_layout()
    spans = []
    for li, d in enumerate(lay):
        for name in ('kernel', 'bias', 'gamma', 'beta'):
            if name in d:
                spans.append(d[name])
        ks = R.kernel_shape(R.LAYERS[li])
        assert d['kernel'][1] == int(np.prod(ks))
    spans.sort()
    pos = 0
    for off, cnt in spans:
        assert off == pos
        pos += cnt
    assert pos == lib.itmo_train_param_count() == R.param_count()['trainable']
    assert sum(d['bn'][1] for d in lay if 'bn' in d) == lib.itmo_train_bn_count() == 3904
    off, cnt = ctypes.c_longlong(), ctypes.c_longlong()
    assert lib.itmo_train_layout(18, 2, ctypes.byref(off), ctypes.byref(cnt)) != 0   # the head has no BN
    assert lib.itmo_train_layout(99, 0, ctypes.byref(off), ctypes.byref(cnt)) != 0


def test_model_training_surface_errors_without_gpu():
    import cnn_itmo_b200 as M
    m = M.Model(M.initial_weights(0))
    with pytest.raises(ValueError):
        m.train_on_batch(np.zeros((1, 64, 64, 3), np.float32), np.zeros((1, 64, 64, 3), np.float32))
    with pytest.raises(ValueError):
        m.train_on_batch(np.zeros((1, 512, 512, 3), np.float32), np.zeros((1, 512, 512, 1), np.float32))
    with pytest.raises(ValueError):
        m.fit_generator(iter([]))   # steps_per_epoch required
    if not torch.cuda.is_available():
        m.compile()
        with pytest.raises(Exception):   # no GPU -> loud failure, never a CPU fallback
            m.train_on_batch(np.zeros((1, 512, 512, 3), np.float32), np.zeros((1, 512, 512, 3), np.float32))


# ---- data-parallel host logic over two gloo ranks (the N > 1 path without GPUs) ---------------------------------------
def _dp_worker(rank, world, port, q):
    import os
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from cnn_itmo_b200 import train
        ws = _small_weights(6)
        rng = np.random.default_rng(5)
        x = rng.uniform(0, 1, (4, 16, 16, 3))
        y = rng.uniform(0, 1, (4, 16, 16, 3))
        b, e = train.shard_range(4, world, rank)
        count = train.global_loss_count(e - b, 16, 16, world)
        res = T.train_forward_backward(ws, x[b:e], y[b:e], None, dtype=torch.float64, loss_count=count)
        flat = torch.from_numpy(train.flatten_layers(res['grads']).astype(np.float64))
        train.allreduce_flat(flat)
        sq = torch.tensor([res['sq_err_sum']], dtype=torch.float64)
        train.allreduce_flat(sq)
        q.put((rank, flat.numpy(), float(sq[0]) / count, count))
    finally:
        dist.destroy_process_group()


def test_data_parallel_exchange_over_two_gloo_ranks():
    """Each rank runs its shard with the GLOBAL loss_count; the all-reduced flat gradient equals the sum of the
    per-shard gradients (BN statistics are per replica) and the all-reduced loss the global mse."""
    import multiprocessing as mp
    import socket
    from cnn_itmo_b200 import train
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = sorted([q.get(timeout=300) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(outs[0][1], outs[1][1])            # every rank ends with the same gradient
    assert outs[0][3] == 4 * 16 * 16 * 3
    ws = _small_weights(6)
    rng = np.random.default_rng(5)
    x = rng.uniform(0, 1, (4, 16, 16, 3))
    y = rng.uniform(0, 1, (4, 16, 16, 3))
    want = np.zeros_like(outs[0][1])
    loss = 0.0
    for r in range(2):
        b, e = train.shard_range(4, 2, r)
        res = T.train_forward_backward(ws, x[b:e], y[b:e], None, dtype=torch.float64, loss_count=float(y.size))
        want += train.flatten_layers(res['grads']).astype(np.float64)
        loss += res['sq_err_sum'] / y.size
    assert np.allclose(outs[0][1], want, rtol=1e-6, atol=1e-12)
    assert np.isclose(outs[0][2], loss)
    with pytest.raises(ValueError):
        train.shard_range(5, 2, 0)
    assert train.shard_range(512, 8, 3) == (192, 256)


def test_fused_exchange_shards_partition_the_flat_buffer():
    """train.shard_bounds (what itmo_exchange_step computes on the device): the ranks' shards tile [0, n) exactly, start
    on 4-float boundaries (128-bit peer loads) and differ by at most one vector."""
    from cnn_itmo_b200 import train
    n = 11159011
    for world in range(1, 9):
        cuts = [train.shard_bounds(n, world, r) for r in range(world)]
        assert cuts[0][0] == 0 and cuts[-1][1] == n
        for (b0, e0), (b1, e1) in zip(cuts, cuts[1:]):
            assert e0 == b1 and b1 % 4 == 0
        sizes = [e - b for b, e in cuts]
        assert max(sizes) - min(sizes) <= 4 * world


def test_prefetch_identity_of_host_batches():
    """Trainer.train_on_batch(prefetch=...) recognises the announced batch by the identity of the arrays AS PASSED (address,
    shape, dtype), before any dtype conversion copies them; the conversion itself keeps uint8 inputs and makes everything
    else float32 (main.py:71's rescaled floats)."""
    from cnn_itmo_b200 import train
    key, host = train.Trainer._batch_key, train.Trainer._host_batch
    x8 = np.zeros((2, 16, 16, 3), np.uint8)
    y32 = np.zeros((2, 16, 16, 3), np.float32)
    x64 = np.zeros((2, 16, 16, 3), np.float64)
    assert key(x8, y32) == key(x8, y32)
    assert key(x8, y32) != key(x8.copy(), y32) and key(x8, y32) != key(x8, y32.copy())
    assert key(x8, y32) != key(x8[:1], y32[:1])                 # same address, other shape
    assert key(x64, y32) == key(x64, y32)                       # float64 input: converted later, recognised by the original
    hx, hy = host(x8, y32)
    assert hx.dtype == np.uint8 and hy.dtype == np.float32 and hx.ctypes.data == x8.ctypes.data and hy.ctypes.data == y32.ctypes.data
    hx, hy = host(x64, x64)
    assert hx.dtype == np.float32 and hy.dtype == np.float32
    hx, _ = host(x8[:, :, ::2], y32)                            # non-contiguous view -> contiguous copy
    assert hx.flags['C_CONTIGUOUS']
