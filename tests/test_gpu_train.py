"""GPU parity tests of the training step (run with ``-m gpu`` on a B200): the CUDA path through the C ABI
(itmo_train_* / itmo_wgrad3x3 / itmo_dgrad3x3 of include/itmo_b200.h) against the CPU oracle
(oracle/unet_train_ref.py: torch-CPU autograd over the graph of model.py:204-278 in training mode, main.py:126-132).

Tolerances. The device keeps activations and activation gradients in bf16 (8 mantissa bits) with fp32 accumulation, so:
  * stateless wgrad / dgrad on bf16-exact inputs: fp32-accumulation error only, relative L2 <= 2e-3 (dgrad output is
    stored in bf16: <= 6e-3);
  * whole step against the fp32 oracle: loss within 2 %, prediction PSNR >= 25 dB, activations within 15 % relative L2
    at the deepest point. Gradients: the 18-layer ReLU->BN stack amplifies rounding noise on this small problem — the
    oracle run twice on the CPU, once in fp32 and once with bf16 storage emulated (emulate_bf16=True), already differs
    by 1 % (conv2d_14), 10-14 % (conv2d_13 / transpose_4) and 40-70 % (everything earlier) in relative L2, the same
    figures the device shows. So the whole-network test holds the last two layers to 8 %, every tensor to cosine >= 0.3
    and a norm ratio in [0.5, 2] (sign / scale / layout errors), and tests/test_gpu_train_local.py checks every kernel of
    every layer in isolation at 1e-2;
  * optimizer update against the restated Keras RMSprop / Adam on the device's own gradients: 2e-5 relative.
Nothing here reads /root/reference.
"""
import ctypes

import numpy as np
import pytest

from oracle import unet_ref as R
from oracle import unet_train_ref as T

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')


def _lib():
    from cnn_itmo_b200 import _lib as L
    return L, L.load()


@pytest.fixture(scope='module')
def dev():
    assert torch.cuda.is_available(), 'the gpu tests need a CUDA device'
    cap = torch.cuda.get_device_capability(0)
    assert cap[0] == 10, f'libitmo_b200 only contains sm_100a code; device is sm_{cap[0]}{cap[1]}'
    return torch.device('cuda:0')


def _rel(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def _cos(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(a @ b / max(np.linalg.norm(a) * np.linalg.norm(b), 1e-30))


# ---- stateless gradient operators ---------------------------------------------------------------------------------
WG_SHAPES = [  # (cin, cout, n, h, w)
    (32, 32, 2, 32, 16), (32, 64, 1, 16, 24), (64, 64, 2, 16, 16), (64, 128, 1, 32, 8), (128, 128, 3, 16, 8),
    (128, 256, 1, 16, 16), (256, 256, 2, 16, 8), (256, 512, 1, 16, 8), (512, 512, 1, 16, 16), (512, 32, 1, 16, 8),
]


@pytest.mark.parametrize('cin,cout,n,h,w', WG_SHAPES)
def test_wgrad3x3_matches_autograd(dev, cin, cout, n, h, w):
    L, lib = _lib()
    g = torch.Generator().manual_seed(cin * 1000 + cout)
    x = torch.randn(n, h, w, cin, generator=g).to(torch.bfloat16)
    gy = torch.randn(n, h, w, cout, generator=g).to(torch.bfloat16)
    xd, gd = x.to(dev), gy.to(dev)
    dw = torch.zeros(3, 3, cin, cout, dtype=torch.float32, device=dev)
    for _ in range(2):   # the operator ACCUMULATES: two launches = 2x
        L.check(lib.itmo_wgrad3x3(xd.data_ptr(), cin, gd.data_ptr(), cout, n, h, w, dw.data_ptr(), None))
    torch.cuda.synchronize()
    k = torch.zeros(cout, cin, 3, 3, dtype=torch.float64, requires_grad=True)
    y = torch.nn.functional.conv2d(x.double().permute(0, 3, 1, 2), k, padding=1)
    y.backward(gy.double().permute(0, 3, 1, 2))
    ref = 2.0 * k.grad.permute(2, 3, 1, 0).numpy()   # OIHW -> HWIO
    got = dw.cpu().numpy()
    assert _rel(got, ref) <= 2e-3, f'wgrad rel err {_rel(got, ref):.3e}'


@pytest.mark.parametrize('cin,cout,n,h,w', WG_SHAPES[:8])
def test_dgrad3x3_matches_autograd(dev, cin, cout, n, h, w):
    L, lib = _lib()
    g = torch.Generator().manual_seed(cin * 77 + cout)
    gy = torch.randn(n, h, w, cout, generator=g).to(torch.bfloat16)
    k = (torch.randn(3, 3, cin, cout, generator=g) / np.sqrt(9 * cout)).to(torch.bfloat16).float()
    gd, kd = gy.to(dev), k.to(dev)
    dx = torch.zeros(n, h, w, cin, dtype=torch.bfloat16, device=dev)
    L.check(lib.itmo_dgrad3x3(gd.data_ptr(), cout, n, h, w, kd.data_ptr(), cin, dx.data_ptr(), None))
    torch.cuda.synchronize()
    xin = torch.zeros(n, cin, h, w, dtype=torch.float64, requires_grad=True)
    y = torch.nn.functional.conv2d(xin, k.double().permute(3, 2, 0, 1), padding=1)
    y.backward(gy.double().permute(0, 3, 1, 2))
    ref = xin.grad.permute(0, 2, 3, 1).numpy()
    got = dx.float().cpu().numpy()
    assert _rel(got, ref) <= 6e-3, f'dgrad rel err {_rel(got, ref):.3e}'


# ---- whole training step ------------------------------------------------------------------------------------------
H, W, NB = 256, 128, 2   # smallest image the tile shapes allow (deepest level 16x8)


def _weights(seed=3):
    """he_normal-like weights with non-trivial BN affine parameters (so dgamma / dbeta paths are exercised)."""
    ws = R.synthetic_weights(seed)
    rng = np.random.default_rng(seed)
    for w in ws:
        if 'gamma' in w:
            w['gamma'] = (1.0 + 0.1 * rng.standard_normal(w['gamma'].shape)).astype(np.float32)
            w['beta'] = (0.2 + 0.1 * rng.standard_normal(w['beta'].shape)).astype(np.float32)
    return ws


@pytest.fixture(scope='module')
def step_case(dev):
    from cnn_itmo_b200 import train
    ws = _weights()
    x, y = T.synthetic_pair(NB, H, W, seed=11)
    tr = train.Trainer(ws, micro_batch=NB, image_hw=(H, W), device=0)
    return ws, x, y, tr


def _run_backward(tr, x, y, dev, seed, dropout):
    tr.zero_grad()
    tr.metrics(reset=True)
    xd = torch.from_numpy(x).to(dev)
    yd = torch.from_numpy(y).to(dev)
    pred = torch.empty(y.shape, dtype=torch.float32, device=dev)
    tr.backward(xd, yd, float(y.size), seed=seed, dropout=dropout, freeze_moving=True, pred_dev=pred)
    sq, hits = tr.metrics(reset=True)
    return sq, hits, pred.cpu().numpy(), tr.get_grads()


def _check_grads(got, ref):
    bad = []
    for li, (g, r) in enumerate(zip(got, ref)):
        for name in ('kernel', 'bias', 'gamma', 'beta'):
            if name not in r:
                continue
            rel, cos = _rel(g[name], r[name]), _cos(g[name], r[name])
            ratio = float(np.linalg.norm(g[name]) / max(np.linalg.norm(r[name]), 1e-30))
            ok = cos >= 0.3 and 0.5 <= ratio <= 2.0
            if li >= 17 and name in ('kernel', 'gamma', 'beta'):
                ok = ok and rel <= 0.08
            if not ok:
                bad.append(f'layer {li} {name}: rel {rel:.3e} cos {cos:.5f} norm ratio {ratio:.3f}')
    assert not bad, '\n'.join(bad)


def test_param_layout_matches_layers_txt():
    L, lib = _lib()
    assert lib.itmo_train_param_count() == R.param_count()['trainable'] == 11159011
    assert 2 * lib.itmo_train_bn_count() == R.param_count()['non_trainable'] == 7808


def test_train_forward_activations(step_case, dev):
    ws, x, y, tr = step_case
    _run_backward(tr, x, y, dev, seed=0, dropout=False)
    ref = T.train_forward_backward(ws, x.astype(np.float32) / 255.0, y, None, return_acts=True)
    for li in (0, 1, 2, 7, 9, 10, 11, 16, 17):
        for which in ('z', 'a'):
            got = tr.read_activation(li, which, NB)
            r = ref[which][li]
            assert _rel(got, r) <= 0.15, f'layer {li} {which}: rel {_rel(got, r):.3e}'


def test_train_step_gradients_no_dropout(step_case, dev):
    ws, x, y, tr = step_case
    sq, hits, pred, grads = _run_backward(tr, x, y, dev, seed=0, dropout=False)
    ref = T.train_forward_backward(ws, x.astype(np.float32) / 255.0, y, None)
    assert abs(sq - ref['sq_err_sum']) <= 0.02 * ref['sq_err_sum']
    assert abs(hits / (NB * H * W) - ref['acc']) <= 0.02
    assert R.psnr(pred, ref['pred']) >= 25.0   # ~30 dB measured: 11 % activation noise on these random weights
    _check_grads(grads, ref['grads'])


def test_train_step_gradients_with_dropout(step_case, dev):
    ws, x, y, tr = step_case
    sq, hits, pred, grads = _run_backward(tr, x, y, dev, seed=1234, dropout=True)
    ref = T.train_forward_backward(ws, x.astype(np.float32) / 255.0, y, 1234)
    assert abs(sq - ref['sq_err_sum']) <= 0.02 * ref['sq_err_sum']
    _check_grads(grads, ref['grads'])
    # and the mask really was applied: the same step without dropout differs
    sq2, _, _, _ = _run_backward(tr, x, y, dev, seed=1234, dropout=False)
    assert abs(sq2 - sq) > 1e-6 * sq


def test_backward_is_bit_reproducible(step_case, dev):
    """No floating-point atomics anywhere in the step (BN statistics, BN backward sums, bias / gamma / beta / head
    gradients, split-K weight gradients are fixed-order reductions): two runs of the same micro-batch give
    bit-identical gradient buffers, also from a second engine instance, with Dropout on."""
    from cnn_itmo_b200 import train
    ws, x, y, tr = step_case
    xd, yd = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    runs = []
    for _ in range(3):
        tr.zero_grad()
        tr.backward(xd, yd, float(y.size), seed=77, dropout=True, freeze_moving=True)
        torch.cuda.synchronize()
        runs.append(tr.grads.clone())
    assert torch.equal(runs[0], runs[1]) and torch.equal(runs[0], runs[2])
    assert float(runs[0].abs().sum()) > 0
    tr2 = train.Trainer(ws, micro_batch=NB, image_hw=(H, W), device=0)
    tr2.zero_grad()
    tr2.backward(xd, yd, float(y.size), seed=77, dropout=True, freeze_moving=True)
    torch.cuda.synchronize()
    assert torch.equal(tr2.grads, runs[0])
    tr2.close()


def test_micro_batches_on_one_rank_equal_ranks_summed(step_case, dev):
    """Data-parallel invariant on one GPU: accumulating two micro-batches into one gradient buffer == computing each
    in its own (zeroed) buffer and adding the buffers, bit for bit — every tensor of the gradient receives exactly ONE
    floating-point addition per micro-batch, which is what makes N ranks + all-reduce reproduce the one-rank run
    (tools/dp_check.py asserts the same across two real GPUs)."""
    ws, x, y, tr = step_case
    xs = [torch.from_numpy(x[i:i + 1]).to(dev) for i in range(NB)]
    ys = [torch.from_numpy(y[i:i + 1]).to(dev) for i in range(NB)]
    tr.zero_grad()
    for i in range(NB):
        tr.backward(xs[i], ys[i], float(y.size), seed=i, dropout=True, freeze_moving=True)
    torch.cuda.synchronize()
    together = tr.grads.clone()
    parts = []
    for i in range(NB):
        tr.zero_grad()
        tr.backward(xs[i], ys[i], float(y.size), seed=i, dropout=True, freeze_moving=True)
        torch.cuda.synchronize()
        parts.append(tr.grads.clone())
    assert torch.equal(parts[0] + parts[1], together)


def test_loss_trajectory_20_steps_against_fp32_oracle(dev):
    """20 RMSprop steps (Keras defaults, model.py:281) on one batch at 256x128 against the fp32 autograd oracle stepping
    the same way. RMSprop divides by sqrt(mean g^2): small gradient components become full-size updates, so on a
    2-image batch the trajectory is chaotic — two CORRECT CPU restatements (fp32 and bf16 storage emulated) already
    differ by 0.03 / 1.2 / 4.6 / 1.0 / 2.9 % over the first five steps and 8-22 % in steps 14-20, with update-direction
    cosines 0.87 / 0.96 / 0.997 for conv2d_transpose_4 / conv2d_14 / the head and ~0.4 in the encoder (the device
    measured 0.02-4 % over the first six steps, 12-18 % later, one step at 37 %). The bars here are what a bf16-storage
    pipeline can be held to: early steps within 8 %, the median step within 15 %, every step within 50 %, a
    geometric-mean loss ratio within 8 %, a 10x loss reduction, and the cumulative update of the last three layers
    pointing the same way (cosine >= 0.8 / 0.9 / 0.98)."""
    from cnn_itmo_b200 import train
    ws0 = _weights()
    x, y = T.synthetic_pair(NB, H, W, seed=11)
    xf = x.astype(np.float32) / 255.0
    tr = train.Trainer(ws0, micro_batch=NB, image_hw=(H, W), device=0, optimizer='rmsprop')
    dev_losses = [tr.train_on_batch(x, y, dropout=False)[0] for _ in range(20)]
    dev_w = tr.get_weights()
    tr.close()
    ws = [{k: v.astype(np.float64).copy() for k, v in w.items()} for w in ws0]
    acc = [{k: np.zeros_like(v) for k, v in w.items() if k in ('kernel', 'bias', 'gamma', 'beta')} for w in ws]
    ref_losses = []
    for _ in range(20):
        r = T.train_forward_backward([{k: v.astype(np.float32) for k, v in w.items()} for w in ws], xf, y, None)
        ref_losses.append(r['loss'])
        for li, g in enumerate(r['grads']):
            for k in g:
                ws[li][k], acc[li][k] = T.rmsprop_step(ws[li][k], g[k], acc[li][k], **T.RMSPROP_DEFAULTS)
    d, r = np.asarray(dev_losses), np.asarray(ref_losses)
    rel = np.abs(d - r) / r
    assert rel[:5].max() <= 0.08, rel
    assert np.median(rel) <= 0.15 and rel.max() <= 0.50, rel
    assert abs(np.exp(np.mean(np.log(d / r))) - 1.0) <= 0.08, (d, r)
    assert d[-1] <= 0.1 * d[0]
    for li, bar in ((16, 0.8), (17, 0.9), (18, 0.98)):
        du = dev_w[li]['kernel'].astype(np.float64) - ws0[li]['kernel']
        ru = ws[li]['kernel'] - ws0[li]['kernel']
        assert _cos(du, ru) >= bar, (li, _cos(du, ru))


def test_last_layers_gradients_at_512(dev):
    """BASELINE config 5's image size: one 512x512 image, gradients of the last four weight layers against the fp32
    oracle (head and conv2d_14 within 8 % relative L2; conv2d_transpose_4 / conv2d_13, whose gradients pass through
    two more bf16-stored tensors, within 30 % with cosine >= 0.95)."""
    from cnn_itmo_b200 import train
    ws = _weights(6)
    x, y = T.synthetic_pair(1, 512, 512, seed=19)
    tr = train.Trainer(ws, micro_batch=1, image_hw=(512, 512), device=0)
    tr.zero_grad()
    tr.backward(torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev), float(y.size), seed=0, dropout=False,
                freeze_moving=True)
    sq, _ = tr.metrics()
    got = tr.get_grads()
    tr.close()
    ref = T.train_forward_backward(ws, x.astype(np.float32) / 255.0, y, None)
    assert abs(sq - ref['sq_err_sum']) <= 0.02 * ref['sq_err_sum']
    for li in (18, 17):
        for name in ('kernel', 'bias'):
            assert _rel(got[li][name], ref['grads'][li][name]) <= 0.08, (li, name, _rel(got[li][name], ref['grads'][li][name]))
    for li in (16, 15):
        rel, cos = _rel(got[li]['kernel'], ref['grads'][li]['kernel']), _cos(got[li]['kernel'], ref['grads'][li]['kernel'])
        assert rel <= 0.30 and cos >= 0.95, (li, rel, cos)


def test_gradient_accumulates_over_micro_batches(step_case, dev):
    """Two micro-batches with the whole batch's loss_count accumulate; BN statistics are per micro-batch."""
    ws, x, y, tr = step_case
    tr.zero_grad()
    for i in range(NB):
        tr.backward(torch.from_numpy(x[i:i + 1]).to(dev), torch.from_numpy(y[i:i + 1]).to(dev), float(y.size), seed=0,
                    dropout=False, freeze_moving=True)
    got = tr.get_grads()
    ref = [None] * NB
    for i in range(NB):
        ref[i] = T.train_forward_backward(ws, x[i:i + 1].astype(np.float32) / 255.0, y[i:i + 1], None,
                                          loss_count=float(y.size))['grads']
    summed = [{k: ref[0][li][k] + ref[1][li][k] for k in ref[0][li]} for li in range(len(ref[0]))]
    _check_grads(got, summed)


@pytest.mark.parametrize('opt', ['rmsprop', 'adam'])
def test_optimizer_update_and_moving_stats(dev, opt):
    from cnn_itmo_b200 import train
    ws = _weights(5)
    x, y = T.synthetic_pair(NB, H, W, seed=12)
    tr = train.Trainer(ws, micro_batch=NB, image_hw=(H, W), device=0, optimizer=opt)
    p0 = tr.params.cpu().numpy().astype(np.float64)
    mov0 = tr.bn_moving.cpu().numpy().astype(np.float64)
    tr.zero_grad()
    tr.backward(torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev), float(y.size), seed=0, dropout=False)
    g = tr.grads.cpu().numpy().astype(np.float64)
    tr.update()
    p1 = tr.params.cpu().numpy().astype(np.float64)
    if opt == 'rmsprop':
        want, _ = T.rmsprop_step(p0, g, np.zeros_like(p0), **T.RMSPROP_DEFAULTS)
    else:
        want, _, _ = T.adam_step(p0, g, np.zeros_like(p0), np.zeros_like(p0), 1, **T.ADAM_COMMENT_DEFAULTS)
    # fp32 arithmetic on the device; updates are O(lr) so compare the deltas
    # fp32 parameters: the update of a parameter of magnitude ~1 is rounded to its ulp (1.2e-7)
    assert np.abs(p1 - want).max() <= 2.5e-7
    # moving statistics of the first and a deep layer against the oracle's batch statistics
    ref = T.train_forward_backward(ws, x.astype(np.float32) / 255.0, y, None)
    mov1 = tr.bn_moving.cpu().numpy().astype(np.float64)
    nbn = tr.n_bn
    for li in (0, 9, 17):
        off, cnt = tr.layout[li]['bn']
        l = R.LAYERS[li]
        scale = 2 if l.kind == 'convT' else 1
        count = NB * (scale * l.hin * H // 512) * (scale * l.hin * W // 512)
        mm, mv = T.moving_update(mov0[off:off + cnt], mov0[nbn + off:nbn + off + cnt], ref['batch_mean'][li],
                                 ref['batch_var'][li], count)
        assert np.abs(mov1[off:off + cnt] - mm).max() <= 2e-3 * max(1.0, np.abs(mm).max())
        assert np.abs(mov1[nbn + off:nbn + off + cnt] - mv).max() <= 2e-3 * max(1.0, np.abs(mv).max())
    tr.close()


def test_loss_decreases_and_model_api(dev):
    """Model.train_on_batch (the Keras surface, main.py:126-132) drives the same engine; a few steps on one batch must
    reduce the mse, and the trained weights flow back into predict()."""
    import cnn_itmo_b200 as M
    m = M.Model(M.initial_weights(7), precision='bf16', max_batch=2)
    m.compile(optimizer='rmsprop', loss='mse', metrics=['accuracy'])
    m.train_micro_batch = 2
    x, y = T.synthetic_pair(2, 512, 512, seed=21)
    w_before = m.get_weights()
    out_before = m.predict(x.astype(np.float32) / 255.0)
    losses = [m.train_on_batch(x, y)[0] for _ in range(6)]
    assert all(np.isfinite(losses)), losses
    assert losses[-1] < losses[0], losses
    # predict / get_weights after train_on_batch must see the TRAINED parameters without an explicit pull
    # (the reference's per-epoch makePrediction callback, main.py:36-63, calls model.predict between training steps)
    w_after = m.get_weights()
    assert any(not np.array_equal(a, b) for a, b in zip(w_before, w_after))
    out = m.predict(x.astype(np.float32) / 255.0)
    assert out.shape == (2, 512, 512, 3) and np.isfinite(out).all()
    assert np.abs(out - out_before).max() > 1e-3
    m.train_on_batch(x, y)
    out2 = m.predict(x.astype(np.float32) / 255.0)          # dirty again -> pulled again
    assert np.abs(out2 - out).max() > 0
    m.close()


def test_prefetched_next_batch_changes_nothing_but_the_overlap(dev):
    """Trainer.train_on_batch(prefetch=next batch) uploads the next step's first micro-batch under this step's compute
    (what Model.fit_generator's look-ahead does, main.py:126-132): losses and weights are bit-identical to plain calls,
    a prefetch that is not followed by the announced batch is dropped, and fit_generator pulls exactly
    epochs * steps_per_epoch batches from the generator."""
    from cnn_itmo_b200 import train
    import cnn_itmo_b200 as M
    ws0 = _weights()
    batches = [T.synthetic_pair(4, H, W, seed=40 + i) for i in range(3)]

    def run(prefetching):
        tr = train.Trainer(ws0, micro_batch=2, image_hw=(H, W), device=0, optimizer='rmsprop')
        out = []
        for i, (x, y) in enumerate(batches):
            nxt = batches[i + 1] if prefetching and i + 1 < len(batches) else None
            out.append(tr.train_on_batch(x, y, prefetch=nxt))
        if prefetching:   # announced batch 0, then trained on batch 2: the staged copy must not be used
            out.append(tr.train_on_batch(*batches[1], prefetch=batches[0]))
            out.append(tr.train_on_batch(*batches[2]))
        else:
            out.append(tr.train_on_batch(*batches[1]))
            out.append(tr.train_on_batch(*batches[2]))
        w = tr.get_weights()
        tr.close()
        return out, w
    a, wa = run(False)
    b, wb = run(True)
    assert a == b, (a, b)
    for la, lb in zip(wa, wb):
        for k in la:
            assert np.array_equal(la[k], lb[k]), k
    # the Keras surface: look-ahead of one batch, no extra pull at the end
    m = M.Model(M.initial_weights(7), precision='bf16', max_batch=2)
    m.compile(optimizer='rmsprop', loss='mse', metrics=['accuracy'])
    m.train_micro_batch = 2
    m.train_seed = 5          # the Dropout mask stream (None = fresh entropy per engine)
    pulled = []
    full = [T.synthetic_pair(2, 512, 512, seed=50 + i) for i in range(4)]

    def gen():
        for i, b_ in enumerate(full):
            pulled.append(i)
            yield b_
    h = m.fit_generator(gen(), steps_per_epoch=2, epochs=2, verbose=0)
    assert pulled == [0, 1, 2, 3] and len(h['loss']) == 2 and all(np.isfinite(h['loss']))
    m2 = M.Model(M.initial_weights(7), precision='bf16', max_batch=2)
    m2.compile(optimizer='rmsprop', loss='mse', metrics=['accuracy'])
    m2.train_micro_batch = 2
    m2.train_seed = 5
    plain = [m2.train_on_batch(x, y)[0] for x, y in full]
    assert h['loss'] == [float(np.mean(plain[:2])), float(np.mean(plain[2:]))]
    m.close()
    m2.close()


def test_adam_resume_continues_bias_correction(dev, tmp_path):
    """save after 2 Adam steps, load, 1 more step == 3 uninterrupted steps, bit for bit (the step counter
    behind Adam's lr_t = lr sqrt(1-b2^t)/(1-b1^t) travels through /optimizer_weights and itmo_train_set_step)."""
    import cnn_itmo_b200 as M
    x, y = T.synthetic_pair(2, 256, 256, seed=41)
    x = np.tile(x, (1, 2, 2, 1))
    y = np.tile(y, (1, 2, 2, 1))

    def fresh():
        m = M.Model(M.initial_weights(9), precision='fp16', max_batch=2)
        m.compile(optimizer='adam', loss='mse', metrics=['accuracy'])
        m.train_micro_batch = 2
        m.train_seed = 5
        return m
    a = fresh()
    for _ in range(3):
        a.train_on_batch(x, y)
    want = a._trainer.params.cpu().numpy().copy()
    assert a._trainer.step_index == 3
    a.close()
    b = fresh()
    for _ in range(2):
        b.train_on_batch(x, y)
    p = str(tmp_path / 'adam.hdf5')
    b.save(p)
    b.close()
    c = M.load_model(p, precision='fp16', max_batch=2)
    c.train_micro_batch = 2
    c.train_seed = 5
    assert c.optimizer == 'adam'
    c.train_on_batch(x, y)
    assert c._trainer.step_index == 3
    got = c._trainer.params.cpu().numpy()
    # the checkpoint stores fp32 parameters and moments exactly, the kernels are deterministic: identical
    np.testing.assert_array_equal(got, want)
    c.close()


def test_checkpoint_resumes_optimizer_state(dev, tmp_path):
    """Model.save after training writes /optimizer_weights; load_model + the next train_on_batch starts
    from the same RMSprop accumulators and parameters (what ModelCheckpoint + keras.models.load_model give)."""
    import cnn_itmo_b200 as M
    m = M.Model(M.initial_weights(7), precision='bf16', max_batch=2)
    m.compile(optimizer='rmsprop', loss='mse', metrics=['accuracy'])
    m.train_micro_batch = 2
    x, y = T.synthetic_pair(2, 512, 512, seed=33)
    for _ in range(2):
        m.train_on_batch(x, y)
    p = str(tmp_path / 'saved7-model-02-0.50.hdf5')
    m.save(p)
    state = m._trainer.opt_state.cpu().numpy().copy()
    params = m._trainer.params.cpu().numpy().copy()
    assert np.abs(state[:m._trainer.n_params]).max() > 0
    m.close()
    r = M.load_model(p, precision='bf16', max_batch=2)
    r.train_micro_batch = 2
    assert len(r.meta['optimizer_weights']) == 74
    tr = r._ensure_trainer()
    np.testing.assert_array_equal(tr.opt_state.cpu().numpy(), state)
    np.testing.assert_array_equal(tr.params.cpu().numpy(), params)
    loss = r.train_on_batch(x, y)[0]
    assert np.isfinite(loss)
    r.close()


def test_fused_exchange_step_equals_plain_update_on_one_rank(dev):
    """itmo_exchange_step (reduce-scatter -> optimizer -> all-gather over peer memory, ONE kernel) with a single rank is
    the plain optimizer step: same parameters and optimizer state bit for bit, RMSprop and Adam, over two steps (the
    window-backed flat buffers, the cooperative launch and the flag barriers are all exercised; tools/dp_check.py
    asserts the N-rank property on real peers)."""
    from cnn_itmo_b200 import train
    x, y = T.synthetic_pair(2, 256, 128, seed=5)
    ws = R.synthetic_weights(4)
    xd, yd = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    for opt in ('rmsprop', 'adam'):
        a = train.Trainer(ws, micro_batch=2, image_hw=(256, 128), device=0, optimizer=opt, exchange='fused')
        b = train.Trainer(ws, micro_batch=2, image_hw=(256, 128), device=0, optimizer=opt)
        for step in range(2):
            for tr in (a, b):
                tr.zero_grad()
                tr.backward(xd, yd, float(y.size), seed=step, dropout=True)
                tr.exchange_and_update()
        torch.cuda.synchronize()
        assert torch.equal(a.params, b.params), opt
        assert torch.equal(a.opt_state, b.opt_state), opt
        assert a.step_index == b.step_index == 2
        assert float((a.params - torch.from_numpy(train.flatten_layers(ws)).to(dev)).abs().max()) > 0
        a.close()
        b.close()
