"""Layer-local parity of the training step on a B200 (``-m gpu``).

Whole-network gradient comparisons of a bf16 pipeline against an fp32 oracle are dominated by rounding noise that the
18-layer ReLU->BN stack amplifies (two CORRECT CPU restatements, fp32 and bf16-storage, already differ by 40-70 % in the
early layers' gradients on this problem — measured in tests/test_gpu_train.py's docstring). So every kernel of the step is
also checked in isolation: for each layer the device's OWN stored tensors (layer inputs, z = ReLU(conv+b), the incoming
gradient dL/da) are fed to a float64 torch restatement of that one layer (model.py:195-200 in training mode), and the
device's outputs of that layer — z, a, dL/d(pre-activation), dW, db, dgamma, dbeta and the gradient handed to the producing
layers (through the max-pool routing and the skip-connection sum) — are compared with it. The only differences left are
the bf16 rounding of each stored output and fp32 summation order: relative L2 <= 1e-2 (6e-3 for forward tensors).
"""
import numpy as np
import pytest

from oracle import unet_ref as R
from oracle import unet_train_ref as T

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')
F = torch.nn.functional

H, W, NB = 256, 128, 2


def _rel(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def _bf16(t):
    return t.to(torch.bfloat16).to(torch.float64)


def _nchw(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(torch.float64).permute(0, 3, 1, 2).contiguous()


def _nhwc(t):
    return t.detach().permute(0, 2, 3, 1).contiguous().numpy()


def _weights(seed=3):
    ws = R.synthetic_weights(seed)
    rng = np.random.default_rng(seed)
    for w in ws:
        if 'gamma' in w:
            w['gamma'] = (1.0 + 0.1 * rng.standard_normal(w['gamma'].shape)).astype(np.float32)
            w['beta'] = (0.2 + 0.1 * rng.standard_normal(w['beta'].shape)).astype(np.float32)
    return ws


@pytest.mark.parametrize('dropout', [False, True])
def test_every_layer_against_local_float64_restatement(dropout):
    assert torch.cuda.is_available()
    from cnn_itmo_b200 import train
    dev = torch.device('cuda:0')
    ws = _weights()
    x, y = T.synthetic_pair(NB, H, W, seed=11)
    seed = 4321
    tr = train.Trainer(ws, micro_batch=NB, image_hw=(H, W), device=0)
    tr.zero_grad()
    tr.backward(torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev), float(y.size), seed=seed, dropout=dropout,
                freeze_moving=True, keep_grads=True)
    grads = tr.get_grads()
    rd = lambda li, which: tr.read_activation(li, which, NB)
    bad = []

    def check(name, got, ref, tol):
        r = _rel(got, ref)
        if not r <= tol:
            bad.append(f'{name}: rel {r:.3e} > {tol}')

    # ---- head (layer 18): loss gradient into a[17], head parameter gradients
    a17 = _nchw(rd(17, 'a')).requires_grad_(True)
    k18 = torch.from_numpy(ws[18]['kernel']).double().requires_grad_(True)
    b18 = torch.from_numpy(ws[18]['bias']).double().requires_grad_(True)
    pred = torch.sigmoid(F.conv2d(a17, k18.permute(3, 2, 0, 1), b18))
    (((pred - _nchw(y)) ** 2).sum() / float(y.size)).backward()
    check('head dA[17]', rd(17, 'dA'), _nhwc(a17.grad), 1e-2)
    check('head kernel', grads[18]['kernel'], k18.grad.numpy(), 5e-3)
    check('head bias', grads[18]['bias'], b18.grad.numpy(), 5e-3)

    contrib = {}   # (source layer, 'a' | 'pooled') -> expected bf16 gradient contribution (NCHW float64)
    for li in range(17, -1, -1):
        s = R.LAYERS[li]
        w = ws[li]
        srcs = []
        for j, si in enumerate(s.src):
            if si < 0:
                srcs.append(_nchw(x.astype(np.float32) / np.float32(255.0)))
            else:
                srcs.append(_nchw(rd(si, 'pooled' if (s.src_pooled and j == 0) else 'a')))
        inp = torch.cat(srcs, dim=1).requires_grad_(True)
        kern = torch.from_numpy(w['kernel']).double()
        if li != 0:
            kern = _bf16(kern)          # tensor-core operand copy
        kern.requires_grad_(True)
        bias = torch.from_numpy(w['bias']).double().requires_grad_(True)
        if s.kind == 'conv':
            pre = F.conv2d(inp, kern.permute(3, 2, 0, 1), bias, padding=1)
        else:
            pre = F.conv_transpose2d(inp, kern.permute(3, 2, 0, 1), bias, stride=2)
        z_dev = _nchw(rd(li, 'z'))
        check(f'layer {li} z', _nhwc(z_dev), _nhwc(torch.relu(pre)), 6e-3)
        # BN (+dropout) forward / backward on the device's own z
        zl = z_dev.clone().requires_grad_(True)
        gamma = torch.from_numpy(w['gamma']).double().requires_grad_(True)
        beta = torch.from_numpy(w['beta']).double().requires_grad_(True)
        m = zl.mean(dim=(0, 2, 3))
        v = zl.var(dim=(0, 2, 3), unbiased=False)
        a = gamma[None, :, None, None] * (zl - m[None, :, None, None]) / torch.sqrt(v[None, :, None, None] + R.BN_EPS) \
            + beta[None, :, None, None]
        if s.dropout and dropout:
            n, c, h, w_ = a.shape
            keep = T.dropout_keep_mask((n, h, w_, c), T.layer_dropout_seed(seed, li))
            a = a * torch.from_numpy(np.ascontiguousarray(keep.transpose(0, 3, 1, 2))).double() * 2.0
        check(f'layer {li} a', rd(li, 'a'), _nhwc(a), 6e-3)
        if s.pool:
            check(f'layer {li} pooled', rd(li, 'pooled'), _nhwc(F.max_pool2d(_nchw(rd(li, 'a')), 2, 2)), 1e-6)
        dA_dev = _nchw(rd(li, 'dA'))
        # what the consumers of this layer should have handed back
        if li < 17:
            if s.pool:
                al = _nchw(rd(li, 'a')).requires_grad_(True)
                F.max_pool2d(al, 2, 2).backward(contrib[(li, 'pooled')])
                exp = _bf16(contrib[(li, 'a')] + al.grad)
            else:
                exp = contrib[(li, 'a')]
            check(f'layer {li} dA (from consumers)', _nhwc(dA_dev), _nhwc(exp), 1e-2)
        a.backward(dA_dev)
        dpre_ref = zl.grad * (z_dev > 0).double()
        dpre_dev = _nchw(rd(li, 'dpre'))
        check(f'layer {li} dpre', _nhwc(dpre_dev), _nhwc(dpre_ref), 1e-2)
        check(f'layer {li} dgamma', grads[li]['gamma'], gamma.grad.numpy(), 1e-2)
        check(f'layer {li} dbeta', grads[li]['beta'], beta.grad.numpy(), 1e-2)
        # conv backward on the device's own (rounded) dpre
        pre.backward(dpre_dev)
        check(f'layer {li} dW', grads[li]['kernel'], kern.grad.numpy(), 1e-2 if li == 0 else 5e-3)
        check(f'layer {li} db', grads[li]['bias'], bias.grad.numpy(), 5e-3)
        if li > 0:
            off = 0
            for j, si in enumerate(s.src):
                c = s.cin[j]
                key = (si, 'pooled' if (s.src_pooled and j == 0) else 'a')
                contrib[key] = _bf16(inp.grad[:, off:off + c])
                off += c
    tr.close()
    assert not bad, '\n'.join(bad)
