"""GPU parity tests (run with ``-m gpu`` on a B200): the CUDA path, called through the C ABI of
include/itmo_b200.h, against the CPU oracle (oracle/unet_ref.py) on the same seeded inputs.

Tolerances are the ones BASELINE.json's north_star states:
  * TF32 / fp32 mode: max abs error <= 1e-3 on the [0,1] network output (and on O(1) layer activations);
  * BF16 mode: PSNR >= 45 dB against the fp32 oracle on the network output; per layer the only error is the
    bf16 rounding of the stored result, |err| <= 2^-8 |ref| + 1e-2;
  * uint8 output (predict.py:64 truncation): +-1 LSB where sigmoid*255 sits within float error of an integer.
Nothing here reads /root/reference.
"""
import ctypes
import os

import numpy as np
import pytest

from oracle import unet_ref as R

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


PRECS = ['fp32', 'tf32', 'fp16', 'bf16']


def _quant(t, prec):
    """Round a float32 CPU tensor to what the device stores / the tensor core reads (fp32 mode: nothing —
    the 3xTF32 path consumes full fp32 values)."""
    if prec == 'bf16':
        return t.to(torch.bfloat16).float()
    if prec == 'fp16':
        return t.to(torch.float16).float()
    if prec == 'tf32':
        return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)
    return t


def _dt(prec):
    return {'bf16': torch.bfloat16, 'fp16': torch.float16}.get(prec, torch.float32)


def _P(L, prec):
    return {'bf16': L.PREC_BF16, 'fp16': L.PREC_FP16, 'tf32': L.PREC_TF32, 'fp32': L.PREC_FP32}[prec]


def _tol(ref, prec):
    """Per-layer tolerance with operands the device represents exactly: what is left is the rounding of the
    stored result (bf16 2^-9, fp16 / tf32 2^-11 relative) plus fp32 accumulation noise; fp32 mode stores fp32."""
    if prec == 'bf16':
        return (2.0 ** -8) * np.abs(ref) + 1e-2
    if prec in ('fp16', 'tf32'):
        return (2.0 ** -10) * np.abs(ref) + 1e-3
    return np.full_like(ref, 2e-4)


# every 3x3 ConvBN shape of the graph: (hin, c0, c1, cout, pool)   (layers.txt:6-136)
CONV_SHAPES = [
    (512, 32, 0, 32, True), (256, 32, 0, 64, False), (256, 64, 0, 64, True), (128, 64, 0, 128, False),
    (128, 128, 0, 128, True), (64, 128, 0, 256, False), (64, 256, 0, 256, True), (32, 256, 0, 512, False),
    (32, 512, 0, 512, False), (64, 256, 512, 512, False), (128, 128, 256, 256, False), (256, 64, 128, 128, False),
    (512, 32, 64, 64, False),
]
CONVT_SHAPES = [(32, 512, 512), (64, 512, 256), (128, 256, 128), (256, 128, 64)]


def _spec(kind, index=5):
    return R.LayerSpec(index, kind, 'x', 'bn', (1,), 1, 1, False, False, (0,), False)


def _bn_params(cout, g):
    """Non-trivial BatchNormalization statistics incl. negative gamma (breaks a pool-before-BN fusion)."""
    gamma = torch.randn(cout, generator=g) * 0.25 + 1.0
    gamma[::7] *= -1
    beta = torch.randn(cout, generator=g) * 0.1
    mean = torch.rand(cout, generator=g) * 0.4
    var = torch.rand(cout, generator=g) * 0.9 + 0.1
    scale = gamma / torch.sqrt(var + R.BN_EPS)
    shift = beta - mean * scale
    return dict(gamma=gamma.double(), beta=beta.double(), mean=mean.double(), var=var.double()), scale, shift


def _oracle_block(kind, x_nhwc, k, bias, bn):
    """ConvBN / ConvBNTranspose of the oracle (model.py:195-200) in float64 on NHWC input."""
    w = dict(bn)
    w['bias'] = bias.double()
    w['kernel'] = k.double().permute(3, 2, 0, 1).contiguous()
    y = R.block_torch(_spec(kind), w, x_nhwc.double().permute(0, 3, 1, 2).contiguous())
    return y


@pytest.mark.parametrize('prec', PRECS)
@pytest.mark.parametrize('shape', CONV_SHAPES, ids=lambda s: f'{s[0]}x{s[1]}+{s[2]}to{s[3]}{"p" if s[4] else ""}')
def test_conv3x3_bn_every_shape(dev, prec, shape):
    """BASELINE config 2: every ConvBN channel width at its native spatial size (batch 2 here; the oracle is
    CPU), incl. the two-source (skip, up) operand loader and the fused 2x2 max-pool dual store."""
    L, lib = _lib()
    h, c0, c1, cout, pool = shape
    n = 2
    g = torch.Generator().manual_seed(h * 1000 + c0 + c1 + cout)
    dt, P = _dt(prec), _P(L, prec)
    cin = c0 + c1
    x0 = _quant(torch.randn(n, h, h, c0, generator=g), prec)
    x1 = _quant(torch.randn(n, h, h, c1, generator=g), prec) if c1 else None
    k = _quant(torch.randn(3, 3, cin, cout, generator=g) * (2.0 / (9 * cin)) ** 0.5, prec)
    bias = torch.rand(cout, generator=g) * 0.2 - 0.1
    bn, scale, shift = _bn_params(cout, g)
    ref = _oracle_block('conv', torch.cat([x0, x1], 3) if c1 else x0, k, bias, bn)
    ref_pool = torch.nn.functional.max_pool2d(ref, 2, 2).permute(0, 2, 3, 1).numpy() if pool else None
    ref = ref.permute(0, 2, 3, 1).numpy()

    d = lambda t: t.to(dev).contiguous()
    x0d, x1d = d(x0.to(dt)), (d(x1.to(dt)) if c1 else None)
    kd, bd, sd, td = d(k), d(bias), d(scale), d(shift)
    out = torch.full((n, h, h, cout), float('nan'), dtype=dt, device=dev)
    pooled = torch.full((n, h // 2, h // 2, cout), float('nan'), dtype=dt, device=dev) if pool else None
    L.check(lib.itmo_conv3x3_bn(P, x0d.data_ptr(), c0, x1d.data_ptr() if c1 else None, c1, n, h, h, kd.data_ptr(),
                                bd.data_ptr(), sd.data_ptr(), td.data_ptr(), cout, out.data_ptr(),
                                pooled.data_ptr() if pool else None, None))
    torch.cuda.synchronize()
    got = out.float().cpu().double().numpy()
    assert not np.isnan(got).any()
    assert (np.abs(got - ref) <= _tol(ref, prec)).all(), float(np.abs(got - ref).max())
    if pool:
        gp = pooled.float().cpu().double().numpy()
        assert (np.abs(gp - ref_pool) <= _tol(ref_pool, prec)).all(), float(np.abs(gp - ref_pool).max())


@pytest.mark.parametrize('prec', PRECS)
@pytest.mark.parametrize('shape', CONVT_SHAPES, ids=lambda s: f'{s[0]}x{s[1]}to{s[2]}')
def test_convT2x2_bn_every_shape(dev, prec, shape):
    L, lib = _lib()
    h, cin, cout = shape
    n = 2
    g = torch.Generator().manual_seed(h * 77 + cin + cout)
    dt, P = _dt(prec), _P(L, prec)
    x = _quant(torch.randn(n, h, h, cin, generator=g), prec)
    k = _quant(torch.randn(2, 2, cout, cin, generator=g) * (2.0 / (4 * cin)) ** 0.5, prec)
    bias = torch.rand(cout, generator=g) * 0.2 - 0.1
    bn, scale, shift = _bn_params(cout, g)
    ref = _oracle_block('convT', x, k, bias, bn).permute(0, 2, 3, 1).numpy()
    d = lambda t: t.to(dev).contiguous()
    xd, kd, bd, sd, td = d(x.to(dt)), d(k), d(bias), d(scale), d(shift)
    out = torch.full((n, 2 * h, 2 * h, cout), float('nan'), dtype=dt, device=dev)
    L.check(lib.itmo_convT2x2_bn(P, xd.data_ptr(), cin, n, h, h, kd.data_ptr(), bd.data_ptr(), sd.data_ptr(),
                                 td.data_ptr(), cout, out.data_ptr(), None))
    torch.cuda.synchronize()
    got = out.float().cpu().double().numpy()
    assert not np.isnan(got).any()
    assert (np.abs(got - ref) <= _tol(ref, prec)).all(), float(np.abs(got - ref).max())


@pytest.mark.parametrize('prec', PRECS)
@pytest.mark.parametrize('in_dtype', ['u8', 'f32'])
def test_first_layer_from_raw_pixels(dev, prec, in_dtype):
    """conv2d_1 (3 -> 32, model.py:208) with the /255 of predict.py:59 fused into the load."""
    L, lib = _lib()
    n, h = 2, 512
    g = torch.Generator().manual_seed(11)
    dt, P = _dt(prec), _P(L, prec)
    if in_dtype == 'u8':
        xin = torch.randint(0, 256, (n, h, h, 3), generator=g, dtype=torch.uint8)
        xf = torch.from_numpy(R.preprocess_u8(xin.numpy()))
    else:
        xin = _quant(torch.rand(n, h, h, 3, generator=g), prec)   # 16-bit modes read the floats as 16-bit operands
        xf = xin
    k = torch.randn(3, 3, 3, 32, generator=g) * (2.0 / 27) ** 0.5
    bias = torch.rand(32, generator=g) * 0.2 - 0.1
    bn, scale, shift = _bn_params(32, g)
    ref = _oracle_block('conv', xf, k, bias, bn).permute(0, 2, 3, 1).numpy()
    d = lambda t: t.to(dev).contiguous()
    xd, kd, bd, sd, td = d(xin), d(k), d(bias), d(scale), d(shift)
    out = torch.full((n, h, h, 32), float('nan'), dtype=dt, device=dev)
    L.check(lib.itmo_conv3x3_bn_rgb(P, xd.data_ptr(), L.DT_U8 if in_dtype == 'u8' else L.DT_F32, n, h, h,
                                    kd.data_ptr(), bd.data_ptr(), sd.data_ptr(), td.data_ptr(), 32, out.data_ptr(),
                                    None))
    torch.cuda.synchronize()
    got = out.float().cpu().double().numpy()
    tol = np.full_like(ref, 1e-5) if prec == 'fp32' else _tol(ref, prec)   # layer 0 is plain fp32 FMA
    assert (np.abs(got - ref) <= tol).all(), float(np.abs(got - ref).max())


@pytest.mark.parametrize('offset', [0, 1, 2, 3])
def test_first_layer_uint8_pointer_alignment(dev, offset):
    """conv2d_1 from uint8 pixels at every byte alignment of the input pointer: 4-byte aligned pointers take the
    word-wise prefetched halo path of rgb_conv_tc_kernel, the others its byte-wise path; both must give the same
    result (bit-exact: same operands, same MMA)."""
    L, lib = _lib()
    n, h = 3, 512
    g = torch.Generator().manual_seed(23 + offset)
    xin = torch.randint(0, 256, (n, h, h, 3), generator=g, dtype=torch.uint8)
    k = torch.randn(3, 3, 3, 32, generator=g) * (2.0 / 27) ** 0.5
    bias = torch.rand(32, generator=g) * 0.2 - 0.1
    bn, scale, shift = _bn_params(32, g)
    ref = _oracle_block('conv', torch.from_numpy(R.preprocess_u8(xin.numpy())), k, bias, bn).permute(0, 2, 3, 1).numpy()
    d = lambda t: t.to(dev).contiguous()
    kd, bd, sd, td = d(k), d(bias), d(scale), d(shift)
    buf = torch.zeros(xin.numel() + 8, dtype=torch.uint8, device=dev)
    assert buf.data_ptr() % 4 == 0
    buf[offset:offset + xin.numel()] = xin.to(dev).reshape(-1)
    out = torch.full((n, h, h, 32), float('nan'), dtype=torch.bfloat16, device=dev)
    L.check(lib.itmo_conv3x3_bn_rgb(L.PREC_BF16, buf.data_ptr() + offset, L.DT_U8, n, h, h, kd.data_ptr(),
                                    bd.data_ptr(), sd.data_ptr(), td.data_ptr(), 32, out.data_ptr(), None))
    torch.cuda.synchronize()
    got = out.float().cpu().double().numpy()
    assert (np.abs(got - ref) <= _tol(ref, 'bf16')).all(), float(np.abs(got - ref).max())
    if offset:
        aligned = torch.full_like(out, float('nan'))
        xd = xin.to(dev).contiguous()
        L.check(lib.itmo_conv3x3_bn_rgb(L.PREC_BF16, xd.data_ptr(), L.DT_U8, n, h, h, kd.data_ptr(), bd.data_ptr(),
                                        sd.data_ptr(), td.data_ptr(), 32, aligned.data_ptr(), None))
        torch.cuda.synchronize()
        assert torch.equal(aligned.view(torch.int16), out.view(torch.int16))


@pytest.mark.parametrize('prec', PRECS)
def test_last_conv_with_fused_head(dev, prec):
    """conv2d_14 + Conv2D(3,1,sigmoid) (model.py:262,276) + the uint8 truncation of predict.py:64."""
    L, lib = _lib()
    n, h, c0, c1 = 1, 512, 32, 64
    g = torch.Generator().manual_seed(5)
    dt, P = _dt(prec), _P(L, prec)
    x0 = _quant(torch.randn(n, h, h, c0, generator=g), prec)
    x1 = _quant(torch.randn(n, h, h, c1, generator=g), prec)
    k = _quant(torch.randn(3, 3, 96, 64, generator=g) * (2.0 / (9 * 96)) ** 0.5, prec)
    bias = torch.rand(64, generator=g) * 0.2 - 0.1
    bn, scale, shift = _bn_params(64, g)
    hk = torch.randn(1, 1, 64, 3, generator=g) * 0.2
    hb = torch.randn(3, generator=g) * 0.1
    y = _oracle_block('conv', torch.cat([x0, x1], 3), k, bias, bn)
    z = torch.nn.functional.conv2d(y, hk.double().permute(3, 2, 0, 1), hb.double())
    ref = torch.sigmoid(z).permute(0, 2, 3, 1).numpy()
    d = lambda t: t.to(dev).contiguous()
    x0d, x1d, kd, bd, sd, td, hkd, hbd = d(x0.to(dt)), d(x1.to(dt)), d(k), d(bias), d(scale), d(shift), d(hk), d(hb)
    of = torch.zeros(n, h, h, 3, dtype=torch.float32, device=dev)
    ou = torch.zeros(n, h, h, 3, dtype=torch.uint8, device=dev)
    for o, odt in ((of, L.DT_F32), (ou, L.DT_U8)):
        L.check(lib.itmo_conv3x3_bn_head(P, x0d.data_ptr(), c0, x1d.data_ptr(), c1, n, h, h, kd.data_ptr(),
                                         bd.data_ptr(), sd.data_ptr(), td.data_ptr(), hkd.data_ptr(), hbd.data_ptr(),
                                         o.data_ptr(), odt, None))
    torch.cuda.synchronize()
    gf = of.cpu().double().numpy()
    assert np.abs(gf - ref).max() <= 1e-4          # the head runs in fp32 on fp32 accumulators in every mode
    want = R.postprocess_u8(ref.astype(np.float32)).astype(np.int32)
    du = np.abs(ou.cpu().numpy().astype(np.int32) - want)
    assert du.max() <= 1 and (du != 0).mean() < 1e-3


# ragged / minimum tiles: one 16x8 M block, tiles narrower than the preferred 4-block width, odd tile counts
SMALL_SHAPES = [(1, 16, 8, 64, 0, 64, False), (1, 16, 16, 32, 0, 32, True), (3, 16, 24, 64, 0, 128, False),
                (1, 32, 8, 128, 0, 256, True), (2, 16, 40, 32, 64, 64, False), (1, 48, 16, 128, 256, 512, False),
                # 64 output channels, width a multiple of 128: conv_rows.cuh's row tiles (one tile wide / three wide, odd counts)
                (1, 16, 128, 32, 0, 64, False), (3, 48, 384, 64, 0, 64, False), (2, 32, 128, 64, 0, 64, True)]


@pytest.mark.parametrize('prec', PRECS)
@pytest.mark.parametrize('shape', SMALL_SHAPES, ids=lambda s: 'x'.join(str(v) for v in s[:6]))
def test_conv3x3_bn_small_and_ragged(dev, prec, shape):
    L, lib = _lib()
    n, h, w, c0, c1, cout, pool = shape
    g = torch.Generator().manual_seed(h * 31 + w * 7 + c0 + c1 + cout)
    dt, P = _dt(prec), _P(L, prec)
    cin = c0 + c1
    x0 = _quant(torch.randn(n, h, w, c0, generator=g), prec)
    x1 = _quant(torch.randn(n, h, w, c1, generator=g), prec) if c1 else None
    k = _quant(torch.randn(3, 3, cin, cout, generator=g) * (2.0 / (9 * cin)) ** 0.5, prec)
    bias = torch.rand(cout, generator=g) * 0.2 - 0.1
    bn, scale, shift = _bn_params(cout, g)
    ref = _oracle_block('conv', torch.cat([x0, x1], 3) if c1 else x0, k, bias, bn)
    ref_pool = torch.nn.functional.max_pool2d(ref, 2, 2).permute(0, 2, 3, 1).numpy() if pool else None
    ref = ref.permute(0, 2, 3, 1).numpy()
    d = lambda t: t.to(dev).contiguous()
    x0d, x1d = d(x0.to(dt)), (d(x1.to(dt)) if c1 else None)
    kd, bd, sd, td = d(k), d(bias), d(scale), d(shift)
    out = torch.full((n, h, w, cout), float('nan'), dtype=dt, device=dev)
    pooled = torch.full((n, h // 2, w // 2, cout), float('nan'), dtype=dt, device=dev) if pool else None
    L.check(lib.itmo_conv3x3_bn(P, x0d.data_ptr(), c0, x1d.data_ptr() if c1 else None, c1, n, h, w, kd.data_ptr(),
                                bd.data_ptr(), sd.data_ptr(), td.data_ptr(), cout, out.data_ptr(),
                                pooled.data_ptr() if pool else None, None))
    torch.cuda.synchronize()
    got = out.float().cpu().double().numpy()
    assert not np.isnan(got).any()
    assert (np.abs(got - ref) <= _tol(ref, prec)).all(), float(np.abs(got - ref).max())
    if pool:
        gp = pooled.float().cpu().double().numpy()
        assert (np.abs(gp - ref_pool) <= _tol(ref_pool, prec)).all(), float(np.abs(gp - ref_pool).max())


@pytest.mark.parametrize('env', [{'ITMO_ROWS': '0'}, {'ITMO_ROWS_POOL': '0'}], ids=['igemm', 'no_rows_pool'])
def test_row_tile_layers_on_the_other_kernel(dev, env):
    """conv2d_3 / conv2d_4 / conv2d_14 + head shapes run on conv_rows.cuh's row tiles by default. The switches are read once
    per process, so the same parity cases run in a fresh process on the other kernel: conv_igemm_kernel for all three
    (ITMO_ROWS=0), and for the pooled layer alone (ITMO_ROWS_POOL=0)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, '-m', 'pytest', os.path.join(root, 'tests', 'test_gpu_parity.py'), '-m', 'gpu', '-q', '-x',
                        '-p', 'no:cacheprovider', '-k',
                        '(every_shape and (256x32+0to64 or 256x64+0to64p or 512x32+64to64)) or fused_head'],
                       cwd=root, env=e, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert ' passed' in r.stdout and 'failed' not in r.stdout, r.stdout[-2000:]
