"""GPU parity of the data-synthesis / augmentation kernels (``-m gpu``): itmo_synth_hdr_field, itmo_virtual_camera,
itmo_reinhard_target, itmo_augment through the C ABI against oracle/datagen_ref.py (the numpy restatement of
m-files/virtual_camera.m, m-files/Reinhard.m and main.py:71-76, the augmentation pinned against scipy on the CPU)."""
import numpy as np
import pytest

from oracle import datagen_ref as D

pytestmark = pytest.mark.gpu
torch = pytest.importorskip('torch')


@pytest.fixture(scope='module')
def dg():
    assert torch.cuda.is_available(), 'the gpu tests need a CUDA device'
    from cnn_itmo_b200 import datagen
    return datagen


def test_hdr_field_matches_restatement(dg):
    got = dg.synth_hdr_field_device(3, 40, 56, seed=11).cpu().numpy()
    want = D.synth_hdr_field(3, 40, 56, seed=11)
    assert got.shape == want.shape and got.dtype == np.float32
    assert np.allclose(got, want, rtol=2e-6, atol=0)          # double math on both sides; last-ulp freedom of log / cos / exp


def test_virtual_camera_and_reinhard_match_restatement(dg):
    n, h, w = 3, 64, 48
    hdr = D.synth_hdr_field(n, h, w, seed=5)
    hdr[1, 7, 9] = 0.0                                         # a black pixel: 0/0 -> NaN -> 0 like imwrite
    cams = [dict(v=-3.0, n=0.6, y=0.9), dict(v=0.5, n=0.31, y=1.3), dict(v=4.0, n=1.1, y=0.55)]
    hd = torch.from_numpy(hdr).cuda()
    sdr = dg.virtual_camera_device(hd, cams).cpu().numpy()
    tgt = dg.reinhard_target_device(hd, as_float=False).cpu().numpy()
    tgt_f = dg.reinhard_target_device(hd, as_float=True).cpu().numpy()
    for i in range(n):
        want = D.virtual_camera(hdr[i], **cams[i])
        d = np.abs(sdr[i].astype(np.int32) - want.astype(np.int32))
        assert d.max() <= 1 and (d != 0).mean() <= 1e-4, (i, d.max(), (d != 0).mean())   # pow()'s last ulp at a .5 boundary
        wt = D.reinhard_target(hdr[i])
        d = np.abs(tgt[i].astype(np.int32) - wt.astype(np.int32))
        assert d.max() <= 1 and (d != 0).mean() <= 1e-4
    np.testing.assert_array_equal(tgt_f, tgt.astype(np.float32) / np.float32(255.0))   # the float output IS the PNG / 255
    assert (sdr[1, 7, 9] == 0).all() and (tgt[1, 7, 9] == 0).all()
    assert sdr[2].mean() > sdr[0].mean()                        # +4 stops is brighter than -3 stops


CASES = [dict(theta=0.0, zx=1.0, zy=1.0, flip_h=False, flip_v=False),
         dict(theta=0.0, zx=1.0, zy=1.0, flip_h=True, flip_v=False),
         dict(theta=0.0, zx=1.0, zy=1.0, flip_h=True, flip_v=True),
         dict(theta=90.0, zx=1.0, zy=1.0, flip_h=False, flip_v=False),
         dict(theta=-28.0, zx=0.85, zy=1.15, flip_h=False, flip_v=True),
         dict(theta=63.0, zx=1.2, zy=0.8, flip_h=True, flip_v=False)]


def test_augment_matches_restatement(dg):
    rng = np.random.default_rng(2)
    x = rng.integers(0, 256, (len(CASES), 96, 80, 3), dtype=np.uint8)
    y = rng.random((len(CASES), 96, 80, 3), dtype=np.float32)
    xa, ya = dg.augment_pair_device(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), CASES)
    xa, ya = xa.cpu().numpy(), ya.cpu().numpy()
    for i, case in enumerate(CASES):
        wx = D.augment_np(x[i], **case)                        # rescale 1/255 for the uint8 input
        wy = D.augment_np(y[i], scale=1.0, **case)
        if case['theta'] == 0 and case['zx'] == 1 and case['zy'] == 1:
            np.testing.assert_array_equal(xa[i], wx)           # identity / flips: exact
            np.testing.assert_array_equal(ya[i], wy)
        else:
            assert np.abs(xa[i] - wx).max() <= 1e-6 and np.abs(ya[i] - wy).max() <= 1e-6
    # and against the real scipy call for one rotated + zoomed case
    ws = D.augment_scipy(x[4], **CASES[4])
    assert np.abs(xa[4] - ws).max() <= 1e-6


def test_pairs_stay_aligned_under_augmentation(dg):
    """main.py:78-100 gives the input and the target generator the same seed so that both get the same transform: a
    pair (x, y = x / 255) must still satisfy y == x after augmentation."""
    x = torch.randint(0, 256, (4, 64, 64, 3), dtype=torch.uint8, device='cuda')
    y = x.float() / 255.0
    params = dg.sample_transform_params(4, seed=1)
    xa, ya = dg.augment_pair_device(x, y, params)
    assert float((xa - ya).abs().max()) <= 1e-6


def test_synthetic_pairs_feed_the_trainer(dg):
    x, y = dg.synthetic_pairs_device(2, 256, 128, seed=3)
    assert x.dtype == torch.uint8 and y.dtype == torch.float32 and x.shape == y.shape == (2, 256, 128, 3)
    assert 0.0 <= float(y.min()) and float(y.max()) <= 1.0 and 5 < float(x.float().mean()) < 250
    from cnn_itmo_b200 import train
    import cnn_itmo_b200 as M
    tr = train.Trainer(M.initial_weights(3), micro_batch=2, image_hw=(256, 128), device=0)
    tr.zero_grad()
    tr.backward(x, y, float(y.numel()), seed=0)
    sq, _ = tr.metrics()
    assert np.isfinite(sq) and sq > 0
    xa, ya = dg.synthetic_pairs_device(2, 256, 128, seed=3, augment=True)
    tr.backward(xa, ya, float(ya.numel()), seed=1)              # augmented pairs are float32: the F32 input path
    assert np.isfinite(tr.metrics()[0])
    tr.close()
