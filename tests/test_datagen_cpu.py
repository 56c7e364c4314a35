"""CPU tests of the data-synthesis / augmentation restatement (oracle/datagen_ref.py) and of the host logic of
cnn-itmo_b200/datagen.py. The augmentation is pinned against the real third-party call Keras makes
(scipy.ndimage.affine_transform, order=1, mode='nearest'); the MATLAB tone curves are unpinned (no MATLAB, no EXR
inputs, no golden PNGs in the reference) and are checked through the identities the scripts imply."""
import numpy as np
import pytest

from oracle import datagen_ref as D

CASES = [dict(theta=0.0, zx=1.0, zy=1.0, flip_h=False, flip_v=False),
         dict(theta=0.0, zx=1.0, zy=1.0, flip_h=True, flip_v=True),
         dict(theta=37.5, zx=1.0, zy=1.0, flip_h=False, flip_v=False),
         dict(theta=-90.0, zx=0.8, zy=1.2, flip_h=True, flip_v=False),
         dict(theta=90.0, zx=1.0, zy=1.0, flip_h=False, flip_v=False),
         dict(theta=12.0, zx=1.17, zy=0.83, flip_h=False, flip_v=True)]


@pytest.mark.parametrize('case', CASES)
def test_augment_restatement_matches_scipy(case):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (48, 40, 3), dtype=np.uint8)
    want = D.augment_scipy(img, **case)
    got = D.augment_np(img, **case)
    assert got.dtype == np.float32 and got.shape == img.shape
    assert np.abs(got - want).max() <= 1e-6
    if case['theta'] == 0 and case['zx'] == 1 and case['zy'] == 1:
        ref = img.astype(np.float32)
        if case['flip_h']:
            ref = ref[:, ::-1]
        if case['flip_v']:
            ref = ref[::-1]
        np.testing.assert_array_equal(got, ref * np.float32(1 / 255.0))     # flips are exact


def test_product_affine_matrix_is_keras_matrix():
    from cnn_itmo_b200 import datagen
    for theta, zx, zy in ((0.0, 1.0, 1.0), (33.0, 1.0, 1.0), (0.0, 0.9, 1.1), (-71.0, 1.2, 0.8)):
        M, off = datagen.affine_matrix(theta, zx, zy, 512, 384)
        Mr, offr = D.keras_affine(theta, zx, zy, 512, 384)
        np.testing.assert_allclose(M, Mr, rtol=0, atol=1e-12)
        np.testing.assert_allclose(off, offr, rtol=0, atol=1e-9)
    # the centre keras_preprocessing rotates about is (h/2 + 0.5, w/2 + 0.5): a fixed point of every transform
    M, off = datagen.affine_matrix(45.0, 1.1, 0.9, 64, 64)
    c = np.array([32.5, 32.5])
    np.testing.assert_allclose(M @ c + off, c, atol=1e-9)


def test_transform_sampler_ranges_and_pairing():
    from cnn_itmo_b200 import datagen
    a = datagen.sample_transform_params(200, seed=1)
    b = datagen.sample_transform_params(200, seed=1)
    assert a == b                                                  # same seed -> the same transforms for inputs and targets
    th = np.array([p['theta'] for p in a])
    z = np.array([[p['zx'], p['zy']] for p in a])
    assert th.min() >= -90 and th.max() <= 90 and th.std() > 30
    assert z.min() >= 0.8 and z.max() <= 1.2
    assert 0.3 < np.mean([p['flip_h'] for p in a]) < 0.7 and 0.3 < np.mean([p['flip_v'] for p in a]) < 0.7
    cams = datagen.sample_camera_params(500, seed=2)
    v = np.array([c['v'] for c in cams])
    assert v.min() >= -4 and v.max() <= 4
    assert min(c['n'] for c in cams) > 0.05 and min(c['y'] for c in cams) > 0.05
    assert abs(np.mean([c['n'] for c in cams]) - 0.6) < 0.08 and abs(np.mean([c['y'] for c in cams]) - 0.9) < 0.08


def test_hdr_field_generator_known_answers():
    f = D.synth_hdr_field(2, 16, 24, seed=7)
    assert f.shape == (2, 16, 24, 3) and f.dtype == np.float32 and np.isfinite(f).all() and f.min() > 0
    g = D.synth_hdr_field(2, 16, 24, seed=7)
    np.testing.assert_array_equal(f, g)
    assert not np.array_equal(f, D.synth_hdr_field(2, 16, 24, seed=8))
    big = D.synth_hdr_field(1, 256, 256, seed=1).astype(np.float64)
    lg = np.log(big)
    assert abs(lg.mean()) < 0.1 and 1.0 < lg.std() < 1.4            # exp(1.2 z1 + 0.15 z2): sigma = 1.209
    # the 8-pixel block structure: neighbours inside a block are far closer than across the field
    assert np.abs(lg[0, 0, 0, 0] - lg[0, 1, 1, 0]) < 1.0


def test_tone_curves_identities():
    hdr = D.synth_hdr_field(1, 64, 64, seed=3)[0]
    Y = D.rgb2lum(hdr).astype(np.float64)
    G = D.geomean_lum(D.rgb2lum(hdr))
    assert abs(np.exp(np.log(Y).mean()) - G) < 1e-9 * G
    tgt = D.reinhard_target(hdr)
    assert tgt.dtype == np.uint8 and tgt.shape == hdr.shape
    # Reinhard.m:20-25: the target's luminance is L = X/(1+X), X = 0.18/G * Y  (up to the 8-bit rounding)
    L = (0.18 / G * Y) / (1 + 0.18 / G * Y)
    lum_t = tgt.astype(np.float64) / 255.0 @ np.array([0.2126, 0.7152, 0.0722])
    sat = (tgt == 255).any(-1)
    assert np.abs(lum_t - L)[~sat].max() < 2.5 / 255
    # virtual camera: more exposure never darkens; clipping at 1
    a = D.virtual_camera(hdr, v=-2.0, n=0.6, y=0.9).astype(np.int32)
    b = D.virtual_camera(hdr, v=+2.0, n=0.6, y=0.9).astype(np.int32)
    assert (b >= a).all() and b.max() == 255 and a.mean() < b.mean()
    # a black pixel: 0/0 -> NaN -> imwrite writes 0
    hdr0 = hdr.copy()
    hdr0[5, 5] = 0
    assert (D.virtual_camera(hdr0, 0.0, 0.6, 0.9)[5, 5] == 0).all() and (D.reinhard_target(hdr0)[5, 5] == 0).all()
