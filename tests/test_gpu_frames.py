"""GPU parity of the frame path (``-m gpu``): tiled prediction of frames larger than 512x512 against the CPU
oracle's fully-convolutional whole-frame output, and the inverse-Reinhard stage against its numpy restatement."""
import numpy as np
import pytest

from oracle import unet_ref as R

pytestmark = pytest.mark.gpu
torch = pytest.importorskip('torch')


@pytest.fixture(scope='module')
def model(calibrated_weights):
    assert torch.cuda.is_available(), 'the gpu tests need a CUDA device'
    import cnn_itmo_b200 as M
    m = M.Model(calibrated_weights, precision='fp32', max_batch=4)
    yield m
    m.close()


@pytest.fixture(scope='module')
def frame():
    return np.ascontiguousarray(R.synthetic_images_u8(1, seed=7, size=640)[0][:528, :640])


def test_tiled_frame_matches_whole_frame_oracle(model, calibrated_weights, frame):
    from cnn_itmo_b200 import frames
    ref = R.forward_torch(R.TorchWeights(calibrated_weights), R.preprocess_u8(frame)[None])[0]
    out = frames.predict_frames(model, frame, out='f32')[0]
    assert out.shape == ref.shape
    assert np.abs(out - ref).max() <= 1e-3          # fp32 mode bar of north_star, across the tile seams too
    u8 = frames.predict_frames(model, frame, out='u8')[0]
    want = R.postprocess_u8(ref)
    diff = np.abs(u8.astype(np.int32) - want.astype(np.int32))
    assert diff.max() <= 1 and (diff != 0).mean() < 0.02   # truncation (predict.py:64) flips at most the last bit


def test_tiles_overlap_consistently_and_batching_is_invisible(calibrated_weights, frame):
    """Two frames in one call == each alone; a context whose arena holds 1 tile == one that holds 4."""
    import cnn_itmo_b200 as M
    from cnn_itmo_b200 import frames
    two = np.stack([frame, frame[::-1, ::-1].copy()])
    a = M.Model(calibrated_weights, precision='bf16', max_batch=1)
    b = M.Model(calibrated_weights, precision='bf16', max_batch=4)
    ya = frames.predict_frames(a, two, out='u8')
    yb = frames.predict_frames(b, two, out='u8')
    assert (ya == yb).all()
    assert (frames.predict_frames(b, two[1], out='u8')[0] == yb[1]).all()
    a.close()
    b.close()


@pytest.mark.parametrize('as_written', [False, True])
def test_inverse_reinhard_matches_restatement(as_written):
    from cnn_itmo_b200 import frames
    img = R.synthetic_images_u8(2, seed=9, size=128)
    img[1, :4, :4] = 0                                  # zero-luminance pixels: PB1 > 0 (inverse_Reinhard.m:13)
    got = frames.inverse_reinhard(img, as_written=as_written, key=0.18)
    for i in range(2):
        want = R.inverse_reinhard_np(img[i], as_written=as_written, key=0.18).astype(np.float32)
        nan_w, nan_g = np.isnan(want), np.isnan(got[i])
        assert (nan_w == nan_g).all()
        assert np.allclose(got[i][~nan_g], want[~nan_w], rtol=2e-5, atol=1e-30)


def test_inverse_reinhard_saturated_and_black_pixels():
    """The repaired inverse must survive what inverse tone mapping is for: saturated highlights (255,255,255) and black
    pixels in the same frame (one saturated pixel used to zero the whole frame, 1 % black pixels used to
    push it into the 2^32 clip)."""
    from cnn_itmo_b200 import frames
    img = R.synthetic_images_u8(1, seed=21, size=128)
    img[0, :16, :16] = 255
    img[0, 100:113, 7:107] = 0                      # ~8 % black pixels
    got = frames.inverse_reinhard(img, as_written=False, key=0.18)[0]
    want = R.inverse_reinhard_np(img[0], as_written=False, key=0.18).astype(np.float32)
    assert np.isfinite(got).all() and got.max() < 2.0 ** 31
    assert np.allclose(got, want, rtol=2e-5, atol=1e-30)
    assert (got[100:113, 7:107] == 0).all() and got[:16, :16].min() > 1.0     # black stays black, highlights expand
    lum = got.astype(np.float64) @ np.array([0.2126, 0.7152, 0.0722])
    nz = lum > 0
    assert abs(np.exp(np.log(lum[nz]).mean()) - 0.18) < 1e-3                  # key of the non-black pixels
    black = np.zeros((1, 64, 64, 3), np.uint8)
    assert (frames.inverse_reinhard(black, as_written=False)[0] == 0).all()   # an all-black frame stays black


def test_fused_hdr_output_equals_u8_then_inverse(model, frame):
    """out='hdr' (statistics fused into the tile scatter) == inverse Reinhard applied to the uint8 frame."""
    from cnn_itmo_b200 import frames
    u8 = frames.predict_frames(model, frame, out='u8')
    hdr = frames.predict_frames(model, frame, out='hdr', key=0.18)[0]
    want = R.inverse_reinhard_np(u8[0], key=0.18).astype(np.float32)
    assert np.isfinite(hdr).all()
    assert np.allclose(hdr, want, rtol=2e-5, atol=1e-30)
    bad = frames.predict_frames(model, frame, out='hdr_as_written')[0]
    wantb = R.inverse_reinhard_np(u8[0], as_written=True).astype(np.float32)
    assert (np.isnan(bad) == np.isnan(wantb)).all()
    assert np.allclose(bad[~np.isnan(bad)], wantb[~np.isnan(wantb)], rtol=2e-5, atol=1e-30)


def test_half_float_hdr_output_is_the_float_output_rounded(model, frame):
    """out='hdr_f16' (ITMO_OUT_HDR_INTENDED_F16): the float32 result of out='hdr', clamped to 65504 and rounded to half."""
    from cnn_itmo_b200 import frames
    hdr = frames.predict_frames(model, frame, out='hdr', key=0.18)
    half = frames.predict_frames(model, frame, out='hdr_f16', key=0.18)
    assert half.dtype == np.float16 and half.shape == hdr.shape
    assert np.array_equal(half, np.minimum(hdr, 65504.0).astype(np.float16))


def test_4k_frame_properties(calibrated_weights):
    """BASELINE config 4 at full size: one 3840x2160 frame = 84 tiles. Properties instead of an oracle run:
    finite, equal to the 512x512 model on a tile that touches no seam, and flip-consistent tile plan."""
    import cnn_itmo_b200 as M
    from cnn_itmo_b200 import frames
    rng = np.random.default_rng(11)
    small = R.synthetic_images_u8(1, seed=12, size=512)[0]
    frame4k = np.tile(small, (5, 8, 1))[:2160, :3840].copy()
    frame4k[700:1212, 900:1412] = rng.integers(0, 256, (512, 512, 3), dtype=np.uint8)
    m = M.Model(calibrated_weights, precision='bf16', max_batch=32)
    out = frames.predict_frames(m, frame4k, out='u8')[0]
    assert out.shape == (2160, 3840, 3)
    # the top-left tile of the plan is the image's own 512x512 window: its valid region must equal predict_u8
    alone = m.predict_u8(frame4k[None, :512, :512])[0]
    assert (out[:416, :416] == alone[:416, :416]).all()
    assert m.launch_count() > 0
    m.close()


def test_hdr_output_is_bit_reproducible(model, frame):
    """The per-frame log-luminance statistic is accumulated with INTEGER atomics (fixed point), so the fused HDR output does
    not depend on the order in which blocks finish: repeated runs return identical bits."""
    from cnn_itmo_b200 import frames
    a = frames.predict_frames(model, frame, out='hdr', key=0.18)
    for _ in range(3):
        assert np.array_equal(frames.predict_frames(model, frame, out='hdr', key=0.18), a)
    u8 = frames.predict_frames(model, frame, out='u8')
    b = frames.inverse_reinhard(u8, as_written=False)
    for _ in range(3):
        assert np.array_equal(frames.inverse_reinhard(u8, as_written=False), b)


def test_streamed_frame_batches_equal_one_call_per_batch(model):
    """frames.predict_frame_batches (upload / kernels / download on three streams, double-buffered) yields, batch by batch,
    exactly what predict_frames returns for that batch — five different batches through the two buffers, pinned and
    ordinary host arrays, and a shape change is refused."""
    from cnn_itmo_b200 import frames
    from cnn_itmo_b200.model import pinned_empty
    batches = []
    for i in range(5):
        b = pinned_empty((2, 528, 640, 3), np.uint8) if i % 2 == 0 else np.empty((2, 528, 640, 3), np.uint8)
        b[...] = R.synthetic_images_u8(2, seed=30 + i, size=640)[:, :528, :640]
        batches.append(b)
    for kind in ('hdr', 'u8'):
        got = [np.array(r, copy=True) for r in frames.predict_frame_batches(model, iter(batches), out=kind)]
        assert len(got) == len(batches)
        for b, g in zip(batches, got):
            want = frames.predict_frames(model, b, out=kind)
            assert g.dtype == want.dtype and np.array_equal(g, want)
    assert list(frames.predict_frame_batches(model, iter([]), out='hdr')) == []
    with pytest.raises(ValueError):
        list(frames.predict_frame_batches(model, iter([batches[0], batches[0][:1]]), out='u8'))
