"""Whole-graph GPU parity (``-m gpu``): Model.predict / predict_u8 / itmo_forward against the CPU oracle, the
committed fixtures, and size-independent properties at BASELINE.json's batch sizes."""
import ctypes
import os

import numpy as np
import pytest

from oracle import unet_ref as R

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')


@pytest.fixture(scope='module')
def pkg():
    assert torch.cuda.is_available(), 'the gpu tests need a CUDA device'
    import cnn_itmo_b200 as M
    return M


@pytest.fixture(scope='module')
def images():
    return R.synthetic_images_u8(3, seed=5)


@pytest.fixture(scope='module')
def oracle_out(calibrated_weights, images):
    outs, pooled = R.forward_torch(R.TorchWeights(calibrated_weights), R.preprocess_u8(images), return_all=True)
    return outs, pooled


def test_predict_fp32_matches_oracle(pkg, calibrated_weights, images, oracle_out):
    """north_star: max abs error <= 1e-3 on [0,1] outputs in the fp32-accurate mode (3xTF32 products, fp32
    accumulation in TMEM); every intermediate tensor too. The checkpoint is the calibrated synthetic one: full
    [0,1] output range, random he-normal kernels, BN with negative gammas — a stress case that amplifies operand
    rounding ~20x from input to output."""
    m = pkg.Model(calibrated_weights, precision='fp32', max_batch=3)
    out = m.predict(R.preprocess_u8(images).astype(np.float64))   # the reference feeds float64 (predict.py:59)
    ref = oracle_out[0][18]
    assert out.dtype == np.float32 and out.shape == (3, 512, 512, 3)
    assert np.abs(out - ref).max() <= 1e-3, float(np.abs(out - ref).max())
    for li in range(17):
        act = m.read_activation(li)
        r = oracle_out[0][li]
        assert np.abs(act - r).max() <= 1e-3 * max(1.0, float(np.abs(r).max())), li
        if R.LAYERS[li].pool:
            pa = m.read_activation(li, pooled=True)
            assert np.abs(pa - oracle_out[1][li]).max() <= 1e-3 * max(1.0, float(np.abs(r).max())), li
    m.close()


@pytest.mark.parametrize('prec', ['tf32', 'fp16'])
def test_predict_11bit_modes_psnr(pkg, calibrated_weights, images, oracle_out, prec):
    """north_star: PSNR >= 45 dB against the reference in the reduced-precision tensor-core mode. On the stress
    checkpoint the 11-bit-significand operand formats (TF32 with round-to-nearest storage, fp16) reach ~50 dB."""
    m = pkg.Model(calibrated_weights, precision=prec, max_batch=2)   # 3 images -> two arena chunks
    out = m.predict(R.preprocess_u8(images))
    ref = oracle_out[0][18]
    assert R.psnr(out, ref) >= 45.0, R.psnr(out, ref)
    for i in range(3):
        assert R.psnr(out[i], ref[i]) >= 45.0
    m.close()


def test_default_precision_is_fp16_and_meets_the_bar(pkg, calibrated_weights, images, oracle_out):
    """The mode a user gets without asking (Model(), load_model(), the predict CLI, bench.py) is fp16 and holds the
    north-star's PSNR >= 45 dB on the calibrated checkpoint, with no saturation."""
    import warnings
    m = pkg.Model(calibrated_weights, max_batch=3)
    assert m.precision == 'fp16'
    with warnings.catch_warnings():
        warnings.simplefilter('error', RuntimeWarning)   # neither the load-time range check nor the device counter may fire
        out = m.predict(R.preprocess_u8(images))
    assert R.psnr(out, oracle_out[0][18]) >= 45.0
    assert m.saturation_count() == 0
    m.close()


def test_fp16_saturation_is_counted_not_silent(pkg, calibrated_weights, images):
    """A checkpoint whose activations leave fp16's range: the load-time check flags it and the device counter sees the
    clamped stores (the output is still finite: the pack saturates instead of producing inf)."""
    import warnings
    ws = [dict(w) for w in calibrated_weights]
    ws[3] = dict(ws[3], gamma=(ws[3]['gamma'] * 3.0e5).astype(np.float32))     # conv2d_4: |y| ~ 1e5 > 65504
    m = pkg.Model(ws, precision='fp16', max_batch=1)
    with warnings.catch_warnings(record=True) as rec:
        warnings.simplefilter('always')
        out = m.predict(R.preprocess_u8(images[:1]))
    msgs = [str(r.message) for r in rec]
    assert any('conv2d_4' in t and '65504' in t for t in msgs), msgs       # static range report
    assert any('saturated' in t for t in msgs), msgs                         # device counter
    assert np.isfinite(out).all()
    m.close()
    # the same weights in an fp32-range mode do not saturate
    m = pkg.Model(ws, precision='tf32', max_batch=1)
    m.predict(R.preprocess_u8(images[:1]))
    assert m.saturation_count() == 0
    m.close()


def test_predict_bf16_psnr_on_the_network_as_built(pkg, images):
    """bf16 (8-bit significand, opt-in): PSNR >= 45 dB on the network exactly as the reference builds it (U_net():
    he_normal kernels, identity BatchNormalization; measured ~76 dB)."""
    init = pkg.initial_weights(seed=218)
    x = R.preprocess_u8(images[:2])
    ref_init = R.forward_torch(R.TorchWeights(init), x)
    m = pkg.Model(init, precision='bf16', max_batch=2)
    out = m.predict(x)
    assert R.psnr(out, ref_init) >= 45.0, R.psnr(out, ref_init)
    m.close()


@pytest.mark.xfail(reason='FAILED BAR, kept visible: bf16 storage reaches ~32 dB on the calibrated stress checkpoint '
                          '(operand rounding alone, emulated on the CPU, costs 33 dB). bf16 is opt-in; the default is '
                          'fp16, which meets 45 dB (DESIGN.md 4.2, BASELINE.md)', strict=False)
def test_predict_bf16_psnr_on_calibrated_checkpoint(pkg, calibrated_weights, images, oracle_out):
    m = pkg.Model(calibrated_weights, precision='bf16', max_batch=2)
    out = m.predict(R.preprocess_u8(images))
    m.close()
    assert R.psnr(out, oracle_out[0][18]) >= 45.0, R.psnr(out, oracle_out[0][18])


@pytest.mark.parametrize('prec', ['fp32', 'tf32', 'fp16'])
def test_predict_u8_is_predict_py_flow(pkg, calibrated_weights, images, prec):
    """predict.py:59-64: u8/255 -> predict -> trunc(p*255). The fused u8-in/u8-out path against (a) the float path
    of the same precision (same kernels, only the load/store differ) and (b) the oracle's predict.py flow, at the
    north-star tolerances: +-1 LSB in the fp32 mode, PSNR >= 45 dB in the reduced-precision modes."""
    m = pkg.Model(calibrated_weights, precision=prec, max_batch=4)
    u8 = m.predict_u8(images)
    via_float = R.postprocess_u8(m.predict(R.preprocess_u8(images)))
    assert u8.dtype == np.uint8 and u8.shape == images.shape
    if prec in ('fp32', 'tf32'):
        np.testing.assert_array_equal(u8, via_float)
    else:  # fp16: raw pixels enter the first layer exactly, normalised floats are rounded to fp16 first
        assert R.psnr(u8 / 255.0, via_float / 255.0) >= 45.0
    ref = R.predict_like_reference(R.TorchWeights(calibrated_weights), images, one_at_a_time=False)
    d = np.abs(u8.astype(np.int32) - ref.astype(np.int32))
    if prec == 'fp32':
        assert d.max() <= 1 and (d != 0).mean() < 0.02
    else:
        assert R.psnr(u8 / 255.0, ref / 255.0) >= 45.0
    m.close()


def test_reference_fixture_images(pkg, calibrated_weights, golden_dir):
    """The reference's 12 committed inputs (images_to_predict/input) through the predict.py flow. Their committed
    outputs belong to the stripped checkpoint, so the comparison is against the oracle with the same synthetic
    weights; the shapes/dtypes of the committed outputs are checked as the format contract."""
    from PIL import Image
    from cnn_itmo_b200 import predict as P
    files = P.list_inputs(os.path.join(golden_dir, 'images_to_predict', 'input'))
    assert len(files) == 12
    imgs = np.stack([P.load_png(f) for f in files])
    m = pkg.Model(calibrated_weights, precision='fp32', max_batch=12)
    got = m.predict_u8(imgs)
    ref = R.predict_like_reference(R.TorchWeights(calibrated_weights), imgs, one_at_a_time=False)
    d = np.abs(got.astype(np.int32) - ref.astype(np.int32))
    assert d.max() <= 1 and (d != 0).mean() < 0.02
    with Image.open(os.path.join(golden_dir, 'images_to_predict', 'output', os.path.basename(files[0]))) as im:
        assert np.asarray(im).shape == got[0].shape and np.asarray(im).dtype == got.dtype
    m.close()


def test_checkpoint_file_to_prediction(pkg, calibrated_weights, images, oracle_out, tmp_path):
    """load_model(path).predict(X) (predict.py:24,62) from a Keras-layout HDF5 file written by our writer."""
    p = str(tmp_path / 'saved7-model-218-0.73.hdf5')
    pkg.write_checkpoint(p, calibrated_weights)
    m = pkg.load_model(p, precision='fp32', max_batch=1)
    out = m.predict(R.preprocess_u8(images[:1]))
    assert np.abs(out - oracle_out[0][18][:1]).max() <= 1e-3
    m.close()


@pytest.mark.parametrize('prec', ['bf16', 'fp16', 'tf32', 'fp32'])
def test_properties_at_full_batch(pkg, calibrated_weights, prec):
    """Size-independent properties at BASELINE batch sizes (the oracle cannot follow here): determinism, batch
    independence (image i of a batch of 64 == the same image alone), arena-chunking invariance, range."""
    n = 64
    base = R.synthetic_images_u8(4, seed=9)
    imgs = np.concatenate([base] * (n // 4))
    imgs[7] = 0
    imgs[8] = 255
    dev = torch.device('cuda:0')
    x = torch.from_numpy(imgs).to(dev)
    m = pkg.Model(calibrated_weights, precision=prec, max_batch=n)
    o1 = torch.empty_like(x)
    o2 = torch.empty_like(x)
    m.forward_device(x.data_ptr(), 'u8', o1.data_ptr(), 'u8', n)
    m.forward_device(x.data_ptr(), 'u8', o2.data_ptr(), 'u8', n)
    torch.cuda.synchronize()
    assert torch.equal(o1, o2)                                      # deterministic
    assert torch.equal(o1[0:3], o1[12:15]) and torch.equal(o1[1], o1[61])   # batch independence
    m2 = pkg.Model(calibrated_weights, precision=prec, max_batch=5)  # ragged chunks 5,5,...,4
    o3 = torch.empty_like(x)
    m2.forward_device(x.data_ptr(), 'u8', o3.data_ptr(), 'u8', n)
    torch.cuda.synchronize()
    assert torch.equal(o1, o3)                                      # chunking invariance
    single = m2.predict_u8(imgs[7:9])
    np.testing.assert_array_equal(single, o1[7:9].cpu().numpy())
    of = torch.empty((n, 512, 512, 3), dtype=torch.float32, device=dev)
    m.forward_device(x.data_ptr(), 'u8', of.data_ptr(), 'f32', n)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(of).all()) and float(of.min()) >= 0.0 and float(of.max()) <= 1.0
    assert torch.equal((of * 255.0).to(torch.uint8), o1)            # u8 store == trunc(f32 store * 255)
    assert m.launch_count() > 0 and m.launch_count() % 3 == 0       # three identical forward calls
    m.close()
    m2.close()


def test_empty_batch_and_errors(pkg, calibrated_weights):
    from cnn_itmo_b200 import _lib as L
    m = pkg.Model(calibrated_weights, precision='bf16', max_batch=1)
    assert m.predict(np.zeros((0, 512, 512, 3), np.float32)).shape == (0, 512, 512, 3)
    with pytest.raises(ValueError):
        m.predict(np.zeros((1, 256, 256, 3), np.float32))
    lib = L.load()
    ctx = ctypes.c_void_p()
    L.check(lib.itmo_create(ctypes.byref(ctx), 0, L.PREC_BF16, 1))
    x = torch.zeros((1, 512, 512, 3), dtype=torch.uint8, device='cuda:0')
    rc = lib.itmo_forward(ctx, x.data_ptr(), L.DT_U8, x.data_ptr(), L.DT_U8, 1, 0, None)
    assert rc == -3 and b'itmo_load_weights' in lib.itmo_last_error()      # ITMO_ERR_STATE, message set
    assert lib.itmo_create(ctypes.byref(ctypes.c_void_p()), 0, 7, 1) == -1  # bad precision
    lib.itmo_destroy(ctx)
    # the stateless operators validate shapes instead of launching garbage
    rc = lib.itmo_conv3x3_bn(L.PREC_BF16, x.data_ptr(), 24, None, 0, 1, 16, 16, x.data_ptr(), x.data_ptr(),
                             x.data_ptr(), x.data_ptr(), 32, x.data_ptr(), None, None)
    assert rc == -1
    m.close()


def test_cuda_graph_replay_equals_eager_launches(pkg, calibrated_weights, images):
    """itmo_set_graph_capture: predict_host through a replayed CUDA graph returns the bytes the 18 eager launches return —
    for every chunk size / dtype pair (each gets its own graph), across repeated calls, and after new weights are loaded
    (the graphs carry by-value affine vectors and must be dropped)."""
    imgs = np.concatenate([images, images[::-1]])[:5]
    eager = pkg.Model(calibrated_weights, max_batch=2)
    eager.use_cuda_graph = False
    graph_m = pkg.Model(calibrated_weights, max_batch=2)
    graph_m.use_cuda_graph = True
    for n in (1, 2, 5, 1):                       # 5 = chunks 2, 2, 1 through both staging slots
        a = eager.predict_u8(imgs[:n])
        b = graph_m.predict_u8(imgs[:n])
        np.testing.assert_array_equal(a, b)
        b2 = graph_m.predict_u8(imgs[:n])        # replay
        np.testing.assert_array_equal(a, b2)
    x = R.preprocess_u8(imgs[:3])
    np.testing.assert_array_equal(eager.predict(x), graph_m.predict(x))
    # new weights: stale graphs would keep producing the old image
    ws2 = [dict(w) for w in calibrated_weights]
    ws2[17] = dict(ws2[17], bias=ws2[17]['bias'] + np.float32(0.5))
    e2 = pkg.Model(ws2, max_batch=2)
    e2.use_cuda_graph = False
    graph_m._install_weights(ws2)
    want = e2.predict_u8(imgs[:2])
    got = graph_m.predict_u8(imgs[:2])
    np.testing.assert_array_equal(want, got)
    assert not np.array_equal(got, eager.predict_u8(imgs[:2]))
    for m in (eager, graph_m, e2):
        m.close()


def test_pipelined_predict_loop_equals_reference_loop(pkg, calibrated_weights, golden_dir, tmp_path):
    """predict.run_pipelined (thread-pool PNG codec + double-buffered pinned batches) writes the same pixels as the
    reference-shaped loop predict.run (predict.py:50-72), ragged last chunk included."""
    from cnn_itmo_b200 import predict as P
    files = P.list_inputs(os.path.join(golden_dir, 'images_to_predict', 'input'))
    m = pkg.Model(calibrated_weights, max_batch=8)
    wa = P.run(m, files, str(tmp_path / 'a'), batch=1, verbose=False)
    stats = {}
    wb = P.run_pipelined(m, files, str(tmp_path / 'b'), batch=5, workers=4, stats=stats)
    assert [os.path.basename(p) for p in wa] == [os.path.basename(p) for p in wb] and len(wb) == 12
    for pa, pb in zip(wa, wb):
        np.testing.assert_array_equal(P.load_png(pa), P.load_png(pb))
    assert stats['files'] == 12 and stats['gpu_seconds'] > 0
    m.close()


def test_fused_stem_kernel_is_bit_identical_to_the_two_layers(pkg, calibrated_weights, images, tmp_path):
    """conv2d_1 + conv2d_2 as one kernel (csrc/stem_fused.cuh, the default for uint8 input in the 16-bit modes) against the
    two layers launched separately (ITMO_STEM_FUSED=0, read once per process -> a second interpreter): same operand split,
    same affine sequence, same packing, so every output byte and every later activation is identical."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r)\n"
        "import cnn_itmo_b200 as M\n"
        "from oracle import unet_ref as R\n"
        "ws = R.calibrated_weights(218); imgs = R.synthetic_images_u8(3, seed=5)\n"
        "out = {}\n"
        "for prec in ('fp16', 'bf16'):\n"
        "    m = M.Model(ws, precision=prec, max_batch=2)\n"
        "    out[prec] = m.predict_u8(imgs); out[prec + '_act1'] = m.read_activation(1); out[prec + '_pool1'] = m.read_activation(1, pooled=True); m.close()\n"
        "np.savez(sys.argv[1], **out)\n" % root)
    outs = {}
    for flag in ('1', '0'):
        path = str(tmp_path / f'stem{flag}.npz')
        env = dict(os.environ, ITMO_STEM_FUSED=flag)
        r = subprocess.run([sys.executable, '-c', code, path], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs[flag] = np.load(path)
    for key in outs['1'].files:
        assert np.array_equal(outs['1'][key], outs['0'][key]), key
    assert outs['1']['fp16'].std() > 5       # not a constant image
