"""CPU tests of the frame path's host logic: the tile plan (library vs independent restatement), the oracle's
tiled == whole-frame property, the inverse-Reinhard restatement, and the sharding / gather of independent units over
ranks (world_size 2, gloo)."""
import os
import socket
import sys

import numpy as np
import pytest

import cnn_itmo_b200 as M
from cnn_itmo_b200 import frames
from oracle import unet_ref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize('extent', [512, 528, 640, 768, 1024, 1088, 2160, 3840, 4096])
def test_tile_plan_library_matches_restatement_and_covers(extent):
    plan = frames.plan_axis(extent)
    assert plan == R.plan_axis_ref(extent)
    covered = 0
    for (o, lo, hi) in plan:
        assert o % 16 == 0 and 0 <= o and o + frames.TILE <= extent          # inside the frame, on the pool grid
        assert lo == covered and hi > lo                                      # valid ranges partition the axis
        assert lo == 0 or lo - o >= frames.HALO                               # halo on every interior cut
        assert hi == extent or o + frames.TILE - hi >= frames.HALO
        assert (o == 0) == (lo == 0) and ((o + frames.TILE == extent) == (hi == extent))
        covered = hi
    assert covered == extent


def test_tile_plan_4k_is_84_tiles_and_bad_extents_raise():
    tiles = frames.plan_tiles(2160, 3840)
    assert len(tiles) == 7 * 12 == 84
    assert sum((y1 - y0) * (x1 - x0) for (_, _, y0, y1, x0, x1) in tiles) == 2160 * 3840
    for bad in (0, 256, 500, 520):
        with pytest.raises(ValueError):
            frames.plan_axis(bad)
    with pytest.raises(ValueError):
        frames._check_frames(np.zeros((1, 600, 512, 3), np.uint8))
    with pytest.raises(ValueError):
        frames._check_frames(np.zeros((1, 512, 512, 3), np.float32))


def test_oracle_tiled_equals_whole_frame(calibrated_weights):
    """The property the 4K path rests on: with tiles inside the frame, on the 16-pixel grid and a 96-pixel halo the
    stitched output IS the fully-convolutional output (fp32 rounding only)."""
    tw = R.TorchWeights(calibrated_weights)
    frame = R.preprocess_u8(R.synthetic_images_u8(1, seed=3, size=640)[0][:528, :640])
    whole = R.forward_torch(tw, frame[None])[0]
    tiled = R.predict_tiled_torch(tw, frame)
    assert np.abs(whole - tiled).max() <= 2e-6
    # and a halo that is too small does break it (the check is not vacuous): shift a tile's valid edge into the halo
    t = R.forward_torch(tw, frame[None, 0:512, 0:512])[0]
    assert np.abs(t[:, 500:512] - whole[0:512, 500:512]).max() > 1e-4


def test_inverse_reinhard_restatement():
    img = R.synthetic_images_u8(1, seed=4, size=64)[0]
    hdr = R.inverse_reinhard_np(img, key=0.18)
    assert np.isfinite(hdr).all() and hdr.min() >= 0
    lum = 0.2126 * hdr[..., 0] + 0.7152 * hdr[..., 1] + 0.0722 * hdr[..., 2]
    I = (img.astype(np.float64) / 255) ** 2.2 @ np.array([0.2126, 0.7152, 0.0722])
    ok = (I > 0) & (I < 1)
    # recovered luminance has geometric mean `key` (Reinhard.m:19-21 normalisation) when nothing is clamped
    if ok.all():
        assert abs(np.exp(np.log(lum).mean()) - 0.18) < 1e-9
    # forward Reinhard (Reinhard.m:22 L = X/(1+X)) of the unscaled inverse gives the luminance back
    X = I[ok] / (1 - I[ok])
    assert np.allclose(X / (1 + X), I[ok])
    # as written: X = I/(I-1) <= 0 everywhere, so G_X = realmin and the result collapses to ~0 / NaN
    bad = R.inverse_reinhard_np(img, as_written=True)
    assert np.nanmax(np.abs(bad)) < 1e-300


def test_inverse_reinhard_restatement_saturated_and_black():
    """Repaired variant on a frame with saturated and black pixels: finite, black stays black, key preserved."""
    img = R.synthetic_images_u8(1, seed=21, size=64)[0]
    img[:8, :8] = 255
    img[40:44, :] = 0
    hdr = R.inverse_reinhard_np(img, key=0.18)
    assert np.isfinite(hdr).all() and hdr.max() < 2.0 ** 31
    assert (hdr[40:44] == 0).all() and hdr[:8, :8].min() > 1.0
    lum = hdr @ np.array([0.2126, 0.7152, 0.0722])
    assert abs(np.exp(np.log(lum[lum > 0]).mean()) - 0.18) < 1e-9
    assert (R.inverse_reinhard_np(np.zeros((16, 16, 3), np.uint8)) == 0).all()


def test_shard_functions():
    for n in (0, 1, 7, 16, 84):
        for world in (1, 2, 3, 8):
            parts = [frames.shard_slice(n, r, world) for r in range(world)]
            assert [i for p in parts for i in range(n)[p]] == list(range(n))
            sizes = [len(range(n)[p]) for p in parts]
            assert max(sizes) - min(sizes) <= 1
            rr = [frames.shard_round_robin(n, r, world) for r in range(world)]
            assert sorted(i for p in rr for i in p) == list(range(n))
    with pytest.raises(ValueError):
        frames.shard_slice(4, 2, 2)
    out = frames.gather_units(np.arange(6).reshape(3, 2), [2, 0, 1], 3)        # single process
    assert out.tolist() == [[2, 3], [4, 5], [0, 1]]


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, n_units, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import cnn_itmo_b200  # noqa: F401
    from cnn_itmo_b200 import frames as F
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        units = np.arange(n_units * 12, dtype=np.uint8).reshape(n_units, 2, 2, 3)
        mine = F.shard_round_robin(n_units, rank, world)
        local = (units[mine].astype(np.int32) * 2 + 1).astype(np.int32)      # stand-in for the per-rank forward
        out = F.gather_units(local, mine, n_units)
        if rank == 0:
            q.put(bool((out == units.astype(np.int32) * 2 + 1).all()))
        else:
            assert out is None
        sl = F.shard_slice(n_units, rank, world)
        out2 = F.gather_units(local[:0] if False else (units[sl].astype(np.int32)), list(range(n_units))[sl], n_units)
        if rank == 0:
            q.put(bool((out2 == units).all()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n_units', [5, 1])
def test_gather_over_two_gloo_ranks(n_units):
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, n_units, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True and q.get(timeout=5) is True
