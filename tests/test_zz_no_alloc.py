"""GPU test of itmo_reinhard_target_static (``-m gpu``; collected LAST on purpose).

The entry point has no counterpart in the reference: it is
itmo_reinhard_target for callers that may not allocate — its scratch is a file-scope ``static __device__`` array of the
library. Called through the C ABI (ctypes, plain pointers) and compared with the existing entry point on two calls in a
row that share the array (different inputs, the second SMALLER than the first so stale contents would show), plus the
size limit and the claim that the call allocates no device memory.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip('torch')


@pytest.fixture(scope='module')
def lib():
    assert torch.cuda.is_available(), 'the gpu tests need a CUDA device'
    from cnn_itmo_b200 import _lib
    return _lib


def _field(n, h, w, seed):
    rng = np.random.default_rng(seed)
    hdr = np.exp(rng.normal(0.0, 1.6, (n, h, w, 3))).astype(np.float32)   # log-normal radiance, like synth_hdr_field
    hdr[0, 3, 5] = 0.0                                                      # a black pixel: 0/0 -> NaN -> 0
    return torch.from_numpy(hdr).cuda()


def _call(L, name, hdr, out, key=0.18):
    n, h, w, _ = hdr.shape
    dt = L.DT_F32 if out.dtype == torch.float32 else L.DT_U8
    fn = getattr(L.load(), name)
    return fn(hdr.data_ptr(), n, h, w, ctypes.c_double(key), out.data_ptr(), dt, torch.cuda.current_stream().cuda_stream)


def test_static_scratch_entry_point_equals_the_allocating_one_twice_in_a_row(lib):
    L = lib
    cases = [(_field(5, 72, 88, seed=1), torch.uint8, 0.18), (_field(3, 40, 56, seed=2), torch.float32, 0.25)]
    # the existing entry point first (its outputs are the expectation), on buffers allocated before anything is measured
    want, got = [], []
    for hdr, dtype, key in cases:
        o = torch.empty(hdr.shape, dtype=dtype, device='cuda')
        L.check(_call(L, 'itmo_reinhard_target', hdr, o, key))
        want.append(o)
        got.append(torch.full(hdr.shape, 77, dtype=dtype, device='cuda'))
    # set-up call (resolves the symbol's device address once per device), then the two calls that are checked
    warm = torch.empty(cases[0][0].shape, dtype=torch.uint8, device='cuda')
    L.check(_call(L, 'itmo_reinhard_target_static', cases[0][0], warm))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    torch_alloc0 = torch.cuda.memory_allocated()
    for (hdr, _, key), o in zip(cases, got):                    # back to back on one stream, sharing the static array
        L.check(_call(L, 'itmo_reinhard_target_static', hdr, o, key))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, 'the call changed the free device memory'
    assert torch.cuda.memory_allocated() == torch_alloc0
    for i, (w_, g_) in enumerate(zip(want, got)):
        # tests/test_gpu_datagen.py allows one 8-bit step against the numpy restatement; the two entry points run the
        # same kernels in the same order, so here the tolerance is zero
        assert torch.equal(w_, g_), (i, (w_.float() - g_.float()).abs().max().item())
    assert float(got[0].float().mean()) > 1.0 and bool((got[0][0, 3, 5] == 0).all())   # real output, NaN -> 0 kept


def test_static_scratch_entry_point_rejects_more_images_than_its_array_holds(lib):
    L = lib
    nmax = L.REINHARD_STATIC_MAX_IMAGES
    hdr = torch.ones((nmax + 1, 2, 2, 3), dtype=torch.float32, device='cuda')
    out = torch.full(hdr.shape, 9, dtype=torch.uint8, device='cuda')
    rc = _call(L, 'itmo_reinhard_target_static', hdr, out)
    assert rc != 0
    msg = L.load().itmo_last_error().decode()
    assert str(nmax) in msg and 'maximum' in msg, msg
    torch.cuda.synchronize()
    assert bool((out == 9).all())                               # nothing was launched
    # the maximum itself is accepted and equals the allocating entry point
    hdr = _field(nmax, 4, 6, seed=3)
    a = torch.empty(hdr.shape, dtype=torch.uint8, device='cuda')
    b = torch.empty_like(a)
    L.check(_call(L, 'itmo_reinhard_target', hdr, a))
    L.check(_call(L, 'itmo_reinhard_target_static', hdr, b))
    assert torch.equal(a, b)
    with pytest.raises(L.ItmoError):
        from cnn_itmo_b200 import datagen
        datagen.reinhard_target_static_into(torch.ones((nmax + 1, 2, 2, 3), device='cuda'),
                                            torch.empty((nmax + 1, 2, 2, 3), device='cuda'))
