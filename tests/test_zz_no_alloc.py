"""fdsm_combine_score_static (include/fdsm.h): the score-form combine with its per-sample scratch in a file-scope
`static __device__` array, for callers that may not allocate.  No counterpart in the reference.
Called straight through the C ABI and checked against the existing entry point
fdsm_combine_score on the same inputs -- the two launch the same kernels, so loss and ds must agree bit for bit (the
bound written below is the suite's 1e-6 relative, and equality is asserted on top).  Collected last (file name), passes
alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL = 1e-6


def _ptr(t):
    return None if t is None else C.cast(C.c_void_p(t.data_ptr()), C.POINTER(C.c_float))


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _both(lib, s, v, sigma, labels, power, k, projected, with_ds):
    """(loss, ds) from the existing entry point and from the no-allocation one; every buffer is made by the caller."""
    B, D = v.shape
    lab = None if labels is None else C.c_void_p(labels.data_ptr())
    f = dict(device=s.device, dtype=torch.float32)
    loss0, per0 = torch.full((), -7.0, **f), torch.empty(B, **f)
    ds0 = torch.zeros(2 * B, D, **f) if with_ds else None
    rc = lib.fdsm_combine_score(_ptr(s), D, _ptr(v), D, _ptr(sigma), lab, power, B, D, k, projected, None, _ptr(per0), _ptr(loss0),
                                _ptr(ds0), D, _stream())
    assert rc == 0, lib.fdsm_last_error_string()
    loss1 = torch.full((), -9.0, **f)
    ds1 = torch.zeros(2 * B, D, **f) if with_ds else None
    rc = lib.fdsm_combine_score_static(_ptr(s), D, _ptr(v), D, _ptr(sigma), lab, power, B, D, k, projected, None, _ptr(loss1),
                                       _ptr(ds1), D, _stream())
    assert rc == 0, lib.fdsm_last_error_string()
    torch.cuda.synchronize()
    return (loss0, ds0), (loss1, ds1)


def _check(a, b):
    (l0, d0), (l1, d1) = a, b
    assert torch.isfinite(l0).item() and torch.isfinite(l1).item()
    assert abs(float(l1) - float(l0)) <= RTOL * abs(float(l0)), (float(l0), float(l1))
    assert torch.equal(l0, l1)
    if d0 is not None:
        assert float((d1 - d0).abs().max()) <= RTOL * float(d0.abs().max())
        assert torch.equal(d0, d1)


def test_static_scratch_matches_allocating_entry_point_two_calls_in_a_row(fdsm):
    lib = fdsm._lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator(device=dev).manual_seed(77)
    # first call: the ncsnv2 annealed form (sigma table + labels, power 2), config-5 batch shape, loss + ds
    B, D = 128, 3 * 32 * 32
    s = torch.randn(2 * B, D, device=dev, generator=g)
    v = torch.randn(B, D, device=dev, generator=g) * 0.01
    sig = torch.linspace(1.0, 0.01, 232, device=dev)
    lab = torch.randint(0, 232, (B,), device=dev, generator=g)
    _check(*_both(lib, s, v, sig, lab, 2.0, 1.0 / 0.01 ** 2 * 0.5, 0, True))
    # second call straight after, another shape (longer batch, ragged D -> scalar path, projected form, forward only):
    # the array still holds the first call's values, which must not leak into this result
    B2, D2 = 1000, 11
    s2 = torch.randn(2 * B2, D2, device=dev, generator=g)
    v2 = torch.randn(B2, D2, device=dev, generator=g) * 0.1
    _check(*_both(lib, s2, v2, None, None, 2.0, 0.5, 1, False))
    # and the first inputs again (shorter than what the array now holds)
    _check(*_both(lib, s, v, sig, lab, 2.0, 1.0 / 0.01 ** 2 * 0.5, 0, True))


def test_static_scratch_allocates_nothing_and_rejects_oversize(fdsm):
    lib = fdsm._lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    B, D = 64, 256
    s, v = torch.randn(2 * B, D, device=dev), torch.randn(B, D, device=dev)
    loss, ds = torch.empty((), device=dev), torch.empty(2 * B, D, device=dev)
    call = lambda b: lib.fdsm_combine_score_static(_ptr(s), D, _ptr(v), D, None, None, 2.0, b, D, 0.5, 0, None, _ptr(loss), _ptr(ds), D,
                                                   _stream())
    assert call(B) == 0          # includes the one-time address look-up on this device
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    n0 = torch.cuda.memory_stats()["allocation.all.allocated"]
    for _ in range(10):
        assert call(B) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0 and torch.cuda.memory_stats()["allocation.all.allocated"] == n0
    # under stream capture: a cudaMalloc would invalidate the capture
    gs = torch.cuda.Stream()
    gs.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(gs):
        with torch.cuda.graph(graph, stream=gs):
            assert call(B) == 0
    want = float(loss)
    loss.fill_(0.0)
    graph.replay()
    torch.cuda.synchronize()
    assert float(loss) == want
    # beyond the static array: a clear error, nothing launched
    limit = 65536
    assert call(limit + 1) == -1
    msg = lib.fdsm_last_error_string().decode()
    assert "FDSM_COMBINE_SCORE_STATIC_MAX_B" in msg and str(limit) in msg
    torch.cuda.synchronize()
