"""cip_w_histogram_static -- the |w| histogram whose scratch (the per-bin counters) is a static device array, for
callers that cannot allocate device memory -- against cip_w_histogram on the same device inputs.  Counts are integers:
the two must agree exactly.  Two calls in a row with different inputs and bin counts check that the shared array is
cleared between calls.  Collected last on purpose (file name), so that a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

from conftest import random_problem

pytestmark = pytest.mark.gpu


def _histograms(wg, cabi, prob, nbins, w_slab=None):
    import torch
    uvw, freq, vis, wgt, nx, px = prob
    dev = lambda a: torch.as_tensor(np.ascontiguousarray(a)).cuda()
    d = [dev(uvw), dev(freq), dev(vis), dev(wgt)]
    torch.cuda.synchronize()
    args = wg.make_args(d[0], d[1], d[2], d[3], None, nx, nx, px, px, 1e-4, True, False, 1, w_slab=w_slab)
    aw = np.abs(uvw[:, 2:3]) * (freq[None, :] / 299792458.0)
    wmin, wmax = float(aw.min()), float(aw.max())
    ctx = cabi.Context.get(-1)
    out = []
    for fn in (cabi.lib().cip_w_histogram, cabi.lib().cip_w_histogram_static):
        counts = np.zeros(nbins, dtype=np.int64)
        with ctx.lock:
            cabi.check(fn(ctx.handle, C.byref(args), wmin, wmax, nbins, counts.ctypes.data))
        out.append(counts)
    del d
    return out[0], out[1], int(np.count_nonzero(wgt))


def test_static_histogram_matches_on_two_calls_in_a_row(built_lib):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from ska_sdp_cip import _cabi, wgridder
    # first call: many bins; second call: other samples and FEWER bins -- stale counts of the first call would show
    big = random_problem(nrow=3000, nchan=8, nx=128, wscale=300.0, seed=7)
    ref, got, nactive = _histograms(wgridder, _cabi, big, 4096)
    assert ref.sum() == nactive > 0
    assert np.array_equal(got, ref)
    small = random_problem(nrow=500, nchan=4, nx=64, wscale=50.0, seed=8)
    ref2, got2, nactive2 = _histograms(wgridder, _cabi, small, 37)
    assert ref2.sum() == nactive2 > 0
    assert np.array_equal(got2, ref2)
    # the same inputs again, restricted to a |w| slab
    uvw, freq = small[0], small[1]
    aw = np.abs(uvw[:, 2:3]) * (freq[None, :] / 299792458.0)
    ref3, got3, _ = _histograms(wgridder, _cabi, small, 37, w_slab=(float(np.quantile(aw, 0.25)), float(np.quantile(aw, 0.75))))
    assert 0 < ref3.sum() < ref2.sum()
    assert np.array_equal(got3, ref3)


def test_static_histogram_rejects_what_it_cannot_hold(built_lib):
    import torch
    from ska_sdp_cip import _cabi, wgridder
    uvw, freq, vis, wgt, nx, px = random_problem(nrow=50, nchan=2, nx=32, seed=9)
    dev = lambda a: torch.as_tensor(np.ascontiguousarray(a)).cuda()
    d = [dev(uvw), dev(freq), dev(vis), dev(wgt)]
    ctx = _cabi.Context.get(-1)
    counts = np.zeros(8192, dtype=np.int64)
    args = wgridder.make_args(d[0], d[1], d[2], d[3], None, nx, nx, px, px, 1e-4, True, False, 1)
    with pytest.raises(ValueError, match="exceeds the 4096 bins"):
        _cabi.check(_cabi.lib().cip_w_histogram_static(ctx.handle, C.byref(args), 0.0, 1.0, 4097, counts.ctypes.data))
    host = wgridder.make_args(uvw, freq, vis, wgt, None, nx, nx, px, px, 1e-4, True, False, 0)
    with pytest.raises(ValueError, match="device inputs only"):
        _cabi.check(_cabi.lib().cip_w_histogram_static(ctx.handle, C.byref(host), 0.0, 1.0, 16, counts.ctypes.data))
    assert not counts.any()
