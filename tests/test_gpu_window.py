"""GPU parity: WISDM window transform (har_window_segments / har_window_features, har_synth_stream)
through the C ABI against the CPU oracle.  Integer outputs and discrete decisions (window starts, bin
fractions, peak selection) are bit-exact; fp64 moments within 1e-6 relative (north_star tolerance;
observed ~1e-15: only the summation order differs)."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def _gpu_stream(n, first=0, per_user=20000, seed=2018):
    from har_b200 import synth
    return synth.synth_stream(n, first_sample=first, per_user=per_user, seed=seed)


def test_synth_stream_bit_identical_to_oracle():
    n, first = 70001, 12345
    g = [t.cpu().numpy() for t in _gpu_stream(n, first)]
    o = orc.synth_stream(n, first_sample=first, per_user=20000, seed=2018)
    for a, b, name in zip(g, o, ("user", "activity", "ts", "x", "y", "z")):
        assert a.dtype == b.dtype and np.array_equal(a.view(np.uint8), b.view(np.uint8)), name


@pytest.mark.parametrize("n,per_user,L", [(60000, 20000, 200), (2048 * 3, 1000, 200), (5000, 700, 64), (199, 199, 200),
                                          (200, 200, 200), (100003, 33333, 200)])
def test_segments_match_oracle(n, per_user, L):
    from har_b200 import core
    u, a, ts, x, y, z = _gpu_stream(n, 0, per_user)
    ws, wu, wa = core.window_segments(u, a, ts, window_len=L)
    uh, ah, th = u.cpu().numpy(), a.cpu().numpy(), ts.cpu().numpy()
    want = orc.window_segments(uh, ah, th, L=L)
    assert np.array_equal(ws.cpu().numpy(), want)
    assert np.array_equal(wu.cpu().numpy(), uh[want]) and np.array_equal(wa.cpu().numpy(), ah[want])


def test_segments_gap_breaks_and_unaligned_and_empty():
    import torch
    from har_b200 import core
    u, a, ts, x, y, z = _gpu_stream(30000, 0, 10000)
    th = ts.cpu().numpy().copy()
    th[7000:] += 10 ** 10          # a pause inside a bout
    th[15011:] += 10 ** 10
    ws, _, _ = core.window_segments(u, a, torch.from_numpy(th).cuda(), gap_ns=500_000_000)
    want = orc.window_segments(u.cpu().numpy(), a.cpu().numpy(), th, gap_ns=500_000_000)
    assert np.array_equal(ws.cpu().numpy(), want)
    assert len(want) != len(orc.window_segments(u.cpu().numpy(), a.cpu().numpy(), th))
    # pointers that are not 16-byte aligned (a shard cut at an odd sample)
    ws2, _, _ = core.window_segments(u[3:], a[3:], ts[3:])
    assert np.array_equal(ws2.cpu().numpy(), orc.window_segments(u.cpu().numpy()[3:], a.cpu().numpy()[3:], None))
    e = torch.empty(0, dtype=torch.int32, device="cuda")
    ws0, _, _ = core.window_segments(e, e, None)
    assert ws0.numel() == 0


@pytest.mark.parametrize("n,first,L", [(60000, 0, 200), (50001, 777, 200), (20000, 3, 50), (9000, 1, 256)])
def test_window_features_match_oracle(n, first, L):
    from har_b200 import core
    u, a, ts, x, y, z = _gpu_stream(n, first)
    ws, _, _ = core.window_segments(u, a, ts, window_len=L)
    got = core.window_features(x, y, z, ts, ws, window_len=L).cpu().numpy()
    want = orc.window_features(x.cpu().numpy(), y.cpu().numpy(), z.cpu().numpy(), ts.cpu().numpy(), ws.cpu().numpy(), L=L)
    assert got.shape == want.shape and got.shape[0] > 10
    assert np.array_equal(got[:, :30], want[:, :30]), "10-bin distributions must be identical"
    assert np.array_equal(np.isnan(got[:, 33:36]), np.isnan(want[:, 33:36]))
    assert np.array_equal(np.nan_to_num(got[:, 33:36], nan=-1.0), np.nan_to_num(want[:, 33:36], nan=-1.0)), "peak intervals"
    for cols in (slice(30, 33), slice(36, 43)):
        np.testing.assert_allclose(got[:, cols], want[:, cols], rtol=RTOL, atol=0)


def test_window_features_unknown_peak_and_odd_offsets():
    """Flat / monotone windows have < 3 local maxima -> NaN ('?' in the reference CSV); odd window starts
    exercise the 16-byte re-alignment of the bulk copies and the tail-window path."""
    import torch
    from har_b200 import core
    L, nw = 200, 9
    n = L * nw + 1
    rng = np.random.RandomState(3)
    x = rng.randn(n); y = np.linspace(0, 1, n); z = np.full(n, 2.5)
    ts = np.arange(n, dtype=np.int64) * 50_000_000
    ws = np.arange(nw, dtype=np.int64) * L + 1          # odd starts; last window ends at n
    got = core.window_features(x, y, z, ts, ws).cpu().numpy()
    want = orc.window_features(x, y, z, ts, ws)
    assert np.all(np.isnan(got[:, 34])) and np.all(np.isnan(got[:, 35])) and not np.any(np.isnan(got[:, 33]))
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[:, :30], want[:, :30])
    np.testing.assert_allclose(np.nan_to_num(got), np.nan_to_num(want), rtol=RTOL, atol=1e-300)
    assert np.all(got[:, 20] == 1.0)                    # constant axis: everything in bin 0
    # unaligned base pointers (views offset by one double) still work
    xo = torch.from_numpy(np.concatenate([[0.0], x])).cuda()[1:]
    got2 = core.window_features(xo, y, z, ts, ws).cpu().numpy()
    assert np.array_equal(np.nan_to_num(got2), np.nan_to_num(got))


def test_window_pipeline_properties_at_scale():
    """Size-independent properties on 20M samples (oracle too slow to be the checker here)."""
    import torch
    from har_b200 import core
    n = 20_000_000
    u, a, ts, x, y, z = _gpu_stream(n, 0, 100000)
    ws, wu, wa = core.window_segments(u, a, ts)
    f = core.window_features(x, y, z, ts, ws)
    assert ws.numel() > 0.95 * n / 200
    assert bool(torch.all(ws[1:] - ws[:-1] >= 200))                                   # non-overlapping, sorted
    assert bool(torch.all(u[ws] == u[ws + 199])) and bool(torch.all(a[ws] == a[ws + 199]))
    assert bool(torch.all((f[:, 0:10].sum(1) - 1.0).abs() < 1e-12))
    assert bool(torch.all(f[:, 36:39] <= f[:, 39:42] + 1e-12))
    # a sampled slice against the oracle
    k0 = ws.numel() // 2
    sl = slice(k0, k0 + 50)
    want = orc.window_features(x.cpu().numpy(), y.cpu().numpy(), z.cpu().numpy(), ts.cpu().numpy(), ws[sl].cpu().numpy())
    np.testing.assert_allclose(np.nan_to_num(f[sl].cpu().numpy()), np.nan_to_num(want), rtol=RTOL, atol=0)


@pytest.mark.parametrize("n,L", [(2048 * 100 + 77, 200), (2048 * 70, 64), (300_001, 2), (2048 * 40 + 5, 5000), (2047, 200), (700_001, 70_000)])
def test_segments_long_bouts_span_many_tiles(n, L):
    """One bout running over a hundred tiles (no boundary for the look-back to find for dozens of tiles), then a few
    boundaries at awkward places: the windows of a bout open at a tile's start are placed in closed form."""
    import torch
    from har_b200 import core
    u = np.zeros(n, np.int32); a = np.zeros(n, np.int32)
    for cut in (n - 4097, n - 2049, n - 2048, n - 1):
        if cut > 0:
            a[cut:] += 1
    ws, wu, wa = core.window_segments(torch.from_numpy(u).cuda(), torch.from_numpy(a).cuda(), None, window_len=L)
    want = orc.window_segments(u, a, None, L=L)
    assert np.array_equal(ws.cpu().numpy(), want) and len(want) >= n // L - 4
    assert np.array_equal(wa.cpu().numpy(), a[want])


def test_window_transform_equals_the_two_calls():
    """har_window_transform (one call, window count taken from device memory) == segments then features."""
    from har_b200 import core
    for n, L in ((150_003, 200), (40_000, 50)):
        u, a, ts, x, y, z = _gpu_stream(n, 5, 20000)
        ws, wu, wa = core.window_segments(u, a, ts, window_len=L)
        f = core.window_features(x, y, z, ts, ws, window_len=L)
        f2, wu2, wa2, ws2 = core.window_transform(u, a, ts, x, y, z, window_len=L)
        assert np.array_equal(ws.cpu().numpy(), ws2.cpu().numpy()) and np.array_equal(wu.cpu().numpy(), wu2.cpu().numpy())
        assert np.array_equal(wa.cpu().numpy(), wa2.cpu().numpy())
        assert np.array_equal(np.nan_to_num(f.cpu().numpy(), nan=-1.0).view(np.int64), np.nan_to_num(f2.cpu().numpy(), nan=-1.0).view(np.int64))


def test_window_features_quantised_samples_with_ties():
    """Readings with one decimal (the real WISDM files have two): plateaus, repeated maxima, windows whose largest
    value sits on the first or last sample (then the highest PEAK is not the window maximum), equal peaks on the
    threshold ladder.  Bins and peak intervals stay bit-exact."""
    from har_b200 import core
    n = 80_000
    u, a, ts, x, y, z = _gpu_stream(n, 11, 20000)
    xq = (x * 2).round() / 2
    yq = y.round()
    zq = (z * 10).round() / 10
    ws, _, _ = core.window_segments(u, a, ts)
    wsh = ws.cpu().numpy()
    xh = xq.cpu().numpy().copy()
    for k, w in enumerate(wsh[:40]):                 # plant the maximum on a window edge / a plateau at the start
        if k % 4 == 0: xh[w] = 1e3
        elif k % 4 == 1: xh[w + 199] = 1e3
        elif k % 4 == 2: xh[w] = xh[w + 1] = 1e3
        else: xh[w + 198] = xh[w + 199] = 1e3
    import torch
    xq = torch.from_numpy(xh).cuda()
    got = core.window_features(xq, yq, zq, ts, ws).cpu().numpy()
    want = orc.window_features(xh, yq.cpu().numpy(), zq.cpu().numpy(), ts.cpu().numpy(), wsh)
    assert np.array_equal(got[:, :30], want[:, :30])
    assert np.array_equal(np.nan_to_num(got[:, 33:36], nan=-1.0), np.nan_to_num(want[:, 33:36], nan=-1.0))
    for cols in (slice(30, 33), slice(36, 43)):
        np.testing.assert_allclose(got[:, cols], want[:, cols], rtol=RTOL, atol=0)
