"""GPU parity of the raw WISDM text ingest against the strtod/strtoll oracle: columns bit for bit,
the same records dropped, literals outside the exact fast path refused loudly; and raw text -> windows end to end."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

SAMPLE = (b"33,Jogging,49105962326000,-0.6946377,12.680544,0.50395286;\n"
          b"33,Jogging,49106062271000,5.012288,11.264028,0.95342433;\n"
          b" 17,Walking,1,1e-3,2.5E2,-.5;11,Walking,1867172313000,4.4,4.4,;\nbad;;  \n"
          b"5,Flying,1,1,1,1;8,Standing, 12 ,0, -0.0 ,7.;9,Upstairs,3,1.2.3,4,5;10,Downstairs,3,--1,4,5;7,Sitting,9,1,2,3")


def _cmp(text):
    from har_b200 import core
    (g, gd), (o, od) = core.parse_wisdm_raw(text), orc.parse_wisdm_raw(text)
    assert gd == od
    for a, b in zip(g, o):
        a = a.cpu().numpy()
        assert a.shape == b.shape and np.array_equal(a.view(np.uint8), np.ascontiguousarray(b).view(np.uint8))
    return g, gd


def test_known_records_and_malformed_ones():
    g, dropped = _cmp(SAMPLE)
    assert g[0].cpu().tolist() == [33, 33, 17, 8, 7] and g[1].cpu().tolist() == [1, 1, 0, 5, 4] and dropped == 5
    assert g[3].cpu().tolist() == [-0.6946377, 5.012288, 0.001, 0.0, 1.0]
    assert np.signbit(g[4].cpu().numpy()[3])                       # "-0.0" keeps its sign
    for empty in (b"", b"  \n;;\n ; "):
        g, dropped = _cmp(empty)
        assert g[0].numel() == 0 and dropped == 0


def test_generated_stream_round_trip_is_bit_exact():
    rng = np.random.RandomState(5)
    n = 200_000
    u, a, ts, x, y, z = orc.synth_stream(n, 0, 20_000, 2018)
    names = orc.WISDM_ACTIVITIES
    recs = []
    for i in range(n):
        k = i & 3
        if k == 0:
            vals = [str(np.float32(v)) for v in (x[i], y[i], z[i])]          # float-precision decimals, as in the raw release
        elif k == 1:
            vals = ["%.6f" % v for v in (x[i], y[i], z[i])]
        elif k == 2:
            vals = ["%.8e" % v for v in (x[i], y[i], z[i])]
        else:
            vals = ["%d" % round(v) for v in (x[i], y[i], z[i])]
        sep = b";\n" if rng.rand() < 0.7 else b";"
        recs.append(("%d,%s,%d,%s,%s,%s" % (u[i], names[a[i]], ts[i], *vals)).encode() + sep)
        if rng.rand() < 0.01:
            recs.append(b"%d,%s,%d,1.0,2.0;\n" % (u[i], names[a[i]].encode(), ts[i]))     # five fields: dropped
    g, dropped = _cmp(b"".join(recs))
    assert g[0].numel() == n and dropped > 0
    assert np.array_equal(g[0].cpu().numpy(), u) and np.array_equal(g[2].cpu().numpy(), ts)


def test_long_literals_are_refused_not_approximated():
    from har_b200 import core
    from har_b200._native import HarError
    with pytest.raises(HarError, match="more than 15 significant digits"):
        core.parse_wisdm_raw(b"1,Walking,5,0.12345678901234567890123,1,2;")
    with pytest.raises(HarError, match="exact fast path"):
        core.parse_wisdm_raw(b"1,Walking,5,1e-30,1,2;")


def test_raw_text_to_window_features():
    from har_b200.window import WisdmWindowFeaturizer, read_wisdm_raw
    u, a, ts, x, y, z = orc.synth_stream(30_000, 0, 10_000, 2018)
    names = orc.WISDM_ACTIVITIES
    sx, sy, sz = ([str(np.float32(v)) for v in col] for col in (x, y, z))
    text = "".join("%d,%s,%d,%s,%s,%s;\n" % (u[i], names[a[i]], ts[i], sx[i], sy[i], sz[i]) for i in range(len(u))).encode()
    df = read_wisdm_raw(text)
    assert df.dropped_records == 0 and df.count() == 30_000
    feats = WisdmWindowFeaturizer(outputCol="features").transform(df)
    xs, ys, zs = (np.array([float(t) for t in col]) for col in (sx, sy, sz))
    ws = orc.window_segments(u, a, ts, 200)
    want = np.nan_to_num(orc.window_features(xs, ys, zs, ts, ws, 200), nan=0.0)      # unknown peaks -> unknownPeakValue
    got = feats._cols["features"].dense
    assert got.shape == want.shape and np.allclose(got, want, rtol=1e-6, atol=1e-9, equal_nan=True)
