"""har_window_segments_static — segmentation for callers that may not allocate CUDA memory.

Who needs it: a process that has given all device memory to its own allocator, a loop with a hard
latency budget, or a plug-in host that only allows pre-approved buffers cannot call the regular
har_window_segments, because that one takes its per-tile status words from cudaMallocAsync on every call.
The static variant keeps the scratch in a 4 MiB array inside the library.  This test drives the new
entry point directly through the C ABI (ctypes), twice in a row on different inputs so the second
call runs over whatever the first one left in the shared array, and requires bit-identical window
lists to the regular entry point (integer output, so the tolerance is zero).  It is collected last
(file name sorts after every other test file) and works both alone in a fresh process and after the
rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(fn, user, act, ts, window_len, gap_ns=0):
    import torch
    n = user.numel()
    cap = max(n // window_len, 1)
    ws = torch.full((cap,), -7, dtype=torch.int64, device="cuda")
    wu = torch.full((cap,), -7, dtype=torch.int32, device="cuda")
    wa = torch.full((cap,), -7, dtype=torch.int32, device="cuda")
    count = C.c_int64(-1)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = fn(C.c_void_p(user.data_ptr()), C.c_void_p(act.data_ptr()), None if ts is None else C.c_void_p(ts.data_ptr()),
            n, window_len, gap_ns, C.c_void_p(ws.data_ptr()), C.c_void_p(wu.data_ptr()), C.c_void_p(wa.data_ptr()),
            C.byref(count), stream)
    torch.cuda.synchronize()
    return rc, count.value, ws.cpu().numpy(), wu.cpu().numpy(), wa.cpu().numpy()


def test_static_segments_match_regular_entry_point_twice_in_a_row():
    import torch
    from har_b200 import _native, synth
    lib = _native.lib()
    assert torch.cuda.is_available()
    # first call: a long stream (many tiles) ; second call: a shorter, different stream with another
    # window length and timestamp-gap breaks, reusing the scratch the first call dirtied
    cases = [(3_000_000, 0, 75_000, 200, 0), (333_333, 5_000_000, 41_000, 64, 500_000_000)]
    for n, first, per_user, L, gap in cases:
        user, act, ts, x, y, z = synth.synth_stream(n, first_sample=first, per_user=per_user, seed=99)
        if gap:
            ts = ts.clone()
            ts[n // 3:] += 10 ** 9           # one pause inside a bout
        rc_a, cnt_a, ws_a, wu_a, wa_a = _call(lib.har_window_segments, user, act, ts, L, gap)
        rc_b, cnt_b, ws_b, wu_b, wa_b = _call(lib.har_window_segments_static, user, act, ts, L, gap)
        assert rc_a == 0 and rc_b == 0, lib.har_last_error()
        assert cnt_b == cnt_a and cnt_a > 100
        assert np.array_equal(ws_b, ws_a) and np.array_equal(wu_b, wu_a) and np.array_equal(wa_b, wa_a)
        assert np.all(ws_b[cnt_b:] == -7)                      # nothing written past the reported count


def test_static_segments_rejects_inputs_beyond_the_built_in_scratch():
    import torch
    from har_b200 import _native
    lib = _native.lib()
    n = 170_000_000                                            # limit at window_len 200 is 165,169,152
    user = torch.zeros(n, dtype=torch.int32, device="cuda")
    act = torch.zeros(n, dtype=torch.int32, device="cuda")
    rc, cnt, *_ = _call(lib.har_window_segments_static, user, act, None, 200)
    assert rc == _native_code("HAR_E_UNSUPPORTED") and cnt == 0
    msg = lib.har_last_error().decode()
    assert "165169152" in msg and "har_window_segments" in msg
    # just under the limit still works and agrees with the regular entry point
    m = 165_169_152
    rc2, cnt2, ws2, _, _ = _call(lib.har_window_segments_static, user[:m], act[:m], None, 200)
    rc3, cnt3, ws3, _, _ = _call(lib.har_window_segments, user[:m], act[:m], None, 200)
    assert rc2 == 0 and rc3 == 0 and cnt2 == cnt3 == m // 200 and np.array_equal(ws2, ws3)


def _native_code(name):
    return {"HAR_E_INVALID": -1, "HAR_E_CUDA": -2, "HAR_E_NOMEM": -3, "HAR_E_UNSUPPORTED": -4}[name]
