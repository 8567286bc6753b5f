"""Synthetic tri-axial accelerometer stream of SURVEY.md §8(d), generated in HBM by har_synth_stream.

Users 0..U-1 in order, `per_user` samples each; inside a user consecutive bouts of
600 + h % 5400 samples (h = splitmix64 of (seed, user, bout index)) with activity (h >> 32) % 6,
never equal to the previous bout's; samples are mu[c,a] + amp[c,a] * sine[(k*f[c] + 85a) & 1023] +
sigma[c,a] * noise with integer noise from splitmix64 of the global sample index.
"""
import ctypes as C

import numpy as np
import torch

from . import _native as N
from .core import hptr, ptr, require_cuda, stream

MU = np.array([[0.6, 9.3, 0.4], [0.5, 9.1, 0.7], [0.3, 9.2, 0.9], [0.5, 9.2, 0.5],
               [1.1, 8.2, 2.1], [0.4, 9.4, 0.6]], dtype=np.float64)
AMP = np.array([[3.0, 4.2, 2.6], [4.4, 5.6, 3.4], [3.1, 4.3, 3.0], [2.9, 4.4, 2.7],
                [0.6, 0.7, 0.5], [0.8, 0.9, 0.7]], dtype=np.float64)
SIGMA = np.array([[2.2, 2.6, 2.1], [2.6, 3.0, 2.3], [2.3, 2.7, 2.2], [2.4, 2.6, 2.2],
                  [1.1, 1.2, 1.0], [1.2, 1.3, 1.1]], dtype=np.float64)
FREQ = np.array([97, 113, 89, 93, 41, 47], dtype=np.int32)
SINE = np.sin(2.0 * np.pi * np.arange(1024) / 1024.0)

_M64 = (1 << 64) - 1


def _splitmix64(z):
    z = (z + 0x9E3779B97F4A7C15) & _M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def bout_table(first_sample, n, per_user, seed):
    """Bouts overlapping global samples [first_sample, first_sample + n)."""
    starts, users, acts, ts0, gains, tilts = [], [], [], [], [], []
    last_end = first_sample
    u0, u1 = first_sample // per_user, (first_sample + n - 1) // per_user
    for u in range(u0, u1 + 1):
        ubeg, uend = u * per_user, (u + 1) * per_user
        b0, bi, prev = ubeg, 0, -1
        while b0 < uend:
            h = _splitmix64(seed ^ ((u * 0xD1B54A32D192ED03) & _M64) ^ ((bi * 0x9E3779B97F4A7C15) & _M64))
            a = (h >> 32) % 6
            if a == prev:
                a = (a + 1) % 6
            b1 = min(b0 + 600 + h % 5400, uend)
            if b1 > first_sample and b0 < first_sample + n:
                starts.append(b0); users.append(u); acts.append(a)
                ts0.append((ubeg + b0) * 50000000 + bi * 1000000000)
                gains.append(0.35 + 1.3 * (float((h >> 8) & 0xFFFF) / 65535.0))
                tilts.append(float((h >> 24) & 0xFFFF) / 65535.0 - 0.5)
                last_end = b1
            prev, b0, bi = a, b1, bi + 1
    starts.append(last_end)
    return (np.array(starts, np.int64), np.array(users, np.int32), np.array(acts, np.int32), np.array(ts0, np.int64),
            np.array(gains, np.float64), np.array(tilts, np.float64))


def synth_stream(n, first_sample=0, per_user=100000, seed=2018):
    """-> (user i32, activity i32, ts i64, x, y, z f64) CUDA tensors of length n."""
    require_cuda()
    bs, bu, ba, bt, bg, bl = bout_table(first_sample, n, per_user, seed)
    user = torch.empty(n, dtype=torch.int32, device="cuda"); act = torch.empty_like(user)
    ts = torch.empty(n, dtype=torch.int64, device="cuda")
    x = torch.empty(n, dtype=torch.float64, device="cuda"); y = torch.empty_like(x); z = torch.empty_like(x)
    N.check(N.lib().har_synth_stream(n, first_sample, C.c_uint64(seed), hptr(bs), hptr(bu), hptr(ba), hptr(bt), hptr(bg), hptr(bl), len(bu),
                                     hptr(SINE), hptr(MU), hptr(AMP), hptr(SIGMA), hptr(FREQ),
                                     ptr(user), ptr(act), ptr(ts), ptr(x), ptr(y), ptr(z), stream()))
    return user, act, ts, x, y, z
