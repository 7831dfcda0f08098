"""The score kernel divides by p(x,z) through one correctly rounded reciprocal per state and two FMA residual
steps (csrc/tnk_enum.cu: div_by).  The reference divides with '/' (stein_utils.py:134); this checks on the CPU,
with libm's fma, that the two give identical bits - random operands in the probability range plus mantissas
next to the rounding boundaries (all ones, all zeros)."""
import ctypes
import ctypes.util

import numpy as np


def _fma():
    libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    libm.fma.restype = ctypes.c_double
    libm.fma.argtypes = [ctypes.c_double] * 3
    return libm.fma


def _div_by(fma, a, p):
    rp = 1.0 / p
    q = a * rp
    q = fma(fma(-q, p, a), rp, q)
    q = fma(fma(-q, p, a), rp, q)
    return q


def test_reciprocal_division_is_correctly_rounded():
    fma = _fma()
    rng = np.random.default_rng(0)
    n = 60000
    mant = rng.integers(0, 1 << 52, size=(n, 2), dtype=np.uint64)
    edge = np.concatenate([np.arange(16, dtype=np.uint64), (np.uint64((1 << 52) - 1) - np.arange(16, dtype=np.uint64))])
    mant[::5, 1] = rng.choice(edge, size=mant[::5, 1].shape)          # divisor mantissa at a boundary
    mant[1::7, 0] = rng.choice(edge, size=mant[1::7, 0].shape)        # dividend mantissa at a boundary
    ea = rng.integers(-200, 1, size=n)
    ep = rng.integers(-39, 1, size=n)                                 # |p| >= 1e-12 on the kernel's live branch
    a = np.ldexp(1.0 + mant[:, 0].astype(np.float64) / 2.0 ** 52, ea)
    p = np.ldexp(1.0 + mant[:, 1].astype(np.float64) / 2.0 ** 52, ep)
    bad = sum(1 for x, y in zip(a.tolist(), p.tolist()) if _div_by(fma, x, y) != x / y)
    assert bad == 0
    assert _div_by(fma, 0.0, 0.3) == 0.0 and _div_by(fma, 0.3, 0.3) == 1.0
