"""CPU checks of the tensor-core variant's host planner (tnk_tc_plan): the MMA K-steps must cover
every chunk product of the bracket exactly once with increasing chunk offsets, and the feature
layout it assumes (csrc/tnk_tc.cu header) must reproduce u = K_p q of the oracle when the steps,
the 3xTF32 split and the epilogue's Hamming weights are emulated in numpy."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import ksd_oracle as orc
from tensornetworks_b200 import _lib
from tensornetworks_b200.synthetic import random_binary_bn, random_q

def plan(n):
    lib = _lib.load()
    np_, k32, kb, ns, ns16 = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    steps = (ctypes.c_int32 * 128)()
    _lib.check(lib.tnk_tc_plan(n, ctypes.byref(np_), ctypes.byref(k32), ctypes.byref(kb), ctypes.byref(ns),
                               ctypes.byref(ns16), steps), "tnk_tc_plan")
    return np_.value, k32.value, kb.value, ns16.value, [tuple(steps[4 * k:4 * k + 4]) for k in range(ns.value)]


@pytest.mark.parametrize("n", range(7, 25))
def test_steps_cover_the_bracket(n):
    np_, k32, kb, ns16, steps = plan(n)
    nc = np_ // 4
    assert np_ == (n + 3) // 4 * 4 and k32 == 2 * nc + 2
    assert ns16 == (3 * np_ + 15) // 16 and kb == k32 + 2 * ns16
    # tf32 part: s_hi x t_hi, bits x r_hi (same chunk index on both sides), the constant chunk and the zero pad
    want = sorted([(c, c) for c in range(2 * nc + 2)])
    got = []
    for a0, alb, b0, blb in steps:
        assert alb > 0 and blb > 0 and a0 + alb < k32 and b0 + blb < k32
        assert alb * 128 < (1 << 14) and blb * 128 < (1 << 14)      # 14-bit LBO field in 16-byte units
        got += [(a0, b0), (a0 + alb, b0 + blb)]
    assert sorted(got) == want
    assert len(steps) == nc + 1 and len(steps) + ns16 <= 32
    assert 256 + 8 * (len(steps) + ns16) <= 512                    # TMEM columns: 2 accumulator stages + A


def test_unsupported_sizes_are_rejected():
    lib = _lib.load()
    for n in (0, 3, 6, 25, 30):
        assert lib.tnk_tc_plan(n, None, None, None, None, None, None) != 0


def _tf_hi(x):
    return (np.asarray(x, dtype=np.float32).view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


@pytest.mark.parametrize("n", [7, 9])
def test_emulated_tiles_match_oracle(n):
    """numpy emulation of tc_pack_kernel + the K-step list + the epilogue weights."""
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    P = orc.joint_table(bn, x, lat)
    S = orc.score_table(P, n)                      # [N, n], latent-variable order
    q64 = random_q(1 << n, seed=n).numpy()
    Kref, gref = orc.ksd_dense(q64, S, n)
    N = 1 << n
    np_, k32, kb, ns16, steps = plan(n)
    nc = np_ // 4
    a = np.exp(-1.0 / n)
    alpha, gamma = 1.0 - a, a - 1.0 / a
    z = np.arange(N)
    s = np.zeros((N, np_), dtype=np.float32)
    s[:, :n] = S[:, ::-1].astype(np.float32)       # bit position p = latent variable n-1-p
    bit = np.zeros((N, np_), dtype=np.float32)
    bit[:, :n] = ((z[:, None] >> np.arange(n)[None, :]) & 1)
    valid = np.zeros(np_, dtype=np.float32); valid[:n] = 1
    rho = (1 - s) * (1 - 2 * bit) * valid
    R = ((1 - s) * valid).sum(1)
    C = (alpha * R + gamma * ((1 - s) * bit).sum(1)).astype(np.float32)
    q = q64.astype(np.float32)
    hi, lo = _tf_hi, lambda v: np.asarray(v, dtype=np.float32) - _tf_hi(v)
    bf = lambda v: torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(torch.bfloat16).to(torch.float32).numpy()
    g = np.float32(gamma) * rho
    t = q[:, None] * s
    r = (np.float32(gamma) * q)[:, None] * rho
    zero = np.zeros(N, dtype=np.float32)
    zc = np.zeros((N, 4), dtype=np.float32)
    A32 = np.concatenate([hi(s), bit, np.stack([zero + 1, zero + 1, zero, zero], 1), zc], 1)
    B32 = np.concatenate([hi(t), hi(r), np.stack([hi(q * C), lo(q * C), zero, zero], 1), zc], 1)
    assert A32.shape[1] == 4 * k32 == B32.shape[1]
    pad16 = np.zeros((N, 16 * ns16 - 3 * np_), dtype=np.float32)
    A16 = np.concatenate([bf(s), bf(lo(s)), bit, pad16], 1)          # second-order cross terms in bf16
    B16 = np.concatenate([bf(lo(t)), bf(t), bf(lo(r)), pad16], 1)
    A64, B64 = A32.astype(np.float64), B32.astype(np.float64)
    D = A16.astype(np.float64) @ B16.astype(np.float64).T            # tensor-core part: q_j [s.s' + g T2 + C(j)]
    for a0, alb, b0, blb in steps:
        for ca, cb in ((a0, b0), (a0 + alb, b0 + blb)):
            D += A64[:, 4 * ca:4 * ca + 4] @ B64[:, 4 * cb:4 * cb + 4].T
    # ALU part per (row, tile J, group): w_group * (E*Q0 + Q1), E = C_i + G_hi(i,J) + G_mid(i,group)
    g64, q_64, C64 = g.astype(np.float64), q.astype(np.float64), C.astype(np.float64)
    h = np.array([[bin(i ^ j).count("1") for j in range(N)] for i in range(N)])
    wgt = np.exp(-h / n)
    u = (wgt * D).sum(1)
    cc = np.arange(16)
    for i in range(0, N, 37):                      # a sample of rows keeps the test fast
        W = np.exp(-np.array([bin((i ^ c) & 15).count("1") for c in cc]) / n)
        Glo = np.array([sum(g64[i, p] for p in range(4) if (c >> p) & 1) for c in cc])
        alu = 0.0
        for J in range(N // 128):
            Ehi = C64[i] + sum(g64[i, p] for p in range(7, n) if (J >> (p - 7)) & 1)
            for grp in range(8):
                qg = q_64[J * 128 + grp * 16:J * 128 + grp * 16 + 16]
                E = Ehi + sum(g64[i, 4 + b] for b in range(3) if (grp >> b) & 1)
                hh = bin((i >> 7) ^ J).count("1") + bin(((i >> 4) & 7) ^ grp).count("1")
                alu += np.exp(-hh / n) * (E * (W * qg).sum() + (W * Glo * qg).sum())
        u[i] += alu
    rows = np.arange(0, N, 37)
    uref = gref / 2
    assert np.max(np.abs(u[rows] - uref[rows])) <= 1e-5 * np.max(np.abs(uref))
