"""CPU check of the symmetric tile variant's operand layout (csrc/tnk_tc.cu, "K3-TCS"): a numpy emulation of
the q-free feature chunks (tf32 high parts + bf16 cross terms), of the tile schedule (row tile I meets column
tiles J >= I, mirrored half from the symmetry of k_p) and of the striped ownership of row tiles must reproduce
u = K_p q of the oracle; the workspace query must accept the sizes the kernel is instantiated for."""
import numpy as np
import pytest
import torch

from oracle import ksd_oracle as orc
from tensornetworks_b200 import _lib
from tensornetworks_b200.synthetic import random_binary_bn, random_q


def _tf_hi(x):
    return (np.asarray(x, dtype=np.float32).view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


def _bf(v):
    return torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def _features(S, n, length_scale=1.0):
    """Row and column feature matrices of the symmetric variant, tf32 part and bf16 part."""
    N = 1 << n
    np_ = (n + 3) // 4 * 4
    a = np.exp(-1.0 / (n * length_scale))
    alpha, gamma = 1.0 - a, a - 1.0 / a
    z = np.arange(N)
    s = np.zeros((N, np_), dtype=np.float32)
    s[:, :n] = S[:, ::-1].astype(np.float32)                  # bit position p = latent variable n-1-p
    bit = np.zeros((N, np_), dtype=np.float32)
    bit[:, :n] = (z[:, None] >> np.arange(n)[None, :]) & 1
    valid = np.zeros(np_, dtype=np.float32)
    valid[:n] = 1
    grho = (np.float32(gamma) * (1 - s)) * (1 - 2 * bit) * valid
    C = (alpha * ((1 - s) * valid).sum(1) + gamma * ((1 - s) * bit).sum(1)).astype(np.float32)
    hi, lo = _tf_hi, lambda v: np.asarray(v, dtype=np.float32) - _tf_hi(v)
    one = np.ones(N, dtype=np.float32)
    nc = np_ // 4
    k32 = (3 * nc + 2) & ~1
    pad = np.zeros((N, 4 * (k32 - 3 * nc - 1)), dtype=np.float32)
    A32 = np.concatenate([hi(s), hi(grho), bit, np.stack([hi(C), lo(C), one, one], 1), pad], 1)
    B32 = np.concatenate([hi(s), bit, hi(grho), np.stack([one, one, hi(C), lo(C)], 1), pad], 1)
    A16 = np.concatenate([_bf(s), _bf(lo(s)), _bf(lo(grho)), bit], 1)
    B16 = np.concatenate([_bf(lo(s)), _bf(s), bit, _bf(lo(grho))], 1)
    assert A32.shape[1] == 4 * k32 and A16.shape[1] == 16 * nc
    return A32, B32, A16, B16, a


def _features_bf16x2(S, n, length_scale=1.0):
    """Operand format 1: every real factor as two bf16 pieces; all products in bf16."""
    N = 1 << n
    np_ = (n + 3) // 4 * 4
    a = np.exp(-1.0 / (n * length_scale))
    alpha, gamma = 1.0 - a, a - 1.0 / a
    z = np.arange(N)
    s = np.zeros((N, np_), dtype=np.float32)
    s[:, :n] = S[:, ::-1].astype(np.float32)
    bit = np.zeros((N, np_), dtype=np.float32)
    bit[:, :n] = (z[:, None] >> np.arange(n)[None, :]) & 1
    valid = np.zeros(np_, dtype=np.float32)
    valid[:n] = 1
    g = (np.float32(gamma) * (1 - s)) * (1 - 2 * bit) * valid
    C = (alpha * ((1 - s) * valid).sum(1) + gamma * ((1 - s) * bit).sum(1)).astype(np.float32)
    s1 = _bf(s); s2 = _bf(s - s1)
    g1 = _bf(g); g2 = _bf(g - g1)
    c1 = _bf(C); c2 = _bf(C - c1); c3 = _bf(C - c1 - c2)
    one = np.ones(N, dtype=np.float32)
    A = np.concatenate([s1, s1, s2, g1, g2, bit, bit, np.stack([c1, c2, c3, one, one, one], 1)], 1)
    B = np.concatenate([s1, s2, s1, bit, bit, g1, g2, np.stack([one, one, one, c1, c2, c3], 1)], 1)
    assert A.shape[1] == 7 * np_ + 6
    return A, B, a


@pytest.mark.parametrize("n", [8, 9])
def test_emulated_bf16x2_operands_match_oracle(n):
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q64 = random_q(1 << n, seed=n).numpy()
    Kref, gref = orc.ksd_dense(q64, S, n)
    N = 1 << n
    A, B, a = _features_bf16x2(S, n)
    h = np.array([[bin(i ^ j).count("1") for j in range(N)] for i in range(N)])
    u = (a ** h * (A.astype(np.float64) @ B.astype(np.float64).T)) @ q64.astype(np.float32).astype(np.float64)
    assert np.max(np.abs(u - gref / 2)) <= 3e-5 * np.max(np.abs(gref / 2))


@pytest.mark.parametrize("n,stride", [(8, 1), (9, 1), (9, 3)])
def test_emulated_symmetric_tiles_match_oracle(n, stride):
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q64 = random_q(1 << n, seed=n).numpy()
    Kref, gref = orc.ksd_dense(q64, S, n)
    N, T = 1 << n, (1 << n) // 128
    A32, B32, A16, B16, a = _features(S, n)
    q = q64.astype(np.float32).astype(np.float64)
    h = np.array([[bin(i ^ j).count("1") for j in range(N)] for i in range(N)])
    u = np.zeros(N)
    for offset in range(stride):                              # the calls of `stride` ranks
        for I in range(offset, T, stride):
            ri = slice(128 * I, 128 * I + 128)
            for J in range(I, T):                             # chunk c of A meets chunk c of B: plain products
                cj = slice(128 * J, 128 * J + 128)
                D = A32[ri].astype(np.float64) @ B32[cj].astype(np.float64).T + A16[ri].astype(np.float64) @ B16[cj].astype(np.float64).T
                Y = a ** h[ri, cj] * D
                u[ri] += Y @ q[cj]
                if J != I:
                    u[cj] += Y.T @ q[ri]
    uref = gref / 2
    assert np.max(np.abs(u - uref)) <= 1e-5 * np.max(np.abs(uref))
    assert abs(q @ u - Kref) <= 1e-5 * abs(Kref)


def test_patch_weights_factorise():
    """The epilogue's weight of a (row slot k, column slot m, e) of a thread's 4 x 16 patch is a per-thread
    constant times a^popc((k ^ m) & 3) times a^((lqi & 1) ^ (m >> 2)): check against popc of the in-tile bits."""
    for lqi in range(4):
        for half in range(2):
            for t in range(32):
                i0 = (t >> 2) & 1
                c0 = bin(((t >> 3) & 3) ^ (t & 3)).count("1") + (((lqi >> 1) ^ half) & 1)
                for k in range(4):
                    row = 32 * lqi + (t >> 2) + 8 * k
                    for m in range(8):
                        for e in range(2):
                            col = 64 * half + 8 * m + 2 * (t & 3) + e
                            d = (k ^ m) & 3
                            want = bin(row ^ col).count("1")
                            got = c0 + (d & 1) + (d >> 1) + ((lqi & 1) ^ (m >> 2)) + (i0 ^ e)
                            assert got == want


def test_transposing_exchange_lands_on_columns_2t():
    """Three exchange levels (lane xor 16, 8, 4) leave thread t with the sums of columns 2t, 2t+1 of the half."""
    rng = np.random.default_rng(0)
    vals = rng.standard_normal((32, 8, 2))                     # [thread][m][e] partial sums over the thread's 4 rows
    cur = {t: {m: vals[t, m].copy() for m in range(8)} for t in range(32)}
    for bitmask, width in ((16, 4), (8, 2), (4, 1)):
        nxt = {}
        for t in range(32):
            b = 1 if t & bitmask else 0
            ms = sorted(cur[t])
            keep = ms[width:] if b else ms[:width]
            nxt[t] = {m: cur[t][m] + cur[t ^ bitmask][m] for m in keep}
        cur = nxt
    for t in range(32):
        (m, v), = cur[t].items()
        assert m == t >> 2                                      # columns 8m + 2(t&3) + e = 2t + e
        want = sum(vals[tt, m] for tt in range(32) if (tt & 3) == (t & 3))
        assert np.allclose(v, want)


def test_workspace_query():
    lib = _lib.load()
    for split in (0, 1, 2, 3):
        for n in range(7, 25):
            assert lib.tnk_ksd_sym_workspace_bytes(n, split, 0, 1) > 0
            assert lib.tnk_ksd_sym_workspace_bytes(n, split, 3, 8) > 0
        for n in (0, 6, 25, 30):
            assert lib.tnk_ksd_sym_workspace_bytes(n, split, 0, 1) == 0
    assert lib.tnk_ksd_sym_workspace_bytes(12, 0, 0, 0) == 0 and lib.tnk_ksd_sym_workspace_bytes(12, 0, -1, 2) == 0
    assert lib.tnk_ksd_sym_workspace_bytes(12, 4, 0, 1) == 0                    # unknown operand split
    # n = 20: q (4 MB) + row partials + column sums (8 MB) + column table 2^20 * (26 | 20) chunks * 16 B
    assert lib.tnk_ksd_sym_workspace_bytes(20, 0, 0, 1) >= (1 << 20) * 26 * 16
    assert (1 << 20) * 20 * 16 <= lib.tnk_ksd_sym_workspace_bytes(20, 1, 0, 1) <= lib.tnk_ksd_sym_workspace_bytes(20, 0, 0, 1) - (1 << 20) * 6 * 16


def _features_tile(S, n, length_scale=1.0):
    """Operand format 2 ("tf32_tile"): only the T1/T2 terms of bit positions < 7 stay on the MMAs."""
    N = 1 << n
    np_ = (n + 3) // 4 * 4
    a = np.exp(-1.0 / (n * length_scale))
    alpha, gamma = 1.0 - a, a - 1.0 / a
    z = np.arange(N)
    s = np.zeros((N, np_), dtype=np.float32)
    s[:, :n] = S[:, ::-1].astype(np.float32)
    bit = np.zeros((N, np_), dtype=np.float32)
    bit[:, :n] = (z[:, None] >> np.arange(n)[None, :]) & 1
    valid = np.zeros(np_, dtype=np.float32)
    valid[:n] = 1
    g = (np.float32(gamma) * (1 - s)) * (1 - 2 * bit) * valid
    C = (alpha * ((1 - s) * valid).sum(1) + gamma * ((1 - s) * bit).sum(1)).astype(np.float32)
    hi, lo = _tf_hi, lambda v: np.asarray(v, dtype=np.float32) - _tf_hi(v)
    one = np.ones(N, dtype=np.float32)
    low = np.zeros(8, dtype=np.float32)
    low[:7] = 1                                                   # positions 0..6 of the low fields
    gl, bl = g[:, :8] * low, bit[:, :8] * low
    nc = np_ // 4
    k32 = (nc + 5 + 1) & ~1
    pad32 = np.zeros((N, 4 * (k32 - nc - 5)), dtype=np.float32)
    A32 = np.concatenate([hi(s), hi(gl), bl, np.stack([hi(C), lo(C), one, one], 1), pad32], 1)
    B32 = np.concatenate([hi(s), bl, hi(gl), np.stack([one, one, hi(C), lo(C)], 1), pad32], 1)
    s16 = (2 * np_ + 16 + 15) // 16
    pad16 = np.zeros((N, 16 * s16 - 2 * np_ - 16), dtype=np.float32)
    A16 = np.concatenate([_bf(s), _bf(lo(s)), _bf(lo(gl)), bl, pad16], 1)
    B16 = np.concatenate([_bf(lo(s)), _bf(s), bl, _bf(lo(gl)), pad16], 1)
    return A32, B32, A16, B16, g.astype(np.float64), a, k32 // 2 + s16


@pytest.mark.parametrize("n", [8, 9, 10])
def test_emulated_tile_format_matches_oracle(n):
    """Format 2: MMA part (dot product, C, T terms of bits < 7) + the tile-constant terms R_i(J), c_j(I) folded into the
    row and column sums through W q_J, W (c q_J), W q_I, W (R q_I) with W the in-tile Hamming weights."""
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q64 = random_q(1 << n, seed=n).numpy()
    Kref, gref = orc.ksd_dense(q64, S, n)
    N, T = 1 << n, (1 << n) // 128
    A32, B32, A16, B16, g, a, steps = _features_tile(S, n)
    assert steps == {8: 6, 9: 7, 10: 7}[n]                       # 13 at n = 20 becomes 9 (tnk_tc.cu: tcs_k32, tcs_steps16)
    q = q64.astype(np.float32).astype(np.float64)
    lowidx = np.arange(128)
    W = a ** np.array([[bin(i ^ j).count("1") for j in lowidx] for i in lowidx], dtype=np.float64)
    u = np.zeros(N)
    for I in range(T):
        ri = slice(128 * I, 128 * I + 128)
        for J in range(I, T):
            cj = slice(128 * J, 128 * J + 128)
            wg = a ** bin(I ^ J).count("1")
            D = A32[ri].astype(np.float64) @ B32[cj].astype(np.float64).T + A16[ri].astype(np.float64) @ B16[cj].astype(np.float64).T
            Y = wg * W * D                                        # per-pair path of the epilogue
            u[ri] += Y @ q[cj]
            if J != I:
                u[cj] += Y.T @ q[ri]
            hib = [p for p in range(7, n)]
            R = sum(((J >> (p - 7)) & 1) * g[ri, p] for p in hib) if hib else np.zeros(128)
            c = sum(((I >> (p - 7)) & 1) * g[cj, p] for p in hib) if hib else np.zeros(128)
            u[ri] += wg * (R * (W @ q[cj]) + W @ (c * q[cj]))     # helper warps: rows
            if J != I:
                u[cj] += wg * (c * (W @ q[ri]) + W @ (R * q[ri]))   # helper warps: mirrored columns
    uref = gref / 2
    assert np.max(np.abs(u - uref)) <= 1e-5 * np.max(np.abs(uref))
    assert abs(q @ u - Kref) <= 1e-5 * abs(Kref)
