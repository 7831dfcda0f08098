"""ctypes wrapper of oracle/libksd_oracle.so (C restatement).  TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libksd_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(HERE, "ksd_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "-B" if force else "-s"], check=True, capture_output=True)
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(LIB)
        vp, ci, i64, dbl = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double
        lib.orc_max_threads.restype = ci
        lib.orc_joint_table.restype = None
        lib.orc_joint_table.argtypes = [ci, ci, ci, vp, vp, vp, vp, vp, vp, vp]
        lib.orc_score_table.restype = None
        lib.orc_score_table.argtypes = [ci, vp, vp]
        lib.orc_stein_kernel_value.restype = dbl
        lib.orc_stein_kernel_value.argtypes = [ci, vp, i64, i64, vp, vp]
        lib.orc_ksd_rows.restype = dbl
        lib.orc_ksd_rows.argtypes = [ci, dbl, vp, vp, i64, i64, vp, ci]
        lib.orc_ksd_modeproduct.restype = dbl
        lib.orc_ksd_modeproduct.argtypes = [ci, dbl, vp, vp, vp, ci]
        _lib = lib
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def max_threads():
    return load().orc_max_threads()


def joint_table(flat):
    """flat: tensornetworks_b200.flatten.FlatBN (host arrays only)."""
    P = np.empty(1 << flat.n_latent, dtype=np.float64)
    pa_idx = flat.pa_idx if flat.pa_idx.size else np.zeros(1, dtype=np.int32)
    cpt = flat.cpt if flat.cpt.size else np.zeros((1, 2))
    load().orc_joint_table(flat.n_latent, flat.n_hidden, flat.n_nodes, _p(flat.node_kind), _p(flat.node_arg),
                           _p(flat.pa_off), _p(pa_idx), _p(flat.cpt_off), _p(np.ascontiguousarray(cpt)), _p(P))
    return P


def score_table(P, n):
    S = np.empty((P.shape[0], n), dtype=np.float64)
    load().orc_score_table(n, _p(np.ascontiguousarray(P)), _p(S))
    return S


def ksd_rows(q, S, n, r0, r1, length_scale=1.0, nthreads=None):
    """(u[r0:r1], partial K) by the literal four-term pair evaluation on `nthreads` host threads
    (default: all online cores)."""
    nthreads = max_threads() if nthreads is None else int(nthreads)
    q = np.ascontiguousarray(q, dtype=np.float64)
    S = np.ascontiguousarray(S, dtype=np.float64)
    u = np.empty(r1 - r0, dtype=np.float64)
    part = load().orc_ksd_rows(n, float(length_scale), _p(S), _p(q), r0, r1, _p(u), nthreads)
    return u, part


def ksd_modeproduct(q, P, n, length_scale=1.0, nthreads=None):
    """(K, dK/dq = 2 K_p q) for all rows through the Kronecker form (O2 in C), from the joint table P;
    the oracle for n = 20 / 24 where the numpy forms do not finish in test time."""
    nthreads = max_threads() if nthreads is None else int(nthreads)
    q = np.ascontiguousarray(q, dtype=np.float64)
    P = np.ascontiguousarray(P, dtype=np.float64)
    u = np.empty(1 << n, dtype=np.float64)
    K = load().orc_ksd_modeproduct(n, float(length_scale), _p(P), _p(q), _p(u), nthreads)
    return K, 2.0 * u
