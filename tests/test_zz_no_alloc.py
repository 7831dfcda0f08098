"""tnk_ksd_rows_static (include/tnk.h): the allocation-free variant of tnk_ksd_rows whose scratch memory is a static
device array inside the library.  Called through the C-ABI (ctypes) and compared with tnk_ksd_rows on the same inputs,
twice in a row (the array keeps the previous call's contents), in the tolerance class of the other GPU tests
(bit-identical here: both run the same kernels, only the workspace address differs).  Collected last on purpose."""
import ctypes

import pytest
import torch

from tensornetworks_b200 import _lib, ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

pytestmark = pytest.mark.gpu


def _static_call(tables, q, row_begin, row_end):
    lib = _lib.load()
    u = torch.zeros(tables.N, dtype=torch.float64, device=tables.device)
    part = torch.zeros(1, dtype=torch.float64, device=tables.device)
    launched = ctypes.c_int(0)
    with torch.cuda.device(tables.device):
        rc = lib.tnk_ksd_rows_static(tables.n, tables.length_scale, tables.prec_code, q.data_ptr(),
                                     0 if q.dtype == torch.float64 else 1, tables.ST.data_ptr(), tables.Cst.data_ptr(),
                                     row_begin, row_end, u.data_ptr(), part.data_ptr(), ctypes.byref(launched),
                                     torch.cuda.current_stream(tables.device).cuda_stream)
    return rc, u, part, launched.value


@pytest.mark.parametrize("precision,n,tol", [("fp64", 12, 1e-12), ("fp32", 12, 1e-12), ("tc32", 10, 1e-12)])
def test_static_scratch_matches_workspace_entry_point(precision, n, tol):
    bn, lat, obs, x = random_binary_bn(n, seed=3)
    tables = ops.SteinTables.from_network(bn, x, lat, precision=precision)
    N = 1 << n
    for call, (seed, rows) in enumerate([(1, (0, N)), (2, (128, N - 256))]):      # two calls in a row, different q and rows
        q = random_q(N, seed=seed).to("cuda", torch.float64)
        u_ref, part_ref = tables.ksd_rows(q, rows[0], rows[1])
        rc, u, part, launched = _static_call(tables, q, rows[0], rows[1])
        assert rc == 0, _lib.load().tnk_last_error().decode()
        assert launched >= 3
        torch.cuda.synchronize()
        sl = slice(rows[0], rows[1])
        scale = u_ref[sl].abs().max().item()
        assert (u[sl] - u_ref[sl]).abs().max().item() <= tol * scale, f"call {call}"
        assert abs(part.item() - part_ref.item()) <= tol * abs(part_ref.item()), f"call {call}"


def test_static_scratch_rejects_problems_beyond_the_array():
    lib = _lib.load()
    cap = lib.tnk_ksd_rows_static_capacity()
    assert cap == 16 << 20
    n = 20
    assert lib.tnk_ksd_workspace_bytes(n, _lib.PREC_FP64, 0, 1 << n) > cap
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = lib.tnk_ksd_rows_static(n, 1.0, _lib.PREC_FP64, dummy.data_ptr(), 0, dummy.data_ptr(), dummy.data_ptr(),
                                 0, 1 << n, dummy.data_ptr(), dummy.data_ptr(), None, None)
    assert rc != 0                                                   # refused before anything is enqueued
    msg = lib.tnk_last_error().decode()
    assert "TNK_STATIC_SCRATCH_BYTES" in msg and str(cap) in msg
