"""Symmetric fp64 kernel (tnk_ksd_sym64) against the fp64 all-pairs kernel: errors, unit stripes, reproducibility, times."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

sizes = [int(a) for a in sys.argv[1:]] or [7, 8, 9, 10, 11, 12, 13, 14, 16, 17, 18, 20]
for n in sizes:
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    q = random_q(1 << n, seed=n).cuda()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    u0, k0 = tabs.ksd_rows(q)
    torch.cuda.synchronize()
    e0.record(); u0, k0 = tabs.ksd_rows(q); e1.record(); torch.cuda.synchronize()
    t_rows = e0.elapsed_time(e1)
    u1, k1 = tabs.ksd_sym64(q)
    torch.cuda.synchronize()
    e0.record(); u2, k2 = tabs.ksd_sym64(q); e1.record(); torch.cuda.synchronize()
    t_sym = e0.elapsed_time(e1)
    err_u = ((u1 - u0).abs().max() / u0.abs().max()).item()
    err_k = abs(k1.item() - k0.item()) / abs(k0.item())
    same = bool(torch.equal(u1, u2) and torch.equal(k1, k2))
    stripes = ""
    if n <= 16:
        ua = sum(tabs.ksd_sym64(q, g, 3)[0] for g in range(3))
        stripes = f" stripes3 {((ua - u1).abs().max() / u1.abs().max()).item():.1e}"
    print(f"n={n:2d} rows {t_rows:9.3f} ms  sym64 {t_sym:9.3f} ms  x{t_rows / t_sym:.2f}  err_u {err_u:.2e} err_K {err_k:.2e} "
          f"reproducible {same}{stripes}", flush=True)
    del tabs
