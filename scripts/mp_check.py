"""Developer probe: mode-product path against the fp64 pair kernel (parity), then timing under an HBM roofline.

usage: mp_check.py [parity sizes, comma separated] [timing sizes]
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

sizes = [int(a) for a in sys.argv[1].split(",") if a] if len(sys.argv) > 1 else [1, 2, 3, 5, 8, 12, 13, 14, 15, 16]
time_sizes = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 and sys.argv[2] else []
for n in sizes:
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    q = random_q(1 << n, seed=n).cuda()
    ref = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    u0, K0 = ref.ksd_rows(q)
    u1, K1 = ref.ksd_modeproduct(q)
    cut = n // 3
    ua, Ka = ref.ksd_modeproduct(q, 0, cut)
    ub, Kb = ref.ksd_modeproduct(q, cut, n)
    torch.cuda.synchronize()
    den = u0.abs().max()
    print(json.dumps({"n": n, "K_pairs": K0.item(), "K_mp": K1.item(), "rel_K": (abs(K1 - K0) / abs(K0)).item(),
                      "normwise_u": ((u1 - u0).abs().max() / den).item(),
                      "normwise_u_split": ((ua + ub - u0).abs().max() / den).item(),
                      "rel_K_split": (abs(Ka + Kb - K0) / abs(K0)).item()}), flush=True)
for n in time_sizes:
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    q = random_q(1 << n, seed=0).cuda()
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="fp32", keep_scores=False)
    N = 1 << n
    for storage, elem in (("fp64", 8.0), ("fp32", 4.0)):
        tabs.ksd_modeproduct(q, storage=storage); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        for _ in range(5):
            e0.record(); u, part = tabs.ksd_modeproduct(q, storage=storage); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        nbytes = elem * N * (2 * n + 2 * (n + 2)) + 8.0 * N * 4      # S twice, V written and read once; q twice, r_sum, u
        print(json.dumps({"n": n, "algorithm": "modeproduct", "storage": storage, "ms": best,
                          "effective_pairs_per_s": 4.0 ** n / (best * 1e-3), "algorithmic_GB": nbytes / 1e9,
                          "GBps": nbytes / (best * 1e-3) / 1e9, "K": part.item()}), flush=True)
        tabs._ws = {}
