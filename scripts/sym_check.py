"""Developer probe: symmetric tile kernel (tnk_ksd_sym) against the fp64 pair kernel (small n) or the
mode-product kernels (large n), striped calls that must add up, then timing."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

split = os.environ.get("SYM_SPLIT", "tf32")
sizes = [int(a) for a in sys.argv[1].split(",") if a] if len(sys.argv) > 1 else [7, 8, 9, 10, 12, 14]
time_sizes = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 and sys.argv[2] else []
for n in sizes:
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    q = random_q(1 << n, seed=n).cuda()
    ref = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    u0, K0 = ref.ksd_rows(q) if n <= 16 else ref.ksd_modeproduct(q)
    tc = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    u1, K1 = tc.ksd_sym(q, split=split)
    u2, K2 = tc.ksd_sym(q, split=split)            # second call reuses the column table
    ua, Ka = tc.ksd_sym(q, 0, 3, split=split)
    ub, Kb = tc.ksd_sym(q, 1, 3, split=split)
    uc, Kc = tc.ksd_sym(q, 2, 3, split=split)
    torch.cuda.synchronize()
    scale = u0.abs().max()
    print(json.dumps({"n": n, "K_ref": K0.item(), "K_sym": K1.item(), "rel_K": (abs(K1 - K0) / abs(K0)).item(),
                      "normwise_u": ((u1 - u0).abs().max() / scale).item(),
                      "rerun_u": ((u2 - u1).abs().max() / scale).item(),
                      "stripes3_u": ((ua + ub + uc - u0).abs().max() / scale).item(),
                      "stripes3_K": (abs(Ka + Kb + Kc - K0) / abs(K0)).item()}), flush=True)
for n in time_sizes:
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    q = random_q(1 << n, seed=0).cuda()
    tc = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    tc.ksd_sym(q, split=split); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record(); k1.record()
    best, bestk = 1e30, 1e30
    for _ in range(3):
        ops.time_pair_kernel(k0, k1)
        e0.record(); u, part = tc.ksd_sym(q, split=split); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1)); bestk = min(bestk, k0.elapsed_time(k1))
    ops.time_pair_kernel(None, None)
    e0.record(); u, part = tc.ksd_rows(q); e1.record(); torch.cuda.synchronize()
    print(json.dumps({"n": n, "kernel": "sym-" + split, "ms_total": best, "ms_tile_kernel": bestk, "pairs_per_s": 4.0 ** n / (best * 1e-3),
                      "K": part.item(), "ms_tc32_full": e0.elapsed_time(e1)}), flush=True)
