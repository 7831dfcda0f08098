"""Developer probe: tensor-core variant (precision tc32) against the fp64 pair kernel, then timing."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

sizes = [int(a) for a in sys.argv[1].split(",") if a] if len(sys.argv) > 1 else [7, 8, 10, 12, 14]
time_sizes = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 and sys.argv[2] else []
for n in sizes:
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    q = random_q(1 << n, seed=n).cuda()
    ref = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    u0, K0 = ref.ksd_rows(q)
    tc = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    u1, K1 = tc.ksd_rows(q)
    torch.cuda.synchronize()
    err_u = ((u1 - u0).abs().max() / u0.abs().max()).item()
    err_K = (abs(K1 - K0) / abs(K0)).item()
    print(json.dumps({"n": n, "K_fp64": K0.item(), "K_tc32": K1.item(), "rel_K": err_K, "normwise_u": err_u}), flush=True)
peak = None
for n in time_sizes:
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    q = random_q(1 << n, seed=0).cuda()
    tc = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    tc.ksd_rows(q); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record(); k1.record()
    best, bestk = 1e30, 1e30
    for _ in range(3):
        ops.time_pair_kernel(k0, k1)
        e0.record(); u, part = tc.ksd_rows(q); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1)); bestk = min(bestk, k0.elapsed_time(k1))
    ops.time_pair_kernel(None, None)
    pairs = 4.0 ** n
    print(json.dumps({"n": n, "prec": "tc32", "ms_total": best, "ms_tile_kernel": bestk, "pairs_per_s": pairs / (best * 1e-3),
                      "K": part.item()}), flush=True)
