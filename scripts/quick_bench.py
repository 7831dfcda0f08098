"""Developer timing probe (not the contract bench): ALU peaks and pair-kernel rates."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

sizes = [int(a) for a in sys.argv[1].split(",")] if len(sys.argv) > 1 else [14, 16]
precs = sys.argv[2].split(",") if len(sys.argv) > 2 else ["fp32", "fp64"]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
peaks = {}
for prec in ("fp32", "fp32x2", "fp64"):
    rate, ms = ops.alu_peak(prec, iters=8192)
    peaks[prec] = rate
    print(json.dumps({"alu_peak": prec, "lane_ops_per_s": rate, "ms": ms}), flush=True)
for n in sizes:
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    q = random_q(1 << n, seed=0).cuda()
    for prec in precs:
        tabs = ops.SteinTables.from_network(bn, x, lat, precision=prec)
        tabs.ksd_rows(q)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        for _ in range(reps):
            e0.record(); u, part = tabs.ksd_rows(q); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        pairs = 4.0 ** n
        print(json.dumps({"n": n, "prec": prec, "ms": best, "pairs_per_s": pairs / (best * 1e-3),
                          "frac_alu": pairs * (n + 10) / (best * 1e-3) / peaks[prec], "K": part.item()}), flush=True)
