"""Ablations of the symmetric tcgen05 tile kernel at n = 20: what the mirrored part, its fp64
REDs and the number of MMA K-steps are worth.  TNK_TCS_ABLATE makes the kernel skip work, so its results are wrong -
this script only times it.  Run each setting in its own process (the flag is read per call, the table is rebuilt)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
only = sys.argv[2].split(",") if len(sys.argv) > 2 else None      # e.g. tf32,tf32_dot
bn, lat, obs, x = random_binary_bn(n, seed=0)
q = random_q(1 << n, seed=0).cuda()
tc = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); b.record()
out = {}
for name, flag, split in (("full kernel (13 K-steps)", "0", "tf32"), ("no mirrored part", "1", "tf32"),
                          ("mirrored sums computed, not stored (no fp64 RED)", "2", "tf32"),
                          ("bf16x2 operands (10 K-steps)", "0", "bf16x2"), ("bf16x2, no mirrored part", "1", "bf16x2"),
                          ("tf32_tile operands (9 K-steps)", "0", "tf32_tile"), ("tf32_tile, no mirrored part", "1", "tf32_tile"),
                          ("tf32_dot operands (6 K-steps)", "0", "tf32_dot"), ("tf32_dot, no mirrored part", "1", "tf32_dot")):
    if only is not None and split not in only:
        continue
    os.environ["TNK_TCS_ABLATE"] = flag
    tc.ksd_sym(q, split=split)
    times = []
    for _ in range(3):
        ops.time_pair_kernel(a, b)
        tc.ksd_sym(q, split=split)
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    ops.time_pair_kernel(None, None)
    out[name] = min(times)
os.environ["TNK_TCS_ABLATE"] = "0"
print(json.dumps({"n": n, "kernel_ms": out}))
