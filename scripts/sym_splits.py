"""Column splits of the symmetric tile kernel on one rank's stripe (rank 0 of G: row tiles 0, G, 2G, ...), timed on one
GPU: how many CTAs the triangle needs to balance over 148 SMs.  TNK_SYM_SPLITS overrides the planner."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
strides = [int(v) for v in (sys.argv[2] if len(sys.argv) > 2 else "1,8").split(",")]
bn, lat, obs, x = random_binary_bn(n, seed=0)
q = random_q(1 << n, seed=0).cuda()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); b.record()
out = {}
for stride in strides:
    for splits in ("planner", 1, 2, 4, 8, 16, 32):
        if splits == "planner":
            os.environ.pop("TNK_SYM_SPLITS", None)
        else:
            os.environ["TNK_SYM_SPLITS"] = str(splits)
        tc = ops.SteinTables.from_network(bn, x, lat, precision="tc32")     # fresh workspace per setting
        tc.ksd_sym(q, 0, stride)
        times = []
        for _ in range(3):
            ops.time_pair_kernel(a, b)
            tc.ksd_sym(q, 0, stride)
            torch.cuda.synchronize()
            times.append(a.elapsed_time(b))
        ops.time_pair_kernel(None, None)
        out[f"stride {stride}, splits {splits}"] = round(min(times), 3)
        del tc
os.environ.pop("TNK_SYM_SPLITS", None)
print(json.dumps({"n": n, "kernel_ms": out}))
