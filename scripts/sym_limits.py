"""Which unit bounds the symmetric tile kernel: times tnk_ksd_sym with parts of the epilogue switched
off (TNK_TCS_ABLATE bits: 1 no mirrored part, 4 accumulators read from TMEM but not folded, 8 accumulators not read),
for the 6-step operand format (the compile-time instances exist for n = 17..20, 4 row tiles per CTA).  Results are WRONG when a bit is set - timing only."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
bn, lat, obs, x = random_binary_bn(n, seed=0)
q = random_q(1 << n, seed=0).cuda()
tc = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); b.record()
out = {}
for split in ("tf32_dot",):
    for name, flag in (("full", 0), ("no mirrored part", 1), ("read, no fold, no mirrored part", 5), ("MMAs only (no read, no fold)", 13),
                       ("epilogue only (no MMAs)", 16), ("epilogue only, no mirrored part", 17),
                       ("read-out loop only (no MMAs, no fold, no mirrored part)", 21)):
        os.environ["TNK_TCS_ABLATE"] = str(flag)
        tc.ksd_sym(q, split=split)
        times = []
        for _ in range(3):
            ops.time_pair_kernel(a, b)
            tc.ksd_sym(q, split=split)
            torch.cuda.synchronize()
            times.append(a.elapsed_time(b))
        ops.time_pair_kernel(None, None)
        out[f"{split}: {name}"] = round(min(times), 2)
os.environ["TNK_TCS_ABLATE"] = "0"
print(json.dumps({"n": n, "row_tiles_env": os.environ.get("TNK_TCS_ROWTILES"), "kernel_ms": out}))
