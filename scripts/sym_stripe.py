"""Tile rate of the symmetric tcgen05 kernel on one rank's stripe (row tiles 0, 8, 16, ... = rank 0 of 8) at n = 20 and
n = 24: explains the per-pair rate of BASELINE.json config #5 (n = 24 on 8 GPUs) against the n = 20 headline."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

stride = int(sys.argv[1]) if len(sys.argv) > 1 else 8
out = {}
for n in (20, 24):
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    q = random_q(1 << n, seed=0).cuda()
    tc = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); b.record()
    tc.ksd_sym(q, 0, stride)                      # builds the column table
    ops.time_pair_kernel(a, b)
    tc.ksd_sym(q, 0, stride)
    torch.cuda.synchronize()
    ops.time_pair_kernel(None, None)
    ms = a.elapsed_time(b)
    T = (1 << n) // 128
    tiles = sum(T - I for I in range(0, T, stride))
    nc = (n + 3) // 4
    s32, s16 = ((3 * nc + 2) & ~1) // 2, nc
    out[n] = {"kernel_ms": ms, "tiles": tiles, "tiles_per_s": tiles / (ms * 1e-3), "k_steps": s32 + s16,
              "tf32_equiv_TFLOPs": tiles * 128 * 128 * 16 * (s32 + s16) / (ms * 1e-3) / 1e12,
              "ns_per_tile_per_SM": ms * 1e6 / (tiles / 148)}
    del tc
    torch.cuda.empty_cache()
out["tile_rate_ratio_n24_over_n20"] = out[24]["tiles_per_s"] / out[20]["tiles_per_s"]
out["k_step_ratio_n20_over_n24"] = out[20]["k_steps"] / out[24]["k_steps"]
print(json.dumps(out))
