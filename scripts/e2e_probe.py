"""Where the end-to-end step of the public op spends its time (host q in, host gradient out), per operand format."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
bn, lat, obs, x = random_binary_bn(n, seed=0)
tables = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
q_host = random_q(1 << n, seed=0).pin_memory()
q_dev = q_host.cuda()
for split in ("tf32", "tf32_dot"):
    ops.DEFAULT_SYM_SPLIT = split
    for rep in range(6):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        qh = q_host.requires_grad_(True); qh.grad = None
        K = ops.ksd_quadratic_form(qh, tables, algorithm="pairs_sym")
        t1 = time.perf_counter()
        K.backward()
        t2 = time.perf_counter()
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); tables.ksd_sym(q_dev, split=split); b.record(); torch.cuda.synchronize()
        if rep >= 2:
            print(f"{split:9s} forward {1e3*(t1-t0):7.2f} ms  backward {1e3*(t2-t1):6.2f} ms  tail {1e3*(t3-t2):5.2f} ms  | device-only call {a.elapsed_time(b):7.2f} ms", flush=True)
