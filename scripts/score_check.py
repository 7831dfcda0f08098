"""Developer probe: time the score kernel (K2) alone, back to back, at a few sizes."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops, _lib
from tensornetworks_b200.synthetic import random_binary_bn

for n in [int(a) for a in (sys.argv[1] if len(sys.argv) > 1 else "16,20,22").split(",")]:
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    for prec, keep in (("fp32", True), ("fp32", False), ("fp64", False)):
        t = ops.SteinTables.from_network(bn, x, lat, precision=prec, keep_scores=keep)
        lib = _lib.load()
        st = torch.cuda.current_stream().cuda_stream
        call = lambda: _lib.check(lib.tnk_score(n, t.length_scale, t.P.data_ptr(), t.S.data_ptr() if keep else None, t.prec_code,
                                                t.ST.data_ptr(), t.Cst.data_ptr(), st), "tnk_score")
        call(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(8):
            call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 8
        N = 1 << n
        nbytes = 8.0 * N + (8.0 * N * n if keep else 0) + t.ST.element_size() * (t.ST.numel() + t.Cst.numel())
        print(json.dumps({"n": n, "precision": prec, "keep_scores": keep, "ms": ms, "GB": nbytes / 1e9, "GBps": nbytes / ms / 1e6}), flush=True)
