"""Developer probe: streaming Born-machine kernels (K6) - time and effective GB/s of forward and backward."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops

for n in [int(a) for a in (sys.argv[1] if len(sys.argv) > 1 else "20,24,26").split(",")]:
    N = 1 << n
    logits = torch.randn(N, device="cuda").requires_grad_(True)
    w = torch.randn(N, device="cuda", dtype=torch.float64)
    def step():
        q, H = ops.born_softmax_entropy(logits)
        return q, H
    q, H = step(); torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    best_f, best_b = 1e30, 1e30
    for _ in range(5):
        logits.grad = None
        e[0].record(); q, H = step(); e[1].record()
        torch.autograd.backward([q, H], [w.to(q.dtype), torch.tensor(-0.01, device="cuda")]); e[2].record()
        torch.cuda.synchronize()
        best_f = min(best_f, e[0].elapsed_time(e[1])); best_b = min(best_b, e[1].elapsed_time(e[2]))
    fbytes = 4.0 * N * 4            # max pass, sum-exp pass, probs pass (read) + q write
    bbytes = 4.0 * N * 5            # dot pass (q, g), backward pass (q, g) + gradient write
    print(json.dumps({"n": n, "forward_ms": best_f, "forward_GBps": fbytes / best_f / 1e6,
                      "backward_ms": best_b, "backward_GBps": bbytes / best_b / 1e6}), flush=True)
