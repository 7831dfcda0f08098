"""Run under torchrun on >= 2 GPUs of one box: the row-sharded evaluation with the exchange fused
into the reduction kernels (tnk_ksd_rows_peers, NVLink peer stores) must give every rank the same
u and K as the NCCL all_gather path and as a single-rank evaluation.  Prints one JSON line."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 12
    precision = sys.argv[2] if len(sys.argv) > 2 else "fp64"
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dist.init_process_group("nccl", device_id=dev)
    N = 1 << n
    bn, lat, obs, x = random_binary_bn(n, seed=1)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision=precision, device=dev)
    q = random_q(N, seed=1).to(dev)
    u_ref, K_ref = tabs.ksd_rows(q)                       # all rows on this rank
    r0, r1 = ops.shard_rows(N, rank, world)
    u_sh, part = tabs.ksd_rows(q, r0, r1)
    u_nccl, K_nccl = ops.combine_row_shards(u_sh[r0:r1], part, N, rank, world)
    ex = ops.PeerExchange(N, dev)
    worst = 0.0
    for it in range(5):                                   # alternates the two sink buffers
        qi = q if it % 2 == 0 else torch.roll(q, 1)
        u_f, K_f = ops.ksd_rows_fused(tabs, qi, ex)
        if it % 2 == 0:
            assert torch.equal(u_f, u_nccl), "fused u differs from the NCCL path"
            # a shard may use another column-split count than the full evaluation: same math, other summation order
            assert (u_f[r0:r1] - u_ref[r0:r1]).abs().max().item() <= 1e-6 * u_ref.abs().max().item()
            worst = max(worst, (abs(K_f - K_ref) / abs(K_ref)).item())
            assert abs(K_f.item() - K_nccl.item()) <= 1e-12 * abs(K_nccl.item())
    qa = q.clone().requires_grad_(True)
    Ka = ops.ksd_quadratic_form(qa, tabs, exchange=ex)
    Ka.backward()
    assert (qa.grad - 2 * u_ref).abs().max().item() <= 2e-6 * u_ref.abs().max().item()
    gathered = [torch.zeros_like(K_f) for _ in range(world)]
    dist.all_gather(gathered, K_f)
    assert all(torch.equal(g, gathered[0]) for g in gathered), "K differs between ranks"
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        print(json.dumps({"fused_exchange": "ok", "world": world, "n": n, "precision": precision,
                          "K": K_f.item(), "rel_vs_single_rank": worst}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
