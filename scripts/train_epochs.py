"""BASELINE.json config #5: full KSDVariationalInference.train epochs on a synthetic random binary BN
(default n = 24) on 1..8 B200s.  One rank per GPU under torchrun; prints one JSON line with s/epoch.

    python scripts/train_epochs.py [--n 24] [--epochs 2] [--precision tc32] [--algorithm pairs|pairs_sym|modeproduct]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from tensornetworks_b200.ksd_vi import KSDVariationalInference
from tensornetworks_b200.synthetic import random_binary_bn


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--latent-bits", dest="n", type=int, default=24)
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--precision", default="tc32")
    ap.add_argument("--algorithm", default="pairs")
    ap.add_argument("--cuda-graph", action="store_true", help="replay the epoch as CUDA graphs (single GPU, Adam)")
    ap.add_argument("--profile", action="store_true", help="print the top CUDA ops of the timed epochs (torch.profiler)")
    args = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    torch.manual_seed(0)                                   # replicated Born machine: same parameters on every rank
    bn, lat, obs, x = random_binary_bn(args.n, seed=0)
    t0 = time.perf_counter()
    vi = KSDVariationalInference(bn, lat, obs, {"use_logits": True, "conditioning_dim": 0}, device=dev,
                                 precision=args.precision, compute_device=dev, algorithm=args.algorithm,
                                 cuda_graph=args.cuda_graph, distributed=world > 1)
    build_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    hist = vi.train(x, num_epochs=1, lr_born_machine=0.05, verbose=False)      # warm-up epoch: tables, workspaces
    torch.cuda.synchronize()
    first_s = time.perf_counter() - t0
    if args.profile and rank == 0:
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            vi.train(x, num_epochs=args.epochs, lr_born_machine=0.05, verbose=False)
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60), file=sys.stderr)
    t0 = time.perf_counter()
    hist = vi.train(x, num_epochs=args.epochs, lr_born_machine=0.05, verbose=False)
    torch.cuda.synchronize()
    per_epoch = (time.perf_counter() - t0) / args.epochs
    if world > 1:
        t = torch.tensor([per_epoch], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        per_epoch = t.item()
    if rank == 0:
        print(json.dumps({"workload": f"KSDVariationalInference.train, synthetic random binary BN, {args.n} latent vars",
                          "n_gpus": world, "precision": args.precision, "algorithm": args.algorithm, "cuda_graph": args.cuda_graph,
                          "s_per_epoch": per_epoch, "pairs_per_s": 4.0 ** args.n / per_epoch if args.algorithm.startswith("pairs") else None,
                          "epochs_timed": args.epochs, "first_train_call_s": first_s, "constructor_s": build_s,
                          "loss_ksd": hist["loss_ksd"][-1]}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
