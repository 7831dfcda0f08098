"""Run under torchrun on >= 2 GPUs of one box: tnk_comm (our all-reduce over peer-mapped buffers, include/tnk.h) against
the NCCL sum-all-reduce on the SAME partial vectors (bit-identical at world 2, where the two summation orders coincide;
<= 1e-15 relative otherwise), every rank bit-identical to every other, and the sharded evaluations (pairs_sym,
pairs_sym64, modeproduct) through it against the single-rank results.  Prints one JSON line."""
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
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dist.init_process_group("nccl", device_id=dev)
    N = 1 << n
    comm = ops.PeerComm(N + 1, dev)
    # 1. raw all-reduce: odd and even counts, several rounds (barrier epochs, buffer reuse)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    worst = 0.0
    for rnd, count in enumerate([N + 1, 1, 7, N, 1000, N + 1]):
        x = torch.randn(N + 1, generator=g, device=dev, dtype=torch.float64) * (10.0 ** (rnd - 2))
        comm.sendbuf.copy_(x)
        got = comm.allreduce_sum(count).clone()
        ref = x[:count].clone()
        dist.all_reduce(ref, op=dist.ReduceOp.SUM)
        if world == 2:
            assert torch.equal(got, ref), f"round {rnd}: tnk_comm differs from NCCL bitwise"
        worst = max(worst, ((got - ref).abs().max() / ref.abs().max()).item())
        assert worst <= 1e-14
        all_ranks = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(all_ranks, got)
        assert all(torch.equal(a, all_ranks[0]) for a in all_ranks), "ranks hold different sums"
    # 2. the sharded evaluations through the public op
    bn, lat, obs, x_ev = random_binary_bn(n, seed=1)
    q = random_q(N, seed=1).to(dev)
    res = {}
    for precision, algorithm, tol in (("tc32", "pairs_sym", 1e-4), ("fp64", "pairs_sym64", 1e-12), ("fp64", "modeproduct", 1e-12)):
        tabs = ops.SteinTables.from_network(bn, x_ev, lat, precision=precision, device=dev)
        qa = q.clone().requires_grad_(True)
        K1 = ops.ksd_quadratic_form(qa, tabs, algorithm=algorithm)            # no group: whole evaluation on this GPU
        K1.backward()
        qb = q.clone().requires_grad_(True)
        Kc = ops.ksd_quadratic_form(qb, tabs, algorithm=algorithm, exchange=comm)
        Kc.backward()
        qc = q.clone().requires_grad_(True)
        Kn = ops.ksd_quadratic_form(qc, tabs, algorithm=algorithm, group=dist.group.WORLD)   # NCCL collective
        Kn.backward()
        scale = qa.grad.abs().max().item()
        assert abs(Kc.item() - K1.item()) <= tol * abs(K1.item()), (algorithm, Kc.item(), K1.item())
        assert (qb.grad - qa.grad).abs().max().item() <= tol * scale, algorithm
        assert abs(Kn.item() - K1.item()) <= tol * abs(K1.item())
        assert (qc.grad - qa.grad).abs().max().item() <= tol * scale
        if algorithm != "pairs_sym" and world == 2:      # deterministic partials: the two exchanges must agree bitwise
            assert torch.equal(qb.grad, qc.grad) and Kc.item() == Kn.item(), algorithm
        res[algorithm] = abs(Kc.item() - K1.item()) / abs(K1.item())
    # 3. the fused C entry point
    import ctypes
    from tensornetworks_b200 import _lib
    tabs = ops.SteinTables.from_network(bn, x_ev, lat, precision="tc32", device=dev)
    lib = _lib.load()
    ws = torch.empty(lib.tnk_ksd_sym_workspace_bytes(n, 0, rank, world), dtype=torch.uint8, device=dev)
    u = torch.empty(N, dtype=torch.float64, device=dev)
    K = torch.empty(1, dtype=torch.float64, device=dev)
    _lib.check(lib.tnk_ksd_sym_allreduce(n, 1.0, 0, q.data_ptr(), 0, tabs.ST.data_ptr(), tabs.Cst.data_ptr(), 0, comm._comm,
                                         u.data_ptr(), K.data_ptr(), ws.data_ptr(), ws.numel(), None,
                                         ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "tnk_ksd_sym_allreduce")
    u1, K1 = tabs.ksd_sym(q)
    assert abs(K.item() - K1.item()) <= 1e-6 * abs(K1.item())
    assert (u - u1).abs().max().item() <= 1e-6 * u1.abs().max().item()
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        print(json.dumps({"tnk_comm": "ok", "world": world, "n": n, "allreduce_vs_nccl_max_rel": worst, "K_rel": res}), flush=True)
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
