#!/usr/bin/env python
"""Benchmark of the exact-KSD hot path (BASELINE.json metric: ordered state-pairs/s of one
KSD^2 + dKSD^2/dq evaluation on a synthetic random binary BN with n latent bits).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 20] [--precision tc32|fp32|fp64]

A step = one evaluation of K = q^T K_p q and 2 K_p q for a fixed network and a given q.
Prints ONE JSON line on rank 0.  See DESIGN.md "Measurement" for how each field is obtained.
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

METRIC = "exact KSD+grad state-pairs/s"
UNIT = "state-pairs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--latent-bits", dest="n", type=int, default=20, help="latent bits (BASELINE.json quotes the metric at n=20)")
    ap.add_argument("--precision", default="tc32", choices=["fp32", "fp64", "tc32"],
                    help="pair-kernel arithmetic of the headline: tc32 = tcgen05 tf32 tiles with 3xTF32 split (1e-4 mode), "
                         "fp64 = the 1e-6 mode, fp32 = FFMA2 ALU kernel")
    ap.add_argument("--algorithm", default="auto", choices=["auto", "pairs", "pairs_sym"],
                    help="pairs_sym evaluates every unordered pair of 128-state tiles once (tnk_ksd_sym / tnk_ksd_sym64); "
                         "pairs evaluates all 4^n ordered pairs (tnk_ksd_rows); auto = pairs_sym where it exists (7 <= n <= 24)")
    ap.add_argument("--split", default=None, choices=["tf32", "bf16x2", "tf32_tile", "tf32_dot"],
                    help="operand format of the symmetric tcgen05 kernel (default: ops.DEFAULT_SYM_SPLIT)")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="N>1: our own exchange over peer-mapped buffers (tnk_comm / tnk_ksd_rows_peers) or one NCCL collective")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-context", action="store_true", help="skip the secondary objects (other kernels, configs #3 and #5)")
    return ap.parse_args()


def workload_name(n):
    return f"synthetic random binary BN, {n} latent vars, in-degree<=3, exact KSD+grad (seed 0, n_obs={max(1, n // 3)})"


def config_of(n):
    """The `config` object: identical on both arms (the driver compares them key by key)."""
    return {"workload": workload_name(n), "n_latent": n, "pairs_per_step": 4.0 ** n}


# ------------------------------------------------------------------ CPU side (oracle port)
def cpu_tables(n, seed):
    """Joint and score tables of the workload from the C oracle (independent of the CUDA path)."""
    from oracle import c_oracle
    from tensornetworks_b200.flatten import flatten_bn
    from tensornetworks_b200.synthetic import random_binary_bn, random_q
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    P = c_oracle.joint_table(flatten_bn(bn, x, lat))
    S = c_oracle.score_table(P, n)
    q = random_q(1 << n, seed=seed).numpy()
    return P, S, q


def cpu_sample(n, S, q, target_seconds, row_start=0):
    """Time the oracle's literal four-term pair loop (reference stein_utils.py:161-196 per pair,
    ksd_vi.py:114-131 over pairs) on all host threads over `rows` full rows of the pair space."""
    from oracle import c_oracle
    threads = c_oracle.max_threads()
    N = 1 << n
    rows = min(N - row_start, threads)
    t0 = time.perf_counter()
    c_oracle.ksd_rows(q, S, n, row_start, row_start + rows)
    dt = time.perf_counter() - t0
    rate = rows * N / dt
    rows = int(min(N - row_start, max(threads, (target_seconds * rate / N) // threads * threads)))
    t0 = time.perf_counter()
    u, part = c_oracle.ksd_rows(q, S, n, row_start, row_start + rows)
    dt = time.perf_counter() - t0
    return {"pairs": rows * N, "seconds": dt, "rate": rows * N / dt, "rows": rows, "threads": threads, "u": u,
            "row_start": row_start}


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for the path.  The reference is pure Python
    (4^n interpreter iterations, ~1e3 pairs/s) and is not present on the GPU box; its algorithm is
    timed through the C oracle port on all host cores - a far faster stand-in than the original."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.n
    P, S, q = cpu_tables(n, args.seed)
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    rates, info = [], None
    for step in range(args.warmup + args.steps):
        info = cpu_sample(n, S, q, per_step, row_start=0)
        if step >= args.warmup:
            rates.append(info["rate"])
    value = float(np.mean(rates))
    pairs_per_step = 4.0 ** n
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": pairs_per_step / value * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(n),
        "run": {"note": "ms_per_step extrapolated from the sample rate to the full 4^n pairs"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": info["threads"], "kind": "port",
                         "sample": f"{info['rows']} full rows x 2^{n} columns = {info['pairs']:.3g} ordered pairs per step"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------ clocks during the timed region
class ClockSampler:
    FIELDS = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.proc, self.gpu_index = None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self, t_begin=None, t_end=None):
        """Summarise the samples taken between the two time.time() stamps (all samples if None)."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        import datetime
        rows_all = []
        for ln in out.strip().splitlines():
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) >= 8:
                rows_all.append(parts)

        def in_window(parts):
            try:
                stamp = datetime.datetime.strptime(parts[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
            except ValueError:
                return False
            return t_begin is None or (t_begin - 0.05 <= stamp <= t_end + 0.05)
        chosen = [r for r in rows_all if in_window(r)] or rows_all[-5:]   # clock-skewed stamps: keep the last samples
        for parts in chosen:
            try:
                sm.append(float(parts[1])); smax.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def load_measured_peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


def ncu_traffic(kernel_tag, n):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from the committed ncu --set full
    summary of the same kernel and size (profiles/), or None when no capture exists for this n."""
    import glob
    import re
    best = None
    paths = glob.glob(os.path.join(REPO, "profiles", f"r*_ncu_full_{kernel_tag}_n{n}_v*.txt"))
    for path in sorted(paths, key=lambda f: (os.path.basename(f)[:3], int(re.search(r"_v(\d+)\.txt$", f).group(1)))):   # newest capture last
        text = open(path).read()
        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = 0.0
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            m = re.search(rf"^{re.escape(key)} = ([0-9.,]+) (\w+)", text, re.M)
            if not m:
                break
            tot += float(m.group(1).replace(",", "")) * unit.get(m.group(2), 1.0)
        else:
            best = (tot, os.path.basename(path))
    return best


def sym_tiles(n, rank, world):
    """128 x 128 tiles the symmetric schedule executes on `rank` (row tile I meets column tiles I..T-1)."""
    T = (1 << n) // 128
    return sum(T - I for I in range(rank, T, world))


def sym_steps(n, split):
    """(tf32 K-steps, bf16 K-steps) per 128 x 128 tile of the symmetric tile kernel (csrc/tnk_tc.cu: tcs_k32, tcs_steps16)."""
    nc = (n + 3) // 4
    return {"tf32": (((3 * nc + 2) & ~1) // 2, nc), "bf16x2": (0, (28 * nc + 21) // 16),
            "tf32_tile": (((nc + 6) & ~1) // 2, (8 * nc + 31) // 16), "tf32_dot": (((nc + 1) & ~1) // 2, (8 * nc + 15) // 16)}[split]


def sym64_tiles(n, rank, world):
    """Tiles of the panel schedule of tnk_ksd_sym64 on `rank` (units dealt round-robin)."""
    from tensornetworks_b200 import _lib
    T = (1 << n) // 128
    units = _lib.load().tnk_ksd_sym64_units(n)
    B = int(round(((8 * units + 1) ** 0.5 - 1) / 2))
    s = T // B
    k, tiles = 0, 0
    for pa in range(B):
        for pb in range(pa, B):
            if k % world == rank:
                tiles += s * (s + 1) // 2 if pa == pb else s * s
            k += 1
    return tiles


# ------------------------------------------------------------------ GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from tensornetworks_b200 import ops
    from tensornetworks_b200.synthetic import random_binary_bn, random_q

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
            os.environ.pop("NCCL_DEBUG")           # these levels print NCCL's version banner on stdout
        dist.init_process_group("nccl", device_id=device)
    group = dist.group.WORLD if world > 1 else None
    n, N = args.n, 1 << args.n
    pairs = 4.0 ** n
    peaks = load_measured_peaks()

    def ev_pair():
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); b.record()                      # materialise the CUDA handles
        return a, b

    def max_over_ranks(x):
        t = torch.tensor([float(x)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    bn, lat, obs, x = random_binary_bn(n, seed=args.seed)
    tables = ops.SteinTables.from_network(bn, x, lat, precision=args.precision, device=device)
    q_host = random_q(N, seed=args.seed).pin_memory()          # fp64, as ksd_vi.py:112 hands it over
    q_dev = q_host.to(device)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=device)  # > 126 MB L2
    r0, r1 = ops.shard_rows(N, rank, world)
    sym_ok = args.algorithm in ("auto", "pairs_sym") and 7 <= n <= 24 and args.precision in ("tc32", "fp64")
    if args.algorithm == "pairs_sym" and not sym_ok:
        raise SystemExit("--algorithm pairs_sym needs --precision tc32 or fp64 and 7 <= n <= 24")
    algorithm = {"tc32": "pairs_sym", "fp64": "pairs_sym64"}[args.precision] if sym_ok else "pairs"
    sym = algorithm != "pairs"
    sym_split = args.split or ops.DEFAULT_SYM_SPLIT
    if args.split:
        ops.DEFAULT_SYM_SPLIT = args.split           # the public op (e2e) follows

    exchange, exchange_name = None, "none (single GPU)"
    if world > 1:
        exchange_name = "one NCCL collective (all_reduce of partial u for the symmetric schedules, all_gather of row blocks otherwise)"
        if args.exchange == "fused":
            try:
                if sym:
                    exchange = ops.PeerComm(N + 1, device, group=group)
                    exchange_name = "tnk_comm (own barrier + reduce-scatter/all-gather kernel over CUDA-IPC peer buffers, no NCCL)"
                else:
                    exchange = ops.PeerExchange(N, device, group=group)
                    exchange_name = "tnk_ksd_rows_peers (rows stored into every rank's buffer by the reduction kernel)"
            except Exception as exc:      # no P2P mapping between these GPUs: keep the NCCL exchange, say so
                exchange, exchange_name = None, f"NCCL collective (peer buffers unavailable: {type(exc).__name__}: {exc})"

    def device_step():
        """One evaluation with q resident in HBM; returns (u [N], K [1]) complete on every rank."""
        if sym:
            out = (exchange.sendbuf[:N], exchange.sendbuf[N:N + 1]) if exchange is not None else None
            if algorithm == "pairs_sym":
                u, part = tables.ksd_sym(q_dev, *ops.sym_tile_stripe(rank, world), split=sym_split, out=out)   # row tiles rank, rank + world, ...
            else:
                u, part = tables.ksd_sym64(q_dev, rank, world, out=out)                        # units rank, rank + world, ...
            if exchange is not None:
                res = exchange.allreduce_sum(N + 1)
                return res[:N], res[N:]
            if world > 1:
                u, part = ops.combine_partial_sums(u, part, group)
            return u, part
        if exchange is not None:
            return ops.ksd_rows_fused(tables, q_dev, exchange)   # K3 + K4 with peer stores + barrier
        u, part = tables.ksd_rows(q_dev, r0, r1)
        if world > 1:
            u, part = ops.combine_row_shards(u[r0:r1], part, N, rank, world, group)
        return u, part

    def e2e_step():
        qh = q_host.requires_grad_(True)
        qh.grad = None
        K = ops.ksd_quadratic_form(qh, tables, group=group, exchange=exchange, algorithm=algorithm)   # H2D of q inside
        K.backward()                               # D2H of dK/dq inside
        return float(K.detach()), qh.grad

    # roofline denominators measured in this run: FMA-chain rate of the ALU kernels' arithmetic type
    probe_prec = "fp64" if args.precision == "fp64" else "fp32"
    peak_rate, _ = ops.alu_peak(probe_prec, iters=8192, device=device)

    sampler = ClockSampler(local_rank)
    sampler.start()                                 # nvidia-smi needs a moment to start: launch it before the warm-up
    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        device_step()
    barrier()

    # ---- timed region 1: inputs resident in HBM, per-step CUDA events, L2 flushed between steps
    ev = [ev_pair() for _ in range(args.steps)]
    kev = [ev_pair() for _ in range(args.steps)]
    ops.launch_counter["kernels"] = 0
    barrier()
    t_region_begin = time.time()
    for (e0, e1), (k0, k1) in zip(ev, kev):
        flush.zero_()
        ops.time_pair_kernel(k0, k1)
        e0.record()
        u, part = device_step()
        e1.record()
    barrier()
    ops.time_pair_kernel(None, None)
    launches = ops.launch_counter["kernels"]
    ms_per_step = max_over_ranks(sum(a.elapsed_time(b) for a, b in ev)) / args.steps
    kernel_ms = max_over_ranks(float(np.mean([a.elapsed_time(b) for a, b in kev])))
    value = pairs / (ms_per_step * 1e-3)
    u = u.clone()
    K_value = float(part.item())

    # ---- timed region 2: end to end through the public op with HOST q and HOST gradient
    for _ in range(3):
        K_e2e, grad_host = e2e_step()               # held like in the timed loop: two pinned gradient blocks stay in rotation
    barrier()
    t0 = time.perf_counter()
    e2e_ms = []
    for _ in range(args.steps):
        t_step = time.perf_counter()
        K_e2e, grad_host = e2e_step()
        e2e_ms.append((time.perf_counter() - t_step) * 1e3)
    barrier()
    e2e_value = pairs / (max_over_ranks(time.perf_counter() - t0) / args.steps)
    clocks = sampler.stop(t_region_begin, time.time())   # samples under load: device-timed and end-to-end regions
    clocks["window"] = "both timed regions (device-resident steps and end-to-end steps)"

    def time_kernel(fn, reps=2, warm=1):
        """Mean duration (ms, max over ranks) of the pair kernel inside fn(), L2 flushed before every call."""
        a, b = ev_pair()
        for _ in range(warm):
            fn()
        times = []
        for _ in range(reps):
            flush.zero_()
            ops.time_pair_kernel(a, b)
            out = fn()
            torch.cuda.synchronize()
            times.append(a.elapsed_time(b))
        ops.time_pair_kernel(None, None)
        return max_over_ranks(float(np.mean(times))), out

    ctx = {}
    if not args.no_context:
        # ---- the other operand formats of the symmetric tile kernel on this rank's stripe
        if algorithm == "pairs_sym":
            variants = {}
            for sp in ("tf32", "tf32_tile", "tf32_dot"):
                if sp == sym_split:
                    variants[sp] = kernel_ms
                    continue
                variants[sp], _ = time_kernel(lambda: tables.ksd_sym(q_dev, *ops.sym_tile_stripe(rank, world), split=sp))
                tables._ws.pop(("sym", rank, world, sp), None)
                tables._sym_tables.discard(("sym", rank, world, sp))
            ctx["variants"] = variants
            # ---- what bounds the headline kernel: its timing-only instances with parts of the pipeline compiled out
            # (tnk_tc.cu DIAG, n = 17..20, 6-step format; results are wrong, nothing here is reported as throughput)
            if sym_split == "tf32_dot" and 17 <= n <= 20 and not os.environ.get("TNK_TCS_ROWTILES"):
                limits = {}
                for name, flag in (("mma_only_ms", 13), ("read_no_fold_ms", 5), ("epilogue_only_ms", 16), ("epilogue_only_no_mirrored_ms", 17),
                                   ("readout_loop_only_ms", 21), ("no_mirrored_part_ms", 1)):
                    os.environ["TNK_TCS_ABLATE"] = str(flag)
                    limits[name], _ = time_kernel(lambda: tables.ksd_sym(q_dev, *ops.sym_tile_stripe(rank, world), split=sym_split))
                os.environ.pop("TNK_TCS_ABLATE", None)
                ctx["limits"] = limits
        # ---- the all-pairs tile kernel (tnk_ksd_rows, precision tc32) on this rank's rows
        if args.precision == "tc32" and sym:
            full_ms, (u_full, _) = time_kernel(lambda: tables.ksd_rows(q_dev, r0, r1))
            ctx["full_ms"] = full_ms
            ctx["full_err"] = float(((u_full[r0:r1] - u[r0:r1]).abs().max() / u.abs().max()).item())
            del u_full
            tables._ws.pop((r0, r1), None)
        # ---- the fp32 ALU pair kernel on this rank's rows
        if args.precision == "tc32":
            t_alu = ops.SteinTables(tables.P, n, precision="fp32")
            ctx["alu_ms"], _ = time_kernel(lambda: t_alu.ksd_rows(q_dev, r0, r1))
            ctx["alu_peak"] = peak_rate
            del t_alu
        # ---- the fp64 mode (1e-6 contract): symmetric DMMA kernel on this rank's units, all-pairs DFMA kernel on its rows
        if args.precision != "fp64" and 7 <= n <= 24:
            t64 = ops.SteinTables(tables.P, n, precision="fp64")
            ctx["fp64_peak"], _ = ops.alu_peak("fp64", iters=8192, device=device)
            ctx["fp64_sym_ms"], (u64p, k64p) = time_kernel(lambda: t64.ksd_sym64(q_dev, rank, world), reps=2, warm=1)
            if world > 1:
                u64p, k64p = ops.combine_partial_sums(u64p, k64p, group)
            ctx["fp64_u"], ctx["fp64_K"] = u64p.clone(), float(k64p.item())
            t64._ws.clear()
            ctx["fp64_rows_ms"], _ = time_kernel(lambda: t64.ksd_rows(q_dev, r0, r1), reps=1, warm=0)
            del t64
        # ---- the same K and gradient through the Kronecker (mode-product) factorisation: HBM-bound, O(n^2 2^n)
        if world == 1:
            m0, m1 = ev_pair()
            mp_times = {}
            for mp_storage in ("fp32", "fp64"):
                tables.ksd_modeproduct(q_dev, storage=mp_storage)   # builds the column-major score copies once
                times = []
                for _ in range(5):
                    flush.zero_()
                    m0.record()
                    u_mp, part_mp = tables.ksd_modeproduct(q_dev, storage=mp_storage)
                    m1.record()
                    torch.cuda.synchronize()
                    times.append(m0.elapsed_time(m1))
                mp_times[mp_storage] = float(np.mean(times))
            ctx["mp_times"], ctx["mp_K"] = mp_times, float(part_mp.item())
            ctx["mp_u"] = u_mp.clone()
            tables._ws = {k: v for k, v in tables._ws.items() if not (isinstance(k, tuple) and k and k[0] == "mp")}
            del u_mp
        # ---- enumeration kernels alone: K2 (score tables) and K1 (joint table), back to back between two events
        from tensornetworks_b200 import _lib as tnk_lib
        from tensornetworks_b200.flatten import flatten_bn
        lib = tnk_lib.load()
        en0, en1 = ev_pair()
        enum_reps = 8       # one call is ~60 us of GPU work, less than its host path
        t_tmp = ops.SteinTables(tables.P, n, precision="fp32", keep_scores=True)
        stream = torch.cuda.current_stream(device).cuda_stream
        enum_ms = []
        for _ in range(3):
            flush.zero_()
            en0.record()
            for _ in range(enum_reps):
                tnk_lib.check(lib.tnk_score(n, t_tmp.length_scale, t_tmp.P.data_ptr(), t_tmp.S.data_ptr(), t_tmp.prec_code,
                                            t_tmp.ST.data_ptr(), t_tmp.Cst.data_ptr(), stream), "tnk_score")
            en1.record()
            torch.cuda.synchronize()
            enum_ms.append(en0.elapsed_time(en1) / enum_reps)
        ctx["enum_ms"], ctx["enum_np"], ctx["enum_reps"] = min(enum_ms), t_tmp.np, enum_reps
        del t_tmp
        flat = flatten_bn(bn, x, lat)
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
        fk, fa, fpo, fco = dev(flat.node_kind), dev(flat.node_arg), dev(flat.pa_off), dev(flat.cpt_off)
        fpi = dev(flat.pa_idx) if flat.pa_idx.size else torch.zeros(1, dtype=torch.int32, device=device)
        fcpt = dev(flat.cpt)
        P_tmp = torch.empty(N, dtype=torch.float64, device=device)
        joint_ms = []
        for _ in range(3):
            flush.zero_()
            en0.record()
            for _ in range(enum_reps):
                tnk_lib.check(lib.tnk_bn_joint(flat.n_latent, flat.n_hidden, flat.n_nodes, fk.data_ptr(), fa.data_ptr(), fpo.data_ptr(),
                                               fpi.data_ptr(), fco.data_ptr(), fcpt.data_ptr(), P_tmp.data_ptr(), stream), "tnk_bn_joint")
            en1.record()
            torch.cuda.synchronize()
            joint_ms.append(en0.elapsed_time(en1) / enum_reps)
        ctx["joint_ms"], ctx["joint_nodes"] = min(joint_ms), int(flat.n_nodes)
        ctx["joint_exact"] = bool(torch.equal(P_tmp, tables.P))
        del P_tmp
        # ---- BASELINE.json config #3: n = 16 on one B200 (every pair kernel), reference CPU path timed beside it
        if world == 1 and n != 16:
            ctx["config3"] = bench_config3(ops, device, flush, time_kernel, args)
        # ---- BASELINE.json config #5: KSDVariationalInference.train epochs at n = 24 on 8 B200s
        if world == 8 and n == 20:
            tables._ws.clear()
            torch.cuda.empty_cache()
            ctx["config5"] = bench_config5(device, world, rank, max_over_ranks)
    return dict(world=world, rank=rank, n=n, pairs=pairs, value=value, ms_per_step=ms_per_step, kernel_ms=kernel_ms,
                K_value=K_value, u=u, e2e_value=e2e_value, e2e_ms=e2e_ms, step_ms=[a.elapsed_time(b) for a, b in ev], launches=launches, clocks=clocks, peak_rate=peak_rate,
                algorithm=algorithm, sym=sym, sym_split=sym_split, exchange_name=exchange_name, ctx=ctx, peaks=peaks,
                exchange=exchange)


def bench_config3(ops, device, flush, time_kernel, args):
    """configs[2]: synthetic random binary BN, 16 latent vars, exact KSD+grad, 1 B200 vs reference CPU."""
    import torch
    from oracle import c_oracle
    from tensornetworks_b200.synthetic import random_binary_bn, random_q
    n = 16
    N = 1 << n
    bn, lat, obs, x = random_binary_bn(n, seed=args.seed)
    q = random_q(N, seed=args.seed).to(device)
    out = {"workload": workload_name(n), "pairs_per_step": 4.0 ** n, "kernel_ms": {}, "value": {}}
    P, S, qh = cpu_tables(n, args.seed)
    Kref, gref = c_oracle.ksd_modeproduct(qh, P, n)
    errs = {}
    for name, precision, fn in (("fp64_pairs", "fp64", lambda t: t.ksd_rows(q)), ("fp64_pairs_sym64", "fp64", lambda t: t.ksd_sym64(q)),
                                ("fp32_pairs", "fp32", lambda t: t.ksd_rows(q)), ("tc32_pairs", "tc32", lambda t: t.ksd_rows(q)),
                                ("tc32_pairs_sym", "tc32", lambda t: t.ksd_sym(q))):
        t = ops.SteinTables.from_network(bn, x, lat, precision=precision, device=device)
        ms, (uu, kk) = time_kernel(lambda: fn(t), reps=3, warm=2)
        out["kernel_ms"][name] = ms
        out["value"][name] = 4.0 ** n / (ms * 1e-3)
        errs[name] = {"K_rel": abs(float(kk.item()) - Kref) / abs(Kref),
                      "u_max_rel": float(np.max(np.abs(uu.cpu().numpy() - gref / 2)) / np.max(np.abs(gref / 2)))}
        del t
    out["vs_cpu_oracle_all_rows"] = errs
    threads = c_oracle.max_threads()
    # the literal four-term loop over the full 4^16 pairs costs ~4.3e9 / (1e7 x threads) s: all rows when that stays
    # under ~25 s, else a stated sample of full rows
    info = cpu_sample(n, S, qh, min(25.0, max(args.cpu_seconds, 10.0)))
    out["cpu_baseline"] = {"value": info["rate"], "unit": UNIT, "cores": threads, "kind": "port",
                           "sample": (f"{info['rows']} of {N} full rows x 2^{n} columns = {info['pairs']:.3g} ordered pairs in "
                                      f"{info['seconds']:.1f} s (oracle/ksd_oracle.c, literal four-term pair loop)"),
                           "seconds_for_all_pairs_at_this_rate": 4.0 ** n / info["rate"]}
    best = min(out["kernel_ms"].values())
    out["gpu_vs_cpu"] = {"best_kernel_ms": best, "cpu_ms_all_pairs": 4.0 ** n / info["rate"] * 1e3,
                         "dense_reference_matrix": "OOM: 3 live 2^16 x 2^16 fp64 matrices = 103 GB (SURVEY 8d ii)"}
    return out


def bench_config5(device, world, rank, max_over_ranks):
    """configs[4]: full KSDVariationalInference.train epochs, n = 24, 8 B200s (pairs_sym and mode-product)."""
    import torch
    from tensornetworks_b200.ksd_vi import KSDVariationalInference
    from tensornetworks_b200.synthetic import random_binary_bn
    n = 24
    out = {"workload": f"KSDVariationalInference.train, synthetic random binary BN, {n} latent vars", "n_gpus": world}
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    for algorithm, precision, epochs in (("pairs_sym", "tc32", 2), ("modeproduct", "fp64", 5)):
        torch.manual_seed(0)
        vi = KSDVariationalInference(bn, lat, obs, {"use_logits": True, "conditioning_dim": 0}, device=str(device),
                                     precision=precision, compute_device=str(device), algorithm=algorithm, distributed=True)
        vi.train(x, num_epochs=1, lr_born_machine=0.05, verbose=False)          # warm-up epoch: tables, workspaces, comm
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        hist = vi.train(x, num_epochs=epochs, lr_born_machine=0.05, verbose=False)
        torch.cuda.synchronize()
        per_epoch = max_over_ranks((time.perf_counter() - t0) / epochs)
        out[algorithm] = {"precision": precision, "s_per_epoch": per_epoch, "epochs_timed": epochs,
                          "pairs_per_s": 4.0 ** n / per_epoch, "loss_ksd": hist["loss_ksd"][-1],
                          "note": "includes the per-call table rebuild of train() (score tables, column table)"}
        del vi
        torch.cuda.empty_cache()
    return out


def emit(line):
    """The one JSON line of the contract, written to the process's ORIGINAL stdout."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def tensor_peak(peaks):
    bf16s, bf16b = peaks.get("bf16_tflops_sustained"), peaks.get("bf16_tflops")
    sustained = (bf16s or bf16b or 1590.0) / 2.0
    burst = (bf16b or 1590.0) / 2.0
    src = ("half of MEASURED_PEAKS.json bf16_tflops_sustained (kind::tf32 runs at half the bf16 rate; the kernel is timed "
           "inside a long step)" if bf16s else "half of the fallback 1.59 PFLOP/s bf16 figure of B200_PROFILING.md")
    return sustained, burst, src


def main():
    global _REAL_STDOUT
    args = parse_args()
    # Library banners (NCCL, torchrun children) must not share stdout with the JSON line: keep a private
    # copy of stdout for emit() and point fd 1 at stderr for everything else.
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
        return
    r = run_ours(args)
    import torch.distributed as dist
    world, rank, n, ctx, peaks = r["world"], r["rank"], args.n, r["ctx"], r["peaks"]
    if rank == 0:
        pairs = r["pairs"]
        N = 1 << n
        w_lane = n + 10                                   # lane-ops per executed pair
        kernel_s = r["kernel_ms"] * 1e-3
        hbm = peaks.get("hbm_gbs") or 6650.0
        if r["algorithm"] == "pairs_sym":
            tiles = sym_tiles(n, 0, world)
            s32, s16 = sym_steps(n, r["sym_split"])
            flops_tile = 128 * 128 * 2 * 8 * (s32 + s16)
            # row tiles sharing one column-operand stage (tnk_tc.cu launch_sym): 4 when their A operands fit in TMEM
            row_tiles_per_cta = 4 if 256 + 4 * 8 * (s32 + s16) <= 512 else 2
            if os.environ.get("TNK_TCS_ROWTILES") in ("1", "2"):
                row_tiles_per_cta = int(os.environ["TNK_TCS_ROWTILES"])
            peak_tf, burst_tf, src = tensor_peak(peaks)
            ach_tf = tiles * flops_tile / kernel_s / 1e12
            roofline = {"bound": "tensor", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s",
                        "frac": ach_tf / peak_tf, "traffic": None,
                        "kernel": "tcs_tiles_kernel" if row_tiles_per_cta == 1 else "tcs_tilesN_kernel",
                        "row_tiles_per_cta": row_tiles_per_cta,
                        "kernel_ms": r["kernel_ms"], "executed_tiles_128x128": tiles,
                        "executed_pairs_per_s": tiles * 16384.0 / kernel_s,
                        "ordered_pairs_per_executed_pair": (pairs / world) / (tiles * 16384.0),
                        "k_steps_tf32": s32, "k_steps_bf16": s16, "peak_source": src,
                        "frac_of_burst_peak": ach_tf / burst_tf, "burst_peak": burst_tf,
                        "operand_format": r["sym_split"],
                        "note": ("format tf32_dot keeps only the rank-n score dot product on the MMAs (3 tf32 + 3 bf16 K-steps: the 3xTF32 "
                                 "split of ceil(n/8) algorithmic K-steps); C(z)+C(z') and the T1/T2 terms are added exactly by n+2 fp64 "
                                 "Kronecker transforms before the tile kernel. With 6 K-steps the kernel is bound by its epilogue - the fp32 fold "
                                 "of every accumulator element into the row and mirrored column sums on the ALU (kernel_limits) - not by the "
                                 "MMA rate or the TMEM read-out: see tile_kernel_variants for the 13- and 9-step formats")}
            if roofline["frac"] > 1.0:
                roofline["frac_note"] = "above 1: the sustained peak is a matmul-loop average; see frac_of_burst_peak"
            # second limit of the same kernel: every accumulator tile (64 KB of fp32) leaves TMEM once through
            # tcgen05.ld.16x256b.x8: 1024 B/clk/SM measured by scripts/probes/tmem_ld_rate.cu (profiles/r02_tmem_ld_rate_probe.log)
            sm_mhz = (r["clocks"] or {}).get("sm_mhz")
            if sm_mhz:
                tmem_peak = 148 * 1024.0 * sm_mhz * 1e6                    # B/s at the clock measured under load
                tmem_ach = tiles * 65536.0 / kernel_s
                roofline["binding_limit"] = "epilogue (ALU fold of the accumulators)" if sum(sym_steps(n, r["sym_split"])) <= 6 else "tensor"
                roofline["tmem_readout"] = {"achieved_TB/s": tmem_ach / 1e12, "peak_TB/s": tmem_peak / 1e12,
                                            "frac": tmem_ach / tmem_peak, "bytes_per_tile": 65536,
                                            "peak_source": "148 SMs x 1024 B/clk (probe, .16x256b.x8) x median SM clock of this run"}
            tag = "tcs_tiles" if row_tiles_per_cta == 1 else "tcs_tilesN"
        elif r["algorithm"] == "pairs_sym64":
            tiles = sym64_tiles(n, 0, world)
            steps = (n + 3) // 4
            lane_pair = 4 * steps + 2.75
            achieved = tiles * 16384.0 * lane_pair / kernel_s
            roofline = {"bound": "alu", "achieved": achieved / 1e12, "peak": r["peak_rate"] / 1e12, "unit": "Tlane-op/s",
                        "frac": achieved / r["peak_rate"], "traffic": None, "kernel": "sym64_kernel",
                        "kernel_ms": r["kernel_ms"], "executed_tiles_128x128": tiles,
                        "executed_pairs_per_s": tiles * 16384.0 / kernel_s,
                        "lane_ops_per_executed_pair": lane_pair,
                        "lane_ops_note": f"{steps} DMMA K-steps x 4 (score dot product) + 0.75 DMUL + 2 DFMA; DMMA and DFMA share one fp64 pipe (ncu pipe_shared); "
                                         "C(z)+C(z') and the T1/T2 terms come from n+2 fp64 Kronecker transforms outside the pair kernel",
                        "frac_at_survey_w": tiles * 16384.0 * w_lane / kernel_s / r["peak_rate"],
                        "peak_source": "tnk_alu_peak DFMA-chain probe measured in this run (MEASURED_PEAKS.json has no fp64 peak)"}
            tag = "sym64"
        elif args.precision == "tc32":
            from tensornetworks_b200 import ops as _ops
            plan = _ops.tc_plan(n)
            executed = pairs / world
            flops_pair = 2 * 8 * (plan["nsteps"] + plan["nsteps16"])
            peak_tf, burst_tf, src = tensor_peak(peaks)
            ach_tf = executed * flops_pair / kernel_s / 1e12
            roofline = {"bound": "tensor", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s",
                        "frac": ach_tf / peak_tf, "traffic": None, "kernel": "tc_rows_kernel",
                        "kernel_ms": r["kernel_ms"], "mma_flops_per_pair": flops_pair, "executed_pairs_per_s": executed / kernel_s,
                        "k_steps_tf32": plan["nsteps"], "k_steps_bf16": plan["nsteps16"], "peak_source": src,
                        "frac_of_burst_peak": ach_tf / burst_tf, "burst_peak": burst_tf}
            tag = "tc_rows"
        else:
            executed = pairs / world                          # this kernel evaluates every ordered pair of its rows
            achieved = executed * w_lane / kernel_s
            roofline = {"bound": "alu", "achieved": achieved / 1e12, "peak": r["peak_rate"] / 1e12,
                        "unit": "Tlane-op/s", "frac": achieved / r["peak_rate"], "traffic": None,
                        "kernel": "ksd_rows_kernel", "kernel_ms": r["kernel_ms"], "executed_pairs_per_s": executed / kernel_s,
                        "lane_ops_per_pair": w_lane,
                        "peak_source": "tnk_alu_peak FMA-chain probe measured in this run (MEASURED_PEAKS.json has no ALU peak)"}
            tag = "ksd_rows"
        traffic = ncu_traffic(tag, n) if world == 1 else None
        if traffic:
            roofline["traffic"], roofline["traffic_source"] = traffic[0], f"profiles/{traffic[1]} (bytes per launch)"
        line = {
            "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": {"fp32": "f32", "fp64": "f64", "tc32": "tf32x3"}[args.precision], "data": "synthetic",
            "config": config_of(n),
            "run": {"precision_mode": {"fp32": "fp32-compute/fp64-accumulate", "fp64": "fp64",
                                       "tc32": "3xTF32 tensor-core tiles (fp32 accumulate per tile, fp64 across tiles)"}[args.precision],
                    "algorithm": r["algorithm"] + (f" (operand format {r['sym_split']})" if r["algorithm"] == "pairs_sym" else ""),
                    "parallelism": ((f"row tiles / units dealt round-robin x{world}" if r["sym"] else f"row-sharded x{world}")
                                    if world > 1 else "single GPU"),
                    "exchange": r["exchange_name"],
                    "l2": "256 MB flush written between timed steps",
                    "executed_pairs": ("N(N+128)/2: every unordered pair of 128-state tiles evaluated once, the symmetry "
                                       "k_p(z,z') = k_p(z',z) supplies the mirrored half (SURVEY.md 8d); value = 4^n / t"
                                       if r["sym"] else "4^n (symmetry not exploited)")},
            "executed_pairs_per_s": roofline.get("executed_pairs_per_s"),
            "roofline": roofline,
            "e2e": {"value": r["e2e_value"], "unit": UNIT, "h2d_bytes_per_step": 8 * N * world,
                    "d2h_bytes_per_step": (8 * N + 8) * world, "frac_of_device_timed": r["e2e_value"] / r["value"],
                    "ms_per_step_rank0": r["e2e_ms"], "device_timed_ms_per_step_rank0": r["step_ms"]},
            "gpu_launches": r["launches"], "clocks": r["clocks"], "ksd_sq": r["K_value"],
        }
        if "variants" in ctx:
            tiles_v = sym_tiles(n, 0, world)
            pk, _, _ = tensor_peak(peaks)
            line["tile_kernel_variants"] = {
                sp: {"kernel_ms": ms, "k_steps": sum(sym_steps(n, sp)), "value": pairs / (ms * 1e-3),
                     "issued_TFLOP/s": tiles_v * 128 * 128 * 16 * sum(sym_steps(n, sp)) / (ms * 1e-3) / 1e12,
                     "frac_of_tensor_peak": tiles_v * 128 * 128 * 16 * sum(sym_steps(n, sp)) / (ms * 1e-3) / 1e12 / pk}
                for sp, ms in ctx["variants"].items()}
            line["tile_kernel_variants"]["note"] = ("tf32: whole bracket on the MMAs; tf32_tile: T1/T2 terms of bits >= 7 through the "
                                                    "Kronecker pass; tf32_dot: only the score dot product on the MMAs (headline)")
        if r["algorithm"] == "pairs_sym":
            # the epilogue's own roofline: 3 fp32 lane-operations per executed pair (weight, row sum, mirrored sum) against the
            # FMA rate measured in this run (packed FFMA2 occupies the FMA pipe for two cycles: no gain over FFMA in lane-ops)
            fold_ops = line["roofline"]["executed_tiles_128x128"] * 16384.0 * 3.0
            line["roofline"]["epilogue_alu"] = {
                "lane_ops_per_executed_pair": 3, "achieved": fold_ops / (r["kernel_ms"] * 1e-3) / 1e12, "peak": r["peak_rate"] / 1e12,
                "unit": "Tlane-op/s", "frac": fold_ops / (r["kernel_ms"] * 1e-3) / r["peak_rate"],
                "peak_source": "tnk_alu_peak FFMA2-chain probe measured in this run"}
        if "limits" in ctx:
            lim = ctx["limits"]
            line["roofline"]["kernel_limits"] = dict(lim, note=(
                "the same kernel with parts compiled out (timing only): MMAs alone / accumulators read from TMEM but not folded / "
                "epilogue alone without MMAs / full kernel without the mirrored column sums. The read-out rides on the MMA time, the "
                "epilogue alone takes about as long as the whole kernel, and it is the sum of the read-out loop alone and the fold "
                "(epilogue_only_no_mirrored = readout_loop_only + fold at the FMA-pipe floor): TMEM read-out and fp32 fold of a "
                "scheduler do not overlap; the tensor pipe waits for them"),
                mma_only_frac_of_tensor_peak=line["roofline"]["achieved"] * r["kernel_ms"] / lim["mma_only_ms"] / line["roofline"]["peak"],
                mma_only_frac_note="above 1 is possible: the sustained peak is a long matmul-loop average under power limits, "
                                   "the MMA-only instance draws less power than a full GEMM")
        if "full_ms" in ctx:
            line["full_tile_kernel"] = {"kernel": "tc_rows_kernel (all 4^n ordered pairs, tnk_ksd_rows precision tc32)",
                                        "kernel_ms": ctx["full_ms"], "value": pairs / (ctx["full_ms"] * 1e-3), "unit": UNIT,
                                        "u_vs_headline_kernel_max_rel": ctx["full_err"]}
        if "alu_ms" in ctx:
            alu_s = ctx["alu_ms"] * 1e-3
            alu_ach = pairs / world * w_lane / alu_s
            line["alu_kernel"] = {"kernel": "ksd_rows_kernel (fp32 FFMA2 pair kernel, precision fp32)", "kernel_ms": ctx["alu_ms"],
                                  "value": pairs / alu_s, "unit": UNIT, "bound": "alu", "achieved": alu_ach / 1e12,
                                  "peak": ctx["alu_peak"] / 1e12, "roofline_unit": "Tlane-op/s",
                                  "frac": alu_ach / ctx["alu_peak"], "lane_ops_per_pair": w_lane,
                                  "peak_source": "tnk_alu_peak FMA-chain probe measured in this run"}
        if "fp64_sym_ms" in ctx:
            tiles64 = sym64_tiles(n, 0, world)
            steps = (n + 3) // 4
            lane_pair = 4 * steps + 2.75
            s64, r64 = ctx["fp64_sym_ms"] * 1e-3, ctx["fp64_rows_ms"] * 1e-3
            line["fp64_kernel"] = {
                "mode": "fp64 (the north_star's 1e-6 mode)", "kernel": "sym64_kernel (tnk_ksd_sym64: symmetric tiles, DMMA score dot product, deterministic)",
                "kernel_ms": ctx["fp64_sym_ms"], "value": pairs / s64, "unit": UNIT, "executed_pairs_per_s": tiles64 * 16384.0 / s64,
                "bound": "alu", "achieved": tiles64 * 16384.0 * lane_pair / s64 / 1e12, "peak": ctx["fp64_peak"] / 1e12,
                "roofline_unit": "Tlane-op/s", "frac": tiles64 * 16384.0 * lane_pair / s64 / ctx["fp64_peak"],
                "lane_ops_per_executed_pair": lane_pair, "frac_at_survey_w": tiles64 * 16384.0 * w_lane / s64 / ctx["fp64_peak"],
                "peak_source": "tnk_alu_peak DFMA-chain probe measured in this run",
                "all_pairs_kernel": {"kernel": "ksd_rows_kernel<double> (tnk_ksd_rows, all 4^n ordered pairs)", "kernel_ms": ctx["fp64_rows_ms"],
                                     "value": pairs / r64, "frac": pairs / world * w_lane / r64 / ctx["fp64_peak"]},
                "ksd_sq": ctx["fp64_K"]}
        if "mp_times" in ctx:
            # S twice, n+2 vectors written and read once, q twice, r_sum, u - all fp64 in this variant
            mp_ms = ctx["mp_times"]["fp64"]
            mp_bytes = 8.0 * N * (2 * n + 2 * (n + 2) + 4)
            mp_gbs = mp_bytes / (mp_ms * 1e-3) / 1e9
            line["modeproduct"] = {"kernels": "mp_first_kernel + mp_last_kernel (fp64 arithmetic, O(n^2 2^n), same K and gradient)",
                                   "storage": "fp64", "ms_fp32_storage": ctx["mp_times"]["fp32"],
                                   "ms": mp_ms, "effective_value": pairs / (mp_ms * 1e-3), "unit": UNIT,
                                   "bound": "hbm", "algorithmic_bytes": mp_bytes, "achieved": mp_gbs, "peak": hbm,
                                   "roofline_unit": "GB/s", "frac": mp_gbs / hbm, "ksd_sq": ctx["mp_K"]}
        if "enum_ms" in ctx:
            enum_bytes = 8.0 * N * (n + 1) + 4.0 * N * (ctx["enum_np"] + 1)   # read P, write S64, ST, Cst
            line["enumeration"] = {"score_kernel_ms": ctx["enum_ms"], "score_kernel_algorithmic_bytes": enum_bytes,
                                   "score_kernel_GB/s": enum_bytes / (ctx["enum_ms"] * 1e-3) / 1e9, "peak_GB/s": hbm,
                                   "score_kernel_frac": enum_bytes / (ctx["enum_ms"] * 1e-3) / 1e9 / hbm,
                                   "timing": f"{ctx['enum_reps']} back-to-back calls between two events, L2 flushed before",
                                   "joint_kernel_ms": ctx["joint_ms"], "joint_kernel_algorithmic_bytes": 8.0 * N,
                                   "joint_kernel_GB/s": 8.0 * N / (ctx["joint_ms"] * 1e-3) / 1e9,
                                   "joint_kernel_frac": 8.0 * N / (ctx["joint_ms"] * 1e-3) / 1e9 / hbm,
                                   "joint_kernel_note": f"{ctx['joint_nodes']} CPT gathers + IEEE multiplies per state in bn.nodes order (bit-exact): "
                                                        "dependent-chain bound, not HBM bound", "joint_bit_exact_rerun": ctx["joint_exact"]}
        if world == 1:
            from oracle import c_oracle
            P, S, q = cpu_tables(n, args.seed)
            # parity at the metric's own size: ALL rows and K against the C mode-product oracle
            t0 = time.perf_counter()
            Kref, gref = c_oracle.ksd_modeproduct(q, P, n)
            oracle_s = time.perf_counter() - t0
            uref, uscale = gref / 2.0, float(np.max(np.abs(gref / 2.0)))
            par = {"oracle": "oracle/ksd_oracle.c:orc_ksd_modeproduct, all 2^n rows", "oracle_seconds": oracle_s,
                   "headline": {"K_rel": abs(r["K_value"] - Kref) / abs(Kref),
                                "u_max_rel": float(np.max(np.abs(r["u"].cpu().numpy() - uref))) / uscale}}
            if "fp64_u" in ctx:
                par["fp64_kernel"] = {"K_rel": abs(ctx["fp64_K"] - Kref) / abs(Kref),
                                      "u_max_rel": float(np.max(np.abs(ctx["fp64_u"].cpu().numpy() - uref))) / uscale}
            if "mp_u" in ctx:
                par["modeproduct"] = {"K_rel": abs(ctx["mp_K"] - Kref) / abs(Kref),
                                      "u_max_rel": float(np.max(np.abs(ctx["mp_u"].cpu().numpy() - uref))) / uscale}
            # the timed CPU baseline: the literal four-term loop on rows of the first and of the last row tile
            half = args.cpu_seconds / 2.0
            first = cpu_sample(n, S, q, half, row_start=0)
            last = cpu_sample(n, S, q, half, row_start=N - first["rows"])      # ends inside the last row tile
            rate = (first["pairs"] + last["pairs"]) / (first["seconds"] + last["seconds"])
            lit_err = max(float(np.max(np.abs(first["u"] - uref[:first["rows"]]))),
                          float(np.max(np.abs(last["u"] - uref[last["row_start"]:last["row_start"] + last["rows"]])))) / uscale
            par["oracle_vs_literal_loop_rows_max_rel"] = lit_err
            line["parity"] = par
            line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": first["threads"], "kind": "port",
                                    "sample": f"{first['rows']} full rows of the first row tile + {last['rows']} of the last x 2^{n} columns = "
                                              f"{first['pairs'] + last['pairs']:.3g} ordered pairs in {first['seconds'] + last['seconds']:.1f} s "
                                              f"(oracle/ksd_oracle.c, literal four-term pair loop)",
                                    "gpu_vs_oracle_rows_max_rel_err": par["headline"]["u_max_rel"]}
        if "config3" in ctx:
            line["config3_n16"] = ctx["config3"]
        if "config5" in ctx:
            line["config5_n24_train"] = ctx["config5"]
        emit(line)
    if r["exchange"] is not None and hasattr(r["exchange"], "close"):
        r["exchange"].close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
