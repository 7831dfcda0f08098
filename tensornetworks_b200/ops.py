"""Batched entry points of the exact-KSD path on top of the C-ABI (include/tnk.h).

  joint_table      2^n calls of compute_prob_joint_xz          (reference stein_utils.py:58-112)
  SteinTables      _precompute_all_s_p                         (reference ksd_vi.py:55-60)
  ksd_quadratic_form   the O(4^n) loop + autograd backward     (reference ksd_vi.py:110-131,145)

All arithmetic runs in libtnk_b200.so on the current CUDA device; torch only owns the
memory, the stream and (for several GPUs) the process group.  No CPU fallback exists.
"""
import ctypes

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .flatten import flatten_bn

PRECISIONS = {"fp64": _lib.PREC_FP64, "fp32": _lib.PREC_FP32, "fp64_simple": _lib.PREC_FP64_SIMPLE,
              "tc32": _lib.PREC_TC32}
DEFAULT_SYM_SPLIT = "tf32_dot"     # operand format of algorithm "pairs_sym" (SteinTables.ksd_sym)
launch_counter = {"kernels": 0}  # kernels of ours enqueued since last reset (bench.py's gpu_launches)


def _require_cuda(device):
    device = torch.device(device)
    if device.type != "cuda" or not torch.cuda.is_available():
        raise RuntimeError("the exact-KSD path runs on a CUDA device only (B200, sm_100a); there is no CPU fallback")
    return device


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def joint_table(bn, x_dict, latent_vars_names, device="cuda", n_latent=None):
    """p(x,z) for every latent state, fp64 tensor [2^n] on `device`, enumeration order
    (bit-exact with compute_prob_joint_xz, reference stein_utils.py:58-112)."""
    device = _require_cuda(device)
    flat = flatten_bn(bn, x_dict, latent_vars_names, n_latent=n_latent)
    lib = _lib.load()
    with torch.cuda.device(device):
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
        kind, arg, pa_off, cpt_off = dev(flat.node_kind), dev(flat.node_arg), dev(flat.pa_off), dev(flat.cpt_off)
        pa_idx = dev(flat.pa_idx) if flat.pa_idx.size else torch.zeros(1, dtype=torch.int32, device=device)
        cpt = dev(flat.cpt) if flat.cpt.size else torch.zeros((1, 2), dtype=torch.float64, device=device)
        P = torch.empty(1 << flat.n_latent, dtype=torch.float64, device=device)
        _lib.check(lib.tnk_bn_joint(flat.n_latent, flat.n_hidden, flat.n_nodes, _ptr(kind), _ptr(arg), _ptr(pa_off),
                                    _ptr(pa_idx), _ptr(cpt_off), _ptr(cpt), _ptr(P), _stream_ptr(device)), "tnk_bn_joint")
        launch_counter["kernels"] += 1
        if np.isnan(flat.cpt).any() and bool(torch.isnan(P).any()):
            raise ValueError("CPT entry for a node/parent configuration that the enumeration reaches was not found.")
    return P


class SteinTables:
    """Device-resident, per-(network, evidence) operands of the pair kernel.

    P    fp64 [N]        joint table p(x,z)
    S    fp64 [N, n]     score table in latent-variable order (None unless keep_scores)
    ST   [n_pad, np]     packed scores by bit position in the compute dtype
    Cst  [n_pad]         alpha*R(z) + gamma*rho0(z)
    """

    def __init__(self, P, n, length_scale=1.0, precision="fp32", keep_scores=False):
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
        device = _require_cuda(P.device)
        if P.dtype != torch.float64 or P.numel() != (1 << n):
            raise ValueError("P must be a float64 tensor with 2^n entries")
        self.n, self.length_scale, self.precision = int(n), float(length_scale), precision
        self.prec_code = PRECISIONS[precision]
        self.device, self.N = device, 1 << n
        self.P = P.contiguous()
        lib = _lib.load()
        np_, n_pad, elem = _lib.layout(self.n, self.prec_code)
        dtype = torch.float32 if elem == 4 else torch.float64
        self.np, self.n_pad, self.dtype = np_, n_pad, dtype
        with torch.cuda.device(device):
            self.S = torch.empty((self.N, self.n), dtype=torch.float64, device=device) if keep_scores else None
            self.ST = torch.empty((n_pad, np_), dtype=dtype, device=device)
            self.Cst = torch.empty(n_pad, dtype=dtype, device=device)
            _lib.check(lib.tnk_score(self.n, self.length_scale, _ptr(self.P), _ptr(self.S), self.prec_code,
                                     _ptr(self.ST), _ptr(self.Cst), _stream_ptr(device)), "tnk_score")
            launch_counter["kernels"] += 1
        self._ws = {}

    @classmethod
    def from_network(cls, bn, x_dict, latent_vars_names, length_scale=1.0, precision="fp32",
                     device="cuda", keep_scores=False, n_latent=None):
        n = len(latent_vars_names) if n_latent is None else int(n_latent)
        P = joint_table(bn, x_dict, latent_vars_names, device=device, n_latent=n)
        return cls(P, n, length_scale=length_scale, precision=precision, keep_scores=keep_scores)

    def scores(self):
        """fp64 [N, n] score table in latent-variable order (what get_score_function_sp_for_z
        returns per state); computed on first use when the tables were built without it."""
        if self.S is None:
            with torch.cuda.device(self.device):
                self.S = torch.empty((self.N, self.n), dtype=torch.float64, device=self.device)
                _lib.check(_lib.load().tnk_score(self.n, self.length_scale, _ptr(self.P), _ptr(self.S), self.prec_code,
                                                 None, None, _stream_ptr(self.device)), "tnk_score")
                launch_counter["kernels"] += 1
        return self.S

    def posterior(self):
        """p(z|x) = P / sum(P) as an fp64 device tensor (tensor form of get_true_posterior)."""
        return self.P / self.P.sum()

    def workspace(self, row_begin, row_end):
        key = (row_begin, row_end)
        if key not in self._ws:
            nbytes = _lib.load().tnk_ksd_workspace_bytes(self.n, self.prec_code, row_begin, row_end)
            if nbytes == 0:
                raise RuntimeError("tnk_ksd_workspace_bytes rejected the row range")
            self._ws[key] = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        return self._ws[key]

    def ksd_rows(self, q, row_begin=0, row_end=None, u=None):
        """u[row_begin:row_end] = (K_p q)[rows] and the partial sum_i q_i u_i over those rows.
        q: fp32/fp64 CUDA tensor [N].  Returns (u [N] fp64 - only the row range is written, part [1] fp64)."""
        row_end = self.N if row_end is None else row_end
        if q.device != self.device or q.dim() != 1 or q.numel() != self.N:
            raise ValueError(f"q must be a 1-D tensor of {self.N} entries on {self.device}")
        if q.dtype not in (torch.float32, torch.float64):
            raise ValueError("q must be float32 or float64")
        q = q.contiguous()
        lib = _lib.load()
        with torch.cuda.device(self.device):
            if u is None:
                u = torch.empty(self.N, dtype=torch.float64, device=self.device)
            part = torch.empty(1, dtype=torch.float64, device=self.device)
            ws = self.workspace(row_begin, row_end)
            launched = ctypes.c_int(0)
            _lib.check(lib.tnk_ksd_rows(self.n, self.length_scale, self.prec_code, _ptr(q),
                                        0 if q.dtype == torch.float64 else 1, _ptr(self.ST), _ptr(self.Cst),
                                        row_begin, row_end, _ptr(u), _ptr(part), _ptr(ws), ws.numel(),
                                        ctypes.byref(launched), _stream_ptr(self.device)), "tnk_ksd_rows")
            launch_counter["kernels"] += launched.value
        return u, part


    def ksd_sym(self, q, tile_offset=0, tile_stride=1, split=None, out=None):
        """Symmetric tensor-core evaluation (tnk_ksd_sym): every unordered pair of 128-state tiles is evaluated
        once and feeds both its row tile and, mirrored, its column tile.  This call owns the row tiles
        tile_offset + k*tile_stride; returns (u_partial [N] fp64, part [1] fp64) whose sums over all offsets of
        one stride are u = K_p q and KSD^2.  Needs the fp32 tables (precision "tc32" or "fp32"), 7 <= n <= 24.
        split: operand format of the tile kernel (default DEFAULT_SYM_SPLIT = "tf32_dot"):
          "tf32"      the whole bracket on the MMAs, tf32 high parts + bf16 cross terms (13 K-steps at n = 20, errors ~1e-6)
          "bf16x2"    the whole bracket, two bf16 pieces per factor (10 K-steps, errors ~1e-5)
          "tf32_tile" the T1/T2 terms of the bit positions >= 7 leave the pair kernel (9 K-steps)
          "tf32_dot"  only the score dot product stays in the pair kernel (6 K-steps); for both, the removed terms are
                      added for all rows by a pass of fp64 Kronecker transforms (O(n^2 2^n); the calls of one stride split it between them).
        out: optional (u [N], part [1]) fp64 device tensors to write into (e.g. views of PeerComm.sendbuf)."""
        split = DEFAULT_SYM_SPLIT if split is None else split
        if split not in _lib.SYM_SPLIT:
            raise ValueError(f"split must be one of {tuple(_lib.SYM_SPLIT)}")
        code = _lib.SYM_SPLIT[split]
        if self.dtype != torch.float32:
            raise ValueError('the symmetric tile kernel reads the fp32 tables: build SteinTables with precision "tc32"')
        if q.device != self.device or q.dim() != 1 or q.numel() != self.N:
            raise ValueError(f"q must be a 1-D tensor of {self.N} entries on {self.device}")
        if q.dtype not in (torch.float32, torch.float64):
            raise ValueError("q must be float32 or float64")
        q = q.contiguous()
        lib = _lib.load()
        with torch.cuda.device(self.device):
            key = ("sym", tile_offset, tile_stride, split)
            built = getattr(self, "_sym_tables", None)
            if built is None:
                built = self._sym_tables = set()      # workspaces whose column table a successful call has built
            reuse = key in built and key in self._ws
            if key not in self._ws:
                nbytes = lib.tnk_ksd_sym_workspace_bytes(self.n, code, tile_offset, tile_stride)
                if nbytes == 0:
                    raise RuntimeError(f"tnk_ksd_sym_workspace_bytes rejected n={self.n}, tiles ({tile_offset}, {tile_stride})")
                self._ws[key] = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            ws = self._ws[key]
            u, part = out if out is not None else (torch.empty(self.N, dtype=torch.float64, device=self.device),
                                                   torch.empty(1, dtype=torch.float64, device=self.device))
            launched = ctypes.c_int(0)
            _lib.check(lib.tnk_ksd_sym(self.n, self.length_scale, code, _ptr(q), 0 if q.dtype == torch.float64 else 1,
                                       _ptr(self.ST), _ptr(self.Cst), tile_offset, tile_stride, 1 if reuse else 0,
                                       _ptr(u), _ptr(part), _ptr(ws), ws.numel(), ctypes.byref(launched),
                                       _stream_ptr(self.device)), "tnk_ksd_sym")
            built.add(key)
            launch_counter["kernels"] += launched.value
        return u, part

    def ksd_sym64(self, q, unit_offset=0, unit_stride=1, out=None):
        """Symmetric fp64 evaluation (tnk_ksd_sym64): every unordered pair of 128-state tiles once, fp64 throughout
        (the 1e-6 mode), bit-reproducible.  This call owns the work units unit_offset + k*unit_stride; returns
        (u_partial [N] fp64, part [1] fp64) whose sums over all offsets of one stride are u = K_p q and KSD^2.
        Needs the fp64 tables (precision "fp64") and 7 <= n <= 24."""
        if self.dtype != torch.float64:
            raise ValueError('the symmetric fp64 kernel reads the fp64 tables: build SteinTables with precision "fp64"')
        if q.device != self.device or q.dim() != 1 or q.numel() != self.N:
            raise ValueError(f"q must be a 1-D tensor of {self.N} entries on {self.device}")
        if q.dtype not in (torch.float32, torch.float64):
            raise ValueError("q must be float32 or float64")
        q = q.contiguous()
        lib = _lib.load()
        with torch.cuda.device(self.device):
            key = ("sym64", unit_offset, unit_stride)
            if key not in self._ws:
                nbytes = lib.tnk_ksd_sym64_workspace_bytes(self.n, unit_offset, unit_stride)
                if nbytes == 0:
                    raise RuntimeError(f"tnk_ksd_sym64_workspace_bytes rejected n={self.n}, units ({unit_offset}, {unit_stride})")
                self._ws[key] = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            ws = self._ws[key]
            u, part = out if out is not None else (torch.empty(self.N, dtype=torch.float64, device=self.device),
                                                   torch.empty(1, dtype=torch.float64, device=self.device))
            launched = ctypes.c_int(0)
            _lib.check(lib.tnk_ksd_sym64(self.n, self.length_scale, _ptr(q), 0 if q.dtype == torch.float64 else 1,
                                         _ptr(self.ST), _ptr(self.Cst), unit_offset, unit_stride, _ptr(u), _ptr(part),
                                         _ptr(ws), ws.numel(), ctypes.byref(launched), _stream_ptr(self.device)),
                       "tnk_ksd_sym64")
            launch_counter["kernels"] += launched.value
        return u, part

    def _st64(self):
        """fp64 packed score table [n_pad, np] (bit order) for the mode-product path; the fp64-mode
        tables already hold it, fp32-mode tables build it on first use."""
        if self.dtype == torch.float64:
            return self.ST
        if getattr(self, "_ST64", None) is None:
            np_, n_pad, _ = _lib.layout(self.n, _lib.PREC_FP64)
            with torch.cuda.device(self.device):
                self._ST64 = torch.empty((n_pad, np_), dtype=torch.float64, device=self.device)
                cst = torch.empty(n_pad, dtype=torch.float64, device=self.device)
                _lib.check(_lib.load().tnk_score(self.n, self.length_scale, _ptr(self.P), None, _lib.PREC_FP64,
                                                 _ptr(self._ST64), _ptr(cst), _stream_ptr(self.device)), "tnk_score")
                launch_counter["kernels"] += 1
        return self._ST64

    def ksd_modeproduct(self, q, bit_begin=0, bit_end=None, storage=None):
        """Partial of u = K_p q over score components (bit positions) [bit_begin, bit_end) for all rows,
        and the matching partial of q.u, through the Kronecker factorisation (O(n^2 2^n), fp64 arithmetic).
        Partials over disjoint bit ranges add up to the pair kernel's result.
        storage: "fp64" keeps the score columns and the transformed vectors in fp64 in HBM (results to ~1e-15
        of the fp64 pair kernel), "fp32" in fp32 (half the traffic, errors of the fp32 pair kernel's size);
        default: "fp64" for fp64-mode tables, "fp32" for fp32/tc32-mode tables."""
        bit_end = self.n if bit_end is None else bit_end
        if storage is None:
            storage = "fp64" if self.dtype == torch.float64 else "fp32"
        if storage not in _lib.MP_STORE:
            raise ValueError(f"storage must be one of {tuple(_lib.MP_STORE)}")
        code = _lib.MP_STORE[storage]
        if q.device != self.device or q.dim() != 1 or q.numel() != self.N:
            raise ValueError(f"q must be a 1-D tensor of {self.N} entries on {self.device}")
        if q.dtype not in (torch.float32, torch.float64):
            raise ValueError("q must be float32 or float64")
        q = q.contiguous()
        lib = _lib.load()
        with torch.cuda.device(self.device):
            key = ("mp", bit_begin, bit_end, storage)
            if key not in self._ws:
                nbytes = lib.tnk_ksd_modeproduct_workspace_bytes(self.n, bit_begin, bit_end, code)
                if nbytes == 0:
                    raise RuntimeError("tnk_ksd_modeproduct_workspace_bytes rejected the bit range")
                # one-off per table and bit range: column-major score components and the row sums of r
                s_cols = torch.empty((2, max(bit_end - bit_begin, 1), self.N), device=self.device,
                                     dtype=torch.float64 if storage == "fp64" else torch.float32)
                r_sum = torch.empty(self.N, dtype=torch.float64, device=self.device)
                _lib.check(lib.tnk_modeproduct_tables(self.n, _ptr(self._st64()), bit_begin, bit_end, code, _ptr(s_cols),
                                                      _ptr(r_sum), _stream_ptr(self.device)), "tnk_modeproduct_tables")
                launch_counter["kernels"] += 1
                if self.dtype != torch.float64:
                    self._ST64 = None        # the row-major fp64 copy was only needed for the transposition
                self._ws[key] = (torch.empty(nbytes, dtype=torch.uint8, device=self.device), s_cols, r_sum)
            ws, s_cols, r_sum = self._ws[key]
            u = torch.empty(self.N, dtype=torch.float64, device=self.device)
            part = torch.empty(1, dtype=torch.float64, device=self.device)
            launched = ctypes.c_int(0)
            _lib.check(lib.tnk_ksd_modeproduct(self.n, self.length_scale, _ptr(q), 0 if q.dtype == torch.float64 else 1, code,
                                               _ptr(s_cols), _ptr(r_sum), bit_begin, bit_end, _ptr(u), _ptr(part),
                                               _ptr(ws), ws.numel(), ctypes.byref(launched), _stream_ptr(self.device)),
                       "tnk_ksd_modeproduct")
            launch_counter["kernels"] += launched.value
        return u, part


def shard_rows(num_states, rank, world):
    """Contiguous row block of rank `rank` among `world` ranks."""
    return (num_states * rank) // world, (num_states * (rank + 1)) // world


def combine_row_shards(u_rows, part, num_states, rank, world, group=None):
    """Exchange step of the sharded evaluation: every rank contributes its rows of u = K_p q and
    its partial of q.u, every rank receives the full u [N] and K [1].  One all_gather of
    [rows | partial] blocks when the rows divide evenly, else one sum-all_reduce of a padded
    vector.  Backend-neutral (NCCL on the GPUs, gloo in the CPU tests)."""
    if world == 1:
        return u_rows, part.clone()
    r0, r1 = shard_rows(num_states, rank, world)
    if u_rows.numel() != r1 - r0:
        raise ValueError("u_rows does not match this rank's row block")
    if num_states % world == 0:
        block = torch.cat([u_rows, part.reshape(1)])
        gathered = torch.empty(world * block.numel(), dtype=block.dtype, device=block.device)
        dist.all_gather_into_tensor(gathered, block, group=group)
        gathered = gathered.view(world, -1)
        return gathered[:, :-1].reshape(-1), gathered[:, -1].sum().reshape(1)
    buf = torch.zeros(num_states + 1, dtype=u_rows.dtype, device=u_rows.device)
    buf[r0:r1] = u_rows
    buf[num_states] = part.reshape(())
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf[:num_states], buf[num_states:].clone()


def sym_tile_stripe(rank, world):
    """(tile_offset, tile_stride) of rank `rank` for the symmetric evaluation: 128-state row tiles dealt round-robin,
    so every rank owns the same share of the triangle J >= I."""
    return rank, world


def combine_partial_sums(u_partial, part, group=None):
    """Exchange step of the symmetric evaluation: every rank holds a full-length partial of u = K_p q (its row
    tiles' rows and mirrored columns) and of q.u; one sum-all-reduce of [u | part] gives every rank u [N] and K [1].
    Backend-neutral (NCCL on the GPUs, gloo in the CPU tests)."""
    buf = torch.cat([u_partial, part.reshape(1)])
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf[:-1], buf[-1:].clone()


class PeerExchange:
    """Sinks of the fused exchange (tnk_ksd_rows_peers): per rank a symmetric-memory buffer
    [buffers][N + world] of fp64 - u rows followed by one partial-K slot per rank - mapped into
    every rank of the box (torch symmetric memory: cuMem + NVLink P2P).  The kernels that finish
    a rank's rows store them into all ranks' buffers; `wait()` is the cross-rank barrier that
    makes them visible.  Two buffers alternate so that a fast rank's next evaluation never
    overwrites rows a slow rank is still reading."""

    def __init__(self, num_states, device, group=None, buffers=2):
        import torch.distributed._symmetric_memory as symm_mem
        device = _require_cuda(device)
        group = dist.group.WORLD if group is None else group
        self.group, self.device = group, device
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > _lib.MAX_PEERS:
            raise ValueError(f"the fused exchange covers one NVSwitch box (<= {_lib.MAX_PEERS} ranks)")
        self.N, self.buffers = int(num_states), int(buffers)
        self.stride = (self.N + self.world + 31) // 32 * 32            # doubles per buffer, 256-byte multiple
        with torch.cuda.device(device):
            self.local = symm_mem.empty(self.buffers * self.stride, dtype=torch.float64, device=device)
            self.handle = symm_mem.rendezvous(self.local, group)
        self.base = [int(p) for p in self.handle.buffer_ptrs]
        self.step = 0

    def sinks(self):
        """(peer_u, peer_parts) ctypes pointer tables of the buffer the next evaluation writes, and
        the local views (u [N], parts [world]) of that buffer."""
        b = self.step % self.buffers
        self.step += 1
        off = b * self.stride * 8
        VP = ctypes.c_void_p * self.world
        peer_u = VP(*[base + off for base in self.base])
        peer_parts = VP(*[base + off + self.N * 8 for base in self.base])
        view = self.local[b * self.stride:(b + 1) * self.stride]
        return peer_u, peer_parts, view[:self.N], view[self.N:self.N + self.world]

    def wait(self):
        """Cross-rank barrier on the current stream (signal pads over NVLink, no NCCL)."""
        with torch.cuda.device(self.device):
            self.handle.barrier(channel=0)


def ksd_rows_fused(tables, q, exchange):
    """Row-sharded u = K_p q with the exchange fused into the reduction kernels: returns the full
    u [N] (fp64 view of the symmetric buffer) and K [1] on every rank.  SURVEY.md section 8(e)."""
    N, rank, world = tables.N, exchange.rank, exchange.world
    if q.device != tables.device or q.dim() != 1 or q.numel() != N:
        raise ValueError(f"q must be a 1-D tensor of {N} entries on {tables.device}")
    if q.dtype not in (torch.float32, torch.float64):
        raise ValueError("q must be float32 or float64")
    q = q.contiguous()
    r0, r1 = shard_rows(N, rank, world)
    lib = _lib.load()
    with torch.cuda.device(tables.device):
        peer_u, peer_parts, u, parts = exchange.sinks()
        part = torch.empty(1, dtype=torch.float64, device=tables.device)
        ws = tables.workspace(r0, r1)
        launched = ctypes.c_int(0)
        _lib.check(lib.tnk_ksd_rows_peers(tables.n, tables.length_scale, tables.prec_code, _ptr(q),
                                          0 if q.dtype == torch.float64 else 1, _ptr(tables.ST), _ptr(tables.Cst),
                                          r0, r1, world, rank, peer_u, peer_parts, _ptr(part), _ptr(ws), ws.numel(),
                                          ctypes.byref(launched), _stream_ptr(tables.device)), "tnk_ksd_rows_peers")
        launch_counter["kernels"] += launched.value
        exchange.wait()
        K = parts.sum(dim=0, keepdim=True)          # fixed slot order: identical bits on every rank
    return u, K


class _DevicePointer:
    """A raw device allocation as a CUDA array (torch.as_tensor wraps it without copying)."""

    def __init__(self, ptr, count, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class PeerComm:
    """tnk_comm of include/tnk.h: the sum-all-reduce of the sharded evaluations through our own kernels over
    peer-mapped (CUDA IPC / NVLink) buffers - barrier, one reduce-scatter + all-gather kernel, barrier; no NCCL on
    the data path.  torch.distributed only carries the 64-byte IPC handles once, at construction (collective over
    `group`; raises on every rank if any rank fails, so the ranks cannot disagree about the exchange path).
    sendbuf / recvbuf: fp64 device views [max_doubles] of this rank's send and receive regions."""

    def __init__(self, max_doubles, device, group=None):
        device = _require_cuda(device)
        group = dist.group.WORLD if group is None else group
        self.group, self.device = group, device
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > _lib.MAX_PEERS:
            raise ValueError(f"tnk_comm covers one NVSwitch box (<= {_lib.MAX_PEERS} ranks)")
        lib = _lib.load()
        self._comm = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        err = None
        with torch.cuda.device(device):
            try:
                _lib.check(lib.tnk_comm_create(self.rank, self.world, int(max_doubles), ctypes.byref(self._comm), handle),
                           "tnk_comm_create")
            except RuntimeError as exc:
                err = str(exc)
            gathered = [None] * self.world
            dist.all_gather_object(gathered, None if err else bytes(handle), group=group)
            if any(h is None for h in gathered):
                self.close()
                raise RuntimeError(f"tnk_comm_create failed on some rank ({err or 'a peer'})")
            try:
                _lib.check(lib.tnk_comm_connect(self._comm, b"".join(gathered)), "tnk_comm_connect")
            except RuntimeError as exc:
                err = str(exc)
            flags = [None] * self.world
            dist.all_gather_object(flags, err, group=group)
            if any(f is not None for f in flags):
                self.close()
                raise RuntimeError(f"tnk_comm_connect failed on some rank ({[f for f in flags if f][0]})")
        self.max_doubles = int(max_doubles)
        self.sendbuf = torch.as_tensor(_DevicePointer(lib.tnk_comm_sendbuf(self._comm), max_doubles), device=device)
        self.recvbuf = torch.as_tensor(_DevicePointer(lib.tnk_comm_recvbuf(self._comm), max_doubles), device=device)

    def allreduce_sum(self, count):
        """Sum of the first `count` doubles of every rank's sendbuf; returns the view recvbuf[:count] (rewritten by
        the next all-reduce: clone what has to live longer)."""
        launched = ctypes.c_int(0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().tnk_comm_allreduce_sum(self._comm, int(count), None, ctypes.byref(launched),
                                                          _stream_ptr(self.device)), "tnk_comm_allreduce_sum")
        launch_counter["kernels"] += launched.value
        return self.recvbuf[:count]

    def close(self):
        if getattr(self, "_comm", None) is not None and self._comm.value:
            with torch.cuda.device(self.device):
                torch.cuda.synchronize(self.device)
                _lib.load().tnk_comm_destroy(self._comm)
            self._comm = ctypes.c_void_p()
            self.sendbuf = self.recvbuf = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _host_result(t, like):
    """Device tensor -> CPU tensor of `like`'s dtype through pinned memory (torch's caching host allocator reuses the
    block), or a device-to-device cast when the caller's tensor lives on a GPU."""
    if like[1].type == "cuda":
        return t.to(device=like[1], dtype=like[0])
    host = torch.empty(t.shape, dtype=like[0], pin_memory=True)
    host.copy_(t if t.dtype == like[0] else t.to(like[0]), non_blocking=True)
    torch.cuda.current_stream(t.device).synchronize()
    return host


class _KsdQuadraticForm(torch.autograd.Function):
    """K(q) = q^T K_p q with dK/dq = 2 K_p q, one pass over the pair space for both."""

    @staticmethod
    def forward(ctx, q, tables, group, algorithm="pairs", exchange=None):
        q_dev = q.detach()
        if q_dev.dtype not in (torch.float32, torch.float64):
            q_dev = q_dev.to(torch.float64)
        q_dev = q_dev.to(tables.device, non_blocking=True)
        # sharding is opt-in: an explicit group, or an exchange object that carries one
        if group is None and exchange is not None:
            group = exchange.group
        world = dist.get_world_size(group) if group is not None else 1
        rank = dist.get_rank(group) if world > 1 else 0
        comm = exchange if isinstance(exchange, PeerComm) and world > 1 else None
        N = tables.N
        if algorithm == "modeproduct":
            b0, b1 = shard_rows(tables.n, rank, world)      # score components split over the ranks
            u, part = tables.ksd_modeproduct(q_dev, b0, b1)
            if comm is not None:
                comm.sendbuf[:N].copy_(u)
                comm.sendbuf[N:N + 1].copy_(part)
                res = comm.allreduce_sum(N + 1).clone()
                u, part = res[:N], res[N:]
            elif world > 1:
                u, part = combine_partial_sums(u, part, group)
        elif algorithm in ("pairs_sym", "pairs_sym64"):
            out = (comm.sendbuf[:N], comm.sendbuf[N:N + 1]) if comm is not None else None
            if algorithm == "pairs_sym":
                u, part = tables.ksd_sym(q_dev, *sym_tile_stripe(rank, world), out=out)
            else:
                u, part = tables.ksd_sym64(q_dev, rank, world, out=out)
            if comm is not None:
                res = comm.allreduce_sum(N + 1).clone()     # the recv region is rewritten by the next evaluation
                u, part = res[:N], res[N:]
            elif world > 1:
                u, part = combine_partial_sums(u, part, group)
        elif isinstance(exchange, PeerExchange) and world > 1:
            u, part = ksd_rows_fused(tables, q_dev, exchange)
            u = u.clone()     # the symmetric buffer is rewritten two evaluations later
        else:
            r0, r1 = shard_rows(N, rank, world)
            u, part = tables.ksd_rows(q_dev, r0, r1)
            if world > 1:
                u, part = combine_row_shards(u[r0:r1], part, N, rank, world, group)
        ctx.save_for_backward(u)
        ctx.q_meta = (q.dtype, q.device)
        ctx.grad_host = None
        if q.device.type == "cpu" and ctx.needs_input_grad[0]:
            # q lives on the host and wants a gradient: dK/dq = 2u goes to pinned memory behind K on the same stream, so
            # the one synchronisation below covers both and backward() finds its result on the host (no second kernel,
            # copy and synchronisation after the autograd dispatch)
            g = u * 2.0
            if g.dtype != q.dtype:
                g = g.to(q.dtype)
            ctx.grad_host = torch.empty(g.shape, dtype=q.dtype, pin_memory=True)
            ctx.grad_host.copy_(g, non_blocking=True)
        return _host_result(part.reshape(()), (torch.float64, q.device))

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.grad_host is not None and not grad_out.is_cuda:
            g, ctx.grad_host = ctx.grad_host, None        # handed over once (autograd may adopt it as q.grad without a copy)
            scale = float(grad_out)
            return (g if scale == 1.0 else g.mul_(scale)), None, None, None, None
        (u,) = ctx.saved_tensors
        scale = grad_out.to(u.device, torch.float64) if grad_out.is_cuda else float(grad_out) * 1.0
        grad = u * (2.0 * scale)
        return _host_result(grad, ctx.q_meta), None, None, None, None


ALGORITHMS = ("pairs", "pairs_sym", "pairs_sym64", "modeproduct")


def ksd_quadratic_form(q, tables, group=None, algorithm="pairs", exchange=None):
    """KSD^2 = sum_ij q_i q_j k_p(z_i, z_j) as an fp64 scalar tensor on q's device, differentiable
    w.r.t. q (any float dtype, CPU or CUDA; e.g. PennyLane probabilities).  Drop-in for the double
    loop of reference ksd_vi.py:110-131 / ksd_vi_quantum.py:122-142.
    algorithm="pairs" evaluates all 4^n pairs on the fly (the reference's loop, tiled);
    algorithm="pairs_sym" evaluates every unordered pair of 128-state tiles once on the tensor cores and uses
    k_p(z,z') = k_p(z',z) for the mirrored half (fp32-grade, tables of precision "tc32");
    algorithm="pairs_sym64" is the same schedule in fp64 (tables of precision "fp64", the 1e-6 mode, bit-reproducible);
    algorithm="modeproduct" returns the same value through the Kronecker factorisation of the
    Hamming kernel in O(n^2 2^n) (fp64).
    Sharding over several GPUs is OPT-IN: pass `group` (a torch.distributed process group whose ranks all call with
    the same q - the caller's responsibility, see KSDVariationalInference(distributed=True)) and/or `exchange`:
    a PeerComm (tnk_comm: our own all-reduce over peer-mapped buffers, every algorithm) or a PeerExchange (rows of
    the "pairs" evaluation stored straight into every rank's buffer, tnk_ksd_rows_peers).  Without an exchange
    object the partial results go through one NCCL/gloo collective.  With neither, the whole evaluation runs on
    this process's GPU even when torch.distributed is initialised."""
    if algorithm not in ALGORITHMS:
        raise ValueError(f"algorithm must be one of {ALGORITHMS}")
    if q.dim() != 1 or q.numel() != tables.N:
        raise ValueError(f"Probabilities shape mismatch: {tuple(q.shape)}")
    return _KsdQuadraticForm.apply(q, tables, group, algorithm, exchange)


class _BornSoftmaxEntropy(torch.autograd.Function):
    """(q, H) = (softmax(logits), -sum q log clamp(q, 1e-10)) for one row of 2^n fp32 logits on the GPU:
    tnk_born_forward / tnk_born_backward (four streaming passes per epoch instead of the autograd graph of
    reference born_machine_classical_sim.py:74-99,177-181)."""

    @staticmethod
    def forward(ctx, logits):
        ctx.set_materialize_grads(False)
        lib = _lib.load()
        logits = logits.contiguous()
        N = logits.numel()
        with torch.cuda.device(logits.device):
            m = logits.max().reshape(1)
            blocks = lib.tnk_born_blocks(N)
            q = torch.empty_like(logits)
            buf = torch.empty(2 * blocks + 1, dtype=torch.float64, device=logits.device)
            _lib.check(lib.tnk_born_forward(_ptr(logits), N, _ptr(m), _ptr(q), _ptr(buf), _ptr(buf[2 * blocks:]),
                                            _ptr(buf[blocks:]), _stream_ptr(logits.device)), "tnk_born_forward")
            launch_counter["kernels"] += 2
            H = buf[:blocks].sum().to(logits.dtype)
        ctx.save_for_backward(q)
        return q, H

    @staticmethod
    def backward(ctx, grad_q, grad_H):
        (q,) = ctx.saved_tensors
        if grad_q is None and grad_H is None:
            return None
        lib = _lib.load()
        N = q.numel()
        with torch.cuda.device(q.device):
            if grad_q is not None:
                if grad_q.dtype not in (torch.float32, torch.float64):
                    grad_q = grad_q.to(torch.float32)
                grad_q = grad_q.contiguous()
            gH = None if grad_H is None else grad_H.detach().to(torch.float64).reshape(1)
            blocks = lib.tnk_born_blocks(N)
            scratch = torch.empty(blocks, dtype=torch.float64, device=q.device)
            g_logits = torch.empty_like(q)
            _lib.check(lib.tnk_born_backward(_ptr(q), None if grad_q is None else _ptr(grad_q),
                                             1 if grad_q is None or grad_q.dtype == torch.float32 else 0,
                                             None if gH is None else _ptr(gH), N, _ptr(g_logits), _ptr(scratch),
                                             _stream_ptr(q.device)), "tnk_born_backward")
            launch_counter["kernels"] += 2
        return g_logits


def born_softmax_entropy(logits):
    """q [N] and H (0-dim) of one unconditional Born machine from its fp32 CUDA logits [N];
    differentiable w.r.t. the logits through both outputs."""
    if not (logits.is_cuda and logits.dtype == torch.float32 and logits.dim() == 1):
        raise ValueError("born_softmax_entropy expects a 1-D float32 CUDA tensor of logits")
    return _BornSoftmaxEntropy.apply(logits)


def tc_plan(n):
    """K-step plan of the tensor-core variant for n latent bits (tnk_tc_plan)."""
    lib = _lib.load()
    np_, k32, kb, ns, ns16 = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    steps = (ctypes.c_int32 * 128)()
    _lib.check(lib.tnk_tc_plan(n, ctypes.byref(np_), ctypes.byref(k32), ctypes.byref(kb), ctypes.byref(ns),
                               ctypes.byref(ns16), steps), "tnk_tc_plan")
    return {"np": np_.value, "k32": k32.value, "kb": kb.value, "nsteps": ns.value, "nsteps16": ns16.value,
            "steps": [tuple(steps[4 * k:4 * k + 4]) for k in range(ns.value)]}


def alu_peak(precision="fp32", iters=4096, device="cuda"):
    """Measured FMA-chain rate of the device in lane-ops/s (the pair kernel's roofline denominator)."""
    device = _require_cuda(device)
    lib = _lib.load()
    with torch.cuda.device(device):
        scratch = torch.zeros(8, dtype=torch.float64, device=device)
        rate, ms = ctypes.c_double(0.0), ctypes.c_float(0.0)
        _lib.check(lib.tnk_alu_peak(3 if precision == "fp32x2" else PRECISIONS[precision], iters, _ptr(scratch), ctypes.byref(rate),
                                    ctypes.byref(ms), _stream_ptr(device)), "tnk_alu_peak")
    return rate.value, ms.value


def time_pair_kernel(start_event=None, stop_event=None):
    """Ask tnk_ksd_rows to record the given torch.cuda.Event pair around the pair kernel alone
    (both must have been recorded once so that their CUDA handles exist); None, None switches off."""
    lib = _lib.load()
    if start_event is None or stop_event is None:
        _lib.check(lib.tnk_ksd_timing_events(None, None), "tnk_ksd_timing_events")
    else:
        _lib.check(lib.tnk_ksd_timing_events(ctypes.c_void_p(start_event.cuda_event),
                                             ctypes.c_void_p(stop_event.cuda_event)), "tnk_ksd_timing_events")
