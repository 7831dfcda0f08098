"""Build and load libtnk_b200.so (the C-ABI of include/tnk.h) through ctypes.

The library is built in-tree (tensornetworks_b200/libtnk_b200.so) with
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo
so it travels with the repo snapshot to the GPU box.  There is no CPU fallback: if the
library cannot be loaded every compute entry point raises.
"""
import ctypes
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_DIR = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
INCLUDE = os.path.join(REPO_DIR, "include")
LIB_PATH = os.path.join(PKG_DIR, "libtnk_b200.so")
SOURCES = ["tnk_enum.cu", "tnk_ksd.cu", "tnk_peak.cu", "tnk_modeprod.cu", "tnk_tc.cu", "tnk_born.cu", "tnk_sym64.cu", "tnk_comm.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-I", INCLUDE]

PREC_FP64, PREC_FP32, PREC_FP64_SIMPLE, PREC_TC32 = 0, 1, 2, 4
MAX_PEERS = 8  # TNK_MAX_PEERS
MP_STORE = {"fp64": 0, "fp32": 1}  # TNK_MP_STORE_*
SYM_SPLIT = {"tf32": 0, "bf16x2": 1, "tf32_tile": 2, "tf32_dot": 3}  # TNK_SYM_SPLIT_*

# every symbol include/tnk.h declares
EXPORTS = ["tnk_last_error", "tnk_version", "tnk_bn_joint", "tnk_ksd_layout", "tnk_score",
           "tnk_ksd_workspace_bytes", "tnk_ksd_rows", "tnk_ksd_rows_static", "tnk_ksd_rows_static_capacity", "tnk_ksd_rows_peers", "tnk_ksd_timing_events", "tnk_alu_peak",
           "tnk_ksd_modeproduct_workspace_bytes", "tnk_modeproduct_tables", "tnk_modeproduct_plan_limit", "tnk_ksd_modeproduct", "tnk_tc_plan",
           "tnk_born_blocks", "tnk_born_forward", "tnk_born_backward",
           "tnk_ksd_sym_workspace_bytes", "tnk_ksd_sym",
           "tnk_ksd_sym64_workspace_bytes", "tnk_ksd_sym64_units", "tnk_ksd_sym64",
           "tnk_comm_create", "tnk_comm_connect", "tnk_comm_destroy", "tnk_comm_rank", "tnk_comm_world",
           "tnk_comm_sendbuf", "tnk_comm_recvbuf", "tnk_comm_allreduce_sum", "tnk_ksd_sym_allreduce"]


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    built = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(INCLUDE, "tnk.h")]
    return any(os.path.getmtime(d) > built for d in deps)


ABI_VERSION = 201   # tnk_version() of the library these ctypes signatures were written for (bumped with include/tnk.h)


def build(force=False, verbose=False):
    """Compile the CUDA sources for sm_100a and link the shared library (no GPU needed).  Serialised across
    processes with a file lock: under torchrun every rank may find the library stale at the same time."""
    import fcntl
    obj_dir = os.path.join(PKG_DIR, "build")
    os.makedirs(obj_dir, exist_ok=True)
    with open(os.path.join(obj_dir, ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not _stale():      # another process built it while this one waited
                return LIB_PATH
            return _build_locked(obj_dir, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(obj_dir, verbose=False):

    def compile_one(src):
        obj = os.path.join(obj_dir, src.replace(".cu", ".o"))
        cmd = [_nvcc(), *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
        if verbose:
            print(" ".join(cmd))
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as pool:
        objs = list(pool.map(compile_one, SOURCES))
    tmp = LIB_PATH + ".tmp"
    cmd = [_nvcc(), "-shared", "-o", tmp, *objs, "-lcudart"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    os.replace(tmp, LIB_PATH)                  # readers never see a half-written library
    return LIB_PATH


_lib = None


def load():
    """Return the ctypes handle, building first if the .so is missing or stale."""
    global _lib
    if _lib is not None:
        return _lib
    if _stale():
        try:
            build()
        except Exception as exc:  # no toolchain: refuse loudly, never fall back
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(f"libtnk_b200.so is missing and could not be built: {exc}") from exc
            import warnings
            warnings.warn(f"libtnk_b200.so is older than its sources and could not be rebuilt ({exc}); "
                          "using the existing library after checking its ABI version")
    lib = ctypes.CDLL(LIB_PATH)
    missing = [name for name in EXPORTS if not hasattr(lib, name)]
    lib.tnk_version.restype = ctypes.c_int
    if missing or lib.tnk_version() != ABI_VERSION:
        raise RuntimeError(f"libtnk_b200.so does not match include/tnk.h (version {lib.tnk_version()} vs {ABI_VERSION}, "
                           f"missing symbols {missing}): rebuild it with python -c 'import __graft_entry__ as g; g.build()'")
    c_int, c_i64, c_dbl, c_vp, c_sz = ctypes.c_int, ctypes.c_int64, ctypes.c_double, ctypes.c_void_p, ctypes.c_size_t
    lib.tnk_last_error.restype = ctypes.c_char_p
    lib.tnk_last_error.argtypes = []
    lib.tnk_version.restype = c_int
    lib.tnk_bn_joint.restype = c_int
    lib.tnk_bn_joint.argtypes = [c_int, c_int, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]
    lib.tnk_ksd_layout.restype = c_int
    lib.tnk_ksd_layout.argtypes = [c_int, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_i64), ctypes.POINTER(c_int)]
    lib.tnk_score.restype = c_int
    lib.tnk_score.argtypes = [c_int, c_dbl, c_vp, c_vp, c_int, c_vp, c_vp, c_vp]
    lib.tnk_ksd_workspace_bytes.restype = c_sz
    lib.tnk_ksd_workspace_bytes.argtypes = [c_int, c_int, c_i64, c_i64]
    lib.tnk_ksd_rows.restype = c_int
    lib.tnk_ksd_rows.argtypes = [c_int, c_dbl, c_int, c_vp, c_int, c_vp, c_vp, c_i64, c_i64, c_vp, c_vp,
                                 c_vp, c_sz, ctypes.POINTER(c_int), c_vp]
    lib.tnk_ksd_rows_static_capacity.restype = c_sz
    lib.tnk_ksd_rows_static_capacity.argtypes = []
    lib.tnk_ksd_rows_static.restype = c_int
    lib.tnk_ksd_rows_static.argtypes = [c_int, c_dbl, c_int, c_vp, c_int, c_vp, c_vp, c_i64, c_i64, c_vp, c_vp,
                                        ctypes.POINTER(c_int), c_vp]
    lib.tnk_ksd_rows_peers.restype = c_int
    lib.tnk_ksd_rows_peers.argtypes = [c_int, c_dbl, c_int, c_vp, c_int, c_vp, c_vp, c_i64, c_i64,
                                       c_int, c_int, ctypes.POINTER(c_vp), ctypes.POINTER(c_vp), c_vp,
                                       c_vp, c_sz, ctypes.POINTER(c_int), c_vp]
    lib.tnk_ksd_sym_workspace_bytes.restype = c_sz
    lib.tnk_ksd_sym_workspace_bytes.argtypes = [c_int, c_int, c_int, c_int]
    lib.tnk_ksd_sym.restype = c_int
    lib.tnk_ksd_sym.argtypes = [c_int, c_dbl, c_int, c_vp, c_int, c_vp, c_vp, c_int, c_int, c_int, c_vp, c_vp,
                                c_vp, c_sz, ctypes.POINTER(c_int), c_vp]
    lib.tnk_ksd_sym64_workspace_bytes.restype = c_sz
    lib.tnk_ksd_sym64_workspace_bytes.argtypes = [c_int, c_int, c_int]
    lib.tnk_ksd_sym64_units.restype = c_int
    lib.tnk_ksd_sym64_units.argtypes = [c_int]
    lib.tnk_ksd_sym64.restype = c_int
    lib.tnk_ksd_sym64.argtypes = [c_int, c_dbl, c_vp, c_int, c_vp, c_vp, c_int, c_int, c_vp, c_vp,
                                  c_vp, c_sz, ctypes.POINTER(c_int), c_vp]
    lib.tnk_comm_create.restype = c_int
    lib.tnk_comm_create.argtypes = [c_int, c_int, c_i64, ctypes.POINTER(c_vp), c_vp]
    lib.tnk_comm_connect.restype = c_int
    lib.tnk_comm_connect.argtypes = [c_vp, c_vp]
    lib.tnk_comm_destroy.restype = c_int
    lib.tnk_comm_destroy.argtypes = [c_vp]
    lib.tnk_comm_rank.restype = c_int
    lib.tnk_comm_rank.argtypes = [c_vp]
    lib.tnk_comm_world.restype = c_int
    lib.tnk_comm_world.argtypes = [c_vp]
    lib.tnk_comm_sendbuf.restype = c_vp
    lib.tnk_comm_sendbuf.argtypes = [c_vp]
    lib.tnk_comm_recvbuf.restype = c_vp
    lib.tnk_comm_recvbuf.argtypes = [c_vp]
    lib.tnk_comm_allreduce_sum.restype = c_int
    lib.tnk_comm_allreduce_sum.argtypes = [c_vp, c_i64, c_vp, ctypes.POINTER(c_int), c_vp]
    lib.tnk_ksd_sym_allreduce.restype = c_int
    lib.tnk_ksd_sym_allreduce.argtypes = [c_int, c_dbl, c_int, c_vp, c_int, c_vp, c_vp, c_int, c_vp, c_vp, c_vp,
                                          c_vp, c_sz, ctypes.POINTER(c_int), c_vp]
    lib.tnk_ksd_modeproduct_workspace_bytes.restype = c_sz
    lib.tnk_ksd_modeproduct_workspace_bytes.argtypes = [c_int, c_int, c_int, c_int]
    lib.tnk_ksd_modeproduct.restype = c_int
    lib.tnk_modeproduct_plan_limit.restype = c_int
    lib.tnk_modeproduct_plan_limit.argtypes = [c_int]
    lib.tnk_modeproduct_tables.restype = c_int
    lib.tnk_modeproduct_tables.argtypes = [c_int, c_vp, c_int, c_int, c_int, c_vp, c_vp, c_vp]
    lib.tnk_ksd_modeproduct.argtypes = [c_int, c_dbl, c_vp, c_int, c_int, c_vp, c_vp, c_int, c_int, c_vp, c_vp, c_vp, c_sz,
                                        ctypes.POINTER(c_int), c_vp]
    lib.tnk_tc_plan.restype = c_int
    lib.tnk_tc_plan.argtypes = [c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                                ctypes.POINTER(c_int), ctypes.POINTER(c_int), ctypes.POINTER(ctypes.c_int32)]
    lib.tnk_born_blocks.restype = c_int
    lib.tnk_born_blocks.argtypes = [ctypes.c_int64]
    lib.tnk_born_forward.restype = c_int
    lib.tnk_born_forward.argtypes = [c_vp, ctypes.c_int64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]
    lib.tnk_born_backward.restype = c_int
    lib.tnk_born_backward.argtypes = [c_vp, c_vp, c_int, c_vp, ctypes.c_int64, c_vp, c_vp, c_vp]
    lib.tnk_ksd_timing_events.restype = c_int
    lib.tnk_ksd_timing_events.argtypes = [c_vp, c_vp]
    lib.tnk_alu_peak.restype = c_int
    lib.tnk_alu_peak.argtypes = [c_int, c_int, c_vp, ctypes.POINTER(c_dbl), ctypes.POINTER(ctypes.c_float), c_vp]
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().tnk_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def layout(n, precision):
    """(np, n_pad, elem_bytes) of the packed score table for (n, precision)."""
    lib = load()
    np_, n_pad, elem = ctypes.c_int(), ctypes.c_int64(), ctypes.c_int()
    check(lib.tnk_ksd_layout(n, precision, ctypes.byref(np_), ctypes.byref(n_pad), ctypes.byref(elem)), "tnk_ksd_layout")
    return np_.value, n_pad.value, elem.value
