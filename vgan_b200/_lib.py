"""ctypes binding of the C ABI in include/vgan_b200.h (libvgan_b200.so, built in-tree).

There is deliberately no fallback: if the library is missing and cannot be built, or a
call fails (e.g. no sm_100 device), a RuntimeError is raised."""
from __future__ import annotations

import ctypes as C
import hashlib
import os
import shutil
import subprocess
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libvgan_b200.so")
DBG_LIB_PATH = os.path.join(HERE, "libvgan_b200_dbg.so")
CSRC = os.path.join(HERE, "csrc")

PREC_BF16, PREC_FP32, CENTER, NEED_DX, LADDER, BF16X3 = 0, 1, 2, 4, 8, 64
MAX_SCALES = 8
SUBSPACE_STATIC_SCRATCH_BYTES = 1 << 20      # VGAN_B200_SUBSPACE_STATIC_SCRATCH_BYTES

EXPORTS = [
    "vgan_b200_abi_version", "vgan_b200_last_error", "vgan_b200_create", "vgan_b200_destroy",
    "vgan_b200_set_cluster_limit", "vgan_b200_launch_count", "vgan_b200_workspace_layout", "vgan_b200_stats", "vgan_b200_scales",
    "vgan_b200_pack", "vgan_b200_mmd2_fwd_bwd", "vgan_b200_finalize", "vgan_b200_mmd2_eval",
    "vgan_b200_upper_softmax", "vgan_b200_upper_softmax_bwd", "vgan_b200_mmd2_matrix",
    "vgan_b200_permutation_scratch_bytes", "vgan_b200_permutation_stats",
    "vgan_b200_subspace_scratch_bytes", "vgan_b200_subspace_dist", "vgan_b200_subspace_dist_static",
]
DEBUG_EXPORTS = ["vgan_b200_debug_tile_sums", "vgan_b200_debug_timeline", "vgan_b200_debug_variant",
                 "vgan_b200_umma_probe", "vgan_b200_pair_probe", "vgan_b200_tma_rate_probe",
                 "vgan_b200_umma_rate_probe"]


class Layout(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("Nx", "Ny", "D", "D_u", "Dp", "Nxp", "Nyp")] + \
               [("flags", C.c_int32), ("n_scales", C.c_int32)] + \
               [(n, C.c_size_t) for n in ("off_par", "off_acc", "off_sums", "off_colsum", "off_stats", "off_cmax",
                                          "off_misc", "acc_bytes", "off_mean", "off_cm", "off_nrm", "off_nrms",
                                          "off_xchg", "off_zq", "off_zf", "total")]


class EvalArgs(C.Structure):
    _fields_ = [("layout", C.POINTER(Layout)), ("ws", C.c_void_p), ("X", C.c_void_p), ("U", C.c_void_p),
                ("Y", C.c_void_p), ("bandwidth", C.c_void_p), ("frozen", C.c_int),
                ("mults_host", C.POINTER(C.c_float)), ("weight", C.c_float), ("dY", C.c_void_p),
                ("dY_mul", C.c_void_p), ("dX", C.c_void_p), ("dU_penalty", C.c_void_p), ("loss", C.c_void_p),
                ("stream", C.c_void_p)]


def _source_hash() -> str:
    h = hashlib.sha1()
    names = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh")) or f == "Makefile")
    for path in [os.path.join(CSRC, f) for f in names] + [os.path.join(os.path.dirname(HERE), "include", "vgan_b200.h")]:
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _stale(path: str) -> bool:
    """True when `path` was not built from the sources that are on disk now (content hash, not mtime: the tree is
    copied to the GPU box and copies do not keep times)."""
    try:
        with open(path + ".stamp") as f:
            return f.read().strip() != _source_hash()
    except OSError:
        return True


def build(force: bool = False, debug: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into libvgan_b200.so (nvcc cross-compiles without a GPU).  ``debug=True`` builds
    libvgan_b200_dbg.so (measurement entry points, `make DEBUG=1`)."""
    path = DBG_LIB_PATH if debug else LIB_PATH
    if os.path.exists(path) and not force and not _stale(path):
        return path
    if shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        raise RuntimeError(f"vgan_b200: {os.path.basename(path)} is missing/stale and nvcc is not available to build it")
    env = dict(os.environ)
    env["PATH"] = "/usr/local/cuda/bin:" + env.get("PATH", "")
    cmd = ["make", "-j8", "-C", CSRC] + (["DEBUG=1"] if debug else [])
    r = subprocess.run(cmd, env=env, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("vgan_b200: building the CUDA library failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    with open(path + ".stamp", "w") as f:
        f.write(_source_hash())
    return path


_lock = threading.Lock()
_lib = None
_dbg = None
_tls = threading.local()


def _declare(L: C.CDLL) -> None:
    i64, p, i, f = C.c_int64, C.c_void_p, C.c_int, C.c_float
    LP = C.POINTER(Layout)
    L.vgan_b200_abi_version.restype = i
    L.vgan_b200_last_error.restype = C.c_char_p
    L.vgan_b200_create.argtypes = [i, C.POINTER(p)]
    L.vgan_b200_destroy.argtypes = [p]
    L.vgan_b200_set_cluster_limit.argtypes = [p, i]
    L.vgan_b200_launch_count.argtypes = [p, C.POINTER(C.c_ulonglong)]
    L.vgan_b200_workspace_layout.argtypes = [i64, i64, i64, i64, i, i, LP]
    L.vgan_b200_stats.argtypes = [p, LP, p, p, p, p, i64, i64, p]
    L.vgan_b200_scales.argtypes = [p, LP, p, p, i, C.POINTER(C.c_float), f, p]
    L.vgan_b200_pack.argtypes = [p, LP, p, p, p, p, i64, i64, p]
    L.vgan_b200_mmd2_fwd_bwd.argtypes = [p, LP, p, p, p, p, i64, i64, i, i, p, p, p, p]
    L.vgan_b200_finalize.argtypes = [p, LP, p, f, p, i64, i64, i, p, p]
    L.vgan_b200_mmd2_eval.argtypes = [p, C.POINTER(EvalArgs)]
    L.vgan_b200_upper_softmax.argtypes = [p, p, i64, i64, p, p, p, p]
    L.vgan_b200_upper_softmax_bwd.argtypes = [p, p, p, p, p, i64, i64, p, p]
    L.vgan_b200_mmd2_matrix.argtypes = [p, p, i64, i64, p, i, f, p, p]
    L.vgan_b200_subspace_scratch_bytes.argtypes = [i64, i64, C.POINTER(C.c_size_t)]
    L.vgan_b200_permutation_scratch_bytes.argtypes = [i64, C.POINTER(C.c_size_t)]
    L.vgan_b200_permutation_stats.argtypes = [p, p, i64, i64, p, i64, p, p, p]
    L.vgan_b200_subspace_dist.argtypes = [p, p, i, i64, i64, f, p, p, p, p, p]
    L.vgan_b200_subspace_dist_static.argtypes = [p, p, i, i64, i64, f, p, p, p, p]
    for name in EXPORTS:
        if name != "vgan_b200_last_error":
            getattr(L, name).restype = i


def lib() -> C.CDLL:
    global _lib
    with _lock:
        if _lib is None:
            build()                          # no-op when the library matches the sources on disk
            L = C.CDLL(LIB_PATH)
            _declare(L)
            _lib = L
        return _lib


def lib_debug() -> C.CDLL:
    """The debug build (tools/, one GPU test): the product entry points plus the measurement ones."""
    global _dbg
    with _lock:
        if _dbg is None:
            build(debug=True)
            L = C.CDLL(DBG_LIB_PATH)
            _declare(L)
            i64, p, i = C.c_int64, C.c_void_p, C.c_int
            L.vgan_b200_debug_tile_sums.argtypes = [p, p, i]
            L.vgan_b200_debug_timeline.argtypes = [p, p]
            L.vgan_b200_debug_variant.argtypes = [p, i, i]
            L.vgan_b200_umma_probe.argtypes = [p, p, p, p, p, p]
            L.vgan_b200_pair_probe.argtypes = [p, p, p, p, p, p, p, p]
            L.vgan_b200_tma_rate_probe.argtypes = [p, p, i64, i64, i, i, i, i, i, i, i, p, p]
            L.vgan_b200_umma_rate_probe.argtypes = [p, i, i, i, p, i64, i, i, p, p]
            for name in DEBUG_EXPORTS:
                getattr(L, name).restype = i
            _dbg = L
        return _dbg


class _Ctx:
    """Owns one vgan_b200_ctx; destroyed with the thread-local table that holds it."""

    def __init__(self, L: C.CDLL, device: int):
        self.L, self.handle = L, C.c_void_p()
        check(L.vgan_b200_create(device, C.byref(self.handle)), "create", L)

    def __del__(self):
        try:
            if self.handle:
                self.L.vgan_b200_destroy(self.handle)
        except Exception:
            pass


def ctx(device_index: int, L: C.CDLL | None = None) -> C.c_void_p:
    """The calling thread's context for `device_index` (contexts are per host thread: distinct contexts are
    thread-safe against each other, one context is not re-entrant)."""
    L = lib() if L is None else L
    table = getattr(_tls, "ctx", None)
    if table is None:
        table = _tls.ctx = {}
    key = (id(L), int(device_index))
    c = table.get(key)
    if c is None:
        c = table[key] = _Ctx(L, int(device_index))
    return c.handle


def launch_count(device_index: int) -> int:
    """Kernels launched so far through the calling thread's context for `device_index`."""
    n = C.c_ulonglong()
    check(lib().vgan_b200_launch_count(ctx(device_index), C.byref(n)), "launch_count")
    return int(n.value)


def check(status: int, what: str, L: C.CDLL | None = None) -> None:
    if status != 0:
        msg = (lib() if L is None else L).vgan_b200_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"vgan_b200.{what} failed (status {status}): {msg}")


def layout(Nx: int, Ny: int, D: int, D_u: int, n_scales: int, flags: int) -> Layout:
    out = Layout()
    check(lib().vgan_b200_workspace_layout(Nx, Ny, D, D_u, n_scales, flags, C.byref(out)), "workspace_layout")
    return out
