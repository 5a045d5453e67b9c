"""ctypes binding of the C ABI in include/vgan_b200.h (libvgan_b200.so, built in-tree).

There is deliberately no fallback: if the library is missing and cannot be built, or a
call fails (e.g. no sm_100 device), a RuntimeError is raised."""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libvgan_b200.so")
CSRC = os.path.join(HERE, "csrc")

PREC_BF16, PREC_FP32, CENTER, NEED_DX, LADDER, NO_CLUSTER, NO_PAIR, BF16X3 = 0, 1, 2, 4, 8, 16, 32, 64
MAX_SCALES = 8

EXPORTS = [
    "vgan_b200_abi_version", "vgan_b200_last_error", "vgan_b200_workspace_layout",
    "vgan_b200_prologue", "vgan_b200_scales", "vgan_b200_mmd2_fwd_bwd",
    "vgan_b200_colmax_penalty", "vgan_b200_finalize_loss", "vgan_b200_mmd2_matrix",
    "vgan_b200_umma_probe", "vgan_b200_debug_tile_sums", "vgan_b200_umma_rate_probe", "vgan_b200_pair_probe",
    "vgan_b200_debug_variant", "vgan_b200_tma_rate_probe",
    "vgan_b200_subspace_scratch_bytes", "vgan_b200_subspace_dist",
    "vgan_b200_permutation_scratch_bytes", "vgan_b200_permutation_stats", "vgan_b200_debug_timeline",
]


class Layout(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("Nx", "Ny", "D", "Dp", "Nxp", "Nyp")] + \
               [(n, C.c_size_t) for n in ("off_zq", "off_zf", "off_nrm", "off_nrms", "off_mean",
                                          "off_stats", "off_cmax", "off_par", "off_xchg", "total")]


def build(force: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into libvgan_b200.so (nvcc cross-compiles without a GPU)."""
    if os.path.exists(LIB_PATH) and not force:
        srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))]
        srcs.append(os.path.join(os.path.dirname(HERE), "include", "vgan_b200.h"))
        if all(os.path.getmtime(s) <= os.path.getmtime(LIB_PATH) for s in srcs):
            return LIB_PATH
    if shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        raise RuntimeError("vgan_b200: libvgan_b200.so is missing/stale and nvcc is not available to build it")
    env = dict(os.environ)
    env["PATH"] = "/usr/local/cuda/bin:" + env.get("PATH", "")
    r = subprocess.run(["make", "-j8", "-C", CSRC], env=env, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("vgan_b200: building the CUDA library failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return LIB_PATH


_lock = threading.Lock()
_lib = None


def lib() -> C.CDLL:
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                build()
            L = C.CDLL(LIB_PATH)
            i64, p, i, f = C.c_int64, C.c_void_p, C.c_int, C.c_float
            L.vgan_b200_abi_version.restype = i
            L.vgan_b200_last_error.restype = C.c_char_p
            L.vgan_b200_workspace_layout.argtypes = [i64, i64, i64, i64, i, C.POINTER(Layout)]
            L.vgan_b200_prologue.argtypes = [p, p, p, i64, i64, i64, i, p, p]
            L.vgan_b200_scales.argtypes = [p, i64, i64, i64, p, i, C.POINTER(C.c_float), i, p, p]
            L.vgan_b200_mmd2_fwd_bwd.argtypes = [p, i64, i64, i64, i64, i64, i, i, p, i, i, p, p, p, p, p]
            L.vgan_b200_colmax_penalty.argtypes = [p, i64, i64, f, p, p, p, i64, i64, p]
            L.vgan_b200_finalize_loss.argtypes = [p, i64, i64, p, p, p]
            L.vgan_b200_mmd2_matrix.argtypes = [p, i64, i64, p, i, f, p, p]
            L.vgan_b200_umma_probe.argtypes = [p, p, p, p, p]
            L.vgan_b200_debug_tile_sums.argtypes = [p, i]
            L.vgan_b200_umma_rate_probe.argtypes = [i, i, i, p, i64, i, i, p, p]
            L.vgan_b200_pair_probe.argtypes = [p, p, p, p, p, p, p]
            L.vgan_b200_debug_variant.argtypes = [i]
            L.vgan_b200_debug_timeline.argtypes = [p]
            L.vgan_b200_subspace_scratch_bytes.argtypes = [i64, i64, C.POINTER(C.c_size_t)]
            L.vgan_b200_permutation_scratch_bytes.argtypes = [i64, C.POINTER(C.c_size_t)]
            L.vgan_b200_permutation_stats.argtypes = [p, i64, i64, p, i64, p, p, p]
            L.vgan_b200_subspace_dist.argtypes = [p, i, i64, i64, f, p, p, p, p, p]
            L.vgan_b200_tma_rate_probe.argtypes = [p, i64, i64, i, i, i, i, i, i, i, p, p]
            for name in EXPORTS:
                if name not in ("vgan_b200_last_error",):
                    getattr(L, name).restype = i
            if os.environ.get("VGAN_B200_VARIANT"):
                L.vgan_b200_debug_variant(int(os.environ["VGAN_B200_VARIANT"]))
            _lib = L
        return _lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib().vgan_b200_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"vgan_b200.{what} failed (status {status}): {msg}")


def layout(Nx: int, Ny: int, D: int, D_u: int, flags: int) -> Layout:
    out = Layout()
    check(lib().vgan_b200_workspace_layout(Nx, Ny, D, D_u, flags, C.byref(out)), "workspace_layout")
    return out
