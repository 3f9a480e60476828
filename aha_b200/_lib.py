"""ctypes binding of libaha_b200.so (include/aha_b200.h) and its in-tree build.

The library is the product; there is no Python or CPU fallback: every scan entry point raises
when the shared object or a CUDA device is missing.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# AHA_B200_SO: load another build of the library (tools/ab.py compares builds with different -D flags on one box)
SO_PATH = os.environ.get("AHA_B200_SO") or os.path.join(_HERE, "libaha_b200.so")
CSRC = os.path.join(_HERE, "csrc")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-fwrapv,-Wno-deprecated-declarations", "-Wno-deprecated-declarations", "-shared",
]


def _sources():
    out = [os.path.join(_HERE, "..", "include", "aha_b200.h"), os.path.join(_HERE, "..", "include", "aha_synth.h")]
    for f in sorted(os.listdir(CSRC)):
        out.append(os.path.join(CSRC, f))
    return out


def _units():
    """translation units of the library: the scan paths (api.cu) and the allocation-free entry point (noalloc.cu)"""
    return [os.path.join(CSRC, "api.cu"), os.path.join(CSRC, "noalloc.cu")]


def build(force=False, verbose=False, out=None, extra_flags=()):
    """nvcc cross-compiles for sm_100a (no GPU needed).  Rebuilds when any source is newer."""
    if out is not None:
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        subprocess.check_call([nvcc] + NVCC_FLAGS + list(extra_flags) + ["-o", out] + _units(),
                              stderr=subprocess.DEVNULL)
        return out
    if os.environ.get("AHA_B200_SO"):
        return SO_PATH
    if not force and os.path.exists(SO_PATH):
        t = os.path.getmtime(SO_PATH)
        if all(os.path.getmtime(s) <= t for s in _sources()):
            return SO_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("AHA_B200_NVCC_EXTRA", "").split()  # e.g. -DAHA_B200_PHASES (per-phase cycle counters on stderr)
    cmd = [nvcc] + NVCC_FLAGS + extra + ["-o", SO_PATH] + _units()
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return SO_PATH


class Hit(C.Structure):
    _fields_ = [("start", C.c_int32), ("end", C.c_int32), ("value", C.c_int32)]


class Hits(C.Structure):
    _fields_ = [
        ("hits", C.c_void_p), ("n_hits", C.c_int64), ("doc_hit_off", C.c_void_p), ("n_docs", C.c_int64),
        ("on_device", C.c_int32), ("device", C.c_int32), ("scan_ms", C.c_double), ("walk_ms", C.c_double), ("expand_ms", C.c_double),
        ("opaque", C.c_void_p),
    ]


EXPORTS = [
    "aha_b200_init", "aha_b200_last_error", "aha_b200_version", "aha_b200_kernel_launches", "aha_b200_builder_new", "aha_b200_builder_insert",
    "aha_b200_builder_lookup", "aha_b200_builder_key", "aha_b200_builder_compile", "aha_b200_builder_free",
    "aha_b200_dfa_from_cedar", "aha_b200_dfa_compile", "aha_b200_flatten_host", "aha_b200_blob_free",
    "aha_b200_dfa_blob", "aha_b200_dfa_from_blob", "aha_b200_dfa_from_device_blob", "aha_b200_dfa_info",
    "aha_b200_dfa_free", "aha_b200_scan", "aha_b200_scan_device", "aha_b200_hits_free",
    "aha_b200_gen_text_host", "aha_b200_gen_text_device", "aha_b200_debug_char_scan", "aha_b200_debug_set",
    "aha_b200_scan_device_noalloc", "aha_b200_debug_folded_longest",
]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()  # no-op when the library is newer than every source: a stale binary is never loaded silently
    L = C.CDLL(SO_PATH)
    vp, i32, i64, u32, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_uint64
    P = C.POINTER
    L.aha_b200_last_error.restype = C.c_char_p
    L.aha_b200_kernel_launches.restype = C.c_int64
    L.aha_b200_init.argtypes = [P(i32), i32]
    L.aha_b200_builder_new.argtypes = [P(vp)]
    L.aha_b200_builder_insert.argtypes = [vp, C.c_char_p, i32, P(i32)]
    L.aha_b200_builder_lookup.argtypes = [vp, C.c_char_p, i32, P(i32)]
    L.aha_b200_builder_key.argtypes = [vp, i32, C.c_char_p, i32, P(i32)]
    L.aha_b200_builder_compile.argtypes = [vp, P(vp), P(i64), P(vp), P(vp), P(vp), P(i64)]
    L.aha_b200_builder_free.argtypes = [vp]
    L.aha_b200_builder_free.restype = None
    L.aha_b200_dfa_from_cedar.argtypes = [vp, i64, vp, vp, vp, i64, P(vp)]
    L.aha_b200_dfa_compile.argtypes = [vp, vp, i64, P(vp)]
    L.aha_b200_flatten_host.argtypes = [vp, i64, vp, vp, vp, i64, P(vp), P(i64)]
    L.aha_b200_blob_free.argtypes = [vp]
    L.aha_b200_blob_free.restype = None
    L.aha_b200_dfa_blob.argtypes = [vp, P(vp), P(i64)]
    L.aha_b200_dfa_from_blob.argtypes = [vp, i64, P(vp)]
    L.aha_b200_dfa_from_device_blob.argtypes = [vp, i64, P(vp)]
    L.aha_b200_dfa_info.argtypes = [vp, i32]
    L.aha_b200_dfa_info.restype = i64
    L.aha_b200_dfa_free.argtypes = [vp]
    L.aha_b200_dfa_free.restype = None
    L.aha_b200_scan.argtypes = [vp, vp, vp, i64, u32, C.c_char_p, i32, P(P(Hits))]
    L.aha_b200_scan_device.argtypes = [vp, vp, vp, i64, i64, u32, C.c_char_p, i32, vp, P(P(Hits))]
    L.aha_b200_scan_device_noalloc.argtypes = [vp, vp, vp, i64, i64, vp, i64, vp, vp, vp]
    L.aha_b200_hits_free.argtypes = [P(Hits)]
    L.aha_b200_hits_free.restype = None
    L.aha_b200_debug_char_scan.argtypes = [vp, i64, vp, i64, i32, i64, vp, i64, P(i64), P(i64)]
    L.aha_b200_debug_folded_longest.argtypes = [vp, i64, vp, i64, u32, C.c_char_p, i32, vp, i64, P(i64), P(i64)]
    L.aha_b200_debug_set.argtypes = [i32, i32]
    L.aha_b200_gen_text_host.argtypes = [i32, u64, i64, i64, vp]
    L.aha_b200_gen_text_device.argtypes = [i32, u64, i64, i64, vp, vp]
    _lib = L
    return L


class AhaError(RuntimeError):
    """Raised where the reference raises a String (ac.cr:66,322; cedar.cr:235,756) or on CUDA errors."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def check(rc):
    if rc != 0:
        raise AhaError(rc, lib().aha_b200_last_error().decode("utf-8", "replace"))
