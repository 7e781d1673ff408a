"""ctypes loader for libb200attn.so (the C ABI declared in include/b200attn.h).

There is NO fallback: if the library is missing or does not export a declared symbol the
import of the product package fails loudly (north_star: "no CPU fallback").
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200attn.so")
CSRC_DIR = os.path.join(_HERE, "csrc")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "b200attn.h")

_c = ctypes
_vp, _i32, _i64, _sz, _f32 = _c.c_void_p, _c.c_int32, _c.c_int64, _c.c_size_t, _c.c_float

# name -> (restype, argtypes); must list every function of include/b200attn.h
SIGNATURES = {
    "b200_abi_version": (_i32, []),
    "b200_last_error": (_c.c_char_p, []),
    "b200_num_sms": (_i32, []),
    "b200_store_kv": (_i32, [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i32, _i32, _i32, _vp]),
    "b200_slot_table_update": (_i32, [_vp, _i64, _vp, _i32, _i32, _vp]),
    "b200_build_meta": (_i32, [_vp, _i64, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _i64, _vp, _vp,
                               _i64, _vp, _vp]),
    "b200_decode_workspace_bytes": (_sz, [_i32, _i32, _i32, _i32]),
    "b200_decode": (_i32, [_vp, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _i32, _i32,
                           _i32, _i32, _i32, _f32, _i32, _vp]),
    "b200_decode_noalloc": (_i32, [_vp, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32,
                                   _i32, _f32, _i32, _vp]),
    "b200_prefill_workspace_bytes": (_sz, [_i32, _i32, _i32, _i32, _i32]),
    "b200_prefill": (_i32, [_vp, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _i32,
                            _i32, _i32, _i32, _i32, _i32, _f32, _i32, _i32, _vp]),
    "b200_qk_norm_rope_store": (_i32, [_vp, _i64, _vp, _vp, _vp, _f32, _vp, _i64, _i32, _vp, _i64, _vp, _i64,
                                       _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200_sample": (_i32, [_vp, _i64, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _i64, _c.c_uint64, _c.c_uint64, _vp, _vp]),
    "b200_launch_count": (_i64, []),
    "b200_reset_launch_count": (None, []),
}

_lib = None


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a with nvcc (cross-compiles without a GPU)."""
    res = subprocess.run(["make", "-C", CSRC_DIR, "-j8"], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("building libb200attn.so failed:\n" + res.stdout[-4000:] + res.stderr[-4000:])
    if verbose:
        print(res.stdout[-2000:])
    return LIB_PATH


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C {CSRC_DIR}` (or __graft_entry__.build()). "
            "The b200attn ops have no CPU/eager fallback by design.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # pragma: no cover
            raise ImportError(f"libb200attn.so does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    if lib.b200_abi_version() != 1:
        raise ImportError("libb200attn.so ABI version mismatch")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed: {msg}")
