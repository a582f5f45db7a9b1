"""ctypes binding of libb200attn.so (the C ABI declared in include/b200attn.h).

There is deliberately no fallback: if the shared library is missing or a call
fails, the error is raised to the caller.  The library is built in-tree by
``sglang_b200/csrc/Makefile`` (see ``__graft_entry__.build``).
"""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import Dict, List

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200attn.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "b200attn.h")

ABI_VERSION = 21  # b200attn_abi_version() of the library this binding matches (api.cu)
BF16, FP16, FP32 = 0, 1, 2
I32, I64 = 0, 1

_vp, _i32, _i64, _f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float

class RadixTable(C.Structure):
    """b200attn_radix_t (include/b200attn.h): device pointers + sizes of the page-trie hash table."""

    _fields_ = [("parent", _vp), ("tokens", _vp), ("values", _vp), ("ehash", _vp), ("tkeys", _vp), ("tvals", _vp),
                ("rev", _vp), ("freelist", _vp), ("ctrs", _vp), ("cap", _i32), ("tcap", _i32), ("page", _i32),
                ("num_pages", _i32)]


# name -> (restype, argtypes); must match include/b200attn.h
SIGNATURES = {
    "b200attn_abi_version": (C.c_int, []),
    "b200attn_last_error": (C.c_char_p, []),
    "b200attn_sm_count": (C.c_int, []),
    "b200attn_kv_store": (
        C.c_int,
        [_vp, _vp, _vp, _vp, _vp, C.c_int, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _vp, _vp],
    ),
    "b200attn_build_kv_indices": (
        C.c_int,
        [_vp, _i64, _vp, C.c_int, _vp, C.c_int, _vp, C.c_int, _vp, C.c_int, _i32, _vp, _vp, C.c_int, _i64, _vp, _vp,
         C.c_int, _vp],
    ),
    "b200attn_translate_slots": (C.c_int, [_vp, C.c_int, _vp, C.c_int, _i64, _vp, _vp]),
    "b200attn_stage_copy": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _vp]),
    "b200attn_allreduce_flag_bytes": (_i64, []),
    "b200attn_allreduce_small_bytes": (_i64, []),
    "b200attn_allreduce_nvls": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i64, C.c_int, _vp]),
    "b200attn_build_page_table": (
        C.c_int,
        [_vp, _i64, _vp, C.c_int, _vp, C.c_int, _i32, _i32, _vp, _i64, _i32, _vp],
    ),
    "b200attn_decode_workspace_bytes": (_i64, [_i32, _i32, _i32, _i32]),
    "b200attn_decode_ctas_per_sm": (C.c_int, [C.c_int, C.c_int]),
    "b200attn_decode_noalloc_max_workspace_bytes": (_i64, []),
    "b200attn_decode_noalloc": (
        C.c_int,
        [_vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, C.c_int, C.c_float,
         _i32, _i32, _i32, _i32, _i32, C.c_float, _vp, _vp, _vp],
    ),
    "b200attn_decode": (
        C.c_int,
        [_vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, C.c_int,
         _f32, _i32, _i32, _i32, _i32, _i32, _f32, _vp, _vp, _vp, _i64, _vp],
    ),
    "b200attn_build_token_items": (C.c_int, [_vp, _vp, _i32, _i64, C.c_int, _vp, _vp, _vp]),
    "b200attn_decode_sk_workspace_bytes": (_i64, [_i32, _i32, _i32, _i32]),
    "b200attn_decode_sk": (
        C.c_int,
        [_vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, C.c_int,
         _f32, _i32, _i32, _f32, _vp, _vp, _i64, _vp],
    ),
    "b200attn_decode_ws_workspace_bytes": (_i64, [_i32, _i32, _i32]),
    "b200attn_decode_ws": (
        C.c_int,
        [_vp, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, C.c_int,
         _f32, _i32, _i32, _f32, _vp, _vp, _vp, _i64, _i64, _vp, _i64, _vp],
    ),
    "b200attn_decode_sk_fp8kv": (
        C.c_int,
        [_vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, C.c_int,
         _f32, _f32, _f32, _i32, _i32, _f32, _vp, _vp, _i64, _vp],
    ),
    "b200attn_decode_fp8kv": (
        C.c_int,
        [_vp, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, C.c_int,
         _f32, _f32, _f32, _i32, _i32, _i32, _i32, _f32, _vp, _vp, _i64, _vp],
    ),
    "b200attn_kv_store_fp8": (
        C.c_int,
        [_vp, _vp, _vp, _vp, _vp, C.c_int, _i64, _i64, _i64, _i64, _i64, _i64, _i64, C.c_int, _f32, _f32,
         _i64, _i64, _vp, _vp],
    ),
    "b200attn_kv_dequant_gather": (
        C.c_int,
        [_vp, _vp, _i64, _vp, _i64, _i64, _vp, _vp, _i64, C.c_int, _f32, _f32, _vp, _vp],
    ),
    "b200attn_kv_dequant_gather_unique": (
        C.c_int,
        [_vp, _vp, _i64, _vp, _i64, _i64, _vp, _vp, _i64, C.c_int, _f32, _f32, _vp, _vp, _i64, _vp, _vp],
    ),
    "b200attn_extend_workspace_bytes": (_i64, [_i32, _i64, _i32]),
    "b200attn_extend": (
        C.c_int,
        [_vp, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _i32, _i64, _i32, _i32, _i32,
         _i32, _i32, C.c_int, C.c_int, _f32, _i32, _f32, _vp, _vp, _vp, C.c_int, _vp, _i64, _vp],
    ),
    "b200attn_extend_fp8kv": (
        C.c_int,
        [_vp, _i64, _vp, _vp, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _i32, _i64, _i32, _i32, _i32,
         _i32, _i32, C.c_int, C.c_int, _f32, _f32, _f32, _i32, _f32, _vp, _vp, _vp, C.c_int, _vp, _i64, _vp],
    ),
    "b200attn_merge_state": (
        C.c_int,
        [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, C.c_int, _vp],
    ),
    "b200attn_radix_match": (C.c_int, [C.POINTER(RadixTable), _vp, C.c_int, _vp, _i32, _vp, _i64, _vp, _vp]),
    "b200attn_radix_insert": (C.c_int, [C.POINTER(RadixTable), _vp, C.c_int, _vp, _vp, _i32, _vp, _vp]),
    "b200attn_radix_erase_slots": (C.c_int, [C.POINTER(RadixTable), _vp, C.c_int, _i64, _vp]),
    "b200attn_alloc_extend": (
        C.c_int, [_vp, _vp, _vp, C.c_int, _vp, C.c_int, _i64, _vp, C.c_int, _i32, _i32, _vp, _vp],
    ),
    "b200attn_alloc_decode": (
        C.c_int, [_vp, _vp, C.c_int, _vp, C.c_int, _i64, _vp, C.c_int, _i32, _i32, _vp, _vp],
    ),
    "b200attn_rope_store": (
        C.c_int,
        [_vp, _vp, _vp, _vp, _vp, _vp, C.c_int, _vp, C.c_int, _i64, _vp, C.c_int, _i64, _vp, C.c_int, _i64,
         _i32, _i32, _i32, _i32, _i32, C.c_int, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _vp, C.c_int, C.c_int,
         _vp],
    ),
    "b200attn_rope_store_fp8": (
        C.c_int,
        [_vp, _vp, _vp, _vp, _vp, _vp, C.c_int, _vp, C.c_int, _i64, _vp, C.c_int, _i64, _vp, C.c_int, _i64,
         _i32, _i32, _i32, _i32, _i32, C.c_int, _i64, _i64, _i64, _i64, _i64, _i64, _i64, _vp, C.c_int, C.c_int,
         _f32, _f32, _vp],
    ),
}


class B200AttnError(RuntimeError):
    """A C-ABI call returned nonzero (message from b200attn_last_error)."""


_lib = None


def header_symbols() -> List[str]:
    """Every function name declared in include/b200attn.h."""
    with open(HEADER_PATH) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200attn_[a-z0-9_]+)\s*\(", text)))


def load() -> C.CDLL:
    """Load the shared library (once) and type every entry point."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C sglang_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). "
            "There is no CPU / PyTorch fallback for this backend."
        )
    lib = C.CDLL(LIB_PATH)
    # a stale .so (it is git-ignored and ships out of band) would be called with mismatched argtypes: refuse it
    # before any signature is typed or any call is made
    lib.b200attn_abi_version.restype = C.c_int
    lib.b200attn_abi_version.argtypes = []
    found = lib.b200attn_abi_version()
    if found != ABI_VERSION:
        raise ImportError(f"{LIB_PATH} has ABI version {found}, this binding needs {ABI_VERSION}: rebuild it with "
                          "`make -C sglang_b200/csrc`")
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def load_variant(path: str) -> C.CDLL:
    """A second build of the same C ABI (e.g. the checker build libb200attn_serial.so), typed like the product one."""
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.b200attn_abi_version() != ABI_VERSION:
        raise ImportError(f"{path}: ABI version {lib.b200attn_abi_version()} != {ABI_VERSION}")
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200attn_last_error().decode("utf-8", "replace")
        kind = {1: "invalid argument", 2: "unsupported", 3: "CUDA error"}.get(rc, f"error {rc}")
        if rc == 2:
            raise NotImplementedError(f"{what}: {msg}")
        raise B200AttnError(f"{what}: {kind}: {msg}")


def exported() -> Dict[str, bool]:
    lib = load()
    return {name: hasattr(lib, name) for name in header_symbols()}
