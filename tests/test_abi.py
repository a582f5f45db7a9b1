"""The C-ABI library loads and exports every symbol include/b200attn.h declares (no compute)."""
import ctypes
import os

from sglang_b200 import _lib


def test_header_symbols_are_typed():
    assert set(_lib.header_symbols()) == set(_lib.SIGNATURES), "keep _lib.SIGNATURES in sync with the header"


def test_library_exports_every_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build with: make -C sglang_b200/csrc"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _lib.header_symbols():
        assert hasattr(lib, name), f"{name} not exported"


def test_abi_version_and_error_string():
    lib = _lib.load()
    assert lib.b200attn_abi_version() == _lib.ABI_VERSION == 9
    assert isinstance(lib.b200attn_last_error(), bytes)
    assert lib.b200attn_decode_workspace_bytes(4, 32, 128, 1) == 0
    assert lib.b200attn_decode_workspace_bytes(4, 32, 128, 3) == 4 * 32 * 3 * 129 * 4 + 4 * 32 * 4


def test_argument_validation_without_gpu():
    lib = _lib.load()
    # negative T is rejected before any CUDA call
    rc = lib.b200attn_kv_store(None, None, None, None, None, 0, -1, 16, 16, 16, 16, 16, 16, 8, 0, None, None)
    assert rc == 1 and b"negative" in lib.b200attn_last_error()
    rc = lib.b200attn_merge_state(None, None, None, None, None, None, 0, 1, 8, 0, None)
    assert rc == 0  # empty problem is a no-op


def test_ops_refuse_cpu_tensors():
    import pytest
    import torch

    from sglang_b200 import ops

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.store_cache(torch.zeros(2, 8), torch.zeros(2, 8), torch.zeros(4, 8), torch.zeros(4, 8),
                        torch.tensor([1, 2]))
