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
    assert lib.b200attn_abi_version() == _lib.ABI_VERSION == 21
    assert isinstance(lib.b200attn_last_error(), bytes)
    assert lib.b200attn_decode_workspace_bytes(4, 32, 128, 1) == 0
    # fixed-size arrival-counter header (65536 int32) + fp32 partials and their LSE
    assert lib.b200attn_decode_workspace_bytes(4, 32, 128, 3) == 65536 * 4 + 4 * 32 * 3 * 129 * 4
    assert lib.b200attn_decode_workspace_bytes(4, 32, 80, 3) == lib.b200attn_decode_workspace_bytes(4, 32, 128, 3)
    assert lib.b200attn_decode_ctas_per_sm(0, 2) == 4 and lib.b200attn_decode_ctas_per_sm(0, 4) == 2


def test_argument_validation_without_gpu():
    lib = _lib.load()
    # negative T is rejected before any CUDA call
    rc = lib.b200attn_kv_store(None, None, None, None, None, 0, -1, 16, 16, 16, 16, 16, 16, 8, 0, None, None)
    assert rc == 1 and b"negative" in lib.b200attn_last_error()
    rc = lib.b200attn_merge_state(None, None, None, None, None, None, 0, 1, 8, 0, None)
    assert rc == 0  # empty problem is a no-op
    # the fp8 / stream-k entry points validate the same way (host checks first, nothing launched)
    rc = lib.b200attn_kv_store_fp8(None, None, None, None, None, 0, -1, 16, 16, 16, 16, 16, 16, 0, 1.0, 1.0, 8, 0,
                                   None, None)
    assert rc == 1 and b"negative" in lib.b200attn_last_error()
    rc = lib.b200attn_kv_store_fp8(None, None, None, None, None, 0, 4, 16, 16, 16, 16, 16, 16, 0, 1.0, 1.0, 8, 0,
                                   None, None)
    assert rc == 1 and b"null pointer" in lib.b200attn_last_error()
    rc = lib.b200attn_kv_dequant_gather(None, None, 16, None, 0, 16, None, None, 16, 0, 1.0, 1.0, None, None)
    assert rc == 0
    rc = lib.b200attn_decode_sk(None, 0, None, None, 0, None, 0, None, None, None, None, None, 4, 32, 8, 128, 0,
                                1.0, 256, -1, 0.0, None, None, 0, None)
    assert rc == 1 and b"null pointer" in lib.b200attn_last_error()
    rc = lib.b200attn_decode_fp8kv(None, 0, None, None, 0, None, 0, None, None, None, None, None, 0, 32, 8, 128, 0,
                                   1.0, 1.0, 1.0, 1, 256, -1, -1, 0.0, None, None, 0, None)
    assert rc == 0  # empty batch
    assert lib.b200attn_decode_sk_workspace_bytes(0, 32, 8, 128) == 0


def test_ops_refuse_cpu_tensors():
    import pytest
    import torch

    from sglang_b200 import ops

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.store_cache(torch.zeros(2, 8), torch.zeros(2, 8), torch.zeros(4, 8), torch.zeros(4, 8),
                        torch.tensor([1, 2]))


def test_every_declared_entry_point_is_in_the_integration_index():
    """INTEGRATION.md §5 is the maintainer's map from C-ABI entry points to reference call sites: it must name every
    function include/b200attn.h declares."""
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "b200attn.h")).read()
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    declared = sorted(set(re.findall(r"\b(b200attn_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 30
    missing = [s for s in declared if s not in doc]
    assert not missing, f"INTEGRATION.md does not mention: {missing}"
