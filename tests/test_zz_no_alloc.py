"""b200attn_decode_noalloc (no counterpart in the reference): the decode
kernel with its split-KV scratch in a static __device__ array inside the library, for callers that cannot allocate.

Called through the C ABI (ctypes), compared with b200attn_decode + a caller workspace on the same inputs, on two calls
in a row (the second call finds the first call's partials and counters in the shared array), with a different shape
in between, and under torch's allocator being watched: the entry point must not allocate.  Collected last on purpose
(the file name)."""
import ctypes

import pytest
import torch

from sglang_b200 import _lib, ops
from tests.util import assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _inputs(lens, Hq, Hkv, D, page, seed):
    B = len(lens)
    sc = make_scenario([l - 1 for l in lens], [1] * B, Hq, Hkv, D, page, seed=seed).to(DEV)
    kc, vc = sc.k_cache, sc.v_cache
    ops.store_cache(sc.k, sc.v, kc, vc, sc.out_cache_loc, reserved_skip_index=0)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(sc.req_to_token, sc.req_pool_indices, torch.tensor(lens, device=DEV), kv_indptr, kv_indices)
    q = sc.q.view(B, Hq, D).contiguous()
    return sc, q, kc, vc, kv_indptr, kv_indices


def _call(lib, name, q, kc, vc, o, lse, kv_indptr, kv_indices, scaling, max_splits, warps, workspace=None):
    B, Hq, D = q.shape
    Hkv = kc.shape[1]
    head = [q.data_ptr(), q.stride(0), kc.data_ptr(), vc.data_ptr(), kc.stride(0), o.data_ptr(), o.stride(0),
            lse.data_ptr(), kv_indptr.data_ptr(), kv_indices.data_ptr(), None, None, B, Hq, Hkv, D,
            0 if q.dtype == torch.bfloat16 else 1, ctypes.c_float(scaling), max_splits, 64, -1, warps, -1,
            ctypes.c_float(0.0), None, None]
    stream = torch.cuda.current_stream().cuda_stream
    if name == "b200attn_decode":
        return lib.b200attn_decode(*head, workspace.data_ptr(), workspace.numel(), stream)
    return lib.b200attn_decode_noalloc(*head, stream)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_noalloc_decode_matches_decode_two_calls_in_a_row(dtype):
    lib = _lib.load()
    assert lib.b200attn_decode_noalloc_max_workspace_bytes() == 65536 * 4 + (8 << 20)
    shapes = [([1, 17, 300, 1500, 4096, 777], 32, 8, 128, 16, 8, 4), ([2500, 64, 65], 14, 2, 64, 1, 16, 2),
              ([1, 17, 300, 1500, 4096, 777], 32, 8, 128, 16, 8, 4)]
    for n, (lens, Hq, Hkv, D, page, ms, warps) in enumerate(shapes):
        sc, q, kc, vc, kv_indptr, kv_indices = _inputs(lens, Hq, Hkv, D, page, seed=900 + n)
        q, kc, vc = q.to(dtype), kc.to(dtype), vc.to(dtype)
        B = len(lens)
        ws = torch.zeros(ops.decode_workspace_bytes(B, Hq, D, ms), dtype=torch.uint8, device=DEV)
        o_ref = torch.empty_like(q)
        lse_ref = torch.empty((B, Hq), dtype=torch.float32, device=DEV)
        rc = _call(lib, "b200attn_decode", q, kc, vc, o_ref, lse_ref, kv_indptr, kv_indices, sc.scaling, ms, warps, ws)
        assert rc == 0, lib.b200attn_last_error()
        for rep in range(2):  # two calls in a row on the shared scratch
            o = torch.full_like(q, float("nan"))
            lse = torch.full_like(lse_ref, float("nan"))
            torch.cuda.synchronize()
            before = torch.cuda.memory_stats()["allocation.all.allocated"]
            rc = _call(lib, "b200attn_decode_noalloc", q, kc, vc, o, lse, kv_indptr, kv_indices, sc.scaling, ms, warps)
            assert rc == 0, lib.b200attn_last_error()
            assert torch.cuda.memory_stats()["allocation.all.allocated"] == before  # torch's allocator was not touched
            torch.cuda.synchronize()
            assert_close_attn(o.view(B, -1), o_ref.view(B, -1), f"noalloc decode shape {n} call {rep}")
            torch.testing.assert_close(lse, lse_ref, atol=1e-3, rtol=1e-3)


def test_noalloc_decode_rejects_shapes_beyond_the_builtin_scratch():
    lib = _lib.load()
    lens = [300] * 64
    sc, q, kc, vc, kv_indptr, kv_indices = _inputs(lens, 32, 8, 128, 16, seed=950)
    o = torch.empty_like(q)
    lse = torch.empty((64, 32), dtype=torch.float32, device=DEV)
    # 64 * 32 * 16 * 129 * 4 B = 16.9 MB of partials > 8 MiB
    rc = _call(lib, "b200attn_decode_noalloc", q, kc, vc, o, lse, kv_indptr, kv_indices, sc.scaling, 16, 4)
    assert rc == 1 and b"built-in" in lib.b200attn_last_error()
    # max_splits = 1 needs no scratch at all and always fits
    rc = _call(lib, "b200attn_decode_noalloc", q, kc, vc, o, lse, kv_indptr, kv_indices, sc.scaling, 1, 4)
    assert rc == 0, lib.b200attn_last_error()
    ws = torch.zeros(16, dtype=torch.uint8, device=DEV)
    o_ref = torch.empty_like(q)
    rc = _call(lib, "b200attn_decode", q, kc, vc, o_ref, lse, kv_indptr, kv_indices, sc.scaling, 1, 4, ws)
    assert rc == 0
    torch.cuda.synchronize()
    assert_close_attn(o.view(64, -1), o_ref.view(64, -1), "noalloc decode, one split")
