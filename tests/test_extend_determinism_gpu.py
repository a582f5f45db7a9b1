"""Run-to-run determinism of the extend kernel and the PV -> QK TMEM-reuse ordering it relies on.

extend.cu reuses the TMEM columns of P_i(c) for S_i(c+2): the MMA issuer thread issues PV_i(c) (reads P as the A
operand from TMEM) and then QK_i(c+2) (writes the same columns) with no wait in between, relying on the tensor pipe
executing one thread's tcgen05.mma in issue order.  A violation would show as (rare) corrupted P values.  This test
  * launches the cfg5 chunked-prefill shard (prefix 8192 + 8192 new, 128 K/V tiles per item) 100 times and requires
    every output to be bit-identical, and
  * requires them to equal, bit for bit, the CHECKER build libb200attn_serial.so, in which the issuer waits for
    PV_i(c) to complete before issuing QK_i(c+2) (csrc/extend.cu, EXT_SERIAL_PVQK).
"""
import os

import pytest
import torch

from sglang_b200 import _lib, ops
from tests.util import make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"
SERIAL = os.path.join(os.path.dirname(_lib.LIB_PATH), "libb200attn_serial.so")


def _launch(sc, tensors, causal=True):
    q, kc, vc, o, qo_indptr, kv_indptr, kv_indices = tensors
    ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=causal,
                         page_size=sc.page_size, max_qo_len=max(sc.extend_lens))
    return o


def _prepare(sc):
    B = len(sc.seq_lens)
    kc, vc = sc.k_cache.to(DEV), sc.v_cache.to(DEV)
    ops.store_cache(sc.k.to(DEV), sc.v.to(DEV), kc, vc, sc.out_cache_loc.to(DEV), reserved_skip_index=0)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    qo_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(sc.seq_lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(sc.req_to_token.to(DEV), sc.req_pool_indices.to(DEV), torch.tensor(sc.seq_lens, device=DEV),
                         kv_indptr, kv_indices, qo_lens=torch.tensor(sc.extend_lens, dtype=torch.int32, device=DEV),
                         qo_indptr=qo_indptr)
    q = sc.q.to(DEV).view(-1, sc.Hq, sc.D)
    return [q, kc, vc, torch.empty_like(q), qo_indptr, kv_indptr, kv_indices]


@pytest.mark.parametrize("shape", ["cfg5_shard_hq8_hkv1", "ragged_mixed_lengths_g7"])
def test_extend_is_bit_deterministic_and_equals_serialized_build(shape):
    if not os.path.exists(SERIAL):
        pytest.skip("libb200attn_serial.so not built (make -C sglang_b200/csrc)")
    if shape == "cfg5_shard_hq8_hkv1":
        sc = make_scenario([8192], [8192], 8, 1, 128, 16, seed=501, pool_slack=1.05)
        reps = 100
    else:
        sc = make_scenario([0, 700, 33, 4000], [1280, 513, 2047, 96], 28, 4, 128, 16, seed=502, pool_slack=1.05)
        reps = 40
    t = _prepare(sc)
    first = _launch(sc, t).clone()
    torch.cuda.synchronize()
    assert torch.isfinite(first.float()).all()
    for i in range(reps - 1):
        t[3] = torch.empty_like(first)
        again = _launch(sc, t)
        assert torch.equal(again, first), f"launch {i + 1} differs from launch 0"
    product = _lib.load()
    try:
        _lib._lib = _lib.load_variant(SERIAL)
        t[3] = torch.empty_like(first)
        serial = _launch(sc, t).clone()
        torch.cuda.synchronize()
    finally:
        _lib._lib = product
    assert torch.equal(serial, first), "product build differs from the build that waits for PV before the next QK"
