"""Deterministic (batch-invariant) mode: a request's output must not depend on what else is in the batch
(python/sglang/srt/layers/attention/triton_backend.py:286-304,377-386: fixed 256-key decode split tiles;
test/registered/attention/test_deterministic.py).  Bit-identical, not merely close."""
import types

import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import (B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention,
                         ReqToTokenPool)
from tests.util import assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"
Hq, Hkv, D, PAGE = 32, 8, 128, 16


def _backend(sc, deterministic):
    req_pool = ReqToTokenPool(sc.req_to_token.shape[0] - 1, sc.req_to_token.shape[1], DEV)
    req_pool.req_to_token.copy_(sc.req_to_token)
    kv_pool = MHATokenToKVPool(sc.k_cache.shape[0] - PAGE, PAGE, sc.k_cache.dtype, Hkv, D, 1, DEV)
    kv_pool.k_buffer[0].copy_(sc.k_cache)
    kv_pool.v_buffer[0].copy_(sc.v_cache)
    runner = types.SimpleNamespace(device=DEV, page_size=PAGE, req_to_token_pool=req_pool, token_to_kv_pool=kv_pool,
                                   sliding_window_size=None,
                                   server_args=types.SimpleNamespace(enable_deterministic_inference=deterministic))
    return B200AttnBackend(runner)


def _decode(sc, backend, rows):
    rt = torch.tensor(rows)
    lens = torch.tensor([sc.seq_lens[b] for b in rows])
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=len(rows),
                      req_pool_indices=sc.req_pool_indices[rt].to(DEV), seq_lens=lens.to(DEV),
                      out_cache_loc=sc.out_cache_loc[rt].to(DEV), seq_lens_sum=int(lens.sum()), seq_lens_cpu=lens,
                      attn_backend=backend)
    backend.init_forward_metadata(fb)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    return layer(sc.q[rt].to(DEV), sc.k[rt].to(DEV), sc.v[rt].to(DEV), fb)


def test_decode_is_batch_invariant_in_deterministic_mode():
    g = torch.Generator().manual_seed(7)
    lens = [3000, 17, 4096, 255, 256, 257, 1500] + torch.randint(1, 4097, (33,), generator=g).tolist()
    sc = make_scenario([l - 1 for l in lens], [1] * len(lens), Hq, Hkv, D, PAGE, seed=701)
    backend = _backend(sc, True)
    assert backend.deterministic
    whole = _decode(sc, backend, list(range(len(lens))))
    ref = oracle.attention_fp32(sc.q.to(DEV), backend.token_to_kv_pool.k_buffer[0], backend.token_to_kv_pool.v_buffer[0],
                                sc.req_to_token.to(DEV), sc.req_pool_indices.to(DEV), torch.tensor(lens),
                                torch.ones(len(lens), dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
    assert_close_attn(whole, ref, "deterministic decode, whole batch")
    for rows in ([0], [2], [5, 0], [6, 3, 2, 1], list(range(0, len(lens), 3))):
        part = _decode(sc, backend, rows)
        assert torch.equal(part, whole[torch.tensor(rows)]), f"rows {rows}: output depends on the batch"


def test_extend_is_batch_invariant():
    prefix, ext = [300, 0, 77, 2048, 5], [200, 333, 140, 512, 1000]
    sc = make_scenario(prefix, ext, Hq, Hkv, D, PAGE, seed=702)
    backend = _backend(sc, True)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    starts = [0]
    for e in ext:
        starts.append(starts[-1] + e)

    def run(rows):
        idx = torch.cat([torch.arange(starts[b], starts[b + 1]) for b in rows])
        loc = torch.cat([sc.out_cache_loc[starts[b]:starts[b + 1]] for b in rows])
        fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=len(rows),
                          req_pool_indices=sc.req_pool_indices[torch.tensor(rows)].to(DEV),
                          seq_lens=torch.tensor([sc.seq_lens[b] for b in rows], device=DEV), out_cache_loc=loc.to(DEV),
                          seq_lens_sum=sum(sc.seq_lens[b] for b in rows),
                          extend_seq_lens=torch.tensor([ext[b] for b in rows], dtype=torch.int32, device=DEV),
                          extend_prefix_lens=torch.tensor([prefix[b] for b in rows], dtype=torch.int32, device=DEV),
                          extend_seq_lens_cpu=[ext[b] for b in rows], extend_prefix_lens_cpu=[prefix[b] for b in rows],
                          attn_backend=backend)
        backend.init_forward_metadata(fb)
        return layer(sc.q[idx].to(DEV), sc.k[idx].to(DEV), sc.v[idx].to(DEV), fb)

    whole = run([0, 1, 2, 3, 4])
    for rows in ([3], [1, 4], [4, 0]):
        part = run(rows)
        want = torch.cat([whole[starts[b]:starts[b + 1]] for b in rows])
        assert torch.equal(part, want), f"extend rows {rows}: output depends on the batch"
