"""Hybrid sliding-window KV pools: SWA layers live in a smaller pool addressed through
`full_to_swa_index_mapping` (python/sglang/srt/mem_cache/swa_memory_pool.py:20-270); the backend translates the
read indices and the write slots (KVWriteLoc.swa_loc, memory_pool.py:1550-1590) once per forward
(triton_backend.py:978-982).  Layer 0 = full attention, layer 1 = sliding window."""
import types

import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import B200AttnBackend, ForwardBatch, ForwardMode, RadixAttention, ReqToTokenPool, SWAKVPool
from tests.util import assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"
Hq, Hkv, D, PAGE, W = 8, 2, 128, 16, 100


def _setup(sc):
    slots = sc.k_cache.shape[0]
    req_pool = ReqToTokenPool(sc.req_to_token.shape[0] - 1, sc.req_to_token.shape[1], DEV)
    req_pool.req_to_token.copy_(sc.req_to_token)
    # the SWA pool is a permuted copy of the full pool: swa slot = perm[full slot] (page granular, page 0 fixed)
    g = torch.Generator().manual_seed(5)
    pages = slots // PAGE
    pperm = torch.cat([torch.zeros(1, dtype=torch.int64), torch.randperm(pages - 1, generator=g) + 1])
    mapping = (pperm[:, None] * PAGE + torch.arange(PAGE)[None]).reshape(-1)
    mapping = torch.cat([mapping, torch.tensor([-1])])  # sentinel (swa_memory_pool.py:230)
    pool = SWAKVPool(slots - PAGE, slots - PAGE, PAGE, sc.k_cache.dtype, Hkv, D, [1], [0], DEV)
    pool.full_to_swa_index_mapping = mapping.to(DEV)
    pool.full_kv_pool.k_buffer[0].copy_(sc.k_cache)
    pool.full_kv_pool.v_buffer[0].copy_(sc.v_cache)
    inv = torch.empty(slots, dtype=torch.int64)
    inv[mapping[:-1]] = torch.arange(slots)
    # different CONTENT in the SWA pool (another layer's K/V), laid out through the mapping
    swa_k = torch.randn(sc.k_cache.shape, generator=g).to(sc.k_cache.dtype)
    swa_v = torch.randn(sc.v_cache.shape, generator=g).to(sc.v_cache.dtype)
    swa_k[:PAGE] = 0
    swa_v[:PAGE] = 0
    pool.swa_kv_pool.k_buffer[0].copy_(swa_k[inv])
    pool.swa_kv_pool.v_buffer[0].copy_(swa_v[inv])
    runner = types.SimpleNamespace(device=DEV, page_size=PAGE, req_to_token_pool=req_pool, token_to_kv_pool=pool,
                                   sliding_window_size=W)
    return B200AttnBackend(runner, skip_padding_slot=False), pool, swa_k, swa_v, mapping


@pytest.mark.parametrize("mode", ["EXTEND", "DECODE"])
def test_hybrid_swa_pool_layers(mode):
    if mode == "EXTEND":
        prefix, ext = [300, 0, 77], [200, 333, 140]
    else:
        prefix, ext = [99, 100, 101, 450], [1, 1, 1, 1]
    sc = make_scenario(prefix, ext, Hq, Hkv, D, PAGE, seed=41)
    backend, pool, swa_k, swa_v, mapping = _setup(sc)
    B = len(prefix)
    fb = ForwardBatch(forward_mode=ForwardMode[mode], batch_size=B, req_pool_indices=sc.req_pool_indices.to(DEV),
                      seq_lens=torch.tensor(sc.seq_lens, dtype=torch.int64, device=DEV),
                      out_cache_loc=sc.out_cache_loc.to(DEV), seq_lens_sum=sum(sc.seq_lens),
                      seq_lens_cpu=torch.tensor(sc.seq_lens), attn_backend=backend)
    if mode == "EXTEND":
        fb.extend_seq_lens = torch.tensor(ext, dtype=torch.int32, device=DEV)
        fb.extend_prefix_lens = torch.tensor(prefix, dtype=torch.int32, device=DEV)
        fb.extend_seq_lens_cpu, fb.extend_prefix_lens_cpu = list(ext), list(prefix)
    backend.init_forward_metadata(fb)
    md = backend.forward_metadata
    assert torch.equal(md.swa_kv_indices.cpu().long(), mapping[md.kv_indices.cpu().long()])
    assert torch.equal(md.swa_out_cache_loc.cpu(), mapping[sc.out_cache_loc])
    q, k, v = sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV)
    full = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    swa = RadixAttention(Hq, D, sc.scaling, Hkv, 1, sliding_window_size=W)
    o_full, o_swa = full(q, k, v, fb), swa(q, k, v, fb)
    torch.cuda.synchronize()
    # layer 0: the full pool, no window
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    oracle.store_kv(kc, vc, sc.out_cache_loc, sc.k, sc.v)
    lens, extl = torch.tensor(sc.seq_lens), torch.tensor(ext)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, lens, extl, Hq, Hkv, D,
                                sc.scaling, causal=True)
    assert_close_attn(o_full, ref, f"hybrid pool, full layer, {mode}")
    assert torch.equal(pool.full_kv_pool.k_buffer[0].cpu(), kc)
    # layer 1: the SWA pool's content (expressed in full-slot order), window W
    oracle.store_kv(swa_k, swa_v, sc.out_cache_loc, sc.k, sc.v)
    ref = oracle.attention_fp32(sc.q, swa_k, swa_v, sc.req_to_token, sc.req_pool_indices, lens, extl, Hq, Hkv, D,
                                sc.scaling, causal=True, window_left=W)
    assert_close_attn(o_swa, ref, f"hybrid pool, SWA layer, {mode}")
    inv = torch.empty(mapping.numel() - 1, dtype=torch.int64)
    inv[mapping[:-1]] = torch.arange(mapping.numel() - 1)
    assert torch.equal(pool.swa_kv_pool.k_buffer[0].cpu(), swa_k[inv])  # the write went through swa_loc
