"""Head dims other than 64 / 128 (the reference's kernel tests sweep 256, 128, 96, 80, 13:
test/registered/attention/test_triton_attention_kernels.py:169-322): every multiple of 8 up to 256 runs - decode on
the padded mma.sync kernel (64 / 128 / 256 instantiations), extend token-by-token through the same kernel (B200AttnBackend.
_forward_extend_generic_head_dim).  Parity vs the fp32 oracle at BASELINE's tolerance; the store stays bit-exact."""
import types

import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import (AttentionType, B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention,
                         ReqToTokenPool, ops)
from tests.util import assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _backend(sc):
    req_pool = ReqToTokenPool(sc.req_to_token.shape[0] - 1, sc.req_to_token.shape[1], DEV)
    req_pool.req_to_token.copy_(sc.req_to_token)
    kv_pool = MHATokenToKVPool(sc.k_cache.shape[0] - sc.page_size, sc.page_size, sc.k_cache.dtype, sc.Hkv, sc.D, 1, DEV)
    kv_pool.k_buffer[0].copy_(sc.k_cache)
    kv_pool.v_buffer[0].copy_(sc.v_cache)
    runner = types.SimpleNamespace(device=DEV, page_size=sc.page_size, req_to_token_pool=req_pool,
                                   token_to_kv_pool=kv_pool, sliding_window_size=None)
    return B200AttnBackend(runner, skip_padding_slot=False), kv_pool


@pytest.mark.parametrize("D", [16, 32, 80, 96, 112, 192, 256])
@pytest.mark.parametrize("Hq,Hkv,page", [(8, 2, 16), (14, 2, 1)])
def test_decode_generic_head_dim(D, Hq, Hkv, page):
    lens = [1, 17, 64, 65, 300, 1500, 4096]
    B = len(lens)
    sc = make_scenario([l - 1 for l in lens], [1] * B, Hq, Hkv, D, page, seed=D + Hq)
    backend, kv_pool = _backend(sc)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=B, req_pool_indices=sc.req_pool_indices.to(DEV),
                      seq_lens=torch.tensor(lens, device=DEV), out_cache_loc=sc.out_cache_loc.to(DEV),
                      seq_lens_sum=sum(lens), seq_lens_cpu=torch.tensor(lens), attn_backend=backend)
    backend.init_forward_metadata(fb)
    o = layer(sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV), fb)
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    oracle.store_kv(kc, vc, sc.out_cache_loc, sc.k, sc.v)
    assert torch.equal(kv_pool.k_buffer[0].cpu(), kc) and torch.equal(kv_pool.v_buffer[0].cpu(), vc)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens),
                                torch.ones(B, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
    assert_close_attn(o, ref, f"decode D={D}")
    # forced splits on the same shape
    d = sc.to(DEV)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, torch.tensor(lens, device=DEV), kv_indptr, kv_indices)
    for warps in (2, 4):
        o2 = torch.empty((B, Hq, D), dtype=torch.bfloat16, device=DEV)
        ws = torch.zeros(ops.decode_workspace_bytes(B, Hq, D, 8), dtype=torch.uint8, device=DEV)
        ops.decode_attention(sc.q.to(DEV).view(B, Hq, D), kv_pool.k_buffer[0], kv_pool.v_buffer[0], o2, kv_indptr,
                             kv_indices, sc.scaling, max_splits=8, min_chunk=64, workspace=ws, warps=warps)
        assert_close_attn(o2.view(B, -1), ref, f"decode D={D} 8 splits warps={warps}")


@pytest.mark.parametrize("D", [80, 96, 32, 256])
@pytest.mark.parametrize("causal", [True, False])
def test_extend_generic_head_dim(D, causal):
    prefix, ext = [0, 130, 33, 500], [70, 129, 1, 300]
    Hq, Hkv = 8, 2
    sc = make_scenario(prefix, ext, Hq, Hkv, D, 16, seed=D)
    backend, kv_pool = _backend(sc)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0,
                           attn_type=AttentionType.DECODER if causal else AttentionType.ENCODER_ONLY)
    fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=4, req_pool_indices=sc.req_pool_indices.to(DEV),
                      seq_lens=torch.tensor(sc.seq_lens, device=DEV), out_cache_loc=sc.out_cache_loc.to(DEV),
                      seq_lens_sum=sum(sc.seq_lens),
                      extend_seq_lens=torch.tensor(ext, dtype=torch.int32, device=DEV),
                      extend_prefix_lens=torch.tensor(prefix, dtype=torch.int32, device=DEV),
                      extend_seq_lens_cpu=list(ext), extend_prefix_lens_cpu=list(prefix), attn_backend=backend)
    backend.init_forward_metadata(fb)
    o = layer(sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV), fb)
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    oracle.store_kv(kc, vc, sc.out_cache_loc, sc.k, sc.v)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                                torch.tensor(ext), Hq, Hkv, D, sc.scaling, causal=causal)
    assert_close_attn(o, ref, f"extend D={D} causal={causal}")
