"""BASELINE configs[3] end to end WITH its M-RoPE positions (Qwen2.5-VL-7B LLM side, scaled to 4 + 4 requests):
token layout -> `mrope_rope_index` (get_rope_index) -> `mrope_positions_extend / _decode`
(forward_batch_info.py:1129-1254) -> fused M-RoPE + KV store (`MRotaryEmbedding`, mrope_section [16, 24, 24]) ->
MIXED attention through B200AttnBackend with save_kv_cache=False, against the oracle chain
rope_port.rope_store (fp32 rotation, bit-exact) + attention_fp32."""
import types

import pytest
import torch

from oracle import rope_port
from oracle import torch_native_port as oracle
from sglang_b200 import (B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention,
                         ReqToTokenPool, rotary)
from sglang_b200.rotary import FusedSetKVBufferArg, MRotaryEmbedding
from tests.util import assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"
IMG, VSTART = 151655, 151652


def test_cfg4_mixed_with_mrope_positions_end_to_end():
    Hq, Hkv, D, page = 28, 4, 128, 16
    n_ext, n_dec = 4, 4
    # extend requests: 3 text + vision_start + 32x32 image tokens + 252 text = 1280 tokens (cfg4's layout)
    ids = torch.tensor([100, 101, 102, VSTART] + [IMG] * 1024 + list(range(200, 452)), dtype=torch.int64)
    grid = torch.tensor([[1, 64, 64]])
    mm_pos, delta = rotary.mrope_rope_index(ids, grid, image_token_id=IMG, vision_start_token_id=VSTART)
    assert mm_pos.shape == (3, 1280) and int(mm_pos[1, 4:1028].max()) == 4 + 31
    prefix = [0] * n_ext + [1279] * n_dec
    ext = [1280] * n_ext + [1] * n_dec
    pos_ext = rotary.mrope_positions_extend(prefix[:n_ext], ext[:n_ext], [mm_pos] * n_ext, [delta] * n_ext)
    # running decodes of the same multimodal prompt: token 1280 (seq_len 1280): (delta - 1) + seq_len on all axes
    pos_dec = rotary.mrope_positions_decode(torch.tensor([1280] * n_dec), [delta] * n_dec)
    positions = torch.cat([pos_ext, pos_dec], dim=1)  # [3, T] in batch order (MIXED: prefill rows, then decode rows)
    T = sum(ext)
    assert positions.shape == (3, T)

    sc = make_scenario(prefix, ext, Hq, Hkv, D, page, seed=44)
    req_pool = ReqToTokenPool(sc.req_to_token.shape[0] - 1, sc.req_to_token.shape[1], DEV)
    req_pool.req_to_token.copy_(sc.req_to_token)
    kv_pool = MHATokenToKVPool(sc.k_cache.shape[0] - page, page, sc.k_cache.dtype, Hkv, D, 1, DEV)
    kv_pool.k_buffer[0].copy_(sc.k_cache)
    kv_pool.v_buffer[0].copy_(sc.v_cache)
    runner = types.SimpleNamespace(device=DEV, page_size=page, req_to_token_pool=req_pool, token_to_kv_pool=kv_pool,
                                   sliding_window_size=None)
    backend = B200AttnBackend(runner)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    rope = MRotaryEmbedding(D, D, 4096, 1e6, True, torch.bfloat16, mrope_section=[16, 24, 24])

    # the model's fused qkv activation: q / k / v are column views with the fused token stride
    qkv = torch.cat([sc.q, sc.k.view(T, -1), sc.v.view(T, -1)], dim=1).to(DEV)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    fb = ForwardBatch(forward_mode=ForwardMode.MIXED, batch_size=len(prefix), req_pool_indices=sc.req_pool_indices.to(DEV),
                      seq_lens=torch.tensor(sc.seq_lens, dtype=torch.int64, device=DEV),
                      out_cache_loc=sc.out_cache_loc.to(DEV), seq_lens_sum=sum(sc.seq_lens),
                      extend_seq_lens=torch.tensor(ext, dtype=torch.int32, device=DEV),
                      extend_prefix_lens=torch.tensor(prefix, dtype=torch.int32, device=DEV),
                      extend_seq_lens_cpu=list(ext), extend_prefix_lens_cpu=list(prefix),
                      mrope_positions=positions.to(DEV), attn_backend=backend)
    backend.init_forward_metadata(fb)
    kb, vb = kv_pool.get_kv_buffer(0)
    rope(fb.mrope_positions, q, k, fused_set_kv_buffer_arg=FusedSetKVBufferArg(v, kb, vb, fb.out_cache_loc))
    o = layer(q, k, v, fb, save_kv_cache=False)
    torch.cuda.synchronize()

    cache = rope_port.cos_sin_cache(D, 4096, 1e6)
    q_ref, k_ref, kc, vc = rope_port.rope_store(positions, sc.q, sc.k.view(T, -1), sc.v.view(T, -1), sc.k_cache,
                                                sc.v_cache, sc.out_cache_loc, cache, D, D, True,
                                                mrope_section=[16, 24, 24], native=False, skip_index=0)
    assert torch.equal(q.cpu(), q_ref) and torch.equal(k.cpu(), k_ref), "fused M-RoPE rotation is not bit-exact"
    assert torch.equal(kb.cpu(), kc) and torch.equal(vb.cpu(), vc), "fused store is not bit-exact"
    ref = oracle.attention_fp32(q_ref, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                                torch.tensor(ext), Hq, Hkv, D, sc.scaling, causal=True)
    assert_close_attn(o, ref, "cfg4 MIXED with M-RoPE positions")
