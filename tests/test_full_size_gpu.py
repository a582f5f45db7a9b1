"""Parity at the FULL sizes of BASELINE.json configs[3] and configs[4] (the bench legs that had no output check):

  cfg4  Qwen2.5-VL-7B LLM side: 16 x 1280 new tokens + 16 decode rows (prefix 1279) in one MIXED batch, Hq 28 / Hkv 4,
        causal and bidirectional (ENCODER_ONLY);
  cfg5  Llama-3.1-70B shape, KV-head shards Hkv 8 / 4 / 2 / 1 (Hq 64 / 32 / 16 / 8): decode batch 256 x 8192 ctx and
        the 8192-token chunked prefill over an 8192-token cached prefix.

Through B200AttnBackend (store + attention) against the blocked fp32-softmax reference on the GPU, tolerance
1e-2 relative / 2e-2 absolute (BASELINE north_star); the store is checked bit-exact."""
import types

import pytest
import torch

from sglang_b200 import (AttentionType, B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention,
                         ReqToTokenPool)
from tests.util import assert_close_attn, attention_fp32_blocked, make_scenario_gpu

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _backend(sc):
    req_pool = ReqToTokenPool(sc.req_to_token.shape[0] - 1, sc.req_to_token.shape[1], DEV)
    req_pool.req_to_token = sc.req_to_token
    kv_pool = MHATokenToKVPool.__new__(MHATokenToKVPool)  # adopt the scenario's buffers (no second multi-GB copy)
    kv_pool.size, kv_pool.page_size = sc.k_cache.shape[0] - sc.page_size, sc.page_size
    kv_pool.dtype = kv_pool.store_dtype = sc.k_cache.dtype
    kv_pool.head_num, kv_pool.head_dim, kv_pool.v_head_dim = sc.Hkv, sc.D, sc.D
    kv_pool.layer_num, kv_pool.device, kv_pool.start_layer = 1, DEV, 0
    kv_pool.skip_padding_slot, kv_pool.layer_transfer_counter = True, None
    kv_pool.k_buffer, kv_pool.v_buffer = [sc.k_cache], [sc.v_cache]
    runner = types.SimpleNamespace(device=DEV, page_size=sc.page_size, req_to_token_pool=req_pool,
                                   token_to_kv_pool=kv_pool, sliding_window_size=None)
    return B200AttnBackend(runner), kv_pool


def _run(sc, mode, causal=True):
    backend, kv_pool = _backend(sc)
    B = len(sc.seq_lens)
    layer = RadixAttention(sc.Hq, sc.D, sc.scaling, sc.Hkv, 0,
                           attn_type=AttentionType.DECODER if causal else AttentionType.ENCODER_ONLY)
    fb = ForwardBatch(forward_mode=mode, batch_size=B, req_pool_indices=sc.req_pool_indices,
                      seq_lens=torch.tensor(sc.seq_lens, dtype=torch.int64, device=DEV), out_cache_loc=sc.out_cache_loc,
                      seq_lens_sum=sum(sc.seq_lens), seq_lens_cpu=torch.tensor(sc.seq_lens), attn_backend=backend)
    if not mode.is_decode():
        fb.extend_seq_lens = torch.tensor(sc.extend_lens, dtype=torch.int32, device=DEV)
        fb.extend_prefix_lens = torch.tensor(sc.prefix_lens, dtype=torch.int32, device=DEV)
        fb.extend_seq_lens_cpu, fb.extend_prefix_lens_cpu = list(sc.extend_lens), list(sc.prefix_lens)
    backend.init_forward_metadata(fb)
    o = layer(sc.q, sc.k, sc.v, fb)
    torch.cuda.synchronize()
    assert torch.equal(sc.k_cache[sc.out_cache_loc], sc.k) and torch.equal(sc.v_cache[sc.out_cache_loc], sc.v)
    single_sees_all = [1 if e == 1 else 0 for e in sc.extend_lens]  # a single row sees every key either way
    ref = attention_fp32_blocked(sc.q, sc.k_cache, sc.v_cache, sc.req_to_token, sc.req_pool_indices.tolist(),
                                 sc.seq_lens, sc.extend_lens, sc.Hq, sc.Hkv, sc.D, sc.scaling, causal=causal)
    del single_sees_all
    return o, ref


@pytest.mark.parametrize("causal", [True, False], ids=["causal", "bidirectional"])
def test_cfg4_full_mixed_batch(causal):
    prefix = [0] * 16 + [1279] * 16
    ext = [1280] * 16 + [1] * 16
    sc = make_scenario_gpu(prefix, ext, 28, 4, 128, 16, seed=404)
    o, ref = _run(sc, ForwardMode.MIXED, causal)
    assert_close_attn(o, ref, f"cfg4 full MIXED causal={causal}")


@pytest.mark.parametrize("hq,hkv", [(64, 8), (32, 4), (16, 2), (8, 1)])
def test_cfg5_full_decode_shards(hq, hkv):
    B, ctx = 256, 8192
    g = torch.Generator().manual_seed(505 + hkv)
    # the full-width case runs ragged lengths (last pages partial), the shards the exact BASELINE shape
    lens = torch.randint(4096, ctx + 1, (B,), generator=g).tolist() if hkv == 8 else [ctx] * B
    sc = make_scenario_gpu([l - 1 for l in lens], [1] * B, hq, hkv, 128, 16, seed=505 + hkv, pool_slack=1.02)
    o, ref = _run(sc, ForwardMode.DECODE)
    assert_close_attn(o, ref, f"cfg5 decode Hq {hq} / Hkv {hkv}")


@pytest.mark.parametrize("hq,hkv", [(64, 8), (8, 1)])
def test_cfg5_full_chunked_prefill(hq, hkv):
    sc = make_scenario_gpu([8192], [8192], hq, hkv, 128, 16, seed=515 + hkv)
    o, ref = _run(sc, ForwardMode.EXTEND)
    assert_close_attn(o, ref, f"cfg5 chunked prefill Hq {hq} / Hkv {hkv}")
