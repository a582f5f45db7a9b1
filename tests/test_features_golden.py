"""Sliding window and TARGET_VERIFY custom masks against fixtures produced by the REFERENCE kit's own fp32
references (oracle/gen_golden_features.py: dense_attention.py:802-855 window rule :831-840, :858-905 custom mask).

CPU: the oracle restatement (`oracle.attention_fp32` window_left / custom_masks) reproduces them.
GPU: B200AttnBackend (store + decode / extend kernels with window_left / custom_mask) reproduces them.
"""
import types

import pytest
import torch

from oracle import torch_native_port as oracle
from tests.util import assert_close_attn, feature_golden_files, feature_pool, load_golden

FILES = feature_golden_files()
IDS = [p.split("/")[-1][:-3] for p in FILES]


def _mask_blocks(g):
    """Per-request [E, L] bool blocks out of the flat backend mask (triton_backend.py:911-917 layout)."""
    if g["custom_mask"] is None:
        return None
    out, off = [], 0
    for L, E in zip(g["seq_lens"], g["extend_lens"]):
        out.append(g["custom_mask"][off:off + E * L].view(E, L))
        off += E * L
    assert off == g["custom_mask"].numel()
    return out


def test_feature_fixtures_present():
    assert len(FILES) >= 9, "run oracle/gen_golden_features.py"


@pytest.mark.parametrize("path", FILES, ids=IDS)
def test_oracle_matches_reference_kit_features(path):
    g = load_golden(path)
    kc, vc = feature_pool(g)
    Hkv, D = g["num_kv_heads"], g["head_dim"]
    oracle.store_kv(kc, vc, g["out_cache_loc"].long(), g["k"].view(-1, Hkv, D), g["v"].view(-1, Hkv, D))
    W = g["window"]
    o = oracle.attention_fp32(g["q"], kc, vc, g["req_to_token"], g["req_pool_indices"].long(),
                              torch.tensor(g["seq_lens"]), torch.tensor(g["extend_lens"]), g["num_heads"],
                              Hkv, D, g["scaling"], causal=g["custom_mask"] is None,
                              window_left=-1 if W is None else W, custom_masks=_mask_blocks(g),
                              skip_prefix_custom_mask=True)
    # both are fp32-softmax references; the kit rounds its output to bf16
    torch.testing.assert_close(o.to(torch.bfloat16).float(), g["o_ref"].float(), rtol=1.6e-2, atol=2e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=IDS)
def test_backend_matches_reference_kit_features(path):
    from sglang_b200 import (B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention,
                             ReqToTokenPool)

    dev = "cuda"
    g = load_golden(path)
    Hq, Hkv, D, page = g["num_heads"], g["num_kv_heads"], g["head_dim"], g["page_size"]
    W = g["window"]
    table = g["req_to_token"]
    kc, vc = feature_pool(g)
    req_pool = ReqToTokenPool(table.shape[0] - 1, table.shape[1], dev)
    req_pool.req_to_token.copy_(table)
    kv_pool = MHATokenToKVPool(kc.shape[0] - page, page, kc.dtype, Hkv, D, 1, dev)
    kv_pool.k_buffer[0].copy_(kc)
    kv_pool.v_buffer[0].copy_(vc)
    runner = types.SimpleNamespace(device=dev, page_size=page, req_to_token_pool=req_pool, token_to_kv_pool=kv_pool,
                                   sliding_window_size=W)
    backend = B200AttnBackend(runner)
    layer = RadixAttention(Hq, D, g["scaling"], Hkv, 0, sliding_window_size=-1 if W is None else W)
    mode = ForwardMode[g["mode"]]
    B = len(g["seq_lens"])
    verify = mode.is_target_verify()
    # TARGET_VERIFY: seq_lens counts the cached tokens only (speculative_target_verify_runner.py:
    # _prepare_target_verify_batch); the draft tokens arrive as this forward's q/k/v
    lens = g["prefix_lens"] if verify else g["seq_lens"]
    fb = ForwardBatch(forward_mode=mode, batch_size=B, req_pool_indices=g["req_pool_indices"].to(dev),
                      seq_lens=torch.tensor(lens, dtype=torch.int32, device=dev),
                      out_cache_loc=g["out_cache_loc"].to(dev), seq_lens_sum=sum(lens),
                      seq_lens_cpu=torch.tensor(lens, dtype=torch.int32), positions=g["positions"].to(dev),
                      attn_backend=backend)
    if mode.is_extend():
        fb.extend_prefix_lens = torch.tensor(g["prefix_lens"], dtype=torch.int32, device=dev)
        fb.extend_seq_lens = torch.tensor(g["extend_lens"], dtype=torch.int32, device=dev)
        fb.extend_prefix_lens_cpu, fb.extend_seq_lens_cpu = g["prefix_lens"], g["extend_lens"]
        fb.extend_num_tokens = sum(g["extend_lens"])
    if verify:
        fb.spec_info = types.SimpleNamespace(draft_token_num=g["extend_lens"][0], custom_mask=g["custom_mask"].to(dev))
    backend.init_forward_metadata(fb)
    o = layer(g["q"].to(dev), g["k"].to(dev), g["v"].to(dev), fb)
    assert o.shape == g["o_ref"].shape and o.dtype == g["o_ref"].dtype
    assert_close_attn(o, g["o_ref"], g["name"])
