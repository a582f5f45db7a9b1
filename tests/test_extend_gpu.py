"""(a5) ragged extend / prefill parity on the tcgen05 kernel (through the C ABI)."""
import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import ops
from tests.util import assert_close_attn, golden_files, load_golden, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run_extend(sc, causal, lse=False, store=True):
    d = sc.to(DEV)
    B = len(sc.seq_lens)
    if store:
        ops.store_cache(d.k, d.v, d.k_cache, d.v_cache, d.out_cache_loc, reserved_skip_index=0)
    seq = torch.tensor(sc.seq_lens, dtype=torch.int64, device=DEV)
    ext = torch.tensor(sc.extend_lens, dtype=torch.int32, device=DEV)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    qo_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(max(sum(sc.seq_lens), 1), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, seq, kv_indptr, kv_indices, qo_lens=ext,
                         qo_indptr=qo_indptr)
    T = sum(sc.extend_lens)
    q = d.q.view(T, sc.Hq, sc.D)
    o = torch.full_like(q, float("nan"))
    lse_t = torch.empty((T, sc.Hq), dtype=torch.float32, device=DEV) if lse else None
    ops.extend_attention(q, d.k_cache, d.v_cache, o, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=causal,
                         page_size=sc.page_size, max_qo_len=max(sc.extend_lens), lse=lse_t)
    torch.cuda.synchronize()
    return o.view(T, -1), lse_t, d


def _fp32_ref(sc, d, causal, lse=False):
    return oracle.attention_fp32(d.q, d.k_cache, d.v_cache, d.req_to_token, d.req_pool_indices,
                                 torch.tensor(sc.seq_lens), torch.tensor(sc.extend_lens), sc.Hq, sc.Hkv, sc.D,
                                 sc.scaling, causal, return_lse=lse)


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("Hq,Hkv,D,page", [(8, 2, 128, 16), (4, 4, 128, 1), (14, 2, 64, 1), (28, 4, 128, 16),
                                           (4, 1, 64, 32)])
def test_extend_matches_oracle_small(causal, Hq, Hkv, D, page):
    prefix = [0, 16, 37, 0, 128, 300, 5]
    ext = [1, 7, 128, 129, 100, 260, 31]
    sc = make_scenario(prefix, ext, Hq, Hkv, D, page, seed=Hq * 7 + page)
    o, _, d = _run_extend(sc, causal)
    # the oracle proper (torch_native restatement, CPU SDPA) ...
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    oracle.store_kv(kc, vc, sc.out_cache_loc, sc.k, sc.v)
    ref = oracle.extend(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                        torch.tensor(prefix), torch.tensor(ext), Hq, Hkv, D, sc.scaling, causal=causal)
    assert_close_attn(o, ref, f"extend vs oracle causal={causal} Hq{Hq} D{D} page{page}")
    # ... and the independent fp32 reference
    assert_close_attn(o, _fp32_ref(sc, d, causal), "extend vs fp32")


def test_extend_lse_and_fp16():
    sc = make_scenario([0, 50], [200, 77], 8, 2, 128, 16, dtype=torch.float16, seed=4)
    o, lse, d = _run_extend(sc, True, lse=True)
    ref, lse_ref = _fp32_ref(sc, d, True, lse=True)
    assert_close_attn(o, ref, "extend fp16")
    torch.testing.assert_close(lse, lse_ref, rtol=2e-3, atol=2e-3)


@pytest.mark.parametrize("path", [p for p in golden_files() if "/decode_" not in p],
                         ids=lambda p: p.split("/")[-1][:-3])
def test_extend_matches_reference_golden(path):
    g = load_golden(path)
    Hq, Hkv, D = g["num_heads"], g["num_kv_heads"], g["head_dim"]
    B, T = len(g["seq_lens"]), g["q"].shape[0]
    kc, vc = g["k_pool_before"].to(DEV), g["v_pool_before"].to(DEV)
    ops.store_cache(g["k"].view(-1, Hkv, D).to(DEV), g["v"].view(-1, Hkv, D).to(DEV), kc, vc,
                    g["out_cache_loc"].to(DEV), reserved_skip_index=-1)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    qo_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(g["seq_lens"]), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(g["req_to_token"].to(DEV), g["req_pool_indices"].to(DEV),
                         torch.tensor(g["seq_lens"], device=DEV), kv_indptr, kv_indices,
                         qo_lens=torch.tensor(g["extend_lens"], dtype=torch.int32, device=DEV), qo_indptr=qo_indptr)
    q = g["q"].to(DEV).view(T, Hq, D)
    o = torch.empty_like(q)
    ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, g["scaling"], causal=g["causal"],
                         page_size=g["page_size"], max_qo_len=max(g["extend_lens"]))
    assert_close_attn(o.view(T, -1), g["o_ref"], g["name"])


def test_extend_full_size_cfg3_radix_hit():
    # Llama-3.1-8B extend: 32 requests sharing a 2048-token prefix (same 128 physical pages) + 512 new tokens
    B, P, E = 32, 2048, 512
    sc = make_scenario([P] * B, [E] * B, 32, 8, 128, 16, seed=20261107, shared_prefix_pages=P // 16)
    o, _, d = _run_extend(sc, True)
    ref = _fp32_ref(sc, d, True)
    assert_close_attn(o, ref, "cfg3 extend")
    # radix-hit property: identical prefixes + identical new tokens => identical outputs
    sc2 = make_scenario([P] * 2, [E] * 2, 32, 8, 128, 16, seed=5, shared_prefix_pages=P // 16)
    sc2.q[E:] = sc2.q[:E]
    sc2.k[E:] = sc2.k[:E]
    sc2.v[E:] = sc2.v[:E]
    o2, _, _ = _run_extend(sc2, True)
    assert torch.equal(o2[:E], o2[E:])


def test_extend_mixed_batch_cfg4_shape():
    # Qwen2.5-VL-7B LLM side (scaled): extend rows + decode rows (extend_len 1) in one batch, G = 7
    prefix = [0] * 4 + [1279] * 4
    ext = [1280] * 4 + [1] * 4
    sc = make_scenario(prefix, ext, 28, 4, 128, 16, seed=44)
    for causal in (True, False):
        o, _, d = _run_extend(sc, causal)
        assert_close_attn(o, _fp32_ref(sc, d, causal), f"cfg4 mixed causal={causal}")
