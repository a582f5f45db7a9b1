"""(a4) paged GQA split-KV decode parity (through the C ABI)."""
import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import ops
from tests.util import assert_close_attn, golden_files, load_golden, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run_decode(sc, max_splits, min_chunk=64, lse=False):
    d = sc.to(DEV)
    B = len(sc.seq_lens)
    seq = torch.tensor(sc.seq_lens, dtype=torch.int64, device=DEV)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(max(sum(sc.seq_lens), 1), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, seq, kv_indptr, kv_indices)
    q = d.q.view(B, sc.Hq, sc.D)
    o = torch.full_like(q, float("nan"))
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, sc.Hq, sc.D, max_splits), 16), dtype=torch.uint8, device=DEV)
    lse_t = torch.empty((B, sc.Hq), dtype=torch.float32, device=DEV) if lse else None
    ops.decode_attention(q, d.k_cache, d.v_cache, o, kv_indptr, kv_indices, sc.scaling, max_splits=max_splits,
                         min_chunk=min_chunk, workspace=ws, lse=lse_t)
    torch.cuda.synchronize()
    return o.view(B, -1), lse_t


def _oracle_decode(sc):
    return oracle.decode(sc.q, sc.k_cache, sc.v_cache, sc.req_to_token, sc.req_pool_indices,
                         torch.tensor(sc.seq_lens), sc.Hq, sc.Hkv, sc.D, sc.scaling)


@pytest.mark.parametrize("Hq,Hkv,D,page", [(32, 8, 128, 16), (14, 2, 64, 1), (28, 4, 128, 16), (8, 8, 64, 32),
                                           (16, 1, 128, 1), (64, 2, 128, 16)])
@pytest.mark.parametrize("max_splits", [1, 4, 16])
def test_decode_matches_oracle(Hq, Hkv, D, page, max_splits):
    lens = [1, 2, 15, 16, 17, 63, 64, 65, 200, 777]
    sc = make_scenario(lens, [0] * len(lens), Hq, Hkv, D, page, seed=Hq + page)
    sc.q = torch.randn((len(lens), Hq * D), generator=torch.Generator().manual_seed(1)).to(torch.bfloat16)
    o, _ = _run_decode(sc, max_splits)
    assert_close_attn(o, _oracle_decode(sc), f"decode Hq{Hq} Hkv{Hkv} D{D} page{page} splits{max_splits}")


def test_decode_fp16_and_lse():
    lens = [5, 130, 1000]
    sc = make_scenario(lens, [0] * 3, 8, 2, 128, 16, dtype=torch.float16, seed=3)
    sc.q = torch.randn((3, 8 * 128), generator=torch.Generator().manual_seed(2)).to(torch.float16)
    for ms in (1, 8):
        o, lse = _run_decode(sc, ms, lse=True)
        ref, lse_ref = oracle.attention_fp32(sc.q, sc.k_cache, sc.v_cache, sc.req_to_token, sc.req_pool_indices,
                                             torch.tensor(lens), torch.ones(3, dtype=torch.int64), 8, 2, 128,
                                             sc.scaling, causal=False, return_lse=True)
        assert_close_attn(o, ref, f"decode fp16 splits{ms}")
        torch.testing.assert_close(lse.cpu(), lse_ref, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("path", [p for p in golden_files() if "/decode_" in p], ids=lambda p: p.split("/")[-1][:-3])
def test_decode_matches_reference_golden(path):
    g = load_golden(path)
    Hq, Hkv, D = g["num_heads"], g["num_kv_heads"], g["head_dim"]
    B = len(g["seq_lens"])
    kc, vc = g["k_pool_before"].to(DEV), g["v_pool_before"].to(DEV)
    ops.store_cache(g["k"].view(-1, Hkv, D).to(DEV), g["v"].view(-1, Hkv, D).to(DEV), kc, vc,
                    g["out_cache_loc"].to(DEV), reserved_skip_index=-1)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(g["seq_lens"]), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(g["req_to_token"].to(DEV), g["req_pool_indices"].to(DEV),
                         torch.tensor(g["seq_lens"], device=DEV), kv_indptr, kv_indices)
    q = g["q"].to(DEV).view(B, Hq, D)
    for ms in (1, 3):
        o = torch.empty_like(q)
        ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
        ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, g["scaling"], max_splits=ms, min_chunk=16,
                             workspace=ws)
        assert_close_attn(o.view(B, -1), g["o_ref"], f"{g['name']} splits{ms}")


def test_decode_full_size_cfg2_vs_fp32_reference():
    # Llama-3.1-8B decode: B=64 x 4096 ctx (ragged variant), Hq 32 / Hkv 8 / D 128, page 16
    g = torch.Generator().manual_seed(2)
    lens = torch.randint(2048, 4097, (64,), generator=g).tolist()
    lens[0] = 4096
    sc = make_scenario(lens, [0] * 64, 32, 8, 128, 16, seed=20261106)
    sc.q = torch.randn((64, 32 * 128), generator=g).to(torch.bfloat16)
    o, _ = _run_decode(sc, 4, min_chunk=256)
    d = sc.to(DEV)
    ref = oracle.attention_fp32(d.q, d.k_cache, d.v_cache, d.req_to_token, d.req_pool_indices,
                                torch.tensor(lens), torch.ones(64, dtype=torch.int64), 32, 8, 128, sc.scaling,
                                causal=False)
    assert_close_attn(o, ref, "cfg2 decode")
    # size-independent property: split count must not change the result beyond rounding
    o1, _ = _run_decode(sc, 1)
    assert_close_attn(o, o1, "cfg2 decode split-invariance", rtol=1e-2, atol=1e-2)


def test_decode_q_rows_and_req_ids_routing():
    lens = [40, 300, 9]
    sc = make_scenario(lens, [0] * 3, 8, 2, 128, 16, seed=9)
    d = sc.to(DEV)
    seq = torch.tensor(lens, device=DEV)
    kv_indptr = torch.zeros(4, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, seq, kv_indptr, kv_indices)
    T = 7
    q = torch.randn((T, 8, 128), device=DEV).to(torch.bfloat16)
    o = torch.zeros_like(q)
    q_rows = torch.tensor([5, 0], dtype=torch.int32, device=DEV)   # work item 0 -> row 5, item 1 -> row 0
    req_ids = torch.tensor([2, 1], dtype=torch.int32, device=DEV)  # ... using request 2 / request 1
    ops.decode_attention(q, d.k_cache, d.v_cache, o, kv_indptr, kv_indices, sc.scaling, q_rows=q_rows,
                         req_ids=req_ids)
    for row, req in ((5, 2), (0, 1)):
        ref = oracle.attention_fp32(q[row].reshape(1, -1).cpu(), sc.k_cache, sc.v_cache, sc.req_to_token,
                                    sc.req_pool_indices[req:req + 1], torch.tensor([lens[req]]),
                                    torch.ones(1, dtype=torch.int64), 8, 2, 128, sc.scaling, causal=False)
        assert_close_attn(o[row].reshape(1, -1), ref, f"routed row {row}")
    assert float(o[[1, 2, 3, 4, 6]].abs().sum()) == 0


@pytest.mark.parametrize("tail_begin,k", [(0, 4), (5, 3), (16, 8), (1000, 4)])
def test_decode_tail_split_matches_oracle(tail_begin, k):
    """Only the units of rank >= tail_begin are split (ops.pick_tail_split); results must not depend on it."""
    from sglang_b200 import ops
    Hq, Hkv, D = 8, 2, 128
    lens = [700, 64, 1, 2049, 513, 300, 1500, 128, 65, 900]
    sc = make_scenario([l - 1 for l in lens], [1] * len(lens), Hq, Hkv, D, 16, seed=5)
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    kc[sc.out_cache_loc] = sc.k
    vc[sc.out_cache_loc] = sc.v
    B = len(lens)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(sc.req_to_token.to(DEV), sc.req_pool_indices.to(DEV), torch.tensor(lens, device=DEV), kv_indptr,
                         kv_indices)
    q = sc.q.to(DEV).view(B, Hq, D)
    o = torch.empty_like(q)
    ws = torch.zeros(ops.decode_workspace_bytes(B, Hq, D, k), dtype=torch.uint8, device=DEV)
    ops.decode_attention(q, kc.to(DEV), vc.to(DEV), o, kv_indptr, kv_indices, sc.scaling, max_splits=k, min_chunk=64,
                         workspace=ws, tail_begin=tail_begin)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens),
                                torch.ones(B, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
    assert_close_attn(o.view(B, -1), ref, f"tail split {tail_begin}/{k}")


def test_pick_tail_split_policy():
    from sglang_b200 import ops
    assert ops.pick_tail_split(64, 8, 4, num_sms=148) == (-1, 1)      # 512 units: tail 216/296 too full to split
    assert ops.pick_tail_split(74, 8, 4, num_sms=148) == (-1, 1)      # exactly two waves
    assert ops.pick_tail_split(40, 8, 4, num_sms=148) == (296, 8)     # 24-unit tail
    assert ops.pick_tail_split(8, 8, 4, num_sms=148) == (-1, 1)       # below one wave: length-based policy
