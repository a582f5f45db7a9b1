"""(a4) decode through the warp-specialised persistent TMA kernel (b200attn_decode_ws): parity against the oracle
over page sizes 1 / 16 / 32 (gather4 and 16-row box paths), ragged lengths, features, row maps, and the fused
new-token store (bit-exact against the oracle store)."""
import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import ops
from tests.util import assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _meta(sc, lens):
    d = sc.to(DEV)
    B = len(lens)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(max(sum(lens), 1), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, torch.tensor(lens, device=DEV), kv_indptr, kv_indices)
    return d, kv_indptr, kv_indices


def _run(sc, lens, q, min_chunk=64, lse=False, dtype=torch.bfloat16, d=None, **kw):
    if d is None:
        d, kv_indptr, kv_indices = _meta(sc, lens)
    else:
        d, kv_indptr, kv_indices = d
    B = len(lens)
    o = torch.full((B, sc.Hq, sc.D), float("nan"), dtype=dtype, device=DEV)
    ws = torch.zeros(ops.decode_ws_workspace_bytes(sc.Hq, sc.Hkv, sc.D), dtype=torch.uint8, device=DEV)
    lse_t = torch.empty((B, sc.Hq), dtype=torch.float32, device=DEV) if lse else None
    for _ in range(2):  # the second launch must find the arrival counters at zero again
        o.fill_(float("nan"))
        ops.decode_attention(q.to(DEV).view(B, sc.Hq, sc.D), d.k_cache, d.v_cache, o, kv_indptr, kv_indices,
                             sc.scaling, min_chunk=min_chunk, workspace=ws, lse=lse_t, ws_kernel=True, **kw)
    torch.cuda.synchronize()
    assert int(ws[:4 * 65536].view(torch.int32).abs().sum()) == 0
    return o.view(B, -1), lse_t


@pytest.mark.parametrize("Hq,Hkv,D,page", [(32, 8, 128, 16), (14, 2, 64, 1), (28, 4, 128, 16), (8, 8, 64, 32),
                                           (16, 1, 128, 1), (64, 2, 128, 16)])
@pytest.mark.parametrize("lens", [[1, 2, 15, 16, 17, 63, 64, 65, 200, 777], [5000], [1], [300] * 37,
                                  [4096, 1, 4096, 7, 2500]])
def test_decode_ws_matches_oracle(Hq, Hkv, D, page, lens):
    B = len(lens)
    sc = make_scenario(lens, [0] * B, Hq, Hkv, D, page, seed=Hq + page + B)
    q = torch.randn((B, Hq * D), generator=torch.Generator().manual_seed(1)).to(torch.bfloat16)
    ref, lse_ref = oracle.attention_fp32(q, sc.k_cache, sc.v_cache, sc.req_to_token, sc.req_pool_indices,
                                         torch.tensor(lens), torch.ones(B, dtype=torch.int64), Hq, Hkv, D,
                                         sc.scaling, causal=False, return_lse=True)
    for mc in (64, 256):
        o, lse = _run(sc, lens, q, min_chunk=mc, lse=True)
        assert_close_attn(o, ref, f"decode_ws Hq{Hq} D{D} page{page} B{B} min_chunk{mc}")
        torch.testing.assert_close(lse.cpu(), lse_ref, rtol=1e-3, atol=1e-3)


def test_decode_ws_large_batch_many_units_per_cta():
    g = torch.Generator().manual_seed(3)
    lens = torch.randint(1, 160, (700,), generator=g).tolist()
    sc = make_scenario(lens, [0] * 700, 32, 8, 128, 16, seed=5)
    q = torch.randn((700, 32 * 128), generator=g).to(torch.bfloat16)
    ref = oracle.attention_fp32(q, sc.k_cache, sc.v_cache, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens),
                                torch.ones(700, dtype=torch.int64), 32, 8, 128, sc.scaling, causal=False)
    o, _ = _run(sc, lens, q)
    assert_close_attn(o, ref, "decode_ws B=700")


def test_decode_ws_features_fp16_and_row_maps():
    lens = [5, 130, 1000, 2049]
    Hq, Hkv, D = 8, 2, 128
    sc = make_scenario(lens, [0] * 4, Hq, Hkv, D, 16, dtype=torch.float16, seed=3)
    q = torch.randn((4, Hq * D), generator=torch.Generator().manual_seed(2)).to(torch.float16)
    sinks = torch.randn(Hq, generator=torch.Generator().manual_seed(5))
    for kw in (dict(), dict(window_left=100), dict(logit_cap=30.0), dict(sinks=sinks)):
        ref = oracle.attention_fp32(q, sc.k_cache, sc.v_cache, sc.req_to_token, sc.req_pool_indices,
                                    torch.tensor(lens), torch.ones(4, dtype=torch.int64), Hq, Hkv, D, sc.scaling,
                                    causal=False, **kw)
        gkw = {k: (v.to(DEV) if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
        o, _ = _run(sc, lens, q, dtype=torch.float16, **gkw)
        assert_close_attn(o, ref, f"decode_ws fp16 {kw}")
    d, kv_indptr, kv_indices = _meta(sc, lens)
    o = torch.zeros((4, Hq, D), dtype=torch.float16, device=DEV)
    ws = torch.zeros(ops.decode_ws_workspace_bytes(Hq, Hkv, D), dtype=torch.uint8, device=DEV)
    qg = q.to(DEV).view(4, Hq, D)
    ops.decode_attention(qg, d.k_cache, d.v_cache, o, kv_indptr, kv_indices, sc.scaling, workspace=ws,
                         q_rows=torch.tensor([2, 0], dtype=torch.int32, device=DEV),
                         req_ids=torch.tensor([3, 1], dtype=torch.int32, device=DEV), num_items=2, ws_kernel=True)
    full = oracle.attention_fp32(q[[0, 0, 2, 2]], sc.k_cache, sc.v_cache, sc.req_to_token,
                                 sc.req_pool_indices[[1, 1, 3, 3]], torch.tensor([lens[1], lens[1], lens[3], lens[3]]),
                                 torch.ones(4, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
    assert_close_attn(o.view(4, -1)[0:1], full[0:1], "decode_ws row map 0")
    assert_close_attn(o.view(4, -1)[2:3], full[2:3], "decode_ws row map 2")
    assert float(o[1].abs().sum()) == 0 and float(o[3].abs().sum()) == 0


@pytest.mark.parametrize("Hq,Hkv,D,page", [(32, 8, 128, 16), (14, 2, 64, 1), (32, 2, 128, 16)])
def test_decode_ws_fused_store_is_bit_exact(Hq, Hkv, D, page):
    """new_k / new_v: the kernel stores row b into the slot of request b's last key and then attends over it."""
    lens = [1, 64, 65, 1000, 4096, 17, 129]
    B = len(lens)
    sc = make_scenario([l - 1 for l in lens], [1] * B, Hq, Hkv, D, page, seed=77 + page)
    d, kv_indptr, kv_indices = _meta(sc, lens)
    # fused-qkv layout: k / v are column views of one [B, (Hq + 2 Hkv) * D] activation (token stride > row)
    W = (Hq + 2 * Hkv) * D
    qkv = torch.randn((B, W), generator=torch.Generator().manual_seed(9)).to(torch.bfloat16).to(DEV)
    qkv[:, Hq * D:(Hq + Hkv) * D] = sc.k.view(B, -1).to(DEV)
    qkv[:, (Hq + Hkv) * D:] = sc.v.view(B, -1).to(DEV)
    q = qkv[:, :Hq * D]
    k_pool0 = d.k_cache.clone()
    o, _ = _run(sc, lens, q.contiguous().cpu(), d=(d, kv_indptr, kv_indices),
                new_k=qkv[:, Hq * D:(Hq + Hkv) * D], new_v=qkv[:, (Hq + Hkv) * D:], skip_slot=0)
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    oracle.store_kv(kc, vc, sc.out_cache_loc, sc.k, sc.v)
    assert not torch.equal(d.k_cache, k_pool0)
    assert torch.equal(d.k_cache.cpu(), kc) and torch.equal(d.v_cache.cpu(), vc), "fused store is not bit-exact"
    ref = oracle.attention_fp32(q.cpu(), kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens),
                                torch.ones(B, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
    assert_close_attn(o, ref, "decode_ws fused store")
