"""Logit cap, attention sinks, the extend sliding window and the e4m3 KV READ path against outputs of the REFERENCE's
production Triton kernels, generated on a B200 by oracle/gen_golden_triton.py (`decode_attention_fwd`,
decode_attention.py:1163-1239; `extend_attention_fwd_unified`, extend_attention.py:1271-1410).  torch_native cannot
express these features, so this is what pins them.  Inputs are regenerated from the recorded seeds.

CPU: the oracle restatement (`oracle.attention_fp32`) agrees with the reference kernels.
GPU: the sm_100a kernels agree with them (BASELINE tolerance 1e-2 rel / 2e-2 abs).

e4m3 pools: the reference kernels cast the probabilities to the VALUE dtype before the second contraction
(`p.to(v.dtype)`, decode_attention.py:742, extend_attention.py:1248) and, in extend, the query to the KEY dtype
(`q.to(k.dtype)`, extend_attention.py:523/1174), i.e. P and Q themselves are rounded to e4m3 (3 mantissa bits).
  * decode and the backend's "dequant" extend path keep Q and P in f16 / bf16: they differ from the reference by the
    reference's own rounding - compared at 1e-1 abs (measured max 7.8e-2) and ALSO against the exact fp32 math over
    the dequantised pool at the normal tolerance (the sm_100a kernel is the more accurate of the two);
  * the default "in_kernel" extend path (b200attn_extend_fp8kv) computes in the reference's arithmetic: measured
    max 3.1e-2 / mean 1.4e-3 from the reference kernel's output (the 16-bit path: 7.8e-2 / 3.4e-3), asserted at
    5e-2 abs; `oracle.fp8_port.extend_attention_e4m3_arith` restates that arithmetic on the CPU and is pinned here
    against the same recorded output (measured max 2.0e-2, asserted 3e-2)."""

FP8_ATOL = 1e-1
import os
import types

import pytest
import torch

from oracle import torch_native_port as oracle
from tests.util import GOLDEN_DIR, assert_close_attn, load_golden, make_scenario

PATH = os.path.join(GOLDEN_DIR, "triton_cases.pt")
CASES = load_golden(PATH)["cases"] if os.path.exists(PATH) else []
IDS = [f"{c['kind']}_{c['name']}" for c in CASES]


def _inputs(c):
    Hq, Hkv, D, page = c["Hq"], c["Hkv"], c["D"], c["page_size"]
    if c["kind"] == "decode":
        lens = c["lens"]
        sc = make_scenario(lens, [0] * len(lens), Hq, Hkv, D, page, seed=c["scenario_seed"])
        q = torch.randn((len(lens), Hq, D), generator=torch.Generator().manual_seed(2)).to(torch.bfloat16).view(len(lens), -1)
        kc, vc = sc.k_cache, sc.v_cache
        seq, ext = lens, [1] * len(lens)
    else:
        sc = make_scenario(c["prefix_lens"], c["extend_lens"], Hq, Hkv, D, page, seed=c["scenario_seed"])
        q = sc.q
        kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
        kc[sc.out_cache_loc] = sc.k
        vc[sc.out_cache_loc] = sc.v
        seq, ext = sc.seq_lens, c["extend_lens"]
    k8 = v8 = None
    if c["k_scale"] != 1.0:
        k8 = (kc.float() / c["k_scale"]).to(torch.float8_e4m3fn)
        v8 = (vc.float() / c["v_scale"]).to(torch.float8_e4m3fn)
        kc, vc = k8.float() * c["k_scale"], v8.float() * c["v_scale"]  # what the read path must see
    return sc, q, kc, vc, k8, v8, seq, ext


def test_triton_fixture_present():
    assert len(CASES) == 9, "generate on the GPU box: python oracle/gen_golden_triton.py gpurun_out/golden"


@pytest.mark.parametrize("c", CASES, ids=IDS)
def test_oracle_matches_reference_triton_kernels(c):
    sc, q, kc, vc, k8, v8, seq, ext = _inputs(c)
    o = oracle.attention_fp32(q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(seq), torch.tensor(ext),
                              c["Hq"], c["Hkv"], c["D"], c["scaling"], causal=True,
                              window_left=c.get("window", -1), logit_cap=c["logit_cap"], sinks=c["sinks"])
    tol = dict(rtol=1e-2, atol=FP8_ATOL) if c["k_scale"] != 1.0 else {}
    assert_close_attn(o.view(c["o_ref"].shape), c["o_ref"], f"oracle vs triton {c['kind']} {c['name']}", **tol)
    if c["kind"] == "extend" and c["k_scale"] != 1.0:  # the e4m3-arithmetic restatement: 3x closer than exact math
        from oracle import fp8_port

        em = fp8_port.extend_attention_e4m3_arith(q, k8.view(torch.uint8), v8.view(torch.uint8), sc.req_to_token,
                                                  sc.req_pool_indices, seq, ext, c["Hq"], c["Hkv"], c["D"],
                                                  c["scaling"], c["k_scale"], c["v_scale"])
        ref = c["o_ref"].float().view(em.shape)
        assert_close_attn(em, ref, "e4m3-arithmetic oracle vs triton extend_fp8", rtol=1e-2, atol=3e-2)
        assert float((em - ref).abs().mean()) < 0.5 * float((o.view(em.shape).float() - ref).abs().mean())
        # the feature-carrying restatement (attention_fp32(e4m3_arith=True)) is the same arithmetic
        em2 = oracle.attention_fp32(q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(seq),
                                    torch.tensor(ext), c["Hq"], c["Hkv"], c["D"], c["scaling"], causal=True,
                                    e4m3_arith=True)
        torch.testing.assert_close(em2, em, atol=1e-5, rtol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("fp8_extend", ["in_kernel", "dequant"])
@pytest.mark.parametrize("c", CASES, ids=IDS)
def test_kernels_match_reference_triton_kernels(c, fp8_extend):
    if fp8_extend == "dequant" and not (c["kind"] == "extend" and c["k_scale"] != 1.0):
        pytest.skip("the extend path switch only matters for the e4m3 extend case")
    from sglang_b200 import (B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention,
                             ReqToTokenPool)

    dev = "cuda"
    sc, q, kc, vc, k8, v8, seq, ext = _inputs(c)
    Hq, Hkv, D, page = c["Hq"], c["Hkv"], c["D"], c["page_size"]
    fp8 = k8 is not None
    pool_src_k, pool_src_v = (k8.view(torch.uint8), v8.view(torch.uint8)) if fp8 else (kc, vc)
    req_pool = ReqToTokenPool(sc.req_to_token.shape[0] - 1, sc.req_to_token.shape[1], dev)
    req_pool.req_to_token.copy_(sc.req_to_token)
    kv_pool = MHATokenToKVPool(pool_src_k.shape[0] - page, page, torch.float8_e4m3fn if fp8 else torch.bfloat16, Hkv, D,
                               1, dev)
    kv_pool.k_buffer[0].copy_(pool_src_k)
    kv_pool.v_buffer[0].copy_(pool_src_v)
    W = c.get("window", -1)
    runner = types.SimpleNamespace(device=dev, page_size=page, req_to_token_pool=req_pool, token_to_kv_pool=kv_pool,
                                   sliding_window_size=W if W > 0 else None)
    backend = B200AttnBackend(runner, fp8_extend=fp8_extend)
    layer = RadixAttention(Hq, D, c["scaling"], Hkv, 0, logit_cap=c["logit_cap"], sliding_window_size=W)
    if fp8:
        layer.k_scale, layer.v_scale = c["k_scale"], c["v_scale"]
        layer.k_scale_float, layer.v_scale_float = c["k_scale"], c["v_scale"]
    B = len(seq)
    decode = c["kind"] == "decode"
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE if decode else ForwardMode.EXTEND, batch_size=B,
                      req_pool_indices=sc.req_pool_indices.to(dev), seq_lens=torch.tensor(seq, device=dev),
                      out_cache_loc=torch.zeros(sum(ext), dtype=torch.int64, device=dev), seq_lens_sum=sum(seq),
                      seq_lens_cpu=torch.tensor(seq), attn_backend=backend)
    if not decode:
        fb.extend_seq_lens = torch.tensor(ext, dtype=torch.int32, device=dev)
        fb.extend_prefix_lens = torch.tensor(c["prefix_lens"], dtype=torch.int32, device=dev)
        fb.extend_seq_lens_cpu, fb.extend_prefix_lens_cpu = list(ext), list(c["prefix_lens"])
    backend.init_forward_metadata(fb)
    kw = {} if c["sinks"] is None else dict(sinks=c["sinks"].to(dev))
    # the pool already holds every key (the reference kernels were given the same pool): nothing to store
    o = layer(q.to(dev), None, None, fb, save_kv_cache=False, **kw)
    ref_arith = fp8 and not decode and fp8_extend == "in_kernel"  # computed in the reference's e4m3 arithmetic
    tol = dict(rtol=1e-2, atol=5e-2 if ref_arith else FP8_ATOL) if fp8 else {}
    assert_close_attn(o.view(c["o_ref"].shape), c["o_ref"], f"b200 vs triton {c['kind']} {c['name']}", **tol)
    if fp8 and not ref_arith:  # and at the normal tolerance against the exact math over the dequantised pool
        exact = oracle.attention_fp32(q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(seq),
                                      torch.tensor(ext), Hq, Hkv, D, c["scaling"], causal=True)
        assert_close_attn(o.view(exact.shape), exact, f"b200 vs exact fp32 over the e4m3 pool, {c['kind']}")
