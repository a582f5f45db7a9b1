"""Feature flags on the in-kernel e4m3 extend path (b200attn_extend_fp8kv): sliding window, logit cap, sinks and
TARGET_VERIFY tree masks share the softmax code of the 16-bit kernel, but the operands are e4m3 - checked against the
reference Triton kernel's arithmetic restated by oracle.attention_fp32(e4m3_arith=True) (pinned through
fp8_port.extend_attention_e4m3_arith, tests/test_triton_golden.py) with the exact math beside it."""
import pytest
import torch

from oracle import fp8_port
from oracle import torch_native_port as oracle
from sglang_b200 import ops
from tests.test_features_gpu import _filled, _meta
from tests.util import assert_close_e4m3_arith, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"
KS, VS = 0.6, 0.9

FEATS = [dict(window_left=3), dict(window_left=64), dict(window_left=200), dict(logit_cap=30.0), dict(sinks=True),
         dict(sinks=True, window_left=127, logit_cap=50.0)]


def _pools(sc):
    kc, vc = _filled(sc)
    k8, v8 = fp8_port.quantize(kc * 2, KS), fp8_port.quantize(vc * 2, VS)
    return k8, v8, fp8_port.dequantize(k8, KS), fp8_port.dequantize(v8, VS)


@pytest.mark.parametrize("feat", FEATS, ids=lambda f: "-".join(f"{k}{v}" for k, v in f.items()))
@pytest.mark.parametrize("Hq,Hkv,D,causal", [(8, 2, 128, True), (14, 2, 64, True), (8, 2, 128, False)])
def test_extend_fp8_features(feat, Hq, Hkv, D, causal):
    pre, ext = [0, 100, 700, 5, 1000], [300, 1, 260, 129, 64]
    sc = make_scenario(pre, ext, Hq, Hkv, D, 16, seed=13)
    k8, v8, kf, vf = _pools(sc)
    kv_indptr, kv_indices, qo_indptr = _meta(sc, ops)
    f = dict(feat)
    sinks = torch.linspace(-2.0, 3.0, Hq) if f.pop("sinks", False) else None
    q = sc.q.to(DEV).view(-1, Hq, D)
    o = torch.zeros_like(q)
    ops.extend_attention(q, k8.to(DEV), v8.to(DEV), o, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=causal,
                         page_size=16, max_qo_len=max(ext), sinks=None if sinks is None else sinks.to(DEV),
                         k_descale=KS, v_descale=VS, **f)
    common = (sc.q, kf, vf, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens), torch.tensor(ext), Hq, Hkv,
              D, sc.scaling)
    exact = oracle.attention_fp32(*common, causal=causal, sinks=sinks, **f)
    emul = oracle.attention_fp32(*common, causal=causal, sinks=sinks, e4m3_arith=True, **f)
    assert_close_e4m3_arith(o.view(sum(ext), -1), emul, exact, f"e4m3 extend {feat} D={D} causal={causal}")


@pytest.mark.parametrize("skip_prefix", [True, False])
@pytest.mark.parametrize("Hq,Hkv,D", [(8, 2, 128), (14, 2, 64)])
def test_extend_fp8_custom_mask_target_verify(skip_prefix, Hq, Hkv, D):
    pre, ext = [37, 0, 300, 1000, 129], [8, 8, 8, 8, 8]
    sc = make_scenario(pre, ext, Hq, Hkv, D, 16, seed=23)
    k8, v8, kf, vf = _pools(sc)
    kv_indptr, kv_indices, qo_indptr = _meta(sc, ops)
    g = torch.Generator().manual_seed(7)
    masks, flat = [], []
    for p_, e in zip(pre, ext):
        m = torch.rand((e, p_ + e), generator=g) < 0.6
        m[:, p_:] &= torch.tril(torch.ones(e, e, dtype=torch.bool))
        m[torch.arange(e), p_ + torch.arange(e)] = True
        masks.append(m)
        flat.append(m.reshape(-1))
    flat = torch.cat(flat)
    mask_indptr = torch.tensor([0] + [int(x) for x in torch.cumsum(torch.tensor([m.numel() for m in masks]), 0)],
                               dtype=torch.int32)
    q = sc.q.to(DEV).view(-1, Hq, D)
    o = torch.zeros_like(q)
    ops.extend_attention(q, k8.to(DEV), v8.to(DEV), o, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=False,
                         page_size=16, max_qo_len=max(ext), custom_mask=flat.to(DEV), mask_indptr=mask_indptr.to(DEV),
                         skip_prefix_custom_mask=skip_prefix, k_descale=KS, v_descale=VS)
    common = (sc.q, kf, vf, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens), torch.tensor(ext), Hq, Hkv,
              D, sc.scaling)
    kw = dict(causal=False, custom_masks=masks, skip_prefix_custom_mask=skip_prefix)
    exact = oracle.attention_fp32(*common, **kw)
    emul = oracle.attention_fp32(*common, e4m3_arith=True, **kw)
    assert_close_e4m3_arith(o.view(sum(ext), -1), emul, exact, f"e4m3 extend tree mask skip_prefix={skip_prefix} D={D}")
