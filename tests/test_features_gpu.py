"""Sliding window, logit cap and attention sinks on the decode and extend kernels, against the
fp32 oracle restating the reference kit's window rule and the production kernels' cap / sink arithmetic.
Tolerance: BASELINE's 1e-2 relative / 2e-2 absolute for bf16 attention outputs."""
import pytest
import torch

from oracle import torch_native_port as oracle
from tests.util import assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _meta(sc, ops):
    B = len(sc.seq_lens)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    qo_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(max(sum(sc.seq_lens), 1), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(sc.req_to_token.to(DEV), sc.req_pool_indices.to(DEV), torch.tensor(sc.seq_lens, device=DEV),
                         kv_indptr, kv_indices, qo_lens=torch.tensor(sc.extend_lens, dtype=torch.int32, device=DEV),
                         qo_indptr=qo_indptr)
    return kv_indptr, kv_indices, qo_indptr


def _filled(sc):
    """Pool with the new tokens already stored (single KV source)."""
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    kc[sc.out_cache_loc] = sc.k
    vc[sc.out_cache_loc] = sc.v
    return kc, vc


FEATS = [
    dict(window_left=0), dict(window_left=3), dict(window_left=63), dict(window_left=64), dict(window_left=200),
    dict(window_left=5000), dict(logit_cap=30.0), dict(logit_cap=5.0, window_left=100), dict(sinks=True),
    dict(sinks=True, window_left=127, logit_cap=50.0),
]


@pytest.mark.parametrize("feat", FEATS, ids=lambda f: "-".join(f"{k}{v}" for k, v in f.items()))
@pytest.mark.parametrize("Hq,Hkv,D", [(8, 2, 128), (14, 2, 64)])
def test_decode_features(feat, Hq, Hkv, D):
    from sglang_b200 import ops
    lens = [1, 64, 65, 300, 1023, 2500]
    sc = make_scenario([l - 1 for l in lens], [1] * len(lens), Hq, Hkv, D, 16, seed=11)
    kc, vc = _filled(sc)
    kv_indptr, kv_indices, _ = _meta(sc, ops)
    f = dict(feat)
    sinks = None
    if f.pop("sinks", False):
        sinks = torch.linspace(-2.0, 3.0, Hq)
    q = sc.q.to(DEV).view(-1, Hq, D)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens),
                                torch.ones(len(lens), dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False,
                                sinks=sinks, **f)
    for max_splits in (1, 8):
        o = torch.empty_like(q)
        ws = torch.zeros(max(ops.decode_workspace_bytes(len(lens), Hq, D, max_splits), 16), dtype=torch.uint8,
                         device=DEV)
        ops.decode_attention(q, kc.to(DEV), vc.to(DEV), o, kv_indptr, kv_indices, sc.scaling, max_splits=max_splits,
                             min_chunk=128, workspace=ws, sinks=None if sinks is None else sinks.to(DEV), **f)
        assert_close_attn(o.view(len(lens), -1), ref, f"decode {feat} splits {max_splits}")


@pytest.mark.parametrize("feat", FEATS, ids=lambda f: "-".join(f"{k}{v}" for k, v in f.items()))
@pytest.mark.parametrize("causal", [True, False])
def test_extend_features(feat, causal):
    from sglang_b200 import ops
    Hq, Hkv, D = 8, 2, 128
    pre, ext = [0, 100, 700, 5, 1000], [300, 1, 260, 129, 64]
    sc = make_scenario(pre, ext, Hq, Hkv, D, 16, seed=13)
    kc, vc = _filled(sc)
    kv_indptr, kv_indices, qo_indptr = _meta(sc, ops)
    f = dict(feat)
    sinks = None
    if f.pop("sinks", False):
        sinks = torch.linspace(-2.0, 3.0, Hq)
    q = sc.q.to(DEV).view(-1, Hq, D)
    o = torch.zeros_like(q)
    ops.extend_attention(q, kc.to(DEV), vc.to(DEV), o, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=causal,
                         page_size=16, max_qo_len=max(ext), sinks=None if sinks is None else sinks.to(DEV), **f)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                                torch.tensor(ext), Hq, Hkv, D, sc.scaling, causal=causal, sinks=sinks, **f)
    assert_close_attn(o.view(sum(ext), -1), ref, f"extend {feat} causal={causal}")


def test_backend_layer_flags_decode_equals_mixed_row():
    """RadixAttention(sliding_window_size, logit_cap) + sinks kwarg through B200AttnBackend: a running decode gives
    the same row in a DECODE batch and as an extend_len == 1 row of a MIXED batch."""
    from sglang_b200 import ForwardBatch, ForwardMode, RadixAttention
    from tests.test_backend_gpu import _setup
    Hq, Hkv, D, page = 8, 2, 128, 16
    pre, ext = [400, 0, 777], [1, 200, 1]
    sc = make_scenario(pre, ext, Hq, Hkv, D, page, seed=17)
    sinks = torch.linspace(-1.0, 1.0, Hq, device=DEV)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0, logit_cap=20.0, sliding_window_size=128)
    backend, req_pool, kv_pool = _setup(sc, False)
    fb = ForwardBatch(forward_mode=ForwardMode.MIXED, batch_size=3, req_pool_indices=sc.req_pool_indices.to(DEV),
                      seq_lens=torch.tensor(sc.seq_lens, device=DEV), out_cache_loc=sc.out_cache_loc.to(DEV),
                      seq_lens_sum=sum(sc.seq_lens), extend_seq_lens=torch.tensor(ext, dtype=torch.int32, device=DEV),
                      extend_prefix_lens=torch.tensor(pre, device=DEV), extend_seq_lens_cpu=ext,
                      extend_prefix_lens_cpu=pre, attn_backend=backend)
    backend.init_forward_metadata(fb)
    o_mixed = layer(sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV), fb, sinks=sinks)
    kc, vc = kv_pool.k_buffer[0].cpu(), kv_pool.v_buffer[0].cpu()
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                                torch.tensor(ext), Hq, Hkv, D, sc.scaling, causal=True, window_left=128, logit_cap=20.0,
                                sinks=sinks.cpu())
    assert_close_attn(o_mixed, ref, "mixed with flags")
    # the two decode rows again, as a DECODE batch
    rows = [0, 201]
    reqs = [0, 2]
    fb2 = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=2,
                       req_pool_indices=sc.req_pool_indices[reqs].to(DEV),
                       seq_lens=torch.tensor([sc.seq_lens[r] for r in reqs], device=DEV),
                       out_cache_loc=sc.out_cache_loc[rows].to(DEV), seq_lens_sum=sum(sc.seq_lens[r] for r in reqs),
                       attn_backend=backend)
    backend.init_forward_metadata(fb2)
    o_dec = layer(sc.q[rows].to(DEV), sc.k[rows].to(DEV), sc.v[rows].to(DEV), fb2, sinks=sinks)
    assert torch.equal(o_dec, o_mixed[rows])


@pytest.mark.parametrize("skip_prefix", [True, False])
@pytest.mark.parametrize("Hq,Hkv,D", [(8, 2, 128), (14, 2, 64)])
def test_extend_custom_mask_target_verify(skip_prefix, Hq, Hkv, D):
    """TARGET_VERIFY tree masks: request b owns the row-major block custom_mask[mask_indptr[b] + j * L + n]."""
    from sglang_b200 import ops
    pre, ext = [37, 0, 300, 1000, 129], [8, 8, 8, 8, 8]   # 8 draft tokens per request
    sc = make_scenario(pre, ext, Hq, Hkv, D, 16, seed=23)
    kc, vc = _filled(sc)
    kv_indptr, kv_indices, qo_indptr = _meta(sc, ops)
    g = torch.Generator().manual_seed(7)
    masks, flat = [], []
    for p_, e in zip(pre, ext):
        L = p_ + e
        m = torch.rand((e, L), generator=g) < 0.6
        tri = torch.tril(torch.ones(e, e, dtype=torch.bool))          # a draft token never sees later draft tokens
        m[:, p_:] &= tri
        m[torch.arange(e), p_ + torch.arange(e)] = True               # ... and always sees itself
        masks.append(m)
        flat.append(m.reshape(-1))
    flat = torch.cat(flat)
    mask_indptr = torch.tensor([0] + [int(x) for x in torch.cumsum(torch.tensor([m.numel() for m in masks]), 0)],
                               dtype=torch.int32)
    q = sc.q.to(DEV).view(-1, Hq, D)
    o = torch.zeros_like(q)
    ops.extend_attention(q, kc.to(DEV), vc.to(DEV), o, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=False,
                         page_size=16, max_qo_len=max(ext), custom_mask=flat.to(DEV), mask_indptr=mask_indptr.to(DEV),
                         skip_prefix_custom_mask=skip_prefix)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                                torch.tensor(ext), Hq, Hkv, D, sc.scaling, causal=False, custom_masks=masks,
                                skip_prefix_custom_mask=skip_prefix)
    assert_close_attn(o.view(sum(ext), -1), ref, f"custom mask skip_prefix={skip_prefix}")


def test_backend_target_verify_tree_mask():
    """ForwardMode.TARGET_VERIFY through B200AttnBackend: seq_lens counts cached tokens only, the draft tokens are
    stored at out_cache_loc and appended to the KV index list, the tree mask replaces the causal mask."""
    from types import SimpleNamespace

    from sglang_b200 import ForwardBatch, ForwardMode, RadixAttention
    from tests.test_backend_gpu import _setup
    Hq, Hkv, D, page, nd = 8, 2, 128, 16, 6
    pre = [50, 0, 333, 16]
    sc = make_scenario(pre, [nd] * len(pre), Hq, Hkv, D, page, seed=29)
    B = len(pre)
    g = torch.Generator().manual_seed(9)
    masks = []
    for p_ in pre:
        m = torch.ones((nd, p_ + nd), dtype=torch.bool)
        tree = torch.tril(torch.rand((nd, nd), generator=g) < 0.5)
        tree[torch.arange(nd), torch.arange(nd)] = True
        m[:, p_:] = tree
        masks.append(m)
    flat = torch.cat([m.reshape(-1) for m in masks])
    backend, req_pool, kv_pool = _setup(sc, False)
    # the verify batch has NOT written its draft slots into req_to_token: wipe them to prove out_cache_loc is used
    for b, p_ in enumerate(pre):
        req_pool.req_to_token[b + 1, p_:p_ + nd] = 0
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    fb = ForwardBatch(forward_mode=ForwardMode.TARGET_VERIFY, batch_size=B, req_pool_indices=sc.req_pool_indices.to(DEV),
                      seq_lens=torch.tensor(pre, device=DEV), out_cache_loc=sc.out_cache_loc.to(DEV),
                      seq_lens_sum=sum(pre), attn_backend=backend,
                      spec_info=SimpleNamespace(custom_mask=flat.to(DEV), draft_token_num=nd))
    backend.init_forward_metadata(fb)
    o = layer(sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV), fb)
    kc, vc = _filled(sc)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                                torch.tensor([nd] * B), Hq, Hkv, D, sc.scaling, causal=False, custom_masks=masks)
    assert_close_attn(o, ref, "target verify")
    assert torch.equal(kv_pool.k_buffer[0][sc.out_cache_loc.to(DEV)], sc.k.to(DEV))
