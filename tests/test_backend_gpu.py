"""B200AttnBackend behind the reference plugin API: golden cases end to end + CUDA-graph decode."""
import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import (AttentionType, B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool,
                         RadixAttention, ReqToTokenPool)
from tests.util import assert_close_attn, golden_files, load_golden, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"


class _Runner:
    def __init__(self, req_pool, kv_pool, page_size):
        self.device = DEV
        self.page_size = page_size
        self.req_to_token_pool = req_pool
        self.token_to_kv_pool = kv_pool
        self.sliding_window_size = None


def _setup(g_or_sc, golden):
    if golden:
        g = g_or_sc
        Hq, Hkv, D, page = g["num_heads"], g["num_kv_heads"], g["head_dim"], g["page_size"]
        table, kc, vc = g["req_to_token"], g["k_pool_before"], g["v_pool_before"]
    else:
        sc = g_or_sc
        Hq, Hkv, D, page = sc.Hq, sc.Hkv, sc.D, sc.page_size
        table, kc, vc = sc.req_to_token, sc.k_cache, sc.v_cache
    req_pool = ReqToTokenPool(table.shape[0] - 1, table.shape[1], DEV)
    req_pool.req_to_token.copy_(table)
    kv_pool = MHATokenToKVPool(kc.shape[0] - page, page, kc.dtype, Hkv, D, 1, DEV, skip_padding_slot=False)
    kv_pool.k_buffer[0].copy_(kc)
    kv_pool.v_buffer[0].copy_(vc)
    backend = B200AttnBackend(_Runner(req_pool, kv_pool, page), skip_padding_slot=False)
    return backend, req_pool, kv_pool


@pytest.mark.parametrize("path", golden_files(), ids=lambda p: p.split("/")[-1][:-3])
def test_backend_matches_reference_backend(path):
    g = load_golden(path)
    Hq, Hkv, D = g["num_heads"], g["num_kv_heads"], g["head_dim"]
    backend, req_pool, kv_pool = _setup(g, True)
    layer = RadixAttention(Hq, D, g["scaling"], Hkv, 0,
                           attn_type=AttentionType.DECODER if g["causal"] else AttentionType.ENCODER_ONLY)
    mode = ForwardMode[g["mode"]]
    B = len(g["seq_lens"])
    fb = ForwardBatch(forward_mode=mode, batch_size=B, req_pool_indices=g["req_pool_indices"].to(DEV),
                      seq_lens=torch.tensor(g["seq_lens"], dtype=torch.int32, device=DEV),
                      out_cache_loc=g["out_cache_loc"].to(DEV), seq_lens_sum=sum(g["seq_lens"]),
                      seq_lens_cpu=torch.tensor(g["seq_lens"], dtype=torch.int32), positions=g["positions"].to(DEV),
                      attn_backend=backend)
    if mode.is_extend():
        fb.extend_prefix_lens = torch.tensor(g["prefix_lens"], dtype=torch.int32, device=DEV)
        fb.extend_seq_lens = torch.tensor(g["extend_lens"], dtype=torch.int32, device=DEV)
        fb.extend_prefix_lens_cpu, fb.extend_seq_lens_cpu = g["prefix_lens"], g["extend_lens"]
        fb.extend_num_tokens = sum(g["extend_lens"])
    backend.init_forward_metadata(fb)
    o = layer(g["q"].to(DEV), g["k"].to(DEV), g["v"].to(DEV), fb)
    assert o.shape == g["o_ref"].shape and o.dtype == g["o_ref"].dtype
    assert_close_attn(o, g["o_ref"], g["name"])
    loc = g["out_cache_loc"].long()
    assert torch.equal(kv_pool.k_buffer[0].cpu()[loc], g["k_rows_after"])  # store is bit-exact
    assert torch.equal(kv_pool.v_buffer[0].cpu()[loc], g["v_rows_after"])


def test_backend_idle_and_unsupported_features():
    sc = make_scenario([4], [0], 8, 2, 128, 16)
    backend, _, _ = _setup(sc, False)
    layer = RadixAttention(8, 128, 0.1, 2, 0)
    fb = ForwardBatch(forward_mode=ForwardMode.IDLE, batch_size=0, req_pool_indices=torch.zeros(0, device=DEV),
                      seq_lens=torch.zeros(0, device=DEV), out_cache_loc=torch.zeros(0, device=DEV),
                      attn_backend=backend)
    backend.init_forward_metadata(fb)
    q = torch.zeros((0, 8 * 128), dtype=torch.bfloat16, device=DEV)
    assert layer(q, None, None, fb).shape == (0, 8 * 128)
    fb.forward_mode = ForwardMode.DECODE
    for kw in (dict(is_cross_attention=True), dict(v_head_dim=64)):
        bad = RadixAttention(8, 128, 0.1, 2, 0, **kw)
        with pytest.raises(NotImplementedError):
            backend.forward_decode(q, None, None, bad, fb)


def test_backend_cuda_graph_decode_replay():
    """Capture one decode step, then replay it with new lengths / slots (base_attn_backend.py:73-107)."""
    Hq, Hkv, D, page, B = 32, 8, 128, 16, 8
    lens0 = [100, 37, 512, 64, 900, 3, 258, 777]
    sc = make_scenario([l - 1 for l in lens0], [1] * B, Hq, Hkv, D, page, seed=77, pool_slack=3.0)
    backend, req_pool, kv_pool = _setup(sc, False)
    backend.init_cuda_graph_state(B, B, num_q_heads=Hq, head_dim=D)
    assert backend.get_cuda_graph_seq_len_fill_value() == 1
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    # static input buffers, as the reference graph runner keeps them (runner_utils/buffers.py:113-117)
    s_req = sc.req_pool_indices.to(DEV)
    s_seq = torch.tensor(lens0, dtype=torch.int64, device=DEV)
    s_loc = sc.out_cache_loc.to(DEV)
    s_q = sc.q.to(DEV)
    s_k, s_v = sc.k.to(DEV), sc.v.to(DEV)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=B, req_pool_indices=s_req, seq_lens=s_seq,
                      out_cache_loc=s_loc, seq_lens_sum=sum(lens0), attn_backend=backend)
    stream = torch.cuda.Stream()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        backend.init_forward_metadata_out_graph(fb, in_capture=True)
        layer(s_q, s_k, s_v, fb)  # warm-up outside capture
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            backend.init_forward_metadata_in_graph(fb)
            s_out = layer(s_q, s_k, s_v, fb)
    torch.cuda.current_stream().wait_stream(stream)

    def check(lens, seed):
        sc2 = make_scenario([l - 1 for l in lens], [1] * B, Hq, Hkv, D, page, seed=seed, pool_slack=1.0)
        # new request state: table rows, slots, inputs (pool contents stay: slots are re-used)
        n = min(req_pool.req_to_token.shape[1], sc2.req_to_token.shape[1])
        req_pool.req_to_token.zero_()
        req_pool.req_to_token[:, :n].copy_(sc2.req_to_token[:, :n] % kv_pool.k_buffer[0].shape[0])
        s_seq.copy_(torch.tensor(lens))
        loc = torch.stack([req_pool.req_to_token[b + 1, lens[b] - 1] for b in range(B)]).long()
        s_loc.copy_(loc)
        s_q.copy_(sc2.q)
        s_k.copy_(sc2.k)
        s_v.copy_(sc2.v)
        backend.init_forward_metadata_out_graph(fb)  # replay-time refill, no capture
        graph.replay()
        torch.cuda.synchronize()
        kc, vc = kv_pool.k_buffer[0], kv_pool.v_buffer[0]
        ref = oracle.attention_fp32(s_q, kc, vc, req_pool.req_to_token, s_req, torch.tensor(lens),
                                    torch.ones(B, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
        assert_close_attn(s_out, ref, f"graph replay seed {seed}")
        assert torch.equal(kc[loc], s_k.view(B, Hkv, D))  # the captured store ran with the new slots

    check(lens0, 5)
    check([1, 900, 17, 16, 333, 640, 2, 45], 6)
    md = backend.forward_metadata
    assert int(md.kv_indptr[0]) == 0 and bool((md.kv_indptr[1:B + 1] >= md.kv_indptr[:B]).all())


def _extend_fb(sc, backend, mode):
    B = len(sc.seq_lens)
    return ForwardBatch(forward_mode=mode, batch_size=B, req_pool_indices=sc.req_pool_indices.to(DEV),
                        seq_lens=torch.tensor(sc.seq_lens, dtype=torch.int64, device=DEV),
                        out_cache_loc=sc.out_cache_loc.to(DEV), seq_lens_sum=sum(sc.seq_lens),
                        extend_seq_lens=torch.tensor(sc.extend_lens, dtype=torch.int32, device=DEV),
                        extend_prefix_lens=torch.tensor(sc.prefix_lens, dtype=torch.int32, device=DEV),
                        extend_seq_lens_cpu=list(sc.extend_lens), extend_prefix_lens_cpu=list(sc.prefix_lens),
                        attn_backend=backend)


@pytest.mark.parametrize("causal", [True, False])
def test_backend_mixed_routes_single_rows_to_decode(causal):
    """MIXED = chunked prefill rows + running decodes as extend_len == 1 rows (schedule_batch.py:2758-2788):
    the backend sends the single rows through the decode kernel (q_rows / req_ids) and the extend kernel skips
    them; interleaved order and G = 7 (Qwen2.5-VL-7B head layout, BASELINE configs[3] scaled down)."""
    prefix = [0, 700, 300, 1279, 0, 63]
    ext = [320, 1, 200, 1, 1, 1]
    sc = make_scenario(prefix, ext, 28, 4, 128, 16, seed=91)
    backend, req_pool, kv_pool = _setup(sc, False)
    layer = RadixAttention(28, 128, sc.scaling, 4, 0,
                           attn_type=AttentionType.DECODER if causal else AttentionType.ENCODER_ONLY)
    fb = _extend_fb(sc, backend, ForwardMode.MIXED)
    backend.init_forward_metadata(fb)
    md = backend.forward_metadata
    assert md.num_dec == 4 and md.max_extend_len == 320
    assert md.dec_q_rows.tolist() == [320, 521, 522, 523] and md.dec_req_ids.tolist() == [1, 3, 4, 5]
    o = layer(sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV), fb)
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    oracle.store_kv(kc, vc, sc.out_cache_loc, sc.k, sc.v)
    assert torch.equal(kv_pool.k_buffer[0].cpu(), kc) and torch.equal(kv_pool.v_buffer[0].cpu(), vc)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                                torch.tensor(ext), 28, 4, 128, sc.scaling, causal=causal)
    assert_close_attn(o, ref, f"mixed routed causal={causal}")


def test_backend_extend_all_single_rows():
    """EXTEND batch in which every request brings one token: nothing left for the tensor-core kernel."""
    prefix, ext = [5, 130, 0, 64], [1, 1, 1, 1]
    sc = make_scenario(prefix, ext, 8, 2, 128, 16, seed=92)
    backend, _, _ = _setup(sc, False)
    layer = RadixAttention(8, 128, sc.scaling, 2, 0)
    fb = _extend_fb(sc, backend, ForwardMode.EXTEND)
    backend.init_forward_metadata(fb)
    assert backend.forward_metadata.max_extend_len == 0 and backend.forward_metadata.num_dec == 4
    o = layer(sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV), fb)
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    oracle.store_kv(kc, vc, sc.out_cache_loc, sc.k, sc.v)
    ref = oracle.attention_fp32(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(sc.seq_lens),
                                torch.tensor(ext), 8, 2, 128, sc.scaling, causal=True)
    assert_close_attn(o, ref, "all-single extend")


def test_backend_cuda_graph_extend_replay():
    """Full-graph prefill: capture ONE extend step over static buffers, replay it with other ragged batches
    (fewer real requests, zero-length padded rows, a single-token row) — prefill_cuda_graph_runner.py:985-1030."""
    Hq, Hkv, D, page, B, TCAP = 8, 2, 128, 16, 4, 1024
    pre0, ext0 = [64, 0, 300, 17], [200, 128, 40, 333]
    sc = make_scenario([0] * B, [TCAP] * B, Hq, Hkv, D, page, seed=91, pool_slack=1.5)  # sizes the table and the pool
    backend, req_pool, kv_pool = _setup(sc, False)
    backend.init_cuda_graph_state(B, TCAP, num_q_heads=Hq, head_dim=D)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    g = torch.Generator().manual_seed(1)
    s_req = sc.req_pool_indices.to(DEV)
    s_seq = torch.zeros(B, dtype=torch.int64, device=DEV)
    s_pre = torch.zeros(B, dtype=torch.int64, device=DEV)
    s_ext = torch.zeros(B, dtype=torch.int32, device=DEV)
    s_loc = torch.zeros(TCAP, dtype=torch.int64, device=DEV)   # padded tokens write the reserved slot 0
    s_q = torch.zeros((TCAP, Hq * D), dtype=torch.bfloat16, device=DEV)
    s_k = torch.zeros((TCAP, Hkv, D), dtype=torch.bfloat16, device=DEV)
    s_v = torch.zeros((TCAP, Hkv, D), dtype=torch.bfloat16, device=DEV)
    fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=B, req_pool_indices=s_req, seq_lens=s_seq,
                      out_cache_loc=s_loc, seq_lens_sum=0, extend_seq_lens=s_ext, extend_prefix_lens=s_pre,
                      attn_backend=backend)

    def fill(pre, ext, seed):
        sc2 = make_scenario(pre, ext, Hq, Hkv, D, page, seed=seed, pool_slack=1.0)
        n = min(req_pool.req_to_token.shape[1], sc2.req_to_token.shape[1])
        req_pool.req_to_token.zero_()
        req_pool.req_to_token[:, :n].copy_(sc2.req_to_token[:, :n] % kv_pool.k_buffer[0].shape[0])
        T = sum(ext)
        s_seq.copy_(torch.tensor([p + e for p, e in zip(pre, ext)]))
        s_pre.copy_(torch.tensor(pre))
        s_ext.copy_(torch.tensor(ext, dtype=torch.int32))
        loc = torch.cat([req_pool.req_to_token[b + 1, pre[b]:pre[b] + ext[b]] for b in range(B)]).long()
        s_loc.zero_()
        s_loc[:T].copy_(loc)
        s_q.zero_(); s_k.zero_(); s_v.zero_()
        s_q[:T].copy_(sc2.q); s_k[:T].copy_(sc2.k); s_v[:T].copy_(sc2.v)
        return T, loc

    fill(pre0, ext0, 91)
    stream = torch.cuda.Stream()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        md0 = backend.init_forward_metadata_for_breakable_cuda_graph_capture(fb)
        layer(s_q, s_k, s_v, fb)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            backend.init_forward_metadata_in_graph(fb)
            s_out = layer(s_q, s_k, s_v, fb)
    torch.cuda.current_stream().wait_stream(stream)
    assert backend.use_captured_forward_metadata_for_breakable_cuda_graph

    for pre, ext, seed in [(pre0, ext0, 92), ([0, 5, 0, 0], [700, 1, 0, 0], 93), ([511, 0, 40, 0], [1, 0, 129, 600], 94),
                           ([10, 20, 30, 40], [128, 256, 128, 512], 95)]:
        T, loc = fill(pre, ext, seed)
        backend.prepare_forward_metadata_for_breakable_cuda_graph_replay(md0, fb)
        assert backend.forward_metadata is md0
        graph.replay()
        torch.cuda.synchronize()
        kc, vc = kv_pool.k_buffer[0], kv_pool.v_buffer[0]
        ref = oracle.attention_fp32(s_q[:T], kc, vc, req_pool.req_to_token, s_req, s_seq.cpu(), s_ext.cpu(), Hq, Hkv, D,
                                    sc.scaling, causal=True)
        assert_close_attn(s_out[:T], ref, f"extend graph replay {ext}")
        assert torch.equal(kc[loc], s_k[:T])


def test_baseline_config0_qwen2_0p5b_decode_page1():
    """BASELINE configs[0] at its exact shape: Qwen2-0.5B (Hq 14 / Hkv 2 / D 64), decode batch 8 x 1K ctx,
    page_size 1, through the backend (store + decode) against the torch_native port."""
    Hq, Hkv, D, B, ctx = 14, 2, 64, 8, 1024
    sc = make_scenario([ctx - 1] * B, [1] * B, Hq, Hkv, D, 1, seed=20261104)
    backend, req_pool, kv_pool = _setup(sc, False)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=B, req_pool_indices=sc.req_pool_indices.to(DEV),
                      seq_lens=torch.full((B,), ctx, dtype=torch.int64, device=DEV),
                      out_cache_loc=sc.out_cache_loc.to(DEV), seq_lens_sum=B * ctx,
                      seq_lens_cpu=torch.full((B,), ctx, dtype=torch.int64), attn_backend=backend)
    backend.init_forward_metadata(fb)
    o = layer(sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV), fb)
    torch.cuda.synchronize()
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    oracle.store_kv(kc, vc, sc.out_cache_loc, sc.k.view(B, Hkv, D), sc.v.view(B, Hkv, D))
    ref = oracle.decode(sc.q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.full((B,), ctx), Hq, Hkv, D, sc.scaling)
    assert torch.equal(kv_pool.k_buffer[0].cpu(), kc) and torch.equal(kv_pool.v_buffer[0].cpu(), vc)
    assert_close_attn(o, ref, "configs[0] qwen2-0.5b decode")
