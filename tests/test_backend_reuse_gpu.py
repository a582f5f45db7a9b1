"""One B200AttnBackend object across forwards of different shapes, the way a server uses it.

Covers what the per-shape tests cannot: the split-decode workspace (arrival counters + fp32 partials) is reused
across batch sizes / split counts (eager and captured graphs sharing one buffer), and CUDA-graph replay with a live
batch SMALLER than the captured bucket, padded the way the reference runner pads it
(python/sglang/srt/model_executor/runner_utils/buffers.py:113-115,236-238: seq_lens <- fill value 1,
req_pool_indices <- 0 = the padding row of req_to_token, out_cache_loc <- 0 = the reserved slot).
"""
import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import (B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention,
                         ReqToTokenPool)
from tests.util import assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"
Hq, Hkv, D, PAGE = 32, 8, 128, 16


class _Runner:
    def __init__(self, req_pool, kv_pool, page_size):
        self.device = DEV
        self.page_size = page_size
        self.req_to_token_pool = req_pool
        self.token_to_kv_pool = kv_pool
        self.sliding_window_size = None


def _setup(sc, **backend_kw):
    req_pool = ReqToTokenPool(sc.req_to_token.shape[0] - 1, sc.req_to_token.shape[1], DEV)
    req_pool.req_to_token.copy_(sc.req_to_token)
    kv_pool = MHATokenToKVPool(sc.k_cache.shape[0] - PAGE, PAGE, sc.k_cache.dtype, Hkv, D, 1, DEV)
    kv_pool.k_buffer[0].copy_(sc.k_cache)
    kv_pool.v_buffer[0].copy_(sc.v_cache)
    return B200AttnBackend(_Runner(req_pool, kv_pool, PAGE), **backend_kw), req_pool, kv_pool


def _decode_rows(sc, rows, backend, req_pool, kv_pool, layer, what):
    """Eager decode of the requests `rows` of scenario `sc` (their new token is stored first), vs the fp32 oracle."""
    rows_t = torch.tensor(rows)
    lens = torch.tensor([sc.seq_lens[b] for b in rows])
    req = sc.req_pool_indices[rows_t]
    loc = sc.out_cache_loc[rows_t]
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=len(rows), req_pool_indices=req.to(DEV),
                      seq_lens=lens.to(DEV), out_cache_loc=loc.to(DEV), seq_lens_sum=int(lens.sum()),
                      seq_lens_cpu=lens, attn_backend=backend)
    backend.init_forward_metadata(fb)
    q, k, v = sc.q[rows_t].to(DEV), sc.k[rows_t].to(DEV), sc.v[rows_t].to(DEV)
    o = layer(q, k, v, fb)
    torch.cuda.synchronize()
    ref = oracle.attention_fp32(q, kv_pool.k_buffer[0], kv_pool.v_buffer[0], req_pool.req_to_token, req.to(DEV), lens,
                                torch.ones(len(rows), dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
    assert_close_attn(o, ref, what)


def test_backend_eager_decode_reuses_workspace_across_batch_sizes():
    """B = 8 (4 splits per unit) -> B = 1 (16+ splits) -> B = 3 -> B = 8 on ONE backend / ONE workspace: the arrival
    counters must not move with the launch shape (they used to sit behind the partials, whose size depends on
    B * max_splits, so a smaller batch found stale fp32 partials where it expected zeroed counters)."""
    lens = [4000, 2100, 3333, 4096, 1500, 2900, 4090, 777]
    sc = make_scenario([l - 1 for l in lens], [1] * 8, Hq, Hkv, D, PAGE, seed=311)
    backend, req_pool, kv_pool = _setup(sc)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    for rows in ([0, 1, 2, 3, 4, 5, 6, 7], [3], [6, 0, 2], [0, 1, 2, 3, 4, 5, 6, 7], [7], [3]):
        _decode_rows(sc, rows, backend, req_pool, kv_pool, layer, f"eager reuse rows {rows}")


def test_backend_cuda_graph_buckets_share_workspace_and_padded_replay():
    """Graphs for bs 8 and bs 1 captured on one backend (shared static workspace) and replayed alternately; then the
    bs-8 graph replayed with 5 live requests padded to the bucket.  Production default skip_padding_slot=True: the
    padded tokens (out_cache_loc 0) must not touch slot 0, and the live rows must match the oracle."""
    BMAX = 8
    lens0 = [4000, 2100, 3333, 4096, 1500, 2900, 4090, 777]
    sc = make_scenario([l - 1 for l in lens0], [1] * BMAX, Hq, Hkv, D, PAGE, seed=312, pool_slack=1.5)
    backend, req_pool, kv_pool = _setup(sc)
    assert backend.skip_padding_slot
    backend.init_cuda_graph_state(BMAX, BMAX, num_q_heads=Hq, head_dim=D)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    s_req = torch.zeros(BMAX, dtype=torch.int64, device=DEV)
    s_seq = torch.ones(BMAX, dtype=torch.int64, device=DEV)
    s_loc = torch.zeros(BMAX, dtype=torch.int64, device=DEV)
    s_q = torch.zeros((BMAX, Hq * D), dtype=torch.bfloat16, device=DEV)
    s_k = torch.zeros((BMAX, Hkv, D), dtype=torch.bfloat16, device=DEV)
    s_v = torch.zeros((BMAX, Hkv, D), dtype=torch.bfloat16, device=DEV)

    def fill(rows):
        """Live requests `rows` first, the rest of the bucket padded as the runner pads it."""
        n = len(rows)
        rt = torch.tensor(rows)
        s_req.zero_(); s_seq.fill_(backend.get_cuda_graph_seq_len_fill_value()); s_loc.zero_()
        s_q.normal_(); s_k.normal_(); s_v.normal_()   # padded rows carry garbage on purpose
        s_req[:n].copy_(sc.req_pool_indices[rt]); s_seq[:n].copy_(torch.tensor([sc.seq_lens[b] for b in rows]))
        s_loc[:n].copy_(sc.out_cache_loc[rt])
        s_q[:n].copy_(sc.q[rt]); s_k[:n].copy_(sc.k[rt]); s_v[:n].copy_(sc.v[rt])

    graphs = {}
    stream = torch.cuda.Stream()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        for bs in (BMAX, 1):  # large bucket first, as the reference captures (reversed capture_bs)
            fill(list(range(bs)))
            fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs, req_pool_indices=s_req[:bs],
                              seq_lens=s_seq[:bs], out_cache_loc=s_loc[:bs], seq_lens_sum=0, attn_backend=backend)
            backend.init_forward_metadata_out_graph(fb, in_capture=True)
            layer(s_q[:bs], s_k[:bs], s_v[:bs], fb)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                backend.init_forward_metadata_in_graph(fb)
                out = layer(s_q[:bs], s_k[:bs], s_v[:bs], fb)
            graphs[bs] = (g, fb, out)
    torch.cuda.current_stream().wait_stream(stream)

    def replay(bs, rows, what):
        g, fb, out = graphs[bs]
        fill(rows)
        backend.init_forward_metadata_out_graph(fb)
        g.replay()
        torch.cuda.synchronize()
        n = len(rows)
        kc, vc = kv_pool.k_buffer[0], kv_pool.v_buffer[0]
        assert not kc[:PAGE].any() and not vc[:PAGE].any(), f"{what}: the padding page was written"
        ref = oracle.attention_fp32(s_q[:n], kc, vc, req_pool.req_to_token, s_req[:n], s_seq[:n].cpu(),
                                    torch.ones(n, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
        assert_close_attn(out[:n], ref, what)
        assert torch.isfinite(out.float()).all(), f"{what}: padded rows produced non-finite values"
        assert torch.equal(kc[s_loc[:n]], s_k[:n])
        md = backend.forward_metadata
        assert int(md.kv_indptr[0]) == 0 and bool((md.kv_indptr[1:bs + 1] >= md.kv_indptr[:bs]).all())

    replay(BMAX, list(range(8)), "bucket 8 full")
    replay(1, [3], "bucket 1 after bucket 8")
    replay(BMAX, list(range(8)), "bucket 8 after bucket 1")
    replay(1, [6], "bucket 1 again")
    replay(BMAX, [5, 1, 7, 0, 2], "bucket 8 with 5 live requests + 3 padded rows")
    replay(BMAX, [4], "bucket 8 with 1 live request + 7 padded rows")
