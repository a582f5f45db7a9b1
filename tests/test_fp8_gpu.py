"""(f4) FP8 (e4m3) KV cache: quantising store (bit-exact), fp8 decode kernel, extend over an fp8 pool, backend."""
import types

import pytest
import torch

from oracle import fp8_port
from oracle import torch_native_port as oracle
from sglang_b200 import (B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention,
                         ReqToTokenPool, ops)
from tests.util import assert_close_e4m3_arith, assert_close_attn, make_scenario

pytestmark = pytest.mark.gpu
DEV = "cuda"
FP8 = torch.float8_e4m3fn


def _all_16bit(dtype):
    x = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(dtype)
    return x[~torch.isnan(x)]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("scale", [None, 0.5, 0.37, 2.3])
def test_store_fp8_exhaustive_bit_exact(dtype, scale):
    """Every finite 16-bit input value, with and without a scale, against the reference expression on the CPU."""
    x = _all_16bit(dtype)
    n = x.numel() // 64 * 64
    x = x[:n].view(-1, 1, 64)
    T = x.shape[0]
    loc = torch.randperm(T, generator=torch.Generator().manual_seed(1)) + 1
    kc = torch.zeros((T + 1, 1, 64), dtype=torch.uint8)
    vc = torch.zeros_like(kc)
    fp8_port.store_kv_fp8(kc, vc, loc, x, x.flip(0), scale, None)
    kg, vg = torch.zeros_like(kc, device=DEV), torch.zeros_like(vc, device=DEV)
    ops.store_cache_fp8(x.to(DEV), x.flip(0).to(DEV), kg, vg, loc.to(DEV), scale, None)
    assert torch.equal(kg.cpu(), kc)
    assert torch.equal(vg.cpu(), vc)


@pytest.mark.parametrize("width,stride", [(96, 96), (100, 104), (7, 9), (2048, 6144)])
def test_store_fp8_odd_widths_strides_and_skip(width, stride):
    g = torch.Generator().manual_seed(width)
    T = 37
    buf = (torch.randn((T, stride), generator=g) * 3).to(torch.bfloat16)
    buf2 = (buf * 1.5).to(torch.bfloat16)
    k, v = buf[:, :width], buf2[:, :width]
    loc = torch.randperm(64, generator=g)[:T]      # contains slot 0 with high probability
    loc[3] = 0
    kc = torch.full((64, width), 7, dtype=torch.uint8)
    vc = torch.full((64, width), 9, dtype=torch.uint8)
    kg, vg = kc.to(DEV).view(FP8), vc.to(DEV).view(FP8)
    fp8_port.store_kv_fp8(kc, vc, loc, k, v, 0.8, 1.3, skip_index=0)
    ops.store_cache_fp8(buf.to(DEV)[:, :width], buf2.to(DEV)[:, :width], kg, vg, loc.to(DEV).int(), 0.8, 1.3)
    assert torch.equal(kg.view(torch.uint8).cpu(), kc) and torch.equal(vg.view(torch.uint8).cpu(), vc)


def _fp8_scenario(lens, ext, Hq, Hkv, D, page, seed, dtype=torch.bfloat16, ks=0.6, vs=0.9):
    """bf16 scenario whose pool is re-quantised to e4m3 bytes; returns (scenario, k8, v8, dequantised fp32 pools)."""
    pre = [l - e for l, e in zip(lens, ext)]
    sc = make_scenario(pre, ext, Hq, Hkv, D, page, dtype=dtype, seed=seed)
    k8 = fp8_port.quantize(sc.k_cache * 2, ks)
    v8 = fp8_port.quantize(sc.v_cache * 2, vs)
    return sc, k8, v8, fp8_port.dequantize(k8, ks), fp8_port.dequantize(v8, vs)


@pytest.mark.parametrize("Hq,Hkv,D,page", [(32, 8, 128, 16), (14, 2, 64, 1), (28, 4, 128, 16), (8, 8, 64, 32),
                                           (16, 1, 128, 1), (64, 2, 128, 16)])
@pytest.mark.parametrize("max_splits", [1, 4, 16])
def test_decode_fp8_matches_oracle(Hq, Hkv, D, page, max_splits):
    lens = [1, 2, 15, 16, 17, 63, 64, 65, 200, 777, 1500]
    B = len(lens)
    sc, k8, v8, kf, vf = _fp8_scenario(lens, [0] * B, Hq, Hkv, D, page, seed=Hq + page)
    q = torch.randn((B, Hq * D), generator=torch.Generator().manual_seed(1)).to(torch.bfloat16)
    ref, lse_ref = oracle.attention_fp32(q, kf, vf, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens),
                                         torch.ones(B, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False,
                                         return_lse=True)
    d = sc.to(DEV)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, torch.tensor(lens, device=DEV), kv_indptr, kv_indices)
    o = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    lse = torch.empty((B, Hq), dtype=torch.float32, device=DEV)
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, max_splits), 16), dtype=torch.uint8, device=DEV)
    for _ in range(2):  # second launch: the split arrival counters must have been left at zero
        ops.decode_attention(q.to(DEV).view(B, Hq, D), k8.to(DEV).view(FP8), v8.to(DEV), o, kv_indptr, kv_indices,
                             sc.scaling, max_splits=max_splits, min_chunk=64, workspace=ws, lse=lse,
                             k_descale=0.6, v_descale=0.9)
    torch.cuda.synchronize()
    assert_close_attn(o.view(B, -1), ref, f"fp8 decode Hq{Hq} D{D} page{page} splits{max_splits}")
    torch.testing.assert_close(lse.cpu(), lse_ref, rtol=2e-3, atol=2e-3)


def test_decode_fp8_fp16_queries_window_cap_sinks():
    lens = [5, 130, 1000]
    Hq, Hkv, D = 8, 2, 128
    sc, k8, v8, kf, vf = _fp8_scenario(lens, [0] * 3, Hq, Hkv, D, 16, seed=3, dtype=torch.float16)
    q = torch.randn((3, Hq * D), generator=torch.Generator().manual_seed(2)).to(torch.float16)
    sinks = torch.randn(Hq, generator=torch.Generator().manual_seed(5))
    d = sc.to(DEV)
    kv_indptr = torch.zeros(4, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, torch.tensor(lens, device=DEV), kv_indptr, kv_indices)
    for kw in (dict(), dict(window_left=100), dict(logit_cap=30.0), dict(sinks=sinks)):
        ref = oracle.attention_fp32(q, kf, vf, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens),
                                    torch.ones(3, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False, **kw)
        for ms in (1, 8):
            o = torch.empty((3, Hq, D), dtype=torch.float16, device=DEV)
            ws = torch.zeros(max(ops.decode_workspace_bytes(3, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
            gkw = {k: (v.to(DEV) if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
            ops.decode_attention(q.to(DEV).view(3, Hq, D), k8.to(DEV), v8.to(DEV), o, kv_indptr, kv_indices,
                                 sc.scaling, max_splits=ms, min_chunk=64, workspace=ws, k_descale=0.6,
                                 v_descale=0.9, **gkw)
            assert_close_attn(o.view(3, -1), ref, f"fp8 decode fp16 {kw} splits{ms}")


def test_dequant_gather_exact():
    g = torch.Generator().manual_seed(0)
    k8 = torch.randint(0, 256, (300, 4, 128), dtype=torch.uint8, generator=g)
    k8[(k8 & 0x7F) == 0x7F] = 0      # no NaN patterns
    v8 = k8.flip(0).contiguous()
    idx = torch.randint(0, 300, (1000,), generator=g).int()
    ok = torch.empty((1000, 4, 128), dtype=torch.bfloat16, device=DEV)
    ov = torch.empty_like(ok)
    ops.kv_dequant_gather(k8.to(DEV), v8.to(DEV).view(FP8), idx.to(DEV), 1000, ok, ov, 1.0, 0.75)
    assert torch.equal(ok.cpu().float(), fp8_port.dequantize(k8)[idx.long()])     # every e4m3 value is a bf16 value
    assert torch.equal(ov.cpu(), (fp8_port.dequantize(v8)[idx.long()] * 0.75).to(torch.bfloat16))


def _backend(sc, Hkv, D, page, k8, v8, fp8_extend=None):
    B = len(sc.seq_lens)
    ctx = sc.req_to_token.shape[1]
    req_pool = ReqToTokenPool(B, ctx, DEV)
    req_pool.req_to_token.copy_(sc.req_to_token)
    slots = sc.k_cache.shape[0]
    pool = MHATokenToKVPool(slots - page, page, FP8, Hkv, D, 1, DEV, skip_padding_slot=False)
    pool.k_buffer[0].copy_(k8)
    pool.v_buffer[0].copy_(v8)

    class Runner:
        device, page_size, sliding_window_size = DEV, page, None
        req_to_token_pool, token_to_kv_pool = req_pool, pool

    return B200AttnBackend(Runner(), skip_padding_slot=False, fp8_extend=fp8_extend), pool


def _check_fp8_extend_rows(o, q, k8, v8, sc, lens, ext, Hq, Hkv, D, ks, vs, fp8_extend, what, single_rows_decode=True):
    """Rows of requests with one query row go through the decode kernel when the host lengths are known (16-bit q and
    p: exact-math tolerance on either path); extend rows: exact-math tolerance on the dequantising path, the reference Triton kernel's e4m3
    arithmetic (oracle.fp8_port.extend_attention_e4m3_arith) on the in-kernel path."""
    exact = oracle.attention_fp32(q, fp8_port.dequantize(k8, ks), fp8_port.dequantize(v8, vs), sc.req_to_token,
                                  sc.req_pool_indices, torch.tensor(lens), torch.tensor(ext), Hq, Hkv, D, D ** -0.5,
                                  causal=True)
    o = o.detach().float().cpu().view(exact.shape)
    single = torch.cat([torch.full((e,), e == 1 and single_rows_decode) for e in ext]) if sum(ext) else \
        torch.zeros(0, dtype=torch.bool)
    if fp8_extend == "dequant":
        assert_close_attn(o, exact, what)
        return
    if single.any():
        assert_close_attn(o[single], exact[single], what + " (decode rows)")
    if (~single).any():
        emul = fp8_port.extend_attention_e4m3_arith(q, k8, v8, sc.req_to_token, sc.req_pool_indices, lens, ext, Hq, Hkv, D,
                                                    D ** -0.5, ks, vs, causal=True)
        assert_close_e4m3_arith(o[~single], emul[~single], exact[~single], what + " (extend rows)")


@pytest.mark.parametrize("D", [128, 64])
@pytest.mark.parametrize("fp8_extend", ["in_kernel", "dequant"])
@pytest.mark.parametrize("mode", ["extend", "mixed", "decode"])
def test_backend_fp8_pool(mode, fp8_extend, D):
    """Store (quantising) + attention through B200AttnBackend on a float8_e4m3fn pool with layer k/v scales."""
    Hq, Hkv, page = 16, 4, 16
    ks, vs = 0.6, 0.9
    if mode == "decode":
        pre, ext = [300, 17, 1024], [1, 1, 1]
    elif mode == "mixed":
        pre, ext = [40, 700, 0, 129], [200, 1, 333, 1]
    else:
        pre, ext = [40, 0, 513], [200, 333, 90]
    lens = [p + e for p, e in zip(pre, ext)]
    sc, k8, v8, _, _ = _fp8_scenario(lens, ext, Hq, Hkv, D, page, seed=11, ks=ks, vs=vs)
    backend, pool = _backend(sc, Hkv, D, page, k8, v8, fp8_extend=fp8_extend)
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    layer.k_scale, layer.v_scale = torch.tensor(ks), torch.tensor(vs)
    layer.k_scale_float, layer.v_scale_float = ks, vs
    T, B = sum(ext), len(ext)
    fmode = ForwardMode.DECODE if mode == "decode" else (ForwardMode.MIXED if mode == "mixed" else ForwardMode.EXTEND)
    kw = dict(forward_mode=fmode, batch_size=B, req_pool_indices=sc.req_pool_indices.to(DEV),
              seq_lens=torch.tensor(lens, device=DEV), out_cache_loc=sc.out_cache_loc.to(DEV), seq_lens_sum=sum(lens),
              seq_lens_cpu=torch.tensor(lens), attn_backend=backend)
    if mode != "decode":
        kw.update(extend_seq_lens=torch.tensor(ext, dtype=torch.int32, device=DEV),
                  extend_prefix_lens=torch.tensor(pre, dtype=torch.int32, device=DEV),
                  extend_seq_lens_cpu=ext, extend_prefix_lens_cpu=pre)
    fb = ForwardBatch(**kw)
    backend.init_forward_metadata(fb)
    o = layer(sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV), fb)
    torch.cuda.synchronize()
    # oracle: quantising store, then fp32 attention over the dequantised pool
    fp8_port.store_kv_fp8(k8, v8, sc.out_cache_loc, sc.k.view(T, Hkv, D), sc.v.view(T, Hkv, D), ks, vs)
    assert torch.equal(pool.k_buffer[0].cpu(), k8) and torch.equal(pool.v_buffer[0].cpu(), v8), "fp8 store not bit-exact"
    _check_fp8_extend_rows(o, sc.q, k8, v8, sc, lens, ext, Hq, Hkv, D, ks, vs, fp8_extend, f"backend fp8 {mode} {fp8_extend}")


def test_backend_fp8_cuda_graph_decode_replay():
    """One captured decode step over an e4m3 pool (quantising store + fp8 decode), replayed with new lengths/slots."""
    Hq, Hkv, D, page, B = 32, 8, 128, 16, 8
    ks, vs = 0.7, 1.1
    lens0 = [100, 37, 512, 64, 900, 3, 258, 777]
    sc, k8, v8, _, _ = _fp8_scenario(lens0, [1] * B, Hq, Hkv, D, page, seed=77, ks=ks, vs=vs)
    backend, pool = _backend(sc, Hkv, D, page, k8, v8)
    req_pool = backend.req_to_token_pool
    backend.init_cuda_graph_state(B, B, num_q_heads=Hq, head_dim=D)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    layer.k_scale, layer.v_scale, layer.k_scale_float, layer.v_scale_float = torch.tensor(ks), torch.tensor(vs), ks, vs
    s_req = sc.req_pool_indices.to(DEV)
    s_seq = torch.tensor(lens0, dtype=torch.int64, device=DEV)
    s_loc = sc.out_cache_loc.to(DEV)
    s_q, s_k, s_v = sc.q.to(DEV), sc.k.to(DEV), sc.v.to(DEV)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=B, req_pool_indices=s_req, seq_lens=s_seq,
                      out_cache_loc=s_loc, seq_lens_sum=sum(lens0), attn_backend=backend)
    stream = torch.cuda.Stream()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        backend.init_forward_metadata_out_graph(fb, in_capture=True)
        layer(s_q, s_k, s_v, fb)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            backend.init_forward_metadata_in_graph(fb)
            s_out = layer(s_q, s_k, s_v, fb)
    torch.cuda.current_stream().wait_stream(stream)

    def check(lens, seed):
        sc2 = make_scenario([l - 1 for l in lens], [1] * B, Hq, Hkv, D, page, seed=seed, pool_slack=1.0)
        n = min(req_pool.req_to_token.shape[1], sc2.req_to_token.shape[1])
        req_pool.req_to_token.zero_()
        req_pool.req_to_token[:, :n].copy_(sc2.req_to_token[:, :n] % pool.k_buffer[0].shape[0])
        s_seq.copy_(torch.tensor(lens))
        loc = torch.stack([req_pool.req_to_token[b + 1, lens[b] - 1] for b in range(B)]).long()
        s_loc.copy_(loc)
        s_q.copy_(sc2.q)
        s_k.copy_(sc2.k)
        s_v.copy_(sc2.v)
        backend.init_forward_metadata_out_graph(fb)
        graph.replay()
        torch.cuda.synchronize()
        kc, vc = pool.k_buffer[0], pool.v_buffer[0]
        assert torch.equal(kc[loc].cpu(), fp8_port.quantize(s_k.cpu().view(B, Hkv, D), ks))  # captured quantising store
        ref = oracle.attention_fp32(s_q, fp8_port.dequantize(kc, ks), fp8_port.dequantize(vc, vs),
                                    req_pool.req_to_token, s_req, torch.tensor(lens),
                                    torch.ones(B, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False)
        assert_close_attn(s_out, ref, f"fp8 graph replay seed {seed}")

    check(lens0, 5)
    check([1, 900, 17, 16, 333, 640, 2, 45], 6)


def test_rotary_with_fp8_pool_store():
    """RotaryEmbedding(..., fused_set_kv_buffer_arg=) over an e4m3 pool: rotate, then the quantising store."""
    from sglang_b200.rotary import FusedSetKVBufferArg, RotaryEmbedding

    T, Hq, Hkv, D = 50, 8, 2, 128
    g = torch.Generator().manual_seed(4)
    q = torch.randn((T, Hq * D), generator=g).to(torch.bfloat16)
    k = torch.randn((T, Hkv * D), generator=g).to(torch.bfloat16)
    v = torch.randn((T, Hkv * D), generator=g).to(torch.bfloat16)
    pos = torch.randint(0, 4096, (T,), generator=g)
    loc = torch.randperm(200, generator=g)[:T] + 1
    rope = RotaryEmbedding(D, D, 4096, 10000.0, True, torch.bfloat16, device=DEV, native_rounding=True)
    k8 = torch.zeros((256, Hkv, D), dtype=FP8, device=DEV)
    v8 = torch.zeros_like(k8)
    qg, kg = q.to(DEV), k.to(DEV)
    rope(pos.to(DEV), qg, kg, fused_set_kv_buffer_arg=FusedSetKVBufferArg(
        value=v.to(DEV), k_buffer=k8, v_buffer=v8, cache_loc=loc.to(DEV), k_scale=0.5, v_scale=None))
    q_ref, k_ref = q.clone(), k.clone()
    rope2 = RotaryEmbedding(D, D, 4096, 10000.0, True, torch.bfloat16, device=DEV, native_rounding=True)
    rope2(pos.to(DEV), (q_g := q_ref.to(DEV)), (k_g := k_ref.to(DEV)))
    assert torch.equal(qg, q_g) and torch.equal(kg, k_g)
    kc = torch.zeros((256, Hkv, D), dtype=torch.uint8)
    vc = torch.zeros_like(kc)
    fp8_port.store_kv_fp8(kc, vc, loc, kg.cpu().view(T, Hkv, D), v.view(T, Hkv, D), 0.5, None)
    assert torch.equal(k8.view(torch.uint8).cpu(), kc) and torch.equal(v8.view(torch.uint8).cpu(), vc)


@pytest.mark.parametrize("T,Hq,Hkv,D,rot,neox,dtype,native,mrope", [
    (3, 8, 2, 128, 128, True, torch.bfloat16, False, False),      # decode-sized batch, 1 token per thread
    (700, 32, 8, 128, 128, True, torch.bfloat16, True, False),    # 2 tokens per thread
    (50, 4, 4, 64, 32, False, torch.float16, False, False),       # gpt-j pairs, partial rotary (pass-through columns)
    (40, 14, 2, 64, 64, True, torch.bfloat16, False, False),      # Qwen2-0.5B geometry
    (33, 28, 4, 128, 128, True, torch.bfloat16, True, True),      # M-RoPE [16, 24, 24]
    (20, 2, 1, 100, 20, True, torch.bfloat16, False, False),      # narrow vectors (4-element alignment)
])
def test_rope_store_fp8_fused_equals_rotate_then_quantising_store(T, Hq, Hkv, D, rot, neox, dtype, native, mrope):
    """b200attn_rope_store_fp8 (one launch) against rotate-only + the store oracle: q/k bit-equal, pool bytes equal."""
    from sglang_b200.rotary import compute_cos_sin_cache

    g = torch.Generator().manual_seed(T + D)
    q = (torch.randn((T, Hq * D), generator=g) * 2).to(dtype)
    k = (torch.randn((T, Hkv * D), generator=g) * 2).to(dtype)
    v = (torch.randn((T, Hkv * D), generator=g) * 2).to(dtype)
    cache = compute_cos_sin_cache(rot, 4096, 10000.0).to(DEV)
    if mrope:
        pos = torch.randint(0, 4096, (3, T), generator=g)
        sec = [16, 24, 24]
    else:
        pos = torch.randint(0, 4096, (T,), generator=g)
        sec = None
    loc = torch.randperm(1000, generator=g)[:T]
    loc[0] = 0                                           # reserved slot: skipped
    k8 = torch.full((1024, Hkv, D), 3, dtype=torch.uint8, device=DEV)
    v8 = torch.full((1024, Hkv, D), 5, dtype=torch.uint8, device=DEV)
    kw = dict(head_size=D, is_neox=neox, rope_dim=rot, mrope_section=sec, native_rounding=native)
    q1, k1 = q.to(DEV), k.to(DEV)
    ops.apply_rope_with_cos_sin_cache_inplace(pos.to(DEV), q1, k1, cache, v=v.to(DEV), k_buffer=k8.view(FP8),
                                              v_buffer=v8.view(FP8), cache_loc=loc.to(DEV), k_scale=0.7,
                                              v_scale=None, **kw)
    q2, k2 = q.to(DEV), k.to(DEV)
    ops.apply_rope_with_cos_sin_cache_inplace(pos.to(DEV), q2, k2, cache, **kw)
    assert torch.equal(q1, q2) and torch.equal(k1, k2)
    kc = torch.full((1024, Hkv, D), 3, dtype=torch.uint8)
    vc = torch.full((1024, Hkv, D), 5, dtype=torch.uint8)
    fp8_port.store_kv_fp8(kc, vc, loc, k2.cpu().view(T, Hkv, D), v.view(T, Hkv, D), 0.7, None, skip_index=0)
    assert torch.equal(k8.cpu(), kc), "fused rope + e4m3 store: K pool bytes differ"
    assert torch.equal(v8.cpu(), vc), "fused rope + e4m3 store: V pool bytes differ"


@pytest.mark.parametrize("fp8_extend", ["in_kernel", "dequant"])
def test_backend_fp8_cuda_graph_extend_replay(fp8_extend):
    """Full-graph prefill over an e4m3 pool: ONE captured step (quantising store, [device-bounded dequantising gather,]
    extend) replayed over other ragged batches."""
    Hq, Hkv, D, page, B, TCAP = 8, 2, 128, 16, 4, 1024
    ks, vs = 0.8, 1.2
    pre0, ext0 = [64, 0, 300, 17], [200, 128, 40, 333]
    sc, k8, v8, _, _ = _fp8_scenario([TCAP] * B, [TCAP] * B, Hq, Hkv, D, page, seed=91, ks=ks, vs=vs)
    backend, pool = _backend(sc, Hkv, D, page, k8, v8, fp8_extend=fp8_extend)
    req_pool = backend.req_to_token_pool
    backend.init_cuda_graph_state(B, TCAP, num_q_heads=Hq, head_dim=D)
    layer = RadixAttention(Hq, D, sc.scaling, Hkv, 0)
    layer.k_scale, layer.v_scale, layer.k_scale_float, layer.v_scale_float = torch.tensor(ks), torch.tensor(vs), ks, vs
    s_req = sc.req_pool_indices.to(DEV)
    s_seq = torch.zeros(B, dtype=torch.int64, device=DEV)
    s_pre = torch.zeros(B, dtype=torch.int64, device=DEV)
    s_ext = torch.zeros(B, dtype=torch.int32, device=DEV)
    s_loc = torch.zeros(TCAP, dtype=torch.int64, device=DEV)
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
        req_pool.req_to_token[:, :n].copy_(sc2.req_to_token[:, :n] % pool.k_buffer[0].shape[0])
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
        layer(s_q, s_k, s_v, fb)      # eager warm-up: allocates the 16-bit operand scratch
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            backend.init_forward_metadata_in_graph(fb)
            s_out = layer(s_q, s_k, s_v, fb)
    torch.cuda.current_stream().wait_stream(stream)

    for pre, ext, seed in [(pre0, ext0, 92), ([0, 5, 0, 0], [700, 1, 0, 0], 93), ([511, 0, 40, 0], [1, 0, 129, 600], 94)]:
        T, loc = fill(pre, ext, seed)
        backend.prepare_forward_metadata_for_breakable_cuda_graph_replay(md0, fb)
        graph.replay()
        torch.cuda.synchronize()
        kc, vc = pool.k_buffer[0], pool.v_buffer[0]
        assert torch.equal(kc[loc].cpu(), fp8_port.quantize(s_k[:T].cpu(), ks))
        view = types.SimpleNamespace(req_to_token=req_pool.req_to_token.cpu(), req_pool_indices=s_req.cpu())
        _check_fp8_extend_rows(s_out[:T], s_q[:T].cpu(), kc.cpu(), vc.cpu(), view, s_seq.cpu().tolist(),
                               s_ext.cpu().tolist(), Hq, Hkv, D, ks, vs, fp8_extend, f"fp8 extend graph replay {ext}",
                               single_rows_decode=False)  # no host lengths under the graph: every row takes the extend kernel


def test_dequant_gather_unique_shares_rows():
    """Every referenced slot is converted once, into the row of its first reference; out_indices names that row."""
    g = torch.Generator().manual_seed(0)
    k8 = torch.randint(0, 256, (300, 4, 128), dtype=torch.uint8, generator=g)
    k8[(k8 & 0x7F) == 0x7F] = 0
    v8 = k8.flip(0).contiguous()
    idx = torch.randint(0, 300, (1000,), generator=g).int()           # many repeated slots
    ok = torch.full((1000, 4, 128), float("nan"), dtype=torch.bfloat16, device=DEV)
    ov = torch.full_like(ok, float("nan"))
    first = torch.empty(300, dtype=torch.int32, device=DEV)
    out_idx = torch.full((1000,), -1, dtype=torch.int32, device=DEV)
    for _ in range(2):   # the call resets its own first-reference map
        ops.kv_dequant_gather_unique(k8.to(DEV), v8.to(DEV).view(FP8), idx.to(DEV), 1000, ok, ov, first, out_idx,
                                     1.0, 0.75)
    torch.cuda.synchronize()
    want_first = {}
    for i, s_ in enumerate(idx.tolist()):
        want_first.setdefault(s_, i)
    want_idx = torch.tensor([want_first[s_] for s_ in idx.tolist()], dtype=torch.int32)
    assert torch.equal(out_idx.cpu(), want_idx)
    rows = out_idx.long()
    assert torch.equal(ok[rows].cpu().float(), fp8_port.dequantize(k8)[idx.long()])
    assert torch.equal(ov[rows].cpu(), (fp8_port.dequantize(v8)[idx.long()] * 0.75).to(torch.bfloat16))
    owners = torch.zeros(1000, dtype=torch.bool)
    owners[want_idx.long()] = True
    assert torch.isnan(ok.cpu()[~owners].float()).all()               # non-owner rows were not written
    # device-side bound: only the first 10 rows are live
    out_idx.fill_(-1)
    ops.kv_dequant_gather_unique(k8.to(DEV), v8.to(DEV), idx.to(DEV), 1000, ok, ov, first, out_idx, 1.0, 0.75,
                                 n_dev=torch.tensor([10], dtype=torch.int32, device=DEV))
    assert (out_idx[10:] == -1).all() and (out_idx[:10] >= 0).all()


@pytest.mark.parametrize("Hq,Hkv,D,page", [(32, 8, 128, 16), (14, 2, 64, 1), (64, 2, 128, 16)])
@pytest.mark.parametrize("lens", [[1, 2, 15, 16, 17, 63, 64, 65, 200, 777, 1500], [5000], [300] * 37])
def test_decode_sk_fp8_matches_oracle(Hq, Hkv, D, page, lens):
    """Flat tile-stream scheduling over the e4m3 pool (b200attn_decode_sk_fp8kv)."""
    B = len(lens)
    sc, k8, v8, kf, vf = _fp8_scenario(lens, [0] * B, Hq, Hkv, D, page, seed=Hq + page + B)
    q = torch.randn((B, Hq * D), generator=torch.Generator().manual_seed(1)).to(torch.bfloat16)
    ref, lse_ref = oracle.attention_fp32(q, kf, vf, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens),
                                         torch.ones(B, dtype=torch.int64), Hq, Hkv, D, sc.scaling, causal=False,
                                         return_lse=True)
    d = sc.to(DEV)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, torch.tensor(lens, device=DEV), kv_indptr, kv_indices)
    o = torch.full((B, Hq, D), float("nan"), dtype=torch.bfloat16, device=DEV)
    lse = torch.empty((B, Hq), dtype=torch.float32, device=DEV)
    ws = torch.zeros(ops.decode_sk_workspace_bytes(B, Hq, Hkv, D), dtype=torch.uint8, device=DEV)
    for _ in range(2):
        ops.decode_attention(q.to(DEV).view(B, Hq, D), k8.to(DEV).view(FP8), v8.to(DEV), o, kv_indptr, kv_indices,
                             sc.scaling, min_chunk=64, workspace=ws, lse=lse, k_descale=0.6, v_descale=0.9,
                             stream_k=True)
    torch.cuda.synchronize()
    assert_close_attn(o.view(B, -1), ref, f"fp8 decode_sk Hq{Hq} D{D} page{page} B{B}")
    torch.testing.assert_close(lse.cpu(), lse_ref, rtol=2e-3, atol=2e-3)
