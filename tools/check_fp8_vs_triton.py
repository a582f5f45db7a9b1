"""Dev (GPU): the two e4m3-pool extend paths of the backend against the REFERENCE Triton kernel's recorded output
(tests/golden/triton_cases.pt, extend_fp8) and against the exact fp32 math over the dequantised pool."""
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool, RadixAttention, ReqToTokenPool
from tests.test_triton_golden import CASES, _inputs

dev = "cuda"
for c in CASES:
    if c["kind"] != "extend" or c["k_scale"] == 1.0:
        continue
    sc, q, kc, vc, k8, v8, seq, ext = _inputs(c)
    Hq, Hkv, D, page = c["Hq"], c["Hkv"], c["D"], c["page_size"]
    exact = oracle.attention_fp32(q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(seq), torch.tensor(ext),
                                  Hq, Hkv, D, c["scaling"], causal=True)
    ref = c["o_ref"].float().view(exact.shape)
    print(f"case {c['name']}: D={D} Hq={Hq} Hkv={Hkv} | reference Triton vs exact: max {float((ref - exact).abs().max()):.4f} "
          f"mean {float((ref - exact).abs().mean()):.5f}")
    for mode in ("in_kernel", "dequant"):
        req_pool = ReqToTokenPool(sc.req_to_token.shape[0] - 1, sc.req_to_token.shape[1], dev)
        req_pool.req_to_token.copy_(sc.req_to_token)
        kv_pool = MHATokenToKVPool(k8.shape[0] - page, page, torch.float8_e4m3fn, Hkv, D, 1, dev)
        kv_pool.k_buffer[0].copy_(k8.view(torch.uint8))
        kv_pool.v_buffer[0].copy_(v8.view(torch.uint8))
        runner = types.SimpleNamespace(device=dev, page_size=page, req_to_token_pool=req_pool, token_to_kv_pool=kv_pool,
                                       sliding_window_size=None)
        backend = B200AttnBackend(runner, fp8_extend=mode)
        layer = RadixAttention(Hq, D, c["scaling"], Hkv, 0, logit_cap=c["logit_cap"])
        layer.k_scale, layer.v_scale = c["k_scale"], c["v_scale"]
        layer.k_scale_float, layer.v_scale_float = c["k_scale"], c["v_scale"]
        B = len(seq)
        fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=B, req_pool_indices=sc.req_pool_indices.to(dev),
                          seq_lens=torch.tensor(seq, device=dev),
                          out_cache_loc=torch.zeros(sum(ext), dtype=torch.int64, device=dev), seq_lens_sum=sum(seq),
                          seq_lens_cpu=torch.tensor(seq), attn_backend=backend)
        fb.extend_seq_lens = torch.tensor(ext, dtype=torch.int32, device=dev)
        fb.extend_prefix_lens = torch.tensor(c["prefix_lens"], dtype=torch.int32, device=dev)
        fb.extend_seq_lens_cpu, fb.extend_prefix_lens_cpu = list(ext), list(c["prefix_lens"])
        backend.init_forward_metadata(fb)
        o = layer(q.to(dev), None, None, fb, save_kv_cache=False).float().cpu().view(exact.shape)
        print(f"  {mode:9s}: vs reference Triton max {float((o - ref).abs().max()):.4f} mean {float((o - ref).abs().mean()):.5f}"
              f" | vs exact max {float((o - exact).abs().max()):.4f} mean {float((o - exact).abs().mean()):.5f}")
