"""fp8-KV decode kernel: time cfg2 / cfg5-shard shapes (run on the GPU box): python tools/sweep_decode_fp8.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from sglang_b200 import _lib, ops  # noqa: E402

if os.environ.get("B200ATTN_LIB"):  # dev only: a variant build (sglang_b200/variants/, -DB200ATTN_FP8_STAGES/_CTAS)
    _lib.LIB_PATH = os.path.abspath(os.environ["B200ATTN_LIB"])
    print("lib:", _lib.LIB_PATH)

dev = "cuda:0"
gen = torch.Generator().manual_seed(1)
peaks = bench.measured_peaks()
print(bench.run_cfg2_decode_fp8(dev, gen, peaks))
for B, ctx in ((1, 8192), (8, 8192), (32, 2048), (64, 4096), (111, 4096), (256, 8192)):
    c = dict(B=B, ctx=ctx, Hq=32, Hkv=8, D=128, page=16)
    req_pool, kv_pool, host = bench.build_decode_state(c, dev, gen)
    k8, v8 = (b.to(torch.float8_e4m3fn) for b in kv_pool.get_kv_buffer(0))
    kbuf, vbuf = kv_pool.get_kv_buffer(0)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    kv_indices = torch.zeros(B * ctx, dtype=torch.int32, device=dev)
    ops.build_kv_indices(req_pool.req_to_token, host["req_pool_indices"].to(dev), host["seq_lens"].to(dev), kv_indptr,
                         kv_indices)
    q = host["q"].to(dev).view(B, 32, 128)
    o = torch.empty_like(q)
    ms = ops.pick_max_splits(B, 8, 4, ctx, 256)
    ms8 = ops.pick_max_splits(B, 8, 4, ctx, 256, ctas_per_sm=3)
    tb8 = -1
    if ms8 == 1:
        tb8, k = ops.pick_tail_split(B, 8, 4, ctas_per_sm=3)
        ms8 = max(k, 1)
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, 32, 128, max(ms, ms8)), 16), dtype=torch.uint8, device=dev)
    t8 = bench.time_steps(lambda: ops.decode_attention(q, k8, v8, o, kv_indptr, kv_indices, 128 ** -0.5, max_splits=ms8,
                                                       workspace=ws, tail_begin=tb8, k_descale=0.5, v_descale=0.5),
                          20, 5) / 20
    ws3 = torch.zeros(ops.decode_sk_workspace_bytes(B, 32, 8, 128), dtype=torch.uint8, device=dev)
    tsk = bench.time_steps(lambda: ops.decode_attention(q, k8, v8, o, kv_indptr, kv_indices, 128 ** -0.5, min_chunk=256,
                                                        workspace=ws3, k_descale=0.5, v_descale=0.5, stream_k=True),
                           20, 5) / 20
    t16 = bench.time_steps(lambda: ops.decode_attention(q, kbuf, vbuf, o, kv_indptr, kv_indices, 128 ** -0.5,
                                                        max_splits=ms, workspace=ws), 20, 5) / 20
    kvb = 2 * B * ctx * 8 * 128
    print(f"B={B} ctx={ctx} splits={ms}/{ms8}: fp8 stream-k {tsk * 1e3:.1f} us ({kvb / tsk / 1e6:.0f} GB/s) | fp8 {t8 * 1e3:.1f} us ({kvb / t8 / 1e6:.0f} GB/s)  bf16 {t16 * 1e3:.1f} us "
          f"({2 * kvb / t16 / 1e6:.0f} GB/s)  speedup {t16 / t8:.2f}x")
    del req_pool, kv_pool, k8, v8, kbuf, vbuf
    torch.cuda.empty_cache()
