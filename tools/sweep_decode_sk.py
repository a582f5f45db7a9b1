"""Split-KV decode vs flat tile-stream decode over batch shapes (run on the GPU box): python tools/sweep_decode_sk.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from sglang_b200 import ops  # noqa: E402

dev = "cuda:0"
gen = torch.Generator().manual_seed(1)
shapes = [(1, 8192), (1, 32768), (4, 8192), (8, 8192), (16, 4096), (16, 16384), (32, 8192), (40, 4096), (64, 4096),
          (111, 4096), (128, 2048), (256, 8192)]
if len(sys.argv) > 1:
    shapes = [tuple(int(x) for x in a.split("x")) for a in sys.argv[1:]]
for B, ctx in shapes:
    c = dict(B=B, ctx=ctx, Hq=32, Hkv=8, D=128, page=16)
    req_pool, kv_pool, host = bench.build_decode_state(c, dev, gen)
    kbuf, vbuf = kv_pool.get_kv_buffer(0)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    kv_indices = torch.zeros(B * ctx, dtype=torch.int32, device=dev)
    ops.build_kv_indices(req_pool.req_to_token, host["req_pool_indices"].to(dev), host["seq_lens"].to(dev), kv_indptr,
                         kv_indices)
    q = host["q"].to(dev).view(B, 32, 128)
    o = torch.empty_like(q)
    o2 = torch.empty_like(q)
    ms = ops.pick_max_splits(B, 8, 4, ctx, 256)
    tb = -1
    if ms == 1:
        tb, k = ops.pick_tail_split(B, 8, 4)
        ms = max(k, 1)
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, 32, 128, ms), 16), dtype=torch.uint8, device=dev)
    ws2 = torch.zeros(ops.decode_sk_workspace_bytes(B, 32, 8, 128), dtype=torch.uint8, device=dev)
    f1 = lambda: ops.decode_attention(q, kbuf, vbuf, o, kv_indptr, kv_indices, 128 ** -0.5, max_splits=ms,
                                      workspace=ws, tail_begin=tb)
    res = {}
    for mc in (256, 512, 1024):
        f2 = lambda: ops.decode_attention(q, kbuf, vbuf, o2, kv_indptr, kv_indices, 128 ** -0.5, min_chunk=mc,
                                          workspace=ws2, stream_k=True)
        res[mc] = bench.time_steps(f2, 30, 5) / 30
    t1 = bench.time_steps(f1, 30, 5) / 30
    err = float((o.float() - o2.float()).abs().max())
    kvb = 4 * B * ctx * 8 * 128
    best = min(res.values())
    print(f"B={B} ctx={ctx}: split(ms={ms},tail={tb}) {t1 * 1e3:.1f} us ({kvb / t1 / 1e6:.0f} GB/s) | stream-k "
          + " ".join(f"mc{m}={t * 1e3:.1f}" for m, t in res.items())
          + f" us ({kvb / best / 1e6:.0f} GB/s) | {t1 / best:.3f}x | max|diff| {err:.2e}", flush=True)
    del req_pool, kv_pool, kbuf, vbuf
    torch.cuda.empty_cache()
