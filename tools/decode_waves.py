"""Dev tool (GPU): decode GB/s vs batch size around the 296-CTA wave boundary (2 CTAs/SM x 148)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from sglang_b200 import ops

dev = "cuda:0"
Hq, Hkv, D, ctx, page = 32, 8, 128, 4096, 16
for B in (37, 64, 74, 111, 148):
    n_pages = B * ctx // page + 8
    perm = torch.randperm(n_pages - 1) + 1
    table = perm[: B * ctx // page].view(B, ctx // page)
    kv_indices = (table[:, :, None] * page + torch.arange(page)).reshape(-1).int().to(dev)
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=dev)
    kc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
    q = torch.randn((B, Hq, D), device=dev).bfloat16()
    o = torch.empty_like(q)
    f = lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5, max_splits=1)
    t = bench.time_steps(f, 100, 10) / 100
    nbytes = B * ctx * Hkv * D * 4 + 2 * B * Hq * D * 2 + B * ctx * 4
    print(f"B={B:4d} ctas={B*Hkv:5d} waves={B*Hkv/296:5.2f}  {t*1e3:8.1f} us  {nbytes / t / 1e6:8.1f} GB/s", flush=True)
    del kc, vc

print("# tail splitting (only the units of the last partial wave are split)")
for B in (64, 100, 40):
    n_pages = B * ctx // page + 8
    perm = torch.randperm(n_pages - 1) + 1
    table = perm[: B * ctx // page].view(B, ctx // page)
    kv_indices = (table[:, :, None] * page + torch.arange(page)).reshape(-1).int().to(dev)
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=dev)
    kc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
    q = torch.randn((B, Hq, D), device=dev).bfloat16()
    o = torch.empty_like(q)
    nbytes = B * ctx * Hkv * D * 4 + 2 * B * Hq * D * 2 + B * ctx * 4
    for tb, k in [(-1, 1), ops.pick_tail_split(B, Hkv, Hq // Hkv), (296 * (B * Hkv // 296), 2), (296 * (B * Hkv // 296), 3), (296 * (B * Hkv // 296), 4), (296 * (B * Hkv // 296), 6)]:
        ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, k), 16), dtype=torch.uint8, device=dev)
        f = lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5, max_splits=k, min_chunk=64, workspace=ws, tail_begin=tb)
        t = bench.time_steps(f, 100, 10) / 100
        print(f"B={B:4d} tail_begin={tb:5d} k={k}  {t*1e3:8.1f} us  {nbytes / t / 1e6:8.1f} GB/s", flush=True)
    del kc, vc
