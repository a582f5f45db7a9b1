"""Dev tool (GPU): extend kernel time vs (items, 64-key chunk-tiles) to separate per-item from per-chunk cost."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from sglang_b200 import ops

dev = "cuda:0"
page = 16


def run(B, P, E, Hq, Hkv, D=128, causal=True):
    ctx = P + E
    n_pages = B * (-(-ctx // page)) + 8
    perm = torch.randperm(n_pages - 1) + 1
    ppr = -(-ctx // page)
    table = perm[: B * ppr].view(B, ppr)
    slots = (table[:, :, None] * page + torch.arange(page)).reshape(B, -1)[:, :ctx]
    kv_indices = slots.reshape(-1).int().to(dev)
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=dev)
    qo_indptr = torch.arange(0, (B + 1) * E, E, dtype=torch.int32, device=dev)
    kc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
    q = torch.randn((B * E, Hq, D), device=dev).bfloat16()
    o = torch.empty_like(q)
    f = lambda: ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=causal,
                                     page_size=page, max_qo_len=E)
    t = bench.time_steps(f, 20, 5) / 20
    G = Hq // Hkv
    qt = -(-E // 128)
    items = B * Hkv * ((qt * G + 1) // 2)
    chunks = 0
    for j in range(qt):
        kvend = min(ctx, P + (j + 1) * 128) if causal else ctx
        chunks += G * (-(-kvend // 64))
    chunks *= B * Hkv
    flops = 4 * D * Hq * B * ((E * P + E * (E + 1) // 2) if causal else E * ctx)
    print(f"B={B:3d} P={P:5d} E={E:5d} Hq={Hq} Hkv={Hkv} causal={int(causal)}  {t*1e3:8.1f} us  items={items:5d} "
          f"chunk_tiles={chunks:7d}  per_cta: items={items/148:6.1f} chunks={chunks/148:7.1f}  "
          f"{flops/t/1e9:7.1f} TF  us/chunk_tile_per_cta={t*1e3/(chunks/148):.3f}", flush=True)


run(16, 0, 1280, 28, 4)
run(16, 4096, 1280, 28, 4)
run(16, 0, 1280, 28, 4, causal=False)
run(16, 0, 1280, 32, 8)
run(32, 2048, 512, 32, 8)
run(32, 0, 512, 32, 8)
run(64, 0, 256, 32, 8)
run(128, 0, 128, 32, 8)
run(1, 8192, 8192, 64, 8)
run(4, 0, 4096, 32, 8)
run(1, 8192, 8192, 8, 1)
run(1, 8192, 8192, 16, 2)
