"""Dev (GPU): decode time over warps-per-CTA x max_splits for a few batch shapes (Llama-3.1-8B heads), next to the
plan ops.plan_decode picks.  usage: python tools/sweep_decode_grid.py [BxCTX ...]"""
import json
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import ops  # noqa: E402
from tools.compare_flashinfer import timed  # noqa: E402

DEV = "cuda"
Hq, Hkv, D, page = 32, 8, 128, 16
shapes = [(64, 4096), (16, 4096), (16, 16384), (32, 8192), (4, 8192), (4, 32768), (1, 32768), (128, 2048), (256, 8192)]
if len(sys.argv) > 1:
    shapes = [tuple(int(x) for x in a.split("x")) for a in sys.argv[1:]]
for B, ctx in shapes:
    n_pages = B * ctx // page + 8
    perm = torch.randperm(n_pages - 1) + 1
    table = perm[: B * ctx // page].view(B, ctx // page).int().to(DEV)
    kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    q = torch.randn((B, Hq, D), device=DEV).bfloat16()
    o = torch.empty_like(q)
    res = dict(B=B, ctx=ctx, units=B * Hkv, plan=list(ops.plan_decode(B, Hkv, Hq // Hkv, ctx, 256)))
    for w in (2, 4):
        for ms in (1, 2, 3, 4, 6, 8, 12, 16, 24, 32):
            if ctx // ms < 256 or B * Hkv * ms > 16384:
                continue
            ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
            t = min(timed(lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5, max_splits=ms,
                                                       min_chunk=256, workspace=ws, warps=w, tail_begin=-1))
                    for _ in range(2))
            res[f"w{w}_ms{ms}"] = round(t * 1e3, 1)
    best = min((v, k) for k, v in res.items() if k.startswith("w"))
    res["best"] = best[1]
    print(json.dumps(res), flush=True)
    del kc, vc
