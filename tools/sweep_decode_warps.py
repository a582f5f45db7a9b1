"""Dev: 2-warp (four CTAs per SM, one wave) vs 4-warp (two CTAs per SM, split cap / tail split) decode plans
around the 297..592-unit band, ctx 4096, Llama-3.1-8B heads (GPU box)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import ops  # noqa: E402
from tools.compare_flashinfer import timed  # noqa: E402

DEV = "cuda"
Hq, Hkv, D, page = 32, 8, 128, 16


def main():
    for B, ctx in [(36, 4096), (38, 4096), (40, 4096), (44, 4096), (48, 4096), (56, 4096), (64, 4096), (72, 4096),
                   (74, 4096), (76, 4096), (80, 4096), (96, 4096), (64, 1024), (64, 16384)]:
        n_pages = B * ctx // page + 8
        perm = torch.randperm(n_pages - 1) + 1
        table = perm[: B * ctx // page].view(B, ctx // page).int().to(DEV)
        kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
        kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
        kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
        vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
        q = torch.randn((B, Hq, D), device=DEV).bfloat16()
        o = torch.empty_like(q)
        res = dict(B=B, ctx=ctx, units=B * Hkv)
        ms4 = ops.pick_max_splits(B, Hkv, Hq // Hkv, ctx, 256, ctas_per_sm=2)
        tb, k = (-1, 1)
        if ms4 == 1:
            tb, k = ops.pick_tail_split(B, Hkv, Hq // Hkv, ctas_per_sm=2)
        plans = {"w4": (4, max(ms4, k), tb if k > 1 else -1), "w2_ms1": (2, 1, -1)}
        ms2 = ops.pick_max_splits(B, Hkv, Hq // Hkv, ctx, 256, ctas_per_sm=4)
        if ms2 > 1:
            plans["w2_split"] = (2, ms2, -1)
        tb2, k2 = ops.pick_tail_split(B, Hkv, Hq // Hkv, ctas_per_sm=4)
        if k2 > 1:
            plans["w2_tail"] = (2, k2, tb2)
        for name, (w, ms, t_) in plans.items():
            ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
            t = timed(lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5, max_splits=ms,
                                                   min_chunk=256, workspace=ws, warps=w, tail_begin=t_))
            res[name] = [round(t * 1e3, 1), ms, t_]
        res["plan"] = list(ops.plan_decode(B, Hkv, Hq // Hkv, ctx, 256))
        print(json.dumps(res), flush=True)
        del kc, vc


if __name__ == "__main__":
    main()
