"""Dev (GPU): does the NHD pool layout (a kv head's 256-byte row every 2 KB) cost DRAM efficiency?  Same bytes, same
512 units, same kernel: (a) cfg2 (B=64, 8 kv heads interleaved per slot) vs (b) B=512 requests over a 1-kv-head pool
(rows contiguous) vs (c) cfg2 with contiguous (unshuffled) pages."""
import json
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import ops  # noqa: E402
from tools.compare_flashinfer import timed  # noqa: E402

DEV = "cuda"
page, D, ctx = 16, 128, 4096


def run(name, B, Hq, Hkv, shuffle=True):
    n_pages = B * ctx // page + 8
    perm = (torch.randperm(n_pages - 1) + 1) if shuffle else (torch.arange(n_pages - 1) + 1)
    table = perm[: B * ctx // page].view(B, ctx // page).int().to(DEV)
    kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    q = torch.randn((B, Hq, D), device=DEV).bfloat16()
    o = torch.empty_like(q)
    warps, ms, tb = ops.plan_decode(B, Hkv, Hq // Hkv, ctx, 256)
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
    t = min(timed(lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5, max_splits=ms,
                                               workspace=ws, warps=warps, tail_begin=tb), n=30, warm=5)
            for _ in range(3))
    nbytes = 2 * B * ctx * Hkv * D * 2 + kv_indices.numel() * 4
    print(json.dumps(dict(case=name, B=B, Hkv=Hkv, plan=[warps, ms, tb], us=round(t * 1e3, 1),
                          gbs=round(nbytes / t / 1e6, 0))), flush=True)
    del kc, vc


import glob
import os

from sglang_b200 import _lib  # noqa: E402

product = _lib.load()
libs = [("product", product)] + [(os.path.basename(p_)[len("libb200attn_"):-3], _lib.load_variant(p_))
                                 for p_ in sorted(glob.glob("sglang_b200/variants/libb200attn_*.so"))]
for lname, lib in libs:
    _lib._lib = lib
    print("== library:", lname, flush=True)
    run("cfg2 NHD 8 kv heads, shuffled pages", 64, 32, 8)
    run("cfg2 NHD 8 kv heads, contiguous pages", 64, 32, 8, shuffle=False)
    run("16 req (split), shuffled pages", 16, 32, 8)
    run("256 req, shuffled pages", 256, 32, 8)
_lib._lib = product
