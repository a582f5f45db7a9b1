"""Dev (GPU): time the extend kernel of every build under sglang_b200/variants/ (+ the product build) on cfg3
(radix-hit layout) and cfg4 (16 x 1280, G = 7, causal)."""
import glob
import json
import os
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import _lib, ops  # noqa: E402
from tools.compare_flashinfer import timed  # noqa: E402

DEV = "cuda"
page = 16


def make(B, P, E, Hq, Hkv, D=128, shared=True):
    ctx = P + E
    sp = P // page if shared else 0
    own = -(-ctx // page) - sp
    n_pages = sp + B * own + 8
    perm = (torch.randperm(n_pages - 1) + 1)
    sh = perm[:sp]
    rows = []
    for b in range(B):
        pages = torch.cat([sh, perm[sp + b * own: sp + (b + 1) * own]])
        rows.append((pages[:, None] * page + torch.arange(page)).reshape(-1)[:ctx])
    kv_indices = torch.cat(rows).int().to(DEV)
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    qo_indptr = torch.arange(0, (B + 1) * E, E, dtype=torch.int32, device=DEV)
    kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    q = torch.randn((B * E, Hq, D), device=DEV).bfloat16()
    o = torch.empty_like(q)
    flops = 4 * D * Hq * B * (E * P + E * (E + 1) // 2)
    f = lambda: ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True,
                                     page_size=page, max_qo_len=E)
    return f, flops, o


def main():
    product = _lib.load()
    libs = [("product", None)] + [(os.path.basename(p)[len("libb200attn_"):-3], p)
                                  for p in sorted(glob.glob("sglang_b200/variants/libb200attn_*.so"))]
    cases = {"cfg3": make(32, 2048, 512, 32, 8), "cfg4": make(16, 0, 1280, 28, 4, shared=False),
             "cfg5s": make(1, 8192, 8192, 8, 1, shared=False)}
    ref = {}
    for name, path in libs:
        _lib._lib = product if path is None else _lib.load_variant(path)
        res = dict(variant=name)
        for cname, (f, flops, o) in cases.items():
            t = min(timed(f, n=20, warm=5) for _ in range(3))
            res[cname + "_us"] = round(t * 1e3, 1)
            res[cname + "_tf"] = round(flops / t / 1e9, 1)
            if name == "product":
                ref[cname] = o.clone()
            else:
                res[cname + "_maxdiff"] = float((o.float() - ref[cname].float()).abs().max())
        print(json.dumps(res), flush=True)
    _lib._lib = product


if __name__ == "__main__":
    main()
