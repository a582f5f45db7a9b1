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
    """Variants are timed in ROTATION (several rounds, order reversed every other round): the kernel is power capped,
    so whatever runs later in a process sees a lower clock; only per-variant medians over the rounds mean anything."""
    product = _lib.load()
    libs = [("product", product)] + [(os.path.basename(p)[len("libb200attn_"):-3], _lib.load_variant(p))
                                     for p in sorted(glob.glob("sglang_b200/variants/libb200attn_*.so"))]
    cases = {"cfg3": make(32, 2048, 512, 32, 8), "cfg4": make(16, 0, 1280, 28, 4, shared=False),
             "cfg5s": make(1, 8192, 8192, 8, 1, shared=False)}
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    times = {name: {c: [] for c in cases} for name, _ in libs}
    ref = {}
    for r in range(rounds):
        for name, lib in (libs if r % 2 == 0 else libs[::-1]):
            _lib._lib = lib
            for cname, (f, flops, o) in cases.items():
                times[name][cname].append(timed(f, n=20, warm=5))
                if r == 0:
                    if name == "product":
                        ref[cname] = o.clone()
                    else:
                        assert torch.equal(o, ref[cname]) or (o.float() - ref[cname].float()).abs().max() < 2e-2
    _lib._lib = product
    for name, _ in libs:
        res = dict(variant=name)
        for cname, (f, flops, o) in cases.items():
            ts = sorted(times[name][cname])
            med = ts[len(ts) // 2]
            res[cname + "_us"] = round(med * 1e3, 1)
            res[cname + "_tf"] = round(flops / med / 1e9, 1)
            res[cname + "_all_us"] = [round(t * 1e3, 1) for t in times[name][cname]]
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
