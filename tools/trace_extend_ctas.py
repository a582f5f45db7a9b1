"""Dev (GPU): start / end wall time of EVERY CTA of the persistent extend kernel (EXT_TRACE build), cfg3 / cfg4 /
cfg5 shard: how evenly do the 148 CTAs finish?"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sglang_b200 import _lib

_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libb200attn_trace.so")
from sglang_b200 import ops  # noqa: E402

dev = "cuda:0"
page = 16


def run(name, B, P, E, Hq, Hkv, D=128, shared=False):
    ctx = P + E
    sp = P // page if shared else 0
    own = -(-ctx // page) - sp
    n_pages = sp + B * own + 8
    perm = torch.randperm(n_pages - 1) + 1
    rows = [(torch.cat([perm[:sp], perm[sp + b * own: sp + (b + 1) * own]])[:, None] * page + torch.arange(page)).reshape(-1)[:ctx]
            for b in range(B)]
    kv_indices = torch.cat(rows).int().to(dev)
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=dev)
    qo_indptr = torch.arange(0, (B + 1) * E, E, dtype=torch.int32, device=dev)
    kc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
    q = torch.randn((B * E, Hq, D), device=dev).bfloat16()
    o = torch.empty_like(q)
    ws = torch.zeros(6 * 32 * 8 + 2 * 160, dtype=torch.int64, device=dev)
    for _ in range(6):
        ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True, page_size=page,
                             max_qo_len=E, workspace=ws)
    torch.cuda.synchronize()
    t = ws.cpu()[6 * 32 * 8:].view(160, 2)[:148].double()
    t0 = t[:, 0].min()
    st, en = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
    qs = torch.tensor([0.0, 0.1, 0.25, 0.5, 0.75, 0.9, 1.0], dtype=torch.float64)
    print(f"{name}: start max {float(st.max()):.1f} us | end quantiles " +
          " ".join(f"{float(x):.1f}" for x in torch.quantile(en, qs)) + f" | mean {float(en.mean()):.1f} us "
          f"(idle tail {100 * (1 - float(en.mean()) / float(en.max())):.1f} %)")


run("cfg3", 32, 2048, 512, 32, 8, shared=True)
run("cfg4", 16, 0, 1280, 28, 4)
run("cfg5s", 1, 8192, 8192, 8, 1)
run("cfg5", 1, 8192, 8192, 64, 8)
