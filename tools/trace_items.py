"""Dev tool (GPU): item-level timeline of CTA 0 of the extend kernel (EXT_TRACE build: make -C sglang_b200/csrc trace).
WG0: [item top, first S seen, last P arrived, OFinal seen, epilogue done]; MMA0: [item top, Q seen, K0 seen, last PV issued]."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sglang_b200 import _lib

_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libb200attn_trace.so")
from sglang_b200 import ops

dev = "cuda:0"
page = 16
B, P, E, Hq, Hkv, D = (int(x) for x in (sys.argv[1:7] if len(sys.argv) > 6 else (64, 0, 256, 32, 8, 128)))
ctx = P + E
ppr = -(-ctx // page)
n_pages = B * ppr + 8
perm = torch.randperm(n_pages - 1) + 1
table = perm[: B * ppr].view(B, ppr)
slots = (table[:, :, None] * page + torch.arange(page)).reshape(B, -1)[:, :ctx]
kv_indices = slots.reshape(-1).int().to(dev)
kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=dev)
qo_indptr = torch.arange(0, (B + 1) * E, E, dtype=torch.int32, device=dev)
kc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
vc = torch.randn((n_pages * page, Hkv, D), device=dev).bfloat16()
q = torch.randn((B * E, Hq, D), device=dev).bfloat16()
o = torch.empty_like(q)
ws = torch.zeros(6 * 32 * 8 + 2 * 160, dtype=torch.int64, device=dev)
for _ in range(10):
    ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True, page_size=page,
                         max_qo_len=E, workspace=ws)
torch.cuda.synchronize()
t = ws.cpu()[:6 * 32 * 8].view(6, 32, 8)
c0, g0, c1, g1 = (int(t[3, 0, k]) for k in range(4))
t0 = c0
print(f"B={B} P={P} E={E}: CTA 0 lifetime {c1 - c0} cycles, {g1 - g0} ns, clock {(c1 - c0) / max(g1 - g0, 1) * 1e3:.0f} MHz")
for it in range(16):
    w = [int(t[4, it, k]) - t0 for k in range(5)]
    m = [int(t[5, it, k]) - t0 for k in range(4)]
    if w[0] < 0:
        break
    print(f"item {it:2d} | WG0 top {w[0]:7d} S0 {w[1]:7d} lastP {w[2]:7d} OFinal {w[3]:7d} epi_done {w[4]:7d} "
          f"| MMA0 top {m[0]:7d} Q {m[1]:7d} K0 {m[2]:7d} lastPV {m[3]:7d}")
