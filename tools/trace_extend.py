"""Dev tool (GPU): per-step timeline of CTA (0,0,0) of the extend kernel on cfg3, from the EXT_TRACE build
(`make -C sglang_b200/csrc trace`).  Prints clock64 deltas per step for softmax WG0/WG1 and the MMA thread."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sglang_b200 import _lib

_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libb200attn_trace.so")
import bench
from sglang_b200 import ops

dev = "cuda:0"
torch.cuda.set_device(0)
gen = torch.Generator().manual_seed(3)
c = bench.CFG3
B, P, E, Hq, Hkv, D, page = c["B"], c["prefix"], c["ext"], c["Hq"], c["Hkv"], c["D"], c["page"]
num_pages = (P + B * E) // page + 8
perm = torch.randperm(num_pages - 1, generator=gen) + 1
table = torch.zeros((B + 1, P + E), dtype=torch.int32)
offs = torch.arange(page)
pre = (perm[:P // page, None] * page + offs[None]).reshape(-1)
own = E // page
for b in range(B):
    mine = perm[P // page + b * own: P // page + (b + 1) * own]
    table[b + 1, :P] = pre
    table[b + 1, P:] = (mine[:, None] * page + offs[None]).reshape(-1)
table = table.to(dev)
kc = torch.randn((num_pages * page, Hkv, D), device=dev).to(torch.bfloat16)
vc = torch.randn((num_pages * page, Hkv, D), device=dev).to(torch.bfloat16)
q = torch.randn((B * E, Hq, D), device=dev).to(torch.bfloat16)
o = torch.empty_like(q)
req = torch.arange(1, B + 1, device=dev)
seq = torch.full((B,), P + E, device=dev)
extl = torch.full((B,), E, dtype=torch.int32, device=dev)
kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
qo_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
kv_indices = torch.zeros(B * (P + E), dtype=torch.int32, device=dev)
ops.build_kv_indices(table, req, seq, kv_indptr, kv_indices, qo_lens=extl, qo_indptr=qo_indptr)
ws = torch.zeros(6 * 32 * 8 + 2 * 160, dtype=torch.int64, device=dev)
for _ in range(30):
    ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True, page_size=page,
                         max_qo_len=E, workspace=ws)
torch.cuda.synchronize()
t = ws.cpu()[:6 * 32 * 8].view(6, 32, 8)
t0 = int(t[2, 0, 0])
print("softmax WG i: [wait_start, S_seen, ld_done, max_done, exp_done, arrived]; "
      "MMA: [w0, P0_seen, issued0, w1, P1_seen, issued1]")
for j in range(20):
    row = []
    for r in range(3):
        row.append(" ".join(f"{int(t[r, j, k]) - t0:6d}" for k in range(6)))
    print(f"j={j:2d} | WG0 {row[0]} | WG1 {row[1]} | MMA {row[2]}")

c0, g0, c1, g1 = (int(t[3, 0, k]) for k in range(4))
print(f"CTA 0 lifetime: {c1 - c0} cycles in {g1 - g0} ns -> SM clock {(c1 - c0) / max(g1 - g0, 1) * 1e3:.0f} MHz (30th back-to-back launch)")
