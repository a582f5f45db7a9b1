"""Dev (GPU): cfg3 over an e4m3 pool, the in-kernel path only - the workload for an ncu capture of
extend_kernel<128, T, kFp8>:  ncu --set full -k regex:extend_kernel -s 5 -c 1 python tools/prof_extend_fp8.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sglang_b200 import ops

DEV, page = "cuda", 16
B, P, E, Hq, Hkv, D = 32, 2048, 512, 32, 8, 128
ctx, sp, own = P + E, P // page, E // page
n_pages = sp + B * own + 8
perm = torch.randperm(n_pages - 1) + 1
rows = [(torch.cat([perm[:sp], perm[sp + b * own: sp + (b + 1) * own]])[:, None] * page + torch.arange(page)).reshape(-1)
        for b in range(B)]
kv_indices = torch.cat(rows).int().to(DEV)
kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
qo_indptr = torch.arange(0, (B + 1) * E, E, dtype=torch.int32, device=DEV)
k8 = torch.randn((n_pages * page, Hkv, D), device=DEV).to(torch.float8_e4m3fn).view(torch.uint8)
v8 = torch.randn((n_pages * page, Hkv, D), device=DEV).to(torch.float8_e4m3fn).view(torch.uint8)
q = torch.randn((B * E, Hq, D), device=DEV).bfloat16()
o = torch.empty_like(q)
for _ in range(8):
    ops.extend_attention(q, k8, v8, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True, page_size=page,
                         max_qo_len=E, k_descale=0.6, v_descale=0.9)
torch.cuda.synchronize()
print("done")
