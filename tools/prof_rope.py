"""One fused rope+store launch at the cfg3 shape (for ncu).  GPU box only."""
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import ops  # noqa: E402
from sglang_b200.rotary import compute_cos_sin_cache  # noqa: E402

T, Hq, Hkv, D = 16384, 32, 8, 128
dev = "cuda"
qkv = torch.randn((T, (Hq + 2 * Hkv) * D), device=dev).bfloat16()
q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
kc = torch.zeros((T + 64, Hkv, D), device=dev).bfloat16()
vc = torch.zeros_like(kc)
loc = torch.randperm(T + 48, device=dev)[:T] + 16
pos = torch.randint(0, 8192, (T,), device=dev)
cache = compute_cos_sin_cache(D, 8192, 5e5).to(dev)
for _ in range(3):
    ops.apply_rope_with_cos_sin_cache_inplace(pos, q, k, cache, head_size=D, v=v, k_buffer=kc, v_buffer=vc, cache_loc=loc)
torch.cuda.synchronize()
