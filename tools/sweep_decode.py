"""Dev tool (GPU): time the decode kernel on cfg2 across split caps / chunk floors."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from sglang_b200 import ops

dev = "cuda:0"
c = bench.CFG2
gen = torch.Generator().manual_seed(1)
req_pool, kv_pool, host = bench.build_decode_state(c, dev, gen)
B, Hq, Hkv, D = c["B"], c["Hq"], c["Hkv"], c["D"]
seq = host["seq_lens"].to(dev)
kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
kv_indices = torch.zeros(B * c["ctx"], dtype=torch.int32, device=dev)
ops.build_kv_indices(req_pool.req_to_token, host["req_pool_indices"].to(dev), seq, kv_indptr, kv_indices)
q = host["q"].to(dev).view(B, Hq, D)
o = torch.empty_like(q)
kbuf, vbuf = kv_pool.get_kv_buffer(0)
nbytes = bench.decode_bytes(c) - 2 * 2 * B * Hkv * D * 2
for ms in (1, 2, 3, 4, 5, 6, 7, 8, 11, 12, 16):
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=dev)
    f = lambda: ops.decode_attention(q, kbuf, vbuf, o, kv_indptr, kv_indices, D ** -0.5, max_splits=ms, min_chunk=64, workspace=ws)
    t = bench.time_steps(f, 200, 10) / 200
    print(f"max_splits={ms:3d}  {t*1e3:8.1f} us  {nbytes / t / 1e6:8.1f} GB/s", flush=True)
