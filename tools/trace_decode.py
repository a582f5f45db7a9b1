"""Dev (GPU): per-CTA phase timeline of decode_kernel (DEC_TRACE build, libb200attn_trace.so).
Stamps (globaltimer, ns): 0 CTA start, 1 after griddepcontrol.wait, 2 prologue done (Q fragments, first loads issued),
3 first tile landed, 4 last tile consumed, 5 partial / output written, 6 ticket drawn, 7 end (merging CTA only).
usage: python tools/trace_decode.py B ctx [max_splits]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sglang_b200 import _lib

_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libb200attn_trace.so")
import bench  # noqa: E402
from sglang_b200 import ops  # noqa: E402

dev = "cuda:0"
B, ctx = int(sys.argv[1]), int(sys.argv[2])
gen = torch.Generator().manual_seed(1)
c = dict(B=B, ctx=ctx, Hq=32, Hkv=8, D=128, page=16)
req_pool, kv_pool, host = bench.build_decode_state(c, dev, gen)
kbuf, vbuf = kv_pool.get_kv_buffer(0)
kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
kv_indices = torch.zeros(B * ctx, dtype=torch.int32, device=dev)
ops.build_kv_indices(req_pool.req_to_token, host["req_pool_indices"].to(dev), host["seq_lens"].to(dev), kv_indptr,
                     kv_indices)
q = host["q"].to(dev).view(B, 32, 128)
o = torch.empty_like(q)
warps, ms, tb = ops.plan_decode(B, 8, 4, ctx, 256)
if len(sys.argv) > 3:
    ms = int(sys.argv[3])
ws = torch.zeros(max(ops.decode_workspace_bytes(B, 32, 128, ms), 8), dtype=torch.uint8, device=dev)
f = lambda: ops.decode_attention(q, kbuf, vbuf, o, kv_indptr, kv_indices, 128 ** -0.5, max_splits=ms, tail_begin=tb,
                                 warps=warps, workspace=ws)
for _ in range(5):
    f()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
f()
e1.record()
torch.cuda.synchronize()
n_cta = ms * 8 * B
buf = (ctypes.c_longlong * (8 * 16384))()
lib = ctypes.CDLL(_lib.LIB_PATH)
assert lib.b200attn_decode_trace_read(buf, ctypes.c_int64(8 * 8 * min(n_cta, 16384))) == 0
t = torch.tensor(list(buf[: 8 * min(n_cta, 16384)]), dtype=torch.float64).view(-1, 8)
t0 = t[:, 0].min()
t = (t - t0) / 1e3
qs = torch.tensor([0.0, 0.5, 0.9, 1.0], dtype=torch.float64)
names = ["start", "dep-wait", "prologue", "1st tile", "loop end", "written", "ticket", "merged"]
print(f"B={B} ctx={ctx} warps={warps} max_splits={ms} ctas={n_cta}: event time {e0.elapsed_time(e1) * 1e3:.1f} us")
for k, nm in enumerate(names):
    col = t[:, k]
    col = col[col >= 0] if k < 7 else col[(col > 0) & (col < 1e6)]
    if col.numel() == 0:
        continue
    print(f"  {nm:9s} min/med/p90/max " + " ".join(f"{float(x):7.2f}" for x in torch.quantile(col, qs)) + f"  (n={col.numel()})")
d = t[:, 3] - t[:, 0]
print(f"  start -> first tile: median {float(d.median()):.2f} us; loop: median {float((t[:, 4] - t[:, 3]).median()):.2f} us; "
      f"loop end -> written: median {float((t[:, 5] - t[:, 4]).median()):.2f} us")
