"""Dev tool (GPU): run the cfg3 extend kernel back to back for a few seconds while sampling SM clock / power."""
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from sglang_b200 import _lib

if os.environ.get("B200ATTN_VARIANT"):
    _lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libb200attn_%s.so" % os.environ["B200ATTN_VARIANT"])
import bench
from sglang_b200 import ops

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 3.0
torch.cuda.set_device(0)
dev = "cuda:0"
gen = torch.Generator().manual_seed(3)
c = bench.CFG3
B, P, E, Hq, Hkv, D, page = c["B"], c["prefix"], c["ext"], c["Hq"], c["Hkv"], c["D"], c["page"]
num_pages = (P + B * E) // page + 8
table = torch.zeros((B + 1, P + E), dtype=torch.int32)
offs = torch.arange(page)
perm = torch.randperm(num_pages - 1, generator=gen) + 1
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
f = lambda: ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True,
                                 page_size=page, max_qo_len=E)
for _ in range(5):
    f()
torch.cuda.synchronize()
# burst: 20 launches
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    f()
e1.record()
torch.cuda.synchronize()
fl = bench.extend_flops(c)
print(f"burst 20 launches: {fl / (e0.elapsed_time(e1) / 20 * 1e-3) / 1e12:.1f} TFLOP/s")
smi = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,"
                        "clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown",
                        "--format=csv,noheader", "-lms", "100"], stdout=subprocess.PIPE, text=True)
t0 = time.time()
n = 0
e0.record()
while time.time() - t0 < secs:
    for _ in range(50):
        f()
    n += 50
    torch.cuda.synchronize()
e1.record()
torch.cuda.synchronize()
smi.terminate()
out = smi.stdout.read().strip().splitlines()
print(f"sustained {n} launches: {fl / (e0.elapsed_time(e1) / n * 1e-3) / 1e12:.1f} TFLOP/s")
print("\n".join(out[2:-1:3][:14]))
# per-launch times in a burst after idling (clock / power-cap dynamics)
time.sleep(1.5)
evs = [torch.cuda.Event(enable_timing=True) for _ in range(41)]
evs[0].record()
for k in range(40):
    f()
    evs[k + 1].record()
torch.cuda.synchronize()
print("per-launch TFLOP/s after 1.5 s idle:", " ".join(f"{fl / (evs[k].elapsed_time(evs[k + 1]) * 1e-3) / 1e12:.0f}" for k in range(40)))
