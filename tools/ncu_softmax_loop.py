"""Dev (CPU): stall samples of the extend kernel's hot softmax loop from an .ncu-rep source page (first kernel of the
report): the instructions executed most often, their stall reasons summed, and the top individual instructions.
usage: python tools/ncu_softmax_loop.py file.ncu-rep [min_exec_fraction=0.9]"""
import csv
import subprocess
import sys

path = sys.argv[1]
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.9
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ix = {k: i for i, k in enumerate(hdr)}
body = []
for r in rows[2:]:
    if len(r) != len(hdr):
        continue
    if r == hdr:
        break
    body.append(r)
stalls = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
S = lambda r: int(r[ix["# Samples"]] or 0)
E = lambda r: int(r[ix["Instructions Executed"]] or 0)
tot = sum(S(r) for r in body)
# the hot loop = instructions executed at least `frac` x the count of the most executed MUFU.EX2
ex2 = max(E(r) for r in body if "MUFU.EX2" in r[ix["Source"]])
hot = [r for r in body if E(r) >= frac * ex2]
hs = sum(S(r) for r in hot)
print(f"{len(body)} instructions, {tot} samples; hot loop: {len(hot)} instructions (executed >= {frac} x {ex2}), "
      f"{hs} samples = {100 * hs / tot:.1f} %")
agg = {k[6:]: sum(int(r[ix[k]] or 0) for r in hot) for k in stalls}
print("hot-loop stalls:", {k: f"{100 * v / hs:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > hs * 0.01})
for i, r in enumerate(body):
    if E(r) >= frac * ex2 and S(r) >= hs * 0.008:
        st = {k[6:]: int(r[ix[k]]) for k in stalls if int(r[ix[k]] or 0) > 0}
        st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{i:5d} {S(r):6d}  {r[ix['Source']][:64]:64s} {st}")
