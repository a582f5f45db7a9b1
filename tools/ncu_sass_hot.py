"""Dev tool (CPU): list the SASS instructions with the most stall samples from an .ncu-rep source page.
usage: python tools/ncu_sass_hot.py file.ncu-rep [topN]"""
import csv
import subprocess
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ix = {k: i for i, k in enumerate(hdr)}
stalls = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
body = [r for r in rows[2:] if len(r) == len(hdr)]
tot = sum(int(r[ix["# Samples"]] or 0) for r in body)
print("total samples", tot)
agg = {k: sum(int(r[ix[k]] or 0) for r in body) for k in stalls}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
order = sorted(range(len(body)), key=lambda i: -int(body[i][ix["# Samples"]] or 0))[:top]
for i in sorted(order):
    r = body[i]
    st = {k[6:]: int(r[ix[k]]) for k in stalls if int(r[ix[k]] or 0) > 0}
    print(f"{i:5d} {r[ix['# Samples']]:>6s} x{r[ix['Instructions Executed']]:>9s}  {r[ix['Source']][:70]:70s} {st}")
