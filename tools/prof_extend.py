"""Dev tool (GPU): run only the cfg3 extend leg of bench.py (short, for ncu captures and quick A/B timing)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=40)
ap.add_argument("--warmup", type=int, default=3)
args = ap.parse_args()
torch.cuda.set_device(0)
gen = torch.Generator().manual_seed(3)
print(json.dumps(bench.run_extend_leg(args, "cuda:0", gen, bench.measured_peaks())))
