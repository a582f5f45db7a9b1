"""cfg4 MIXED (Qwen2.5-VL shapes) layer-call for kernel-level profiling: ncu --metrics gpu__time_duration.sum python tools/prof_cfg4.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

dev = "cuda:0"
torch.cuda.set_device(dev)
gen = torch.Generator().manual_seed(4)
causal = (sys.argv[1] if len(sys.argv) > 1 else "causal") == "causal"
r = bench.run_backend_case(dev, gen, bench.measured_peaks(), "cfg4", [0] * 16 + [1279] * 16, [1280] * 16 + [1] * 16,
                           28, 4, 128, 16, causal, "MIXED", steps=4, warmup=2)
print(r)
