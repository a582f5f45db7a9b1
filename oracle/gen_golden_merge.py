"""Generate tests/golden/merge_state_*.pt from the REFERENCE's own merge_state kernels (run on the GPU box).

  * merge_state_v2 CUDA — python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu:31-206, compiled from
    where it lies by oracle/make_ref.sh into oracle/_ref/merge_state_ref.so;
  * merge_state_triton — python/sglang/kernels/ops/attention/merge_state.py:8-96, imported from the staged
    tree oracle/_ref/python (cross-check; both outputs are stored).

Cases cover bf16 / fp16 / fp32, head sizes 64 / 128 / 256, and the LSE edge values the reference handles:
+inf (treated as -inf, merge_attn_states.cu:66-67), one side -inf (the other side passes through), and both
sides -inf (the reference computes -inf - -inf = NaN; recorded as it is).

Run on the GPU box:  python oracle/gen_golden_merge.py gpurun_out/golden   (then copy the files to tests/golden/)
"""
from __future__ import annotations

import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_env  # noqa: E402


def main(out_dir: str) -> None:
    so = ref_env.merge_state_ref_path()
    assert so is not None, "oracle/_ref/merge_state_ref.so missing: run oracle/make_ref.sh in the build container"
    torch.ops.load_library(so)
    ref_env.install()
    from sglang.kernels.ops.attention.merge_state import merge_state_triton

    os.makedirs(out_dir, exist_ok=True)
    dev = "cuda"
    cases = [("bf16_t37_h8_d128", torch.bfloat16, 37, 8, 128), ("fp16_t5_h32_d64", torch.float16, 5, 32, 64),
             ("fp32_t16_h4_d256", torch.float32, 16, 4, 256), ("bf16_t300_h28_d128", torch.bfloat16, 300, 28, 128)]
    for name, dt, T, H, Dh in cases:
        g = torch.Generator().manual_seed(len(name) + T)
        va = torch.randn((T, H, Dh), generator=g).to(dt)
        vb = torch.randn((T, H, Dh), generator=g).to(dt)
        sa = torch.randn((T, H), generator=g) * 3
        sb = torch.randn((T, H), generator=g) * 3
        flat_a, flat_b = sa.view(-1), sb.view(-1)
        flat_a[0::11] = math.inf          # +inf -> -inf (empty side a)
        flat_b[3::13] = -math.inf         # empty side b
        flat_a[5::17] = -math.inf
        flat_b[5::17] = -math.inf         # both empty (reference: NaN)
        flat_b[7::19] = math.inf
        flat_a[2::23] = 80.0              # large but finite spread
        flat_b[2::23] = -80.0
        d = {k: t.to(dev) for k, t in dict(va=va, vb=vb, sa=sa, sb=sb).items()}
        vo = torch.empty_like(d["va"])
        so_ = torch.empty_like(d["sa"])
        torch.ops.b200ref.merge_state_v2(d["va"], d["sa"], d["vb"], d["sb"], vo, so_)
        vt, st = merge_state_triton(d["va"], d["sa"], d["vb"], d["sb"])
        torch.cuda.synchronize()
        rec = dict(name=name, v_a=va, s_a=sa, v_b=vb, s_b=sb, v_cuda=vo.cpu(), s_cuda=so_.cpu(), v_triton=vt.cpu(),
                   s_triton=st.cpu(), torch_version=torch.__version__, gpu=torch.cuda.get_device_name(0),
                   source="merge_attn_states.cu:31-206 (merge_state_v2) + kernels/ops/attention/merge_state.py:8-96")
        path = os.path.join(out_dir, f"merge_state_{name}.pt")
        torch.save(rec, path)
        fin = torch.isfinite(vo.float()).all(dim=-1)
        agree = torch.allclose(vo.float()[fin], vt.float()[fin], rtol=1e-2, atol=1e-2)
        print(f"{name}: {os.path.getsize(path) / 1024:.0f} KiB, rows with NaN from the reference: "
              f"{int((~fin).sum())}/{fin.numel()}, cuda~triton on finite rows: {agree}")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
