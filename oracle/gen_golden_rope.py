"""Generate tests/golden/rope_*.pt by running the REAL reference rotary classes on CPU (this container only).

Imports RotaryEmbedding / MRotaryEmbedding from /root/reference/python (stubs of oracle/gen_golden.py) and records
inputs, the reference's own cos_sin_cache and the outputs of ``forward_native``
(python/sglang/srt/layers/rotary_embedding/base.py:236-276, mrope.py:173-217).

Run:  python oracle/gen_golden_rope.py        (needs /root/reference; NOT run on the GPU box)
"""
from __future__ import annotations

import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gen_golden import OUT, _install_stubs  # noqa: E402

# name, dtype, D, rot, neox, Hq, Hkv, T, max_pos, base, mrope_section, interleaved
CASES = [
    ("rope_llama_neox", torch.bfloat16, 128, 128, True, 8, 2, 37, 512, 500000.0, None, False),
    ("rope_gptj_fp16", torch.float16, 64, 64, False, 14, 2, 29, 512, 10000.0, None, False),
    ("rope_partial_neox", torch.bfloat16, 128, 64, True, 4, 4, 33, 512, 10000.0, None, False),
    ("rope_partial_gptj", torch.bfloat16, 96, 32, False, 6, 2, 19, 256, 10000.0, None, False),
    ("rope_mrope_qwen25vl", torch.bfloat16, 128, 128, True, 7, 1, 48, 512, 1000000.0, [16, 24, 24], False),
    ("rope_mrope_interleaved", torch.bfloat16, 128, 128, True, 4, 2, 40, 512, 5000000.0, [24, 20, 20], True),
]


def main() -> None:
    _install_stubs()
    from sglang.srt.layers.rotary_embedding.base import RotaryEmbedding
    from sglang.srt.layers.rotary_embedding.mrope import MRotaryEmbedding

    index = {}
    for i, (name, dt, D, rot, neox, Hq, Hkv, T, max_pos, base, sec, inter) in enumerate(CASES):
        g = torch.Generator().manual_seed(20261104 + 100 + i)
        if sec is None:
            rope = RotaryEmbedding(D, rot, max_pos, base, neox, dt)
            pos = torch.randint(0, max_pos, (T,), generator=g, dtype=torch.int64)
        else:
            rope = MRotaryEmbedding(D, rot, max_pos, base, neox, dt, mrope_section=sec, mrope_interleaved=inter)
            pos = torch.randint(0, max_pos, (3, T), generator=g, dtype=torch.int64)
            pos[:, T // 2:] = pos[0:1, T // 2:]  # text rows: t == h == w
        q = torch.randn((T, Hq * D), generator=g).to(dt)
        k = torch.randn((T, Hkv * D), generator=g).to(dt)
        oq, ok = rope.forward_native(pos, q.clone(), k.clone())
        cache32 = rope._compute_cos_sin_cache().float()  # the reference's own table before its dtype cast
        torch.save(dict(name=name, D=D, rot=rot, neox=neox, Hq=Hq, Hkv=Hkv, T=T, base=base, max_pos=max_pos,
                        mrope_section=sec, interleaved=inter, positions=pos, q=q, k=k, out_q=oq, out_k=ok,
                        cos_sin_cache=cache32), os.path.join(OUT, name + ".pt"))
        index[name] = dict(dtype=str(dt), D=D, rot=rot, neox=neox, T=T, mrope_section=sec, interleaved=inter)
        print(name, tuple(oq.shape), tuple(ok.shape))
    with open(os.path.join(OUT, "INDEX_rope.json"), "w") as f:
        json.dump(index, f, indent=1)


if __name__ == "__main__":
    main()
