"""Generate tests/golden/mrope_positions.pt from the REAL M-RoPE position builders (this container only).

  * `get_rope_index` (python/sglang/srt/layers/rotary_embedding/mrope_rope_index.py:46-230) for Qwen2-VL /
    Qwen2.5-VL token layouts (text + image grids): the per-request `mrope_positions [3, L]` and
    `mrope_position_delta` the multimodal processor attaches to a request;
  * `ForwardBatch._compute_mrope_positions_extend / _decode`
    (python/sglang/srt/model_executor/forward_batch_info.py:1129-1254), called unbound on SimpleNamespace
    stand-ins for the batch, including a chunked-prefill slice, a text-only request, and the
    `_expand_mrope_from_input` fallback (:1103-1127).

Run:  python oracle/gen_golden_mrope.py      (needs /root/reference; NOT run on the GPU box)
"""
from __future__ import annotations

import os
import sys
from types import SimpleNamespace

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_env  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
IMG, VID, VSTART = 151655, 151656, 151652


def layout_ids(layout):
    """layout: list of ("text", n) | ("image", t, h, w) in GRID units (pre spatial-merge) -> input ids."""
    ids = []
    for item in layout:
        if item[0] == "text":
            ids += list(range(100, 100 + item[1]))
        else:
            _, t, h, w = item
            ids += [VSTART] + [IMG] * (t * (h // 2) * (w // 2))
    return ids


def main() -> None:
    ref_env.install()
    from sglang.srt.layers.rotary_embedding.mrope_rope_index import get_rope_index
    from sglang.srt.model_executor import forward_batch_info as fbi

    layouts = {
        "cfg4_image_then_text": [("text", 3), ("image", 1, 64, 64), ("text", 252)],   # 3 + 1 + 1024 + 252 = 1280
        "two_images": [("text", 5), ("image", 1, 8, 12), ("text", 7), ("image", 1, 4, 4), ("text", 2)],
        "image_last": [("text", 9), ("image", 1, 6, 6)],
    }
    reqs = {}
    for model_type in ("qwen2_5_vl", "qwen2_vl"):
        for name, layout in layouts.items():
            ids = torch.tensor([layout_ids(layout)], dtype=torch.int64)
            grids = torch.tensor([[it[1], it[2], it[3]] for it in layout if it[0] == "image"], dtype=torch.int64)
            pos, delta = get_rope_index(2, IMG, VID, VSTART, model_type, tokens_per_second=2, input_ids=ids,
                                        image_grid_thw=grids)
            reqs[f"{model_type}/{name}"] = dict(layout=layout, model_type=model_type, input_ids=ids[0],
                                                mrope_positions=pos[:, 0].clone(), mrope_position_delta=delta.clone())

    def mm(key):
        r = reqs[key]
        return SimpleNamespace(mrope_positions=r["mrope_positions"], mrope_position_delta=r["mrope_position_delta"],
                               mrope_position_delta_repeated_cache=None)

    # the deterministic-mode switch the builders read (forward_batch_info.py: get_exec().deterministic)
    fbi.get_exec = lambda: SimpleNamespace(deterministic=SimpleNamespace(rl_on_policy_target=None))
    runner = SimpleNamespace(device="cpu")
    batches = []
    keys = ["qwen2_5_vl/cfg4_image_then_text", None, "qwen2_5_vl/two_images", "qwen2_vl/image_last"]
    L = [None if k is None else reqs[k]["mrope_positions"].shape[1] for k in keys]
    ext_cases = [
        ("extend_full", [0, 0, 0, 0], [L[0], 17, L[2], L[3]]),
        ("extend_chunked", [512, 5, 10, 3], [L[0] - 512, 40, L[2] - 10, L[3] - 3]),
        ("extend_past_prompt", [L[0], 0, L[2] + 4, L[3]], [1, 3, 2, 1]),
    ]
    for name, pre, ext in ext_cases:
        seq = [p + e for p, e in zip(pre, ext)]
        self_ = SimpleNamespace(seq_lens_cpu=torch.tensor(seq), mrope_positions=None)
        self_._expand_mrope_from_input = lambda mm_in, sl, _s=self_: fbi.ForwardBatch._expand_mrope_from_input(_s, mm_in, sl)
        batch = SimpleNamespace(multimodal_inputs=[None if k is None else mm(k) for k in keys], extend_lens=ext,
                                prefix_lens=pre)
        fbi.ForwardBatch._compute_mrope_positions_extend(self_, runner, batch)
        batches.append(dict(name=name, kind="extend", keys=keys, prefix_lens=pre, extend_lens=ext, seq_lens=seq,
                            mrope_positions=self_.mrope_positions.clone()))
    for name, seq, use in [("decode_mm", [L[0] + 1, 30, L[2] + 7, L[3] + 2], keys),
                           ("decode_text_only", [5, 30, 9, 100], [None] * 4)]:
        self_ = SimpleNamespace(seq_lens_cpu=torch.tensor(seq), seq_lens=torch.tensor(seq, dtype=torch.int32),
                                mrope_positions=None)
        batch = SimpleNamespace(multimodal_inputs=[None if k is None else mm(k) for k in use])
        for m in batch.multimodal_inputs:  # requests past their prompt: positions come from the delta
            if m is not None:
                m.mrope_positions = None
        fbi.ForwardBatch._compute_mrope_positions_decode(self_, runner, batch)
        batches.append(dict(name=name, kind="decode", keys=use, seq_lens=seq,
                            mrope_positions=self_.mrope_positions.clone()))
    path = os.path.join(OUT, "mrope_positions.pt")
    torch.save(dict(requests=reqs, batches=batches, spatial_merge_size=2, tokens_per_second=2,
                    image_token_id=IMG, vision_start_token_id=VSTART,
                    generator="oracle/gen_golden_mrope.py", torch_version=torch.__version__), path)
    print(f"{path}: {os.path.getsize(path) / 1024:.0f} KiB, {len(reqs)} requests, {len(batches)} batches")
    for b in batches:
        print(" ", b["name"], tuple(b["mrope_positions"].shape))


if __name__ == "__main__":
    main()
