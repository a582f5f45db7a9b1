"""M-RoPE position builders against fixtures generated from the REAL reference functions
(oracle/gen_golden_mrope.py: `get_rope_index`, mrope_rope_index.py:46-230, and
`ForwardBatch._compute_mrope_positions_extend/_decode`, forward_batch_info.py:1129-1254).  Integer work: bit-exact."""
import os

import torch

from sglang_b200 import rotary
from tests.util import GOLDEN_DIR, load_golden

G = load_golden(os.path.join(GOLDEN_DIR, "mrope_positions.pt"))


def test_rope_index_matches_reference():
    assert len(G["requests"]) == 6
    for key, r in G["requests"].items():
        grids = torch.tensor([[it[1], it[2], it[3]] for it in r["layout"] if it[0] == "image"], dtype=torch.int64)
        pos, delta = rotary.mrope_rope_index(r["input_ids"], grids, image_token_id=G["image_token_id"],
                                             vision_start_token_id=G["vision_start_token_id"],
                                             spatial_merge_size=G["spatial_merge_size"])
        assert pos.dtype == torch.int64 and torch.equal(pos, r["mrope_positions"]), key
        assert torch.equal(delta.view(-1), r["mrope_position_delta"].view(-1)), key


def test_batch_positions_match_reference():
    reqs = G["requests"]
    for b in G["batches"]:
        mm = [None if k is None else reqs[k]["mrope_positions"] for k in b["keys"]]
        deltas = [None if k is None else reqs[k]["mrope_position_delta"] for k in b["keys"]]
        if b["kind"] == "extend":
            got = rotary.mrope_positions_extend(b["prefix_lens"], b["extend_lens"], mm, deltas)
        else:
            got = rotary.mrope_positions_decode(torch.tensor(b["seq_lens"], dtype=torch.int32),
                                                None if all(k is None for k in b["keys"]) else deltas)
        assert got.dtype == torch.int64, b["name"]
        assert torch.equal(got, b["mrope_positions"]), b["name"]
