"""b200attn_stage_copy: host <-> device staging by a kernel (pinned memory is mapped), also inside a CUDA graph."""
import pytest
import torch

from sglang_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_two_segments_h2d_and_d2h():
    a = torch.randn(64, 6144).to(torch.bfloat16).pin_memory()
    b = torch.randint(0, 1 << 40, (3, 64), dtype=torch.int64).pin_memory()
    da, db = torch.zeros_like(a, device=DEV), torch.zeros_like(b, device=DEV)
    ops.stage_copy(db, b, da, a)
    torch.cuda.synchronize()
    assert torch.equal(da.cpu(), a) and torch.equal(db.cpu(), b)
    back = torch.zeros_like(a).pin_memory()
    ops.stage_copy(back, da)
    torch.cuda.synchronize()
    assert torch.equal(back, a)
    # device -> device and a large, odd multiple of 16 bytes
    x = torch.randint(0, 255, (16 * 100003,), dtype=torch.uint8, device=DEV)
    y = torch.zeros_like(x)
    ops.stage_copy(y, x)
    assert torch.equal(x, y)


def test_inside_a_graph_sees_new_host_data_every_replay():
    h = torch.zeros(4096, dtype=torch.int32).pin_memory()
    out = torch.zeros(4096, dtype=torch.int32).pin_memory()
    d = torch.zeros(4096, dtype=torch.int32, device=DEV)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ops.stage_copy(d, h)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            ops.stage_copy(d, h)
            d.add_(1)
            ops.stage_copy(out, d)
    for i in range(5):
        h.fill_(i * 7)
        g.replay()
        torch.cuda.synchronize()
        assert int(out.min()) == int(out.max()) == i * 7 + 1


def test_rejects_pageable_host_memory_and_bad_sizes():
    d = torch.zeros(64, dtype=torch.uint8, device=DEV)
    with pytest.raises(RuntimeError, match="pinned"):
        ops.stage_copy(d, torch.zeros(64, dtype=torch.uint8))
    with pytest.raises(RuntimeError):
        ops.stage_copy(d[:8], torch.zeros(8, dtype=torch.uint8).pin_memory())  # 8 bytes: not a multiple of 16
