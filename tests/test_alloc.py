"""Paged allocator: oracle vs reference-generated fixtures (CPU) and device kernels vs both (GPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import alloc_port
from tests.util import GOLDEN_DIR

CASES = torch.load(os.path.join(GOLDEN_DIR, "alloc_cases.pt"), weights_only=False)
IDS = [f"B{c['B']}p{c['page_size']}" for c in CASES]


@pytest.mark.parametrize("c", CASES, ids=IDS)
def test_oracle_matches_reference_naive(c):
    got = alloc_port.alloc_extend(c["prefix_lens"], c["seq_lens"], c["last_loc"], c["free_pages"], c["page_size"])
    assert np.array_equal(got, c["out_indices"].numpy())
    assert len(np.unique(got)) == len(got)  # paged.py:204-205 debug invariant


def test_oracle_decode_is_extend_by_one():
    for c in CASES:
        seq = c["prefix_lens"] + 1
        ok = c["prefix_lens"] > 0
        seq, last, pre = seq[ok], c["last_loc"][ok], c["prefix_lens"][ok]
        if len(seq) == 0:
            continue
        a = alloc_port.alloc_decode(seq, last, c["free_pages"], c["page_size"])
        b = alloc_port.alloc_extend(pre, seq, last, c["free_pages"], c["page_size"])
        assert np.array_equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("c", CASES, ids=IDS)
@pytest.mark.parametrize("dt", [torch.int64, torch.int32])
def test_device_alloc_extend_bit_exact(c, dt):
    from sglang_b200 import ops
    dev = "cuda"
    out = torch.full_like(c["out_indices"], -7).to(dt).to(dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    ops.alloc_extend(c["prefix_lens"].to(dt).to(dev), c["seq_lens"].to(dt).to(dev), c["last_loc"].to(dt).to(dev),
                     c["free_pages"].to(dev), out, c["page_size"], status)
    assert torch.equal(out.cpu().long(), c["out_indices"])
    assert int(status) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("B,page", [(1, 16), (64, 16), (255, 1), (256, 32), (257, 16), (1000, 64), (4096, 16)])
def test_device_alloc_decode_bit_exact(B, page):
    from sglang_b200 import ops
    g = torch.Generator().manual_seed(B * 7 + page)
    seq = torch.randint(1, 5000, (B,), generator=g)
    seq[::3] = (seq[::3] // page) * page + 1  # force a fresh page on a third of the rows
    perm = torch.randperm(4 * B + 8, generator=g) + 1
    last = perm[:B] * page + (seq - 2) % page
    free = perm[B:2 * B + 4]
    want = alloc_port.alloc_decode(seq, last, free, page)
    out = torch.empty(B, dtype=torch.int64, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.alloc_decode(seq.cuda(), last.cuda(), free.cuda(), out, page, status)
    assert np.array_equal(out.cpu().numpy(), want) and int(status) == 0


@pytest.mark.gpu
def test_allocator_mirror_lifecycle_feeds_backend_tables():
    """alloc_extend -> req_to_token rows -> alloc_decode; exhausted free list returns None like the reference."""
    from sglang_b200.allocator import PagedTokenToKVPoolAllocator
    page = 16
    a = PagedTokenToKVPoolAllocator(size=64 * page, page_size=page)
    pre = torch.tensor([0, 0, 0])
    seq = torch.tensor([40, 16, 5])
    last = torch.tensor([-1, -1, -1])
    out = a.alloc_extend(pre.cuda(), pre, seq.cuda(), seq, last.cuda(), int(seq.sum()))
    want = alloc_port.alloc_extend(pre, seq, last, np.arange(1, 65), page)
    assert np.array_equal(out.cpu().numpy(), want)
    assert a.available_size() == (64 - 5) * page
    ends = torch.cumsum(seq, 0) - 1
    last2 = out[ends.cuda()]
    seq2 = seq + 1
    out2 = a.alloc_decode(seq2.cuda(), seq2, last2)
    want2 = alloc_port.alloc_decode(seq2, last2.cpu(), np.arange(6, 65), page)
    assert np.array_equal(out2.cpu().numpy(), want2)
    assert a.available_size() == (64 - 6) * page  # only the 16-token request opened a page
    a.free(out[:40])
    assert a.available_size() == (64 - 3) * page
    big = torch.tensor([64 * page])
    assert a.alloc_extend(torch.tensor([0]).cuda(), torch.tensor([0]), big.cuda(), big, torch.tensor([-1]).cuda(),
                          int(big)) is None


@pytest.mark.gpu
def test_device_alloc_reports_exhaustion():
    from sglang_b200 import ops
    out = torch.full((48,), -7, dtype=torch.int64, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.alloc_extend(torch.tensor([0], device="cuda"), torch.tensor([48], device="cuda"),
                     torch.tensor([-1], device="cuda"), torch.tensor([5, 6], device="cuda"), out, 16, status)
    assert int(status) == 1 and int(out[0]) == -7


@pytest.mark.gpu
def test_device_alloc_extend_large_ragged_batch():
    """4096 requests, ragged prefixes / extends, page 16: bit-exact against the numpy oracle; every slot unique."""
    from sglang_b200 import ops
    g = torch.Generator().manual_seed(99)
    B, page = 4096, 16
    pre = torch.randint(0, 3000, (B,), generator=g)
    ext = torch.randint(1, 200, (B,), generator=g)
    seq = pre + ext
    n_new = int(((seq + page - 1) // page - (pre + page - 1) // page).sum())
    perm = torch.randperm(B + n_new + 64, generator=g) + 1
    last = torch.where(pre > 0, perm[:B] * page + (pre - 1) % page, torch.full_like(pre, -1))
    free = perm[B:B + n_new + 8]
    want = alloc_port.alloc_extend(pre, seq, last, free, page)
    out = torch.full((int(ext.sum()),), -7, dtype=torch.int64, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.alloc_extend(pre.cuda(), seq.cuda(), last.cuda(), free.cuda(), out, page, status)
    assert np.array_equal(out.cpu().numpy(), want) and int(status) == 0
    assert len(np.unique(want)) == len(want)
