"""Generate tests/golden/alloc_cases.pt with the REAL reference allocator rule on CPU (this container only):
``alloc_extend_naive`` (python/sglang/srt/mem_cache/allocator/paged.py:44-102).  Run: python oracle/gen_golden_alloc.py"""
from __future__ import annotations

import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gen_golden import OUT, _install_stubs  # noqa: E402


def make_case(g, B, page, max_pre, max_ext):
    pre = torch.randint(0, max_pre + 1, (B,), generator=g)
    ext = torch.randint(1, max_ext + 1, (B,), generator=g)
    seq = pre + ext
    n_pages_total = int(((seq + page - 1) // page).sum()) + 8
    perm = torch.randperm(4 * n_pages_total, generator=g) + 1
    # each request's last prefix token sits at offset (pre-1) % page of some already-owned page
    own = perm[:B]
    last_loc = torch.where(pre > 0, own * page + (pre - 1) % page, torch.full_like(pre, -1))
    free = perm[B:B + n_pages_total].clone()
    return pre, seq, last_loc, free


def main() -> None:
    _install_stubs()
    from sglang.srt.mem_cache.allocator.paged import alloc_extend_naive

    g = torch.Generator().manual_seed(20261104 + 300)
    cases = []
    for B, page, max_pre, max_ext in [(1, 16, 0, 40), (4, 16, 50, 70), (7, 1, 9, 9), (16, 64, 300, 500), (33, 32, 0, 5),
                                      (5, 16, 64, 1), (64, 16, 4096, 1), (3, 128, 1000, 3000), (300, 16, 40, 40)]:
        pre, seq, last_loc, free = make_case(g, B, page, max_pre, max_ext)
        out = torch.full((int((seq - pre).sum()),), -1, dtype=torch.int64)
        alloc_extend_naive(pre, seq, last_loc, free, out, page, "cpu")
        cases.append(dict(B=B, page_size=page, prefix_lens=pre, seq_lens=seq, last_loc=last_loc, free_pages=free,
                          out_indices=out))
        print(B, page, out.numel(), int(out.min()))
    torch.save(cases, os.path.join(OUT, "alloc_cases.pt"))


if __name__ == "__main__":
    main()
