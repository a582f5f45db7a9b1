"""TEST INFRASTRUCTURE ONLY — numpy restatement of the paged allocator's extend / decode rule.

Follows python/sglang/srt/mem_cache/allocator/paged.py:44-102 (alloc_extend_naive) and
python/sglang/kernels/ops/memory/allocator.py:106-140 (alloc_decode_kernel).  Parity PINNED: tests/golden/alloc_cases.pt
holds outputs of the REAL alloc_extend_naive (oracle/gen_golden_alloc.py)."""
from __future__ import annotations

import numpy as np


def alloc_extend(prefix_lens, seq_lens, last_loc, free_pages, page_size: int) -> np.ndarray:
    pre, seq, last, free = (np.asarray(x, dtype=np.int64) for x in (prefix_lens, seq_lens, last_loc, free_pages))
    ext = seq - pre
    end = np.cumsum(ext)
    start = end - ext
    new_pages = (seq + page_size - 1) // page_size - (pre + page_size - 1) // page_size
    page_end = np.cumsum(new_pages)
    page_start = page_end - new_pages
    out = np.full(int(ext.sum()), -1, dtype=np.int64)
    lane = np.arange(page_size, dtype=np.int64)
    for i in range(len(pre)):
        up = (pre[i] + page_size - 1) // page_size * page_size
        n1 = min(seq[i], up) - pre[i]
        out[start[i]:start[i] + n1] = last[i] + 1 + lane[:n1]
        if pre[i] + n1 == seq[i]:
            continue
        n2 = seq[i] // page_size * page_size - up
        full = n2 // page_size
        pages = free[page_start[i]:page_start[i] + full]
        out[start[i] + n1:start[i] + n1 + n2] = (pages[:, None] * page_size + lane[None]).reshape(-1)
        n3 = seq[i] - pre[i] - n1 - n2
        if n3:
            out[end[i] - n3:end[i]] = free[page_end[i] - 1] * page_size + lane[:n3]
    return out


def alloc_decode(seq_lens, last_loc, free_pages, page_size: int) -> np.ndarray:
    seq, last, free = (np.asarray(x, dtype=np.int64) for x in (seq_lens, last_loc, free_pages))
    need = (seq + page_size - 1) // page_size - (seq - 1 + page_size - 1) // page_size
    before = np.cumsum(need) - need
    out = np.where(need == 0, last + 1, free[np.minimum(before, len(free) - 1)] * page_size)
    return out.astype(np.int64)
