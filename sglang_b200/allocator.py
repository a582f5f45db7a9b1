"""Host-side mirror of ``PagedTokenToKVPoolAllocator`` (python/sglang/srt/mem_cache/allocator/paged.py:105-330)
for the part of it that sits on the hot path: the free-page list lives on the device, ``alloc_extend`` /
``alloc_decode`` run the device kernels of csrc/alloc.cu (no torch fallback), the list is advanced with host
arithmetic on the CPU copies of the lengths exactly as the reference does (paged.py:207-217, 246-256)."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops


def get_num_new_pages(seq_lens: torch.Tensor, page_size: int, prefix_lens: Optional[torch.Tensor] = None,
                      decode: bool = False) -> int:
    """python/sglang/srt/utils (get_num_new_pages): host-side count of pages an extend / decode step opens."""
    after = (seq_lens + page_size - 1) // page_size
    if decode:
        before = (seq_lens - 1 + page_size - 1) // page_size
    else:
        before = (prefix_lens + page_size - 1) // page_size
    return int((after - before).sum())


class PagedTokenToKVPoolAllocator:
    def __init__(self, size: int, page_size: int, device: str = "cuda"):
        self.size, self.page_size, self.device = size, page_size, device
        self.num_pages = size // page_size
        self.status = torch.zeros(1, dtype=torch.int32, device=device)
        self.clear()

    def clear(self) -> None:
        # page 0 is the padding page (paged.py:322-330)
        self.free_pages = torch.arange(1, self.num_pages + 1, dtype=torch.int64, device=self.device)
        self.release_pages = torch.empty((0,), dtype=torch.int64, device=self.device)

    def available_size(self) -> int:
        return (len(self.free_pages) + len(self.release_pages)) * self.page_size

    def merge_and_sort_free(self) -> None:
        if len(self.release_pages) > 0:
            self.free_pages = torch.sort(torch.cat((self.free_pages, self.release_pages)))[0]
            self.release_pages = torch.empty((0,), dtype=torch.int64, device=self.device)

    def alloc(self, need_size: int) -> Optional[torch.Tensor]:
        """Whole pages (paged.py:150-170)."""
        assert need_size % self.page_size == 0
        n = need_size // self.page_size
        if n > len(self.free_pages):
            self.merge_and_sort_free()
        if n > len(self.free_pages):
            return None
        pages, self.free_pages = self.free_pages[:n], self.free_pages[n:]
        return (pages[:, None] * self.page_size + torch.arange(self.page_size, device=self.device)).reshape(-1)

    def alloc_extend(self, prefix_lens, prefix_lens_cpu, seq_lens, seq_lens_cpu, last_loc, extend_num_tokens: int,
                     num_new_pages: Optional[int] = None) -> Optional[torch.Tensor]:
        bs = len(prefix_lens)
        if extend_num_tokens // self.page_size + bs + 1 > len(self.free_pages):
            self.merge_and_sort_free()
        if num_new_pages is None:
            num_new_pages = get_num_new_pages(seq_lens_cpu, self.page_size, prefix_lens=prefix_lens_cpu)
        if num_new_pages > len(self.free_pages):
            return None
        out = torch.empty((extend_num_tokens,), dtype=torch.int64, device=self.device)
        ops.alloc_extend(prefix_lens, seq_lens, last_loc, self.free_pages, out, self.page_size, self.status)
        self.free_pages = self.free_pages[num_new_pages:]
        return out

    def alloc_decode(self, seq_lens, seq_lens_cpu, last_loc) -> Optional[torch.Tensor]:
        bs = len(seq_lens)
        if bs > len(self.free_pages):
            self.merge_and_sort_free()
        num_new_pages = get_num_new_pages(seq_lens_cpu, self.page_size, decode=True)
        if num_new_pages > len(self.free_pages):
            return None
        out = torch.empty((bs,), dtype=torch.int64, device=self.device)
        ops.alloc_decode(seq_lens, last_loc, self.free_pages, out, self.page_size, self.status)
        self.free_pages = self.free_pages[num_new_pages:]
        return out

    def free(self, free_index: torch.Tensor) -> None:
        if free_index.numel() == 0:
            return
        self.release_pages = torch.cat((torch.unique(free_index // self.page_size), self.release_pages))
