"""Device-side radix prefix cache (the lookup side of the reference's RadixCache) for the attention path.

Mirrors, for batches of requests, the three calls the scheduler makes on
python/sglang/srt/mem_cache/radix_cache.py::RadixCache:

  match_prefix(MatchPrefixParams(key))  :377-423  ->  DeviceRadixCache.match_prefix(keys)   (device_indices per request)
  insert(InsertParams(key, value))      :425-445  ->  DeviceRadixCache.insert(keys, values) (prefix_len per request)
  evict(...) -> allocator.free(values)  :593-621  ->  DeviceRadixCache.free(slots)          (the LRU choice of WHICH
                                                      leaves to evict stays the scheduler's; it hands over their slots)

The tree lives in HBM as a page trie in a hash table (csrc/radix.cu); keys are truncated to whole pages like
``RadixKey.page_aligned``.  Results are bit-identical to the reference on the fixtures generated from it
(tests/test_radix_gpu.py).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import I32, I64, check


def _tok_dtype(t: torch.Tensor) -> int:
    if t.dtype == torch.int32:
        return I32
    if t.dtype == torch.int64:
        return I64
    raise TypeError(f"radix: token / slot tensors must be int32 or int64, got {t.dtype}")


def _flatten(keys: Sequence[torch.Tensor], device) -> Tuple[torch.Tensor, torch.Tensor]:
    lens = [int(k.numel()) for k in keys]
    indptr = torch.zeros(len(keys) + 1, dtype=torch.int64)
    if lens:
        indptr[1:] = torch.cumsum(torch.tensor(lens, dtype=torch.int64), 0)
    parts = [k.reshape(-1).to(device=device, dtype=torch.int64, non_blocking=True) for k in keys]
    flat = torch.cat(parts + [torch.zeros(1, dtype=torch.int64, device=device)])  # never empty
    return flat, indptr.to(device, non_blocking=True)


class DeviceRadixCache:
    """``max_cached_pages`` bounds the number of cached pages (entries); ``num_kv_pages`` = pages of the KV pool."""

    def __init__(self, page_size: int, max_cached_pages: int, num_kv_pages: int, device: str = "cuda"):
        if not torch.cuda.is_available():
            raise AssertionError("DeviceRadixCache needs a CUDA device; there is no CPU fallback")
        _lib.load()
        self.page_size = int(page_size)
        self.device = torch.device(device)
        cap = int(max_cached_pages) + 1  # entry 0 is the root
        tcap = 1
        while tcap < 4 * cap:
            tcap *= 2
        dev = self.device
        self.parent = torch.zeros(cap, dtype=torch.int32, device=dev)
        self.tokens = torch.zeros(cap * self.page_size, dtype=torch.int32, device=dev)
        self.values = torch.zeros(cap * self.page_size, dtype=torch.int64, device=dev)
        self.ehash = torch.zeros(cap, dtype=torch.int64, device=dev)
        self.tkeys = torch.zeros(tcap, dtype=torch.int64, device=dev)
        self.tvals = torch.zeros(tcap, dtype=torch.int32, device=dev)
        self.rev = torch.full((int(num_kv_pages),), -1, dtype=torch.int32, device=dev)
        self.freelist = torch.zeros(cap, dtype=torch.int32, device=dev)
        self.ctrs = torch.tensor([1, 0, 0, 0], dtype=torch.int32, device=dev)
        self._table = _lib.RadixTable(
            self.parent.data_ptr(), self.tokens.data_ptr(), self.values.data_ptr(), self.ehash.data_ptr(),
            self.tkeys.data_ptr(), self.tvals.data_ptr(), self.rev.data_ptr(), self.freelist.data_ptr(),
            self.ctrs.data_ptr(), cap, tcap, self.page_size, int(num_kv_pages))

    def _stream(self) -> int:
        return torch.cuda.current_stream().cuda_stream

    # ------------------------------------------------------------------ flat (sync-free) interface
    def match_prefix_flat(self, tokens: torch.Tensor, indptr: torch.Tensor, out_indices: torch.Tensor,
                          out_len: torch.Tensor) -> None:
        """``tokens`` flat int32/int64, ``indptr`` int64 [B+1] (device); fills ``out_indices`` int64 [B, stride] and
        ``out_len`` int32 [B] (matched tokens per request).  No allocation, no sync."""
        B = indptr.numel() - 1
        rc = _lib.load().b200attn_radix_match(C.byref(self._table), tokens.data_ptr(), _tok_dtype(tokens),
                                              indptr.data_ptr(), B, out_indices.data_ptr(), out_indices.stride(0),
                                              out_len.data_ptr(), self._stream())
        check(rc, "b200attn_radix_match")

    def insert_flat(self, tokens: torch.Tensor, indptr: torch.Tensor, values: torch.Tensor,
                    out_prefix_len: torch.Tensor) -> None:
        B = indptr.numel() - 1
        if values.dtype != torch.int64 or values.numel() != tokens.numel():
            raise ValueError("radix insert: values must be int64, one KV slot per token")
        rc = _lib.load().b200attn_radix_insert(C.byref(self._table), tokens.data_ptr(), _tok_dtype(tokens),
                                               indptr.data_ptr(), values.data_ptr(), B, out_prefix_len.data_ptr(),
                                               self._stream())
        check(rc, "b200attn_radix_insert")

    # ------------------------------------------------------------------ reference-shaped interface
    def match_prefix(self, keys: Sequence[torch.Tensor]) -> List[torch.Tensor]:
        """Per request: int64 KV slots of the longest cached prefix (``MatchResult.device_indices``)."""
        B = len(keys)
        if B == 0:
            return []
        flat, indptr = _flatten(keys, self.device)
        width = max(max(int(k.numel()) for k in keys), 1)
        out = torch.empty((B, width), dtype=torch.int64, device=self.device)
        out_len = torch.empty(B, dtype=torch.int32, device=self.device)
        self.match_prefix_flat(flat, indptr, out, out_len)
        lens = out_len.tolist()
        return [out[b, :lens[b]] for b in range(B)]

    def insert(self, keys: Sequence[torch.Tensor], values: Sequence[torch.Tensor]) -> List[int]:
        """Per request: tokens that were already cached (``InsertResult.prefix_len``)."""
        B = len(keys)
        if B == 0:
            return []
        flat, indptr = _flatten(keys, self.device)
        vflat, _ = _flatten(values, self.device)
        out = torch.empty(B, dtype=torch.int32, device=self.device)
        self.insert_flat(flat, indptr, vflat, out)
        res = out.tolist()
        if int(self.ctrs[3]) != 0:
            raise RuntimeError("DeviceRadixCache: out of entries (raise max_cached_pages)")
        return res

    def free(self, slots: torch.Tensor) -> None:
        """KV slots released by the allocator for evicted leaves (``allocator.free(node.value)``)."""
        if slots.numel() == 0:
            return
        slots = slots.to(self.device).contiguous()
        rc = _lib.load().b200attn_radix_erase_slots(C.byref(self._table), slots.data_ptr(), _tok_dtype(slots),
                                                    slots.numel(), self._stream())
        check(rc, "b200attn_radix_erase_slots")

    def cached_pages(self) -> int:
        return int(self.ctrs[2])
