"""Host-side mirror of the reference types the attention path exchanges.

These are *not* a re-implementation of SGLang: they are the minimum data
carriers a caller needs to drive ``B200AttnBackend`` without SGLang installed
(tests, bench, smoke) — same names, fields and semantics as the reference so the
backend is duck-type compatible with the real objects:

  ForwardMode            python/sglang/srt/model_executor/forward_batch_info.py:100-198
  ForwardBatch           python/sglang/srt/model_executor/forward_batch_info.py:379-604
  compute_position       python/sglang/srt/model_executor/forward_batch_info.py:1771-1808
  AttentionType          python/sglang/srt/layers/radix_attention.py:77-88
  RadixAttention         python/sglang/srt/layers/radix_attention.py:91-287
  ReqToTokenPool         python/sglang/srt/mem_cache/memory_pool.py:256-332
  KVWriteLoc             python/sglang/srt/mem_cache/memory_pool.py:1550-1590
  MHATokenToKVPool       python/sglang/srt/mem_cache/memory_pool.py:1759-2456
  SWAKVPool              python/sglang/srt/mem_cache/swa_memory_pool.py:20-270
  alloc_extend_naive     python/sglang/srt/mem_cache/allocator/paged.py:44-102
"""
from __future__ import annotations

import dataclasses
from enum import Enum, IntEnum, auto
from typing import List, Optional, Sequence, Tuple

import torch

from . import ops


class ForwardMode(IntEnum):
    EXTEND = auto()
    DECODE = auto()
    MIXED = auto()
    IDLE = auto()
    TARGET_VERIFY = auto()
    DRAFT_EXTEND_V2 = auto()
    PREBUILT = auto()
    SPLIT_PREFILL = auto()
    DLLM_EXTEND = auto()

    def is_extend(self, include_draft_extend_v2: bool = False) -> bool:
        return (
            self == ForwardMode.EXTEND
            or self == ForwardMode.MIXED
            or (include_draft_extend_v2 and self == ForwardMode.DRAFT_EXTEND_V2)
            or self == ForwardMode.TARGET_VERIFY
            or self == ForwardMode.SPLIT_PREFILL
            or self == ForwardMode.DLLM_EXTEND
        )

    is_prefill = is_extend

    def is_decode(self) -> bool:
        return self == ForwardMode.DECODE

    def is_mixed(self) -> bool:
        return self == ForwardMode.MIXED

    def is_idle(self) -> bool:
        return self == ForwardMode.IDLE

    def is_decode_or_idle(self) -> bool:
        return self in (ForwardMode.DECODE, ForwardMode.IDLE)

    def is_target_verify(self) -> bool:
        return self == ForwardMode.TARGET_VERIFY

    def is_draft_extend_v2(self) -> bool:
        return self == ForwardMode.DRAFT_EXTEND_V2

    def is_cuda_graph(self) -> bool:
        return self in (ForwardMode.DECODE, ForwardMode.TARGET_VERIFY, ForwardMode.IDLE, ForwardMode.DLLM_EXTEND)


class AttentionType(Enum):
    DECODER = "decoder"
    DECODER_BIDIRECTIONAL = "decoder_bidirectional"
    ENCODER_ONLY = "encoder_only"


@dataclasses.dataclass
class KVWriteLoc:
    loc: torch.Tensor
    swa_loc: Optional[torch.Tensor] = None
    full_loc: Optional[torch.Tensor] = None


class ReqToTokenPool:
    """``req_to_token: int32[size + 1, max_context_len]``; row 0 is the padding row."""

    def __init__(self, size: int, max_context_len: int, device: str):
        self.size = size
        self.max_context_len = max_context_len
        self.device = device
        self.req_to_token = torch.zeros((size + 1, max_context_len), dtype=torch.int32, device=device)
        self.free_slots = list(range(1, size + 1))

    def write(self, indices, values):
        self.req_to_token[indices] = values

    def available_size(self) -> int:
        return len(self.free_slots)

    def alloc(self, n: int = 1) -> Optional[List[int]]:
        if n > len(self.free_slots):
            return None
        out, self.free_slots = self.free_slots[:n], self.free_slots[n:]
        return out

    def free(self, idx) -> None:
        self.free_slots.extend([idx] if isinstance(idx, int) else list(idx))

    def clear(self) -> None:
        self.free_slots = list(range(1, self.size + 1))


class MHATokenToKVPool:
    """Per-layer ``k_buffer[l], v_buffer[l]: [size + page_size, Hkv, D]`` (NHD), zero-initialised.

    ``set_kv_buffer`` is the store boundary: on CUDA it launches the sm_100a
    ``kv_store`` kernel (no torch index_put), exactly where the reference launches
    its JIT ``store_cache`` (memory_pool.py:141-192).  ``skip_padding_slot``
    selects the reference CUDA behaviour (rows with loc == 0 are skipped,
    kvcache.cuh:215); ``False`` reproduces the naive ``k_cache[loc] = k`` branch.
    """

    def __init__(self, size: int, page_size: int, dtype: torch.dtype, head_num: int, head_dim: int,
                 layer_num: int, device: str, start_layer: int = 0, v_head_dim: Optional[int] = None,
                 skip_padding_slot: bool = True):
        self.size = size
        self.page_size = page_size
        self.dtype = dtype
        # float8 pools are stored as bytes and viewed on read (memory_pool.py:2036-2040, 2292-2329)
        self.store_dtype = torch.uint8 if dtype in (torch.float8_e4m3fn, torch.float8_e5m2) else dtype
        self.head_num = head_num
        self.head_dim = head_dim
        self.v_head_dim = head_dim if v_head_dim is None else v_head_dim
        self.layer_num = layer_num
        self.device = device
        self.start_layer = start_layer
        self.skip_padding_slot = skip_padding_slot
        self.layer_transfer_counter = None
        n = size + page_size
        sd = self.store_dtype
        self.k_buffer = [torch.zeros((n, head_num, head_dim), dtype=sd, device=device) for _ in range(layer_num)]
        self.v_buffer = [torch.zeros((n, head_num, self.v_head_dim), dtype=sd, device=device) for _ in range(layer_num)]

    def get_key_buffer(self, layer_id: int) -> torch.Tensor:
        b = self.k_buffer[layer_id - self.start_layer]
        return b.view(self.dtype) if self.store_dtype != self.dtype else b

    def get_value_buffer(self, layer_id: int) -> torch.Tensor:
        b = self.v_buffer[layer_id - self.start_layer]
        return b.view(self.dtype) if self.store_dtype != self.dtype else b

    def get_kv_buffer(self, layer_id: int) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def set_kv_buffer(self, layer, loc, cache_k: torch.Tensor, cache_v: torch.Tensor,
                      k_scale: Optional[float] = None, v_scale: Optional[float] = None) -> None:
        if isinstance(loc, KVWriteLoc) or hasattr(loc, "loc"):
            loc = loc.loc
        layer_id = layer.layer_id if hasattr(layer, "layer_id") else int(layer)
        if self.dtype == torch.float8_e4m3fn and cache_k.dtype != self.dtype:  # memory_pool.py:2365-2375
            ops.store_cache_fp8(
                cache_k, cache_v, self.get_key_buffer(layer_id), self.get_value_buffer(layer_id), loc,
                None if k_scale is None else float(k_scale), None if v_scale is None else float(v_scale),
                reserved_skip_index=0 if self.skip_padding_slot else -1,
            )
            return
        if k_scale is not None or v_scale is not None:
            raise NotImplementedError("b200 KV pool: k_scale / v_scale need a float8_e4m3fn pool")
        if cache_k.dtype != self.dtype:  # memory_pool.py:2365-2371
            cache_k = cache_k.to(self.dtype)
            cache_v = cache_v.to(self.dtype)
        ops.store_cache(
            cache_k, cache_v, self.get_key_buffer(layer_id), self.get_value_buffer(layer_id), loc,
            reserved_skip_index=0 if self.skip_padding_slot else -1,
        )


class SWAKVPool:
    """Hybrid full / sliding-window cache: two MHA pools, the SWA one smaller and addressed through
    ``full_to_swa_index_mapping`` (swa_memory_pool.py:20-270).  ``layers_mapping[layer_id] = (index in its pool,
    is_swa_layer)``.  The attention backend translates slot ids once per forward (``KVWriteLoc.swa_loc``)."""

    def __init__(self, size: int, size_swa: int, page_size: int, dtype: torch.dtype, head_num: int, head_dim: int,
                 swa_attention_layer_ids: List[int], full_attention_layer_ids: List[int], device: str):
        self.size, self.size_swa, self.page_size, self.dtype, self.device = size, size_swa, page_size, dtype, device
        self.head_num, self.head_dim = head_num, head_dim
        self.full_kv_pool = MHATokenToKVPool(size, page_size, dtype, head_num, head_dim,
                                             len(full_attention_layer_ids), device)
        self.swa_kv_pool = MHATokenToKVPool(size_swa, page_size, dtype, head_num, head_dim,
                                            len(swa_attention_layer_ids), device)
        self.layers_mapping = {}
        for i, lid in enumerate(full_attention_layer_ids):
            self.layers_mapping[lid] = (i, False)
        for i, lid in enumerate(swa_attention_layer_ids):
            self.layers_mapping[lid] = (i, True)
        self.full_to_swa_index_mapping: Optional[torch.Tensor] = None  # int64 [size + page_size + 1], set by the allocator
        self.start_layer = 0
        self.layer_transfer_counter = None

    def _pool(self, layer_id: int):
        idx, is_swa = self.layers_mapping[layer_id]
        return (self.swa_kv_pool if is_swa else self.full_kv_pool), idx

    def get_key_buffer(self, layer_id: int) -> torch.Tensor:
        pool, idx = self._pool(layer_id)
        return pool.get_key_buffer(idx)

    def get_value_buffer(self, layer_id: int) -> torch.Tensor:
        pool, idx = self._pool(layer_id)
        return pool.get_value_buffer(idx)

    def get_kv_buffer(self, layer_id: int) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def translate_loc_from_full_to_swa(self, kv_indices: torch.Tensor) -> torch.Tensor:
        return ops.translate_slots(kv_indices, self.full_to_swa_index_mapping)

    def set_kv_buffer(self, layer, loc_info, cache_k, cache_v, k_scale=None, v_scale=None) -> None:
        pool, idx = self._pool(layer.layer_id)
        loc = loc_info
        if hasattr(loc_info, "loc"):
            loc = loc_info.swa_loc if self.layers_mapping[layer.layer_id][1] else loc_info.loc
            assert loc is not None, "SWA layers need KVWriteLoc.swa_loc"
        pool.set_kv_buffer(idx, loc, cache_k, cache_v, k_scale, v_scale)


class RadixAttention:
    """Attribute carrier + dispatcher with the reference layer's field names."""

    def __init__(self, num_heads: int, head_dim: int, scaling: float, num_kv_heads: int, layer_id: int,
                 logit_cap: float = 0.0, v_head_dim: int = -1, sliding_window_size: int = -1,
                 is_cross_attention: bool = False, attn_type: AttentionType = AttentionType.DECODER):
        self.tp_q_head_num = num_heads
        self.tp_k_head_num = num_kv_heads
        self.tp_v_head_num = num_kv_heads
        self.head_dim = head_dim
        self.qk_head_dim = head_dim
        self.v_head_dim = v_head_dim if v_head_dim != -1 else head_dim
        self.scaling = scaling
        self.layer_id = layer_id
        self.logit_cap = logit_cap
        self.sliding_window_size = sliding_window_size or -1
        self.is_cross_attention = is_cross_attention
        self.attn_type = attn_type
        self.k_scale = None
        self.v_scale = None
        self.k_scale_float = None  # radix_attention.py: python floats of the scales (fp8 KV descale)
        self.v_scale_float = None

    def forward(self, q, k, v, forward_batch, save_kv_cache: bool = True, **kwargs):
        # radix_attention.py:160-167: k/v are viewed [T, Hkv, D] before the backend call
        if k is not None:
            k = k.view(-1, self.tp_k_head_num, self.qk_head_dim)
            v = v.view(-1, self.tp_v_head_num, self.v_head_dim)
        return forward_batch.attn_backend.forward(q, k, v, self, forward_batch, save_kv_cache, **kwargs)

    __call__ = forward


@dataclasses.dataclass
class ForwardBatch:
    """The fields of the reference ForwardBatch that attention backends consume."""

    forward_mode: ForwardMode
    batch_size: int
    req_pool_indices: torch.Tensor
    seq_lens: torch.Tensor
    out_cache_loc: torch.Tensor
    seq_lens_sum: int = 0
    seq_lens_cpu: Optional[torch.Tensor] = None
    positions: Optional[torch.Tensor] = None
    extend_num_tokens: Optional[int] = None
    extend_seq_lens: Optional[torch.Tensor] = None
    extend_prefix_lens: Optional[torch.Tensor] = None
    extend_start_loc: Optional[torch.Tensor] = None
    extend_seq_lens_cpu: Optional[List[int]] = None
    extend_prefix_lens_cpu: Optional[List[int]] = None
    mrope_positions: Optional[torch.Tensor] = None
    encoder_lens: Optional[torch.Tensor] = None
    spec_info: Optional[object] = None
    req_to_token_pool: Optional[ReqToTokenPool] = None
    token_to_kv_pool: Optional[MHATokenToKVPool] = None
    attn_backend: Optional[object] = None
    _attn_output: Optional[torch.Tensor] = None


def compute_position(extend_prefix_lens: Sequence[int], extend_seq_lens: Sequence[int],
                     device: str = "cpu") -> Tuple[torch.Tensor, torch.Tensor]:
    """positions[t] = prefix_len + j ; extend_start_loc = exclusive cumsum(extend_seq_lens).

    Follows compute_position_torch (forward_batch_info.py:1790-1808)."""
    pos = torch.cat(
        [torch.arange(p, p + e, dtype=torch.int64) for p, e in zip(extend_prefix_lens, extend_seq_lens)]
        or [torch.zeros(0, dtype=torch.int64)]
    )
    start = torch.zeros(len(extend_seq_lens), dtype=torch.int32)
    if len(extend_seq_lens) > 1:
        start[1:] = torch.cumsum(torch.tensor(list(extend_seq_lens[:-1]), dtype=torch.int32), 0)
    return pos.to(device), start.to(device)


def alloc_extend_naive(prefix_lens: Sequence[int], seq_lens: Sequence[int], last_loc: Sequence[int],
                       free_pages: List[int], page_size: int) -> torch.Tensor:
    """Pure-python restatement of the paged allocator's extend rule
    (allocator/paged.py:44-102): fill the tail of the last partial page
    (``last_loc + 1 ...``), then whole fresh pages, then a final partial page.
    Returns the flat out_cache_loc (int64) and pops pages from ``free_pages``."""
    out: List[int] = []
    for pre, seq, last in zip(prefix_lens, seq_lens, last_loc):
        n = seq - pre
        pos = pre
        # 1. tail of the current partial page
        while n > 0 and pos % page_size != 0:
            last += 1
            out.append(last)
            pos += 1
            n -= 1
        # 2./3. fresh pages
        while n > 0:
            page = free_pages.pop(0)
            take = min(page_size, n)
            out.extend(range(page * page_size, page * page_size + take))
            last = page * page_size + take - 1
            pos += take
            n -= take
    return torch.tensor(out, dtype=torch.int64)
