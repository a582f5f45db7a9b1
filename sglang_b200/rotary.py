"""Host-side mirror of the reference rotary layer for the fused RoPE + KV-store step.

``RotaryEmbedding`` / ``MRotaryEmbedding`` keep the reference constructor arguments and call convention
(python/sglang/srt/layers/rotary_embedding/base.py:78-143,236-276,365-436; mrope.py:54-136,173-251):
``rope(positions, query, key, fused_set_kv_buffer_arg=None) -> (query, key)``, rotating IN PLACE like the
reference CUDA path.  All arithmetic runs in ``b200attn_rope_store``; there is no torch fallback.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch

from . import ops


@dataclass
class FusedSetKVBufferArg:
    """python/sglang/kernels/ops/attention/rope.py:98-113."""

    value: torch.Tensor
    k_buffer: torch.Tensor
    v_buffer: torch.Tensor
    cache_loc: torch.Tensor
    k_scale: Optional[float] = None  # e4m3 pools only (memory_pool.py:2365-2371)
    v_scale: Optional[float] = None


def compute_cos_sin_cache(rotary_dim: int, max_position: int, base: float) -> torch.Tensor:
    """fp32 ``[max_position, rotary_dim]`` = (cos | sin); base.py:153-182 (computed on the host like the
    reference's HF-exact path, then moved)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


class RotaryEmbedding:
    def __init__(self, head_size: int, rotary_dim: int, max_position_embeddings: int, base: float,
                 is_neox_style: bool, dtype: torch.dtype, device: str = "cuda", native_rounding: bool = False):
        self.head_size, self.rotary_dim = head_size, rotary_dim
        self.max_position_embeddings, self.base = max_position_embeddings, base
        self.is_neox_style, self.dtype = is_neox_style, dtype
        # fp32 cache as on CUDA in the reference (base.py:107-109); native_rounding reproduces forward_native
        self.native_rounding = native_rounding
        self.cos_sin_cache = compute_cos_sin_cache(rotary_dim, max_position_embeddings, base).to(device)
        self.mrope_section: Optional[List[int]] = None
        self.mrope_interleaved = False

    def __call__(self, positions, query, key, offsets=None, fused_set_kv_buffer_arg=None):
        return self.forward(positions, query, key, offsets, fused_set_kv_buffer_arg)

    def forward(self, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                offsets: Optional[torch.Tensor] = None,
                fused_set_kv_buffer_arg: Optional[FusedSetKVBufferArg] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        if offsets is not None:
            positions = positions + offsets
        f = fused_set_kv_buffer_arg
        if positions.dim() == 2 and not self.mrope_section:
            raise AssertionError("2D positions (multimodal RoPE) need an MRotaryEmbedding")  # base.py:216-219
        ops.apply_rope_with_cos_sin_cache_inplace(
            positions, query, key, self.cos_sin_cache, head_size=self.head_size, is_neox=self.is_neox_style,
            rope_dim=self.rotary_dim, mrope_section=self.mrope_section if positions.dim() == 2 else None,
            mrope_interleaved=self.mrope_interleaved, v=None if f is None else f.value,
            k_buffer=None if f is None else f.k_buffer, v_buffer=None if f is None else f.v_buffer,
            cache_loc=None if f is None else f.cache_loc, native_rounding=self.native_rounding,
            k_scale=None if f is None else f.k_scale, v_scale=None if f is None else f.v_scale)
        return query, key


class MRotaryEmbedding(RotaryEmbedding):
    """Sectioned / interleaved multimodal RoPE (mrope.py:54-136).  The GLM axis-map variant is not on the path."""

    def __init__(self, head_size: int, rotary_dim: int, max_position_embeddings: int, base: float,
                 is_neox_style: bool, dtype: torch.dtype, mrope_section: Optional[Sequence[int]] = None,
                 mrope_interleaved: bool = False, mrope_interleaved_glm: bool = False, device: str = "cuda",
                 native_rounding: bool = False):
        super().__init__(head_size, rotary_dim, max_position_embeddings, base, is_neox_style, dtype, device,
                         native_rounding)
        if mrope_interleaved_glm:
            raise NotImplementedError("b200 rotary: mrope_interleaved_glm axis map")
        self.mrope_section = list(mrope_section) if mrope_section else None
        if self.mrope_section and sum(self.mrope_section) != rotary_dim // 2:
            raise ValueError(f"mrope_section {self.mrope_section} must sum to rotary_dim // 2 = {rotary_dim // 2}")
        self.mrope_interleaved = mrope_interleaved


def mrope_rope_index(input_ids: torch.Tensor, image_grid_thw: Optional[torch.Tensor], *, image_token_id: int,
                     vision_start_token_id: int, spatial_merge_size: int = 2) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-request M-RoPE positions ``int64[3, L]`` and ``mrope_position_delta`` for a Qwen2-VL / Qwen2.5-VL prompt
    of text runs and still-image blocks (``get_rope_index``, python/sglang/srt/layers/rotary_embedding/
    mrope_rope_index.py:46-230, image branch): a text run advances all three axes together from ``max + 1`` of what
    came before; an image block of grid ``(t, h, w)`` (``h, w`` divided by ``spatial_merge_size``) holds the t axis
    at its frame index (0 for still images in both model types) and walks the h / w axes over the grid, all offset
    by the block's start.  ``delta = max position + 1 - L``."""
    ids = input_ids.tolist()
    L = len(ids)
    out, st, img, nxt = [], 0, 0, 0
    starts = [i for i, t in enumerate(ids) if t == vision_start_token_id and i + 1 < L and ids[i + 1] == image_token_id]
    for _ in starts:
        ed = ids.index(image_token_id, st)
        t, h, w = (int(x) for x in image_grid_thw[img])
        img += 1
        gh, gw = h // spatial_merge_size, w // spatial_merge_size
        text_len = ed - st
        out.append(torch.arange(text_len, dtype=torch.int64).view(1, -1).expand(3, -1) + nxt)
        ti = torch.arange(t, dtype=torch.int64).view(-1, 1, 1).expand(t, gh, gw).reshape(-1)
        if t == 1:
            ti = torch.zeros_like(ti)
        hi = torch.arange(gh, dtype=torch.int64).view(1, -1, 1).expand(t, gh, gw).reshape(-1)
        wi = torch.arange(gw, dtype=torch.int64).view(1, 1, -1).expand(t, gh, gw).reshape(-1)
        out.append(torch.stack([ti, hi, wi]) + text_len + nxt)
        nxt = int(out[-1].max()) + 1
        st = ed + t * gh * gw
    if st < L:
        out.append(torch.arange(L - st, dtype=torch.int64).view(1, -1).expand(3, -1) + nxt)
    pos = torch.cat(out, dim=1) if out else torch.zeros((3, 0), dtype=torch.int64)
    delta = torch.tensor([[int(pos.max()) + 1 - L if L else 0]], dtype=torch.int64)
    return pos, delta


def _expand_mrope_from_input(mm_positions: Optional[torch.Tensor], delta, seq_len: int) -> torch.Tensor:
    """ONE position column for a request that is past its precomputed table (forward_batch_info.py:1103-1127)."""
    if mm_positions is not None and mm_positions.shape[1] >= seq_len:
        return mm_positions[:, seq_len - 1:seq_len].to(torch.int64)
    d = torch.as_tensor(delta, dtype=torch.int64).flatten()
    return (d - 1).unsqueeze(0).repeat(3, 1) + seq_len


def mrope_positions_extend(prefix_lens: Sequence[int], extend_lens: Sequence[int],
                           mm_positions: Sequence[Optional[torch.Tensor]],
                           deltas: Optional[Sequence] = None) -> torch.Tensor:
    """``int64[3, T]`` for an extend batch: text-only requests get ``arange(prefix, prefix+extend)`` on all
    three axes, multimodal ones the slice of their precomputed ``mrope_positions``; a request whose slice is empty
    (already past its table) takes the single delta-based column of ``_expand_mrope_from_input``, as the reference
    does (forward_batch_info.py:1213-1254)."""
    out = []
    for b, (pre, ext, mm) in enumerate(zip(prefix_lens, extend_lens, mm_positions)):
        if mm is None:
            out.append(torch.arange(pre, pre + ext, dtype=torch.int64).unsqueeze(0).repeat(3, 1))
            continue
        sl = mm[:, pre:pre + ext].to(torch.int64)
        if sl.numel() == 0:
            if deltas is None:
                raise ValueError("mrope_positions_extend: a request past its position table needs its delta")
            sl = _expand_mrope_from_input(mm, deltas[b], pre + ext)
        out.append(sl)
    return torch.cat(out, dim=1) if out else torch.zeros((3, 0), dtype=torch.int64)


def mrope_positions_decode(seq_lens: torch.Tensor, deltas: Optional[Sequence[int]] = None) -> torch.Tensor:
    """``(delta - 1) + seq_len`` on all three axes (forward_batch_info.py:1166-1184)."""
    p = seq_lens.to(torch.int64) - 1
    if deltas is not None:
        p = p + torch.as_tensor([int(torch.as_tensor(d).flatten()[0]) if d is not None else 0 for d in deltas],
                                dtype=torch.int64, device=seq_lens.device)
    return p.unsqueeze(0).repeat(3, 1)
