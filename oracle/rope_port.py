"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference rotary layer (+ fused KV store).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; nothing under sglang_b200/ does.
Parity PINNED: tests/golden/rope_*.pt are outputs of the REAL reference classes (oracle/gen_golden_rope.py).

Follows
  python/sglang/srt/layers/rotary_embedding/base.py:153-182 (cos_sin_cache), :236-276 (forward_native)
  python/sglang/srt/layers/rotary_embedding/utils.py:36-62  (apply_rotary_emb)
  python/sglang/srt/layers/rotary_embedding/mrope.py:47-51,173-217 (M-RoPE select)
  python/sglang/kernels/jit/csrc/elementwise/rope.cuh:56-110 (fp32 arithmetic of the production kernel)
  python/sglang/srt/mem_cache/memory_pool.py:190-192 (naive store)
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import numpy as np
import torch


def cos_sin_cache(rotary_dim: int, max_position: int, base: float) -> torch.Tensor:
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


def select_cos_sin(cache: torch.Tensor, positions: torch.Tensor, mrope_section: Optional[Sequence[int]],
                   interleaved: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    """cos, sin: [T, rot/2] (mrope.py:183-199)."""
    cos, sin = cache[positions].chunk(2, dim=-1)
    if positions.ndim == 2:
        sec = list(mrope_section)
        if interleaved:
            def pick(x):
                x_t = x[0].clone()
                x_t[..., 1:sec[1] * 3:3] = x[1, ..., 1:sec[1] * 3:3]
                x_t[..., 2:sec[2] * 3:3] = x[2, ..., 2:sec[2] * 3:3]
                return x_t
            cos, sin = pick(cos), pick(sin)
        else:
            cos = torch.cat([m[i] for i, m in enumerate(cos.split(sec, dim=-1))], dim=-1)
            sin = torch.cat([m[i] for i, m in enumerate(sin.split(sec, dim=-1))], dim=-1)
    return cos, sin


def _pairs(x: torch.Tensor, neox: bool):
    if neox:
        return torch.chunk(x, 2, dim=-1)
    return x[..., ::2], x[..., 1::2]


def _unpairs(o1: torch.Tensor, o2: torch.Tensor, neox: bool) -> torch.Tensor:
    if neox:
        return torch.cat((o1, o2), dim=-1)
    return torch.stack((o1, o2), dim=-1).flatten(-2)


def rope_native(positions, x, cache, head_size, rotary_dim, neox, mrope_section=None, interleaved=False):
    """torch arithmetic in x's dtype, as forward_native does it.  x: [T, H*head_size] -> same shape."""
    T = x.shape[0]
    cos, sin = select_cos_sin(cache, positions, mrope_section, interleaved)
    xs = x.view(T, -1, head_size)
    rot, pas = xs[..., :rotary_dim], xs[..., rotary_dim:]
    c, s = cos.unsqueeze(-2).to(x.dtype), sin.unsqueeze(-2).to(x.dtype)
    x1, x2 = _pairs(rot, neox)
    o = _unpairs(x1 * c - x2 * s, x2 * c + x1 * s, neox)
    return torch.cat((o, pas), dim=-1).reshape(x.shape)


def rope_f32(positions, x, cache, head_size, rotary_dim, neox, mrope_section=None, interleaved=False):
    """fp32 products and sum (numpy: no FMA contraction), one rounding to x's dtype."""
    T = x.shape[0]
    cos, sin = select_cos_sin(cache.float(), positions, mrope_section, interleaved)
    xs = x.view(T, -1, head_size)
    rot, pas = xs[..., :rotary_dim].float(), xs[..., rotary_dim:]
    c, s = cos.unsqueeze(-2).numpy(), sin.unsqueeze(-2).numpy()
    x1, x2 = _pairs(rot, neox)
    x1, x2 = np.ascontiguousarray(x1.numpy()), np.ascontiguousarray(x2.numpy())
    o1 = (x1 * c).astype(np.float32) - (x2 * s).astype(np.float32)
    o2 = (x2 * c).astype(np.float32) + (x1 * s).astype(np.float32)
    o = _unpairs(torch.from_numpy(o1), torch.from_numpy(o2), neox).to(x.dtype)
    return torch.cat((o, pas), dim=-1).reshape(x.shape)


def rope_store(positions, q, k, v, k_cache, v_cache, loc, cache, head_size, rotary_dim, neox, mrope_section=None,
               interleaved=False, native=True, skip_index: Optional[int] = None):
    """Returns rotated (q, k); scatters rotated k and v into clones of the pool which are returned too."""
    f = rope_native if native else rope_f32
    q2 = f(positions, q, cache, head_size, rotary_dim, neox, mrope_section, interleaved)
    k2 = f(positions, k, cache, head_size, rotary_dim, neox, mrope_section, interleaved)
    kc, vc = k_cache.clone(), v_cache.clone()
    keep = torch.ones_like(loc, dtype=torch.bool) if skip_index is None else loc != skip_index
    kc[loc[keep]] = k2.view(k2.shape[0], *kc.shape[1:])[keep]
    vc[loc[keep]] = v.view(v.shape[0], *vc.shape[1:])[keep]
    return q2, k2, kc, vc
