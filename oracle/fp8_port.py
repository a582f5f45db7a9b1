"""TEST INFRASTRUCTURE ONLY (oracle): CPU restatement of the reference's FP8 KV-cache path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.

Store (python/sglang/srt/mem_cache/memory_pool.py:2365-2375):
    if cache_k.dtype != self.dtype:
        if k_scale is not None: cache_k.div_(k_scale)
        cache_k = cache_k.to(self.dtype)             # float8_e4m3fn
    if self.store_dtype != self.dtype: cache_k = cache_k.view(self.store_dtype)   # uint8
    k_cache[loc] = cache_k                             # memory_pool.py:190-192 (naive branch)
Read side (python/sglang/srt/layers/attention/triton_backend.py:1418-1420 and the kernels behind it):
    logits = (q . k8) * sm_scale * k_descale;  out = (softmax . v8) * v_descale,  k_descale = layer.k_scale_float.

Pinning: `encode_e4m3_bits` is an independent bit-level restatement of c10's fp8e4m3fn_from_fp32_value
(round to nearest even, values that would round above 448 become the NaN pattern 0x7f, no saturation);
tests/test_fp8_oracle.py checks it against torch's own `.to(torch.float8_e4m3fn)` on ALL 65,536 bf16 bit patterns,
so the store oracle is pinned to the arithmetic the reference executes (torch), not to this file alone.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

FP8 = torch.float8_e4m3fn


def quantize(x: torch.Tensor, scale: Optional[float]) -> torch.Tensor:
    """memory_pool.py:2365-2375 for one tensor: div_ in the input dtype, cast to e4m3, uint8 view."""
    x = x.clone()
    if scale is not None:
        x.div_(scale)
    return x.to(FP8).view(torch.uint8)


def store_kv_fp8(k_cache_u8: torch.Tensor, v_cache_u8: torch.Tensor, loc: torch.Tensor, k: torch.Tensor,
                 v: torch.Tensor, k_scale: Optional[float], v_scale: Optional[float], skip_index: int = -1) -> None:
    """``k_cache[loc] = quantize(k, k_scale)`` (same for v); ``skip_index`` mirrors kvcache.cuh:215."""
    keep = loc != skip_index
    k_cache_u8[loc[keep]] = quantize(k, k_scale)[keep]
    v_cache_u8[loc[keep]] = quantize(v, v_scale)[keep]


def dequantize(cache_u8: torch.Tensor, descale: float = 1.0) -> torch.Tensor:
    """fp32 values of an e4m3 byte pool times the descale (what the attention kernels must reproduce)."""
    return cache_u8.view(FP8).float() * descale


def extend_attention_e4m3_arith(q: torch.Tensor, k_cache_u8: torch.Tensor, v_cache_u8: torch.Tensor,
                                req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens, extend_lens,
                                num_q_heads: int, num_kv_heads: int, head_dim: int, scaling: float, k_descale: float,
                                v_descale: float, causal: bool = True) -> torch.Tensor:
    """The ARITHMETIC of the reference's Triton extend kernel over an e4m3 pool, restated in torch
    (python/sglang/kernels/ops/attention/extend_attention.py: ``qk = tl.dot(q.to(k.dtype), k)`` :523/:1174 - the
    query is cast to e4m3; ``p = p.to(v.dtype)`` :596/:1248 - the probabilities are cast to e4m3 before the second
    contraction, the denominator sums the un-rounded p; k / v descales as in triton_backend.py:1418-1420).
    Pinned by tests/test_triton_golden.py against the kernel's recorded output (extend_fp8).
    q [T, Hq*D] -> fp32 [T, Hq*D]."""
    Hq, Hkv, D = num_q_heads, num_kv_heads, head_dim
    G = Hq // Hkv
    outs, t0 = [], 0
    for b, (L, E) in enumerate(zip([int(x) for x in seq_lens], [int(x) for x in extend_lens])):
        if E == 0:
            continue
        slots = req_to_token[int(req_pool_indices[b]), :L].long()
        qb = q[t0:t0 + E].view(E, Hq, D).float().to(FP8).float()
        kk = k_cache_u8[slots].view(FP8).float().view(L, Hkv, D).repeat_interleave(G, dim=1)
        vv = v_cache_u8[slots].view(FP8).float().view(L, Hkv, D).repeat_interleave(G, dim=1)
        s_ = torch.einsum("ehd,lhd->hel", qb, kk) * (scaling * k_descale)
        if causal:
            pos = torch.arange(E)[:, None] + (L - E)
            s_ = s_.masked_fill(torch.arange(L)[None, :] > pos, float("-inf"))
        p_ = torch.exp(s_ - s_.max(dim=-1, keepdim=True).values)
        den = p_.sum(-1, keepdim=True)
        o = torch.einsum("hel,lhd->ehd", p_.to(FP8).float(), vv) * v_descale / den.permute(1, 0, 2)
        outs.append(o.reshape(E, Hq * D))
        t0 += E
    return torch.cat(outs) if outs else torch.zeros((0, Hq * D))


def encode_e4m3_bits(x: np.ndarray) -> np.ndarray:
    """Independent fp32 -> e4m3fn encoder (uint8), restating c10/util/Float8_e4m3fn.h: RNE on the 3-bit
    mantissa, exponent bias 7, denormals below 2^-6 (step 2^-9), no infinities, |x| rounding above 448 -> 0x7f."""
    x = np.asarray(x, dtype=np.float32)
    bits = x.view(np.uint32)
    sign = ((bits >> 24) & 0x80).astype(np.uint8)
    a = np.abs(x).astype(np.float64)
    out = np.zeros(x.shape, dtype=np.uint8)
    nan_or_big = ~(a <= 464.0)          # NaN, inf and everything past the 448/480 midpoint (tie -> even = 448)
    a = np.where(nan_or_big, 0.0, a)
    den = a < 2.0 ** -6
    # denormal range: multiples of 2^-9, round half to even (np.rint)
    q_den = np.rint(a / 2.0 ** -9)
    with np.errstate(divide="ignore", invalid="ignore"):
        e = np.floor(np.log2(np.where(a > 0, a, 1.0)))
    e = np.clip(e, -6, 8)
    m = np.rint(a / 2.0 ** e * 8.0) - 8.0           # mantissa in eighths, may round up to 8
    carry = m >= 8
    e = np.where(carry, e + 1, e)
    m = np.where(carry, 0, m)
    norm = ((e + 7).astype(np.int64) << 3 | m.astype(np.int64)).astype(np.uint8)
    den_bits = q_den.astype(np.uint8)               # q_den == 8 is exactly the smallest normal (0x08)
    out = np.where(den, den_bits, norm)
    out = np.where(nan_or_big, np.uint8(0x7F), out)
    return (out | sign).astype(np.uint8)
