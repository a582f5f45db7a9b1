"""ORACLE (test infrastructure, not product code).

CPU restatement of the reference's attention hot path, i.e. SGLang's own
``torch_native`` backend plus the integer bookkeeping around it.  Only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this module; nothing under ``sglang_b200/``
does.  The floating-point arithmetic is ``torch.nn.functional.
scaled_dot_product_attention`` on CPU — the same third-party arithmetic the
reference calls (torch_native_backend.py:159-168, 261-270; the reference pins
torch==2.13.0 in python/pyproject.toml:82, this image has torch 2.11.0 — SDPA's
``is_causal`` / ``enable_gqa`` / ``scale`` semantics are the same).

Parity is PINNED: ``oracle/gen_golden.py`` imported the real reference
(``/root/reference/python``, TorchNativeAttnBackend + MHATokenToKVPool +
ReqToTokenPool driven through the reference's own MockModelRunner kit) and wrote
``tests/golden/*.pt``; ``tests/test_oracle_golden.py`` checks every function
below against those fixtures bit-for-bit (integer paths) / to 0 ulp (same SDPA).

Each function cites the reference lines it follows (paths relative to the
SGLang tree).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import numpy as np
import torch
from torch.nn.functional import scaled_dot_product_attention


# ----------------------------------------------------------------------------- a1
def store_kv(k_cache: torch.Tensor, v_cache: torch.Tensor, loc: torch.Tensor, k: torch.Tensor,
             v: torch.Tensor, skip_index: Optional[int] = None) -> None:
    """``k_cache[loc] = k; v_cache[loc] = v``.

    python/sglang/srt/mem_cache/memory_pool.py:190-192 (the naive branch the
    reference takes on CPU).  ``skip_index=0`` restates the CUDA JIT kernel's
    reserved-slot rule instead (python/sglang/kernels/jit/csrc/elementwise/
    kvcache.cuh:215): rows whose slot equals ``skip_index`` are not written.
    Dtype cast before the store follows memory_pool.py:2365-2371.
    """
    if k.dtype != k_cache.dtype:
        k = k.to(k_cache.dtype)
        v = v.to(v_cache.dtype)
    loc = loc.long()
    if skip_index is not None:
        keep = loc != skip_index
        loc, k, v = loc[keep], k[keep], v[keep]
    k_cache[loc] = k
    v_cache[loc] = v


# ----------------------------------------------------------------------------- a3
def kv_indptr_indices(req_to_token: np.ndarray, req_pool_indices: Sequence[int], kv_lens: Sequence[int],
                      kv_start: Optional[Sequence[int]] = None) -> Tuple[np.ndarray, np.ndarray]:
    """``kv_indptr = [0, cumsum(lens)]``; ``kv_indices = concat_b req_to_token[req_b, start_b : start_b+len_b]``.

    python/sglang/srt/layers/attention/triton_backend.py:438-456 and
    python/sglang/kernels/ops/kvcache/kv_indices.py:8-44; the python gather the
    reference's own test uses is test/registered/attention/test_create_kvindices.py:44-52.
    """
    req_to_token = np.asarray(req_to_token)
    lens = np.asarray(kv_lens, dtype=np.int64)
    indptr = np.zeros(len(lens) + 1, dtype=np.int32)
    indptr[1:] = np.cumsum(lens)
    out = np.empty(int(indptr[-1]), dtype=np.int64)
    for b, (r, n) in enumerate(zip(req_pool_indices, lens)):
        s = 0 if kv_start is None else int(kv_start[b])
        out[indptr[b]:indptr[b + 1]] = req_to_token[int(r), s:s + int(n)]
    return indptr, out


def qo_indptr(extend_seq_lens: Sequence[int]) -> np.ndarray:
    """``qo_indptr[1:] = cumsum(extend_seq_lens)`` (triton_backend.py:964-966)."""
    out = np.zeros(len(extend_seq_lens) + 1, dtype=np.int32)
    out[1:] = np.cumsum(np.asarray(extend_seq_lens, dtype=np.int64))
    return out


def page_table(req_to_token: np.ndarray, req_pool_indices: Sequence[int], seq_lens: Sequence[int],
               page_size: int, max_pages: int, fill: int = 0) -> np.ndarray:
    """``page_table[b, j] = req_to_token[req_b, j*page] // page`` for ``j < ceil(len_b/page)``.

    python/sglang/kernels/ops/kvcache/trtllm_mha_page_table.py:40-95; entries past
    the page count are not written by the reference (left at ``fill``).
    """
    req_to_token = np.asarray(req_to_token)
    out = np.full((len(seq_lens), max_pages), fill, dtype=np.int32)
    for b, (r, n) in enumerate(zip(req_pool_indices, seq_lens)):
        pages = min(-(-int(n) // page_size), max_pages)
        pos = np.arange(pages) * page_size
        out[b, :pages] = req_to_token[int(r), pos] // page_size
    return out


def positions(extend_prefix_lens: Sequence[int], extend_seq_lens: Sequence[int]) -> Tuple[np.ndarray, np.ndarray]:
    """``positions[t] = prefix + j``; ``extend_start_loc = exclusive cumsum(extend_seq_lens)``.

    python/sglang/srt/model_executor/forward_batch_info.py:1790-1808."""
    pos = np.concatenate([np.arange(p, p + e, dtype=np.int64) for p, e in zip(extend_prefix_lens, extend_seq_lens)]
                         or [np.zeros(0, dtype=np.int64)])
    start = np.zeros(len(extend_seq_lens), dtype=np.int32)
    start[1:] = np.cumsum(np.asarray(extend_seq_lens, dtype=np.int64))[:-1]
    return pos, start


# ----------------------------------------------------------------------------- a4
def decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
           req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, num_q_heads: int, num_kv_heads: int,
           head_dim: int, scaling: float) -> torch.Tensor:
    """One query row per request over ``seq_lens[b]`` gathered keys, non-causal.

    python/sglang/srt/layers/attention/torch_native_backend.py:176-277 (loop body
    :227-275) called from forward_decode :338-398.  ``q`` is ``[B, Hq*D]``; caches
    ``[slots, Hkv, D]``; returns ``[B, Hq*D]`` in q's dtype.
    """
    B = q.shape[0]
    out = torch.empty_like(q).view(B, num_q_heads, head_dim)
    query = q.view(B, num_q_heads, head_dim).movedim(0, 1)  # [Hq, B, D]
    enable_gqa = num_q_heads != num_kv_heads
    for i in range(B):
        n = int(seq_lens[i])
        tokens = req_to_token[int(req_pool_indices[i]), :n].long()
        key = k_cache[tokens].movedim(0, 1)      # [Hkv, n, D]
        value = v_cache[tokens].movedim(0, 1)
        qi = query[:, i:i + 1, :]
        if not (qi.dtype == key.dtype == value.dtype):
            key, value = key.to(qi.dtype), value.to(qi.dtype)
        o = scaled_dot_product_attention(qi.unsqueeze(0), key.unsqueeze(0), value.unsqueeze(0),
                                         enable_gqa=enable_gqa, scale=scaling, is_causal=False)
        out[i] = o.squeeze(0).movedim(1, 0)[0]
    return out.view(B, num_q_heads * head_dim)


# ----------------------------------------------------------------------------- a5
def extend(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
           req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, extend_prefix_lens: torch.Tensor,
           extend_seq_lens: torch.Tensor, num_q_heads: int, num_kv_heads: int, head_dim: int,
           scaling: float, causal: bool = True) -> torch.Tensor:
    """Prefix + new tokens, "redundant query" padded to ``seq_len_kv`` rows.

    python/sglang/srt/layers/attention/torch_native_backend.py:61-174: the real
    ``extend_len`` query rows sit at the tail of a ``seq_len_kv``-row query
    (:128-134), all keys come from the pool (:138-141), SDPA runs with
    ``is_causal=causal`` on the square problem (:159-168) and rows
    ``[prefix_len:]`` are kept (:172).  (The padded rows are uninitialised in the
    reference; zeros here — they never reach the output.)
    """
    T = q.shape[0]
    out = torch.empty_like(q).view(T, num_q_heads, head_dim)
    query = q.view(T, num_q_heads, head_dim).movedim(0, 1)  # [Hq, T, D]
    enable_gqa = num_q_heads != num_kv_heads
    start_q = 0
    for i in range(seq_lens.shape[0]):
        e, p, n = int(extend_seq_lens[i]), int(extend_prefix_lens[i]), int(seq_lens[i])
        end_q = start_q + e
        per_q = query[:, start_q:end_q, :]
        red = torch.zeros((num_q_heads, n, head_dim), dtype=q.dtype)
        red[:, p:, :] = per_q
        tokens = req_to_token[int(req_pool_indices[i]), :n].long()
        key = k_cache[tokens].movedim(0, 1)
        value = v_cache[tokens].movedim(0, 1)
        if not (red.dtype == key.dtype == value.dtype):
            key, value = key.to(red.dtype), value.to(red.dtype)
        o = scaled_dot_product_attention(red.unsqueeze(0), key.unsqueeze(0), value.unsqueeze(0),
                                         enable_gqa=enable_gqa, scale=scaling, is_causal=causal)
        out[start_q:end_q] = o.squeeze(0).movedim(1, 0)[p:]
        start_q = end_q
    return out.view(T, num_q_heads * head_dim)


# ----------------------------------------------------------------------------- a6
def merge_state(v_a: torch.Tensor, s_a: torch.Tensor, v_b: torch.Tensor, s_b: torch.Tensor
                ) -> Tuple[torch.Tensor, torch.Tensor]:
    """LSE-weighted 2-way merge, natural log, ``+inf`` LSE treated as ``-inf``.

    python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu:58-104 (the
    per-(token, head) scalar part :62-77, the vector part :86-99, the LSE :101-104).
    Computed in fp32, output cast back to ``v_a.dtype``.
    """
    sa = s_a.float().clone()
    sb = s_b.float().clone()
    sa[torch.isposinf(sa)] = -math.inf
    sb[torch.isposinf(sb)] = -math.inf
    m = torch.maximum(sa, sb)
    safe_m = torch.where(torch.isneginf(m), torch.zeros_like(m), m)
    wa = torch.exp(sa - safe_m)
    wb = torch.exp(sb - safe_m)
    wa = torch.where(torch.isneginf(m), torch.zeros_like(wa), wa)
    wb = torch.where(torch.isneginf(m), torch.zeros_like(wb), wb)
    tot = wa + wb
    inv = torch.where(tot > 0, 1.0 / tot, torch.zeros_like(tot))
    v = (v_a.float() * (wa * inv).unsqueeze(-1) + v_b.float() * (wb * inv).unsqueeze(-1)).to(v_a.dtype)
    s = torch.where(tot > 0, torch.log(tot) + m, torch.full_like(m, -math.inf))
    return v, s


# ------------------------------------------------------------- independent fp32 check
def attention_fp32(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
                   req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, extend_seq_lens: torch.Tensor,
                   num_q_heads: int, num_kv_heads: int, head_dim: int, scaling: float, causal: bool,
                   return_lse: bool = False, window_left: int = -1, logit_cap: float = 0.0, sinks=None,
                   custom_masks=None, skip_prefix_custom_mask: bool = True, e4m3_arith: bool = False):
    """Independent fp32-softmax reference (works on any device, used at large sizes).

    Restates the reference kit's ``_dense_attention_reference``
    (python/sglang/test/kits/attention_unittest/attention_methods/dense_attention.py:802-855)
    and ``decode_attention_fwd_torch`` (test/registered/attention/
    test_triton_attention_kernels.py:106-151): per request, scores in fp32,
    query j (absolute position ``L - E + j``) sees keys ``<= L - E + j`` when causal.

    Feature flags restate the production kernels: ``window_left`` = the kit's extend-window rule
    ``kv_id >= q_id - window`` (dense_attention.py:831-840, extend_attention.py:480-486); ``logit_cap``:
    ``cap * tanh(qk / cap)`` (extend_attention.py:546-547); ``sinks`` [Hq]: ``exp(sink - max)`` joins the
    denominator (extend_attention.py:721-723, decode_attention.py:986-988).  ``custom_masks``: per request a bool
    ``[E, L]`` block (True = visible), the TARGET_VERIFY tree mask of ``dense_attention_reference_with_custom_mask``
    (dense_attention.py:858-905); with ``skip_prefix_custom_mask`` the prefix keys stay visible
    (extend_attention.py:467-479).

    ``e4m3_arith``: the arithmetic of the reference's Triton EXTEND kernel over an e4m3 pool (``k_cache`` / ``v_cache``
    are then the DEQUANTISED pools): the query is cast to e4m3 (``q.to(k.dtype)``, extend_attention.py:523/1174) and
    the un-normalised probabilities are cast to e4m3 before the second contraction (``p.to(v.dtype)``, :596/1248) while
    the denominator sums them un-rounded.  Agrees with ``fp8_port.extend_attention_e4m3_arith`` (which is pinned to
    the kernel's recorded output) and adds the feature flags.
    """
    T = q.shape[0]
    G = num_q_heads // num_kv_heads
    out = torch.empty((T, num_q_heads, head_dim), dtype=torch.float32, device=q.device)
    lse = torch.empty((T, num_q_heads), dtype=torch.float32, device=q.device)
    qf = q.view(T, num_q_heads, head_dim).float()
    if e4m3_arith:
        qf = qf.to(torch.float8_e4m3fn).float()
    start = 0
    for i in range(seq_lens.shape[0]):
        L, E = int(seq_lens[i]), int(extend_seq_lens[i])
        tokens = req_to_token[int(req_pool_indices[i]), :L].long()
        k = k_cache[tokens].float().repeat_interleave(G, dim=1)   # [L, Hq, D]
        v = v_cache[tokens].float().repeat_interleave(G, dim=1)
        s = torch.einsum("ehd,lhd->hel", qf[start:start + E], k) * scaling
        if logit_cap > 0:
            s = logit_cap * torch.tanh(s / logit_cap)
        qpos = torch.arange(L - E, L, device=q.device).view(1, E, 1)
        kpos = torch.arange(L, device=q.device).view(1, 1, L)
        if causal:
            s = s.masked_fill(kpos > qpos, -math.inf)
        if window_left >= 0:
            s = s.masked_fill(kpos < qpos - window_left, -math.inf)
        if custom_masks is not None:
            allowed = custom_masks[i].to(q.device).bool().view(1, E, L).clone()
            if skip_prefix_custom_mask:
                allowed[..., :L - E] = True
            s = s.masked_fill(~allowed, -math.inf)
        if e4m3_arith:
            sk = sinks.to(s.device).float().view(-1, 1, 1).expand(-1, E, 1) if sinks is not None else None
            full = s if sk is None else torch.cat([s, sk], dim=-1)
            m = full.max(dim=-1, keepdim=True).values
            m = torch.where(torch.isfinite(m), m, torch.zeros_like(m))
            pu = torch.exp(full - m)
            den = pu.sum(dim=-1, keepdim=True)
            lse[start:start + E] = (m + torch.log(den)).squeeze(-1).transpose(0, 1)
            p = pu[..., :L].to(torch.float8_e4m3fn).float() / den
        elif sinks is not None:
            sk = sinks.to(s.device).float().view(-1, 1, 1).expand(-1, E, 1)
            full = torch.cat([s, sk], dim=-1)
            lse[start:start + E] = torch.logsumexp(full, dim=-1).transpose(0, 1)
            p = torch.softmax(full, dim=-1)[..., :L]
        else:
            lse[start:start + E] = torch.logsumexp(s, dim=-1).transpose(0, 1)
            p = torch.softmax(s, dim=-1)
        out[start:start + E] = torch.einsum("hel,lhd->ehd", p, v)
        start += E
    out = out.view(T, num_q_heads * head_dim)
    return (out, lse) if return_lse else out
