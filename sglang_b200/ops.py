"""Tensor-level wrappers over the C ABI (include/b200attn.h).

Each wrapper mirrors the reference operator it replaces (same argument meaning
and error behaviour) and only turns ``torch.Tensor`` arguments into raw device
pointers + the current CUDA stream.  No wrapper allocates on the hot path
except for the output when the caller did not pass one; none synchronises.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import BF16, FP16, FP32, I32, I64, check


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _fdtype(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return BF16
    if t.dtype == torch.float16:
        return FP16
    if t.dtype == torch.float32:
        return FP32
    raise NotImplementedError(f"b200attn: unsupported dtype {t.dtype}")


def _idtype(t: torch.Tensor) -> int:
    if t.dtype == torch.int32:
        return I32
    if t.dtype == torch.int64:
        return I64
    raise TypeError(f"b200attn: index tensor must be int32 or int64, got {t.dtype}")


def _require_cuda(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "b200attn kernels need CUDA tensors (there is no CPU fallback); got a "
                f"{t.device} tensor"
            )


def _sinks_ptr(sinks: Optional[torch.Tensor], Hq: int) -> Optional[int]:
    """Attention sinks: float32 [Hq] on the device (triton_backend.py:1416, decode_attention.py:986-988)."""
    if sinks is None:
        return None
    _require_cuda(sinks)
    if sinks.dtype != torch.float32 or sinks.numel() != Hq or not sinks.is_contiguous():
        raise ValueError("sinks must be a contiguous float32 tensor with one entry per q head")
    return sinks.data_ptr()


def _mask_ptr(custom_mask: Optional[torch.Tensor], mask_indptr: Optional[torch.Tensor], B: int) -> Optional[int]:
    """TARGET_VERIFY custom mask: flat bool / uint8 on the device + int32 mask_indptr[B+1] (triton_backend.py:911-917)."""
    if custom_mask is None:
        return None
    _require_cuda(custom_mask, mask_indptr)
    if custom_mask.dtype not in (torch.bool, torch.uint8) or not custom_mask.is_contiguous():
        raise TypeError("custom_mask must be a contiguous bool / uint8 tensor")
    if mask_indptr is None or mask_indptr.dtype != torch.int32 or mask_indptr.numel() < B + 1:
        raise ValueError("custom_mask needs an int32 mask_indptr with B + 1 entries")
    return custom_mask.data_ptr()


def sm_count() -> int:
    return _lib.load().b200attn_sm_count()


# --------------------------------------------------------------------------- store
def store_cache(
    k: torch.Tensor,
    v: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    indices: torch.Tensor,
    *,
    reserved_skip_index: int = 0,
    oob_count: Optional[torch.Tensor] = None,
) -> None:
    """``k_cache[indices] = k; v_cache[indices] = v`` (bit-exact row scatter).

    Mirrors ``store_cache`` (python/sglang/kernels/ops/kvcache/kvcache.py:56-110):
    ``k``/``v`` are ``[T, ...]`` with contiguous rows (the token stride may be
    larger than the row, e.g. views of the fused qkv output); caches are
    ``[slots, ...]``; rows whose index equals ``reserved_skip_index`` are skipped
    (pass ``-1`` to write every row, the reference's naive branch).
    """
    _require_cuda(k, v, k_cache, v_cache, indices)
    T = k.shape[0]
    if v.shape[0] != T or indices.numel() != T:
        raise ValueError("store_cache: k, v and indices disagree on the number of tokens")
    if T == 0:
        return
    if k.dtype != k_cache.dtype or v.dtype != v_cache.dtype:
        raise TypeError("store_cache: cache dtype must equal the row dtype (cast before the store)")
    k2, v2 = k.reshape(T, -1), v.reshape(T, -1)
    kc, vc = k_cache.reshape(k_cache.shape[0], -1), v_cache.reshape(v_cache.shape[0], -1)
    if k2.stride(1) != 1 or v2.stride(1) != 1 or kc.stride(1) != 1 or vc.stride(1) != 1:
        raise ValueError("store_cache: rows must be contiguous")
    if k2.shape[1] != kc.shape[1] or v2.shape[1] != vc.shape[1]:
        raise ValueError("store_cache: row width differs between input and cache")
    if not indices.is_contiguous():
        indices = indices.contiguous()
    ke, ve = k.element_size(), v.element_size()
    rc = _lib.load().b200attn_kv_store(
        k2.data_ptr(), v2.data_ptr(), kc.data_ptr(), vc.data_ptr(), indices.data_ptr(),
        _idtype(indices), T, k2.shape[1] * ke, v2.shape[1] * ve, k2.stride(0) * ke,
        v2.stride(0) * ve, kc.stride(0) * ke, vc.stride(0) * ve,
        min(kc.shape[0], vc.shape[0]), reserved_skip_index, _ptr(oob_count), _stream(),
    )
    check(rc, "b200attn_kv_store")


_FP8 = torch.float8_e4m3fn


def _as_bytes(t: torch.Tensor) -> torch.Tensor:
    """float8 pools are handled as raw bytes (the reference's store dtype, memory_pool.py:2373-2375)."""
    return t.view(torch.uint8) if t.dtype == _FP8 else t


def store_cache_fp8(
    k: torch.Tensor,
    v: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    indices: torch.Tensor,
    k_scale: Optional[float] = None,
    v_scale: Optional[float] = None,
    *,
    reserved_skip_index: int = 0,
    oob_count: Optional[torch.Tensor] = None,
) -> None:
    """Quantising store into a float8_e4m3fn (or uint8-viewed) pool.

    ``k_cache[indices] = (k / k_scale).to(float8_e4m3fn)`` with the reference's rounding order
    (memory_pool.py:2365-2371: ``div_`` in the input dtype, then the cast); bit-exact against that
    expression evaluated by torch on the CPU.  ``None`` scales skip the division.
    """
    _require_cuda(k, v, k_cache, v_cache, indices)
    T = k.shape[0]
    if v.shape[0] != T or indices.numel() != T:
        raise ValueError("store_cache_fp8: k, v and indices disagree on the number of tokens")
    if T == 0:
        return
    if k.dtype not in (torch.bfloat16, torch.float16) or v.dtype != k.dtype:
        raise TypeError("store_cache_fp8: k/v must be bf16 or fp16")
    kc, vc = _as_bytes(k_cache), _as_bytes(v_cache)
    if kc.dtype != torch.uint8 or vc.dtype != torch.uint8:
        raise TypeError("store_cache_fp8: the pool must be float8_e4m3fn or its uint8 view")
    k2, v2 = k.reshape(T, -1), v.reshape(T, -1)
    kc, vc = kc.reshape(kc.shape[0], -1), vc.reshape(vc.shape[0], -1)
    if k2.stride(1) != 1 or v2.stride(1) != 1 or kc.stride(1) != 1 or vc.stride(1) != 1:
        raise ValueError("store_cache_fp8: rows must be contiguous")
    if k2.shape[1] != kc.shape[1] or v2.shape[1] != vc.shape[1]:
        raise ValueError("store_cache_fp8: row width differs between input and cache")
    if not indices.is_contiguous():
        indices = indices.contiguous()
    rc = _lib.load().b200attn_kv_store_fp8(
        k2.data_ptr(), v2.data_ptr(), kc.data_ptr(), vc.data_ptr(), indices.data_ptr(), _idtype(indices), T,
        k2.shape[1], v2.shape[1], k2.stride(0), v2.stride(0), kc.stride(0), vc.stride(0), _fdtype(k),
        float(k_scale) if k_scale is not None else 0.0, float(v_scale) if v_scale is not None else 0.0,
        min(kc.shape[0], vc.shape[0]), reserved_skip_index, _ptr(oob_count), _stream(),
    )
    check(rc, "b200attn_kv_store_fp8")


def kv_dequant_gather(k_cache: torch.Tensor, v_cache: torch.Tensor, kv_indices: torch.Tensor, n: int,
                      out_k: torch.Tensor, out_v: torch.Tensor, k_mul: float = 1.0, v_mul: float = 1.0,
                      n_dev: Optional[torch.Tensor] = None) -> None:
    """``out_k[i] = k_cache[kv_indices[i]] * k_mul`` (same for v) from an fp8 pool into 16-bit rows ``[n, Hkv, D]``:
    the K/V operand scratch of extend over an fp8 pool (b200attn_kv_dequant_gather)."""
    _require_cuda(k_cache, v_cache, kv_indices, out_k, out_v)
    kc, vc = _as_bytes(k_cache), _as_bytes(v_cache)
    kc, vc = kc.reshape(kc.shape[0], -1), vc.reshape(vc.shape[0], -1)
    ok, ov = out_k.reshape(out_k.shape[0], -1), out_v.reshape(out_v.shape[0], -1)
    if kv_indices.dtype != torch.int32 or (n_dev is not None and n_dev.dtype != torch.int32):
        raise TypeError("kv_dequant_gather: kv_indices / n_dev must be int32")
    if ok.shape[0] < n or ov.shape[0] < n or kv_indices.numel() < n:
        raise ValueError("kv_dequant_gather: outputs / indices shorter than n")
    if kc.stride() != vc.stride() or ok.stride() != ov.stride() or ok.shape[1] != kc.shape[1] or ok.dtype != ov.dtype:
        raise ValueError("kv_dequant_gather: K and V must share layouts")
    rc = _lib.load().b200attn_kv_dequant_gather(
        kc.data_ptr(), vc.data_ptr(), kc.stride(0), kv_indices.data_ptr(), int(n), kc.shape[1],
        ok.data_ptr(), ov.data_ptr(), ok.stride(0), _fdtype(out_k), float(k_mul), float(v_mul), _ptr(n_dev),
        _stream(),
    )
    check(rc, "b200attn_kv_dequant_gather")


def kv_dequant_gather_unique(k_cache: torch.Tensor, v_cache: torch.Tensor, kv_indices: torch.Tensor, n: int,
                             out_k: torch.Tensor, out_v: torch.Tensor, first_ws: torch.Tensor,
                             out_indices: torch.Tensor, k_mul: float = 1.0, v_mul: float = 1.0,
                             n_dev: Optional[torch.Tensor] = None) -> None:
    """De-duplicating ``kv_dequant_gather``: every referenced slot is converted once, into the scratch row of its
    first referencing position; ``out_indices[i]`` names that row (b200attn_kv_dequant_gather_unique)."""
    _require_cuda(k_cache, v_cache, kv_indices, out_k, out_v, first_ws, out_indices)
    kc, vc = _as_bytes(k_cache), _as_bytes(v_cache)
    kc, vc = kc.reshape(kc.shape[0], -1), vc.reshape(vc.shape[0], -1)
    ok, ov = out_k.reshape(out_k.shape[0], -1), out_v.reshape(out_v.shape[0], -1)
    if any(t.dtype != torch.int32 for t in (kv_indices, first_ws, out_indices)) or \
            (n_dev is not None and n_dev.dtype != torch.int32):
        raise TypeError("kv_dequant_gather_unique: index tensors must be int32")
    if ok.shape[0] < n or ov.shape[0] < n or kv_indices.numel() < n or out_indices.numel() < n:
        raise ValueError("kv_dequant_gather_unique: outputs / indices shorter than n")
    if first_ws.numel() < kc.shape[0]:
        raise ValueError("kv_dequant_gather_unique: first_ws must have one int32 per pool slot")
    if kc.stride() != vc.stride() or ok.stride() != ov.stride() or ok.shape[1] != kc.shape[1] or ok.dtype != ov.dtype:
        raise ValueError("kv_dequant_gather_unique: K and V must share layouts")
    rc = _lib.load().b200attn_kv_dequant_gather_unique(
        kc.data_ptr(), vc.data_ptr(), kc.stride(0), kv_indices.data_ptr(), int(n), kc.shape[1],
        ok.data_ptr(), ov.data_ptr(), ok.stride(0), _fdtype(out_k), float(k_mul), float(v_mul), _ptr(n_dev),
        first_ws.data_ptr(), kc.shape[0], out_indices.data_ptr(), _stream(),
    )
    check(rc, "b200attn_kv_dequant_gather_unique")


# -------------------------------------------------------------------- rope + store
ROPE_F32, ROPE_NATIVE = 0, 1


def apply_rope_with_cos_sin_cache_inplace(
    positions: torch.Tensor,
    q: torch.Tensor,
    k: torch.Tensor,
    cos_sin_cache: torch.Tensor,
    *,
    head_size: int,
    is_neox: bool = True,
    rope_dim: int = 0,
    mrope_section: Optional[Sequence[int]] = None,
    mrope_interleaved: bool = False,
    v: Optional[torch.Tensor] = None,
    k_buffer: Optional[torch.Tensor] = None,
    v_buffer: Optional[torch.Tensor] = None,
    cache_loc: Optional[torch.Tensor] = None,
    reserved_skip_index: int = 0,
    oob_count: Optional[torch.Tensor] = None,
    native_rounding: bool = False,
    k_scale: Optional[float] = None,
    v_scale: Optional[float] = None,
) -> None:
    """Rotate ``q`` and ``k`` in place; with ``k_buffer/v_buffer/cache_loc/v`` also scatter the rotated k
    and v into the pool (one kernel).

    Mirrors ``apply_rope_with_cos_sin_cache_inplace(q, k, cos_sin_cache, positions, is_neox=, rope_dim=,
    fused_args=FusedSetKVBufferArg(value, k_buffer, v_buffer, cache_loc))``
    (python/sglang/kernels/ops/attention/rope.py:98-113,183-225) and adds the M-RoPE select of
    ``MRotaryEmbedding.forward_native`` (python/sglang/srt/layers/rotary_embedding/mrope.py:173-217):
    ``positions`` is ``[T]`` or ``[3, T]`` (then ``mrope_section`` is required).  ``q``/``k`` are
    ``[T, H*head_size]`` or ``[T, H, head_size]`` with contiguous rows (token stride free).
    ``native_rounding`` selects torch's element-dtype arithmetic (bit-exact with ``forward_native``),
    otherwise fp32 arithmetic with one rounding (the reference CUDA kernel's).  A float8_e4m3fn (or uint8-viewed)
    pool takes b200attn_rope_store_fp8: the rotated k and v are quantised on the way into the pool with
    ``k_scale`` / ``v_scale`` (memory_pool.py:2365-2375), same bytes as rotate-then-``store_cache_fp8``.
    """
    _require_cuda(positions, q, k, cos_sin_cache, v, k_buffer, v_buffer, cache_loc)
    T = q.shape[0]
    if k.shape[0] != T:
        raise ValueError("rope: q and k disagree on the number of tokens")
    rot = rope_dim or cos_sin_cache.shape[-1]
    if cos_sin_cache.dim() != 2 or cos_sin_cache.shape[1] != rot or not cos_sin_cache.is_contiguous():
        raise ValueError("rope: cos_sin_cache must be contiguous [max_position, rope_dim]")
    if q.dtype != k.dtype:
        raise TypeError("rope: q and k must share a dtype")
    if cos_sin_cache.dtype not in (torch.float32, q.dtype):
        raise TypeError("rope: cos_sin_cache must be float32 or q's dtype")
    q2, k2 = q.reshape(T, -1) if q.dim() != 2 else q, k.reshape(T, -1) if k.dim() != 2 else k
    if T and (q2.data_ptr() != q.data_ptr() or k2.data_ptr() != k.data_ptr()):
        raise ValueError("rope: q/k rows must be viewable as [T, H*head_size] without a copy (in-place op)")
    if q2.stride(1) != 1 or k2.stride(1) != 1:
        raise ValueError("rope: rows must be contiguous")
    if q2.shape[1] % head_size or k2.shape[1] % head_size:
        raise ValueError("rope: row width is not a multiple of head_size")
    Hq, Hkv = q2.shape[1] // head_size, k2.shape[1] // head_size
    if positions.dim() == 2:
        if positions.shape[0] != 3 or mrope_section is None:
            raise ValueError("rope: 2-D positions must be [3, T] and need mrope_section")
        if positions.stride(1) != 1:
            positions = positions.contiguous()
        if positions.shape[1] != T:
            raise ValueError("rope: positions and q disagree on the number of tokens")
        # stride(0) == 0 is a valid axis stride (text-only batches: pos.unsqueeze(0).expand(3, -1)); anything
        # else below T would alias rows, so it is made contiguous
        if positions.stride(0) != 0 and positions.stride(0) < T:
            positions = positions.contiguous()
        pos_axis_stride = positions.stride(0)
        import ctypes as C
        sec = (C.c_int32 * 3)(*[int(x) for x in mrope_section])
    else:
        if positions.numel() != T:
            raise ValueError("rope: positions and q disagree on the number of tokens")
        if not positions.is_contiguous():
            positions = positions.contiguous()
        pos_axis_stride, sec = 0, None
        mrope_interleaved = False
    store = k_buffer is not None
    Dv = 0
    kc = vc = v2 = None
    if store:
        if v is None or v_buffer is None or cache_loc is None:
            raise ValueError("rope: fused store needs v, k_buffer, v_buffer and cache_loc")
        pool_fp8 = k_buffer.dtype in (_FP8, torch.uint8)
        if pool_fp8:
            k_buffer, v_buffer = _as_bytes(k_buffer), _as_bytes(v_buffer)
            if v_buffer.dtype != torch.uint8 or v.dtype != k.dtype:
                raise TypeError("rope: K and V pools must both be float8_e4m3fn and k/v share a dtype")
        elif k.dtype != k_buffer.dtype or v.dtype != v_buffer.dtype or v.dtype != k.dtype:
            raise TypeError("rope: cache dtype must equal the row dtype")
        v2 = v.reshape(T, -1)
        kc, vc = k_buffer.reshape(k_buffer.shape[0], -1), v_buffer.reshape(v_buffer.shape[0], -1)
        if v2.stride(1) != 1 or kc.stride(1) != 1 or vc.stride(1) != 1:
            raise ValueError("rope: rows must be contiguous")
        if kc.shape[1] != k2.shape[1] or vc.shape[1] != v2.shape[1] or v2.shape[1] % Hkv:
            raise ValueError("rope: row width differs between input and cache")
        Dv = v2.shape[1] // Hkv
        if cache_loc.numel() != T:
            raise ValueError("rope: cache_loc and q disagree on the number of tokens")
        if not cache_loc.is_contiguous():
            cache_loc = cache_loc.contiguous()
    if T == 0:
        return
    if store and pool_fp8:
        rc = _lib.load().b200attn_rope_store_fp8(
            q2.data_ptr(), k2.data_ptr(), _ptr(v2), _ptr(kc), _ptr(vc), _ptr(cache_loc), _idtype(cache_loc),
            cos_sin_cache.data_ptr(), _fdtype(cos_sin_cache), cos_sin_cache.shape[0], positions.data_ptr(),
            _idtype(positions), pos_axis_stride, sec, int(bool(mrope_interleaved)), T, Hq, Hkv, head_size, Dv, rot,
            int(bool(is_neox)), q2.stride(0), k2.stride(0), v2.stride(0), kc.stride(0), vc.stride(0),
            min(kc.shape[0], vc.shape[0]), reserved_skip_index, _ptr(oob_count), _fdtype(q),
            ROPE_NATIVE if native_rounding else ROPE_F32, float(k_scale) if k_scale is not None else 0.0,
            float(v_scale) if v_scale is not None else 0.0, _stream(),
        )
        check(rc, "b200attn_rope_store_fp8")
        return
    rc = _lib.load().b200attn_rope_store(
        q2.data_ptr(), k2.data_ptr(), _ptr(v2), _ptr(kc), _ptr(vc), _ptr(cache_loc),
        _idtype(cache_loc) if store else I32, cos_sin_cache.data_ptr(), _fdtype(cos_sin_cache),
        cos_sin_cache.shape[0], positions.data_ptr(), _idtype(positions), pos_axis_stride, sec,
        int(bool(mrope_interleaved)), T, Hq, Hkv, head_size, Dv, rot, int(bool(is_neox)), q2.stride(0),
        k2.stride(0), v2.stride(0) if store else 0, kc.stride(0) if store else 0,
        vc.stride(0) if store else 0, min(kc.shape[0], vc.shape[0]) if store else 0, reserved_skip_index,
        _ptr(oob_count), _fdtype(q), ROPE_NATIVE if native_rounding else ROPE_F32, _stream(),
    )
    check(rc, "b200attn_rope_store")


# ---------------------------------------------------------------------- allocator
def alloc_extend(prefix_lens: torch.Tensor, seq_lens: torch.Tensor, last_loc: torch.Tensor,
                 free_pages: torch.Tensor, out_indices: torch.Tensor, page_size: int,
                 status: Optional[torch.Tensor] = None) -> None:
    """Fill ``out_indices`` (``[sum(seq_lens - prefix_lens)]``) with the KV slots of every new token.

    Mirrors ``alloc_extend_kernel[(bs,)](prefix_lens, seq_lens, last_loc, free_pages, out_indices, ...)``
    (python/sglang/kernels/ops/memory/allocator.py:15-103); ``status`` (int32[1]) is set to 1 when the free
    list is too short."""
    _require_cuda(prefix_lens, seq_lens, last_loc, free_pages, out_indices, status)
    B = seq_lens.numel()
    if prefix_lens.numel() != B or last_loc.numel() != B:
        raise ValueError("alloc_extend: prefix_lens, seq_lens and last_loc disagree on the batch size")
    if not (prefix_lens.dtype == seq_lens.dtype == last_loc.dtype):
        raise TypeError("alloc_extend: prefix_lens, seq_lens and last_loc must share a dtype")
    for t in (prefix_lens, seq_lens, last_loc, free_pages, out_indices):
        if not t.is_contiguous():
            raise ValueError("alloc_extend: tensors must be contiguous")
    rc = _lib.load().b200attn_alloc_extend(
        prefix_lens.data_ptr(), seq_lens.data_ptr(), last_loc.data_ptr(), _idtype(seq_lens), free_pages.data_ptr(),
        _idtype(free_pages), free_pages.numel(), out_indices.data_ptr(), _idtype(out_indices), B, page_size,
        _ptr(status), _stream())
    check(rc, "b200attn_alloc_extend")


def alloc_decode(seq_lens: torch.Tensor, last_loc: torch.Tensor, free_pages: torch.Tensor,
                 out_indices: torch.Tensor, page_size: int, status: Optional[torch.Tensor] = None) -> None:
    """One new slot per running request (allocator.py:106-140)."""
    _require_cuda(seq_lens, last_loc, free_pages, out_indices, status)
    B = seq_lens.numel()
    if last_loc.numel() != B or out_indices.numel() != B:
        raise ValueError("alloc_decode: seq_lens, last_loc and out_indices disagree on the batch size")
    if seq_lens.dtype != last_loc.dtype:
        raise TypeError("alloc_decode: seq_lens and last_loc must share a dtype")
    for t in (seq_lens, last_loc, free_pages, out_indices):
        if not t.is_contiguous():
            raise ValueError("alloc_decode: tensors must be contiguous")
    rc = _lib.load().b200attn_alloc_decode(
        seq_lens.data_ptr(), last_loc.data_ptr(), _idtype(seq_lens), free_pages.data_ptr(), _idtype(free_pages),
        free_pages.numel(), out_indices.data_ptr(), _idtype(out_indices), B, page_size, _ptr(status), _stream())
    check(rc, "b200attn_alloc_decode")


# ------------------------------------------------------------------------ metadata
def build_kv_indices(
    req_to_token: torch.Tensor,
    req_pool_indices: torch.Tensor,
    kv_lens: torch.Tensor,
    kv_indptr: torch.Tensor,
    kv_indices: Optional[torch.Tensor],
    *,
    kv_start: Optional[torch.Tensor] = None,
    qo_lens: Optional[torch.Tensor] = None,
    qo_indptr: Optional[torch.Tensor] = None,
    slot_map: Optional[torch.Tensor] = None,
) -> None:
    """Fill ``kv_indptr`` (int32 ``[B+1]``), ``kv_indices`` (flat slot ids) and optionally ``qo_indptr``.

    Mirrors ``create_flashinfer_kv_indices_triton`` + the cumsum around it
    (python/sglang/srt/layers/attention/triton_backend.py:438-456).  ``slot_map`` translates every slot id
    (``full_to_swa_index_mapping`` of a hybrid sliding-window pool, swa_memory_pool.py:228-231).
    """
    _require_cuda(req_to_token, req_pool_indices, kv_lens, kv_indptr, kv_indices, kv_start, qo_lens, qo_indptr,
                  slot_map)
    B = kv_lens.numel()
    if req_to_token.dtype != torch.int32 or req_to_token.stride(1) != 1:
        raise TypeError("build_kv_indices: req_to_token must be int32 with contiguous rows")
    if kv_indptr.dtype != torch.int32 or kv_indptr.numel() < B + 1:
        raise ValueError("build_kv_indices: kv_indptr must be int32[B+1]")
    if qo_indptr is not None and (qo_indptr.dtype != torch.int32 or qo_indptr.numel() < B + 1):
        raise ValueError("build_kv_indices: qo_indptr must be int32[B+1]")
    rc = _lib.load().b200attn_build_kv_indices(
        req_to_token.data_ptr(), req_to_token.stride(0),
        req_pool_indices.data_ptr(), _idtype(req_pool_indices),
        kv_lens.data_ptr(), _idtype(kv_lens),
        _ptr(kv_start), _idtype(kv_start) if kv_start is not None else I32,
        _ptr(qo_lens), _idtype(qo_lens) if qo_lens is not None else I32,
        B, kv_indptr.data_ptr(), _ptr(kv_indices),
        _idtype(kv_indices) if kv_indices is not None else I32,
        kv_indices.numel() if kv_indices is not None else 0, _ptr(qo_indptr),
        _ptr(slot_map), _idtype(slot_map) if slot_map is not None else I32, _stream(),
    )
    check(rc, "b200attn_build_kv_indices")


def stage_copy(dst: torch.Tensor, src: torch.Tensor, dst1: Optional[torch.Tensor] = None,
               src1: Optional[torch.Tensor] = None) -> None:
    """``dst[...] = src`` (and ``dst1[...] = src1``) by ONE kernel launch on the current stream.  Each tensor is a
    contiguous CUDA tensor or a contiguous PINNED host tensor (mapped into the device address space): the kernel
    reads / writes host memory itself, so a CUDA graph of a host-driven step holds no memcpy nodes
    (csrc/stage.cu; the buffer refills of cuda_graph_runner.py replay_prepare)."""
    segs = []
    for d, s_ in ((dst, src), (dst1, src1)):
        if d is None:
            segs += [0, 0, 0]
            continue
        for t in (d, s_):
            if not (t.is_cuda or t.is_pinned()):
                raise RuntimeError("b200attn.stage_copy: tensors must be CUDA or pinned host tensors (no CPU fallback)")
            if not t.is_contiguous():
                raise RuntimeError("b200attn.stage_copy: tensors must be contiguous")
        nb = d.numel() * d.element_size()
        if nb != s_.numel() * s_.element_size():
            raise RuntimeError("b200attn.stage_copy: size mismatch")
        segs += [d.data_ptr(), s_.data_ptr(), nb]
    rc = _lib.load().b200attn_stage_copy(*segs, _stream())
    check(rc, "b200attn_stage_copy")


def translate_slots(loc: torch.Tensor, slot_map: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``slot_map[loc]`` as int64: ``KVWriteLoc.swa_loc`` of a hybrid sliding-window pool
    (``SWAKVPool.translate_loc_from_full_to_swa``, swa_memory_pool.py:228-231)."""
    _require_cuda(loc, slot_map, out)
    loc = loc.contiguous()
    if out is None:
        out = torch.empty(loc.numel(), dtype=torch.int64, device=loc.device)
    rc = _lib.load().b200attn_translate_slots(loc.data_ptr(), _idtype(loc), slot_map.data_ptr(), _idtype(slot_map),
                                              loc.numel(), out.data_ptr(), _stream())
    check(rc, "b200attn_translate_slots")
    return out


def build_page_table(
    req_to_token: torch.Tensor,
    req_pool_indices: torch.Tensor,
    seq_lens: torch.Tensor,
    page_table: torch.Tensor,
    page_size: int,
) -> None:
    """``page_table[b, j] = req_to_token[req[b], j*page] // page`` for ``j < ceil(len/page)``.

    Mirrors ``build_trtllm_mha_page_table``
    (python/sglang/kernels/ops/kvcache/trtllm_mha_page_table.py:96-120).
    """
    _require_cuda(req_to_token, req_pool_indices, seq_lens, page_table)
    if page_table.dtype != torch.int32 or page_table.dim() != 2 or page_table.stride(1) != 1:
        raise TypeError("build_page_table: page_table must be int32 [B, max_pages] with contiguous rows")
    B = seq_lens.numel()
    rc = _lib.load().b200attn_build_page_table(
        req_to_token.data_ptr(), req_to_token.stride(0), req_pool_indices.data_ptr(),
        _idtype(req_pool_indices), seq_lens.data_ptr(), _idtype(seq_lens), B, page_size,
        page_table.data_ptr(), page_table.stride(0), page_table.shape[1], _stream(),
    )
    check(rc, "b200attn_build_page_table")


def build_token_items(qo_indptr: torch.Tensor, kv_indptr: torch.Tensor, B: int, T: int, causal: bool,
                      req_ids: Optional[torch.Tensor] = None, item_lens: Optional[torch.Tensor] = None
                      ) -> Tuple[torch.Tensor, torch.Tensor]:
    """One decode item per query token of an extend batch (b200attn_build_token_items): ``(req_ids, item_lens)``
    int32 [T] for ``decode_attention(..., req_ids=, item_lens=)`` - the generic-head-dim extend path."""
    _require_cuda(qo_indptr, kv_indptr, req_ids, item_lens)
    if req_ids is None:
        req_ids = torch.empty(max(T, 1), dtype=torch.int32, device=qo_indptr.device)
    if item_lens is None:
        item_lens = torch.empty(max(T, 1), dtype=torch.int32, device=qo_indptr.device)
    rc = _lib.load().b200attn_build_token_items(qo_indptr.data_ptr(), kv_indptr.data_ptr(), B, T, int(bool(causal)),
                                                req_ids.data_ptr(), item_lens.data_ptr(), _stream())
    check(rc, "b200attn_build_token_items")
    return req_ids, item_lens


# -------------------------------------------------------------------------- decode
def decode_workspace_bytes(B: int, Hq: int, D: int, max_splits: int) -> int:
    return _lib.load().b200attn_decode_workspace_bytes(B, Hq, D, max_splits)


def decode_sk_workspace_bytes(B: int, Hq: int, Hkv: int, D: int) -> int:
    return int(_lib.load().b200attn_decode_sk_workspace_bytes(B, Hq, Hkv, D))


def decode_ws_workspace_bytes(Hq: int, Hkv: int, D: int) -> int:
    return int(_lib.load().b200attn_decode_ws_workspace_bytes(Hq, Hkv, D))


def decode_ctas_per_sm(fp8_kv: bool = False, warps: int = 4) -> int:
    """Resident CTAs per SM of the decode kernel that will run (its __launch_bounds__): 2 x 96 KB rings (4 warps) or
    4 x 48 KB rings of 64-thread CTAs (2 warps) for 16-bit pools, 3 x 64 KB for e4m3 pools (csrc/decode.cu); the wave
    arithmetic of the split heuristics below needs it."""
    return int(_lib.load().b200attn_decode_ctas_per_sm(1 if fp8_kv else 0, warps))


def plan_decode(B: int, Hkv: int, G: int, max_kv_len: int, min_chunk: int = 256,
                num_sms: Optional[int] = None) -> Tuple[int, int, int]:
    """(warps, max_splits, tail_begin) of a 16-bit decode launch.  A pure function of the batch SHAPE (graph-safe).

    Units = B * Hkv * ceil(G / 16) equal-priority CTAs.  Measured on B200 (profiles/r02_decode_warps.jsonl): a 4-warp
    CTA streams twice as fast as a 2-warp one and the 4-warp grid is 3.5 % faster in steady state (cfg5: 1248 vs
    1291 us), but only 2 fit on an SM; with 2 warps four do, and a batch of ~415..592 units (cfg2: 512) runs as ONE
    wave in which every CTA streams at the same rate to the end (170.5 -> 166.5 us) instead of a full wave plus a
    partial one.  Everything else keeps 4 warps with the split cap / tail split of pick_max_splits / pick_tail_split."""
    if num_sms is None:
        num_sms = sm_count()
    units = B * Hkv * (-(-G // 16))
    # measured crossover (tools/sweep_decode_warps.py, ctx 4096): 384 units 140.9 (2 warps) vs 130.1 us (4 warps +
    # tail split), 448 units 152.5 vs 160.6, 512 units 166.5 vs 170.5, 592 units 187.0 vs 189.2
    if 2.8 * num_sms < units <= 4 * num_sms:
        return 2, 1, -1
    ms = pick_max_splits(B, Hkv, G, max_kv_len, min_chunk, ctas_per_sm=2, num_sms=num_sms)
    tail_begin = -1
    if ms == 1:
        tail_begin, k = pick_tail_split(B, Hkv, G, num_sms=num_sms, ctas_per_sm=2)
        if k > 1:
            ms = k
        else:
            tail_begin = -1
    return 4, ms, tail_begin


def pick_max_splits(B: int, Hkv: int, G: int, max_kv_len: int, min_chunk: int = 256,
                    ctas_per_sm: int = 2, num_sms: Optional[int] = None) -> int:
    """Host-side choice of the split cap.  The decode grid is (splits x kv-head groups x batch) CTAs over
    ``W = ctas_per_sm * num_sms`` resident slots.  Measured on B200 (tools/compare_small_decode.py, 8 kv heads,
    W = 296): the best cap is the LARGEST one that keeps every CTA in the first wave, ``floor(W / units)`` —
    B=4 x 8K: 9 -> 33.5 us (10 spills into a second wave: 41.8 us); B=8 x 8K: 4 -> 51.8 us (5: 68.7); B=16 x 4K:
    2 -> 53.0 us (3: 64.2); units >= W/2: 1 (B=32 x 8K: 167 us vs 178 at 2) — and never finer than ``min_chunk``
    keys.  The per-request split count itself is computed on the device (split_policy in csrc/common.cuh); a nearly
    empty LAST wave is handled by pick_tail_split."""
    if num_sms is None:
        num_sms = sm_count()
    groups = (G + 15) // 16
    items = max(1, B * Hkv * groups)
    slots = ctas_per_sm * num_sms
    cap = int(max(1, min(-(-max_kv_len // min_chunk), 64)))
    return int(max(1, min(cap, slots // items)))


def pick_tail_split(B: int, Hkv: int, G: int, num_sms: Optional[int] = None, ctas_per_sm: int = 2,
                    max_k: int = 8) -> Tuple[int, int]:
    """Tail splitting for batches whose units do not fill whole waves of resident CTAs.

    Units = B * Hkv * ceil(G / 16) equal-priority CTAs, ``W = ctas_per_sm * SMs`` run at a time.  A last wave of
    only R << W units leaves most SMs idle for a whole unit, so ONLY those R units are split k ways (k = W // R:
    the split tail is one wave).  Measured on B200 (tools/decode_waves.py, ctx 4096): R/W = 0.08 -> 158.5 us ->
    117.0 us; R/W = 0.73 -> 171 us -> 177 us (the shorter CTAs and the merge launch cost more than the idle
    slots), hence the R <= W/2 gate.  Returns ``(tail_begin, k)``; ``(-1, 1)`` when nothing is gained."""
    W = ctas_per_sm * (num_sms or sm_count())
    U = B * Hkv * (-(-G // 16))
    F, R = divmod(U, W)
    if F == 0 or R == 0 or 2 * R > W:
        return -1, 1
    k = min(max_k, W // R)
    return (F * W, k) if k >= 2 else (-1, 1)


def decode_attention(
    q: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    o: torch.Tensor,
    kv_indptr: torch.Tensor,
    kv_indices: torch.Tensor,
    sm_scale: float,
    *,
    max_splits: int = 1,
    min_chunk: int = 256,
    workspace: Optional[torch.Tensor] = None,
    lse: Optional[torch.Tensor] = None,
    q_rows: Optional[torch.Tensor] = None,
    req_ids: Optional[torch.Tensor] = None,
    num_items: Optional[int] = None,
    window_left: int = -1,
    logit_cap: float = 0.0,
    sinks: Optional[torch.Tensor] = None,
    tail_begin: int = -1,
    k_descale: float = 1.0,
    v_descale: float = 1.0,
    stream_k: bool = False,
    ws_kernel: bool = False,
    warps: int = 4,
    item_lens: Optional[torch.Tensor] = None,
    new_k: Optional[torch.Tensor] = None,
    new_v: Optional[torch.Tensor] = None,
    skip_slot: int = 0,
) -> torch.Tensor:
    """Paged GQA decode: ``q [B,Hq,D]``, caches ``[slots,Hkv,D]`` -> ``o [B,Hq,D]``.

    A float8_e4m3fn (or uint8-viewed) pool takes the fp8 kernel (b200attn_decode_fp8kv) with
    ``k_descale`` / ``v_descale`` (triton_backend.py:1418-1420); 16-bit pools ignore them.

    Mirrors ``decode_attention_fwd`` (python/sglang/kernels/ops/attention/
    decode_attention.py:1163-1239) with token-granular ``kv_indices`` (int32).
    """
    _require_cuda(q, k_cache, v_cache, o, kv_indptr, kv_indices, workspace, lse, q_rows, req_ids)
    if q.dim() != 3 or k_cache.dim() != 3:
        raise ValueError("decode_attention: q must be [B,Hq,D] and caches [slots,Hkv,D]")
    _, Hq, D = q.shape
    B = num_items if num_items is not None else (q_rows.numel() if q_rows is not None else q.shape[0])
    Hkv = k_cache.shape[1]
    if k_cache.shape[2] != D or v_cache.shape[2] != D:
        raise NotImplementedError("decode_attention: v_head_dim != head_dim is not supported")
    fp8 = k_cache.dtype in (_FP8, torch.uint8)
    if fp8:
        k_cache, v_cache = _as_bytes(k_cache), _as_bytes(v_cache)
    if k_cache.dtype != v_cache.dtype or (not fp8 and q.dtype != k_cache.dtype):
        raise TypeError("decode_attention: q and the KV pool must share a dtype (or the pool is float8_e4m3fn)")
    if q.stride(2) != 1 or q.stride(1) != D or o.stride(2) != 1 or o.stride(1) != D:
        raise ValueError("decode_attention: q/o heads must be densely packed")
    if k_cache.stride(2) != 1 or k_cache.stride(1) != D or v_cache.stride() != k_cache.stride():
        raise ValueError("decode_attention: KV pool must be NHD contiguous")
    if kv_indptr.dtype != torch.int32 or kv_indices.dtype != torch.int32:
        raise TypeError("decode_attention: kv_indptr / kv_indices must be int32")
    ws_bytes = 0 if workspace is None else workspace.numel() * workspace.element_size()
    if ws_kernel:
        # warp-specialised persistent TMA kernel (b200attn_decode_ws); with new_k / new_v it also performs the
        # step's KV store (row b -> the slot of request b's last key) before reading the pool
        if fp8:
            raise NotImplementedError("decode_attention: ws_kernel takes 16-bit pools")
        nk = nv = None
        nstride = 0
        if new_k is not None:
            _require_cuda(new_k, new_v)
            nk, nv = new_k.reshape(new_k.shape[0], -1), new_v.reshape(new_v.shape[0], -1)
            if nk.dtype != k_cache.dtype or nv.dtype != v_cache.dtype or nk.stride(1) != 1 or nv.stride(1) != 1 \
                    or nk.stride(0) != nv.stride(0) or nk.shape[1] != Hkv * D or nv.shape[1] != Hkv * D:
                raise ValueError("decode_attention: new_k / new_v must be [B, Hkv*D] rows of the pool dtype "
                                 "with one common token stride")
            nstride = nk.stride(0)
        rc = _lib.load().b200attn_decode_ws(
            q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
            min(k_cache.shape[0], v_cache.shape[0]), o.data_ptr(), o.stride(0), _ptr(lse), kv_indptr.data_ptr(),
            kv_indices.data_ptr(), _ptr(q_rows), _ptr(req_ids), B, Hq, Hkv, D, _fdtype(q), float(sm_scale),
            min_chunk, int(window_left), float(logit_cap), _sinks_ptr(sinks, Hq), _ptr(nk), _ptr(nv), nstride,
            int(skip_slot), _ptr(workspace), ws_bytes, _stream(),
        )
        check(rc, "b200attn_decode_ws")
        return o
    if fp8 and stream_k:
        rc = _lib.load().b200attn_decode_sk_fp8kv(
            q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
            o.data_ptr(), o.stride(0), _ptr(lse), kv_indptr.data_ptr(), kv_indices.data_ptr(),
            _ptr(q_rows), _ptr(req_ids), B, Hq, Hkv, D, _fdtype(q), float(sm_scale), float(k_descale),
            float(v_descale), min_chunk, int(window_left), float(logit_cap), _sinks_ptr(sinks, Hq), _ptr(workspace),
            ws_bytes, _stream(),
        )
        check(rc, "b200attn_decode_sk_fp8kv")
        return o
    if fp8:
        rc = _lib.load().b200attn_decode_fp8kv(
            q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
            o.data_ptr(), o.stride(0), _ptr(lse), kv_indptr.data_ptr(), kv_indices.data_ptr(),
            _ptr(q_rows), _ptr(req_ids), B, Hq, Hkv, D, _fdtype(q), float(sm_scale), float(k_descale),
            float(v_descale), max_splits, min_chunk, int(tail_begin), int(window_left), float(logit_cap),
            _sinks_ptr(sinks, Hq), _ptr(workspace), ws_bytes, _stream(),
        )
        check(rc, "b200attn_decode_fp8kv")
        return o
    if stream_k:  # flat tile-stream scheduling (b200attn_decode_sk): max_splits / tail_begin do not apply
        rc = _lib.load().b200attn_decode_sk(
            q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
            o.data_ptr(), o.stride(0), _ptr(lse), kv_indptr.data_ptr(), kv_indices.data_ptr(),
            _ptr(q_rows), _ptr(req_ids), B, Hq, Hkv, D, _fdtype(q), float(sm_scale), min_chunk,
            int(window_left), float(logit_cap), _sinks_ptr(sinks, Hq), _ptr(workspace), ws_bytes, _stream(),
        )
        check(rc, "b200attn_decode_sk")
        return o
    rc = _lib.load().b200attn_decode(
        q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
        o.data_ptr(), o.stride(0), _ptr(lse), kv_indptr.data_ptr(), kv_indices.data_ptr(),
        _ptr(q_rows), _ptr(req_ids), B, Hq, Hkv, D, _fdtype(q), float(sm_scale), max_splits,
        min_chunk, int(tail_begin), int(warps), int(window_left), float(logit_cap), _sinks_ptr(sinks, Hq),
        _ptr(item_lens), _ptr(workspace), ws_bytes, _stream(),
    )
    check(rc, "b200attn_decode")
    return o


# -------------------------------------------------------------------------- extend
def extend_workspace_bytes(B: int, total_q: int, Hq: int) -> int:
    return _lib.load().b200attn_extend_workspace_bytes(B, total_q, Hq)


def extend_attention(
    q: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    o: torch.Tensor,
    qo_indptr: torch.Tensor,
    kv_indptr: torch.Tensor,
    kv_indices: torch.Tensor,
    sm_scale: float,
    *,
    causal: bool,
    page_size: int,
    max_qo_len: int,
    total_q: Optional[int] = None,
    workspace: Optional[torch.Tensor] = None,
    lse: Optional[torch.Tensor] = None,
    skip_single_rows: bool = False,
    window_left: int = -1,
    logit_cap: float = 0.0,
    sinks: Optional[torch.Tensor] = None,
    custom_mask: Optional[torch.Tensor] = None,
    mask_indptr: Optional[torch.Tensor] = None,
    skip_prefix_custom_mask: bool = True,
    k_descale: float = 1.0,
    v_descale: float = 1.0,
) -> torch.Tensor:
    """Ragged prefix+extend attention: ``q [T,Hq,D]`` -> ``o [T,Hq,D]``.

    uint8 / float8_e4m3fn caches are an e4m3 pool: the kernel casts Q and P to e4m3 and contracts on the pool's bytes
    (``b200attn_extend_fp8kv``, what the reference's Triton kernel does: extend_attention.py:523,596), head_dim 64 / 128.

    Mirrors ``extend_attention_fwd_unified`` (python/sglang/kernels/ops/attention/
    extend_attention.py:963-1410): one KV source (the pool) indexed by
    ``kv_indices`` over prefix + new tokens; causal by absolute position.
    ``skip_single_rows`` leaves requests with one query row untouched (MIXED batches:
    the caller runs them through ``decode_attention`` with ``q_rows`` / ``req_ids``).
    """
    _require_cuda(q, k_cache, v_cache, o, qo_indptr, kv_indptr, kv_indices, workspace, lse)
    if q.dim() != 3 or k_cache.dim() != 3:
        raise ValueError("extend_attention: q must be [T,Hq,D] and caches [slots,Hkv,D]")
    T, Hq, D = q.shape
    B = qo_indptr.numel() - 1
    Hkv = k_cache.shape[1]
    if k_cache.shape[2] != D or v_cache.shape[2] != D:
        raise NotImplementedError("extend_attention: v_head_dim != head_dim is not supported")
    fp8 = k_cache.dtype in (torch.uint8, torch.float8_e4m3fn)
    if (not fp8 and q.dtype != k_cache.dtype) or k_cache.dtype != v_cache.dtype:
        raise TypeError("extend_attention: q and the KV pool must share a dtype (or the pool is e4m3 bytes)")
    if q.stride(2) != 1 or q.stride(1) != D or o.stride(2) != 1 or o.stride(1) != D:
        raise ValueError("extend_attention: q/o heads must be densely packed")
    if k_cache.stride(2) != 1 or k_cache.stride(1) != D or v_cache.stride() != k_cache.stride():
        raise ValueError("extend_attention: KV pool must be NHD contiguous")
    if qo_indptr.dtype != torch.int32 or kv_indptr.dtype != torch.int32 or kv_indices.dtype != torch.int32:
        raise TypeError("extend_attention: indptr / indices must be int32")
    ws_bytes = 0 if workspace is None else workspace.numel() * workspace.element_size()
    if fp8:
        rc = _lib.load().b200attn_extend_fp8kv(
            q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
            min(k_cache.shape[0], v_cache.shape[0]), o.data_ptr(), o.stride(0), _ptr(lse),
            qo_indptr.data_ptr(), kv_indptr.data_ptr(), kv_indices.data_ptr(), B,
            T if total_q is None else total_q, max_qo_len, Hq, Hkv, D, page_size, _fdtype(q),
            (1 if causal else 0) | (2 if skip_single_rows else 0), float(sm_scale), float(k_descale),
            float(v_descale), int(window_left), float(logit_cap), _sinks_ptr(sinks, Hq),
            _mask_ptr(custom_mask, mask_indptr, B), _ptr(mask_indptr), int(bool(skip_prefix_custom_mask)),
            _ptr(workspace), ws_bytes, _stream(),
        )
        check(rc, "b200attn_extend_fp8kv")
        return o
    rc = _lib.load().b200attn_extend(
        q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
        min(k_cache.shape[0], v_cache.shape[0]), o.data_ptr(), o.stride(0), _ptr(lse),
        qo_indptr.data_ptr(), kv_indptr.data_ptr(), kv_indices.data_ptr(), B,
        T if total_q is None else total_q, max_qo_len, Hq, Hkv, D, page_size, _fdtype(q),
        (1 if causal else 0) | (2 if skip_single_rows else 0), float(sm_scale), int(window_left),
        float(logit_cap), _sinks_ptr(sinks, Hq), _mask_ptr(custom_mask, mask_indptr, B), _ptr(mask_indptr),
        int(bool(skip_prefix_custom_mask)), _ptr(workspace), ws_bytes,
        _stream(),
    )
    check(rc, "b200attn_extend")
    return o


# --------------------------------------------------------------------------- merge
def merge_state(
    v_a: torch.Tensor,
    s_a: torch.Tensor,
    v_b: torch.Tensor,
    s_b: torch.Tensor,
    v_merged: Optional[torch.Tensor] = None,
    s_merged: Optional[torch.Tensor] = None,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Mirrors ``merge_state`` (python/sglang/srt/layers/attention/merge_state.py:26-46):
    ``v_* [T,H,D]``, ``s_* [T,H]`` fp32 natural-log LSE; returns ``(v_merged, s_merged)``."""
    _require_cuda(v_a, s_a, v_b, s_b, v_merged, s_merged)
    if v_a.shape != v_b.shape or s_a.shape != s_b.shape or v_a.dim() != 3:
        raise ValueError("merge_state: shape mismatch")
    if s_a.dtype != torch.float32 or s_b.dtype != torch.float32:
        raise TypeError("merge_state: LSE tensors must be float32")
    v_a, v_b, s_a, s_b = v_a.contiguous(), v_b.contiguous(), s_a.contiguous(), s_b.contiguous()
    T, H, D = v_a.shape
    if v_merged is None:
        v_merged = torch.empty_like(v_a)
    if s_merged is None:
        s_merged = torch.empty_like(s_a)
    rc = _lib.load().b200attn_merge_state(
        v_a.data_ptr(), s_a.data_ptr(), v_b.data_ptr(), s_b.data_ptr(), v_merged.data_ptr(),
        s_merged.data_ptr(), T, H, D, _fdtype(v_a), _stream(),
    )
    check(rc, "b200attn_merge_state")
    return v_merged, s_merged
