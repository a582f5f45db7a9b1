"""B200AttnBackend — the drop-in for SGLang's ``AttentionBackend`` plugin API.

Reference interface mirrored here (same method names, argument meaning, return
shapes and error behaviour):
  python/sglang/srt/layers/attention/base_attn_backend.py:36-308   (AttentionBackend ABC)
  python/sglang/srt/layers/attention/torch_native_backend.py:19-401 (oracle semantics)
  python/sglang/srt/layers/attention/triton_backend.py:766-1115     (metadata + graph hooks)

Every compute step is a call into libb200attn.so (hand-written sm_100a CUDA);
there is no Triton, no per-arch dispatch and no CPU / PyTorch fallback.
"""
from __future__ import annotations

import contextlib
import dataclasses
import os
from types import SimpleNamespace
from typing import Dict, Optional

import torch

from . import ops

try:  # inside SGLang the real ABC is used so isinstance checks in the runners pass
    from sglang.srt.layers.attention.base_attn_backend import AttentionBackend as _Base  # type: ignore
except Exception:  # standalone: same contract, defined locally

    class _Base:  # type: ignore
        """Stand-in with the reference ABC's dispatcher (base_attn_backend.py:215-258)."""

        def init_forward_metadata(self, forward_batch):
            self.init_forward_metadata_out_graph(forward_batch)
            self.init_forward_metadata_in_graph(forward_batch)

        def init_forward_metadata_in_graph(self, forward_batch):
            pass

        def on_after_cuda_graph_warmup(self):  # base_attn_backend.py:190-197
            pass

        def forward(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
            mode = forward_batch.forward_mode
            if mode.is_idle():
                return q.new_empty(q.shape[0], layer.tp_q_head_num * layer.v_head_dim)
            if mode.is_decode():
                return self.forward_decode(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)
            return self.forward_extend(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)


@dataclasses.dataclass
class ForwardMetadata:
    """Per-forward metadata; field names chosen to satisfy the reference's
    metadata invariants kit (runner_modes/metadata_invariants.py:22-103)."""

    kv_indptr: torch.Tensor                  # int32 [B+1]
    kv_indices: torch.Tensor                 # int32 [sum kv_len], token-granular slots
    cache_seqlens_int32: Optional[torch.Tensor] = None
    qo_indptr: Optional[torch.Tensor] = None  # int32 [B+1] (extend only)
    page_table: Optional[torch.Tensor] = None  # int32 [B, max_pages] (exported, page_size > 1)
    max_extend_len: int = 0
    num_extend_tokens: int = 0
    max_splits: int = 1
    min_chunk: int = 256
    workspace: Optional[torch.Tensor] = None
    batch_size: int = 0
    # MIXED batches: requests with extend_len == 1 (running decodes, schedule_batch.py:2758-2788)
    # are routed to the HBM-bound decode kernel instead of a 128-row tensor-core tile
    dec_q_rows: Optional[torch.Tensor] = None   # int32 [num_dec] row of q/o
    dec_req_ids: Optional[torch.Tensor] = None  # int32 [num_dec] request index (kv_indptr entry)
    num_dec: int = 0
    dec_max_kv_len: int = 0
    # TARGET_VERIFY (speculative decoding): flat tree mask + int32 offsets (triton_backend.py:911-917)
    custom_mask: Optional[torch.Tensor] = None
    mask_indptr: Optional[torch.Tensor] = None
    # hybrid sliding-window pools (swa_memory_pool.py): the same positions in the SWA pool's index space, used by
    # the SWA layers for both the read (kv_indices) and the write (KVWriteLoc.swa_loc, memory_pool.py:1550-1590)
    swa_kv_indices: Optional[torch.Tensor] = None
    swa_out_cache_loc: Optional[torch.Tensor] = None


def _attn_type_value(layer) -> str:
    t = getattr(layer, "attn_type", None)
    return str(getattr(t, "value", t))


class B200AttnBackend(_Base):
    """Attention backend for B200 behind the reference plugin API."""

    # decode kv splits never finer than this many keys
    MIN_CHUNK = 256

    # deterministic (batch-invariant) mode: decode splits are fixed 256-key tiles, a function of the request's own
    # length only (triton_backend.py:286-304,377-386: SGLANG_TRITON_DECODE_SPLIT_TILE_SIZE = 256)
    DETERMINISTIC_SPLIT_TILE = 256

    def __init__(self, model_runner, skip_padding_slot: bool = True, export_page_table: bool = False,
                 deterministic: Optional[bool] = None, fp8_extend: Optional[str] = None):
        super().__init__()
        if not torch.cuda.is_available():
            raise AssertionError("B200AttnBackend needs a CUDA device (sm_100a); no CPU fallback")
        from . import _lib

        _lib.load()  # fail loudly now if libb200attn.so is missing
        self.device = model_runner.device
        self.page_size = int(getattr(model_runner, "page_size", 1) or 1)
        self.req_to_token_pool = model_runner.req_to_token_pool
        self.token_to_kv_pool = model_runner.token_to_kv_pool
        self.req_to_token = self.req_to_token_pool.req_to_token
        self.max_context_len = int(self.req_to_token.shape[1])
        # Sliding-window models (model_runner.sliding_window_size set, load_model_utils.py:135-148): the per-layer
        # window comes from layer.sliding_window_size.  With a hybrid pool (SWAKVPool: separate, smaller pool for
        # the SWA layers) those layers address the SWA pool through full_to_swa_index_mapping.
        self.sliding_window_size = getattr(model_runner, "sliding_window_size", None)
        self._hybrid_swa = (hasattr(self.token_to_kv_pool, "layers_mapping")
                            and hasattr(self.token_to_kv_pool, "full_to_swa_index_mapping"))
        self.skip_padding_slot = skip_padding_slot
        self.export_page_table = export_page_table
        # --enable-deterministic-inference (server_args): outputs must not depend on what else is in the batch.  The
        # extend kernel is batch-invariant by construction (one pass over a row's keys in order, per-row lazy
        # rescale); decode becomes so by splitting every request into fixed 256-key tiles merged in index order,
        # instead of a split cap chosen from the batch shape.
        if deterministic is None:
            deterministic = bool(getattr(getattr(model_runner, "server_args", None),
                                         "enable_deterministic_inference", False))
        self.deterministic = bool(deterministic)
        # Extend over an e4m3 pool.  "in_kernel" (default): the tcgen05 kernel contracts on the pool's
        # bytes with Q and P cast to e4m3 - the arithmetic of the reference's Triton kernels for such a pool
        # (extend_attention.py:523,596) - no scratch copy, faster than the 16-bit pool.  "dequant": the batch's rows
        # are dequantised once into a 16-bit scratch and Q / P stay 16-bit: closer to the exact math over the
        # dequantised pool than the reference itself, ~5 % slower.
        self.fp8_extend = fp8_extend or os.environ.get("B200ATTN_FP8_EXTEND", "in_kernel")
        if self.fp8_extend not in ("in_kernel", "dequant"):
            raise ValueError("fp8_extend must be 'in_kernel' or 'dequant'")
        self.forward_metadata: Optional[ForwardMetadata] = None
        self.num_sms = ops.sm_count()
        # CUDA-graph state (init_cuda_graph_state)
        self._graph: Optional[SimpleNamespace] = None
        self._graph_splits: Dict[object, object] = {}
        # eager-mode scratch that survives across forwards (grown on demand)
        self._ws: Optional[torch.Tensor] = None
        self._fp8_scratch: Optional[torch.Tensor] = None  # [2, n, Hkv, D] 16-bit K/V operands of extend over an fp8 pool
        self._fp8_iota: Optional[torch.Tensor] = None     # int32 [n]: scratch row of every kv position
        self._fp8_first: Optional[torch.Tensor] = None    # int32 [slots]: first kv position referencing a slot
        # MIXED batches: the decode rows run on a side stream so that they fill the tail of the persistent extend kernel
        self._side: Optional[torch.cuda.Stream] = None
        self._ws_dec: Optional[torch.Tensor] = None

    # ------------------------------------------------------------------ helpers
    def support_triton(self) -> bool:  # base_attn_backend.py:298
        return False

    def _workspace(self, nbytes: int) -> Optional[torch.Tensor]:
        if nbytes <= 0:
            return None
        if self._ws is None or self._ws.numel() < nbytes:
            # zero-initialised ONCE: the decode kernel's split arrival counters live here and are left at zero by
            # every launch (include/b200attn.h, b200attn_decode)
            self._ws = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
        return self._ws

    def _check_layer(self, layer, kwargs) -> None:
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("b200 backend: cross attention is not supported")
        if getattr(layer, "v_head_dim", layer.head_dim) != layer.head_dim:
            raise NotImplementedError("b200 backend: v_head_dim != head_dim is not supported")
        if ((getattr(layer, "k_scale", None) is not None or getattr(layer, "v_scale", None) is not None)
                and not self._fp8_pool()):
            raise NotImplementedError("b200 backend: k_scale / v_scale need a float8_e4m3fn KV pool")

    def _fp8_pool(self) -> bool:
        return getattr(self.token_to_kv_pool, "dtype", None) == torch.float8_e4m3fn

    @staticmethod
    def _descales(layer) -> dict:
        """k_descale / v_descale of an fp8 pool (triton_backend.py:1418-1420: layer.k_scale_float)."""
        def one(name):
            f = getattr(layer, name + "_float", None)
            if f is None:
                t = getattr(layer, name, None)
                if t is None:
                    return 1.0
                if isinstance(t, torch.Tensor) and t.is_cuda:
                    # the reference fills k_scale_float when the scales are loaded (radix_attention.py); without
                    # it this is one device sync per layer, once: the float is kept on the layer
                    if torch.cuda.is_current_stream_capturing():
                        raise RuntimeError("b200 backend: set layer.k_scale_float / v_scale_float before graph capture")
                    f = float(t)
                    setattr(layer, name + "_float", f)
                else:
                    f = float(t)
            return float(f)
        return dict(k_descale=one("k_scale"), v_descale=one("v_scale"))

    @staticmethod
    def _features(layer, kwargs) -> dict:
        """Per-layer flags of the kernels (triton_backend.py:1405-1416,1515-1530): sliding window, logit cap,
        attention sinks.  The window rule is the extend kernel's for BOTH modes — a query at q_id sees keys
        kv_id >= q_id - sliding_window_size (window + 1 keys, extend_attention.py:480-486; the same rule as
        flashinfer's decode metadata, dense_attention.py:40-48) — so that a running decode computes the same row
        whether it arrives in a DECODE batch or as an extend_len == 1 row of a MIXED batch."""
        w = getattr(layer, "sliding_window_size", None)
        sinks = kwargs.get("sinks")
        if sinks is not None and sinks.dtype != torch.float32:
            sinks = sinks.float()
        return dict(window_left=int(w) if w is not None and w > -1 else -1,
                    logit_cap=float(getattr(layer, "logit_cap", 0.0) or 0.0), sinks=sinks)

    def _store(self, layer, forward_batch, k, v) -> None:
        """KV pool write path (memory_pool.py:2331-2456) through the sm_100a store kernel."""
        pool = self.token_to_kv_pool
        kbuf, vbuf = pool.get_kv_buffer(layer.layer_id)
        k = k.view(-1, layer.tp_k_head_num, layer.head_dim)
        v = v.view(-1, layer.tp_v_head_num, layer.v_head_dim)
        loc = forward_batch.out_cache_loc
        if self._is_swa_layer(layer):  # KVWriteLoc.swa_loc: translated once per forward (triton_backend.py:978-982)
            loc = self.forward_metadata.swa_out_cache_loc
        if self._fp8_pool():  # quantising store: div by the scale in the input dtype, then e4m3 (memory_pool.py:2365-2375)
            sc = self._descales(layer)
            ops.store_cache_fp8(k, v, kbuf, vbuf, loc,
                                sc["k_descale"] if getattr(layer, "k_scale", None) is not None else None,
                                sc["v_descale"] if getattr(layer, "v_scale", None) is not None else None,
                                reserved_skip_index=0 if self.skip_padding_slot else -1)
            return
        if k.dtype != kbuf.dtype:  # memory_pool.py:2365-2371
            k, v = k.to(kbuf.dtype), v.to(vbuf.dtype)
        ops.store_cache(k, v, kbuf, vbuf, loc, reserved_skip_index=0 if self.skip_padding_slot else -1)

    def _forward_extend_generic_head_dim(self, q2, o, layer, md, kbuf, vbuf, causal, fp8, kwargs):
        """Head dims the tcgen05 extend kernel is not instantiated for (it takes 64 and 128; the reference's kernel
        tests also sweep 96, 80, ... - test_triton_attention_kernels.py:169-322): every query TOKEN becomes one item
        of the mma.sync decode kernel (any head dim that is a multiple of 8 up to 256) and sees its first
        prefix + j + 1 keys (all keys when not causal).  Same semantics (torch_native_backend.py:61-174), HBM-bound
        instead of tensor-core tiled: a compatibility path, not the fast one."""
        if fp8 or md.custom_mask is not None:
            raise NotImplementedError("b200 backend: head_dim not in (64, 128) with an fp8 pool / custom mask")
        Hq, Hkv, D = layer.tp_q_head_num, layer.tp_k_head_num, layer.head_dim
        T = q2.shape[0]
        if T == 0:
            return o
        req_ids, item_lens = ops.build_token_items(md.qo_indptr, md.kv_indptr, md.batch_size, T, causal)
        kv_idx = self._kv_indices(layer, md)
        q3, o3 = q2.view(-1, Hq, D), o.view(-1, Hq, D)
        step = 32768  # grid.z limit of the decode launch
        for t0 in range(0, T, step):
            n = min(step, T - t0)
            warps, ms, tail = ops.plan_decode(n, Hkv, Hq // Hkv, self.max_context_len, md.min_chunk,
                                              num_sms=self.num_sms)
            ws = self._workspace(ops.decode_workspace_bytes(n, Hq, D, ms))
            ops.decode_attention(q3[t0:t0 + n], kbuf, vbuf, o3[t0:t0 + n], md.kv_indptr, kv_idx, layer.scaling,
                                 max_splits=ms, min_chunk=md.min_chunk, workspace=ws, num_items=n,
                                 req_ids=req_ids[t0:t0 + n], item_lens=item_lens[t0:t0 + n], warps=warps,
                                 tail_begin=tail, **self._features(layer, kwargs))
        return o

    def _deterministic_splits(self) -> int:
        """Split cap of the deterministic mode: with min_chunk = 256 the device-side policy then cuts a request of L
        keys into ceil(L / 256) pieces whatever the batch (csrc/common.cuh split_policy)."""
        return max(1, min(1024, -(-self.max_context_len // self.DETERMINISTIC_SPLIT_TILE)))

    def _is_swa_layer(self, layer) -> bool:
        return bool(self._hybrid_swa and self.token_to_kv_pool.layers_mapping[layer.layer_id][1])

    def _kv_indices(self, layer, md) -> torch.Tensor:
        return md.swa_kv_indices if self._is_swa_layer(layer) else md.kv_indices

    def _build_swa(self, fb, md, kv_lens, qo_lens=None, out_indices=None, out_loc=None) -> None:
        """Hybrid pool: the batch's kv positions and write slots in the SWA pool's index space."""
        if not self._hybrid_swa:
            return
        mapping = self.token_to_kv_pool.full_to_swa_index_mapping
        if mapping is None:
            raise RuntimeError("b200 backend: hybrid SWA pool without full_to_swa_index_mapping")
        if out_indices is None:
            out_indices = torch.empty_like(md.kv_indices)
        scratch = torch.empty_like(md.kv_indptr)
        ops.build_kv_indices(self.req_to_token, fb.req_pool_indices, kv_lens, scratch, out_indices,
                             qo_lens=qo_lens, qo_indptr=torch.empty_like(scratch) if qo_lens is not None else None,
                             slot_map=mapping)
        md.swa_kv_indices = out_indices
        loc = fb.out_cache_loc
        if loc is not None:
            md.swa_out_cache_loc = ops.translate_slots(loc, mapping, out=None if out_loc is None else out_loc[:loc.numel()])

    # ------------------------------------------------------------------ metadata
    def _build(self, fb, kv_lens, kv_indptr, kv_indices, qo_lens=None, qo_indptr=None) -> None:
        ops.build_kv_indices(self.req_to_token, fb.req_pool_indices, kv_lens, kv_indptr, kv_indices,
                             qo_lens=qo_lens, qo_indptr=qo_indptr)

    def _page_table(self, fb, B: int) -> Optional[torch.Tensor]:
        if not self.export_page_table or self.page_size <= 1:
            return None
        max_pages = -(-self.max_context_len // self.page_size)
        pt = torch.zeros((B, max_pages), dtype=torch.int32, device=self.device)
        ops.build_page_table(self.req_to_token, fb.req_pool_indices, fb.seq_lens, pt, self.page_size)
        return pt

    def init_forward_metadata(self, forward_batch) -> None:
        """Eager entry point (base_attn_backend.py:65-71)."""
        fb = forward_batch
        mode = fb.forward_mode
        if mode.is_idle():
            self.forward_metadata = None
            return
        B = int(fb.batch_size)
        dev = self.device
        total = getattr(fb, "seq_lens_sum", 0) or 0
        if total <= 0:
            cpu = getattr(fb, "seq_lens_cpu", None)
            total = int(cpu.sum()) if cpu is not None else B * self.max_context_len
        kv_indptr = torch.empty(B + 1, dtype=torch.int32, device=dev)
        kv_indices = torch.empty(max(int(total), 1), dtype=torch.int32, device=dev)
        if mode.is_decode():
            self._build(fb, fb.seq_lens, kv_indptr, kv_indices)
            md = ForwardMetadata(kv_indptr=kv_indptr, kv_indices=kv_indices, batch_size=B,
                                 cache_seqlens_int32=fb.seq_lens.to(torch.int32),
                                 page_table=self._page_table(fb, B), min_chunk=self.MIN_CHUNK)
            cpu = getattr(fb, "seq_lens_cpu", None)
            md._max_kv_len = int(cpu.max()) if cpu is not None else self.max_context_len
            self._build_swa(fb, md, fb.seq_lens)
            self.forward_metadata = md
            return
        if mode.is_target_verify():
            return self._init_target_verify(fb, B, kv_indptr)
        if mode.is_draft_extend_v2() or getattr(fb, "spec_info", None) is not None:
            raise NotImplementedError("b200 backend: draft-extend speculative modes are not supported")
        if getattr(fb, "encoder_lens", None) is not None:
            raise NotImplementedError("b200 backend: encoder_lens (cross attention) is not supported")
        qo_indptr = torch.empty(B + 1, dtype=torch.int32, device=dev)
        self._build(fb, fb.seq_lens, kv_indptr, kv_indices, qo_lens=fb.extend_seq_lens, qo_indptr=qo_indptr)
        ext_cpu = getattr(fb, "extend_seq_lens_cpu", None)
        if ext_cpu is not None:
            max_ext, tot_ext = max(ext_cpu), sum(ext_cpu)
        else:  # gpu-only mode: one sync, eager path only
            max_ext, tot_ext = int(fb.extend_seq_lens.max()), int(fb.extend_seq_lens.sum())
        md = ForwardMetadata(
            kv_indptr=kv_indptr, kv_indices=kv_indices, qo_indptr=qo_indptr, batch_size=B,
            cache_seqlens_int32=fb.seq_lens.to(torch.int32), page_table=self._page_table(fb, B),
            max_extend_len=int(max_ext), num_extend_tokens=int(tot_ext), min_chunk=self.MIN_CHUNK)
        if ext_cpu is not None and 1 in ext_cpu:
            rows, reqs, off = [], [], 0
            for b, e in enumerate(ext_cpu):
                if e == 1:
                    rows.append(off)
                    reqs.append(b)
                off += e
            both = torch.tensor([rows, reqs], dtype=torch.int32).pin_memory().to(dev, non_blocking=True)
            md.dec_q_rows, md.dec_req_ids, md.num_dec = both[0], both[1], len(rows)
            md.max_extend_len = max([e for e in ext_cpu if e != 1], default=0)
            pre_cpu = getattr(fb, "extend_prefix_lens_cpu", None)
            md.dec_max_kv_len = (max(pre_cpu[b] for b in reqs) + 1) if pre_cpu is not None else self.max_context_len
        self._build_swa(fb, md, fb.seq_lens, qo_lens=fb.extend_seq_lens)
        self.forward_metadata = md

    def _init_target_verify(self, fb, B: int, kv_indptr: torch.Tensor) -> None:
        """TARGET_VERIFY (triton_backend.py:859-917): every request verifies ``draft_token_num`` tokens under the
        tree mask ``spec_info.custom_mask`` (row-major ``[draft, seq_len + draft]`` blocks, prefix keys always
        visible).  ``seq_lens`` counts the cached tokens only; the draft tokens' K/V are stored by forward_extend at
        ``out_cache_loc`` and appended to ``kv_indices`` here, so the kernels still see one KV source."""
        if self._hybrid_swa:
            raise NotImplementedError("b200 backend: TARGET_VERIFY over a hybrid sliding-window pool is not supported")
        spec = fb.spec_info
        nd = int(spec.draft_token_num)
        dev = self.device
        seq = fb.seq_lens.to(torch.int64)
        kv_lens = seq + nd
        total = getattr(fb, "seq_lens_sum", None)
        total = (int(total) if total else B * self.max_context_len) + B * nd
        kv_indices = torch.zeros(total, dtype=torch.int32, device=dev)
        qo_lens = torch.full((B,), nd, dtype=torch.int32, device=dev)
        qo_indptr = torch.empty(B + 1, dtype=torch.int32, device=dev)
        self._build(fb, kv_lens, kv_indptr, kv_indices, qo_lens=qo_lens, qo_indptr=qo_indptr)
        tail = (kv_indptr[:B].to(torch.int64) + seq)[:, None] + torch.arange(nd, device=dev)[None, :]
        kv_indices[tail.reshape(-1)] = fb.out_cache_loc[:B * nd].to(torch.int32)
        mask_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
        mask_indptr[1:] = torch.cumsum(nd * kv_lens, 0).to(torch.int32)
        mask = spec.custom_mask
        if mask.dtype not in (torch.bool, torch.uint8):
            mask = mask != 0
        self.forward_metadata = ForwardMetadata(
            kv_indptr=kv_indptr, kv_indices=kv_indices, qo_indptr=qo_indptr, batch_size=B,
            cache_seqlens_int32=kv_lens.to(torch.int32), max_extend_len=nd, num_extend_tokens=B * nd,
            min_chunk=self.MIN_CHUNK, custom_mask=mask.contiguous(), mask_indptr=mask_indptr)

    # --- CUDA-graph hooks (base_attn_backend.py:73-107,159-197; triton_backend.py:1014-1115)
    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int, *, num_q_heads: Optional[int] = None,
                              head_dim: Optional[int] = None, max_splits_cap: int = 16) -> None:
        dev = self.device
        if self.deterministic:
            max_splits_cap = max(max_splits_cap, self._deterministic_splits())
        g = SimpleNamespace()
        g.max_bs = max_bs
        g.kv_indptr = torch.zeros(max_bs + 1, dtype=torch.int32, device=dev)
        g.kv_indices = torch.zeros(max_bs * self.max_context_len, dtype=torch.int32, device=dev)
        g.cache_seqlens_int32 = torch.zeros(max_bs, dtype=torch.int32, device=dev)
        g.qo_indptr = torch.zeros(max_bs + 1, dtype=torch.int32, device=dev)  # full-graph / breakable-graph extend
        g.max_num_tokens = max_num_tokens
        g.max_splits_cap = max_splits_cap
        g.swa_kv_indices = g.swa_out_cache_loc = None
        if self._hybrid_swa:  # triton_backend.py:1100-1108
            g.swa_kv_indices = torch.zeros_like(g.kv_indices)
            g.swa_out_cache_loc = torch.zeros(max(max_num_tokens, max_bs), dtype=torch.int64, device=dev)
        g.workspace = None
        if num_q_heads is not None and head_dim is not None:
            nbytes = ops.decode_workspace_bytes(max_bs, num_q_heads, head_dim, max_splits_cap)
            g.workspace = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=dev)
        self._graph = g

    def get_cuda_graph_seq_len_fill_value(self) -> int:
        return 1  # padded rows read slot 0 with one key (triton_backend.py:1257-1258)

    def init_forward_metadata_out_graph(self, forward_batch, in_capture: bool = False) -> None:
        """Refill the static decode metadata before capture / replay (sync-free)."""
        fb = forward_batch
        mode = fb.forward_mode
        spec = mode.is_target_verify() or mode.is_draft_extend_v2() or getattr(fb, "spec_info", None) is not None
        if self._graph is None or spec or getattr(fb, "encoder_lens", None) is not None:
            if in_capture:
                raise NotImplementedError("b200 backend: only DECODE / EXTEND / MIXED can be captured in a CUDA graph")
            return B200AttnBackend.init_forward_metadata(self, fb)
        if mode.is_idle():
            self.forward_metadata = None
            return
        g = self._graph
        B = int(fb.batch_size)
        if B > g.max_bs:
            raise ValueError(f"b200 backend: batch {B} exceeds cuda-graph max_bs {g.max_bs}")
        if not mode.is_decode():
            # Full-graph / breakable-graph prefill (prefill_cuda_graph_runner.py:985-1030): the persistent extend
            # kernel derives its work list from qo_indptr ON THE DEVICE, so one captured launch serves any ragged
            # batch that fits the static buffers.  Padded requests carry extend_len 0 and are skipped.  No row is
            # routed to the decode kernel here (that split is a host decision).
            self._build(fb, fb.seq_lens, g.kv_indptr, g.kv_indices, qo_lens=fb.extend_seq_lens,
                        qo_indptr=g.qo_indptr)
            g.cache_seqlens_int32[:B].copy_(fb.seq_lens[:B])
            md = ForwardMetadata(
                kv_indptr=g.kv_indptr, kv_indices=g.kv_indices, qo_indptr=g.qo_indptr, batch_size=B,
                cache_seqlens_int32=g.cache_seqlens_int32[:B], max_extend_len=int(g.max_num_tokens),
                num_extend_tokens=0, min_chunk=self.MIN_CHUNK)
            md._graph = True
            self._build_swa(fb, md, fb.seq_lens, qo_lens=fb.extend_seq_lens, out_indices=g.swa_kv_indices,
                            out_loc=g.swa_out_cache_loc)
            self.forward_metadata = md
            return
        self._build(fb, fb.seq_lens, g.kv_indptr, g.kv_indices)
        g.cache_seqlens_int32[:B].copy_(fb.seq_lens[:B])
        md = ForwardMetadata(
            kv_indptr=g.kv_indptr, kv_indices=g.kv_indices, batch_size=B,
            cache_seqlens_int32=g.cache_seqlens_int32[:B], min_chunk=self.MIN_CHUNK,
            workspace=g.workspace)
        md._graph = True
        self._build_swa(fb, md, fb.seq_lens, out_indices=g.swa_kv_indices, out_loc=g.swa_out_cache_loc)
        self.forward_metadata = md

    def init_forward_metadata_in_graph(self, forward_batch) -> None:
        pass  # nothing is recomputed inside the graph: the kernels read the static buffers

    # Breakable-CUDA-graph prefill with captured metadata (base_attn_backend.py:124-129,163-188): the metadata
    # object created at capture keeps its tensor addresses; replay refills them in place.
    use_captured_forward_metadata_for_breakable_cuda_graph = True

    def init_forward_metadata_for_breakable_cuda_graph_capture(self, forward_batch):
        if self._graph is None:
            raise RuntimeError("b200 backend: call init_cuda_graph_state before graph capture")
        self.init_forward_metadata_out_graph(forward_batch, in_capture=True)
        return self.forward_metadata

    def prepare_forward_metadata_for_breakable_cuda_graph_replay(self, capture_metadata, forward_batch, *,
                                                                 static_forward_batch=None) -> None:
        fb = static_forward_batch if static_forward_batch is not None else forward_batch
        self.init_forward_metadata_out_graph(fb)
        md = self.forward_metadata
        if md is not None and capture_metadata is not None:
            # same static tensors; refresh the per-replay python fields on the captured object
            capture_metadata.batch_size = md.batch_size
            capture_metadata.cache_seqlens_int32 = md.cache_seqlens_int32
            self.forward_metadata = capture_metadata

    # ------------------------------------------------------------------ forward
    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        self._check_layer(layer, kwargs)
        md = self.forward_metadata
        Hq, Hkv, D = layer.tp_q_head_num, layer.tp_k_head_num, layer.head_dim
        q2 = q.reshape(-1, Hq * D)
        o = q2.new_empty(q2.shape[0], Hq * D)
        if save_kv_cache and k is not None:  # store BEFORE the read: the new token attends to itself
            self._store(layer, forward_batch, k, v)
        kbuf, vbuf = self.token_to_kv_pool.get_kv_buffer(layer.layer_id)
        B = md.batch_size
        G = Hq // Hkv
        fp8 = self._fp8_pool()
        graph = getattr(md, "_graph", False)
        warps, tail_begin = 4, -1
        if self.deterministic:
            ms = self._deterministic_splits()
        elif fp8:
            cps = ops.decode_ctas_per_sm(True)
            if graph:  # the split cap is baked into the captured grid: fix it per batch size
                ms = self._graph_splits.get(B)
                if ms is None:
                    ms = min(self._graph.max_splits_cap, ops.pick_max_splits(
                        B, Hkv, G, self.max_context_len, md.min_chunk, num_sms=self.num_sms, ctas_per_sm=cps))
                    self._graph_splits[B] = ms
            else:
                ms = ops.pick_max_splits(B, Hkv, G, getattr(md, "_max_kv_len", self.max_context_len), md.min_chunk,
                                         num_sms=self.num_sms, ctas_per_sm=cps)
            if ms == 1:  # whole waves stay unsplit; a nearly empty last wave is split (ops.pick_tail_split)
                tail_begin, k = ops.pick_tail_split(B, Hkv, G, num_sms=self.num_sms, ctas_per_sm=cps)
                ms, tail_begin = (k, tail_begin) if k > 1 else (1, -1)
        else:
            # 16-bit pools: (warps per CTA, split cap, tail split) from the batch shape (ops.plan_decode); under a
            # graph the plan is a function of the bucket size only (kv lengths bounded by the context length)
            key = ("plan", B, Hkv, G)
            plan = self._graph_splits.get(key) if graph else None
            if plan is None:
                plan = ops.plan_decode(B, Hkv, G, self.max_context_len if graph else
                                       getattr(md, "_max_kv_len", self.max_context_len), md.min_chunk,
                                       num_sms=self.num_sms)
                if graph:
                    plan = (plan[0], min(plan[1], self._graph.max_splits_cap), plan[2])
                    self._graph_splits[key] = plan
            warps, ms, tail_begin = plan
        need = ops.decode_workspace_bytes(B, Hq, D, ms)
        if graph:
            ws = md.workspace
            if need > 0 and (ws is None or ws.numel() < need):
                if torch.cuda.is_current_stream_capturing():
                    raise RuntimeError("b200 backend: pass num_q_heads/head_dim to init_cuda_graph_state")
                ws = self._graph.workspace = torch.zeros(
                    ops.decode_workspace_bytes(self._graph.max_bs, Hq, D, self._graph.max_splits_cap),
                    dtype=torch.uint8, device=self.device)
                md.workspace = ws
        else:
            ws = self._workspace(need)
        ops.decode_attention(q2.view(-1, Hq, D), kbuf, vbuf, o.view(-1, Hq, D), md.kv_indptr,
                             self._kv_indices(layer, md),
                             layer.scaling, max_splits=ms, min_chunk=md.min_chunk, workspace=ws, num_items=B,
                             tail_begin=tail_begin, warps=warps, **self._features(layer, kwargs),
                             **(self._descales(layer) if fp8 else {}))
        return o

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        self._check_layer(layer, kwargs)
        md = self.forward_metadata
        Hq, D = layer.tp_q_head_num, layer.head_dim
        q2 = q.reshape(-1, Hq * D)
        o = getattr(forward_batch, "_attn_output", None)
        if o is None or o.shape != q2.shape or o.dtype != q2.dtype:
            o = q2.new_empty(q2.shape[0], Hq * D)
        if save_kv_cache and k is not None:  # torch_native_backend.py:298-301
            self._store(layer, forward_batch, k, v)
        # torch_native_backend.py:308-310 — DECODER_BIDIRECTIONAL stays causal in the oracle
        causal = not (getattr(layer, "is_cross_attention", False) or _attn_type_value(layer) == "encoder_only")
        mask_kw = {}
        if md.custom_mask is not None:  # TARGET_VERIFY: the tree mask replaces the causal mask (triton_backend.py:1376-1379)
            causal = False
            mask_kw = dict(custom_mask=md.custom_mask, mask_indptr=md.mask_indptr, skip_prefix_custom_mask=True)
        kbuf, vbuf = self.token_to_kv_pool.get_kv_buffer(layer.layer_id)
        B = md.batch_size
        fp8 = self._fp8_pool()
        dsc = self._descales(layer) if fp8 else {}
        if D not in (64, 128):
            return self._forward_extend_generic_head_dim(q2, o, layer, md, kbuf, vbuf, causal, fp8, kwargs)
        overlap = md.max_extend_len > 0 and md.num_dec > 0 and os.environ.get("B200ATTN_MIXED_OVERLAP", "1") != "0"
        if overlap:
            # fork AFTER the store and BEFORE the extend launch: the extend kernel's 148 persistent CTAs take the SMs
            # first, the decode rows' CTAs (independent rows of o) start as extend CTAs retire
            cur = torch.cuda.current_stream()
            if self._side is None:
                self._side = torch.cuda.Stream()
            self._side.wait_stream(cur)
        if md.max_extend_len > 0:
            ws = self._workspace(ops.extend_workspace_bytes(B, q2.shape[0], Hq))
            kv_idx = self._kv_indices(layer, md)
            ek, ev, eidx, escale = kbuf, vbuf, kv_idx, layer.scaling
            ext_dsc = {}
            if fp8 and self.fp8_extend == "in_kernel":  # head_dim 64 and 128 (the two the tcgen05 kernel has)
                ext_dsc = dsc  # b200attn_extend_fp8kv: contracts on the e4m3 bytes of the pool itself
            elif fp8:
                # the tcgen05 kernel takes 16-bit K/V operands: gather the batch's rows once into a 16-bit scratch
                # (K exact, k_descale on the score scale; v_descale folded into the V rows), slot ids become 0..n-1
                n = kv_idx.numel()
                Hkv = layer.tp_k_head_num
                graph = getattr(md, "_graph", False)
                if self._fp8_scratch is None or self._fp8_scratch.shape[1] < n or self._fp8_scratch.dtype != q2.dtype \
                        or self._fp8_scratch.shape[2:] != (Hkv, D):
                    if torch.cuda.is_current_stream_capturing():
                        raise RuntimeError("b200 backend: run one eager extend before capturing over an fp8 KV pool")
                    self._fp8_scratch = torch.empty((2, n, Hkv, D), dtype=q2.dtype, device=self.device)
                    self._fp8_iota = torch.empty(n, dtype=torch.int32, device=self.device)
                slots = kbuf.shape[0]
                if self._fp8_first is None or self._fp8_first.numel() < slots:
                    if torch.cuda.is_current_stream_capturing():
                        raise RuntimeError("b200 backend: run one eager extend before capturing over an fp8 KV pool")
                    self._fp8_first = torch.empty(slots, dtype=torch.int32, device=self.device)
                ek, ev, eidx = self._fp8_scratch[0], self._fp8_scratch[1], self._fp8_iota
                # a slot shared by several requests (radix hit) is converted once; eidx[i] = the scratch row holding
                # row i's slot.  Graph mode: the launch covers the static kv_indices buffer, the live row count is
                # kv_indptr[B] on the device
                ops.kv_dequant_gather_unique(kbuf, vbuf, kv_idx, n, ek, ev, self._fp8_first, eidx, 1.0,
                                             dsc["v_descale"], n_dev=md.kv_indptr[B:B + 1] if graph else None)
                escale = layer.scaling * dsc["k_descale"]
            ops.extend_attention(q2.view(-1, Hq, D), ek, ev, o.view(-1, Hq, D), md.qo_indptr, md.kv_indptr,
                                 eidx, escale, causal=causal, page_size=self.page_size,
                                 max_qo_len=md.max_extend_len,
                                 total_q=q2.shape[0] if getattr(md, "_graph", False) else md.num_extend_tokens,
                                 workspace=ws,
                                 skip_single_rows=md.num_dec > 0, **mask_kw, **self._features(layer, kwargs),
                                 **ext_dsc)
        if md.num_dec > 0:  # a single query row sees all of its keys, causal or not
            Hkv = layer.tp_k_head_num
            dwarps, dtail = 4, -1
            if self.deterministic:
                ms = self._deterministic_splits()
            elif fp8:
                ms = ops.pick_max_splits(md.num_dec, Hkv, Hq // Hkv, md.dec_max_kv_len, md.min_chunk,
                                         num_sms=self.num_sms, ctas_per_sm=ops.decode_ctas_per_sm(True))
            else:
                dwarps, ms, dtail = ops.plan_decode(md.num_dec, Hkv, Hq // Hkv, md.dec_max_kv_len, md.min_chunk,
                                                    num_sms=self.num_sms)
            need = ops.decode_workspace_bytes(md.num_dec, Hq, D, ms)
            with (torch.cuda.stream(self._side) if overlap else contextlib.nullcontext()):
                if overlap:  # its own workspace: the extend kernel is using self._ws concurrently
                    if need > 0 and (self._ws_dec is None or self._ws_dec.numel() < need):
                        self._ws_dec = torch.zeros(need, dtype=torch.uint8, device=self.device)
                    ws = self._ws_dec if need > 0 else None
                else:
                    ws = self._workspace(need)
                ops.decode_attention(q2.view(-1, Hq, D), kbuf, vbuf, o.view(-1, Hq, D), md.kv_indptr,
                                     self._kv_indices(layer, md),
                                     layer.scaling, max_splits=ms, min_chunk=md.min_chunk, workspace=ws,
                                     q_rows=md.dec_q_rows, req_ids=md.dec_req_ids, num_items=md.num_dec,
                                     warps=dwarps, tail_begin=dtail, **self._features(layer, kwargs), **dsc)
            if overlap:
                cur.wait_stream(self._side)
        return o


def register(name: str = "b200") -> None:
    """Register with SGLang's backend registry (attention_registry.py:31-39) and CLI
    choices (server_args.py:416-417).  Call from an ``sglang.srt.plugins`` entry point."""
    from sglang.srt.layers.attention.attention_registry import register_attention_backend  # type: ignore
    from sglang.srt.server_args import add_attention_backend_choices  # type: ignore

    add_attention_backend_choices([name])

    @register_attention_backend(name)
    def _create(runner):
        return B200AttnBackend(runner)
