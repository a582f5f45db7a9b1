"""The one collective on the attention path: the sum all-reduce of the o_proj output (RowParallelLinear after
RadixAttention -> ``GroupCoordinator.all_reduce``, python/sglang/srt/distributed/parallel_state.py:648-757).

``NvlsAllReduce`` mirrors the reference's symmetric-memory communicator
(python/sglang/srt/distributed/device_communicators/torch_symm_mem.py:34-200: same constructor arguments, ``disabled``
/ ``max_size`` / ``buffer`` attributes, ``should_torch_symm_mem_allreduce`` and ``all_reduce(inp, out=)``), so a
``GroupCoordinator`` can hold it as its ``torch_symm_mem_comm``.  The difference is underneath: the reference stages the
input into the symmetric buffer with one copy kernel, calls torch's multimem op and copies the result out (three
launches); here ``b200attn_allreduce_nvls`` (csrc/allreduce.cu) does stage-in, the in-switch reduce-scatter +
all-gather and stage-out in ONE launch.  PyTorch is plumbing only: it allocates the symmetric memory and exchanges the
handles (``torch.distributed._symmetric_memory``).

No fallback: without a CUDA device, symmetric memory or an NVLS multicast address the communicator stays ``disabled``
and ``all_reduce`` returns None, exactly as the reference's does (the caller then takes its next method, NCCL).
"""
import ctypes
from typing import Optional, Union

import torch
import torch.distributed as dist

from . import _lib
from .ops import _stream, check

__all__ = ["NvlsAllReduce"]


class NvlsAllReduce:
    """Sum all-reduce over NVLS for one process group (one process per GPU, single node)."""

    DEFAULT_MAX_SIZE = 64 * 1024 * 1024  # bytes of payload; a [8192, 4096] bf16 prefill chunk

    def __init__(self, group, device: Union[int, str, torch.device], max_size: Optional[int] = None,
                 dtype: torch.dtype = torch.bfloat16):
        self.disabled = True
        self.buffer = None
        self.max_size = 0
        self.group = group
        self.dtype = dtype
        if isinstance(device, int):
            device = torch.device(f"cuda:{device}")
        elif isinstance(device, str):
            device = torch.device(device)
        self.device = device
        if device.type != "cuda" or not torch.cuda.is_available():
            return
        try:
            import torch.distributed._symmetric_memory as symm
        except ImportError:
            return
        self.world_size = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world_size < 2 or self.world_size > 16:
            return
        lib = _lib.load()
        self._flag_bytes = int(lib.b200attn_allreduce_flag_bytes())
        self._small_bytes = int(lib.b200attn_allreduce_small_bytes())
        self.max_size = int(max_size or self.DEFAULT_MAX_SIZE)
        # one symmetric allocation: [flag block | one-shot region (two halves) | payload]
        self._head = self._flag_bytes + self._small_bytes
        self._raw = symm.empty(self._head + self.max_size, dtype=torch.uint8, device=device)
        self._raw.zero_()
        torch.cuda.synchronize(device)
        handle = symm.rendezvous(self._raw, group.group_name)
        if not getattr(handle, "multicast_ptr", 0):
            self._raw = None
            self.max_size = 0
            return
        self._handle = handle
        handle.barrier()  # every rank's flag block is zero before anyone signals
        self._flag_ptrs = (ctypes.c_void_p * self.world_size)(*[int(p) for p in handle.buffer_ptrs])
        self._local = int(handle.buffer_ptrs[self.rank]) + self._head
        self._mc = int(handle.multicast_ptr) + self._head
        self._small_local = int(handle.buffer_ptrs[self.rank]) + self._flag_bytes
        self._small_mc = int(handle.multicast_ptr) + self._flag_bytes
        self.buffer = self._raw[self._head:].view(dtype)
        self.disabled = False

    # ---- reference interface (torch_symm_mem.py:133-200) -------------------------------------------------------
    def should_torch_symm_mem_allreduce(self, inp: torch.Tensor) -> bool:
        if self.disabled or inp.device != self.device or inp.dtype not in _DTYPE_CODE_AR:
            return False
        nbytes = inp.numel() * inp.element_size()
        return nbytes % 16 == 0 and inp.is_contiguous() and inp.data_ptr() % 16 == 0 and nbytes <= self.max_size

    def all_reduce(self, inp: torch.Tensor, *, out: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """Sum of ``inp`` over the group (every rank calls it with the same shape); None when not eligible."""
        if not self.should_torch_symm_mem_allreduce(inp):
            return None
        if out is None:
            out = torch.empty_like(inp)
        self._launch(inp.data_ptr(), out.data_ptr(), inp.numel() * inp.element_size(), inp.dtype)
        return out

    # ---- producer-side fusion: let the o_proj GEMM write its output straight into the symmetric payload --------
    def payload(self, shape, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
        """A view of the symmetric payload to be used as the GEMM's ``out=``; follow with all_reduce_payload()."""
        dtype = dtype or self.dtype
        n = 1
        for s in shape:
            n *= int(s)
        if self.disabled or n * dtype.itemsize > self.max_size:
            raise RuntimeError("NvlsAllReduce.payload: communicator disabled or payload too large")
        return self._raw[self._head: self._head + n * dtype.itemsize].view(dtype).view(*shape)

    def all_reduce_payload(self, t: torch.Tensor) -> torch.Tensor:
        """In-place all-reduce of a tensor returned by payload(): no staging at either end."""
        nbytes = t.numel() * t.element_size()
        if t.data_ptr() != self._local or nbytes % 16 != 0:
            raise RuntimeError("NvlsAllReduce.all_reduce_payload: not a payload() tensor / size not a multiple of 16 B")
        self._launch(0, 0, nbytes, t.dtype)
        return t

    def _launch(self, in_ptr: int, out_ptr: int, nbytes: int, dtype: torch.dtype) -> None:
        rc = _lib.load().b200attn_allreduce_nvls(in_ptr or None, out_ptr or None, self._local, self._mc,
                                                 self._small_local, self._small_mc, self._flag_ptrs, self.rank,
                                                 self.world_size, nbytes, _DTYPE_CODE_AR[dtype], _stream())
        check(rc, "b200attn_allreduce_nvls")


_DTYPE_CODE_AR = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}
