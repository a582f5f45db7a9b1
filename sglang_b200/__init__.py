"""sglang_b200 — B200-native (sm_100a) attention hot path behind SGLang's AttentionBackend API.

Layout:
  csrc/        hand-written CUDA kernels + the C ABI (include/b200attn.h) -> libb200attn.so
  _lib.py      ctypes binding of the C ABI (fails loudly if the library is missing)
  ops.py       tensor-level operator wrappers (store_cache, decode/extend attention, merge_state, ...)
  host.py      mirror of the reference data carriers (ForwardBatch, pools, RadixAttention)
  backend.py   B200AttnBackend (init_forward_metadata / forward_decode / forward_extend / graph hooks)
"""
from .host import (  # noqa: F401
    AttentionType,
    ForwardBatch,
    ForwardMode,
    KVWriteLoc,
    MHATokenToKVPool,
    RadixAttention,
    ReqToTokenPool,
    SWAKVPool,
)
from .backend import B200AttnBackend, ForwardMetadata  # noqa: F401

__version__ = "0.1.0"
