"""Test infrastructure: make the UNMODIFIED reference importable.

In the build container the reference lives at /root/reference/python; on the GPU box (no /root/reference)
the tree staged by oracle/make_ref.sh under oracle/_ref/python is used instead.  Two third-party modules the
reference imports are absent from this image (orjson, IPython: SURVEY.md 9.5) and are replaced by json-backed stubs.
Only tests/, oracle/ generators and tools/ comparators call this; nothing under sglang_b200/ does.
"""
from __future__ import annotations

import json
import os
import sys
import types
from typing import Optional

HERE = os.path.dirname(os.path.abspath(__file__))
CANDIDATES = ("/root/reference/python", os.path.join(HERE, "_ref", "python"))


def reference_root() -> Optional[str]:
    for p in CANDIDATES:
        if os.path.isdir(os.path.join(p, "sglang", "srt")):
            return p
    return None


def install() -> str:
    """Put the reference first on sys.path (+ the stubs); returns the root used.  Raises ImportError if absent."""
    root = reference_root()
    if root is None:
        raise ImportError("reference not available: neither /root/reference/python nor oracle/_ref/python exists "
                          "(run oracle/make_ref.sh in the build container)")
    if root not in sys.path:
        sys.path.insert(0, root)
    if "orjson" not in sys.modules:
        try:
            import orjson  # noqa: F401
        except ImportError:
            o = types.ModuleType("orjson")
            o.dumps = lambda x, *a, **k: json.dumps(x).encode()
            o.loads = json.loads
            o.OPT_NON_STR_KEYS = o.OPT_SERIALIZE_NUMPY = o.OPT_INDENT_2 = o.OPT_SORT_KEYS = 0
            o.JSONDecodeError = ValueError
            o.JSONEncodeError = TypeError
            sys.modules["orjson"] = o
    if "IPython" not in sys.modules:
        try:
            import IPython  # noqa: F401
        except ImportError:
            ip = types.ModuleType("IPython")
            ip.__path__ = []
            sys.modules["IPython"] = ip
            ipd = types.ModuleType("IPython.display")
            ipd.HTML = object
            ipd.display = lambda *a, **k: None
            sys.modules["IPython.display"] = ipd
    import torch

    if torch.cuda.is_available():
        # with a CUDA device the reference imports its AOT wheel at module scope throughout srt/ (guarded by
        # `_is_cuda`); without one it never does, so the stand-in is only installed here
        try:
            import sgl_kernel  # noqa: F401
        except ImportError:
            if not any(isinstance(f, _MissingFinder) for f in sys.meta_path):
                sys.meta_path.append(_MissingFinder("sgl_kernel"))
    return root


class _Absent:
    """Placeholder for any object of a package that is not installed: attribute access yields another placeholder
    (module-scope constant tables such as `scalar_types.uint4b8` import fine), CALLING it raises."""

    def __init__(self, name):
        self.__dict__["_name"] = name

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Absent(f"{self._name}.{name}")

    def __call__(self, *a, **k):
        raise NotImplementedError(f"{self._name} is not available in this image")

    def __repr__(self):
        return f"<absent {self._name}>"


class _Missing(types.ModuleType):
    """Stand-in for a compiled package the reference imports at module scope but that this image lacks
    (sgl_kernel: the reference's own AOT wheel, not built here).  Every attribute is an `_Absent` placeholder, so
    importing the reference works and only kernels that are really absent fail, loudly, when CALLED."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Absent(f"{self.__name__}.{name}")


class _MissingFinder:
    """Resolves `<pkg>` and every `<pkg>.<sub>` import to a _Missing module."""

    def __init__(self, pkg):
        self.pkg = pkg

    def find_spec(self, fullname, path=None, target=None):
        import importlib.machinery

        if fullname == self.pkg or fullname.startswith(self.pkg + "."):
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _Missing(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


def import_runner_modes():
    """Import the kit's runner-mode helpers (cuda-graph decode runner, TARGET_VERIFY runner).  Their package pulls in
    every attention-method kit, one of which imports the reference's own AOT wheel `sgl_kernel` (not built in this
    image) at module scope; the dense path never calls it.  The stand-in is installed ONLY while that import runs,
    after the dense kit (and with it the backends' `try: import sgl_kernel` probes) has been imported for real."""
    import importlib

    importlib.import_module("sglang.test.kits.attention_unittest.attention_methods.dense_attention")
    finder = None
    try:
        import sgl_kernel  # noqa: F401
    except ImportError:
        if not any(isinstance(f, _MissingFinder) for f in sys.meta_path):
            finder = _MissingFinder("sgl_kernel")
            sys.meta_path.append(finder)
    try:
        dec = importlib.import_module("sglang.test.kits.attention_unittest.runner_modes.cuda_graph_decode_runner")
        tv = importlib.import_module("sglang.test.kits.attention_unittest.runner_modes.speculative_target_verify_runner")
    finally:
        if finder is not None:
            sys.meta_path.remove(finder)
            for k in [k for k in sys.modules if k == "sgl_kernel" or k.startswith("sgl_kernel.")]:
                del sys.modules[k]
    return dec, tv


def merge_state_ref_path() -> Optional[str]:
    p = os.path.join(HERE, "_ref", "merge_state_ref.so")
    return p if os.path.exists(p) else None
