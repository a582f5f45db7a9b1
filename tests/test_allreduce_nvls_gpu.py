"""b200attn_allreduce_nvls through NvlsAllReduce on 2 GPUs (skipped on a one-GPU box): tools/allreduce_nvls_check.py
under torchrun compares it with NCCL - bit-equal on integer-valued inputs in bf16 / fp16 / fp32, within rounding on
random ones, 300 back-to-back calls, the payload() in-place form and CUDA-graph replay - and exits non-zero on a
mismatch.  The argument checks of the C entry point need no GPU pair."""
import os
import subprocess
import sys

import pytest
import torch

from sglang_b200 import _lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_argument_checks():
    lib = _lib.load()
    assert lib.b200attn_allreduce_flag_bytes() == 16384 and lib.b200attn_allreduce_small_bytes() == 2 * 128 * 32768
    assert lib.b200attn_allreduce_nvls(None, None, None, None, None, None, None, 0, 2, 0, 0, None) == 0  # empty
    assert lib.b200attn_allreduce_nvls(None, None, None, None, None, None, None, 3, 2, 64, 0, None) == 1
    assert b"rank" in lib.b200attn_last_error()
    assert lib.b200attn_allreduce_nvls(None, None, None, None, None, None, None, 0, 2, 24, 0, None) == 1


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_nvls_allreduce_matches_nccl_on_two_gpus():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tools", "allreduce_nvls_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and ("RESULT PASS" in r.stdout or "unavailable" in r.stdout), r.stdout[-3000:] + r.stderr[-3000:]
