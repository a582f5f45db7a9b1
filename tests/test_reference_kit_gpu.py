"""The reference's own dense-attention test kit against the registered "b200" backend (tools/run_reference_kit.py).

Runs in a subprocess: the kit installs process-wide overrides (server args, parallel state) and B200AttnBackend
must be imported AFTER the reference so that it subclasses the real AttentionBackend ABC.  Skipped cleanly when
neither /root/reference nor the staged oracle/_ref/python tree is present."""
import os
import subprocess
import sys

import pytest

from oracle import ref_env

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_reference_dense_kit_passes_on_registered_backend():
    if ref_env.reference_root() is None:
        pytest.skip("reference tree not staged (oracle/make_ref.sh)")
    out = os.path.join(ROOT, "gpurun_out", "reference_kit.txt")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "run_reference_kit.py"), "--out", out],
                       capture_output=True, text=True, timeout=1500)
    tail = (r.stdout + r.stderr)[-6000:]
    assert r.returncode == 0, tail
    assert " 0 failed" in r.stdout, tail
