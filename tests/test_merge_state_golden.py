"""merge_state against outputs of the REFERENCE's own kernels, generated on a B200 by oracle/gen_golden_merge.py:
merge_state_v2 (python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu:31-206, compiled from where it lies)
and merge_state_triton (python/sglang/kernels/ops/attention/merge_state.py:8-96).

Rows where both input states are empty (-inf, -inf) are NaN in the reference (-inf - -inf); this build defines them as
zeros with LSE -inf (what the N-way decode merge needs) — those rows are checked for that, all others against the
reference: the oracle within fp32 rounding, the CUDA kernel BIT-EXACT (same operations in the same order)."""
import glob
import os

import pytest
import torch

from oracle import torch_native_port as oracle
from tests.util import GOLDEN_DIR, load_golden

FILES = sorted(glob.glob(os.path.join(GOLDEN_DIR, "merge_state_*.pt")))
IDS = [os.path.basename(p)[:-3] for p in FILES]


def _rows(g):
    a, b = g["s_a"].clone(), g["s_b"].clone()
    a[torch.isinf(a)] = float("-inf")
    b[torch.isinf(b)] = float("-inf")
    both_empty = torch.isneginf(a) & torch.isneginf(b)
    return both_empty


def test_merge_fixtures_present():
    assert len(FILES) >= 4, "generate them on the GPU box: python oracle/gen_golden_merge.py gpurun_out/golden"


@pytest.mark.parametrize("path", FILES, ids=IDS)
def test_oracle_matches_reference_merge_kernels(path):
    g = load_golden(path)
    empty = _rows(g)
    assert torch.isnan(g["v_cuda"][empty].float()).all() or not empty.any()   # documents the reference's NaN rows
    v, s = oracle.merge_state(g["v_a"], g["s_a"], g["v_b"], g["s_b"])
    tol = dict(rtol=1e-5, atol=1e-6) if g["v_a"].dtype == torch.float32 else dict(rtol=8e-3, atol=1e-3)
    torch.testing.assert_close(v[~empty].float(), g["v_cuda"][~empty].float(), **tol)
    torch.testing.assert_close(s[~empty], g["s_cuda"][~empty], rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(v[~empty].float(), g["v_triton"][~empty].float(), **tol)
    assert not v[empty].any() and torch.isneginf(s[empty]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=IDS)
def test_kernel_matches_reference_merge_kernels_bit_exact(path):
    from sglang_b200 import ops

    g = load_golden(path)
    empty = _rows(g)
    v, s = ops.merge_state(g["v_a"].cuda(), g["s_a"].cuda(), g["v_b"].cuda(), g["s_b"].cuda())
    v, s = v.cpu(), s.cpu()
    assert torch.equal(v[~empty], g["v_cuda"][~empty]), "merged values differ from merge_state_v2"
    assert torch.equal(s[~empty], g["s_cuda"][~empty]), "merged LSE differs from merge_state_v2"
    assert not v[empty].any() and torch.isneginf(s[empty]).all()


@pytest.mark.gpu
def test_kernel_matches_compiled_reference_live():
    """When oracle/_ref/merge_state_ref.so travelled to the box: a large random case straight against it."""
    from oracle import ref_env
    from sglang_b200 import ops

    so = ref_env.merge_state_ref_path()
    if so is None:
        pytest.skip("oracle/_ref/merge_state_ref.so not built")
    torch.ops.load_library(so)
    g = torch.Generator(device="cuda").manual_seed(9)
    for dt in (torch.bfloat16, torch.float16, torch.float32):
        va = torch.randn((4096, 32, 128), device="cuda", generator=g).to(dt)
        vb = torch.randn((4096, 32, 128), device="cuda", generator=g).to(dt)
        sa = torch.randn((4096, 32), device="cuda", generator=g) * 5
        sb = torch.randn((4096, 32), device="cuda", generator=g) * 5
        vo, so_ = torch.empty_like(va), torch.empty_like(sa)
        torch.ops.b200ref.merge_state_v2(va, sa, vb, sb, vo, so_)
        v, s = ops.merge_state(va, sa, vb, sb)
        assert torch.equal(v, vo) and torch.equal(s, so_), f"{dt}: not bit-identical to merge_state_v2"
