"""(a6) merge_state parity vs the oracle (tolerance: fp32 arithmetic, output rounding of the dtype)."""
import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("dtype,tol", [(torch.bfloat16, 1e-2), (torch.float16, 2e-3), (torch.float32, 1e-5)])
@pytest.mark.parametrize("T,H,D", [(1, 1, 64), (77, 14, 64), (512, 32, 128), (33, 28, 128)])
def test_merge_matches_oracle(dtype, tol, T, H, D):
    g = torch.Generator().manual_seed(T * H)
    va = torch.randn((T, H, D), generator=g).to(dtype)
    vb = torch.randn((T, H, D), generator=g).to(dtype)
    sa = torch.randn((T, H), generator=g) * 4
    sb = torch.randn((T, H), generator=g) * 4
    sa[0, 0] = float("inf")        # +inf -> -inf rule
    if T > 2:
        sb[1, 0] = float("-inf")   # empty state b
        sa[2, 0] = sb[2, 0] = float("-inf")  # both empty
    v_ref, s_ref = oracle.merge_state(va, sa, vb, sb)
    v, s = ops.merge_state(va.to(DEV), sa.to(DEV), vb.to(DEV), sb.to(DEV))
    torch.testing.assert_close(v.cpu().float(), v_ref.float(), rtol=tol, atol=tol)
    torch.testing.assert_close(s.cpu(), s_ref, rtol=1e-5, atol=1e-5)


def test_merge_full_size_linearity():
    # cfg3 size; merging a state with itself is the identity and adds ln 2 to the LSE
    T, H, D = 16384, 32, 128
    v = torch.randn((T, H, D), device=DEV).to(torch.bfloat16)
    s = torch.randn((T, H), device=DEV)
    vm, sm = ops.merge_state(v, s, v, s)
    assert torch.equal(vm, v)
    torch.testing.assert_close(sm, s + 0.6931471805599453, rtol=1e-6, atol=1e-6)
