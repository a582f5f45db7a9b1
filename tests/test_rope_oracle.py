"""CPU: the rotary oracle (oracle/rope_port.py) against fixtures produced by the REAL reference classes."""
import pytest
import torch

from oracle import rope_port
from tests.util import load_golden, rope_golden_files

FILES = rope_golden_files()


def test_fixtures_present():
    assert len(FILES) >= 6


@pytest.mark.parametrize("path", FILES, ids=lambda p: p.split("/")[-1][:-3])
def test_native_restatement_bit_exact(path):
    g = load_golden(path)
    cache = g["cos_sin_cache"]
    for x, want in ((g["q"], g["out_q"]), (g["k"], g["out_k"])):
        got = rope_port.rope_native(g["positions"], x, cache, g["D"], g["rot"], g["neox"], g["mrope_section"],
                                    g["interleaved"])
        assert torch.equal(got.view(torch.int16), want.view(torch.int16))


@pytest.mark.parametrize("path", FILES, ids=lambda p: p.split("/")[-1][:-3])
def test_f32_restatement_within_one_rounding(path):
    """fp32 arithmetic differs from torch's element-dtype arithmetic only by intermediate roundings."""
    g = load_golden(path)
    got = rope_port.rope_f32(g["positions"], g["q"], g["cos_sin_cache"], g["D"], g["rot"], g["neox"],
                             g["mrope_section"], g["interleaved"])
    torch.testing.assert_close(got.float(), g["out_q"].float(), rtol=2e-2, atol=2e-2)


def test_cos_sin_cache_matches_reference_table():
    for path in FILES:
        g = load_golden(path)
        mine = rope_port.cos_sin_cache(g["rot"], g["max_pos"], g["base"])
        torch.testing.assert_close(mine, g["cos_sin_cache"], rtol=0, atol=2e-6)


def test_store_skips_reserved_slot():
    T, H, D = 5, 2, 16
    g = torch.Generator().manual_seed(3)
    cache = rope_port.cos_sin_cache(D, 32, 10000.0)
    q = torch.randn(T, H * D, generator=g).bfloat16()
    k = torch.randn(T, H * D, generator=g).bfloat16()
    v = torch.randn(T, H * D, generator=g).bfloat16()
    kc = torch.zeros(8, H, D).bfloat16()
    vc = torch.zeros(8, H, D).bfloat16()
    loc = torch.tensor([3, 0, 5, 1, 7])
    pos = torch.arange(T)
    _, k2, kc2, vc2 = rope_port.rope_store(pos, q, k, v, kc, vc, loc, cache, D, D, True, skip_index=0)
    assert torch.equal(kc2[3].reshape(-1), k2[0]) and torch.equal(vc2[7].reshape(-1), v[4])
    assert kc2[0].abs().sum() == 0
