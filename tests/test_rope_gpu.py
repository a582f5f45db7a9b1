"""GPU: b200attn_rope_store through the C ABI against the oracle and the reference-generated fixtures.
Bit-exact in both arithmetic modes (integer/byte bar for the store, exact rounding emulation for the rotation)."""
import pytest
import torch

from oracle import rope_port
from tests.util import load_golden, rope_golden_files

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bits(t):
    return t.contiguous().view(torch.int16)


@pytest.mark.parametrize("path", rope_golden_files(), ids=lambda p: p.split("/")[-1][:-3])
@pytest.mark.parametrize("cache_f32", [True, False])
def test_golden_native_mode_bit_exact(path, cache_f32):
    from sglang_b200 import ops
    g = load_golden(path)
    q, k = g["q"].to(DEV), g["k"].to(DEV)
    cache = g["cos_sin_cache"].to(DEV)
    if not cache_f32:
        cache = cache.to(q.dtype)  # the reference's CPU-side cache dtype (base.py:107-109)
    ops.apply_rope_with_cos_sin_cache_inplace(
        g["positions"].to(DEV), q, k, cache, head_size=g["D"], is_neox=g["neox"], rope_dim=g["rot"],
        mrope_section=g["mrope_section"], mrope_interleaved=g["interleaved"], native_rounding=True)
    assert torch.equal(_bits(q.cpu()), _bits(g["out_q"]))
    assert torch.equal(_bits(k.cpu()), _bits(g["out_k"]))


@pytest.mark.parametrize("path", rope_golden_files(), ids=lambda p: p.split("/")[-1][:-3])
def test_golden_f32_mode_bit_exact_vs_oracle(path):
    from sglang_b200 import ops
    g = load_golden(path)
    q, k = g["q"].to(DEV), g["k"].to(DEV)
    pos32 = g["positions"].to(torch.int32).to(DEV)  # int32 positions accepted too
    ops.apply_rope_with_cos_sin_cache_inplace(
        pos32, q, k, g["cos_sin_cache"].to(DEV), head_size=g["D"], is_neox=g["neox"], rope_dim=g["rot"],
        mrope_section=g["mrope_section"], mrope_interleaved=g["interleaved"])
    for got, x in ((q, g["q"]), (k, g["k"])):
        want = rope_port.rope_f32(g["positions"], x, g["cos_sin_cache"], g["D"], g["rot"], g["neox"],
                                  g["mrope_section"], g["interleaved"])
        assert torch.equal(_bits(got.cpu()), _bits(want))
    torch.testing.assert_close(q.cpu().float(), g["out_q"].float(), rtol=2e-2, atol=2e-2)


# D, rot, neox, Hq, Hkv, Dv, T, sec, interleaved, dtype
FUSED = [
    (128, 128, True, 32, 8, 128, 64, None, False, torch.bfloat16),       # llama-3.1-8b decode step
    (128, 128, True, 28, 4, 128, 333, [16, 24, 24], False, torch.bfloat16),  # qwen2.5-vl
    (128, 128, True, 4, 2, 128, 50, [24, 20, 20], True, torch.bfloat16),
    (64, 64, False, 14, 2, 64, 17, None, False, torch.float16),
    (128, 64, True, 4, 4, 128, 21, None, False, torch.bfloat16),         # partial rotary: pass-through stored
    (96, 32, False, 6, 2, 80, 9, None, False, torch.bfloat16),           # odd widths, Dv != D
    (100, 20, True, 3, 1, 100, 7, None, False, torch.bfloat16),          # rot/2 = 10 -> 4-byte vectors
    (66, 6, False, 3, 1, 66, 5, None, False, torch.float16),             # rot/2 = 3 -> scalar path
]


@pytest.mark.parametrize("case", FUSED, ids=lambda c: f"D{c[0]}r{c[1]}{'n' if c[2] else 'g'}H{c[3]}T{c[6]}")
@pytest.mark.parametrize("native", [True, False])
@pytest.mark.parametrize("loc_dtype", [torch.int64, torch.int32])
def test_fused_rope_store_bit_exact(case, native, loc_dtype):
    from sglang_b200 import ops
    D, rot, neox, Hq, Hkv, Dv, T, sec, inter, dt = case
    g = torch.Generator().manual_seed(T * 131 + D)
    max_pos, slots = 300, 4 * T + 16
    cache = rope_port.cos_sin_cache(rot, max_pos, 10000.0)
    # q/k/v are views of one fused qkv row, as the model hands them over (token stride > row)
    qkv = torch.randn((T, (Hq + Hkv) * D + Hkv * Dv), generator=g).to(dt)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * Dv], dim=-1)
    pos = torch.randint(0, max_pos, (3, T) if sec else (T,), generator=g)
    loc = torch.randperm(slots - 1, generator=g)[:T] + 1
    if T > 4:
        loc[T // 2] = 0  # reserved slot: skipped
    kc = torch.randn((slots, Hkv, D), generator=g).to(dt)
    vc = torch.randn((slots, Hkv, Dv), generator=g).to(dt)
    wq, wk, wkc, wvc = rope_port.rope_store(pos, q.contiguous(), k.contiguous(), v.contiguous(), kc, vc, loc, cache,
                                            D, rot, neox, sec, inter, native=native, skip_index=0)
    dqkv = qkv.to(DEV)
    dq, dk, dv = dqkv.split([Hq * D, Hkv * D, Hkv * Dv], dim=-1)
    dkc, dvc = kc.to(DEV), vc.to(DEV)
    ops.apply_rope_with_cos_sin_cache_inplace(
        pos.to(DEV), dq, dk, cache.to(DEV), head_size=D, is_neox=neox, rope_dim=rot, mrope_section=sec,
        mrope_interleaved=inter, v=dv, k_buffer=dkc, v_buffer=dvc, cache_loc=loc.to(loc_dtype).to(DEV),
        native_rounding=native)
    assert torch.equal(_bits(dq.cpu()), _bits(wq))
    assert torch.equal(_bits(dk.cpu()), _bits(wk))
    assert torch.equal(_bits(dv.cpu()), _bits(v))  # v untouched
    assert torch.equal(_bits(dkc.cpu()), _bits(wkc))
    assert torch.equal(_bits(dvc.cpu()), _bits(wvc))


def test_rope_then_store_equals_fused_and_matches_layer_classes():
    """RotaryEmbedding mirror + FusedSetKVBufferArg == separate rope then store_cache."""
    from sglang_b200 import ops
    from sglang_b200.rotary import FusedSetKVBufferArg, MRotaryEmbedding, RotaryEmbedding
    T, Hq, Hkv, D = 97, 8, 2, 128
    g = torch.Generator(device=DEV).manual_seed(5)
    q = torch.randn((T, Hq * D), device=DEV, generator=g).bfloat16()
    k = torch.randn((T, Hkv * D), device=DEV, generator=g).bfloat16()
    v = torch.randn((T, Hkv * D), device=DEV, generator=g).bfloat16()
    loc = (torch.randperm(400, device=DEV)[:T] + 16)
    pos = torch.randint(0, 1000, (T,), device=DEV)
    rope = RotaryEmbedding(D, D, 1024, 500000.0, True, torch.bfloat16)
    kc1, vc1 = torch.zeros((512, Hkv, D), device=DEV).bfloat16(), torch.zeros((512, Hkv, D), device=DEV).bfloat16()
    kc2, vc2 = kc1.clone(), vc1.clone()
    q1, k1 = q.clone(), k.clone()
    rope(pos, q1, k1)
    ops.store_cache(k1.view(T, Hkv, D), v.view(T, Hkv, D), kc1, vc1, loc)
    q2, k2 = q.clone(), k.clone()
    rope(pos, q2, k2, fused_set_kv_buffer_arg=FusedSetKVBufferArg(v, kc2, vc2, loc))
    assert torch.equal(q1, q2) and torch.equal(k1, k2) and torch.equal(kc1, kc2) and torch.equal(vc1, vc2)
    m = MRotaryEmbedding(D, D, 1024, 1e6, True, torch.bfloat16, mrope_section=[16, 24, 24])
    q3, q4 = q.clone(), q.clone()
    k3, k4 = k.clone(), k.clone()
    m(pos.unsqueeze(0).repeat(3, 1), q3, k3)   # t == h == w  ==> identical to 1-D rope with the same base
    m(pos, q4, k4)
    assert torch.equal(q3, q4) and torch.equal(k3, k4)
    # text-only batches hand the three axes as ONE expanded row (stride 0 between axes): no out-of-bounds reads,
    # same result as the materialised [3, T] positions
    q5, k5 = q.clone(), k.clone()
    expanded = pos.unsqueeze(0).expand(3, -1)
    assert expanded.stride(0) == 0
    m(expanded, q5, k5)
    assert torch.equal(q3, q5) and torch.equal(k3, k5)
    with pytest.raises(AssertionError):
        rope(pos.unsqueeze(0).repeat(3, 1), q3, k3)


def test_errors_and_empty():
    from sglang_b200 import ops
    from sglang_b200._lib import B200AttnError
    cache = rope_port.cos_sin_cache(64, 16, 1e4).to(DEV)
    q = torch.zeros((0, 128), device=DEV).bfloat16()
    ops.apply_rope_with_cos_sin_cache_inplace(torch.zeros(0, dtype=torch.int64, device=DEV), q, q.clone(), cache,
                                              head_size=64)
    q = torch.zeros((4, 128), device=DEV).bfloat16()
    pos = torch.zeros(4, dtype=torch.int64, device=DEV)
    with pytest.raises(RuntimeError):
        ops.apply_rope_with_cos_sin_cache_inplace(pos.cpu(), q, q.clone(), cache, head_size=64)
    with pytest.raises(B200AttnError):
        ops.apply_rope_with_cos_sin_cache_inplace(pos.unsqueeze(0).repeat(3, 1), q, q.clone(), cache, head_size=64,
                                                  mrope_section=[8, 8, 8])
    with pytest.raises(TypeError):
        ops.apply_rope_with_cos_sin_cache_inplace(pos, q, q.clone(), cache.double(), head_size=64)


def test_full_size_cfg3_properties():
    """BASELINE configs[2] token count (16,384 tokens, Llama-3.1-8B heads) through the bulk-copy pipeline:
    (1) sampled rows bit-exact against the oracle, (2) rotation preserves every pair's norm up to the output rounding,
    (3) position 0 is the identity, (4) the pool holds exactly the rotated k and the untouched v."""
    from sglang_b200 import ops
    T, Hq, Hkv, D = 16384, 32, 8, 128
    g = torch.Generator(device=DEV).manual_seed(11)
    qkv = torch.randn((T, (Hq + 2 * Hkv) * D), device=DEV, generator=g).bfloat16()
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    q0, k0, v0 = q.clone(), k.clone(), v.clone()
    pos = torch.randint(0, 8192, (T,), device=DEV, generator=g)
    pos[::97] = 0
    cache = rope_port.cos_sin_cache(D, 8192, 500000.0).to(DEV)
    slots = 2 * T + 16
    kc = torch.zeros((slots, Hkv, D), device=DEV).bfloat16()
    vc = torch.zeros_like(kc)
    loc = torch.randperm(slots - 16, device=DEV, generator=g)[:T] + 16
    ops.apply_rope_with_cos_sin_cache_inplace(pos, q, k, cache, head_size=D, v=v, k_buffer=kc, v_buffer=vc, cache_loc=loc)
    rows = torch.arange(0, T, 331, device=DEV)
    want_q = rope_port.rope_f32(pos[rows].cpu(), q0[rows].cpu().contiguous(), cache.cpu(), D, D, True)
    want_k = rope_port.rope_f32(pos[rows].cpu(), k0[rows].cpu().contiguous(), cache.cpu(), D, D, True)
    assert torch.equal(_bits(q[rows].cpu()), _bits(want_q)) and torch.equal(_bits(k[rows].cpu()), _bits(want_k))
    a, b = q0.view(T, Hq, D).float(), q.view(T, Hq, D).float()
    n0 = a[..., :64] ** 2 + a[..., 64:] ** 2
    n1 = b[..., :64] ** 2 + b[..., 64:] ** 2
    assert float(((n1 - n0).abs() / (n0 + 1e-3)).max()) < 3e-2            # bf16 output rounding of both components
    ident = pos == 0
    assert torch.equal(q[ident], q0[ident]) and torch.equal(k[ident], k0[ident])
    assert torch.equal(kc[loc].view(T, -1), k) and torch.equal(vc[loc].view(T, -1), v0) and torch.equal(v, v0)
