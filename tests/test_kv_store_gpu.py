"""(a1) KV store parity: bit-exact against the oracle / reference pool (through the C ABI)."""
import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import ops
from tests.util import golden_files, load_golden

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run(T, width_k, width_v, dtype, idx_dtype, slots=4096, strided=False, skip=-1, seed=0):
    g = torch.Generator().manual_seed(seed)
    mk = lambda *s: (torch.randn(*s, generator=g) * 3).to(dtype)
    if strided:  # views into a fused qkv-like buffer (token stride > row width)
        buf = mk(T, width_k + width_v + 64)
        k, v = buf[:, 32:32 + width_k], buf[:, 32 + width_k:32 + width_k + width_v]
    else:
        k, v = mk(T, width_k), mk(T, width_v)
    kc, vc = mk(slots, width_k), mk(slots, width_v)
    loc = torch.randperm(slots - 1, generator=g)[:T] + 1
    if skip >= 0 and T > 2:
        loc[::3] = skip
    kc_ref, vc_ref = kc.clone(), vc.clone()
    oracle.store_kv(kc_ref, vc_ref, loc, k, v, skip_index=skip if skip >= 0 else None)
    kd, vd = kc.to(DEV), vc.to(DEV)
    kin, vin = k.to(DEV), v.to(DEV)
    if strided:
        bufd = buf.to(DEV)
        kin, vin = bufd[:, 32:32 + width_k], bufd[:, 32 + width_k:32 + width_k + width_v]
    ops.store_cache(kin, vin, kd, vd, loc.to(idx_dtype).to(DEV), reserved_skip_index=skip)
    torch.cuda.synchronize()
    assert torch.equal(kd.cpu().view(torch.int16 if dtype != torch.float32 else torch.int32),
                       kc_ref.view(torch.int16 if dtype != torch.float32 else torch.int32))
    assert torch.equal(vd.cpu().view(torch.int16 if dtype != torch.float32 else torch.int32),
                       vc_ref.view(torch.int16 if dtype != torch.float32 else torch.int32))


@pytest.mark.parametrize("T", [1, 2, 3, 63, 64, 65, 1023, 1024, 4097])
@pytest.mark.parametrize("width", [64, 96, 98, 100, 128, 256, 1024])
def test_store_bit_exact_sizes(T, width):
    # sizes follow test/registered/kernels/ops/kvcache/test_store_cache.py:29-46
    _run(T, width, width, torch.bfloat16, torch.int64, slots=8192)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("idx", [torch.int32, torch.int64])
def test_store_dtypes_and_index_types(dtype, idx):
    _run(257, 128, 128, dtype, idx)


def test_store_asymmetric_widths_and_strided_views():
    _run(300, 1024, 512, torch.bfloat16, torch.int64, strided=True)   # llama-8B row + narrower V
    _run(300, 256, 256, torch.bfloat16, torch.int64, strided=True)    # 70B TP8 row


def test_store_skips_reserved_slot_with_nan_rows():
    # test_store_cache.py:98-131: padded rows carry NaN garbage and target slot 0 -> must not be written
    T, W = 64, 256
    k = torch.full((T, W), float("nan"), dtype=torch.bfloat16, device=DEV)
    v = torch.full((T, W), float("nan"), dtype=torch.bfloat16, device=DEV)
    kc = torch.zeros((128, W), dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros((128, W), dtype=torch.bfloat16, device=DEV)
    loc = torch.zeros(T, dtype=torch.int64, device=DEV)
    loc[:8] = torch.arange(1, 9)
    k[:8], v[:8] = 1.0, 2.0
    ops.store_cache(k, v, kc, vc, loc, reserved_skip_index=0)
    torch.cuda.synchronize()
    assert torch.isfinite(kc).all() and torch.isfinite(vc).all()
    assert (kc[1:9] == 1).all() and (vc[1:9] == 2).all() and (kc[0] == 0).all() and (kc[9:] == 0).all()
    _run(100, 128, 128, torch.bfloat16, torch.int64, skip=0)


def test_store_out_of_range_is_counted_not_written():
    W = 64
    k = torch.ones((4, W), dtype=torch.bfloat16, device=DEV)
    kc = torch.zeros((16, W), dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros((16, W), dtype=torch.bfloat16, device=DEV)
    cnt = torch.zeros(1, dtype=torch.int32, device=DEV)
    loc = torch.tensor([1, 16, -3, 2], dtype=torch.int64, device=DEV)
    ops.store_cache(k, k, kc, vc, loc, reserved_skip_index=-1, oob_count=cnt)
    torch.cuda.synchronize()
    assert int(cnt) == 4  # two rows x (K and V)
    assert float(kc.sum()) == 2 * W


@pytest.mark.parametrize("path", golden_files(), ids=lambda p: p.split("/")[-1][:-3])
def test_store_matches_reference_pool(path):
    g = load_golden(path)
    Hkv, D = g["num_kv_heads"], g["head_dim"]
    kc, vc = g["k_pool_before"].to(DEV), g["v_pool_before"].to(DEV)
    ops.store_cache(g["k"].view(-1, Hkv, D).to(DEV), g["v"].view(-1, Hkv, D).to(DEV), kc, vc,
                    g["out_cache_loc"].to(DEV), reserved_skip_index=-1)
    loc = g["out_cache_loc"].long()
    assert torch.equal(kc.cpu()[loc], g["k_rows_after"]) and torch.equal(vc.cpu()[loc], g["v_rows_after"])
    mask = torch.ones(kc.shape[0], dtype=torch.bool)
    mask[loc] = False
    assert torch.equal(kc.cpu()[mask], g["k_pool_before"][mask])  # untouched slots stay untouched


def test_store_full_size_roundtrip():
    # cfg3 size: T = 16384 rows of 2048 B; scatter then gather must be the identity
    T, Hkv, D = 16384, 8, 128
    k = torch.randn((T, Hkv, D), device=DEV).to(torch.bfloat16)
    v = torch.randn((T, Hkv, D), device=DEV).to(torch.bfloat16)
    kc = torch.zeros((T + 4096, Hkv, D), dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros_like(kc)
    loc = (torch.randperm(T + 4095, device=DEV)[:T] + 1)
    ops.store_cache(k, v, kc, vc, loc, reserved_skip_index=0)
    assert torch.equal(kc[loc], k) and torch.equal(vc[loc], v)
    assert int((kc != 0).any(dim=(1, 2)).sum()) <= T
