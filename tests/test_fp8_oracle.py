"""(f4) FP8 KV oracle pinned to torch's own conversion (CPU, no GPU): every bf16 bit pattern."""
import numpy as np
import torch

from oracle import fp8_port


def _all_bf16():
    return torch.arange(65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)


def test_encoder_matches_torch_on_every_bf16_value():
    x = _all_bf16()
    want = x.to(fp8_port.FP8).view(torch.uint8).numpy()
    got = fp8_port.encode_e4m3_bits(x.float().numpy())
    nan = torch.isnan(x).numpy()
    assert np.array_equal(got[~nan], want[~nan])
    assert np.all((got[nan] & 0x7F) == 0x7F) and np.all((want[nan] & 0x7F) == 0x7F)


def test_quantize_is_div_in_input_dtype_then_cast():
    x = _all_bf16()
    x = x[torch.isfinite(x)]
    for scale in (0.5, 0.37, 1.7, 3.0):
        want = fp8_port.quantize(x, scale).numpy()
        q = (x.float() / scale).to(torch.bfloat16).float().numpy()     # quotient rounded to bf16 first
        assert np.array_equal(fp8_port.encode_e4m3_bits(q), want), scale


def test_no_saturation_above_464():
    x = torch.tensor([448.0, 464.0, 466.0, 480.0, 1e4, -466.0], dtype=torch.bfloat16)
    assert fp8_port.quantize(x, None).tolist() == [0x7E, 0x7E, 0x7F, 0x7F, 0x7F, 0xFF]


def test_store_and_dequant_round_trip():
    g = torch.Generator().manual_seed(0)
    k = torch.randn((5, 2, 16), generator=g).to(torch.bfloat16)
    v = torch.randn((5, 2, 16), generator=g).to(torch.bfloat16)
    kc = torch.zeros((8, 2, 16), dtype=torch.uint8)
    vc = torch.zeros_like(kc)
    loc = torch.tensor([3, 0, 7, 1, 5])
    fp8_port.store_kv_fp8(kc, vc, loc, k, v, 0.5, None, skip_index=0)
    assert kc[0].eq(0).all() and not kc[3].eq(0).all()
    back = fp8_port.dequantize(kc, 0.5)[3]
    assert torch.allclose(back, k[0].float(), rtol=0.07, atol=2e-3)   # e4m3: 3 mantissa bits
