"""Achieved HBM GB/s of the fused RoPE + KV-store kernel vs (separate rope, then store).  GPU box only.
Algorithmic bytes per token: read (Hq+2Hkv)*D*2, write in place (Hq+Hkv)*D*2, write pool 2*Hkv*D*2 (+ pos/loc 16 B)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import ops  # noqa: E402
from sglang_b200.rotary import compute_cos_sin_cache  # noqa: E402


def timeit(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n


def main():
    dev = "cuda"
    res = []
    for name, T, Hq, Hkv, D, sec in [("llama8b_decode_T64", 64, 32, 8, 128, None), ("llama8b_extend_T16384", 16384, 32, 8, 128, None),
                                     ("qwen25vl_mrope_T20496", 20496, 28, 4, 128, [16, 24, 24]),
                                     ("llama8b_extend_T131072", 131072, 32, 8, 128, None)]:
        qkv = torch.randn((T, (Hq + 2 * Hkv) * D), device=dev).bfloat16()
        q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        slots = T + 64
        kc = torch.zeros((slots, Hkv, D), device=dev).bfloat16()
        vc = torch.zeros_like(kc)
        loc = torch.randperm(slots - 16, device=dev)[:T] + 16
        pos = torch.randint(0, 8192, (3, T) if sec else (T,), device=dev)
        cache = compute_cos_sin_cache(D, 8192, 5e5).to(dev)
        fused = lambda: ops.apply_rope_with_cos_sin_cache_inplace(pos, q, k, cache, head_size=D, mrope_section=sec, v=v,
                                                                  k_buffer=kc, v_buffer=vc, cache_loc=loc)

        def separate():
            ops.apply_rope_with_cos_sin_cache_inplace(pos, q, k, cache, head_size=D, mrope_section=sec)
            ops.store_cache(k.view(T, Hkv, D), v.view(T, Hkv, D), kc, vc, loc)

        rope_only = lambda: ops.apply_rope_with_cos_sin_cache_inplace(pos, q, k, cache, head_size=D, mrope_section=sec)
        store_only = lambda: ops.store_cache(k.view(T, Hkv, D), v.view(T, Hkv, D), kc, vc, loc)
        inplace = lambda: qkv.mul_(1.0)
        tr, tso, ti = timeit(rope_only), timeit(store_only), timeit(inplace)
        print(json.dumps(dict(case=name, rope_only_us=tr * 1e3, rope_only_gbs=T * 2 * (Hq + Hkv) * D * 2 / tr / 1e6,
                              store_only_us=tso * 1e3, store_only_gbs=T * 4 * Hkv * D * 2 / tso / 1e6,
                              torch_inplace_mul_qkv_us=ti * 1e3, torch_inplace_gbs=2 * qkv.numel() * 2 / ti / 1e6)), flush=True)
        by = T * ((Hq + 2 * Hkv) * D * 2 + (Hq + Hkv) * D * 2 + 2 * Hkv * D * 2 + 16)
        tf, ts = timeit(fused), timeit(separate)
        res.append(dict(case=name, bytes=by, fused_us=tf * 1e3, fused_gbs=by / tf / 1e6, separate_us=ts * 1e3))
        print(json.dumps(res[-1]), flush=True)


if __name__ == "__main__":
    main()
