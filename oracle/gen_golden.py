"""Generate tests/golden/*.pt by running the REAL reference on CPU (this container only).

Imports SGLang from /root/reference/python (read-only) with two stub modules
(orjson, IPython) and drives its own unit-test kit
(python/sglang/test/kits/attention_unittest/attention_methods/dense_attention.py):
MockModelRunner -> real ReqToTokenPool + MHATokenToKVPool, real
TorchNativeAttnBackend, real RadixAttention, real ForwardBatch.  For each case
it records the backend-boundary tensors (q, k, v, pool before/after, metadata)
and the reference's attention output, so the fixtures pin

  * the KV store   (pool_after  == reference set_kv_buffer result, bit-exact),
  * the bookkeeping (req_to_token / out_cache_loc / positions as the kit built them),
  * decode / extend / mixed attention outputs of TorchNativeAttnBackend.

Run:  python oracle/gen_golden.py        (needs /root/reference; NOT run on the GPU box)
"""
from __future__ import annotations

import json
import os
import sys
import types
import unittest

import torch

REF = "/root/reference/python"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _install_stubs() -> None:
    sys.path.insert(0, REF)
    o = types.ModuleType("orjson")
    o.dumps = lambda x, *a, **k: json.dumps(x).encode()
    o.loads = json.loads
    o.OPT_NON_STR_KEYS = o.OPT_SERIALIZE_NUMPY = 0
    o.JSONDecodeError = ValueError
    sys.modules["orjson"] = o
    ip = types.ModuleType("IPython")
    ip.__path__ = []
    sys.modules["IPython"] = ip
    ipd = types.ModuleType("IPython.display")
    ipd.HTML = object
    ipd.display = lambda *a, **k: None
    sys.modules["IPython.display"] = ipd


def main() -> None:
    _install_stubs()
    from sglang.srt.layers.radix_attention import AttentionType
    from sglang.srt.model_executor.forward_batch_info import ForwardMode
    from sglang.srt.model_executor.forward_context import ForwardContext, forward_context
    from sglang.test.kits.attention_unittest.attention_methods import dense_attention as da

    tc = unittest.TestCase()
    E, D_, M = ForwardMode.EXTEND, ForwardMode.DECODE, ForwardMode.MIXED
    # (name, mode, Hq, Hkv, head_dim, page, prefix_lens, extend_lens, layout, encoder_only)
    cases = [
        ("decode_p1_qwen05", D_, 14, 2, 64, 1, (17, 40, 5, 63), (), "shuffled_pages", False),
        ("decode_p16_llama", D_, 8, 2, 128, 16, (15, 16, 47, 90), (), "shuffled_pages", False),
        ("decode_p16_interleaved", D_, 8, 1, 128, 16, (31, 64, 1), (), "interleaved_pages", False),
        ("decode_p32_mha", D_, 4, 4, 64, 32, (33, 95), (), "contiguous", False),
        ("extend_noprefix_p16", E, 8, 2, 128, 16, (0, 0, 0), (40, 16, 7), "shuffled_pages", False),
        ("extend_prefix_p16", E, 8, 2, 128, 16, (32, 17, 48, 0), (24, 30, 1, 50), "shuffled_pages", False),
        ("extend_prefix_p1", E, 14, 2, 64, 1, (9, 30), (20, 11), "shuffled_pages", False),
        ("extend_nonmono_p16", E, 4, 1, 128, 16, (16, 5), (33, 40), "non_monotonic_extend", False),
        ("extend_bidir_p16", E, 8, 2, 128, 16, (16, 0, 35), (30, 44, 9), "shuffled_pages", True),
        ("mixed_p16", M, 8, 2, 128, 16, (0, 20, 57, 33), (36, 25, 1, 1), "shuffled_pages", False),
        ("extend_gqa7_p16", E, 28, 4, 128, 16, (0, 12), (45, 20), "shuffled_pages", False),
    ]
    os.makedirs(OUT, exist_ok=True)
    index = []
    for (name, mode, hq, hkv, hd, page, prefix, ext, layout, enc_only) in cases:
        case = da.DenseAttentionCase(name=name, backend="torch_native", forward_mode=mode, num_heads=hq,
                                     num_kv_heads=hkv, page_size=page, prefix_lens=prefix, extend_lens=ext)
        ctx = 128 if page > 1 else 96
        fx = da.build_dense_attention_fixture(tc, case, head_dim=hd, hidden_size=hq * hd // 2,
                                              max_context_len=ctx, dtype=torch.bfloat16, device="cpu",
                                              loc_layout=layout)
        layer = fx.actual_module.attn
        if enc_only:
            layer.attn_type = AttentionType.ENCODER_ONLY
        fb = fx.forward_batch
        pool = fx.runner.token_to_kv_pool
        with torch.no_grad():
            q, k, v = fx.actual_module.project_qkv(fx.input_hidden)
        k_before = pool.get_key_buffer(layer.layer_id).clone()
        v_before = pool.get_value_buffer(layer.layer_id).clone()
        with torch.no_grad(), forward_context(ForwardContext(attn_backend=fx.backend)):
            fx.backend.init_forward_metadata(fb)
            o = layer(q, k, v, fb)
        k_after = pool.get_key_buffer(layer.layer_id)
        v_after = pool.get_value_buffer(layer.layer_id)
        loc = fb.out_cache_loc.long()
        # the store touches only out_cache_loc rows
        mask = torch.ones(k_after.shape[0], dtype=torch.bool)
        mask[loc] = False
        assert torch.equal(k_after[mask], k_before[mask]) and torch.equal(v_after[mask], v_before[mask])
        rec = dict(
            name=name, mode=mode.name, num_heads=hq, num_kv_heads=hkv, head_dim=hd, page_size=page,
            scaling=float(layer.scaling), causal=not enc_only, prefix_lens=list(prefix),
            extend_lens=list(case.input_lens), seq_lens=list(case.seq_lens), layout=layout,
            q=q.contiguous(), k=k.contiguous(), v=v.contiguous(),
            req_to_token=fx.runner.req_to_token_pool.req_to_token[: case.batch_size + 1].clone(),
            req_pool_indices=fb.req_pool_indices.clone(), out_cache_loc=fb.out_cache_loc.clone(),
            positions=fb.positions.clone(), k_pool_before=k_before, v_pool_before=v_before,
            k_rows_after=k_after[loc].clone(), v_rows_after=v_after[loc].clone(),
            o_ref=o.contiguous(), torch_version=torch.__version__,
        )
        path = os.path.join(OUT, f"{name}.pt")
        torch.save(rec, path)
        index.append(dict(name=name, file=f"{name}.pt", bytes=os.path.getsize(path)))
        print(f"{name}: o {tuple(o.shape)} {os.path.getsize(path) / 1024:.0f} KiB")
    with open(os.path.join(OUT, "INDEX.json"), "w") as f:
        json.dump(dict(generator="oracle/gen_golden.py", reference="/root/reference/python "
                       "(TorchNativeAttnBackend via the dense attention unit-test kit)",
                       torch=torch.__version__, cases=index), f, indent=1)


if __name__ == "__main__":
    main()
