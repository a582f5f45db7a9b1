#!/usr/bin/env python
"""Run the REFERENCE's own dense-attention unit-test kit against the registered "b200" backend (GPU box).

The reference looks a backend up by name in ATTENTION_BACKENDS and drives it through its real objects
(python/sglang/test/kits/attention_unittest/attention_methods/dense_attention.py:307-449,1003,1269-1293):
MockModelRunner(ModelRunner) -> real ReqToTokenPool + MHATokenToKVPool, real RadixAttention + ForwardBatch +
ForwardContext, q/k/v from real nn.Linear projections, expected output from the kit's independent fp32 reference,
`assert_close(atol = rtol = 3e-2)` in fp16.  This script registers sglang_b200.backend.register("b200") — the
entry point INTEGRATION.md describes — and runs

  * make_dense_cases("b200") x {contiguous, shuffled_pages, interleaved_pages, non_monotonic_extend}
    at head_dim 128, 64, 256 and the kit's own default 16;
  * the kit's sliding-window cases (dense_attention.py:212-260) — plain pool, layer.sliding_window_size;
  * the CUDA-graph decode runner mode (runner_modes/cuda_graph_decode_runner.py: capture-bucket 4, padded replay,
    metadata invariants) and the EAGLE TARGET_VERIFY mode (speculative_target_verify_runner.py, chain + tree masks).

The reference tree comes from oracle/_ref/python (staged by oracle/make_ref.sh; /root/reference in the build
container).  Test infrastructure: nothing here is on the product path.

  python tools/run_reference_kit.py [--out profiles/r02_reference_kit.txt]
"""
from __future__ import annotations

import argparse
import os
import sys
import traceback
import unittest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch

    from oracle import ref_env

    root = ref_env.install()
    from sglang.srt.layers.attention.attention_registry import ATTENTION_BACKENDS
    from sglang.srt.layers.attention.base_attn_backend import AttentionBackend
    from sglang.srt.model_executor.forward_batch_info import ForwardMode
    from sglang.test.kits.attention_unittest.attention_methods import dense_attention as da

    dec, tv = ref_env.import_runner_modes()
    import sglang_b200.backend as bk  # imported AFTER the reference: B200AttnBackend subclasses the real ABC

    bk.register("b200")
    assert "b200" in ATTENTION_BACKENDS and issubclass(bk.B200AttnBackend, AttentionBackend)
    # the kit wants every SWA-capable backend classified by its decode window rule (dense_attention.py:53-68):
    # ours is the extend kernel's (window + 1 keys), like torch_native / fa3 / flashinfer
    da._SWA_DECODE_EXTEND_WINDOW = frozenset(da._SWA_DECODE_EXTEND_WINDOW | {"b200"})

    tc = unittest.TestCase()
    results = []
    notes = []
    # The kit fills the prefix K/V through the REFERENCE pool's set_kv_buffer, which on CUDA JIT-compiles the
    # reference's store_cache kernel through tvm-ffi (memory_pool.py:141-192).  If that toolchain does not work in
    # this image the pool's own naive branch (k_cache[indices] = k) is selected instead; either way it is the
    # reference's code that writes the prefix, and the backend under test writes the new tokens itself.
    try:
        from sglang.srt.mem_cache import memory_pool as mp

        probe = mp.MHATokenToKVPool(size=64, page_size=16, dtype=torch.float16, head_num=2, head_dim=64, layer_num=1,
                                    device="cuda", enable_memory_saver=False, enable_alt_stream=False)
        from types import SimpleNamespace as NS

        probe.set_kv_buffer(NS(layer_id=0), torch.tensor([17, 18], device="cuda"),
                            torch.ones((2, 2, 64), dtype=torch.float16, device="cuda"),
                            torch.ones((2, 2, 64), dtype=torch.float16, device="cuda"))
        torch.cuda.synchronize()
        assert float(probe.get_key_buffer(0)[17].sum()) == 128.0
        notes.append("reference pool store: JIT store_cache kernel (tvm-ffi) works in this image")
    except Exception as e:  # noqa: BLE001
        mp.can_use_store_cache = lambda *a, **k: False
        notes.append(f"reference pool store: JIT store_cache unavailable here ({type(e).__name__}: {str(e)[:120]}); "
                     "the pool's naive branch k_cache[indices] = k is used for the kit's prefix fill")

    def run(label, fn):
        try:
            fn()
            results.append((label, "PASS", ""))
        except unittest.SkipTest as e:
            results.append((label, "SKIP", str(e)))
        except Exception as e:  # noqa: BLE001
            results.append((label, "FAIL", f"{type(e).__name__}: {str(e)[:300]}"))
            traceback.print_exc()

    layouts = ("contiguous", "shuffled_pages", "interleaved_pages", "non_monotonic_extend")
    for hd in (128, 64, 16, 256):  # 256 = Gemma-style heads; 16 = the kit's own DEFAULT_HEAD_DIM (generic-head-dim path of the backend)
        kw = dict(head_dim=hd, hidden_size=64, device="cuda", dtype=torch.float16)
        for case in da.make_dense_cases("b200"):
            for layout in layouts:
                if layout == "non_monotonic_extend" and case.forward_mode.is_decode():
                    continue  # decode has no extend tokens to scatter (same skip as the reference's tests)
                if layout == "non_monotonic_extend" and case.page_size > 1:
                    pass  # token-granular indices: no page contiguity assumed
                run(f"dense/{case.name}/d{hd}/{layout}",
                    lambda c=case, l=layout: da.run_dense_attention_case(tc, c, loc_layout=l, **kw))
        for case in da.make_swa_no_prefix_input_config_cases("b200") + da.make_swa_prefix_input_config_cases("b200"):
            run(f"swa/{case.name}/d{hd}", lambda c=case: da.run_dense_attention_case(tc, c, **kw))
        swa_dec = da.DenseAttentionCase(name="swa_decode_window_edges", backend="b200", forward_mode=ForwardMode.DECODE,
                                        num_heads=4, num_kv_heads=4, page_size=16, prefix_lens=(2, 3, 4, 5, 40),
                                        sliding_window_size=4)
        run(f"swa/{swa_dec.name}/d{hd}", lambda c=swa_dec: da.run_dense_attention_case(tc, c, **kw))
        # runner modes
        for name, heads, kvh, pre in (("runner_cuda_graph_decode_padded", 4, 4, (4, 7)),
                                      ("runner_cuda_graph_decode_gqa_full_bucket", 8, 2, (14, 15, 16, 33))):
            case = da.DenseAttentionCase(name=name, backend="b200", forward_mode=ForwardMode.DECODE, num_heads=heads,
                                         num_kv_heads=kvh, page_size=16, prefix_lens=pre)
            run(f"cuda_graph_decode/{name}/d{hd}", lambda c=case: dec.run_dense_cuda_graph_decode_case(tc, c, **kw))
        if hd not in (64, 128):
            continue  # tree masks need the tcgen05 extend kernel (64 / 128): the generic-head-dim path raises
        for name, topk, nd, pre in (("verify_chain", 1, 4, (4, 7, 30)), ("verify_tree", 2, 3, (4, 7))):
            case = da.DenseAttentionCase(name=name, backend="b200", forward_mode=ForwardMode.TARGET_VERIFY,
                                         num_heads=4, num_kv_heads=4, page_size=16, prefix_lens=pre,
                                         extend_lens=(nd,) * len(pre))
            run(f"target_verify/{name}/d{hd}",
                lambda c=case, t=topk: tv.run_dense_eagle_verify_case(tc, c, topk=t, **kw))

    n_pass = sum(r[1] == "PASS" for r in results)
    n_fail = sum(r[1] == "FAIL" for r in results)
    n_skip = sum(r[1] == "SKIP" for r in results)
    lines = [f"reference dense-attention kit vs registered 'b200' backend ({torch.cuda.get_device_name(0) if torch.cuda.is_available() else 'no GPU'})",
             f"reference tree: {root}; torch {torch.__version__}",
             f"{n_pass} passed, {n_fail} failed, {n_skip} skipped of {len(results)}"] + notes + [""]
    lines += [f"{st:4s} {label}" + (f"   {msg}" if msg else "") for label, st, msg in results]
    text = "\n".join(lines)
    print(text)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            f.write(text + "\n")
    return 0 if n_fail == 0 and n_pass > 0 else 1


if __name__ == "__main__":
    sys.exit(main())
