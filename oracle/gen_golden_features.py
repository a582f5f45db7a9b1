"""Generate tests/golden/swa_*.pt and verify_*.pt from the reference kit's OWN CPU references (this container only).

torch_native has no sliding window and no custom masks, so the reference outputs here come from the two
independent fp32 references the kit ships and every production backend is tested against:

  * `_dense_attention_reference`  (python/sglang/test/kits/attention_unittest/attention_methods/
    dense_attention.py:802-855; window rule :831-840, SWA cases :212-260) for sliding-window extend / decode;
  * `dense_attention_reference_with_custom_mask` (:858-905) with the masks of
    runner_modes/speculative_target_verify_runner.py:199-263 for TARGET_VERIFY tree / chain masks.

The kit's fixture builder (real MockModelRunner, ReqToTokenPool, MHATokenToKVPool, ForwardBatch, projections)
produces the inputs; the reference's o_proj is replaced by the identity so `o_ref` is the attention output itself.
Only the pool rows a request references are stored (`pool_slots`, `k_pool_rows`, `v_pool_rows`).

Run:  python oracle/gen_golden_features.py     (needs /root/reference; NOT run on the GPU box)
"""
from __future__ import annotations

import json
import os
import sys
import unittest

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_env  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main() -> None:
    ref_env.install()
    from sglang.srt.model_executor.forward_batch_info import ForwardMode
    from sglang.test.kits.attention_unittest.attention_methods import dense_attention as da

    _, tv = ref_env.import_runner_modes()

    tc = unittest.TestCase()
    E, D_, V = ForwardMode.EXTEND, ForwardMode.DECODE, ForwardMode.TARGET_VERIFY
    # (name, mode, Hq, Hkv, head_dim, page, prefix_lens, extend_lens, window, topk, ctx)
    cases = [
        # the kit's own SWA cases (dense_attention.py:212-260) at a head_dim the sm_100a kernels take
        ("swa_extend_no_prefix_window_edges", E, 4, 4, 128, 16, (0, 0, 0), (3, 4, 5), 4, 0, 64),
        ("swa_extend_prefix_window_edges", E, 4, 4, 128, 16, (3, 4, 5), (2, 2, 2), 4, 0, 64),
        # windows that cross K/V tile (128) and chunk (64) boundaries, GQA, head_dim 64
        ("swa_extend_w100_gqa", E, 8, 2, 128, 16, (300, 0, 77), (200, 333, 140), 100, 0, 512),
        ("swa_extend_w64_d64", E, 4, 2, 64, 16, (129, 10), (130, 255), 64, 0, 384),
        ("swa_decode_w4_edges", D_, 4, 4, 128, 16, (2, 3, 4, 5, 40), (), 4, 0, 64),
        ("swa_decode_w100", D_, 8, 2, 128, 16, (99, 100, 101, 450), (), 100, 0, 512),
        # TARGET_VERIFY: chain mask (topk 1) and the kit's two-branch tree (topk 2, draft 3)
        ("verify_chain_draft4", V, 8, 2, 128, 16, (17, 64, 5), (4, 4, 4), None, 1, 128),
        ("verify_tree_draft3", V, 4, 4, 128, 16, (31, 2, 100), (3, 3, 3), None, 2, 128),
        ("verify_tree_draft3_swa", V, 4, 2, 64, 16, (31, 2, 100), (3, 3, 3), 16, 2, 128),
    ]
    index = []
    for (name, mode, hq, hkv, hd, page, prefix, ext, window, topk, ctx) in cases:
        case = da.DenseAttentionCase(name=name, backend="torch_native", forward_mode=mode, num_heads=hq,
                                     num_kv_heads=hkv, page_size=page, prefix_lens=prefix, extend_lens=ext,
                                     sliding_window_size=window)
        fx = da.build_dense_attention_fixture(tc, case, head_dim=hd, hidden_size=hq * hd // 2, max_context_len=ctx,
                                              dtype=torch.bfloat16, device="cpu", loc_layout="shuffled_pages")
        fx.reference_module.reconstruct_output = lambda x: x  # o_ref = the attention output, not o_proj(attention)
        fb = fx.forward_batch
        pool = fx.runner.token_to_kv_pool
        layer = fx.actual_module.attn
        with torch.no_grad():
            q, k, v = fx.actual_module.project_qkv(fx.input_hidden)
            mask_flat = None
            if mode.is_target_verify():
                masks_by_req, mask_flat = tv._make_custom_masks(case, topk=topk, device="cpu")
                o_ref = da.dense_attention_reference_with_custom_mask(fx.reference_module, case, fx.prefix_hidden,
                                                                      fx.input_hidden, masks_by_req)
            else:
                o_ref = da._dense_attention_reference(fx.reference_module, case, fx.prefix_hidden, fx.input_hidden)
        table = fx.runner.req_to_token_pool.req_to_token[: case.batch_size + 1].clone()
        kb, vb = pool.get_key_buffer(layer.layer_id), pool.get_value_buffer(layer.layer_id)
        used = torch.unique(torch.cat([table[b, : case.seq_lens[b]].long() for b in range(case.batch_size)]))
        rec = dict(
            name=name, mode=mode.name, num_heads=hq, num_kv_heads=hkv, head_dim=hd, page_size=page,
            scaling=float(layer.scaling), prefix_lens=list(prefix), extend_lens=list(case.input_lens),
            seq_lens=list(case.seq_lens), window=window, topk=topk,
            q=q.contiguous(), k=k.contiguous(), v=v.contiguous(), req_to_token=table,
            req_pool_indices=fb.req_pool_indices.clone(), out_cache_loc=fb.out_cache_loc.clone(),
            positions=fb.positions.clone(), pool_slots_total=int(kb.shape[0]), pool_slots=used,
            k_pool_rows=kb[used].clone(), v_pool_rows=vb[used].clone(),
            custom_mask=mask_flat, o_ref=o_ref.contiguous(), torch_version=torch.__version__,
        )
        path = os.path.join(OUT, f"{name}.pt")
        torch.save(rec, path)
        index.append(dict(name=name, file=f"{name}.pt", bytes=os.path.getsize(path)))
        print(f"{name}: o {tuple(o_ref.shape)} {os.path.getsize(path) / 1024:.0f} KiB")
    with open(os.path.join(OUT, "INDEX_features.json"), "w") as f:
        json.dump(dict(generator="oracle/gen_golden_features.py",
                       reference="/root/reference/python: dense_attention.py _dense_attention_reference / "
                                 "dense_attention_reference_with_custom_mask on the kit's own fixtures",
                       torch=torch.__version__, cases=index), f, indent=1)


if __name__ == "__main__":
    main()
