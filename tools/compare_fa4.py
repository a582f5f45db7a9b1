"""Same-box comparator (iii) of SURVEY §6: the reference's FA4 kernel (CuTe-DSL,
python/sglang/kernels/ops/attention/flash_attn/cute/flash_fwd_sm100.py:214 through interface.flash_attn_varlen_func),
imported from the staged tree oracle/_ref/python, on cfg3 (extend, paged KV through a page table) and cfg2 (decode).
The image ships nvidia-cutlass-dsl 4.5.0 against the reference's pinned 4.6.2: a leg that does not compile here is
reported as unavailable with the reason.  Comparator only; GPU box only.

    python tools/compare_fa4.py > gpurun_out/compare_fa4.jsonl
"""
import json
import sys
import traceback

import torch

sys.path.insert(0, ".")
from oracle import ref_env  # noqa: E402
from sglang_b200 import ops  # noqa: E402
from tools.compare_flashinfer import timed  # noqa: E402

DEV = "cuda"
page = 16
Hq, Hkv, D = 32, 8, 128


def leg(name, fn, work, unit, ref=None, out=None, n=20):
    try:
        fn()
        torch.cuda.synchronize()
        t = timed(fn, n=n)
        rec = dict(leg=name, ms=t, value=work / t / (1e6 if unit == "GB/s" else 1e9), unit=unit)
        if ref is not None:
            rec["max_abs_err_vs_b200attn"] = float((out().float() - ref.float()).abs().max())
    except Exception as e:  # noqa: BLE001
        rec = dict(leg=name, unavailable=f"{type(e).__name__}: {str(e)[:300]}")
        traceback.print_exc()
    print(json.dumps(rec), flush=True)


def main():
    ref_env.install()
    try:
        from sglang.kernels.ops.attention.flash_attn.cute.interface import flash_attn_varlen_func
    except Exception as e:  # noqa: BLE001
        print(json.dumps(dict(leg="fa4_import", unavailable=f"{type(e).__name__}: {str(e)[:300]}")), flush=True)
        traceback.print_exc()
        return
    # ---- cfg3 extend: 32 x (2048 shared prefix + 512 new)
    B, P, E = 32, 2048, 512
    ctx = P + E
    sp, own = P // page, E // page
    n_pages = sp + B * own + 8
    perm = torch.randperm(n_pages - 1) + 1
    table = torch.stack([torch.cat([perm[:sp], perm[sp + b * own: sp + (b + 1) * own]]) for b in range(B)]).int().to(DEV)
    kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    qo_indptr = torch.arange(0, (B + 1) * E, E, dtype=torch.int32, device=DEV)
    kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    q = torch.randn((B * E, Hq, D), device=DEV).bfloat16()
    o = torch.empty_like(q)
    flops = 4 * D * Hq * B * (E * P + E * (E + 1) // 2)
    ours = lambda: ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True,
                                        page_size=page, max_qo_len=E)
    leg("cfg3_extend_b200attn", ours, flops, "TFLOP/s")
    k4, v4 = kc.view(n_pages, page, Hkv, D), vc.view(n_pages, page, Hkv, D)
    seqk = torch.full((B,), ctx, dtype=torch.int32, device=DEV)
    res = {}

    def fa4():
        r = flash_attn_varlen_func(q, k4, v4, cu_seqlens_q=qo_indptr, max_seqlen_q=E, max_seqlen_k=ctx, seqused_k=seqk,
                                   page_table=table, softmax_scale=D ** -0.5, causal=True)
        res["o"] = r[0] if isinstance(r, tuple) else r

    leg("cfg3_extend_reference_fa4", fa4, flops, "TFLOP/s", ref=o, out=lambda: res["o"])
    del kc, vc
    # ---- cfg2 decode: 64 x 4096
    B, ctx = 64, 4096
    n_pages = B * ctx // page + 8
    perm = torch.randperm(n_pages - 1) + 1
    table = perm[: B * ctx // page].view(B, ctx // page).int().to(DEV)
    kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    q = torch.randn((B, Hq, D), device=DEV).bfloat16()
    o = torch.empty_like(q)
    nbytes = B * ctx * Hkv * D * 4
    warps, ms, tb = ops.plan_decode(B, Hkv, Hq // Hkv, ctx, 256)
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
    leg("cfg2_decode_b200attn", lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5,
                                                             max_splits=ms, workspace=ws, warps=warps, tail_begin=tb),
        nbytes, "GB/s")
    k4, v4 = kc.view(n_pages, page, Hkv, D), vc.view(n_pages, page, Hkv, D)
    seqk = torch.full((B,), ctx, dtype=torch.int32, device=DEV)
    cuq = torch.arange(0, B + 1, dtype=torch.int32, device=DEV)

    def fa4d():
        r = flash_attn_varlen_func(q, k4, v4, cu_seqlens_q=cuq, max_seqlen_q=1, max_seqlen_k=ctx, seqused_k=seqk,
                                   page_table=table, softmax_scale=D ** -0.5, causal=True)
        res["o"] = r[0] if isinstance(r, tuple) else r

    leg("cfg2_decode_reference_fa4", fa4d, nbytes, "GB/s", ref=o, out=lambda: res["o"])


if __name__ == "__main__":
    main()
