"""Same-box comparators (ii) of SURVEY §6: the reference's own Triton kernels, imported from the staged tree
oracle/_ref/python, timed with the same CUDA-event loop and paged layout as tools/compare_flashinfer.py on cfg2
(decode) and cfg3 (extend).  Comparator only; GPU box only.

    python tools/compare_reference_kernels.py > gpurun_out/compare_triton.jsonl
"""
import json
import sys

import torch

sys.path.insert(0, ".")
from oracle import ref_env  # noqa: E402
from sglang_b200 import ops  # noqa: E402
from tools.compare_flashinfer import timed  # noqa: E402

DEV = "cuda"
page = 16


def main():
    ref_env.install()
    from sglang.kernels.ops.attention.decode_attention import decode_attention_fwd
    from sglang.kernels.ops.attention.extend_attention import extend_attention_fwd_unified

    Hq, Hkv, D = 32, 8, 128
    # ---- cfg2 decode
    B, ctx = 64, 4096
    n_pages = B * ctx // page + 8
    perm = torch.randperm(n_pages - 1) + 1
    table = perm[: B * ctx // page].view(B, ctx // page).to(DEV)
    kv_indices = (table[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1)
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    q = torch.randn((B, Hq, D), device=DEV).bfloat16()
    o, o2 = torch.empty_like(q), torch.empty_like(q)
    nbytes = B * ctx * Hkv * D * 4
    warps, ms, tb = ops.plan_decode(B, Hkv, Hq // Hkv, ctx, 256)
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
    i32 = kv_indices.int()
    t = timed(lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, i32, D ** -0.5, max_splits=ms, workspace=ws,
                                           warps=warps, tail_begin=tb))
    print(json.dumps(dict(leg="cfg2_decode_b200attn", ms=t, value=nbytes / t / 1e6, unit="GB/s")), flush=True)
    for splits in (1, 2, 4, 8):
        logits = torch.empty((B, Hq, splits, D), dtype=torch.float32, device=DEV)
        lse = torch.empty((B, Hq, splits), dtype=torch.float32, device=DEV)
        nks = torch.full((B,), splits, dtype=torch.int32, device=DEV)
        f = lambda: decode_attention_fwd(q, kc, vc, o2, kv_indptr, kv_indices, logits, lse, nks, splits, D ** -0.5,
                                         1.0, 1.0)
        t = timed(f)
        print(json.dumps(dict(leg=f"cfg2_decode_reference_triton_splits{splits}", ms=t, value=nbytes / t / 1e6,
                              unit="GB/s", max_abs_err_vs_b200attn=float((o.float() - o2.float()).abs().max()))),
              flush=True)
    del kc, vc
    # ---- cfg3 extend (radix-hit layout)
    B, P, E = 32, 2048, 512
    ctx = P + E
    sp, own = P // page, E // page
    n_pages = sp + B * own + 8
    perm = torch.randperm(n_pages - 1) + 1
    rows = []
    for b in range(B):
        pages = torch.cat([perm[:sp], perm[sp + b * own: sp + (b + 1) * own]])
        rows.append((pages[:, None] * page + torch.arange(page)).reshape(-1))
    kv_indices = torch.cat(rows).to(DEV)
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    qo_indptr = torch.arange(0, (B + 1) * E, E, dtype=torch.int32, device=DEV)
    kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    q = torch.randn((B * E, Hq, D), device=DEV).bfloat16()
    o, o2 = torch.empty_like(q), torch.empty_like(q)
    flops = 4 * D * Hq * B * (E * P + E * (E + 1) // 2)
    t = timed(lambda: ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices.int(), D ** -0.5, causal=True,
                                           page_size=page, max_qo_len=E), n=20)
    print(json.dumps(dict(leg="cfg3_extend_b200attn", ms=t, value=flops / t / 1e9, unit="TFLOP/s")), flush=True)
    pre = torch.full((B,), P, dtype=torch.int32, device=DEV)
    f = lambda: extend_attention_fwd_unified(q, o2, kc, vc, 1.0, 1.0, qo_indptr, kv_indptr, kv_indices, pre, E,
                                             sm_scale=D ** -0.5, is_causal=True)
    t = timed(f, n=10)
    print(json.dumps(dict(leg="cfg3_extend_reference_triton_unified", ms=t, value=flops / t / 1e9, unit="TFLOP/s",
                          max_abs_err_vs_b200attn=float((o.float() - o2.float()).abs().max()))), flush=True)


if __name__ == "__main__":
    main()
