"""Small-batch / long-context decode latency: b200attn (length-based split policy) vs trtllm-gen (library comparator).
GPU box only."""
import json
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import ops  # noqa: E402
from tools.compare_flashinfer import timed  # noqa: E402

DEV = "cuda"
Hq, Hkv, D, page = 32, 8, 128, 16


def main():
    from flashinfer.decode import trtllm_batch_decode_with_kv_cache as trt
    ws_t = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV)
    for B, ctx in ([(4, 8192), (8, 8192), (16, 4096), (16, 16384), (24, 8192), (32, 8192), (36, 4096)] if len(sys.argv) > 1 else [(1, 8192), (1, 32768), (1, 131072), (4, 8192), (4, 32768), (16, 4096), (16, 16384), (32, 8192)]):
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
        res = dict(B=B, ctx=ctx)
        best = None
        for ms in sorted({ops.pick_max_splits(B, Hkv, Hq // Hkv, ctx, 256), 1, 2, max(1, 296 // (B * Hkv)), 2 * max(1, 296 // (B * Hkv)), 4, 8} if len(sys.argv) > 1 else {ops.pick_max_splits(B, Hkv, Hq // Hkv, ctx, 256), 4, 8, 16, 32, 64}):
            ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
            t = timed(lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5, max_splits=ms,
                                                   min_chunk=256, workspace=ws))
            res[f"b200_ms{ms}_us"] = round(t * 1e3, 1)
            if best is None or t < best:
                best = t
        res["b200_policy_splits"] = ops.pick_max_splits(B, Hkv, Hq // Hkv, ctx, 256)
        res["b200_best_gbs"] = round(nbytes / best / 1e6)
        seq = torch.full((B,), ctx, dtype=torch.int32, device=DEV)
        k4, v4 = kc.view(n_pages, page, Hkv, D), vc.view(n_pages, page, Hkv, D)
        try:
            t = timed(lambda: trt(q, (k4, v4), ws_t, table, seq, ctx, bmm1_scale=D ** -0.5, bmm2_scale=1.0, kv_layout="NHD"))
            res["trtllm_us"] = round(t * 1e3, 1)
            res["trtllm_gbs"] = round(nbytes / t / 1e6)
        except Exception as ex:  # noqa: BLE001
            res["trtllm"] = str(ex)[:100]
        print(json.dumps(res), flush=True)
        del kc, vc, k4, v4


if __name__ == "__main__":
    main()
