"""Dev timing: decode_ws vs the split-KV kernel vs trtllm-gen on the comparator shapes (GPU box)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import ops  # noqa: E402
from tools.compare_flashinfer import timed  # noqa: E402

DEV = "cuda"
Hq, Hkv, D, page = 32, 8, 128, 16


def main():
    try:
        from flashinfer.decode import trtllm_batch_decode_with_kv_cache as trt
    except Exception:  # noqa: BLE001
        trt = None
    ws_t = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV)
    shapes = [(64, 4096), (1, 8192), (1, 32768), (1, 131072), (4, 8192), (4, 32768), (16, 4096), (16, 16384),
              (32, 8192), (256, 8192)]
    for B, ctx in shapes:
        n_pages = B * ctx // page + 8
        perm = torch.randperm(n_pages - 1) + 1
        table = perm[: B * ctx // page].view(B, ctx // page).int().to(DEV)
        kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
        kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
        kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
        vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
        q = torch.randn((B, Hq, D), device=DEV).bfloat16()
        o, o2 = torch.empty_like(q), torch.empty_like(q)
        nbytes = B * ctx * Hkv * D * 4
        res = dict(B=B, ctx=ctx)
        warps, ms, tb = ops.plan_decode(B, Hkv, Hq // Hkv, ctx, 256)
        ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms), 16), dtype=torch.uint8, device=DEV)
        t = timed(lambda: ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5, max_splits=ms,
                                               min_chunk=256, workspace=ws, warps=warps, tail_begin=tb))
        res["split_us"] = round(t * 1e3, 1)
        res["plan"] = [warps, ms, tb]
        ws2 = torch.zeros(ops.decode_ws_workspace_bytes(Hq, Hkv, D), dtype=torch.uint8, device=DEV)
        for mc in (64, 256, 512):
            t = timed(lambda: ops.decode_attention(q, kc, vc, o2, kv_indptr, kv_indices, D ** -0.5, min_chunk=mc,
                                                   workspace=ws2, ws_kernel=True))
            res[f"ws_mc{mc}_us"] = round(t * 1e3, 1)
        res["ws_gbs"] = round(nbytes / min(res["ws_mc64_us"], res["ws_mc256_us"], res["ws_mc512_us"]) / 1e3)
        res["max_abs_diff"] = float((o.float() - o2.float()).abs().max())
        if trt is not None:
            seq = torch.full((B,), ctx, dtype=torch.int32, device=DEV)
            k4, v4 = kc.view(n_pages, page, Hkv, D), vc.view(n_pages, page, Hkv, D)
            try:
                t = timed(lambda: trt(q, (k4, v4), ws_t, table, seq, ctx, bmm1_scale=D ** -0.5, bmm2_scale=1.0,
                                      kv_layout="NHD"))
                res["trtllm_us"] = round(t * 1e3, 1)
            except Exception as ex:  # noqa: BLE001
                res["trtllm"] = str(ex)[:80]
        print(json.dumps(res), flush=True)
        del kc, vc


if __name__ == "__main__":
    main()
