"""Dev: per-CTA timeline of decode_ws_kernel (libb200attn_trace.so, `make -C sglang_b200/csrc trace`).
Stamps (%globaltimer ns, per CTA): 0 entry, 1 after griddepcontrol.wait, 2 partition + barrier init done,
3 producer: first tile issued, 4 consumers: first tile landed, 5 consumers: last tile computed, 6 exit."""
import os
import sys

import torch

sys.path.insert(0, ".")
from sglang_b200 import _lib, ops  # noqa: E402

DEV = "cuda"
Hq, Hkv, D, page = 32, 8, 128, 16


def main():
    _lib.load()
    _lib._lib = _lib.load_variant(os.path.join(os.path.dirname(_lib.LIB_PATH), "libb200attn_trace.so"))
    for B, ctx in [(64, 4096), (16, 4096), (256, 8192)]:
        n_pages = B * ctx // page + 8
        perm = torch.randperm(n_pages - 1) + 1
        table = perm[: B * ctx // page].view(B, ctx // page).int().to(DEV)
        kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
        kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
        kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
        vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
        q = torch.randn((B, Hq, D), device=DEV).bfloat16()
        o = torch.empty_like(q)
        ws = torch.zeros(ops.decode_ws_workspace_bytes(Hq, Hkv, D), dtype=torch.uint8, device=DEV)
        ctas = 2 * ops.sm_count()
        trace = torch.zeros((max(ctas * 8, B * Hq) * 2,), dtype=torch.float32, device=DEV)  # viewed as int64 [ctas][8]
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(4):
            trace.zero_()
            ev0.record()
            ops.decode_attention(q, kc, vc, o, kv_indptr, kv_indices, D ** -0.5, min_chunk=256, workspace=ws,
                                 lse=trace, ws_kernel=True)
            ev1.record()
            torch.cuda.synchronize()
        t = trace.view(torch.int64)[: ctas * 8].view(ctas, 8).cpu().double()
        live = t[:, 6] > 0
        t = t[live]
        t0 = t[:, 0].min()
        rel = (t - t0) / 1e3
        names = ["entry", "after_wait", "prologue_done", "first_issue", "first_landed", "last_computed", "exit"]
        print(f"B={B} ctx={ctx}: event time {ev0.elapsed_time(ev1) * 1e3:.1f} us, {int(live.sum())} CTAs")
        for j, nm in enumerate(names):
            col = rel[:, j]
            qs = torch.quantile(col, torch.tensor([0.0, 0.1, 0.25, 0.5, 0.75, 0.9, 1.0], dtype=torch.float64))
            print(f"   {nm:14s} " + "  ".join(f"{float(x):8.2f}" for x in qs) + f"   mean {float(col.mean()):8.2f}")
        fin = rel[:, 5]
        print("   last_computed by CTA index (first 40):", [round(float(x)) for x in fin[:40]])
        print("   last_computed, CTA c vs c+148 (same SM?) corr:",
              float(torch.corrcoef(torch.stack([fin[:148], fin[148:296]]))[0, 1]) if fin.numel() == 296 else None)
        del kc, vc


if __name__ == "__main__":
    main()
