"""Same-box GPU comparators: the attention libraries SGLang's production backends call on B200 —
flashinfer's paged decode / prefill wrappers (flashinfer_backend.py) and the trtllm-gen cubins
(trtllm_mha_backend.py:1113,1430) — timed with the SAME CUDA-event loop, shapes and paged layout as bench.py's
cfg2 (decode) and cfg3 (extend), beside b200attn.  LIBRARY code, comparator only: nothing here is on the product path.
GPU box only; flashinfer JIT-compiles its launchers on first use (minutes), each leg has its own timeout.

    python tools/compare_flashinfer.py > gpurun_out/compare.jsonl
"""
import json
import signal
import sys
import time

import torch

sys.path.insert(0, ".")
from sglang_b200 import ops  # noqa: E402

DEV = "cuda"


class Timeout(Exception):
    pass


def _alarm(*_):
    raise Timeout()


def timed(fn, n=30, warm=5):
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


def leg(name, build, work, unit, budget_s=420):
    """build() -> (fn, check) ; reports ms and work/ms."""
    signal.signal(signal.SIGALRM, _alarm)
    signal.alarm(budget_s)
    t0 = time.time()
    try:
        fn, err = build()
        ms = timed(fn)
        out = dict(leg=name, ms=ms, value=work / (ms * 1e-3) / (1e9 if unit == "GB/s" else 1e12), unit=unit,
                   max_abs_err_vs_b200attn=err, setup_s=round(time.time() - t0, 1))
    except Timeout:
        out = dict(leg=name, unavailable=f"timeout after {budget_s}s (JIT compile)")
    except Exception as ex:  # noqa: BLE001
        out = dict(leg=name, unavailable=f"{type(ex).__name__}: {str(ex)[:200]}")
    finally:
        signal.alarm(0)
    print(json.dumps(out), flush=True)


def paged_state(B, ctx, Hkv, D, page, shared_prefix=0, gen=None):
    """Random page table [B, ctx/page] (page 0 reserved) + NHD pools viewed as [pages, page, Hkv, D]."""
    ppr = ctx // page
    sh = shared_prefix // page
    n_pages = sh + B * (ppr - sh) + 8
    perm = torch.randperm(n_pages - 1, generator=gen) + 1
    table = torch.empty((B, ppr), dtype=torch.int32)
    for b in range(B):
        table[b, :sh] = perm[:sh]
        table[b, sh:] = perm[sh + b * (ppr - sh): sh + (b + 1) * (ppr - sh)]
    kc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    vc = torch.randn((n_pages * page, Hkv, D), device=DEV).bfloat16()
    return table.to(DEV), kc, vc, n_pages


def main():
    import flashinfer
    g = torch.Generator().manual_seed(20261104)
    # ---------------- cfg2: decode B=64 x 4096, 32q/8kv, D=128, page 16
    B, ctx, Hq, Hkv, D, page = 64, 4096, 32, 8, 128, 16
    table, kc, vc, n_pages = paged_state(B, ctx, Hkv, D, page, gen=g)
    q = torch.randn((B, Hq, D), device=DEV).bfloat16()
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
    o_ref = torch.empty_like(q)
    nbytes = B * ctx * Hkv * D * 2 * 2 + 2 * B * Hq * D * 2 + B * ctx * 4
    scale = D ** -0.5

    pwarps, pms, ptail = ops.plan_decode(B, Hkv, Hq // Hkv, ctx, 256)  # the launch plan the backend would pick
    pws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, pms), 16), dtype=torch.uint8, device=DEV)

    def b200_decode():
        return (lambda: ops.decode_attention(q, kc, vc, o_ref, kv_indptr, kv_indices, scale, max_splits=pms,
                                             warps=pwarps, tail_begin=ptail, workspace=pws)), 0.0

    leg("cfg2_decode_b200attn", b200_decode, nbytes, "GB/s")
    seq = torch.full((B,), ctx, dtype=torch.int32, device=DEV)
    k4, v4 = kc.view(n_pages, page, Hkv, D), vc.view(n_pages, page, Hkv, D)
    ws = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV)

    def fi_decode(tensor_cores):
        def build():
            w = flashinfer.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD", use_tensor_cores=tensor_cores)
            indptr = torch.arange(0, (B + 1) * (ctx // page), ctx // page, dtype=torch.int32, device=DEV)
            w.plan(indptr, table.reshape(-1), torch.full((B,), page, dtype=torch.int32, device=DEV), Hq, Hkv, D, page,
                   q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16, sm_scale=scale)
            o = w.run(q, (k4, v4))
            return (lambda: w.run(q, (k4, v4))), float((o.float() - o_ref.float()).abs().max())
        return build

    leg("cfg2_decode_flashinfer_wrapper_tensor_cores", fi_decode(True), nbytes, "GB/s")
    leg("cfg2_decode_flashinfer_wrapper_cuda_cores", fi_decode(False), nbytes, "GB/s")

    def trt_decode():
        from flashinfer.decode import trtllm_batch_decode_with_kv_cache as f
        call = lambda: f(q, (k4, v4), ws, table, seq, ctx, bmm1_scale=scale, bmm2_scale=1.0, kv_layout="NHD")
        o = call()
        return call, float((o.float() - o_ref.float()).abs().max())

    leg("cfg2_decode_trtllm_gen", trt_decode, nbytes, "GB/s")
    # the same two kernels on other page placements (the random permutation above is the worst case): every request's
    # pages in one ascending run (a fresh prefill), and page j of request b at b + j * B (64 requests decoding in step)
    ppr = ctx // page
    for pname, pages in (("contiguous", 1 + torch.arange(B)[:, None] * ppr + torch.arange(ppr)[None, :]),
                         ("interleaved", 1 + torch.arange(B)[:, None] + torch.arange(ppr)[None, :] * B)):
        table2 = pages.to(torch.int32).to(DEV).contiguous()
        idx2 = (table2.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()

        def b200_decode2():
            return (lambda: ops.decode_attention(q, kc, vc, o_ref, kv_indptr, idx2, scale, max_splits=pms,
                                                 warps=pwarps, tail_begin=ptail, workspace=pws)), 0.0

        leg(f"cfg2_decode_b200attn_{pname}_pages", b200_decode2, nbytes, "GB/s")

        def trt_decode2():
            from flashinfer.decode import trtllm_batch_decode_with_kv_cache as f
            call = lambda: f(q, (k4, v4), ws, table2, seq, ctx, bmm1_scale=scale, bmm2_scale=1.0, kv_layout="NHD")
            o = call()
            return call, float((o.float() - o_ref.float()).abs().max())

        leg(f"cfg2_decode_trtllm_gen_{pname}_pages", trt_decode2, nbytes, "GB/s")
    del kc, vc, k4, v4, kv_indices
    torch.cuda.empty_cache()

    # ---------------- cfg3: extend 32 x (2048 shared prefix + 512 new), causal
    B, P, E = 32, 2048, 512
    table, kc, vc, n_pages = paged_state(B, P + E, Hkv, D, page, shared_prefix=P, gen=g)
    T = B * E
    q = torch.randn((T, Hq, D), device=DEV).bfloat16()
    o_ref = torch.empty_like(q)
    qo_indptr = torch.arange(0, T + 1, E, dtype=torch.int32, device=DEV)
    kv_indptr = torch.arange(0, (B + 1) * (P + E), P + E, dtype=torch.int32, device=DEV)
    kv_indices = (table.long()[:, :, None] * page + torch.arange(page, device=DEV)).reshape(-1).int()
    flops = 4 * D * Hq * B * (E * P + E * (E + 1) // 2)

    def b200_extend():
        return (lambda: ops.extend_attention(q, kc, vc, o_ref, qo_indptr, kv_indptr, kv_indices, scale, causal=True,
                                             page_size=page, max_qo_len=E)), 0.0

    leg("cfg3_extend_b200attn", b200_extend, flops, "TFLOP/s")
    k4, v4 = kc.view(n_pages, page, Hkv, D), vc.view(n_pages, page, Hkv, D)
    ppr = (P + E) // page

    def fi_prefill():
        w = flashinfer.BatchPrefillWithPagedKVCacheWrapper(ws, "NHD")
        indptr = torch.arange(0, (B + 1) * ppr, ppr, dtype=torch.int32, device=DEV)
        w.plan(qo_indptr, indptr, table.reshape(-1), torch.full((B,), page, dtype=torch.int32, device=DEV), Hq, Hkv, D,
               page, causal=True, sm_scale=scale, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
        o = w.run(q, (k4, v4))
        return (lambda: w.run(q, (k4, v4))), float((o.float() - o_ref.float()).abs().max())

    leg("cfg3_extend_flashinfer_wrapper", fi_prefill, flops, "TFLOP/s")

    def trt_context():
        from flashinfer.prefill import trtllm_batch_context_with_kv_cache as f
        seq = torch.full((B,), P + E, dtype=torch.int32, device=DEV)
        cum_kv = torch.arange(0, (B + 1) * (P + E), P + E, dtype=torch.int32, device=DEV)
        call = lambda: f(q, (k4, v4), ws, table, seq, E, P + E, scale, 1.0, B, qo_indptr, cum_kv, kv_layout="NHD")
        o = call()
        return call, float((o.float() - o_ref.float()).abs().max())

    leg("cfg3_extend_trtllm_gen", trt_context, flops, "TFLOP/s")


if __name__ == "__main__":
    main()
