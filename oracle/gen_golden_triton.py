"""Generate tests/golden/triton_*.pt from the REFERENCE's production Triton kernels (run on the GPU box).

The features torch_native cannot express - logit cap, attention sinks, the sliding window, an e4m3 KV pool read
with k/v descale - are pinned here against the kernels the reference ships for them, imported from the staged tree
oracle/_ref/python:

  decode : `decode_attention_fwd`          python/sglang/kernels/ops/attention/decode_attention.py:1163-1239
  extend : `extend_attention_fwd_unified`  python/sglang/kernels/ops/attention/extend_attention.py:1271-1410
           (one KV source indexed by kv_indices over prefix + new tokens: this build's layout)

Each record holds the inputs (pool, slot table, q) and the reference output.  bf16, seeded.

Run on the GPU box:  python oracle/gen_golden_triton.py gpurun_out/golden   (then copy the files to tests/golden/)
"""
from __future__ import annotations

import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_env  # noqa: E402
from tests.util import make_scenario  # noqa: E402

DEV = "cuda"


def _indices(sc, lens):
    kv_indptr = torch.zeros(len(lens) + 1, dtype=torch.int32)
    kv_indptr[1:] = torch.cumsum(torch.tensor(lens), 0)
    kv_indices = torch.cat([sc.req_to_token[int(sc.req_pool_indices[b]), :lens[b]] for b in range(len(lens))]).to(torch.int64)
    return kv_indptr, kv_indices


def main(out_dir: str) -> None:
    ref_env.install()
    from sglang.kernels.ops.attention.decode_attention import decode_attention_fwd
    from sglang.kernels.ops.attention.extend_attention import extend_attention_fwd_unified

    os.makedirs(out_dir, exist_ok=True)
    Hq, Hkv, D, page = 8, 2, 128, 16
    sinks = torch.randn(Hq, generator=torch.Generator().manual_seed(5))
    variants = [("plain", {}), ("cap30", dict(logit_cap=30.0)), ("sinks", dict(sinks=sinks)),
                ("fp8", dict(fp8=True, k_scale=0.5, v_scale=0.25)), ("window100", dict(window=100))]

    # ---------------- decode
    lens = [5, 130, 1000, 2049]
    B = len(lens)
    sc = make_scenario(lens, [0] * B, Hq, Hkv, D, page, seed=61)
    q = torch.randn((B, Hq, D), generator=torch.Generator().manual_seed(2)).to(torch.bfloat16)
    kv_indptr, kv_indices = _indices(sc, lens)
    for name, kw in variants:
        if "window" in kw:
            continue  # the reference's decode window is a metadata rule (min(seq_len, W) keys), not a kernel flag
        kc, vc = sc.k_cache, sc.v_cache
        ks = vs = 1.0
        if kw.get("fp8"):
            ks, vs = kw["k_scale"], kw["v_scale"]
            kc = (kc.float() / ks).to(torch.float8_e4m3fn)
            vc = (vc.float() / vs).to(torch.float8_e4m3fn)
        splits = 8
        o = torch.empty((B, Hq, D), dtype=torch.bfloat16, device=DEV)
        logits = torch.empty((B, Hq, splits, D), dtype=torch.float32, device=DEV)
        lse = torch.empty((B, Hq, splits), dtype=torch.float32, device=DEV)
        nks = torch.full((B,), splits, dtype=torch.int32, device=DEV)
        s_dev = kw["sinks"].to(DEV) if "sinks" in kw else None
        decode_attention_fwd(q.to(DEV), kc.to(DEV), vc.to(DEV), o, kv_indptr.to(DEV), kv_indices.to(DEV), logits, lse,
                             nks, splits, sc.scaling, ks, vs, logit_cap=kw.get("logit_cap", 0.0), sinks=s_dev)
        torch.cuda.synchronize()
        rec = dict(kind="decode", name=name, Hq=Hq, Hkv=Hkv, D=D, page_size=page, lens=lens, scaling=sc.scaling,
                   q=q, k_cache=kc.cpu(), v_cache=vc.cpu(), req_to_token=sc.req_to_token,
                   req_pool_indices=sc.req_pool_indices, logit_cap=kw.get("logit_cap", 0.0),
                   sinks=kw.get("sinks"), k_scale=ks, v_scale=vs, o_ref=o.cpu(),
                   source="decode_attention.py:1163-1239 decode_attention_fwd", torch_version=torch.__version__)
        torch.save(rec, os.path.join(out_dir, f"triton_decode_{name}.pt"))
        print("decode", name, float(o.float().abs().mean()))

    # ---------------- extend (unified: prefix + new tokens through kv_indices)
    prefix, ext = [300, 0, 77], [200, 333, 140]
    sc = make_scenario(prefix, ext, Hq, Hkv, D, page, seed=62)
    T = sum(ext)
    # the new tokens' K/V are part of the pool (stored before the kernel runs)
    kc0, vc0 = sc.k_cache.clone(), sc.v_cache.clone()
    kc0[sc.out_cache_loc] = sc.k
    vc0[sc.out_cache_loc] = sc.v
    kv_indptr, kv_indices = _indices(sc, sc.seq_lens)
    qo_indptr = torch.zeros(len(ext) + 1, dtype=torch.int32)
    qo_indptr[1:] = torch.cumsum(torch.tensor(ext), 0)
    for name, kw in variants:
        kc, vc = kc0, vc0
        ks = vs = 1.0
        if kw.get("fp8"):
            ks, vs = kw["k_scale"], kw["v_scale"]
            kc = (kc.float() / ks).to(torch.float8_e4m3fn)
            vc = (vc.float() / vs).to(torch.float8_e4m3fn)
        o = torch.empty((T, Hq, D), dtype=torch.bfloat16, device=DEV)
        s_dev = kw["sinks"].to(DEV) if "sinks" in kw else None
        extend_attention_fwd_unified(
            sc.q.view(T, Hq, D).to(DEV), o, kc.to(DEV), vc.to(DEV), ks, vs, qo_indptr.to(DEV), kv_indptr.to(DEV),
            kv_indices.to(DEV), torch.tensor(prefix, dtype=torch.int32, device=DEV), max(ext), sm_scale=sc.scaling,
            logit_cap=kw.get("logit_cap", 0.0), is_causal=True, sliding_window_size=kw.get("window", -1), sinks=s_dev)
        torch.cuda.synchronize()
        rec = dict(kind="extend", name=name, Hq=Hq, Hkv=Hkv, D=D, page_size=page, prefix_lens=prefix, extend_lens=ext,
                   seq_lens=sc.seq_lens, scaling=sc.scaling, q=sc.q, k_cache=kc.cpu(), v_cache=vc.cpu(),
                   req_to_token=sc.req_to_token, req_pool_indices=sc.req_pool_indices,
                   logit_cap=kw.get("logit_cap", 0.0), sinks=kw.get("sinks"), window=kw.get("window", -1),
                   k_scale=ks, v_scale=vs, o_ref=o.cpu().view(T, Hq * D),
                   source="extend_attention.py:1271-1410 extend_attention_fwd_unified", torch_version=torch.__version__)
        torch.save(rec, os.path.join(out_dir, f"triton_extend_{name}.pt"))
        print("extend", name, float(o.float().abs().mean()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
