"""Dev (GPU): the in-kernel e4m3 extend (b200attn_extend_fp8kv) against
  (a) the fp32 reference over the DEQUANTISED pool with exact q and p (what the dequantising-gather path computes), and
  (b) an emulation of the reference Triton kernel's arithmetic for an e4m3 pool: q -> e4m3, p -> e4m3
      (extend_attention.py:523,596), fp32 accumulation.
Prints max / mean abs errors; --time adds cfg3 timing vs the bf16 kernel."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from oracle import fp8_port
from sglang_b200 import ops
from tests.util import make_scenario

DEV = "cuda"
FP8 = torch.float8_e4m3fn


def emulate(q, k8f, v8f, lens, ext, Hq, Hkv, D, scale, ks, vs, causal, emulate_fp8=True):
    """q [T, Hq*D] (bf16 values as float), k8f / v8f: per-request lists of [L, Hkv, D] RAW e4m3 values as float."""
    outs = []
    t0 = 0
    G = Hq // Hkv
    for b, (L, E) in enumerate(zip(lens, ext)):
        qb = q[t0:t0 + E].view(E, Hq, D).float()
        if emulate_fp8:
            qb = qb.to(FP8).float()
        k, v = k8f[b], v8f[b]
        kk = k.repeat_interleave(G, dim=1)  # [L, Hq, D]
        vv = v.repeat_interleave(G, dim=1)
        s = torch.einsum("ehd,lhd->hel", qb, kk) * (scale * ks)
        if causal:
            pos = torch.arange(E, device=q.device)[:, None] + (L - E)
            s = s.masked_fill(torch.arange(L, device=q.device)[None, :] > pos, float("-inf"))
        m = s.max(dim=-1, keepdim=True).values
        p = torch.exp(s - m)
        den = p.sum(-1, keepdim=True)
        if emulate_fp8:
            p = p.to(FP8).float()
        o = torch.einsum("hel,lhd->ehd", p, vv) * vs / den.permute(1, 0, 2)
        outs.append(o.reshape(E, Hq * D))
        t0 += E
    return torch.cat(outs)


def run(name, pre, ext, Hq, Hkv, D, page, causal=True, seed=1, ks=0.6, vs=0.9):
    lens = [a + b for a, b in zip(pre, ext)]
    sc = make_scenario(pre, ext, Hq, Hkv, D, page, seed=seed)
    # fill the pool (old rows and the new tokens' rows) and quantise it
    kc, vc = sc.k_cache.clone(), sc.v_cache.clone()
    kc[sc.out_cache_loc] = sc.k.view(-1, Hkv, D)
    vc[sc.out_cache_loc] = sc.v.view(-1, Hkv, D)
    k8 = fp8_port.quantize(kc * 2, ks)
    v8 = fp8_port.quantize(vc * 2, vs)
    d = sc.to(DEV)
    B = len(lens)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    qo_indptr = torch.zeros(B + 1, dtype=torch.int32, device=DEV)
    kv_indices = torch.zeros(sum(lens), dtype=torch.int32, device=DEV)
    ops.build_kv_indices(d.req_to_token, d.req_pool_indices, torch.tensor(lens, device=DEV), kv_indptr, kv_indices,
                         qo_lens=torch.tensor(ext, dtype=torch.int32, device=DEV), qo_indptr=qo_indptr)
    q = d.q.view(-1, Hq, D)
    o = torch.full_like(q, float("nan"))
    k8d, v8d = k8.to(DEV), v8.to(DEV)
    ops.extend_attention(q, k8d, v8d, o, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=causal, page_size=page,
                         max_qo_len=max(ext), k_descale=ks, v_descale=vs)
    torch.cuda.synchronize()
    # per-request raw e4m3 values
    idx = kv_indices.long()
    ip = kv_indptr.tolist()
    k8f = [k8d.view(FP8)[idx[ip[b]:ip[b + 1]]].float() for b in range(B)]
    v8f = [v8d.view(FP8)[idx[ip[b]:ip[b + 1]]].float() for b in range(B)]
    qf = d.q.float()
    ref_exact = emulate(qf, k8f, v8f, lens, ext, Hq, Hkv, D, sc.scaling, ks, vs, causal, emulate_fp8=False)
    ref_emul = emulate(qf, k8f, v8f, lens, ext, Hq, Hkv, D, sc.scaling, ks, vs, causal, emulate_fp8=True)
    of = o.view(-1, Hq * D).float()
    fin = bool(torch.isfinite(of).all())
    e_exact = (of - ref_exact).abs()
    e_emul = (of - ref_emul).abs()
    e_ee = (ref_emul - ref_exact).abs()
    print(f"{name}: finite {fin} | vs exact max {float(e_exact.max()):.4f} mean {float(e_exact.mean()):.5f} | vs e4m3 "
          f"emulation max {float(e_emul.max()):.4f} mean {float(e_emul.mean()):.5f} | emulation vs exact max "
          f"{float(e_ee.max()):.4f} mean {float(e_ee.mean()):.5f} | |ref| mean {float(ref_exact.abs().mean()):.4f}",
          flush=True)


run("one tile, no prefix", [0], [128], 4, 4, 128, 16)
run("one request, prefix", [300], [200], 8, 2, 128, 16)
run("ragged, G=4", [0, 700, 33, 4000], [1280, 513, 2047, 96], 32, 8, 128, 16)
run("ragged, G=7, bidirectional, page 1", [17, 0, 300], [5, 3000, 129], 28, 4, 128, 1, causal=False)
run("head_dim 64: one tile", [0], [128], 4, 4, 64, 16)
run("head_dim 64: ragged, G=7, page 1", [17, 0, 300, 2000], [5, 3000, 129, 700], 14, 2, 64, 1)
run("head_dim 64: ragged, G=2, bidirectional", [0, 700, 33], [1280, 513, 2047], 8, 4, 64, 16, causal=False)
if "--time" in sys.argv:
    from tools.compare_flashinfer import timed
    from tools.sweep_extend_variants import make

    f16, flops, _ = make(32, 2048, 512, 32, 8)
    t16 = min(timed(f16, n=20, warm=5) for _ in range(3))
    page = 16
    B, P, E, Hq, Hkv, D = 32, 2048, 512, 32, 8, 128
    ctx = P + E
    sp, own = P // page, E // page
    n_pages = sp + B * own + 8
    perm = torch.randperm(n_pages - 1) + 1
    rows = [(torch.cat([perm[:sp], perm[sp + b * own: sp + (b + 1) * own]])[:, None] * page + torch.arange(page)).reshape(-1)
            for b in range(B)]
    kv_indices = torch.cat(rows).int().to(DEV)
    kv_indptr = torch.arange(0, (B + 1) * ctx, ctx, dtype=torch.int32, device=DEV)
    qo_indptr = torch.arange(0, (B + 1) * E, E, dtype=torch.int32, device=DEV)
    k8 = torch.randn((n_pages * page, Hkv, D), device=DEV).to(FP8).view(torch.uint8)
    v8 = torch.randn((n_pages * page, Hkv, D), device=DEV).to(FP8).view(torch.uint8)
    q = torch.randn((B * E, Hq, D), device=DEV).bfloat16()
    o = torch.empty_like(q)
    f8 = lambda: ops.extend_attention(q, k8, v8, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True,
                                      page_size=page, max_qo_len=E, k_descale=0.6, v_descale=0.9)
    t8 = min(timed(f8, n=20, warm=5) for _ in range(3))
    print(f"cfg3: 16-bit pool {t16 * 1e3:.1f} us ({flops / t16 / 1e9:.0f} TFLOP/s) | e4m3 pool in-kernel {t8 * 1e3:.1f} us "
          f"({flops / t8 / 1e9:.0f} TFLOP/s)")
