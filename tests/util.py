"""Shared scenario builders for the parity tests (seeded, device agnostic)."""
from __future__ import annotations

import glob
import os
from dataclasses import dataclass
from typing import List, Sequence

import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_files() -> List[str]:
    """Attention-backend fixtures (oracle/gen_golden.py)."""
    other = ("rope_", "alloc_", "swa_", "verify_", "mrope_", "merge_", "radix_", "triton_")
    return sorted(p for p in glob.glob(os.path.join(GOLDEN_DIR, "*.pt")) if not os.path.basename(p).startswith(other))


def feature_golden_files() -> List[str]:
    """Sliding-window and TARGET_VERIFY custom-mask fixtures from the reference kit's own fp32 references
    (oracle/gen_golden_features.py)."""
    return sorted(glob.glob(os.path.join(GOLDEN_DIR, "swa_*.pt")) + glob.glob(os.path.join(GOLDEN_DIR, "verify_*.pt")))


def feature_pool(g: dict):
    """Rebuild the (k, v) pool of a feature fixture from its referenced rows (all other slots are zero)."""
    shape = (g["pool_slots_total"],) + tuple(g["k_pool_rows"].shape[1:])
    kc = torch.zeros(shape, dtype=g["k_pool_rows"].dtype)
    vc = torch.zeros(shape, dtype=g["v_pool_rows"].dtype)
    kc[g["pool_slots"]] = g["k_pool_rows"]
    vc[g["pool_slots"]] = g["v_pool_rows"]
    return kc, vc


def rope_golden_files() -> List[str]:
    """Rotary-layer fixtures (oracle/gen_golden_rope.py)."""
    return sorted(glob.glob(os.path.join(GOLDEN_DIR, "rope_*.pt")))


def load_golden(path: str) -> dict:
    return torch.load(path, map_location="cpu", weights_only=False)


@dataclass
class Scenario:
    """A paged KV pool + req_to_token table + one forward's worth of inputs (CPU tensors)."""

    Hq: int
    Hkv: int
    D: int
    page_size: int
    prefix_lens: List[int]
    extend_lens: List[int]
    seq_lens: List[int]
    req_to_token: torch.Tensor      # int32 [B+1, ctx]
    req_pool_indices: torch.Tensor  # int64 [B]
    out_cache_loc: torch.Tensor     # int64 [T]
    k_cache: torch.Tensor           # [slots, Hkv, D]
    v_cache: torch.Tensor
    q: torch.Tensor                 # [T, Hq*D]
    k: torch.Tensor                 # [T, Hkv, D]
    v: torch.Tensor
    scaling: float

    def to(self, device):
        out = Scenario(**self.__dict__)
        for name, val in self.__dict__.items():
            if isinstance(val, torch.Tensor):
                setattr(out, name, val.to(device))
        return out


def make_scenario(prefix_lens: Sequence[int], extend_lens: Sequence[int], Hq: int, Hkv: int, D: int,
                  page_size: int, dtype=torch.bfloat16, seed: int = 0, shared_prefix_pages: int = 0,
                  pool_slack: float = 1.25) -> Scenario:
    """Random pages per request (the reference kit's `shuffled_pages` layout,
    dense_attention.py:663-674); slot 0 / page 0 is reserved and never referenced.
    `shared_prefix_pages` > 0 makes the first pages of every request the SAME
    physical pages (a radix-cache hit)."""
    g = torch.Generator().manual_seed(seed)
    B = len(prefix_lens)
    seq_lens = [p + e for p, e in zip(prefix_lens, extend_lens)]
    ctx = max(max(seq_lens), 1)
    ctx = -(-ctx // page_size) * page_size
    pages_per_req = [-(-s // page_size) for s in seq_lens]
    need = sum(pages_per_req) + shared_prefix_pages
    num_pages = int(need * pool_slack) + 2
    perm = (torch.randperm(num_pages - 1, generator=g) + 1).tolist()  # page 0 reserved
    shared = [perm.pop() for _ in range(shared_prefix_pages)]
    req_to_token = torch.zeros((B + 1, ctx), dtype=torch.int32)
    offs = torch.arange(page_size, dtype=torch.int32)
    for b in range(B):
        pages = []
        for j in range(pages_per_req[b]):
            pages.append(shared[j] if j < shared_prefix_pages else perm.pop())
        for j, pg in enumerate(pages):
            lo = j * page_size
            hi = min(lo + page_size, ctx)
            req_to_token[b + 1, lo:hi] = pg * page_size + offs[: hi - lo]
    req_pool_indices = torch.arange(1, B + 1, dtype=torch.int64)
    loc = []
    for b in range(B):
        loc.append(req_to_token[b + 1, prefix_lens[b]:seq_lens[b]].long())
    out_cache_loc = torch.cat(loc) if loc else torch.zeros(0, dtype=torch.int64)
    slots = num_pages * page_size
    k_cache = torch.randn((slots, Hkv, D), generator=g, dtype=torch.float32).to(dtype)
    v_cache = torch.randn((slots, Hkv, D), generator=g, dtype=torch.float32).to(dtype)
    k_cache[:page_size] = 0
    v_cache[:page_size] = 0
    T = sum(extend_lens)
    q = torch.randn((T, Hq * D), generator=g, dtype=torch.float32).to(dtype)
    k = torch.randn((T, Hkv, D), generator=g, dtype=torch.float32).to(dtype)
    v = torch.randn((T, Hkv, D), generator=g, dtype=torch.float32).to(dtype)
    return Scenario(Hq=Hq, Hkv=Hkv, D=D, page_size=page_size, prefix_lens=list(prefix_lens),
                    extend_lens=list(extend_lens), seq_lens=seq_lens, req_to_token=req_to_token,
                    req_pool_indices=req_pool_indices, out_cache_loc=out_cache_loc, k_cache=k_cache,
                    v_cache=v_cache, q=q, k=k, v=v, scaling=D ** -0.5)


def assert_close_attn(actual: torch.Tensor, expected: torch.Tensor, what: str = "",
                      rtol: float = 1e-2, atol: float = 2e-2) -> None:
    """BASELINE tolerance for bf16 attention outputs: 1e-2 relative / 2e-2 absolute."""
    a, e = actual.float().cpu(), expected.float().cpu()
    assert a.shape == e.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(e.shape)}"
    assert torch.isfinite(a).all(), f"{what}: non-finite output"
    err = (a - e).abs()
    tol = atol + rtol * e.abs()
    bad = err > tol
    assert not bad.any(), (f"{what}: {int(bad.sum())}/{bad.numel()} elements out of tolerance, "
                           f"max abs err {float(err.max()):.4e}")


def attention_fp32_blocked(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, req_to_token: torch.Tensor,
                           req_pool_indices, seq_lens: Sequence[int], extend_lens: Sequence[int], Hq: int, Hkv: int,
                           D: int, scaling: float, causal: bool, qblock: int = 1024) -> torch.Tensor:
    """fp32-softmax reference for FULL-SIZE configs on the GPU: the same arithmetic as `oracle.attention_fp32`
    (the kit's `_dense_attention_reference`, dense_attention.py:802-855), evaluated per (request, kv head, block of
    query rows) so that an 8192 x 16384 score matrix never exists at once.  Test infrastructure only."""
    T = q.shape[0]
    G = Hq // Hkv
    out = torch.empty((T, Hq, D), dtype=torch.float32, device=q.device)
    q3 = q.view(T, Hq, D)
    start = 0
    for i, (L, E) in enumerate(zip(seq_lens, extend_lens)):
        L, E = int(L), int(E)
        tokens = req_to_token[int(req_pool_indices[i]), :L].long()
        K, V = k_cache[tokens], v_cache[tokens]          # [L, Hkv, D]
        kpos = torch.arange(L, device=q.device).view(1, 1, L)
        for h in range(Hkv):
            Kh, Vh = K[:, h].float(), V[:, h].float()    # [L, D]
            for r0 in range(0, E, qblock):
                r1 = min(E, r0 + qblock)
                qb = q3[start + r0:start + r1, h * G:(h + 1) * G].float()           # [qb, G, D]
                s = torch.einsum("egd,ld->gel", qb, Kh) * scaling
                if causal:
                    qpos = torch.arange(L - E + r0, L - E + r1, device=q.device).view(1, -1, 1)
                    s = s.masked_fill(kpos > qpos, float("-inf"))
                p = torch.softmax(s, dim=-1)
                out[start + r0:start + r1, h * G:(h + 1) * G] = torch.einsum("gel,ld->egd", p, Vh)
        start += E
    return out.view(T, Hq * D)


def make_scenario_gpu(prefix_lens: Sequence[int], extend_lens: Sequence[int], Hq: int, Hkv: int, D: int,
                      page_size: int, device: str = "cuda", seed: int = 0, shared_prefix_pages: int = 0,
                      pool_slack: float = 1.1) -> Scenario:
    """`make_scenario` for full-size configs: the page permutation is drawn on the CPU (small), the pool and the
    activations are filled on the device (GBs of N(0,1) values)."""
    g = torch.Generator().manual_seed(seed)
    B = len(prefix_lens)
    seq_lens = [p + e for p, e in zip(prefix_lens, extend_lens)]
    ctx = -(-max(seq_lens) // page_size) * page_size
    pages_per_req = [-(-s // page_size) for s in seq_lens]
    need = sum(pages_per_req) - (B - 1) * shared_prefix_pages if shared_prefix_pages else sum(pages_per_req)
    num_pages = int(need * pool_slack) + 2
    perm = torch.randperm(num_pages - 1, generator=g) + 1
    shared = perm[:shared_prefix_pages]
    nxt = shared_prefix_pages
    offs = torch.arange(page_size, dtype=torch.int32)
    table = torch.zeros((B + 1, ctx), dtype=torch.int32)
    for b in range(B):
        own = pages_per_req[b] - shared_prefix_pages
        pages = torch.cat([shared, perm[nxt:nxt + own]]).to(torch.int32)
        nxt += own
        row = (pages[:, None] * page_size + offs[None]).reshape(-1)
        table[b + 1, :row.numel()] = row
    loc = torch.cat([table[b + 1, prefix_lens[b]:seq_lens[b]].long() for b in range(B)])
    gg = torch.Generator(device=device).manual_seed(seed)
    slots = num_pages * page_size
    kc = torch.empty((slots, Hkv, D), dtype=torch.bfloat16, device=device).normal_(generator=gg)
    vc = torch.empty((slots, Hkv, D), dtype=torch.bfloat16, device=device).normal_(generator=gg)
    kc[:page_size] = 0
    vc[:page_size] = 0
    T = sum(extend_lens)
    q = torch.empty((T, Hq * D), dtype=torch.bfloat16, device=device).normal_(generator=gg)
    k = torch.empty((T, Hkv, D), dtype=torch.bfloat16, device=device).normal_(generator=gg)
    v = torch.empty((T, Hkv, D), dtype=torch.bfloat16, device=device).normal_(generator=gg)
    return Scenario(Hq=Hq, Hkv=Hkv, D=D, page_size=page_size, prefix_lens=list(prefix_lens),
                    extend_lens=list(extend_lens), seq_lens=seq_lens, req_to_token=table.to(device),
                    req_pool_indices=torch.arange(1, B + 1, dtype=torch.int64, device=device),
                    out_cache_loc=loc.to(device), k_cache=kc, v_cache=vc, q=q, k=k, v=v, scaling=D ** -0.5)


def assert_close_e4m3_arith(actual: torch.Tensor, emul: torch.Tensor, exact: torch.Tensor, what: str = "") -> None:
    """Check a kernel that computes in the REFERENCE's e4m3 arithmetic (q and p cast to e4m3:
    oracle.fp8_port.extend_attention_e4m3_arith = `emul`) when an exact-math result (`exact`) is also at hand.  Two
    implementations of that arithmetic differ wherever a probability sits at an e4m3 rounding boundary (a 6 % step on
    one term), so the check is statistical, with the size of the arithmetic's own error as the yardstick:
      * mean |actual - emul| <= 0.75 * mean |emul - exact| + 5e-4  (closer to the reference arithmetic than that
        arithmetic is to exact math: measured 0.15 .. 0.65 of it), and
      * max |actual - exact| <= 2 * max |emul - exact| + 2e-2      (no outlier beyond the arithmetic's own)."""
    a, e, x = actual.detach().float().cpu(), emul.float().cpu().view(actual.shape), exact.float().cpu().view(actual.shape)
    assert torch.isfinite(a).all(), f"{what}: non-finite output"
    arith_mean, arith_max = float((e - x).abs().mean()), float((e - x).abs().max())
    got_mean, got_max = float((a - e).abs().mean()), float((a - x).abs().max())
    assert got_mean <= 0.75 * arith_mean + 5e-4, (f"{what}: mean |kernel - e4m3 arithmetic| {got_mean:.3e} vs the "
                                                  f"arithmetic's own mean error {arith_mean:.3e}")
    assert got_max <= 2 * arith_max + 2e-2, f"{what}: max |kernel - exact| {got_max:.3e} vs the arithmetic's {arith_max:.3e}"
