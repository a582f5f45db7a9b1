#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 attention hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one attention-layer call of the decode hot path (KV store of the new
tokens + paged GQA split-KV decode + split merge) on BASELINE.json configs[1]:
Llama-3.1-8B shape, batch 64 x 4096 ctx, GQA 32q/8kv, head_dim 128, page_size 16,
bf16.  The KV working set of one call (1.07 GB) is 8.5x the 126 MB L2, so every
timed iteration streams from HBM ("inputs larger than L2").  `value` = algorithmic
bytes of the call / device time (CUDA events, max over ranks).  The same JSON line
also carries the extend leg (configs[2], TFLOP/s on tcgen05), the roofline of the
dominant kernel, the end-to-end number through the AttentionBackend API with host
buffers, and the CPU baseline (oracle port of torch_native on the host cores).

For --gpus N > 1 (torchrun, one rank per GPU) the batch is head-sharded the way
tensor parallelism does it: every rank owns the same number of KV heads (8) of a
model N times as wide and runs the same kernels on its own pool; no collective is
on the data path (o_proj's all-reduce stays the reference's), scaling is weak.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# ---- workload (BASELINE.json configs[1] and configs[2]) ------------------------
CFG2 = dict(B=64, ctx=4096, Hq=32, Hkv=8, D=128, page=16, layers=32)
CFG3 = dict(B=32, prefix=2048, ext=512, Hq=32, Hkv=8, D=128, page=16)
SEED = 20261104
WORKLOAD = ("llama-3.1-8b decode attention layer-call (KV store + paged GQA split-KV decode + merge): "
            "batch 64 x 4096 ctx, GQA 32q/8kv, head_dim 128, page_size 16 (BASELINE configs[1])")


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d["bf16_tflops_sustained"],
                    src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src="fallback")


def ncu_traffic(key):
    """DRAM bytes per launch of the named kernel from the committed ncu --set full capture (profiles/), or None."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(path) as f:
            return json.load(f)[key]["bytes"]
    except (OSError, KeyError, ValueError):
        return None


def decode_bytes(c, B=None):
    """Algorithmic bytes of one decode layer-call: K+V rows read once,
    q read, o written, int32 slot ids read once per kv head group... counted once, new-token store."""
    B = c["B"] if B is None else B
    kv = B * c["ctx"] * c["Hkv"] * c["D"] * 2 * 2
    qo = 2 * B * c["Hq"] * c["D"] * 2
    idx = B * c["ctx"] * 4
    store = 2 * 2 * B * c["Hkv"] * c["D"] * 2
    return kv + qo + idx + store


def extend_flops(c):
    pairs = c["B"] * (c["ext"] * c["prefix"] + c["ext"] * (c["ext"] + 1) // 2)
    return 4 * c["D"] * c["Hq"] * pairs


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            t0 = time.time()
            while not self.rows and time.time() - t0 < 3.0:  # first sample before the timed region starts
                time.sleep(0.02)
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower() == "active"})
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=reasons, samples=len(sm))


def sampled_check(q, o, kc, vc, table, req_rows, seq_lens, ext_lens, Hq, Hkv, D, causal, k_mul=1.0, v_mul=1.0,
                  max_reqs=3, max_rows=48, atol=2e-2, rtol=1e-2):
    """Checker for the side legs (never timed): a torch fp32-softmax reference on the GPU for a SAMPLE of the batch
    (first / middle / last request, the first and last `max_rows / 2` query rows of each), compared with the kernel
    output at BASELINE's tolerance (1e-2 relative / 2e-2 absolute; the e4m3-arithmetic extend leg passes the bound
    it is held to against the reference Triton kernel's output, 5e-2).  Returns dict(max_abs_err, within_tolerance, rows)."""
    G = Hq // Hkv
    B = len(seq_lens)
    picks = sorted({0, B // 2, B - 1})[:max_reqs]
    starts = [0]
    for e in ext_lens:
        starts.append(starts[-1] + int(e))
    worst, ok, nrows = 0.0, True, 0
    for b in picks:
        L, E = int(seq_lens[b]), int(ext_lens[b])
        rows = sorted(set(list(range(min(E, max_rows // 2))) + list(range(max(0, E - max_rows // 2), E))))
        tok = table[int(req_rows[b]), :L].long()
        K = kc[tok].float().mul_(k_mul).repeat_interleave(G, dim=1)   # [L, Hq, D]
        V = vc[tok].float().mul_(v_mul).repeat_interleave(G, dim=1)
        r = torch.tensor(rows, device=q.device)
        qq = q.view(-1, Hq, D)[starts[b] + r].float()
        sc = torch.einsum("ehd,lhd->hel", qq, K) * (D ** -0.5)
        if causal:
            qpos = (L - E + r).view(1, -1, 1)
            sc = sc.masked_fill(torch.arange(L, device=q.device).view(1, 1, L) > qpos, float("-inf"))
        ref = torch.einsum("hel,lhd->ehd", torch.softmax(sc, dim=-1), V)
        got = o.view(-1, Hq, D)[starts[b] + r].float()
        err = (got - ref).abs()
        worst = max(worst, float(err.max()))
        ok = ok and bool((err <= atol + rtol * ref.abs()).all()) and bool(torch.isfinite(got).all())
        nrows += len(rows)
    return dict(max_abs_err=worst, within_tolerance=ok, rows_checked=nrows, requests_checked=picks,
                tolerance=dict(atol=atol, rtol=rtol))


# ---- synthetic state ------------------------------------------------------------
def build_decode_state(c, device, gen):
    """Pool, req_to_token (random pages per request, page 0 reserved) and one step's inputs."""
    from sglang_b200 import MHATokenToKVPool, ReqToTokenPool

    B, ctx, page = c["B"], c["ctx"], c["page"]
    pages_per_req = ctx // page
    num_pages = int(B * pages_per_req * 1.25) + 1
    req_pool = ReqToTokenPool(B, ctx, device)
    perm = (torch.randperm(num_pages - 1, generator=gen) + 1)[: B * pages_per_req].view(B, pages_per_req)
    table = (perm[:, :, None] * page + torch.arange(page)[None, None, :]).view(B, ctx).to(torch.int32)
    req_pool.req_to_token[1:] = table.to(device)
    kv_pool = MHATokenToKVPool(num_pages * page - page, page, torch.bfloat16, c["Hkv"], c["D"], 1, device)
    for buf in (kv_pool.k_buffer[0], kv_pool.v_buffer[0]):
        buf.normal_(0, 1, generator=None)
        buf[:page].zero_()
    host = dict(
        q=torch.randn((B, c["Hq"] * c["D"]), generator=gen).to(torch.bfloat16).pin_memory(),
        k=torch.randn((B, c["Hkv"] * c["D"]), generator=gen).to(torch.bfloat16).pin_memory(),
        v=torch.randn((B, c["Hkv"] * c["D"]), generator=gen).to(torch.bfloat16).pin_memory(),
        seq_lens=torch.full((B,), ctx, dtype=torch.int64).pin_memory(),
        req_pool_indices=torch.arange(1, B + 1, dtype=torch.int64).pin_memory(),
        out_cache_loc=table[:, ctx - 1].to(torch.int64).pin_memory(),
    )
    return req_pool, kv_pool, host


class Runner:
    def __init__(self, device, req_pool, kv_pool, page):
        self.device, self.page_size = device, page
        self.req_to_token_pool, self.token_to_kv_pool = req_pool, kv_pool
        self.sliding_window_size = None


SIDE_STEPS, SIDE_WARMUP = 20, 5  # every side leg: >= 20 timed launches after >= 5 warm-ups


def max_over_ranks(ms, dev):
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return ms
    t = torch.tensor([ms], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def time_steps(fn, steps, warmup, barrier=None):
    """W warm-up calls, then K calls bracketed by (barrier +) synchronize; CUDA events on the
    launching (current) stream.  Returns total milliseconds on this rank."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if barrier is not None:
        barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if barrier is not None:
        barrier()
    return e0.elapsed_time(e1)


def cpu_decode_sample(c, sample_B, reps, threads):
    """The oracle port of torch_native forward_decode on the host cores over `sample_B`
    requests of the cfg2 workload (full 4096-key contexts).  Returns (seconds per call, bytes)."""
    from oracle import torch_native_port as oracle

    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(SEED)
    ctx, page, Hkv, D, Hq = c["ctx"], c["page"], c["Hkv"], c["D"], c["Hq"]
    pages = ctx // page
    num_pages = sample_B * pages + 1
    perm = (torch.randperm(num_pages - 1, generator=g) + 1).view(sample_B, pages)
    table = torch.zeros((sample_B + 1, ctx), dtype=torch.int32)
    table[1:] = (perm[:, :, None] * page + torch.arange(page)[None, None, :]).view(sample_B, ctx).to(torch.int32)
    kc = torch.randn((num_pages * page, Hkv, D), generator=g).to(torch.bfloat16)
    vc = torch.randn((num_pages * page, Hkv, D), generator=g).to(torch.bfloat16)
    q = torch.randn((sample_B, Hq * D), generator=g).to(torch.bfloat16)
    k = torch.randn((sample_B, Hkv, D), generator=g).to(torch.bfloat16)
    v = torch.randn((sample_B, Hkv, D), generator=g).to(torch.bfloat16)
    req = torch.arange(1, sample_B + 1)
    seq = torch.full((sample_B,), ctx)
    loc = table[1:, ctx - 1].long()

    def call():
        oracle.store_kv(kc, vc, loc, k, v)
        return oracle.decode(q, kc, vc, table, req, seq, Hq, Hkv, D, D ** -0.5)

    call()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        call()
        ts.append(time.perf_counter() - t0)
    return statistics.median(ts), decode_bytes(c, sample_B)


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (torch_native; the oracle
    port, since /root/reference does not exist on the GPU box) on all host threads."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample_B = 64
    for _ in range(max(args.warmup - 1, 0)):
        cpu_decode_sample(CFG2, sample_B, 1, threads)
    sec, nbytes = cpu_decode_sample(CFG2, sample_B, max(args.steps, 1), threads)
    gbs = nbytes / sec / 1e9
    line = dict(
        impl="reference", metric="decode_attn_hbm_gbs", value=gbs, unit="GB/s", n_gpus=args.gpus, steps=args.steps,
        warmup=args.warmup, ms_per_step=sec * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None,
        dtype="bf16", data="synthetic",
        config=dict(workload=WORKLOAD),
        cpu_baseline=dict(value=gbs, unit="GB/s", cores=threads, kind="port",
                          sample=f"{sample_B} of 64 requests x 4096 ctx per call, median of {max(args.steps, 1)}"),
        e2e=dict(value=gbs, unit="GB/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
    )
    print(json.dumps(line), flush=True)


def bind_to_gpu_numa_node(dev):
    """Run this process (and so its pinned host buffers and copy submissions) on the CPUs local to the GPU's PCIe
    root, as SGLang's launcher does with its numa binding; returns (original affinity, description)."""
    try:
        pr = torch.cuda.get_device_properties(dev)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as f:
            txt = f.read().strip()
        cpus = set()
        for part in txt.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        old = os.sched_getaffinity(0)
        cpus &= old
        if cpus and cpus != old:
            os.sched_setaffinity(0, cpus)
            return old, f"bound to the GPU-local cpus {txt} ({bdf})"
        return old, f"already on the GPU-local cpus {txt} ({bdf})"
    except Exception as e:  # no sysfs / no permission: run unbound
        return None, f"unbound ({type(e).__name__})"


def run_b200(args, rank, world, local_rank):
    from sglang_b200 import B200AttnBackend, ForwardBatch, ForwardMode, RadixAttention, ops

    dev = f"cuda:{local_rank}"
    torch.cuda.set_device(dev)
    old_affinity, numa_note = bind_to_gpu_numa_node(dev)
    barrier = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device(dev))
        barrier = lambda: dist.barrier()
    peaks = measured_peaks()
    gen = torch.Generator().manual_seed(SEED + rank)
    c = CFG2
    B, Hq, Hkv, D = c["B"], c["Hq"], c["Hkv"], c["D"]
    req_pool, kv_pool, host = build_decode_state(c, dev, gen)
    backend = B200AttnBackend(Runner(dev, req_pool, kv_pool, c["page"]))
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    devt = {k: v.to(dev) for k, v in host.items()}
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=B, req_pool_indices=devt["req_pool_indices"],
                      seq_lens=devt["seq_lens"], out_cache_loc=devt["out_cache_loc"], seq_lens_sum=B * c["ctx"],
                      seq_lens_cpu=host["seq_lens"], attn_backend=backend)
    backend.init_forward_metadata(fb)
    md = backend.forward_metadata
    kbuf, vbuf = kv_pool.get_kv_buffer(0)
    warps, max_splits, tail_begin = ops.plan_decode(B, Hkv, Hq // Hkv, c["ctx"], md.min_chunk)
    launches_per_step = 2  # kv_store_kernel + decode_kernel (split partials are merged inside the decode kernel)

    # (1) the step: store + decode attention (+ in-kernel merge), inputs resident in HBM.  The layer-call is captured
    #     ONCE into a CUDA graph and replayed per step - the way SGLang's decode runner drives a backend
    #     (decode_cuda_graph_runner.py:1148-1363) - so the host launch gap between the store and the decode kernel
    #     (two eager ctypes launches from Python) is not part of the number.  `eager_ms_per_step` keeps the eager loop.
    def step():
        return layer(devt["q"], devt["k"], devt["v"], fb)

    o_eager = step().clone()
    side0 = torch.cuda.Stream()
    side0.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side0):
        step()
        torch.cuda.synchronize()
        step_graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(step_graph, stream=side0):
            o_graph = step()
    torch.cuda.current_stream().wait_stream(side0)
    step_graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(o_graph, o_eager), "captured layer-call disagrees with the eager one"
    head_check = sampled_check(devt["q"], o_graph, kbuf, vbuf, req_pool.req_to_token, host["req_pool_indices"],
                               [c["ctx"]] * B, [1] * B, Hq, Hkv, D, causal=False)
    assert head_check["within_tolerance"], f"cfg2 decode output out of tolerance: {head_check}"

    clk = ClockSampler(local_rank).__enter__()
    eager_ms = time_steps(step, args.steps, args.warmup, barrier) / args.steps
    total_ms = time_steps(step_graph.replay, args.steps, args.warmup, barrier)
    if world > 1:
        t = torch.tensor([total_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t)
    ms_step = total_ms / args.steps
    nbytes = decode_bytes(c)
    value = world * nbytes / (ms_step * 1e-3) / 1e9

    # (2) the dominant kernel alone (decode attention, without the store) for the roofline
    q3 = devt["q"].view(B, Hq, D)
    o3 = torch.empty_like(q3)
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, max_splits), 16), dtype=torch.uint8, device=dev)

    def kern():
        ops.decode_attention(q3, kbuf, vbuf, o3, md.kv_indptr, md.kv_indices, layer.scaling, max_splits=max_splits,
                             min_chunk=md.min_chunk, workspace=ws, warps=warps, tail_begin=tail_begin)

    kern_ms = time_steps(kern, args.steps, args.warmup) / args.steps
    kern_bytes = nbytes - 2 * 2 * B * Hkv * D * 2  # without the store's read+write
    achieved = kern_bytes / (kern_ms * 1e-3) / 1e9

    # (2b) the same kernel on the same pool with other PAGE PLACEMENTS (rank 0, N = 1; side key, not the headline):
    #      the headline uses a random permutation of the pool's pages - the worst case.  "contiguous" = every
    #      request's pages in one ascending run (what alloc_extend hands a fresh prefill); "interleaved" = page j of
    #      request b at b + j * B (what 64 requests decoding in step get from alloc_decode, one page each per 16 steps).
    placement = None
    if world == 1:
        placement = {"random_pages (headline)": dict(kernel_ms=kern_ms, gbs=achieved)}
        ppr = c["ctx"] // c["page"]
        for name, pages in (("contiguous_pages", 1 + torch.arange(B)[:, None] * ppr + torch.arange(ppr)[None, :]),
                            ("interleaved_pages", 1 + torch.arange(B)[:, None] + torch.arange(ppr)[None, :] * B)):
            idx2 = (pages[:, :, None] * c["page"] + torch.arange(c["page"])[None, None, :]).reshape(-1).to(torch.int32).to(dev)
            t2 = time_steps(lambda: ops.decode_attention(q3, kbuf, vbuf, o3, md.kv_indptr, idx2, layer.scaling,
                                                         max_splits=max_splits, min_chunk=md.min_chunk, workspace=ws,
                                                         warps=warps, tail_begin=tail_begin),
                            args.steps, args.warmup) / args.steps
            placement[name] = dict(kernel_ms=t2, gbs=kern_bytes / (t2 * 1e-3) / 1e9)

    # (3) end to end through the AttentionBackend API with HOST buffers (pinned): per step the new
    #     q/k/v rows and the batch metadata go host->device, metadata is rebuilt, the layer runs,
    #     and the attention output comes back device->host.
    o_host = torch.empty((B, Hq * D), dtype=torch.bfloat16).pin_memory()
    h2d = sum(t.numel() * t.element_size() for t in host.values())
    d2h = o_host.numel() * o_host.element_size()

    def e2e_step():
        for k_, t in host.items():
            devt[k_].copy_(t, non_blocking=True)
        backend.init_forward_metadata(fb)
        o_host.copy_(layer(devt["q"], devt["k"], devt["v"], fb), non_blocking=True)

    E2E_REPS = 5

    def median_of(fn, reps=E2E_REPS):
        """K timed steps (max over ranks), repeated `reps` times; returns (median, min, max) ms per step.  The MEDIAN
        repetition is what is reported."""
        out = []
        for _ in range(reps):
            ms = time_steps(fn, args.steps, args.warmup, barrier)
            if world > 1:
                t = torch.tensor([ms], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t)
            out.append(ms / args.steps)
        return statistics.median(out), min(out), max(out)

    e2e_ms, e2e_lo, e2e_hi = median_of(e2e_step)
    e2e_val = world * nbytes / (e2e_ms * 1e-3) / 1e9

    # (3b) the same step the way SGLang's decode runner drives a backend (decode_cuda_graph_runner.py:1148-1363):
    #      static buffers, ONE captured graph holding the pinned-host H2D copies, the sync-free metadata refill
    #      (init_forward_metadata_out_graph), store + decode, and the D2H read; one replay per step.
    backend.init_cuda_graph_state(B, B, num_q_heads=Hq, head_dim=D)
    fbg = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=B, req_pool_indices=devt["req_pool_indices"],
                       seq_lens=devt["seq_lens"], out_cache_loc=devt["out_cache_loc"], seq_lens_sum=B * c["ctx"],
                       attn_backend=backend)

    aux = torch.cuda.Stream()
    fork = os.environ.get("B200ATTN_BENCH_FORK", "0") == "1"  # measured: no gain (234 vs 235 us), copy latency dominates

    # Host layout of a step's inputs.  "fused" is the layout the model hands a backend: ONE [B, (Hq+2Hkv)*D] qkv
    # activation (q/k/v are column views with the fused token stride) and one int64 [3, B] block of
    # index vectors -> two H2D copies per step.  "separate" keeps six tensors -> six copies.  Same bytes either way.
    W = (Hq + 2 * Hkv) * D
    host_qkv = torch.cat([host["q"], host["k"], host["v"]], dim=1).pin_memory()
    host_idx = torch.stack([host["seq_lens"], host["req_pool_indices"], host["out_cache_loc"]]).pin_memory()
    dev_qkv = torch.empty((B, W), dtype=torch.bfloat16, device=dev)
    dev_idx = torch.empty((3, B), dtype=torch.int64, device=dev)
    assert host_qkv.numel() * 2 + host_idx.numel() * 8 == h2d
    fbp = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=B, req_pool_indices=dev_idx[1], seq_lens=dev_idx[0],
                       out_cache_loc=dev_idx[2], seq_lens_sum=B * c["ctx"], attn_backend=backend)

    def graph_body(fused):
        cur = torch.cuda.current_stream()
        if fused == "kernel_staging":
            # host <-> device moves by the library's staging kernel (pinned memory is mapped): kernel nodes only
            ops.stage_copy(dev_idx, host_idx, dev_qkv, host_qkv)
            backend.init_forward_metadata_out_graph(fbp)
            o = layer(dev_qkv[:, :Hq * D], dev_qkv[:, Hq * D:(Hq + Hkv) * D], dev_qkv[:, (Hq + Hkv) * D:], fbp)
            ops.stage_copy(o_host, o)
            return
        if fused:
            dev_idx.copy_(host_idx, non_blocking=True)
            backend.init_forward_metadata_out_graph(fbp)
            dev_qkv.copy_(host_qkv, non_blocking=True)
            o = layer(dev_qkv[:, :Hq * D], dev_qkv[:, Hq * D:(Hq + Hkv) * D], dev_qkv[:, (Hq + Hkv) * D:], fbp)
            o_host.copy_(o, non_blocking=True)
            return
        for k_ in ("seq_lens", "req_pool_indices", "out_cache_loc"):
            devt[k_].copy_(host[k_], non_blocking=True)
        if fork:  # the metadata refill only needs the index vectors: a graph branch beside the q/k/v copies
            aux.wait_stream(cur)
            with torch.cuda.stream(aux):
                backend.init_forward_metadata_out_graph(fbg)
        else:
            backend.init_forward_metadata_out_graph(fbg)
        for k_ in ("q", "k", "v"):
            devt[k_].copy_(host[k_], non_blocking=True)
        if fork:
            cur.wait_stream(aux)
        o_host.copy_(layer(devt["q"], devt["k"], devt["v"], fbg), non_blocking=True)

    def time_graph(fused):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph_body(fused)
            torch.cuda.synchronize()
            g_ = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_, stream=side):
                graph_body(fused)
        torch.cuda.current_stream().wait_stream(side)
        return median_of(g_.replay), g_

    (e2e_sep_ms, sep_lo, sep_hi), graph = time_graph(False)
    o_sep = o_host.clone()
    # HEADLINE e2e driver, fixed a priori: ONE captured graph per step over the model's fused-qkv layout (the path
    # SGLang's decode runner takes), ONE capture, median of E2E_REPS repetitions.  The other two drivers (six-tensor
    # graph, eager loop) are side keys.  Round 1 saw this graph come out bimodal ACROSS captures (225 vs 275-290 us);
    # that is reported, not selected away: two further captures are timed and listed in `capture_spread_ms` (N = 1).
    (e2e_memcpy_ms, memcpy_lo, memcpy_hi), graph2 = time_graph(True)
    torch.cuda.synchronize()
    assert torch.equal(o_sep, o_host), "fused-layout e2e step disagrees with the separate-tensor step"
    # Round 2: the memcpy-node graph stayed bimodal ACROSS PROCESSES (225 us in one, 275 us in the next, same box,
    # whatever the kernels did): every memcpy <-> kernel edge is a copy-engine / SM hand-over and the engine a node
    # lands on is decided at instantiation.  The headline driver moves the same bytes between the same pinned host
    # tensors and the same static device buffers with the library's staging kernel (b200attn_stage_copy: pinned
    # memory is mapped, the kernel reads / writes it over PCIe), so the graph is kernel nodes only.
    o_host.zero_()
    (e2e_graph_ms, fused_lo, fused_hi), graph3 = time_graph("kernel_staging")
    torch.cuda.synchronize()
    assert torch.equal(o_sep, o_host), "kernel-staging e2e step disagrees with the separate-tensor step"
    capture_spread = [e2e_graph_ms]
    if world == 1:
        for _ in range(2):
            (ms_, _, _), g_ = time_graph("kernel_staging")
            capture_spread.append(ms_)
            del g_
    e2e_sep_val = world * nbytes / (e2e_sep_ms * 1e-3) / 1e9
    e2e_memcpy_val = world * nbytes / (e2e_memcpy_ms * 1e-3) / 1e9
    fused_ms = e2e_graph_ms
    e2e_driver = "graph_fused_layout_kernel_staging"
    e2e_graph_val = world * nbytes / (e2e_graph_ms * 1e-3) / 1e9

    # (4) extend leg (configs[2]): 32 requests sharing a 2048-token prefix + 512 new tokens, causal
    ext = run_extend_leg(args, dev, gen, peaks)
    extra = {}
    if not args.headline_only:
        del req_pool, kv_pool, backend, fb, fbg, fbp, graph, graph2, graph3, md, kbuf, vbuf
        torch.cuda.empty_cache()
        # configs[3]: Qwen2.5-VL-7B LLM side, 16 x (1024 image + 256 text) extend rows + 16 decode rows, MIXED
        for causal in (True, False):
            extra["cfg4_mixed_" + ("causal" if causal else "bidirectional")] = run_backend_case(
                dev, gen, peaks, "qwen2.5-vl-7b MIXED: 16 req x 1280 new tokens + 16 decode rows (prefix 1279), "
                "Hq 28 / Hkv 4 (BASELINE configs[3])", [0] * 16 + [1279] * 16, [1280] * 16 + [1] * 16, 28, 4, 128,
                16, causal, "MIXED")
        # configs[0]: the reference's own CPU-runnable case, timed on the host cores beside the GPU path
        if rank == 0 and not args.no_cpu_baseline:
            if old_affinity is not None:
                os.sched_setaffinity(0, old_affinity)  # the CPU legs use every host core
                old_affinity = None
            extra["cfg1_qwen2_0p5b_decode"] = run_cfg1(dev, gen)
        # configs[1] again over a float8_e4m3fn pool: half the KV bytes per call
        extra["cfg2_decode_fp8kv"] = run_cfg2_decode_fp8(dev, gen, peaks)
        torch.cuda.empty_cache()
        # configs[2] over the e4m3 pool: quantising store + extend on the pool's e4m3 bytes per layer-call
        extra["cfg3_extend_fp8kv"] = run_backend_case(
            dev, gen, peaks, "llama-3.1-8b extend over a float8_e4m3fn KV pool: 32 req x (2048 shared prefix + 512 new), "
            "causal; layer-call = quantising store + extend contracting on the pool's e4m3 bytes (kind::f8f6f4)",
            [2048] * 32, [512] * 32, 32, 8, 128, 16, True, "EXTEND", shared_prefix=2048, fp8_kv=True)
        torch.cuda.empty_cache()
        # configs[4]: Llama-3.1-70B attention shape, this rank's KV-head shard
        # STRONG scaling (the partitioning north_star names): the SAME total work - 8 kv heads / 64 q heads - cut into
        # Hkv = 8 / N per rank; every leg barrier-to-barrier, max over ranks.  t1 (the unsharded shape) is measured
        # in the same run on every rank, so efficiency = t1 / (N * tN) needs no second invocation.
        extra["cfg5_decode"] = run_cfg5_decode(dev, gen, peaks, world, barrier=barrier)
        torch.cuda.empty_cache()
        extra["cfg5_chunked_prefill"] = run_backend_case(
            dev, gen, peaks, f"llama-3.1-70b chunked prefill shard: prefix 8192 + 8192 new, causal, Hq {64 // world} "
            f"/ Hkv {max(1, 8 // world)} per GPU of {world} (BASELINE configs[4])", [8192], [8192], 64 // world,
            max(1, 8 // world), 128, 16, True, "EXTEND", barrier=barrier)
        torch.cuda.empty_cache()
        if world > 1:
            full_dec = run_cfg5_decode(dev, gen, peaks, 1, barrier=barrier)
            torch.cuda.empty_cache()
            full_pre = run_backend_case(dev, gen, peaks, "llama-3.1-70b chunked prefill, unsharded (Hq 64 / Hkv 8)",
                                        [8192], [8192], 64, 8, 128, 16, True, "EXTEND", barrier=barrier)
            torch.cuda.empty_cache()
        else:
            full_dec, full_pre = extra["cfg5_decode"], extra["cfg5_chunked_prefill"]
        extra["scaling_strong"] = dict(
            workload="llama-3.1-70b attention (BASELINE configs[4]) head-sharded over N GPUs: decode 256 x 8192 ctx + "
                     "8192-token chunked prefill over an 8192-token prefix; total work fixed (8 kv heads)",
            n_gpus=world, timing="barrier-to-barrier, max over ranks; t1 = the unsharded shape timed in this same run",
            decode=dict(t1_ms=full_dec["kernel_ms"], tN_ms=extra["cfg5_decode"]["kernel_ms"],
                        efficiency=full_dec["kernel_ms"] / (world * extra["cfg5_decode"]["kernel_ms"]),
                        aggregate_gbs=world * extra["cfg5_decode"]["gbs"]),
            chunked_prefill=dict(t1_ms=full_pre["ms_per_step"], tN_ms=extra["cfg5_chunked_prefill"]["ms_per_step"],
                                 efficiency=full_pre["ms_per_step"] / (world * extra["cfg5_chunked_prefill"]["ms_per_step"]),
                                 aggregate_tflops=world * extra["cfg5_chunked_prefill"]["tflops"]))
    clk.__exit__()
    clocks = clk.summary()

    if rank != 0:
        return
    # (5) CPU baseline on the host cores (rank 0, N=1 only): bounded sample of the same workload
    cpu = None
    if old_affinity is not None:
        os.sched_setaffinity(0, old_affinity)
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        sec, sb = cpu_decode_sample(c, 64, 3, threads)
        cpu = dict(value=sb / sec / 1e9, unit="GB/s", cores=threads, kind="port",
                   sample="all 64 requests x 4096 ctx per call (oracle port of torch_native forward_decode), median of 3")
    line = dict(
        metric="decode_attn_hbm_gbs", value=value, unit="GB/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
        ms_per_step=ms_step, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16", data="synthetic",
        config=dict(workload=WORKLOAD,
                    parallelism=f"kv-head sharded x{world} (8 kv heads per GPU), no collective",
                    l2="inputs larger than L2 (1.07 GB KV per call vs 126 MB)", max_splits=max_splits,
                    decode_warps_per_cta=warps,
                    step_driver="one CUDA graph replay per step (store + decode captured once)"),
        eager_ms_per_step=eager_ms, output_check=head_check,
        roofline=dict(bound="hbm", achieved=achieved, peak=peaks["hbm"], unit="GB/s", frac=achieved / peaks["hbm"],
                      traffic=ncu_traffic("decode_kernel<128,bf16> cfg2") if world == 1 else None,
                      kernel="decode_kernel<128,bf16> (+merge)", kernel_ms=kern_ms,
                      peak_source=peaks["src"], frac_of_8tbs=achieved / 8000.0),
        decode_page_placement=placement,
        cpu_baseline=cpu,
        e2e=dict(value=e2e_graph_val, unit="GB/s", h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h,
                 ms_per_step=e2e_graph_ms, ms_min=fused_lo, ms_max=fused_hi, repetitions=E2E_REPS,
                 host_placement=numa_note, driver=e2e_driver, capture_spread_ms=capture_spread,
                 mode="B200AttnBackend behind one CUDA graph per step (SGLang's decode-runner path): the step's inputs "
                      "are read from PINNED HOST tensors in the model's layout (one fused [B, (Hq+2Hkv)*D] qkv activation "
                      "+ one int64 [3, B] index block) into the runner's static device buffers and the result is "
                      "written to a pinned host tensor by the library's staging kernel (b200attn_stage_copy: mapped "
                      "pinned memory, the bytes cross PCIe inside the timed graph), then metadata refill, store + "
                      "decode; kernel nodes only.  ONE capture, value = MEDIAN of 5 repetitions of K steps (min / max "
                      "beside it); `capture_spread_ms` = the medians of this capture and two more captures",
                 memcpy_nodes=dict(value=e2e_memcpy_val, ms_per_step=e2e_memcpy_ms, ms_min=memcpy_lo, ms_max=memcpy_hi,
                                   mode="same graph with the three host <-> device moves as cudaMemcpyAsync nodes "
                                        "(copy engines); bimodal across processes, 225 vs 275 us"),
                 separate_tensors=dict(value=e2e_sep_val, ms_per_step=e2e_sep_ms, ms_min=sep_lo, ms_max=sep_hi,
                                       mode="same graph with q, k, v and the three index vectors as six host "
                                            "tensors = 6 H2D copies"),
                 eager=dict(value=e2e_val, ms_per_step=e2e_ms, ms_min=e2e_lo, ms_max=e2e_hi,
                            mode="same step launched eagerly from Python (CPU-launch-bound)")),
        gpu_launches=launches_per_step * args.steps,
        clocks=clocks,
        extend=ext,
        other_configs=extra,
        # weak scaling of a head-sharded model N times as wide: every rank serves the SAME B tokens
        tokens_per_s_attention_only=B / (c["layers"] * ms_step * 1e-3),
    )
    print(json.dumps(line), flush=True)


def run_extend_leg(args, dev, gen, peaks):
    from sglang_b200 import ops

    c = CFG3
    B, P, E, Hq, Hkv, D, page = c["B"], c["prefix"], c["ext"], c["Hq"], c["Hkv"], c["D"], c["page"]
    # The batch's KV layout is produced the way production produces it: the shared prefix is a
    # radix-cache HIT - one earlier request left its 2048 tokens in the (device-side) radix cache, every request of
    # this batch matches them and so maps its first 128 logical pages to the SAME physical pages - and the new
    # tokens' slots come from the paged allocator's alloc_extend (fragmented free list: shuffled page order).
    from sglang_b200.allocator import PagedTokenToKVPoolAllocator
    from sglang_b200.radix import DeviceRadixCache

    shared = P // page
    own = E // page
    num_pages = shared + B * own + 8
    alloc = PagedTokenToKVPoolAllocator(num_pages * page, page, dev)
    alloc.free_pages = alloc.free_pages[torch.randperm(alloc.free_pages.numel(), generator=gen).to(dev)]
    radix = DeviceRadixCache(page, max_cached_pages=num_pages, num_kv_pages=num_pages + 1, device=dev)
    prefix_tokens = torch.randint(0, 128000, (P,), generator=gen)
    prefix_slots = alloc.alloc(P)
    assert radix.insert([prefix_tokens], [prefix_slots]) == [0]
    keys = [torch.cat([prefix_tokens, torch.randint(0, 128000, (E,), generator=gen)]) for _ in range(B)]
    hits = radix.match_prefix(keys)
    assert all(h.numel() == P and torch.equal(h, prefix_slots) for h in hits), "radix match must return the shared prefix"
    pre_cpu, seq_cpu = torch.full((B,), P, dtype=torch.int64), torch.full((B,), P + E, dtype=torch.int64)
    last_loc = torch.stack([h[-1] for h in hits])
    new_slots = alloc.alloc_extend(pre_cpu.to(dev), pre_cpu, seq_cpu.to(dev), seq_cpu, last_loc, B * E)
    assert new_slots is not None and int(alloc.status) == 0
    table = torch.zeros((B + 1, P + E), dtype=torch.int32, device=dev)
    table[1:, :P] = torch.stack(hits).to(torch.int32)
    table[1:, P:] = new_slots.view(B, E).to(torch.int32)
    num_pages += 1  # slots of the pool: page 0 is the padding page
    kc = torch.randn((num_pages * page, Hkv, D), device=dev).to(torch.bfloat16)
    vc = torch.randn((num_pages * page, Hkv, D), device=dev).to(torch.bfloat16)
    T = B * E
    q = torch.randn((T, Hq, D), device=dev).to(torch.bfloat16)
    k = torch.randn((T, Hkv, D), device=dev).to(torch.bfloat16)
    v = torch.randn((T, Hkv, D), device=dev).to(torch.bfloat16)
    o = torch.empty_like(q)
    loc = new_slots
    req = torch.arange(1, B + 1, device=dev)
    seq = torch.full((B,), P + E, device=dev)
    extl = torch.full((B,), E, dtype=torch.int32, device=dev)
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    qo_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    kv_indices = torch.zeros(B * (P + E), dtype=torch.int32, device=dev)
    ops.build_kv_indices(table, req, seq, kv_indptr, kv_indices, qo_lens=extl, qo_indptr=qo_indptr)

    def step():
        ops.store_cache(k, v, kc, vc, loc, reserved_skip_index=0)
        ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True,
                             page_size=page, max_qo_len=E)

    def kern():
        ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, D ** -0.5, causal=True,
                             page_size=page, max_qo_len=E)

    steps = max(args.steps, SIDE_STEPS)
    ms = time_steps(step, steps, max(args.warmup, SIDE_WARMUP)) / steps
    kms = time_steps(kern, steps, max(args.warmup, SIDE_WARMUP)) / steps
    fl = extend_flops(c)
    tf = fl / (kms * 1e-3) / 1e12
    check = sampled_check(q, o, kc, vc, table, list(range(1, B + 1)), [P + E] * B, [E] * B, Hq, Hkv, D, causal=True)
    assert check["within_tolerance"], f"cfg3 extend output out of tolerance: {check}"
    return dict(workload="llama-3.1-8b extend: 32 req x (2048 shared prefix + 512 new), causal (BASELINE configs[2])",
                ms_per_step=ms, kernel_ms=kms, tflops=tf, flops=fl, tokens_per_s_attention_only=T / (32 * ms * 1e-3),
                timed_launches=steps, output_check=check,
                roofline=dict(bound="tensor", achieved=tf, peak=peaks["tf_burst"], unit="TFLOP/s",
                              frac=tf / peaks["tf_burst"], frac_of_2250=tf / 2250.0, peak_source=peaks["src"],
                              traffic=ncu_traffic("extend_kernel<128,bf16> cfg3")))


def run_backend_case(dev, gen, peaks, label, prefix_lens, ext_lens, Hq, Hkv, D, page, causal, mode_name,
                     shared_prefix=0, steps=SIDE_STEPS, warmup=SIDE_WARMUP, fp8_kv=False, barrier=None):
    """One ragged extend / MIXED layer-call through B200AttnBackend (store + attention), inputs resident in
    HBM.  `shared_prefix` leading tokens of every request map to the SAME physical pages (radix hit).
    FLOPs are algorithmic: 4*D per (q head, query, visible key); rows with extend_len 1 count as bytes."""
    from sglang_b200 import (AttentionType, B200AttnBackend, ForwardBatch, ForwardMode, MHATokenToKVPool,
                             RadixAttention, ReqToTokenPool)

    B = len(prefix_lens)
    seq = [p_ + e for p_, e in zip(prefix_lens, ext_lens)]
    ctx = -(-max(seq) // page) * page
    own_pages = [-(-(s - shared_prefix) // page) for s in seq]
    num_pages = shared_prefix // page + sum(own_pages) + 8
    perm = (torch.randperm(num_pages - 1, generator=gen) + 1).tolist()
    shared = [perm.pop() for _ in range(shared_prefix // page)]
    table = torch.zeros((B + 1, ctx), dtype=torch.int32)
    offs = torch.arange(page, dtype=torch.int32)
    for b in range(B):
        pages = torch.tensor(shared + [perm.pop() for _ in range(own_pages[b])], dtype=torch.int32)
        row = (pages[:, None] * page + offs[None]).reshape(-1)
        table[b + 1, :row.numel()] = row[:ctx]
    req_pool = ReqToTokenPool(B, ctx, dev)
    req_pool.req_to_token.copy_(table.to(dev))
    kv_pool = MHATokenToKVPool(num_pages * page - page, page, torch.float8_e4m3fn if fp8_kv else torch.bfloat16, Hkv,
                               D, 1, dev)
    for buf in (kv_pool.k_buffer[0], kv_pool.v_buffer[0]):
        if fp8_kv:  # e4m3 bytes of N(0, 1) values (scale 0.5)
            buf.copy_(torch.randn(buf.shape, device=dev).mul_(2).to(torch.float8_e4m3fn).view(torch.uint8))
        else:
            buf.normal_(0, 1)
        buf[:page].zero_()
    T = sum(ext_lens)
    q = torch.randn((T, Hq * D), device=dev).to(torch.bfloat16)
    k = torch.randn((T, Hkv * D), device=dev).to(torch.bfloat16)
    v = torch.randn((T, Hkv * D), device=dev).to(torch.bfloat16)
    loc = torch.cat([table[b + 1, prefix_lens[b]:seq[b]] for b in range(B)]).long().to(dev)
    backend = B200AttnBackend(Runner(dev, req_pool, kv_pool, page))
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0,
                           attn_type=AttentionType.DECODER if causal else AttentionType.ENCODER_ONLY)
    if fp8_kv:
        layer.k_scale, layer.v_scale, layer.k_scale_float, layer.v_scale_float = 0.5, 0.5, 0.5, 0.5
    fb = ForwardBatch(forward_mode=ForwardMode[mode_name], batch_size=B,
                      req_pool_indices=torch.arange(1, B + 1, device=dev),
                      seq_lens=torch.tensor(seq, dtype=torch.int64, device=dev), out_cache_loc=loc,
                      seq_lens_sum=sum(seq), extend_seq_lens=torch.tensor(ext_lens, dtype=torch.int32, device=dev),
                      extend_prefix_lens=torch.tensor(prefix_lens, dtype=torch.int32, device=dev),
                      extend_seq_lens_cpu=list(ext_lens), extend_prefix_lens_cpu=list(prefix_lens),
                      attn_backend=backend)
    backend.init_forward_metadata(fb)
    o = layer(q, k, v, fb)
    kb, vb = kv_pool.get_kv_buffer(0)
    check = sampled_check(q, o, kb, vb, req_pool.req_to_token, list(range(1, B + 1)), seq, ext_lens, Hq, Hkv, D,
                          causal=causal, k_mul=0.5 if fp8_kv else 1.0, v_mul=0.5 if fp8_kv else 1.0,
                          # e4m3 pool: the kernel computes in the reference Triton kernel's arithmetic (q and p cast
                          # to e4m3); vs EXACT math over the dequantised pool that arithmetic itself is ~1e-2 off
                          atol=5e-2 if fp8_kv else 2e-2)
    assert check["within_tolerance"], f"{label}: output out of tolerance: {check}"
    ms = time_steps(lambda: layer(q, k, v, fb), steps, warmup, barrier) / steps
    if barrier is not None:
        ms = max_over_ranks(ms, dev)
    pairs = 0
    dec_bytes = 0
    for p_, e in zip(prefix_lens, ext_lens):
        if e == 1:
            dec_bytes += (p_ + 1) * Hkv * D * 2 * 2
        else:
            pairs += e * p_ + e * (e + 1) // 2 if causal else e * (p_ + e)
    fl = 4 * D * Hq * pairs
    tf = fl / (ms * 1e-3) / 1e12
    return dict(workload=label, ms_per_step=ms, flops=fl, tflops=tf, decode_row_kv_bytes=dec_bytes, tokens=T,
                frac_of_measured_bf16=tf / peaks["tf_burst"], frac_of_2250=tf / 2250.0, timed_launches=steps,
                output_check=check)


def run_cfg1(dev, gen, steps=50, warmup=5):
    """BASELINE configs[0]: Qwen2-0.5B shape, decode batch 8 x 1K ctx, page_size 1.  The torch_native port runs on the
    host cores (the reference's own CPU case); the same tensors go through B200AttnBackend (store + decode) on the GPU
    and the two outputs are compared (1e-2 rel / 2e-2 abs)."""
    from oracle import torch_native_port as oracle
    from sglang_b200 import B200AttnBackend, ForwardBatch, ForwardMode, RadixAttention

    c = dict(B=8, ctx=1024, Hq=14, Hkv=2, D=64, page=1)
    B, Hq, Hkv, D = c["B"], c["Hq"], c["Hkv"], c["D"]
    req_pool, kv_pool, host = build_decode_state(c, dev, gen)
    backend = B200AttnBackend(Runner(dev, req_pool, kv_pool, 1))
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    devt = {k: v.to(dev) for k, v in host.items()}
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=B, req_pool_indices=devt["req_pool_indices"],
                      seq_lens=devt["seq_lens"], out_cache_loc=devt["out_cache_loc"], seq_lens_sum=B * c["ctx"],
                      seq_lens_cpu=host["seq_lens"], attn_backend=backend)
    backend.init_forward_metadata(fb)
    kc, vc = (b.cpu().clone() for b in kv_pool.get_kv_buffer(0))
    table = req_pool.req_to_token.cpu()
    step = lambda: layer(devt["q"], devt["k"], devt["v"], fb)
    o = step()
    t_gpu = time_steps(step, steps, warmup) / steps
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    k3, v3 = host["k"].view(B, Hkv, D), host["v"].view(B, Hkv, D)

    def cpu_call():
        oracle.store_kv(kc, vc, host["out_cache_loc"], k3, v3)
        return oracle.decode(host["q"], kc, vc, table, host["req_pool_indices"], host["seq_lens"], Hq, Hkv, D, D ** -0.5)

    ref = cpu_call()
    ts = []
    for _ in range(20):
        t0 = time.perf_counter()
        cpu_call()
        ts.append(time.perf_counter() - t0)
    t_cpu = statistics.median(ts) * 1e3
    err = (o.float().cpu() - ref.float()).abs()
    ok = bool((err <= 2e-2 + 1e-2 * ref.float().abs()).all())
    nbytes = decode_bytes(c)
    return dict(workload="qwen2-0.5b shape decode: batch 8 x 1024 ctx, Hq 14 / Hkv 2, head_dim 64, page_size 1 "
                         "(BASELINE configs[0])", bytes=nbytes,
                cpu_torch_native_port=dict(ms=t_cpu, gbs=nbytes / t_cpu / 1e6, cores=threads),
                gpu_layer_call=dict(ms=t_gpu, gbs=nbytes / t_gpu / 1e6, note="store + decode through B200AttnBackend, "
                                    "eager launches; 4 MB of KV: launch-latency-bound, not an HBM number"),
                max_abs_err=float(err.max()), within_tolerance=ok)


def run_cfg2_decode_fp8(dev, gen, peaks, steps=SIDE_STEPS, warmup=SIDE_WARMUP):
    """BASELINE configs[1] shape with an e4m3 KV pool (k/v descale 0.5): decode_fp8_kernel alone, CUDA events."""
    from sglang_b200 import ops

    c = dict(B=64, ctx=4096, Hq=32, Hkv=8, D=128, page=16)
    req_pool, kv_pool, host = build_decode_state(c, dev, gen)
    B, Hq, Hkv, D = c["B"], c["Hq"], c["Hkv"], c["D"]
    k8, v8 = (b.mul(2).to(torch.float8_e4m3fn) for b in kv_pool.get_kv_buffer(0))
    del kv_pool
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    kv_indices = torch.zeros(B * c["ctx"], dtype=torch.int32, device=dev)
    ops.build_kv_indices(req_pool.req_to_token, host["req_pool_indices"].to(dev), host["seq_lens"].to(dev), kv_indptr,
                         kv_indices)
    q = host["q"].to(dev).view(B, Hq, D)
    o = torch.empty_like(q)
    cps = ops.decode_ctas_per_sm(True)
    tail_begin, ms_ = ops.pick_tail_split(B, Hkv, Hq // Hkv, ctas_per_sm=cps)  # 512 units on 444 slots: split the tail
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms_), 16), dtype=torch.uint8, device=dev)
    f = lambda: ops.decode_attention(q, k8, v8, o, kv_indptr, kv_indices, D ** -0.5, max_splits=ms_, min_chunk=256,
                                     workspace=ws, tail_begin=tail_begin, k_descale=0.5, v_descale=0.5)
    f()
    check = sampled_check(q, o, k8, v8, req_pool.req_to_token, list(range(1, B + 1)), [c["ctx"]] * B, [1] * B, Hq, Hkv, D,
                          causal=False, k_mul=0.5, v_mul=0.5)
    assert check["within_tolerance"], f"cfg2 fp8 decode output out of tolerance: {check}"
    t = time_steps(f, steps, warmup) / steps
    nbytes = 2 * B * c["ctx"] * Hkv * D + 2 * B * Hq * D * 2 + 4 * B * c["ctx"]
    gbs = nbytes / (t * 1e-3) / 1e9
    return dict(workload="llama-3.1-8b decode over a float8_e4m3fn KV pool: batch 64 x 4096 ctx, GQA 32q/8kv, "
                         "head_dim 128, page_size 16", kernel="decode_fp8_kernel<128,bf16>", kernel_ms=t, bytes=nbytes,
                tail_split=[tail_begin, ms_],
                gbs=gbs, frac_of_measured_hbm=gbs / peaks["hbm"], frac_of_8tbs=gbs / 8000.0, timed_launches=steps,
                output_check=check, bf16_pool_equivalent_gbs=(nbytes + 2 * B * c["ctx"] * Hkv * D) / (t * 1e-3) / 1e9)


def run_cfg5_decode(dev, gen, peaks, world, steps=SIDE_STEPS, warmup=SIDE_WARMUP, barrier=None):
    """BASELINE configs[4](i): Llama-3.1-70B decode, batch 256 x 8192 ctx, this rank's KV-head shard (Hkv 8/N).
    With `barrier`: barrier-to-barrier, max over ranks."""
    from sglang_b200 import ops

    c = dict(B=256, ctx=8192, Hq=64 // world, Hkv=max(1, 8 // world), D=128, page=16, layers=80)
    req_pool, kv_pool, host = build_decode_state(c, dev, gen)
    B, Hq, Hkv, D = c["B"], c["Hq"], c["Hkv"], c["D"]
    kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    kv_indices = torch.zeros(B * c["ctx"], dtype=torch.int32, device=dev)
    ops.build_kv_indices(req_pool.req_to_token, host["req_pool_indices"].to(dev), host["seq_lens"].to(dev), kv_indptr,
                         kv_indices)
    q = host["q"].to(dev).view(B, Hq, D)
    o = torch.empty_like(q)
    kbuf, vbuf = kv_pool.get_kv_buffer(0)
    warps, ms_, tb_ = ops.plan_decode(B, Hkv, Hq // Hkv, c["ctx"], 256)
    ws = torch.zeros(max(ops.decode_workspace_bytes(B, Hq, D, ms_), 16), dtype=torch.uint8, device=dev)
    f = lambda: ops.decode_attention(q, kbuf, vbuf, o, kv_indptr, kv_indices, D ** -0.5, max_splits=ms_,
                                     min_chunk=256, workspace=ws, warps=warps, tail_begin=tb_)
    f()
    check = sampled_check(q, o, kbuf, vbuf, req_pool.req_to_token, list(range(1, B + 1)), [c["ctx"]] * B, [1] * B,
                          Hq, Hkv, D, causal=False)
    assert check["within_tolerance"], f"cfg5 decode output out of tolerance: {check}"
    t = time_steps(f, steps, warmup, barrier) / steps
    if barrier is not None:
        t = max_over_ranks(t, dev)
    nbytes = decode_bytes(c) - 2 * 2 * B * Hkv * D * 2
    gbs = nbytes / (t * 1e-3) / 1e9
    return dict(workload=f"llama-3.1-70b decode shard: batch 256 x 8192 ctx, Hq {Hq} / Hkv {Hkv} per GPU of {world} "
                         "(BASELINE configs[4])", kernel_ms=t, bytes=nbytes, gbs=gbs, max_splits=ms_,
                timed_launches=steps, output_check=check,
                frac_of_measured_hbm=gbs / peaks["hbm"], frac_of_8tbs=gbs / 8000.0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--headline-only", action="store_true", help="skip the configs[3]/[4] legs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback (use --impl reference)")
    run_b200(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


if __name__ == "__main__":
    main()
