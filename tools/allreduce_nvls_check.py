"""torchrun worker (GPU, N >= 2): NvlsAllReduce vs NCCL - results and time.
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/allreduce_nvls_check.py
Prints one JSON line per message size from rank 0; exits non-zero on a mismatch."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from sglang_b200 import _lib

TRACE = "--trace" in sys.argv
if TRACE:
    _lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libb200attn_trace.so")
from sglang_b200.collective import NvlsAllReduce  # noqa: E402


def timed(fn, n=20, reps=5, warm=3):
    """Device time per call: n calls captured in ONE CUDA graph (no host launch cost in the number), replayed `reps`
    times; max over ranks."""
    s_ = torch.cuda.Stream()
    s_.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s_):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s_):
            for _ in range(n):
                fn()
        g.replay()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    torch.cuda.current_stream().wait_stream(s_)
    t = torch.tensor([e0.elapsed_time(e1) / (n * reps)], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e3


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl")
    dev = torch.device("cuda", torch.cuda.current_device())
    comm = NvlsAllReduce(dist.group.WORLD, dev, max_size=64 << 20)
    if comm.disabled:
        if rank == 0:
            print(json.dumps(dict(unavailable="no NVLS multicast address on this box")))
        dist.destroy_process_group()
        return
    ok = True
    g = torch.Generator(device="cuda").manual_seed(1234 + rank)
    if TRACE:  # per-phase times of the two-shot kernel (stage-in, barrier, switch reduce + store, barrier, stage-out)
        import ctypes

        lib = ctypes.CDLL(_lib.LIB_PATH)
        for shape in ((2048, 4096), (8192, 4096)):
            x = torch.randn(shape, device=dev, generator=g).to(torch.bfloat16)
            out = torch.empty_like(x)
            for _ in range(5):
                comm.all_reduce(x, out=out)
            torch.cuda.synchronize()
            dist.barrier()
            comm.all_reduce(x, out=out)
            torch.cuda.synchronize()
            buf = (ctypes.c_longlong * (128 * 8))()
            assert lib.b200attn_allreduce_trace_read(buf, ctypes.c_int64(128 * 8 * 8)) == 0
            t = torch.tensor(list(buf), dtype=torch.float64).view(128, 8)[:, :6]
            t = (t - t[:, 0].min()) / 1e3
            if rank == 0:
                names = ["start", "staged", "barrier1", "reduced", "barrier2", "done"]
                print(f"two-shot {shape} bf16, rank 0, us since first CTA start (median / max over 128 CTAs):", flush=True)
                for k, nm in enumerate(names):
                    print(f"  {nm:9s} {float(t[:, k].median()):8.2f} {float(t[:, k].max()):8.2f}", flush=True)
        dist.destroy_process_group()
        return
    # ---- results: integers in bf16 / fp16 / fp32 sum exactly, random values within rounding of the NCCL result ----
    for dtype in (torch.bfloat16, torch.float16, torch.float32):
        for numel in (8, 4096, 64 * 4096, 1000 * 1000 + 8, 8 << 20):
            if numel * dtype.itemsize % 16:
                continue
            x = torch.randint(-8, 9, (numel,), device=dev, generator=g).to(dtype)
            ref = x.clone()
            dist.all_reduce(ref)
            out = comm.all_reduce(x)
            assert out is not None
            if not torch.equal(out, ref):
                ok = False
                print(f"rank {rank}: MISMATCH ints {dtype} numel {numel}", flush=True)
            x = torch.randn(numel, device=dev, generator=g).to(dtype)
            ref = x.float().clone()
            dist.all_reduce(ref)
            out = comm.all_reduce(x)
            tol = 2e-2 if dtype != torch.float32 else 1e-5
            if not torch.allclose(out.float(), ref, atol=tol * world, rtol=tol):
                ok = False
                print(f"rank {rank}: MISMATCH randn {dtype} numel {numel}", flush=True)
    # ---- back-to-back calls reuse the flag slots; the payload()-in-place form; inside a CUDA graph ----
    x = torch.randint(-4, 5, (64, 4096), device=dev, generator=g).to(torch.bfloat16)
    ref = x.clone()
    dist.all_reduce(ref)
    for _ in range(300):
        out = comm.all_reduce(x)
    if not torch.equal(out, ref):
        ok = False
        print(f"rank {rank}: MISMATCH after 300 back-to-back calls", flush=True)
    p = comm.payload((64, 4096))
    p.copy_(x)
    comm.all_reduce_payload(p)
    if not torch.equal(p, ref):
        ok = False
        print(f"rank {rank}: MISMATCH payload form", flush=True)
    out_g = torch.empty_like(x)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        comm.all_reduce(x, out=out_g)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            comm.all_reduce(x, out=out_g)
    torch.cuda.current_stream().wait_stream(s)
    for i in range(5):
        x.fill_(i + rank)
        graph.replay()
        torch.cuda.synchronize()
        want = float(sum(i + r for r in range(world)))
        if not bool((out_g.float() == want).all()):
            ok = False
            print(f"rank {rank}: MISMATCH graph replay {i}", flush=True)
    # ---- time: this kernel vs NCCL vs the reference communicator's sequence (copy in, torch multimem op, copy out) ----
    import torch.distributed._symmetric_memory as symm

    ref_buf = symm.empty(64 << 19, dtype=torch.bfloat16, device=dev)
    symm.rendezvous(ref_buf, dist.group.WORLD.group_name)
    for shape, what in (((1, 4096), "1 token x 4096"), ((8, 4096), "8 x 4096"), ((16, 4096), "16 x 4096"),
                        ((32, 4096), "32 x 4096"), ((64, 4096), "cfg2 decode: 64 tokens x 4096"),
                        ((64, 8192), "cfg5 decode step: 64 x 8192"), ((2048, 4096), "prefill chunk 2048 x 4096"),
                        ((8192, 4096), "prefill chunk 8192 x 4096")):
        x = torch.randn(shape, device=dev, generator=g).to(torch.bfloat16)
        out = torch.empty_like(x)
        n = x.numel()
        t_ours = timed(lambda: comm.all_reduce(x, out=out))
        t_nccl = timed(lambda: dist.all_reduce(x))

        def reference_sequence():
            ref_buf[:n].copy_(x.view(-1))
            torch.ops.symm_mem.multimem_all_reduce_(ref_buf[:n], "sum", dist.group.WORLD.group_name)
            out.copy_(ref_buf[:n].view(out.shape))

        try:
            t_ref = timed(reference_sequence)
        except Exception as e:  # noqa: BLE001
            t_ref = None
            if rank == 0:
                print("reference sequence unavailable:", type(e).__name__, e, flush=True)
        if rank == 0:
            nbytes = n * 2
            print(json.dumps(dict(leg="o_proj_allreduce", message=what, bytes=nbytes, n_gpus=world,
                                  b200attn_us=round(t_ours, 2), nccl_us=round(t_nccl, 2),
                                  reference_symm_mem_sequence_us=None if t_ref is None else round(t_ref, 2),
                                  b200attn_busbw_gbs=round(2 * (world - 1) / world * nbytes / t_ours / 1e3, 1))), flush=True)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print("RESULT", "PASS" if int(flag) == 0 else "FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 0 else 1)


if __name__ == "__main__":
    main()
