"""N>1 path on CPU: kv-head sharding (tensor parallelism) needs no collective inside attention.

world_size 2, gloo: each rank takes its contiguous block of kv heads (and the matching q heads) of the
same batch — same req_to_token / out_cache_loc on both ranks, own pool shard — runs the oracle on its
shard, and the all-gathered result must equal the unsharded oracle bit for bit."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import torch_native_port as oracle
from tests.util import make_scenario


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def shard_heads(sc, rank, world):
    """Per-rank view of a scenario (what QKVParallelLinear + a per-rank pool give, llama.py:158-171)."""
    hq, hkv = sc.Hq // world, sc.Hkv // world
    T = sc.q.shape[0]
    q = sc.q.view(T, sc.Hq, sc.D)[:, rank * hq:(rank + 1) * hq].reshape(T, hq * sc.D)
    kc = sc.k_cache[:, rank * hkv:(rank + 1) * hkv].contiguous()
    vc = sc.v_cache[:, rank * hkv:(rank + 1) * hkv].contiguous()
    return q, kc, vc, hq, hkv


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    lens = [33, 120, 7]
    sc = make_scenario(lens, [0] * 3, 8, 4, 64, 16, seed=11)
    sc.q = torch.randn((3, 8 * 64), generator=torch.Generator().manual_seed(3)).to(torch.bfloat16)
    q, kc, vc, hq, hkv = shard_heads(sc, rank, world)
    o = oracle.decode(q, kc, vc, sc.req_to_token, sc.req_pool_indices, torch.tensor(lens), hq, hkv, 64, sc.scaling)
    parts = [torch.empty_like(o) for _ in range(world)]
    dist.all_gather(parts, o.contiguous())
    dist.barrier()
    if rank == 0:
        full = oracle.decode(sc.q, sc.k_cache, sc.v_cache, sc.req_to_token, sc.req_pool_indices,
                             torch.tensor(lens), 8, 4, 64, sc.scaling)
        got = torch.cat([p.view(3, hq, 64) for p in parts], dim=1).view(3, -1)
        out.put(bool(torch.equal(got, full)))
    dist.destroy_process_group()


def test_head_sharding_equals_unsharded_world2():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(timeout=5) is True
