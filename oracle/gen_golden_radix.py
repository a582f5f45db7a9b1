"""Generate tests/golden/radix_cases.pt by driving the REAL RadixCache (this container only).

python/sglang/srt/mem_cache/radix_cache.py: `RadixCache.create_simulated(page_size=...)`, then a seeded script of
  insert(InsertParams(key=RadixKey(tokens), value=kv slots))      :425-445
  match_prefix(MatchPrefixParams(key=RadixKey(tokens)))           :377-423   -> device_indices (int64)
  evict(EvictParams(num_tokens=n))                                :593-621   -> slots handed to allocator.free_segment
over token sequences that share prefixes of assorted lengths (page-aligned and not), for page sizes 1, 16 and 64.
Every operation's inputs and the reference's outputs are recorded in order; the device-side matcher
(sglang_b200/radix.py) replays the script and must reproduce every output bit for bit.

Run:  python oracle/gen_golden_radix.py     (needs /root/reference; NOT run on the GPU box)
"""
from __future__ import annotations

import os
import random
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_env  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


class RecordingAllocator:
    """Stands in for the paged allocator: evict() hands the evicted nodes' slots to free_segment."""

    device = torch.device("cpu")

    def __init__(self):
        self.freed = []

    def free_segment(self, indices, start_pos=0):
        self.freed.append(indices.clone())

    def free(self, indices):
        self.freed.append(indices.clone())

    def free_segments(self, segs):
        for ind, _ in segs:
            self.freed.append(ind.clone())


def main() -> None:
    ref_env.install()
    from sglang.srt.mem_cache.base_prefix_cache import EvictParams, InsertParams, MatchPrefixParams
    from sglang.srt.mem_cache.radix_cache import RadixCache, RadixKey

    cases = []
    for page, seed, n_ops, vocab in [(1, 11, 60, 50), (16, 12, 120, 30000), (64, 13, 80, 30000), (16, 14, 200, 7)]:
        rng = random.Random(seed)
        alloc = RecordingAllocator()
        cache = RadixCache.create_simulated(mock_allocator=alloc, page_size=page)
        next_page = [1]  # KV page 0 is the reserved padding page
        stems = [[rng.randrange(vocab) for _ in range(rng.choice([page * 3, page * 8 + 5, page * 20]))] for _ in range(4)]
        ops = []

        def new_key():
            stem = rng.choice(stems)
            cut = rng.randrange(0, len(stem) + 1)
            tail = [rng.randrange(vocab) for _ in range(rng.randrange(0, page * 6 + 3))]
            return stem[:cut] + tail

        def fresh_slots(n):
            pages = -(-n // page)
            base = next_page[0]
            next_page[0] += pages
            return (torch.arange(pages * page, dtype=torch.int64) + base * page)[:n]

        for _ in range(n_ops):
            kind = rng.choices(["insert", "match", "evict"], weights=[5, 6, 1])[0]
            if kind == "insert":
                toks = new_key()
                vals = fresh_slots(len(toks))
                res = cache.insert(InsertParams(key=RadixKey(toks, None), value=vals.clone()))
                ops.append(dict(op="insert", tokens=torch.tensor(toks, dtype=torch.int64), values=vals,
                                prefix_len=int(res.prefix_len)))
            elif kind == "match":
                toks = new_key()
                res = cache.match_prefix(MatchPrefixParams(key=RadixKey(toks, None)))
                ops.append(dict(op="match", tokens=torch.tensor(toks, dtype=torch.int64),
                                indices=res.device_indices.clone()))
            else:
                alloc.freed.clear()
                n = rng.randrange(1, page * 10 + 2)
                res = cache.evict(EvictParams(num_tokens=n))
                freed = torch.cat(alloc.freed) if alloc.freed else torch.zeros(0, dtype=torch.int64)
                ops.append(dict(op="evict", num_tokens=n, freed=freed, num_evicted=int(res.num_tokens_evicted)))
        # a final batch of lookups over every stem and every insert seen, against the final tree
        finals = [torch.tensor(s, dtype=torch.int64) for s in stems] + [o["tokens"] for o in ops if o["op"] == "insert"]
        for t in finals:
            res = cache.match_prefix(MatchPrefixParams(key=RadixKey(t.tolist(), None)))
            ops.append(dict(op="match", tokens=t, indices=res.device_indices.clone()))
        cases.append(dict(page_size=page, seed=seed, num_kv_pages=next_page[0] + 1, ops=ops))
        kinds = [o["op"] for o in ops]
        hits = sum(int(o["indices"].numel() > 0) for o in ops if o["op"] == "match")
        print(f"page {page}: {len(ops)} ops ({kinds.count('insert')} insert, {kinds.count('match')} match [{hits} hits], "
              f"{kinds.count('evict')} evict), {next_page[0]} KV pages")
    path = os.path.join(OUT, "radix_cases.pt")
    torch.save(dict(cases=cases, generator="oracle/gen_golden_radix.py",
                    reference="python/sglang/srt/mem_cache/radix_cache.py RadixCache (simulated, CPU)"), path)
    print(path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
