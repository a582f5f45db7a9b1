"""(f3) Device-side radix prefix match against scripts recorded from the REAL RadixCache
(oracle/gen_golden_radix.py: python/sglang/srt/mem_cache/radix_cache.py insert / match_prefix / evict, page sizes
1, 16, 64).  Integer work: every match result and every insert's prefix_len must be bit-identical."""
import os

import pytest
import torch

from tests.util import GOLDEN_DIR, load_golden

PATH = os.path.join(GOLDEN_DIR, "radix_cases.pt")


def test_radix_fixture_present():
    assert os.path.exists(PATH), "run oracle/gen_golden_radix.py"
    g = load_golden(PATH)
    assert len(g["cases"]) == 4 and all(len(c["ops"]) > 50 for c in g["cases"])


@pytest.mark.gpu
@pytest.mark.parametrize("ci", range(4))
def test_device_radix_replays_reference_script(ci):
    from sglang_b200.radix import DeviceRadixCache

    case = load_golden(PATH)["cases"][ci]
    page = case["page_size"]
    cache = DeviceRadixCache(page, max_cached_pages=case["num_kv_pages"] + 8, num_kv_pages=case["num_kv_pages"] + 8)
    for i, op in enumerate(case["ops"]):
        if op["op"] == "insert":
            got = cache.insert([op["tokens"]], [op["values"]])
            assert got == [op["prefix_len"]], f"op {i}: insert prefix_len {got} != {op['prefix_len']}"
        elif op["op"] == "match":
            got = cache.match_prefix([op["tokens"]])[0].cpu()
            assert got.dtype == torch.int64
            assert torch.equal(got, op["indices"]), f"op {i}: match differs ({got.numel()} vs {op['indices'].numel()})"
        else:
            cache.free(op["freed"])


@pytest.mark.gpu
def test_device_radix_batched_match_and_int32_tokens():
    """All lookups of a script in ONE launch (one warp per request), int32 token ids, against the final tree."""
    from sglang_b200.radix import DeviceRadixCache

    case = load_golden(PATH)["cases"][1]
    page = case["page_size"]
    cache = DeviceRadixCache(page, max_cached_pages=case["num_kv_pages"] + 8, num_kv_pages=case["num_kv_pages"] + 8)
    ops = case["ops"]
    last_mut = max(i for i, o in enumerate(ops) if o["op"] != "match")
    for op in ops[:last_mut + 1]:
        if op["op"] == "insert":
            cache.insert([op["tokens"]], [op["values"]])
        elif op["op"] == "evict":
            cache.free(op["freed"])
    tail = [o for o in ops[last_mut + 1:]]
    assert len(tail) > 20
    got = cache.match_prefix([o["tokens"].to(torch.int32) for o in tail])
    for o, g in zip(tail, got):
        assert torch.equal(g.cpu(), o["indices"])
