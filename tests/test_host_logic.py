"""Host-side mirror of the reference data contract (no GPU)."""
import numpy as np
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import host, ops


def test_forward_mode_predicates():
    M = host.ForwardMode
    assert M.DECODE.is_decode() and not M.DECODE.is_extend()
    assert M.MIXED.is_extend() and M.MIXED.is_mixed()
    assert M.EXTEND.is_extend() and M.TARGET_VERIFY.is_extend()
    assert M.IDLE.is_idle() and M.IDLE.is_decode_or_idle()
    assert not M.DRAFT_EXTEND_V2.is_extend() and M.DRAFT_EXTEND_V2.is_extend(include_draft_extend_v2=True)


def test_compute_position_matches_oracle():
    pre, ext = [0, 5, 17, 3], [4, 1, 9, 0]
    pos, start = host.compute_position(pre, ext)
    opos, ostart = oracle.positions(pre, ext)
    assert np.array_equal(pos.numpy(), opos) and np.array_equal(start.numpy(), ostart)


def test_alloc_extend_naive_page_invariant():
    page = 16
    free = list(range(1, 40))
    prefix, seq = [0, 20, 32], [37, 45, 33]
    last = [-1, 50 * page + 3, 51 * page + 15]  # last slot of each prefix
    loc = host.alloc_extend_naive(prefix, seq, last, free, page).tolist()
    assert len(loc) == sum(s - p for p, s in zip(prefix, seq))
    t = 0
    for p, s, l in zip(prefix, seq, last):
        seg = loc[t:t + s - p]
        t += s - p
        for j, slot in enumerate(seg):
            assert slot % page == (p + j) % page          # slot = page*P + pos % P
            if (p + j) % page != 0:
                prev = seg[j - 1] if j else l
                assert slot == prev + 1                   # continues the partial page
    assert len(set(loc)) == len(loc)


def test_pool_shapes_and_padding_row():
    r = host.ReqToTokenPool(4, 32, "cpu")
    assert r.req_to_token.shape == (5, 32) and r.req_to_token.dtype == torch.int32
    assert r.alloc(2) == [1, 2] and r.available_size() == 2
    p = host.MHATokenToKVPool(64, 16, torch.bfloat16, 2, 8, 3, "cpu", start_layer=1)
    assert p.get_key_buffer(1).shape == (80, 2, 8) and float(p.get_key_buffer(3).abs().sum()) == 0


def test_pick_max_splits_bounds():
    assert ops.pick_max_splits(64, 8, 4, 4096, num_sms=148) == 1     # 512 CTAs already cover 296 slots
    assert ops.pick_max_splits(8, 8, 4, 4096, num_sms=148) == 4      # 64 units -> 256 CTAs stay in the first wave
    assert ops.pick_max_splits(4096, 8, 4, 4096, num_sms=148) == 1   # enough CTAs already
    assert ops.pick_max_splits(1, 1, 4, 100, num_sms=148) == 1       # never finer than min_chunk
    assert ops.pick_max_splits(1, 8, 4, 1 << 20, num_sms=148) == 37
    assert ops.pick_max_splits(16, 8, 4, 16384, num_sms=148) == 2


def test_split_heuristics_follow_the_resident_cta_count():
    """fp8 pools run 3 CTAs per SM (decode_fp8_kernel), 16-bit pools 2: the wave arithmetic must use the right one."""
    from sglang_b200 import ops

    assert ops.decode_ctas_per_sm(False) == 2 and ops.decode_ctas_per_sm(True) == 3
    # cfg2: 512 units.  2 CTAs/SM: 296 slots, 216-unit tail is 73 % of a wave -> left alone; 3 CTAs/SM: 444 slots,
    # 68-unit tail -> split 6 ways (6 * 68 = 408 <= 444)
    assert ops.pick_tail_split(64, 8, 4, num_sms=148, ctas_per_sm=2) == (-1, 1)
    assert ops.pick_tail_split(64, 8, 4, num_sms=148, ctas_per_sm=3) == (444, 6)
    # split cap = floor(slots / units), bounded by ceil(len / min_chunk)
    assert ops.pick_max_splits(8, 8, 4, 8192, 256, ctas_per_sm=2, num_sms=148) == 4
    assert ops.pick_max_splits(8, 8, 4, 8192, 256, ctas_per_sm=3, num_sms=148) == 6
    assert ops.pick_max_splits(1, 8, 4, 512, 256, ctas_per_sm=3, num_sms=148) == 2


def test_plan_decode_invariants_over_a_shape_sweep():
    """plan_decode is a pure function of the batch shape; whatever it returns must be launchable: 2-warp CTAs never
    split, a whole-grid split stays inside the first wave of resident CTAs, a tail split starts on a wave boundary
    inside the grid and its k copies of the tail fit one wave."""
    sms, slots = 148, 296
    for B in (1, 2, 3, 4, 7, 8, 16, 31, 32, 37, 48, 52, 56, 64, 74, 80, 128, 256, 1000):
        for Hkv, G in ((8, 4), (1, 4), (2, 7), (8, 1), (4, 16), (8, 17)):
            for ctx in (1, 255, 256, 4096, 131072):
                warps, ms, tail = ops.plan_decode(B, Hkv, G, ctx, 256, num_sms=sms)
                units = B * Hkv * (-(-G // 16))
                assert warps in (2, 4) and ms >= 1
                assert (warps, ms, tail) == ops.plan_decode(B, Hkv, G, ctx, 256, num_sms=sms)
                if warps == 2:
                    assert (ms, tail) == (1, -1) and 2.8 * sms < units <= 4 * sms
                elif tail >= 0:
                    assert 2 <= ms <= 8 and tail % slots == 0 and 0 < tail < units
                    assert ms * (units - tail) <= slots
                elif ms > 1:
                    assert ms * units <= slots and ms <= -(-ctx // 256)
                assert ops.decode_workspace_bytes(B, Hkv * G, 128, ms) >= ops.decode_workspace_bytes(B, Hkv * G, 128, 1)
