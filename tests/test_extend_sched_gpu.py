"""The extend kernel's dynamic item schedule (csrc/extend.cu: ticket counter pairs in g_ext_sched).

  * a CUDA graph holding extend launches is replayed many times: every replay must find its counter pair re-armed
    (the last CTA out of a launch zeroes it) and give the eager result bit for bit;
  * more launches than counter pairs (1024) in a row, alternating two ragged batches, so slots are reused;
  * two streams launching at the same time use different pairs and do not disturb one another;
  * a grid larger than the item list (the host only knows an upper bound): CTAs without work retire at once.
"""
import pytest
import torch

from sglang_b200 import ops
from tests.test_extend_determinism_gpu import _launch, _prepare
from tests.util import make_scenario

pytestmark = pytest.mark.gpu


def _two_batches():
    a = make_scenario([0, 700, 33, 4000], [1280, 513, 2047, 96], 28, 4, 128, 16, seed=601, pool_slack=1.05)
    b = make_scenario([17, 0, 300], [5, 3000, 129], 32, 8, 128, 16, seed=602, pool_slack=1.05)
    return (a, _prepare(a)), (b, _prepare(b))


def test_graph_replay_rearms_counters():
    (sa, ta), (sb, tb) = _two_batches()
    ea, eb = _launch(sa, ta).clone(), _launch(sb, tb).clone()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            _launch(sa, ta)
            _launch(sb, tb)
    for _ in range(50):
        ta[3].zero_()
        tb[3].zero_()
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(ta[3], ea) and torch.equal(tb[3], eb)


def test_more_launches_than_counter_pairs():
    (sa, ta), (sb, tb) = _two_batches()
    ea, eb = _launch(sa, ta).clone(), _launch(sb, tb).clone()
    for i in range(1100):
        _launch(sa, ta)
        _launch(sb, tb)
        if i % 275 == 274:
            torch.cuda.synchronize()
            assert torch.equal(ta[3], ea) and torch.equal(tb[3], eb)


def test_two_streams_at_once():
    (sa, ta), (sb, tb) = _two_batches()
    ea, eb = _launch(sa, ta).clone(), _launch(sb, tb).clone()
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(30):
        with torch.cuda.stream(s1):
            _launch(sa, ta)
        with torch.cuda.stream(s2):
            _launch(sb, tb)
    torch.cuda.synchronize()
    assert torch.equal(ta[3], ea) and torch.equal(tb[3], eb)


def test_counters_in_the_callers_workspace():
    """With a workspace the ticket counters live at its start: two streams with a workspace each, many launches, the
    counters are back at zero afterwards."""
    (sa, ta), (sb, tb) = _two_batches()
    ea, eb = _launch(sa, ta).clone(), _launch(sb, tb).clone()
    torch.cuda.synchronize()
    assert ops.extend_workspace_bytes(4, 100, 8) == 64
    wa, wb = (torch.zeros(64, dtype=torch.uint8, device="cuda") for _ in range(2))
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def go(sc, t, ws):
        q, kc, vc, o, qo_indptr, kv_indptr, kv_indices = t
        ops.extend_attention(q, kc, vc, o, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=True,
                             page_size=sc.page_size, max_qo_len=max(sc.extend_lens), workspace=ws)

    for _ in range(40):
        with torch.cuda.stream(s1):
            go(sa, ta, wa)
        with torch.cuda.stream(s2):
            go(sb, tb, wb)
    torch.cuda.synchronize()
    assert torch.equal(ta[3], ea) and torch.equal(tb[3], eb)
    assert int(wa.sum()) == 0 and int(wb.sum()) == 0


def test_grid_larger_than_item_list():
    # max_qo_len is only an upper bound: claim 4096 rows per request while the real ones have 1..40
    sc = make_scenario([100, 0, 900], [1, 40, 7], 8, 2, 128, 16, seed=603, pool_slack=1.05)
    t = _prepare(sc)
    exact = _launch(sc, t).clone()
    q, kc, vc, o, qo_indptr, kv_indptr, kv_indices = t
    o2 = torch.zeros_like(o)
    for _ in range(3):
        ops.extend_attention(q, kc, vc, o2, qo_indptr, kv_indptr, kv_indices, sc.scaling, causal=True,
                             page_size=sc.page_size, max_qo_len=4096)
    torch.cuda.synchronize()
    assert torch.equal(o2, exact)
