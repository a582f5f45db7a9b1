"""(a3) metadata builders: bit-exact against the oracle (through the C ABI)."""
import numpy as np
import pytest
import torch

from oracle import torch_native_port as oracle
from sglang_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _table(rows, ctx, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 1 << 30, (rows, ctx), generator=g, dtype=torch.int32)


@pytest.mark.parametrize("B,ctx", [(1, 17), (8, 1024), (64, 4096), (300, 2050), (1500, 130)])
@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64])
@pytest.mark.parametrize("out_dtype", [torch.int32, torch.int64])
def test_kv_indices_exact(B, ctx, idx_dtype, out_dtype):
    g = torch.Generator().manual_seed(B)
    table = _table(B + 3, ctx, B)
    req = torch.randperm(B + 3, generator=g)[:B]
    lens = torch.randint(0, ctx + 1, (B,), generator=g)
    lens[0] = ctx
    qo = torch.randint(0, 50, (B,), generator=g)
    indptr_ref, idx_ref = oracle.kv_indptr_indices(table.numpy(), req.tolist(), lens.tolist())
    kv_indptr = torch.full((B + 1,), -7, dtype=torch.int32, device=DEV)
    qo_indptr = torch.full((B + 1,), -7, dtype=torch.int32, device=DEV)
    kv_indices = torch.full((int(lens.sum()) + 5,), -1, dtype=out_dtype, device=DEV)
    ops.build_kv_indices(table.to(DEV), req.to(idx_dtype).to(DEV), lens.to(idx_dtype).to(DEV), kv_indptr,
                         kv_indices, qo_lens=qo.to(torch.int32).to(DEV), qo_indptr=qo_indptr)
    assert np.array_equal(kv_indptr.cpu().numpy(), indptr_ref)
    assert np.array_equal(kv_indices.cpu().numpy()[: len(idx_ref)], idx_ref)
    assert (kv_indices.cpu()[len(idx_ref):] == -1).all()  # nothing written past the end
    assert np.array_equal(qo_indptr.cpu().numpy(), oracle.qo_indptr(qo.tolist()))


def test_kv_indices_with_start_offsets_and_capacity():
    table = _table(6, 512, 1)
    req = torch.tensor([5, 1, 3])
    lens = torch.tensor([100, 0, 257])
    start = torch.tensor([12, 0, 255])
    indptr_ref, idx_ref = oracle.kv_indptr_indices(table.numpy(), req.tolist(), lens.tolist(), start.tolist())
    kv_indptr = torch.zeros(4, dtype=torch.int32, device=DEV)
    kv_indices = torch.full((300,), -1, dtype=torch.int32, device=DEV)  # capacity < total
    ops.build_kv_indices(table.to(DEV), req.to(DEV), lens.to(DEV), kv_indptr, kv_indices, kv_start=start.to(DEV))
    assert np.array_equal(kv_indptr.cpu().numpy(), indptr_ref)
    assert np.array_equal(kv_indices.cpu().numpy(), idx_ref[:300])


@pytest.mark.parametrize("page", [1, 16, 32, 64])
def test_page_table_exact(page):
    B, ctx = 37, 1024
    g = torch.Generator().manual_seed(page)
    table = _table(B + 1, ctx, page)
    req = torch.randperm(B + 1, generator=g)[:B]
    lens = torch.randint(1, ctx + 1, (B,), generator=g)
    max_pages = ctx // page
    ref = oracle.page_table(table.numpy(), req.tolist(), lens.tolist(), page, max_pages, fill=-5)
    pt = torch.full((B, max_pages), -5, dtype=torch.int32, device=DEV)
    ops.build_page_table(table.to(DEV), req.to(DEV), lens.to(DEV), pt, page)
    assert np.array_equal(pt.cpu().numpy(), ref)
