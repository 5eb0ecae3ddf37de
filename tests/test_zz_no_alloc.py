"""`gf_csr_to_csc_static`: the minibatch transpose for hosts that may not allocate device memory.

Some programs that load this library through its C ABI cannot call cudaMalloc: their own pool took
the whole device at start-up, or they sit in a latency-critical loop where an allocation would
synchronise the device.  For them the library carries the scratch (the per-word write cursors) in a
fixed device array.  This test drives the entry point the way such a host would — straight through
ctypes, passing only inputs and outputs — and checks it against `gf_csr_to_csc`, whose scratch the
caller allocates.  It runs the pair twice in a row on different minibatches, because the built-in
array is shared between calls and still holds the first call's cursors when the second one starts.

The file name sorts last on purpose: if this test fails, a `pytest -x` run has already executed
every other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _canonical(colptr, csc_row, csc_src, nnz, V):
    """Order inside a column is arbitrary (atomic cursors), so compare sorted (column, source) keys."""
    counts = (colptr[1:] - colptr[:-1]).long()
    col_of_pos = torch.repeat_interleave(torch.arange(V, device=colptr.device), counts)
    key = col_of_pos * max(nnz, 1) + csc_src[:nnz].long()
    order = torch.argsort(key)
    return key[order], csc_row[:nnz].long()[order]


def _minibatch(B, V, per_doc, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.zeros(B, V, device="cuda")
    idx = torch.randint(0, V, (B, per_doc), device="cuda", generator=g)
    x.scatter_add_(1, idx, torch.ones(B, per_doc, device="cuda"))
    from gfedntm_b200.ops.decoder import CsrBatch

    return CsrBatch.from_dense(x)


def test_static_scratch_transpose_matches_allocating_entry_point():
    from gfedntm_b200.ops import _lib

    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    # two different shapes back to back: the second call starts from the first call's leftovers
    for B, V, per_doc, seed in ((96, 5000, 40, 1), (200, 1234, 25, 2)):
        batch = _minibatch(B, V, per_doc, seed)
        nnz = batch.nnz
        i32 = dict(dtype=torch.int32, device="cuda")
        # reference result: the caller allocates the scratch
        colptr_a = torch.full((V + 1,), -1, **i32)
        row_a, src_a = torch.full((nnz,), -1, **i32), torch.full((nnz,), -1, **i32)
        scratch = torch.empty(V, **i32)
        rc = lib.gf_csr_to_csc(batch.indptr.data_ptr(), batch.indices.data_ptr(), B, V, nnz,
                               colptr_a.data_ptr(), scratch.data_ptr(), row_a.data_ptr(), src_a.data_ptr(),
                               stream)
        assert rc == 0
        # entry point under test: inputs and outputs only
        colptr_b = torch.full((V + 1,), -1, **i32)
        row_b, src_b = torch.full((nnz,), -1, **i32), torch.full((nnz,), -1, **i32)
        rc = lib.gf_csr_to_csc_static(batch.indptr.data_ptr(), batch.indices.data_ptr(), B, V, nnz,
                                      colptr_b.data_ptr(), row_b.data_ptr(), src_b.data_ptr(), stream)
        assert rc == 0
        torch.cuda.synchronize()
        assert torch.equal(colptr_a, colptr_b)
        key_a, rows_a = _canonical(colptr_a, row_a, src_a, nnz, V)
        key_b, rows_b = _canonical(colptr_b, row_b, src_b, nnz, V)
        assert torch.equal(key_a, key_b)
        assert torch.equal(rows_a, rows_b)


def test_static_scratch_rejects_oversized_vocabulary():
    from gfedntm_b200.ops import _lib

    lib = _lib.load()
    i32 = dict(dtype=torch.int32, device="cuda")
    indptr = torch.zeros(2, **i32)
    dummy = torch.zeros(4, **i32)
    too_big = (1 << 20) + 1
    rc = lib.gf_csr_to_csc_static(indptr.data_ptr(), dummy.data_ptr(), 1, too_big, 0, dummy.data_ptr(),
                                  dummy.data_ptr(), dummy.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    torch.cuda.synchronize()
