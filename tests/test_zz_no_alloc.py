"""-m gpu test of vgpu_selection_indices_noalloc (include/vgpu.h): the allocation-free variant of the bitmap -> row
numbers step of vgpu_filter_project, whose scratch is a fixed __device__ array.  Checked bit-exactly against
vgpu_filter_project's own sel_idx / num_selected for the same bitmap, on calls in a row of different sizes (the scratch
is shared and keeps the previous call's contents), and that sizes past the fixed maximum are refused.
Collected last on purpose."""
import numpy as np
import pytest

import helpers as H
from helpers import abi, ex, D, I32

pytestmark = pytest.mark.gpu

MAX_ROWS = 16 * 1024 * 1024  # VGPU_NOALLOC_MAX_ROWS


def reference_selection(engine, n, seed, keep):
    """A selection made by the existing entry point: rows with value < keep pass."""
    rng = np.random.default_rng(seed)
    v = rng.random(n)
    prog = engine.Program(ex.call("lt", ex.col(0, D), ex.lit(keep, D)), [ex.col(0, D)], n_columns=1)
    res = engine.filter_project(prog, [engine.flat_column(D, v)], n, want_nulls=False)
    return res, np.flatnonzero(v < keep).astype(np.int32)


def no_alloc_indices(engine, sel_bits, n):
    import torch
    idx = torch.full((max(n, 1),), -1, dtype=torch.int32, device="cuda")
    nsel = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    abi.check(abi.lib().vgpu_selection_indices_noalloc(sel_bits.data_ptr(), n, idx.data_ptr(), nsel.data_ptr(), engine.stream_ptr()))
    torch.cuda.synchronize()
    return idx, int(nsel.item())


# sizes: a large batch first, then smaller ones whose tiles reuse scratch written by the larger call; ragged tails
@pytest.mark.parametrize("sizes", [[(3_000_017, 0.37), (70_001, 0.9), (1023, 0.5), (1, 1.1), (2_000_000, 0.01)]])
def test_matches_filter_project_on_calls_in_a_row(engine, sizes):
    import torch
    for k, (n, keep) in enumerate(sizes):
        ref, want = reference_selection(engine, n, 100 + k, keep)
        assert ref.num_selected == len(want)
        for _ in range(2):  # twice in a row on the same input as well
            idx, nsel = no_alloc_indices(engine, ref.sel_bits, n)
            assert nsel == ref.num_selected
            assert torch.equal(idx[:nsel], ref.sel_idx[:nsel])
            assert np.array_equal(idx[:nsel].cpu().numpy(), want)
            assert bool((idx[nsel:] == -1).all())  # nothing written past the selected rows


def test_empty_and_all_set(engine):
    import torch
    n = 5000
    for fill, count in ((0, 0), (-1, n)):
        bits = torch.full(((n + 63) // 64,), fill, dtype=torch.int64, device="cuda")  # all-ones includes bits past n
        idx, nsel = no_alloc_indices(engine, bits, n)
        assert nsel == count
        assert torch.equal(idx[:nsel].cpu(), torch.arange(count, dtype=torch.int32))
    idx, nsel = no_alloc_indices(engine, torch.zeros(1, dtype=torch.int64, device="cuda"), 0)
    assert nsel == 0


def test_maximum_size_and_refusal_beyond_it(engine):
    import torch
    n = MAX_ROWS
    bits = torch.zeros(n // 64, dtype=torch.int64, device="cuda")
    bits[::1000] = 0x8000000000000001 - (1 << 64)  # rows 0 and 63 of every 1000th word
    idx, nsel = no_alloc_indices(engine, bits, n)
    words = torch.arange(0, n // 64, 1000, dtype=torch.int64)
    want = torch.stack([words * 64, words * 64 + 63], dim=1).reshape(-1).to(torch.int32)
    assert nsel == len(want) and torch.equal(idx[:nsel].cpu(), want)
    with pytest.raises(abi.VgpuError) as e:
        no_alloc_indices(engine, bits, n + 1)
    assert e.value.status == abi.ERR_UNSUPPORTED and "at most" in str(e.value)
    # the refusal left the library usable
    idx, nsel = no_alloc_indices(engine, bits, 64_000)
    assert nsel == 2
