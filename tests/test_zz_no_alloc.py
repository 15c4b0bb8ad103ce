"""ps_compute_start_array_static: the allocation-free variant of ps_compute_start_array.

Some callers of the library cannot afford (or are not allowed) a device allocation inside a call: a
60 fps render loop where an allocation can stall the GPU, or a process whose memory pool already owns
all of HBM so that a fresh cudaMalloc may fail.  For them the library keeps the scan's partial sums
in a fixed 64 KiB array inside its own image.  This test drives that entry point through the C ABI
and requires bit-identical output to the allocating entry point (integer work: no tolerance), on two
calls in a row with different sizes and contents, because both calls share the one built-in array and
the second must not be disturbed by what the first left behind.  It also checks the size limit.

The file sorts last on purpose: under `pytest -x` everything else has already run when it starts.
It makes no assumption about earlier tests, so it passes alone in a fresh process as well.
"""
import ctypes

import numpy as np
import pytest

from particle_simulator_b200 import capi

pytestmark = pytest.mark.gpu

MAX_CELLS = 16384 * 4096          # PS_STATIC_SCAN_MAX_CELLS in include/particles_cuda.h


def _run(fn, grid: np.ndarray):
    import torch

    cells = grid.size
    d_in = torch.from_numpy(grid.view(np.int32)).cuda()
    d_out = torch.full((cells + 1,), -1, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = fn(ctypes.c_void_p(d_in.data_ptr()), ctypes.c_void_p(d_out.data_ptr()), cells, ctypes.c_void_p(stream))
    assert rc == capi.PS_OK
    torch.cuda.synchronize()
    return d_out.cpu().numpy().view(np.uint32)


def test_static_variant_equals_allocating_variant_twice_in_a_row():
    L = capi.load()
    rng = np.random.default_rng(7)
    # first call: the 16 Mi-particle scene's grid (6144^2 cells, 9216 tiles); second call: a smaller, ragged
    # size with different contents, so stale partial sums from the first call would show up as wrong offsets
    first = rng.poisson(0.444, size=6144 * 6144).astype(np.uint32)
    second = rng.integers(0, 200, size=1_000_003, dtype=np.uint32)
    for grid in (first, second):
        want = _run(L.ps_compute_start_array, grid)
        got = _run(L.ps_compute_start_array_static, grid)
        assert np.array_equal(got, want)
        assert int(got[-1]) == int(grid.sum(dtype=np.uint64) & 0xFFFFFFFF)


def test_static_variant_takes_no_device_memory():
    """The point of the entry point: after its first call on the device (which looks up the built-in array), further
    calls leave the amount of free device memory exactly where it was, also for a size it has not seen before."""
    import torch

    L = capi.load()
    rng = np.random.default_rng(3)
    d_in = torch.from_numpy(rng.integers(0, 9, size=3_000_001, dtype=np.uint32).view(np.int32)).cuda()
    d_out = torch.empty(d_in.numel() + 1, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    call = lambda cells: L.ps_compute_start_array_static(ctypes.c_void_p(d_in.data_ptr()), ctypes.c_void_p(d_out.data_ptr()), cells, ctypes.c_void_p(stream))
    assert call(1000) == capi.PS_OK                      # set-up call
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for cells in (3_000_001, 12_345, 2_000_000):
        assert call(cells) == capi.PS_OK
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0
    want = np.concatenate([[0], np.cumsum(d_in[:2_000_000].cpu().numpy().view(np.uint32), dtype=np.uint64)]).astype(np.uint32)
    assert np.array_equal(d_out[:2_000_001].cpu().numpy().view(np.uint32), want)


def test_header_exports_the_limit():
    import os
    import re

    text = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "particles_cuda.h")).read()
    m = re.search(r"#define\s+PS_STATIC_SCAN_MAX_CELLS\s+(\d+)ull", text)
    assert m and int(m.group(1)) == MAX_CELLS


def test_static_variant_rejects_sizes_beyond_its_scratch():
    import torch

    L = capi.load()
    buf = torch.zeros(16, dtype=torch.int32, device="cuda")
    rc = L.ps_compute_start_array_static(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(buf.data_ptr()), MAX_CELLS + 1, None)
    assert rc == capi.PS_ERR_CAPACITY
    assert b"67108864" in L.ps_last_error(None)
    torch.cuda.synchronize()
    assert not buf.any()          # nothing was launched


def test_static_variant_at_its_exact_limit():
    L = capi.load()
    grid = np.ones(MAX_CELLS, dtype=np.uint32)
    got = _run(L.ps_compute_start_array_static, grid)
    assert got[0] == 0 and got[-1] == MAX_CELLS and got[12345] == 12345
