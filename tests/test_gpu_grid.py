"""GPU parity of the raw grid kernels exposed on caller-owned device memory."""
import ctypes

import numpy as np
import pytest

from particle_simulator_b200 import capi

pytestmark = pytest.mark.gpu


def start_array_reference(grid_array: np.ndarray) -> np.ndarray:
    """computeStartArray, src/core.zig:137-142, plus the particle total as the extra last entry (:168)."""
    out = np.zeros(grid_array.size + 1, dtype=np.uint64)
    np.cumsum(grid_array, dtype=np.uint64, out=out[1:])
    return out.astype(np.uint32)


@pytest.mark.parametrize("cells", [0, 1, 2, 4095, 4096, 4097, 28_800, 1_000_003, 37_748_736])
def test_compute_start_array_matches_the_serial_prefix_sum(cells):
    import torch

    L = capi.load()
    rng = np.random.default_rng(cells)
    # cell populations like the reference density (Poisson, mean 0.444), with a few crowded cells
    grid = rng.poisson(0.444, size=cells).astype(np.uint32)
    if cells > 10:
        grid[rng.integers(0, cells, size=5)] = 255
    want = start_array_reference(grid)
    d_in = torch.from_numpy(grid.view(np.int32)).cuda()
    d_out = torch.full((cells + 1,), -1, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.ps_compute_start_array(ctypes.c_void_p(d_in.data_ptr() if cells else 0), ctypes.c_void_p(d_out.data_ptr()),
                                  cells, ctypes.c_void_p(stream))
    assert rc == capi.PS_OK
    torch.cuda.synchronize()
    got = d_out.cpu().numpy().view(np.uint32)
    assert np.array_equal(got, want)


def test_compute_start_array_in_place():
    import torch

    L = capi.load()
    cells = 180 * 160  # the default scene's grid (src/main.zig:25-26,74)
    grid = np.random.default_rng(1).poisson(0.45, size=cells).astype(np.uint32)
    buf = torch.zeros(cells + 1, dtype=torch.int32, device="cuda")
    buf[:cells] = torch.from_numpy(grid.view(np.int32)).cuda()
    rc = L.ps_compute_start_array(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(buf.data_ptr()), cells, None)
    assert rc == capi.PS_OK
    torch.cuda.synchronize()
    assert np.array_equal(buf.cpu().numpy().view(np.uint32), start_array_reference(grid))


def test_compute_start_array_rejects_null():
    L = capi.load()
    assert L.ps_compute_start_array(None, None, 10, None) == capi.PS_ERR_BAD_ARG
