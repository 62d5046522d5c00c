"""Worker: finalize_global_grid() followed by a second init_global_grid() in the same process."""
import torch

import rocm_mpi_b200 as igg

for rep, (nx, ny) in enumerate(((8, 6), (10, 7))):
    me, dims, nprocs, coords, _ = igg.init_global_grid(nx, ny, 1, device_type="cpu", quiet=True)
    A = torch.full((nx, ny), float(me))
    igg.update_halo(A)
    g = igg.global_grid()
    if g.neighbors[1][0] >= 0:
        assert float(A[-1, 1]) == float(g.neighbors[1][0])
    igg.finalize_global_grid()
    assert not igg.grid_is_initialized()
print(f"rank {me} ok")
