"""rocm_mpi_b200 -- a Blackwell-native distributed-stencil framework.

Public surface (what the reference scripts call through ImplicitGlobalGrid, e.g.
``scripts/diffusion_2D_ap.jl:17,19,28,42,46,48`` and ``scripts/diffusion_2D_perf.jl:48,53``):

``init_global_grid  finalize_global_grid  update_halo  gather  nx_g ny_g nz_g  x_g y_g z_g
tic toc  select_device  hide_communication``
"""
from .parallel import hide_communication  # noqa: F401
from .parallel import (PROC_NULL, VirtualGrid, a_eff, coords_g, dims_create, finalize_global_grid,  # noqa: F401
                       gather, gather_interior, global_grid, grid_is_initialized, init_global_grid, inner,
                       nx_g, ny_g, nz_g, ol, select_device, t_eff, tic, toc, update_halo, x_g, y_g, z_g)

__version__ = "0.1.0"


def zeros(shape, dtype=None, symmetric: bool = True):
    """Zero-filled field on this rank's device.

    On CUDA with ``symmetric=True`` (default) the field lives in peer-addressable fabric memory (a
    collective call), which is what the fused kernels and the ``peer`` transport of ``update_halo`` use
    without a registration step; otherwise it is an ordinary ``torch.zeros`` tensor."""
    import torch
    from .parallel.fabric import fabric_zeros
    g = global_grid()
    dtype = dtype or torch.float64
    if g.device_type == "cuda" and symmetric:
        return fabric_zeros(shape, dtype)
    return torch.zeros(tuple(shape), dtype=dtype, device=g.device)
