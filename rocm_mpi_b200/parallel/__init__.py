"""Parallel runtime: topology, global grid, halo exchange, gather, timers, peer fabric."""
from .topology import PROC_NULL, CartTopology, cart_coords, cart_rank, cart_shift, dims_create  # noqa: F401
from .grid import (CartComm, GlobalGrid, coords_g, finalize_global_grid, global_grid,  # noqa: F401
                   grid_is_initialized, init_global_grid, nx_g, ny_g, nz_g, ol, select_device, x_g, y_g, z_g)
from .halo import update_halo, halo_planes  # noqa: F401
from .gather import gather, gather_interior, gather_interior_device, inner  # noqa: F401
from .timing import tic, toc, t_eff, a_eff  # noqa: F401
from .virtual import VirtualGrid  # noqa: F401
from .hide import hide_communication, frame_regions, interior_region  # noqa: F401
