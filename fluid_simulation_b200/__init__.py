"""fluid_simulation_b200 -- B200-native (sm_100a) lid-driven-cavity projection solver.

One hot path of danieljvickers/fluid_simulation's ``fluid_cpp`` library, rebuilt from scratch:
fused tentative velocity + divergence kernel, temporally blocked Jacobi pressure-Poisson sweeps
with the boundary rows fused in, fused velocity correction. The product is the C-ABI library
``libfluidb200.so`` (``include/fluidb200.h``); this package is its thin Python face.
"""
from ._lib import FluidB200Error, load  # noqa: F401
from .solver import MultiGpuNavierStokes, ParallelNavierStokes, TiledNavierStokes, solve_navier_stokes  # noqa: F401

__all__ = ["ParallelNavierStokes", "TiledNavierStokes", "MultiGpuNavierStokes", "solve_navier_stokes", "FluidB200Error", "load"]
