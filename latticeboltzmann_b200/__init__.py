"""latticeboltzmann_b200 -- B200-native D2Q9 BGK timestep path (drop-in for the hot path of
jodavies/latticeboltzmann).  CUDA kernels + C ABI live in ``csrc/`` and ``include/``; this
package is the thin host-side mirror used by tests and bench.  No CPU fallback exists."""
from . import geometry, slab                      # noqa: F401
from ._capi import LbmError, load                 # noqa: F401
from .solver import Lattice, group_run, make_ring, print_lattice_group, tb_schedule, tb_schedule_chains  # noqa: F401

__all__ = ["Lattice", "LbmError", "load", "group_run", "make_ring", "print_lattice_group", "tb_schedule", "tb_schedule_chains", "geometry", "slab"]
