"""imd_b200 -- B200-native EAM + TTM hot path behind IMD/lasMD's call sites.

Only what the path needs: ``csrc/`` (CUDA kernels + the C ABI of include/imdcu.h, built
into ``libimdcu.so``), ``capi`` (ctypes binding of that ABI), ``host`` (mirror of the
reference's calc_forces / move_atoms / check_nblist interface) and ``synth`` (synthetic
potentials, lattices and IMD input files).
"""
from . import synth  # noqa: F401

__all__ = ["synth", "capi", "host"]
