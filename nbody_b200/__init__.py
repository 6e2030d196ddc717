"""nbody_b200 — B200-native (sm_100a) gravity and density hot path behind the reference's
numpy-in / numpy-out entry points.  See DESIGN.md; C-ABI in include/nbody_b200.h."""
from .snapshot import find_hdf5_files, read_attrs, read_boxsize, read_momenta, read_positions  # noqa: F401

__all__ = ["api", "engine", "tables", "snapshot"]
