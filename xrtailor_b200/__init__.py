"""xrtailor_b200 — B200-native Swift-mode cloth hot path of XRTailor.

The product is the C-ABI shared library ``libxrt_b200.so`` (``include/xrt_b200.h``)
built from ``csrc/``.  This package is the thin host-side mirror of the
reference's solver component (``ClothSolverBase``,
xrtailor/include/xrtailor/pipeline/base/ClothSolverBase.hpp:55-189) used by the
tests and the bench.  There is no CPU fallback: importing works without a GPU,
creating a solver does not.
"""
from ._abi import lib, XrtParams, XrtQualityParams, XrtError, default_params, default_quality_params  # noqa: F401
from .solver import (ClothSolver, Batch, build_constraints, batch_profile_substep, batch_run, bvh_build_standalone, morton_codes_noalloc, ObjSequenceWriter, greedy_color, traverse_pairs, fem_solve,  # noqa: F401
                     persist_info)  # noqa: F401
from . import scene  # noqa: F401
