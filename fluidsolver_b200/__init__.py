"""fluidsolver_b200 -- B200-native (sm_100a) implementation of FluidSolver's per-timestep stepping
path behind the reference's FS.hpp / Operators.hpp / LinearSolver_StructHypre call surface.
The product is libfsgpu.so (C ABI: include/fsgpu.h); this package only binds it."""
from .lib import (FS, LinearSolver, FsgpuError, load, npy_write_host, launch_count, device_count, field_shape,  # noqa: F401
                  FIELDS, FIELD_ID, SYMBOLS, LIB_PATH, PROF_CLASSES, prof_enable, prof_read,
                  DIRICHLET, NEUMANN, PERIODIC, SYMMETRY, LEFT, RIGHT, BOTTOM, TOP,
                  PSD_NONE, PSD_LEFT, PSD_RIGHT, PSD_BOTTOM, PSD_TOP, RELAX_JACOBI,
                  FLOW_FORCE_TOTAL, FLOW_OUTFLOW_FIX, FLOW_OUTFLOW_FIX_DY, STEP_OUTFLOW_FIX,
                  STEP_REBUILD_OPERATOR, STEP_DENSITY, STEP_CAPILLARY, TRANSPORT_NCCL, TRANSPORT_LOCAL,
                  COMM_ID_BYTES, comm_id, slab_range, slab_coarse_range, slab_rows, selftest_division)
