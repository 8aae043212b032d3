"""eqp-cluster_b200: B200-native compute-counts hot path of EQP-cluster behind a C ABI (include/eqp_count.h).

Import name: `eqp_cluster_b200` (the root shim eqp_cluster_b200.py maps it onto this directory, whose name has a hyphen).
"""
from . import build  # noqa: F401
from . import sharding  # noqa: F401
from . import combine  # noqa: F401
from . import fpkm  # noqa: F401
from .native import (CountContext, ChunkCounts, EqpError, Params, default_params, annotation_paths, load_library,  # noqa: F401
                     EXPORTS, MODES, EQP_FIXED_SHIFT, comm_id, reduce_sample_local, wide_to_double)
