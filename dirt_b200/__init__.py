"""dirt_b200 -- the `rasterise` forward op and `rasterise_grad` backward of yyuting/dirt as
hand-written sm_100a CUDA kernels behind a C ABI (include/dirt_b200.h, libdirt_b200.so).

The public names mirror the reference package (`import dirt`): dirt/__init__.py:1-2 re-exports
dirt/rasterise_ops.py.
"""
from .rasterise_ops import rasterise, rasterise_batch, rasterise_batch_shared, rasterise_grad  # noqa: F401
from .scene_ops import oceanic_horizon, oceanic_horizon_batch  # noqa: F401
from .deferred import dilate_silhouette, rasterise_gbuffer  # noqa: F401
from ._lib import DirtError, InvalidArgumentError  # noqa: F401
