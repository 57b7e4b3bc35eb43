"""dirt_b200 -- the `rasterise` forward op and `rasterise_grad` backward of yyuting/dirt as
hand-written sm_100a CUDA kernels behind a C ABI (include/dirt_b200.h, libdirt_b200.so).

The public names mirror the reference package (`import dirt`): dirt/__init__.py:1-2 re-exports
dirt/rasterise_ops.py.
"""
from .rasterise_ops import rasterise, rasterise_batch, rasterise_batch_shared, rasterise_grad  # noqa: F401
from .scene_ops import (hill, hill_batch, oceanic, oceanic_batch, oceanic_horizon, oceanic_horizon_batch,  # noqa: F401
                        oceanic_no_cloud, oceanic_no_cloud_batch, oceanic_opt_flow, oceanic_opt_flow_batch,
                        oceanic_simple_proxy, oceanic_simple_proxy_batch, oceanic_still_cloud, oceanic_still_cloud_batch)
from .deferred import dilate_silhouette, rasterise_gbuffer  # noqa: F401
from ._lib import DirtError, InvalidArgumentError  # noqa: F401
