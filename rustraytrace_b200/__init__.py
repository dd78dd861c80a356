"""rustraytrace_b200 — B200 (sm_100a) implementation of RustRayTrace's per-pixel path-tracing loop.

Layout: csrc/ (CUDA kernels + C ABI -> librrt.so), host/ (C++ mirror of the reference's Rust host
surface -> librrt_host.so, bin/rrt), and this thin ctypes layer.  Importing the package does not
load the libraries; the first call does, and fails loudly if they are missing.
"""
from . import _ffi
from .api import (BookScene, Context, DescBuilder, GpuScene, RrtError, copy_camera, flatten_check, flatten_grid, flatten_rects,
                  make_camera, make_overrides, scene_features, shard_plan, shipped_overrides, tile_scatter_host)

__all__ = ["BookScene", "Context", "DescBuilder", "GpuScene", "RrtError", "copy_camera",
           "flatten_check", "flatten_grid", "flatten_rects", "make_camera", "make_overrides", "scene_features", "shard_plan", "shipped_overrides", "tile_scatter_host", "_ffi"]
