"""Gridding functions (the sdp_GridderUvwEsFft path only)."""

from .gridder_uvw_es_fft import GridderUvwEsFft
from .sharded_gridder_uvw_es_fft import ShardedGridderUvwEsFft

__all__ = ["GridderUvwEsFft", "ShardedGridderUvwEsFft"]
