"""Gridding functions (the sdp_GridderUvwEsFft path only)."""

from .gridder_uvw_es_fft import GridderUvwEsFft

__all__ = ["GridderUvwEsFft"]
