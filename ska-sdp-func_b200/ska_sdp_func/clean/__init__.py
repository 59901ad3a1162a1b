"""Image-plane deconvolution (same module layout as the reference's
``ska_sdp_func.clean``): Hogbom CLEAN."""
from .hogbom_clean import hogbom_clean

__all__ = ["hogbom_clean"]
