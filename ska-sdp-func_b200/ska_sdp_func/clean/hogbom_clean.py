"""Hogbom CLEAN.

Same function, argument order and meaning as the reference's
``src/ska_sdp_func/clean/hogbom_clean.py:25-84``, bound to
``sdp_hogbom_clean`` of this library. Images may be numpy arrays (staged
through the GPU inside the call) or any ``__cuda_array_interface__`` object
(processed in place); ``cbeam_details`` is a numpy array in both cases, as
with the reference.
"""

import ctypes

from ..utility import Lib, Mem

Lib.wrap_func(
    "sdp_hogbom_clean",
    restype=None,
    argtypes=[
        Mem.handle_type(),
        Mem.handle_type(),
        Mem.handle_type(),
        ctypes.c_double,
        ctypes.c_double,
        ctypes.c_int,
        Mem.handle_type(),
        Mem.handle_type(),
        Mem.handle_type(),
    ],
    check_errcode=True,
)


def hogbom_clean(
    dirty_img,
    psf,
    cbeam_details,
    loop_gain,
    threshold,
    cycle_limit,
    clean_model,
    residual,
    skymodel,
):
    """Deconvolves ``dirty_img`` with ``psf`` by Hogbom's CLEAN.

    :param dirty_img: dirty image ``[n, n]``, fp64 or fp32
    :param psf: point spread function ``[2n, 2n]``, peak at ``[n, n]``
    :param cbeam_details: ``[sigma_x, sigma_y, theta_degrees, size]`` of the
        Gaussian CLEAN beam (numpy array of the images' data type)
    :param loop_gain: fraction of the peak moved per cycle (typically 0.1)
    :param threshold: the loop stops when the peak falls below it
    :param cycle_limit: maximum number of cycles
    :param clean_model: output ``[n, n]``: unconvolved components
    :param residual: output ``[n, n]``: remaining residual image
    :param skymodel: output ``[n, n]``: components convolved with the CLEAN
        beam, plus the residual
    """
    Lib.sdp_hogbom_clean(
        Mem(dirty_img),
        Mem(psf),
        Mem(cbeam_details),
        loop_gain,
        threshold,
        cycle_limit,
        Mem(clean_model),
        Mem(residual),
        Mem(skymodel),
    )
