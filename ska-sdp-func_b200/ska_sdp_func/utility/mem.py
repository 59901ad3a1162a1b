"""Wrap arrays as ``sdp_Mem`` handles without copying.

Accepts what the reference's ``src/ska_sdp_func/utility/mem.py:38-121``
accepts -- numpy arrays (host) and cupy arrays (device) -- and, because cupy
is only one of several owners of device memory, any object that implements
``__cuda_array_interface__`` (cupy, numba, torch CUDA tensors ...). Nothing
here imports cupy or torch.
"""

import ctypes

import numpy

from .lib import Lib
from .struct_wrapper import StructWrapper


class Mem(StructWrapper):
    """A non-owning ``sdp_Mem`` view of an array."""

    class MemType:
        """sdp_MemType values (sdp_mem.h)."""

        SDP_MEM_VOID = 0
        SDP_MEM_CHAR = 1
        SDP_MEM_INT = 2
        SDP_MEM_FLOAT = 4
        SDP_MEM_DOUBLE = 8
        SDP_MEM_COMPLEX_FLOAT = 36
        SDP_MEM_COMPLEX_DOUBLE = 40

    class MemLocation:
        """sdp_MemLocation values (sdp_mem.h)."""

        SDP_MEM_CPU = 0
        SDP_MEM_GPU = 1

    _TYPE_OF = {
        numpy.dtype(numpy.int8): MemType.SDP_MEM_CHAR,
        numpy.dtype(numpy.int32): MemType.SDP_MEM_INT,
        numpy.dtype(numpy.float32): MemType.SDP_MEM_FLOAT,
        numpy.dtype(numpy.float64): MemType.SDP_MEM_DOUBLE,
        numpy.dtype(numpy.complex64): MemType.SDP_MEM_COMPLEX_FLOAT,
        numpy.dtype(numpy.complex128): MemType.SDP_MEM_COMPLEX_DOUBLE,
    }

    def __init__(self, *args):
        obj = args[0] if len(args) == 1 else None
        self._keepalive = obj
        if isinstance(obj, numpy.ndarray):
            mem_type = self._lookup(obj.dtype, "numpy")
            self._wrap(obj.ctypes.data, mem_type, self.MemLocation.SDP_MEM_CPU,
                       obj.shape, obj.strides, not obj.flags.writeable)
        elif hasattr(obj, "__cuda_array_interface__"):
            cai = obj.__cuda_array_interface__
            dtype = numpy.dtype(cai["typestr"])
            mem_type = self._lookup(dtype, "device")
            shape = tuple(cai["shape"])
            strides = cai.get("strides")
            if strides is None:  # C-contiguous
                strides, step = [], dtype.itemsize
                for extent in reversed(shape):
                    strides.insert(0, step)
                    step *= extent
            ptr, read_only = cai["data"]
            self._wrap(ptr, mem_type, self.MemLocation.SDP_MEM_GPU, shape,
                       tuple(strides), bool(read_only))
        elif obj is None:
            self._wrap(None, self.MemType.SDP_MEM_VOID,
                       self.MemLocation.SDP_MEM_CPU, (), (), False)
        else:
            raise TypeError("Unsupported argument type")

    def _lookup(self, dtype, kind):
        try:
            return self._TYPE_OF[numpy.dtype(dtype)]
        except KeyError as err:
            raise TypeError(f"Unsupported type of {kind} array") from err

    def _wrap(self, ptr, mem_type, location, shape, strides, read_only):
        ndim = len(shape)
        c_shape = (ctypes.c_int64 * max(ndim, 1))(*shape)
        c_strides = (ctypes.c_int64 * max(ndim, 1))(*strides)
        super().__init__(
            Lib.sdp_mem_create_wrapper,
            (ctypes.c_void_p(ptr), mem_type, location, ndim, c_shape, c_strides),
            Lib.sdp_mem_free,
        )
        Lib.sdp_mem_set_read_only(self, int(read_only))


Lib.wrap_func(
    "sdp_mem_create_wrapper",
    restype=Mem.handle_type(),
    argtypes=[
        ctypes.c_void_p,
        ctypes.c_int,
        ctypes.c_int,
        ctypes.c_int32,
        ctypes.POINTER(ctypes.c_int64),
        ctypes.POINTER(ctypes.c_int64),
    ],
    check_errcode=True,
)

Lib.wrap_func(
    "sdp_mem_set_read_only",
    restype=None,
    argtypes=[Mem.handle_type(), ctypes.c_int32],
)

Lib.wrap_func("sdp_mem_free", restype=None, argtypes=[Mem.handle_type()])
