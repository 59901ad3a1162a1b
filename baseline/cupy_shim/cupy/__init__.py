"""Minimal stand-in for the ``cupy`` module (harness only, not product code).

The reference's Python package accepts device memory only as
``cupy.ndarray`` (``src/ska_sdp_func/utility/mem.py:75-105``,
``grid_data/gridder_uvw_es_fft.py:90-106``) and cupy is not installed in this
image. This module gives the UNMODIFIED reference package the few things it
touches -- ``cupy.ndarray`` with ``dtype / ndim / shape / strides / data.ptr``,
indexing, ``amin / amax / abs`` and scalar arithmetic -- on top of torch CUDA
tensors, so that ``baseline/ref_gpu.py`` can drive the reference's own CUDA
implementation through its own public API on the same device arrays as the
product. Wrap a tensor with ``cupy.asarray(tensor)``.
"""
import numpy
import torch

int8 = numpy.int8
byte = numpy.byte
int32 = numpy.int32
float32 = numpy.float32
float64 = numpy.float64
complex64 = numpy.complex64
complex128 = numpy.complex128

_NUMPY_OF = {
    torch.int8: numpy.int8, torch.int32: numpy.int32,
    torch.float32: numpy.float32, torch.float64: numpy.float64,
    torch.complex64: numpy.complex64, torch.complex128: numpy.complex128,
}


class _MemoryPointer:
    def __init__(self, ptr):
        self.ptr = ptr


class ndarray:  # pylint: disable=invalid-name
    """The subset of cupy.ndarray the reference package uses."""

    def __init__(self, tensor):
        assert tensor.is_cuda
        self.tensor = tensor

    @property
    def dtype(self):
        return numpy.dtype(_NUMPY_OF[self.tensor.dtype])

    @property
    def ndim(self):
        return self.tensor.dim()

    @property
    def shape(self):
        return tuple(self.tensor.shape)

    @property
    def strides(self):
        size = self.tensor.element_size()
        return tuple(s * size for s in self.tensor.stride())

    @property
    def data(self):
        return _MemoryPointer(self.tensor.data_ptr())

    def __getitem__(self, index):
        return ndarray(self.tensor[index])

    def __float__(self):
        return float(self.tensor)

    def _other(self, other):
        return other.tensor if isinstance(other, ndarray) else other

    def __mul__(self, other):
        return ndarray(self.tensor * self._other(other))

    __rmul__ = __mul__

    def __truediv__(self, other):
        return ndarray(self.tensor / self._other(other))


def asarray(tensor):
    return tensor if isinstance(tensor, ndarray) else ndarray(tensor)


def amin(a):
    return ndarray(torch.amin(a.tensor))


def amax(a):
    return ndarray(torch.amax(a.tensor))


def abs(a):  # pylint: disable=redefined-builtin
    return ndarray(torch.abs(a.tensor))
