"""Base class for Python objects that own a handle to an opaque C struct
(the role of the reference's ``src/ska_sdp_func/utility/struct_wrapper.py``)."""

import ctypes
import weakref


class StructWrapper:
    """Creates the C object, keeps its handle, frees it on garbage collection.

    Every subclass gets its own empty ``ctypes.Structure`` type so that ctypes
    refuses a handle of the wrong kind. Instances can be passed straight to
    wrapped C functions (``_as_parameter_``).
    """

    _HANDLE_CLASS = None

    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        cls._HANDLE_CLASS = type(f"{cls.__name__}Handle", (ctypes.Structure,), {})

    def __init__(self, create_func, create_args, free_func):
        if not callable(free_func):
            raise ValueError("free_func must be callable")
        self._handle = None
        handle = create_func(*create_args)
        if not handle:
            raise RuntimeError(
                "Cannot initialise struct wrapper: creation function for "
                f"{type(self).__name__} handle returned a null pointer"
            )
        self._handle = handle
        weakref.finalize(self, free_func, handle)

    @property
    def _as_parameter_(self):
        return self._handle

    @classmethod
    def handle_type(cls):
        """ctypes pointer type to use in ``argtypes`` / ``restype``."""
        return ctypes.POINTER(cls._HANDLE_CLASS)
