"""Locate ``libska_sdp_func*`` and expose its C functions lazily.

Same contract as the reference's ``src/ska_sdp_func/utility/lib.py:59-161``:
``Lib.wrap_func(name, restype=..., argtypes=[...], check_errcode=bool)``
registers a signature, ``Lib.<name>`` binds it on first use; the library is
looked for in ``$SKA_SDP_FUNC_LIB_DIR``, the in-tree build directory, the
current directory, the package root and ``/usr/local/lib``.
There is no fallback implementation: if the shared library (and with it the
CUDA kernels) cannot be loaded, every wrapped function raises.
"""

import ctypes
import glob
import os
import threading

from .error_checking import ERROR_CODE_ARGTYPE, error_checking


class _LazyFunctions(type):
    """``Lib.some_function`` resolves (and caches) the ctypes function."""

    def __getattr__(cls, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return cls._bind(name)


class Lib(metaclass=_LazyFunctions):
    """Holder of the shared-library handle and the wrapped functions."""

    name = "libska_sdp_func"
    env_name = "SKA_SDP_FUNC_LIB_DIR"
    this_dir = os.path.dirname(os.path.abspath(__file__))
    package_root = os.path.normpath(os.path.join(this_dir, "..", ".."))
    search_dirs = [
        os.path.join(package_root, "lib"),  # in-tree build (Makefile)
        ".",
        package_root,
        "/usr/local/lib",
    ]
    lib = None
    lib_path = None
    mutex = threading.Lock()
    _signatures = {}

    @staticmethod
    def find_lib(search_dirs=None):
        """Path of the first ``libska_sdp_func*`` found, or ''."""
        dirs = list(Lib.search_dirs if search_dirs is None else search_dirs)
        env_dir = os.environ.get(Lib.env_name)
        if env_dir:
            dirs.insert(0, env_dir)
        for directory in dirs:
            hits = sorted(glob.glob(os.path.join(directory, Lib.name) + "*"))
            if hits:
                return os.path.abspath(hits[0])
        return ""

    @staticmethod
    def handle():
        """The ``ctypes.CDLL`` of the library (loaded once)."""
        if Lib.lib is None:
            with Lib.mutex:
                if Lib.lib is None:
                    path = Lib.find_lib()
                    try:
                        Lib.lib = ctypes.CDLL(path) if path else None
                        Lib.lib_path = path
                    except OSError:
                        Lib.lib = None
            if Lib.lib is None:
                raise RuntimeError(
                    f"Cannot find {Lib.name} in {Lib.search_dirs}. "
                    f"Try setting the environment variable {Lib.env_name}"
                )
        return Lib.lib

    @staticmethod
    def wrap_func(func_name, *, restype, argtypes, check_errcode=False):
        """Register the signature of a C function; binding is deferred until
        the function is first used. With ``check_errcode`` the C function is
        assumed to take a trailing ``int* status``, which is supplied and
        checked automatically."""
        Lib._signatures[func_name] = (restype, list(argtypes), check_errcode)

    @classmethod
    def _bind(cls, func_name):
        try:
            restype, argtypes, check_errcode = cls._signatures[func_name]
        except KeyError as err:
            raise KeyError(
                f"The wrapping details for {func_name!r} have not been defined"
            ) from err
        try:
            func = getattr(cls.handle(), func_name)
        except AttributeError as err:
            raise AttributeError(
                f"The C library does not expose a function named {func_name!r}"
            ) from err
        func.restype = restype
        func.argtypes = argtypes + ([ERROR_CODE_ARGTYPE] if check_errcode else [])
        if check_errcode:
            func = error_checking(func)
        setattr(cls, func_name, func)
        return func
