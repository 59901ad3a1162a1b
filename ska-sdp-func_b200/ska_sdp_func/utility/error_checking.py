"""Turn the library's trailing ``sdp_Error* status`` argument into Python
exceptions (same messages as the reference's
``src/ska_sdp_func/utility/error_checking.py:11-46``)."""

import ctypes

ERROR_CODE_ARGTYPE = ctypes.POINTER(ctypes.c_int)

# sdp_errors.h: value -> text shown to the user
ERROR_CODE_MEANING = {
    0: "No error",
    1: "Generic runtime error",
    2: "Invalid function argument",
    3: "Unsupported data type(s)",
    4: "Memory allocation failure",
    5: "Memory copy failure",
    6: "Memory location mismatch",
}


class CError(Exception):
    """A library call finished with a non-zero status code."""


def error_checking(lib_func):
    """Wrap ``lib_func`` so that it is called with a fresh status variable
    appended and raises :class:`CError` when the status comes back set."""

    def call_and_check(*args):
        status = ctypes.c_int(0)
        result = lib_func(*args, ctypes.byref(status))
        if status.value != 0:
            text = ERROR_CODE_MEANING.get(status.value, "Unknown error")
            raise CError(f"Error {status.value}: {text}")
        return result

    return call_and_check
