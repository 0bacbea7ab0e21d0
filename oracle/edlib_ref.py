"""TEST INFRASTRUCTURE: ctypes view of oracle/_ref/libedlib_ref.so (the reference's vendored edlib, built by oracle/Makefile)."""
import ctypes as C
import os

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libedlib_ref.so")


def available() -> bool:
    return os.path.exists(_PATH)


_lib = None


def distance(a: str, b: str) -> int:
    global _lib
    if _lib is None:
        _lib = C.CDLL(_PATH)
        _lib.edlib_ref_distance.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int]
        _lib.edlib_ref_distance.restype = C.c_int
    x, y = a.encode(), b.encode()
    return _lib.edlib_ref_distance(x, len(x), y, len(y))
