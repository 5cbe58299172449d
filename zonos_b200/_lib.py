"""ctypes binding of libzonos_b200.so (the C ABI declared in include/zonos_b200.h).

There is no CPU fallback: if the library is missing or a call fails, an exception is raised.
"""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
# ZB200_LIB_SUFFIX selects an experiment build made with the same variable (zonos_b200/build.py); default: the product library
LIB_PATH = os.path.join(_HERE, "lib", f"libzonos_b200{os.environ.get('ZB200_LIB_SUFFIX', '')}.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "zonos_b200.h")

ZB200_OK, ZB200_ERR_INVALID, ZB200_ERR_CUDA, ZB200_ERR_STATE = 0, -1, -2, -3

_lib = None


class Zb200Error(RuntimeError):
    pass


def declared_symbols():
    """Every function name declared in include/zonos_b200.h."""
    src = open(HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(zb200_[a-z0-9_]+)\s*\(", src)))


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Zb200Error(
                f"{LIB_PATH} not found: build it with `python -m zonos_b200.build` (there is no CPU fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.zb200_last_error.restype = ctypes.c_char_p
        _lib.zb200_packed_weight_bytes.restype = ctypes.c_int64
    return _lib


def check(status: int, what: str = ""):
    if status == ZB200_OK:
        return
    msg = lib().zb200_last_error().decode(errors="replace")
    text = f"{what}: {msg}" if what else msg
    if status == ZB200_ERR_INVALID:
        raise ValueError(text)
    if status == ZB200_ERR_STATE:
        raise AssertionError(text)
    raise Zb200Error(text)


def ptr(t):
    """data_ptr of a torch tensor (or None) as void*."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr(stream=None):
    import torch

    s = torch.cuda.current_stream() if stream is None else stream
    return ctypes.c_void_p(s.cuda_stream)
