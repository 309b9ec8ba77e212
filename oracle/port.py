"""oracle/port.py -- TEST INFRASTRUCTURE, NOT PRODUCT.

ctypes binding of oracle/_build/libjpeg_oracle.so, the plain-C restatement (oracle/jpeg_oracle.c) of the reference's
baseline JPEG decode/encode arithmetic.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "libjpeg_oracle.so")
FMT_I_RGB, FMT_I_BGR, FMT_Y, FMT_P_RGB, FMT_P_BGR = 0, 1, 2, 3, 4
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", HERE, "port"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.jo_info.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]
        L.jo_decode.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_int64]
        L.jo_read_coefficients.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_int)]
        L.jo_read_plane.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_int)]
        L.jo_idct_islow.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.jo_encode.restype = C.c_long
        L.jo_encode.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                C.c_size_t]
        _lib = L
    return _lib


def info(data):
    out = (C.c_int * 8)()
    rc = lib().jo_info(data, len(data), out)
    if rc != 0:
        raise RuntimeError("jo_info rc=%d" % rc)
    keys = ("width", "height", "ncomp", "hs", "vs", "restart_interval", "sof", "scan_bytes")
    return dict(zip(keys, out))


def decode(data, fmt=FMT_I_RGB, fancy=True):
    i = info(data)
    w, h = i["width"], i["height"]
    if fmt in (FMT_I_RGB, FMT_I_BGR):
        out = np.zeros((h, w, 3), np.uint8)
        stride = 3 * w
    elif fmt == FMT_Y:
        out = np.zeros((h, w), np.uint8)
        stride = w
    else:
        out = np.zeros((3, h, w), np.uint8)
        stride = w
    rc = lib().jo_decode(data, len(data), fmt, int(fancy), out.ctypes.data, stride)
    if rc != 0:
        raise RuntimeError("jo_decode rc=%d" % rc)
    return out


def coefficients(data, comp):
    i = info(data)
    cap = ((i["width"] + 15) // 8 + 2) * ((i["height"] + 15) // 8 + 2)
    out = np.zeros(cap * 64, np.int16)
    dims = (C.c_int * 2)()
    rc = lib().jo_read_coefficients(data, len(data), comp, out.ctypes.data, cap, dims)
    if rc != 0:
        raise RuntimeError("jo_read_coefficients rc=%d" % rc)
    return out[:dims[0] * dims[1] * 64].reshape(dims[1], dims[0], 64)


def plane(data, comp):
    i = info(data)
    cap = (i["width"] + 32) * (i["height"] + 32)
    out = np.zeros(cap, np.uint8)
    dims = (C.c_int * 2)()
    rc = lib().jo_read_plane(data, len(data), comp, out.ctypes.data, cap, dims)
    if rc != 0:
        raise RuntimeError("jo_read_plane rc=%d" % rc)
    return out[:dims[0] * dims[1]].reshape(dims[1], dims[0])


def encode(pixels, quality=90, samp=(2, 2), restart_interval=0):
    pixels = np.ascontiguousarray(pixels, dtype=np.uint8)
    h, w = pixels.shape[:2]
    ch = 1 if pixels.ndim == 2 else 3
    cap = h * w * ch * 2 + 65536
    out = np.empty(cap, np.uint8)
    n = lib().jo_encode(pixels.ctypes.data, w, h, ch, w * ch, quality, samp[0], samp[1], restart_interval, out.ctypes.data, cap)
    if n < 0:
        raise RuntimeError("jo_encode failed")
    return out[:n].tobytes()
