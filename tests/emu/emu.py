"""tests/emu/emu.py -- TEST INFRASTRUCTURE: build + ctypes binding of the CPU emulation of the b200jpeg kernels."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SO = os.path.join(HERE, "libb200jpeg_emu.so")
CSRC = os.path.join(ROOT, "nvimagecodec_b200", "csrc")
_lib = None


def build(force=False):
    srcs = [os.path.join(HERE, "emu_decode.cpp"), os.path.join(HERE, "emu_encode.cpp"), os.path.join(CSRC, "batch_plan.cpp"),
            os.path.join(CSRC, "jpeg_parse.cpp"), os.path.join(CSRC, "enc_plan.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("jpeg_math.cuh", "huffman.cuh", "device_types.h", "batch_plan.h", "jpeg_parse.h", "host_shim.h", "enc_math.cuh", "enc_types.h",
                                                      "enc_plan.h")]
    if force or not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fwrapv", "-fno-strict-aliasing", "-fPIC", "-shared", "-Wno-unknown-pragmas", "-o", SO] + srcs)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(SO)
        L.emu_decode_batch.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_void_p),
                                       C.POINTER(C.c_int64), C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.emu_encode.restype = C.c_long
        L.emu_encode.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        _lib = L
    return _lib


def encode(pixels, quality=90, subsampling=3, restart_interval=0, planar=False):
    """pixels: (H,W,3) RGB, (3,H,W) planar RGB with planar=True, or (H,W) gray uint8 -> JPEG bytes (emulated kernels)"""
    a = np.ascontiguousarray(pixels, dtype=np.uint8)
    if a.ndim == 2:
        h, w = a.shape; fmt = 7; stride = w; subsampling = 7
    elif planar:
        _, h, w = a.shape; fmt = 3; stride = w
    else:
        h, w = a.shape[:2]; fmt = 4; stride = 3 * w
    cap = h * w * 6 + 65536
    out = np.empty(cap, np.uint8)
    n = lib().emu_encode(a.ctypes.data, w, h, fmt, stride, subsampling, quality, restart_interval, out.ctypes.data, cap)
    if n < 0:
        raise RuntimeError("emu_encode rc=%d" % n)
    return out[:n].tobytes()


def decode_batch(datas, dims, fmt=4, fancy=True, subseq_bytes=128, max_rounds=1 << 30, jacobi=True, want_coef=False, slots=None, rois=None, orients=None,
                 fused=False, halo=0):
    """dims: list of (w, h).  Returns (outputs, statuses, rounds, coefs).  rois: per image None or (x, y, w, h): the output
    holds that region only.  fused: emulate k_huff_fused (decoder option fused=1) instead of the separate kernels;
    halo: its halo length (0 = the kernel's kFusedHalo; 1 makes the chain validation between CTAs fail often)"""
    n = len(datas)
    if rois is not None:
        dims = [(r[2], r[3]) if r else d for d, r in zip(dims, rois)]
    fmts = fmt if isinstance(fmt, (list, tuple)) else [fmt] * n
    if orients is not None:  # (rotated, flip_x, flip_y) per image: odd quarter turns swap the output's width and height
        dims = [(d[1], d[0]) if o and (o[0] // 90) & 1 else d for d, o in zip(dims, orients)]
    outs, strides = [], []
    for (w, h), f in zip(dims, fmts):
        if f in (4, 6, 2):
            outs.append(np.zeros((h, w, 3), np.uint8)); strides.append(3 * w)
        elif f == 7:
            outs.append(np.zeros((1, h, w), np.uint8)); strides.append(w)
        else:
            outs.append(np.zeros((3, h, w), np.uint8)); strides.append(w)
    coefs = [np.zeros(s, np.int16) for s in slots] if want_coef else None
    st = (C.c_int * n)()
    rounds = C.c_int(0)
    stats = (C.c_uint64 * 8)()
    lib().emu_decode_batch(n, (C.c_char_p * n)(*datas), (C.c_size_t * n)(*[len(d) for d in datas]),
                           (C.c_void_p * n)(*[o.ctypes.data for o in outs]), (C.c_int64 * n)(*strides), (C.c_int * n)(*fmts),
                           int(fancy), subseq_bytes, max_rounds, (2 | (halo << 8)) if fused else 1, st, C.byref(rounds),
                           (C.c_void_p * n)(*[c.ctypes.data for c in coefs]) if want_coef else None, stats,
                           (C.c_int * (4 * n))(*[v for r in rois for v in (r if r else (0, 0, 0, 0))]) if rois is not None else None,
                           (C.c_int * n)(*[((o[0] // 90) & 3 | (4 if o[1] else 0) | (8 if o[2] else 0)) if o else 0 for o in orients])
                           if orients is not None else None)
    decode_batch.last_stats = {"rounds": int(stats[0]), "redone": int(stats[1]), "subsequences": int(stats[2]), "skipped": int(stats[3]), "chain_redo": int(stats[4])}
    return outs, list(st), rounds.value, coefs
