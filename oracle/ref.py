"""oracle/ref.py -- TEST INFRASTRUCTURE, NOT PRODUCT.

ctypes binding of oracle/_ref/libnvimgcodec_ref.so: the *unmodified* reference (nvImageCodec core +
extensions/libjpeg_turbo + libjpeg-turbo 3.0.3, all compiled from /root/reference by oracle/Makefile) behind the
small C driver oracle/harness.cpp.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libnvimgcodec_ref.so")

DEVICE_CPU_ONLY = -99999
# nvimgcodecBackendKind_t (include/nvimgcodec.h:543-549) as mask bits (1 << (kind-1))
BACKEND_CPU_ONLY, BACKEND_GPU_ONLY, BACKEND_HYBRID, BACKEND_HW = 1, 2, 4, 8
# nvimgcodecSampleFormat_t (include/nvimgcodec.h:384-392)
P_UNCHANGED, I_UNCHANGED, P_RGB, I_RGB, P_BGR, I_BGR, P_Y, P_YUV = 1, 2, 3, 4, 5, 6, 7, 9
BUFFER_DEVICE, BUFFER_HOST = 1, 2
# nvimgcodecChromaSubsampling_t (include/nvimgcodec.h:360-372)
SAMPLING_444, SAMPLING_422, SAMPLING_420, SAMPLING_440, SAMPLING_411, SAMPLING_410, SAMPLING_GRAY = 0, 2, 3, 4, 5, 6, 7
STATUS_SUCCESS = 1
# nvimgcodecProcessingStatus_t (include/nvimgcodec.h:583-612) and nvimgcodecSampleDataType_t (:340-356)
PS_FAIL, PS_IMAGE_CORRUPTED, PS_CODEC_UNSUPPORTED, PS_BACKEND_UNSUPPORTED, PS_ENCODING_UNSUPPORTED = 0x3, 0x7, 0xB, 0x13, 0x23
PS_ORIENTATION_UNSUPPORTED, PS_ROI_UNSUPPORTED, PS_SAMPLE_TYPE_UNSUPPORTED, PS_SAMPLE_FORMAT_UNSUPPORTED = 0x9, 0x11, 0x41, 0x81
PS_NUM_PLANES_UNSUPPORTED, PS_NUM_CHANNELS_UNSUPPORTED = 0x101, 0x201
TYPE_UINT8, TYPE_UINT16 = 0x0802, 0x1004

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        L.refh_create.restype = C.c_void_p
        L.refh_create.argtypes = [C.c_char_p, C.c_int, C.c_int]
        L.refh_destroy.argtypes = [C.c_void_p]
        L.refh_decoder_create.restype = C.c_void_p
        L.refh_decoder_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_char_p]
        L.refh_decoder_destroy.argtypes = [C.c_void_p]
        L.refh_encoder_create.restype = C.c_void_p
        L.refh_encoder_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_char_p]
        L.refh_encoder_destroy.argtypes = [C.c_void_p]
        L.refh_get_info.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]
        L.refh_decode.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_void_p),
                                  C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64), C.c_int,
                                  C.c_void_p, C.c_int, C.POINTER(C.c_uint), C.POINTER(C.c_double)]
        L.refh_decode_roi.argtypes = L.refh_decode.argtypes + [C.POINTER(C.c_int)]
        L.refh_decode_ex.argtypes = L.refh_decode_roi.argtypes + [C.POINTER(C.c_int), C.c_int]
        L.refh_decode_pipelined.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_void_p),
                                            C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                            C.POINTER(C.c_int64), C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_size_t, C.POINTER(C.c_uint), C.POINTER(C.c_double)]
        L.refh_can_decode_ex.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_int),
                                         C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                         C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint)]
        L.refh_can_decode.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_int),
                                      C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int, C.c_int, C.POINTER(C.c_uint)]
        L.refh_encode.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                  C.POINTER(C.c_int), C.POINTER(C.c_int64), C.c_int, C.c_void_p, C.c_float, C.POINTER(C.c_int),
                                  C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), C.c_int,
                                  C.POINTER(C.c_uint), C.POINTER(C.c_double)]
        L.refh_ljt_encode.restype = C.c_long
        L.refh_ljt_encode.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        L.refh_ljt_read_coefficients.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_int)]
        L.refh_ljt_decode.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64]
        L.refh_ljt_decode_raw.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.POINTER(C.c_size_t), C.POINTER(C.c_int)]
        _lib = L
    return _lib


def _arr(ctype, values):
    return (ctype * len(values))(*values)


def ljt_encode(pixels, quality=90, samp=(2, 2), restart_interval=0, restart_in_rows=0, optimize=False, progressive=False,
               arith=False):
    """libjpeg-turbo jpeg_compress of an (H,W,3) RGB or (H,W) gray uint8 array -> bytes (std Huffman tables, ISLOW)."""
    pixels = np.ascontiguousarray(pixels, dtype=np.uint8)
    h, w = pixels.shape[:2]
    ch = 1 if pixels.ndim == 2 else pixels.shape[2]
    cap = h * w * ch * 2 + 65536
    out = np.empty(cap, dtype=np.uint8)
    n = lib().refh_ljt_encode(pixels.ctypes.data, w, h, ch, quality, samp[0], samp[1], restart_interval, restart_in_rows,
                              int(optimize), int(progressive), int(arith), out.ctypes.data, cap)
    if n < 0:
        raise RuntimeError("libjpeg-turbo encode failed")
    return out[:n].tobytes()


def ljt_decode(data, channels=3, bgr=False, fancy=True):
    """plain libjpeg-turbo decode (ISLOW) -> (H,W,C) uint8"""
    w, h = Session.shared().get_info(data)[:2]
    out = np.zeros((h, w, channels), dtype=np.uint8)
    rc = lib().refh_ljt_decode(data, len(data), channels, int(bgr), int(fancy), out.ctypes.data, w * channels)
    if rc != 0:
        raise RuntimeError("libjpeg-turbo decode failed")
    return out


def ljt_coefficients(data, comp, blocks_w, blocks_h):
    """jpeg_read_coefficients for one component -> (hib, wib, 64) int16 natural order over the MCU-padded block grid"""
    out = np.zeros((blocks_h * blocks_w * 64,), dtype=np.int16)
    dims = (C.c_int * 2)()
    rc = lib().refh_ljt_read_coefficients(data, len(data), comp, out.ctypes.data, blocks_h * blocks_w, dims)
    if rc != 0:
        raise RuntimeError("jpeg_read_coefficients failed rc=%d (grid %dx%d)" % (rc, dims[0], dims[1]))
    assert dims[0] == blocks_w and dims[1] == blocks_h, (dims[0], dims[1], blocks_w, blocks_h)
    return out.reshape(blocks_h, blocks_w, 64)


class Session:
    """nvimgcodecInstance_t with builtin parsers (+ reference libjpeg_turbo_ext, + extensions found in ext_dir)."""
    _shared = None

    def __init__(self, ext_dir=None, with_ljt=True, verbose=0):
        self.h = lib().refh_create(ext_dir.encode() if ext_dir else None, int(with_ljt), verbose)
        if not self.h:
            raise RuntimeError("nvimgcodecInstanceCreate failed")

    @classmethod
    def shared(cls):
        if cls._shared is None:
            cls._shared = cls()
        return cls._shared

    def close(self):
        if self.h:
            lib().refh_destroy(self.h)
            self.h = None

    def get_info(self, data):
        out = (C.c_int * 6)()
        rc = lib().refh_get_info(self.h, data, len(data), out)
        if rc != 0:
            raise RuntimeError("nvimgcodecCodeStreamGetImageInfo failed: %d" % rc)
        return tuple(out)  # width, height, num_planes, channels, chroma_subsampling, jpeg_encoding

    def decoder(self, device_id=DEVICE_CPU_ONLY, backends=BACKEND_CPU_ONLY, num_threads=0, options=":fancy_upsampling=1"):
        return Decoder(self, device_id, backends, num_threads, options)

    def encoder(self, device_id=0, backends=0, num_threads=0, options=""):
        return Encoder(self, device_id, backends, num_threads, options)


def _channels(fmt, unchanged=3):
    return {P_Y: 1, I_RGB: 3, I_BGR: 3, P_RGB: 3, P_BGR: 3}.get(fmt, unchanged)


def _planar(fmt):
    return fmt in (P_RGB, P_BGR, P_Y, P_UNCHANGED, P_YUV)


class Decoder:
    def __init__(self, session, device_id, backends, num_threads, options):
        self.session = session
        self.h = lib().refh_decoder_create(session.h, device_id, backends, num_threads, options.encode())
        if not self.h:
            raise RuntimeError("nvimgcodecDecoderCreate failed")

    def close(self):
        if self.h:
            lib().refh_decoder_destroy(self.h)
            self.h = None

    def can_decode(self, datas, fmt=I_RGB, buffer_kind=BUFFER_HOST, force_format=1, dims=None):
        n = len(datas)
        if dims is None:
            dims = [self.session.get_info(d)[:2] for d in datas]
        fmts = fmt if isinstance(fmt, (list, tuple)) else [fmt] * n
        st = (C.c_uint * n)()
        rc = lib().refh_can_decode(self.h, n, _arr(C.c_char_p, datas), _arr(C.c_size_t, [len(d) for d in datas]),
                                   _arr(C.c_int, [d[0] for d in dims]), _arr(C.c_int, [d[1] for d in dims]),
                                   _arr(C.c_int, fmts), buffer_kind, force_format, st)
        if rc != 0:
            raise RuntimeError("refh_can_decode rc=%d" % rc)
        return list(st)

    def can_decode_ex(self, datas, dims, fmts, planes, channels, types=None, orient=None, apply_exif=0, buffer_kind=BUFFER_HOST,
                      force_format=1):
        """nvimgcodecDecoderCanDecode with the target's image info spelled out per image -> status list"""
        n = len(datas)
        types = types or [TYPE_UINT8] * n
        st = (C.c_uint * n)()
        flat = _arr(C.c_int, [v for o in orient for v in o]) if orient else None
        rc = lib().refh_can_decode_ex(self.h, n, _arr(C.c_char_p, datas), _arr(C.c_size_t, [len(d) for d in datas]),
                                      _arr(C.c_int, [d[0] for d in dims]), _arr(C.c_int, [d[1] for d in dims]), _arr(C.c_int, fmts),
                                      _arr(C.c_int, planes), _arr(C.c_int, channels), _arr(C.c_int, types), flat, apply_exif,
                                      buffer_kind, force_format, st)
        if rc != 0:
            raise RuntimeError("refh_can_decode_ex rc=%d" % rc)
        return list(st)

    def decode_pipelined(self, datas, buffers_a, buffers_b, dims, fmts, strides, buffer_kind, stream=None, rounds=2, depth=2, d2h=None, other=None):
        """`rounds` decode calls with `depth` in flight (2: batch r + 1 is handed in before the future of batch r is waited
        for), alternating between the two buffer sets; d2h = (host_a, host_b, dev_a, dev_b, nbytes): one device -> host copy
        per finished batch on the stream; other = a second Decoder that takes the odd rounds (two batches in flight need two
        decoders with the reference's own plugins) -> (statuses, seconds for all rounds)"""
        n = len(datas)
        st = (C.c_uint * n)()
        secs = (C.c_double * 1)()
        rc = lib().refh_decode_pipelined(self.h, other.h if other else None, n, _arr(C.c_char_p, datas), _arr(C.c_size_t, [len(d) for d in datas]),
                                         _arr(C.c_void_p, buffers_a), _arr(C.c_void_p, buffers_b), _arr(C.c_int, [d[0] for d in dims]),
                                         _arr(C.c_int, [d[1] for d in dims]), _arr(C.c_int, fmts), _arr(C.c_int64, strides), buffer_kind,
                                         stream, rounds, depth, *(d2h if d2h else (None, None, None, None, 0)), st, secs)
        if rc != 0:
            raise RuntimeError("refh_decode_pipelined rc=%d" % rc)
        return list(st), secs[0]

    def decode_host(self, datas, fmt=I_RGB, reps=1, dims=None, row_strides=None, rois=None):
        """Decode to host numpy buffers. Returns (list of arrays, status list, seconds list).
        Interleaved -> (H, W, C); planar -> (C, H, W).  rois: per image (x, y, w, h) or None: decode that region only."""
        n = len(datas)
        if rois is not None:
            dims = [(r[2], r[3]) if r else self.session.get_info(d)[:2] for d, r in zip(datas, rois)]
        if dims is None:
            dims = [self.session.get_info(d)[:2] for d in datas]
        fmts = fmt if isinstance(fmt, (list, tuple)) else [fmt] * n
        outs, strides = [], []
        for i, ((w, h), f) in enumerate(zip(dims, fmts)):
            ch = _channels(f)
            stride = row_strides[i] if row_strides else (w if _planar(f) else w * ch)
            planes = ch if _planar(f) else 1
            outs.append(np.zeros(stride * h * planes, dtype=np.uint8))
            strides.append(stride)
        st = (C.c_uint * n)()
        secs = (C.c_double * reps)()
        args = (self.h, n, _arr(C.c_char_p, datas), _arr(C.c_size_t, [len(d) for d in datas]),
                _arr(C.c_void_p, [o.ctypes.data for o in outs]), _arr(C.c_int, [d[0] for d in dims]),
                _arr(C.c_int, [d[1] for d in dims]), _arr(C.c_int, fmts), _arr(C.c_int64, strides), BUFFER_HOST, None, reps, st, secs)
        if rois is None:
            rc = lib().refh_decode(*args)
        else:
            flat = [v for r in rois for v in (r if r else (0, 0, 0, 0))]
            rc = lib().refh_decode_roi(*args, _arr(C.c_int, flat))
        if rc != 0:
            raise RuntimeError("refh_decode rc=%d" % rc)
        res = []
        for o, (w, h), f, s in zip(outs, dims, fmts, strides):
            ch = _channels(f)
            if _planar(f):
                res.append(o.reshape(ch, h, s)[:, :, :w])
            else:
                res.append(o.reshape(h, s)[:, :w * ch].reshape(h, w, ch))
        return res, list(st), list(secs)

    def decode_raw(self, datas, buffers, dims, fmts, strides, buffer_kind, stream=None, reps=1, rois=None, orient=None, apply_exif=0):
        """Decode into caller-provided raw pointers (ints); used with device buffers for the b200jpeg plugin.
        rois: per image (x, y, w, h) or None; dims are then the regions' sizes.  orient: per image (rotated, flip_x, flip_y)
        with apply_exif=1: dims are the oriented sizes."""
        n = len(datas)
        st = (C.c_uint * n)()
        secs = (C.c_double * reps)()
        args = (self.h, n, _arr(C.c_char_p, datas), _arr(C.c_size_t, [len(d) for d in datas]),
                _arr(C.c_void_p, buffers), _arr(C.c_int, [d[0] for d in dims]), _arr(C.c_int, [d[1] for d in dims]),
                _arr(C.c_int, fmts), _arr(C.c_int64, strides), buffer_kind, stream, reps, st, secs)
        roi_arr = _arr(C.c_int, [v for r in rois for v in (r if r else (0, 0, 0, 0))]) if rois is not None else None
        if orient is not None or apply_exif:
            rc = lib().refh_decode_ex(*args, roi_arr, _arr(C.c_int, [v for o in orient for v in o]) if orient else None, apply_exif)
        elif rois is None:
            rc = lib().refh_decode(*args)
        else:
            rc = lib().refh_decode_roi(*args, roi_arr)
        if rc != 0:
            raise RuntimeError("refh_decode rc=%d" % rc)
        return list(st), list(secs)


class Encoder:
    def __init__(self, session, device_id, backends, num_threads, options):
        self.session = session
        self.h = lib().refh_encoder_create(session.h, device_id, backends, num_threads, options.encode())
        if not self.h:
            raise RuntimeError("nvimgcodecEncoderCreate failed")

    def close(self):
        if self.h:
            lib().refh_encoder_destroy(self.h)
            self.h = None

    def encode_raw(self, buffers, dims, fmts, strides, buffer_kind, quality, samplings, stream=None, reps=1, cap_per_image=None):
        n = len(buffers)
        caps = [cap_per_image or (w * h * 3 + 65536) for (w, h) in dims]
        outs = [np.zeros(c, dtype=np.uint8) for c in caps]
        sizes = (C.c_size_t * n)()
        st = (C.c_uint * n)()
        secs = (C.c_double * reps)()
        rc = lib().refh_encode(self.h, n, _arr(C.c_void_p, buffers), _arr(C.c_int, [d[0] for d in dims]),
                               _arr(C.c_int, [d[1] for d in dims]), _arr(C.c_int, fmts), _arr(C.c_int64, strides), buffer_kind, stream,
                               float(quality), _arr(C.c_int, samplings), _arr(C.c_void_p, [o.ctypes.data for o in outs]),
                               _arr(C.c_size_t, caps), sizes, reps, st, secs)
        if rc != 0:
            raise RuntimeError("refh_encode rc=%d" % rc)
        return [o[:s].tobytes() for o, s in zip(outs, sizes)], list(st), list(secs)
