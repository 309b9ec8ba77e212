"""ctypes binding of libb200jpeg.so (include/b200jpeg.h).  Fails loudly when the CUDA library is missing: there is no
CPU fallback anywhere in this package."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200JPEG_LIB") or os.path.join(HERE, "lib", "libb200jpeg.so")  # override: kernel-variant experiments
EXT_DIR = os.path.join(HERE, "extensions")
EXT_PATH = os.path.join(EXT_DIR, "libb200jpeg_ext.so")

STATUS_SUCCESS, STATUS_INVALID_PARAMETER, STATUS_BAD_JPEG, STATUS_UNSUPPORTED, STATUS_CORRUPT_STREAM = 0, 1, 2, 3, 4
STATUS_CUDA_ERROR, STATUS_ARCH_MISMATCH, STATUS_ALLOC_FAILED, STATUS_BUFFER_TOO_SMALL, STATUS_INTERNAL_ERROR = 5, 6, 7, 8, 9
FMT_P_UNCHANGED, FMT_I_UNCHANGED, FMT_P_RGB, FMT_I_RGB, FMT_P_BGR, FMT_I_BGR, FMT_P_Y = 1, 2, 3, 4, 5, 6, 7
CSS_444, CSS_422, CSS_420, CSS_440, CSS_GRAY = 0, 2, 3, 4, 7


class StreamInfo(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("num_components", C.c_int32), ("h_samp", C.c_int32 * 3),
                ("v_samp", C.c_int32 * 3), ("subsampling", C.c_int32), ("sof_marker", C.c_int32), ("restart_interval", C.c_int32),
                ("precision", C.c_int32), ("supported", C.c_int32), ("scan_offset", C.c_uint64), ("scan_bytes", C.c_uint64)]


class Output(C.Structure):
    _fields_ = [("buffer", C.c_void_p), ("row_stride", C.c_int64), ("format", C.c_int32), ("on_host", C.c_int32),
                ("roi_x", C.c_int32), ("roi_y", C.c_int32), ("roi_width", C.c_int32), ("roi_height", C.c_int32),
                ("cuda_stream", C.c_void_p), ("has_cuda_stream", C.c_int32), ("orientation", C.c_int32)]


MALLOC_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_void_p), C.c_size_t, C.c_void_p)
FREE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)


class Allocators(C.Structure):
    _fields_ = [("device_malloc", MALLOC_FN), ("device_free", FREE_FN), ("device_ctx", C.c_void_p), ("device_padding", C.c_size_t),
                ("pinned_malloc", MALLOC_FN), ("pinned_free", FREE_FN), ("pinned_ctx", C.c_void_p), ("pinned_padding", C.c_size_t)]


class EncodeInput(C.Structure):
    _fields_ = [("buffer", C.c_void_p), ("row_stride", C.c_int64), ("width", C.c_int32), ("height", C.c_int32), ("format", C.c_int32),
                ("on_host", C.c_int32), ("subsampling", C.c_int32), ("quality", C.c_int32), ("restart_interval", C.c_int32)]


EXPORTS = ["b200jpeg_version", "b200jpeg_status_string", "b200jpeg_set_log_callback", "b200jpeg_decode_submit", "b200jpeg_decode_wait", "b200jpeg_shard_batch",
           "b200jpeg_multi_decoder_create", "b200jpeg_multi_decoder_destroy", "b200jpeg_multi_decode_batch", "b200jpeg_parse_header", "b200jpeg_decoder_create", "b200jpeg_decoder_destroy",
           "b200jpeg_decode_batch", "b200jpeg_decode_prepare", "b200jpeg_decode_run", "b200jpeg_decoder_get_stats", "b200jpeg_decoder_get_stage_times", "b200jpeg_debug_fetch",
           "b200jpeg_encoder_create", "b200jpeg_encoder_destroy", "b200jpeg_encode_batch", "b200jpeg_encoder_get_stats",
           "b200jpeg_decoder_set_allocators", "b200jpeg_encoder_set_allocators",
           "b200jpeg_unstuff_noalloc", "b200jpeg_unstuff_noalloc_max_bytes"]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libb200jpeg.so is missing (%s): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                               "this package has no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.b200jpeg_version.restype = C.c_char_p
        L.b200jpeg_status_string.restype = C.c_char_p
        L.b200jpeg_status_string.argtypes = [C.c_int]
        L.b200jpeg_parse_header.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(StreamInfo)]
        L.b200jpeg_decoder_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_char_p]
        L.b200jpeg_decoder_destroy.argtypes = [C.c_void_p]
        L.b200jpeg_decode_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(Output),
                                            C.c_void_p, C.POINTER(C.c_int32)]
        L.b200jpeg_decode_submit.argtypes = L.b200jpeg_decode_batch.argtypes
        L.b200jpeg_decode_wait.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
        L.b200jpeg_shard_batch.argtypes = [C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int, C.POINTER(C.c_int32)]
        L.b200jpeg_multi_decoder_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_int), C.c_char_p]
        L.b200jpeg_multi_decoder_destroy.argtypes = [C.c_void_p]
        L.b200jpeg_multi_decode_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(Output),
                                                  C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.b200jpeg_decode_prepare.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(Output),
                                              C.POINTER(C.c_int32)]
        L.b200jpeg_decode_run.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]
        L.b200jpeg_decoder_get_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.b200jpeg_decoder_get_stage_times.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_int]
        L.b200jpeg_debug_fetch.restype = C.c_int64
        L.b200jpeg_debug_fetch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        L.b200jpeg_encoder_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_char_p]
        L.b200jpeg_encoder_destroy.argtypes = [C.c_void_p]
        L.b200jpeg_encode_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(EncodeInput), C.POINTER(C.c_void_p), C.POINTER(C.c_size_t),
                                            C.POINTER(C.c_size_t), C.c_void_p, C.POINTER(C.c_int32)]
        L.b200jpeg_encoder_get_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.b200jpeg_decoder_set_allocators.argtypes = [C.c_void_p, C.POINTER(Allocators)]
        L.b200jpeg_encoder_set_allocators.argtypes = [C.c_void_p, C.POINTER(Allocators)]
        L.b200jpeg_unstuff_noalloc_max_bytes.restype = C.c_size_t
        L.b200jpeg_unstuff_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def status_string(code):
    return lib().b200jpeg_status_string(code).decode()


def parse_header(data):
    info = StreamInfo()
    rc = lib().b200jpeg_parse_header(data, len(data), C.byref(info))
    if rc != 0:
        raise ValueError("b200jpeg_parse_header: %s" % status_string(rc))
    return info
