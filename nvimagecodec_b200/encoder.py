"""Python host mirror of the reference's `nvimgcodec.Encoder` for the JPEG path (reference python/encoder.cpp:31-60 ctor,
:150-260 batch encode; parameters python/encode_params.cpp / jpeg_encode_params.cpp).  Same names and argument meaning:

    enc = Encoder(device_id=0)
    data = enc.encode(image, ".jpg", EncodeParams(quality=90, chroma_subsampling="4:2:0"))   # -> bytes
    datas = enc.encode([img0, img1, ...], ".jpg", params)                                     # -> list of bytes / None

Images are (H, W, 3) RGB or (H, W) / (H, W, 1) gray uint8: CUDA torch tensors (anything with __cuda_array_interface__)
or host numpy arrays.  All encode work happens in libb200jpeg.so (CUDA, sm_100a); a missing library or GPU raises."""
import ctypes as C

import numpy as np
import torch

from . import _capi

_SUBSAMPLING = {"4:4:4": _capi.CSS_444, "4:2:2": _capi.CSS_422, "4:2:0": _capi.CSS_420, "4:4:0": _capi.CSS_440, "gray": _capi.CSS_GRAY}


class EncodeParams:
    """quality 1..100, chroma_subsampling "4:4:4" | "4:2:2" | "4:2:0" | "4:4:0", restart_interval in MCUs (0 = none);
    optimized_huffman / progressive are outside this codec's scope (python/jpeg_encode_params.cpp)."""

    def __init__(self, quality=95, chroma_subsampling="4:4:4", restart_interval=0, optimized_huffman=False, progressive=False):
        if optimized_huffman or progressive:
            raise ValueError("b200jpeg encodes baseline JPEG with the standard Huffman tables only")
        self.quality = int(quality)
        self.chroma_subsampling = chroma_subsampling
        self.restart_interval = int(restart_interval)


class Encoder:
    def __init__(self, device_id=0, max_num_cpu_threads=0, backend_kinds=None, options=""):
        if not torch.cuda.is_available():
            raise RuntimeError("b200jpeg needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device_id = torch.cuda.current_device() if device_id is None or device_id < 0 else device_id
        self._h = C.c_void_p()
        rc = _capi.lib().b200jpeg_encoder_create(C.byref(self._h), self.device_id, options.encode())
        if rc != 0:
            raise RuntimeError("b200jpeg_encoder_create: %s" % _capi.status_string(rc))
        self.last_statuses = []

    def close(self):
        if getattr(self, "_h", None):
            _capi.lib().b200jpeg_encoder_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def encode_raw(self, inputs, capacities, cuda_stream=0):
        """inputs: list of dicts(buffer, row_stride, width, height, format, on_host, subsampling, quality, restart_interval)
        -> (list of bytes or None, statuses)"""
        n = len(inputs)
        arr = (_capi.EncodeInput * n)()
        for i, d in enumerate(inputs):
            for k, v in d.items():
                setattr(arr[i], k, v)
        bufs = [np.empty(c, np.uint8) for c in capacities]
        outs = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        caps = (C.c_size_t * n)(*capacities)
        sizes = (C.c_size_t * n)()
        st = (C.c_int32 * n)()
        rc = _capi.lib().b200jpeg_encode_batch(self._h, n, arr, outs, caps, sizes, C.c_void_p(cuda_stream), st)
        if rc != 0:
            raise RuntimeError("b200jpeg_encode_batch: %s" % _capi.status_string(rc))
        self.last_statuses = list(st)
        return [bufs[i][:sizes[i]].tobytes() if st[i] == 0 else None for i in range(n)], self.last_statuses

    def encode_into(self, inputs, out_ptrs, capacities, cuda_stream=0):
        """lowest level: caller-owned output buffers (host pointers; page-locked ones are filled without a bounce copy)
        -> (sizes, statuses)"""
        n = len(inputs)
        arr = (_capi.EncodeInput * n)()
        for i, d in enumerate(inputs):
            for k, v in d.items():
                setattr(arr[i], k, v)
        sizes = (C.c_size_t * n)()
        st = (C.c_int32 * n)()
        rc = _capi.lib().b200jpeg_encode_batch(self._h, n, arr, (C.c_void_p * n)(*out_ptrs), (C.c_size_t * n)(*capacities), sizes,
                                               C.c_void_p(cuda_stream), st)
        if rc != 0:
            raise RuntimeError("b200jpeg_encode_batch: %s" % _capi.status_string(rc))
        self.last_statuses = list(st)
        return list(sizes), self.last_statuses

    def stats(self):
        s = (C.c_uint64 * 8)()
        _capi.lib().b200jpeg_encoder_get_stats(self._h, s)
        keys = ("launches", "blocks", "in_bytes", "out_bytes", "stream_bytes", "dct_us", "entropy_us", "stuff_us")
        return dict(zip(keys, [int(v) for v in s]))

    def encode(self, images, codec=".jpg", params=None, cuda_stream=0):
        if codec.lower().lstrip(".") not in ("jpg", "jpeg"):
            raise ValueError("b200jpeg only writes JPEG")
        params = params or EncodeParams()
        single = not isinstance(images, (list, tuple))
        imgs = [images] if single else list(images)
        keep, inputs, caps = [], [], []
        for im in imgs:
            on_host = isinstance(im, np.ndarray)
            if on_host:
                a = np.ascontiguousarray(im, dtype=np.uint8)
                ptr, shape, stride0 = a.ctypes.data, a.shape, a.strides[0]
            else:
                t = im.torch() if hasattr(im, "torch") else im
                if not isinstance(t, torch.Tensor):
                    t = torch.as_tensor(t, device=torch.device("cuda", self.device_id))
                a = t.contiguous()
                ptr, shape, stride0 = a.data_ptr(), tuple(a.shape), a.stride(0)
            keep.append(a)
            gray = len(shape) == 2 or shape[2] == 1
            h, w = shape[0], shape[1]
            inputs.append(dict(buffer=ptr, row_stride=stride0, width=w, height=h, format=_capi.FMT_P_Y if gray else _capi.FMT_I_RGB,
                               on_host=int(on_host), subsampling=_capi.CSS_GRAY if gray else _SUBSAMPLING[params.chroma_subsampling],
                               quality=params.quality, restart_interval=params.restart_interval))
            caps.append(h * w * (1 if gray else 3) * 2 + 4096)
        if not single or not isinstance(imgs[0], np.ndarray):
            torch.cuda.current_stream(torch.device("cuda", self.device_id)).synchronize() if cuda_stream == 0 else None
        out, _ = self.encode_raw(inputs, caps, cuda_stream)
        return out[0] if single else out
