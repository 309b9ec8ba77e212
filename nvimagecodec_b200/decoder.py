"""Python host mirror of the reference's `nvimgcodec.Decoder` for the JPEG path (reference python/decoder.cpp:31-58 ctor,
:148-252 batch decode, :283-326 defaults; python/image.cpp for Image).  Same names and argument meaning:

    dec = Decoder(device_id=0, options=":fancy_upsampling=0")
    img = dec.decode(jpeg_bytes)              # -> Image or None
    imgs = dec.decode([bytes, bytes, ...])    # -> list of Image / None

Outputs are device-resident interleaved RGB uint8 (H, W, 3) exactly as the reference's Python layer allocates them
(python/decoder.cpp:148-225: STRIDED_DEVICE, I_RGB).  PyTorch is used only to own device memory and streams.
All decode work happens in libb200jpeg.so (CUDA, sm_100a); a missing library or GPU raises."""
import ctypes as C

import numpy as np
import torch

from . import _capi


class Image:
    """Device image, mirrors nvimgcodec.Image (python/image.cpp:434-481): shape/width/height/dtype,
    __cuda_array_interface__, cpu(), cuda(), plus torch() for zero-copy hand-off."""

    def __init__(self, tensor):
        self._t = tensor

    @property
    def shape(self):
        return tuple(self._t.shape)

    @property
    def height(self):
        return self._t.shape[0]

    @property
    def width(self):
        return self._t.shape[1]

    @property
    def ndim(self):
        return self._t.dim()

    @property
    def dtype(self):
        return np.uint8

    @property
    def __cuda_array_interface__(self):
        return self._t.__cuda_array_interface__

    def __dlpack__(self, stream=None):
        return self._t.__dlpack__(stream=stream)

    def __dlpack_device__(self):
        return self._t.__dlpack_device__()

    def torch(self):
        return self._t

    def cuda(self):
        return self

    def cpu(self):
        return self._t.cpu().numpy()


_FORMATS = {"rgb": _capi.FMT_I_RGB, "bgr": _capi.FMT_I_BGR, "gray": _capi.FMT_P_Y, "planar_rgb": _capi.FMT_P_RGB,
            "planar_bgr": _capi.FMT_P_BGR}


class Decoder:
    def __init__(self, device_id=0, max_num_cpu_threads=0, backend_kinds=None, options=":fancy_upsampling=0"):
        if not torch.cuda.is_available():
            raise RuntimeError("b200jpeg needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device_id = torch.cuda.current_device() if device_id is None or device_id < 0 else device_id
        self._h = C.c_void_p()
        rc = _capi.lib().b200jpeg_decoder_create(C.byref(self._h), self.device_id, options.encode())
        if rc != 0:
            raise RuntimeError("b200jpeg_decoder_create: %s" % _capi.status_string(rc))
        self.last_statuses = []

    def close(self):
        if getattr(self, "_h", None):
            _capi.lib().b200jpeg_decoder_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- low level: caller-provided buffers -------------------------------------------------------------------
    def _outputs(self, buffers, strides, formats, on_host, rois=None, streams=None, orientations=None):
        n = len(buffers)
        outs = (_capi.Output * n)()
        for i in range(n):
            outs[i].buffer = buffers[i]
            outs[i].row_stride = strides[i]
            outs[i].format = formats[i]
            outs[i].on_host = int(on_host)
            if streams is not None and streams[i] is not None:
                outs[i].cuda_stream = streams[i]
                outs[i].has_cuda_stream = 1
            if orientations is not None and orientations[i]:
                rotated, flip_x, flip_y = orientations[i]
                outs[i].orientation = (rotated // 90) & 3 | (4 if flip_x else 0) | (8 if flip_y else 0)
            if rois is not None and rois[i]:
                outs[i].roi_x, outs[i].roi_y, outs[i].roi_width, outs[i].roi_height = rois[i]
        return outs

    def decode_into(self, datas, buffers, strides, formats, on_host=False, cuda_stream=0, rois=None, streams=None, orientations=None):
        """decode a batch into raw pointers (device pointers, or host pointers with on_host=True) -> status list.
        rois: per image None or (x, y, width, height): the buffer receives that region of the image only"""
        n = len(datas)
        st = (C.c_int32 * n)()
        rc = _capi.lib().b200jpeg_decode_batch(self._h, n, (C.c_char_p * n)(*datas), (C.c_size_t * n)(*[len(d) for d in datas]),
                                               self._outputs(buffers, strides, formats, on_host, rois, streams, orientations), C.c_void_p(cuda_stream), st)
        if rc != 0:
            raise RuntimeError("b200jpeg_decode_batch: %s" % _capi.status_string(rc))
        self.last_statuses = list(st)
        return self.last_statuses

    def submit(self, datas, buffers, strides, formats, on_host=False, cuda_stream=0, rois=None, streams=None):
        """first half of decode_into: returns when the host work is done and the device work is enqueued (up to two batches
        in flight); -> the statuses the host already knows"""
        n = len(datas)
        st = (C.c_int32 * n)()
        rc = _capi.lib().b200jpeg_decode_submit(self._h, n, (C.c_char_p * n)(*datas), (C.c_size_t * n)(*[len(d) for d in datas]),
                                                self._outputs(buffers, strides, formats, on_host, rois, streams), C.c_void_p(cuda_stream), st)
        if rc != 0:
            raise RuntimeError("b200jpeg_decode_submit: %s" % _capi.status_string(rc))
        self._pending = getattr(self, "_pending", [])
        self._pending.append(n)
        return list(st)

    def wait(self):
        """second half: blocks until the oldest submitted batch is done -> final status list"""
        n = self._pending.pop(0)
        st = (C.c_int32 * max(n, 1))()
        rc = _capi.lib().b200jpeg_decode_wait(self._h, st)
        if rc != 0:
            raise RuntimeError("b200jpeg_decode_wait: %s" % _capi.status_string(rc))
        self.last_statuses = list(st)[:n]
        return self.last_statuses

    def prepare(self, datas, buffers, strides, formats, on_host=False):
        n = len(datas)
        st = (C.c_int32 * n)()
        self._keep = (datas, (C.c_char_p * n)(*datas))
        rc = _capi.lib().b200jpeg_decode_prepare(self._h, n, self._keep[1], (C.c_size_t * n)(*[len(d) for d in datas]),
                                                 self._outputs(buffers, strides, formats, on_host), st)
        if rc != 0:
            raise RuntimeError("b200jpeg_decode_prepare: %s" % _capi.status_string(rc))
        self._n = n
        return list(st)

    def run(self, cuda_stream=0):
        st = (C.c_int32 * self._n)()
        rc = _capi.lib().b200jpeg_decode_run(self._h, C.c_void_p(cuda_stream), st)
        if rc != 0:
            raise RuntimeError("b200jpeg_decode_run: %s" % _capi.status_string(rc))
        self.last_statuses = list(st)
        return self.last_statuses

    def stats(self):
        s = (C.c_uint64 * 16)()
        _capi.lib().b200jpeg_decoder_get_stats(self._h, s)
        keys = ("launches", "sync_rounds", "scan_bytes", "blocks", "subsequences", "scratch_bytes", "tail_redone", "out_bytes")
        d = dict(zip(keys, [int(v) for v in s[:8]]))
        d["record_bytes"] = int(s[8])
        return d

    STAGES = ("unstuff", "huff_sync", "huff_tail", "huff_write", "dc_scan", "idct", "color")

    def stage_times(self):
        """per-stage device ms of the last run (needs option profile=1) or None"""
        ms = (C.c_float * 7)()
        if _capi.lib().b200jpeg_decoder_get_stage_times(self._h, ms, 7) < 0:
            return None
        return dict(zip(self.STAGES, [float(v) for v in ms]))

    def debug_fetch(self, what, image, component, nbytes, dtype=np.uint8):
        buf = np.zeros(nbytes, np.uint8)
        got = _capi.lib().b200jpeg_debug_fetch(self._h, what, image, component, buf.ctypes.data, nbytes)
        if got < 0:
            raise RuntimeError("b200jpeg_debug_fetch failed: %d" % got)
        return buf[:got].view(dtype)

    # ---- reference-style API ----------------------------------------------------------------------------------
    def decode(self, data, params=None, cuda_stream=0, color="rgb"):
        single = isinstance(data, (bytes, bytearray, memoryview, np.ndarray))
        datas = [data] if single else list(data)
        datas = [d.tobytes() if isinstance(d, np.ndarray) else bytes(d) for d in datas]
        fmt = _FORMATS[color]
        dev = torch.device("cuda", self.device_id)
        tensors, bufs, strides, ok = [], [], [], []
        for d in datas:
            try:
                info = _capi.parse_header(d)
            except ValueError:
                tensors.append(None)
                continue
            # nothing is allocated for a stream the decoder will not take (the header is untrusted: a 65535 x 65535 SOF
            # must not turn into a 12.9 GB allocation)
            if not info.supported or info.width * info.height > (1 << 28):
                tensors.append(None)
                continue
            if fmt in (_capi.FMT_I_RGB, _capi.FMT_I_BGR):
                t = torch.empty((info.height, info.width, 3), dtype=torch.uint8, device=dev)
                stride = info.width * 3
            elif fmt == _capi.FMT_P_Y:
                t = torch.empty((info.height, info.width, 1), dtype=torch.uint8, device=dev)
                stride = info.width
            else:
                t = torch.empty((3, info.height, info.width), dtype=torch.uint8, device=dev)
                stride = info.width
            tensors.append(t)
        live = [i for i, t in enumerate(tensors) if t is not None]
        if live:
            torch.cuda.current_stream(dev).synchronize() if cuda_stream == 0 else None
            st = self.decode_into([datas[i] for i in live], [tensors[i].data_ptr() for i in live],
                                  [tensors[i].shape[1] * tensors[i].shape[2] if tensors[i].dim() == 3 and fmt in (
                                      _capi.FMT_I_RGB, _capi.FMT_I_BGR, _capi.FMT_P_Y) else tensors[i].shape[2] for i in live],
                                  [fmt] * len(live), False, cuda_stream)
            for i, s in zip(live, st):
                if s != 0:
                    tensors[i] = None
        imgs = [Image(t) if t is not None else None for t in tensors]
        return imgs[0] if single else imgs


class MultiDeviceDecoder:
    """One batch over several GPUs of a box: one decoder, host thread and stream set per device, the batch dealt by cost, no
    collective (reference hook: one nvimgcodecDecoder_t per device_id, include/nvimgcodec.h:845).  Outputs stay
    on the GPU that decoded them.

        md = MultiDeviceDecoder([0, 1, 2, 3])
        imgs = md.decode(list_of_bytes)      # -> list of Image (on different devices) / None
    """

    def __init__(self, device_ids, options=":fancy_upsampling=0"):
        if not torch.cuda.is_available():
            raise RuntimeError("b200jpeg needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device_ids = list(device_ids)
        self._h = C.c_void_p()
        rc = _capi.lib().b200jpeg_multi_decoder_create(C.byref(self._h), len(self.device_ids), (C.c_int * len(self.device_ids))(*self.device_ids),
                                                       options.encode())
        if rc != 0:
            raise RuntimeError("b200jpeg_multi_decoder_create: %s" % _capi.status_string(rc))

    def close(self):
        if getattr(self, "_h", None):
            _capi.lib().b200jpeg_multi_decoder_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def shard(self, datas, infos):
        from .sharding import shard_batch_native
        return shard_batch_native([len(d) for d in datas], [i.width for i in infos], [i.height for i in infos], len(self.device_ids))

    def decode_into(self, datas, buffers, strides, formats, device_of_image, on_host=False):
        n = len(datas)
        outs = (_capi.Output * n)()
        for i in range(n):
            outs[i].buffer, outs[i].row_stride, outs[i].format, outs[i].on_host = buffers[i], strides[i], formats[i], int(on_host)
        st = (C.c_int32 * max(n, 1))()
        rc = _capi.lib().b200jpeg_multi_decode_batch(self._h, n, (C.c_char_p * n)(*datas), (C.c_size_t * n)(*[len(d) for d in datas]), outs,
                                                     (C.c_int32 * max(n, 1))(*device_of_image), st)
        if rc != 0:
            raise RuntimeError("b200jpeg_multi_decode_batch: %s" % _capi.status_string(rc))
        return list(st)[:n]

    def decode(self, datas):
        datas = [bytes(d) for d in datas]
        infos, live = [], []
        for i, d in enumerate(datas):
            try:
                info = _capi.parse_header(d)
            except ValueError:
                continue
            if info.supported and info.width * info.height <= (1 << 28):
                infos.append(info)
                live.append(i)
        where = self.shard([datas[i] for i in live], infos)
        tensors = [torch.empty((info.height, info.width, 3), dtype=torch.uint8, device=torch.device("cuda", self.device_ids[w]))
                   for info, w in zip(infos, where)]
        for dev in set(self.device_ids):
            torch.cuda.synchronize(dev)
        st = self.decode_into([datas[i] for i in live], [t.data_ptr() for t in tensors], [info.width * 3 for info in infos],
                              [_capi.FMT_I_RGB] * len(live), where)
        out = [None] * len(datas)
        for i, t, s in zip(live, tensors, st):
            if s == 0:
                out[i] = Image(t)
        return out
