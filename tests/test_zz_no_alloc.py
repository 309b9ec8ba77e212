"""b200jpeg_unstuff_noalloc (include/b200jpeg.h): the unstuffing stage with its scratch in a fixed __device__ array, no
allocation inside the call.  Checked through the C ABI against what the existing entry point (b200jpeg_decode_batch, tapped
with b200jpeg_debug_fetch what=2) produces for the same files, byte for byte (the bar of every byte-level test here), on
two calls in a row -- the second call sees the scratch the first one left behind.  Collected last on purpose (file
name)."""
import ctypes as C

import numpy as np
import pytest

from util import make_jpeg

pytestmark = pytest.mark.gpu


def _existing_clean(torch, datas):
    """unstuffed streams as the existing decode path produces them"""
    from nvimagecodec_b200 import parse_header
    from nvimagecodec_b200.decoder import Decoder
    dec = Decoder(0, options=":fancy_upsampling=1")
    outs, ptrs, strides = [], [], []
    for d in datas:
        h = parse_header(d)
        t = torch.zeros((h.height, h.width * 3), dtype=torch.uint8, device="cuda")
        outs.append(t)
        ptrs.append(t.data_ptr())
        strides.append(h.width * 3)
    st = dec.decode_into(datas, ptrs, strides, [4] * len(datas))
    torch.cuda.synchronize()
    assert all(s == 0 for s in st), st
    res = [dec.debug_fetch(2, i, 0, len(d) + 64).tobytes() for i, d in enumerate(datas)]
    dec.close()
    return res


def _scan_bytes(data):
    from nvimagecodec_b200 import parse_header
    h = parse_header(data)
    return np.frombuffer(data, np.uint8)[h.scan_offset:h.scan_offset + h.scan_bytes].copy()


def _noalloc_clean(torch, data, capacity_slack=64):
    from nvimagecodec_b200 import _capi
    scan = _scan_bytes(data)
    d_raw = torch.from_numpy(scan).cuda()
    d_clean = torch.full((scan.size + capacity_slack,), 0xA5, dtype=torch.uint8, device="cuda")
    d_res = torch.full((4,), 0x7FFFFFFF, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = _capi.lib().b200jpeg_unstuff_noalloc(0, d_raw.data_ptr(), scan.size, d_clean.data_ptr(), d_clean.numel(), d_res.data_ptr(),
                                              C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, _capi.status_string(rc)
    assert torch.cuda.memory_allocated() == before
    torch.cuda.synchronize()
    res = d_res.cpu().numpy().astype(np.int64)
    assert res[3] == 0
    out = d_clean.cpu().numpy()
    assert (out[res[0]:] == 0xA5).all(), "wrote past the reported length"
    return out[:res[0]].tobytes(), int(res[1]), int(res[2])


def test_unstuff_noalloc_matches_decode_path():
    import torch
    assert torch.cuda.is_available()
    # a big stream first, so that later calls find many stale chunk counts in the scratch / restart markers (RSTn removal)
    # / a tiny stream / trailing data behind the first EOI with a second EOI in it (the data is cut at the first marker)
    big = make_jpeg(5, 1920, 1080, (2, 2), 92, 0)
    dri = make_jpeg(6, 640, 427, (2, 1), 85, 7)
    small = make_jpeg(7, 97, 61, (1, 1), 75, 0)
    appended = make_jpeg(8, 320, 240, (2, 2), 80, 4) + b"\x00" * 37 + make_jpeg(9, 64, 48, (2, 2), 80, 0)
    datas = [big, dri, small, appended]
    want = _existing_clean(torch, datas)
    for round_ in range(2):  # two calls in a row per file: the shared scratch keeps the previous call's contents
        for i, (d, w) in enumerate(zip(datas, want)):
            got, markers, consumed = _noalloc_clean(torch, d)
            assert len(got) == len(w) and got == w, (round_, i, len(got), len(w))
            assert (markers > 0) == (i in (1, 3))
    got, _, consumed = _noalloc_clean(torch, appended)
    assert consumed < len(_scan_bytes(appended))  # stopped at the first EOI


def test_unstuff_noalloc_limits():
    import torch
    from nvimagecodec_b200 import _capi
    L = _capi.lib()
    limit = L.b200jpeg_unstuff_noalloc_max_bytes()
    assert limit == 64 << 20
    t = torch.zeros(64, dtype=torch.uint8, device="cuda")
    r = torch.zeros(4, dtype=torch.int32, device="cuda")
    # beyond the fixed scratch: refused, nothing launched
    assert L.b200jpeg_unstuff_noalloc(0, t.data_ptr(), limit + 1, t.data_ptr(), 64, r.data_ptr(), None) == _capi.STATUS_INVALID_PARAMETER
    assert L.b200jpeg_unstuff_noalloc(0, None, 16, t.data_ptr(), 64, r.data_ptr(), None) == _capi.STATUS_INVALID_PARAMETER
    # output too small: flagged on the device, nothing written past the capacity
    scan = _scan_bytes(make_jpeg(11, 320, 240, (2, 2), 90, 0))
    assert scan.size > 3 * 4096
    d_raw = torch.from_numpy(scan).cuda()
    d_clean = torch.full((scan.size,), 0xA5, dtype=torch.uint8, device="cuda")
    cap = 4096 + 100
    assert L.b200jpeg_unstuff_noalloc(0, d_raw.data_ptr(), scan.size, d_clean.data_ptr(), cap, r.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert int(r.cpu()[3]) == 1
    assert (d_clean.cpu().numpy()[cap:] == 0xA5).all()
