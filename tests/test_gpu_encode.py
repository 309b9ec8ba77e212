"""GPU parity tests of the encoder: b200jpeg_encode_batch (through the C ABI) must write exactly the files libjpeg-turbo
writes for the same pixels and parameters (oracle/_ref when present, else the plain-C port)."""
import hashlib
import json
import os

import numpy as np
import pytest

import util
from util import GOLDEN, port, ref, synth_image

pytestmark = pytest.mark.gpu

CSS = {(2, 2): "4:2:0", (1, 1): "4:4:4", (2, 1): "4:2:2", (1, 2): "4:4:0"}


def want(img, q, samp, ri):
    if util.have_ref():
        return ref.ljt_encode(img, q, samp, restart_interval=ri)
    return port.encode(img, q, samp, ri)


@pytest.fixture(scope="module")
def enc():
    from nvimagecodec_b200.encoder import Encoder
    e = Encoder(0)
    yield e
    e.close()


def test_encode_batch_byte_exact_all_samplings(enc):
    import torch
    from nvimagecodec_b200.encoder import EncodeParams
    sizes = [(64, 48), (227, 149), (17, 9), (8, 8), (1, 1), (33, 65), (100, 31), (1100, 40), (640, 426)]
    imgs = [synth_image(7 + w, w, h) for (w, h) in sizes]
    for samp, name in CSS.items():
        for q, ri in [(90, 0), (75, 3), (100, 0), (5, 2)]:
            p = EncodeParams(quality=q, chroma_subsampling=name, restart_interval=ri)
            got_host = enc.encode(imgs, ".jpg", p)
            got_dev = enc.encode([torch.from_numpy(i).cuda() for i in imgs], ".jpg", p)
            for im, a, b in zip(imgs, got_host, got_dev):
                w = want(im, q, samp, ri)
                assert a == w, (im.shape, samp, q, ri, "host input")
                assert b == w, (im.shape, samp, q, ri, "device input")


def test_encode_gray_noise_and_golden_md5(enc):
    from nvimagecodec_b200.encoder import EncodeParams
    img = synth_image(3, 131, 77)
    gray = np.ascontiguousarray(img[:, :, 1])
    for q, ri in [(90, 0), (50, 5)]:
        assert enc.encode(gray, ".jpg", EncodeParams(quality=q, restart_interval=ri)) == want(gray, q, (1, 1), ri)
    noise = np.random.default_rng(1).integers(0, 256, (95, 130, 3), dtype=np.uint8)
    for samp, name in CSS.items():
        assert enc.encode(noise, ".jpg", EncodeParams(quality=100, chroma_subsampling=name)) == want(noise, 100, samp, 0)
        # > 512 bits per block: the emit kernel's tiles do not fit their shared-memory window and write to global memory
        assert enc.encode(noise, ".jpg", EncodeParams(quality=98, chroma_subsampling=name, restart_interval=2)) == want(noise, 98, samp, 2)
    md5 = json.load(open(os.path.join(GOLDEN, "golden.json")))["ljt_md5"]
    raw = open(os.path.join(GOLDEN, "testorig.ppm"), "rb").read()
    t = np.frombuffer(raw.split(b"\n", 3)[3], dtype=np.uint8).reshape(149, 227, 3)
    assert hashlib.md5(enc.encode(t, ".jpg", EncodeParams(quality=75, chroma_subsampling="4:2:0"))).hexdigest() == md5["MD5_JPEG_420_ISLOW"]
    assert hashlib.md5(enc.encode(t, ".jpg", EncodeParams(quality=75, chroma_subsampling="4:4:0"))).hexdigest() == md5["MD5_JPEG_440_ISLOW"]


def test_encode_1080p_batch_and_round_trip(enc):
    """BASELINE config C5 shape (reduced count): 1080p RGB -> 4:2:0 q90, byte-exact; and the GPU decoder reads it back to
    exactly what the reference decoder makes of the reference's file"""
    import torch
    from nvimagecodec_b200.decoder import Decoder
    from nvimagecodec_b200.encoder import EncodeParams
    imgs = [synth_image(1000 + i, 1920, 1080) for i in range(6)]
    dev = [torch.from_numpy(i).cuda() for i in imgs]
    got = enc.encode(dev, ".jpg", EncodeParams(quality=90, chroma_subsampling="4:2:0"))
    for im, g in zip(imgs, got):
        assert g == want(im, 90, (2, 2), 0)
    dec = Decoder(0, options=":fancy_upsampling=1")
    back = dec.decode(got)
    for g, b in zip(got, back):
        assert np.array_equal(b.cpu(), port.decode(g, port.FMT_I_RGB, True))
    dec.close()


def test_encode_rejects_out_of_scope(enc):
    from nvimagecodec_b200 import _capi
    img = synth_image(1, 32, 32)
    out, st = enc.encode_raw([dict(buffer=img.ctypes.data, row_stride=96, width=32, height=32, format=_capi.FMT_I_RGB, on_host=1,
                                   subsampling=5, quality=90, restart_interval=0),
                              dict(buffer=img.ctypes.data, row_stride=96, width=32, height=32, format=_capi.FMT_I_RGB, on_host=1,
                                   subsampling=_capi.CSS_420, quality=0, restart_interval=0),
                              dict(buffer=img.ctypes.data, row_stride=96, width=32, height=32, format=_capi.FMT_I_RGB, on_host=1,
                                   subsampling=_capi.CSS_420, quality=90, restart_interval=0)], [100000, 100000, 64])
    assert st[0] == 3 and st[1] == 1 and st[2] == 8  # UNSUPPORTED, INVALID_PARAMETER, BUFFER_TOO_SMALL
    assert out == [None, None, None]


def test_encode_randomized_batches(enc):
    """eight calls on one encoder object (buffers reused with different shapes): 40 random images each (sides 1..600
    log-uniform, photo / noise / flat, a few gray), random quality, sampling and restart interval per call; every file
    byte-exact against libjpeg-turbo"""
    import torch
    from nvimagecodec_b200.encoder import EncodeParams
    rng = np.random.default_rng(4242)
    for call in range(8):
        samp = list(CSS)[int(rng.integers(0, 4))]
        q, ri = int(rng.integers(1, 101)), int(rng.choice([0, 0, 1, 4, 37]))
        gray = call % 4 == 3
        imgs = []
        for i in range(40):
            w, h = int(round(2 ** rng.uniform(0, np.log2(600)))), int(round(2 ** rng.uniform(0, np.log2(600))))
            imgs.append(synth_image(9000 + 100 * call + i, w, h, 1 if gray else 3, ["photo", "photo", "noise", "flat"][int(rng.integers(0, 4))]))
        p = EncodeParams(quality=q, chroma_subsampling="4:4:4" if gray else CSS[samp], restart_interval=ri)
        got = enc.encode([torch.from_numpy(i).cuda() for i in imgs] if call % 2 else imgs, ".jpg", p)
        for i, (im, a) in enumerate(zip(imgs, got)):
            assert a == want(im, q, (1, 1) if gray else samp, ri), (call, i, im.shape, samp, q, ri)
