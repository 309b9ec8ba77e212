"""CPU tests of the encoder logic: the per-thread bodies of the CUDA encode kernels (enc_math.cuh), the planner and the
header writer (enc_plan.cpp) run in a CPU emulation (tests/emu/emu_encode.cpp) that keeps the kernels' data layout
(tiles, zigzag coefficient array, per-block bit offsets, OR-ed stream windows, stuffing) and must produce the oracle's
files byte for byte."""
import hashlib
import os
import sys

import numpy as np
import pytest

import util
from util import GOLDEN, port, synth_image

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import emu  # noqa: E402

CSS = {(2, 2): 3, (1, 1): 0, (2, 1): 2, (1, 2): 4}


@pytest.mark.parametrize("samp", list(CSS))
def test_emulated_encoder_byte_exact(samp):
    for (w, h) in [(64, 48), (227, 149), (17, 9), (8, 8), (1, 1), (33, 65), (100, 31), (1100, 40)]:
        img = synth_image(7 + w, w, h)
        for q, ri in [(90, 0), (75, 3), (100, 0), (5, 2), (95, 1)]:
            assert emu.encode(img, q, CSS[samp], ri) == port.encode(img, q, samp, ri), (w, h, samp, q, ri)


def test_emulated_encoder_gray_planar_noise():
    img = synth_image(3, 131, 77)
    gray = np.ascontiguousarray(img[:, :, 1])
    for q, ri in [(90, 0), (50, 5)]:
        assert emu.encode(gray, q, 7, ri) == port.encode(gray, q, (1, 1), ri)
    planar = np.ascontiguousarray(img.transpose(2, 0, 1))
    assert emu.encode(planar, 85, 3, 0, planar=True) == port.encode(img, 85, (2, 2), 0)
    noise = np.random.default_rng(1).integers(0, 256, (95, 130, 3), dtype=np.uint8)
    for samp in CSS:  # quality 100 on noise: the longest codes, ZRL runs and plenty of 0xFF bytes to stuff
        assert emu.encode(noise, 100, CSS[samp], 0) == port.encode(noise, 100, samp, 0)


def test_emulated_encoder_reproduces_libjpeg_turbo_golden_md5():
    import json
    md5 = json.load(open(os.path.join(GOLDEN, "golden.json")))["ljt_md5"]
    raw = open(os.path.join(GOLDEN, "testorig.ppm"), "rb").read()
    img = np.frombuffer(raw.split(b"\n", 3)[3], dtype=np.uint8).reshape(149, 227, 3)
    assert hashlib.md5(emu.encode(img, 75, 3)).hexdigest() == md5["MD5_JPEG_420_ISLOW"]
    assert hashlib.md5(emu.encode(img, 75, 4)).hexdigest() == md5["MD5_JPEG_440_ISLOW"]


@pytest.mark.skipif(not util.have_ref(), reason="oracle/_ref not built")
def test_emulated_encoder_output_decodes_in_reference():
    from util import ref
    img = synth_image(21, 333, 211)
    data = emu.encode(img, 90, 3, 7)
    assert data == ref.ljt_encode(img, 90, (2, 2), 7)
    rgb, st, _ = ref.Session.shared().decoder(options=":fancy_upsampling=1").decode_host([data])
    assert st == [ref.STATUS_SUCCESS] and rgb[0].shape == (211, 333, 3)
