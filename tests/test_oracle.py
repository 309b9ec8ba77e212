"""CPU tests: pin the oracle.  (a) plain-C port vs libjpeg-turbo's own golden MD5s and the committed fixtures,
(b) port vs the compiled reference (when oracle/_ref is present) on a seeded corpus incl. coefficients."""
import hashlib
import io
import json
import os

import numpy as np
import pytest

import util
from util import GOLDEN, make_jpeg, port, ref, synth_image

G = json.load(open(os.path.join(GOLDEN, "golden.json")))


def _fixture(name):
    return open(os.path.join(GOLDEN, name), "rb").read()


def test_port_reproduces_libjpeg_turbo_known_answers():
    md5 = G["ljt_md5"]
    j440, j420, jgray = _fixture("testorig_440_islow.jpg"), _fixture("testorig_420_islow.jpg"), _fixture("testorig_gray_islow.jpg")
    assert hashlib.md5(j440).hexdigest() == md5["MD5_JPEG_440_ISLOW"]
    assert hashlib.md5(j420).hexdigest() == md5["MD5_JPEG_420_ISLOW"]
    assert hashlib.md5(jgray).hexdigest() == md5["MD5_JPEG_GRAY_ISLOW"]
    rgb = port.decode(j440, port.FMT_I_RGB, True)
    assert hashlib.md5(b"P6\n227 149\n255\n" + rgb.tobytes()).hexdigest() == md5["MD5_PPM_440_ISLOW"]
    y = port.decode(jgray, port.FMT_Y, True)
    assert hashlib.md5(b"P5\n227 149\n255\n" + y.tobytes()).hexdigest() == md5["MD5_PPM_GRAY_ISLOW"]
    rgb = port.decode(jgray, port.FMT_I_RGB, True)
    assert hashlib.md5(b"P6\n227 149\n255\n" + rgb.tobytes()).hexdigest() == md5["MD5_PPM_GRAY_ISLOW_RGB"]


def test_port_matches_reference_plugin_fixture_hashes():
    for name, want in G["decoded_md5"].items():
        rgb = port.decode(_fixture(name), port.FMT_I_RGB, True)
        assert hashlib.md5(b"P6\n227 149\n255\n" + rgb.tobytes()).hexdigest() == want["ppm_rgb"], name


@pytest.mark.skipif(not util.have_ref(), reason="oracle/_ref not built")
def test_synthetic_golden_hashes_and_port_vs_reference():
    d1 = ref.Session.shared().decoder(options=":fancy_upsampling=1")
    d0 = ref.Session.shared().decoder(options=":fancy_upsampling=0")
    for e in G["synthetic_sha256"]:
        data = make_jpeg(e["seed"], e["w"], e["h"], tuple(e["samp"]), e["quality"], restart_interval=e["restart_interval"], gray=e["gray"])
        assert hashlib.sha256(data).hexdigest() == e["jpeg_sha256"], "corpus generator / libjpeg-turbo encoder drifted"
        for fancy, dec in ((1, d1), (0, d0)):
            rgb, st, _ = dec.decode_host([data])
            assert st == [ref.STATUS_SUCCESS]
            assert hashlib.sha256(rgb[0].tobytes()).hexdigest() == e["rgb_sha256_fancy%d" % fancy]
            assert np.array_equal(port.decode(data, port.FMT_I_RGB, fancy), rgb[0])
        ncomp = 1 if e["gray"] else 3
        for c in range(ncomp):
            pc = port.coefficients(data, c)
            assert np.array_equal(pc, ref.ljt_coefficients(data, c, pc.shape[1], pc.shape[0]))


@pytest.mark.skipif(not util.have_ref(), reason="oracle/_ref not built")
def test_port_output_formats_vs_reference():
    d1 = ref.Session.shared().decoder(options=":fancy_upsampling=1")
    color = make_jpeg(77, 120, 90, (2, 2), 90)
    gray = make_jpeg(78, 99, 55, gray=True)
    for pf, rf in ((port.FMT_P_RGB, ref.P_RGB), (port.FMT_I_BGR, ref.I_BGR), (port.FMT_P_BGR, ref.P_BGR), (port.FMT_Y, ref.P_Y)):
        for data in (color, gray):
            o, st, _ = d1.decode_host([data], fmt=rf)
            assert st == [ref.STATUS_SUCCESS]
            assert np.array_equal(np.squeeze(o[0]), np.squeeze(port.decode(data, pf, True)))


@pytest.mark.skipif(not util.have_ref(), reason="oracle/_ref not built")
def test_pil_encoder_is_byte_identical_to_reference_libjpeg_turbo():
    """bench.py builds its corpus with PIL's bundled libjpeg-turbo; it must be the same bitstream the reference's makes"""
    from PIL import Image
    img = synth_image(5, 320, 240)
    for ss, samp in (("4:2:0", (2, 2)), ("4:2:2", (2, 1)), ("4:4:4", (1, 1))):
        b = io.BytesIO()
        Image.fromarray(img).save(b, "JPEG", quality=90, subsampling=ss, optimize=False)
        assert b.getvalue() == ref.ljt_encode(img, 90, samp)


def test_idct_stage_helper_matches_full_decode():
    data = _fixture("testorig_420_islow.jpg")
    coef = port.coefficients(data, 0)
    plane = port.plane(data, 0)
    import ctypes as C
    # luma quant table of quality 75 (jcparam.c std table scaled by 50%)
    base = [16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22,
            37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99]
    q = np.array([max(1, min(255, (b * 50 + 50) // 100)) for b in base], np.uint16)
    out = np.zeros((8, 8), np.uint8)
    for by, bx in ((0, 0), (3, 5), (10, 20)):
        blk = np.ascontiguousarray(coef[by, bx])
        port.lib().jo_idct_islow(blk.ctypes.data, q.ctypes.data, out.ctypes.data, 8)
        assert np.array_equal(out, plane[by * 8:by * 8 + 8, bx * 8:bx * 8 + 8])


def _testorig_ppm():
    raw = _fixture("testorig.ppm")
    return np.frombuffer(raw.split(b"\n", 3)[3], dtype=np.uint8).reshape(149, 227, 3)


def test_port_encoder_reproduces_libjpeg_turbo_known_answers():
    """cjpeg -dct int on testimages/testorig.ppm (ljt/CMakeLists.txt:1273-1345): the encode mirror must emit these files"""
    md5 = G["ljt_md5"]
    img = _testorig_ppm()
    assert hashlib.md5(port.encode(img, 75, (2, 2))).hexdigest() == md5["MD5_JPEG_420_ISLOW"]
    assert hashlib.md5(port.encode(img, 75, (1, 2))).hexdigest() == md5["MD5_JPEG_440_ISLOW"]
    assert port.encode(img, 75, (2, 2)) == _fixture("testorig_420_islow.jpg")
    # cjpeg -gray: luma of the RGB->YCC conversion (ljt/jccolor.c rgb_gray_convert), then a one-component stream
    r, g, b = (img[:, :, i].astype(np.int64) for i in range(3))
    y = ((19595 * r + 38470 * g + 7471 * b + 32768) >> 16).astype(np.uint8)
    assert hashlib.md5(port.encode(y, 75, (1, 1))).hexdigest() == md5["MD5_JPEG_GRAY_ISLOW"]


@pytest.mark.skipif(not util.have_ref(), reason="oracle/_ref not built")
def test_port_encoder_vs_reference_libjpeg_turbo():
    rng = np.random.default_rng(5)
    for (w, h) in [(64, 48), (227, 149), (17, 9), (8, 8), (1, 1), (33, 65), (100, 31)]:
        img = synth_image(7 + w, w, h)
        for samp in [(2, 2), (1, 1), (2, 1), (1, 2)]:
            for q, ri in [(90, 0), (75, 3), (100, 0), (5, 2)]:
                assert port.encode(img, q, samp, ri) == ref.ljt_encode(img, q, samp, ri), (w, h, samp, q, ri)
        gray = np.ascontiguousarray(img[:, :, 1])
        for q, ri in [(90, 0), (50, 5)]:
            assert port.encode(gray, q, (1, 1), ri) == ref.ljt_encode(gray, q, (1, 1), ri)
    noise = rng.integers(0, 256, (95, 130, 3), dtype=np.uint8)
    for samp in [(2, 2), (1, 1), (2, 1)]:
        for q in [100, 50]:
            assert port.encode(noise, q, samp, 0) == ref.ljt_encode(noise, q, samp, 0)
