"""CPU tests of the kernel logic: the per-thread bodies of the CUDA kernels (jpeg_math.cuh) + the batch planner run in a
CPU emulation (tests/emu) and must match the oracle bit-for-bit, including the self-synchronising Huffman decode with
Jacobi rounds, restart segments, every sampling mode and output format."""
import os
import sys

import numpy as np
import pytest

import util
from util import GOLDEN, make_jpeg, port, ref

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import emu  # noqa: E402

pytestmark = pytest.mark.skipif(not util.have_ref(), reason="corpus needs oracle/_ref's libjpeg-turbo encoder")

CASES = [(227, 149, (2, 2), 75, 0, False), (64, 48, (1, 1), 90, 0, False), (131, 77, (2, 1), 85, 0, False), (131, 77, (1, 2), 85, 0, False),
         (333, 211, (2, 2), 90, 5, False), (17, 9, (2, 2), 90, 0, False), (3, 5, (2, 2), 90, 0, False), (2, 2, (2, 1), 90, 0, False),
         (1, 1, (1, 1), 90, 0, False), (640, 426, (2, 2), 75, 0, False), (100, 100, (1, 1), 95, 7, False), (500, 375, (2, 2), 75, 1, False),
         (400, 300, (2, 1), 90, 50, False), (201, 99, (1, 1), 85, 0, True), (227, 149, (4, 1), 85, 0, False), (330, 77, (4, 1), 90, 3, False)]


def corpus():
    return [make_jpeg(1000 + i, w, h, samp, q, restart_interval=ri, gray=g) for i, (w, h, samp, q, ri, g) in enumerate(CASES)]


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("fancy", [1, 0])
@pytest.mark.parametrize("subseq", [32, 128, 512])
def test_emulated_pipeline_bit_exact(fancy, subseq, fused):
    datas = corpus()
    dims = [(c[0], c[1]) for c in CASES]
    outs, st, rounds, _ = emu.decode_batch(datas, dims, 4, fancy, subseq, fused=fused)
    assert st == [0] * len(datas)
    for d, o in zip(datas, outs):
        assert np.array_equal(o, port.decode(d, port.FMT_I_RGB, fancy))
    assert rounds < 64


def test_emulated_fused_chain_with_short_halo():
    """k_huff_fused hands the decoder state from CTA to CTA: a CTA whose halo did not deliver the true entry state walks its
    range again.  With the kernel's halo (8 subsequences) that is rare; a halo of one subsequence makes it the rule, and
    the result must still be exact -- with and without restart markers, 4:2:0 (slow to lock on) and 4:4:4."""
    datas = [make_jpeg(7000, 1920, 1080, (2, 2), 90), make_jpeg(7001, 1600, 1200, (1, 1), 85), make_jpeg(7002, 1280, 960, (2, 2), 92, restart_interval=40),
             make_jpeg(7003, 2048, 1536, (2, 1), 75)]
    dims = [(1920, 1080), (1600, 1200), (1280, 960), (2048, 1536)]
    want = [port.decode(d, port.FMT_I_RGB, True) for d in datas]
    redo = {}
    for halo in (1, 2, 0):
        outs, st, rounds, _ = emu.decode_batch(datas, dims, 4, 1, 128, fused=True, halo=halo)
        assert st == [0] * 4
        for o, w in zip(outs, want):
            assert np.array_equal(o, w)
        redo[halo] = emu.decode_batch.last_stats["chain_redo"]
    assert redo[1] > 0 and redo[1] >= redo[2] >= redo[0]
    assert redo[0] <= 1, redo  # the kernel's halo: a mismatch is a once-in-thousands event


def test_emulated_coefficients_match_reference():
    datas = corpus()
    dims = [(c[0], c[1]) for c in CASES]
    slots = [util.mcu_order_coefficients(d).size for d in datas]
    _, st, _, coefs = emu.decode_batch(datas, dims, 4, 1, 128, want_coef=True, slots=slots)
    for d, c in zip(datas, coefs):
        assert np.array_equal(c, util.mcu_order_coefficients(d))


@pytest.mark.parametrize("fmt,pfmt", [(3, port.FMT_P_RGB), (5, port.FMT_P_BGR), (6, port.FMT_I_BGR), (7, port.FMT_Y)])
def test_emulated_output_formats(fmt, pfmt):
    datas = corpus()
    dims = [(c[0], c[1]) for c in CASES]
    outs, st, _, _ = emu.decode_batch(datas, dims, fmt)
    for d, o in zip(datas, outs):
        assert np.array_equal(np.squeeze(o), np.squeeze(port.decode(d, pfmt, True)))


def test_emulated_1080p_converges_with_worklist_rounds():
    d = make_jpeg(1000, 1920, 1080, (2, 2), 90)
    want = port.decode(d, port.FMT_I_RGB, True)
    for subseq in (128, 256):
        outs, st, rounds, _ = emu.decode_batch([d], [(1920, 1080)], 4, 1, subseq)
        assert st == [0] and np.array_equal(outs[0], want)
        stats = emu.decode_batch.last_stats
        assert rounds <= 40
        # rounds 0 and 1 decode everything; what the work-list rounds re-decode must stay a fraction of one pass
        assert stats["redone"] < stats["subsequences"]


def test_emulated_corrupt_and_unsupported():
    good = make_jpeg(5000, 320, 240)
    prog = make_jpeg(5001, 320, 240, arith=True)  # arithmetic coding: not ours
    truncated = good[:len(good) // 2]
    bad_rst = bytearray(make_jpeg(5002, 320, 240, restart_interval=4))
    i = bad_rst.find(b"\xff\xd1")
    bad_rst[i + 1] = 0xD5  # wrong restart marker number
    datas = [good, prog, truncated, b"nope", bytes(bad_rst)]
    outs, st, _, _ = emu.decode_batch(datas, [(320, 240)] * 5, 4)
    assert st == [0, 3, 4, 2, 4]


def test_golden_fixtures_through_emulation():
    for name in ("testorig.jpg", "testorig_420_islow.jpg", "testorig_440_islow.jpg", "testorig_gray_islow.jpg"):
        d = open(os.path.join(GOLDEN, name), "rb").read()
        outs, st, _, _ = emu.decode_batch([d], [(227, 149)], 4)
        assert st == [0] and np.array_equal(outs[0], port.decode(d, port.FMT_I_RGB, True)), name


def test_emulated_many_short_restart_segments():
    """Thousands of short restart segments: subsequences begin in every possible place relative to a segment end, in
    particular inside the pad bits of a segment whose blocks are already complete (the bits there plus the start of the
    next segment may look like an invalid code; that is padding, not corruption)."""
    datas = [make_jpeg(3020, 1536, 1024, (1, 1), 90, restart_interval=1), make_jpeg(3021, 1280, 960, (2, 2), 92, restart_interval=2),
             make_jpeg(3022, 1024, 1024, (1, 1), 97, restart_interval=3)]
    dims = [(1536, 1024), (1280, 960), (1024, 1024)]
    for subseq in (32, 128):
        outs, st, rounds, _ = emu.decode_batch(datas, dims, 4, 1, subseq)
        assert st == [0, 0, 0]
        for d, o in zip(datas, outs):
            assert np.array_equal(o, port.decode(d, port.FMT_I_RGB, True))


def test_emulated_region_of_interest():
    """Region-of-interest decode: the output is exactly the same pixels of the reference's ROI decode (which equals the
    crop of its full decode), for every sampling mode, fancy and plain upsampling, interleaved and planar outputs."""
    rng = np.random.default_rng(11)
    sess = ref.Session.shared()
    for fancy in (1, 0):
        rdec = sess.decoder(options=":fancy_upsampling=%d" % fancy)
        for (w, h, samp, gray) in [(227, 149, (2, 2), False), (131, 77, (2, 1), False), (131, 77, (1, 2), False), (100, 100, (1, 1), False),
                                   (201, 99, (1, 1), True), (640, 426, (2, 2), False)]:
            d = make_jpeg(77, w, h, samp, 85, gray=gray)
            rois = [(0, 0, w, h), (0, 0, 1, 1), (w - 1, h - 1, 1, 1), (1, 1, w - 2, h - 2), (16, 16, 32, 32), (15, 17, 33, 31)]
            for _ in range(6):
                x, y = int(rng.integers(0, w - 1)), int(rng.integers(0, h - 1))
                rois.append((x, y, int(rng.integers(1, w - x + 1)), int(rng.integers(1, h - y + 1))))
            for fmt, rfmt in ((4, ref.I_RGB), (3, ref.P_RGB)):
                want, rst, _ = rdec.decode_host([d] * len(rois), fmt=rfmt, rois=rois)
                assert all(s == ref.STATUS_SUCCESS for s in rst)
                outs, st, _, _ = emu.decode_batch([d] * len(rois), [(w, h)] * len(rois), fmt, fancy, 128, rois=rois)
                assert st == [0] * len(rois)
                for r, o, wv in zip(rois, outs, want):
                    assert np.array_equal(o, wv), (fancy, w, h, samp, gray, fmt, r)


def test_emulated_extended_sequential_sof1():
    """An 8-bit extended-sequential stream (SOF1) is the baseline process with room for four tables per class: same path."""
    d = bytearray(make_jpeg(55, 211, 133, (2, 2), 85))
    k = d.find(b"\xff\xc0")
    assert k > 0
    d[k + 1] = 0xC1
    d = bytes(d)
    want, st, _ = ref.Session.shared().decoder(options=":fancy_upsampling=1").decode_host([d])
    assert st == [ref.STATUS_SUCCESS]
    outs, st, _, _ = emu.decode_batch([d], [(211, 133)], 4, 1, 128)
    assert st == [0] and np.array_equal(outs[0], want[0])


def test_emulated_randomized_batch():
    """random sizes (1..160 log-uniform), samplings, qualities, restart intervals and content through the emulated pipeline
    in one batch, twice (interleaved RGB with fancy upsampling, planar RGB without): bit-exact against the port"""
    rng = np.random.default_rng(99)
    specs = []
    for _ in range(24):
        gray = rng.random() < 0.15
        specs.append(dict(w=int(round(2 ** rng.uniform(0, np.log2(160)))), h=int(round(2 ** rng.uniform(0, np.log2(160)))),
                          samp=(1, 1) if gray else [(1, 1), (2, 1), (1, 2), (2, 2), (4, 1)][int(rng.integers(0, 5))],
                          quality=int(rng.integers(25, 101)), restart_interval=int(rng.choice([0, 0, 1, 3, 17])), gray=gray,
                          kind=["photo", "noise", "flat"][int(rng.integers(0, 3))]))
    datas = [make_jpeg(7000 + i, **s) for i, s in enumerate(specs)]
    dims = [(s["w"], s["h"]) for s in specs]
    for fmt, pfmt, fancy in ((4, port.FMT_I_RGB, 1), (3, port.FMT_P_RGB, 0)):
        outs, st, rounds, _ = emu.decode_batch(datas, dims, fmt, fancy, 64)
        assert st == [0] * len(datas), [(i, specs[i]) for i, x in enumerate(st) if x]
        for i, (d, o) in enumerate(zip(datas, outs)):
            assert np.array_equal(o, port.decode(d, pfmt, fancy)), (i, specs[i], fmt)


def test_emulated_flat_images_with_optimised_tables():
    """Flat and mostly-flat images with optimised Huffman tables: the end-of-block and zero-DC codes are one bit long, so a
    block is two symbols in two bits -- the record reservation must cover one record per bit (such files came
    back as CORRUPT_STREAM)."""
    rng = np.random.default_rng(5)
    flat = np.full((512, 512, 3), 200, np.uint8)
    mostly = flat.copy()
    mostly[:48] = rng.integers(0, 255, (48, 512, 3))
    gray = np.full((256, 512), 17, np.uint8)
    datas, dims = [], []
    for img in (flat, mostly, gray):
        for samp in ((1, 1), (2, 2)):
            if img.ndim == 2 and samp != (1, 1):
                continue
            datas.append(ref.ljt_encode(img, 90, samp, optimize=True))
            dims.append((img.shape[1], img.shape[0]))
    outs, st, rounds, _ = emu.decode_batch(datas, dims, 4, 1, 128)
    assert st == [0] * len(datas)
    for d, o in zip(datas, outs):
        assert np.array_equal(o, port.decode(d, port.FMT_I_RGB, True))


def test_emulated_periodic_streams_converge():
    """A flat image is a periodic bit stream: a decoder that starts at the wrong phase never locks on, so the repair has
    to travel the whole image.  The run-ahead rounds walk it (without decoding where the period divides the subsequence
    length) instead of spending one work-list round per subsequence (4096x4096 hit the old cap of 512 rounds)."""
    flat = np.full((2048, 2048, 3), 200, np.uint8)
    half = flat.copy()
    half[:256] = np.random.default_rng(6).integers(0, 255, (256, 2048, 3))
    for img, samp, opt in ((flat, (2, 2), False), (half, (2, 2), False), (flat, (1, 1), True), (half, (2, 1), True)):
        d = ref.ljt_encode(img, 90, samp, optimize=opt)
        for fused in (False, True):  # k_huff_fused: its rounds from kFusedRunHeads on walk the runs the same way inside the CTA
            outs, st, rounds, _ = emu.decode_batch([d], [(2048, 2048)], 4, 1, 128, fused=fused)
            assert st == [0] and np.array_equal(outs[0], port.decode(d, port.FMT_I_RGB, True))
            assert rounds <= 2 + 4 + 8, rounds  # Jacobi rounds, then a few run-ahead rounds
    d = ref.ljt_encode(flat, 90, (2, 2))
    emu.decode_batch([d], [(2048, 2048)], 4, 1, 128)
    assert emu.decode_batch.last_stats["skipped"] > 100  # the periodic stretch was walked without decoding it


def test_emulated_data_after_first_eoi():
    """Files that carry more data after the image's EOI -- a second JPEG (MPO, gain maps, motion photos), a thumbnail, stray
    bytes with an FFD9 in them: the scan ends at the first marker, like for the reference (the parser took the
    last EOI of the file and the image came back corrupt)."""
    a = make_jpeg(91, 320, 200, (2, 2), 85)
    b2 = make_jpeg(92, 64, 64, (1, 1), 85)
    c = make_jpeg(93, 211, 97, (2, 1), 90, restart_interval=3)
    datas = [a + b"\x00" * 37 + b2, a + b2, c + b"\xff\xd9\x12\x34\xff\xd9", a + b"\xff\xe1\x00\x08abcdef\xff\xd9"]
    dims = [(320, 200), (320, 200), (211, 97), (320, 200)]
    wants = [port.decode(x, port.FMT_I_RGB, True) for x in (a, a, c, a)]
    rwant, rst, _ = ref.Session.shared().decoder(options=":fancy_upsampling=1").decode_host(datas)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    outs, st, _, _ = emu.decode_batch(datas, dims, 4, 1, 128)
    assert st == [0] * len(datas)
    for o, w, rw in zip(outs, wants, rwant):
        assert np.array_equal(o, w) and np.array_equal(o, rw)


def test_emulated_exif_orientation():
    """the eight EXIF orientations applied while the pixels are written (f3): equal to the reference's decode rotated /
    mirrored with numpy; EXIF 5 must come out as the transpose and EXIF 7 as the anti-transpose"""
    for (w, h, samp, gray) in [(227, 149, (2, 2), False), (131, 77, (2, 1), False), (100, 64, (1, 1), False), (201, 99, (1, 1), True)]:
        d = make_jpeg(31, w, h, samp, 85, gray=gray)
        want = port.decode(d, port.FMT_I_RGB, True)
        n = len(util.EXIF_ORIENTATIONS)
        outs, st, _, _ = emu.decode_batch([d] * n, [(w, h)] * n, 4, 1, 128, orients=util.EXIF_ORIENTATIONS)
        assert st == [0] * n
        for o, got in zip(util.EXIF_ORIENTATIONS, outs):
            assert np.array_equal(got, util.apply_orientation(want, *o)), (w, h, o)
        assert np.array_equal(outs[4], want.transpose(1, 0, 2)) and np.array_equal(outs[6], want[::-1, ::-1].transpose(1, 0, 2))
        pl, st, _, _ = emu.decode_batch([d] * n, [(w, h)] * n, 3, 1, 128, orients=util.EXIF_ORIENTATIONS)
        for o, got in zip(util.EXIF_ORIENTATIONS, pl):
            assert np.array_equal(got.transpose(1, 2, 0), util.apply_orientation(want, *o)), ("planar", w, h, o)


PROG_CASES = [(64, 48, (2, 2), False, 0), (227, 149, (2, 2), False, 0), (131, 77, (2, 1), False, 0), (131, 77, (1, 2), False, 0), (100, 100, (1, 1), False, 0),
              (201, 99, (1, 1), True, 0), (333, 211, (2, 2), False, 5), (17, 9, (2, 2), False, 0), (1, 1, (1, 1), False, 0), (330, 77, (4, 1), False, 3),
              (227, 149, (4, 2), False, 0), (500, 375, (2, 2), False, 0)]


def test_emulated_progressive_streams():
    """f2: progressive streams (SOF2: spectral selection + successive approximation, DC/AC first and refinement scans, EOB
    runs, restart intervals) and 4:1:0 sampling, decoded by the body of k_prog_decode (progressive.cuh) into dense
    coefficients and reconstructed by the usual kernels: bit-exact against the compiled reference, pixels and coefficients"""
    import nvimagecodec_b200 as nb
    for fancy in (1, 0):
        rdec = ref.Session.shared().decoder(options=":fancy_upsampling=%d" % fancy)
        datas = [make_jpeg(70 + i, w, h, samp, 85, gray=g, progressive=True, restart_interval=ri) for i, (w, h, samp, g, ri) in enumerate(PROG_CASES)]
        datas.append(make_jpeg(99, 227, 149, (4, 2), 90))  # 4:1:0 baseline
        dims = [(c[0], c[1]) for c in PROG_CASES] + [(227, 149)]
        assert all(nb.parse_header(d).supported for d in datas)
        want, rst, _ = rdec.decode_host(datas)
        assert all(s == ref.STATUS_SUCCESS for s in rst)
        slots = [util.mcu_order_coefficients(d).size for d in datas]
        outs, st, _, coefs = emu.decode_batch(datas, dims, 4, fancy, 128, want_coef=True, slots=slots)
        assert st == [0] * len(datas), st
        for i, (o, w) in enumerate(zip(outs, want)):
            assert np.array_equal(o, w), (fancy, i)
        for d, c in zip(datas, coefs):
            assert np.array_equal(c, util.mcu_order_coefficients(d))


def test_emulated_progressive_damage_is_contained():
    """a progressive stream cut short or damaged comes back failed (the framework then hands it to the CPU decoder), never
    hangs or writes outside its coefficients"""
    good = make_jpeg(5, 160, 120, (2, 2), 85, progressive=True)
    rng = np.random.default_rng(3)
    bad = []
    for k in range(12):
        d = bytearray(good)
        lo = d.find(b"\xff\xda") + 14
        if k % 3 == 0:
            d[int(rng.integers(lo, len(d) - 2))] ^= 1 << int(rng.integers(0, 8))
        elif k % 3 == 1:
            p = int(rng.integers(lo, len(d) - 40))
            d[p:p + 16] = bytes(rng.integers(0, 256, 16, dtype=np.uint8))
        else:
            d = d[:int(rng.integers(lo, len(d) - 2))]
        bad.append(bytes(d))
    outs, st, _, _ = emu.decode_batch([good] + bad + [good], [(160, 120)] * (len(bad) + 2), 4, 1, 128)
    assert st[0] == 0 and st[-1] == 0 and np.array_equal(outs[0], outs[-1])
    assert all(s in (0, 2, 3, 4) for s in st), st
