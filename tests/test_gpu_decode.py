"""GPU parity tests proper: the CUDA path (through the C ABI) against the compiled reference (oracle/_ref) and the
plain-C port, bit-exact on RGB bytes, DCT coefficients and intermediate planes."""
import numpy as np
import pytest

import util
from util import make_jpeg, port, ref

pytestmark = pytest.mark.gpu

FMT_REF = {4: ref.I_RGB, 6: ref.I_BGR, 3: ref.P_RGB, 5: ref.P_BGR, 7: ref.P_Y}


@pytest.fixture(scope="module")
def torch_mod():
    import torch
    return torch


@pytest.fixture(scope="module")
def refdec():
    s = ref.Session.shared()
    return {1: s.decoder(options=":fancy_upsampling=1"), 0: s.decoder(options=":fancy_upsampling=0")}


def gpu_decode(torch, datas, fmt=4, options=":fancy_upsampling=1", pad_stride=0, want_decoder=False):
    from nvimagecodec_b200.decoder import Decoder
    dec = Decoder(0, options=options + " :dump_coef=1")  # keep dense coefficients for the debug tap
    outs, ptrs, strides = [], [], []
    for d in datas:
        i = port.info(d)
        w, h = i["width"], i["height"]
        if fmt in (4, 6):
            stride = w * 3 + pad_stride
            t = torch.zeros((h, stride), dtype=torch.uint8, device="cuda")
        elif fmt == 7:
            stride = w + pad_stride
            t = torch.zeros((h, stride), dtype=torch.uint8, device="cuda")
        else:
            stride = w + pad_stride
            t = torch.zeros((3 * h, stride), dtype=torch.uint8, device="cuda")
        outs.append(t)
        ptrs.append(t.data_ptr())
        strides.append(stride)
    st = dec.decode_into(datas, ptrs, strides, [fmt] * len(datas))
    torch.cuda.synchronize()
    res = []
    for d, t in zip(datas, outs):
        i = port.info(d)
        w, h = i["width"], i["height"]
        a = t.cpu().numpy()
        if fmt in (4, 6):
            res.append(a[:, :w * 3].reshape(h, w, 3))
        elif fmt == 7:
            res.append(a[:, :w].reshape(1, h, w))
        else:
            res.append(a[:, :w].reshape(3, h, w))
    if want_decoder:
        return res, st, dec
    dec.close()
    return res, st


def diagnose(dec, datas, idx):
    """stage-wise comparison for image idx of the last batch -> string"""
    d = datas[idx]
    info = port.info(d)
    msgs = []
    hdr = __import__("nvimagecodec_b200").parse_header(d)
    scan = d[hdr.scan_offset:hdr.scan_offset + hdr.scan_bytes]
    clean_ref, marks = util.unstuff_reference(scan)
    clean = dec.debug_fetch(2, idx, 0, len(scan) + 64).tobytes()
    msgs.append("clean stream: %s (len %d vs %d)" % ("OK" if clean == clean_ref else "MISMATCH", len(clean), len(clean_ref)))
    want = util.mcu_order_coefficients(d)
    got = dec.debug_fetch(0, idx, 0, want.size * 2, np.int16)
    if got.size == want.size:
        bad = np.nonzero(got != want)[0]
        msgs.append("coefficients: %d mismatches%s" % (bad.size, "" if not bad.size else " first at slot %d (block %d)" % (bad[0], bad[0] // 64)))
    else:
        msgs.append("coefficients: size %d vs %d" % (got.size, want.size))
    for c in range(info["ncomp"]):
        wp = port.plane(d, c)
        gp = dec.debug_fetch(1, idx, c, wp.size).reshape(wp.shape)
        msgs.append("plane %d: %d mismatches" % (c, int((gp != wp).sum())))
    return "; ".join(msgs)


CASES = [
    # w, h, samp, quality, restart_interval, gray
    (227, 149, (2, 2), 75, 0, False),
    (64, 48, (1, 1), 90, 0, False),
    (131, 77, (2, 1), 85, 0, False),
    (131, 77, (1, 2), 85, 0, False),
    (333, 211, (2, 2), 90, 5, False),
    (17, 9, (2, 2), 90, 0, False),
    (3, 5, (2, 2), 90, 0, False),
    (2, 2, (2, 1), 90, 0, False),
    (1, 1, (1, 1), 90, 0, False),
    (640, 426, (2, 2), 75, 0, False),
    (100, 100, (1, 1), 95, 7, False),
    (500, 375, (2, 2), 75, 1, False),
    (800, 600, (2, 1), 90, 50, False),
    (201, 99, (1, 1), 85, 0, True),
    (227, 149, (4, 1), 85, 0, False),
    (330, 77, (4, 1), 90, 3, False),
    (1920, 1080, (2, 2), 90, 0, False),
    (1280, 720, (2, 2), 50, 0, False),
]


@pytest.mark.parametrize("fancy", [1, 0])
def test_mixed_batch_rgb_bit_exact(torch_mod, refdec, fancy):
    datas = [make_jpeg(1000 + i, w, h, samp, q, restart_interval=ri, gray=g) for i, (w, h, samp, q, ri, g) in enumerate(CASES)]
    got, st, dec = gpu_decode(torch_mod, datas, 4, ":fancy_upsampling=%d" % fancy, want_decoder=True)
    want, rst, _ = refdec[fancy].decode_host(datas)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    problems = []
    for i, (g, w) in enumerate(zip(got, want)):
        if st[i] != 0 or not np.array_equal(g, w):
            problems.append("case %d %s status %d diff %d | %s" % (i, CASES[i], st[i], int((g != w).sum()), diagnose(dec, datas, i)))
    print(dec.stats())
    dec.close()
    assert not problems, "\n".join(problems)


@pytest.mark.parametrize("fmt", [6, 3, 5, 7])
def test_output_formats(torch_mod, refdec, fmt):
    datas = [make_jpeg(2000 + i, w, h, samp, q, restart_interval=ri, gray=g) for i, (w, h, samp, q, ri, g) in enumerate(CASES[:14])]
    got, st = gpu_decode(torch_mod, datas, fmt, pad_stride=13)
    want, rst, _ = refdec[1].decode_host(datas, fmt=FMT_REF[fmt])
    for i, (g, w) in enumerate(zip(got, want)):
        assert st[i] == 0, (i, st[i])
        assert np.array_equal(g, w), (i, CASES[i], fmt)


def test_coefficients_bit_exact(torch_mod):
    datas = [make_jpeg(3000 + i, w, h, samp, q, restart_interval=ri, gray=g) for i, (w, h, samp, q, ri, g) in enumerate(CASES)]  # every case, 4:1:1 included
    got, st, dec = gpu_decode(torch_mod, datas, 4, want_decoder=True)
    for i, d in enumerate(datas):
        want = util.mcu_order_coefficients(d)
        g = dec.debug_fetch(0, i, 0, want.size * 2, np.int16)
        assert np.array_equal(g, want), (i, CASES[i])
    dec.close()


def test_subsequence_sizes_and_rounds(torch_mod, refdec):
    datas = [make_jpeg(4000 + i, 640, 480, (2, 2), 90) for i in range(4)] + [make_jpeg(4100, 1920, 1080, (2, 2), 90)]
    want, _, _ = refdec[1].decode_host(datas)
    for opts in (":fancy_upsampling=1 :subseq_bytes=64", ":fancy_upsampling=1 :subseq_bytes=32",
                 ":fancy_upsampling=1 :subseq_bytes=1024", "b200jpeg_decoder:subseq_bytes=512 other_plugin:subseq_bytes=32",
                 # the work-list rounds every way they can run: all in the cooperative kernel / first ones as ordinary
                 # launches, items dealt thread-major / lane-major, one stream / groups
                 ":tail_launches=0 :tail_spread=0", ":tail_launches=0 :tail_spread=1 :groups=1", ":tail_launches=3 :tail_spread=1",
                 ":tail_launches=32 :tail_spread=0 :subseq_bytes=64", ":tail_launches=1 :profile=1"):
        got, st = gpu_decode(torch_mod, datas, 4, opts)
        for i, (g, w) in enumerate(zip(got, want)):
            assert st[i] == 0 and np.array_equal(g, w), (opts, i)


def test_fused_entropy_kernel(torch_mod, refdec):
    """fused=1: the whole entropy-decode stage as one kernel (k_huff_fused: rounds inside the CTA over a halo, CTAs chained by
    look-back, block pass at the end).  Every sampling mode, restart intervals, tiny images, big images whose streams span
    many CTAs, flat images (periodic streams that never lock on: the chain between CTAs has to repair them) and damaged
    data -- pixels and coefficients as the reference has them."""
    datas = [make_jpeg(1000 + i, w, h, samp, q, restart_interval=ri, gray=g) for i, (w, h, samp, q, ri, g) in enumerate(CASES)]
    flat = np.full((1536, 2048, 3), 131, np.uint8)
    half = flat.copy()
    half[:256] = np.random.default_rng(6).integers(0, 255, (256, 2048, 3))
    datas += [ref.ljt_encode(flat, 90, (1, 1), optimize=True), ref.ljt_encode(half, 90, (2, 2), optimize=False), ref.ljt_encode(half, 90, (2, 1), optimize=True),
              make_jpeg(4200, 3840, 2160, (1, 1), 90), make_jpeg(4201, 3840, 2160, (2, 2), 90, restart_interval=240)]
    want, _, _ = refdec[1].decode_host(datas)
    for opts in (":fancy_upsampling=1 :fused=1", ":fancy_upsampling=1 :fused=1 :groups=1", ":fancy_upsampling=1 :fused=1 :subseq_bytes=64",
                 ":fancy_upsampling=1 :fused=1 :subseq_bytes=256 :profile=1"):
        got, st, dec = gpu_decode(torch_mod, datas, 4, opts, want_decoder=True)
        for i, (g, w) in enumerate(zip(got, want)):
            assert st[i] == 0 and np.array_equal(g, w), (opts, i, diagnose(dec, datas, i))
        assert dec.stats()["launches"] < 12 * 8  # one Huffman launch per group instead of six
        for i in (0, 4, 16):
            wantc = util.mcu_order_coefficients(datas[i])
            assert np.array_equal(dec.debug_fetch(0, i, 0, wantc.size * 2, np.int16), wantc), (opts, i)
        dec.close()
    # damaged data: contained exactly as by the separate kernels (same statuses, good images untouched)
    good = make_jpeg(4300, 640, 480, (2, 2), 90)
    rng = np.random.default_rng(3)
    bad = bytearray(make_jpeg(4301, 1280, 960, (2, 2), 90))
    for pos in rng.integers(len(bad) // 2, len(bad) - 4, 40):
        bad[pos] ^= 1 << int(rng.integers(0, 8))
    batch = [good, bytes(bad), good[:len(good) // 2], good]
    a, sa = gpu_decode(torch_mod, batch, 4, ":fancy_upsampling=1 :fused=1")
    b, sb = gpu_decode(torch_mod, batch, 4, ":fancy_upsampling=1 :fused=0")
    assert sa[0] == 0 and sa[3] == 0 and list(sa) == list(sb)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[3], b[3])


def test_bad_and_unsupported_streams(torch_mod):
    good = make_jpeg(5000, 320, 240)
    prog = make_jpeg(5001, 320, 240, progressive=True)
    arith = make_jpeg(5002, 320, 240, arith=True)
    truncated = good[:len(good) // 2]
    garbage = b"not a jpeg at all"
    datas = [good, prog, arith, truncated, garbage, good]
    from nvimagecodec_b200.decoder import Decoder
    dec = Decoder(0, options=":fancy_upsampling=1")
    outs = [torch_mod.zeros((240, 960), dtype=torch_mod.uint8, device="cuda") for _ in datas]
    st = dec.decode_into(datas, [o.data_ptr() for o in outs], [960] * len(datas), [4] * len(datas))
    assert st[0] == 0 and st[5] == 0
    assert st[1] == 0                      # complete progressive streams are decoded (f2)
    want = ref.Session.shared().decoder(options=":fancy_upsampling=1").decode_host([prog])[0][0]
    assert np.array_equal(outs[1].cpu().numpy().reshape(240, 320, 3), want)
    assert st[2] == 3                      # UNSUPPORTED -> the framework falls back to the CPU decoder
    assert st[3] == 4                      # CORRUPT_STREAM (entropy data ends early)
    assert st[4] == 2                      # BAD_JPEG
    assert torch_mod.equal(outs[0], outs[5])
    dec.close()


def test_host_output_buffers(torch_mod, refdec):
    datas = [make_jpeg(6000 + i, 400 + 3 * i, 300 + i, (2, 2), 80) for i in range(5)]
    from nvimagecodec_b200.decoder import Decoder
    dec = Decoder(0, options=":fancy_upsampling=1")
    outs = [np.zeros((300 + i, (400 + 3 * i) * 3), np.uint8) for i in range(5)]
    st = dec.decode_into(datas, [o.ctypes.data for o in outs], [o.shape[1] for o in outs], [4] * 5, on_host=True)
    want, _, _ = refdec[1].decode_host(datas)
    for i in range(5):
        assert st[i] == 0
        assert np.array_equal(outs[i].reshape(want[i].shape), want[i])
    dec.close()


def test_python_mirror_api(torch_mod, refdec):
    from nvimagecodec_b200 import Decoder
    datas = [make_jpeg(7000 + i, 320, 200 + i) for i in range(3)]
    dec = Decoder(device_id=0)  # reference Python default: ":fancy_upsampling=0"
    imgs = dec.decode(datas)
    want, _, _ = refdec[0].decode_host(datas)
    for im, w in zip(imgs, want):
        assert im is not None and im.shape == w.shape
        assert np.array_equal(im.cpu(), w)
        assert im.__cuda_array_interface__["shape"] == w.shape
    one = dec.decode(datas[0])
    assert np.array_equal(one.cpu(), want[0])
    assert dec.decode(b"garbage") is None


def test_region_of_interest(torch_mod, refdec):
    """ROI decode through the C ABI: device and host outputs, interleaved / planar / Y, regions that start and end inside
    the 512-pixel tiles of the colour kernel; bit-exact against the reference's ROI decode; invalid regions are refused."""
    torch = torch_mod
    from nvimagecodec_b200.decoder import Decoder
    rng = np.random.default_rng(12)
    dec = Decoder(0, options=":fancy_upsampling=1")
    for (w, h, samp, gray) in [(1300, 420, (2, 2), False), (1111, 333, (2, 1), False), (700, 300, (1, 2), False), (1030, 200, (1, 1), False),
                               (900, 260, (1, 1), True)]:
        d = make_jpeg(99, w, h, samp, 88, gray=gray)
        rois = [(0, 0, w, h), (w - 1, h - 1, 1, 1), (511, 0, 2, h), (512, 7, 513, 100), (1, 1, w - 2, h - 2), (3, 5, 1021, 17)]
        for _ in range(6):
            x, y = int(rng.integers(0, w - 1)), int(rng.integers(0, h - 1))
            rois.append((x, y, int(rng.integers(1, w - x + 1)), int(rng.integers(1, h - y + 1))))
        rois = [(x, y, min(rw, w - x), min(rh, h - y)) for (x, y, rw, rh) in rois]
        n = len(rois)
        for fmt in (4, 3, 7):
            want, rst, _ = refdec[1].decode_host([d] * n, fmt=FMT_REF[fmt], rois=rois)
            assert all(s == ref.STATUS_SUCCESS for s in rst)
            for on_host in (False, True):
                outs, ptrs, strides = [], [], []
                for (x, y, rw, rh) in rois:
                    stride = rw * 3 + 5 if fmt == 4 else rw + 3
                    rows = rh if fmt in (4, 7) else 3 * rh
                    t = torch.zeros((rows, stride), dtype=torch.uint8, device="cpu" if on_host else "cuda")
                    outs.append(t.pin_memory() if on_host else t)
                    ptrs.append(outs[-1].data_ptr())
                    strides.append(stride)
                st = dec.decode_into([d] * n, ptrs, strides, [fmt] * n, on_host, rois=rois)
                torch.cuda.synchronize()
                assert st == [0] * n
                for (x, y, rw, rh), t, wv in zip(rois, outs, want):
                    a = t.cpu().numpy()
                    got = a[:, :rw * 3].reshape(rh, rw, 3) if fmt == 4 else a[:, :rw].reshape(1 if fmt == 7 else 3, rh, rw)
                    assert np.array_equal(got, wv), (w, h, samp, gray, fmt, on_host, (x, y, rw, rh))
                    assert not a[:, (rw * 3 if fmt == 4 else rw):].any(), ("wrote past the region's row", fmt, on_host, (x, y, rw, rh))
    d = make_jpeg(5, 64, 48, (2, 2), 90)
    t = torch.zeros((48, 64 * 3), dtype=torch.uint8, device="cuda")
    bad = [(60, 0, 8, 8), (0, 40, 8, 16), (-1, 0, 4, 4), (0, 0, 0, 5)]
    st = dec.decode_into([d] * len(bad), [t.data_ptr()] * len(bad), [64 * 3] * len(bad), [4] * len(bad), False, rois=bad)
    assert all(s != 0 for s in st[:3]), st
    dec.close()


def test_custom_allocators(torch_mod, refdec):
    """b200jpeg_decoder_set_allocators / b200jpeg_encoder_set_allocators (the framework's device / pinned allocator
    callbacks): every device allocation goes through the callbacks, frees pass the allocated size back, results unchanged"""
    import ctypes as C
    torch = torch_mod
    from nvimagecodec_b200 import _capi
    from nvimagecodec_b200.decoder import Decoder
    from nvimagecodec_b200.encoder import Encoder, EncodeParams
    live, log = {}, {"malloc": 0, "free": 0, "bad_free": 0}

    def dev_malloc(ctx, out, size, stream):
        p = torch.cuda.caching_allocator_alloc(size)
        live[p] = size
        out[0] = p
        log["malloc"] += 1
        return 0

    def dev_free(ctx, p, size, stream):
        if live.pop(p, None) != size:
            log["bad_free"] += 1
        torch.cuda.caching_allocator_delete(p)
        log["free"] += 1
        return 0

    al = _capi.Allocators()
    al.device_malloc, al.device_free, al.device_padding = _capi.MALLOC_FN(dev_malloc), _capi.FREE_FN(dev_free), 512
    datas = [make_jpeg(31 + i, 320, 200, samp, 90) for i, samp in enumerate([(2, 2), (1, 1)])]
    dec = Decoder(0, options=":fancy_upsampling=1")
    assert _capi.lib().b200jpeg_decoder_set_allocators(dec._h, C.byref(al)) == 0
    imgs = dec.decode(datas)
    want, _, _ = refdec[1].decode_host(datas)
    for im, w in zip(imgs, want):
        assert np.array_equal(im.cpu(), w)
    assert log["malloc"] > 5 and all(s % 512 == 0 for s in live.values())
    assert _capi.lib().b200jpeg_decoder_set_allocators(dec._h, C.byref(al)) != 0  # too late: buffers exist
    dec.close()
    enc = Encoder(0)
    assert _capi.lib().b200jpeg_encoder_set_allocators(enc._h, C.byref(al)) == 0
    px = util.synth_image(3, 200, 120)
    assert enc.encode(px, ".jpg", EncodeParams(quality=90, chroma_subsampling="4:2:0")) == port.encode(px, 90, (2, 2))
    enc.close()
    torch.cuda.synchronize()
    assert not live and log["free"] == log["malloc"] and log["bad_free"] == 0, (live, log)


def _corruptions(rng, good, hdr):
    """damaged copies of a good stream: bit flips, byte runs, inserted markers, truncation -- entropy-coded part only"""
    lo, hi = hdr.scan_offset, hdr.scan_offset + hdr.scan_bytes
    out = []
    for k in range(40):
        d = bytearray(good)
        kind = k % 5
        if kind == 0:
            for _ in range(int(rng.integers(1, 6))):
                d[int(rng.integers(lo, hi))] ^= 1 << int(rng.integers(0, 8))
        elif kind == 1:
            p = int(rng.integers(lo, hi - 40))
            d[p:p + 32] = bytes(rng.integers(0, 256, 32, dtype=np.uint8))
        elif kind == 2:
            p = int(rng.integers(lo, hi - 2))
            d[p:p + 2] = bytes([0xFF, int(rng.integers(0xC0, 0xFF))])
        elif kind == 3:
            d = d[:int(rng.integers(lo + 1, hi))] + b"\xff\xd9"
        else:
            p = int(rng.integers(lo, hi - 300))
            d[p:p + 256] = b"\xff" * 256 if k % 2 else b"\x00" * 256
        out.append(bytes(d))
    return out


@pytest.mark.timeout(300)
def test_damaged_streams_are_contained(torch_mod, refdec):
    """Damaged entropy-coded data must neither hang nor disturb the good images of the same batch; every damaged image
    comes back as decoded-with-garbage (status 0, like libjpeg's warnings) or CORRUPT_STREAM, never anything else."""
    torch = torch_mod
    import nvimagecodec_b200 as nb
    from nvimagecodec_b200.decoder import Decoder
    rng = np.random.default_rng(77)
    dec = Decoder(0, options=":fancy_upsampling=1")
    for (w, h, samp, ri) in [(320, 240, (2, 2), 0), (333, 211, (1, 1), 4), (640, 200, (2, 1), 0)]:
        good = make_jpeg(6000 + w, w, h, samp, 88, restart_interval=ri)
        bad = _corruptions(rng, good, nb.parse_header(good))
        datas = [good] + bad + [good]
        outs = [torch.zeros((h + 2, w * 3), dtype=torch.uint8, device="cuda") for _ in datas]  # two guard rows per image
        st = dec.decode_into(datas, [o.data_ptr() for o in outs], [w * 3] * len(datas), [4] * len(datas))
        torch.cuda.synchronize()
        assert st[0] == 0 and st[-1] == 0
        assert all(s in (0, 2, 4) for s in st), st
        want = refdec[1].decode_host([good])[0][0]
        for k in (0, len(datas) - 1):
            assert np.array_equal(outs[k][:h].cpu().numpy().reshape(h, w, 3), want)
        for o in outs:
            assert not o[h:].any(), "wrote past the image"
    dec.close()


def _random_specs(rng, n, max_side):
    specs = []
    for _ in range(n):
        w = int(round(2 ** rng.uniform(0, np.log2(max_side))))
        h = int(round(2 ** rng.uniform(0, np.log2(max_side))))
        gray = rng.random() < 0.15
        samp = [(1, 1), (2, 1), (1, 2), (2, 2), (4, 1)][int(rng.integers(0, 5))]
        ri = int(rng.choice([0, 0, 0, 1, 2, 5, 11, 64]))
        specs.append(dict(w=w, h=h, samp=(1, 1) if gray else samp, quality=int(rng.integers(25, 101)), restart_interval=ri, gray=gray,
                          kind=["photo", "photo", "noise", "flat"][int(rng.integers(0, 4))]))
    return specs


@pytest.mark.parametrize("fancy", [1, 0])
def test_randomized_batches_bit_exact(torch_mod, refdec, fancy):
    """Three batches of 64 random images (sizes 1..700 log-uniform, every supported sampling, qualities 25..100, restart
    intervals, photo / noise / flat content, a random output format and row slack per image) through ONE decoder object,
    so that every buffer is reused with different contents: every image bit-exact against the reference decoder."""
    from nvimagecodec_b200.decoder import Decoder
    torch = torch_mod
    rng = np.random.default_rng(77 + fancy)
    dec = Decoder(0, options=":fancy_upsampling=%d" % fancy)
    for batch in range(3):
        specs = _random_specs(rng, 64, 700)
        datas = [make_jpeg(5000 + 100 * batch + i, **s) for i, s in enumerate(specs)]
        fmts, outs, ptrs, strides = [], [], [], []
        for s in specs:
            f = int(rng.choice([7, 4, 6])) if s["gray"] else int(rng.choice([4, 6, 3, 5, 7]))
            pad = int(rng.choice([0, 0, 1, 13, 64]))
            row = s["w"] * (3 if f in (4, 6) else 1) + pad
            rows = s["h"] * (1 if f in (4, 6, 7) else 3)
            t = torch.full((rows, row), 0xA5, dtype=torch.uint8, device="cuda")
            fmts.append(f)
            outs.append(t)
            ptrs.append(t.data_ptr())
            strides.append(row)
        st = dec.decode_into(datas, ptrs, strides, fmts)
        torch.cuda.synchronize()
        assert all(x == 0 for x in st), (batch, [(i, specs[i]) for i, x in enumerate(st) if x])
        want, rst, _ = refdec[fancy].decode_host(datas, fmt=[FMT_REF[f] for f in fmts])
        assert all(x == ref.STATUS_SUCCESS for x in rst)
        for i, s in enumerate(specs):
            a = outs[i].cpu().numpy()
            f, w, h = fmts[i], s["w"], s["h"]
            used = w * (3 if f in (4, 6) else 1)
            got = a[:, :used].reshape(h, w, 3) if f in (4, 6) else a[:, :used].reshape(1 if f == 7 else 3, h, w)
            assert np.array_equal(got, want[i]), "batch %d image %d %s fmt %d differs from the reference decoder" % (batch, i, s, f)
            assert (a[:, used:] == 0xA5).all(), "batch %d image %d %s fmt %d: wrote into the row slack" % (batch, i, s, f)
    dec.close()
