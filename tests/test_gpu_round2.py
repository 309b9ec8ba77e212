"""Round-2 GPU parity tests: every accepted stream class and every north-star configuration at its full size, the plugin's
canDecode matrix and buffer kinds through the unmodified reference framework, per-image CUDA streams, two batches in flight,
the files ADVICE r1 found broken, and the multi-device decoder.  Everything is compared bit for bit with the compiled
reference (oracle/_ref: nvimgcodecDecoderDecode -> libjpeg_turbo_ext, libjpeg-turbo 3.0.3)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import util
from util import make_jpeg, port, ref, synth_image

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not util.have_ref(), reason="oracle/_ref not built")]


def _threads():
    return ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1))


@pytest.fixture(scope="module")
def sess():
    import nvimagecodec_b200 as nb
    s = ref.Session(ext_dir=nb.EXT_DIR, with_ljt=True)
    yield s
    s.close()


@pytest.fixture(scope="module")
def cpu():
    return ref.Session.shared().decoder(options=":fancy_upsampling=1")


def _to_sof1(d):
    d = bytearray(d)
    k = d.find(b"\xff\xc0")
    assert k > 0
    d[k + 1] = 0xC1
    return bytes(d)


def _decode_c_abi(datas, dims, fmt=4, options=":fancy_upsampling=1", **kw):
    import torch
    from nvimagecodec_b200.decoder import Decoder
    dec = Decoder(0, options=options)
    outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
    st = dec.decode_into(datas, [o.data_ptr() for o in outs], [w * 3 for (w, h) in dims], [fmt] * len(datas), **kw)
    torch.cuda.synchronize()
    stats = dec.stats()
    dec.close()
    return [o.cpu().numpy().reshape(h, w, 3) for o, (w, h) in zip(outs, dims)], st, stats


def test_extended_sequential_sof1_on_the_gpu(sess, cpu):
    """SOF1 (8-bit extended sequential, Huffman) is accepted by the parser: it must decode on the GPU like SOF0 -- through
    the C ABI and, with a GPU_ONLY backend list, through the framework"""
    import torch
    cases = [(211, 133, (2, 2), 85, 0), (640, 426, (2, 1), 90, 7), (97, 301, (1, 1), 75, 0), (333, 211, (4, 1), 90, 0)]
    datas = [_to_sof1(make_jpeg(55 + i, w, h, samp, q, restart_interval=ri)) for i, (w, h, samp, q, ri) in enumerate(cases)]
    dims = [(c[0], c[1]) for c in cases]
    import nvimagecodec_b200 as nb
    assert all(nb.parse_header(d).sof_marker == 0xC1 and nb.parse_header(d).supported for d in datas)
    want, rst, _ = cpu.decode_host(datas)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    got, st, _ = _decode_c_abi(datas, dims)
    assert st == [0] * len(datas)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    dec = sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=":fancy_upsampling=1")
    outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
    torch.cuda.synchronize()
    fst, _ = dec.decode_raw(datas, [o.data_ptr() for o in outs], dims, [ref.I_RGB] * len(datas), [w * 3 for (w, h) in dims], ref.BUFFER_DEVICE)
    torch.cuda.synchronize()
    assert fst == [ref.STATUS_SUCCESS] * len(datas)
    for o, w in zip(outs, want):
        assert np.array_equal(o.cpu().numpy().reshape(w.shape), w)
    dec.close()


def test_c1_full_batch_64x1080p(sess, cpu):
    """North-star config 1 at its full size: 64 distinct 1920x1080 4:2:0 q90 JPEGs without restart markers -> interleaved RGB;
    every image against the reference, through the C ABI and through nvimgcodecDecoderDecode (device buffers)"""
    import torch
    with _threads() as ex:
        datas = list(ex.map(lambda i: make_jpeg(1000 + i, 1920, 1080, (2, 2), 90), range(64)))
    dims = [(1920, 1080)] * 64
    want, rst, _ = cpu.decode_host(datas, dims=dims)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    got, st, stats = _decode_c_abi(datas, dims)
    assert st == [0] * 64
    for i in range(64):
        assert np.array_equal(got[i], want[i]), "C1 image %d differs from the reference decoder" % i
    assert stats["blocks"] == 64 * 48960
    dec = sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=":fancy_upsampling=1")
    outs = torch.zeros((64, 1080, 5760), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    fst, _ = dec.decode_raw(datas, [outs[i].data_ptr() for i in range(64)], dims, [ref.I_RGB] * 64, [5760] * 64, ref.BUFFER_DEVICE)
    torch.cuda.synchronize()
    assert fst == [ref.STATUS_SUCCESS] * 64
    host = outs.cpu().numpy().reshape(64, 1080, 1920, 3)
    for i in range(64):
        assert np.array_equal(host[i], want[i]), "C1 image %d (through the framework) differs" % i
    dec.close()


def test_c5_full_batch_128x1080p_encode():
    """North-star config 5 at its full size: 128x 1920x1080 RGB -> 4:2:0 q90 baseline, every file byte-exact against
    libjpeg-turbo (16 distinct images, each 8 times: the repeats must equal their first occurrence)"""
    import torch
    from nvimagecodec_b200.encoder import Encoder, EncodeParams
    with _threads() as ex:
        base = list(ex.map(lambda i: synth_image(2000 + i, 1920, 1080), range(16)))
        want = list(ex.map(lambda im: ref.ljt_encode(im, 90, (2, 2)), base))
    dev = [torch.from_numpy(base[i % 16]).cuda() for i in range(128)]
    enc = Encoder(0)
    got = enc.encode(dev, ".jpg", EncodeParams(quality=90, chroma_subsampling="4:2:0"))
    assert len(got) == 128
    for i in range(128):
        assert got[i] == want[i % 16], "C5 image %d differs from libjpeg-turbo" % i
    enc.close()


def _exif_app1(orientation):
    """a minimal EXIF APP1 segment carrying only the orientation tag (0x0112)"""
    tiff = b"II*\x00\x08\x00\x00\x00" + b"\x01\x00" + b"\x12\x01\x03\x00\x01\x00\x00\x00" + bytes([orientation, 0, 0, 0]) + b"\x00\x00\x00\x00"
    body = b"Exif\x00\x00" + tiff
    return b"\xff\xe1" + (len(body) + 2).to_bytes(2, "big") + body


def test_can_decode_matrix_matches_the_cpu_plugin_and_falls_back(sess):
    """Every rejection of the plugin's canDecode (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:118-173 is the model):
    same nvimgcodecProcessingStatus_t as the CPU plugin for the same target description; what b200jpeg declines goes
    down the framework's fallback chain and is decoded all the same."""
    import torch
    import nvimagecodec_b200 as nb
    d = make_jpeg(1, 64, 48)
    # b200jpeg alone in a session of its own: nvimgcodecDecoderCanDecode walks the fallback chain, and the reference
    # dereferences a null decoder when the last worker of the chain has no allowed backend (src/image_generic_decoder.cpp:110
    # after src/decoder_worker.cpp:63-93) -- so the plugin is asked where it is the only JPEG decoder there is
    only = ref.Session(ext_dir=nb.EXT_DIR, with_ljt=False)
    gpu = only.decoder(device_id=0, backends=0, options=":fancy_upsampling=1")
    cpu_only = ref.Session.shared().decoder(options=":fancy_upsampling=1")
    # (format, planes, channels, sample type, orientation, expected status)
    rows = [(ref.P_YUV, 3, 1, ref.TYPE_UINT8, (0, 0, 0), ref.PS_SAMPLE_FORMAT_UNSUPPORTED),
            (ref.I_RGB, 1, 3, ref.TYPE_UINT16, (0, 0, 0), ref.PS_SAMPLE_TYPE_UNSUPPORTED),
            (ref.I_RGB, 2, 3, ref.TYPE_UINT8, (0, 0, 0), ref.PS_NUM_PLANES_UNSUPPORTED),
            (ref.I_RGB, 1, 2, ref.TYPE_UINT8, (0, 0, 0), ref.PS_NUM_CHANNELS_UNSUPPORTED),
            (ref.I_RGB, 1, 3, ref.TYPE_UINT8, (0, 0, 0), ref.STATUS_SUCCESS),
            (ref.P_RGB, 3, 1, ref.TYPE_UINT8, (0, 0, 0), ref.STATUS_SUCCESS),
            (ref.P_Y, 1, 1, ref.TYPE_UINT8, (0, 0, 0), ref.STATUS_SUCCESS),
            (ref.I_BGR, 1, 3, ref.TYPE_UINT8, (0, 0, 0), ref.STATUS_SUCCESS)]
    n = len(rows)
    args = dict(dims=[(64, 48)] * n, fmts=[r[0] for r in rows], planes=[r[1] for r in rows], channels=[r[2] for r in rows],
                types=[r[3] for r in rows], orient=[r[4] for r in rows], apply_exif=1)
    st_gpu = gpu.can_decode_ex([d] * n, buffer_kind=ref.BUFFER_DEVICE, **args)
    st_cpu = cpu_only.can_decode_ex([d] * n, buffer_kind=ref.BUFFER_HOST, **args)
    assert st_cpu == [r[5] for r in rows], [hex(x) for x in st_cpu]
    assert st_gpu == st_cpu, ([hex(x) for x in st_gpu], [hex(x) for x in st_cpu])
    # host buffers are the framework's business (it stages them for GPU backends after canDecode, src/work.h:144-190)
    assert gpu.can_decode_ex([d], [(64, 48)], [ref.I_RGB], [1], [3], buffer_kind=ref.BUFFER_HOST) == [ref.STATUS_SUCCESS]
    # a target that is not the image's size: the CPU plugin does not look, b200jpeg declines (it never resizes)
    st = gpu.can_decode_ex([d], [(32, 24)], [ref.I_RGB], [1], [3], buffer_kind=ref.BUFFER_DEVICE)
    assert st == [ref.PS_ROI_UNSUPPORTED]
    # progressive streams are ours; arithmetic coding: ENCODING_UNSUPPORTED from b200jpeg, decoded by libjpeg_turbo_ext through the chain
    prog, arith = make_jpeg(2, 64, 48, progressive=True), make_jpeg(3, 64, 48, arith=True)
    st = gpu.can_decode_ex([prog, arith], [(64, 48)] * 2, [ref.I_RGB] * 2, [1] * 2, [3] * 2, buffer_kind=ref.BUFFER_DEVICE)
    assert st == [ref.STATUS_SUCCESS, ref.PS_ENCODING_UNSUPPORTED]
    both = sess.decoder(device_id=0, backends=0, options=":fancy_upsampling=1")
    outs = [torch.zeros((48, 192), dtype=torch.uint8, device="cuda") for _ in range(3)]
    torch.cuda.synchronize()
    fst, _ = both.decode_raw([d, prog, arith], [o.data_ptr() for o in outs], [(64, 48)] * 3, [ref.I_RGB] * 3, [192] * 3, ref.BUFFER_DEVICE)
    torch.cuda.synchronize()
    want, _, _ = ref.Session.shared().decoder(options=":fancy_upsampling=1").decode_host([d, prog, arith])
    assert fst == [ref.STATUS_SUCCESS] * 3, fst
    for k in (0, 1, 2):
        assert np.array_equal(outs[k].cpu().numpy().reshape(48, 64, 3), want[k])
    for x in (gpu, cpu_only, both):
        x.close()
    only.close()


def test_framework_with_host_user_buffers(sess, cpu):
    """The user hands HOST buffers to a decoder whose first backend is GPU: the framework stages each image in a temporary
    device buffer and copies back on the image's stream (src/work.h:144-190).  Result = the reference's."""
    cases = [(640, 426, (2, 2), 90, 0), (333, 211, (2, 1), 85, 5), (227, 149, (1, 1), 75, 0), (1920, 1080, (2, 2), 90, 0)]
    datas = [make_jpeg(700 + i, w, h, samp, q, restart_interval=ri) for i, (w, h, samp, q, ri) in enumerate(cases)]
    want, _, _ = cpu.decode_host(datas)
    dec = sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=":fancy_upsampling=1")
    got, st, _ = dec.decode_host(datas)  # STRIDED_HOST numpy buffers
    assert st == [ref.STATUS_SUCCESS] * len(datas), st
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    dec.close()


def test_coefficients_bit_exact_411():
    """quantised DCT coefficients of the 4:1:1 streams (four luma blocks in a row per MCU)"""
    import torch
    from nvimagecodec_b200.decoder import Decoder
    cases = [(227, 149, 85, 0), (330, 77, 90, 3), (64, 64, 75, 0)]
    datas = [make_jpeg(3100 + i, w, h, (4, 1), q, restart_interval=ri) for i, (w, h, q, ri) in enumerate(cases)]
    dec = Decoder(0, options=":fancy_upsampling=1 :dump_coef=1")
    outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h, _, _) in cases]
    st = dec.decode_into(datas, [o.data_ptr() for o in outs], [c[0] * 3 for c in cases], [4] * len(datas))
    assert st == [0] * len(datas)
    for i, d in enumerate(datas):
        want = util.mcu_order_coefficients(d)
        assert np.array_equal(dec.debug_fetch(0, i, 0, want.size * 2, np.int16), want), cases[i]
    dec.close()


def test_fast_idct_is_left_to_the_cpu_plugin(sess):
    """fast_idct=1 selects JDCT_IFAST in the reference (libjpeg_turbo_decoder.cpp:272-273,352).  b200jpeg does not offer it:
    creating the decoder with that option fails, the framework treats the plugin as absent for the session and the CPU
    plugin produces the reference's (IFAST) pixels; with GPU backends only nothing decodes."""
    import torch
    from nvimagecodec_b200 import _capi
    import ctypes as C
    h = C.c_void_p()
    assert _capi.lib().b200jpeg_decoder_create(C.byref(h), 0, b":fast_idct=1") == _capi.STATUS_UNSUPPORTED and not h.value
    d = make_jpeg(8, 320, 240, (2, 2), 85)
    opts = ":fancy_upsampling=1 :fast_idct=1"
    want, wst, _ = ref.Session.shared().decoder(options=opts).decode_host([d])
    islow, _, _ = ref.Session.shared().decoder(options=":fancy_upsampling=1").decode_host([d])
    assert wst == [ref.STATUS_SUCCESS] and not np.array_equal(want[0], islow[0])  # the option does change the pixels
    dec = sess.decoder(device_id=0, backends=0, options=opts)
    out = torch.zeros((240, 960), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    st, _ = dec.decode_raw([d], [out.data_ptr()], [(320, 240)], [ref.I_RGB], [960], ref.BUFFER_DEVICE)
    torch.cuda.synchronize()
    assert st == [ref.STATUS_SUCCESS]
    assert np.array_equal(out.cpu().numpy().reshape(240, 320, 3), want[0])
    dec.close()
    try:  # GPU backends only: nobody is left to decode (whether that shows at create or at decode time is the framework's business)
        g = sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=opts)
    except RuntimeError:
        return
    st, _ = g.decode_raw([d], [out.data_ptr()], [(320, 240)], [ref.I_RGB], [960], ref.BUFFER_DEVICE)
    g.close()
    assert st != [ref.STATUS_SUCCESS]


def test_images_on_two_user_streams(cpu):
    """Two images of one batch on two different CUDA streams (nvimgcodecImageInfo_t::cuda_stream per image): the decoder
    waits for what each stream has in flight on its buffer, and each stream waits for the decode -- without any host
    synchronisation by the caller in between."""
    import torch
    from nvimagecodec_b200.decoder import Decoder
    datas = [make_jpeg(40 + i, 1280, 720, (2, 2), 90) for i in range(2)]
    want, _, _ = cpu.decode_host(datas)
    dec = Decoder(0, options=":fancy_upsampling=1")
    s = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [torch.zeros((720, 3840), dtype=torch.uint8, device="cuda") for _ in range(2)]
    copies = [torch.zeros_like(o) for o in outs]
    torch.cuda.synchronize()
    for k in range(2):
        with torch.cuda.stream(s[k]):
            torch.cuda._sleep(200_000_000)  # ~0.1 s of work in flight on the stream ...
            outs[k].fill_(0x77)             # ... that ends with a write to the output buffer
    st = dec.submit(datas, [o.data_ptr() for o in outs], [3840] * 2, [4] * 2, streams=[x.cuda_stream for x in s])
    for k in range(2):
        with torch.cuda.stream(s[k]):
            copies[k].copy_(outs[k])        # ordered after the decode by the decoder's event
    assert dec.wait() == [0, 0]
    for x in s:
        x.synchronize()
    for k in range(2):
        assert np.array_equal(copies[k].cpu().numpy().reshape(720, 1280, 3), want[k]), "stream %d" % k
    dec.close()


def test_two_batches_in_flight(sess, cpu):
    """b200jpeg_decode_submit / _wait with two batches in flight (each has its own scratch), and two batches in flight
    through the framework"""
    import torch
    from nvimagecodec_b200.decoder import Decoder
    a = [make_jpeg(60 + i, 640, 480, (2, 2), 85) for i in range(6)]
    b = [make_jpeg(80 + i, 500, 375, (2, 1), 90, restart_interval=4) for i in range(5)]
    wa, _, _ = cpu.decode_host(a)
    wb, _, _ = cpu.decode_host(b)
    dec = Decoder(0, options=":fancy_upsampling=1")
    oa = [torch.zeros((480, 1920), dtype=torch.uint8, device="cuda") for _ in a]
    ob = [torch.zeros((375, 1500), dtype=torch.uint8, device="cuda") for _ in b]
    torch.cuda.synchronize()
    for rep in range(3):
        dec.submit(a, [o.data_ptr() for o in oa], [1920] * 6, [4] * 6)
        dec.submit(b, [o.data_ptr() for o in ob], [1500] * 5, [4] * 5)
        with pytest.raises(RuntimeError):
            dec.submit(a, [o.data_ptr() for o in oa], [1920] * 6, [4] * 6)  # a third one is refused
        assert dec.wait() == [0] * 6 and dec.wait() == [0] * 5
        torch.cuda.synchronize()
        for o, w in zip(oa + ob, wa + wb):
            assert np.array_equal(o.cpu().numpy().reshape(w.shape), w)
            o.zero_()
    dec.close()
    # through the framework: two decoders (one batch in flight each, which is what the framework promises), batch r + 1 handed
    # to the other decoder before the future of batch r is waited for; decode() of the plugin returns with the device busy
    fdec = [sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=":fancy_upsampling=1") for _ in range(2)]
    oa2 = [torch.zeros_like(o) for o in oa]
    torch.cuda.synchronize()
    st, secs = fdec[0].decode_pipelined(a, [o.data_ptr() for o in oa], [o.data_ptr() for o in oa2], [(640, 480)] * 6, [ref.I_RGB] * 6,
                                        [1920] * 6, ref.BUFFER_DEVICE, rounds=6, other=fdec[1])
    torch.cuda.synchronize()
    assert st == [ref.STATUS_SUCCESS] * 6
    for o, o2, w in zip(oa, oa2, wa):
        assert np.array_equal(o.cpu().numpy().reshape(w.shape), w) and torch.equal(o, o2)
    for x in fdec:
        x.close()


def test_flat_optimised_periodic_and_appended_files(cpu):
    """The files ADVICE r1 found returned CORRUPT_STREAM: flat / mostly flat images with optimised Huffman tables (1-bit
    codes: one record per bit), large flat images (periodic streams: no self-synchronisation, 4096x4096 used to hit the
    round cap) and files with data after the first EOI (MPO-style second image, thumbnails, stray FFD9)."""
    rng = np.random.default_rng(5)
    flat = np.full((1024, 1024, 3), 200, np.uint8)
    mostly = flat.copy()
    mostly[:64] = rng.integers(0, 255, (64, 1024, 3))
    big = np.full((4096, 4096, 3), 128, np.uint8)
    big2 = big.copy()
    big2[:512] = rng.integers(0, 255, (512, 4096, 3))
    a = make_jpeg(91, 320, 200, (2, 2), 85)
    b2 = make_jpeg(92, 64, 64, (1, 1), 85)
    datas = [ref.ljt_encode(flat, 90, (1, 1), optimize=True), ref.ljt_encode(flat, 90, (2, 2), optimize=True),
             ref.ljt_encode(mostly, 90, (2, 2), optimize=True), ref.ljt_encode(flat, 90, (2, 2)),
             ref.ljt_encode(big, 90, (2, 2)), ref.ljt_encode(big2, 90, (2, 2)), ref.ljt_encode(big, 75, (1, 1), optimize=True),
             a + b"\x00" * 37 + b2, a + b2, a + b"\xff\xd9\x12\x34\xff\xd9"]
    dims = [(1024, 1024)] * 4 + [(4096, 4096)] * 3 + [(320, 200)] * 3
    want, rst, _ = cpu.decode_host(datas)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    got, st, stats = _decode_c_abi(datas, dims)
    assert st == [0] * len(datas), st
    for i, (g, w) in enumerate(zip(got, want)):
        assert np.array_equal(g, w), i
    assert stats["sync_rounds"] < 40, stats  # Jacobi rounds, then a handful of run-ahead rounds: never one round per subsequence


def test_row_stride_shorter_than_a_row_is_refused():
    import torch
    from nvimagecodec_b200.decoder import Decoder
    d = make_jpeg(5, 64, 48)
    dec = Decoder(0, options=":fancy_upsampling=1")
    t = torch.zeros((48, 192), dtype=torch.uint8, device="cuda")
    st = dec.decode_into([d, d, d], [t.data_ptr()] * 3, [0, 191, 192], [4] * 3)
    assert st[0] == 1 and st[1] == 1 and st[2] == 0  # INVALID_PARAMETER twice
    assert dec.decode(b"\xff\xd8\xff\xc0\x00\x11\x08\xff\xff\xff\xff\x03\x01\x22\x00\x02\x11\x01\x03\x11\x01\xff\xd9") is None
    dec.close()


def test_multi_device_decoder(cpu):
    """One batch over every GPU of the box (two decoder contexts on GPU 0 when there is only one): a host thread and stream set
    per device, the batch dealt by cost, outputs stay on the decoding GPU"""
    import torch
    from nvimagecodec_b200.decoder import MultiDeviceDecoder
    ndev = torch.cuda.device_count()
    devices = list(range(ndev)) if ndev > 1 else [0, 0]
    rng = np.random.default_rng(3)
    specs = [(int(rng.integers(320, 641)), int(rng.integers(240, 501))) for _ in range(24)]
    datas = [make_jpeg(2000 + i, w, h, (2, 2), 75) for i, (w, h) in enumerate(specs)] + [b"junk"]
    want, _, _ = cpu.decode_host(datas[:-1])
    md = MultiDeviceDecoder(devices, options=":fancy_upsampling=1")
    for _ in range(2):
        imgs = md.decode(datas)
        assert imgs[-1] is None
        used = set()
        for im, w in zip(imgs[:-1], want):
            assert im is not None and np.array_equal(im.cpu(), w)
            used.add(im.torch().device.index)
        assert used == set(devices)
    md.close()


@pytest.mark.timeout(900)
def test_c2_sharded_over_the_gpus():
    """North-star config 2: 256 ImageNet-shaped JPEGs dealt over the ranks by shard_batch, one process + decoder per rank under
    torch.distributed.run, every image compared with the reference (tools/c2_sharded.py does the work and the checking)"""
    import torch
    ndev = torch.cuda.device_count()
    for world in sorted({2, min(8, max(2, ndev))}):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
               "--master-port", str(29600 + world), os.path.join(util.ROOT, "tools", "c2_sharded.py"), "--images", "256", "--distinct", "32"]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=800)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
        assert "C2 sharded OK" in r.stdout, r.stdout[-2000:]


def test_exif_orientation_applied_by_the_colour_kernel(sess, cpu):
    """f3: the eight EXIF orientations (nvimgcodecImageInfo_t::orientation with apply_exif_orientation, the meaning of
    src/parsers/exif_orientation.h:37-60) through the C ABI and through nvimgcodecDecoderDecode; oracle = the reference's
    decode rotated / mirrored with numpy (the CPU plugin itself declines orientation, libjpeg_turbo_decoder.cpp:164-165)"""
    import torch
    from nvimagecodec_b200.decoder import Decoder
    dec = Decoder(0, options=":fancy_upsampling=1")
    fdec = sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=":fancy_upsampling=1")
    orients = util.EXIF_ORIENTATIONS
    n = len(orients)
    for (w, h, samp, gray) in [(640, 426, (2, 2), False), (333, 211, (2, 1), False), (1030, 200, (1, 1), False), (201, 99, (1, 1), True)]:
        d = make_jpeg(41, w, h, samp, 88, gray=gray)
        want = cpu.decode_host([d])[0][0]
        odims = [(h, w) if (o[0] // 90) & 1 else (w, h) for o in orients]
        for fmt in (4, 3):
            outs = [torch.zeros((oh, ow * 3) if fmt == 4 else (3 * oh, ow), dtype=torch.uint8, device="cuda") for (ow, oh) in odims]
            strides = [ow * 3 if fmt == 4 else ow for (ow, oh) in odims]
            st = dec.decode_into([d] * n, [o.data_ptr() for o in outs], strides, [fmt] * n, orientations=orients)
            torch.cuda.synchronize()
            assert st == [0] * n
            for o, t, (ow, oh) in zip(orients, outs, odims):
                a = t.cpu().numpy()
                got = a.reshape(oh, ow, 3) if fmt == 4 else a.reshape(3, oh, ow).transpose(1, 2, 0)
                assert np.array_equal(got, util.apply_orientation(want, *o)), (w, h, samp, fmt, o)
        outs = [torch.zeros((oh, ow * 3), dtype=torch.uint8, device="cuda") for (ow, oh) in odims]
        torch.cuda.synchronize()
        fst, _ = fdec.decode_raw([d] * n, [o.data_ptr() for o in outs], odims, [ref.I_RGB] * n, [ow * 3 for (ow, oh) in odims], ref.BUFFER_DEVICE,
                                 orient=[(o[0], int(o[1]), int(o[2])) for o in orients], apply_exif=1)
        torch.cuda.synchronize()
        assert fst == [ref.STATUS_SUCCESS] * n, fst
        for o, t, (ow, oh) in zip(orients, outs, odims):
            assert np.array_equal(t.cpu().numpy().reshape(oh, ow, 3), util.apply_orientation(want, *o)), ("framework", w, h, o)
    # orientation together with a region is not offered
    t = torch.zeros((10, 30), dtype=torch.uint8, device="cuda")
    st = dec.decode_into([d], [t.data_ptr()], [30], [4], rois=[(0, 0, 10, 10)], orientations=[(90, False, False)])
    assert st == [3]
    dec.close()
    fdec.close()


def test_many_progressive_images_one_warp_each(cpu):
    """from kProgManyImages (512) progressive images in one launch on, k_prog_decode runs one warp per image instead of a CTA of
    four (the images alone fill the device): 640 small progressive images in one group, every sampling mode, bit-exact"""
    import torch
    from nvimagecodec_b200.decoder import Decoder
    base_cases = [(64, 48, (2, 2), False), (97, 61, (2, 1), False), (80, 80, (1, 1), False), (120, 75, (1, 1), True), (131, 77, (1, 2), False),
                  (75, 50, (4, 1), False), (160, 120, (2, 2), False), (33, 17, (2, 2), False)]
    base = [make_jpeg(900 + i, w, h, samp, 85, gray=g, progressive=True) for i, (w, h, samp, g) in enumerate(base_cases)]
    want, rst, _ = cpu.decode_host(base)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    n = 640
    datas = [base[i % len(base)] for i in range(n)]
    dims = [base_cases[i % len(base)][:2] for i in range(n)]
    for opts in (":fancy_upsampling=1 :groups=1", ":fancy_upsampling=1"):  # one launch of 640 images / four of 160
        dec = Decoder(0, options=opts)
        outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
        st = dec.decode_into(datas, [o.data_ptr() for o in outs], [w * 3 for (w, h) in dims], [4] * n)
        torch.cuda.synchronize()
        assert st == [0] * n
        for i, o in enumerate(outs):
            w = want[i % len(base)]
            assert np.array_equal(o.cpu().numpy().reshape(w.shape), w), (opts, i)
        dec.close()


def test_many_small_baseline_images_in_one_launch(cpu):
    """640 small baseline images in one group: every per-kernel CTA map has more than 64 entries (the warp-wide image lookup
    narrows the range by binary search first), restart intervals and every sampling mode in the mix"""
    import torch
    from nvimagecodec_b200.decoder import Decoder
    base_cases = [(64, 48, (2, 2), False, 0), (97, 61, (2, 1), False, 3), (80, 80, (1, 1), False, 0), (120, 75, (1, 1), True, 0), (131, 77, (1, 2), False, 0),
                  (75, 50, (4, 1), False, 2), (160, 120, (2, 2), False, 1), (33, 17, (2, 2), False, 0), (1, 1, (1, 1), False, 0), (17, 9, (2, 2), False, 0)]
    base = [make_jpeg(950 + i, w, h, samp, 85, gray=g, restart_interval=ri) for i, (w, h, samp, g, ri) in enumerate(base_cases)]
    want, rst, _ = cpu.decode_host(base)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    n = 640
    datas = [base[i % len(base)] for i in range(n)]
    dims = [base_cases[i % len(base)][:2] for i in range(n)]
    for opts in (":fancy_upsampling=1 :groups=1", ":fancy_upsampling=1 :groups=1 :fused=1", ":fancy_upsampling=1 :groups=3"):
        dec = Decoder(0, options=opts)
        outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
        st = dec.decode_into(datas, [o.data_ptr() for o in outs], [w * 3 for (w, h) in dims], [4] * n)
        torch.cuda.synchronize()
        assert st == [0] * n
        for i, o in enumerate(outs):
            w = want[i % len(base)]
            assert np.array_equal(o.cpu().numpy().reshape(w.shape), w), (opts, i)
        dec.close()


def test_progressive_and_410_streams_on_the_gpu(sess, cpu):
    """f2: progressive streams (k_prog_decode: every scan of every image into dense coefficients, the images side by side) and
    4:1:0 sampling, mixed with baseline images in one batch: pixels and quantised coefficients bit-exact against the compiled
    reference; through the C ABI and through the framework with only GPU backends allowed"""
    import torch
    import nvimagecodec_b200 as nb
    from nvimagecodec_b200.decoder import Decoder
    cases = [(64, 48, (2, 2), False, 0, True), (227, 149, (2, 2), False, 0, True), (131, 77, (2, 1), False, 0, True), (640, 426, (2, 2), False, 0, False),
             (131, 77, (1, 2), False, 0, True), (100, 100, (1, 1), False, 0, True), (201, 99, (1, 1), True, 0, True), (333, 211, (2, 2), False, 5, True),
             (17, 9, (2, 2), False, 0, True), (1, 1, (1, 1), False, 0, True), (330, 77, (4, 1), False, 3, True), (227, 149, (4, 2), False, 0, True),
             (227, 149, (4, 2), False, 0, False), (1920, 1080, (2, 2), False, 0, True), (500, 375, (2, 2), False, 0, True), (800, 600, (2, 2), False, 0, False)]
    datas = [make_jpeg(70 + i, w, h, samp, 85, gray=g, progressive=p, restart_interval=ri) for i, (w, h, samp, g, ri, p) in enumerate(cases)]
    dims = [(c[0], c[1]) for c in cases]
    assert all(nb.parse_header(d).supported for d in datas)
    want, rst, _ = cpu.decode_host(datas)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    for opts in (":fancy_upsampling=1", ":fancy_upsampling=1 :groups=1 :dump_coef=1"):
        dec = Decoder(0, options=opts)
        outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
        for rep in range(2):  # twice through one decoder: the coefficient buffer is reused
            st = dec.decode_into(datas, [o.data_ptr() for o in outs], [w * 3 for (w, h) in dims], [4] * len(datas))
            torch.cuda.synchronize()
            assert st == [0] * len(datas), st
            for i, (o, w) in enumerate(zip(outs, want)):
                assert np.array_equal(o.cpu().numpy().reshape(w.shape), w), (opts, rep, i, cases[i])
        for i, d in enumerate(datas):
            if cases[i][5] or "dump_coef" in opts:
                wc = util.mcu_order_coefficients(d)
                assert np.array_equal(dec.debug_fetch(0, i, 0, wc.size * 2, np.int16), wc), (i, cases[i])
        dec.close()
    fdec = sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=":fancy_upsampling=1")
    outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
    torch.cuda.synchronize()
    fst, _ = fdec.decode_raw(datas, [o.data_ptr() for o in outs], dims, [ref.I_RGB] * len(datas), [w * 3 for (w, h) in dims], ref.BUFFER_DEVICE)
    torch.cuda.synchronize()
    assert fst == [ref.STATUS_SUCCESS] * len(datas)
    for o, w in zip(outs, want):
        assert np.array_equal(o.cpu().numpy().reshape(w.shape), w)
    fdec.close()
    # a progression with its last scan missing is declined (the CPU decoder smooths what is missing) and decoded by the chain
    prog = datas[1]
    cut = prog[:prog.rfind(b"\xff\xda")] + b"\xff\xd9"
    assert not nb.parse_header(cut).supported
    both = sess.decoder(device_id=0, backends=0, options=":fancy_upsampling=1")
    out = torch.zeros((149, 227 * 3), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    st, _ = both.decode_raw([cut], [out.data_ptr()], [(227, 149)], [ref.I_RGB], [227 * 3], ref.BUFFER_DEVICE)
    torch.cuda.synchronize()
    wcut, wst, _ = cpu.decode_host([cut])
    assert st == wst == [ref.STATUS_SUCCESS] and np.array_equal(out.cpu().numpy().reshape(149, 227, 3), wcut[0])
    both.close()
