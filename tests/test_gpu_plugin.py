"""Drop-in proof on the GPU: the *unmodified* reference framework (oracle/_ref: nvimgcodec core + libjpeg_turbo_ext)
dlopen()s libb200jpeg_ext.so from an extensions directory and nvimgcodecDecoderDecode / nvimgcodecEncoderEncode route
the JPEG path through it (priority 50 < libjpeg_turbo's 300); out-of-scope streams fall back to the CPU plugin through
the framework's own fallback chain.  Results must equal the CPU reference bit for bit."""
import numpy as np
import pytest

import util
from util import make_jpeg, port, ref, synth_image

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not util.have_ref(), reason="oracle/_ref not built")]


@pytest.fixture(scope="module")
def sess():
    import nvimagecodec_b200 as nb
    s = ref.Session(ext_dir=nb.EXT_DIR, with_ljt=True)
    yield s
    s.close()


def test_decode_through_reference_framework(sess):
    import torch
    cases = [(640, 426, (2, 2), 90, 0, False), (333, 211, (2, 1), 85, 5, False), (227, 149, (1, 1), 75, 0, False), (201, 99, (1, 1), 85, 0, True),
             (1920, 1080, (2, 2), 90, 0, False)]
    datas = [make_jpeg(300 + i, w, h, samp, q, restart_interval=ri, gray=g) for i, (w, h, samp, q, ri, g) in enumerate(cases)]
    datas.append(make_jpeg(399, 320, 240, progressive=True))  # progressive: decoded by b200jpeg too (f2)
    dims = [(c[0], c[1]) for c in cases] + [(320, 240)]
    dec = sess.decoder(device_id=0, backends=0, options=":fancy_upsampling=1")
    outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
    torch.cuda.synchronize()
    st, _ = dec.decode_raw(datas, [o.data_ptr() for o in outs], dims, [ref.I_RGB] * len(datas), [w * 3 for (w, h) in dims], ref.BUFFER_DEVICE)
    torch.cuda.synchronize()
    assert st == [ref.STATUS_SUCCESS] * len(datas), st
    want, wst, _ = ref.Session.shared().decoder(options=":fancy_upsampling=1").decode_host(datas)
    for o, w in zip(outs, want):
        assert np.array_equal(o.cpu().numpy().reshape(w.shape), w)
    dec.close()


def test_region_of_interest_through_reference_framework(sess):
    """nvimgcodecImageInfo_t::region with the GPU_ONLY backend list: only our plugin can have produced the pixels, and they
    equal the region decoded by libjpeg_turbo_ext"""
    import torch
    cases = [(640, 426, (2, 2)), (333, 211, (2, 1)), (227, 149, (1, 1))]
    datas = [make_jpeg(600 + i, w, h, samp, 88) for i, (w, h, samp) in enumerate(cases)]
    rois = [(100, 50, 333, 200), (1, 0, 332, 211), (226, 148, 1, 1)]
    dims = [(r[2], r[3]) for r in rois]
    dec = sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=":fancy_upsampling=1")
    outs = [torch.zeros((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
    torch.cuda.synchronize()
    st, _ = dec.decode_raw(datas, [o.data_ptr() for o in outs], dims, [ref.I_RGB] * 3, [w * 3 for (w, h) in dims], ref.BUFFER_DEVICE, rois=rois)
    torch.cuda.synchronize()
    assert st == [ref.STATUS_SUCCESS] * 3, st
    want, wst, _ = ref.Session.shared().decoder(options=":fancy_upsampling=1").decode_host(datas, rois=rois)
    for o, w in zip(outs, want):
        assert np.array_equal(o.cpu().numpy().reshape(w.shape), w)
    dec.close()


def test_gpu_only_backend_selects_b200jpeg(sess):
    """backends=[GPU_ONLY]: libjpeg_turbo_ext (CPU_ONLY) is filtered out by the framework (src/decoder_worker.cpp:63-93), so a
    successful decode can only have come from our plugin"""
    import torch
    datas = [make_jpeg(410 + i, 320, 240, samp) for i, samp in enumerate([(2, 2), (1, 1)])]
    dec = sess.decoder(device_id=0, backends=ref.BACKEND_GPU_ONLY, options=":fancy_upsampling=1")
    outs = [torch.zeros((240, 960), dtype=torch.uint8, device="cuda") for _ in range(2)]
    torch.cuda.synchronize()
    st, _ = dec.decode_raw(datas, [o.data_ptr() for o in outs], [(320, 240)] * 2, [ref.I_RGB] * 2, [960] * 2, ref.BUFFER_DEVICE)
    torch.cuda.synchronize()
    assert st == [ref.STATUS_SUCCESS] * 2
    for d, o in zip(datas, outs):
        assert np.array_equal(o.cpu().numpy().reshape(240, 320, 3), port.decode(d, port.FMT_I_RGB, True))
    dec.close()


def test_encode_through_reference_framework(sess):
    import torch
    imgs = [synth_image(500 + i, w, h) for i, (w, h) in enumerate([(640, 426), (333, 211), (1920, 1080)])]
    dev = [torch.from_numpy(i).cuda() for i in imgs]
    dims = [(i.shape[1], i.shape[0]) for i in imgs]
    enc = sess.encoder(device_id=0, backends=0)
    torch.cuda.synchronize()
    for css, samp in ((3, (2, 2)), (0, (1, 1)), (2, (2, 1))):
        got, st, _ = enc.encode_raw([d.data_ptr() for d in dev], dims, [ref.I_RGB] * 3, [w * 3 for (w, h) in dims], ref.BUFFER_DEVICE, 90,
                                    [css] * 3)
        assert st == [ref.STATUS_SUCCESS] * 3, st
        for im, g in zip(imgs, got):
            assert g == ref.ljt_encode(im, 90, samp)
    enc.close()
