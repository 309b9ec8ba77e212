"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol include/b200jpeg.h declares, the
header parser agrees with the reference's probe, the option grammar matches, and the nvImageCodec extension module
registers through the unmodified reference framework with the right priority / fallback behaviour.  No GPU compute."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import util
from util import ROOT, make_jpeg, ref

import nvimagecodec_b200 as nb
from nvimagecodec_b200 import _capi


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b200jpeg.h")).read()
    declared = set(re.findall(r"B200JPEG_API\s+[\w\s\*]+?\b(b200jpeg_\w+)\s*\(", hdr))
    assert declared == set(_capi.EXPORTS), declared ^ set(_capi.EXPORTS)
    lib = _capi.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.b200jpeg_version()
    assert _capi.status_string(3) == "unsupported jpeg"


def test_struct_layouts_match_header():
    assert C.sizeof(_capi.Output) == 56
    assert C.sizeof(_capi.StreamInfo) == 72
    assert C.sizeof(_capi.EncodeInput) == 48


@pytest.mark.skipif(not util.have_ref(), reason="oracle/_ref not built")
def test_parse_header_agrees_with_reference_probe():
    sess = ref.Session.shared()
    for samp, css in (((2, 2), 3), ((2, 1), 2), ((1, 1), 0), ((1, 2), 4)):
        d = make_jpeg(1, 123, 77, samp, restart_interval=3)
        i = nb.parse_header(d)
        w, h, planes, ch, rcss, enc = sess.get_info(d)
        assert (i.width, i.height, i.subsampling, i.sof_marker, i.restart_interval, i.supported) == (w, h, rcss, enc, 3, 1)
        assert i.subsampling == css
        scan = d[i.scan_offset:i.scan_offset + i.scan_bytes]
        assert d[i.scan_offset + i.scan_bytes:] == b"\xff\xd9" and len(scan) > 0
    g = make_jpeg(2, 50, 40, gray=True)
    assert nb.parse_header(g).subsampling == 7 and nb.parse_header(g).num_components == 1
    i = nb.parse_header(make_jpeg(3, 64, 64, progressive=True))
    assert i.supported == 1 and i.sof_marker == 0xC2  # complete progressions are decoded (f2)
    i = nb.parse_header(make_jpeg(3, 64, 64, arith=True))
    assert i.supported == 0 and i.sof_marker not in (0xC0, 0xC1, 0xC2)
    prog = make_jpeg(3, 64, 64, progressive=True)
    cut = prog[:prog.rfind(b"\xff\xda")] + b"\xff\xd9"  # the last scan missing: an incomplete progression is the CPU decoder's business
    assert nb.parse_header(cut).supported == 0
    with pytest.raises(ValueError):
        nb.parse_header(b"\x89PNG....")
    with pytest.raises(ValueError):
        nb.parse_header(make_jpeg(4, 64, 64)[:100])


def test_decoder_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = _capi.lib().b200jpeg_decoder_create(C.byref(h), 0, b"")
    assert rc in (_capi.STATUS_CUDA_ERROR, _capi.STATUS_INVALID_PARAMETER) and not h.value
    with pytest.raises(RuntimeError):
        nb.Decoder(0)


@pytest.mark.skipif(not (util.have_ref() and os.path.exists(nb.EXT_PATH)), reason="needs oracle/_ref and the extension module")
def test_extension_module_registers_and_falls_back():
    """dlopen discovery path (src/plugin_framework.cpp:286-351): our .so in the extensions dir is loaded by the unmodified
    framework; with a CPU-only device our GPU_ONLY decoder declines creation and libjpeg_turbo_ext (priority 300) decodes."""
    lib = C.CDLL(nb.EXT_PATH)
    assert hasattr(lib, "nvimgcodecExtensionModuleEntry") and hasattr(lib, "get_b200jpeg_extension_desc")
    sess = ref.Session(ext_dir=nb.EXT_DIR, with_ljt=True)
    d = make_jpeg(9, 64, 48)
    dec = sess.decoder(device_id=ref.DEVICE_CPU_ONLY, backends=0)
    out, st, _ = dec.decode_host([d])
    assert st == [ref.STATUS_SUCCESS]
    want, _, _ = ref.Session.shared().decoder().decode_host([d])
    assert np.array_equal(out[0], want[0])
    # without the CPU plugin nothing can decode on a CPU-only device: our plugin never falls back to the CPU itself
    sess2 = ref.Session(ext_dir=nb.EXT_DIR, with_ljt=False)
    dec2 = sess2.decoder(device_id=ref.DEVICE_CPU_ONLY, backends=0)
    out2, st2, _ = dec2.decode_host([d])
    assert st2 != [ref.STATUS_SUCCESS]
    dec.close(); dec2.close(); sess.close(); sess2.close()
