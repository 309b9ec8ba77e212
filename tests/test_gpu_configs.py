"""BASELINE.json configs C2-C4 at their full batch sizes: the CUDA path through the C ABI against the compiled reference
(libjpeg_turbo_ext), bit-exact.  A batch is built from a few distinct seeded images (encoded by the reference's
libjpeg-turbo) repeated up to the batch size: every distinct image is compared with the reference decoder's output, every
repeat with the first occurrence on the device.  Each case also writes its timings (device-resident inputs, and host
bytes -> host pixels) as one JSON line into gpurun_out/configs_bench.jsonl when that directory exists."""
import json
import os
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import util
from util import make_jpeg, ref

pytestmark = pytest.mark.gpu

FMT_REF = {4: ref.I_RGB, 3: ref.P_RGB, 7: ref.P_Y}


def _encode_all(specs):
    """specs: list of dicts for make_jpeg -> list of bytes (threads: the encoder is a ctypes call that drops the GIL)"""
    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as ex:
        return list(ex.map(lambda s: make_jpeg(**s), specs))


def _run_config(name, distinct, fmts_distinct, batch, options=":fancy_upsampling=1"):
    import torch
    from nvimagecodec_b200.decoder import Decoder
    n0 = len(distinct)
    idx = [i % n0 for i in range(batch)]
    datas = [distinct[i] for i in idx]
    fmts = [fmts_distinct[i] for i in idx]
    import nvimagecodec_b200 as nb
    dims = [dict(width=i.width, height=i.height) for i in (nb.parse_header(d) for d in distinct)]
    outs, ptrs, strides = [], [], []
    for i in idx:
        w, h = dims[i]["width"], dims[i]["height"]
        f = fmts_distinct[i]
        shape = (h, w * 3) if f == 4 else (h, w) if f == 7 else (3 * h, w)
        t = torch.zeros(shape, dtype=torch.uint8, device="cuda")
        outs.append(t)
        ptrs.append(t.data_ptr())
        strides.append(shape[1])
    dec = Decoder(0, options=options)
    st = dec.decode_into(datas, ptrs, strides, fmts)
    torch.cuda.synchronize()
    assert all(s == 0 for s in st), (name, [i for i, s in enumerate(st) if s])
    # reference decode of the distinct images (CPU thread pool)
    rdec = ref.Session.shared().decoder(options=options)
    want, rst, rsecs = rdec.decode_host(distinct, fmt=[FMT_REF[f] for f in fmts_distinct], reps=2)
    assert all(s == ref.STATUS_SUCCESS for s in rst)
    cpu_mpix = sum(d["width"] * d["height"] for d in dims) / 1e6 / min(rsecs)
    for i in range(n0):
        got = outs[i].cpu().numpy()
        w, h = dims[i]["width"], dims[i]["height"]
        f = fmts_distinct[i]
        got = got.reshape(h, w, 3) if f == 4 else got.reshape(1, h, w) if f == 7 else got.reshape(3, h, w)
        assert got.shape == want[i].shape, (name, i, got.shape, want[i].shape)
        assert np.array_equal(got, want[i]), "%s: image %d (%dx%d fmt %d) differs from the reference decoder" % (name, i, w, h, f)
    for k in range(n0, batch):
        assert torch.equal(outs[k], outs[idx[k]]), "%s: repeat %d differs from its first occurrence" % (name, k)
    # timings
    pixels = sum(dims[i]["width"] * dims[i]["height"] for i in idx)
    st = dec.prepare(datas, ptrs, strides, fmts, False)
    for _ in range(3):
        dec.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 5
    e0.record()
    for _ in range(steps):
        dec.run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    stats = dec.stats()
    houts = [torch.empty(o.shape, dtype=torch.uint8).pin_memory() for o in outs]
    hptrs = [o.data_ptr() for o in houts]
    dec.decode_into(datas, hptrs, strides, fmts, True)
    t0 = time.perf_counter()
    for _ in range(3):
        dec.decode_into(datas, hptrs, strides, fmts, True)
    e2e_ms = (time.perf_counter() - t0) / 3 * 1e3
    for i in (0, batch - 1):
        assert torch.equal(houts[i], outs[i].cpu()), "%s: host-output path differs" % name
    dec.close()
    line = {"config": name, "images": batch, "distinct": n0, "mpixel": pixels / 1e6, "jpeg_mb": sum(len(d) for d in datas) / 1e6,
            "ms_per_batch_device_inputs": ms, "mpixel_per_s": pixels / 1e6 / (ms / 1e3), "images_per_s": batch / (ms / 1e3),
            "e2e_ms_host_to_host": e2e_ms, "e2e_mpixel_per_s": pixels / 1e6 / (e2e_ms / 1e3), "sync_rounds": stats["sync_rounds"],
            "subsequences": stats["subsequences"], "tail_redone": stats["tail_redone"], "parity": "bit-exact vs libjpeg_turbo_ext",
            "cpu_reference_mpixel_per_s": cpu_mpix, "cpu_sample": "the %d distinct images, libjpeg_turbo_ext on %d host threads (WITH_SIMD=0)"
                                                                  % (n0, os.cpu_count() or 1)}
    out_dir = os.path.join(util.ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        with open(os.path.join(out_dir, "configs_bench.jsonl"), "a") as f:
            f.write(json.dumps(line) + "\n")
    return line


def test_c2_imagenet_shaped_batch():
    """C2: 256 variable-size ~500x375 4:2:0 q75 JPEGs -> interleaved RGB"""
    rng = np.random.default_rng(2)
    specs = [dict(seed=2000 + i, w=int(rng.integers(320, 641)), h=int(rng.integers(240, 501)), samp=(2, 2), quality=75) for i in range(32)]
    distinct = _encode_all(specs)
    _run_config("C2 256x ~500x375 4:2:0 q75", distinct, [4] * len(distinct), 256)


@pytest.mark.parametrize("samp,tag", [((1, 1), "4:4:4"), ((2, 1), "4:2:2")])
def test_c3_4k_with_and_without_restart(samp, tag):
    """C3: 64x 3840x2160, q90; no DRI / one restart per MCU row / restart every 8 MCUs"""
    specs = []
    for k, (ri, rows) in enumerate([(0, 0), (0, 1), (8, 0)]):
        for s in range(2):
            specs.append(dict(seed=3000 + 10 * k + s, w=3840, h=2160, samp=samp, quality=90, restart_interval=ri, restart_in_rows=rows))
    distinct = _encode_all(specs)
    _run_config("C3 64x 3840x2160 %s q90, DRI none/row/8" % tag, distinct, [4] * len(distinct), 64)


def test_c4_mixed_batch():
    """C4: 256x gray + 4:2:0 + 4:4:4, sizes log-uniform 256-4096 per side (odd sizes included), planar and interleaved"""
    rng = np.random.default_rng(4)
    specs, fmts = [], []
    for i in range(24):
        w = int(round(2 ** rng.uniform(8, 12)))
        h = int(round(2 ** rng.uniform(8, 12)))
        if i % 4 == 0:
            w, h = w | 1, h | 1  # force sizes that are no multiple of the MCU
        kind = i % 3
        if kind == 0:
            specs.append(dict(seed=4000 + i, w=w, h=h, gray=True, quality=85))
            fmts.append(7 if (i // 3) % 2 == 0 else 4)
        else:
            specs.append(dict(seed=4000 + i, w=w, h=h, samp=(2, 2) if kind == 1 else (1, 1), quality=85))
            fmts.append(4 if (i // 3) % 2 == 0 else 3)
    distinct = _encode_all(specs)
    _run_config("C4 256x mixed gray/4:2:0/4:4:4 256-4096 px, planar+interleaved", distinct, fmts, 256)


def test_c2_progressive_imagenet_shaped_batch():
    """C2 as progressive files (the reference corpus is full of progressive ImageNet JPEGs): 256 variable-size
    ~500x375 4:2:0 q75 progressive JPEGs -> interleaved RGB, every scan decoded on the GPU"""
    rng = np.random.default_rng(2)
    specs = [dict(seed=2000 + i, w=int(rng.integers(320, 641)), h=int(rng.integers(240, 501)), samp=(2, 2), quality=75, progressive=True)
             for i in range(32)]
    distinct = _encode_all(specs)
    _run_config("C2p 256x ~500x375 4:2:0 q75 PROGRESSIVE", distinct, [4] * len(distinct), 256)
