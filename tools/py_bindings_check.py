#!/usr/bin/env python
"""f4: the reference's own Python layer (python/*.cpp built by oracle/Makefile `py` into oracle/_ref/py) picks the b200jpeg
plugin up unchanged.  Mirrors the decode part of the reference's test/python/test_python.py:111-205 (single image / batch,
numpy / bytes input, path input via Decoder.read, a cuda_stream, max_num_cpu_threads, backends) on regenerated fixtures,
with a torch-based comparison (cupy is not in the image): every image decoded through nvimgcodec.Decoder on the GPU is
bit-exact with the same call on the CPU backend (libjpeg_turbo_ext), and b200jpeg is shown to have done the work (only
GPU_ONLY backends allowed in one of the decoders).  Run as a script: it sets NVIMGCODEC_EXTENSIONS_PATH before the import.
"""
import os
import shutil
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
PY = os.path.join(ROOT, "oracle", "_ref", "py")


def main():
    import numpy as np
    import torch
    import nvimagecodec_b200 as nb
    from util import make_jpeg
    ext = tempfile.mkdtemp(prefix="nvimgcodec_ext_")
    shutil.copy(nb.EXT_PATH, ext)
    shutil.copy(os.path.join(PY, "nvidia", "nvimgcodec", "extensions", "libjpeg_turbo_ext.so"), ext)
    os.environ["NVIMGCODEC_EXTENSIONS_PATH"] = ext
    import nvidia  # site-packages has a regular package of that name: graft the reference's sub-package onto it
    nvidia.__path__.insert(0, os.path.join(PY, "nvidia"))
    from nvidia import nvimgcodec
    print("nvidia.nvimgcodec", nvimgcodec.__version__, "cuda", nvimgcodec.__cuda_version__)

    cases = [(640, 426, (2, 2), 75, 0, False), (640, 426, (2, 1), 85, 0, False), (640, 426, (1, 1), 90, 0, False), (640, 426, (1, 2), 90, 0, False),
             (640, 426, (4, 1), 90, 0, False), (640, 426, (1, 1), 85, 0, True), (333, 211, (2, 2), 90, 5, False), (1920, 1080, (2, 2), 90, 0, False)]
    datas = [make_jpeg(100 + i, w, h, samp, q, restart_interval=ri, gray=g) for i, (w, h, samp, q, ri, g) in enumerate(cases)]
    datas.append(make_jpeg(198, 320, 240, progressive=True))
    datas.append(make_jpeg(199, 320, 240, arith=True))  # whatever b200jpeg declines comes from the CPU plugin through the fallback chain

    def host(img):
        return np.asarray(img.cpu())

    opts = ":fancy_upsampling=1"
    cpu_dec = nvimgcodec.Decoder(backends=[nvimgcodec.Backend(nvimgcodec.CPU_ONLY)], options=opts)
    want = [host(im) for im in cpu_dec.decode(datas)]
    checked = 0
    for threads in (0, 1, 5):
        for backends in (None, [nvimgcodec.Backend(nvimgcodec.GPU_ONLY, load_hint=0.5), nvimgcodec.Backend(nvimgcodec.HYBRID_CPU_GPU),
                                nvimgcodec.Backend(nvimgcodec.CPU_ONLY)]):
            dec = nvimgcodec.Decoder(max_num_cpu_threads=threads, backends=backends, options=opts) if backends else \
                nvimgcodec.Decoder(max_num_cpu_threads=threads, options=opts)
            for stream in (None, torch.cuda.Stream()):
                for fmt in ("python", "numpy"):
                    enc = [np.frombuffer(d, dtype=np.uint8) for d in datas] if fmt == "numpy" else datas
                    imgs = dec.decode(enc, cuda_stream=0 if stream is None else stream.cuda_stream)
                    if stream is not None:
                        stream.synchronize()
                    torch.cuda.synchronize()
                    for i, (im, w) in enumerate(zip(imgs, want)):
                        assert im is not None, (threads, backends is not None, fmt, i)
                        assert im.buffer_kind == nvimgcodec.ImageBufferKind.STRIDED_DEVICE
                        t = torch.as_tensor(im, device="cuda")  # __cuda_array_interface__: zero-copy hand-off to PyTorch
                        assert tuple(t.shape) == w.shape and np.array_equal(t.cpu().numpy(), w), (threads, fmt, i, cases[i] if i < len(cases) else "prog")
                        checked += 1
                    one = dec.decode(enc[0])
                    assert np.array_equal(host(one), want[0])
    # only GPU backends allowed: what decodes was decoded by b200jpeg (nvjpeg_ext is not in the extensions directory)
    gpu_only = nvimgcodec.Decoder(backends=[nvimgcodec.Backend(nvimgcodec.GPU_ONLY)], options=opts)
    # (without the arithmetic-coded file: when the last allowed decoder declines a sample, the reference's fallback worker has no
    # decoder and fails the whole batch -- src/decoder_worker.cpp:268-272 setAll)
    imgs = gpu_only.decode(datas[:-1])
    assert len(imgs) == len(datas) - 1
    for i, (im, w) in enumerate(zip(imgs, want)):
        assert im is not None and np.array_equal(host(im), w), i
    # path input (Decoder.read) and DLPack
    d = tempfile.mkdtemp(prefix="nvimgcodec_img_")
    paths = []
    for i, data in enumerate(datas[:4]):
        paths.append(os.path.join(d, "img%d.jpg" % i))
        open(paths[-1], "wb").write(data)
    for im, w in zip(gpu_only.read(paths), want):
        assert np.array_equal(host(im), w)
        t = torch.from_dlpack(im.to_dlpack()) if hasattr(im, "to_dlpack") else torch.as_tensor(im, device="cuda")
        assert np.array_equal(t.cpu().numpy(), w)
    # the Python default ":fancy_upsampling=0" (python/decoder.cpp:283-326)
    d0 = nvimgcodec.Decoder()
    c0 = nvimgcodec.Decoder(backends=[nvimgcodec.Backend(nvimgcodec.CPU_ONLY)])
    for a, b in zip(d0.decode(datas[:5]), c0.decode(datas[:5])):
        assert np.array_equal(host(a), host(b))
    shutil.rmtree(ext, ignore_errors=True)
    shutil.rmtree(d, ignore_errors=True)
    print("python bindings OK: %d images through nvidia.nvimgcodec.Decoder bit-exact against its CPU backend" % checked)


if __name__ == "__main__":
    main()
