"""Developer probe / secondary reference: the reference's nvjpeg_ext (oracle/_ref/nvjpeg, `make -C oracle nvjpeg`) through
nvimgcodecDecoderDecode on the C1 batch -- host JPEG bytes in, device pixels out, hybrid (GPU_HYBRID) and hardware (NVJPG
engine) backends -- next to the CPU decoder's pixels (nvJPEG's IDCT is not libjpeg-turbo's: differences are expected, and
reported)."""
import io
import os
import sys
import time

import numpy as np
import torch
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nvimagecodec_b200.corpus import synth_image  # noqa: E402
from oracle import ref  # noqa: E402


def main():
    n, distinct = 64, int(sys.argv[1]) if len(sys.argv) > 1 else 64
    w, h = 1920, 1080
    base = []
    for i in range(distinct):
        b = io.BytesIO()
        Image.fromarray(synth_image(1000 + i, w, h)).save(b, "JPEG", quality=90, subsampling="4:2:0")
        base.append(b.getvalue())
    datas = [base[i % distinct] for i in range(n)]
    dims = [(w, h)] * n
    want, _, _ = ref.Session.shared().decoder().decode_host(datas[:4])
    ext_dir = os.path.join(ROOT, "oracle", "_ref", "nvjpeg")
    if not os.path.exists(os.path.join(ext_dir, "libnvjpeg_ext.so")):
        print("nvjpeg_ext not built")
        return
    sess = ref.Session(ext_dir=ext_dir, with_ljt=False)
    stream = torch.cuda.Stream()
    for name, mask in (("hybrid (GPU_HYBRID: Huffman on the CPU, the rest on the GPU)", ref.BACKEND_HYBRID), ("hardware (NVJPG engine)", ref.BACKEND_HW),
                       ("gpu_only", ref.BACKEND_GPU_ONLY)):
        try:
            decs = [sess.decoder(device_id=0, backends=mask, options=":fancy_upsampling=1") for _ in range(2)]
        except Exception as e:  # noqa: BLE001
            print(name, "decoder create failed:", e)
            continue
        bufs = [torch.zeros((n, h, w * 3), dtype=torch.uint8, device="cuda") for _ in range(2)]
        ptrs = [[b[i].data_ptr() for i in range(n)] for b in bufs]
        try:
            st, _ = decs[0].decode_pipelined(datas, ptrs[0], ptrs[1], dims, [ref.I_RGB] * n, [w * 3] * n, ref.BUFFER_DEVICE, stream.cuda_stream, rounds=2, depth=1)
        except Exception as e:  # noqa: BLE001
            print(name, "decode failed:", e)
            continue
        torch.cuda.synchronize()
        ok = sum(1 for s in st if s == ref.STATUS_SUCCESS)
        if ok != n:
            print(name, "statuses: %d of %d decoded (first: %s)" % (ok, n, [hex(s) for s in st[:4]]))
            if ok == 0:
                continue
        got = bufs[0][:4].cpu().numpy().reshape(4, h, w, 3)
        diff = np.abs(got.astype(np.int16) - np.stack(want).astype(np.int16))
        for depth, other, label in ((1, None, "one call at a time"), (2, decs[1], "two calls in flight")):
            rounds = 10
            st, secs = decs[0].decode_pipelined(datas, ptrs[0], ptrs[1], dims, [ref.I_RGB] * n, [w * 3] * n, ref.BUFFER_DEVICE, stream.cuda_stream, rounds=rounds, depth=depth,
                                                other=other)
            torch.cuda.synchronize()
            ms = secs / rounds * 1e3
            print("nvjpeg_ext %s, %s: %.2f ms per batch of %d, %.0f Mpixel/s; vs libjpeg-turbo pixels: max |diff| %d, mean %.3f, %0.1f%% identical" % (
                name, label, ms, n, n * w * h / 1e6 / (ms / 1e3), int(diff.max()), float(diff.mean()), 100.0 * float((diff == 0).mean())))
        for d in decs:
            d.close()
    sess.close()


if __name__ == "__main__":
    main()
