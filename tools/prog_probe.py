"""Developer probe: throughput of progressive decode (one warp per image) against the batch size."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from util import make_jpeg  # noqa: E402
from nvimagecodec_b200.decoder import Decoder  # noqa: E402


def main():
    rng = np.random.default_rng(2)
    specs = [(int(rng.integers(320, 641)), int(rng.integers(240, 501))) for _ in range(32)]
    base = [make_jpeg(2000 + i, w, h, (2, 2), 75, progressive=True) for i, (w, h) in enumerate(specs)]
    for n in (64, 256, 1024, 4096):
        datas = [base[i % 32] for i in range(n)]
        dims = [specs[i % 32] for i in range(n)]
        outs = [torch.empty((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
        dec = Decoder(0, options=":fancy_upsampling=1")
        dec.prepare(datas, [o.data_ptr() for o in outs], [w * 3 for (w, h) in dims], [4] * n)
        dec.run()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            st = dec.run()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / reps * 1e3
        pix = sum(w * h for w, h in dims)
        print("progressive batch %5d: %.1f ms  %.0f img/s  %.0f Mpixel/s  ok=%s" % (n, ms, n / ms * 1e3, pix / 1e6 / ms * 1e3, all(s == 0 for s in st)))
        dec.close()


if __name__ == "__main__":
    main()
