"""Developer probe: stage times of a batch of 4K images (C3-like) -- the per-image kernels scale with the image."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from util import make_jpeg  # noqa: E402
from nvimagecodec_b200.decoder import Decoder  # noqa: E402

opts = sys.argv[1] if len(sys.argv) > 1 else ":fancy_upsampling=1 :profile=1"
for samp, name in (((2, 1), "4:2:2"), ((1, 1), "4:4:4")):
    base = [make_jpeg(8000 + i, 3840, 2160, samp, 90, restart_interval=ri) for i, ri in enumerate((0, 0, 480 if samp == (1, 1) else 240, 8))]
    n = 32
    datas = [base[i % len(base)] for i in range(n)]
    dec = Decoder(0, options=opts)
    outs = [torch.empty((2160, 3840 * 3), dtype=torch.uint8, device="cuda") for _ in range(n)]
    st = dec.prepare(datas, [o.data_ptr() for o in outs], [3840 * 3] * n, [4] * n)
    for it in range(4):
        torch.cuda.synchronize()
        t0 = time.time()
        st = dec.run()
        dt = time.time() - t0
    print(name, "%.3f ms" % (dt * 1e3), "ok", all(s == 0 for s in st), {k: round(v, 3) for k, v in (dec.stage_times() or {}).items()})
    dec.close()
