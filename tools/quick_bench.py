"""Developer timing probe (not the contract bench): per-stage CUDA-event times of the decode of N 1080p JPEGs."""
import io
import os
import sys
import time

import numpy as np
import torch
from PIL import Image

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nvimagecodec_b200.corpus import synth_image  # noqa: E402
from nvimagecodec_b200.decoder import Decoder  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    distinct = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    opts = sys.argv[3] if len(sys.argv) > 3 else ":fancy_upsampling=1 :profile=1"
    w, h = 1920, 1080
    base = []
    for i in range(distinct):
        b = io.BytesIO()
        Image.fromarray(synth_image(1000 + i, w, h)).save(b, "JPEG", quality=90, subsampling="4:2:0")
        base.append(b.getvalue())
    datas = [base[i % distinct] for i in range(n)]
    print("jpeg bytes", sum(len(d) for d in datas) / n)
    dec = Decoder(0, options=opts)
    outs = [torch.empty((h, w * 3), dtype=torch.uint8, device="cuda") for _ in range(n)]
    t0 = time.time()
    st = dec.prepare(datas, [o.data_ptr() for o in outs], [w * 3] * n, [4] * n)
    print("prepare s", time.time() - t0, "status ok", all(s == 0 for s in st))
    for it in range(6):
        torch.cuda.synchronize()
        t0 = time.time()
        st = dec.run()
        dt = time.time() - t0
        print("run %d: %.3f ms  %.0f img/s  ok=%s  stages=%s stats=%s" % (it, dt * 1e3, n / dt, all(s == 0 for s in st),
              {k: round(v, 3) for k, v in (dec.stage_times() or {}).items()}, dec.stats()))
    # end-to-end (prepare+run) from host bytes
    for it in range(3):
        t0 = time.time()
        st = dec.decode_into(datas, [o.data_ptr() for o in outs], [w * 3] * n, [4] * n)
        dt = time.time() - t0
        print("e2e(device out) %d: %.3f ms %.0f img/s" % (it, dt * 1e3, n / dt))


if __name__ == "__main__":
    main()
