"""Developer probe: phase times of every CTA of k_huff_fused (needs a library built with -DB2J_FUSED_TRACE:
tools/build_variant.sh trace -DB2J_FUSED_TRACE; B200JPEG_LIB=tools/_variants/trace/libb200jpeg.so python tools/fused_trace.py)."""
import io
import os
import sys

import numpy as np
import torch
from PIL import Image

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nvimagecodec_b200.corpus import synth_image  # noqa: E402
from nvimagecodec_b200.decoder import Decoder  # noqa: E402


def main():
    n, distinct = 64, 8
    opts = sys.argv[1] if len(sys.argv) > 1 else ":fancy_upsampling=1 :profile=1"
    w, h = 1920, 1080
    base = []
    for i in range(distinct):
        b = io.BytesIO()
        Image.fromarray(synth_image(1000 + i, w, h)).save(b, "JPEG", quality=90, subsampling="4:2:0")
        base.append(b.getvalue())
    datas = [base[i % distinct] for i in range(n)]
    dec = Decoder(0, options=opts)
    outs = [torch.empty((h, w * 3), dtype=torch.uint8, device="cuda") for _ in range(n)]
    dec.prepare(datas, [o.data_ptr() for o in outs], [w * 3] * n, [4] * n)
    for it in range(4):
        st = dec.run()
    torch.cuda.synchronize()
    print("stages", dec.stage_times(), dec.stats())
    t = dec.debug_fetch(3, 0, 0, 64 * 4096, np.uint64).reshape(-1, 8).astype(np.int64)
    t = t[t[:, 0] > 0]
    t0 = t[:, 0].min()
    names = ["start", "staged", "round0", "round1", "round2", "converged", "linked", "block pass"]
    print("CTAs", len(t), "kernel span us", (t[:, 7].max() - t0) / 1e3)
    d = np.diff(t, axis=1) / 1e3
    for j in range(7):
        print("%-10s -> %-10s: mean %7.1f  p10 %7.1f  p50 %7.1f  p90 %7.1f  max %7.1f us" % (names[j], names[j + 1], d[:, j].mean(), np.percentile(d[:, j], 10),
              np.percentile(d[:, j], 50), np.percentile(d[:, j], 90), d[:, j].max()))
    life = (t[:, 7] - t[:, 0]) / 1e3
    print("CTA life: mean %.1f p50 %.1f max %.1f us; start times p50 %.1f p90 %.1f max %.1f us" % (life.mean(), np.percentile(life, 50), life.max(),
          np.percentile(t[:, 0] - t0, 50) / 1e3, np.percentile(t[:, 0] - t0, 90) / 1e3, (t[:, 0].max() - t0) / 1e3))
    order = np.argsort(t[:, 0])
    for idx in list(order[:3]) + list(order[len(order) // 2: len(order) // 2 + 3]) + list(order[-3:]):
        print("cta", idx, [round((x - t0) / 1e3, 1) for x in t[idx]])


if __name__ == "__main__":
    main()
