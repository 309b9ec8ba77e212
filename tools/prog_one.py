"""Developer probe: one progressive batch of 64 images (for ncu captures of k_prog_decode)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from util import make_jpeg  # noqa: E402
from nvimagecodec_b200.decoder import Decoder  # noqa: E402

rng = np.random.default_rng(2)
specs = [(int(rng.integers(320, 641)), int(rng.integers(240, 501))) for _ in range(32)]
base = [make_jpeg(2000 + i, w, h, (2, 2), 75, progressive=True) for i, (w, h) in enumerate(specs)]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
datas = [base[i % 32] for i in range(n)]
dims = [specs[i % 32] for i in range(n)]
outs = [torch.empty((h, w * 3), dtype=torch.uint8, device="cuda") for (w, h) in dims]
dec = Decoder(0, options=":fancy_upsampling=1")
dec.prepare(datas, [o.data_ptr() for o in outs], [w * 3 for (w, h) in dims], [4] * n)
for _ in range(2):
    st = dec.run()
torch.cuda.synchronize()
print("ok", all(s == 0 for s in st), "bytes", sum(len(d) for d in datas) / n)
