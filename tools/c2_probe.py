"""Developer probe: host-side timeline (B200JPEG_TRACE) of a C2-shaped batch (256 small 4:2:0 images), host bytes -> host
pixels, outputs in one contiguous pinned buffer and in separate pinned buffers."""
import io, os, sys, time
import numpy as np, torch
from PIL import Image
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200JPEG_TRACE"] = "1"
from nvimagecodec_b200.corpus import synth_image
from nvimagecodec_b200.decoder import Decoder

rng = np.random.default_rng(2)
base, dims = [], []
for i in range(32):
    w, h = int(rng.integers(320, 641)), int(rng.integers(240, 501))
    b = io.BytesIO(); Image.fromarray(synth_image(2000 + i, w, h)).save(b, "JPEG", quality=75, subsampling="4:2:0")
    base.append(b.getvalue()); dims.append((w, h))
n = 256
datas = [base[i % 32] for i in range(n)]
sizes = [dims[i % 32][0] * dims[i % 32][1] * 3 for i in range(n)]
strides = [dims[i % 32][0] * 3 for i in range(n)]
host = torch.empty(sum(sizes), dtype=torch.uint8).pin_memory()
offs = np.concatenate([[0], np.cumsum(sizes)])
contig = [host.data_ptr() + int(offs[i]) for i in range(n)]
sep_t = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in sizes]
sep = [t.data_ptr() for t in sep_t]
mpix = sum(sizes) / 3e6
for opts in sys.argv[1:] or [":fancy_upsampling=1"]:
    dec = Decoder(0, options=opts)
    for name, ptrs in (("contiguous", contig), ("separate", sep)):
        ts = []
        for k in range(6):
            sys.stderr.write("---- %s call %d\n" % (name, k))
            t0 = time.perf_counter(); st = dec.decode_into(datas, ptrs, strides, [4] * n, True); ts.append(time.perf_counter() - t0)
            assert all(s == 0 for s in st)
        print(opts, name, "e2e ms: min %.2f med %.2f -> %.0f Mpix/s" % (min(ts) * 1e3, sorted(ts)[len(ts) // 2] * 1e3, mpix / min(ts)))
    dec.close()
