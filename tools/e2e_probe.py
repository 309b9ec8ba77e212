"""Developer probe: PCIe copy rates and the end-to-end (host bytes -> host pixels) decode time for several option sets."""
import io, os, sys, time
import numpy as np, torch
from PIL import Image
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nvimagecodec_b200.corpus import synth_image
from nvimagecodec_b200.decoder import Decoder

n, w, h = 64, 1920, 1080
base = []
for i in range(8):
    b = io.BytesIO(); Image.fromarray(synth_image(1000 + i, w, h)).save(b, "JPEG", quality=90, subsampling="4:2:0"); base.append(b.getvalue())
datas = [base[i % 8] for i in range(n)]
dev = torch.empty(n * h * w * 3, dtype=torch.uint8, device="cuda")
host = torch.empty(n * h * w * 3, dtype=torch.uint8).pin_memory()
for name, a, b_ in (("D2H", host, dev), ("H2D", dev, host)):
    a.copy_(b_, non_blocking=True); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        a.copy_(b_, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 3
    print("%s %.1f MB in %.2f ms = %.1f GB/s" % (name, a.numel() / 1e6, dt * 1e3, a.numel() / dt / 1e9))
houts = [host[i * h * w * 3:(i + 1) * h * w * 3] for i in range(n)]
for opts in sys.argv[1:] or [":fancy_upsampling=1"]:
    dec = Decoder(0, options=opts)
    ptrs = [o.data_ptr() for o in houts]
    for _ in range(3):
        dec.decode_into(datas, ptrs, [w * 3] * n, [4] * n, True)
    ts = []
    for _ in range(8):
        t0 = time.perf_counter(); dec.decode_into(datas, ptrs, [w * 3] * n, [4] * n, True); ts.append(time.perf_counter() - t0)
    dts = [torch.empty((h, w * 3), dtype=torch.uint8, device="cuda") for _ in range(n)]
    dp = [o.data_ptr() for o in dts]
    td = []
    for _ in range(5):
        t0 = time.perf_counter(); dec.decode_into(datas, dp, [w * 3] * n, [4] * n, False); td.append(time.perf_counter() - t0)
    print(opts, "e2e host->host ms: min %.2f med %.2f | host->device ms: min %.2f" % (min(ts) * 1e3, sorted(ts)[len(ts) // 2] * 1e3, min(td) * 1e3))
    dec.close()
