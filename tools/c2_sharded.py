#!/usr/bin/env python
"""North-star config 2, sharded: N ImageNet-shaped JPEGs (variable size ~500x375, 4:2:0, q75) dealt over the ranks of a
torch.distributed.run job by nvimagecodec_b200.sharding.shard_batch -- one process, one decoder, one stream set per rank, no
collective on the data path (gloo carries only the verification and the timings).  Every rank decodes its share on its GPU
(rank % device_count, so the script also runs with more ranks than GPUs), compares every image with the compiled reference
(oracle/_ref, CPU) and rank 0 prints one JSON line with the strong-scaling time (fixed N images, max over ranks).

    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 tools/c2_sharded.py --images 256
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=256)
    ap.add_argument("--distinct", type=int, default=32)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    import torch
    import torch.distributed as dist
    from util import make_jpeg, ref
    from nvimagecodec_b200 import parse_header
    from nvimagecodec_b200.decoder import Decoder
    from nvimagecodec_b200.sharding import estimate_cost, shard_batch
    if world > 1:
        dist.init_process_group("gloo")
    dev_index = int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count()
    torch.cuda.set_device(dev_index)
    rng = np.random.default_rng(2)
    specs = [(int(rng.integers(320, 641)), int(rng.integers(240, 501))) for _ in range(args.distinct)]
    distinct = [make_jpeg(2000 + i, w, h, (2, 2), 75) for i, (w, h) in enumerate(specs)]  # the same bytes on every rank
    datas = [distinct[i % args.distinct] for i in range(args.images)]
    dims = [specs[i % args.distinct] for i in range(args.images)]
    costs = [estimate_cost(len(d), w, h) for d, (w, h) in zip(datas, dims)]
    mine = shard_batch(costs, world)[rank]
    for i in mine[:4]:
        info = parse_header(datas[i])
        assert (info.width, info.height) == dims[i]
    dec = Decoder(dev_index, options=":fancy_upsampling=1")
    outs = [torch.zeros((dims[i][1], dims[i][0] * 3), dtype=torch.uint8, device="cuda") for i in mine]
    my_datas = [datas[i] for i in mine]
    ptrs, strides, fmts = [o.data_ptr() for o in outs], [dims[i][0] * 3 for i in mine], [4] * len(mine)
    st = dec.decode_into(my_datas, ptrs, strides, fmts)
    torch.cuda.synchronize()
    assert all(s == 0 for s in st), st
    bad = 0
    if not args.no_check:
        seen = {}
        rdec = ref.Session.shared().decoder(options=":fancy_upsampling=1")
        for k, i in enumerate(mine):
            j = i % args.distinct
            if j not in seen:
                seen[j] = rdec.decode_host([distinct[j]])[0][0]
            w, h = dims[i]
            if not np.array_equal(outs[k].cpu().numpy().reshape(h, w, 3), seen[j]):
                bad += 1
    # strong scaling: host JPEG bytes -> device pixels of this rank's share, all ranks at once
    for _ in range(2):
        dec.decode_into(my_datas, ptrs, strides, fmts)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        dec.decode_into(my_datas, ptrs, strides, fmts)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / args.steps * 1e3
    report = (rank, len(mine), bad, ms, sum(costs[i] for i in mine))
    gathered = [report]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, report)
    if rank == 0:
        total_bad = sum(g[2] for g in gathered)
        covered = sum(g[1] for g in gathered)
        pixels = sum(w * h for (w, h) in dims)
        step_ms = max(g[3] for g in gathered)
        print(json.dumps({"workload": "C2: %d variable-size ~500x375 4:2:0 q75 JPEGs -> RGB, sharded over %d ranks (%d GPUs)" %
                          (args.images, world, min(world, torch.cuda.device_count())), "scaling": "strong", "ranks": world,
                          "ms_per_batch": step_ms, "images_per_s": args.images / (step_ms / 1e3), "mpixel_per_s": pixels / 1e6 / (step_ms / 1e3),
                          "per_rank_ms": [round(g[3], 3) for g in gathered], "per_rank_images": [g[1] for g in gathered],
                          "cost_imbalance": max(g[4] for g in gathered) / (sum(g[4] for g in gathered) / world),
                          "timing": "wall clock, host JPEG bytes -> device pixels (b200jpeg_decode_batch), max over ranks",
                          "mismatches": total_bad}))
        assert covered == args.images and total_bad == 0, (covered, total_bad)
        print("C2 sharded OK: %d images over %d ranks, every image bit-exact against the reference" % (args.images, world))
    dec.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
