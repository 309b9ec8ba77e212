"""CPU test of the N>1 path: two gloo ranks shard one batch (no data-path collective), each decodes its shard (CPU
emulation of the kernels stands in for the GPU here), results are gathered only to verify coverage and parity."""
import hashlib
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import util
from nvimagecodec_b200.sharding import estimate_cost, shard_batch

pytestmark = pytest.mark.skipif(not util.have_ref(), reason="corpus needs oracle/_ref's libjpeg-turbo encoder")


def test_shard_batch_properties():
    costs = [5, 1, 9, 3, 3, 7, 2, 8]
    for world in (1, 2, 3, 4, 8):
        shards = shard_batch(costs, world)
        assert sorted(i for s in shards for i in s) == list(range(len(costs)))
        loads = [sum(costs[i] for i in s) for s in shards]
        assert max(loads) - min(loads) <= max(costs)
    assert shard_batch([], 2) == [[], []]


def _worker(rank, world, port_no, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
    import emu
    from util import make_jpeg, port
    sizes = [(160 + 16 * i, 120 + 8 * i) for i in range(6)]
    datas = [make_jpeg(900 + i, w, h) for i, (w, h) in enumerate(sizes)]
    costs = [estimate_cost(len(d), w, h) for d, (w, h) in zip(datas, sizes)]
    mine = shard_batch(costs, world)[rank]
    outs, st, _, _ = emu.decode_batch([datas[i] for i in mine], [sizes[i] for i in mine], 4)
    digests = {i: hashlib.sha256(o.tobytes()).hexdigest() for i, o in zip(mine, outs)}
    gathered = [None] * world
    dist.all_gather_object(gathered, (mine, st, digests))
    if rank == 0:
        want = {i: hashlib.sha256(port.decode(d, port.FMT_I_RGB, True).tobytes()).hexdigest() for i, d in enumerate(datas)}
        q.put((gathered, want))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_decode():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port_no = 29500 + (os.getpid() % 1000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port_no, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered, want = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    seen = {}
    for mine, st, digests in gathered:
        assert all(s == 0 for s in st)
        seen.update(digests)
    assert seen == want
