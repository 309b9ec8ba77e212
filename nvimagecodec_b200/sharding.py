"""Batch sharding for multi-GPU decode: images are independent, so a batch is split across ranks with no
data-path collective -- one process, decoder, stream set and pinned staging per GPU.  Greedy longest-processing-time
partition on an estimated cost (compressed bytes + pixels), deterministic on every rank without communication."""


def estimate_cost(jpeg_bytes, width, height):
    # Huffman work scales with compressed bytes (~8 symbols/pixel-byte), reconstruction with pixels
    return 8 * jpeg_bytes + width * height


def shard_batch(costs, world_size):
    """-> list (per rank) of sorted index lists; every index appears exactly once."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    loads = [0] * world_size
    shards = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (loads[k], k))
        shards[r].append(i)
        loads[r] += costs[i]
    return [sorted(s) for s in shards]


def my_shard(costs, world_size, rank):
    return shard_batch(costs, world_size)[rank]


def shard_batch_native(jpeg_sizes, widths, heights, num_devices):
    """the library's deal (b200jpeg_shard_batch, the one b200jpeg_multi_decode_batch callers use) -> device index per image;
    the same greedy partition as shard_batch"""
    import ctypes as C
    from . import _capi
    n = len(jpeg_sizes)
    out = (C.c_int32 * max(n, 1))()
    rc = _capi.lib().b200jpeg_shard_batch(n, (C.c_size_t * max(n, 1))(*jpeg_sizes), (C.c_int32 * max(n, 1))(*widths),
                                          (C.c_int32 * max(n, 1))(*heights), num_devices, out)
    if rc != 0:
        raise RuntimeError("b200jpeg_shard_batch: %s" % _capi.status_string(rc))
    return list(out)[:n]
