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
