"""Multi-GPU plumbing of the hot path: interleaved block ownership and the film sum.

The image is cut into a raster grid of blocks like the reference's tile grid (ImageSplitter, /root/reference/src/common/imagesplitter.cc:37-50), of at most
8 x 8 pixels; block (tx, ty) belongs to rank (tx + ty * stride) % world with stride = the integer nearest 0.618 * world that is coprime with it — the rule
csrc/integrator.cu buildPixelList applies on the device side (it keeps the partitions' costs within about 2 % of each other on the bench frame).  The film is a
linear accumulator (ImageFilm::imageFilmLoadAllInFolder sums col_ and weight_, /root/reference/src/common/imagefilm.cc:1506-1535),
so the merge is ONE sum-reduction of w*h*5 floats — a plain gather would lose filter footprints that spill across tile borders.
"""
import numpy as np


def tile_grid(width, height, tile_size):
    return (width + tile_size - 1) // tile_size, (height + tile_size - 1) // tile_size


PARTITION_UNIT = 8   # YG_PARTITION_UNIT of csrc/integrator.cu


def partition_stride(world):
    import math
    s = max(1, int(math.floor(0.618 * world + 0.5)))
    while math.gcd(s, world) != 1:
        s += 1
    return s % world or 1


def tile_owner(tx, ty, world):
    return (tx + ty * partition_stride(world)) % max(1, world)


def owner_map(width, height, tile_size, world):
    """[h,w] int array: rank that renders each pixel."""
    world = max(1, world)
    unit = min(tile_size, PARTITION_UNIT) if world > 1 else tile_size
    ys, xs = np.mgrid[0:height, 0:width]
    return ((xs // unit) + (ys // unit) * partition_stride(world)) % world


def reduce_film(film, dst=0, group=None):
    """Sum the per-rank film accumulators onto rank `dst` (torch tensor, CPU/gloo or CUDA/NCCL). In place; returns film."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(film, dst, op=dist.ReduceOp.SUM, group=group)
    return film
