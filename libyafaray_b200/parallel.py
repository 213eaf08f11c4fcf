"""Multi-GPU plumbing of the hot path: interleaved tile ownership and the film sum.

The image is cut into the reference's tile grid (ImageSplitter raster order, /root/reference/src/common/imagesplitter.cc:37-50);
tile t belongs to rank t % world (the same rule csrc/integrator.cu buildPixelList applies on the device side).  The film is a
linear accumulator (ImageFilm::imageFilmLoadAllInFolder sums col_ and weight_, /root/reference/src/common/imagefilm.cc:1506-1535),
so the merge is ONE sum-reduction of w*h*5 floats — a plain gather would lose filter footprints that spill across tile borders.
"""
import numpy as np


def tile_grid(width, height, tile_size):
    return (width + tile_size - 1) // tile_size, (height + tile_size - 1) // tile_size


def tile_owner(tile_index, world):
    return tile_index % max(1, world)


def owner_map(width, height, tile_size, world):
    """[h,w] int array: rank that renders each pixel."""
    ntx, _ = tile_grid(width, height, tile_size)
    ys, xs = np.mgrid[0:height, 0:width]
    return ((ys // tile_size) * ntx + (xs // tile_size)) % max(1, world)


def reduce_film(film, dst=0, group=None):
    """Sum the per-rank film accumulators onto rank `dst` (torch tensor, CPU/gloo or CUDA/NCCL). In place; returns film."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(film, dst, op=dist.ReduceOp.SUM, group=group)
    return film
