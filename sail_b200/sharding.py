"""Multi-GPU work split for the path tracer; one process per GPU, torch.distributed for
the plumbing.  The reference is single-GPU (one WebGL context, /root/reference/src/core/webgl.js:215-227);
what makes the split legal is in the shader: a pixel reads only its own accumulator texel
(src/shader/main/fstrace.glsl:13-14) and frames are independent given (matrix, timeSinceStart).

Two splits, neither needs a collective on the data path:
  * sample sharding -- rank g traces frame indices k = g (mod G) into an FP32 *sum*; one
    ``reduce(SUM)`` of W*H*3 floats when an image is wanted, then a scale by 1/N;
  * tile sharding   -- rank g traces all frames for an interleaved set of row bands; one ``gather`` of
    those rows, no arithmetic (bit-identical to a single-GPU render).
"""
import numpy as np


def frames_for_rank(first, count, rank, world):
    """Frame indices in [first, first+count) owned by ``rank``: k = rank (mod world)."""
    return [k for k in range(first, first + count) if k % world == rank]


def row_bands(height, world, band=16):
    """Interleaved row bands: band b (rows [b*band, (b+1)*band)) belongs to rank b % world.  Returns a
    list, per rank, of (row_begin, row_end) tuples.  Interleaving balances glass/metal regions."""
    out = [[] for _ in range(world)]
    b = 0
    for y0 in range(0, height, band):
        out[b % world].append((y0, min(height, y0 + band)))
        b += 1
    return out


def contiguous_rows(height, rank, world, align=8):
    """One contiguous stripe per rank, aligned to the CTA tile height."""
    per = -(-height // world)
    per = -(-per // align) * align
    return min(height, rank * per), min(height, (rank + 1) * per)


def reduce_sum(tensor, dst=0, group=None):
    """In-place SUM-reduce of per-rank partial sums onto ``dst`` (NCCL over NVLink on GPUs, gloo on CPU)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(tensor, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return tensor


def gather_rows(image, my_rows, all_rows, dst=0, group=None):
    """Tile sharding: assemble the full image on ``dst`` from each rank's rows.  ``image`` is this rank's
    full-size (H, W, C) tensor of which only ``my_rows`` are valid; ``all_rows[r]`` lists rank r's
    (begin, end) bands.  Returns the assembled tensor on ``dst`` (None elsewhere)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return image
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    counts = [sum(b - a for a, b in all_rows[r]) for r in range(world)]
    most = max(counts)
    # NCCL's gather wants one size for every rank: ranks with fewer rows pad up to the largest share
    mine = torch.zeros((most,) + tuple(image.shape[1:]), dtype=image.dtype, device=image.device)
    off = 0
    for a, b in my_rows:
        mine[off:off + (b - a)] = image[a:b]
        off += b - a
    if rank == dst:
        bufs = [torch.empty_like(mine) for _ in range(world)]
        dist.gather(mine, bufs, dst=dst, group=group)
        full = torch.empty_like(image)
        for r in range(world):
            off = 0
            for a, b in all_rows[r]:
                full[a:b] = bufs[r][off:off + (b - a)]
                off += b - a
        return full
    dist.gather(mine, None, dst=dst, group=group)
    return None


class DeviceBuffer:
    """``__cuda_array_interface__`` view of a buffer owned by a sail_ctx, so torch can wrap it without a copy:
    ``torch.as_tensor(DeviceBuffer(ctx.device_ptr(AOV_COLOR), (H, W, 3)), device='cuda')``."""

    def __init__(self, ptr, shape, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def mean_from_sum(total, nframes):
    return np.asarray(total, np.float32) * np.float32(1.0 / nframes)
