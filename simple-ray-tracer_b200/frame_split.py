"""Multi-GPU split of one frame: who renders which rows, and the one exchange per frame.

The shader's invocations are independent per pixel (each reads shared read-only scene data and touches only its own
accumulation texel, raytrace_compute.glsl:388,404-406), so a frame shards with no data-path exchange until the
picture is needed:
  * rank r of W renders the 8-row bands b with b % W == r (interleaved bands keep loads even when the model
    covers only part of the picture); srt_set_stripe(r, W) tells the kernel.  Rows a rank does not own stay exactly
    0.0 in its accumulation image.
  * one SUM reduce of the fp32 accumulation image to rank 0 per reported frame; x + 0.0 == x in IEEE arithmetic, so
    the reduced image has the bits of a single-GPU render.  Rank 0 then runs the tonemap (srt_tonemap).
The reduce goes through torch.distributed (NCCL over NVLink on GPUs; gloo in the CPU tests), one process per GPU.
"""
import numpy as np

BAND_ROWS = 8  # the shader's work group height (raytrace_compute.glsl:12) and the kernel's band unit


def band_count(height):
    if height % BAND_ROWS:
        raise ValueError("height must be a multiple of %d" % BAND_ROWS)
    return height // BAND_ROWS


def bands_of_rank(height, rank, world):
    """8-row bands rendered by `rank`: b, b+world, b+2*world, ... (what srt_set_stripe selects)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("rank %d of world %d" % (rank, world))
    return list(range(rank, band_count(height), world))


def rows_of_rank(height, rank, world):
    """[(y0, y1), ...] row ranges of the bands of `rank`."""
    return [(b * BAND_ROWS, (b + 1) * BAND_ROWS) for b in bands_of_rank(height, rank, world)]


def owner_of_row(y, world):
    return (y // BAND_ROWS) % world


def sample_ranges(n_frames, world, first_frame=1):
    """The sample-range half of the split (BASELINE.json configs[3]): rank g renders frames
    [first + g*n/W, first + (g+1)*n/W).  Sums of different ranks are added by the reduce, so this split reorders fp32
    additions (within tolerance, not bit-exact); the band split is the default."""
    edges = [first_frame + (n_frames * g) // world for g in range(world + 1)]
    return [(edges[g], edges[g + 1] - edges[g]) for g in range(world)]


def reduce_accum(accum, dst=0, group=None):
    """SUM-reduces the RGBA32F accumulation image (a torch tensor on this rank's device) to rank `dst`, in place."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def merge_on_host(parts):
    """Reference implementation of the reduce for tests: sum of per-rank images (numpy, fp32)."""
    out = np.zeros_like(parts[0])
    for p in parts:
        out = out + p
    return out
