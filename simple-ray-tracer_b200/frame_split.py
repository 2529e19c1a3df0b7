"""Multi-GPU split of one frame, host-side statement of what the library does (csrc/srt_multi.cu).

The shader's invocations are independent per pixel (each reads shared read-only scene data and touches only its own
accumulation texel, raytrace_compute.glsl:388,404-406) and samples are independent given accumFrames (:400-401,412), so a
frame shards with no data-path exchange until the picture is needed.  The product path is the C-ABI
(srt_group_join / srt_create_multi + SRT_OPT_SPLIT): partition and collective run inside srt_render over NCCL.  This
module states the same two partitions and the same exchange over torch.distributed tensors so that the protocol can be
run on CPUs (gloo, world_size 2) in tests/test_frame_split_gloo.py:
  * image bands (default): rank r of W renders the 8-row bands b with b % W == r (interleaved bands keep loads even when
    the model covers only part of the picture).  Each rank packs ITS rows and rank 0 gathers them (P*16/W bytes per rank)
    into the gathered image: the bits of a single-GPU render.  A rank's own image never receives other ranks' rows, so
    progressive rendering without resets counts nothing twice.
  * sample ranges: rank r renders the whole image for frames [first + n*r/W, first + n*(r+1)/W); the images are summed on
    rank 0 (one reduce): fp32 additions re-ordered, equal within tolerance, not in bits.
"""
import numpy as np

BAND_ROWS = 8  # the shader's work group height (raytrace_compute.glsl:12) and the kernel's band unit


def band_count(height):
    if height % BAND_ROWS:
        raise ValueError("height must be a multiple of %d" % BAND_ROWS)
    return height // BAND_ROWS


def bands_of_rank(height, rank, world):
    """8-row bands rendered by `rank`: b, b+world, b+2*world, ... (what SRT_SPLIT_BANDS / srt_set_stripe select)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("rank %d of world %d" % (rank, world))
    return list(range(rank, band_count(height), world))


def rows_of_rank(height, rank, world):
    """[(y0, y1), ...] row ranges of the bands of `rank`."""
    return [(b * BAND_ROWS, (b + 1) * BAND_ROWS) for b in bands_of_rank(height, rank, world)]


def owner_of_row(y, world):
    return (y // BAND_ROWS) % world


def sample_ranges(n_frames, world, first_frame=1):
    """SRT_SPLIT_SAMPLES (BASELINE.json configs[3] "by image tile and spp"): rank g renders frames
    [first + n*g/W, first + n*(g+1)/W) -- (first frame, count) per rank, the rule of srt_group_before_render."""
    edges = [first_frame + (n_frames * g) // world for g in range(world + 1)]
    return [(edges[g], edges[g + 1] - edges[g]) for g in range(world)]


def pack_rows(accum, rank, world):
    """The rows `rank` owns, packed band after band (pack_bands_kernel): accum [H, W, 4] -> [rows, W, 4]."""
    return np.concatenate([accum[y0:y1] for y0, y1 in rows_of_rank(accum.shape[0], rank, world)], axis=0)


def unpack_rows(gathered, packed, rank, world):
    """Scatters one rank's packed rows into the gathered image (pack_bands_kernel, unpack direction)."""
    at = 0
    for y0, y1 in rows_of_rank(gathered.shape[0], rank, world):
        gathered[y0:y1] = packed[at:at + (y1 - y0)]
        at += y1 - y0
    return gathered


def gather_bands(accum, rank, world, group=None):
    """The band split's exchange: every rank sends its own rows, rank 0 returns the gathered image (others None).
    `accum` is this rank's numpy image; only its own rows are read."""
    import torch
    import torch.distributed as dist
    mine = torch.from_numpy(np.ascontiguousarray(pack_rows(accum, rank, world)))
    if world == 1:
        return unpack_rows(np.zeros_like(accum), mine.numpy(), 0, 1)
    rows = [sum(y1 - y0 for y0, y1 in rows_of_rank(accum.shape[0], r, world)) for r in range(world)]
    if rank == 0:
        gathered = unpack_rows(np.zeros_like(accum), mine.numpy(), 0, world)
        for r in range(1, world):
            buf = torch.empty((rows[r],) + accum.shape[1:], dtype=mine.dtype)
            dist.recv(buf, src=r, group=group)
            unpack_rows(gathered, buf.numpy(), r, world)
        return gathered
    dist.send(mine, dst=0, group=group)
    return None


def reduce_samples(accum, group=None):
    """The sample split's exchange: SUM of the ranks' images on rank 0 (others None); `accum` itself is not changed."""
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(accum.copy())
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(t, dst=0, op=dist.ReduceOp.SUM, group=group)
        return t.numpy() if dist.get_rank(group) == 0 else None
    return t.numpy()


def merge_on_host(parts):
    """Reference implementation of the reduce for tests: sum of per-rank images (numpy, fp32)."""
    out = np.zeros_like(parts[0])
    for p in parts:
        out = out + p
    return out
