"""World-size-2 test of the multi-GPU protocol on CPU (gloo): the two partitions of csrc/srt_multi.cu (interleaved 8-row
bands gathered on rank 0; sample ranges summed on rank 0) as frame_split.py states them, over two progressive calls
without a reset.  The per-rank renderer here is the CPU oracle standing in for the kernels (test infrastructure); the
same protocol over NCCL inside srt_render is tested on GPUs by tests/test_multi_gpu.py."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import scenes


def _split():
    import __graft_entry__ as g
    g.load_package()
    import importlib
    return importlib.import_module("simple_ray_tracer_b200.frame_split")


def test_band_partition_covers_every_row_once():
    fs = _split()
    for H, world in ((1080 // 8 * 8, 8), (56, 2), (64, 3), (8, 4)):
        seen = np.zeros(H, int)
        for r in range(world):
            for y0, y1 in fs.rows_of_rank(H, r, world):
                seen[y0:y1] += 1
                assert all(fs.owner_of_row(y, world) == r for y in range(y0, y1))
        assert (seen == 1).all()
    assert fs.sample_ranges(1024, 8) == [(1 + 128 * g, 128) for g in range(8)]
    assert fs.sample_ranges(3, 2) == [(1, 1), (2, 2)] and fs.sample_ranges(1, 4, first_frame=9) == [(9, 0), (9, 0), (9, 0), (9, 1)]
    assert sum(n for _, n in fs.sample_ranges(10, 4)) == 10
    with pytest.raises(ValueError):
        fs.rows_of_rank(60, 0, 2)


def _worker(rank, world, port, W, H, frames, out_path):
    """Two srt_render-like calls of `frames` frames each WITHOUT a reset in between (the reference's progressive loop),
    through both partitions; rank 0 saves what it would show after each call."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, scenes.ROOT)
    sys.path.insert(0, os.path.join(scenes.ROOT, "tests"))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    fs = _split()
    import __graft_entry__ as g
    O = g.load_oracle()
    unit, uniform = scenes.noise_tables(W, H)
    lights, cam, scene = scenes.model_lights(), scenes.reset_camera(True), scenes.rubik_scene()
    shown = {}
    # image bands: this rank's own image only ever holds its own rows
    acc = np.zeros((H, W, 4), np.float32)
    for call in range(2):
        p = O.make_params(W, H, cam, True, 1 + call * frames, len(lights), 1)
        for rows in fs.rows_of_rank(H, rank, world):
            O.render(p, frames, scene, unit, uniform, lights, accum=acc, rows=rows, want_out=False, want_ids=False, n_threads=1)
        gathered = fs.gather_bands(acc, rank, world)
        if rank == 0:
            shown["bands%d" % call] = gathered.copy()
    own = np.array([fs.owner_of_row(y, world) == rank for y in range(H)])
    assert not acc[~own][..., :3].any()
    # sample ranges: this rank's own image holds the sum of ITS frames over the whole picture
    acc = np.zeros((H, W, 4), np.float32)
    for call in range(2):
        first, n = fs.sample_ranges(frames, world, first_frame=1 + call * frames)[rank]
        if n:
            p = O.make_params(W, H, cam, True, first, len(lights), 1)
            O.render(p, n, scene, unit, uniform, lights, accum=acc, want_out=False, want_ids=False, n_threads=1)
        total = fs.reduce_samples(acc)
        if rank == 0:
            shown["samples%d" % call] = total.copy()
    if rank == 0:
        np.savez(out_path, **shown)
    dist.barrier()
    dist.destroy_process_group()


def _bits(a):
    a = np.ascontiguousarray(a[..., :3], np.float32)
    return np.where(np.isnan(a), np.uint32(0x7FC00000), a.view(np.uint32))


def test_two_ranks_progressive_calls_both_partitions(oracle, tmp_path):
    """World size 2 on CPU (gloo): after each of two calls without a reset, the band split shows the bits of a single
    render of the frames so far (nothing counted twice), the sample split the same picture within 1e-4 relative."""
    W, H, frames, world = 64, 48, 3, 2
    out = str(tmp_path / "shown.npz")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, W, H, frames, out), nprocs=world, join=True)
    shown = np.load(out)
    unit, uniform = scenes.noise_tables(W, H)
    lights, cam, scene = scenes.model_lights(), scenes.reset_camera(True), scenes.rubik_scene()
    p = oracle.make_params(W, H, cam, True, 1, len(lights), 1)
    for call in range(2):
        whole, _, _, _ = oracle.render(p, frames * (call + 1), scene, unit, uniform, lights, want_out=False, want_ids=False)
        assert np.array_equal(_bits(shown["bands%d" % call]), _bits(whole))
        a, b = shown["samples%d" % call][..., :3], whole[..., :3]
        assert np.allclose(a, b, rtol=1e-4, atol=1e-6)
    assert not np.array_equal(_bits(shown["samples1"]), _bits(shown["samples0"]))


def test_pack_and_unpack_rows_round_trip():
    fs = _split()
    rng = np.random.default_rng(0)
    for H, world in ((48, 2), (64, 3), (1080 // 8 * 8, 8)):
        img = rng.random((H, 16, 4), dtype=np.float32)
        out = np.zeros_like(img)
        total = 0
        for r in range(world):
            packed = fs.pack_rows(img, r, world)
            total += packed.shape[0]
            fs.unpack_rows(out, packed, r, world)
        assert total == H and np.array_equal(out, img)
