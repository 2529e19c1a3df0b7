"""World-size-2 test of the multi-GPU host logic on CPU (gloo): interleaved 8-row bands + one SUM reduce of the
accumulation image give the bits of a single-process render.  The per-rank renderer here is the CPU oracle standing in
for the kernel (test infrastructure); the partition and the reduce are the product code of frame_split.py."""
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
    assert sum(n for _, n in fs.sample_ranges(10, 4)) == 10
    with pytest.raises(ValueError):
        fs.rows_of_rank(60, 0, 2)


def _worker(rank, world, port, W, H, frames, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, scenes.ROOT)
    sys.path.insert(0, os.path.join(scenes.ROOT, "tests"))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    fs = _split()
    import __graft_entry__ as g
    O = g.load_oracle()
    unit, uniform = scenes.noise_tables(W, H)
    lights, cam, scene = scenes.model_lights(), scenes.reset_camera(True), scenes.rubik_scene()
    p = O.make_params(W, H, cam, True, 1, len(lights), 1)
    acc = np.zeros((H, W, 4), np.float32)
    for rows in fs.rows_of_rank(H, rank, world):
        O.render(p, frames, scene, unit, uniform, lights, accum=acc, rows=rows, want_out=False, want_ids=False, n_threads=1)
    t = torch.from_numpy(acc)
    fs.reduce_accum(t, dst=0)
    if rank == 0:
        np.save(out_path, t.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_reduce_to_single_render_bits(oracle, tmp_path):
    W, H, frames, world = 64, 48, 2, 2
    out = str(tmp_path / "reduced.npy")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, W, H, frames, out), nprocs=world, join=True)
    reduced = np.load(out)
    unit, uniform = scenes.noise_tables(W, H)
    lights, cam, scene = scenes.model_lights(), scenes.reset_camera(True), scenes.rubik_scene()
    p = oracle.make_params(W, H, cam, True, 1, len(lights), 1)
    whole, _, _, _ = oracle.render(p, frames, scene, unit, uniform, lights, want_out=False, want_ids=False)
    a, b = reduced[..., :3], whole[..., :3]
    assert np.array_equal(np.where(np.isnan(a), 0, a).view(np.uint32), np.where(np.isnan(b), 0, b).view(np.uint32))
