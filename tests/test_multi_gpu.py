"""Several GPUs behind one srt_render (csrc/srt_multi.cu): both ways in, both partitions, against the oracle.

Needs two GPUs (gpurun --gpus 2): skipped on a single-GPU box.  The same protocol is exercised on CPUs with gloo
(world_size 2) by tests/test_frame_split_gloo.py.
  * srt_create_multi: one process, a head context driving two devices
  * srt_group_join:   one process per GPU, the group id handed over by the launcher (a file here, torch.distributed in
                      bench.py)
  * SRT_SPLIT_BANDS: the gathered image has the oracle's BITS after every call of a progressive sequence (no reset in
    between: nothing is counted twice); SRT_SPLIT_SAMPLES: the same picture within 1e-4 relative
"""
import os
import sys
import time

import numpy as np
import pytest

import scenes
from test_parity_gpu import assert_bit_exact, cpu_render

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs (gpurun --gpus 2)")

W, H, FRAMES = 320, 200, 3


def _setup(pkg, ctx, split):
    scene = scenes.rubik_scene()
    unit, uniform = scenes.noise_tables(W, H)
    ctx.set_option(pkg.OPT_SPLIT, split)
    ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_REFERENCE)
    ctx.upload_scene(scene)
    ctx.upload_noise(unit, uniform)
    ctx.set_lights(scenes.model_lights())
    ctx.resize(W, H)
    ctx.set_params(pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
    return scene


def _check(pkg, oracle, split, acc, out, scene, n_frames):
    oacc, oout, _, _ = cpu_render(oracle, W, H, True, n_frames, scene=scene)
    if split == pkg.SPLIT_BANDS:
        assert_bit_exact(acc[..., :3], oacc[..., :3])
        assert np.array_equal(out, oout)
    else:
        assert np.allclose(acc[..., :3], oacc[..., :3], rtol=1e-4, atol=1e-6)
        assert (np.abs(out.astype(int) - oout.astype(int)) <= 1).all()
    assert (acc[..., 3] == 1.0).all()


@needs2
@pytest.mark.parametrize("split", ["bands", "samples"])
def test_head_context_two_devices_progressive(pkg, oracle, split):
    split = pkg.SPLIT_BANDS if split == "bands" else pkg.SPLIT_SAMPLES
    ctx = pkg.Context(devices=[0, 1])
    try:
        scene = _setup(pkg, ctx, split)
        ctx.reset_counters()
        for call in range(2):  # no reset in between: the reference's progressive loop
            ctx.render(FRAMES)
            _check(pkg, oracle, split, ctx.read_accum(), ctx.read_output(), scene, FRAMES * (call + 1))
        cnt = ctx.counters()
        assert cnt["paths"] == W * H * FRAMES * 2  # the two devices' shares add up to the whole job
        ms, n = ctx.collective_ms()
        assert n == 2 and ms > 0.0
        # a reset dispatch, then one more call: the picture starts over on every device
        p = pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1, reset=1)
        ctx.set_params(p)
        ctx.render(1)
        ctx.set_params(pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
        ctx.render(FRAMES)
        _check(pkg, oracle, split, ctx.read_accum(), ctx.read_output(), scene, FRAMES)
    finally:
        ctx.close()


def _rank_worker(rank, world, id_path, out_path, split_name):
    sys.path.insert(0, scenes.ROOT)
    sys.path.insert(0, os.path.join(scenes.ROOT, "tests"))
    import __graft_entry__ as g
    pkg = g.load_package()
    ctx = pkg.Context(rank)
    if rank == 0:
        with open(id_path + ".tmp", "wb") as f:
            f.write(pkg.group_unique_id())
        os.replace(id_path + ".tmp", id_path)
    t0 = time.time()
    while not os.path.exists(id_path):
        if time.time() - t0 > 120:
            raise RuntimeError("no group id")
        time.sleep(0.01)
    with open(id_path, "rb") as f:
        gid = f.read()
    ctx.group_join(rank, world, gid)
    split = pkg.SPLIT_BANDS if split_name == "bands" else pkg.SPLIT_SAMPLES
    _setup(pkg, ctx, split)
    shown = {}
    for call in range(2):
        ctx.render(FRAMES)
        if rank == 0:
            shown["acc%d" % call] = ctx.read_accum()
            shown["out%d" % call] = ctx.read_output()
    if rank == 0:
        np.savez(out_path, **shown)
    ctx.close()


@needs2
@pytest.mark.parametrize("split", ["bands", "samples"])
def test_one_process_per_gpu_group_join(pkg, oracle, tmp_path, split):
    import torch.multiprocessing as mp
    out = str(tmp_path / "shown.npz")
    mp.spawn(_rank_worker, args=(2, str(tmp_path / "group.id"), out, split), nprocs=2, join=True)
    shown = np.load(out)
    scene = scenes.rubik_scene()
    for call in range(2):
        _check(pkg, oracle, pkg.SPLIT_BANDS if split == "bands" else pkg.SPLIT_SAMPLES, shown["acc%d" % call],
               shown["out%d" % call], scene, FRAMES * (call + 1))


def test_group_calls_fail_cleanly_outside_a_group(pkg, ctx):
    with pytest.raises(pkg.SrtError, match="not in a group"):
        ctx.set_option(pkg.OPT_SPLIT, pkg.SPLIT_SAMPLES)
    assert ctx.collective_ms() == (0.0, 0)
    with pytest.raises(pkg.SrtError, match="rank 3 of world 2"):
        ctx.group_join(3, 2, bytes(pkg.GROUP_ID_BYTES))
