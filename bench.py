"""bench.py -- Mrays/s (path segments) of the path-tracing hot path on 1/2/4/8 B200, CPU oracle beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload rubik|spheres]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

A "step" is one pass of the hot path over the workload: 64 progressive frames (64 spp) of the reference's Rubik
scene at 1920x1080 (BASELINE.json metric "1080p 64spp" on configs[1]'s scene; the Airplane OBJ of configs[2]
is absent from the reference mount) rendered by ONE srt_render call (about 36 kernel launches).  With N GPUs the library splits
the frame itself (srt_group_join; --split bands: interleaved 8-row bands, bit-identical to one GPU, each rank's rows
gathered on rank 0; --split samples: every rank renders the whole image for 1/N of the frames, one ncclReduce) with one
NCCL collective per step inside srt_render (strong scaling: total work fixed).
  value : path segments / s, inputs resident in HBM, CUDA events on the launching stream, max over ranks
  e2e   : the same through the C-ABI with HOST buffers: scene + noise + lights + uniforms uploaded, 64 frames
          rendered, (reduced,) RGBA8 image read back -- every step, wall clock, max over ranks
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import scenes  # noqa: E402  (deterministic inputs; no reference access at run time)

WORKLOADS = {
    # name: (show_model, width, height, frames per step, maxDepth (0 = the shader's 5), description)
    "rubik": (True, 1920, 1080, 64, 0,
              "objects/Rubik (1188 tris, reference-built BVH, 6 lights) 1920x1080, 64 spp per step, maxDepth 5"),
    "spheres": (False, 1920, 1080, 64, 0, "built-in 5-sphere scene, 2 lights, 1920x1080, 64 spp per step"),
    # BASELINE.json configs[2]; the Airplane .obj is absent from the reference mount, so a documented procedural mesh
    # with the Airplane .mtl materials stands in (tests/scenes.py:write_airplane_standin) -- never under the Airplane name
    "airplane-standin": (True, 1920, 1080, 64, 0,
                         "airplane-standin (procedural airliner mesh, ~110 k tris, the 6 materials of 11803_Airplane_v1_l1.mtl, "
                         "BVH by the reference algorithm) 1920x1080, 64 spp per step, maxDepth 5"),
    # BASELINE.json configs[3]: the same stand-in at 3840x2160, 1024 spp, meant for 2/4/8 GPUs (image bands + NCCL reduce)
    "airplane-standin-4k": (True, 3840, 2160, 1024, 0,
                            "airplane-standin (procedural airliner mesh, ~110 k tris, the 6 materials of 11803_Airplane_v1_l1.mtl, "
                            "BVH by the reference algorithm) 3840x2160, 1024 spp per step, maxDepth 5"),
    # BASELINE.json configs[4]
    "soup": (True, 1920, 1080, 4, 8,
             "synthetic random triangle soup (PCG32 0xC0FFEE, BVH by the reference algorithm) 1920x1080, 4 spp per step, "
             "maxDepth 8, camera (0,0,120) looking -z"),
}
METRIC = "Mrays/s (path segments) at 1080p 64spp"
UNIT = "Mrays/s"


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload_inputs(name, soup_tris=10_000_000, obj=None):
    import tempfile
    show_model, W, H, spp, max_depth, desc = WORKLOADS[name]
    unit, uniform = scenes.noise_tables(W, H)
    cam = scenes.reset_camera(show_model)
    scene = None
    if obj:
        # --obj <objects dir>/<NAME>/<NAME>.obj (LoadObject's layout, model_loader.cpp:14,24-26): the scene of the chosen
        # workload is replaced by this file, loaded by the product's own loader + BVH builder; the line carries the
        # file's own name (BASELINE.json configs[2]/[3] name objects/11803_Airplane_v1_l1, absent from the reference mount)
        path = os.path.abspath(obj)
        folder, base = os.path.split(path)
        obj_name = os.path.splitext(base)[0]
        if os.path.basename(folder) != obj_name or not os.path.isfile(path):
            raise SystemExit("--obj: expected <objects dir>/%s/%s.obj, got %s" % (obj_name, obj_name, obj))
        with tempfile.TemporaryDirectory() as d:
            scene = scenes.load_obj_scene(os.path.dirname(folder), obj_name, d)
        show_model = True
        desc = "objects/%s (%d tris, %d nodes, BVH by the reference algorithm, 6 lights) %dx%d, %d spp per step, maxDepth %d" % (
            obj_name, len(scene["triangles"]), len(scene["nodes"]), W, H, spp, max_depth or 5)
        cam = scenes.reset_camera(True)
    elif name == "rubik":
        scene = scenes.rubik_scene()
    elif name in ("airplane-standin", "airplane-standin-4k"):
        with tempfile.TemporaryDirectory() as d:
            obj, n_tri = scenes.write_airplane_standin(d, detail=128)
            scene = scenes.load_obj_scene(d, obj, d)
        desc = desc.replace("~110 k tris", "%d tris, %d nodes" % (len(scene["triangles"]), len(scene["nodes"])))
    elif name == "soup":
        with tempfile.TemporaryDirectory() as d:
            scene, info = scenes.soup_scene(soup_tris, d)
        desc = desc.replace("soup (", "soup of %d triangles, %d nodes, depth %d (" % (info["triangles"], info["nodes"], info["depth"]))
        cam = scenes.reset_camera(True, origin=(0.0, 0.0, 120.0))
    return {"show_model": show_model, "W": W, "H": H, "spp": spp, "desc": desc, "unit": unit, "uniform": uniform,
            "scene": scene, "max_depth": max_depth,
            "lights": scenes.model_lights() if show_model else scenes.sphere_lights(), "cam": cam}


def cpu_kind():
    """("reference", "shader_ref") when oracle/_ref/libshader_ref.so -- the reference's OWN shader text compiled as C++
    (oracle/shader_ref.cpp) -- travelled with the repo, else ("port", "oracle"): the hand-written restatement."""
    import __graft_entry__ as g
    O = g.load_oracle()
    return ("reference", "shader_ref") if O.shader_ref_available() else ("port", "oracle")


def oracle_run(inp, first_frame, n_frames, threads=0):
    """TEST-INFRASTRUCTURE leg: the reference's CPU implementation of the path (oracle/) timed on the host cores.
    Returns (Mseg/s, seconds, counters)."""
    import __graft_entry__ as g
    O = g.load_oracle()
    p = O.make_params(inp["W"], inp["H"], inp["cam"], inp["show_model"], first_frame, len(inp["lights"]),
                      1 if inp["show_model"] else 0, inp["max_depth"])
    t0 = time.perf_counter()
    _, _, _, cnt = O.render(p, n_frames, inp["scene"], inp["unit"], inp["uniform"], inp["lights"], want_out=False,
                            want_ids=False, n_threads=threads, impl=cpu_kind()[1])
    dt = time.perf_counter() - t0
    return cnt["segments"] / dt / 1e6, dt, cnt


def run_reference(args):
    """--impl reference: the reference's shader cannot run through a graphics API here (GLSL, no GL stack: "unavailable
    on this box"), so this arm times its CPU form on all host cores: the reference's own shader text compiled as C++
    (oracle/_ref/libshader_ref.so, kind "reference") when that file travelled with the repo, else the hand-written
    restatement pinned to it (kind "port").  Each step is a bounded sample of the workload: as many whole frames (1 spp
    each; per-frame cost is stationary) as fit about five seconds, sized from one calibration frame."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    inp = workload_inputs(args.workload, args.soup_tris, args.obj)
    cores = os.cpu_count() or 1
    _, t1, _ = oracle_run(inp, 1, 1)  # calibration (also the first warm-up)
    per_step = max(1, min(inp["spp"], int(5.0 / max(t1, 1e-3))))
    for w in range(max(args.warmup - 1, 0)):
        oracle_run(inp, 1 + w, per_step)
    secs, segs = 0.0, 0
    for k in range(args.steps):
        _, dt, cnt = oracle_run(inp, 1 + (args.warmup + k) * per_step, per_step)
        secs += dt
        segs += cnt["segments"]
    value = segs / secs / 1e6
    sample = "%d frame(s) (1 spp each) of the %d-spp workload per step, whole %dx%d image" % (per_step, inp["spp"], inp["W"], inp["H"])
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": inp["desc"], "step": sample, "host_threads": cores},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": cpu_kind()[0], "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = g.load_package()
    inp = workload_inputs(args.workload, args.soup_tris, args.obj)
    W, H, spp, show_model = inp["W"], inp["H"], inp["spp"], inp["show_model"]
    mode = pkg.TRAV_REFERENCE if args.traversal == "reference" else pkg.TRAV_ORDERED

    ctx = pkg.Context(local)
    if world > 1:
        # the product's own multi-GPU layer (srt_group_join: partition + one NCCL collective inside srt_render);
        # torch.distributed is only the launcher's channel for the 128-byte group id
        idt = torch.zeros(pkg.GROUP_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.tensor(list(pkg.group_unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(idt, 0)
        ctx.group_join(rank, world, bytes(idt.cpu().numpy().tobytes()))
        ctx.set_option(pkg.OPT_SPLIT, pkg.SPLIT_SAMPLES if args.split == "samples" else pkg.SPLIT_BANDS)
    stream = torch.cuda.Stream()  # a real (non-NULL) stream: kernels, events and NCCL ordering all go through it
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_option(pkg.OPT_SCHEDULE, {"wavefront": pkg.SCHED_WAVEFRONT, "persistent": pkg.SCHED_PERSISTENT, "regen": pkg.SCHED_REGEN, "per-path": pkg.SCHED_PER_PATH}[args.schedule])
    ctx.set_option(pkg.OPT_TRAVERSAL, mode)
    ctx.set_option(pkg.OPT_BLOCK_THREADS, args.block)
    if args.wave_slots:
        ctx.set_option(pkg.OPT_WAVE_SLOTS, args.wave_slots)
    if args.l2_persist >= 0:
        ctx.set_option(pkg.OPT_L2_PERSIST, args.l2_persist)
    if args.refill:
        ctx.set_option(pkg.OPT_REFILL, args.refill)
    if args.tail_paths >= 0:
        ctx.set_option(pkg.OPT_TAIL_PATHS, args.tail_paths)
    if args.finder >= 0:
        ctx.set_option(pkg.OPT_FINDER, args.finder)
    if args.primary_cull >= 0:
        ctx.set_option(pkg.OPT_PRIMARY_CULL, args.primary_cull)
    if args.finder_build >= 0:
        ctx.set_option(pkg.OPT_DEVICE_FINDER_BUILD, args.finder_build)
    if args.overlap >= 0:
        ctx.set_option(pkg.OPT_OVERLAP, args.overlap)
    if args.fused_primary >= 0:
        ctx.set_option(pkg.OPT_FUSED_PRIMARY, args.fused_primary)
    if args.sort_rays >= 0:
        ctx.set_option(pkg.OPT_SORT_RAYS, args.sort_rays)

    # the step's host inputs live in pinned host memory: the two noise tables (the library DMAs them as they are and packs
    # on the device; pageable tables are packed by host threads first) and the five scene arrays (a scene of more than
    # 2 Mi triangles goes to the device as it is and is re-laid out there, csrc/srt_relayout.cu; smaller ones are walked
    # on the host, where pinning changes nothing)
    pinned = [torch.from_numpy(np.ascontiguousarray(inp[k], np.float32)).pin_memory() for k in ("unit", "uniform")]
    noise_unit, noise_uniform = pinned[0].numpy(), pinned[1].numpy()
    if inp["scene"] is not None:
        for k in ("bvhs", "nodes", "materials", "triangles", "vertices"):
            a = np.ascontiguousarray(inp["scene"][k])
            t = torch.from_numpy(a.view(np.uint8).reshape(-1)).pin_memory()
            pinned.append(t)
            inp["scene"][k] = t.numpy().view(a.dtype).reshape(a.shape)
    if args.device_relayout >= 0:
        ctx.set_option(pkg.OPT_DEVICE_RELAYOUT, args.device_relayout)

    def upload_all():
        ctx.upload_noise(noise_unit, noise_uniform)  # queued DMA: runs while the host prepares the scene
        if inp["scene"] is not None:
            ctx.upload_scene(inp["scene"])
        ctx.set_lights(inp["lights"])

    upload_all()
    ctx.resize(W, H)
    if args.stripe_of > 1 and world == 1:
        ctx.set_stripe(0, args.stripe_of)  # study aid: this GPU renders what rank 0 of `stripe_of` GPUs would (no collective)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def params(reset):
        return pkg.make_params(W, H, inp["cam"], show_model, accum_frames=1, bvh_count=1 if show_model else 0, reset=reset,
                               max_depth=inp["max_depth"])

    launches = [0]

    def step(ev_kernel=None):
        """One fresh 64-spp frame: clear (the shader's reset dispatch), then 64 frames in one srt_render -- which, in a
        group, ends with the call's one collective and rank 0's tonemap, all queued on the stream by the library."""
        ctx.set_params(params(1))
        ctx.render_async(1)
        ctx.set_params(params(0))
        if ev_kernel:
            ev_kernel[0].record(stream)
        ctx.render_async(spp)
        if ev_kernel:
            ev_kernel[1].record(stream)
        launches[0] += 2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 0)):
        step()
    barrier()
    if world > 1:
        ctx.collective_ms()  # forget the warm-up's collectives (the first one sets the NCCL connections up)

    # ---- timed region: K steps, device time by CUDA events on the launching stream, L2 flushed between steps
    ctx.reset_counters()
    launches[0] = 0
    wavefront = show_model and args.schedule == "wavefront"
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    evk = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    wall0 = time.perf_counter()
    for k in range(args.steps):
        flush.fill_(k & 0xFF)
        ev[k][0].record(stream)
        step(evk[k])
        ev[k][1].record(stream)
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop() if sampler else None
    step_ms = [a.elapsed_time(b) for a, b in ev]
    kern_ms = [a.elapsed_time(b) for a, b in evk]
    cnt = ctx.counters()
    # ---- per-launch durations for the roofline: the same steps again with every launch bracketed by CUDA events on the
    # stream it is launched on (SRT_OPT_TIME_KERNELS) and with SRT_OPT_OVERLAP off, so that each launch is timed ALONE -- in
    # the timed region above a chunk's tail kernel runs beside the next step's waves, and the sum of two overlapping
    # launches' durations is not the time either of them needs
    trace_ms = logic_ms = 0.0
    n_waves = 0
    roof_steps = 0
    if wavefront:
        roof_steps = min(args.steps, 3)
        ctx.set_option(pkg.OPT_OVERLAP, 0)
        ctx.set_option(pkg.OPT_TIME_KERNELS, 1)
        ctx.kernel_times()
        barrier()
        for k in range(roof_steps):
            flush.fill_(k & 0xFF)
            step()
        barrier()
        trace_ms, logic_ms, n_waves = ctx.kernel_times()
        ctx.set_option(pkg.OPT_TIME_KERNELS, 0)
        ctx.set_option(pkg.OPT_OVERLAP, 1 if args.overlap < 0 else args.overlap)
    coll_ms, coll_n = ctx.collective_ms() if world > 1 else (0.0, 0)
    gpu_launches = cnt["kernel_launches"]  # every kernel this context launched in the timed region (clear, init, waves, resolve, tonemap)

    tot = torch.tensor([sum(step_ms), sum(kern_ms), trace_ms, logic_ms, float(n_waves)], dtype=torch.float64, device="cuda")
    work = torch.tensor([cnt["segments"], cnt["shadow_rays"], cnt["paths"], gpu_launches], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
    total_ms, kernel_ms, trace_ms, logic_ms, n_waves = [float(v) for v in tot.tolist()]
    segments, shadow, paths, gpu_launches = [int(v) for v in work.tolist()]
    ms_per_step = total_ms / args.steps
    value = segments / (total_ms * 1e-3) / 1e6

    # ---- work per ray for the roofline, from two untimed counting passes over the SAME frames as a timed step:
    #   reference: box / triangle tests of the SHADER's walk (Nn, Nt "as executed by the reference traversal
    #              order"; shadow rays stop at the first accepted triangle here, the shader runs them on: a lower bound)
    #   executed : box / triangle tests the default path really makes (hit finder + replay + overflow walk + primary
    #              cull; the counting build makes waves to the end instead of the tail kernel, which carries the last
    #              ~1 % of the rays through the reference BVH)
    ref_nodes_per_ray = ref_tris_per_ray = exe_nodes_per_ray = exe_tris_per_ray = 0.0
    culled_per_step = 0.0
    if show_model and wavefront:
        def count_pass(kind):
            ctx.set_option(pkg.OPT_COUNT_TRAVERSAL, kind)
            if kind == 1:
                ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_REFERENCE)
            ctx.reset_counters()
            ctx.set_params(params(1)); ctx.render(1)
            ctx.set_params(params(0)); ctx.render(spp)
            c = ctx.counters()
            ctx.set_option(pkg.OPT_COUNT_TRAVERSAL, 0)
            ctx.set_option(pkg.OPT_TRAVERSAL, mode)
            w = torch.tensor([c["nodes"], c["tris"], c["segments"] + c["shadow_rays"], c["culled_primaries"]],
                             dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(w, op=dist.ReduceOp.SUM)
            return w.tolist()
        n1, t1, r1, _ = count_pass(1)
        ref_nodes_per_ray, ref_tris_per_ray = n1 / max(r1, 1), t1 / max(r1, 1)
        n2, t2, r2, culled_per_step = count_pass(2)
        exe_nodes_per_ray, exe_tris_per_ray = n2 / max(r2, 1), t2 / max(r2, 1)

    # ---- end to end through the C-ABI with host buffers: upload everything, render, (reduce,) read the image back
    out_host = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory().numpy()
    e2e_steps = args.steps
    barrier()
    ctx.reset_counters()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ta = time.perf_counter()
        upload_all()
        tb = time.perf_counter()
        step()
        tc = time.perf_counter()
        if rank == 0:
            ctx.read_output(out_host)
        else:
            ctx.sync()
        if os.environ.get("SRT_BENCH_DEBUG"):
            print("e2e iteration: upload %.1f ms, queue %.1f ms, finish+read %.1f ms" % (
                (tb - ta) * 1e3, (tc - tb) * 1e3, (time.perf_counter() - tc) * 1e3), file=sys.stderr)
    barrier()
    e2e_s = time.perf_counter() - t0
    ce = ctx.counters()
    e2 = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    es = torch.tensor([ce["segments"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2, op=dist.ReduceOp.MAX)
        dist.all_reduce(es, op=dist.ReduceOp.SUM)
    e2e_value = float(es.item()) / float(e2.item()) / 1e6
    h2d = 24 * W * H + 320 + 64  # the two RGB32F noise tables as the host holds them + light list + uniform block
    if inp["scene"] is not None:
        n_inner = int((inp["scene"]["nodes"]["count"] == 0).sum())
        n_leaf = int((inp["scene"]["nodes"]["count"] > 0).sum())
        relayout_on_device, relayout_ms = ctx.upload_info()
        if relayout_on_device:  # the reference's arrays as they are (re-laid out by kernels on the device)
            h2d += sum(int(inp["scene"][k].nbytes) for k in ("bvhs", "nodes", "triangles", "vertices")) + 32 * len(inp["scene"]["materials"])
        else:                   # the records the host walk made
            h2d += 64 * n_inner + 8 * n_leaf + 48 * len(inp["scene"]["triangles"]) + 32 * len(inp["scene"]["materials"]) + 96
    else:
        relayout_on_device, relayout_ms = False, 0.0
    d2h = W * H * 4

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the kernels that answer rays: which roof binds depends on where the scene lives
    peaks, which = measured_peaks()
    rays = segments + shadow
    roof_rays = rays * (roof_steps / args.steps) if wavefront else rays  # the rays of the steps whose launches were bracketed
    finder_on = bool(show_model and wavefront and ctx.finder_info()[0])
    if wavefront:
        # SRT_OPT_TIME_KERNELS brackets every launch that answers rays: wf_init_kernel's cull, wf_find_kernel,
        # wf_replay_kernel, the overflow wf_trace_kernel, wf_tail_kernel (its shading included); launches of one step
        # differ in size (one per wave), so per-launch figures are averages over the timed region's launches
        kname = "wf_find_kernel+wf_replay_kernel (+wf_trace_kernel overflow, wf_init_kernel cull, wf_tail_kernel)" if finder_on else "wf_trace_kernel (+wf_init_kernel cull, wf_tail_kernel)"
        k_ms, k_launches = trace_ms, max(n_waves, 1.0)
    else:
        kname, k_ms, k_launches = "pathtrace_persistent_kernel" if args.schedule == "persistent" else "pathtrace_kernel", kernel_ms, float(args.steps * world)
    k_s = max(k_ms, 1e-9) * 1e-3
    scene_mb = finder_mb = 0.0
    if inp["scene"] is not None:
        n_inner = int((inp["scene"]["nodes"]["count"] == 0).sum())
        scene_mb = (64 * n_inner + 48 * len(inp["scene"]["triangles"])) / 1e6  # NodePair + TriRecord arrays as the kernels read them
    # roofs measured now, on this GPU, through the C-ABI (srt_measure_gather / srt_measure_issue)
    roofs = {"issue_fmul_fadd_Gop_s": ctx.measure_issue(0), "issue_fmul_fadd_2fmnmx_Gop_s": ctx.measure_issue(1),
             "l1_gather_GBps": ctx.measure_gather(48 << 10, 1), "l2_gather_GBps": ctx.measure_gather(32 << 20, 1),
             "hbm_gather_GBps": ctx.measure_gather(1 << 30, 1), "hbm_copy_GBps": peaks["hbm_gbs"]}
    # algorithmic work of the tests the kernels EXECUTE (SURVEY.md 8d per-test figures: 28 fp32 ops per box test, 50 per
    # triangle test; 32 B per node, 52 B per triangle), and of the shader's own walk for comparison across designs
    exe_flops_per_ray = 28.0 * exe_nodes_per_ray + 50.0 * exe_tris_per_ray
    exe_bytes_per_ray = 32.0 * exe_nodes_per_ray + 52.0 * exe_tris_per_ray
    if not show_model:
        # the sphere scene (configs[0]) has no BVH: CheckHit is SphereHit on the 5 spheres of main() (raytrace_compute.glsl:
        # 93-141), about 30 fp32 operations each (two lengths, two dots, the discriminant, one sqrt, a divide, the interval
        # test) on constants held in registers -- no scene bytes at all, so the bound can only be instruction issue; the
        # one-thread-per-pixel kernel's whole duration (traversal + shading) is the denominator
        exe_flops_per_ray, exe_bytes_per_ray = 5.0 * 30.0, 0.0
    ref_bytes_per_ray = 32.0 * ref_nodes_per_ray + 52.0 * ref_tris_per_ray
    achieved_tflops = roof_rays * exe_flops_per_ray / k_s / 1e12
    achieved_gbs = roof_rays * exe_bytes_per_ray / k_s / 1e9
    issue_peak_tflops = roofs["issue_fmul_fadd_Gop_s"] * world / 1e3
    level = "l1" if scene_mb <= 0.2 else ("l2" if scene_mb <= 100.0 else "hbm")
    mem_peak = {"l1": roofs["l1_gather_GBps"], "l2": roofs["l2_gather_GBps"], "hbm": peaks["hbm_gbs"]}[level] * world
    frac_issue = achieved_tflops / issue_peak_tflops
    frac_mem = achieved_gbs / mem_peak
    # DRAM bytes per launch from the committed ncu pass of this command (never measured in this run: numbers taken under
    # a profiler are not bench values); one file per workload and build
    traffic, traffic_extra = None, {}
    tf = os.path.join(ROOT, "profiles", "r02_traffic_%s.json" % args.workload)
    if wavefront and world == 1 and os.path.exists(tf):
        t = json.load(open(tf))
        traffic = t["dram_bytes_per_launch"]
        algorithmic_dram = (12.0 * W * H + 36.0 * W * H) * spp / max(k_launches / max(roof_steps, 1), 1.0)  # noise gathers + accumulate/output streaming
        traffic_extra = {"traffic_source": "profiles/%s (dram__bytes_read.sum + dram__bytes_write.sum over the launches of one "
                                           "step from an ncu pass of this command, divided by the ray-answering launches)" % os.path.basename(tf),
                         "traffic_per_step": t.get("dram_bytes_per_step"), "traffic_build": t.get("build"),
                         "wasted_ratio_vs_algorithmic_dram": (t.get("dram_bytes_per_step") or 0.0) / ((12.0 + 36.0) * W * H * spp) if show_model else None}
    if frac_mem > frac_issue:
        roofline = {"bound": level if level != "l1" else "l1", "achieved": achieved_gbs, "peak": mem_peak, "unit": "GB/s", "frac": frac_mem}
    else:
        roofline = {"bound": "issue", "achieved": achieved_tflops, "peak": issue_peak_tflops, "unit": "TFLOP/s", "frac": frac_issue}
    roofline.update({
        "traffic": traffic,
        "peak_source": ("measured in this run through the C-ABI: srt_measure_issue(FMUL+FADD, un-fused: the library is built "
                        "-fmad=false) x n_gpus" if roofline["bound"] == "issue" else
                        ("%s (MEASURED_PEAKS.json hbm_gbs, burst copy figure) x n_gpus" % which if level == "hbm" else
                         "measured in this run through the C-ABI: srt_measure_gather(%s-resident array, 64 B records) x n_gpus" % level.upper())),
        "kernel": kname, "kernel_launches": k_launches, "kernel_ms_total": k_ms, "ms_per_launch": k_ms / k_launches,
        "logic_kernel_ms_total": logic_ms,
        "frac_issue": frac_issue, "frac_memory": frac_mem, "memory_level": level,
        "achieved_tflops": achieved_tflops, "achieved_GBps": achieved_gbs,
        "flops_per_launch": roof_rays * exe_flops_per_ray / k_launches, "bytes_per_launch": roof_rays * exe_bytes_per_ray / k_launches,
        "timed_alone_steps": roof_steps,
        "executed_per_ray": {"box_tests": exe_nodes_per_ray, "triangle_tests": exe_tris_per_ray, "sphere_tests": 0.0 if show_model else 5.0,
                             "flops": exe_flops_per_ray, "bytes": exe_bytes_per_ray},
        "reference_walk_per_ray": {"box_tests": ref_nodes_per_ray, "triangle_tests": ref_tris_per_ray, "bytes": ref_bytes_per_ray,
                                   "answered_GBps": roof_rays * ref_bytes_per_ray / k_s / 1e9},
        "rays_per_step": rays / args.steps, "scene_mb": scene_mb, "roofs": roofs,
        "note": "achieved = rays x algorithmic work of the box and triangle tests the kernels EXECUTE (28 fp32 ops / 32 B per "
                "box test, 50 ops / 52 B per triangle test, counted by an untimed pass over the same frames with "
                "SRT_OPT_COUNT_TRAVERSAL = 2) / sum of CUDA-event durations around every ray-answering launch on the "
                "stream it is launched on (max over ranks), taken over `timed_alone_steps` further steps with SRT_OPT_OVERLAP off so "
                "that no two bracketed launches run at the same time.  Both fractions are reported; the larger names the bound.  frac_issue "
                "counts only the tests' own arithmetic against the issue rate of un-fused FP32 instructions: address, "
                "stack, queue and control instructions use the other slots (ncu: profiles/).  reference_walk_per_ray is "
                "what the shader's own traversal does for the same rays (answered_GBps: reference work answered per "
                "second, comparable across designs, not bytes moved)."})
    roofline.update(traffic_extra)

    cpu_baseline = None
    if world == 1 and not args.no_cpu:
        n_cpu = args.cpu_frames
        if n_cpu <= 0:  # about 15 s of CPU work, sized from one calibration frame
            _, t1, _ = oracle_run(inp, 1, 1)
            n_cpu = max(1, min(inp["spp"], int(15.0 / max(t1, 1e-3))))
        v, dt, c = oracle_run(inp, 1, n_cpu)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": cpu_kind()[0],
                        "sample": "%d of the workload's %d frames (1 spp each), whole image, %.1f s" % (n_cpu, inp["spp"], dt),
                        "ms_per_frame": dt / n_cpu * 1e3}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": inp["desc"], "traversal_requested": args.traversal,
                   "traversal": "reference (the shader's visiting order: the hit finder answers the rays, and every kernel beside it walks "
                                "in that order)" if finder_on else args.traversal,
                   "schedule": args.schedule, "block_threads": args.block,
                   "split": ("single GPU" if world == 1 else
                             ("sample ranges: each of %d GPUs renders the whole image for 1/%d of the frames; one ncclReduce(sum) of the "
                              "fp32 accumulation images to rank 0 per step, inside srt_render" % (world, world)) if args.split == "samples" else
                             ("interleaved 8-row bands over %d GPUs; each rank's own rows (P*16/%d bytes) gathered on rank 0 with "
                              "ncclSend/ncclRecv once per step, inside srt_render" % (world, world))),
                   "l2": "512 MB device buffer rewritten between timed steps (L2 flush)",
                   "tail_paths": "library default" if args.tail_paths < 0 else args.tail_paths,
                   "primary_cull": "library default (on)" if args.primary_cull < 0 else args.primary_cull,
                   "sort_rays": "library default (off)" if args.sort_rays < 0 else args.sort_rays,
                   "hit_finder": dict(zip(("active", "cost_reference", "cost_finder"), ctx.finder_info())) if show_model else None,
                   "ms_per_frame": ms_per_step / spp},
        "segments_per_step": segments / args.steps, "shadow_rays_per_step": shadow / args.steps,
        "mrays_incl_shadow": rays / (total_ms * 1e-3) / 1e6,
        # `value` counts every CheckHit of the path loop, as the shader and the CPU arm do; most camera rays of this view
        # miss the model's root box (ended by the primary cull).  The rates of the rays that do reach a BVH:
        "paths_per_step": paths / args.steps, "culled_primaries_per_step": culled_per_step,
        "segments_reaching_bvh_per_s_M": (segments / args.steps - culled_per_step) / (ms_per_step * 1e-3) / 1e6,
        "bounce_segments_per_s_M": (segments - paths) / args.steps / (ms_per_step * 1e-3) / 1e6,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": float(e2.item()) / e2e_steps * 1e3,
                "scene_relayout": ("device kernels, %.1f ms" % relayout_ms) if relayout_on_device else "host walk"},
        "emulated_rank_of": args.stripe_of if (args.stripe_of > 1 and world == 1) else None,
        "collective_ms_per_step": (coll_ms / max(coll_n, 1)) if world > 1 else None,
        "gpu_launches": gpu_launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
        "wall_s_timed_region": wall,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="rubik", choices=sorted(WORKLOADS))
    ap.add_argument("--traversal", default="ordered", choices=["ordered", "reference"])
    ap.add_argument("--block", type=int, default=64)
    ap.add_argument("--schedule", default="wavefront", choices=["wavefront", "persistent", "regen", "per-path"])
    ap.add_argument("--cpu-frames", type=int, default=0,
                    help="frames of the workload the CPU baseline renders (0 = as many as fit about 15 s)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--obj", default=None,
                    help="<objects dir>/<NAME>/<NAME>.obj: render this model instead of the workload's scene (size, spp and depth stay "
                         "the workload's), e.g. the Airplane OBJ of BASELINE.json configs[2] where someone supplies it")
    ap.add_argument("--soup-tris", type=int, default=10_000_000, help="triangles of the soup workload")
    ap.add_argument("--fused-primary", type=int, default=-1, help="0/1: SRT_OPT_FUSED_PRIMARY (default: library default, 0 = init kernel; 1 = the first find kernel makes the camera rays)")
    ap.add_argument("--overlap", type=int, default=-1, help="0/1: SRT_OPT_OVERLAP (default: library default, 1)")
    ap.add_argument("--finder-build", type=int, default=-1, help="0/1: SRT_OPT_DEVICE_FINDER_BUILD (default: library default, 1 = on the device)")
    ap.add_argument("--device-relayout", type=int, default=-1, help="0/1/2: SRT_OPT_DEVICE_RELAYOUT (default: library default, 2)")
    ap.add_argument("--l2-persist", type=int, default=-1, help="1/0: pin the BVH top levels in L2 (default: library default, on)")
    ap.add_argument("--refill", type=int, default=0, help="refill threshold of the wavefront trace kernel (0 = default)")
    ap.add_argument("--primary-cull", type=int, default=-1,
                    help="1/0: primary rays that miss every root box end in the init kernel (-1 = library default, on)")
    ap.add_argument("--sort-rays", type=int, default=-1,
                    help="ray queues regrouped by origin cell before each wave: 0 off, 1 on (-1 = library default, off: measured "
                         "slower on the soup)")
    ap.add_argument("--finder", type=int, default=-1,
                    help="hit finder of the wavefront schedule: 0 off, 1 on, 2 auto (-1 = library default, auto)")
    ap.add_argument("--tail-paths", type=int, default=-1,
                    help="live paths at which the wavefront schedule's tail kernel takes over (-1 = library default, 0 = never)")
    ap.add_argument("--split", default="samples", choices=["bands", "samples"],
                    help="multi-GPU partition (SRT_OPT_SPLIT): sample ranges (default here: every rank renders the whole image for "
                         "its share of the frames; equal to one GPU within the metric's 1e-4, and 10 %% faster at 8 GPUs because "
                         "the ranks' loads are equal: profiles/r02l_*) or interleaved image bands (the library's default: the bits "
                         "of one GPU)")
    ap.add_argument("--stripe-of", type=int, default=1,
                    help="study aid (1 GPU only): render only the bands rank 0 of this many GPUs would own -- the per-rank step of a "
                         "multi-GPU run without its collective; the line is marked emulated_rank_of")
    ap.add_argument("--wave-slots", type=int, default=0, help="path slots of the wavefront schedule (0 = library default)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        if args.warmup < 3:
            args.warmup = 3
        run_gpu(args)


if __name__ == "__main__":
    main()
