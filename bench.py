#!/usr/bin/env python
"""bench.py — measures BASELINE.json's metric (Mrays/s on the 1M-triangle scene at
1080p) for the B200 path, with the roofline of the trace kernel and the CPU
baseline beside it.  One JSON line on stdout (rank 0).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # reference CPU arm (oracle port)

Step = one frame of the workload through the hot path.  `value`: scene resident
in HBM, CUDA-event time of the K frames (L2 flushed between frames), max over
ranks.  `e2e`: the same frames through the C ABI with HOST buffers — scene
upload from pinned host memory + render + composite read-back, wall clock.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[2] — the configuration the metric is quoted on ("1080p, 1M-tri scene")
    "c3": dict(nx=1000, nz=500, w=1920, h=1080, bvh="auto", mode="path", spp=16, depth=4,
               desc="C3: procedural 1M-triangle terrain, 1920x1080, 4 bounces, 16 spp progressive accumulation "
                    "(EXTENSION semantics, SPEC.md: the reference stubs bounces/spp), scene uploaded with the "
                    "reference's chunked BVH (100 tris/leaf)"),
    "c3-primary": dict(nx=1000, nz=500, w=1920, h=1080, bvh="auto", mode="primary", spp=1, depth=0,
                       desc="C3 scene: procedural 1M-triangle terrain, 1920x1080, primary rays + direct shading "
                            "(reference semantics), reference chunked BVH (100 tris/leaf)"),
    "c2": dict(nx=250, nz=200, w=1920, h=1080, bvh="auto", mode="primary", spp=1, depth=0,
               desc="C2: procedural 100K-triangle terrain, 1920x1080, primary rays + direct shading, "
                    "standard-format BVH (1 tri/leaf)"),
    "c3-primary-smallleaf": dict(nx=1000, nz=500, w=1920, h=1080, bvh="small_leaf", mode="primary", spp=1, depth=0,
                                 desc="C3 scene, primary rays + direct shading, small-leaf BVH (<=4 tris/leaf) in "
                                      "the reference node format"),
    "c3-smallleaf": dict(nx=1000, nz=500, w=1920, h=1080, bvh="small_leaf", mode="path", spp=16, depth=4,
                         desc="C3 with the scene uploaded as a small-leaf BVH (<=4 tris/leaf, binned SAH, reference node "
                              "format; SURVEY 8f #1): 1M-triangle terrain, 1920x1080, 4 bounces, 16 spp (EXTENSION semantics)"),
    "field": dict(scene="field", w=1920, h=1080, bvh="auto", mode="path", spp=16, depth=4,
                  desc="Field (not a height field; tree-quality check off the benchmark terrain): 1996 tori at mixed scales and "
                       "orientations scattered through a volume over a ground grid = 1 000 000 triangles, 1920x1080, 4 bounces, "
                       "16 spp (EXTENSION semantics), uploaded with the reference's chunked BVH (100 tris/leaf)"),
    "field-smallleaf": dict(scene="field", w=1920, h=1080, bvh="small_leaf", mode="path", spp=16, depth=4,
                            desc="Field scene uploaded as a small-leaf BVH (<= 4 tris/leaf, host binned SAH, reference node "
                                 "format), 1920x1080, 4 bounces, 16 spp (EXTENSION semantics)"),
    "c4": dict(scene="cornell", w=1024, h=1024, bvh="auto", mode="path", spp=64, depth=8,
               desc="C4: Cornell box (36 triangles), 1024x1024, diffuse bounces to depth 8, 64 spp (EXTENSION "
                    "semantics, SPEC.md) - divergence stress"),
    "c5": dict(nx=2500, nz=2000, w=3840, h=2160, bvh="auto", mode="path", spp=1, depth=4,
               desc="C5: procedural 10M-triangle terrain, 3840x2160, 4 bounces, 1 spp (EXTENSION semantics), "
                    "reference chunked BVH (1000 tris/leaf)"),
}


# The bounded CPU sample of each workload (cpu_baseline leg AND the --impl reference arm use the same one, so the two
# figures are the same measurement): a centred band of rows of the workload's own frame, full spp and depth.
# Sized for ~2-6 s on 16 host threads; the oracle hands 64-pixel spans to its pool, so even 2 rows keep every
# thread busy.  Primary workloads: whole 128-row tile bands, three channel passes (the reference's own loop).
CPU_SAMPLE_ROWS = {"field": 4, "field-smallleaf": 64, "c3": 4, "c3-smallleaf": 64, "c4": 48, "c5": 1, "c3-primary": 128, "c3-primary-smallleaf": 1080,
                   "c2": 1080}


def source_hash():
    """sha256 over the native sources: ties an ncu summary under profiles/ to the build it was captured from."""
    hsh = hashlib.sha256()
    for d, exts in (("gpu_raytracer_b200/csrc", (".cu", ".cuh", ".cpp", "Makefile")), ("include", (".h",))):
        for name in sorted(os.listdir(os.path.join(ROOT, d))):
            if name.endswith(exts):
                with open(os.path.join(ROOT, d, name), "rb") as f:
                    hsh.update(name.encode() + b"\0" + f.read())
    return hsh.hexdigest()[:16]


def ncu_summary_for(workload):
    """profiles/ncu_summary.json: {workload: {"src_hash": ..., "file": ..., per-kernel counters}} written by
    scripts/ncu_extract.py from an `ncu --set full` capture.  Used only when it was captured from THIS build."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_summary.json")) as f:
            rec = json.load(f).get(workload)
    except Exception:
        return None, "no profiles/ncu_summary.json"
    if not rec:
        return None, "no ncu capture of this workload in profiles/ncu_summary.json"
    if rec.get("src_hash") != source_hash():
        return None, "profiles/ncu_summary.json was captured from another build (src_hash %s != %s)" % (rec.get("src_hash"), source_hash())
    return rec, rec.get("file")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-seconds", type=float, default=10.0,
                    help="GPU arm: repeat the fixed cpu_baseline sample (CPU_SAMPLE_ROWS) until this much CPU time has passed")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dump-rows", default=None, help="reference arm: also write the rendered rows of the last step to this .npz "
                                                      "(the GPU arm's cpu_baseline leg runs the reference arm as a child process "
                                                      "and compares its pixels with the GPU frame)")
    ap.add_argument("--ref-seconds", type=float, default=0.0,
                    help="reference arm: keep adding timed steps past --steps until this much time has been measured")
    ap.add_argument("--variant", type=int, default=0, help="path mode engine: 0 auto = streaming wavefront (default), 1 megakernel, "
                                                           "4 streaming wavefront")
    ap.add_argument("--traversal", default="auto", choices=["auto", "literal", "ordered", "rebuilt"])
    ap.add_argument("--balance", type=int, default=4,
                    help="N > 1: load-balancing steps during warm-up (explicit per-rank tile lists adjusted on the previous "
                         "frames' per-rank GPU times, tiles.rebalance); 0 keeps tile i -> rank i %% N")
    ap.add_argument("--pipeline", type=int, default=1,
                    help="1: also measure e2e pipelined over three contexts per GPU (upload + build two steps ahead beside the running "
                         "frames, asynchronous read-back) and report it as `e2e` (the serial figure stays in `e2e_serial`); 0: serial only")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N > 1: how finished tiles reach rank 0: peer = P2P stores from the render kernels into rank 0's "
                         "frame (CUDA IPC) + a 4-byte NCCL all-reduce as barrier; nccl = pack + dist.gather + unpack")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def build_scene(wl):
    import gpu_raytracer_b200 as g
    from gpu_raytracer_b200 import abi
    strat = {"auto": abi.BVH_AUTO, "small_leaf": abi.BVH_SMALL_LEAF, "standard": abi.BVH_STANDARD}[wl["bvh"]]
    if wl.get("scene") == "cornell":
        return g.cornell_scene(strat)
    if wl.get("scene") == "field":
        return g.field_scene(strategy=strat)
    return g.terrain_scene(wl["nx"], wl["nz"], seed=1, strategy=strat)


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cpu_sample_rows(args, wl):
    n = min(CPU_SAMPLE_ROWS.get(args.workload, 8), wl["h"])
    if wl["mode"] != "path":  # whole tile bands (src/compute.rs:194-209)
        bands = max(1, n // 128)
        total = (wl["h"] + 127) // 128
        first = max(0, (total - bands) // 2)
        return first * 128, min(wl["h"], (first + bands) * 128)
    y0 = max(0, (wl["h"] - n) // 2)
    return y0, y0 + n


def cpu_baseline_sample(scene, wl, rows, bound=None, min_seconds=0.0):
    """Times the oracle on a fixed band of rows of the same frame on all host threads, repeating the band until
    `min_seconds` have passed.  Primary workloads: the reference's own loop (literal traversal, 3 channel passes
    per tile).  Path workloads: the extension oracle (SPEC.md) with the literal traversal.  Returns the record
    and the rendered rows (kept: the GPU frame is compared with them)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob
    bound = bound or ob.BoundScene(scene)
    w, h = wl["w"], wl["h"]
    threads = ob.hardware_threads()
    y0, y1 = rows
    reps, dt, r = 0, 0.0, None
    while reps == 0 or dt < min_seconds:
        t0 = time.perf_counter()
        if wl["mode"] == "path":
            r = ob.render_path(bound, w, h, spp=wl["spp"], max_depth=wl["depth"], base_seed=1, rows=(y0, y1), threads=threads)
        else:
            r = ob.render(bound, w, h, rows=(y0, y1), threads=threads)
        dt += time.perf_counter() - t0
        reps += 1
    if wl["mode"] == "path":
        rays = r["counters"]["rays"]
        what = "%d spp, depth %d, literal traversal, %d ray segments" % (wl["spp"], wl["depth"], rays)
        pixels = dict(accum=r["accum"][y0:y1].copy(), resolved=r["resolved"][y0:y1].copy())
    else:
        # the reference traces every pixel three times (one dispatch per colour channel);
        # the metric counts one camera ray per pixel, like the GPU arm
        rays = (y1 - y0) * w
        what = "3 channel passes per 128x128 tile, literal traversal, %d pixels" % rays
        pixels = dict(radiance=r["radiance"][y0:y1].copy(), composite=r["composite"][y0:y1].copy())
    rec = {"value": rays * reps / dt / 1e6, "unit": "Mrays/s", "cores": threads, "threads_used": threads, "kind": "port",
           "sample": "rows %d-%d of the %dx%d frame, %s, %.2f s per pass x %d" % (y0, y1, w, h, what, dt / reps, reps),
           "seconds": dt / reps}
    return rec, pixels


def workload_config(wl, scene):
    """The `config` both arms print: the workload only, nothing arm-specific."""
    return {"workload": wl["desc"], "width": wl["w"], "height": wl["h"], "triangles": int(len(scene.triangles)),
            "bvh_nodes": int(len(scene.bvh_nodes)), "spp": wl["spp"], "max_depth": wl["depth"],
            "semantics": "extension (SPEC.md; the reference stubs bounces/spp)" if wl["mode"] == "path" else "reference",
            "l2": "GPU arm: flushed between timed steps (512 MiB memset)"}


def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and args.impl != "reference":  # the reference arm is rank-0-only CPU work
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = "gloo" if args.impl == "reference" else "nccl"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return world, rank, local


def max_over_ranks(x, world):
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda" if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def run_reference(args, wl, world, rank):
    """Reference arm: the reference's CPU implementation of the path.  The Rust crate cannot be built here (no
    rustc; rust-gpu nightly + crates.io), so this is the oracle port (literal traversal; primary workloads: three
    channel passes per tile) on all host threads.  Each step is the SAME bounded sample the GPU arm's cpu_baseline
    leg times (cpu_sample_rows).  Rank 0 only."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob
    scene = build_scene(wl)
    bound = ob.BoundScene(scene)
    rows = cpu_sample_rows(args, wl)
    vals, last, px, i = [], None, None, 0
    while i < args.warmup + args.steps or sum(x["seconds"] for x in vals) < args.ref_seconds:
        last, px = cpu_baseline_sample(scene, wl, rows, bound)
        if i >= args.warmup:
            vals.append(last)
        i += 1
    if args.dump_rows:
        np.savez(args.dump_rows, rows=np.array(rows), **px)
    secs = float(np.sum([x["seconds"] for x in vals]))
    v = float(np.sum([x["value"] * x["seconds"] for x in vals]) / secs)  # total rays / total time
    line = {"impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": len(vals), "warmup": args.warmup, "ms_per_step": secs / len(vals) * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(wl, scene),
            "traversal": "literal (the reference's own unordered, un-pruned DFS over the uploaded tree)",
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": last["cores"], "threads_used": last["threads_used"],
                             "kind": "port", "sample": last["sample"] + " (%d timed passes)" % len(vals)},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def sum_over_ranks(x, world):
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def run_b200(args, wl, world, rank, local):
    import torch
    import torch.distributed as dist
    import gpu_raytracer_b200 as g
    from gpu_raytracer_b200 import abi, tiles

    torch.cuda.set_device(local)
    scene = build_scene(wl)
    w, h = wl["w"], wl["h"]
    path = wl["mode"] == "path"
    ctx = g.Context(local, w, h)
    bm, cr = g.BufferManager(ctx), g.ComputeRenderer(ctx)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local)  # the context's stream, for events + NCCL ordering
    # host inputs in pinned memory (the e2e leg copies from these every step)
    names = ("spheres", "lights", "bvh_nodes", "triangle_indices", "vertices", "triangles", "materials")
    pinned, keep = {}, []
    for name in names:
        a = np.ascontiguousarray(getattr(scene, name))
        if a.size:
            t = torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).pin_memory()
            keep.append(t)
            pinned[name] = t.numpy().view(a.dtype)
        else:
            pinned[name] = a
    ps = g.Scene(scene.camera, pinned["spheres"], pinned["triangles"], pinned["vertices"], pinned["materials"],
                 pinned["lights"], pinned["bvh_nodes"], pinned["triangle_indices"])
    h2d_bytes = sum(int(pinned[n].nbytes) for n in names)
    bm.upload_scene(ps)
    stride, phase = (world, rank) if world > 1 else (0, 0)
    outputs = abi.OUT_ACCUM | abi.OUT_COMPOSITE
    my_tiles = len(tiles.tiles_of_rank(w, h, world, rank))
    my_pixels = tiles.pixels_of_rank(w, h, world, rank)
    trav = {"auto": abi.TRAVERSAL_AUTO, "literal": abi.TRAVERSAL_LITERAL, "ordered": abi.TRAVERSAL_ORDERED,
            "rebuilt": abi.TRAVERSAL_REBUILT}[args.traversal]
    gather = peer = None
    if world > 1 and args.gather == "peer":
        try:
            peer = tiles.PeerFrame(dist, torch, ctx, stream)
            outputs |= peer.extra_outputs
        except g.RtError as e:  # no P2P between these GPUs: NCCL path
            print("peer frame unavailable (%s): falling back to the NCCL gather" % e, file=sys.stderr)
            peer = None
    lists = None
    if world > 1 and args.balance > 0:  # explicit tile lists, balanced below during warm-up
        lists = tiles.initial_lists(w, h, world)
        ctx.set_tile_list(lists[rank])
        stride, phase = 0, 0

    def make_gather():
        return tiles.TileGather(dist, torch, w, h, tiles.WHAT_COMPOSITE, "cuda", lists=lists)
    if world > 1 and peer is None:
        gather = make_gather()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def render():
        if path:
            cr.render_frame(ps.camera, mode=abi.MODE_PATH, spp=wl["spp"], max_depth=wl["depth"], base_seed=1,
                            outputs=outputs, tile_stride=stride, tile_phase=phase, traversal=trav, variant=args.variant)
        else:
            cr.render_frame(ps.camera, outputs=outputs, tile_stride=stride, tile_phase=phase, traversal=trav)

    def frame():
        """One step: trace this rank's tiles, then (N > 1) gather the finished tiles to rank 0."""
        render()
        if peer is not None:
            peer.finish()  # the tiles are already in rank 0's frame (P2P stores in the kernels); stream-ordered barrier
        elif gather is not None:
            with torch.cuda.stream(stream):  # NCCL is ordered after the kernel on the context's stream
                if lists is not None:
                    gather.run(lambda dst: ctx.pack_tile_list(tiles.WHAT_COMPOSITE, lists[rank], dst.data_ptr()),
                               lambda src, r: ctx.unpack_tile_list(tiles.WHAT_COMPOSITE, lists[r], src.data_ptr()))
                else:
                    gather.run(lambda dst: ctx.pack_tiles(tiles.WHAT_COMPOSITE, world, rank, dst.data_ptr()),
                               lambda src, r: ctx.unpack_tiles(tiles.WHAT_COMPOSITE, world, r, src.data_ptr()))

    clocks = ClockSampler(local)
    clocks.start()  # sampled under the same load from the warm-up on (nvidia-smi takes a few 100 ms to start)
    balance_log = []
    if lists is not None:  # load balancing on the previous frames' per-rank GPU times (untimed warm-up frames)
        for _ in range(args.balance):
            for _ in range(2):
                frame()
            ctx.synchronize()
            t = torch.tensor([ctx.last_kernel_ms], dtype=torch.float64, device="cuda")
            allt = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allt, t)
            times = [float(x.item()) for x in allt]
            balance_log.append({"rank_frame_ms": [round(x, 4) for x in times], "tiles": [len(l) for l in lists]})
            lists = tiles.rebalance(lists, times)
            ctx.set_tile_list(lists[rank])
            if gather is not None:
                gather = make_gather()
        my_tiles = len(lists[rank])
        my_pixels = tiles.pixels_of_tiles(w, h, lists[rank])
    for _ in range(max(args.warmup, 3)):
        frame()
    ctx.synchronize()
    torch.cuda.synchronize()

    # instrumented launch (untimed): algorithmic work per ray segment for the roofline
    ctx.count_next_launch()
    render()
    cnt = ctx.read_counters()
    seg_per_frame_rank = ctx.last_segments if path else my_pixels
    seg_per_frame = sum_over_ranks(float(seg_per_frame_rank), world)

    # ---- timed: K steps, CUDA events on the launching stream, L2 flushed between steps
    barrier(world)
    torch.cuda.synchronize()
    clocks.rows.clear()  # keep the samples taken from here on (timed region + the same frames repeated below)
    launches0 = ctx.launch_count
    step_ms, kernel_ms, stage_ms = [], [], []
    for _ in range(args.steps):
        with torch.cuda.stream(stream):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
        frame()
        e1.record(stream)
        e1.synchronize()
        torch.cuda.synchronize()
        step_ms.append(e0.elapsed_time(e1))
        kernel_ms.append(ctx.last_kernel_ms)  # this rank's own frame (all its launches), without the gather
    launches = ctx.launch_count - launches0
    t_end = time.perf_counter() + 2.0
    while len(clocks.rows) < 3 and time.perf_counter() < t_end:  # short timed region: keep the same load up until sampled
        render()  # rank-local (no collective: ranks may loop a different number of times)
        ctx.synchronize()
    clk = clocks.stop()
    # second timed pass for the per-stage breakdown: the headline pass overlaps the kernels of two
    # sub-frames on two streams, so its launches cannot be timed one by one; here the same frames run as
    # one sub-frame with a CUDA event between consecutive launches
    if path and args.variant in (0, 4):
        os.environ["RTB200_WAVE_SPLIT"] = "1"
        for _ in range(max(3, min(args.steps, 10))):
            with torch.cuda.stream(stream):
                flush.zero_()
            render()
            try:
                stage_ms.append(ctx.last_stage_ms())
            except g.RtError:
                break
        del os.environ["RTB200_WAVE_SPLIT"]
    barrier(world)
    total_ms = max_over_ranks(float(np.sum(step_ms)), world)
    rank_kernel_ms = [float(np.mean(kernel_ms))]
    if world > 1:  # load balance of the tile split: every rank's own mean frame time
        t = torch.tensor(rank_kernel_ms, dtype=torch.float64, device="cuda")
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        rank_kernel_ms = [float(x.item()) for x in allt]
    value = seg_per_frame * args.steps / (total_ms * 1e-3) / 1e6

    # ---- e2e: host buffers -> C ABI -> host image on rank 0, wall clock, every step
    host_img = torch.empty((h, w, 4), dtype=torch.uint8).pin_memory().numpy()
    e2e_ms = []
    barrier(world)
    for i in range(args.warmup + args.steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        bm.mark_all_dirty()
        bm.upload_scene(ps)           # H2D of every scene buffer + device re-layout / LBVH build
        frame()
        if rank == 0:
            ctx.read_composite(host_img)  # D2H of the gathered frame
        else:
            ctx.synchronize()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        if i >= args.warmup:
            e2e_ms.append(dt)
    e2e_total = max_over_ranks(float(np.sum(e2e_ms)), world)
    e2e_value = seg_per_frame * args.steps / (e2e_total * 1e-3) / 1e6

    # ---- e2e, pipelined: the same work per step (upload of every scene buffer from pinned host memory, device
    # re-layout + LBVH build, frame, gather, D2H of the image), over THREE contexts ("lanes") per GPU.  What a host with a
    # scene that changes every frame does with this API (contexts are independent; one thread per context):
    #   * an uploader thread prepares step i+2's scene while step i renders.  Beside a running frame an upload takes one
    #     frame time (its 25 small build launches only get SMs at the frame's kernel boundaries), so the lead has to be
    #     two steps for the host to be able to enqueue frame i+1 before frame i ends (profiles/r2zd_e2e_pipeline_timeline_c3.log);
    #   * frames follow each other on the GPU without a host round trip (a GPU-side event between consecutive frames);
    #   * rank 0 reads the assembled image with an asynchronous copy of Context.composite_cuda_array() (zero-copy view
    #     of rt_device_composite) on the lane's stream, and waits for it one step later.
    # Order that keeps the peer-store epilogue safe: a rank waits (host) for step i-1 to be complete - rank 0: its image
    # has arrived; others: its barrier has completed - after it has enqueued frame i and BEFORE it enqueues barrier i.
    # So a peer launches frame i+3, which stores into the image frame i was read from, only after it has seen barrier
    # i+1 complete, and rank 0 enqueued that barrier only after the image of frame i had arrived.
    pipe = None
    if args.pipeline:
        from concurrent.futures import ThreadPoolExecutor
        N_LANES, LEAD = 3, 2
        lanes = [dict(ctx=ctx, bm=bm, cr=cr, stream=stream, peer=peer, gather=gather, outputs=outputs)]
        for _ in range(N_LANES - 1):
            c_ = g.Context(local, w, h)
            ln = dict(ctx=c_, bm=g.BufferManager(c_), cr=g.ComputeRenderer(c_), stream=torch.cuda.ExternalStream(c_.stream, device=local),
                      peer=None, gather=None, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
            ln["bm"].upload_scene(ps, prepare=True)
            if peer is not None:
                ln["peer"] = tiles.PeerFrame(dist, torch, c_, ln["stream"])
                ln["outputs"] |= ln["peer"].extra_outputs
            elif world > 1:
                ln["gather"] = make_gather()
            if lists is not None:
                c_.set_tile_list(lists[rank])
            lanes.append(ln)

        def lane_upload(ln):
            ln["bm"].mark_all_dirty()
            ln["bm"].upload_scene(ps, prepare=True)  # H2D + host validation + GPU re-layout + LBVH build, waits for them

        def lane_render(ln):
            if path:
                ln["cr"].render_frame(ps.camera, mode=abi.MODE_PATH, spp=wl["spp"], max_depth=wl["depth"], base_seed=1,
                                      outputs=ln["outputs"], tile_stride=stride, tile_phase=phase, traversal=trav, variant=args.variant)
            else:
                ln["cr"].render_frame(ps.camera, outputs=ln["outputs"], tile_stride=stride, tile_phase=phase, traversal=trav)

        def lane_gather(ln):
            if ln["peer"] is not None:
                ln["peer"].finish()
            elif ln["gather"] is not None:
                c_ = ln["ctx"]
                with torch.cuda.stream(ln["stream"]):
                    if lists is not None:
                        ln["gather"].run(lambda dst: c_.pack_tile_list(tiles.WHAT_COMPOSITE, lists[rank], dst.data_ptr()),
                                         lambda src, r: c_.unpack_tile_list(tiles.WHAT_COMPOSITE, lists[r], src.data_ptr()))
                    else:
                        ln["gather"].run(lambda dst: c_.pack_tiles(tiles.WHAT_COMPOSITE, world, rank, dst.data_ptr()),
                                         lambda src, r: c_.unpack_tiles(tiles.WHAT_COMPOSITE, world, r, src.data_ptr()))

        host_t = [torch.empty((h, w, 4), dtype=torch.uint8).pin_memory() for _ in lanes]
        dev_t = [torch.as_tensor(ln["ctx"].composite_cuda_array(), device="cuda:%d" % local) for ln in lanes] if rank == 0 else None
        step_done = [None] * N_LANES   # per lane: rank 0 - the image has arrived; other ranks - the barrier has completed
        render_done = None
        arrived = {}                   # step -> wall clock at which this rank saw it complete
        pool = ThreadPoolExecutor(max_workers=1)
        n_total = args.warmup + args.steps
        barrier(world)
        torch.cuda.synchronize()
        pending = {k: pool.submit(lane_upload, lanes[k % N_LANES]) for k in range(min(LEAD, n_total))}
        for f in list(pending.values()):
            f.result()                                        # the two-step lead exists before the first frame (untimed)
        barrier(world)
        t_begin = time.perf_counter()
        trace = [] if os.environ.get("RTB200_E2E_TRACE") else None  # per step: host time stamps + GPU events (timeline to stderr)
        ms_now = lambda: (time.perf_counter() - t_begin) * 1e3
        if trace is not None:
            ev_base = torch.cuda.Event(enable_timing=True)
            ev_base.record(lanes[0]["stream"])
        for i in range(n_total):
            ln = lanes[i % N_LANES]
            row = [ms_now()] if trace is not None else None
            pending.pop(i).result()                           # step i's scene is resident and re-laid-out
            if row is not None:
                row.append(ms_now())
            if i + LEAD < n_total:
                pending[i + LEAD] = pool.submit(lane_upload, lanes[(i + LEAD) % N_LANES])
            if render_done is not None:
                ln["stream"].wait_event(render_done)          # GPU-side: one frame after the other, no host round trip
            if row is not None:
                ev_start = torch.cuda.Event(enable_timing=True)
                ev_start.record(ln["stream"])
            lane_render(ln)
            render_done = torch.cuda.Event(enable_timing=trace is not None)
            render_done.record(ln["stream"])
            if row is not None:
                row.append(ms_now())
            if i > 0:
                step_done[(i - 1) % N_LANES].synchronize()    # step i-1 is complete (see the order above)
                arrived[i - 1] = time.perf_counter()
            if row is not None:
                row.append(ms_now())
            lane_gather(ln)
            with torch.cuda.stream(ln["stream"]):
                if rank == 0:
                    host_t[i % N_LANES].copy_(dev_t[i % N_LANES], non_blocking=True)
                step_done[i % N_LANES] = torch.cuda.Event(enable_timing=trace is not None)
                step_done[i % N_LANES].record(ln["stream"])
            if row is not None:
                trace.append((row, ev_start, render_done, step_done[i % N_LANES]))
        step_done[(n_total - 1) % N_LANES].synchronize()      # the last image has arrived
        arrived[n_total - 1] = time.perf_counter()
        # K steps in steady state: wall clock from the arrival of the last warm-up step's image to the last image's
        pipe_total = max_over_ranks((arrived[n_total - 1] - (arrived[args.warmup - 1] if args.warmup > 0 else t_begin)) * 1e3, world)
        pool.shutdown()
        for ln in lanes:
            ln["ctx"].synchronize()
        if trace is not None:
            sys.stderr.write("[e2e trace rank %d] step | host: loop start, upload ready, render() returned, step i-1 seen | GPU: frame start, frame end, step complete\n" % rank)
            for i, (row, e_s, e_r, e_d) in enumerate(trace):
                sys.stderr.write("[e2e trace rank %d] %4d | %8.2f %8.2f %8.2f %8.2f | %8.2f %8.2f %8.2f\n" % (
                    rank, i, row[0], row[1], row[2], row[3], ev_base.elapsed_time(e_s), ev_base.elapsed_time(e_r), ev_base.elapsed_time(e_d)))
            del trace
        pipe = {"value": seg_per_frame * args.steps / (pipe_total * 1e-3) / 1e6, "ms_per_step": pipe_total / args.steps,
                "lanes": N_LANES, "uploads_ahead": LEAD,
                "d2h": "asynchronous copy on the lane's stream, waited for one step later",
                "timed": "wall clock between the arrival of the last warm-up image and of the last image (steady state)"}
        # leave the serial lane's image in host_img for the assembled-frame check below
        frame()
        if rank == 0:
            imgs = [host_t[k].numpy() for k in range(min(N_LANES, n_total))]
            pipe["image_equals_serial_lane"] = bool(all(np.array_equal(x, ctx.read_composite()) for x in imgs))
            ctx.read_composite(host_img)
            del imgs
        else:
            ctx.synchronize()
        # torch's pinned-memory allocator records an event on every stream a block was used on when the block is
        # freed: free them while the lanes' streams still exist
        del host_t, dev_t, step_done, render_done
        for ln in lanes[1:]:
            if ln["peer"] is not None:
                ln["peer"].close()
        barrier(world)
        for ln in lanes[1:]:
            ln["ctx"].close()

    # ---- N > 1: the assembled frame on rank 0 must equal rank 0's own render of the whole frame
    assembled_ok = None
    if world > 1 and rank == 0:
        got = host_img.copy()
        ctx.set_tile_list(None)  # the whole frame on this one context
        if path:
            cr.render_frame(ps.camera, mode=abi.MODE_PATH, spp=wl["spp"], max_depth=wl["depth"], base_seed=1,
                            outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE, traversal=trav, variant=args.variant)
        else:
            cr.render_frame(ps.camera, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE, traversal=trav)
        if lists is not None:
            ctx.set_tile_list(lists[0])
        assembled_ok = bool(np.array_equal(got, ctx.read_composite()))
    if peer is not None:
        peer.close()  # ranks != 0 unmap rank 0's frame before anybody frees anything
    barrier(world)
    if rank != 0:
        ctx.close()
        return

    # ---- rank 0, N = 1: the reference-semantics primary frame of the same scene as a sub-record (north_star's
    # ">= 1 Grays/s primary-ray throughput" target), device-timed and end to end like the main figure
    primary = None
    if world == 1 and path and args.workload == "c3":
        pk = dict(outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE, traversal=trav)
        for _ in range(3):
            cr.render_frame(ps.camera, **pk)
        ctx.synchronize()
        p_ms = []
        for _ in range(args.steps):
            with torch.cuda.stream(stream):
                flush.zero_()
            cr.render_frame(ps.camera, **pk)
            ctx.synchronize()
            p_ms.append(ctx.last_kernel_ms)
        p_e2e = []
        for i in range(3 + args.steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            bm.mark_all_dirty()
            bm.upload_scene(ps)
            cr.render_frame(ps.camera, **pk)
            ctx.read_composite(host_img)
            if i >= 3:
                p_e2e.append((time.perf_counter() - t0) * 1e3)
        primary = {"workload": WORKLOADS["c3-primary"]["desc"], "rays_per_frame": w * h,
                   "value": w * h / (float(np.mean(p_ms)) * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": float(np.mean(p_ms)),
                   "e2e": {"value": w * h / (float(np.mean(p_e2e)) * 1e-3) / 1e6, "unit": "Mrays/s",
                           "ms_per_step": float(np.mean(p_e2e)), "h2d_bytes_per_step": h2d_bytes,
                           "d2h_bytes_per_step": int(host_img.nbytes)},
                   "kernel": "rt_trace_shade", "traversal": trav_name(ctx.last_traversal)}
        # leave the path frame in the context's buffers for the parity check below
        render()
        ctx.synchronize()

    # ---- parity at the configured size: the cpu_baseline leg renders a band of rows of this very frame with the
    # literal oracle; its pixels must equal the GPU frame's bit for bit
    cpu_rec, parity = None, None
    if world == 1 and not args.no_cpu_baseline:
        # the leg IS the reference arm, run as a child process (same code, same sample, a process of its own like
        # the driver's reference run), so the two figures are one measurement
        import tempfile
        tmp = os.path.join(tempfile.mkdtemp(), "rows.npz")
        env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
        child = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--workload", args.workload,
                                "--steps", "1", "--warmup", "1", "--ref-seconds", str(args.cpu_seconds), "--dump-rows", tmp],
                               capture_output=True, text=True, env=env)
        if child.returncode != 0:
            raise RuntimeError("cpu_baseline child failed: " + child.stderr[-2000:])
        cpu_rec = json.loads(child.stdout.strip().splitlines()[-1])["cpu_baseline"]
        px = np.load(tmp)
        y0, y1 = (int(x) for x in px["rows"])
        if path:
            acc = ctx.read_accum()[y0:y1]
            ok = bool(np.array_equal(acc.view(np.uint32), px["accum"].view(np.uint32)) and
                      np.array_equal(ctx.read_composite()[y0:y1], px["resolved"]))
        else:
            acc = ctx.read_accum()[y0:y1, :, :3]
            ok = bool(np.array_equal(acc.view(np.uint32), px["radiance"].view(np.uint32)) and
                      np.array_equal(ctx.read_composite()[y0:y1], px["composite"]))
        parity = {"parity_rows_checked": int(y1 - y0), "rows": [int(y0), int(y1)], "bit_exact": ok,
                  "against": "literal-traversal oracle (oracle/rt_oracle.cpp), fp32 accumulation bits + rgba8 pixels"}

    # ---- roofline of the dominant kernel (rank 0)
    n_lights = len(scene.lights)
    seg = max(cnt["segments"], 1)
    V, T, H = cnt["box_tests"] / seg, cnt["tri_tests"] / seg, cnt["shaded"] / seg
    out_bytes_per_pixel = 16 + 4  # float4 accumulation store + rgba8 composite store
    trav_bytes_per_ray = 48 * V + 56 * T + 20 * len(scene.spheres)            # closest hit (48 V + 56 T + 20 S)
    shade_bytes_per_ray = 128 * H + 52 * n_lights * H + out_bytes_per_pixel * my_pixels / seg  # 128 H + 52 L H + O
    bytes_per_ray = trav_bytes_per_ray + shade_bytes_per_ray
    hbm_peak, hbm_src = measured_peaks()
    wave = path and args.variant in (0, 4) and ctx.last_traversal != abi.TRAVERSAL_LITERAL
    if wave and stage_ms:
        # streaming wavefront engine: the dominant kernel is wv_extend (closest hit of every ray segment);
        # its launches are timed by CUDA events between consecutive launches on the context's stream
        ext_ms = float(np.mean([x["extend"] for x in stage_ms]))
        stages = {k: float(np.mean([x[k] for x in stage_ms])) for k in ("generate", "extend", "shade", "resolve")}
        stages["launches_per_step"] = stage_ms[-1]["launches"]
        kernel, mean_ms, kernel_bytes_per_ray = "wv_extend", ext_ms, trav_bytes_per_ray
    else:
        mean_ms = float(np.mean(kernel_ms))
        stages = None
        kernel = ("wave engine frame: wv_extend + wv_shade + wv_resolve on overlapped sub-frame streams"
                  if wave else ("rt_path_trace" if path else "rt_trace_shade"))
        kernel_bytes_per_ray = bytes_per_ray
    algorithmic_gbs = kernel_bytes_per_ray * seg_per_frame_rank / (mean_ms * 1e-3) / 1e9
    # measured denominators of the traversal itself: random independent 32-byte sector gathers
    # (the node-fetch instruction) from an L1-resident, an L2-resident and an HBM-resident set
    gather_bw = {"l1_64KiB": ctx.probe_gather_bandwidth(64 << 10), "l2_32MiB": ctx.probe_gather_bandwidth(32 << 20),
                 "hbm_4GiB": ctx.probe_gather_bandwidth(4 << 30)}
    # bytes the re-laid-out layout actually loads per ray segment: two-child nodes (32 B with 16-bit boxes in the
    # wave engine's rebuilt tree, else 64 B) and 48-B triangles
    node_bytes = 32 if (wave and ctx.last_traversal == abi.TRAVERSAL_REBUILT) else 64
    fetched_per_ray = node_bytes * (V / 2.0) + 48 * T
    fetched_gbs = fetched_per_ray * seg_per_frame_rank / (mean_ms * 1e-3) / 1e9
    # where the traversal's working set lives decides the roof: the re-laid-out scene of C2 / C3 / C4
    # fits the 126 MB L2 (and mostly L1), C5's does not
    # bytes the traversal touches: 48-B triangles + the node array of the policy in use
    scene_bytes = 48 * len(scene.triangles) + (16 * len(scene.triangles) if node_bytes == 32 else 32 * len(scene.bvh_nodes))
    on_chip = scene_bytes < 100e6
    ncu, ncu_src = ncu_summary_for(args.workload) if world == 1 else (None, "multi-GPU run")
    krec = (ncu or {}).get("kernels", {}).get(kernel) if ncu else None
    if on_chip:
        # achieved = node + triangle sectors really fetched per second; peak = this run's measured rate of divergent
        # 32-byte gathers from an L2-resident set (with the capture's L1 hit rate h: the harmonic mix of the L1 and L2 rates)
        peak = gather_bw["l2_32MiB"]
        peak_src = "measured in this run: random 32-B sector gathers (ld.global.nc.v8) over a 32 MiB L2-resident set"
        if krec and krec.get("l1_hit_rate") is not None:
            hit = krec["l1_hit_rate"]
            peak = 1.0 / (hit / gather_bw["l1_64KiB"] + (1.0 - hit) / gather_bw["l2_32MiB"])
            peak_src = ("measured in this run: random 32-B sector gathers, L1-resident (64 KiB) and L2-resident (32 MiB) sets "
                        "mixed by the ncu L1 sector hit rate %.3f of this kernel (%s)" % (hit, ncu_src))
        bound, achieved = "l2", fetched_gbs
    else:
        # HBM-resident scene: the upper tree levels are still served by L1 / L2, the leaves and triangles by HBM at
        # random-sector efficiency.  peak = the three measured gather rates mixed (harmonically) by where this kernel's
        # sectors were served: L1 share = ncu L1 sector hit rate, HBM share = ncu DRAM bytes / fetched bytes, L2 the rest
        bound, achieved, peak = "hbm", fetched_gbs, gather_bw["hbm_4GiB"]
        peak_src = ("measured in this run: random 32-B sector gathers over a 4 GiB HBM-resident set; no ncu capture of this "
                    "build to say which share of the fetches HBM serves, so frac is against the all-from-HBM rate and may exceed 1")
        if krec and krec.get("l1_hit_rate") is not None and krec.get("dram_bytes_per_step"):
            f1 = krec["l1_hit_rate"]
            f3 = min(1.0 - f1, krec["dram_bytes_per_step"] / max(fetched_per_ray * seg_per_frame_rank, 1.0))
            f2 = 1.0 - f1 - f3
            peak = 1.0 / (f1 / gather_bw["l1_64KiB"] + f2 / gather_bw["l2_32MiB"] + f3 / gather_bw["hbm_4GiB"])
            peak_src = ("measured in this run: random 32-B sector gathers from an L1- (64 KiB), L2- (32 MiB) and HBM-resident "
                        "(4 GiB) set, mixed by where this kernel's sectors were served: L1 %.3f (ncu L1 sector hit rate), HBM %.3f "
                        "(ncu DRAM bytes / fetched bytes), L2 %.3f (%s); HBM copy peak %.0f GB/s (%s) for context"
                        % (f1, f3, f2, ncu_src, hbm_peak, hbm_src))
    roof = {"bound": bound, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": krec.get("dram_bytes_per_step") if krec else None,
            "traffic_source": ncu_src,
            "kernel": kernel, "peak_source": peak_src,
            "what": "achieved = node and triangle bytes the kernel fetches per second (%d-B two-child nodes, 48-B triangles; "
                    "V/2 node visits and T triangle tests per ray segment counted by an instrumented launch of the same "
                    "kernels) over its CUDA-event launch time" % node_bytes,
            "algorithmic_gbs": algorithmic_gbs,
            "algorithmic_note": "SURVEY 8d figure in reference-layout bytes (48 V + 56 T + 20 S per segment; megakernels add "
                                "128 H + 52 L H + output stores): context only; it charges 96 B per node pair where the kernel "
                                "loads one 32-B sector, and the scene is served from L1 / L2, so it may exceed the HBM copy peak",
            "hbm_copy_peak_gbs": hbm_peak, "hbm_copy_peak_source": hbm_src,
            "bytes_per_ray": kernel_bytes_per_ray, "step_bytes_per_ray": bytes_per_ray,
            "fetched_bytes_per_ray": fetched_per_ray,
            "box_tests_per_ray": V, "tri_tests_per_ray": T, "kernel_ms": mean_ms,
            "kernel_share_of_step": mean_ms / (total_ms / args.steps) if world == 1 else None,
            "stage_ms": stages,
            "stage_ms_note": "second timed pass, one sub-frame, CUDA events between consecutive launches; the headline "
                             "pass overlaps two sub-frames on two streams" if stages else None,
            "gather_gbs_measured": gather_bw,
            "frac_of_l1_gather": fetched_gbs / gather_bw["l1_64KiB"], "frac_of_l2_gather": fetched_gbs / gather_bw["l2_32MiB"]}
    if krec:
        # the second roof of a divergent traversal: instruction issue.  issue_active x lanes / 32 = useful lane-slots
        roof["issue"] = {"issue_active_pct": krec.get("issue_active_pct"), "lanes_per_inst": krec.get("lanes_per_inst"),
                         "useful_lane_slot_frac": (krec["issue_active_pct"] / 100.0 * krec["lanes_per_inst"] / 32.0)
                         if krec.get("issue_active_pct") and krec.get("lanes_per_inst") else None,
                         "warp_execution_efficiency_pct": (krec["lanes_per_inst"] / 32.0 * 100.0) if krec.get("lanes_per_inst") else None,
                         "alu_pipe_pct": krec.get("alu_pipe_pct"), "fma_pipe_pct": krec.get("fma_pipe_pct"),
                         "source": ncu_src}
    cfg = workload_config(wl, scene)
    line = {
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "run": {"ray_segments_per_frame": int(seg_per_frame), "traversal": trav_name(ctx.last_traversal),
                "engine": ("streaming wavefront (extend / shade per depth, resolve)" if wave else
                           ("megakernel" if path else "fused trace+shade kernel")),
                "literal_fallback_rays": cnt["literal_fallbacks"], "upload_relayout_build_ms": ctx.last_build_ms,
                "tiles_per_rank": my_tiles, "src_hash": source_hash(),
                "tile_balance": ({"steps": balance_log, "tiles_per_rank": [len(l) for l in lists]} if lists is not None else None),
                "parallelism": ("64x64 tile split x%d (%s), scene replicated, %s inside the timed step"
                                % (world, "per-rank tile lists balanced on the warm-up frames' per-rank GPU times" if lists is not None
                                   else "tile i -> rank i % N", "composite pixels stored into rank 0's frame by the render kernels (P2P over NVLink, "
                                          "CUDA IPC) + 4-byte NCCL all-reduce as barrier" if peer is not None else
                                   "NCCL gather of the composite tiles to rank 0")) if world > 1 else "single GPU"},
        "roofline": roof,
        "e2e": ({"value": pipe["value"], "unit": "Mrays/s", "mode": "pipelined", "h2d_bytes_per_step": h2d_bytes,
                 "d2h_bytes_per_step": int(host_img.nbytes), "ms_per_step": pipe["ms_per_step"],
                 "how": "every step: all scene buffers host->device from pinned memory + host validation + GPU re-layout + LBVH "
                        "build + frame + gather + device->host read of the image; pipelined over three contexts per GPU: an "
                        "uploader thread prepares step i+2 while step i renders, frames follow each other through a GPU-side "
                        "event, the image copy is asynchronous and waited for one step later (wall clock between the arrival of "
                        "the last warm-up image and of the last image, max over ranks)",
                 "lanes": pipe.get("lanes"), "image_equals_serial_lane": pipe.get("image_equals_serial_lane")} if (pipe and pipe["value"] >= e2e_value) else
                {"value": e2e_value, "unit": "Mrays/s", "mode": "serial", "h2d_bytes_per_step": h2d_bytes,
                 "d2h_bytes_per_step": int(host_img.nbytes), "ms_per_step": e2e_total / args.steps,
                 "how": "serial: upload, build, frame, read back one after the other in one context (the pipelined "
                        "mode was %s)" % ("not measured" if not pipe else "slower on this workload: %.0f Mrays/s" % pipe["value"])}),
        "e2e_serial": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": e2e_total / args.steps,
                       "how": "the same step with nothing overlapped: upload, build, frame, read-back one after the other in one context"},
        "gpu_launches": int(launches), "clocks": clk, "rank_frame_ms": rank_kernel_ms,
        "assembled_frame_bit_identical_to_single_gpu": assembled_ok,
    }
    if primary:
        line["primary"] = primary
    if cpu_rec:
        line["cpu_baseline"] = cpu_rec
        line.update(parity)
    print(json.dumps(line), flush=True)
    ctx.close()
    if parity and not parity["bit_exact"]:
        print("PARITY FAILURE: GPU frame differs from the oracle on rows %s" % parity["rows"], file=sys.stderr)
        sys.exit(3)


def trav_name(t):
    from gpu_raytracer_b200 import abi
    return {abi.TRAVERSAL_LITERAL: "literal", abi.TRAVERSAL_ORDERED: "ordered (uploaded tree)",
            abi.TRAVERSAL_REBUILT: "rebuilt (internal LBVH, winners certified against the uploaded tree)"}.get(t, str(t))


def main():
    args = parse_args()
    wl = WORKLOADS[args.workload]
    world, rank, local = dist_setup(args)
    if args.impl == "reference":
        run_reference(args, wl, world, rank)
    else:
        run_b200(args, wl, world, rank, local)
    if world > 1 and args.impl != "reference":
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
