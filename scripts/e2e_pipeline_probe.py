"""Timeline of bench.py's double-buffered e2e loop at N = 1 (GPU box): host and GPU time stamps of every step, for the
shipped order (host waits for the frame, copy in flight) and the run-ahead order (GPU-side event between frames).
  python scripts/e2e_pipeline_probe.py c3 [steps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from concurrent.futures import ThreadPoolExecutor
import numpy as np
import torch
import bench
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wl = bench.WORKLOADS[name]
sc = bench.build_scene(wl)
keep = []
def pin(a):
    a = np.ascontiguousarray(a)
    if not a.size:
        return a
    t = torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).pin_memory(); keep.append(t)
    return t.numpy().view(a.dtype)
ps = g.Scene(sc.camera, pin(sc.spheres), pin(sc.triangles), pin(sc.vertices), pin(sc.materials), pin(sc.lights), pin(sc.bvh_nodes), pin(sc.triangle_indices))
w, h = wl["w"], wl["h"]
N_LANES = 3
lanes = []
for k in range(N_LANES):
    c = g.Context(0, w, h)
    lanes.append(dict(ctx=c, bm=g.BufferManager(c), cr=g.ComputeRenderer(c), stream=torch.cuda.ExternalStream(c.stream, device=0)))
    lanes[-1]["bm"].upload_scene(ps, prepare=True)
host_t = [torch.empty((h, w, 4), dtype=torch.uint8).pin_memory() for _ in lanes]
dev_t = [torch.as_tensor(ln["ctx"].composite_cuda_array(), device="cuda:0") for ln in lanes]

def render(ln):
    ln["cr"].render_frame(ps.camera, mode=abi.MODE_PATH, spp=wl["spp"], max_depth=wl["depth"], base_seed=1, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)

def run(n_lanes, ahead):
    """ahead = 0: the host waits for every frame (uploads one step ahead, two lanes);
    ahead = 1: uploads two steps ahead over three lanes, frames chained by a GPU-side event, the host only waits for the
    previous step's image after it has enqueued this step's frame."""
    T0 = [0.0]
    now = lambda: (time.perf_counter() - T0[0]) * 1e3
    ulog = {}
    def upload(i, ln):
        a = now()
        ln["bm"].mark_all_dirty()
        ln["bm"].update_scene_metadata(ps.spheres, ps.lights, ps.bvh_nodes, ps.triangle_indices, ps.vertices)
        b = now()
        ln["bm"].update_triangles(ps.triangles); ln["bm"].update_materials(ps.materials)
        ln["bm"].update_textures(ps.textures); ln["bm"].update_texture_data(ps.texture_data)
        c = now()
        ln["bm"].prepare_scene()
        ulog[i] = (a, b, c, now())
    pool = ThreadPoolExecutor(max_workers=1)
    warm = 4
    n_total = steps + warm
    depth = 2 if ahead else 1  # uploads in flight ahead of the frame
    read_done, frame_done, rows, seen = [None] * n_lanes, None, [], {}
    base = torch.cuda.Event(enable_timing=True)
    gpu_ev = []
    torch.cuda.synchronize()
    T0[0] = time.perf_counter()
    base.record(lanes[0]["stream"])
    pending = {k: pool.submit(upload, k, lanes[k % n_lanes]) for k in range(depth)}
    for f in list(pending.values()):
        f.result()  # prime the pipeline: with a frame running beside it an upload takes one frame time, so the lead has to exist before the first frame
    for i in range(n_total):
        ln = lanes[i % n_lanes]
        r = [i, now()]
        pending.pop(i).result(); r.append(now())
        if i + depth < n_total:
            pending[i + depth] = pool.submit(upload, i + depth, lanes[(i + depth) % n_lanes])
        if ahead and frame_done is not None:
            ln["stream"].wait_event(frame_done)
        e0 = torch.cuda.Event(enable_timing=True); e0.record(ln["stream"])
        render(ln); r.append(now())
        frame_done = torch.cuda.Event(enable_timing=True); frame_done.record(ln["stream"])
        if i > 0:
            read_done[(i - 1) % n_lanes].synchronize()
            seen[i - 1] = now()
        r.append(now())
        if not ahead:
            frame_done.synchronize()
        r.append(now())
        with torch.cuda.stream(ln["stream"]):
            host_t[i % n_lanes].copy_(dev_t[i % n_lanes], non_blocking=True)
            read_done[i % n_lanes] = torch.cuda.Event(enable_timing=True); read_done[i % n_lanes].record(ln["stream"])
        gpu_ev.append((e0, frame_done, read_done[i % n_lanes]))
        rows.append(r)
    read_done[(n_total - 1) % n_lanes].synchronize()
    seen[n_total - 1] = now()
    pool.shutdown()
    per_step = (seen[n_total - 1] - seen[warm - 1]) / steps  # wall clock between the arrivals of image warm-1 and the last image
    print("== %d lanes, %s: %.3f ms per step over %d steps (wall clock between image arrivals)" % (n_lanes, "run-ahead (GPU-side event between frames, uploads two steps ahead)" if ahead else "host waits for the frame", per_step, steps))
    print("   step | host: loop start, upload ready, render() returned, prev image seen, frame seen | GPU: frame start, frame end, copy end | upload thread: start, metadata, rest, prepared")
    for r, ev in zip(rows, gpu_ev):
        i = r[0]
        gs, ge, gc = base.elapsed_time(ev[0]), base.elapsed_time(ev[1]), base.elapsed_time(ev[2])
        u = ulog.get(i, (0, 0, 0, 0))
        print("   %4d | %8.2f %8.2f %8.2f %8.2f %8.2f | %8.2f %8.2f %8.2f | %8.2f %8.2f %8.2f %8.2f" % (i, r[1], r[2], r[3], r[4], r[5], gs, ge, gc, *u))
    imgs = [t.numpy().copy() for t in host_t[:n_lanes]]
    print("   images of all lanes identical:", all(np.array_equal(imgs[0], x) for x in imgs[1:]))

run(2, 0)
run(3, 1)
del host_t, dev_t
for ln in lanes:
    ln["ctx"].close()
