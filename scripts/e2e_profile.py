"""Where the end-to-end step goes (GPU box): wall clock of every C-ABI call of one e2e step, host buffers pinned.
  python scripts/e2e_profile.py c3 [reps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
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
ctx = g.Context(0, wl["w"], wl["h"]); bm = g.BufferManager(ctx); cr = g.ComputeRenderer(ctx)
img = torch.empty((wl["h"], wl["w"], 4), dtype=torch.uint8).pin_memory().numpy()
def render():
    if wl["mode"] == "path":
        cr.render_frame(ps.camera, mode=abi.MODE_PATH, spp=wl["spp"], max_depth=wl["depth"], base_seed=1)
    else:
        cr.render_frame(ps.camera)
acc = {}
def t(label, fn):
    t0 = time.perf_counter(); r = fn(); acc.setdefault(label, []).append((time.perf_counter() - t0) * 1e3); return r
for i in range(reps + 2):
    if i == 2: acc.clear()
    t0 = time.perf_counter()
    bm.mark_all_dirty()
    t("update_scene_metadata", lambda: bm.update_scene_metadata(ps.spheres, ps.lights, ps.bvh_nodes, ps.triangle_indices, ps.vertices))
    t("update_triangles", lambda: bm.update_triangles(ps.triangles))
    t("update_materials", lambda: bm.update_materials(ps.materials))
    t("update_textures+data", lambda: (bm.update_textures(ps.textures), bm.update_texture_data(ps.texture_data)))
    t("render_frame (call returns)", render)
    t("synchronize (frame done)", ctx.synchronize)
    t("read_composite", lambda: ctx.read_composite(img))
    acc.setdefault("step total", []).append((time.perf_counter() - t0) * 1e3)
    acc.setdefault("  of which: relayout+build (GPU ms)", []).append(ctx.last_build_ms)
    acc.setdefault("  of which: frame kernels (GPU ms)", []).append(ctx.last_kernel_ms)
print("e2e profile %s (%d reps, mean ms / min ms)" % (name, reps))
for k, v in acc.items():
    print("  %-40s %8.3f %8.3f" % (k, float(np.mean(v)), float(np.min(v))))
ctx.close()
