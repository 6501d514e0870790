"""Launch table (CUDA events between launches, one sub-frame) of one path frame of a bench workload:
  RTB200_WAVE_SPLIT=1 python scripts/probe_launches.py field"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("RTB200_WAVE_SPLIT", "1")
import bench
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi
wl = bench.WORKLOADS[sys.argv[1]]
sc = bench.build_scene(wl)
ctx = g.Context(0, wl["w"], wl["h"]); g.BufferManager(ctx).upload_scene(sc); cr = g.ComputeRenderer(ctx)
for _ in range(3):
    cr.render_frame(sc.camera, mode=abi.MODE_PATH, spp=wl["spp"], max_depth=wl["depth"], base_seed=1)
    ctx.synchronize()
print("%s: %.3f ms, %d segments" % (sys.argv[1], ctx.last_kernel_ms, ctx.last_segments))
print("   launches:", " ".join("%s=%.3f" % (n[:3], t) for n, t in ctx.last_launch_table()))
ctx.count_next_launch()
cr.render_frame(sc.camera, mode=abi.MODE_PATH, spp=wl["spp"], max_depth=wl["depth"], base_seed=1)
c = ctx.read_counters(); print("   fallbacks", c["literal_fallbacks"], "V %.1f T %.2f" % (c["box_tests"] / c["segments"], c["tri_tests"] / c["segments"]))
wc = ctx.last_wavefront_counters(); print("   rays per depth", list(wc["rays_per_bounce"]))
