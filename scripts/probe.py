"""Tuning probe (GPU box): times path / primary workloads for one library build.
  RTB200_LIB=build/tune/x.so python scripts/probe.py c3:0 c3:2 c3:3 c4:0 c3p c2"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi

SCENES = {}
def scene(name):
    if name not in SCENES:
        SCENES[name] = {"c3": lambda: (g.terrain_scene(1000, 500), 1920, 1080, 4, 4),
                        "c3p": lambda: (g.terrain_scene(1000, 500), 1920, 1080, 0, 0),
                        "c5": lambda: (g.terrain_scene(2500, 2000), 3840, 2160, 1, 4),
                        "field": lambda: (g.field_scene(), 1920, 1080, 4, 4),
                        "fieldsl": lambda: (g.field_scene(strategy=abi.BVH_SMALL_LEAF), 1920, 1080, 4, 4),
                        "c2": lambda: (g.terrain_scene(250, 200), 1920, 1080, 0, 0),
                        "c4": lambda: (g.cornell_scene(), 1024, 1024, 16, 8)}[name]()
    return SCENES[name]

tag = os.path.basename(os.environ.get("RTB200_LIB", "default"))
for arg in sys.argv[1:]:
    name, _, v = arg.partition(":")
    v = int(v or 0)
    sc, w, h, spp, d = scene(name)
    ctx = g.Context(0, w, h); g.BufferManager(ctx).upload_scene(sc); cr = g.ComputeRenderer(ctx)
    def run():
        if spp: cr.render_frame(sc.camera, mode=abi.MODE_PATH, spp=spp, max_depth=d, variant=v)
        else: cr.render_frame(sc.camera)
        ctx.synchronize()
    best = 1e9
    for rep in range(4):
        run(); best = min(best, ctx.last_kernel_ms)
    ctx.count_next_launch(); run(); c = ctx.read_counters()
    seg = c["segments"]
    print("%-16s %-4s v%d  %8.3f ms  %8.1f Mrays/s  V=%.1f T=%.2f fb=%d build=%.2fms" %
          (tag, name, v, best, seg / best / 1e3, c["box_tests"] / seg, c["tri_tests"] / seg, c["literal_fallbacks"], ctx.last_build_ms), flush=True)
    ctx.close()
