"""Emulates one rank's share of the C3 frame on a single GPU: python scripts/probe_rank.py STRIDE PHASE [SPP]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi
stride, phase = int(sys.argv[1]), int(sys.argv[2])
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 16
sc = g.terrain_scene(1000, 500); w, h = 1920, 1080
ctx = g.Context(0, w, h); g.BufferManager(ctx).upload_scene(sc); cr = g.ComputeRenderer(ctx)
best = 1e9
for rep in range(5):
    cr.render_frame(sc.camera, mode=abi.MODE_PATH, spp=spp, max_depth=4, tile_stride=stride, tile_phase=phase,
                    outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
    ctx.synchronize(); best = min(best, ctx.last_kernel_ms)
print("stride %d phase %d spp %d split %s: %.3f ms, %d segments, %.1f Mrays/s" %
      (stride, phase, spp, os.environ.get("RTB200_WAVE_SPLIT", "default"), best, ctx.last_segments, ctx.last_segments / best / 1e3), flush=True)
ctx.count_next_launch()
cr.render_frame(sc.camera, mode=abi.MODE_PATH, spp=spp, max_depth=4, tile_stride=stride, tile_phase=phase, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
cnt = ctx.read_counters()
print("   literal fallbacks %d, V=%.1f T=%.2f" % (cnt["literal_fallbacks"], cnt["box_tests"] / max(1, cnt["segments"]), cnt["tri_tests"] / max(1, cnt["segments"])), flush=True)
cr.render_frame(sc.camera, mode=abi.MODE_PATH, spp=spp, max_depth=4, tile_stride=stride, tile_phase=phase, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
ctx.synchronize()
if os.environ.get("RTB200_WAVE_SPLIT") == "1":
    print("   stages:", {k: (round(v, 3) if isinstance(v, float) else v) for k, v in ctx.last_stage_ms().items()}, flush=True)
    print("   launches:", " ".join("%s=%.3f" % (n[:3], t) for n, t in ctx.last_launch_table()), flush=True)
