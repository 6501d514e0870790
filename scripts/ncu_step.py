"""One short program for ncu (GPU box): upload a workload's scene, render two warm-up frames and `frames` more.
  RTB200_WAVE_SPLIT=1 python scripts/ncu_step.py c3 [frames]
With RTB200_WAVE_SPLIT=1 a path frame is one sub-frame: per depth one wv_extend + two wv_shade launches (one of the
two shade forms returns at once), then wv_resolve."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 1
wl = bench.WORKLOADS[name]
sc = bench.build_scene(wl)
ctx = g.Context(0, wl["w"], wl["h"])
g.BufferManager(ctx).upload_scene(sc)
cr = g.ComputeRenderer(ctx)
for i in range(2 + frames):
    if wl["mode"] == "path":
        cr.render_frame(sc.camera, mode=abi.MODE_PATH, spp=wl["spp"], max_depth=wl["depth"], base_seed=1)
    else:
        cr.render_frame(sc.camera)
    ctx.synchronize()
    print("frame %d: %.3f ms, %d launches so far" % (i, ctx.last_kernel_ms, ctx.launch_count), flush=True)
print("src_hash", bench.source_hash())
ctx.close()
