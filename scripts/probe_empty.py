"""Per-launch overhead of the wave engine: camera looks at the sky, so depth >= 1 queues are empty."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi
os.environ["RTB200_WAVE_SPLIT"] = "1"
sc = g.terrain_scene(1000, 500); w, h = 1920, 1080
cam = sc.camera.copy(); cam["direction"] = [0, 1, 0]; cam["up"] = [0, 0, -1]
ctx = g.Context(0, w, h); g.BufferManager(ctx).upload_scene(sc); cr = g.ComputeRenderer(ctx)
for rep in range(4):
    cr.render_frame(cam, mode=abi.MODE_PATH, spp=2, max_depth=4, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
    ctx.synchronize()
print("sky frame %.3f ms, %d segments" % (ctx.last_kernel_ms, ctx.last_segments), ctx.last_stage_ms())
