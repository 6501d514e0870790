import sys, os; sys.path.insert(0,'/root/repo')
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi
v=int(os.environ.get('VARIANT','0')); spp=int(os.environ.get('SPP','2'))
sc=g.terrain_scene(1000,500); w,h=1920,1080
ctx=g.Context(0,w,h); g.BufferManager(ctx).upload_scene(sc); cr=g.ComputeRenderer(ctx)
for rep in range(2):
    cr.render_frame(sc.camera,mode=abi.MODE_PATH,spp=spp,max_depth=4,variant=v); ctx.synchronize()
print('variant',v,'ms',ctx.last_kernel_ms,'segs',ctx.last_segments)
