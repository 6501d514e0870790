import sys; sys.path.insert(0,'/root/repo')
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi
for name,sc,(w,h),spp,d in (("cornell",g.cornell_scene(),(512,512),16,8),("c3",g.terrain_scene(1000,500),(960,540),4,4)):
    ctx=g.Context(0,w,h); g.BufferManager(ctx).upload_scene(sc); cr=g.ComputeRenderer(ctx)
    for v in (0,1):
        for rep in range(2):
            cr.render_frame(sc.camera,mode=abi.MODE_PATH,spp=spp,max_depth=d,variant=v); ctx.synchronize()
        ms=ctx.last_kernel_ms
        ctx.count_next_launch(); cr.render_frame(sc.camera,mode=abi.MODE_PATH,spp=spp,max_depth=d,variant=v); c=ctx.read_counters()
        print(name,'variant',v,'ms',round(ms,2),'Mrays/s',round(c['segments']/ms/1e3,1),c)
