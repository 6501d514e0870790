import sys; sys.path.insert(0,'/root/repo')
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi
import os
for name,sc,(w,h),spp,d in (("cornell",g.cornell_scene(),(1024,1024),16,8),("c3",g.terrain_scene(1000,500),(1920,1080),4,4)):
    ctx=g.Context(0,w,h); g.BufferManager(ctx).upload_scene(sc); cr=g.ComputeRenderer(ctx)
    for v in ((0,) if os.environ.get('RTB200_LIB') else (1,)):
        for rep in range(2):
            cr.render_frame(sc.camera,mode=abi.MODE_PATH,spp=spp,max_depth=d,variant=v); ctx.synchronize()
        ms=ctx.last_kernel_ms
        ctx.count_next_launch(); cr.render_frame(sc.camera,mode=abi.MODE_PATH,spp=spp,max_depth=d,variant=v); c=ctx.read_counters()
        print(os.environ.get('RTB200_LIB','default')[-24:],name,'variant',v,'ms',round(ms,2),'Mrays/s',round(c['segments']/ms/1e3,1),{k:c[k] for k in ('s_phases','box_tests','tri_tests','literal_fallbacks')})
