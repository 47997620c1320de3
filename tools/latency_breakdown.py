import sys, time; sys.path.insert(0,'/root/repo')
import numpy as np
import orbslam2commentedbyxcm_b200 as pkg
from orbslam2commentedbyxcm_b200 import synth
img=synth.gen_frame((480,640),0)
ex=pkg.ORBextractor(1000,1.2,8,20,7)
out=ex.alloc_outputs(1,480,640)
for _ in range(20): ex.extract_batch(img[None],out=out)
ex.set_profiling(True)
acc={}
for _ in range(50):
    ex.extract_batch(img[None],out=out)
    for k,v in ex.stage_times().items(): acc[k]=acc.get(k,0)+v/50
print({k:round(v*1e3,1) for k,v in acc.items()}, 'us; sum', round(sum(acc.values())*1e3,1))
ex.set_profiling(False)
for g in (True, False):
    ex.set_use_graph(g)
    for _ in range(20): ex.extract_batch(img[None],out=out)
    t=[]
    for _ in range(300):
        t0=time.perf_counter(); ex.extract_batch(img[None],out=out); t.append(time.perf_counter()-t0)
    print('graph',g,'median us',round(np.median(t)*1e6,1),'p10',round(np.percentile(t,10)*1e6,1))
