import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for sub in ('', 'tests', 'tools'):
    sys.path.insert(0, os.path.join(ROOT, sub))
import gpu_ray_tracing_b200 as rt, scenes
from dev_bench import time_variant
W,H=3840,2160
out=[]
ms,rays,_=time_variant(scenes.random_scene(1024), scenes.camera(W,H,30,500), W,H,4,10); out.append(f"r1024 4K d30 {ms:.3f}")
ms,rays,_=time_variant(scenes.default_scene(), scenes.camera(1280,720,30,500), 1280,720,4,30); out.append(f"default 720p d30 {ms:.3f}")
ms,rays,_=time_variant(scenes.default_scene(), scenes.camera(W,H,30,500), W,H,4,10); out.append(f"default 4K d30 {ms:.3f}")
ms,rays,_=time_variant(scenes.random_scene(1024), scenes.camera(W,H,8,500), W,H,4,10); out.append(f"r1024 4K d8 {ms:.3f}")
print(os.path.basename(os.environ.get("RT_B200_LIB","default")), " | ".join(out))
