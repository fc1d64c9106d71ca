import sys, json, subprocess, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import tinypathtracer_b200 as tpt
from tinypathtracer_b200.scene_io import load_scene_npz
from pathlib import Path
REF=Path('oracle/_ref')
ctx=tpt.Context(0)
def ours(scene,W,H,spp,seed,depth=8):
    s=load_scene_npz(f'tests/golden/scenes/{scene}.npz'); ds=tpt.DeviceScene(ctx,s); pt=tpt.PathTracer(ctx,W,H)
    fb=np.zeros((H,W,4),np.uint8); pt.doTrace(ds,tpt.Camera.from_scene(s),fb,spp,max_depth=depth,seed=seed); ds.close(); return fb[...,:3].astype(float)/255
def ref(scene,W,H,spp):
    out='/tmp/r.bgra'
    subprocess.run([str(REF/'tpt_ref_headless'),'--scene',str(REF/'input'/f'{scene}.gltf'),'--env',str(REF/'black.png'),'--width',str(W),'--height',str(H),'--spp',str(spp),'--frames','1','--warmup','0','--out',out],capture_output=True)
    return np.fromfile(out,np.uint8).reshape(H,W,4)[...,:3].astype(float)/255
rm=lambda a,b: ((a-b)**2).sum()/(b**2).sum()
W,H=192,108
for scene,spps in [('box',[1024,4096,16384]),('tir',[4096,16384,65536])]:
    for spp in spps:
        a=ours(scene,W,H,spp,4); b=ours(scene,W,H,spp,5); r=ref(scene,W,H,spp); 
        import time; time.sleep(1.1); r2=ref(scene,W,H,spp)
        print(scene,spp,'ours-ours %.3e ours-ref %.3e ref-ref %.3e means %.5f %.5f'%(rm(a,b),rm(a,r),rm(r,r2),a.mean(),r.mean()))
