import numpy as np, time, sys
sys.path.insert(0,'.')
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from oracle import oracle as O
from tests.parity import s16_stats
ctx=A.Context(0)
rng=np.random.default_rng(0)
for (w,h) in [(127,61),(640,480),(1920,1080)]:
    g=rng.integers(0,256,(h,w,3)).astype(np.uint8)
    src=(rng.random((h,w))*255).astype(np.float32)
    got=A.fastGlobalSmootherFilter(g,src,8000.0,1.5,ctx=ctx); ref=O.fgs_filter(g,src,8000.0,1.5)
    print("fgs",w,h,"maxerr",np.abs(got-ref).max(), "nan", np.isnan(got).sum())
    ch,cv=A.fgs_weights(g,1.5,ctx=ctx); oh,ov=O.fgs_weights(g,1.5); print(" weights eq",(ch==oh).mean(),(cv==ov).mean())
view,dl,dr=synth.make_textured_scene(1920,1080,3,seed=42,max_disp=160)
f=A.createDisparityWLSFilterGeneric(True,ctx=ctx); f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(2)
t=time.time(); got=f.filter(dl,view,None,dr,(160,0,1760,1080)); print("gpu host call s",time.time()-t)
t=time.time(); got=f.filter(dl,view,None,dr,(160,0,1760,1080)); print("gpu host call s",time.time()-t)
t=time.time(); ref=O.wls_filter(dl,view,dr,roi=(160,0,1760,1080),lam=8000.0,sigma=1.5,disc_radius=2,want_mask=True,nstripes=O.hardware_threads(),threaded=True); print("oracle s",time.time()-t)
print("conf maxdiff",np.abs(f.getConfidenceMap()-ref['conf']).max())
print("mask eq",(f.lrc_mask(dl,dr,ROI=(160,0,1760,1080))==ref['mask']).all())
print(s16_stats(got,ref['out'],(160,0,1760,1080)))
