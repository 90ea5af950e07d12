"""Development aid (CPU only): how often is the closest accepted triangle of the whole scene (plain closest-hit search of the
scene-wide culling hierarchy, emu_closest_global) NOT what the reference octree walk returns?  Camera rays and rays from random
origins inside the scene.  Feeds the round-2 design note in profiles/r01_experiments.md.
  python tools/closest_vs_reference.py c4 c2 c3a c3b"""
import ctypes as C, os, sys
os.environ["RRT_GLOBAL_CULL"]="1"
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0,ROOT); sys.path.insert(0,os.path.join(ROOT,'tests'))
import numpy as np, bench
import test_host_emulation as T
L=T._load_emu(T._build_emu()); vp=C.c_void_p
L.emu_closest_global.argtypes=[vp,C.c_uint64,vp,vp,vp]
for name in sys.argv[1:]:
    desc,W,H=bench.WORKLOADS[name]; sd=bench.load_workload(name); es=T.EmuScene(L,sd,parallel=True)
    rng=np.random.default_rng(1); n=400000
    x=rng.uniform(-W/2,W/2,n)/W; y=rng.uniform(-H/2,H/2,n)/H
    o=np.tile(np.array([0.0,2.0,-10.0]),(n,1)); d=np.stack([x,y,np.ones(n)],1)
    # second half: rays from random surface-ish points in random directions (shadow / mirror like)
    o2=rng.uniform(-9,9,(n,3))+np.array([0,2,4.0]); d2=rng.normal(size=(n,3))
    for lab,(oo,dd) in (("camera",(o,d)),("random",(o2,d2))):
        oo=np.ascontiguousarray(oo); dd=np.ascontiguousarray(dd)
        ref,_=es.trace(oo,dd)
        g=np.zeros(n,np.int32); assert L.emu_closest_global(es.h,n,oo.ctypes.data,dd.ctypes.data,g.ctypes.data)==0
        bad=(ref!=g)
        print(f"{name} {lab}: {n} rays, {(ref>=0).mean():.1%} hit; closest-accepted differs from the reference walk on {bad.sum()} rays ({bad.mean():.2e}); of these reference None: {(bad&(ref<0)).sum()}, closest None: {(bad&(g<0)).sum()}")
    es.close()
