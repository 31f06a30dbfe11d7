"""Diagnostics (numpy, CPU): which objects of the 10,000-sphere frame are refreshed on which march trip, and how many
divergent SDF passes the register large-scene kernel needs for them under different object orderings
(profiles/r01_v9_large_scene_analysis.txt).  The march is restated here in numpy; no GPU, no oracle."""
import numpy as np, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracing_engine_b200 import scene, camera
from raytracing_engine_b200.engine import host_level_dims
f32=np.float32
sph=scene.synthetic_spheres(10000, seed=0xB200)
W,H=1920,1080
n=8; dims=host_level_dims(n,W,H)
rots,poss=camera.static_path(1)
ratio=scene.view_ratio(W,H)
RD=f32(1000.0)
rs=np.random.RandomState(1)
C=sph[:,:3].astype(f32); R=sph[:,3].astype(f32)
def sdf(p): 
    d=p[None,:]-C
    return np.sqrt((d[:,0]*d[:,0]+d[:,1]*d[:,1])+d[:,2]*d[:,2]).astype(f32)-R
def march(origin,dirv,thr):
    d=sdf(origin); ln=f32(0); last=f32(0); events=[]
    while ln<RD:
        p=origin+dirv*ln; radius=(ln+f32(1))*thr
        d=d-last
        due=np.nonzero(d<=radius)[0]
        if len(due):
            dd=p[None,:]-C[due]
            d[due]=np.sqrt((dd[:,0]*dd[:,0]+dd[:,1]*dd[:,1])+dd[:,2]*dd[:,2]).astype(f32)-R[due]
        events.append(due)
        dist=min(RD,d.min()); last=max(dist,f32(0)); ln=f32(ln+last)
        if dist<=radius: ln=f32(ln-radius); break
    return ln,events
def ray(level,x,y,start):
    Wl,Hl=dims[level]
    isx=f32(2.0**(n-1-level))/f32(W); isy=f32(2.0**(n-1-level))/f32(H)
    ncx=(f32(2*x+1)*isx-f32(1))*f32(ratio[0]); ncy=(f32(2*y+1)*isy-f32(1))*f32(ratio[1])
    v=np.array([ncx,1,ncy],f32); v=v/np.sqrt((v*v).sum()).astype(f32)
    thr=f32(1.4142135)*f32(8)*isx
    origin=np.array(poss[0],f32)+v*f32(start)
    ln,ev=march(origin,v,thr)
    return max(f32(start+ln),f32(0)),ev
# sample chains
samples=[]
t=time.time()
NS=int(sys.argv[1]) if len(sys.argv)>1 else 60
per_level={l:[] for l in range(n)}
for s in range(NS):
    X=int(rs.randint(0,W)); Y=int(rs.randint(0,H))
    start=f32(1.0)
    for l in range(n):
        x=X>>(n-1-l); y=Y>>(n-1-l)
        val,ev=ray(l,x,y,start)
        per_level[l].append(ev)
        start=val
print("sim time",time.time()-t)
# orderings
cam=np.array(poss[0],f32)
def morton(P,bits=10):
    q=((P-P.min(0))/(P.max(0)-P.min(0)+1e-9)*(2**bits-1)).astype(np.uint64)
    code=np.zeros(len(P),np.uint64)
    for b in range(bits):
        for k in range(3):
            code|=((q[:,k]>>np.uint64(b))&np.uint64(1))<<np.uint64(3*b+k)
    return code
orders={
 "original":np.arange(10000),
 "camdist":np.argsort(np.sqrt(((C-cam)**2).sum(1))-R,kind='stable'),
 "morton":np.argsort(morton(C),kind='stable'),
}
for name,perm in orders.items():
    pos=np.empty(10000,np.int64); pos[perm]=np.arange(10000)   # object -> position
    print("==",name)
    for l in range(n):
        due=0; passes=0; trips=0; grp=0
        for ev in per_level[l]:
            for e in ev:
                trips+=1
                if len(e):
                    p=pos[e]; due+=len(e); passes+=len(np.unique(p//32))
                    # group entries: (warp, group-of-8-registers): site p//32 -> (j=p//256, w=(p%256)//32) ; group = (j//8, w)
                    j=p//256; w=(p%256)//32
                    grp+=len(np.unique((j//8)*8+w))
        print(f" level {l}: rays {len(per_level[l])} trips/ray {trips/len(per_level[l]):.1f} due/trip {due/max(trips,1):.1f} passes/trip {passes/max(trips,1):.1f} lanes/pass {due/max(passes,1):.2f} group-entries/trip {grp/max(trips,1):.1f} (of 40)")
