import sys, time; sys.path.insert(0,'/root/repo')
import numpy as np
import of_dis_b200 as dis
from oracle.bindings import Port, OrcParams
from of_dis_b200.synth import make_pair
port = Port()
print(dis.lib().disb200_version(), 'devices', dis.device_count())

def to_orc(p):
    q = OrcParams()
    for f,_ in OrcParams._fields_:
        src = {'psz':'patchsz','noc':'channels'}.get(f,f)
        setattr(q,f,getattr(p,src))
    return q

def epe(a,b): 
    d = np.sqrt(((a-b)**2).sum(-1)); return d.mean(), d.max()

def run(mode,noc,op,h,w,**kw):
    A,B,gt = make_pair(h,w,channels=noc,mode='OF' if mode==1 else 'DE',seed=0)
    p = dis.operating_point(op,w,mode,noc).copy(verbosity=0,**kw)
    q = to_orc(p)
    ref_full, ref_low = port.run_dense(q,A,B,want_lowres=True)
    with dis.Handle(p,w,h,max_batch=2) as hd:
        out = hd.run_host(np.stack([A,A]),np.stack([B,B]))
        # pyramid parity
        ok=True
        for l in range(p.sc_l,p.sc_f+1):
            for which,img in ((0,A),(1,B)):
                g = hd.fetch_pyramid(which,l,pair=1)
                o = port.pyramid_level(img,p.sc_f,l,p.patchsz)
                for a_,b_ in zip(g,o): ok &= np.array_equal(a_,b_)
        low = hd.fetch_flow(0,p.sc_l,pair=0)
        m1 = epe(low,ref_low); m2 = epe(out[0],ref_full); m3 = epe(out[1],ref_full)
        print(f'mode{mode} noc{noc} op{op} {kw}: pyramid exact={ok}; lowres EPE mean {m1[0]:.2e} max {m1[1]:.2e}; full mean {m2[0]:.2e} max {m2[1]:.2e}; pair1 {m3[0]:.2e}; EPE-GT gpu {epe(out[0],gt)[0]:.3f} cpu {epe(ref_full,gt)[0]:.3f}; launches {hd.launch_count()}', flush=True)

# varref-only parity
def tv(mode,noc,h,w,level):
    A,B,gt = make_pair(h,w,channels=noc,mode='OF' if mode==1 else 'DE',seed=1)
    p = dis.operating_point(2,w,mode,noc).copy(verbosity=0, sc_f=level, sc_l=level)
    q = to_orc(p)
    Wp,Hp,pl,pt = port.pad_dims(w,h,level)
    a = port.pyramid_level(A,level,level,p.patchsz)[0]; b = port.pyramid_level(B,level,level,p.patchsz)[0]
    wl,hl = Wp>>level, Hp>>level
    rng = np.random.default_rng(0)
    nop = 2 if mode==1 else 1
    fl = (rng.standard_normal((hl,wl,nop))*0.7).astype(np.float32)
    if mode==2: fl = -np.abs(fl)
    r = port.varref_level(q,a,b,wl,hl,level,0,fl)
    g = dis.varref_level(p,a,b,wl,hl,level,0,fl)
    print(f'varref mode{mode} noc{noc} {wl}x{hl} L{level}: equal={np.array_equal(r,g)} maxdiff {np.abs(r-g).max():.3e} (change {np.abs(r-fl).max():.3f})', flush=True)

tv(1,1,436,1024,3); tv(1,3,436,1024,3); tv(2,1,436,1024,3); tv(2,3,436,1024,3); tv(1,1,218,500,1)
run(1,1,2,436,1024)
run(1,1,1,436,1024)
run(1,1,2,436,1024,costfct=1)
run(1,1,2,436,1024,costfct=2)
run(1,3,2,436,1024)
run(2,1,2,720,1280)
run(2,3,2,720,1280)
run(1,1,2,436,1024,usefbcon=1)
run(1,1,3,436,1024)
run(1,3,3,436,1024)
run(2,1,3,720,1280,usefbcon=1)
